#!/bin/bash
# contact-free config [1] at 4096 envs: generic vs static instances at 1 / 8 lanes per env
for T in cartpole-swingup acrobot-swingup pendulum-swingup; do
  for L in 1 8; do
    for NS in 1 0; do
      if [ $NS = 1 ]; then export BPHYS_NO_STATIC=1; else unset BPHYS_NO_STATIC; fi
      python bench.py --task $T --lanes $L --steps 50 --warmup 5 --no-extra --no-cpu 2>/dev/null | tail -1 | \
        python -c "import json,sys; r=json.loads(sys.stdin.read()); print('$T lanes $L no_static $NS value %.3e e2e %.3e ms %.4f' % (r['value'], r['e2e']['value'], r['ms_per_step']))"
    done
  done
done
