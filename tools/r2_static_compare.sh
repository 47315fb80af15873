#!/bin/bash
# kernel-only time with and without the static instances, same box, back to back
for t in ${@:-cheetah-run walker-walk humanoid-run quadruped-escape finger-spin}; do
  for v in 0 1; do
    if [ $v = 1 ]; then export BPHYS_NO_STATIC=1; else unset BPHYS_NO_STATIC; fi
    python bench.py --task $t --no-extra --no-cpu --steps 20 --warmup 5 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$t no_static=$v kernel_ms %.4f' % d['roofline']['kernel_ms'], 'env-steps/s %.4e e2e %.4e' % (d['value'], d['e2e']['value']))"
  done
done
