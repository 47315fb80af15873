#!/bin/bash
# generic LEAN / MID kernels with inlined stages (this build) against variants/libbphys_base.so
export BPHYS_NO_STATIC=1
run() { python bench.py --task $1 --steps 20 --warmup 5 --no-extra --no-cpu 2>/dev/null | tail -1 | \
  python -c "import json,sys; r=json.loads(sys.stdin.read()); print('$1 generic lib=$2 value %.4e e2e %.4e ms %.4f' % (r['value'], r['e2e']['value'], r['ms_per_step']))"; }
for T in cheetah-run humanoid-run walker-walk quadruped-walk quadruped-escape; do
export BPHYS_LIB_PATH=$PWD/variants/libbphys_base.so; run $T base; unset BPHYS_LIB_PATH; run $T inlined
done
