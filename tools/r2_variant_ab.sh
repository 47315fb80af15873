#!/bin/bash
# A/B of the default build against variants/libbphys_base.so on one box
run() { python bench.py --task $1 --steps 30 --warmup 5 --no-extra --no-cpu 2>/dev/null | tail -1 | \
  python -c "import json,sys; r=json.loads(sys.stdin.read()); print('$1 lib=${BPHYS_LIB_PATH:-default} value %.4e e2e %.4e ms %.4f' % (r['value'], r['e2e']['value'], r['ms_per_step']))"; }
for T in ${TASKS:-cheetah-run humanoid-run walker-walk quadruped-escape finger-spin}; do
  export BPHYS_LIB_PATH=$PWD/variants/libbphys_base.so; run $T; unset BPHYS_LIB_PATH; run $T
done
