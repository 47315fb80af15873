#!/bin/bash
# field / image alignment experiment (generic kernels: the static instances are built for one layout)
export BPHYS_NO_STATIC=1
run() { python bench.py --task $1 --steps 30 --warmup 5 --no-extra --no-cpu 2>/dev/null | tail -1 | \
  python -c "import json,sys; r=json.loads(sys.stdin.read()); print('$1 ALIGN=${BPHYS_ALIGN:-1} value %.4e ms %.4f stride %d' % (r['value'], r['ms_per_step'], r['config']['smem_bytes_per_env']//8))"; }
for T in cheetah-run humanoid-run; do
  unset BPHYS_ALIGN; run $T
  for A in 2 4 16; do export BPHYS_ALIGN=$A; run $T; done
done
