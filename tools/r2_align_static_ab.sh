#!/bin/bash
# 128-byte field alignment with the static instances: variants/libbphys_align1.so (packed layout) vs the default build
run() { python bench.py --task $1 --steps 30 --warmup 5 --no-extra --no-cpu 2>/dev/null | tail -1 | \
  python -c "import json,sys; r=json.loads(sys.stdin.read()); print('$1 lib=${BPHYS_LIB_PATH:-default} value %.4e e2e %.4e ms %.4f stride %d' % (r['value'], r['e2e']['value'], r['ms_per_step'], r['config']['smem_bytes_per_env']//8))"; }
for T in cheetah-run humanoid-run walker-walk quadruped-escape finger-spin; do
  for V in align1 alignnat; do export BPHYS_LIB_PATH=$PWD/variants/libbphys_$V.so; run $T; done; unset BPHYS_LIB_PATH; run $T
done
