run() { python bench.py --task $1 --steps 30 --warmup 5 --no-extra --no-cpu 2>/dev/null | tail -1 | \
  python -c "import json,sys; r=json.loads(sys.stdin.read()); print('$1 lib=$2 value %.4e e2e %.4e ms %.4f' % (r['value'], r['e2e']['value'], r['ms_per_step']))"; }
for i in 1 2; do
unset BPHYS_LIB_PATH; run cheetah-run default
for V in $(ls variants | sed 's/libbphys_//; s/.so//'); do export BPHYS_LIB_PATH=$PWD/variants/libbphys_$V.so; run cheetah-run $V; done
done
