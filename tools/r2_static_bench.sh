#!/bin/bash
# kernel-only time of the static (lanes = 1) instances; usage: r2_static_bench.sh [tasks...]
for t in ${@:-cheetah-run walker-walk humanoid-run}; do
  python bench.py --task $t --lanes 1 --no-extra --no-cpu --steps 20 --warmup 5 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$t static G=1 kernel_ms %.4f' % d['roofline']['kernel_ms'], 'env-steps/s %.3e e2e %.3e' % (d['value'], d['e2e']['value']))"
done
