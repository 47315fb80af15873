#!/bin/bash
# static instances at narrower lane widths (make EXPERIMENTS=1): cheetah G = 8 / 4 / 2, humanoid G = 16 / 8
run() { python bench.py --task $1 --lanes $2 --steps 30 --warmup 5 --no-extra --no-cpu 2>/dev/null | tail -1 | \
  python -c "import json,sys; r=json.loads(sys.stdin.read()); print('$1 lanes $2 value %.4e e2e %.4e ms %.4f' % (r['value'], r['e2e']['value'], r['ms_per_step']))"; }
run cheetah-run 8; run cheetah-run 4; run cheetah-run 2; run cheetah-run 1
run humanoid-run 16; run humanoid-run 8
