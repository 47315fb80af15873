#!/bin/bash
# Throughput of each execution shape (lanes per env) on the headline workloads.
out=gpurun_out/sweep_lanes.jsonl
: > $out
for task in cheetah-run humanoid-run; do
  for lanes in ${BP_LANES:-32 16 8 4 2 1}; do
    timeout 300 python bench.py --task $task --lanes $lanes --steps ${STEPS:-10} --warmup 3 --no-extra --no-cpu 2>&1 | tail -1 >> $out
  done
done
python - <<'PY'
import json
for l in open('gpurun_out/sweep_lanes.jsonl'):
    try:
        d = json.loads(l)
        print(d['metric'], 'lanes', d['config'].get('lanes_per_env'), 'value %.3e' % d['value'], 'e2e %.3e' % d['e2e']['value'], 'kernel_ms %.3f' % d['roofline']['kernel_ms'], d['warnings'], d['stats'])
    except Exception as e:
        print('bad line', l[:200])
PY
