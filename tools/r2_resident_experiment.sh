#!/bin/bash
# How fast is the step when an SM's working set fits its L1 (would state on
# chip pay)?  The same kernel at 16 / 32 / 64 / 111 envs per SM: time per launch, then ncu
# --set full of one launch at 16 envs per SM (working set 16 x ~16 KB = L1-resident) and at
# the bench shape.
for N in 2368 4736 9472 16384; do
  python bench.py --task cheetah-run --nenv $N --no-extra --no-cpu --steps 20 --warmup 5 2>/dev/null \
    | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('nenv', $N, 'kernel_ms %.4f' % d['roofline']['kernel_ms'], 'env-steps/s %.3e' % d['value'])"
done > gpurun_out/r2_resident_sweep.txt
cat gpurun_out/r2_resident_sweep.txt
for N in 2368 16384; do
  ncu --set full --import-source on --clock-control none -k regex:bp_step_kernel --launch-skip 205 --launch-count 1 \
    -f -o gpurun_out/prof_r2_cheetah_n$N python bench.py --task cheetah-run --nenv $N --steps 3 --warmup 3 --no-extra --no-cpu > gpurun_out/ncu_r2_n$N.log 2>&1
  python tools/ncu_summary.py gpurun_out/prof_r2_cheetah_n$N.ncu-rep > gpurun_out/r2_cheetah_n$N.summary.txt
done
