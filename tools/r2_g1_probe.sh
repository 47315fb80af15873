python bench.py --task cheetah-run --lanes 1 --no-extra --no-cpu --steps 20 --warmup 5 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('G=1 kernel_ms %.4f' % d['roofline']['kernel_ms'], 'env-steps/s %.3e' % d['value'])"
ncu --set full --import-source on --clock-control none -k regex:bp_step_kernel --launch-skip 205 --launch-count 1 -f -o gpurun_out/prof_r2_cheetah_g1 python bench.py --task cheetah-run --lanes 1 --steps 3 --warmup 3 --no-extra --no-cpu > gpurun_out/ncu_r2_g1.log 2>&1
python tools/ncu_summary.py gpurun_out/prof_r2_cheetah_g1.ncu-rep > gpurun_out/r2_cheetah_g1.summary.txt
cat gpurun_out/r2_cheetah_g1.summary.txt
