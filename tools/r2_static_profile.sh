#!/bin/bash
# ncu --set full of one launch of a static instance + per-function / per-line tables
T=${1:-cheetah-run}; N=${2:-r2_static_cheetah}
ncu --set full --import-source on --clock-control none -k regex:bp_step_kernel --launch-skip 205 --launch-count 1 \
    -f -o gpurun_out/prof_$N python bench.py --task $T --lanes 1 --steps 3 --warmup 3 --no-extra --no-cpu > gpurun_out/ncu_$N.log 2>&1
python tools/ncu_summary.py gpurun_out/prof_$N.ncu-rep > gpurun_out/$N.summary.txt
python tools/ncu_by_function.py gpurun_out/prof_$N.ncu-rep dm_control_b200/csrc/_obj/bp_step_static_${3:-cheetah}.o ${3:-cheetah} > gpurun_out/$N.by_function.txt
python tools/ncu_by_line.py gpurun_out/prof_$N.ncu-rep > gpurun_out/$N.by_line.txt 2>/dev/null
cat gpurun_out/$N.summary.txt; head -40 gpurun_out/$N.by_function.txt
