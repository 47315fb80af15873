#!/bin/bash
# Round-2 evidence run on one B200: bench (both arms), launch list, ncu --set full of the
# two timed kernels.  Every ncu pass follows a plain run of the same command that exited 0.
R=${1:-r2p}
mkdir -p gpurun_out
python bench.py 2> gpurun_out/${R}_bench.err | tail -1 > gpurun_out/${R}_bench.json || exit 1
python bench.py --impl reference --steps 5 --warmup 3 2>> gpurun_out/${R}_bench.err | tail -1 > gpurun_out/${R}_bench_reference_arm.json
python bench.py --steps 3 --warmup 3 --no-extra --no-cpu > /dev/null 2>&1 || exit 2
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${R}_launches.csv \
    python bench.py --steps 3 --warmup 3 --no-extra --no-cpu > gpurun_out/ncu_${R}_launches.log 2>&1
python tools/launch_summary.py gpurun_out/${R}_launches.csv > gpurun_out/${R}_launches_summary.txt
for spec in ${SPECS:-"cheetah-run cheetah_g8 cheetah 205" "humanoid-run humanoid_g16 humanoid 103"}; do
  set -- $spec
  python bench.py --task $1 --steps 3 --warmup 3 --no-extra --no-cpu > /dev/null 2>&1 || exit 3
  N=${R}_static_$2
  ncu --set full --import-source on --clock-control none -k regex:bp_step_kernel --launch-skip $4 --launch-count 1 \
      -f -o gpurun_out/prof_$N python bench.py --task $1 --steps 3 --warmup 3 --no-extra --no-cpu > gpurun_out/ncu_$N.log 2>&1
  python tools/ncu_summary.py gpurun_out/prof_$N.ncu-rep > gpurun_out/$N.summary.txt
  python tools/ncu_by_function.py gpurun_out/prof_$N.ncu-rep dm_control_b200/csrc/_obj/bp_static_$2.o $3 > gpurun_out/$N.by_function.txt 2>/dev/null
  python tools/ncu_by_line.py gpurun_out/prof_$N.ncu-rep > gpurun_out/$N.by_line.txt 2>/dev/null
  rm -f gpurun_out/prof_$N.ncu-rep
done
ls -la gpurun_out | tail -20
cat gpurun_out/${R}_bench.json | head -c 1500
