#!/bin/bash
# usage: ncu_profile.sh TASK LANES NAME [SKIP]  -> gpurun_out/prof_NAME.ncu-rep (one step-kernel launch, --set full)
TASK=$1; LANES=$2; NAME=$3; SKIP=${4:-205}
ncu --set full --import-source on --clock-control none -k regex:bp_step_kernel --launch-skip $SKIP --launch-count 1 \
    -f -o gpurun_out/prof_$NAME python bench.py --task $TASK --lanes $LANES --steps 3 --warmup 3 --no-extra --no-cpu > gpurun_out/ncu_$NAME.log 2>&1
tail -2 gpurun_out/ncu_$NAME.log
