#!/bin/bash
# runs bench (no cpu baseline) for each variant library and task; output jsonl
out=gpurun_out/sweep_variants.jsonl; : > $out
for lib in "" $(ls variants/*.so 2>/dev/null); do
  for spec in ${SPECS:-cheetah-run|0|4 humanoid-run|0|8 walker-walk|0|4 cartpole-swingup|0|2}; do
    set -- ${spec//|/ }
    extra=""; [ "$2" != "0" ] && extra="--nenv $2"
    echo "{\"lib\": \"$lib\", \"spec\": \"$spec\"}" >> $out
    if [ -n "$lib" ]; then export BPHYS_LIB_PATH=$PWD/$lib; else unset BPHYS_LIB_PATH; fi
    python bench.py --task $1 --lanes $3 $extra --steps 10 --warmup 3 --no-extra --no-cpu >> $out 2>> gpurun_out/sweep_variants.err
  done
done
