#!/bin/bash
out=gpurun_out/sweep_tpb.jsonl; : > $out
for tpb in ${TPBS:-128 224 256 448 512}; do
  for spec in ${SPECS:-cheetah-run|4 humanoid-run|8 walker-walk|4}; do
    set -- ${spec//|/ }
    echo "{\"lib\": \"tpb$tpb\", \"spec\": \"$spec\"}" >> $out
    BPHYS_BLOCK_THREADS=$tpb python bench.py --task $1 --lanes $2 --steps 10 --warmup 3 --no-extra --no-cpu >> $out 2>> gpurun_out/sweep_tpb.err
  done
done
python tools/show_sweep.py $out
