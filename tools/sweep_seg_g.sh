#!/bin/bash
# usage: sweep_seg_g.sh TASK BASE_G  -> one bench line per (segment, lane width) with the others at BASE_G
TASK=$1; BASE=$2
run() { BPHYS_SPLIT=1 BPHYS_SEG_G=$1 python bench.py --task $TASK --lanes $BASE --steps 10 --warmup 3 --no-extra --no-cpu 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$1', round(d['value']), d['ms_per_step'], d['warnings'])"; }
base=$(printf "$BASE,%.0s" 1 2 3 4 5 6 7 8 9); base=${base%,}
run $base
for k in 0 1 2 3 4 5 6 7 8; do
  for g in 4 8 16; do
    [ $g = $BASE ] && continue
    spec=$(echo $base | awk -F, -v k=$k -v g=$g 'BEGIN{OFS=","} {$(k+1)=g; print}')
    run $spec
  done
done
