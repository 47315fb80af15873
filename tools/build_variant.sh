#!/bin/bash
# usage: build_variant.sh NAME "extra nvcc flags"  -> variants/libbphys_NAME.so (experiments)
set -e
cd "$(dirname "$0")/../dm_control_b200/csrc"
NAME=$1; FLAGS=$2
OBJ=_obj_$NAME; mkdir -p $OBJ ../../variants
for s in bp_step_g32 bp_step_g1 bp_step_g2 bp_step_g4 bp_step_g8 bp_step_g16 bp_step_g4_lean bp_step_g8_lean bp_step_g16_lean bp_task bp_api; do
  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC -Xptxas -v $FLAGS -c $s.cu -o $OBJ/$s.o 2> $OBJ/$s.ptxas.log &
done
wait
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../../variants/libbphys_$NAME.so $OBJ/*.o -lcudart
grep -h "Used" $OBJ/bp_step_g4.ptxas.log $OBJ/bp_step_g8.ptxas.log | tail -2
