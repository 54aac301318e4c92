#!/bin/bash
# A/B of the CTA-chained step graph (TAFL_STEP_CHAIN: 0 off, 1 by batch size, 2 always) against whole-grid dependencies
set -u
mkdir -p gpurun_out
TAG=${1:-c1}
OUT=gpurun_out/chain_$TAG.txt
: > $OUT
run() { label=$1; shift; env "$@" timeout 300 python scripts/step_ab.py "$label" >> $OUT 2>gpurun_out/chain_err_$TAG.txt || { echo "$label FAILED" >> $OUT; tail -5 gpurun_out/chain_err_$TAG.txt >> $OUT; }; }
for n in ${SIZES:-16384 24576 32768 40960 131072 196608}; do
  run "chained, $n games" TAFL_STEP_CHAIN=2 AB_GAMES=$n
  run "whole-grid, $n games" TAFL_STEP_CHAIN=0 AB_GAMES=$n
done
run "default, 32768 games" AB_GAMES=32768
run "default, 65536 games" AB_GAMES=65536
cat $OUT
