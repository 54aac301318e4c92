#!/bin/bash
# Time the step kernel with one block of a ply switched off at a time (libraries built with -DTAFL_ABL_* and
# --suffix=_abl_*): what each block costs in time, which the instruction counts do not tell.  Results are wrong by
# construction (a rule is missing); only the µs per ply are read.
set -u
mkdir -p gpurun_out
TAG=${1:-abl}
D=alphazeroforhnefatafl_b200
OUT=gpurun_out/ablate_$TAG.txt
: > $OUT
run() { label=$1; shift; env "$@" timeout 300 python scripts/step_ab.py "$label" >> $OUT 2>gpurun_out/ablate_err_$TAG.txt || echo "$label FAILED" >> $OUT; }
run "baseline" TAFL_STEP_SINGLE=1
for f in $D/libtafl_b200_abl_*.so; do
  [ -f "$f" ] || continue
  v=${f##*_abl_}; v=${v%.so}
  run "without $v" TAFL_LIB=$PWD/$f
done
run "baseline again" TAFL_STEP_SINGLE=1
cat $OUT
