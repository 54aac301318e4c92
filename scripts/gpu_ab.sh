#!/bin/bash
# A/B of the step-kernel variants (libraries built with build.py --suffix) and knobs, one line each.
set -u
mkdir -p gpurun_out
TAG=${1:-ab}
D=alphazeroforhnefatafl_b200
OUT=gpurun_out/ab_$TAG.txt
: > $OUT
run() { label=$1; shift; env "$@" timeout 300 python scripts/step_ab.py "$label" >> $OUT 2>gpurun_out/ab_err_$TAG.txt || echo "$label FAILED" >> $OUT; }
run "v1 lib, multi-ply kernel"        TAFL_LIB=$PWD/$D/libtafl_b200_v1.so TAFL_STEP_SINGLE=0
run "v1 lib, single-ply kernel"       TAFL_LIB=$PWD/$D/libtafl_b200_v1.so TAFL_STEP_SINGLE=1
run "new prefilter only, single"      TAFL_LIB=$PWD/$D/libtafl_b200_cntold.so TAFL_STEP_SINGLE=1
run "new count only, single"          TAFL_LIB=$PWD/$D/libtafl_b200_swold.so TAFL_STEP_SINGLE=1
run "new lib, multi-ply kernel"       TAFL_STEP_SINGLE=0
run "new lib, single-ply kernel"      TAFL_STEP_SINGLE=1
run "new lib, single, LM_G=16"        TAFL_STEP_SINGLE=1 TAFL_LM_G=16
run "new lib, single, LM_G=1"         TAFL_STEP_SINGLE=1 TAFL_LM_G=1
run "new lib, 2048 games"             AB_GAMES=2048
run "new lib, 2048 games LM_G=16"     AB_GAMES=2048 TAFL_LM_G=16
cat $OUT
