#!/bin/bash
# A/B of the step-kernel variants (libraries built with build.py --suffix) and knobs, one line each.
set -u
mkdir -p gpurun_out
TAG=${1:-ab}
D=alphazeroforhnefatafl_b200
OUT=gpurun_out/ab_$TAG.txt
: > $OUT
run() { label=$1; shift; env "$@" timeout 300 python scripts/step_ab.py "$label" >> $OUT 2>gpurun_out/ab_err_$TAG.txt || echo "$label FAILED" >> $OUT; }
[ -f $D/libtafl_b200_v1.so ] && run "r01 count + index prefilter, multi-ply" TAFL_LIB=$PWD/$D/libtafl_b200_v1.so TAFL_STEP_SINGLE=0
run "new lib (64-thread CTAs)"        TAFL_STEP_SINGLE=1
run "new lib, multi-ply kernel"       TAFL_STEP_SINGLE=0
[ -f $D/libtafl_b200_b32.so ] && run "32-thread CTAs"   TAFL_LIB=$PWD/$D/libtafl_b200_b32.so
[ -f $D/libtafl_b200_b128.so ] && run "128-thread CTAs" TAFL_LIB=$PWD/$D/libtafl_b200_b128.so
run "new lib, no graph"               TAFL_STEP_GRAPH=0
run "new lib, graph without PDL"      TAFL_STEP_PDL=0
run "new lib, 2048 games"             AB_GAMES=2048
cat $OUT
