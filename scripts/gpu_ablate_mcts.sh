#!/bin/bash
# the same timing-only ablation (scripts/gpu_ablate.sh) for the fused MCTS search and the evaluator hook
set -u
mkdir -p gpurun_out
TAG=${1:-m1}
D=alphazeroforhnefatafl_b200
OUT=gpurun_out/ablate_mcts_$TAG.txt
: > $OUT
run() { label=$1; shift; env "$@" timeout 300 python scripts/mcts_ab.py "$label" >> $OUT 2>gpurun_out/ablate_mcts_err_$TAG.txt || echo "$label FAILED" >> $OUT; }
run "baseline" X=1
for f in $D/libtafl_b200_abl_*.so; do
  [ -f "$f" ] || continue
  v=${f##*_abl_}; v=${v%.so}
  run "without $v" TAFL_LIB=$PWD/$f
done
run "baseline again" X=1
run "hook baseline" AB_HOOK=1
cat $OUT
