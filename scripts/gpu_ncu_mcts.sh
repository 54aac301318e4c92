#!/bin/bash
# one full ncu capture of the dominant MCTS kernel at the bench configuration (4096 trees x 800 sims)
mkdir -p gpurun_out
TAG=${1:-mcts}
CMD="python bench.py --steps 2 --warmup 3 --plies 8 --no-cpu-baseline"
timeout 600 $CMD > gpurun_out/plain_$TAG.log 2>&1 || exit 1
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:k_mcts_sims -s 1 -c 1 -o gpurun_out/prof_mcts_$TAG -f $CMD > gpurun_out/ncu_mcts_$TAG.log 2>&1
echo "ncu exit $?"; tail -3 gpurun_out/ncu_mcts_$TAG.log
