#!/bin/bash
# launch list + one full capture of the evaluator-hook round kernel at bench size (4096 trees), short search
set -u
mkdir -p gpurun_out
TAG=${1:-hook}
CMD="python scripts/hook_target.py"
timeout 300 $CMD > gpurun_out/plain_$TAG.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 20 -c 300 --csv --log-file gpurun_out/launches_$TAG.csv $CMD > gpurun_out/ncu_launches_$TAG.log 2>&1
echo "launch list exit $?"
timeout 900 ncu --set full --clock-control none --cache-control none --import-source on -k "regex:k_mcts_hook_round" -s 150 -c 2 -o gpurun_out/prof_hook_$TAG -f $CMD > gpurun_out/ncu_hook_$TAG.log 2>&1
echo "ncu hook exit $?"
python profiles/summarize_launches.py gpurun_out/launches_$TAG.csv
