#!/bin/bash
# One gpurun call: GPU parity tests, the bench line, the ncu launch list and one full capture of
# each dominant kernel.  Outputs land in gpurun_out/ (copied into profiles/ by hand afterwards).
set -u
mkdir -p gpurun_out
TAG=${1:-r01}
nvidia-smi --query-gpu=index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active --format=csv -lms 500 > gpurun_out/clocks_$TAG.csv &
SMI=$!
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_$TAG.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/pytest_gpu_$TAG.log
tail -3 gpurun_out/pytest_gpu_$TAG.log
timeout 900 python bench.py > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; echo "bench exit $?"
cat gpurun_out/bench_$TAG.json
timeout 600 python bench.py --impl reference --steps 20 --warmup 3 > gpurun_out/bench_ref_$TAG.json 2>&1; echo "ref exit $?"
kill $SMI
SHORT="python bench.py --steps 2 --warmup 3 --plies 8 --no-cpu-baseline --mcts-games 1024 --mcts-sims 200"
timeout 600 $SHORT > gpurun_out/plain_$TAG.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_$TAG.csv $SHORT > gpurun_out/ncu_launches_$TAG.log 2>&1
echo "launch list exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_step_random -s 20 -c 2 -o gpurun_out/prof_step_$TAG -f $SHORT > gpurun_out/ncu_step_$TAG.log 2>&1
echo "ncu step exit $?"
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:k_mcts_sims -c 1 -o gpurun_out/prof_mcts_$TAG -f $SHORT > gpurun_out/ncu_mcts_$TAG.log 2>&1
echo "ncu mcts exit $?"
ls -la gpurun_out
