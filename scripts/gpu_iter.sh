#!/bin/bash
# quick iteration call: GPU parity tests, bench (no CPU leg), steady-state ncu capture of the step kernel
set -u
mkdir -p gpurun_out
TAG=${1:-it}
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_$TAG.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/pytest_gpu_$TAG.log
timeout 900 python bench.py --no-cpu-baseline > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; echo "bench exit $?"
python - <<PY
import json
d=json.load(open("gpurun_out/bench_$TAG.json"))
print("env %.4g steps/s kernel_us %.2f e2e %.4g mcts %s"%(d["value"],d["roofline"]["kernel_us"],d["e2e"]["value"],d.get("mcts",{}).get("value")))
PY
if [ "${2:-ncu}" = "ncu" ]; then
SHORT="python bench.py --steps 8 --warmup 3 --plies 64 --no-cpu-baseline --no-mcts"
timeout 600 $SHORT > gpurun_out/plain_$TAG.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_step_random -s 600 -c 2 -o gpurun_out/prof_step_$TAG -f $SHORT > gpurun_out/ncu_step_$TAG.log 2>&1
echo "ncu step exit $?"
fi
