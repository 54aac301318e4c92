#!/bin/bash
# One gpurun call of round 2: GPU parity tests, the bench line (+ reference arm), the ncu launch list and one
# full capture of each hot kernel.  Outputs land in gpurun_out/ (summarised into profiles/ afterwards).
#   scripts/gpu_round2.sh <tag> [tests|notests] [bench|nobench] [ncu kernels...]
set -u
mkdir -p gpurun_out
TAG=${1:-r02a}; DO_TESTS=${2:-tests}; DO_BENCH=${3:-bench}; shift 3 || true
NCU_WHAT="$*"
nvidia-smi --query-gpu=index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap --format=csv -lms 200 > gpurun_out/clocks_$TAG.csv &
SMI=$!
if [ "$DO_TESTS" = "tests" ]; then
  timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_$TAG.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/pytest_gpu_$TAG.log
  tail -5 gpurun_out/pytest_gpu_$TAG.log
fi
if [ "$DO_BENCH" = "bench" ]; then
  timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; echo "bench exit $?"
  tail -c 600 gpurun_out/bench_$TAG.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/bench_$TAG.json"))
    m=d.get("mcts",{}); t=d.get("mcts_tablut",{}); lm=d.get("legal_moves",{})
    print("env %.4g steps/s kernel_us %.2f e2e %.4g | legal %.1f us (2048: %.1f us) frac %.3f | mcts %s e2e %s hook %s reuse %s | tablut %s" % (
        d["value"], d["roofline"]["kernel_us"], d["e2e"]["value"], lm.get("us_per_launch",0), lm.get("us_per_launch_2048_games",0), lm.get("roofline",{}).get("frac",0),
        m.get("value"), (m.get("e2e") or {}).get("value"), (m.get("evaluator_hook") or {}).get("value"), (m.get("tree_reuse") or {}).get("value"), t.get("value")))
    print("cpu", d.get("cpu_baseline",{}).get("value"), "c1", d.get("c1_brandubh_cpu",{}).get("value"), "mcts cpu", (m.get("cpu_baseline") or {}).get("value"), "clocks", d.get("clocks"))
    if "error" in m: print("MCTS ERROR", m["error"])
    if "error" in t: print("TABLUT ERROR", t["error"])
except Exception as ex:
    print("no bench line:", ex)
PY
  timeout 600 python bench.py --impl reference --steps 20 --warmup 3 > gpurun_out/bench_ref_$TAG.json 2> gpurun_out/bench_ref_$TAG.err; echo "ref exit $?"
  cat gpurun_out/bench_ref_$TAG.json | cut -c1-300
fi
kill $SMI
if [ -n "$NCU_WHAT" ]; then
  CMD="python scripts/ncu_target.py $NCU_WHAT"
  timeout 600 $CMD > gpurun_out/plain_$TAG.log 2>&1 &&
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_$TAG.csv $CMD > gpurun_out/ncu_launches_$TAG.log 2>&1
  echo "launch list exit $?"
  for K in $NCU_WHAT; do
    case $K in
      step) PAT=k_step_random; SKIP=3; CNT=2;;
      legal) PAT=k_legal_moves_g; SKIP=1; CNT=2;;
      mcts) PAT=k_mcts_sims_g; SKIP=0; CNT=1;;
      hook) PAT="k_mcts_hook_round"; SKIP=300; CNT=2;;
    esac
    timeout 1500 ncu --set full --clock-control none --import-source on -k "regex:$PAT" -s $SKIP -c $CNT -o gpurun_out/prof_${K}_$TAG -f $CMD > gpurun_out/ncu_${K}_$TAG.log 2>&1
    echo "ncu $K exit $?"
  done
fi
ls -la gpurun_out | tail -30
