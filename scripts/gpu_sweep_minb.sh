#!/bin/bash
# tuning sweep: register cap / residency of k_step_random (TAFL_STEP_MINB), env-step bench only
mkdir -p gpurun_out
for mb in 4 6 8 10 12; do
  echo "== TAFL_STEP_MINB=$mb"
  TAFL_STEP_MINB=$mb timeout 300 python bench.py --no-mcts --no-cpu-baseline --steps 100 --warmup 5 | tee gpurun_out/sweep_minb_$mb.json | python -c "import json,sys; d=json.load(sys.stdin); print(d['value'], d['roofline']['kernel_us'], d['e2e']['value'])"
done
