#!/bin/bash
# parity + speed of k_mcts_sims_g for every group size (TAFL_MCTS_G = lanes per tree; 0 = k_mcts_sims)
mkdir -p gpurun_out
for g in 16 8 32; do
  echo "== TAFL_MCTS_G=$g parity"
  TAFL_MCTS_G=$g timeout 900 python -m pytest tests/test_gpu_mcts.py tests/test_mcts_reference_golden.py tests/test_mcts_eval_golden.py -m gpu -x -q 2>&1 | tail -3
done
for g in 0 32 16 8; do
  TAFL_MCTS_G=$g timeout 600 python bench.py --no-cpu-baseline --steps 10 --warmup 3 | python -c "import json,sys; d=json.load(sys.stdin)['mcts']; print('G=$g', 'sims/s %.4g' % d['value'], 'ms/move %.2f' % d['ms_per_move'], 'reuse %.4g' % d['tree_reuse']['value'])"
done
