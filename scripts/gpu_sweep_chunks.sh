#!/bin/bash
# tuning sweep: chunk count of the pipelined host-buffer path (TAFL_HOST_CHUNKS), e2e number only
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for ch in 1 2 4 8 16 32; do
  echo "== TAFL_HOST_CHUNKS=$ch"
  TAFL_HOST_CHUNKS=$ch timeout 300 python bench.py --no-mcts --no-cpu-baseline --steps 200 --warmup 3 --plies 4 | tee gpurun_out/sweep_chunks_$ch.json | python -c "import json,sys; d=json.load(sys.stdin); print('e2e', d['e2e']['value'], 'env', d['value'])"
done
