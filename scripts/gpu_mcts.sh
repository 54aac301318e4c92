#!/bin/bash
# MCTS iteration: parity tests, TAFL_MCTS_MINB sweep on the bench config, one ncu capture of k_mcts_sims
mkdir -p gpurun_out
TAG=${1:-m}
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for mb in ${MINBS:-4 6 8 10}; do
  echo "== TAFL_MCTS_MINB=$mb"
  TAFL_MCTS_MINB=$mb timeout 300 python bench.py --no-cpu-baseline --steps 4 --warmup 3 --plies 8 | tee gpurun_out/mcts_${TAG}_minb_$mb.json | python -c "import json,sys; d=json.load(sys.stdin); m=d['mcts']; print('mcts sims/s %.4g ms/move %.2f depth %.2f'%(m['value'], m['ms_per_move'], m['mean_depth']))"
done
if [ "${2:-ncu}" = "ncu" ]; then
SHORT="python bench.py --steps 2 --warmup 3 --plies 4 --no-cpu-baseline --mcts-games 1024 --mcts-sims 200"
timeout 600 $SHORT > gpurun_out/plain_mcts_$TAG.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:k_mcts_sims -s 1 -c 1 -o gpurun_out/prof_mcts_$TAG -f $SHORT > gpurun_out/ncu_mcts_$TAG.log 2>&1
echo "ncu mcts exit $?"
fi
