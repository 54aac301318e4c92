"""tafl_gather_records (C ABI, NCCL by dlopen): a one-rank communicator inside pytest, and — when the box
has at least two GPUs — scripts/nccl_gather_check.py under torch.distributed.run with two ranks on ragged
shards: the C-ABI gather against the torch.distributed gather row for row, and SelfPlay's side-stream
overlap path against both."""
import numpy as np
import pytest

from alphazeroforhnefatafl_b200 import _abi as A

pytestmark = pytest.mark.gpu


def test_single_rank_gather(orc, built_lib):
    import torch
    from alphazeroforhnefatafl_b200 import engine as E
    from alphazeroforhnefatafl_b200 import selfplay as SP
    from parity_cases import board_for
    rules, fen = orc.preset_rules(A.PRESET_BRANDUBH), board_for(orc, A.PRESET_BRANDUBH)
    n = 24
    with E.LockstepGames(rules, fen, n, 0, 64) as g:
        g.set_game_id_base(700)
        sp = SP.SelfPlay(g, num_sims=60)
        rows = sp.move(gather=False)
        torch.cuda.synchronize()
        assert rows.shape[0] == n                                        # every game was Ongoing: nothing compacted away
        ng = SP.NcclGather(g, SP.NcclGather.unique_id(), 1, 0)
        try:
            got, counts = ng.gather(rows.contiguous())
            assert list(counts) == [rows.shape[0]]
            assert torch.equal(got.cpu(), rows.cpu())
            recs = SP.records_to_numpy(got)
            assert list(recs["h"]["game_id"]) == list(range(700, 700 + n))
            empty, counts = ng.gather(rows[:0].contiguous())          # a rank may have nothing to send
            assert list(counts) == [0] and empty.shape[0] == 0
            with pytest.raises(E.TaflError):                           # root buffer too small
                ng.gather(rows.contiguous(), recv=torch.empty((3, A.REC_STRIDE), dtype=torch.uint8, device="cuda"))
        finally:
            ng.close()


def test_two_rank_gather_on_ragged_shards(built_lib):
    import os
    import socket
    import subprocess
    import sys
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs (gpurun --gpus 2)")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
                        "--master-port", str(port), os.path.join(root, "scripts", "nccl_gather_check.py")],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "NCCL GATHER CHECK OK" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]
