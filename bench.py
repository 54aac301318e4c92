#!/usr/bin/env python
"""bench.py — Copenhagen 11x11 env steps/s (BASELINE.json configs[1]: 65 536 lockstep games of
random legal play on one B200) plus MCTS simulations/s, through libtafl_b200.so.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

One bench "step" = `--plies` lockstep plies of all resident games (default 64 launches of the
fused step kernel, the state round-trips HBM after every ply); the L2 is flushed between timed
steps.  Prints ONE JSON line (see DESIGN.md §measurement for every field).
"""
import argparse
import ctypes as C
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

ENV_BYTES_PER_STEP = 136          # 64 B state read + 64 B state write + 4 B action + 4 B status
METRIC, UNIT = "copenhagen_11x11_env_steps_per_s", "env steps/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--games", type=int, default=65536)
    ap.add_argument("--plies", type=int, default=64, help="lockstep plies per bench step")
    ap.add_argument("--seed", type=int, default=20261104)
    ap.add_argument("--no-mcts", action="store_true")
    ap.add_argument("--mcts-games", type=int, default=4096)
    ap.add_argument("--mcts-sims", type=int, default=800)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    return ap.parse_args()


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)", d
    return 6650.0, "fallback (B200_PROFILING.md)", {}


# ------------------------------------------------------------------------------------------------
# CPU arm: the oracle (C++ restatement of the reference crate; the Rust crate itself cannot be
# built in this image) on the host cores, one game per thread.
# ------------------------------------------------------------------------------------------------
def oracle_handle():
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_ffi
    return oracle_ffi.oracle()


def cpu_playout_rate(orc, seconds, seed, plies):
    """steps/s of the oracle random-playout loop with every host thread busy, on a bounded sample."""
    from alphazeroforhnefatafl_b200 import _abi as A
    cores = max(1, orc.hw_threads())
    rules, fen = orc.preset_rules(A.PRESET_COPENHAGEN), orc.preset_board(A.PRESET_COPENHAGEN)
    games = cores * 8
    st, _ = orc.random_playout(rules, fen, 128, seed, games, 16, n_threads=cores)   # warm
    steps = 0
    t0 = time.perf_counter()
    ply = 16
    while True:
        st, stats = orc.random_playout(rules, fen, 128, seed, games, plies, ply0=ply, n_threads=cores, states=st)
        steps += stats.steps
        ply += plies
        el = time.perf_counter() - t0
        if el >= seconds:
            break
    return steps / el, cores, "%d games x %d plies (auto-reset), %d threads, %.1f s" % (games, ply - 16, cores, el)


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from alphazeroforhnefatafl_b200 import _abi as A
    orc = oracle_handle()
    cores = max(1, orc.hw_threads())
    rules, fen = orc.preset_rules(A.PRESET_COPENHAGEN), orc.preset_board(A.PRESET_COPENHAGEN)
    # each step is a bounded sample of the 65 536-game workload: `games` games x `plies` plies
    games = cores * 4
    budget = 150.0 / max(1, args.steps + args.warmup)
    st, _ = orc.random_playout(rules, fen, 128, args.seed, games, 8, n_threads=cores)
    t0 = time.perf_counter()
    st, s0 = orc.random_playout(rules, fen, 128, args.seed, games, 32, ply0=8, n_threads=cores, states=st)
    rate = s0.steps / (time.perf_counter() - t0)
    plies = int(max(1, min(args.plies, rate * min(budget, 0.5) / games)))
    ply = 40
    for _ in range(args.warmup):
        st, _ = orc.random_playout(rules, fen, 128, args.seed, games, plies, ply0=ply, n_threads=cores, states=st)
        ply += plies
    total = 0
    t0 = time.perf_counter()
    for _ in range(args.steps):
        st, s = orc.random_playout(rules, fen, 128, args.seed, games, plies, ply0=ply, n_threads=cores, states=st)
        total += s.steps
        ply += plies
    el = time.perf_counter() - t0
    v = total / el
    sample = "%d games x %d plies per step (sample of the 65536-game workload), %d threads" % (games, plies, cores)
    _emit(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * el / max(1, args.steps), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u128 bitboards (2 x u64)", "data": "synthetic",
        "config": {"workload": "Copenhagen 11x11 random-play env stepping (BASELINE configs[1]), CPU oracle port of "
                               "the reference crate (Rust toolchain unavailable)", "sample": sample},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


# ------------------------------------------------------------------------------------------------
# clocks sampler (NVML): SM clock + throttle reasons during the timed region
# ------------------------------------------------------------------------------------------------
class Clocks(threading.Thread):
    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap",
               0x80: "hw_power_brake", 0x2: "applications_clocks_setting", 0x100: "display_clock_setting"}

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz, self.stop_flag = index, [], set(), None, False
        self.power = []
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if not self.nv:
            return
        while not self.stop_flag:
            try:
                self.samples.append(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM))
                self.power.append(self.nv.nvmlDeviceGetPowerUsage(self.h) / 1000.0)
                r = self.nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in self.REASONS.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.02)

    def summary(self):
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "power_w_max": max(self.power) if self.power else None, "samples": len(s)}


def run_b200(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    from alphazeroforhnefatafl_b200 import _abi as A
    from alphazeroforhnefatafl_b200 import engine as E
    from alphazeroforhnefatafl_b200._lib import lib

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — libtafl_b200 has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    hbm_peak, peak_src, _ = peaks()
    n, P = args.games, args.plies
    rules, fen = E.preset.rules.COPENHAGEN, E.preset.boards.COPENHAGEN
    g = E.LockstepGames(rules, fen, n, local, 128)
    g.set_game_id_base(rank * n)           # games shard by id range; no data-path collective
    stream = torch.cuda.ExternalStream(g.stream, device=local)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")    # > 126 MB L2

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def one_step():
        g.step_random(args.seed, P, auto_reset=True, steps_per_launch=1, sync=False)

    with torch.cuda.stream(stream):
        for _ in range(max(3, args.warmup)):
            flush.zero_()
            one_step()
        barrier()
        g.stats_reset()
        clk = Clocks(local)
        clk.start()
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
        l0 = g.launches
        barrier()
        t_wall0 = time.perf_counter()
        for a, b in evs:
            flush.zero_()                  # L2 flush between timed iterations (not inside the events)
            a.record(stream)
            one_step()
            b.record(stream)
        barrier()
        t_wall = time.perf_counter() - t_wall0
        launches = g.launches - l0
        clk.stop_flag = True
        clk.join()
    dev_ms = sum(a.elapsed_time(b) for a, b in evs)
    stats = g.stats()
    t = torch.tensor([dev_ms], dtype=torch.float64, device="cuda")
    steps_done = torch.tensor([stats.steps], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(steps_done, op=dist.ReduceOp.SUM)
    dev_ms_max, total_steps = float(t.item()), float(steps_done.item())
    value = total_steps / (dev_ms_max * 1e-3)
    kernel_us = 1e3 * dev_ms / max(1, args.steps * P)          # average launch duration (this rank)
    achieved = ENV_BYTES_PER_STEP * n / (kernel_us * 1e-6) / 1e9
    mean_legal = stats.legal_moves / max(1, stats.steps)

    # ---- e2e (headline): Game::do_play through the C ABI with HOST buffers -------------------------
    # The reference-facing call of this path is Game::do_play(play) -> Result<GameStatus, PlayInvalid>
    # (game/mod.rs:95-101).  A seeded random-play action sequence is recorded first (untimed: the fused
    # kernel reports the play it chose, tafl_actions_host); the timed region then replays it from the
    # start position with one tafl_step_status call per ply: H2D of that ply's 65 536 plays from pinned
    # memory, validate + apply on the device, D2H of the status and PlayInvalid bytes.
    L = lib()
    e2e_plies = 48
    rounds = max(2, min(8, args.steps // 25))
    plays_rec = torch.empty((e2e_plies, n, 4), dtype=torch.uint8).pin_memory()
    st_host = torch.empty((e2e_plies, n), dtype=torch.uint8).pin_memory()
    err_host = torch.empty((e2e_plies, n), dtype=torch.uint8).pin_memory()
    g.reset()
    for ply in range(e2e_plies):
        g.step_random(args.seed + 1, 1, auto_reset=False)
        moves, _, _ = g.last_actions()
        plays_rec[ply].numpy()[:] = E.moves_to_plays(moves, 128).view(np.uint8).reshape(n, 4)
    final_rec = g.get_states().tobytes()

    # argument pointers taken once: indexing a torch tensor costs microseconds a compiled host does not pay
    e2e_ptrs = [(plays_rec[p].data_ptr(), st_host[p].data_ptr(), err_host[p].data_ptr()) for p in range(e2e_plies)]
    step_status, handle = L.tafl_step_status, g.handle

    def e2e_round():
        for p_in, p_st, p_err in e2e_ptrs:
            rc = step_status(handle, p_in, p_st, p_err)
            if rc != 0:
                raise RuntimeError("tafl_step_status failed: %d" % rc)

    g.reset()
    e2e_round()                                  # warm-up round, also the correctness check of the replay
    if g.get_states().tobytes() != final_rec:
        raise RuntimeError("e2e replay through Game::do_play does not reproduce the recorded games")
    e2e_n_round = int((err_host.numpy() == 0).sum())
    e2e_time = 0.0
    for _ in range(rounds):
        g.reset()
        barrier()
        t0 = time.perf_counter()
        e2e_round()
        barrier()
        e2e_time += time.perf_counter() - t0
    e2e_t = torch.tensor([e2e_time], dtype=torch.float64, device="cuda")
    e2e_n = torch.tensor([float(e2e_n_round * rounds)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(e2e_t, op=dist.ReduceOp.MAX)
        dist.all_reduce(e2e_n, op=dist.ReduceOp.SUM)
    e2e_value = float(e2e_n.item()) / float(e2e_t.item())

    # ---- second e2e figure: by-value states (GameLogic::do_play(play, state) -> state), 64 B each way --
    h_in = torch.empty(n * 64, dtype=torch.uint8).pin_memory()
    h_out = torch.empty(n * 64, dtype=torch.uint8).pin_memory()
    h_in.numpy()[:] = np.frombuffer(g.get_states().tobytes(), np.uint8)
    bv_steps = max(10, min(args.steps, 100))

    def bv_step():
        rc = L.tafl_step_random_host(g.handle, h_in.data_ptr(), h_out.data_ptr(), args.seed, 1, 1)
        if rc != 0:
            raise RuntimeError("tafl_step_random_host failed: %d" % rc)

    for _ in range(3):
        bv_step()
    g.stats_reset()
    barrier()
    t0 = time.perf_counter()
    for _ in range(bv_steps):
        bv_step()
        h_in, h_out = h_out, h_in
    barrier()
    bv_t = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
    bv_n = torch.tensor([float(g.stats().steps)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(bv_t, op=dist.ReduceOp.MAX)
        dist.all_reduce(bv_n, op=dist.ReduceOp.SUM)
    bv_value = float(bv_n.item()) / float(bv_t.item())

    out = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": dev_ms_max / max(1, args.steps), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u128 bitboards (2 x u64)", "data": "synthetic",
        "config": {"workload": "Copenhagen 11x11 random-play env stepping, %d lockstep games per GPU (BASELINE configs[1])" % n,
                   "games_per_gpu": n, "plies_per_step": P, "steps_per_launch": 1, "auto_reset": True,
                   "l2": "flushed between timed iterations (256 MiB memset, outside the events)",
                   "sharding": "game-id ranges, no data-path collective", "mean_legal_moves": mean_legal},
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": hbm_peak, "unit": "GB/s", "frac": achieved / hbm_peak,
                     "traffic": ncu_traffic(), "peak_source": peak_src, "kernel": "k_step_random<BB128>",
                     "binding_limit": ncu_binding_limit(),
                     "kernel_us": kernel_us, "bytes_per_env_step": ENV_BYTES_PER_STEP,
                     "note": "integer-issue bound, not HBM bound: see profiles/ for issue-slot utilisation"},
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": n * 4, "d2h_bytes_per_step": n * 2,
                "api": "tafl_step_status == Game::do_play(play) -> Result<GameStatus, PlayInvalid> for 65536 resident games: "
                       "pinned host plays in, status + PlayInvalid bytes out, one call per ply (the kernel reads / writes the "
                       "pinned buffers over PCIe itself, bytes per step as listed); recorded seeded random-play "
                       "action sequence replayed from the start position (replay checked bit-exact against the recording)",
                "plies_per_round": e2e_plies, "rounds": rounds,
                "by_value_states": {"value": bv_value, "unit": UNIT, "h2d_bytes_per_step": n * 64, "d2h_bytes_per_step": n * 64,
                                    "api": "tafl_step_random_host(1 ply): pinned host states in, successor states out, chunked over streams",
                                    "steps": bv_steps}},
        "gpu_launches": int(launches),
        "clocks": clk.summary(),
        "wall_s_timed_region": t_wall,
    }
    if not args.no_mcts:
        try:
            out["mcts"] = bench_mcts(args, E, rank, world, local, barrier, dist if world > 1 else None, torch)
        except Exception as ex:  # the env-step line must survive an MCTS failure, loudly
            out["mcts"] = {"error": repr(ex)}
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        v, cores, sample = cpu_playout_rate(oracle_handle(), args.cpu_seconds, args.seed, 64)
        out["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample}
    g.close()
    if rank == 0:
        _emit(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


def ncu_summary(kernel="k_step_random"):
    """the committed ncu --set full summary of the kernel (profiles/ncu_summary.json, written by profiles/ncu_summary.py)"""
    p = os.path.join(ROOT, "profiles", "ncu_summary.json")
    try:
        with open(p) as f:
            return json.load(f).get(kernel, {})
    except Exception:
        return {}


def ncu_traffic():
    return ncu_summary().get("dram_bytes_per_launch")


def ncu_binding_limit():
    """What actually bounds the step kernel (SURVEY.md §8d asks for it next to the HBM fraction): ALU-pipe and
    issue-slot utilisation from the last committed ncu capture — evidence copied from profiles/, not measured in this run."""
    d = ncu_summary()
    g = lambda k: (d.get(k) or {}).get("avg")   # noqa: E731
    alu, issue = g("sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active"), g("smsp__issue_active.avg.pct_of_peak_sustained_active")
    if alu is None and issue is None:
        return None
    return {"alu_pipe_frac": None if alu is None else alu / 100.0, "issue_slot_frac": None if issue is None else issue / 100.0,
            "warp_inst_per_launch": g("smsp__inst_executed.sum"), "registers": g("launch__registers_per_thread"),
            "source": "profiles/ncu_summary.json (ncu --set full, last committed capture)"}


def bench_mcts(args, E, rank, world, local, barrier, dist, torch):
    """Copenhagen MCTS, 800 sims/move, stub evaluator, 4096 trees per GPU (BASELINE configs[2])."""
    n = args.mcts_games
    g = E.LockstepGames(E.preset.rules.COPENHAGEN, E.preset.boards.COPENHAGEN, n, local, 128)
    g.set_game_id_base(rank * n)
    g.step_random(args.seed, 20, auto_reset=True)          # start the trees from 20-ply positions
    from alphazeroforhnefatafl_b200 import selfplay as SP
    sp = SP.SelfPlay(g, num_sims=args.mcts_sims, cpuct=1.0)     # search + first-max move + pack + gather to rank 0
    stream = sp.stream
    moves = 3
    for _ in range(2):                                     # warm-up moves (NCCL channels, allocator blocks of the gather)
        sp.move()
    s0 = g.mcts_stats().as_dict()
    l0 = g.launches
    marks = [torch.cuda.Event(enable_timing=True) for _ in range(moves + 1)]
    barrier()
    marks[0].record(stream)
    n_rec = 0
    for i in range(moves):
        out = sp.move()
        n_rec += 0 if out is None else out.shape[0]
        marks[i + 1].record(stream)
    barrier()
    ms = marks[0].elapsed_time(marks[-1])
    ms_moves = [marks[i].elapsed_time(marks[i + 1]) for i in range(moves)]
    s1 = g.mcts_stats().as_dict()
    sims = s1["sims"] - s0["sims"]
    t = torch.tensor([ms], dtype=torch.float64, device="cuda")
    c = torch.tensor([float(sims)], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(c, op=dist.ReduceOp.SUM)
    nodes, depth, edges = s1["nodes"] - s0["nodes"], s1["depth_sum"] - s0["depth_sum"], s1["edges"] - s0["edges"]
    d_bar, l_bar = depth / max(1, sims), edges / max(1, nodes)
    b_sim = d_bar * (16 + 12 * l_bar) + (144 + 18 * l_bar) + 32 * d_bar
    v = float(c.item()) / (float(t.item()) * 1e-3)
    hbm_peak, _, _ = peaks()
    res = {"metric": "copenhagen_11x11_mcts_sims_per_s", "value": v, "unit": "sims/s", "trees_per_gpu": n,
           "sims_per_move": args.mcts_sims, "moves_timed": moves, "ms_per_move": float(t.item()) / moves,
           "ms_each_move_rank0": ms_moves, "mean_depth": d_bar, "mean_edges_per_node": l_bar, "bytes_per_sim_model": b_sim,
           "roofline": {"bound": "hbm", "achieved": v / world * b_sim / 1e9, "peak": hbm_peak, "unit": "GB/s",
                        "frac": v / world * b_sim / 1e9 / hbm_peak},
           "records_gathered_on_rank0": int(n_rec), "record_bytes": 1592,
           "includes": "k_mcts_init + k_mcts_sims + k_mcts_results + k_mcts_play_best + k_pack_records + torch record gather",
           "gpu_launches": int(g.launches - l0)}
    g.close()
    # secondary numbers (rank-local, same trees / sims): the reference's own usage keeps one MCTS
    # object per episode (tree reuse), and a real evaluator plugs in through the split-phase hook
    for key, kw in (("tree_reuse", dict(tree_reuse=True)), ("evaluator_hook", dict(evaluator=("stub",)))):
        try:
            res[key] = _mcts_variant(args, E, SP, rank, local, torch, moves, **kw)
        except Exception as ex:  # noqa: BLE001
            res[key] = {"error": repr(ex)}
    return res


def _mcts_variant(args, E, SP, rank, local, torch, moves, **kw):
    n = args.mcts_games
    g = E.LockstepGames(E.preset.rules.COPENHAGEN, E.preset.boards.COPENHAGEN, n, local, 128)
    try:
        g.set_game_id_base(rank * n)
        g.step_random(args.seed, 20, auto_reset=True)
        sp = SP.SelfPlay(g, num_sims=args.mcts_sims, cpuct=1.0, **kw)
        sp.move(gather=False)
        s0, l0 = g.mcts_stats().as_dict(), g.launches
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(sp.stream)
        for _ in range(moves):
            sp.move(gather=False)
        b.record(sp.stream)
        torch.cuda.synchronize()
        ms = a.elapsed_time(b)
        s1 = g.mcts_stats().as_dict()
        sims = s1["sims"] - s0["sims"]
        return {"value": sims / (ms * 1e-3), "unit": "sims/s (this rank)", "ms_per_move": ms / moves, "moves_timed": moves,
                "nodes_created_per_sim": (s1["nodes"] - s0["nodes"]) / max(1, sims), "mean_depth": (s1["depth_sum"] - s0["depth_sum"]) / max(1, sims),
                "gpu_launches": int(g.launches - l0)}
    finally:
        g.close()


def _emit(line):
    """the ONE JSON line, on the process's real stdout"""
    os.write(_REAL_STDOUT, (line + "\n").encode())


if __name__ == "__main__":
    a = parse()
    # Libraries write to fd 1 behind Python's back (NCCL prints its version banner there when NCCL_DEBUG
    # is set in the environment): everything but the JSON line goes to stderr.
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)
