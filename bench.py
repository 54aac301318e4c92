#!/usr/bin/env python
"""bench.py — Copenhagen 11x11 env steps/s (BASELINE.json configs[1]: 65 536 lockstep games of
random legal play on one B200) plus MCTS simulations/s, through libtafl_b200.so.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload tablut_mcts]

One bench "step" = `plies_per_step` lockstep plies of all resident games (one launch of the fused step
kernel per ply, the state round-trips HBM after every ply); the L2 is flushed between timed steps.
Whatever --steps is, the work per step is scaled so that every timed region (env, do_play e2e, MCTS)
lasts at least --min-region-s seconds: these are sustained figures, not bursts.  Prints ONE JSON line
(DESIGN.md §4 explains every field):

    value / roofline      fused random-play env step, states resident in HBM      (configs[1])
    e2e                   Game::do_play through the C ABI with pinned host buffers (headline)
    legal_moves           get_all_possible_moves, 65 536 games per launch          (north-star K1)
    mcts                  Copenhagen 800 sims/move, 4096 trees, self-play sweep with record gather (configs[2], [4])
    mcts_tablut           the same on Tablut 9x9                                   (configs[3])
    cpu_baseline, c1_brandubh_cpu, mcts.cpu_baseline, mcts_tablut.cpu_baseline     the oracle on the host cores
"""
import argparse
import json
import math
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

ENV_BYTES_PER_STEP = 136          # 64 B state read + 64 B state write + 4 B action + 4 B status
METRIC, UNIT = "copenhagen_11x11_env_steps_per_s", "env steps/s"
DTYPE = "u128 (two u64 bitboard words per side)"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="copenhagen_env", choices=["copenhagen_env", "tablut_mcts"],
                    help="tablut_mcts: BASELINE configs[3] alone, as the headline line (the default run carries it as mcts_tablut)")
    ap.add_argument("--games", type=int, default=65536)
    ap.add_argument("--plies", type=int, default=64, help="lockstep plies per graph launch (a step is a whole number of these)")
    ap.add_argument("--seed", type=int, default=20261104)
    ap.add_argument("--min-region-s", type=float, default=1.2, help="every timed region lasts at least this long")
    ap.add_argument("--no-mcts", action="store_true")
    ap.add_argument("--mcts-games", type=int, default=4096)
    ap.add_argument("--mcts-sims", type=int, default=800)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-seconds", type=float, default=10.0, help="length of each CPU baseline loop")
    return ap.parse_args()


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)", d
    return 6650.0, "fallback (B200_PROFILING.md)", {}


def git_head():
    try:
        return subprocess.check_output(["git", "-C", ROOT, "rev-parse", "--short", "HEAD"], text=True, stderr=subprocess.DEVNULL).strip()
    except Exception:  # noqa: BLE001  (the GPU box has no .git)
        return None


# ------------------------------------------------------------------------------------------------
# CPU arm: the oracle (C++ restatement of the reference crate; the Rust crate itself cannot be
# built in this image: no cargo / rustc) on the host cores, one game / one tree per thread.
# ------------------------------------------------------------------------------------------------
def oracle_handle():
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_ffi
    return oracle_ffi.oracle()


def preset_ids():
    from alphazeroforhnefatafl_b200 import _abi as A
    return {"copenhagen": A.PRESET_COPENHAGEN, "tablut": A.PRESET_TABLUT, "brandubh": A.PRESET_BRANDUBH}


def cpu_playout_rate(orc, seconds, seed, preset="copenhagen", width=128, threads=None, games_per_thread=8, plies=2048):
    """steps/s of the oracle random-playout loop (auto-reset) on a bounded sample.  Each call runs
    games x plies plies (>= 100 ms of work, so thread start-up is noise)."""
    cores = max(1, orc.hw_threads()) if threads is None else threads
    pid = preset_ids()[preset]
    rules, fen = orc.preset_rules(pid), orc.preset_board(pid)
    games = cores * games_per_thread
    st, _ = orc.random_playout(rules, fen, width, seed, games, 64, n_threads=cores)   # warm
    finished, ply, total_el, halves = 0, 64, 0.0, []
    for _ in range(2):   # two halves, the better one is reported: the host is shared, and an under-read CPU flatters the GPU
        steps, t0 = 0, time.perf_counter()
        while True:
            st, stats = orc.random_playout(rules, fen, width, seed, games, plies, ply0=ply, n_threads=cores, states=st)
            steps += stats.steps
            finished += stats.games_finished
            ply += plies
            el = time.perf_counter() - t0
            if el >= seconds / 2:
                break
        halves.append(steps / el)
        total_el += el
    sample = "%d games x %d plies (random legal play to terminal, auto-restart), %d threads, %.1f s in two halves (%.3g / %.3g per s, the better one reported), %d games finished" % (
        games, ply - 64, cores, total_el, halves[0], halves[1], finished)
    return max(halves), cores, sample


def cpu_mcts_rate(orc, seconds, seed, sims, preset="copenhagen"):
    """sims/s of the oracle MCTS (mcts.py restated, stub evaluator), one tree per host thread at a time"""
    cores = max(1, orc.hw_threads())
    pid = preset_ids()[preset]
    rules, fen = orc.preset_rules(pid), orc.preset_board(pid)
    trees = cores * 16
    roots, _ = orc.random_playout(rules, fen, 128, seed, trees, 20, n_threads=cores)     # 20-ply positions, as the GPU leg
    orc.mcts_search(rules, 128, roots[:cores], sims, n_threads=cores)                    # warm
    calls, total_el, halves = 0, 0.0, []
    for _ in range(2):   # two halves, the better one is reported (shared host)
        done, t0 = 0, time.perf_counter()
        while True:
            r = orc.mcts_search(rules, 128, roots, sims, n_threads=cores)
            done += r["stats"].sims
            calls += 1
            el = time.perf_counter() - t0
            if el >= seconds / 2:
                break
        halves.append(done / el)
        total_el += el
    v = max(halves)
    return {"value": v, "unit": "sims/s", "per_core": v / cores, "cores": cores, "kind": "port",
            "sample": "%d calls x %d trees x %d sims (20-ply %s positions, fresh tree per search), %d threads, %.1f s in two halves (%.3g / %.3g per s, the better one reported)" % (
                calls, trees, sims, preset, cores, total_el, halves[0], halves[1])}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    orc = oracle_handle()
    cores = max(1, orc.hw_threads())
    pid = preset_ids()["copenhagen"]
    rules, fen = orc.preset_rules(pid), orc.preset_board(pid)
    # Each step is a bounded sample of the 65 536-game workload: cores x 8 games x `plies` plies, sized
    # from a calibration call so that one step is >= 150 ms of work (thread start-up is noise) and the
    # whole run stays within ~2 minutes.
    games = cores * 8
    st, _ = orc.random_playout(rules, fen, 128, args.seed, games, 64, n_threads=cores)
    t0 = time.perf_counter()
    st, s0 = orc.random_playout(rules, fen, 128, args.seed, games, 512, ply0=64, n_threads=cores, states=st)
    rate = s0.steps / (time.perf_counter() - t0)
    total_calls = max(1, args.steps + args.warmup)
    per_call_s = min(max(0.15, args.min_region_s / max(1, args.steps)), 120.0 / total_calls)
    plies = int(max(64, rate * per_call_s / games))
    ply = 576
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
    sample = "%d games x %d plies per step (sample of the 65536-game workload, auto-restart), %d threads, %.2f s timed" % (
        games, plies, cores, el)
    _emit(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * el / max(1, args.steps), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": DTYPE, "data": "synthetic",
        "config": {"workload": "Copenhagen 11x11 random-play env stepping (BASELINE configs[1]), CPU oracle port of "
                               "the reference crate (Rust toolchain unavailable on this box)", "sample": sample},
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


class Ctx:
    """what every leg needs: ranks, barrier, reductions"""

    def __init__(self, args):
        import torch
        import torch.distributed as dist
        self.args, self.torch, self.dist = args, torch, dist
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        if not torch.cuda.is_available():
            raise SystemExit("bench.py: no CUDA device — libtafl_b200 has no CPU fallback (use --impl reference for the CPU arm)")
        torch.cuda.set_device(self.local)
        if self.world > 1:
            dist.init_process_group("nccl", device_id=torch.device("cuda", self.local))

        self.affinity0, self.bound_cpus = os.sched_getaffinity(0), None
        self.bind_near_gpu()

    def bind_near_gpu(self):
        """Keep this rank's host thread — the one that launches, waits and allocates the pinned buffers — on the CPUs
        NVML names as closest to its GPU (same NUMA node / PCIe root), within what the process may use.  The CPU
        baseline legs undo it (all_cpus) so that they see every core."""
        try:
            import pynvml
            pynvml.nvmlInit()
            h = pynvml.nvmlDeviceGetHandleByIndex(self.local)
            words = pynvml.nvmlDeviceGetCpuAffinity(h, (max(self.affinity0) + 64) // 64)
            near = {64 * i + b for i, w in enumerate(words) for b in range(64) if (int(w) >> b) & 1} & self.affinity0
            if near and near != self.affinity0:
                os.sched_setaffinity(0, near)
                self.bound_cpus = len(near)
        except Exception:  # noqa: BLE001  (no NVML, no such call, not permitted: stay where we are)
            pass

    def all_cpus(self):
        os.sched_setaffinity(0, self.affinity0)

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def reduce(self, x, op="max"):
        t = self.torch.tensor([float(x)], dtype=self.torch.float64, device="cuda")
        if self.world > 1:
            self.dist.all_reduce(t, op={"max": self.dist.ReduceOp.MAX, "min": self.dist.ReduceOp.MIN}.get(op, self.dist.ReduceOp.SUM))
        return float(t.item())


def run_b200(args):
    cx = Ctx(args)
    torch = cx.torch
    if args.workload == "tablut_mcts":
        from alphazeroforhnefatafl_b200 import engine as E
        res = bench_mcts(cx, E, "tablut")
        if cx.rank == 0:
            _emit(json.dumps({
                "metric": res["metric"], "value": res["value"], "unit": "sims/s", "n_gpus": cx.world, "steps": res["moves_timed"],
                "warmup": 2, "ms_per_step": res["ms_per_move"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "f64 PUCT over " + DTYPE, "data": "synthetic",
                "config": {"workload": "Tablut 9x9 MCTS self-play, 800 sims/move, %d trees per GPU, game-sharded (BASELINE configs[3])" % args.mcts_games},
                "roofline": res["roofline"], "cpu_baseline": res.get("cpu_baseline"), "e2e": res.get("e2e"),
                "gpu_launches": res["gpu_launches"], "detail": res}))
        if cx.world > 1:
            cx.dist.destroy_process_group()
        return

    import numpy as np
    from alphazeroforhnefatafl_b200 import engine as E
    from alphazeroforhnefatafl_b200._lib import lib
    rank, world, local = cx.rank, cx.world, cx.local
    hbm_peak, peak_src, _ = peaks()
    n, P = args.games, args.plies
    rules, fen = E.preset.rules.COPENHAGEN, E.preset.boards.COPENHAGEN
    g = E.LockstepGames(rules, fen, n, local, 128)
    g.set_game_id_base(rank * n)           # games shard by id range; no data-path collective
    stream = torch.cuda.ExternalStream(g.stream, device=local)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")    # > 126 MB L2
    barrier = cx.barrier

    def graph_launch():
        g.step_random(args.seed, P, auto_reset=True, steps_per_launch=1, sync=False)

    with torch.cuda.stream(stream):
        # warm-up (also brings the games to steady state: auto-reset keeps a mix of openings, middle games, endings)
        for _ in range(max(3, args.warmup)):
            flush.zero_()
            for _ in range(4):
                graph_launch()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        for _ in range(4):
            graph_launch()
        b.record(stream)
        torch.cuda.synchronize()
        t_graph = cx.reduce(a.elapsed_time(b) * 1e-3 / 4)                         # seconds per P plies
        reps = max(1, int(math.ceil(args.min_region_s / max(1, args.steps) / t_graph)))
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
            for _ in range(reps):
                graph_launch()
            b.record(stream)
        barrier()
        t_wall = time.perf_counter() - t_wall0
        launches = g.launches - l0
        clk.stop_flag = True
        clk.join()
    dev_ms = sum(a.elapsed_time(b) for a, b in evs)
    stats = g.stats()
    dev_ms_max, total_steps = cx.reduce(dev_ms), cx.reduce(stats.steps, "sum")
    value = total_steps / (dev_ms_max * 1e-3)
    plies_per_step = reps * P
    kernel_us = 1e3 * dev_ms / max(1, args.steps * plies_per_step)          # average launch duration (this rank)
    achieved = ENV_BYTES_PER_STEP * n / (kernel_us * 1e-6) / 1e9
    mean_legal = stats.legal_moves / max(1, stats.steps)
    steady = g.get_states()                                                 # steady-state positions for the other legs

    legal = bench_legal_moves(cx, E, g, stream, flush, hbm_peak)

    # ---- e2e (headline): Game::do_play through the C ABI with HOST buffers -------------------------
    # The reference-facing call of this path is Game::do_play(play) -> Result<GameStatus, PlayInvalid>
    # (game/mod.rs:95-101).  A seeded random-play action sequence is recorded first from the STEADY-STATE
    # positions of the leg above (untimed: the fused kernel reports the play it chose, tafl_actions_host),
    # with auto-restart, so the sequence holds captures, wins and draws at their natural rate; the timed
    # region replays it with one tafl_step_status call per ply — H2D of that ply's 65 536 plays from
    # pinned memory, validate + apply on the device, D2H of the status and PlayInvalid bytes; games that had
    # ended are replaced by new ones in the same launch (tafl_step_status_restart == tafl_reset_finished +
    # tafl_step_status: what a host does when it sees a finished status, Game::new).
    L = lib()
    e2e_plies = 256
    plays_rec = torch.empty((e2e_plies, n, 4), dtype=torch.uint8).pin_memory()
    st_host = torch.empty((e2e_plies, n), dtype=torch.uint8).pin_memory()
    err_host = torch.empty((e2e_plies, n), dtype=torch.uint8).pin_memory()
    g.set_states(steady)
    terminals = 0
    for ply in range(e2e_plies):
        g.step_random(args.seed + 1, 1, auto_reset=True)
        moves, status, _ = g.last_actions()
        terminals += int((status != 0).sum())
        plays_rec[ply].numpy()[:] = E.moves_to_plays(moves, 128).view(np.uint8).reshape(n, 4)
    final_rec = g.get_states().tobytes()
    # argument pointers taken once: indexing a torch tensor costs microseconds a compiled host does not pay
    e2e_ptrs = [(plays_rec[p].data_ptr(), st_host[p].data_ptr(), err_host[p].data_ptr()) for p in range(e2e_plies)]
    step_status, handle = L.tafl_step_status_restart, g.handle

    def e2e_round():
        for p_in, p_st, p_err in e2e_ptrs:
            rc = step_status(handle, p_in, p_st, p_err)   # games that ended on the previous ply start over, then the play
            if rc != 0:
                raise RuntimeError("tafl_step_status_restart failed: %d" % rc)

    g.set_states(steady)
    t0 = time.perf_counter()
    e2e_round()                                  # warm-up round, also the correctness check of the replay
    t_round = time.perf_counter() - t0
    if g.get_states().tobytes() != final_rec:
        raise RuntimeError("e2e replay through Game::do_play does not reproduce the recorded games")
    if int((err_host.numpy() != 0).sum()) != 0:
        raise RuntimeError("e2e replay: a recorded play was rejected")
    rounds = max(2, int(math.ceil(1.25 * args.min_region_s / cx.reduce(t_round))))
    e2e_time = 0.0
    for _ in range(rounds):
        g.set_states(steady)
        barrier()
        t0 = time.perf_counter()
        e2e_round()
        barrier()
        e2e_time += time.perf_counter() - t0
    e2e_value = cx.reduce(float(n) * e2e_plies * rounds, "sum") / cx.reduce(e2e_time)

    # ---- second e2e figure: by-value states (GameLogic::do_play(play, state) -> state), 64 B each way --
    h_in = torch.empty(n * 64, dtype=torch.uint8).pin_memory()
    h_out = torch.empty(n * 64, dtype=torch.uint8).pin_memory()
    h_in.numpy()[:] = np.frombuffer(steady.tobytes(), np.uint8)

    def bv_step():
        rc = L.tafl_step_random_host(g.handle, h_in.data_ptr(), h_out.data_ptr(), args.seed, 1, 1)
        if rc != 0:
            raise RuntimeError("tafl_step_random_host failed: %d" % rc)

    t0 = time.perf_counter()
    for _ in range(5):
        bv_step()
    bv_steps = max(10, int(math.ceil(args.min_region_s / cx.reduce((time.perf_counter() - t0) / 5))))
    g.stats_reset()
    barrier()
    t0 = time.perf_counter()
    for _ in range(bv_steps):
        bv_step()
        h_in, h_out = h_out, h_in
    barrier()
    bv_value = cx.reduce(g.stats().steps, "sum") / cx.reduce(time.perf_counter() - t0)

    out = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": dev_ms_max / max(1, args.steps), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": DTYPE, "data": "synthetic",
        "config": {"workload": "Copenhagen 11x11 random-play env stepping, %d lockstep games per GPU (BASELINE configs[1])" % n,
                   "games_per_gpu": n, "plies_per_step": plies_per_step, "steps_per_launch": 1, "auto_reset": True,
                   "l2": "flushed between timed iterations (256 MiB memset, outside the events)",
                   "sharding": "game-id ranges, no data-path collective", "mean_legal_moves": mean_legal,
                   "host_thread_cpus": cx.bound_cpus if cx.bound_cpus else len(cx.affinity0),
                   "timed_region_s": dev_ms_max * 1e-3},
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": hbm_peak, "unit": "GB/s", "frac": achieved / hbm_peak,
                     "traffic": ncu_traffic("k_step_random"), "peak_source": peak_src, "kernel": "k_step_random<BB128, Copenhagen>",
                     "binding_limit": ncu_binding_limit("k_step_random"),
                     "kernel_us": kernel_us, "bytes_per_env_step": ENV_BYTES_PER_STEP,
                     "note": "integer-issue bound, not HBM bound: binding_limit carries the ALU-pipe / issue-slot fractions of the committed ncu capture"},
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": n * 4, "d2h_bytes_per_step": n * 2,
                "api": "tafl_step_status_restart == Game::new for the games that had ended + Game::do_play(play) -> Result<GameStatus, PlayInvalid> for 65536 resident games: "
                       "pinned host plays in, status + PlayInvalid bytes out, one call per ply (the kernel reads / writes the "
                       "pinned buffers over PCIe itself, bytes per step as listed); a recorded seeded "
                       "random-play action sequence over steady-state positions with auto-restart is replayed (checked bit-exact)",
                "plies_per_round": e2e_plies, "rounds": rounds, "timed_region_s": cx.reduce(e2e_time),
                "timed_region_s_fastest_rank": cx.reduce(e2e_time, "min"),
                "games_ended_in_the_sequence": terminals,
                "by_value_states": {"value": bv_value, "unit": UNIT, "h2d_bytes_per_step": n * 64, "d2h_bytes_per_step": n * 64,
                                    "api": "tafl_step_random_host(1 ply): pinned host states in, successor states out, chunked over streams",
                                    "steps": bv_steps}},
        "legal_moves": legal,
        "gpu_launches": int(launches),
        "clocks": clk.summary(),
        "wall_s_timed_region": t_wall,
        "git": git_head(),
    }
    g.close()
    del flush
    if not args.no_mcts:
        for key, preset in (("mcts", "copenhagen"), ("mcts_tablut", "tablut")):
            try:
                out[key] = bench_mcts(cx, E, preset)
            except Exception as ex:  # the env-step line must survive an MCTS failure, loudly
                out[key] = {"error": repr(ex)}
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        orc = oracle_handle()
        cx.all_cpus()
        v, cores, sample = cpu_playout_rate(orc, args.cpu_seconds, args.seed)
        out["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample}
        # C1 (BASELINE configs[0]): Brandubh 7x7, ONE game on ONE core, random legal play to terminal with restart
        v1, _, sample1 = cpu_playout_rate(orc, args.cpu_seconds, args.seed, preset="brandubh", width=64, threads=1,
                                          games_per_thread=1, plies=40000)
        out["c1_brandubh_cpu"] = {"metric": "brandubh_7x7_single_game_env_steps_per_s", "value": v1, "unit": UNIT, "cores": 1,
                                  "kind": "port", "sample": sample1}
    if rank == 0:
        _emit(json.dumps(out))
    if world > 1:
        cx.dist.destroy_process_group()


def bench_legal_moves(cx, E, g, stream, flush, hbm_peak):
    """K1: tafl_legal_moves (get_all_possible_moves, game/main.rs:33-43) on the resident steady-state games —
    device time per launch with the L2 flushed, and the small-batch (2048 games, latency-bound) launch."""
    torch, args = cx.torch, cx.args
    from alphazeroforhnefatafl_b200._lib import lib
    L = lib()
    n = g.n_games
    d_moves, d_counts = g.legal_moves_device()                 # allocates + warms
    with torch.cuda.stream(stream):
        for _ in range(3):
            L.tafl_legal_moves_async(g.handle, None, None)
        reps = 40
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
        for a, b in evs:
            flush.zero_()
            a.record(stream)
            L.tafl_legal_moves_async(g.handle, None, None)
            b.record(stream)
        torch.cuda.synchronize()
    us = sorted(1e3 * a.elapsed_time(b) for a, b in evs)
    us_med = us[len(us) // 2]
    moves, counts = g.legal_moves(0, min(n, 8192))
    l_bar = float(counts.mean())
    bytes_per_game = 64 + 4 + 2 * l_bar
    res = {"kernel": "k_legal_moves_g<BB128, Copenhagen, G=%s>" % os.environ.get("TAFL_LM_G", "8 (16 for the 2048-game launch)"), "games": n, "us_per_launch": us_med,
           "us_min": us[0], "mean_legal_moves": l_bar, "bytes_per_game": bytes_per_game,
           "roofline": {"bound": "hbm", "achieved": bytes_per_game * n / (us_med * 1e-6) / 1e9, "peak": hbm_peak, "unit": "GB/s",
                        "frac": bytes_per_game * n / (us_med * 1e-6) / 1e9 / hbm_peak, "traffic": ncu_traffic("k_legal_moves_g"),
                        "binding_limit": ncu_binding_limit("k_legal_moves_g")},
           "includes": "device time of one launch (CUDA events around tafl_legal_moves_async), L2 flushed before each launch, median of 40"}
    # 2048 games: one launch is latency-bound
    small = E.LockstepGames(g.rules, g.start_fen, 2048, cx.local, 128)
    try:
        small.set_states(g.get_states(0, 2048))
        sst = torch.cuda.ExternalStream(small.stream, device=cx.local)
        small.legal_moves_device()
        with torch.cuda.stream(sst):
            evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(40)]
            for a, b in evs:
                a.record(sst)
                L.tafl_legal_moves_async(small.handle, None, None)
                b.record(sst)
            torch.cuda.synchronize()
        us2 = sorted(1e3 * a.elapsed_time(b) for a, b in evs)
        res["us_per_launch_2048_games"] = us2[len(us2) // 2]
    finally:
        small.close()
    return res


def ncu_summary(kernel):
    """the committed ncu --set full summary of the kernel (profiles/ncu_summary.json, written by profiles/ncu_summary.py)"""
    p = os.path.join(ROOT, "profiles", "ncu_summary.json")
    try:
        with open(p) as f:
            return json.load(f).get(kernel, {})
    except Exception:
        return {}


def ncu_traffic(kernel):
    return ncu_summary(kernel).get("dram_bytes_per_launch")


def ncu_binding_limit(kernel):
    """What actually bounds the kernel (SURVEY.md §8d asks for it next to the HBM fraction): ALU-pipe and issue-slot
    utilisation of the last committed ncu capture — evidence copied from profiles/, NOT measured in this run; `capture`
    names the capture's tag and the commit it was taken on, so a stale capture is visible next to `git` of this run."""
    d = ncu_summary(kernel)
    g = lambda k: (d.get(k) or {}).get("avg")   # noqa: E731
    alu, issue = g("sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active"), g("smsp__issue_active.avg.pct_of_peak_sustained_active")
    if alu is None and issue is None:
        return None
    return {"alu_pipe_frac": None if alu is None else alu / 100.0, "issue_slot_frac": None if issue is None else issue / 100.0,
            "warp_inst_per_launch": g("smsp__inst_executed.sum"), "lanes_per_inst": g("smsp__thread_inst_executed_per_inst_executed.ratio"),
            "warps_active_frac": None if g("sm__warps_active.avg.pct_of_peak_sustained_active") is None else g("sm__warps_active.avg.pct_of_peak_sustained_active") / 100.0,
            "registers": g("launch__registers_per_thread"), "ncu_kernel_us": g("gpu__time_duration.sum"),
            "capture": {"tag": d.get("tag"), "commit": d.get("commit"), "file": "profiles/ncu_summary.json"},
            "source": "ncu --set full --clock-control none (cold cache, serialised): shares and fractions carry over, absolute times do not"}


def bench_mcts(cx, E, preset):
    """MCTS 800 sims/move, stub evaluator, 4096 trees per GPU (BASELINE configs[2] Copenhagen / configs[3] Tablut) as a
    self-play generation sweep (configs[4]): every move = search + first-max move + compacted records + gather to the
    replay-buffer rank (C-ABI NCCL, side stream, overlapped with the next search); finished games restart."""
    args, torch = cx.args, cx.torch
    import numpy as np
    from alphazeroforhnefatafl_b200 import _abi as A
    from alphazeroforhnefatafl_b200 import selfplay as SP
    n = args.mcts_games
    rules = getattr(E.preset.rules, preset.upper())
    fen = getattr(E.preset.boards, preset.upper())
    hbm_peak, _, _ = peaks()

    def fresh():
        gg = E.LockstepGames(rules, fen, n, cx.local, 128)
        gg.set_game_id_base(cx.rank * n)
        gg.step_random(args.seed, 20, auto_reset=True)          # start the trees from 20-ply positions
        return gg

    g = fresh()
    sp = SP.SelfPlay(g, num_sims=args.mcts_sims, cpuct=1.0, overlap=True, restart_finished=True)
    replay = SP.DeviceReplayBuffer(n * cx.world * 64, cx.local) if cx.rank == 0 else None   # HBM-resident rolling window
    stream = sp.stream
    for _ in range(3):                                     # warm-up moves (NCCL channels, allocator blocks)
        sp.move()
    sp.flush()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(4):                                     # calibration: how many moves make min_region_s
        sp.move()
    sp.flush()
    torch.cuda.synchronize()
    moves = max(20, int(math.ceil(1.25 * args.min_region_s / cx.reduce((time.perf_counter() - t0) / 4))))
    s0 = g.mcts_stats().as_dict()
    l0 = g.launches
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    cx.barrier()
    t_w0 = time.perf_counter()
    ev0.record(stream)
    n_rec = 0
    for i in range(moves):
        out = sp.move()
        if out is not None and replay is not None:
            n_rec += out.shape[0]
            replay.extend(out, sp.side)
    out = sp.flush()
    if out is not None and replay is not None:
        n_rec += out.shape[0]
        replay.extend(out, sp.side)
    ev1.record(stream)
    cx.barrier()
    t_wall = time.perf_counter() - t_w0
    ms = ev0.elapsed_time(ev1)
    s1 = g.mcts_stats().as_dict()
    sims = s1["sims"] - s0["sims"]
    t_max, c_sum = cx.reduce(max(ms, t_wall * 1e3)), cx.reduce(sims, "sum")     # the slower of device span and host span
    nodes, depth, edges = s1["nodes"] - s0["nodes"], s1["depth_sum"] - s0["depth_sum"], s1["edges"] - s0["edges"]
    d_bar, l_bar = depth / max(1, sims), edges / max(1, nodes)
    b_sim = d_bar * (16 + 12 * l_bar) + (144 + 18 * l_bar) + 32 * d_bar
    v = c_sum / (t_max * 1e-3)
    kern = "k_mcts_sims_g"
    traffic_launch = ncu_traffic(kern) if preset == "copenhagen" else None      # the committed capture is the Copenhagen bench size
    cap = ncu_summary(kern)
    cap_sims = cap.get("sims_per_launch")
    per_sim = traffic_launch / cap_sims if traffic_launch and cap_sims else None
    res = {"metric": "%s_mcts_sims_per_s" % ("copenhagen_11x11" if preset == "copenhagen" else "tablut_9x9"), "value": v, "unit": "sims/s",
           "trees_per_gpu": n, "sims_per_move": args.mcts_sims, "moves_timed": moves, "ms_per_move": t_max / moves,
           "timed_region_s": t_max * 1e-3, "device_ms_rank0": ms, "wall_ms_rank0": t_wall * 1e3,
           "mean_depth": d_bar, "mean_edges_per_node": l_bar, "bytes_per_sim_model": b_sim,
           "roofline": {"bound": "hbm", "achieved": v / cx.world * b_sim / 1e9, "peak": hbm_peak, "unit": "GB/s",
                        "frac": v / cx.world * b_sim / 1e9 / hbm_peak,
                        "basis": "SURVEY §8d MODEL bytes per simulation (every edge of every visited node read) — the kernel moves far less, see frac_measured",
                        "traffic": per_sim, "traffic_unit": "DRAM bytes per simulation (ncu dram__bytes_read+write of one bench-size launch / its simulations)",
                        "frac_measured": None if per_sim is None else v / cx.world * per_sim / 1e9 / hbm_peak,
                        "binding_limit": ncu_binding_limit(kern) if preset == "copenhagen" else None},
           "selfplay_sweep": {"records_gathered_on_rank0": int(n_rec), "record_bytes": A.REC_STRIDE,
                              "replay_window_rows": None if replay is None else len(replay),
                              "gather": "tafl_gather_records_on (C-ABI NCCL, ragged, side stream, overlapped)" if sp.ng is not None else "single rank: rows stay on the device",
                              "restart_finished": True},
           "includes": "k_mcts_init + k_mcts_sims_g + k_mcts_results + k_mcts_play_best + k_record_rows + k_pack_records + k_reset_finished + record gather + replay-window insert",
           "gpu_launches": int(g.launches - l0)}
    if replay is not None:      # z back-fill of the window (host; outside the timed region) — episodes that ran to a terminal
        t0 = time.perf_counter()
        recs = replay.as_array()
        z = SP.backfill_outcomes(recs)
        res["selfplay_sweep"].update({"backfill_s": time.perf_counter() - t0, "rows_with_outcome": int((~np.isnan(z)).sum()),
                                      "episodes_finished": int(((recs["h"]["status"] & 0x80) != 0).sum())})
    sp.close()
    g.close()

    # ---- e2e through the C ABI with HOST buffers: root states in, visit counts / moves / choice out, every move ----
    g = fresh()
    try:
        res["e2e"] = mcts_e2e(cx, E, g)
    except Exception as ex:  # noqa: BLE001
        res["e2e"] = {"error": repr(ex)}
    finally:
        g.close()
    # secondary numbers (rank-local, same trees / sims): the reference's own usage keeps one MCTS
    # object per episode (tree reuse), and a real evaluator plugs in through the split-phase hook
    if preset == "copenhagen":
        for key, kw in (("tree_reuse", dict(tree_reuse=True)), ("evaluator_hook", dict(evaluator=("stub",)))):
            try:
                res[key] = _mcts_variant(cx, E, SP, fresh, **kw)
            except Exception as ex:  # noqa: BLE001
                res[key] = {"error": repr(ex)}
        try:
            res["virtual_loss"] = _mcts_virtual_loss(cx, fresh)
        except Exception as ex:  # noqa: BLE001
            res["virtual_loss"] = {"error": repr(ex)}
        try:
            res["selfplay_sweep_to_terminal"] = _sweep_to_terminal(cx, E, SP, rules, fen)
        except Exception as ex:  # noqa: BLE001
            res["selfplay_sweep_to_terminal"] = {"error": repr(ex)}
    if cx.rank == 0 and cx.world == 1 and not args.no_cpu_baseline:
        cx.all_cpus()
        res["cpu_baseline"] = cpu_mcts_rate(oracle_handle(), args.cpu_seconds, args.seed, args.mcts_sims, preset)
        cx.bind_near_gpu()
    return res


def mcts_e2e(cx, E, g):
    """MCTS.getActionProb for every game through the C ABI with HOST buffers: pinned root states in (tafl_set_states),
    search, moves + visit counts + counts + choice out (tafl_mcts_results_host), the move played (tafl_mcts_play_best ->
    status bytes), successor states out (tafl_get_states)."""
    import numpy as np
    from alphazeroforhnefatafl_b200 import _abi as A
    from alphazeroforhnefatafl_b200._lib import lib
    args = cx.args
    L, n, h = lib(), g.n_games, g.handle
    g.mcts_configure(args.mcts_sims, 1.0, 1e-4)
    st = E.PinnedArray(n, A.STATE_DTYPE)
    mv, vis = E.PinnedArray((n, A.MAX_MOVES), np.uint16), E.PinnedArray((n, A.MAX_MOVES), np.uint32)
    cnt, best, status = E.PinnedArray(n, np.uint32), E.PinnedArray(n, np.uint32), E.PinnedArray(n, np.uint8)
    st.a[:] = g.get_states()

    def one_move():
        for rc in (L.tafl_set_states(h, A.ptr(st.a), 0, n), L.tafl_mcts_search(h),
                   L.tafl_mcts_results_host(h, 0, n, A.ptr(mv.a), A.ptr(vis.a), None, A.ptr(cnt.a), A.ptr(best.a)),
                   L.tafl_mcts_play_best(h, A.ptr(status.a)), L.tafl_get_states(h, A.ptr(st.a), 0, n)):
            if rc != 0:
                raise RuntimeError("MCTS e2e call failed: %d %s" % (rc, L.tafl_last_error(h).decode()))

    t0 = time.perf_counter()
    for _ in range(2):
        one_move()
    moves = max(10, int(math.ceil(args.min_region_s / cx.reduce((time.perf_counter() - t0) / 2))))
    s0 = g.mcts_stats().as_dict()["sims"]
    cx.barrier()
    t0 = time.perf_counter()
    for _ in range(moves):
        one_move()
    cx.barrier()
    el = cx.reduce(time.perf_counter() - t0)
    sims = cx.reduce(g.mcts_stats().as_dict()["sims"] - s0, "sum")
    out = {"value": sims / el, "unit": "sims/s", "moves": moves, "ms_per_move": 1e3 * el / moves, "timed_region_s": el,
           "h2d_bytes_per_step": n * 64, "d2h_bytes_per_step": n * (A.MAX_MOVES * 6 + 4 + 4 + 1 + 64),
           "api": "tafl_set_states + tafl_mcts_search + tafl_mcts_results_host + tafl_mcts_play_best + tafl_get_states, pinned host buffers, per move"}
    for p in (st, mv, vis, cnt, best, status):
        p.close()
    return out


def _mcts_variant(cx, E, SP, fresh, **kw):
    args, torch = cx.args, cx.torch
    g = fresh()
    try:
        sp = SP.SelfPlay(g, num_sims=args.mcts_sims, cpuct=1.0, restart_finished=True, overlap=True, **kw)
        t0 = time.perf_counter()
        sp.move(gather=False)
        sp.flush(gather=False)
        torch.cuda.synchronize()
        moves = max(5, int(math.ceil(0.6 * args.min_region_s / (time.perf_counter() - t0))))
        s0, l0 = g.mcts_stats().as_dict(), g.launches
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(sp.stream)
        for _ in range(moves):
            sp.move(gather=False)
        sp.flush(gather=False)
        b.record(sp.stream)
        torch.cuda.synchronize()
        ms = a.elapsed_time(b)
        s1 = g.mcts_stats().as_dict()
        sims = s1["sims"] - s0["sims"]
        sp.close()
        return {"value": sims / (ms * 1e-3), "unit": "sims/s (this rank)", "ms_per_move": ms / moves, "moves_timed": moves,
                "nodes_created_per_sim": (s1["nodes"] - s0["nodes"]) / max(1, sims), "mean_depth": (s1["depth_sum"] - s0["depth_sum"]) / max(1, sims),
                "gpu_launches": int(g.launches - l0)}
    finally:
        g.close()


def _mcts_virtual_loss(cx, fresh):
    """sims/s of the fused stub search against the virtual-loss width W (north star: PUCT with virtual loss; W = 1 is
    mcts.py's sequential search, W > 1 the batched schedule of oracle/mcts_oracle.hpp::run_from_root_vl), this rank"""
    torch = cx.torch
    out = {"unit": "sims/s (this rank)", "trees": cx.args.mcts_games, "sims_per_move": cx.args.mcts_sims, "by_width": {}}
    for W in (1, 2, 4, 8, 16):
        g = fresh()
        try:
            g.mcts_configure(cx.args.mcts_sims, 1.0, 1e-4, vl_width=W)
            st = torch.cuda.ExternalStream(g.stream, device=cx.local)
            g.mcts_search()
            s0 = g.mcts_stats().as_dict()["sims"]
            with torch.cuda.stream(st):
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record(st)
                for _ in range(8):
                    g.mcts_search(sync=False)
                b.record(st)
                torch.cuda.synchronize()
            out["by_width"][str(W)] = (g.mcts_stats().as_dict()["sims"] - s0) / (a.elapsed_time(b) * 1e-3)
        finally:
            g.close()
    # the same through the evaluator hook (stub evaluator on the device): W leaves per tree and round, the evaluator
    # sees trees x W rows per call (k_mcts_hook_vl; W = 1 is k_mcts_hook_round)
    from alphazeroforhnefatafl_b200 import _abi as A
    out["hook_by_width"] = {}
    for W in (1, 4, 8, 16):
        g = fresh()
        try:
            g.mcts_configure(cx.args.mcts_sims, 1.0, 1e-4, vl_width=W)
            st = torch.cuda.ExternalStream(g.stream, device=cx.local)
            g.mcts_search_builtin(A.EVAL_STUB)          # captures the graph of rounds
            s0 = g.mcts_stats().as_dict()["sims"]
            with torch.cuda.stream(st):
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record(st)
                for _ in range(3):
                    g.mcts_search_builtin(A.EVAL_STUB)
                b.record(st)
                torch.cuda.synchronize()
            out["hook_by_width"][str(W)] = (g.mcts_stats().as_dict()["sims"] - s0) / (a.elapsed_time(b) * 1e-3)
        finally:
            g.close()
    return out


def _sweep_to_terminal(cx, E, SP, rules, fen):
    """configs[4] evidence that the sweep runs episodes to their end: the same self-play loop started from a steady-state
    MIXTURE of game stages (300 random plies with auto-restart), so games finish, restart (episode + 1) and get their
    outcome back-filled inside a short run; records go through the same gather."""
    import numpy as np
    args, torch = cx.args, cx.torch
    n = args.mcts_games
    g = E.LockstepGames(rules, fen, n, cx.local, 128)
    try:
        g.set_game_id_base(cx.rank * n)
        g.step_random(args.seed + 7, 300, auto_reset=True)
        # endgame positions have few pieces and long rays: more edges per node than the default pool (160 per node) assumes
        sp = SP.SelfPlay(g, num_sims=args.mcts_sims, cpuct=1.0, overlap=True, restart_finished=True,
                         edge_capacity=(args.mcts_sims + 1) * 320)
        replay = SP.DeviceReplayBuffer(n * cx.world * 48, cx.local) if cx.rank == 0 else None
        moves = 40
        for _ in range(2):                                 # warm-up: the gather's communicator connects on first use
            sp.move()
        sp.flush()
        cx.barrier()
        t0 = time.perf_counter()
        for _ in range(moves):
            out = sp.move()
            if out is not None and replay is not None:
                replay.extend(out, sp.side)
        out = sp.flush()
        if out is not None and replay is not None:
            replay.extend(out, sp.side)
        cx.barrier()
        el = time.perf_counter() - t0
        res = {"moves": moves, "seconds": cx.reduce(el), "trees_per_gpu": n}
        if replay is not None:
            t0 = time.perf_counter()
            recs = replay.as_array()
            z = SP.backfill_outcomes(recs)
            ended = (recs["h"]["status"] & 0x80) != 0
            res.update({"rows": int(len(recs)), "episodes_finished": int(ended.sum()), "rows_with_outcome": int((~np.isnan(z)).sum()),
                        "max_episode_index": int(recs["h"]["episode"].max()) if len(recs) else 0, "backfill_s": time.perf_counter() - t0})
        sp.close()
        return res
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
