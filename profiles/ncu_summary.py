#!/usr/bin/env python
"""Summarise `ncu --set full` reports (.ncu-rep) into the small text/JSON files kept under profiles/.

    python profiles/ncu_summary.py <tag> <name>=<report.ncu-rep>[@key=number...] [...]

`@key=number` attaches facts about the captured launch that ncu cannot know (e.g. sims_per_launch of the
MCTS capture); the tag and the commit the capture was taken on (git HEAD, or NCU_COMMIT) are stored with
every entry, so bench.py can show a stale capture for what it is.

For every report: the raw-page metrics that carry the roofline argument (duration, DRAM bytes,
registers, occupancy, issue-slot and ALU-pipe utilisation, divergence, stall reasons), averaged over
the captured launches -> profiles/ncu_<tag>_<name>.txt, plus the per-source-line instruction
hot spots from the source page -> profiles/ncu_<tag>_<name>_hotspots.txt.  All reports of one call
are merged into profiles/ncu_summary.json (read by bench.py for `roofline.traffic`).
"""
import collections
import csv
import io
import json
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
METRICS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__grid_size",
    "launch__block_size", "launch__registers_per_thread", "launch__occupancy_limit_registers",
    "sm__maximum_warps_per_active_cycle_pct", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum", "sm__cycles_active.avg",
    "smsp__thread_inst_executed_per_inst_executed.ratio",
    "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_alu.sum", "sm__inst_executed_pipe_fp64.sum", "sm__inst_executed_pipe_lsu.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct", "smsp__warps_eligible.avg.per_cycle_active",
    "l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum", "l1tex__t_requests_pipe_lsu_mem_global_op_st.sum",
    "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_st.sum", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_xu.sum", "sm__inst_executed_pipe_fma.sum",
    "smsp__inst_executed_op_local_ld.sum", "smsp__inst_executed_op_local_st.sum",
    "smsp__average_warp_latency_issue_stalled_long_scoreboard.ratio",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_dispatch_stall_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_imc_miss_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
]


def ncu_csv(rep, page, extra=()):
    out = subprocess.run(["ncu", "-i", rep, "--page", page, "--csv", *extra], capture_output=True, text=True).stdout
    return list(csv.reader(io.StringIO(out)))


def to_float(x):
    try:
        return float(x.replace(",", ""))
    except ValueError:
        return None


def raw_summary(rep):
    rows = ncu_csv(rep, "raw")
    hdr, units, body = rows[0], rows[1], rows[2:]
    res = {"kernel": body[0][hdr.index("Kernel Name")].split("(")[0], "launches": len(body)}
    for m in METRICS:
        if m in hdr:
            i = hdr.index(m)
            vals = [to_float(r[i]) for r in body]
            vals = [v for v in vals if v is not None]
            if vals:
                res[m] = {"avg": sum(vals) / len(vals), "unit": units[i]}
    return res


def hotspots(rep, top=40):
    rows = ncu_csv(rep, "source", ["--print-source", "cuda,sass"])
    per_line = collections.Counter()
    samples = collections.Counter()
    text = {}
    fname, total_inst, total_samp = None, 0, 0
    cols = None
    for r in rows:
        if not r:
            continue
        if r[0] == "File Path":
            fname = os.path.basename(r[1])
        elif r[0] == "Line No":
            cols = r
        elif cols and r[0].isdigit():
            ii, si = cols.index("Instructions Executed"), cols.index("# Samples")
            n, s = to_float(r[ii]) or 0, to_float(r[si]) or 0
            key = (fname, int(r[0]))
            per_line[key] += n
            samples[key] += s
            text[key] = r[1].strip()
            total_inst += n
            total_samp += s
    lines = ["%-22s %5s %12s %7s %8s %7s  %s" % ("file", "line", "warp_inst", "share", "samples", "share", "source")]
    for key, n in per_line.most_common(top):
        lines.append("%-22s %5d %12d %7.3f %8d %7.3f  %s" % (key[0], key[1], n, n / max(1, total_inst), samples[key],
                                                               samples[key] / max(1, total_samp), text[key][:90]))
    by_file = collections.Counter()
    for (f, _), n in per_line.items():
        by_file[f] += n
    lines.append("")
    lines.append("per file: " + ", ".join("%s %.3f" % (f, n / max(1, total_inst)) for f, n in by_file.most_common()))
    return "\n".join(lines)


def main():
    tag = sys.argv[1]
    merged_path = os.path.join(HERE, "ncu_summary.json")
    merged = json.load(open(merged_path)) if os.path.exists(merged_path) else {}
    commit = os.environ.get("NCU_COMMIT")
    if not commit:
        try:
            commit = subprocess.check_output(["git", "-C", HERE, "rev-parse", "--short", "HEAD"], text=True).strip()
        except Exception:  # noqa: BLE001
            commit = None
    for arg in sys.argv[2:]:
        name, rep = arg.split("=", 1)
        rep, *extras = rep.split("@")
        s = raw_summary(rep)
        s["commit"] = commit
        for kv in extras:
            k, v = kv.split("=", 1)
            s[k] = float(v)
        rd = s.get("dram__bytes_read.sum", {"avg": 0, "unit": "byte"})
        wr = s.get("dram__bytes_write.sum", {"avg": 0, "unit": "byte"})
        scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
        s["dram_bytes_per_launch"] = rd["avg"] * scale.get(rd["unit"], 1) + wr["avg"] * scale.get(wr["unit"], 1)
        s["tag"] = tag
        merged[name] = s
        with open(os.path.join(HERE, "ncu_%s_%s.txt" % (tag, name)), "w") as f:
            f.write("# ncu --set full --clock-control none, %d launch(es) of %s (tag %s)\n" % (s["launches"], s["kernel"], tag))
            for k, v in s.items():
                if isinstance(v, dict):
                    f.write("%-80s %16.4f %s\n" % (k, v["avg"], v["unit"]))
            f.write("%-80s %16.0f byte\n" % ("dram_bytes_per_launch (read+write)", s["dram_bytes_per_launch"]))
        with open(os.path.join(HERE, "ncu_%s_%s_hotspots.txt" % (tag, name)), "w") as f:
            f.write(hotspots(rep) + "\n")
    with open(merged_path, "w") as f:
        json.dump(merged, f, indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
