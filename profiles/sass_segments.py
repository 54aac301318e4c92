#!/usr/bin/env python
"""Walk the SASS of one profiled kernel launch (ncu source page) in equal segments and print, per
segment, executed warp instructions, thread efficiency, stall samples and the dominant source lines.
    python profiles/sass_segments.py <report.ncu-rep> [segments] [launch]"""
import collections, csv, io, subprocess, sys
rep = sys.argv[1]; B = int(sys.argv[2]) if len(sys.argv) > 2 else 24; which = int(sys.argv[3]) if len(sys.argv) > 3 else 0
def page(extra):
    return list(csv.reader(io.StringIO(subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", *extra], capture_output=True, text=True).stdout)))
addr2src, fname, cur = {}, None, None
for r in page(["--print-source", "cuda,sass"]):
    if not r: continue
    if r[0] == "File Path": fname = r[1].split("/")[-1]; continue
    if r[0].isdigit(): cur = (fname, int(r[0]), r[1].strip()[:70]); continue
    if len(r) > 3 and r[2].startswith("0x"): addr2src.setdefault(r[2], cur)
hdr, launches = None, []
for r in page([]):
    if r and r[0] == "Address": hdr = r; launches.append([]); continue
    if hdr and len(r) == len(hdr) and r[0].startswith("0x"): launches[-1].append(r)
L = launches[which]; n = len(L)
ix = {k: hdr.index(k) for k in ["# Samples", "Instructions Executed", "Thread Instructions Executed", "stall_no_inst", "stall_wait", "stall_branch_resolving", "stall_short_sb", "stall_long_sb", "stall_math"]}
f = lambda x: float(x) if x not in ("", "-") else 0.0
print("static instructions", n, "executed", sum(f(r[ix["Instructions Executed"]]) for r in L))
for b in range(B):
    seg = L[b * n // B:(b + 1) * n // B]
    g = lambda k: sum(f(r[ix[k]]) for r in seg)
    ex = g("Instructions Executed")
    c = collections.Counter()
    for r in seg: c[addr2src.get(r[0])] += f(r[ix["Instructions Executed"]])
    top = "; ".join("%s:%d" % (k[0].replace("tafl_", ""), k[1]) for k, _ in c.most_common(3) if k)
    print("seg %2d [%5d] exec %8d thr/inst %5.1f samp %4d no_inst %3d wait %3d br %3d ssb %3d lsb %3d math %3d | %s" % (
        b, b * n // B, ex, g("Thread Instructions Executed") / max(ex, 1), g("# Samples"), g("stall_no_inst"), g("stall_wait"),
        g("stall_branch_resolving"), g("stall_short_sb"), g("stall_long_sb"), g("stall_math"), top))
