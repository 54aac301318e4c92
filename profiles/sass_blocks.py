#!/usr/bin/env python
"""Runs of SASS instructions with the same execution count (~ basic blocks) of one profiled kernel launch:
static range, executions per instruction, active lanes per instruction and the non-inlined source lines they map to.
    python profiles/sass_blocks.py <report.ncu-rep> [min_total_exec]"""
import collections, csv, io, subprocess, sys
rep = sys.argv[1]; floor = int(sys.argv[2]) if len(sys.argv) > 2 else 20000
def page(extra):
    return list(csv.reader(io.StringIO(subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", *extra], capture_output=True, text=True).stdout)))
addr2src, fname, cur = {}, None, None
for r in page(["--print-source", "cuda,sass"]):
    if not r: continue
    if r[0] == "File Path": fname = r[1].split("/")[-1]; continue
    if r[0].isdigit(): cur = (fname, int(r[0])); continue
    if len(r) > 3 and r[2].startswith("0x"): addr2src.setdefault(r[2], cur)
hdr, L = None, []
for r in page([]):
    if r and r[0] == "Address":
        if hdr: break
        hdr = r; continue
    if hdr and len(r) == len(hdr) and r[0].startswith("0x"): L.append(r)
ie, it = hdr.index("Instructions Executed"), hdr.index("Thread Instructions Executed")
f = lambda x: float(x) if x not in ("", "-") else 0.0
tot = sum(f(r[ie]) for r in L)
print("static %d executed %d" % (len(L), tot))
prev, start, acc, thr, lines = None, 0, 0.0, 0.0, collections.Counter()
def flush(i):
    if prev is not None and acc >= floor:
        tops = ["%s:%d" % (k[0].replace("tafl_", ""), k[1]) for k, _ in lines.most_common(8) if k and not k[0].endswith("bb.cuh")][:5]
        print("static %4d-%4d n=%3d exec/instr %7d lanes %4.1f share %4.1f%% | %s" % (start, i, i - start, prev, thr, 100 * acc / tot, " ".join(tops)))
for i, r in enumerate(L):
    e = f(r[ie])
    if prev is None or abs(e - prev) > 0.02 * max(prev, 1):
        flush(i); prev, start, acc, lines = e, i, 0.0, collections.Counter(); thr = f(r[it]) / max(e, 1)
    acc += e; lines[addr2src.get(r[0])] += 1
flush(len(L))
