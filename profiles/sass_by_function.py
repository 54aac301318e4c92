#!/usr/bin/env python
"""Attribute the SASS of one profiled kernel (ncu source page) to the source FUNCTIONS it was inlined from.

    python profiles/sass_by_function.py <report.ncu-rep> [launch]

The bitboard helpers (tafl_bb.cuh) are inlined everywhere, so a per-line view cannot say which rule they were working
for.  This walks the instructions in address order and gives every instruction the context of the nearest preceding
instruction that comes from tafl_core.cuh / tafl_select.cuh / tafl_engine.cu, then sums executed warp instructions,
active lanes and stall samples per enclosing function of that context line."""
import collections, csv, io, os, re, subprocess, sys
rep = sys.argv[1]; which = int(sys.argv[2]) if len(sys.argv) > 2 else 0
CSRC = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "alphazeroforhnefatafl_b200", "csrc")
CTX_FILES = ("tafl_core.cuh", "tafl_select.cuh", "tafl_engine.cu", "tafl_mcts.cu", "tafl_mcts.cuh", "tafl_wide.cuh")
def functions(fname):
    out = []   # (first line, name)
    pat = re.compile(r"^(?:TAFL_HD|TAFL_COLD|__global__|__device__|static|template)?.*?\b([A-Za-z_][A-Za-z0-9_]*)\s*\((?:const|tafl|int|uint|BB|Ctx|Mcts|unsigned|double|float|bool)")
    try:
        lines = open(os.path.join(CSRC, fname)).read().split("\n")
    except OSError:
        return out
    for i, l in enumerate(lines, 1):
        if l and not l[0].isspace() and not l.startswith(("//", "#", "}", "template", "using", "struct", "namespace", "constexpr", "enum")):
            m = pat.match(l)
            if m: out.append((i, m.group(1)))
    return out
FUNCS = {f: functions(f) for f in CTX_FILES}
def func_of(fname, line):
    name = "?"
    for first, n in FUNCS.get(fname, []):
        if first <= line: name = n
        else: break
    return name
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
inst, fname, cur, hdr, launch = [], None, None, None, -1
for r in csv.reader(io.StringIO(out)):
    if not r: continue
    if r[0] == "File Path":
        fname = r[1].split("/")[-1]; continue
    if r[0] == "Function Name":
        continue
    if r[0] == "Line No":
        hdr = r
        if fname is not None and not inst or True: pass
        continue
    if hdr is None: continue
    if r[0].isdigit():
        cur = (fname, int(r[0])); continue
    if len(r) > 7 and r[2].startswith("0x"):
        f = lambda x: float(x) if x not in ("", "-") else 0.0
        inst.append((int(r[2], 16), cur, f(r[6]), f(r[7]), f(r[8])))
# several launches repeat the same addresses: keep launch `which`
by_addr = collections.OrderedDict()
seen = collections.Counter()
for a, c, s, e, t in inst:
    k = seen[a]; seen[a] += 1
    if k == which: by_addr[a] = (c, s, e, t)
ctx, agg, order = None, collections.defaultdict(lambda: [0.0, 0.0, 0.0, 0]), []
addrs = sorted(by_addr)
first_ctx = next((by_addr[a][0] for a in addrs if by_addr[a][0] and by_addr[a][0][0] in CTX_FILES), None)
ctx = first_ctx
for a in addrs:
    c, s, e, t = by_addr[a]
    if c and c[0] in CTX_FILES: ctx = c
    key = (ctx[0], func_of(*ctx)) if ctx else ("?", "?")
    if key not in agg: order.append(key)
    g = agg[key]; g[0] += s; g[1] += e; g[2] += t; g[3] += 1
tot_s = sum(v[0] for v in agg.values()) or 1; tot_e = sum(v[1] for v in agg.values()) or 1
print("%-18s %-28s %7s %10s %7s %7s %7s %6s" % ("file", "function", "static", "warp_inst", "share", "lanes", "samples", "share"))
for key in sorted(agg, key=lambda k: -agg[k][0]):
    s, e, t, n = agg[key]
    print("%-18s %-28s %7d %10d %7.3f %7.1f %7d %6.3f" % (key[0], key[1], n, e, e / tot_e, t / e if e else 0, s, s / tot_s))
