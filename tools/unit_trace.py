#!/usr/bin/env python3
"""Development aid: where does the dataflow kernel's time go for one staged frame?
Replays frame K of the bench clip with per-unit device timestamps (eav1cu_staged_trace) and walks
the chain of last-finishing producers back from the unit that finished last.
    python tools/unit_trace.py [frame=0] [verbose=0]"""
import ctypes as C, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import load_clip
from easyav1_b200.lib import load_engine, load_host, StagedInfo, StreamFrame
K = int(sys.argv[1]) if len(sys.argv) > 1 else 0
verbose = int(sys.argv[2]) if len(sys.argv) > 2 else 0
man, obu = load_clip()
E, host = load_engine(), load_host()
cap = man["w"] * man["h"] * 4
buf = C.create_string_buffer(cap)
keep = C.create_string_buffer(obu, len(obu))
h = host.eav1_stream_open(keep, len(obu), 1, 7, 1)
eng = host.eav1rec_engine(host.eav1_stream_ctx(h))
E.eav1cu_set_capture(eng, 1)
info, used = StreamFrame(), C.c_size_t(0)
i = 0
while i <= K + 2 and host.eav1_stream_next(h, C.byref(info), buf, cap, C.byref(used)) == 0:
    i += 1
E.eav1cu_set_capture(eng, 0)
host.eav1rec_drain(host.eav1_stream_ctx(h))
si = StagedInfo(); E.eav1cu_staged_info(eng, K, C.byref(si))
n = si.n_chains
W = 12
out = (C.c_ulonglong * (W * (n + 43)))()
E.eav1cu_staged_trace.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_ulonglong), C.c_size_t]
prof = (C.c_ulonglong * 64)()
for _ in range(3):
    E.eav1cu_prof_read(prof)
    r = E.eav1cu_staged_trace(eng, K, out, n + 43)
have_prof = E.eav1cu_prof_read(prof)
assert r == n, r
U = [tuple(out[W * u + j] for j in range(6)) for u in range(n)]
X = [tuple(out[W * u + j] for j in range(6, 12)) for u in range(n)]
t0 = min(u[0] for u in U); tend = max(u[2] for u in U)
print("frame %d: %d units, %d cmds, kernel span %.1f us" % (K, n, si.n_cmds, (tend - t0) / 1e3))
tot_exec = sum(u[2] - u[1] for u in U); tot_cmds = sum(u[4] & 0xffffffff for u in U)
print("sum of unit execution %.1f us over %d cmds = %.2f us/cmd; sum of waits %.1f us" %
      (tot_exec / 1e3, tot_cmds, tot_exec / 1e3 / max(tot_cmds, 1), sum(u[1] - u[0] for u in U) / 1e3))
# walk back from the last finisher
last = max(range(n), key=lambda u: U[u][2])
path = []
u = last
while u != 0xffffffffffffffff and u < n:
    path.append(u); u = U[u][3]
path.reverse()
ex = hop = cmds = 0
prev_done = None
for u in path:
    tk, rd, dn, who, fc, tw = U[u]
    c = fc & 0xffffffff
    e = dn - rd
    hp = (rd - prev_done) if prev_done is not None else (rd - t0)
    ex += e; hop += hp; cmds += c
    if verbose:
        print("  unit %5d ticket %5d cta %4d: %2d cmds exec %7.2f us (%.2f/cmd) hop %6.2f us ticket->ready %7.1f"
              % (u, tw >> 32, (tw >> 8) & 0xffffff, c, e / 1e3, e / 1e3 / c, hp / 1e3, (rd - tk) / 1e3))
    prev_done = dn
print("critical path: %d units, %d cmds, exec %.1f us (%.2f us/cmd), hops %.1f us (%.2f us/unit)" %
      (len(path), cmds, ex / 1e3, ex / 1e3 / max(cmds, 1), hop / 1e3, hop / 1e3 / len(path)))
# per-command cost by unit size
import collections
by = collections.defaultdict(lambda: [0, 0, 0])
for tk, rd, dn, who, fc, tw in U:
    c = fc & 0xffffffff
    b = 1 if c == 1 else 2 if c == 2 else 4 if c <= 4 else 8 if c <= 8 else 16 if c <= 16 else 64
    by[b][0] += 1; by[b][1] += c; by[b][2] += dn - rd
for b in sorted(by):
    k, c, t = by[b]
    print("  units with <=%2d cmds: %5d units, %6d cmds, %.2f us/cmd, %.2f us/unit" % (b, k, c, t / 1e3 / c, t / 1e3 / k))
st = sum(x[0] for x in X); ip = sum(x[1] for x in X); it = sum(x[2] for x in X); ot = sum(x[3] for x in X)
nip = sum(x[4] >> 32 for x in X); nit = sum(x[4] & 0xffffffff for x in X); no = sum(x[5] for x in X)
print("cycles: staging %.0f per cmd | ipred %.0f per cmd (%d) | itx %.0f per cmd (%d) | other %.0f per cmd (%d)" %
      (st / max(tot_cmds, 1), ip / max(nip, 1), nip, it / max(nit, 1), nit, ot / max(no, 1), no))
names = ["nop", "mc", "warp", "avg", "wavg", "mask", "wmask", "blend", "blendv", "blendh", "ipred", "cfl", "pal", "itx"]
print("cycles per command by op and log2(w*h):")
modes = ["dc", "vert", "hor", "left_dc", "top_dc", "dc128", "z1", "z2", "z3", "smooth", "smooth_v", "smooth_h", "paeth", "filter"]
for tab, label in ((15, "all sizes"), (14, ">= 256 samples")):
    print("ipred cycles by mode (%s): " % label + " | ".join("%s: %d x %.0f" % (modes[m], out[W * n + (tab * 16 + m) * 2 + 1],
          out[W * n + (tab * 16 + m) * 2] / out[W * n + (tab * 16 + m) * 2 + 1]) for m in range(14) if out[W * n + (tab * 16 + m) * 2 + 1]))
for op in range(14):
    cells = []
    for lg in range(16):
        cyc, cnt = out[W * n + (op * 16 + lg) * 2], out[W * n + (op * 16 + lg) * 2 + 1]
        if cnt:
            cells.append("2^%d: %d x %.0f" % (lg, cnt, cyc / cnt))
    if cells:
        print("  %-6s %s" % (names[op], " | ".join(cells)))
if have_prof:
    lab = {0: "exec_cmd entry->switch", 1: "switch->ipred case", 2: "build_edge", 3: "ipred_block", 4: "switch->itx case",
           5: "itx_add total", 6: "switch->avg case", 7: "comp_avg", 10: "itx reg: row pass", 11: "itx reg: barrier",
           12: "itx reg: col pass", 13: "itx reg: barrier", 20: "edge: entry", 21: "edge: left part", 22: "edge: top part",
           23: "edge: corner", 24: "edge: barrier", 25: "ipred_block: entry", 26: "ipred_block: mode body", 27: "ipred_block: barrier", 30: "probe: volatile global load", 31: "probe: ld.global.cg"}
    for k in sorted(lab):
        if prof[2 * k + 1]:
            print("  prof %-24s %8d x %6.0f cycles" % (lab[k], prof[2 * k + 1], prof[2 * k] / prof[2 * k + 1]))
host.eav1_stream_close(h)
