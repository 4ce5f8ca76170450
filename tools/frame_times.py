#!/usr/bin/env python3
"""Development aid: per-frame device time of each stage for the staged bench clip."""
import ctypes as C, hashlib, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from bench import load_clip
from easyav1_b200.lib import load_engine, load_host, StagedInfo, StreamFrame
man, obu = load_clip(os.environ.get("EAV1_CLIP"))   # default: the bench clip
E, host = load_engine(), load_host()
cap = man["w"] * man["h"] * 4
buf = C.create_string_buffer(cap)
keep = C.create_string_buffer(obu, len(obu))
h = host.eav1_stream_open(keep, len(obu), 1, 7, 1)
eng = host.eav1rec_engine(host.eav1_stream_ctx(h))
E.eav1cu_set_capture(eng, 1)
info, used = StreamFrame(), C.c_size_t(0)
nmax = int(sys.argv[1]) if len(sys.argv) > 1 else 30
i = 0
while i < nmax and host.eav1_stream_next(h, C.byref(info), buf, cap, C.byref(used)) == 0:
    i += 1
# frames may still be in flight in the decoder's threads: let them reach the engine before replaying
host.eav1rec_drain(host.eav1_stream_ctx(h))
E.eav1cu_set_capture(eng, 0)
n = E.eav1cu_staged_count(eng)
E.eav1cu_set_stage_timing(eng, 1)
ms = (C.c_double * 7)(); ln = (C.c_longlong * 7)(); calls = (C.c_longlong * 7)()
for rep in range(2):
    for k in range(n):
        E.eav1cu_stage_times(eng, None, None, None, 1)
        for _ in range(3):
            E.eav1cu_staged_replay(eng, k)
        E.eav1cu_stage_times(eng, ms, ln, calls, 1)
        si = StagedInfo(); E.eav1cu_staged_info(eng, k, C.byref(si))
        if rep == 1:
            print("frame %3d inter %d cmds %6d chains %5d (cta %5d warp %5d) levels %4d | mc %7.1f resid %6.1f recon %8.1f us deblock %6.1f cdef %6.1f lr %6.1f"
                  % (k, si.inter, si.n_cmds, si.n_chains, si.n_big, si.n_small, si.n_levels,
                     1e3 * ms[4] / 3, 1e3 * ms[6] / 3, 1e3 * ms[0] / 3, 1e3 * ms[1] / 3, 1e3 * ms[2] / 3, 1e3 * ms[3] / 3))
host.eav1_stream_close(h)
