#!/usr/bin/env python3
"""Decode the named streams (fixtures of tests/golden/manifest.json and synthetic syn_* streams) in ONE
process with the engine's tool-coverage histogram on, check every frame against its golden MD5 and print
the histogram (struct Eav1Coverage, include/easyav1_cuda.h) as JSON. TEST INFRASTRUCTURE.

    EAV1_COVERAGE=1 python tools/coverage_run.py cuda|emu name [name ...]
"""
import ctypes as C
import hashlib
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ.setdefault("EAV1_COVERAGE", "1")


class Coverage(C.Structure):
    _fields_ = [("frames", C.c_uint64 * 3), ("layout", C.c_uint64 * 4),
                ("itx", ((C.c_uint64 * 17) * 19) * 3), ("itx_eob", (C.c_uint64 * 3) * 3),
                ("mc", (((((C.c_uint64 * 4) * 4) * 4) * 2) * 2) * 2), ("mc_wh", ((C.c_uint64 * 8) * 8) * 2),
                ("intrabc", C.c_uint64), ("warp", (C.c_uint64 * 2) * 2), ("comp", (C.c_uint64 * 4) * 2),
                ("w_mask_ss", C.c_uint64 * 3), ("blend", (C.c_uint64 * 3) * 2), ("ipred", (C.c_uint64 * 14) * 3),
                ("ipred_wh", (C.c_uint64 * 8) * 8), ("cfl", (C.c_uint64 * 14) * 2), ("pal", C.c_uint64 * 2),
                ("lf", (((C.c_uint64 * 3) * 2) * 2) * 2), ("lr", (C.c_uint64 * 5) * 2),
                ("cdef_frames", C.c_uint64 * 2), ("superres_frames", C.c_uint64 * 2), ("grain", (C.c_uint64 * 2) * 2)]


def to_py(x):
    if isinstance(x, C.Array):
        return [to_py(v) for v in x]
    return int(x)


def main():
    which, names = sys.argv[1], sys.argv[2:]
    from easyav1_b200.lib import load_host, load_engine
    from easyav1_b200.decoder import Decoder
    from tests.golden import synth_streams
    from tests.golden.gen_fixtures import read_ivf
    if which == "emu":
        host = load_host(os.path.join(ROOT, "tests/emu/libdav1d_b200_emu.so"))
        E = C.CDLL(os.path.join(ROOT, "tests/emu/libeav1_emu.so"))
    else:
        host = load_host()
        E = load_engine()
    man = json.load(open(os.path.join(ROOT, "tests/golden/manifest.json")))
    sman = synth_streams.load_manifest()
    n_frames = 0
    for name in names:
        if name.startswith("syn_"):
            obu, _ = synth_streams.build(name)
            want = sman[name]["frames"]
        else:
            obu = b"".join(read_ivf(os.path.join(ROOT, "tests/golden", man[name]["file"])))
            want = man[name]["frames"]
        d = Decoder(obu, host=host, max_frame_bytes=max(f["w"] * f["h"] for f in want) * 8)
        got = [hashlib.md5(p).hexdigest() for _, p in d]
        d.close()
        assert got == [f["md5"] for f in want], "%s: decoded frames differ from the golden MD5s" % name
        n_frames += len(got)
    cov = Coverage()
    E.eav1cu_abi_sizeof.restype = C.c_size_t
    assert E.eav1cu_abi_sizeof(6) == C.sizeof(Coverage), (E.eav1cu_abi_sizeof(6), C.sizeof(Coverage))
    E.eav1cu_coverage.argtypes = [C.POINTER(Coverage), C.c_int]
    on = E.eav1cu_coverage(C.byref(cov), 0)
    out = {f: to_py(getattr(cov, f)) for f, _ in Coverage._fields_}
    out["_on"] = on
    out["_streams"] = names
    out["_frames_checked"] = n_frames
    print(json.dumps(out))


if __name__ == "__main__":
    main()
