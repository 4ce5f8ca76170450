#!/usr/bin/env python3
"""Development aid: decode a fixture several times through the product library and report which
frames (if any) differ from the golden MD5s.  EAV1_EMU=1 uses the host emulation engine."""
import sys, os, ctypes as C, hashlib
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
from easyav1_b200.lib import load_host, StreamFrame
import json
name = sys.argv[1] if len(sys.argv) > 1 else "c1_360p"
man = json.load(open(os.path.join(ROOT, "tests/golden/manifest.json")))[name]
from tests.golden.gen_fixtures import read_ivf
obu = b"".join(read_ivf(os.path.join(ROOT, "tests/golden", man["file"])))
host = load_host(os.path.join(ROOT, "tests/emu/libdav1d_b200_emu.so")) if os.environ.get("EAV1_EMU") else load_host()
cap = man["w"] * man["h"] * 4
buf = C.create_string_buffer(cap)
for rep in range(int(sys.argv[2]) if len(sys.argv) > 2 else 5):
    import time
    keep = C.create_string_buffer(obu, len(obu))
    t0 = time.perf_counter()
    h = host.eav1_stream_open(keep, len(obu), 1, 7, 1)
    t1 = time.perf_counter(); tfirst = None
    info, used, n, bad = StreamFrame(), C.c_size_t(0), 0, []
    keepf = []
    fast = bool(os.environ.get("E2E_FAST"))
    while True:
        r = host.eav1_stream_next(h, C.byref(info), buf, cap, C.byref(used))
        if r: break
        if tfirst is None: tfirst = time.perf_counter()
        if fast: keepf.append(buf.raw[:used.value] if (os.environ["E2E_FAST"] == "2" or n % 8 == 7 or n >= len(man["frames"]) - 3) else None)
        elif n < len(man["frames"]) and hashlib.md5(buf.raw[:used.value]).hexdigest() != man["frames"][n]["md5"]: bad.append(n)
        n += 1
    t2 = time.perf_counter()
    host.eav1_stream_close(h)
    t3 = time.perf_counter()
    print("  open %.1f ms, first frame after %.1f ms, decode loop %.1f ms (%.1f fps), close %.1f ms" % (1e3*(t1-t0), 1e3*(tfirst-t1), 1e3*(t2-t1), n/(t2-t1), 1e3*(t3-t2)))
    for i, fb in enumerate(keepf):
        if fb is not None and hashlib.md5(fb).hexdigest() != man["frames"][i]["md5"]: bad.append(i)
    print("pass", rep, "frames", n, "r", r, "bad", bad[:10])
    if bad and os.environ.get("E2E_FAST") == "2":
        import numpy as np
        from tests import refdec
        ref = [p for _, p in refdec.decode_all_big(obu, cap, n_threads=8, max_frame_delay=0)]
        w, hh = man["w"], man["h"]
        cw, ch = (w + 1) // 2, (hh + 1) // 2
        golden = {fr["md5"]: k for k, fr in enumerate(man["frames"])}
        print("  bad frames equal to golden frame:", [(i, golden.get(hashlib.md5(keepf[i]).hexdigest())) for i in bad[:16]])
        a0 = np.frombuffer(keepf[bad[0]], np.uint8)[:w * hh].reshape(hh, w); b0 = np.frombuffer(ref[bad[0]], np.uint8)[:w * hh].reshape(hh, w)
        print("  luma 64x64-block map of frame", bad[0], "(. exact, digit = tenths of samples differing)")
        for by in range(0, hh, 64):
            print("   ", "".join("." if not (a0[by:by+64, bx:bx+64] != b0[by:by+64, bx:bx+64]).any() else str(min(9, int(10 * (a0[by:by+64, bx:bx+64] != b0[by:by+64, bx:bx+64]).mean()))) for bx in range(0, w, 64)))
        try:
            from PIL import Image
            Image.fromarray(a0[::2, ::2]).save(os.path.join(ROOT, "gpurun_out", "bad_frame_y.png"))
            Image.fromarray(b0[::2, ::2]).save(os.path.join(ROOT, "gpurun_out", "good_frame_y.png"))
            print("  unique luma values", len(np.unique(a0)), "mean", a0.mean(), "ref mean", b0.mean())
        except Exception as e:
            print("  png dump failed", e)
        for i in bad[:3]:
            a = np.frombuffer(keepf[i], np.uint8); b = np.frombuffer(ref[i], np.uint8)
            for name, off, pw, ph in (("Y", 0, w, hh), ("U", w * hh, cw, ch), ("V", w * hh + cw * ch, cw, ch)):
                d = (a[off:off + pw * ph] != b[off:off + pw * ph]).reshape(ph, pw)
                if d.any():
                    ys, xs = np.nonzero(d)
                    print("  frame", i, name, "diff px", int(d.sum()), "bbox x", xs.min(), xs.max(), "y", ys.min(), ys.max(),
                          "sample got/want", a[off + ys[0] * pw + xs[0]], b[off + ys[0] * pw + xs[0]])
        break
