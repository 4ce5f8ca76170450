"""SURVEY.md N4 — output conversion on the device as an ADDITIONAL getter (eav1cu_pic_convert,
eav1rec_picture_convert, easyav1_b200_get_video_frame_converted). The reference has no code for these formats (its
consumers convert after decode: tools/easyav1_player.c:1338), so the checker is the numpy restatement below of the
integer formulas documented in include/easyav1_cuda.h: PARITY UNPINNED for this row. Bit-exact against that
restatement for NV12 / P010 / RGBA8 over layouts, bit depths, matrices, ranges and odd sizes, on the host emulation
and on the GPU; and through the easyAV1 API on a WebM fixture (the converted frame must equal the conversion of the
planes easyav1_get_video_frame hands out, which themselves are MD5-pinned to the reference)."""
import ctypes as C
import json
import os
import subprocess
import sys

import numpy as np
import pytest

from tests import ops

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
NV12, P010, RGBA8 = 1, 2, 3


class Convert(C.Structure):
    _fields_ = [("format", C.c_int32), ("matrix", C.c_int32), ("full_range", C.c_int32)]


def coefficients(bpc, matrix, full_range):
    kr = 0.2126 if matrix == 1 else 0.2627 if matrix == 9 else 0.299
    kb = 0.0722 if matrix == 1 else 0.0593 if matrix == 9 else 0.114
    kg = 1.0 - kr - kb
    sh = bpc - 8
    yr = float((255 if full_range else 219) << sh)
    cr = float((255 if full_range else 224) << sh)
    q = lambda v: int(v + 0.5)
    return (q(8192.0 * 255.0 / yr), q(8192.0 * 255.0 * 2.0 * (1.0 - kr) / cr), q(8192.0 * 255.0 * 2.0 * (1.0 - kb) * kb / kg / cr),
            q(8192.0 * 255.0 * 2.0 * (1.0 - kr) * kr / kg / cr), q(8192.0 * 255.0 * 2.0 * (1.0 - kb) / cr),
            0 if full_range else 16 << sh, 128 << sh)


def want_rgba(planes, layout, bpc, matrix, full_range):
    ky, krv, kgu, kgv, kbu, yoff, coff = coefficients(bpc, matrix, full_range)
    y = planes[0].astype(np.int64)
    h, w = y.shape
    if layout == 0:
        u = v = np.zeros_like(y)
    else:
        sx, sy = int(layout != 3), int(layout == 1)
        yy, xx = np.mgrid[0:h, 0:w]
        u = planes[1].astype(np.int64)[yy >> sy, xx >> sx] - coff
        v = planes[2].astype(np.int64)[yy >> sy, xx >> sx] - coff
    base = ky * (y - yoff) + 4096
    r = np.clip((base + krv * v) >> 13, 0, 255)
    g = np.clip((base - kgu * u - kgv * v) >> 13, 0, 255)
    b = np.clip((base + kbu * u) >> 13, 0, 255)
    return np.stack([r, g, b, np.full_like(r, 255)], axis=-1).astype(np.uint8)


def want_semiplanar(planes, bpc):
    h, w = planes[0].shape
    ch, cw = planes[1].shape
    up = 0 if bpc == 8 else 16 - bpc
    dt = np.uint8 if bpc == 8 else np.uint16
    luma = (planes[0].astype(np.uint32) << up).astype(dt)
    uv = np.zeros((ch, 2 * cw), dt)
    uv[:, 0::2] = (planes[1].astype(np.uint32) << up).astype(dt)
    uv[:, 1::2] = (planes[2].astype(np.uint32) << up).astype(dt)
    return luma, uv


def run(engine_so, seeds):
    eng = ops.Engine(engine_so)
    E = eng.E
    E.eav1cu_pic_convert.restype = C.c_int
    E.eav1cu_pic_convert.argtypes = [C.c_void_p, C.c_int, C.POINTER(Convert), C.c_void_p, C.c_size_t, C.c_size_t]
    try:
        for seed in range(seeds):
            rng = np.random.default_rng(900 + seed)
            for layout in (0, 1, 2, 3):
                for bpc in (8, 10, 12):
                    w, h = int(rng.integers(17, 200)), int(rng.integers(9, 90))
                    dt = np.uint16 if bpc > 8 else np.uint8
                    planes = [rng.integers(0, 1 << bpc, size=(ph, pw)).astype(dt) for pw, ph in ops.plane_dims(w, h, layout)]
                    pid = eng.pic_alloc(w, h, layout, bpc)
                    eng.pic_upload(pid, planes)
                    # RGBA8: every layout / depth, a few matrices and both ranges
                    for matrix, full in ((1, 0), (6, 1), (9, 0), (2, int(rng.integers(0, 2)))):
                        stride = 4 * w + 4 * int(rng.integers(0, 3))
                        out = np.zeros((h, stride), np.uint8)
                        cv = Convert(RGBA8, matrix, full)
                        assert E.eav1cu_pic_convert(eng.ctx, pid, C.byref(cv), out.ctypes.data, stride, out.nbytes) == 0
                        got = out[:, :4 * w].reshape(h, w, 4)
                        assert np.array_equal(got, want_rgba(planes, layout, bpc, matrix, full)), (layout, bpc, matrix, full, w, h)
                    # NV12 / P010: 4:2:0 only, refused elsewhere
                    fmt, bps = (NV12, 1) if bpc == 8 else (P010, 2)
                    stride = bps * ((w + 1) & ~1) + 2 * int(rng.integers(0, 3))
                    rows = h + (h + 1) // 2
                    out = np.zeros((rows, stride), np.uint8)
                    cv = Convert(fmt, 0, 0)
                    r = E.eav1cu_pic_convert(eng.ctx, pid, C.byref(cv), out.ctypes.data, stride, out.nbytes)
                    if layout != 1:
                        assert r != 0, "semi-planar output must be refused for layout %d" % layout
                    else:
                        assert r == 0
                        luma, uv = want_semiplanar(planes, bpc)
                        view = out.view(np.uint16) if bps == 2 else out
                        assert np.array_equal(view[:h, :w], luma)
                        assert np.array_equal(view[h:, :uv.shape[1]], uv)
                        other = Convert(P010 if fmt == NV12 else NV12, 0, 0)
                        assert E.eav1cu_pic_convert(eng.ctx, pid, C.byref(other), out.ctypes.data, stride, out.nbytes) != 0
                        assert E.eav1cu_pic_convert(eng.ctx, pid, C.byref(cv), out.ctypes.data, stride, out.nbytes - 1) != 0, "short dst"
                    E.eav1cu_pic_release(eng.ctx, pid)
    finally:
        eng.close()


def test_convert_emulated_device_code():
    run(ops.EMU_SO, 2)


@pytest.mark.gpu
def test_convert_cuda():
    run(ops.CUDA_SO, 4)


API_SCRIPT = """
import ctypes as C, hashlib, json, sys
sys.path.insert(0, %r)
import numpy as np
from tests import easyav1_api as ea
from tests.test_convert import want_semiplanar, want_rgba
L = ea.bind(sys.argv[1]); webm = open(sys.argv[2], 'rb').read(); keep = C.create_string_buffer(webm, len(webm))
L.easyav1_b200_get_video_frame_converted.restype = C.c_int
L.easyav1_b200_get_video_frame_converted.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_size_t, C.c_size_t]
st = L.easyav1_default_settings(); st.enable_audio = 0; st.log_level = 1
h = L.easyav1_init_from_memory(keep, len(webm), C.byref(st))
n, ok = 0, True
while L.easyav1_decode_next(h) == 1:
    if L.easyav1_has_video_frame(h) == 1:
        fr = L.easyav1_get_video_frame(h).contents
        w, hh = fr.width, fr.height
        planes = []
        for pl, (pw, ph) in enumerate([(w, hh), ((w + 1) // 2, (hh + 1) // 2), ((w + 1) // 2, (hh + 1) // 2)]):
            rows = [np.frombuffer(C.string_at(fr.data[pl] + y * fr.stride[pl], pw), np.uint8) for y in range(ph)]
            planes.append(np.stack(rows))
        stride = (w + 1) & ~1
        nv = np.zeros((hh + (hh + 1) // 2, stride), np.uint8)
        ok &= L.easyav1_b200_get_video_frame_converted(h, 1, nv.ctypes.data, stride, nv.nbytes) == 1
        luma, uv = want_semiplanar(planes, 8)
        ok &= bool(np.array_equal(nv[:hh, :w], luma) and np.array_equal(nv[hh:, :uv.shape[1]], uv))
        # the fixture (aom through Pillow's AVIF writer) signals matrix 6 (BT.601) and full range in its sequence header
        rgba = np.zeros((hh, 4 * w), np.uint8)
        ok &= L.easyav1_b200_get_video_frame_converted(h, 3, rgba.ctypes.data, 4 * w, rgba.nbytes) == 1
        ok &= bool(np.array_equal(rgba.reshape(hh, w, 4), want_rgba(planes, 1, 8, fr.matrix_coefficients, 1)))
        ok &= L.easyav1_b200_get_video_frame_converted(h, 2, nv.ctypes.data, stride, nv.nbytes) != 1     # P010 of an 8-bit stream
        n += 1
hp = C.c_void_p(h); L.easyav1_destroy(C.byref(hp))
print(json.dumps({"frames": n, "ok": bool(ok)}))
""" % ROOT


def run_api(lib):
    path = os.path.join(ROOT, "tests", "golden", "c1_360p_webm.webm")
    out = subprocess.run([sys.executable, "-c", API_SCRIPT, lib, path], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    r = json.loads(out.stdout.strip().splitlines()[-1])
    assert r["frames"] == 48 and r["ok"], r


def test_converted_getter_over_emulated_engine():
    lib = os.path.join(ROOT, "tests", "emu", "libeasyav1_b200_emu.so")
    if not os.path.exists(lib):
        pytest.skip("tests/emu not built")
    run_api(lib)


@pytest.mark.gpu
def test_converted_getter_cuda():
    run_api(os.path.join(ROOT, "easyav1_b200", "libeasyav1_b200.so"))
