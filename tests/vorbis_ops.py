"""Vorbis float slice: ctypes access to the product's batched entry points
(libeasyav1_cuda.so, or its host emulation) and to the reference compiled from minivorbis.h
(oracle/_ref/libvorbis_ref.so, oracle/vorbis_harness.c). TEST INFRASTRUCTURE."""
import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
VREF_SO = os.path.join(ROOT, "oracle", "_ref", "libvorbis_ref.so")

_ref = None


def ref():
    global _ref
    if _ref is None:
        if not os.path.exists(VREF_SO):
            raise RuntimeError("oracle/_ref/libvorbis_ref.so missing: run `make -C oracle ref`")
        L = _ref = C.CDLL(VREF_SO)
        L.vref_mdct_backward.argtypes = [C.c_int, C.c_void_p, C.c_void_p]
        L.vref_window.restype = C.c_void_p
        L.vref_window.argtypes = [C.c_int]
        L.vref_open.restype = C.c_void_p
        L.vref_open.argtypes = [C.c_int, C.c_int, C.c_int]
        L.vref_block.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_long, C.c_long]
        L.vref_restart.argtypes = [C.c_void_p]
        L.vref_close.argtypes = [C.c_void_p]
    return _ref


def ref_window(n):
    p = ref().vref_window(n)
    return np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_float)), shape=(n // 2,)).copy()


def ref_mdct_backward(n, X):
    X = np.ascontiguousarray(X, np.float32)
    out = np.zeros(n, np.float32)
    assert ref().vref_mdct_backward(n, X.ctypes.data, out.ctypes.data) == 0
    return out


def make_batch(seed, channels, bs, n_blocks, p_long=0.8, peak=0.5):
    """Config C3: random long/short sequence, Gaussian spectra with a per-band decay,
    scaled so the PCM stays inside full scale like decoded audio."""
    rng = np.random.default_rng(seed)
    flags = (rng.random(n_blocks) < p_long).astype(np.uint8)
    parts = []
    for f in flags:
        m = bs[int(f)] // 2
        decay = 10.0 ** (-(np.arange(m) * 60.0 / m) / 20.0)
        x = rng.standard_normal((channels, m)) * decay * (peak / np.sqrt(m) * 4.0)
        parts.append(x.astype(np.float32).ravel())
    return flags, np.concatenate(parts)


def ref_decode(flags, spectra, channels, bs, handle=None):
    """vorbis_synthesis_blockin/pcmout over the batch; returns planar PCM [channels][n]"""
    L = ref()
    h = handle or L.vref_open(channels, bs[0], bs[1])
    cap = int(sum(bs[int(f)] // 2 for f in flags)) + bs[1]
    out = np.zeros((channels, cap), np.float32)
    pos = off = 0
    for f in flags:
        m = bs[int(f)] // 2
        blk = np.ascontiguousarray(spectra[off:off + channels * m])
        got = L.vref_block(h, int(f), blk.ctypes.data, out.ctypes.data, cap, pos)
        assert got >= 0, got
        pos += got
        off += channels * m
    if handle is None:
        L.vref_close(h)
    return out[:, :pos].copy()


class Vorbis:
    def __init__(self, so_path, channels, bs, windows=None):
        E = self.E = C.CDLL(so_path)
        E.eav1cu_vorbis_open.argtypes = [C.POINTER(C.c_void_p), C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
        E.eav1cu_vorbis_close.argtypes = [C.c_void_p]
        E.eav1cu_vorbis_close.restype = None
        E.eav1cu_vorbis_restart.argtypes = [C.c_void_p]
        E.eav1cu_vorbis_imdct_ola.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_size_t,
                                              C.c_int, C.POINTER(C.c_size_t)]
        self.channels, self.bs = channels, bs
        self.keep = windows
        h = C.c_void_p()
        w0 = windows[0].ctypes.data if windows else None
        w1 = windows[1].ctypes.data if windows else None
        r = E.eav1cu_vorbis_open(C.byref(h), 0, channels, bs[0], bs[1], w0, w1)
        if r:
            raise RuntimeError("eav1cu_vorbis_open failed: %d" % r)
        self.h = h

    def decode(self, flags, spectra, interleave=False):
        flags = np.ascontiguousarray(flags, np.uint8)
        spectra = np.ascontiguousarray(spectra, np.float32)
        cap = int(sum(self.bs[int(f)] // 2 for f in flags)) + self.bs[1]
        out = np.zeros((cap, self.channels) if interleave else (self.channels, cap), np.float32)
        n = C.c_size_t(0)
        r = self.E.eav1cu_vorbis_imdct_ola(self.h, flags.ctypes.data, len(flags), spectra.ctypes.data,
                                           out.ctypes.data, cap, int(interleave), C.byref(n))
        if r:
            raise RuntimeError("eav1cu_vorbis_imdct_ola failed: %d" % r)
        return out[:n.value].copy() if interleave else out[:, :n.value].copy()

    def restart(self):
        self.E.eav1cu_vorbis_restart(self.h)

    def close(self):
        if self.h:
            self.E.eav1cu_vorbis_close(self.h)
            self.h = None
