"""ctypes access to oracle/_ref/libdav1d_ref.so (the reference's own dav1d C build).

TEST INFRASTRUCTURE: only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs may import this.
"""
import ctypes as C
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_SO = os.path.join(ROOT, "oracle", "_ref", "libdav1d_ref.so")


class RefFrameInfo(C.Structure):
    _fields_ = [("w", C.c_int), ("h", C.c_int), ("layout", C.c_int), ("bpc", C.c_int),
                ("frame_type", C.c_int), ("show_frame", C.c_int), ("showable", C.c_int),
                ("offset", C.c_int64)]


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(REF_SO):
            raise RuntimeError("oracle/_ref/libdav1d_ref.so missing: run `make -C oracle ref` "
                               "in the build container (needs /root/reference)")
        _lib = C.CDLL(REF_SO)
        _lib.ref_dec_open.restype = C.c_void_p
        _lib.ref_dec_open.argtypes = [C.c_char_p, C.c_size_t, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int]
        _lib.ref_dec_next.argtypes = [C.c_void_p, C.POINTER(RefFrameInfo), C.c_void_p, C.c_size_t,
                                      C.POINTER(C.c_size_t)]
        _lib.ref_dec_close.argtypes = [C.c_void_p]
        _lib.ref_table.restype = C.c_void_p
        _lib.ref_table.argtypes = [C.c_char_p, C.POINTER(C.c_size_t)]
    return _lib


def decode_all(obu: bytes, n_threads=1, max_frame_delay=1, apply_grain=1, inloop_filters=7):
    """Yield (info, packed visible Y|U|V bytes) for every output frame."""
    L = lib()
    keep = C.create_string_buffer(obu, len(obu))
    d = L.ref_dec_open(keep, len(obu), n_threads, max_frame_delay, apply_grain, inloop_filters, 1)
    if not d:
        raise RuntimeError("ref_dec_open failed")
    try:
        cap = 1 << 20
        buf = C.create_string_buffer(cap)
        while True:
            info = RefFrameInfo()
            used = C.c_size_t(0)
            r = L.ref_dec_next(d, C.byref(info), buf, cap, C.byref(used))
            if r == 1:
                break
            if r < 0:
                raise RuntimeError("reference decoder error %d" % r)
            if used.value > cap:
                raise RuntimeError("internal: frame larger than buffer on first sight; "
                                   "pre-size with frame_bytes()")
            yield info, buf.raw[:used.value]
    finally:
        L.ref_dec_close(d)


def decode_all_big(obu: bytes, cap_bytes: int, **kw):
    """Same as decode_all but with an explicit per-frame capacity (1080p/4K)."""
    L = lib()
    keep = C.create_string_buffer(obu, len(obu))
    d = L.ref_dec_open(keep, len(obu), kw.get("n_threads", 1), kw.get("max_frame_delay", 1),
                       kw.get("apply_grain", 1), kw.get("inloop_filters", 7), 1)
    if not d:
        raise RuntimeError("ref_dec_open failed")
    try:
        buf = C.create_string_buffer(cap_bytes)
        while True:
            info = RefFrameInfo()
            used = C.c_size_t(0)
            r = L.ref_dec_next(d, C.byref(info), buf, cap_bytes, C.byref(used))
            if r == 1:
                break
            if r < 0:
                raise RuntimeError("reference decoder error %d" % r)
            assert used.value <= cap_bytes
            yield info, buf.raw[:used.value]
    finally:
        L.ref_dec_close(d)


def table(name: str) -> bytes:
    L = lib()
    sz = C.c_size_t(0)
    p = L.ref_table(name.encode(), C.byref(sz))
    if not p:
        raise KeyError(name)
    return C.string_at(p, sz.value)
