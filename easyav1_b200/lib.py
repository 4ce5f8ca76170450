"""ctypes loaders for the two in-tree shared libraries (built by __graft_entry__.build())."""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
ENGINE_SO = os.path.join(HERE, "libeasyav1_cuda.so")
HOST_SO = os.path.join(HERE, "libdav1d_b200.so")


class EngineError(RuntimeError):
    pass


_engine = None
_host = None


def load_engine():
    """libeasyav1_cuda.so (include/easyav1_cuda.h). Raises if it is missing: no fallback."""
    global _engine
    if _engine is None:
        if not os.path.exists(ENGINE_SO):
            raise EngineError("libeasyav1_cuda.so is not built (run __graft_entry__.build()); "
                              "there is no CPU fallback")
        _engine = C.CDLL(ENGINE_SO, mode=C.RTLD_GLOBAL)
        E = _engine
        E.eav1cu_launch_count.restype = C.c_longlong
        E.eav1cu_launch_count.argtypes = [C.c_void_p]
        E.eav1cu_counters.argtypes = [C.c_void_p, C.POINTER(C.c_longlong)]
        E.eav1cu_set_capture.argtypes = [C.c_void_p, C.c_int]
        E.eav1cu_staged_count.argtypes = [C.c_void_p]
        E.eav1cu_staged_replay.argtypes = [C.c_void_p, C.c_int]
        E.eav1cu_staged_replay_all.argtypes = [C.c_void_p, C.c_int]
        E.eav1cu_staged_graph_build.argtypes = [C.c_void_p]
        E.eav1cu_staged_graph_build_grain.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        E.eav1cu_staged_graph_build_grain.restype = C.c_int
        E.eav1cu_staged_graph_launch.argtypes = [C.c_void_p, C.c_int]
        E.eav1cu_staged_info.argtypes = [C.c_void_p, C.c_int, C.POINTER(StagedInfo)]
        E.eav1cu_set_stream.argtypes = [C.c_void_p, C.c_void_p]
        E.eav1cu_set_stage_timing.argtypes = [C.c_void_p, C.c_int]
        E.eav1cu_stage_times.argtypes = [C.c_void_p, C.POINTER(C.c_double), C.POINTER(C.c_longlong),
                                         C.POINTER(C.c_longlong), C.c_int]
        E.eav1cu_sync.argtypes = [C.c_void_p]
        E.eav1cu_pic_download.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_void_p),
                                          C.POINTER(C.c_ssize_t)]
        E.eav1cu_last_error.argtypes = [C.c_void_p, C.c_char_p, C.c_size_t]
    return _engine


def load_host(path=None):
    """libdav1d_b200.so: dav1d public API + eav1_stream_* helpers, bound to the CUDA engine."""
    global _host
    if path is not None:
        return _bind_host(C.CDLL(path, mode=C.RTLD_GLOBAL))
    if _host is None:
        load_engine()
        if not os.path.exists(HOST_SO):
            raise EngineError("libdav1d_b200.so is not built (run __graft_entry__.build())")
        _host = _bind_host(C.CDLL(HOST_SO, mode=C.RTLD_GLOBAL))
    return _host


class StagedInfo(C.Structure):
    _fields_ = [("cur", C.c_int), ("w", C.c_int), ("h", C.c_int), ("bpc", C.c_int), ("layout", C.c_int),
                ("n_levels", C.c_int), ("n_chains", C.c_int), ("launches", C.c_int),
                ("n_lf", C.c_int), ("n_lr", C.c_int), ("cdef", C.c_int), ("inter", C.c_int), ("n_big", C.c_int), ("n_small", C.c_int), ("n_mc_items", C.c_int), ("n_resid_items", C.c_int),
                ("n_cmds", C.c_longlong), ("payload_bytes", C.c_longlong),
                ("mc_px", C.c_longlong), ("unit_px", C.c_longlong), ("resid_px", C.c_longlong)]


class StreamFrame(C.Structure):
    _fields_ = [("w", C.c_int), ("h", C.c_int), ("layout", C.c_int), ("bpc", C.c_int),
                ("frame_type", C.c_int), ("show_frame", C.c_int), ("showable", C.c_int),
                ("offset", C.c_int64)]


def _bind_host(L):
    L.eav1_stream_open.restype = C.c_void_p
    L.eav1_stream_open.argtypes = [C.c_char_p, C.c_size_t, C.c_int, C.c_int, C.c_int]
    L.eav1_stream_next.argtypes = [C.c_void_p, C.POINTER(StreamFrame), C.c_void_p, C.c_size_t,
                                   C.POINTER(C.c_size_t)]
    L.eav1_stream_close.argtypes = [C.c_void_p]
    L.eav1_stream_ctx.restype = C.c_void_p
    L.eav1_stream_ctx.argtypes = [C.c_void_p]
    L.eav1rec_frames_submitted.restype = C.c_longlong
    L.eav1rec_frames_submitted.argtypes = [C.c_void_p]
    L.eav1rec_launches.restype = C.c_longlong
    L.eav1rec_launches.argtypes = [C.c_void_p]
    L.eav1rec_error.argtypes = [C.c_void_p]
    L.eav1rec_engine.restype = C.c_void_p
    L.eav1rec_engine.argtypes = [C.c_void_p]
    L.eav1rec_drain.restype = None
    L.eav1rec_drain.argtypes = [C.c_void_p]
    L.eav1_pump_run.argtypes = [C.POINTER(C.c_char_p), C.POINTER(C.c_size_t), C.c_int, C.c_int, C.c_int, C.c_int,
                                C.POINTER(C.c_void_p), C.c_size_t, C.POINTER(C.c_int), C.POINTER(C.c_size_t),
                                C.c_void_p, C.c_void_p]
    return L


def pump_run(host, obus, n_workers, cap, apply_grain=1, inloop_filters=7, bufs=None):
    """host/pump.c: decode len(obus) streams to the end on n_workers threads (every decoder single-threaded, the
    workers multiplex them). Returns (frames per stream, packed planes of each stream's last picture)."""
    n = len(obus)
    keep = [C.create_string_buffer(o, len(o)) if isinstance(o, (bytes, bytearray)) else o for o in obus]
    ptrs = (C.c_char_p * n)(*[C.cast(k, C.c_char_p) for k in keep])
    sizes = (C.c_size_t * n)(*[len(o) for o in obus])
    bufs = bufs or [C.create_string_buffer(cap) for _ in range(n)]
    outs = (C.c_void_p * n)(*[C.addressof(b) for b in bufs])
    frames = (C.c_int * n)()
    used = (C.c_size_t * n)()
    r = host.eav1_pump_run(ptrs, sizes, n, n_workers, apply_grain, inloop_filters, outs, cap, frames, used, None, None)
    if r < 0:
        raise EngineError("eav1_pump_run failed: %d" % r)
    return list(frames), [bufs[i].raw[:used[i]] for i in range(n)]
