"""Operator-level access to an engine library through the C ABI of include/easyav1_cuda.h
(libeasyav1_cuda.so on a GPU box; tests/emu/libeav1_emu.so, the host emulation of the same
device code, in the build container) next to the reference operators of oracle/_ref.
TEST INFRASTRUCTURE."""
import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CUDA_SO = os.path.join(ROOT, "easyav1_b200", "libeasyav1_cuda.so")
EMU_SO = os.path.join(ROOT, "tests", "emu", "libeav1_emu.so")
HOST_SO = os.path.join(ROOT, "easyav1_b200", "libdav1d_b200.so")


class GrainParams(C.Structure):
    """Eav1GrainParams (include/easyav1_cuda.h)"""
    _fields_ = [("seed", C.c_uint32), ("num_y_points", C.c_int32), ("y_points", (C.c_uint8 * 2) * 14),
                ("chroma_scaling_from_luma", C.c_int32), ("num_uv_points", C.c_int32 * 2),
                ("uv_points", ((C.c_uint8 * 2) * 10) * 2), ("scaling_shift", C.c_int32),
                ("ar_coeff_lag", C.c_int32), ("ar_coeffs_y", C.c_int8 * 24), ("ar_coeffs_uv", (C.c_int8 * 28) * 2),
                ("ar_coeff_shift", C.c_int32), ("grain_scale_shift", C.c_int32), ("uv_mult", C.c_int32 * 2),
                ("uv_luma_mult", C.c_int32 * 2), ("uv_offset", C.c_int32 * 2), ("overlap_flag", C.c_int32),
                ("clip_to_restricted_range", C.c_int32), ("mtrx_identity", C.c_int32)]


class RefGrainData(C.Structure):
    """Dav1dFilmGrainData (ext/dav1d/include/dav1d/headers.h:315-333)"""
    _fields_ = [("seed", C.c_uint), ("num_y_points", C.c_int), ("y_points", (C.c_uint8 * 2) * 14),
                ("chroma_scaling_from_luma", C.c_int), ("num_uv_points", C.c_int * 2),
                ("uv_points", ((C.c_uint8 * 2) * 10) * 2), ("scaling_shift", C.c_int), ("ar_coeff_lag", C.c_int),
                ("ar_coeffs_y", C.c_int8 * 24), ("ar_coeffs_uv", (C.c_int8 * 28) * 2), ("ar_coeff_shift", C.c_uint64),
                ("grain_scale_shift", C.c_int), ("uv_mult", C.c_int * 2), ("uv_luma_mult", C.c_int * 2),
                ("uv_offset", C.c_int * 2), ("overlap_flag", C.c_int), ("clip_to_restricted_range", C.c_int)]


def plane_dims(w, h, layout):
    """[(w, h)] of the planes of a picture (layout: 0 I400, 1 I420, 2 I422, 3 I444)"""
    if layout == 0:
        return [(w, h)]
    sx, sy = int(layout != 3), int(layout == 1)
    return [(w, h), ((w + sx) >> sx, (h + sy) >> sy), ((w + sx) >> sx, (h + sy) >> sy)]


class Engine:
    def __init__(self, so_path):
        if not os.path.exists(so_path):
            raise RuntimeError(so_path + " is not built")
        E = self.E = C.CDLL(so_path)
        H = self.H = C.CDLL(HOST_SO)
        for name in ("eav1cu_ctx_create", "eav1cu_set_tables", "eav1cu_set_masks", "eav1cu_pic_alloc",
                     "eav1cu_pic_release", "eav1cu_pic_upload", "eav1cu_pic_download", "eav1cu_apply_grain",
                     "eav1cu_sync", "eav1cu_frame_submit", "eav1cu_arena_upload", "eav1cu_arena_download"):
            getattr(E, name).restype = C.c_int
        E.eav1cu_ctx_create.argtypes = [C.POINTER(C.c_void_p), C.c_int]
        E.eav1cu_ctx_destroy.argtypes = [C.c_void_p]
        E.eav1cu_ctx_destroy.restype = None
        E.eav1cu_set_tables.argtypes = [C.c_void_p, C.c_void_p]
        E.eav1cu_set_masks.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
        E.eav1cu_pic_alloc.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int]
        E.eav1cu_pic_release.argtypes = [C.c_void_p, C.c_int]
        E.eav1cu_pic_upload.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_ssize_t)]
        E.eav1cu_pic_download.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_ssize_t)]
        E.eav1cu_apply_grain.argtypes = [C.c_void_p, C.c_int, C.c_int, C.POINTER(GrainParams)]
        E.eav1cu_sync.argtypes = [C.c_void_p]
        E.eav1cu_last_error.argtypes = [C.c_void_p, C.c_char_p, C.c_size_t]
        ctx = C.c_void_p()
        r = E.eav1cu_ctx_create(C.byref(ctx), 0)
        if r:
            raise RuntimeError("eav1cu_ctx_create failed: %d" % r)
        self.ctx = ctx
        tab = C.create_string_buffer(1 << 16)
        H.eav1rec_fill_tables.argtypes = [C.c_void_p]
        H.eav1rec_fill_tables.restype = None
        H.eav1rec_fill_tables(tab)
        self.check(E.eav1cu_set_tables(ctx, tab))
        blob = C.c_void_p()
        H.eav1rec_masks.argtypes = [C.POINTER(C.c_void_p)]
        H.eav1rec_masks.restype = C.c_size_t
        n = H.eav1rec_masks(C.byref(blob))
        self.check(E.eav1cu_set_masks(ctx, blob, n))

    def check(self, r):
        if r < 0:
            msg = C.create_string_buffer(256)
            self.E.eav1cu_last_error(self.ctx, msg, 256)
            raise RuntimeError("engine call failed (%d): %s" % (r, msg.value.decode()))
        return r

    def close(self):
        if self.ctx:
            self.E.eav1cu_ctx_destroy(self.ctx)
            self.ctx = None

    # pictures are lists of 2-D numpy arrays (uint8 or uint16), one per plane
    def pic_alloc(self, w, h, layout, bpc):
        return self.check(self.E.eav1cu_pic_alloc(self.ctx, w, h, layout, bpc))

    def _ptrs(self, planes):
        data = (C.c_void_p * 3)(*[p.ctypes.data for p in planes] + [None] * (3 - len(planes)))
        stride = (C.c_ssize_t * 2)(planes[0].strides[0], planes[1].strides[0] if len(planes) > 1 else 0)
        return data, stride

    def pic_upload(self, pid, planes):
        data, stride = self._ptrs(planes)
        self.check(self.E.eav1cu_pic_upload(self.ctx, pid, data, stride))

    def pic_download(self, pid, w, h, layout, bpc):
        dt = np.uint16 if bpc > 8 else np.uint8
        planes = [np.zeros((ph, pw), dt) for pw, ph in plane_dims(w, h, layout)]
        data, stride = self._ptrs(planes)
        self.check(self.E.eav1cu_pic_download(self.ctx, pid, data, stride))
        return planes

    def apply_grain(self, src, dst, g):
        self.check(self.E.eav1cu_apply_grain(self.ctx, src, dst, C.byref(g)))
        self.check(self.E.eav1cu_sync(self.ctx))


def ref_apply_grain(planes, w, h, layout, bpc, g, mtrx_identity=0):
    """dav1d_apply_grain_{8,16}bpc of the reference on caller-owned planes (oracle/ref_harness.c)"""
    from tests import refdec
    L = refdec.lib()
    L.ref_apply_grain.restype = None
    L.ref_apply_grain.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(RefGrainData),
                                  C.POINTER(C.c_void_p), C.POINTER(C.c_void_p), C.POINTER(C.c_ssize_t)]
    d = RefGrainData()
    for name, _ in GrainParams._fields_:
        if name == "mtrx_identity":
            continue
        v = getattr(g, name)
        if isinstance(v, C.Array):
            C.memmove(C.addressof(getattr(d, name)), C.addressof(v), C.sizeof(v))
        else:
            setattr(d, name, v)
    # the reference wants strides that are multiples of 32 pixels and pads the luma row in place
    bps = 2 if bpc > 8 else 1
    ins, outs = [], []
    for p in planes:
        pitch = ((p.shape[1] + 63) & ~63) + 64
        a = np.zeros((p.shape[0] + 2, pitch), p.dtype)
        a[:p.shape[0], :p.shape[1]] = p
        ins.append(a)
        outs.append(np.zeros_like(a))
    if len(planes) > 1:
        assert ins[1].strides[0] == ins[2].strides[0]
    di = (C.c_void_p * 3)(*[a.ctypes.data for a in ins] + [None] * (3 - len(ins)))
    do = (C.c_void_p * 3)(*[a.ctypes.data for a in outs] + [None] * (3 - len(outs)))
    st = (C.c_ssize_t * 2)(ins[0].strides[0], ins[1].strides[0] if len(ins) > 1 else 0)
    L.ref_apply_grain(bpc, w, h, layout, mtrx_identity, C.byref(d), do, di, st)
    return [o[:p.shape[0], :p.shape[1]].copy() for o, p in zip(outs, planes)]
