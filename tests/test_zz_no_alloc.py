"""eav1cu_apply_grain_noalloc (include/easyav1_cuda.h): the film-grain entry point whose scratch is a
file-scope static __device__ array, for callers that cannot call cudaMalloc. It has no counterpart in
the reference; the check is against the existing entry point eav1cu_apply_grain (itself bit-exact
against dav1d_apply_grain_{8,16}bpc in tests/test_filmgrain.py) and against the reference operator,
both bit-exact like every other integer test here, on two calls in a row per picture (the second call
finds the first call's contents in the shared array) and across geometries / bit depths (so a call
also finds the leftovers of a different layout). Named test_zz_* so that pytest collects it last.
The allocation claim is checked too: free device memory must not drop across the calls."""
import ctypes as C

import numpy as np
import pytest

from tests import ops
from tests.test_filmgrain import random_grain

CASES = [
    # (w, h, layout, bpc)
    (352, 288, 1, 8), (353, 287, 1, 10), (130, 70, 3, 8), (97, 65, 2, 10), (64, 33, 0, 8),
    (129, 66, 3, 12), (1920, 1080, 1, 8), (3840, 2160, 1, 10),
]


def _free_bytes():
    import torch            # plumbing only: cudaMemGetInfo on device 0's primary context
    return torch.cuda.mem_get_info(0)[0]


@pytest.mark.gpu
def test_apply_grain_noalloc_matches_apply_grain():
    eng = ops.Engine(ops.CUDA_SO)
    E = eng.E
    E.eav1cu_apply_grain_noalloc.restype = C.c_int
    E.eav1cu_apply_grain_noalloc.argtypes = [C.c_void_p, C.c_int, C.c_int, C.POINTER(ops.GrainParams)]
    try:
        for ci, (w, h, layout, bpc) in enumerate(CASES):
            dt = np.uint16 if bpc > 8 else np.uint8
            src = eng.pic_alloc(w, h, layout, bpc)
            want_pic = eng.pic_alloc(w, h, layout, bpc)
            got_pic = eng.pic_alloc(w, h, layout, bpc)
            for call in range(2):       # two calls in a row, different pictures and parameters
                rng = np.random.default_rng(7000 + 10 * ci + call)
                planes = [rng.integers(0, 1 << bpc, size=(ph, pw)).astype(dt)
                          for pw, ph in ops.plane_dims(w, h, layout)]
                g = random_grain(rng, bpc, layout)
                eng.pic_upload(src, planes)
                eng.apply_grain(src, want_pic, g)
                want = eng.pic_download(want_pic, w, h, layout, bpc)
                eng.check(E.eav1cu_sync(eng.ctx))
                before = _free_bytes()
                eng.check(E.eav1cu_apply_grain_noalloc(eng.ctx, src, got_pic, C.byref(g)))
                eng.check(E.eav1cu_sync(eng.ctx))
                if ci or call:   # the very first call may load the module lazily
                    assert _free_bytes() >= before, "device memory was allocated by the no-alloc entry point"
                got = eng.pic_download(got_pic, w, h, layout, bpc)
                ref = ops.ref_apply_grain(planes, w, h, layout, bpc, g, g.mtrx_identity)
                for pl, (x, y, z) in enumerate(zip(got, want, ref)):
                    assert np.array_equal(x, y), "case %s call %d plane %d: %d samples differ from eav1cu_apply_grain" % (
                        (w, h, layout, bpc), call, pl, int((x != y).sum()))
                    assert np.array_equal(x, z), "case %s call %d plane %d differs from the reference" % (
                        (w, h, layout, bpc), call, pl)
            for p in (src, want_pic, got_pic):
                E.eav1cu_pic_release(eng.ctx, p)
    finally:
        eng.close()


@pytest.mark.gpu
def test_apply_grain_noalloc_rejects_oversize():
    """More 32x32 luma blocks than the static array has room for: a clear error, no launch."""
    eng = ops.Engine(ops.CUDA_SO)
    E = eng.E
    E.eav1cu_apply_grain_noalloc.restype = C.c_int
    E.eav1cu_apply_grain_noalloc.argtypes = [C.c_void_p, C.c_int, C.c_int, C.POINTER(ops.GrainParams)]
    try:
        w, h = 16384, 2208          # 512 x 69 = 35,328 blocks > EAV1CU_NOALLOC_MAX_GRAIN_BLOCKS (34,816)
        a = eng.pic_alloc(w, h, 0, 8)
        b = eng.pic_alloc(w, h, 0, 8)
        g = random_grain(np.random.default_rng(1), 8, 0)
        r = E.eav1cu_apply_grain_noalloc(eng.ctx, a, b, C.byref(g))
        assert r == -3              # EAV1CU_ERR_INVAL
        msg = C.create_string_buffer(256)
        E.eav1cu_last_error(eng.ctx, msg, 256)
        assert b"EAV1CU_NOALLOC_MAX_GRAIN_BLOCKS" in msg.value
    finally:
        eng.close()
