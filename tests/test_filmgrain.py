"""Film grain synthesis: eav1cu_apply_grain against the reference's
dav1d_apply_grain_{8,16}bpc (oracle/_ref) on seeded random pictures and Dav1dFilmGrainData,
every layout and bit depth, odd sizes. The CPU variant drives the host emulation of the same
device code (tests/emu), the GPU variant the CUDA kernels; both must be bit-exact."""
import numpy as np
import pytest

from tests import ops


def random_grain(rng, bpc, layout, cfl=None, overlap=None, clip=None):
    g = ops.GrainParams()
    g.seed = int(rng.integers(0, 1 << 16))

    def points(arr, n):
        xs = np.sort(rng.choice(256, size=n, replace=False))
        for i, x in enumerate(xs):
            arr[i][0] = int(x)
            arr[i][1] = int(rng.integers(0, 256))

    g.num_y_points = int(rng.integers(0, 15))
    points(g.y_points, g.num_y_points)
    g.chroma_scaling_from_luma = int(rng.integers(0, 2)) if cfl is None else cfl
    if layout == 0:
        g.chroma_scaling_from_luma = 0
    for i in range(2):
        n = 0 if (g.chroma_scaling_from_luma or layout == 0) else int(rng.integers(0, 11))
        g.num_uv_points[i] = n
        points(g.uv_points[i], n)
        g.uv_mult[i] = int(rng.integers(-128, 128))
        g.uv_luma_mult[i] = int(rng.integers(-128, 128))
        g.uv_offset[i] = int(rng.integers(-256, 256))
        for k in range(25):
            g.ar_coeffs_uv[i][k] = int(rng.integers(-128, 128))
    for k in range(24):
        g.ar_coeffs_y[k] = int(rng.integers(-128, 128))
    g.scaling_shift = int(rng.integers(8, 12))
    g.ar_coeff_lag = int(rng.integers(0, 4))
    g.ar_coeff_shift = int(rng.integers(6, 10))
    g.grain_scale_shift = int(rng.integers(0, 4))
    g.overlap_flag = int(rng.integers(0, 2)) if overlap is None else overlap
    g.clip_to_restricted_range = int(rng.integers(0, 2)) if clip is None else clip
    g.mtrx_identity = int(rng.integers(0, 2))
    return g


CASES = [
    # (w, h, layout, bpc)
    (352, 288, 1, 8), (353, 287, 1, 8), (130, 70, 3, 8), (97, 65, 2, 8), (64, 33, 0, 8),
    (352, 288, 1, 10), (131, 67, 1, 10), (129, 66, 3, 12), (99, 64, 2, 10), (33, 31, 1, 12),
]


def run_cases(engine_so, cases, reps):
    eng = ops.Engine(engine_so)
    try:
        for ci, (w, h, layout, bpc) in enumerate(cases):
            for rep in range(reps):
                rng = np.random.default_rng(1000 * ci + rep)
                dt = np.uint16 if bpc > 8 else np.uint8
                planes = [rng.integers(0, 1 << bpc, size=(ph, pw)).astype(dt) for pw, ph in ops.plane_dims(w, h, layout)]
                g = random_grain(rng, bpc, layout, cfl=rep & 1 if rep < 2 else None)
                a = eng.pic_alloc(w, h, layout, bpc)
                b = eng.pic_alloc(w, h, layout, bpc)
                eng.pic_upload(a, planes)
                eng.apply_grain(a, b, g)
                got = eng.pic_download(b, w, h, layout, bpc)
                want = ops.ref_apply_grain(planes, w, h, layout, bpc, g, g.mtrx_identity)
                for pl, (x, y) in enumerate(zip(got, want)):
                    assert np.array_equal(x, y), "case %s rep %d plane %d: %d samples differ" % (
                        (w, h, layout, bpc), rep, pl, int((x != y).sum()))
                eng.E.eav1cu_pic_release(eng.ctx, a)
                eng.E.eav1cu_pic_release(eng.ctx, b)
    finally:
        eng.close()


def test_filmgrain_emulated_device_code():
    run_cases(ops.EMU_SO, CASES, 3)


@pytest.mark.gpu
def test_filmgrain_cuda():
    run_cases(ops.CUDA_SO, CASES, 6)
    run_cases(ops.CUDA_SO, [(1920, 1080, 1, 8), (3840, 2160, 1, 10)], 1)
