"""Vorbis IMDCT + window + overlap-add against the reference's own
mdct_backward / vorbis_synthesis_blockin compiled from minivorbis.h. Tolerance: 1e-5 max-abs on PCM
inside full scale (the transform uses a different butterfly order than the reference, so this is
a floating-point tolerance, not bit equality)."""
import numpy as np
import pytest

from tests import ops
from tests import vorbis_ops as vo

TOL = 1e-5


def check(engine_so, channels, bs, n_blocks, seed, interleave=False, split=None, ref_windows=True, p_long=0.8):
    flags, spectra = vo.make_batch(seed, channels, bs, n_blocks, p_long=p_long)
    want = vo.ref_decode(flags, spectra, channels, bs)
    assert want.size == 0 or np.abs(want).max() < 4.0
    wins = [vo.ref_window(bs[0]), vo.ref_window(bs[1])] if ref_windows else None
    v = vo.Vorbis(engine_so, channels, bs, wins)
    try:
        if split is None:
            got = v.decode(flags, spectra, interleave)
        else:   # the overlap state must carry across calls like vorbis_dsp_state does
            parts, off = [], 0
            cuts = [0] + list(split) + [n_blocks]
            for a, b in zip(cuts[:-1], cuts[1:]):
                m = int(sum(channels * (bs[int(f)] // 2) for f in flags[a:b]))
                parts.append(v.decode(flags[a:b], spectra[off:off + m], interleave))
                off += m
            got = np.concatenate(parts, axis=0 if interleave else 1)
    finally:
        v.close()
    if interleave:
        got = got.T
    assert got.shape == want.shape, (got.shape, want.shape)
    err = float(np.abs(got - want).max()) if want.size else 0.0
    assert err <= TOL, "max abs error %g (peak %g)" % (err, float(np.abs(want).max()))
    return err


def test_oracle_window_and_transform_definitions():
    """pins what the harness exposes: the Vorbis window formula and the IMDCT definition"""
    for n in (64, 256, 2048, 8192):
        i = np.arange(n // 2)
        f = np.sin(0.5 * np.pi * np.sin((i + 0.5) / n * np.pi) ** 2)
        assert np.abs(vo.ref_window(n) - f).max() < 1e-7
    rng = np.random.default_rng(0)
    for n in (64, 256):
        X = rng.standard_normal(n // 2).astype(np.float32)
        t, k = np.arange(n), np.arange(n // 2)
        y = np.cos(2 * np.pi / n * np.outer(t + 0.5 + n / 4, k + 0.5)) @ X.astype(np.float64)
        assert np.abs(vo.ref_mdct_backward(n, X) - y).max() < 1e-4


@pytest.mark.parametrize("case", [
    dict(channels=2, bs=(256, 2048), n_blocks=40, seed=1),
    dict(channels=6, bs=(256, 2048), n_blocks=33, seed=2, interleave=True),
    dict(channels=1, bs=(64, 64), n_blocks=9, seed=3),
    dict(channels=2, bs=(128, 8192), n_blocks=12, seed=4),
    dict(channels=3, bs=(512, 1024), n_blocks=50, seed=5, p_long=0.5, split=(1, 7, 8, 30)),
    dict(channels=2, bs=(256, 2048), n_blocks=1, seed=6),
    dict(channels=2, bs=(256, 2048), n_blocks=20, seed=7, ref_windows=False),
])
def test_vorbis_emulated_device_code(case):
    check(ops.EMU_SO, **case)


@pytest.mark.gpu
@pytest.mark.parametrize("case", [
    dict(channels=2, bs=(256, 2048), n_blocks=40, seed=1),
    dict(channels=6, bs=(256, 2048), n_blocks=333, seed=2, interleave=True),
    dict(channels=1, bs=(64, 64), n_blocks=9, seed=3),
    dict(channels=2, bs=(128, 8192), n_blocks=12, seed=4),
    dict(channels=3, bs=(512, 1024), n_blocks=50, seed=5, p_long=0.5, split=(1, 7, 8, 30)),
    dict(channels=2, bs=(256, 2048), n_blocks=1, seed=6),
    dict(channels=2, bs=(256, 2048), n_blocks=20, seed=7, ref_windows=False),
    dict(channels=6, bs=(256, 2048), n_blocks=4096, seed=3),          # config C3
])
def test_vorbis_cuda(case):
    check(ops.CUDA_SO, **case)
