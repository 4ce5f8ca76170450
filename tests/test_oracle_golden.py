"""The oracle (oracle/_ref = the reference's own dav1d C build) against the committed golden
vectors: per-frame MD5 of Y|U|V for every fixture (tests/golden/manifest.json). Pins the checker
before any GPU result is compared with it. Also checks determinism across thread counts, which
is what lets bench.py run the reference multi-threaded and still compare MD5s."""
import hashlib
import os

import pytest

from tests import refdec
from tests.conftest import fixture_obu

pytestmark = pytest.mark.skipif(not os.path.exists(refdec.REF_SO), reason="oracle/_ref not built")

SMALL = ["intra_64x64_nofilt", "cif_intra", "cif_inter", "cif_tools_s4", "cif_tools_s2", "c1_360p", "c2_1080p", "cif_10bit", "c4_720p_10bit", "c4_2160p_10bit",
         "cif_grain_t1", "cif_grain_t15", "cif_grain_est", "cif_444", "cif_422", "cif_400", "qcif_lossless", "cif_screen"]


@pytest.mark.parametrize("name", SMALL)
def test_reference_matches_manifest(manifest, name):
    man = manifest[name]
    frames = list(refdec.decode_all_big(fixture_obu(man), man["w"] * man["h"] * 4))
    assert len(frames) == len(man["frames"])
    for (info, planes), want in zip(frames, man["frames"]):
        assert (info.w, info.h, info.bpc, info.frame_type) == (want["w"], want["h"], want["bpc"], want["frame_type"])
        assert hashlib.md5(planes).hexdigest() == want["md5"]


def test_reference_threads_do_not_change_output(manifest):
    man = manifest["cif_tools_s2"]
    obu = fixture_obu(man)
    cap = man["w"] * man["h"] * 4
    a = [p for _, p in refdec.decode_all_big(obu, cap, n_threads=1, max_frame_delay=1)]
    b = [p for _, p in refdec.decode_all_big(obu, cap, n_threads=4, max_frame_delay=0)]
    assert a == b


def test_bench_clip_last_frame(manifest):
    man = manifest["c2_1080p_bench"]
    last = None
    n = 0
    for _, planes in refdec.decode_all_big(fixture_obu(man), man["w"] * man["h"] * 4, n_threads=4, max_frame_delay=0):
        last = planes
        n += 1
    assert n == len(man["frames"]) == 120
    assert hashlib.md5(last).hexdigest() == man["frames"][-1]["md5"]


def test_inloop_filter_mask_changes_output(manifest):
    """Dav1dSettings.inloop_filters (include/dav1d/dav1d.h) really gates deblock / CDEF / LR in
    the oracle, so the per-stage parity tests compare different things."""
    man = manifest["cif_tools_s2"]
    obu = fixture_obu(man)
    cap = man["w"] * man["h"] * 4
    outs = {}
    for f in (0, 1, 3, 7):
        outs[f] = hashlib.md5(b"".join(p for _, p in refdec.decode_all_big(obu, cap, inloop_filters=f))).hexdigest()
    assert len(set(outs.values())) == 4, outs
