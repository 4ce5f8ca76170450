"""Stream-level parity: each fixture decoded through libdav1d_b200.so (dav1d API, CUDA kernels)
must give the per-frame MD5 of Y|U|V that the reference's dav1d C build gave
(tests/golden/manifest.json, produced by tests/golden/gen_fixtures.py from oracle/_ref)."""
import hashlib

import pytest

from tests.conftest import fixture_obu

CLIPS = ["intra_64x64_nofilt", "cif_intra", "cif_inter", "cif_tools_s4", "cif_tools_s2", "c1_360p", "c2_1080p",
         "cif_10bit", "c4_720p_10bit", "c4_2160p_10bit",
         "cif_grain_t1", "cif_grain_t15", "cif_grain_est", "cif_444", "cif_422", "cif_400", "qcif_lossless", "cif_screen"]


@pytest.mark.gpu
@pytest.mark.parametrize("name", CLIPS)
def test_stream_md5(manifest, name):
    from easyav1_b200.decoder import Decoder
    man = manifest[name]
    d = Decoder(fixture_obu(man), max_frame_bytes=man["w"] * man["h"] * 4)
    n = 0
    for i, (info, planes) in enumerate(d):
        want = man["frames"][i]
        assert (info.w, info.h, info.bpc, info.layout) == (want["w"], want["h"], want["bpc"], want["layout"])
        assert hashlib.md5(planes).hexdigest() == want["md5"], "frame %d of %s" % (i, name)
        n += 1
    assert n == len(man["frames"])
    assert d.frames_submitted >= n
    assert d.launches > 0
    d.close()


@pytest.mark.gpu
@pytest.mark.parametrize("filters", [0, 1, 3])
def test_stream_filter_subsets(manifest, filters):
    """Dav1dSettings.inloop_filters subsets against the reference library run on the host."""
    from easyav1_b200.decoder import Decoder
    from tests import refdec
    man = manifest["cif_inter"]
    obu = fixture_obu(man)
    cap = man["w"] * man["h"] * 4
    ref = list(refdec.decode_all_big(obu, cap, inloop_filters=filters))
    got = list(Decoder(obu, max_frame_bytes=cap, inloop_filters=filters))
    assert len(got) == len(ref)
    for i, ((_, a), (_, b)) in enumerate(zip(got, ref)):
        assert a == b, "frame %d differs with inloop_filters=%d" % (i, filters)
