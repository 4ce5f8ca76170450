"""Synthetic streams (tests/golden/obu_synth.py: hand-written sequence / frame headers, seeded random tile
data) through the product and the oracle. They cover what the offline encoders cannot be made to emit:
super-resolution, scaled references, 12-bit, global motion, header-signalled film grain, segmentation
features, ... (list in tests/golden/synth_streams.py). Golden = per-frame MD5 of the reference's dav1d C
build (tests/golden/synth_manifest.json). Bit-exact."""
import hashlib
import json
import os
import subprocess
import sys

import pytest

from tests.golden import synth_streams

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EMU_DIR = os.path.join(ROOT, "tests", "emu")
MAN = synth_streams.load_manifest()
ALL = sorted(synth_streams.STREAMS)
QUICK = [n for n in ALL if synth_streams.STREAMS[n][1]]


def test_manifest_covers_every_stream():
    assert sorted(MAN) == ALL


@pytest.mark.parametrize("name", ALL)
def test_generator_is_deterministic(name):
    """the stream rebuilt here is byte-identical to the one the golden MD5s were made from"""
    data, tus = synth_streams.build(name)
    assert len(data) == MAN[name]["bytes"] and len(tus) == MAN[name]["tus"]
    assert hashlib.sha256(data).hexdigest() == MAN[name]["sha256"]


@pytest.mark.parametrize("name", QUICK)
def test_oracle_matches_manifest(name):
    """oracle/_ref (the reference's own decoder, as built here) still gives the committed MD5s"""
    from tests import refdec
    data, _ = synth_streams.build(name)
    got = [hashlib.md5(p).hexdigest() for _, p in refdec.decode_all_big(data, 1 << 25)]
    assert got == [f["md5"] for f in MAN[name]["frames"]]


@pytest.mark.parametrize("name", QUICK)
def test_recorder_and_device_code_on_host_emulation(name):
    """host recorder + the kernels' source built for the host (tests/emu): checks the host logic and the
    arithmetic without a GPU; never the product path"""
    if not os.path.exists(os.path.join(EMU_DIR, "libdav1d_b200_emu.so")):
        pytest.skip("tests/emu not built (needs the reference sources)")
    env = dict(os.environ, EASYAV1_THREADS="3")
    r = subprocess.run([sys.executable, os.path.join(EMU_DIR, "decode_md5.py"), name], env=env,
                       stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    got = json.loads(r.stdout.strip().splitlines()[-1])
    assert got["error"] is None
    assert got["passes"][0] == [f["md5"] for f in MAN[name]["frames"]]


@pytest.mark.gpu
@pytest.mark.parametrize("name", ALL)
def test_synth_stream_md5_cuda(name):
    from easyav1_b200.decoder import Decoder
    data, _ = synth_streams.build(name)
    want = MAN[name]["frames"]
    cap = max(f["w"] * f["h"] for f in want) * 8
    d = Decoder(data, max_frame_bytes=cap)
    n = 0
    for i, (info, planes) in enumerate(d):
        assert (info.w, info.h, info.bpc, info.layout) == (want[i]["w"], want[i]["h"], want[i]["bpc"], want[i]["layout"])
        assert hashlib.md5(planes).hexdigest() == want[i]["md5"], "frame %d of %s" % (i, name)
        n += 1
    assert n == len(want)
    assert d.launches > 0
    d.close()
