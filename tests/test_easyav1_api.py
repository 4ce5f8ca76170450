"""Drop-in at the top boundary: the reference's src/easyav1.c + nestegg,
compiled unmodified and linked against libdav1d_b200.so, play a WebM file through the public
easyAV1 API (easyav1_init_from_memory / easyav1_decode_next / easyav1_get_video_frame, and
easyav1_init_from_filename / easyav1_play with the video callback) and every frame handed out must carry the
MD5 the reference decoder gave (tests/golden/manifest.json)."""
import json
import os
import subprocess
import sys

import pytest

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SCRIPT = """
import json, sys
sys.path.insert(0, %r)
from tests import easyav1_api as ea
L = ea.bind(sys.argv[1])
md5s, finished = ea.decode_webm(L, open(sys.argv[2], 'rb').read())
print(json.dumps({'md5': md5s, 'finished': finished}))
""" % ROOT


def run(lib, name, manifest):
    man = manifest[name]
    path = os.path.join(GOLDEN, name + ".webm")
    # a process of its own: the library spawns decoder threads and, for the emulation build, must
    # not share a symbol namespace with the CUDA engine
    out = subprocess.run([sys.executable, "-c", SCRIPT, lib, path], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    res = json.loads(out.stdout.strip().splitlines()[-1])
    assert res["finished"]
    assert res["md5"] == [f["md5"] for f in man["frames"]]


PLAY_SCRIPT = """
import json, sys
sys.path.insert(0, %r)
from tests import easyav1_api as ea
L = ea.bind(sys.argv[1])
md5s, finished = ea.play_file(L, sys.argv[2])
print(json.dumps({'md5': md5s, 'finished': finished}))
""" % ROOT


def run_play(lib, name, manifest):
    """easyav1_init_from_filename + easyav1_play: real-time playback on the library's own thread; every frame the
    video callback receives must be one of the golden frames, in stream order, and playback must reach the end."""
    want = [f["md5"] for f in manifest[name]["frames"]]
    path = os.path.join(GOLDEN, name + ".webm")
    out = subprocess.run([sys.executable, "-c", PLAY_SCRIPT, lib, path], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    res = json.loads(out.stdout.strip().splitlines()[-1])
    assert res["finished"]
    got = res["md5"]
    assert len(got) >= len(want) // 2, "playback delivered %d of %d frames" % (len(got), len(want))
    pos = -1
    for m in got:          # in order, each a golden frame (late frames may be dropped by the player, never altered)
        assert m in want[pos + 1:], "callback frame is not the next golden frame"
        pos = want.index(m, pos + 1)


def test_easyav1_play_over_emulated_engine(manifest):
    lib = os.path.join(ROOT, "tests", "emu", "libeasyav1_b200_emu.so")
    if not os.path.exists(lib):
        pytest.skip("tests/emu not built")
    run_play(lib, "c1_360p_webm", manifest)


@pytest.mark.gpu
def test_easyav1_play_cuda(manifest):
    run_play(os.path.join(ROOT, "easyav1_b200", "libeasyav1_b200.so"), "c1_360p_webm", manifest)


def test_easyav1_api_over_emulated_engine(manifest):
    lib = os.path.join(ROOT, "tests", "emu", "libeasyav1_b200_emu.so")
    if not os.path.exists(lib):
        pytest.skip("tests/emu not built")
    run(lib, "c1_360p_webm", manifest)


@pytest.mark.gpu
def test_easyav1_api_cuda(manifest):
    run(os.path.join(ROOT, "easyav1_b200", "libeasyav1_b200.so"), "c1_360p_webm", manifest)


API = """easyav1_decode_for easyav1_decode_next easyav1_decode_until easyav1_default_settings easyav1_destroy
easyav1_get_audio_channels easyav1_get_audio_frame easyav1_get_audio_sample_rate easyav1_get_current_settings
easyav1_get_current_timestamp easyav1_get_duration easyav1_get_status easyav1_get_total_audio_tracks
easyav1_get_total_video_frames_processed easyav1_get_total_video_tracks easyav1_get_video_fps easyav1_get_video_frame
easyav1_get_video_height easyav1_get_video_width easyav1_has_audio_track easyav1_has_video_frame easyav1_has_video_track
easyav1_init_from_custom_stream easyav1_init_from_file easyav1_init_from_filename easyav1_init_from_memory
easyav1_is_audio_buffer_filled easyav1_is_finished easyav1_play easyav1_seek_backward easyav1_seek_forward
easyav1_seek_to_timestamp easyav1_stop easyav1_update_settings""".split()


def test_easyav1_library_exports_the_reference_api():
    """every function src/easyav1.h declares is exported by the drop-in library (symbols only)"""
    import ctypes
    lib = os.path.join(ROOT, "easyav1_b200", "libeasyav1_b200.so")
    assert os.path.exists(lib), "run __graft_entry__.build()"
    L = ctypes.CDLL(lib)
    missing = [n for n in API if not hasattr(L, n)]
    assert not missing, missing
