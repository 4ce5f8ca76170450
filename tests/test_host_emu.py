"""Host-side logic without a GPU: the recorder (reference host code + our DSP recorder), the
job queue / submit thread and the dependency planner, run against tests/emu's HOST build of the
kernel source (one 'thread' per work group). This is test infrastructure: it checks the host
logic and the kernels' arithmetic, never the product path (which needs the CUDA library)."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EMU_DIR = os.path.join(ROOT, "tests", "emu")
EMU_HOST = os.path.join(EMU_DIR, "libdav1d_b200_emu.so")


@pytest.fixture(scope="module")
def emu():
    if os.path.isdir("/root/reference/ext/dav1d/src"):
        subprocess.run(["make", "-C", EMU_DIR], check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    if not os.path.exists(EMU_HOST):
        pytest.skip("tests/emu not built (needs the reference sources)")

    def run(name, threads, repeat=1, truncate=None):
        env = dict(os.environ, EASYAV1_THREADS=str(threads), EMU_REPEAT=str(repeat))
        cmd = [sys.executable, os.path.join(EMU_DIR, "decode_md5.py"), name] + ([str(truncate)] if truncate else [])
        r = subprocess.run(cmd, env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=300)
        assert r.returncode == 0, r.stderr[-2000:]
        return json.loads(r.stdout.strip().splitlines()[-1])
    return run


@pytest.mark.parametrize("name", ["intra_64x64_nofilt", "cif_intra", "cif_inter", "cif_tools_s4", "cif_tools_s2", "cif_10bit",
                                  "cif_grain_t1", "cif_grain_t15", "cif_grain_est", "cif_444", "cif_422", "cif_400", "qcif_lossless", "cif_screen"])
@pytest.mark.parametrize("threads", [1, 4])
def test_recorded_frames_match_golden(emu, manifest, name, threads):
    """single-threaded walk and dav1d's frame/tile threading both give the golden frames"""
    got = emu(name, threads)
    assert got["error"] is None
    assert got["passes"][0] == [f["md5"] for f in manifest[name]["frames"]]


def test_repeated_open_close(emu, manifest):
    """decoders opened back to back reuse pooled buffers; results must not depend on history"""
    want = [f["md5"] for f in manifest["cif_tools_s2"]["frames"]]
    got = emu("cif_tools_s2", 3, repeat=4)
    assert got["passes"] == [want] * 4


def test_truncated_stream_fails_cleanly(emu, manifest):
    """a damaged stream ends with an error or early EOS, never a hang or a crash"""
    got = emu("cif_inter", 4, truncate=0.66)
    assert len(got["passes"][0]) < len(manifest["cif_inter"]["frames"])


def test_fastdiv_is_exact_over_its_documented_range():
    """dev_common.cuh FastDiv (multiply-high division used by the tile loops): exact for divisors
    up to 1024 and dividends below 2^22."""
    import ctypes as C
    import os
    so = os.path.join(os.path.dirname(os.path.abspath(__file__)), "emu", "libeav1_emu.so")
    L = C.CDLL(so)
    L.eav1emu_fastdiv_mismatches.restype = C.c_longlong
    L.eav1emu_fastdiv_mismatches.argtypes = [C.c_int, C.c_int]
    assert L.eav1emu_fastdiv_mismatches(1024, 1 << 22) == 0


def test_failed_frame_does_not_poison_the_decoder(manifest):
    """One frame the recorder cannot express (forced with EAV1_TEST_FAIL_FRAME) fails alone, the frames that
    reference it fail with it (DAV1D_ERR(EIO), as after a damaged frame in dav1d), and the SAME decoder is
    bit-exact again from the next key frame on."""
    import ctypes as C
    import hashlib
    code = r"""
import sys, json, hashlib, ctypes as C
sys.path.insert(0, %r)
from easyav1_b200.lib import load_host, StreamFrame
from tests.golden.gen_fixtures import read_ivf
man = json.load(open(%r))["cif_inter"]
tus = read_ivf(%r)
obu = b"".join(tus + tus)
host = load_host(%r)
keep = C.create_string_buffer(obu, len(obu))
h = host.eav1_stream_open(keep, len(obu), 1, 7, 1)
buf = C.create_string_buffer(1 << 20); info = StreamFrame(); used = C.c_size_t()
res = []
while len(res) < 64:
    r = host.eav1_stream_next(h, C.byref(info), buf, 1 << 20, C.byref(used))
    if r == 1: break
    res.append(r if r < 0 else hashlib.md5(buf.raw[:used.value]).hexdigest())
host.eav1_stream_close(h)
print(json.dumps(res))
""" % (ROOT, os.path.join(ROOT, "tests", "golden", "manifest.json"),
       os.path.join(ROOT, "tests", "golden", "cif_inter.ivf"), EMU_HOST)
    if not os.path.exists(EMU_HOST):
        pytest.skip("tests/emu not built (needs the reference sources)")
    want = [f["md5"] for f in manifest["cif_inter"]["frames"]]
    for threads in ("1", "3"):
        env = dict(os.environ, EAV1_TEST_FAIL_FRAME="2", EASYAV1_THREADS=threads)
        r = subprocess.run([sys.executable, "-c", code], env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=300)
        assert r.returncode == 0, r.stderr[-2000:]
        got = json.loads(r.stdout.strip().splitlines()[-1])
        assert len(got) == 2 * len(want)
        assert got[0] == want[0]
        assert any(isinstance(x, int) and x < 0 for x in got[:len(want)])          # the damaged GOP reports errors ...
        assert all(isinstance(x, int) or x == want[i] for i, x in enumerate(got[:len(want)]))   # ... never wrong pictures
        assert got[len(want):] == want                                               # and the next GOP is exact again
