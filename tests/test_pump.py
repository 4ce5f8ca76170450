"""host/pump.c: many single-threaded decoders multiplexed on a few worker threads, frames
handed to the engine by the worker that walked them, pictures retired a few frames later. Every stream must give
the frame count and the last-frame MD5 of the golden manifests, whatever the interleaving."""
import hashlib
import json
import os
import subprocess
import sys

import pytest

from tests.golden import synth_streams

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
NAMES = ["cif_inter", "cif_tools_s2", "syn_sr_8", "cif_10bit", "cif_inter", "syn_scaled_8", "cif_grain_t1", "cif_444",
         "syn_420_12", "cif_screen"]


def expected(names):
    man = json.load(open(os.path.join(ROOT, "tests", "golden", "manifest.json")))
    sman = synth_streams.load_manifest()
    fr = [(sman if n.startswith("syn_") else man)[n]["frames"] for n in names]
    return [len(f) for f in fr], [f[-1]["md5"] for f in fr]


@pytest.mark.parametrize("workers", [1, 3, 8])
def test_pump_on_host_emulation(workers):
    if not os.path.exists(os.path.join(ROOT, "tests", "emu", "libdav1d_b200_emu.so")):
        pytest.skip("tests/emu not built (needs the reference sources)")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tests", "emu", "pump_md5.py"), str(workers)] + NAMES,
                       stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    got = json.loads(r.stdout.strip().splitlines()[-1])
    frames, md5 = expected(NAMES)
    assert got["frames"] == frames
    assert got["md5"] == md5


@pytest.mark.gpu
@pytest.mark.parametrize("workers", [2, 6])
def test_pump_cuda(workers):
    from easyav1_b200.lib import load_host, pump_run
    from tests.golden.gen_fixtures import read_ivf
    man = json.load(open(os.path.join(ROOT, "tests", "golden", "manifest.json")))
    names = NAMES + ["c2_1080p", "c4_720p_10bit"]
    obus = [synth_streams.build(n)[0] if n.startswith("syn_") else
            b"".join(read_ivf(os.path.join(ROOT, "tests", "golden", man[n]["file"]))) for n in names]
    frames, last = pump_run(load_host(), obus, workers, 1920 * 1080 * 4)
    want_frames, want_md5 = expected(names)
    assert frames == want_frames
    assert [hashlib.md5(x).hexdigest() for x in last] == want_md5
