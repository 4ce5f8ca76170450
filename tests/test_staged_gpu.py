"""Device-resident replay and the launch policies (GPU): the staged frames of a stream replayed with direct launches and
as one CUDA graph give the golden last frame; film grain captured inside the graph gives what eav1cu_apply_grain
gives frame by frame; every setting of the programmatic-launch mask, a narrowed dataflow grid and two decoders open
at once (sleeping waits, no early release) decode bit-exactly."""
import ctypes as C
import hashlib
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def libs():
    from easyav1_b200.lib import load_engine, load_host
    return load_engine(), load_host()


def staged(libs, name):
    import bench
    E, host = libs
    man, obu = bench.load_clip(name)
    return bench.Staged(E, host, man, obu), man      # the constructor checks every decoded frame against the manifest


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["cif_tools_s2", "cif_10bit"])
def test_graph_replay_matches_direct_launches(libs, name):
    st, man = staged(libs, name)
    want = man["frames"][-1]["md5"]
    st.replay(2)
    assert st.last_frame_md5() == want
    st.graph_build()
    st.graph_replay(3)
    assert st.last_frame_md5() == want
    st.replay(1)                    # and back: the graph leaves the context usable for direct launches
    assert st.last_frame_md5() == want
    st.close()


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["cif_tools_s2", "cif_10bit"])
def test_film_grain_inside_the_graph(libs, name):
    import numpy as np
    from tests import ops
    from tests.test_filmgrain import random_grain
    E, _ = libs
    st, man = staged(libs, name)
    si = st.infos[0]
    grain = random_grain(np.random.default_rng(7), si.bpc, 1, cfl=0, overlap=1, clip=0)
    E.eav1cu_pic_alloc.restype = C.c_int
    E.eav1cu_pic_alloc.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int]
    E.eav1cu_apply_grain.argtypes = [C.c_void_p, C.c_int, C.c_int, C.POINTER(ops.GrainParams)]
    busy = {x.cur for x in st.infos}
    gp = -1
    for _ in range(64):
        gp = E.eav1cu_pic_alloc(st.eng, si.w, si.h, si.layout, si.bpc)
        assert gp >= 0
        if gp not in busy:
            break
    assert gp not in busy
    for i in range(st.n_staged):    # frame by frame: replay, then grain into gp
        assert E.eav1cu_staged_replay(st.eng, i) == 0
        r = E.eav1cu_apply_grain(st.eng, st.infos[i].cur, gp, C.byref(grain))
        if r:
            msg = C.create_string_buffer(256)
            E.eav1cu_last_error(C.c_void_p(st.eng), msg, 256)
            raise AssertionError("apply_grain: %d %s" % (r, msg.value.decode()))
    direct = st.last_frame_md5(gp)
    assert direct != st.last_frame_md5()        # the grain did something
    st.graph_build(gp, grain)
    st.graph_replay(2)
    assert st.last_frame_md5(gp) == direct
    assert st.last_frame_md5() == man["frames"][-1]["md5"]
    st.close()


CHILD = r"""
import hashlib, json, sys
sys.path.insert(0, %r)
from tests.conftest import fixture_obu
from easyav1_b200.decoder import Decoder
man = json.load(open(%r))[sys.argv[1]]
d = Decoder(fixture_obu(man), max_frame_bytes=man["w"] * man["h"] * 4)
print(json.dumps([hashlib.md5(planes).hexdigest() for _, planes in d]))
d.close()
"""


@pytest.mark.gpu
@pytest.mark.parametrize("env", [{"EAV1_PDL": "0"}, {"EAV1_PDL": "7"}, {"EAV1_PDL": "15"}, {"EAV1_PDL": "9"},
                                 {"EAV1_RECON_CTAS_PER_SM": "1"}, {"EAV1_COPY_ENGINE": "1"}, {"EAV1_SPIN_WAIT": "0"}])
def test_launch_policies_are_bit_exact(manifest, env):
    """each policy the engine picks by itself (and its A/B override) decodes the same frames"""
    name = "cif_tools_s2"
    code = CHILD % (ROOT, os.path.join(ROOT, "tests", "golden", "manifest.json"))
    r = subprocess.run([sys.executable, "-c", code, name], env=dict(os.environ, **env), capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    import json
    got = json.loads(r.stdout.strip().splitlines()[-1])
    assert got == [f["md5"] for f in manifest[name]["frames"]]


@pytest.mark.gpu
def test_two_decoders_open_at_once(manifest):
    """a second live decoder switches the process to sleeping waits, a narrower dataflow grid and no early release;
    frames of both, pulled alternately, stay golden, and so does a decoder opened after the others have closed"""
    from tests.conftest import fixture_obu
    from easyav1_b200.decoder import Decoder
    names = ["cif_tools_s2", "cif_inter"]
    decs = [Decoder(fixture_obu(manifest[n]), max_frame_bytes=manifest[n]["w"] * manifest[n]["h"] * 4) for n in names]
    its = [iter(d) for d in decs]
    got = [[], []]
    live = [True, True]
    while any(live):
        for k in range(2):
            if not live[k]:
                continue
            try:
                _, planes = next(its[k])
                got[k].append(hashlib.md5(planes).hexdigest())
            except StopIteration:
                live[k] = False
    for k, n in enumerate(names):
        assert got[k] == [f["md5"] for f in manifest[n]["frames"]], n
    for d in decs:
        d.close()
    d = Decoder(fixture_obu(manifest[names[0]]), max_frame_bytes=manifest[names[0]]["w"] * manifest[names[0]]["h"] * 4)
    assert [hashlib.md5(p).hexdigest() for _, p in d] == [f["md5"] for f in manifest[names[0]]["frames"]]
    d.close()
