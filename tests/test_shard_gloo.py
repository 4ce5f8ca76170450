"""N > 1: units (streams / closed-GOP segments) are dealt to ranks with no data-path collective.
world_size-2 gloo run on the CPU: every unit is decoded exactly once, the per-unit digests
gathered from both ranks equal the single-process run, and the elapsed time is the MAX over
ranks. The decoder stand-in on the CPU is the oracle (the product needs a GPU)."""
import hashlib
import os
import socket

import pytest

from easyav1_b200 import shard
from tests import refdec
from tests.conftest import fixture_obu
from tests.golden.gen_fixtures import read_ivf

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.skipif(not os.path.exists(refdec.REF_SO), reason="oracle/_ref not built")


def test_assign_round_robin():
    assert shard.assign(5, 2, 0) == [0, 2, 4]
    assert shard.assign(5, 2, 1) == [1, 3]
    assert shard.assign(0, 4, 3) == []
    got = sorted(i for r in range(8) for i in shard.assign(64, 8, r))
    assert got == list(range(64))
    with pytest.raises(ValueError):
        shard.assign(4, 2, 2)


def test_closed_gop_segments(manifest):
    man = manifest["cif_inter"]
    tus = read_ivf(os.path.join(ROOT, "tests", "golden", man["file"]))
    segs = shard.closed_gop_segments(tus)
    assert sum(len(s) for s in segs) == len(tus)
    assert b"".join(b"".join(s) for s in segs) == b"".join(tus)
    # two copies of the clip back to back: at least two segments, each decodable on its own
    segs2 = shard.closed_gop_segments(tus + tus)
    assert len(segs2) >= 2
    first = [hashlib.md5(p).hexdigest() for _, p in refdec.decode_all_big(b"".join(segs2[0]), man["w"] * man["h"] * 4)]
    assert first == [f["md5"] for f in man["frames"]][: len(first)] and first


def _digest(obu):
    h = hashlib.md5()
    for _, planes in refdec.decode_all(obu):
        h.update(planes)
    return h.hexdigest()


def _worker(rank, world, port, names, q):
    import json
    import torch.distributed as dist
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    man = json.load(open(os.path.join(ROOT, "tests", "golden", "manifest.json")))
    units = [fixture_obu(man[n]) for n in names]
    out, dt, total = shard.run_sharded(units, _digest, rank, world, dist)
    gathered = [None] * world
    dist.all_gather_object(gathered, out)
    if rank == 0:
        merged = {}
        for g in gathered:
            assert not set(g) & set(merged)      # no unit decoded twice
            merged.update(g)
        q.put((merged, dt, total))
    dist.barrier()
    dist.destroy_process_group()


def test_two_ranks_equal_one(manifest):
    import torch.multiprocessing as mp
    names = ["intra_64x64_nofilt", "cif_intra", "cif_inter", "cif_tools_s2", "cif_intra"]
    single = {i: _digest(fixture_obu(manifest[n])) for i, n in enumerate(names)}
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, names, q)) for r in range(2)]
    for p in procs:
        p.start()
    merged, dt, total = q.get(timeout=240)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert merged == single and total == len(names) and dt > 0


def test_open_gop_is_not_split():
    """A sequence header in front of a hidden (forward) key frame or an intra-only frame does not start an
    independently decodable segment: only a shown key frame refreshes all eight reference slots."""
    from tests.golden.obu_synth import Seq, Stream
    st = Stream(Seq(64, 64), 5)
    st.frame(0, qidx=100)                                                  # shown key frame: segment 1
    st.frame(1, offset=1, qidx=100, refresh=1)
    st.frame(0, show_frame=0, showable=1, refresh=2, offset=4, qidx=100)   # forward key frame behind a sequence header
    st.frame(1, offset=2, qidx=100, refresh=4)                             # still references the first GOP
    st.frame(0, qidx=100)                                                  # shown key frame: segment 2
    st.frame(1, offset=1, qidx=100, refresh=1)
    tus = st.samples()
    assert sum(1 for tu in tus if any(t == 1 for t, _ in shard._obus(tu))) == 3
    segs = shard.closed_gop_segments(tus)
    assert [len(s) for s in segs] == [4, 2]
    assert b"".join(sum(segs, [])) == b"".join(tus)
