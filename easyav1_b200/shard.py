"""Multi-GPU scaling of the decode path: independent units only.

A unit is one stream, or one closed-GOP segment of a stream (from a temporal unit whose first frame
is a shown key frame behind a sequence header up to the next such unit: a shown key frame refreshes
all eight reference slots, so nothing crosses; src/easyav1.c:1979,2929-2959 seeks on key frames too). Units are
dealt round-robin to ranks, each rank owns one GPU, and NOTHING is exchanged on the data path:
the only collective is the MAX over ranks of the elapsed time (and, in tests, a gather of
digests). One process per GPU, torch.distributed only as plumbing."""
import time


def assign(n_units: int, world: int, rank: int):
    """indices of the units rank `rank` of `world` processes (config C5: stream i -> GPU i mod n)"""
    if world < 1 or not 0 <= rank < world:
        raise ValueError("bad rank/world")
    return list(range(rank, n_units, world))


def _obus(tu: bytes):
    """yield (obu_type, payload) of one temporal unit (low-overhead format, size fields present)"""
    pos = 0
    while pos < len(tu):
        hdr = tu[pos]
        typ, ext, has_size = (hdr >> 3) & 15, (hdr >> 2) & 1, (hdr >> 1) & 1
        pos += 1 + ext
        if not has_size:
            yield typ, tu[pos:]
            return
        size, shift = 0, 0
        while True:
            b = tu[pos]
            pos += 1
            size |= (b & 0x7F) << shift
            shift += 7
            if not b & 0x80:
                break
        yield typ, tu[pos:pos + size]
        pos += size


def _starts_closed_gop(tu: bytes) -> bool:
    """True if the temporal unit carries a sequence header AND its first frame header is a shown key frame
    (show_existing_frame = 0, frame_type = KEY_FRAME, show_frame = 1: only then are all eight reference slots
    refreshed, AV1 spec 5.9.2 / ext/dav1d/src/obu.c:409-470). Encoders also repeat the sequence header in front of
    forward key frames and intra-only frames of open GOPs, whose leading frames still reference the previous GOP."""
    has_seq = False
    for typ, payload in _obus(tu):
        if typ == 1:
            has_seq = True
        elif typ in (3, 6) and payload:                # OBU_FRAME_HEADER / OBU_FRAME: the first three header fields
            b = payload[0]
            show_existing, frame_type, show_frame = b >> 7, (b >> 5) & 3, (b >> 4) & 1
            return has_seq and not show_existing and frame_type == 0 and show_frame == 1
    return False


def closed_gop_segments(samples):
    """split a list of temporal units into independently decodable segments"""
    segs, cur = [], []
    for tu in samples:
        starts = _starts_closed_gop(tu)
        if starts and cur:
            segs.append(cur)
            cur = []
        cur.append(tu)
    if cur:
        segs.append(cur)
    return segs


def run_sharded(units, decode_fn, rank=0, world=1, dist=None):
    """decode this rank's share of `units`; returns (results {unit index: value}, elapsed seconds
    as the MAX over ranks, units processed by all ranks). decode_fn(unit) -> value."""
    mine = assign(len(units), world, rank)
    if dist is not None and world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    out = {i: decode_fn(units[i]) for i in mine}
    dt = time.perf_counter() - t0
    if dist is not None and world > 1:
        import torch
        t = torch.tensor([dt], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dt = float(t.item())
    return out, dt, len(units)
