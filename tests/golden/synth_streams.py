#!/usr/bin/env python3
"""Named synthetic streams (hand-written headers + seeded random tile data, see obu_synth.py) and their
golden per-frame MD5s. TEST INFRASTRUCTURE.

    python tests/golden/synth_streams.py            # (build container) regenerate synth_manifest.json

The streams are rebuilt from their seeds wherever the tests run (numpy's PCG64 integers are stable); the
manifest pins the SHA-256 of every stream next to the MD5s of the frames the reference's own dav1d C build
(oracle/_ref) decodes from it, so a drifting generator is noticed before a decoder is blamed.

What they add over the encoder-made fixtures: super-resolution (every denominator class, with and without
CDEF / loop restoration), references of another size (scaled motion compensation, up and down), 12-bit,
4:2:2 / 4:4:4 / 4:0:0 at high bit depth, global motion (translation / rot-zoom / affine), every
interpolation filter and the dual filter, intra block copy and palettes, segmentation features, delta q /
delta lf, quantiser matrices, lossless frames, 128x128 superblocks, several tiles, film grain signalled in
the frame header, hidden frames shown later. Because the symbols are drawn from the default CDFs every
transform size / type and prediction mode turns up (tests/test_coverage.py holds the histogram).
"""
import hashlib
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))

from tests.golden.obu_synth import Seq, Stream  # noqa: E402

MANIFEST = os.path.join(HERE, "synth_manifest.json")


def _plain(w, h, bpc=8, layout=1, seed=1, n_inter=3, seqkw=None, kf=None, inf=None):
    def b():
        st = Stream(Seq(w, h, bpc, layout, **(seqkw or {})), seed)
        st.frame(0, **dict(dict(qidx=90, cdef=(4, [(10, 5), (20, 9), (0, 0), (63, 63)]), lr=(1, 1, 1)), **(kf or {})))
        for i in range(n_inter):
            st.frame(1, **dict(dict(offset=i + 1, qidx=100 + 10 * i, refresh=1 << (i % 8),
                                    cdef=(3 + i % 4, [(8 + i, 4), (0, 3)]), lr=((2, 3, 1), (1, 1, 1), (3, 2, 2))[i % 3],
                                    lr_unit=(i % 3, i % 2)), **(inf or {})))
        return st
    return b


def _scaled(seed, bpc=8, layout=1):
    def b():
        st = Stream(Seq(256, 192, bpc, layout), seed)
        st.frame(0, qidx=90, cdef=(4, [(10, 5), (20, 9)]), lr=(1, 1, 1))
        st.frame(1, offset=1, qidx=100, refresh=1, size=(200, 144), lr=(2, 3, 1))      # smaller than its references
        st.frame(1, offset=2, qidx=100, refresh=2, size=(136, 104), refidx=(0, 0, 0, 1, 1, 1, 2))
        st.frame(1, offset=3, qidx=100, refresh=4, size=(256, 192), refidx=(0, 1, 2, 0, 1, 2, 0))  # larger than them
        st.frame(1, offset=4, qidx=100, refresh=8, size_from_ref=1, refidx=(2, 1, 0, 2, 1, 0, 2))
        st.frame(1, offset=5, qidx=100, refresh=16, size=(200, 144), refidx=(0, 1, 2, 3, 0, 1, 2), filter=3)  # bilinear, scaled
        return st
    return b


def _gmv(seed, bpc=8):
    def b():
        st = Stream(Seq(256, 192, bpc, 1), seed)
        st.frame(0, qidx=90)
        g = {0: (2, [3 << 13, -(2 << 13), (1 << 16) + 400, -300, 300, (1 << 16) + 400]),
             1: (1, [5 << 13, -(7 << 13), 1 << 16, 0, 0, 1 << 16]),
             3: (3, [1 << 13, 2 << 13, (1 << 16) - 200, 120, -90, (1 << 16) + 260])}
        st.frame(1, offset=1, qidx=100, refresh=1, gmv=g)
        st.frame(1, offset=2, qidx=100, refresh=2, gmv=g, hp=0, filter=0)
        st.frame(1, offset=3, qidx=100, refresh=4, gmv=g, filter=3, switchable_motion_mode=0, ref_select=0)
        st.frame(1, offset=4, qidx=100, refresh=8, filter=1)
        st.frame(1, offset=5, qidx=100, refresh=16, filter=2, warp_motion=0)
        st.frame(1, offset=6, qidx=100, refresh=32, filter=3)                        # bilinear with compound references
        return st
    return b


def _screen(seed, intrabc):
    def b():
        st = Stream(Seq(192, 128, 8, 3, screen_content_tools=2), seed)
        st.frame(0, qidx=90, screen=1, allow_intrabc=intrabc)
        st.frame(1, offset=1, qidx=100, refresh=1, screen=1, force_integer_mv=1)
        st.frame(2, offset=2, qidx=100, refresh=2, screen=1, allow_intrabc=intrabc, show_frame=1)
        st.frame(1, offset=3, qidx=100, refresh=4, screen=1, force_integer_mv=0)
        return st
    return b


def _grain(seed, bpc, layout, w=200, h=136):
    def b():
        st = Stream(Seq(w, h, bpc, layout, film_grain_present=1), seed)
        g = dict(seed=1234, y=[(0, 20), (80, 60), (160, 90), (255, 120)],
                 uv=[[(0, 30), (128, 70), (255, 50)], [(10, 40), (200, 90)]], cfl=0, scaling_shift=10, lag=3,
                 ar_y=[((i * 7) % 23) - 11 for i in range(24)],
                 ar_uv=[[((i * 5) % 19) - 9 for i in range(25)], [((i * 3) % 17) - 8 for i in range(25)]],
                 ar_shift=7, grain_scale_shift=0, uv_mult=[10, -20], uv_luma_mult=[30, 12], uv_offset=[5, -9],
                 overlap=1, clip=1)
        st.frame(0, qidx=90, grain=g)
        st.frame(1, offset=1, qidx=100, refresh=1, grain=dict(g, seed=999, cfl=1, lag=2, overlap=0, clip=0))
        st.frame(1, offset=2, qidx=100, refresh=2, grain=None)
        st.frame(1, offset=3, qidx=100, refresh=4, grain=dict(g, seed=77, y=[], lag=1))
        return st
    return b


def _segs(seed):
    def b():
        st = Stream(Seq(200, 136, 8, 1), seed)
        sg = [dict(delta_q=-20), dict(delta_q=30, lf_y_v=5), dict(skip=1), dict(ref=0), dict(globalmv=1),
              dict(lf_u=-4, lf_v=9), dict(delta_q=-90), {}]
        st.frame(0, qidx=90, seg=sg, lf_deltas=([1, 0, -1, 2, None, None, 3, -2], [2, -3]), sharp=3)
        st.frame(1, offset=1, qidx=100, refresh=1, seg=sg, lf=(30, 25, 20, 15), sharp=7)
        st.frame(1, offset=2, qidx=100, refresh=2, primary_ref=0, seg=sg, use_ref_frame_mvs=1)
        st.frame(1, offset=5, qidx=100, refresh=4, primary_ref=1, use_ref_frame_mvs=1, show_frame=0, showable=1)
        st.frame(1, offset=3, qidx=100, refresh=8, primary_ref=1, use_ref_frame_mvs=1, refidx=(0, 1, 2, 3, 2, 2, 2))
        st.show_existing(2)
        return st
    return b


def _superres(seed, bpc=8, layout=1, w=264, h=136, cdef=1, lr=1, sb128=0):
    def b():
        st = Stream(Seq(w, h, bpc, layout, super_res=1, sb128=sb128, cdef=cdef, restoration=lr), seed)
        st.frame(0, qidx=90, superres_denom=12, cdef=(4, [(10, 5), (20, 9)]), lr=(1, 1, 1))
        st.frame(1, offset=1, qidx=100, refresh=1, superres_denom=16, lr=(2, 3, 1), lr_unit=(1, 1))
        st.frame(1, offset=2, qidx=100, refresh=2, superres_denom=9, lr=(0, 0, 0))
        st.frame(1, offset=3, qidx=100, refresh=4, superres_denom=8, lr=(3, 2, 2))
        st.frame(1, offset=4, qidx=100, refresh=8, superres_denom=11, size=(w - 64, h - 16), lr=(1, 0, 1),
                 lr_unit=(0 if sb128 else 2, 0))
        return st
    return b


def _uhd_grain_warp(seed):
    """BASELINE.json configs[3] as a parity case: 3840x2160 10-bit with film grain signalled in the frame headers,
    global motion (rot-zoom and affine, i.e. warped prediction of whole blocks) and local warped motion allowed."""
    def b():
        st = Stream(Seq(3840, 2160, 10, 1, film_grain_present=1), seed)
        g = dict(seed=4321, y=[(0, 24), (64, 48), (128, 80), (255, 110)],
                 uv=[[(0, 20), (255, 60)], [(30, 50), (220, 70)]], cfl=0, scaling_shift=10, lag=3,
                 ar_y=[((i * 5) % 21) - 10 for i in range(24)],
                 ar_uv=[[((i * 7) % 17) - 8 for i in range(25)], [((i * 11) % 19) - 9 for i in range(25)]],
                 ar_shift=7, grain_scale_shift=0, uv_mult=[12, -16], uv_luma_mult=[20, 8], uv_offset=[3, -7],
                 overlap=1, clip=0)
        gm = {0: (2, [3 << 13, -(2 << 13), (1 << 16) + 300, -200, 200, (1 << 16) + 300]),
              3: (3, [1 << 13, 2 << 13, (1 << 16) - 150, 100, -80, (1 << 16) + 210])}
        st.frame(0, qidx=110, cdef=(5, [(12, 5), (24, 9), (0, 0), (40, 20)]), lr=(1, 2, 3), grain=g)
        st.frame(1, offset=1, qidx=120, refresh=1, gmv=gm, cdef=(4, [(9, 4), (0, 3)]), lr=(2, 1, 1), grain=dict(g, seed=99, cfl=1))
        st.frame(1, offset=2, qidx=130, refresh=2, gmv=gm, cdef=(3, [(20, 7)]), lr=(3, 3, 2), lr_unit=(1, 1), grain=dict(g, seed=7, lag=2))
        return st
    return b


# name -> (builder, in the quick CPU subset?)
STREAMS = {
    "syn_420_8":        (_plain(352, 288), True),
    "syn_420_10":       (_plain(352, 288, 10, seed=2), False),
    "syn_420_12":       (_plain(352, 288, 12, seed=3), True),
    "syn_444_8":        (_plain(136, 120, 8, 3, seed=4), False),
    "syn_444_12":       (_plain(136, 120, 12, 3, seed=5), False),
    "syn_422_10":       (_plain(128, 128, 10, 2, seed=140, n_inter=2), False),
    "syn_422_12":       (_plain(128, 128, 12, 2, seed=172, n_inter=2), True),
    "syn_400_8":        (_plain(136, 120, 8, 0, seed=8), False),
    "syn_400_12":       (_plain(136, 120, 12, 0, seed=9), False),
    "syn_sb128":        (_plain(392, 264, 8, 1, seed=10, seqkw=dict(sb128=1)), False),
    "syn_tiles":        (_plain(392, 264, 8, 1, seed=11, kf=dict(tile_cols_log2=1, tile_rows_log2=1),
                                inf=dict(tile_cols_log2=2, tile_rows_log2=1)), False),
    "syn_lowq":         (_plain(200, 136, 8, 1, seed=12, kf=dict(qidx=20), inf=dict(qidx=30)), False),
    "syn_highq":        (_plain(200, 136, 8, 1, seed=13, kf=dict(qidx=240), inf=dict(qidx=250)), False),
    "syn_lossless":     (_plain(136, 120, 8, 1, seed=14, kf=dict(qidx=0), inf=dict(qidx=0)), False),
    "syn_lossless_12":  (_plain(136, 120, 12, 3, seed=34, kf=dict(qidx=0), inf=dict(qidx=0)), False),
    "syn_deltaq":       (_plain(200, 136, 8, 1, seed=15, kf=dict(delta_q=1, delta_lf=(1, 1)),
                                inf=dict(delta_q=2, delta_lf=(0, 0))), False),
    "syn_qm":           (_plain(200, 136, 8, 1, seed=16, kf=dict(qm=(3, 5, 5)), inf=dict(qm=(10, 2, 2))), False),
    "syn_reduced_tx":   (_plain(200, 136, 8, 1, seed=17, kf=dict(reduced_txtp=1),
                                inf=dict(reduced_txtp=1, tx_switchable=0)), False),
    "syn_scaled_8":     (_scaled(21), True),
    "syn_scaled_10":    (_scaled(31, 10), False),
    "syn_scaled_444":   (_scaled(32, 8, 3), False),
    "syn_scaled_12":    (_scaled(35, 12, 1), False),
    "syn_gmv_8":        (_gmv(22), False),
    "syn_gmv_10":       (_gmv(33, 10), False),
    "syn_screen":       (_screen(23, 0), False),
    "syn_intrabc":      (_screen(26, 1), False),
    "syn_grain_10":     (_grain(24, 10, 1), False),
    "syn_grain_8_444":  (_grain(27, 8, 3), False),
    "syn_grain_12_400": (_grain(28, 12, 0), False),
    "syn_segs":         (_segs(25), False),
    "syn_sr_8":         (_superres(41), True),
    "syn_sr_10":        (_superres(42, 10), False),
    "syn_sr_12_444":    (_superres(43, 12, 3), False),
    "syn_sr_nocdef":    (_superres(44, 8, 1, cdef=0), False),
    "syn_sr_nolr":      (_superres(45, 8, 1, lr=0), False),
    "syn_sr_plain":     (_superres(46, 8, 1, cdef=0, lr=0), False),
    "syn_sr_400":       (_superres(47, 8, 0), False),
    "syn_sr_sb128":     (_superres(48, 10, 1, w=392, h=264, sb128=1), False),
    # larger frames: many superblock rows / columns, restoration units of every size
    "syn_720p_10":      (_plain(1280, 720, 10, 1, seed=51, n_inter=2, inf=dict(tile_cols_log2=1)), False),
    "syn_sr_1080p":     (_superres(52, 8, 1, w=1920, h=1080), False),
    "syn_2160p_10_grain_warp": (_uhd_grain_warp(61), True),
}


def build(name):
    """-> (OBU bytes of the whole stream, list of temporal units)"""
    st = STREAMS[name][0]()
    tus = st.samples()
    return b"".join(tus), tus


def load_manifest():
    return json.load(open(MANIFEST))


def main():
    from tests import refdec
    only = sys.argv[1:]
    man = load_manifest() if os.path.exists(MANIFEST) else {}
    for name in STREAMS:
        if only and name not in only:
            continue
        data, tus = build(name)
        frames = []
        for info, planes in refdec.decode_all_big(data, 3840 * 2160 * 4):
            frames.append({"md5": hashlib.md5(planes).hexdigest(), "w": info.w, "h": info.h, "bpc": info.bpc,
                           "layout": info.layout})
        man[name] = {"sha256": hashlib.sha256(data).hexdigest(), "bytes": len(data), "tus": len(tus), "frames": frames}
        print(name, len(data), "bytes ->", len(frames), "frames", sorted({(f["w"], f["h"]) for f in frames}))
    json.dump(man, open(MANIFEST, "w"), indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
