#!/usr/bin/env python3
"""Synthetic AV1 streams for tools the offline encoders cannot be made to emit. TEST INFRASTRUCTURE.

No stand-alone AV1 encoder exists in this image and the AVIF writers (Pillow / OpenCV over libaom)
expose neither super-resolution, reference scaling, 12-bit, global motion nor frame-header film
grain. This module writes the sequence and frame headers itself, bit by bit, in the order
ext/dav1d/src/obu.c parses them (parse_seq_hdr :72-300, parse_frame_hdr :409-1150, read_frame_size
:341-396), and fills the tile data with seeded random bytes. An arithmetic decoder never rejects
input, so the reference decodes such a tile as a frame whose symbols are drawn from the default
CDFs: every partition shape, prediction mode, transform size / type, interpolation filter,
compound / OBMC / warp mode the header allows turns up, at any frame size, bit depth and layout.
The golden answer is, as for every other fixture, the per-frame MD5 of the reference's own decode.
"""
import numpy as np


class BitWriter:
    def __init__(self):
        self.bits = []

    def u(self, n, v):
        assert 0 <= v < (1 << n), (n, v)
        for i in range(n - 1, -1, -1):
            self.bits.append((v >> i) & 1)

    def su(self, n, v):          # dav1d_get_sbits: n-bit two's complement
        self.u(n, v & ((1 << n) - 1))

    def uniform(self, mx, v):    # inverse of dav1d_get_uniform
        l = mx.bit_length()
        m = (1 << l) - mx
        if v < m:
            self.u(l - 1, v)
        else:
            self.u(l - 1, (v + m) >> 1)
            self.u(1, (v + m) & 1)

    def subexp_u(self, ref, n, target):   # inverse of get_bits_subexp_u
        # recenter
        if ref * 2 <= n:
            r, t = ref, target
        else:
            r, t = n - ref, n - target
        v = t if t > 2 * r else (2 * (t - r) if t >= r else 2 * (r - t) - 1)
        cur, i = 0, 0
        while True:
            b = 3 + i - 1 if i else 3
            if n < cur + 3 * (1 << b):
                self.uniform(n - cur + 1, v - cur)
                return
            if v < cur + (1 << b):
                self.u(1, 0)
                self.u(b, v - cur)
                return
            self.u(1, 1)
            cur += 1 << b
            i += 1

    def subexp(self, ref, n, target):     # inverse of dav1d_get_bits_subexp
        self.subexp_u(ref + (1 << n), 2 << n, target + (1 << n))

    def align(self):
        while len(self.bits) % 8:
            self.bits.append(0)

    def bytes(self):
        self.align()
        a = np.array(self.bits, dtype=np.uint8).reshape(-1, 8)
        return bytes(np.packbits(a, axis=1).reshape(-1))


def leb128(v):
    out = bytearray()
    while True:
        b = v & 0x7f
        v >>= 7
        out.append(b | (0x80 if v else 0))
        if not v:
            return bytes(out)


def obu(obu_type, payload):
    return bytes([obu_type << 3 | 2]) + leb128(len(payload)) + payload


class Seq:
    """Sequence-header fields (defaults: 8-bit 4:2:0, every tool allowed)."""
    def __init__(self, w, h, bpc=8, layout=1, **kw):
        self.w, self.h, self.bpc, self.layout = w, h, bpc, layout   # layout: 0 I400 1 I420 2 I422 3 I444
        self.sb128 = 0
        self.filter_intra = 1
        self.intra_edge_filter = 1
        self.inter_intra = 1
        self.masked_compound = 1
        self.warped_motion = 1
        self.dual_filter = 1
        self.order_hint = 1
        self.jnt_comp = 1
        self.ref_frame_mvs = 1
        self.screen_content_tools = 0     # 0 off, 1 on, 2 adaptive (per frame)
        self.force_integer_mv = 2         # 2 adaptive
        self.order_hint_n_bits = 7
        self.super_res = 0
        self.cdef = 1
        self.restoration = 1
        self.film_grain_present = 0
        self.separate_uv_delta_q = 0
        self.mtrx_identity = 0
        for k, v in kw.items():
            assert hasattr(self, k), k
            setattr(self, k, v)
        self.mono = layout == 0
        self.ss_hor = int(layout in (0, 1, 2))
        self.ss_ver = int(layout in (0, 1))
        if bpc == 12 or layout == 2:
            self.profile = 2
        elif layout == 3:
            self.profile = 1
        else:
            self.profile = 0

    def write(self):
        b = BitWriter()
        b.u(3, self.profile)
        b.u(1, 0)                     # still_picture
        b.u(1, 0)                     # reduced_still_picture_header
        b.u(1, 0)                     # timing_info_present
        b.u(1, 0)                     # display_model_info_present
        b.u(5, 0)                     # num_operating_points - 1
        b.u(12, 0)                    # idc
        b.u(3, 4)                     # major_level - 2  (6.x: no size limits that matter here)
        b.u(2, 0)                     # minor_level
        b.u(1, 0)                     # tier (major_level > 3)
        wb, hb = max(1, (self.w - 1).bit_length()), max(1, (self.h - 1).bit_length())
        b.u(4, wb - 1); b.u(4, hb - 1)
        b.u(wb, self.w - 1); b.u(hb, self.h - 1)
        self.width_n_bits, self.height_n_bits = wb, hb
        b.u(1, 0)                     # frame_id_numbers_present
        b.u(1, self.sb128); b.u(1, self.filter_intra); b.u(1, self.intra_edge_filter)
        b.u(1, self.inter_intra); b.u(1, self.masked_compound); b.u(1, self.warped_motion)
        b.u(1, self.dual_filter); b.u(1, self.order_hint)
        if self.order_hint:
            b.u(1, self.jnt_comp); b.u(1, self.ref_frame_mvs)
        if self.screen_content_tools == 2:
            b.u(1, 1)
        else:
            b.u(1, 0); b.u(1, self.screen_content_tools)
        if self.screen_content_tools:
            if self.force_integer_mv == 2:
                b.u(1, 1)
            else:
                b.u(1, 0); b.u(1, self.force_integer_mv)
        if self.order_hint:
            b.u(3, self.order_hint_n_bits - 1)
        b.u(1, self.super_res); b.u(1, self.cdef); b.u(1, self.restoration)
        # color config
        b.u(1, int(self.bpc > 8))
        if self.profile == 2 and self.bpc > 8:
            b.u(1, int(self.bpc == 12))
        if self.profile != 1:
            b.u(1, int(self.mono))
        if self.mtrx_identity:        # sRGB / identity: layout 4:4:4 implied, full range
            assert self.layout == 3
            b.u(1, 1); b.u(8, 1); b.u(8, 13); b.u(8, 0)
        else:
            b.u(1, 0)                 # color_description_present
            if self.mono:
                b.u(1, 0)             # color_range
            else:
                b.u(1, 0)             # color_range
                if self.profile == 2 and self.bpc == 12:
                    b.u(1, self.ss_hor)
                    if self.ss_hor:
                        b.u(1, self.ss_ver)
                else:
                    assert (self.profile == 0 and self.layout == 1) or (self.profile == 1 and self.layout == 3) or \
                           (self.profile == 2 and self.layout == 2), "layout not expressible in this profile"
                if self.ss_hor and self.ss_ver:
                    b.u(2, 0)         # chroma_sample_position
        if not self.mono:
            b.u(1, self.separate_uv_delta_q)
        b.u(1, self.film_grain_present)
        b.u(1, 1)                     # trailing one bit
        return b.bytes()


def _poc_diff(bits, a, b):
    if not bits:
        return 0
    m = 1 << (bits - 1)
    d = a - b
    return (d & (m - 1)) - (d & m)


class Stream:
    """Builds a temporal-unit list. Tracks the 8 reference slots the way the decoder does, as far as the
    header syntax depends on them (sizes for size-from-ref, order hints for skip mode)."""

    def __init__(self, seq, seed):
        self.seq = seq
        self.rng = np.random.default_rng(seed)
        self.slots = [None] * 8            # dicts: w (upscaled), h, offset
        self.tus = []
        self.first = True

    def frame(self, ftype, **kw):
        """ftype: 0 key, 1 inter, 2 intra-only. Keyword fields override the defaults below."""
        s = self.seq
        p = dict(show_frame=1, showable=0, error_res=0, disable_cdf_update=0, screen=0, force_integer_mv=0,
                 size=None,                # (w, h) upscaled frame size, None = sequence maximum
                 superres_denom=8, render_size=None, allow_intrabc=0, refresh=0xff, offset=0,
                 primary_ref=7, refidx=(0, 1, 2, 3, 4, 5, 6), size_from_ref=None, hp=1, filter=4,  # 4 = switchable
                 switchable_motion_mode=1, use_ref_frame_mvs=0, refresh_context=1,
                 tile_cols_log2=0, tile_rows_log2=0, qidx=120, ydc=0, udc=0, uac=0, vdc=0, vac=0, qm=None,
                 seg=None, delta_q=None, delta_lf=None,
                 lf=(12, 10, 8, 8), sharp=0, lf_deltas=None, cdef=None, lr=None, lr_unit=None,
                 tx_switchable=1, ref_select=1, skip_mode=1, warp_motion=1, reduced_txtp=0, gmv=None,
                 grain=None, tile_bytes=None)
        for k, v in kw.items():
            assert k in p, k
            p[k] = v
        b = BitWriter()
        key, inter, intra_only = ftype == 0, ftype == 1, ftype == 2
        b.u(1, 0)                                          # show_existing_frame
        b.u(2, ftype); b.u(1, p["show_frame"])
        if not p["show_frame"]:
            b.u(1, p["showable"])
        error_res = 1 if (key and p["show_frame"]) else p["error_res"]
        if not (key and p["show_frame"]):
            b.u(1, error_res)
        b.u(1, p["disable_cdf_update"])
        screen = p["screen"] if s.screen_content_tools == 2 else s.screen_content_tools
        if s.screen_content_tools == 2:
            b.u(1, screen)
        fim = 0
        if screen:
            fim = p["force_integer_mv"] if s.force_integer_mv == 2 else s.force_integer_mv
            if s.force_integer_mv == 2:
                b.u(1, fim)
        if not inter:
            fim = 1
        size = p["size"]
        override = int(size is not None and size != (s.w, s.h)) or int(p["size_from_ref"] is not None)
        b.u(1, override)
        offset = p["offset"] & ((1 << s.order_hint_n_bits) - 1) if s.order_hint else 0
        if s.order_hint:
            b.u(s.order_hint_n_bits, offset)
        primary_ref = 7
        if not error_res and inter:
            primary_ref = p["primary_ref"]
            b.u(3, primary_ref)
        refresh = 0xff if (key and p["show_frame"]) else p["refresh"]

        def frame_size():
            nonlocal size
            if override:
                w, h = size
                b.u(s.width_n_bits, w - 1); b.u(s.height_n_bits, h - 1)
            else:
                size = (s.w, s.h)
            superres(size[0])
            if p["render_size"]:
                b.u(1, 1); b.u(16, p["render_size"][0] - 1); b.u(16, p["render_size"][1] - 1)
            else:
                b.u(1, 0)

        def superres(w1):
            nonlocal w0
            d = p["superres_denom"]
            if s.super_res:
                b.u(1, int(d != 8))
                if d != 8:
                    assert 9 <= d <= 16
                    b.u(3, d - 9)
            else:
                assert d == 8
            w0 = max((w1 * 8 + (d >> 1)) // d, min(16, w1)) if d != 8 else w1

        w0 = None
        refidx = list(p["refidx"])
        if not inter:
            if not (key and p["show_frame"]):
                b.u(8, refresh)
            if refresh != 0xff and error_res and s.order_hint:
                for i in range(8):
                    b.u(s.order_hint_n_bits, self.slots[i]["offset"] if self.slots[i] else 0)
            frame_size()
            if screen and p["superres_denom"] == 8:
                b.u(1, p["allow_intrabc"])
            allow_intrabc = p["allow_intrabc"] if (screen and p["superres_denom"] == 8) else 0
        else:
            allow_intrabc = 0
            b.u(8, refresh)
            if error_res and s.order_hint:
                for i in range(8):
                    b.u(s.order_hint_n_bits, self.slots[i]["offset"] if self.slots[i] else 0)
            if s.order_hint:
                b.u(1, 0)                                  # frame_refs_short_signaling
            for i in range(7):
                assert self.slots[refidx[i]] is not None, "reference slot %d is empty" % refidx[i]
                b.u(3, refidx[i])
            if not error_res and override:
                k = p["size_from_ref"]
                for i in range(7):
                    if k is not None and i == k:
                        b.u(1, 1)
                        rs = self.slots[refidx[i]]
                        size = (rs["w"], rs["h"])
                        superres(size[0])
                        break
                    b.u(1, 0)
                else:
                    frame_size()
            else:
                assert p["size_from_ref"] is None
                frame_size()
            if not fim:
                b.u(1, p["hp"])
            if p["filter"] == 4:
                b.u(1, 1)
            else:
                b.u(1, 0); b.u(2, p["filter"])
            b.u(1, p["switchable_motion_mode"])
            if not error_res and s.ref_frame_mvs and s.order_hint:
                b.u(1, p["use_ref_frame_mvs"])
        if not p["disable_cdf_update"]:
            b.u(1, int(not p["refresh_context"]))
        # tiling (uniform)
        b.u(1, 1)
        sbl = 6 + s.sb128
        sbw = (w0 + (1 << sbl) - 1) >> sbl
        sbh = (size[1] + (1 << sbl) - 1) >> sbl

        def tl2(sz, tgt):
            k = 0
            while (sz << k) < tgt:
                k += 1
            return k
        min_cols = tl2(4096 >> sbl, sbw)
        max_cols = tl2(1, min(sbw, 64))
        max_rows = tl2(1, min(sbh, 64))
        min_tiles = max(tl2(4096 * 2304 >> (2 * sbl), sbw * sbh), min_cols)
        lc = max(min_cols, min(max_cols, p["tile_cols_log2"]))
        for k in range(min_cols, lc):
            b.u(1, 1)
        if lc < max_cols:
            b.u(1, 0)
        tile_w = 1 + ((sbw - 1) >> lc)
        cols = (sbw + tile_w - 1) // tile_w
        min_rows = max(min_tiles - lc, 0)
        lr_ = max(min_rows, min(max_rows, p["tile_rows_log2"]))
        for k in range(min_rows, lr_):
            b.u(1, 1)
        if lr_ < max_rows:
            b.u(1, 0)
        tile_h = 1 + ((sbh - 1) >> lr_)
        rows = (sbh + tile_h - 1) // tile_h
        n_size_bytes = 4
        if lc or lr_:
            b.u(lc + lr_, 0)                               # context_update_tile_id
            b.u(2, n_size_bytes - 1)
        # quant
        b.u(8, p["qidx"])

        def delta(v):
            if v:
                b.u(1, 1); b.su(7, v)
            else:
                b.u(1, 0)
        delta(p["ydc"])
        if not s.mono:
            diff_uv = int(s.separate_uv_delta_q and (p["vdc"] != p["udc"] or p["vac"] != p["uac"]))
            if s.separate_uv_delta_q:
                b.u(1, diff_uv)
            delta(p["udc"]); delta(p["uac"])
            if diff_uv:
                delta(p["vdc"]); delta(p["vac"])
        if p["qm"]:
            b.u(1, 1); b.u(4, p["qm"][0]); b.u(4, p["qm"][1])
            if s.separate_uv_delta_q:
                b.u(4, p["qm"][2])
        else:
            b.u(1, 0)
        # segmentation: seg = list of 8 dicts(delta_q, lf_y_v, lf_y_h, lf_u, lf_v, ref, skip, globalmv) or None
        seg = p["seg"]
        b.u(1, int(seg is not None))
        seg_dq = [0] * 8
        if seg is not None:
            if primary_ref != 7:
                b.u(1, 1); b.u(1, 0); b.u(1, 1)            # update_map, temporal, update_data
            for i in range(8):
                d = seg[i] if i < len(seg) else {}
                for name, nb in (("delta_q", 9), ("lf_y_v", 7), ("lf_y_h", 7), ("lf_u", 7), ("lf_v", 7)):
                    if d.get(name):
                        b.u(1, 1); b.su(nb, d[name])
                    else:
                        b.u(1, 0)
                if d.get("ref") is not None:
                    b.u(1, 1); b.u(3, d["ref"])
                else:
                    b.u(1, 0)
                b.u(1, int(bool(d.get("skip")))); b.u(1, int(bool(d.get("globalmv"))))
                seg_dq[i] = d.get("delta_q", 0) or 0
        delta_lf_present = 0
        if p["qidx"]:
            dq = p["delta_q"]
            b.u(1, int(dq is not None))
            if dq is not None:
                b.u(2, dq)
                if not allow_intrabc:
                    dl = p["delta_lf"]
                    b.u(1, int(dl is not None))
                    if dl is not None:
                        b.u(2, dl[0]); b.u(1, dl[1])
                        delta_lf_present = 1
        delta_lossless = not (p["ydc"] or p["udc"] or p["uac"] or p["vdc"] or p["vac"])
        all_lossless = all(min(255, max(0, p["qidx"] + (seg_dq[i] if seg is not None else 0))) == 0 and delta_lossless
                           for i in range(8))
        # loop filter
        if not (all_lossless or allow_intrabc):
            lf = p["lf"]
            b.u(6, lf[0]); b.u(6, lf[1])
            if not s.mono and (lf[0] or lf[1]):
                b.u(6, lf[2]); b.u(6, lf[3])
            b.u(3, p["sharp"])
            ld = p["lf_deltas"]
            b.u(1, int(ld is not None))
            if ld is not None:
                b.u(1, 1)
                for i in range(8):
                    if ld[0][i] is not None:
                        b.u(1, 1); b.su(7, ld[0][i])
                    else:
                        b.u(1, 0)
                for i in range(2):
                    if ld[1][i] is not None:
                        b.u(1, 1); b.su(7, ld[1][i])
                    else:
                        b.u(1, 0)
        # cdef: (damping 3..6, [(y_strength, uv_strength), ... 1/2/4/8 entries])
        if not all_lossless and s.cdef and not allow_intrabc:
            cd = p["cdef"] or (3, [(0, 0)])
            n = len(cd[1])
            assert n in (1, 2, 4, 8)
            b.u(2, cd[0] - 3); b.u(2, n.bit_length() - 1)
            for y, uv in cd[1]:
                b.u(6, y)
                if not s.mono:
                    b.u(6, uv)
        # restoration: lr = (type_y, type_u, type_v) in header coding: 0 none 1 switchable 2 wiener 3 sgrproj
        if (not all_lossless or p["superres_denom"] != 8) and s.restoration and not allow_intrabc:
            lr = p["lr"] or (0, 0, 0)
            b.u(2, lr[0])
            if not s.mono:
                b.u(2, lr[1]); b.u(2, lr[2])
            if any(lr[:1 if s.mono else 3]):
                lu = p["lr_unit"] or (0, 0)   # (extra log2 of the luma unit 0..2 [sb128: 0..1], chroma one size smaller?)
                if s.sb128:
                    b.u(1, int(lu[0] > 0))
                else:
                    b.u(1, int(lu[0] > 0))
                    if lu[0] > 0:
                        b.u(1, int(lu[0] > 1))
                if not s.mono and (lr[1] or lr[2]) and s.ss_hor and s.ss_ver:
                    b.u(1, lu[1])
        if not all_lossless:
            b.u(1, p["tx_switchable"])
        ref_select = 0
        if inter:
            ref_select = p["ref_select"]
            b.u(1, ref_select)
        skip_allowed = 0
        if ref_select and inter and s.order_hint:
            nb = s.order_hint_n_bits
            off_before = off_after = -1
            for i in range(7):
                refpoc = self.slots[refidx[i]]["offset"]
                d = _poc_diff(nb, refpoc, offset)
                if d > 0:
                    if off_after < 0 or _poc_diff(nb, off_after, refpoc) > 0:
                        off_after = refpoc
                elif d < 0 and (off_before < 0 or _poc_diff(nb, refpoc, off_before) > 0):
                    off_before = refpoc
            if off_before >= 0 and off_after >= 0:
                skip_allowed = 1
            elif off_before >= 0:
                off_before2 = -1
                for i in range(7):
                    refpoc = self.slots[refidx[i]]["offset"]
                    if _poc_diff(nb, refpoc, off_before) < 0:
                        if off_before2 < 0 or _poc_diff(nb, refpoc, off_before2) > 0:
                            off_before2 = refpoc
                if off_before2 >= 0:
                    skip_allowed = 1
        if skip_allowed:
            b.u(1, p["skip_mode"])
        if not error_res and inter and s.warped_motion:
            b.u(1, p["warp_motion"])
        b.u(1, p["reduced_txtp"])
        if inter:
            # gmv: dict ref index -> (type, [m0..m5]); types 1 translation, 2 rot-zoom, 3 affine (dav1d enum);
            # coded against the default parameters (primary_ref_frame none) only
            gmv = p["gmv"] or {}
            assert not gmv or primary_ref == 7, "gmv deltas are coded against the defaults only"
            for i in range(7):
                if i not in gmv:
                    b.u(1, 0)
                    continue
                typ, m = gmv[i]
                b.u(1, 1)
                if typ == 2:
                    b.u(1, 1)
                else:
                    b.u(1, 0); b.u(1, int(typ == 1))
                if typ >= 2:
                    b.subexp(0, 12, (m[2] - (1 << 16)) >> 1)
                    b.subexp(0, 12, m[3] >> 1)
                    bits, shift = 12, 10
                else:
                    hp_eff = 0 if fim else p["hp"]
                    bits, shift = 9 - int(not hp_eff), 13 + int(not hp_eff)
                if typ == 3:
                    b.subexp(0, 12, m[4] >> 1)
                    b.subexp(0, 12, (m[5] - (1 << 16)) >> 1)
                b.subexp(0, bits, m[0] >> shift)
                b.subexp(0, bits, m[1] >> shift)
        if s.film_grain_present and (p["show_frame"] or p["showable"]):
            g = p["grain"]
            b.u(1, int(g is not None))
            if g is not None:
                b.u(16, g["seed"])
                if inter:
                    b.u(1, 1)                              # update_grain
                b.u(4, len(g["y"]))
                for x, y in g["y"]:
                    b.u(8, x); b.u(8, y)
                cfl = g.get("cfl", 0)
                if not s.mono:
                    b.u(1, cfl)
                n_uv = [0, 0]
                if not (s.mono or cfl or (s.ss_ver and s.ss_hor and not g["y"])):
                    for pl in range(2):
                        pts = g["uv"][pl]
                        n_uv[pl] = len(pts)
                        b.u(4, len(pts))
                        for x, y in pts:
                            b.u(8, x); b.u(8, y)
                b.u(2, g["scaling_shift"] - 8)
                lag = g["lag"]
                b.u(2, lag)
                n_y = 2 * lag * (lag + 1)
                if g["y"]:
                    for i in range(n_y):
                        b.u(8, g["ar_y"][i] + 128)
                for pl in range(2):
                    if n_uv[pl] or cfl:
                        for i in range(n_y + int(bool(g["y"]))):
                            b.u(8, g["ar_uv"][pl][i] + 128)
                b.u(2, g["ar_shift"] - 6); b.u(2, g["grain_scale_shift"])
                for pl in range(2):
                    if n_uv[pl]:
                        b.u(8, g["uv_mult"][pl] + 128); b.u(8, g["uv_luma_mult"][pl] + 128); b.u(9, g["uv_offset"][pl] + 256)
                b.u(1, g.get("overlap", 1)); b.u(1, g.get("clip", 0))
        hdr = b.bytes()
        # tile group: every tile's symbols are random bytes; all but the last carry a size prefix
        n_tiles = cols * rows
        per = p["tile_bytes"]
        if per is None:
            bpp = 3.0 if p["qidx"] < 60 else 2.0 if p["qidx"] < 140 else 1.2
            chroma = {0: 1.0, 1: 1.5, 2: 2.0, 3: 3.0}[s.layout]
            per = int(w0 * size[1] * chroma * bpp / 8 / n_tiles) + 4096
        tg = bytearray()
        if n_tiles > 1:
            tg.append(0)                                   # tile_start_and_end_present_flag = 0 (+ alignment)
        for t in range(n_tiles):
            data = self.rng.integers(0, 256, size=per, dtype=np.uint8).tobytes()
            if t + 1 < n_tiles:
                tg += (per - 1).to_bytes(n_size_bytes, "little")
            tg += data
        tu = b""
        tu += obu(2, b"")                                  # temporal delimiter
        if self.first or key:
            tu += obu(1, s.write())
            self.first = False
        tu += obu(6, hdr + bytes(tg))                      # OBU_FRAME
        self.tus.append(tu)
        for i in range(8):
            if refresh >> i & 1:
                self.slots[i] = dict(w=size[0], h=size[1], offset=offset)
        return tu

    def show_existing(self, idx):
        b = BitWriter()
        b.u(1, 1); b.u(3, idx)
        tu = obu(2, b"") + obu(3, self._hdr_only(b))
        self.tus.append(tu)
        return tu

    @staticmethod
    def _hdr_only(b):
        b.u(1, 1)          # trailing bit of a stand-alone frame header OBU
        return b.bytes()

    def samples(self):
        return list(self.tus)
