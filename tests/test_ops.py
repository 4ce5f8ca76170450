"""Operator-level parity through the DSP surface: every operator of dav1d's
Dav1dDSPContext that the engine replaces is driven on its own, through the C ABI, with inputs generated the way
the reference's own harness does (ext/dav1d/tests/checkasm/{itx,mc,ipred,loopfilter,cdef,looprestoration}.c:
random samples over the full range of the bit depth, random coefficients / levels / strengths), and compared
bit-exactly (incl. everything around the written block) with the reference's C operator called through
oracle/_ref (oracle/ref_harness.c: ref_itxfm_add, ref_mc_put[_scaled], ref_mc_prep[_scaled], ref_avg, ...).

Per-call mode = eav1cu_frame_submit with a list of independent commands (one per operator call, each in its own
cell of a large picture) and no post filters, or with post-filter side data only. The same tests run against the
host build of the device code (tests/emu, `-m "not gpu"`) and against the CUDA kernels (`-m gpu`)."""
import ctypes as C
import os
import struct

import numpy as np
import pytest

from tests import ops, refdec

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

OP_MC, OP_WARP, OP_AVG, OP_W_AVG, OP_MASK, OP_W_MASK, OP_BLEND, OP_BLEND_V, OP_BLEND_H, OP_IPRED, OP_CFL, OP_PAL, OP_ITX = range(1, 14)
F_DST_ARENA, F_PREP, F_BLOCK_HEAD, F_MASK_ARENA = 1, 2, 4, 8
FILT_H = [0, 0, 0, 2, 2, 2, 1, 1, 1, 3]          # enum Filter2d -> (horizontal, vertical) 1-D filter, as host/rec_tmpl.c
FILT_V = [0, 1, 2, 0, 1, 2, 0, 1, 2, 3]
TX_DIM = {0: (4, 4), 1: (8, 8), 2: (16, 16), 3: (32, 32), 4: (64, 64), 5: (4, 8), 6: (8, 4), 7: (8, 16), 8: (16, 8),
          9: (16, 32), 10: (32, 16), 11: (32, 64), 12: (64, 32), 13: (4, 16), 14: (16, 4), 15: (8, 32), 16: (32, 8),
          17: (16, 64), 18: (64, 16)}


def cmd(op, plane=0, w=0, h=0, x=0, y=0, dst_off=0, dst_stride=0, flags=0, body=b""):
    hdr = struct.pack("<BBBBhhIHBB", op, plane, w, h, x, y, dst_off, dst_stride, flags | F_BLOCK_HEAD, 0)
    assert len(body) <= 44
    return hdr + body + bytes(44 - len(body))


def mc_body(sx, sy, mx, my, ref, filt_h, filt_v, dx=0, dy=0, iw=1 << 24, ih=1 << 24):
    return struct.pack("<iiiiHHHHBBBB", sx, sy, iw, ih, mx, my, dx, dy, ref, filt_h, filt_v, 0)


class FrameDesc(C.Structure):
    _fields_ = [("w", C.c_int), ("h", C.c_int), ("bpc", C.c_int), ("layout", C.c_int), ("bw4", C.c_int), ("bh4", C.c_int),
                ("sb128", C.c_int), ("cur", C.c_int), ("ref", C.c_int * 7),
                ("lf_recs", C.c_void_p), ("n_lf_recs", C.c_size_t), ("lf_level", C.c_void_p), ("lf_level_bytes", C.c_size_t),
                ("b4_stride", C.c_int), ("lut_e", C.c_uint8 * 64), ("lut_i", C.c_uint8 * 64),
                ("cdef_enable", C.c_int), ("cdef_damping", C.c_int), ("cdef_y", C.c_uint8 * 8), ("cdef_uv", C.c_uint8 * 8),
                ("lf_mask", C.c_void_p), ("sb128w", C.c_int), ("sb128h", C.c_int),
                ("lr_recs", C.c_void_p), ("n_lr_recs", C.c_size_t),
                ("superres", C.c_int), ("sr_pic", C.c_int), ("sr_w", C.c_int),
                ("resize_step", C.c_int * 2), ("resize_start", C.c_int * 2)]


class Rig(ops.Engine):
    """one engine context + helpers to run a list of commands on pictures held as numpy arrays"""

    def __init__(self, so):
        super().__init__(so)
        E = self.E
        E.eav1cu_frame_submit.argtypes = [C.c_void_p, C.POINTER(FrameDesc), C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.c_size_t]
        E.eav1cu_arena_upload.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t]
        E.eav1cu_arena_download.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t]
        E.eav1cu_abi_sizeof.restype = C.c_size_t
        E.eav1cu_abi_sizeof.argtypes = [C.c_int]
        assert E.eav1cu_abi_sizeof(3) in (0, C.sizeof(FrameDesc)), (E.eav1cu_abi_sizeof(3), C.sizeof(FrameDesc))

    def desc(self, w, h, bpc, layout, cur, refs=()):
        fd = FrameDesc()
        fd.w, fd.h, fd.bpc, fd.layout = w, h, bpc, layout
        fd.bw4, fd.bh4, fd.cur = (w + 3) // 4, (h + 3) // 4, cur
        for i in range(7):
            fd.ref[i] = refs[i] if i < len(refs) else -1
        return fd

    def submit(self, fd, cmds=b"", payload=b"", arena_bytes=0):
        pl = C.create_string_buffer(payload, max(len(payload), 1))
        cm = C.create_string_buffer(cmds, max(len(cmds), 1))
        self.check(self.E.eav1cu_frame_submit(self.ctx, C.byref(fd), cm, len(cmds) // 60, pl, len(payload), arena_bytes))
        self.check(self.E.eav1cu_sync(self.ctx))

    def arena_put(self, off, arr):
        a = np.ascontiguousarray(arr)
        self.check(self.E.eav1cu_arena_upload(self.ctx, off, a.ctypes.data, a.nbytes))

    def arena_get(self, off, shape, dtype):
        out = np.zeros(shape, dtype)
        self.check(self.E.eav1cu_arena_download(self.ctx, off, out.ctypes.data, out.nbytes))
        return out


def engines():
    return [pytest.param(ops.EMU_SO, id="emu"), pytest.param(ops.CUDA_SO, id="cuda", marks=pytest.mark.gpu)]


def R():
    L = refdec.lib()
    return L


def ptr(a, y=0, x=0):
    return C.c_void_p(a.ctypes.data + y * a.strides[0] + x * a.itemsize)


def rand_pic(rng, h, w, bpc):
    return rng.integers(0, 1 << bpc, size=(h, w)).astype(np.uint16 if bpc > 8 else np.uint8)


def skip_if_missing(so):
    if not os.path.exists(so):
        pytest.skip(os.path.basename(so) + " not built")


# ------------------------------------------------------------------------------------------- K1 inverse transforms

VALID_ITX = {}
for _tx in (0, 5, 13, 6, 1, 7, 14, 8):
    VALID_ITX[_tx] = list(range(16))
VALID_ITX[0] = list(range(17))
VALID_ITX[2] = list(range(12))
for _tx in (15, 9, 16, 10, 3):
    VALID_ITX[_tx] = [0, 9]
for _tx in (17, 11, 18, 12, 4):
    VALID_ITX[_tx] = [0]


@pytest.mark.parametrize("so", engines())
@pytest.mark.parametrize("bpc", [8, 10, 12])
def test_itxfm_add_all_sizes_and_types(so, bpc):
    """itx.itxfm_add[tx][txtp] (itx_tmpl.c:43-203): every entry of the table, dc-only / sparse / full coefficient
    blocks, on random destination samples; the destination cell and its surroundings must match and so must the
    zeroing contract (checked on the reference side only: our coefficients live in a payload copy)."""
    skip_if_missing(so)
    rng = np.random.default_rng(100 + bpc)
    rig = Rig(so)
    L = R()
    L.ref_itxfm_add.argtypes = [C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_ssize_t, C.c_void_p, C.c_int]
    L.ref_itxfm_add.restype = None
    cases = [(tx, tp, kind) for tx in sorted(VALID_ITX) for tp in VALID_ITX[tx] for kind in range(4)]
    cell, cols = 96, 32
    rows = (len(cases) + cols - 1) // cols
    W, H = cols * cell, rows * cell
    cdt = np.int32 if bpc > 8 else np.int16
    try:
        pic = rand_pic(rng, H, W, bpc)
        want = pic.copy()
        cur = rig.pic_alloc(W, H, 0, bpc)
        rig.pic_upload(cur, [pic])
        cmds, payload = [], bytearray()
        for i, (tx, tp, kind) in enumerate(cases):
            w, h = TX_DIM[tx]
            sw, sh = min(w, 32), min(h, 32)
            x0, y0 = (i % cols) * cell + 16, (i // cols) * cell + 16
            cf = np.zeros((sw, sh), cdt)                       # column-major: cf[x][y]
            amp = 1 << (bpc + (3 if kind != 3 else 6))
            if tp == 16:                                       # WHT (lossless): residuals are small
                cf[:] = rng.integers(-(1 << bpc), 1 << bpc, size=cf.shape)
                eob = sw * sh - 1
            elif kind == 0:                                    # dc only
                cf[0, 0] = rng.integers(-amp, amp)
                eob = 0
            elif kind == 1:                                    # sparse, full eob: every code path except zero skipping
                n = max(2, sw * sh // 8)
                cf.reshape(-1)[rng.choice(sw * sh, n, replace=False)] = rng.integers(-amp, amp, size=n)
                eob = sw * sh - 1
            else:                                              # dense (kind 3: large magnitudes, clipping paths)
                cf[:] = rng.integers(-amp, amp, size=cf.shape)
                eob = sw * sh - 1
            off = len(payload)
            payload += cf.tobytes()
            payload += bytes(-len(payload) % 16)
            cmds.append(cmd(OP_ITX, 0, w, h, x0, y0, body=struct.pack("<IhBB", off, eob, tx, tp)))
            ref_cf = cf.copy()
            L.ref_itxfm_add(bpc, tx, tp, ptr(want, y0, x0), want.strides[0], ptr(ref_cf), eob)
            assert not ref_cf.any(), "reference itxfm_add must zero its coefficients"
        rig.submit(rig.desc(W, H, bpc, 0, cur), b"".join(cmds), bytes(payload))
        got = rig.pic_download(cur, W, H, 0, bpc)[0]
        bad = np.argwhere(got != want)
        if len(bad):
            i = (bad[0][0] // cell) * cols + bad[0][1] // cell
            raise AssertionError("itxfm_add[tx=%d][txtp=%d] kind %d differs (%d samples in all)" % (cases[i] + (len(bad),)))
    finally:
        rig.close()


# ------------------------------------------------------------------------------------------- K2 motion compensation

def mc_cases(rng, scaled):
    sizes = [(2, 2), (2, 4), (4, 2), (4, 4), (4, 8), (8, 4), (8, 8), (8, 16), (16, 8), (16, 16), (16, 32), (32, 16), (32, 32),
             (64, 32), (32, 64), (64, 64), (128, 64), (64, 128), (128, 128), (4, 16), (16, 4), (8, 32), (32, 8), (16, 64), (64, 16)]
    out = []
    for filt in range(10):
        for (w, h) in sizes:
            for prep in (0, 1):
                if prep and (w < 4 or h < 4):
                    continue
                if scaled:
                    for _ in range(2):
                        out.append((filt, w, h, prep, int(rng.integers(0, 1024)), int(rng.integers(0, 1024)),
                                    int(rng.integers(512, 2049)), int(rng.integers(512, 2049))))
                else:
                    r1, r2 = int(rng.integers(1, 16)), int(rng.integers(1, 16))
                    for mx, my in ((0, 0), (r1, 0), (0, r2), (r1, r2)):
                        out.append((filt, w, h, prep, mx, my, 0, 0))
    return out


@pytest.mark.parametrize("so", engines())
@pytest.mark.parametrize("bpc", [8, 10, 12])
@pytest.mark.parametrize("scaled", [0, 1], ids=["unscaled", "scaled"])
def test_mc_put_prep(so, bpc, scaled):
    """mc.mc / mc.mct and mc.mc_scaled / mc.mct_scaled (mc_tmpl.c:130-626): the ten 2-D filters, every block
    size, integer / horizontal / vertical / 2-D sub-pel positions, scaled steps between 1/2 and 2."""
    skip_if_missing(so)
    rng = np.random.default_rng(200 + bpc + 7 * scaled)
    rig = Rig(so)
    L = R()
    for name, n_int in (("ref_mc_put", 4), ("ref_mc_put_scaled", 6)):
        getattr(L, name).argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_ssize_t, C.c_void_p, C.c_ssize_t] + [C.c_int] * n_int
        getattr(L, name).restype = None
    for name, n_int in (("ref_mc_prep", 4), ("ref_mc_prep_scaled", 6)):
        getattr(L, name).argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_ssize_t] + [C.c_int] * n_int
        getattr(L, name).restype = None
    cases = mc_cases(rng, scaled)
    cell, cols = 160, 16
    rows = (len(cases) + cols - 1) // cols
    W, H = cols * cell, rows * cell
    RW, RH = 1024, 768
    try:
        refpic = rand_pic(rng, RH, RW, bpc)
        pic = rand_pic(rng, H, W, bpc)
        want = pic.copy()
        cur, rp = rig.pic_alloc(W, H, 0, bpc), rig.pic_alloc(RW, RH, 0, bpc)
        rig.pic_upload(cur, [pic])
        rig.pic_upload(rp, [refpic])
        cmds, preps, arena = [], [], 0
        for i, (filt, w, h, prep, mx, my, dx, dy) in enumerate(cases):
            x0, y0 = (i % cols) * cell + 16, (i // cols) * cell + 16
            span_x = ((w - 1) * max(dx, 1024) + mx >> 10) + 8 if scaled else w + 8
            span_y = ((h - 1) * max(dy, 1024) + my >> 10) + 8 if scaled else h + 8
            sx, sy = int(rng.integers(8, RW - span_x - 8)), int(rng.integers(8, RH - span_y - 8))
            body = mc_body(sx, sy, mx, my, 0, FILT_H[filt], FILT_V[filt], dx, dy)
            src = ptr(refpic, sy, sx)
            if prep:
                cmds.append(cmd(OP_MC, 0, w, h, dst_off=arena, dst_stride=w, flags=F_DST_ARENA | F_PREP, body=body))
                tmp = np.zeros((h, w), np.int16)
                if scaled:
                    L.ref_mc_prep_scaled(bpc, filt, ptr(tmp), src, refpic.strides[0], w, h, mx, my, dx, dy)
                else:
                    L.ref_mc_prep(bpc, filt, ptr(tmp), src, refpic.strides[0], w, h, mx, my)
                preps.append((arena, tmp, i))
                arena += (w * h * 2 + 15) & ~15
            else:
                cmds.append(cmd(OP_MC, 0, w, h, x0, y0, body=body))
                if scaled:
                    L.ref_mc_put_scaled(bpc, filt, ptr(want, y0, x0), want.strides[0], src, refpic.strides[0], w, h, mx, my, dx, dy)
                else:
                    L.ref_mc_put(bpc, filt, ptr(want, y0, x0), want.strides[0], src, refpic.strides[0], w, h, mx, my)
        rig.submit(rig.desc(W, H, bpc, 0, cur, [rp]), b"".join(cmds), b"", arena)
        got = rig.pic_download(cur, W, H, 0, bpc)[0]
        bad = np.argwhere(got != want)
        if len(bad):
            i = (bad[0][0] // cell) * cols + bad[0][1] // cell
            raise AssertionError("put %s differs: (filter2d, w, h, prep, mx, my, dx, dy) = %s" % ("scaled" if scaled else "", cases[i]))
        for off, tmp, i in preps:
            g = rig.arena_get(off, tmp.shape, np.int16)
            assert np.array_equal(g, tmp), "prep differs: (filter2d, w, h, prep, mx, my, dx, dy) = %s" % (cases[i],)
    finally:
        rig.close()


@pytest.mark.parametrize("so", engines())
@pytest.mark.parametrize("bpc", [8, 10, 12])
def test_mc_edge_emulation(so, bpc):
    """mc.emu_edge + mc.mc (mc_tmpl.c:868-916; recon_tmpl.c:971-1026): blocks whose filter window leaves the
    reference plane. The engine folds emu_edge into clamped coordinates (sx, sy, iw, ih of the command)."""
    skip_if_missing(so)
    rng = np.random.default_rng(300 + bpc)
    rig = Rig(so)
    L = R()
    L.ref_mc_put.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_ssize_t, C.c_void_p, C.c_ssize_t] + [C.c_int] * 4
    L.ref_mc_put.restype = None
    L.ref_emu_edge.argtypes = [C.c_int] + [C.c_ssize_t] * 6 + [C.c_void_p, C.c_ssize_t, C.c_void_p, C.c_ssize_t]
    L.ref_emu_edge.restype = None
    RW, RH, cell, cols = 200, 120, 96, 12
    cases = []
    for (w, h) in ((8, 8), (16, 16), (32, 16), (64, 64)):
        for sx, sy in ((-20, 5), (-3, -3), (RW - 10, 20), (40, RH - 5), (-70, -70), (RW + 30, RH + 40), (RW - w - 2, -1)):
            for mx, my in ((0, 0), (5, 9)):
                cases.append((w, h, sx, sy, mx, my, int(rng.integers(0, 10))))
    rows = (len(cases) + cols - 1) // cols
    W, H = cols * cell, rows * cell
    try:
        refpic = rand_pic(rng, RH, RW, bpc)
        pic = rand_pic(rng, H, W, bpc)
        want = pic.copy()
        cur, rp = rig.pic_alloc(W, H, 0, bpc), rig.pic_alloc(RW, RH, 0, bpc)
        rig.pic_upload(cur, [pic])
        rig.pic_upload(rp, [refpic])
        cmds = []
        for i, (w, h, sx, sy, mx, my, filt) in enumerate(cases):
            x0, y0 = (i % cols) * cell + 16, (i // cols) * cell + 16
            cmds.append(cmd(OP_MC, 0, w, h, x0, y0, body=mc_body(sx, sy, mx, my, 0, FILT_H[filt], FILT_V[filt], iw=RW, ih=RH)))
            # reference: emu_edge builds the (w + 7) x (h + 7) patch around the block, mc filters from it
            patch = np.zeros((h + 7, 192), refpic.dtype)
            L.ref_emu_edge(bpc, w + 7, h + 7, RW, RH, sx - 3, sy - 3, ptr(patch), patch.strides[0], ptr(refpic), refpic.strides[0])
            L.ref_mc_put(bpc, filt, ptr(want, y0, x0), want.strides[0], ptr(patch, 3, 3), patch.strides[0], w, h, mx, my)
        rig.submit(rig.desc(W, H, bpc, 0, cur, [rp]), b"".join(cmds))
        got = rig.pic_download(cur, W, H, 0, bpc)[0]
        bad = np.argwhere(got != want)
        if len(bad):
            i = (bad[0][0] // cell) * cols + bad[0][1] // cell
            raise AssertionError("edge-emulated put differs: (w, h, sx, sy, mx, my, filter2d) = %s" % (cases[i],))
    finally:
        rig.close()


@pytest.mark.parametrize("so", engines())
@pytest.mark.parametrize("bpc", [8, 10, 12])
def test_compound_and_blend(so, bpc):
    """mc.avg / w_avg / mask / w_mask[444, 422, 420] / blend / blend_v / blend_h (mc_tmpl.c:628-797) on random
    prep blocks within the range prep produces."""
    skip_if_missing(so)
    rng = np.random.default_rng(400 + bpc)
    rig = Rig(so)
    L = R()
    vp, ss = C.c_void_p, C.c_ssize_t
    L.ref_avg.argtypes = [C.c_int, vp, ss, vp, vp, C.c_int, C.c_int]
    L.ref_w_avg.argtypes = [C.c_int, vp, ss, vp, vp, C.c_int, C.c_int, C.c_int]
    L.ref_mask.argtypes = [C.c_int, vp, ss, vp, vp, C.c_int, C.c_int, vp]
    L.ref_w_mask.argtypes = [C.c_int, C.c_int, vp, ss, vp, vp, C.c_int, C.c_int, vp, C.c_int]
    L.ref_blend.argtypes = [C.c_int, vp, ss, vp, C.c_int, C.c_int, vp]
    L.ref_blend_v.argtypes = [C.c_int, vp, ss, vp, C.c_int, C.c_int]
    L.ref_blend_h.argtypes = [C.c_int, vp, ss, vp, C.c_int, C.c_int]
    for n in ("ref_avg", "ref_w_avg", "ref_mask", "ref_w_mask", "ref_blend", "ref_blend_v", "ref_blend_h"):
        getattr(L, n).restype = None
    sizes = [(4, 4), (4, 8), (8, 8), (16, 8), (16, 16), (32, 16), (32, 32), (64, 32), (64, 64), (128, 64), (128, 128)]
    cases = []
    for (w, h) in sizes:
        cases += [("avg", w, h, 0), ("w_avg", w, h, int(rng.integers(1, 16))), ("mask", w, h, 0)]
        for ssi in range(3):
            if ssi and (w < 8 or h < 8):
                continue
            cases += [("w_mask", w, h, ssi * 2 + s) for s in (0, 1)]
        if w <= 32 and h <= 32:
            cases += [("blend", w, h, 0)]
        if w <= 64 and h <= 64:                            # OBMC: blend_v up to 32 wide blocks' halves, blend_h likewise
            cases += [("blend_v", max(w // 2, 2) if w > 4 else w, h, 0), ("blend_h", w, max(h // 2, 2), 0)]
    cell, cols = 160, 12
    rows = (len(cases) + cols - 1) // cols
    W, H = cols * cell, rows * cell
    pdt = np.uint16 if bpc > 8 else np.uint8
    ib = 4 if bpc == 8 else 14 - bpc
    lo, hi = (0, (255 << ib) + 1) if bpc == 8 else (-8192, ((1 << bpc) - 1 << ib) - 8192 + 1)
    masks_blob = None
    try:
        pic = rand_pic(rng, H, W, bpc)
        want = pic.copy()
        cur = rig.pic_alloc(W, H, 0, bpc)
        rig.pic_upload(cur, [pic])
        cmds, arena, checks = [], 0, []

        def put(arr):
            nonlocal arena
            off = arena
            rig.arena_put(off, arr)
            arena += (arr.nbytes + 15) & ~15
            return off
        # blend masks come from the table blob the product uploads: find the ii masks through the host library
        blob_p = C.c_void_p()
        nblob = rig.H.eav1rec_masks(C.byref(blob_p))
        masks_blob = np.frombuffer(C.string_at(blob_p, nblob), np.uint8)
        for i, (kind, w, h, arg) in enumerate(cases):
            x0, y0 = (i % cols) * cell + 16, (i // cols) * cell + 16
            d = ptr(want, y0, x0)
            st = want.strides[0]
            if kind in ("avg", "w_avg", "mask", "w_mask"):
                t1 = rng.integers(lo, hi, size=(h, w)).astype(np.int16)
                t2 = rng.integers(lo, hi, size=(h, w)).astype(np.int16)
                o1, o2 = put(t1), put(t2)
                if kind == "avg":
                    cmds.append(cmd(OP_AVG, 0, w, h, x0, y0, body=struct.pack("<IIIhBBB", o1, o2, 0, 0, 0, 0, 0)))
                    L.ref_avg(bpc, d, st, ptr(t1), ptr(t2), w, h)
                elif kind == "w_avg":
                    cmds.append(cmd(OP_W_AVG, 0, w, h, x0, y0, body=struct.pack("<IIIhBBB", o1, o2, 0, arg, 0, 0, 0)))
                    L.ref_w_avg(bpc, d, st, ptr(t1), ptr(t2), w, h, arg)
                elif kind == "mask":
                    m = rng.integers(0, 65, size=(h, w)).astype(np.uint8)
                    om = put(m)
                    cmds.append(cmd(OP_MASK, 0, w, h, x0, y0, flags=F_MASK_ARENA, body=struct.pack("<IIIhBBB", o1, o2, om, 0, 0, 0, 0)))
                    L.ref_mask(bpc, d, st, ptr(t1), ptr(t2), w, h, ptr(m))
                else:
                    ssi, sign = arg >> 1, arg & 1
                    sh, sv = (0, 0) if ssi == 0 else (1, 0) if ssi == 1 else (1, 1)
                    mw, mh = w >> sh, h >> sv
                    om = arena
                    arena += (mw * mh + 15) & ~15
                    cmds.append(cmd(OP_W_MASK, 0, w, h, x0, y0, body=struct.pack("<IIIhBBB", o1, o2, om, 0, sign, sh, sv)))
                    m = np.zeros((mh, mw), np.uint8)
                    L.ref_w_mask(bpc, ssi, d, st, ptr(t1), ptr(t2), w, h, ptr(m), sign)
                    checks.append((om, m, i))
            else:
                tmp = rand_pic(rng, h, w, bpc)
                ot = put(tmp)
                if kind == "blend":
                    moff = int(rng.integers(0, len(masks_blob) - w * h))
                    cmds.append(cmd(OP_BLEND, 0, w, h, x0, y0, body=struct.pack("<II", ot, moff)))
                    m = np.minimum(masks_blob[moff:moff + w * h], 64).copy()
                    if not np.array_equal(m, masks_blob[moff:moff + w * h]):      # keep to real mask values (0..64)
                        m = masks_blob[moff:moff + w * h].copy()
                    L.ref_blend(bpc, d, st, ptr(tmp), w, h, ptr(m))
                elif kind == "blend_v":
                    cmds.append(cmd(OP_BLEND_V, 0, w, h, x0, y0, body=struct.pack("<II", ot, 0)))
                    L.ref_blend_v(bpc, d, st, ptr(tmp), w, h)
                else:
                    cmds.append(cmd(OP_BLEND_H, 0, w, h, x0, y0, body=struct.pack("<II", ot, 0)))
                    L.ref_blend_h(bpc, d, st, ptr(tmp), w, h)
        rig.submit(rig.desc(W, H, bpc, 0, cur), b"".join(cmds), b"", arena)
        got = rig.pic_download(cur, W, H, 0, bpc)[0]
        bad = np.argwhere(got != want)
        if len(bad):
            i = (bad[0][0] // cell) * cols + bad[0][1] // cell
            raise AssertionError("%s %dx%d (arg %d) differs" % cases[i])
        for off, m, i in checks:
            g = rig.arena_get(off, m.shape, np.uint8)
            assert np.array_equal(g, m), "w_mask mask output differs: %s" % (cases[i],)
    finally:
        rig.close()


@pytest.mark.parametrize("so", engines())
@pytest.mark.parametrize("bpc", [8, 10, 12])
def test_warp8x8(so, bpc):
    """mc.warp8x8 / warp8x8t (mc_tmpl.c:799-866) with checkasm's parameter ranges (tests/checkasm/mc.c:565-640:
    abcd, mx, my = (rnd() & 0x1fff) - 0xa00, which keeps every filter index inside dav1d_mc_warp_filter)."""
    skip_if_missing(so)
    rng = np.random.default_rng(500 + bpc)
    rig = Rig(so)
    L = R()
    L.ref_warp8x8.argtypes = [C.c_int, C.c_void_p, C.c_ssize_t, C.c_void_p, C.c_ssize_t, C.c_void_p, C.c_int, C.c_int]
    L.ref_warp8x8t.argtypes = [C.c_int, C.c_void_p, C.c_ssize_t, C.c_void_p, C.c_ssize_t, C.c_void_p, C.c_int, C.c_int]
    L.ref_warp8x8.restype = L.ref_warp8x8t.restype = None
    n, cell, cols = 96, 32, 16
    W, H, RW, RH = cols * cell, (n // cols) * cell, 256, 256
    try:
        refpic = rand_pic(rng, RH, RW, bpc)
        pic = rand_pic(rng, H, W, bpc)
        want = pic.copy()
        cur, rp = rig.pic_alloc(W, H, 0, bpc), rig.pic_alloc(RW, RH, 0, bpc)
        rig.pic_upload(cur, [pic])
        rig.pic_upload(rp, [refpic])
        cmds, preps, arena = [], [], 0
        for i in range(n):
            x0, y0 = (i % cols) * cell + 8, (i // cols) * cell + 8
            abcd = ((rng.integers(0, 1 << 32, size=4) & 0x1fff) - 0xa00).astype(np.int16)      # checkasm/mc.c:586-587
            if i % 5 == 0:
                abcd[:] = 0
            mx, my = int(rng.integers(0, 1 << 32) & 0x1fff) - 0xa00, int(rng.integers(0, 1 << 32) & 0x1fff) - 0xa00   # :578-579
            sx, sy = int(rng.integers(16, RW - 32)), int(rng.integers(16, RH - 32))
            body = struct.pack("<iiiiii4hB3x", sx, sy, 1 << 24, 1 << 24, mx, my, *[int(v) for v in abcd], 0)
            src = ptr(refpic, sy, sx)
            if i & 1:
                cmds.append(cmd(OP_WARP, 0, 8, 8, dst_off=arena, dst_stride=8, flags=F_DST_ARENA | F_PREP, body=body))
                tmp = np.zeros((8, 8), np.int16)
                L.ref_warp8x8t(bpc, ptr(tmp), 8, src, refpic.strides[0], ptr(abcd), mx, my)
                preps.append((arena, tmp))
                arena += 128
            else:
                cmds.append(cmd(OP_WARP, 0, 8, 8, x0, y0, body=body))
                L.ref_warp8x8(bpc, ptr(want, y0, x0), want.strides[0], src, refpic.strides[0], ptr(abcd), mx, my)
        rig.submit(rig.desc(W, H, bpc, 0, cur, [rp]), b"".join(cmds), b"", arena)
        got = rig.pic_download(cur, W, H, 0, bpc)[0]
        assert np.array_equal(got, want), "warp8x8 differs in %d samples" % int((got != want).sum())
        for off, tmp in preps:
            assert np.array_equal(rig.arena_get(off, (8, 8), np.int16), tmp), "warp8x8t differs"
    finally:
        rig.close()


# ------------------------------------------------------------------------------------------- K2 resize (super-resolution)

@pytest.mark.parametrize("so", engines())
@pytest.mark.parametrize("bpc,layout", [(8, 1), (10, 3), (12, 0), (8, 2)])
def test_resize(so, bpc, layout):
    """mc.resize (mc_tmpl.c:918-955) through the engine's super-resolution stage: every denominator 9..16."""
    skip_if_missing(so)
    rng = np.random.default_rng(600 + bpc)
    rig = Rig(so)
    L = R()
    L.ref_resize.argtypes = [C.c_int, C.c_void_p, C.c_ssize_t, C.c_void_p, C.c_ssize_t] + [C.c_int] * 5
    L.ref_resize.restype = None
    try:
        for denom in range(9, 17):
            w1, h = 264 + 8 * denom, 40
            w0 = max((w1 * 8 + (denom >> 1)) // denom, 16)
            bw4 = ((w0 + 7) >> 3) << 1
            planes = [rng.integers(0, 1 << bpc, size=(ph, pw)).astype(np.uint16 if bpc > 8 else np.uint8)
                      for pw, ph in ops.plane_dims(4 * bw4, h, layout)]
            cur = rig.pic_alloc(4 * bw4, h, layout, bpc)
            sr = rig.pic_alloc(w1, h, layout, bpc)
            rig.pic_upload(cur, planes)
            fd = rig.desc(w0, h, bpc, layout, cur)
            fd.bw4 = bw4
            fd.superres, fd.sr_pic, fd.sr_w = 1, sr, w1
            ss_hor = int(layout != 3)

            def scale_fac(a, b):          # decode.c: f->resize_step / resize_start
                return ((a << 14) + (b >> 1)) // b

            def x0_of(in_w, out_w, step):
                err = out_w * step - (in_w << 14)
                x0 = int((-((out_w - in_w) << 13) + (out_w >> 1)) / out_w) if False else None
                num = -((out_w - in_w) << 13) + (out_w >> 1)
                q = abs(num) // out_w * (1 if num >= 0 else -1)          # C division truncates towards zero
                e2 = abs(err) // 2 * (1 if err >= 0 else -1)
                return (q + 128 - e2) & 0x3fff
            in_cw, out_cw = (w0 + ss_hor) >> ss_hor, (w1 + ss_hor) >> ss_hor
            steps = [scale_fac(w0, w1), scale_fac(in_cw, out_cw)]
            starts = [x0_of(w0, w1, steps[0]), x0_of(in_cw, out_cw, steps[1])]
            for i in range(2):
                fd.resize_step[i], fd.resize_start[i] = steps[i], starts[i]
            rig.submit(fd)
            got = rig.pic_download(sr, w1, h, layout, bpc)
            for pl, src in enumerate(planes):
                c = int(pl > 0)
                dw = (w1 + ss_hor) >> ss_hor if pl else w1
                sw = (4 * bw4 + ss_hor) >> ss_hor if pl else 4 * bw4
                want = np.zeros((src.shape[0], dw), src.dtype)
                L.ref_resize(bpc, ptr(want), want.strides[0], ptr(src), src.strides[0], dw, src.shape[0], sw, steps[c], starts[c])
                assert np.array_equal(got[pl], want), "resize denominator %d plane %d differs" % (denom, pl)
            rig.E.eav1cu_pic_release(rig.ctx, cur)
            rig.E.eav1cu_pic_release(rig.ctx, sr)
    finally:
        rig.close()


# ------------------------------------------------------------------------------------------- K4 deblocking

class FilterLUT(C.Structure):      # Av1FilterLUT, src/lf_mask.h:36-40
    _fields_ = [("e", C.c_uint8 * 64), ("i", C.c_uint8 * 64), ("sharp", C.c_uint64 * 2)]


def calc_eih(sharp):
    """dav1d_calc_eih (src/lf_mask.c): limits for filter level 0..63 at a given sharpness"""
    lut = FilterLUT()
    for level in range(64):
        limit = level
        if sharp > 0:
            limit >>= (sharp + 3) >> 2
            limit = min(limit, 9 - sharp)
        limit = max(limit, 1)
        lut.i[level] = limit
        lut.e[level] = 2 * (level + 2) + limit
    lut.sharp[0] = (sharp + 3) >> 2
    lut.sharp[1] = 9 - sharp if sharp else 0xff
    return lut


@pytest.mark.parametrize("so", engines())
@pytest.mark.parametrize("bpc", [8, 10, 12])
def test_loop_filter_sb(so, bpc):
    """lf.loop_filter_sb[plane][dir] (loopfilter_tmpl.c:172-246): random edge masks of every filter size class,
    random levels incl. the 'take the neighbour's level' fallback, smooth content so that the flat branches run."""
    skip_if_missing(so)
    rng = np.random.default_rng(700 + bpc)
    rig = Rig(so)
    L = R()
    L.ref_loop_filter_sb.argtypes = [C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_ssize_t, C.c_void_p, C.c_void_p, C.c_ssize_t,
                                     C.c_void_p, C.c_int]
    L.ref_loop_filter_sb.restype = None
    W, H = 640, 512
    b4_stride = W // 4
    dt = np.uint16 if bpc > 8 else np.uint8
    try:
        for trial in range(6):
            sharp = int(rng.integers(0, 8))
            lut = calc_eih(sharp)
            planes = []
            for pl in range(3):       # smooth ramps + sparse noise: a mix of flat and busy edges
                yy, xx = np.mgrid[0:H, 0:W]
                base = (xx * int(rng.integers(1, 4)) + yy * int(rng.integers(1, 4))) % (1 << bpc)
                noise = rng.integers(-(1 << (bpc - 6)), (1 << (bpc - 6)) + 1, size=(H, W)) * (rng.random((H, W)) < 0.5)
                planes.append(np.clip(base // 4 * 4 + noise, 0, (1 << bpc) - 1).astype(dt))
            level = rng.integers(0, 64, size=(H // 4, b4_stride, 4)).astype(np.uint8)
            level[rng.random(level.shape[:2]) < 0.3] = 0
            want = [p.copy() for p in planes]
            cur = rig.pic_alloc(W, H, 3, bpc)
            rig.pic_upload(cur, planes)
            recs = bytearray()
            n_recs = 0
            for direction in (0, 1):           # all column edges first, then all row edges, as the engine orders them
                # one call = one edge line of 32 four-sample units; lines 24 samples apart never share a sample
                along = [0, 128, 256, 384] if direction == 0 else [0, 128, 256, 384, 512]
                across = list(range(16, (W if direction == 0 else H) - 16, 24))
                slots = [(a, b) for a in across for b in along if (b + 128 <= (H if direction == 0 else W))]
                for plane in range(3):
                    pick = rng.choice(len(slots), 10, replace=False)
                    for k in pick:
                        a, b = slots[k]
                        x, y = (a, b) if direction == 0 else (b, a)
                        masks = np.zeros(3, np.uint32)
                        bits = int(rng.integers(0, 1 << 32, dtype=np.uint64))
                        sel = rng.integers(0, 3 if plane == 0 else 2, size=32)
                        for bb in range(32):
                            if bits >> bb & 1:
                                masks[sel[bb]] |= np.uint32(1 << bb)
                        lvl_off = ((y // 4) * b4_stride + x // 4) * 4 + (direction if plane == 0 else 1 + plane)
                        recs += struct.pack("<HHBBH3II", x, y, plane, direction, 0, int(masks[0]), int(masks[1]), int(masks[2]), lvl_off)
                        n_recs += 1
                        lv = C.c_void_p(level.ctypes.data + lvl_off)
                        L.ref_loop_filter_sb(bpc, int(plane > 0), direction, ptr(want[plane], y, x), want[plane].strides[0], ptr(masks),
                                             lv, b4_stride, C.byref(lut), 32)
            fd = rig.desc(W, H, bpc, 3, cur)
            rbuf = C.create_string_buffer(bytes(recs), len(recs))
            fd.lf_recs, fd.n_lf_recs = C.addressof(rbuf), n_recs
            fd.lf_level, fd.lf_level_bytes = level.ctypes.data, level.nbytes
            fd.b4_stride = b4_stride
            for i in range(64):
                fd.lut_e[i], fd.lut_i[i] = lut.e[i], lut.i[i]
            rig.submit(fd)
            got = rig.pic_download(cur, W, H, 3, bpc)
            rig.E.eav1cu_pic_release(rig.ctx, cur)
            for pl in range(3):
                assert np.array_equal(got[pl], want[pl]), "loop_filter_sb trial %d plane %d differs in %d samples" % (
                    trial, pl, int((got[pl] != want[pl]).sum()))
    finally:
        rig.close()


# ------------------------------------------------------------------------------------------- K6 loop restoration

class LrParams(C.Structure):      # LooprestorationParams, src/looprestoration.h:49-55 (16-byte aligned union)
    _fields_ = [("raw", C.c_int16 * 16)]


@pytest.mark.parametrize("so", engines())
@pytest.mark.parametrize("bpc", [8, 10, 12])
def test_loop_restoration(so, bpc):
    """lr.wiener[7-tap, 5-tap] and lr.sgr[5x5, 3x3, mix] (looprestoration_tmpl.c:250-1384) on restoration-unit stripes
    of every width class with all edge combinations, random taps / weights within the bitstream's ranges."""
    skip_if_missing(so)
    rng = np.random.default_rng(800 + bpc)
    rig = Rig(so)
    L = R()
    for n in ("ref_lr_wiener", "ref_lr_sgr"):
        getattr(L, n).argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_ssize_t, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int]
        getattr(L, n).restype = None
    sgr_params = np.frombuffer(refdec.table("dav1d_sgr_params"), np.uint16).reshape(16, 2)
    W, H = 1024, 512
    dt = np.uint16 if bpc > 8 else np.uint8
    try:
        for trial in range(4):
            pic = rand_pic(rng, H, W, bpc) if trial & 1 else \
                np.clip(np.add.outer(np.arange(H) * 3, np.arange(W) * 2) % (1 << bpc) + rng.integers(-3, 4, size=(H, W)), 0, (1 << bpc) - 1).astype(dt)
            want = pic.copy()
            cur = rig.pic_alloc(W, H, 0, bpc)
            rig.pic_upload(cur, [pic])
            recs = bytearray()
            n = 0
            # units on a grid 288 x 80 apart (<= 256 wide, <= 64 tall, 3 samples of border on every side stay inside)
            for gy in range(6):
                for gx in range(3):
                    kind = int(rng.integers(0, 5))
                    w = int(rng.choice([16, 36, 64, 100, 128, 200, 256]))
                    h = int(rng.choice([8, 16, 56, 64]))
                    x, y = 16 + gx * 320, 8 + gy * 80
                    edges = int(rng.integers(0, 16))                    # LR_HAVE_LEFT 1, RIGHT 2, TOP 4, BOTTOM 8
                    prm = LrParams()
                    if kind <= 1:
                        f = np.zeros((2, 8), np.int16)
                        for a in range(2):
                            c0 = 0 if kind == 1 else int(rng.integers(-5, 11))
                            c1, c2 = int(rng.integers(-23, 9)), int(rng.integers(-17, 47))
                            f[a, 0] = f[a, 6] = c0
                            f[a, 1] = f[a, 5] = c1
                            f[a, 2] = f[a, 4] = c2
                            f[a, 3] = -(c0 + c1 + c2) * 2
                        if bpc > 8:
                            f[0, 3] += 128                                 # lr_apply_tmpl.c:58-62
                        f[1, 3] += 128
                        C.memmove(prm.raw, f.ctypes.data, 32)
                        body = f.tobytes()
                    else:
                        idx = int(rng.choice([i for i in range(16) if (kind == 2 and sgr_params[i][0] and not sgr_params[i][1]) or
                                              (kind == 3 and sgr_params[i][1] and not sgr_params[i][0]) or
                                              (kind == 4 and sgr_params[i][0] and sgr_params[i][1])]))
                        s0, s1 = int(sgr_params[idx][0]), int(sgr_params[idx][1])
                        w0 = int(rng.integers(-96, 32)) if s0 else 0
                        w1r = int(rng.integers(-32, 96)) if s1 else 95
                        w1 = 128 - (w0 + w1r) if (s0 and s1) else (128 - w1r if s1 else 0)
                        if not s1:
                            w1 = 0
                        raw = struct.pack("<IIhh", s0, s1, w0, w1)
                        C.memmove(prm.raw, raw, len(raw))
                        body = raw + bytes(32 - len(raw))
                    recs += struct.pack("<HHHHBBBB", x, y, w, h, 0, edges, kind, 0) + body
                    n += 1
                    # reference call: left = 4 pre-LR samples left of the unit per row, lpf = 2 rows above and 2 below
                    left = np.ascontiguousarray(pic[y:y + h, x - 4:x])
                    lpf = np.zeros((8, W), dt)
                    lpf[0:2] = pic[y - 2:y]
                    lpf[6:8] = pic[y + h:y + h + 2]
                    fn = L.ref_lr_wiener if kind <= 1 else L.ref_lr_sgr
                    fn(bpc, kind if kind <= 1 else kind - 2, ptr(want, y, x), want.strides[0], ptr(left), ptr(lpf, 0, x), w, h,
                       C.byref(prm), edges)
            fd = rig.desc(W, H, bpc, 0, cur)
            rbuf = C.create_string_buffer(bytes(recs), len(recs))
            fd.lr_recs, fd.n_lr_recs = C.addressof(rbuf), n
            rig.submit(fd)
            got = rig.pic_download(cur, W, H, 0, bpc)[0]
            rig.E.eav1cu_pic_release(rig.ctx, cur)
            bad = np.argwhere(got != want)
            if len(bad):
                raise AssertionError("loop restoration trial %d differs in %d samples, first at (y, x) = %s" % (trial, len(bad), bad[0]))
    finally:
        rig.close()


# ------------------------------------------------------------------------------------------- K5 CDEF

@pytest.mark.parametrize("so", engines())
@pytest.mark.parametrize("bpc,layout", [(8, 1), (10, 1), (12, 3), (8, 2), (10, 0)])
def test_cdef(so, bpc, layout):
    """cdef.dir + cdef.fb[8x8, 4x8, 4x4] (cdef_tmpl.c:56-315) as dav1d_cdef_brow drives them
    (cdef_apply_tmpl.c:97-308: strengths from cdef_idx, variance-adjusted primary strength, chroma direction remap
    for 4:2:2, skipped 8x8s, frame-edge flags). The engine drives CDEF itself from the Av1Filter array; the reference
    side walks the same decisions in Python and calls the reference operators block by block."""
    skip_if_missing(so)
    rng = np.random.default_rng(900 + bpc + layout)
    rig = Rig(so)
    L = R()
    L.ref_cdef_dir.argtypes = [C.c_int, C.c_void_p, C.c_ssize_t, C.POINTER(C.c_uint)]
    L.ref_cdef_dir.restype = C.c_int
    L.ref_cdef_fb.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_ssize_t, C.c_void_p, C.c_void_p, C.c_void_p] + [C.c_int] * 5
    L.ref_cdef_fb.restype = None
    W, H = 328, 200                    # not a multiple of 64: partial superblocks at the right / bottom edge
    bw, bh = W // 4, H // 4
    sb128w, sb128h = (W + 127) // 128, (H + 127) // 128
    bd8 = bpc - 8
    ss_hor, ss_ver = int(layout in (1, 2)), int(layout == 1)
    uv_idx = 3 - layout
    uv_dirs = [[0, 1, 2, 3, 4, 5, 6, 7], [7, 0, 2, 4, 5, 6, 6, 6]][int(layout == 2)]
    dt = np.uint16 if bpc > 8 else np.uint8

    def adjust(strength, var):
        if not var:
            return 0
        i = min(int(var >> 6).bit_length() - 1, 12) if var >> 6 else 0
        return (strength * (4 + i) + 8) >> 4
    try:
        for trial in range(3):
            planes = []
            for pw, ph in ops.plane_dims(W, H, layout):     # edges + texture so that directions and variances vary
                yy, xx = np.mgrid[0:ph, 0:pw]
                a = ((xx * int(rng.integers(1, 9)) + yy * int(rng.integers(1, 9))) // 8 % 2) * (1 << (bpc - 2))
                a = a + rng.integers(0, 1 << (bpc - 1), size=(ph, pw))
                planes.append(np.clip(a, 0, (1 << bpc) - 1).astype(dt))
            # picture storage with a margin so that the pointers two samples outside stay valid for the reference
            pad = [np.zeros((p.shape[0] + 8, p.shape[1] + 16), dt) for p in planes]
            for q, p in zip(pad, planes):
                q[4:-4, 8:-8] = p
            orig = [q.copy() for q in pad]
            filt = np.zeros((sb128h * sb128w, 1348), np.uint8)
            y_str = rng.integers(0, 64, size=8)
            uv_str = rng.integers(0, 64, size=8)
            y_str[0], uv_str[0] = 0, 0
            uv_str[1] = 0
            y_str[2] = 0
            damping = int(rng.integers(3, 7))
            for sb in range(sb128h * sb128w):
                idx = rng.integers(-1, 8, size=4).astype(np.int8)
                filt[sb, 1280:1284] = idx.view(np.uint8)
                ns = rng.integers(0, 1 << 16, size=(16, 2)).astype(np.uint16)
                filt[sb, 1284:1348] = ns.view(np.uint8).reshape(-1)
            # reference walk
            for by in range(0, bh, 2):
                for bx in range(0, bw, 2):
                    sbi = (by >> 5) * sb128w + (bx >> 5)
                    ci = int(np.int8(filt[sbi, 1280 + (((by & 16) >> 3) + ((bx >> 4) & 1))]))
                    if ci == -1 or (not y_str[ci] and not uv_str[ci]):
                        continue
                    ns = filt[sbi, 1284:1348].view(np.uint16).reshape(16, 2)[(by & 30) >> 1]
                    mask = int(ns[1]) << 16 | int(ns[0])
                    if not mask & (3 << (bx & 30)):
                        continue
                    edges = (1 if bx > 0 else 0) | (2 if bx + 2 < bw else 0) | (4 if by > 0 else 0) | (8 if by + 2 < bh else 0)
                    y_lvl, uv_lvl = int(y_str[ci]), int(uv_str[ci])
                    y_pri = (y_lvl >> 2) << bd8
                    y_sec = y_lvl & 3
                    y_sec = (y_sec + (y_sec == 3)) << bd8
                    uv_pri = (uv_lvl >> 2) << bd8
                    uv_sec = uv_lvl & 3
                    uv_sec = (uv_sec + (uv_sec == 3)) << bd8
                    d, var = 0, C.c_uint(0)
                    x0, y0 = bx * 4, by * 4
                    if y_pri or uv_pri:
                        d = L.ref_cdef_dir(bpc, ptr(orig[0], y0 + 4, x0 + 8), orig[0].strides[0], C.byref(var))

                    def fb(pl, idx, px, py, bw_, bh_, pri, sec, dr, damp):
                        left = np.ascontiguousarray(orig[pl][py + 4:py + 4 + bh_, px + 6:px + 8])
                        lpad = np.zeros((8, 2), dt)
                        lpad[:bh_] = left
                        L.ref_cdef_fb(bpc, idx, ptr(pad[pl], py + 4, px + 8), pad[pl].strides[0], ptr(lpad),
                                      ptr(orig[pl], py + 2, px + 8), ptr(orig[pl], py + 4 + bh_, px + 8), pri, sec, dr, damp, edges)
                    if y_pri:
                        adj = adjust(y_pri, var.value)
                        if adj or y_sec:
                            fb(0, 0, x0, y0, 8, 8, adj, y_sec, d, damping + bd8)
                    elif y_sec:
                        fb(0, 0, x0, y0, 8, 8, 0, y_sec, 0, damping + bd8)
                    if uv_lvl and layout:
                        uvdir = uv_dirs[d] if uv_pri else 0
                        for pl in (1, 2):
                            fb(pl, uv_idx, x0 >> ss_hor, y0 >> ss_ver, 8 >> ss_hor, 8 >> ss_ver, uv_pri, uv_sec, uvdir, damping + bd8 - 1)
            cur = rig.pic_alloc(W, H, layout, bpc)
            rig.pic_upload(cur, planes)
            fd = rig.desc(W, H, bpc, layout, cur)
            fd.cdef_enable, fd.cdef_damping = 1, damping
            for i in range(8):
                fd.cdef_y[i], fd.cdef_uv[i] = int(y_str[i]), int(uv_str[i])
            fd.lf_mask = filt.ctypes.data
            fd.sb128w, fd.sb128h = sb128w, sb128h
            rig.submit(fd)
            got = rig.pic_download(cur, W, H, layout, bpc)
            rig.E.eav1cu_pic_release(rig.ctx, cur)
            for pl in range(len(planes)):
                want = pad[pl][4:-4, 8:-8]
                bad = np.argwhere(got[pl] != want)
                assert not len(bad), "cdef trial %d plane %d: %d samples differ, first at (y, x) = %s: got %s want %s" % (
                    trial, pl, len(bad), bad[0], np.asarray(got[pl])[bad[:8, 0], bad[:8, 1]].tolist(), np.asarray(want)[bad[:8, 0], bad[:8, 1]].tolist())
    finally:
        rig.close()


# ------------------------------------------------------------------------------------------- K3 intra prediction

NEEDS = {0: 3, 1: 2, 2: 1, 3: 1, 4: 2, 5: 0, 6: 2 | 8 | 4, 7: 7, 8: 1 | 16 | 4, 9: 3, 10: 3, 11: 3, 12: 7, 13: 7}   # host/rec_tmpl.c needs_tab


@pytest.mark.parametrize("so", engines())
@pytest.mark.parametrize("bpc", [8, 10, 12])
def test_intra_pred(so, bpc):
    """dav1d_prepare_intra_edges + ipred.intra_pred[14] (ipred_prepare_tmpl.c:75-204, ipred_tmpl.c:86-655): all 13
    bitstream modes + filter intra, every angle delta, edge filter / upsampling and smooth-neighbour flags, every
    block size, blocks with and without left / top neighbours. The reference builds its edge array with
    dav1d_prepare_intra_edges; the engine gathers the edge on the device from the same picture."""
    skip_if_missing(so)
    rng = np.random.default_rng(1000 + bpc)
    rig = Rig(so)
    L = R()
    L.ref_prepare_intra_edges.argtypes = [C.c_int] * 8 + [C.c_void_p, C.c_ssize_t, C.c_void_p, C.c_int, C.POINTER(C.c_int), C.c_int, C.c_int,
                                                          C.c_int, C.c_void_p]
    L.ref_prepare_intra_edges.restype = C.c_int
    L.ref_intra_pred.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_ssize_t, C.c_void_p] + [C.c_int] * 5
    L.ref_intra_pred.restype = None
    sizes = [(4, 4), (8, 8), (16, 16), (32, 32), (64, 64), (4, 8), (8, 4), (8, 16), (16, 8), (16, 32), (32, 16), (32, 64), (64, 32),
             (4, 16), (16, 4), (8, 32), (32, 8), (16, 64), (64, 16)]
    cases = []
    for (w, h) in sizes:
        for mode in range(13):
            deltas = range(-3, 4) if 1 <= mode <= 8 else [0]
            for delta in deltas:
                for avail in ((1, 1), (1, 0), (0, 1)) if delta in (0, -3, 3) else ((1, 1),):
                    cases.append((w, h, mode, delta, avail, int(rng.integers(0, 4))))
        if w <= 32 and h <= 32:
            for f in range(5):
                cases.append((w, h, 13, f, (1, 1), 0))
    cell, cols = 192, 24
    rows = (len(cases) + cols - 1) // cols
    W, H = cols * cell, rows * cell
    dt = np.uint16 if bpc > 8 else np.uint8
    try:
        pic = rand_pic(rng, H, W, bpc)
        # smoother neighbourhoods in half of the cells so that the directional filters / upsamplers see gradients too
        sm = (np.add.outer(np.arange(H) * 5, np.arange(W) * 3) >> 2) % (1 << bpc)
        half = (np.arange(W) // cell % 2 == 0)[None, :]
        pic = np.where(half, sm, pic).astype(dt)
        want = pic.copy()
        cur = rig.pic_alloc(W, H, 0, bpc)
        rig.pic_upload(cur, [pic])
        cmds = []
        w4f, h4f = W // 4, H // 4
        for i, (w, h, mode, delta, (have_left, have_top), fl) in enumerate(cases):
            x0, y0 = (i % cols) * cell + 64, (i // cols) * cell + 64
            tw, th = w // 4, h // 4
            angle = C.c_int(delta)
            edge_buf = np.zeros(257 + 64, dt)
            edge = C.c_void_p(edge_buf.ctypes.data + 128 * edge_buf.itemsize)
            filter_edge = fl & 1
            m = L.ref_prepare_intra_edges(bpc, x0 // 4, have_left, y0 // 4, have_top, w4f, h4f, 0x3f, ptr(want, y0, x0), want.strides[0],
                                          None, mode, C.byref(angle), tw, th, filter_edge, edge)
            flags = 0
            if 6 <= m <= 8:
                flags = (1024 if filter_edge else 0) | (512 if fl & 2 else 0)
            ang = angle.value | flags if mode != 13 else delta
            needs = NEEDS[m]
            have = have_left | have_top << 1
            n_left = n_bl = n_top = n_tr = 0
            if needs & 1:
                if have_left:
                    n_left = min(th * 4, (h4f - y0 // 4) * 4)
                if needs & 16 and have_left and y0 // 4 + th < h4f:
                    have |= 8
                    n_bl = min(th * 4, (h4f - y0 // 4 - th) * 4)
            if needs & 2:
                if have_top:
                    n_top = min(tw * 4, (w4f - x0 // 4) * 4)
                if needs & 8 and have_top and x0 // 4 + tw < w4f:
                    have |= 16
                    n_tr = min(tw * 4, (w4f - x0 // 4 - tw) * 4)
            if needs & 4 and m == 7 and tw + th >= 6 and filter_edge:
                have |= 4
            body = struct.pack("<hhhBBBBBBBBhhhhbBBB", x0, y0, ang, m, have, min(n_left, 255), min(n_bl, 255), min(n_top, 255), min(n_tr, 255),
                               tw, th, 16383, 16383, 0, 0, 0, 0, 0, 0)
            cmds.append(cmd(OP_IPRED, 0, w, h, x0, y0, body=body))
            L.ref_intra_pred(bpc, m, ptr(want, y0, x0), want.strides[0], edge, w, h, ang, 16383, 16383)
        rig.submit(rig.desc(W, H, bpc, 0, cur), b"".join(cmds))
        got = rig.pic_download(cur, W, H, 0, bpc)[0]
        bad = np.argwhere(got != want)
        if len(bad):
            i = (bad[0][0] // cell) * cols + bad[0][1] // cell
            raise AssertionError("intra_pred differs: (w, h, mode, angle delta / filter, (left, top), flags) = %s; %d samples in all"
                                 % (cases[i], len(bad)))
    finally:
        rig.close()


@pytest.mark.parametrize("so", engines())
@pytest.mark.parametrize("bpc,layout", [(8, 1), (10, 2), (12, 3), (10, 1)])
def test_cfl_and_palette(so, bpc, layout):
    """ipred.cfl_ac[420 / 422 / 444] + ipred.cfl_pred[DC, LEFT_DC, TOP_DC, DC_128] (ipred_tmpl.c:103-218, 657-715) and
    ipred.pal_pred (:717-730): chroma from the co-located luma block with random alpha and padding, palette blocks
    with random palettes and packed indices."""
    skip_if_missing(so)
    rng = np.random.default_rng(1100 + bpc + layout)
    rig = Rig(so)
    L = R()
    L.ref_cfl_ac.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_ssize_t] + [C.c_int] * 4
    L.ref_cfl_pred.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_ssize_t, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int]
    L.ref_pal_pred.argtypes = [C.c_int, C.c_void_p, C.c_ssize_t, C.c_void_p, C.c_void_p, C.c_int, C.c_int]
    L.ref_prepare_intra_edges.argtypes = [C.c_int] * 8 + [C.c_void_p, C.c_ssize_t, C.c_void_p, C.c_int, C.POINTER(C.c_int), C.c_int, C.c_int,
                                                          C.c_int, C.c_void_p]
    L.ref_prepare_intra_edges.restype = C.c_int
    L.ref_cfl_ac.restype = L.ref_cfl_pred.restype = L.ref_pal_pred.restype = None
    ss_hor, ss_ver = int(layout != 3), int(layout == 1)
    dt = np.uint16 if bpc > 8 else np.uint8
    csizes = [(4, 4), (8, 8), (16, 16), (32, 32), (4, 8), (8, 4), (8, 16), (16, 8), (16, 32), (32, 16), (4, 16), (16, 4), (8, 32), (32, 8)]
    cases = []
    for (cw, ch) in csizes:
        if (cw << ss_hor) > 64 or (ch << ss_ver) > 64:
            continue
        for avail in ((1, 1), (1, 0), (0, 1), (0, 0)):
            cases.append(("cfl", cw, ch, avail, int(rng.integers(-16, 17))))
    for (w, h) in [(8, 8), (16, 16), (32, 32), (64, 64), (8, 16), (64, 32), (16, 64)]:
        cases.append(("pal", w, h, (1, 1), 0))
    cell, cols = 192, 8
    rows = (len(cases) + cols - 1) // cols
    W, H = cols * cell, rows * cell
    try:
        planes = [rng.integers(0, 1 << bpc, size=(ph, pw)).astype(dt) for pw, ph in ops.plane_dims(W, H, layout)]
        want = [p.copy() for p in planes]
        cur = rig.pic_alloc(W, H, layout, bpc)
        rig.pic_upload(cur, planes)
        cmds, payload = [], bytearray()
        for i, (kind, w, h, (have_left, have_top), alpha) in enumerate(cases):
            lx0, ly0 = (i % cols) * cell + 64, (i // cols) * cell + 64
            if kind == "pal":
                pal = rng.integers(0, 1 << bpc, size=8).astype(dt)
                idx = (rng.integers(0, 8, size=(h, w // 2)) | rng.integers(0, 8, size=(h, w // 2)) << 4).astype(np.uint8)
                po = len(payload); payload += pal.tobytes(); payload += bytes(-len(payload) % 16)
                io = len(payload); payload += idx.tobytes(); payload += bytes(-len(payload) % 16)
                cmds.append(cmd(OP_PAL, 0, w, h, lx0, ly0, body=struct.pack("<II", po, io)))
                L.ref_pal_pred(bpc, ptr(want[0], ly0, lx0), want[0].strides[0], ptr(pal), ptr(idx), w, h)
                continue
            pl = 1 + (i & 1)
            cx0, cy0 = lx0 >> ss_hor, ly0 >> ss_ver
            w_pad, h_pad = int(rng.integers(0, max(w // 4, 1))) if w > 4 else 0, int(rng.integers(0, max(h // 4, 1))) if h > 4 else 0
            ac = np.zeros(w * h, np.int16)
            L.ref_cfl_ac(bpc, {1: 0, 2: 1, 3: 2}[layout], ptr(ac), ptr(want[0], ly0, lx0), want[0].strides[0], w_pad, h_pad, w, h)
            angle = C.c_int(0)
            edge_buf = np.zeros(257 + 64, dt)
            edge = C.c_void_p(edge_buf.ctypes.data + 128 * edge_buf.itemsize)
            cw4, ch4 = planes[pl].shape[1] // 4, planes[pl].shape[0] // 4
            m = L.ref_prepare_intra_edges(bpc, cx0 // 4, have_left, cy0 // 4, have_top, cw4, ch4, 0, ptr(want[pl], cy0, cx0), want[pl].strides[0],
                                          None, 0, C.byref(angle), w // 4, h // 4, 0, edge)
            needs = NEEDS[m]
            n_left = min(h, (ch4 - cy0 // 4) * 4) if (needs & 1 and have_left) else 0
            n_top = min(w, (cw4 - cx0 // 4) * 4) if (needs & 2 and have_top) else 0
            body = struct.pack("<hhhBBBBBBBBhhhhbBBB", cx0, cy0, 0, m, have_left | have_top << 1, n_left, 0, n_top, 0, w // 4, h // 4,
                               16383, 16383, lx0, ly0, alpha, w_pad, h_pad, 0)
            cmds.append(cmd(OP_CFL, pl, w, h, cx0, cy0, body=body))
            L.ref_cfl_pred(bpc, m, ptr(want[pl], cy0, cx0), want[pl].strides[0], edge, w, h, ptr(ac), alpha)
        rig.submit(rig.desc(W, H, bpc, layout, cur), b"".join(cmds), bytes(payload))
        got = rig.pic_download(cur, W, H, layout, bpc)
        for pl in range(3):
            bad = np.argwhere(got[pl] != want[pl])
            if len(bad):
                sx, sy = (ss_hor, ss_ver) if pl else (0, 0)
                i = ((bad[0][0] << sy) // cell) * cols + (bad[0][1] << sx) // cell
                raise AssertionError("plane %d differs at case %s (%d samples in all)" % (pl, cases[i], len(bad)))
    finally:
        rig.close()
