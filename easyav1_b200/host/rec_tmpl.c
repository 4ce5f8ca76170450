/*
 * rec_tmpl.c — recording implementations of dav1d's per-bitdepth surfaces, compiled twice
 * (-DBITDEPTH=8 / 16) exactly like the reference's *_tmpl.c files:
 *   - the seven Dav1dDSPContext tables (B3): eav1rec_*_dsp_init_{8,16}bpc are
 *     bound in place of dav1d_*_dsp_init_{8,16}bpc (decode.c:3383-3416) by host/Makefile;
 *   - dav1d_prepare_intra_edges (ipred_prepare_tmpl.c:75): mode/angle resolution stays here,
 *     the pixel gathering moves to the device;
 *   - wrappers around recon_b_intra / recon_b_inter (block boundaries) and filter_sbrow
 *     (per-frame submit), dav1d_backup_ipred_edge / dav1d_copy_lpf become no-ops because the
 *     pre-filter frame stays resident on the device.
 */
#include <stdio.h>
#include <string.h>

#include "rec.h"
#include "common/bitdepth.h"
#include "common/intops.h"
#include "src/recon.h"
#include "src/tables.h"
#include "src/scan.h"
#include "src/wedge.h"
#include "src/ipred_prepare.h"

#define REC() Eav1Walk *const r = &eav1_walk
/* a construct the recorder cannot express: only the frame being walked fails (its picture and whatever references it
 * come back as DAV1D_ERR(EIO)); the decoder recovers at the next key frame like dav1d does after a damaged frame */
#define UNSUPPORTED(r) eav1rec_unsupported(r)

#if BITDEPTH == 8
#define BDMAX_ARG
#else
#define BDMAX_ARG , const int bitdepth_max
#endif

/* the reference's walk, renamed at compile time */
void bitfn(eav1ref_recon_b_intra)(Dav1dTaskContext *t, enum BlockSize bs, enum EdgeFlags flags, const Av1Block *b);
int  bitfn(eav1ref_recon_b_inter)(Dav1dTaskContext *t, enum BlockSize bs, const Av1Block *b);
void bitfn(eav1ref_filter_sbrow_deblock_cols)(Dav1dFrameContext *f, int sby);
void bitfn(eav1ref_filter_sbrow_deblock_rows)(Dav1dFrameContext *f, int sby);
void bitfn(eav1ref_filter_sbrow_lr)(Dav1dFrameContext *f, int sby);

/* ---------------------------------------------------------------- helpers */

static void set_dst(Eav1Walk *r, Eav1Cmd *c, const void *dst, int w, int h, int reg, size_t elem)
{
    int x, y;
    const int pl = eav1rec_locate_cur(r, dst, &x, &y);
    c->w = (uint8_t)w; c->h = (uint8_t)h;
    if (pl >= 0) {
        c->plane = (uint8_t)pl; c->x = (int16_t)x; c->y = (int16_t)y;
    } else {
        c->flags |= EAV1_F_DST_ARENA;
        c->dst_off = eav1rec_region_new(r, reg, (size_t)imax(w * h, r->blk_w * r->blk_h) * elem);
        c->dst_stride = (uint16_t)w;
    }
}

/* ---------------------------------------------------------------- itx */

static void rec_itx(const int tx, const int txtp, pixel *dst, const ptrdiff_t stride, coef *cf, const int eob)
{
    REC();
    (void)stride;
    const TxfmInfo *const td = &dav1d_txfm_dimensions[tx];
    const int w = td->w * 4, h = td->h * 4;
    const int n = imin(w, 32) * imin(h, 32);
    Eav1Cmd *c = eav1rec_push(r);
    c->op = EAV1_OP_ITX;
    set_dst(r, c, dst, w, h, REG_II, sizeof(pixel));
    c->u.itx.tx = (uint8_t)tx; c->u.itx.txtp = (uint8_t)txtp; c->u.itx.eob = (int16_t)eob;
    /* Only the rows that can hold a non-zero coefficient travel: eob bounds them exactly as the
     * reference's own transform derives last_nonzero_col (itx_tmpl.c:58-105): a dc-only block is one coefficient, a
     * 1-D class block follows eob directly, everything else the scan-order table. The block is column-major, so the
     * first `rows` entries of each of its min(w,32) columns are copied and, as the contract of itxfm_add demands
     * (itx_tmpl.c:108: the callee clears the coefficients), cleared; the rest of the block is zero already. */
    const int sw = imin(w, 32), sh = imin(h, 32);
    int rows;
    if (txtp == WHT_WHT) rows = sh;
    else if (eob < (txtp == DCT_DCT)) rows = 1;
    else if (txtp == H_DCT || txtp == H_ADST || txtp == H_FLIPADST) rows = imin(sh - 1, eob) + 1;
    else if (txtp == V_DCT || txtp == V_ADST || txtp == V_FLIPADST) rows = (eob >> (td->lw + 2)) + 1;
    else rows = dav1d_last_nonzero_col_from_eob[tx][eob] + 1;
    (void)n;
    if (txtp != WHT_WHT && eob < (txtp == DCT_DCT)) {
        c->u.itx.rows = 1;
        c->u.itx.cf_off = eav1rec_payload(r, cf, sizeof(coef), 4);
        cf[0] = 0;
    } else if (rows >= sh) {
        c->u.itx.rows = (uint8_t)sh;
        c->u.itx.cf_off = eav1rec_payload(r, cf, (size_t)sw * sh * sizeof(coef), 16);
        memset(cf, 0, (size_t)sw * sh * sizeof(coef));
    } else {
        c->u.itx.rows = (uint8_t)rows;
        coef *dstp;
        c->u.itx.cf_off = eav1rec_payload_reserve(r, (size_t)sw * rows * sizeof(coef), 16, (void **)&dstp);
        if (dstp)
            for (int x = 0; x < sw; x++) {
                memcpy(dstp + x * rows, cf + x * sh, (size_t)rows * sizeof(coef));
                memset(cf + x * sh, 0, (size_t)rows * sizeof(coef));
            }
    }
}

#define ITX_FN(tx, tp) \
static void rec_itx_##tx##_##tp(pixel *dst, ptrdiff_t stride, coef *cf, int eob HIGHBD_DECL_SUFFIX) \
{ rec_itx(tx, tp, dst, stride, cf, eob); }
#define ITX_ROW(M, tx) M(tx,0) M(tx,1) M(tx,2) M(tx,3) M(tx,4) M(tx,5) M(tx,6) M(tx,7) M(tx,8) \
    M(tx,9) M(tx,10) M(tx,11) M(tx,12) M(tx,13) M(tx,14) M(tx,15) M(tx,16)
#define ITX_ALL(M) ITX_ROW(M,0) ITX_ROW(M,1) ITX_ROW(M,2) ITX_ROW(M,3) ITX_ROW(M,4) ITX_ROW(M,5) \
    ITX_ROW(M,6) ITX_ROW(M,7) ITX_ROW(M,8) ITX_ROW(M,9) ITX_ROW(M,10) ITX_ROW(M,11) ITX_ROW(M,12) \
    ITX_ROW(M,13) ITX_ROW(M,14) ITX_ROW(M,15) ITX_ROW(M,16) ITX_ROW(M,17) ITX_ROW(M,18)
ITX_ALL(ITX_FN)

void bitfn(eav1rec_itx_dsp_init)(Dav1dInvTxfmDSPContext *const c, int bpc)
{
    (void)bpc;
    dav1d_init_last_nonzero_col_from_eob_tables();
#define ITX_SET(tx, tp) c->itxfm_add[tx][tp] = rec_itx_##tx##_##tp;
    ITX_ALL(ITX_SET)
}

/* ---------------------------------------------------------------- mc */

static const uint8_t filt_h_of[N_2D_FILTERS] = { 0, 0, 0, 2, 2, 2, 1, 1, 1, 3 };
static const uint8_t filt_v_of[N_2D_FILTERS] = { 0, 1, 2, 0, 1, 2, 0, 1, 2, 3 };

/* Resolve a source pointer handed to an mc function: either inside a reference plane or
 * inside the emu_edge scratch patch announced by the preceding emu_edge call. */
static int resolve_src(Eav1Walk *r, const pixel *src, int *ref, int *plane, int *sx, int *sy, int *iw, int *ih)
{
    if (r->emu.valid) {
        const ptrdiff_t off = (const uint8_t *)src - (const uint8_t *)r->emu.dst;
        if (off >= 0 && off < r->emu.dst_stride * r->emu.bh) {
            *ref = r->emu.ref; *plane = r->emu.plane;
            *sy = r->emu.y + (int)(off / r->emu.dst_stride);
            *sx = r->emu.x + (int)((off % r->emu.dst_stride) / (ptrdiff_t)sizeof(pixel));
            *iw = r->emu.iw; *ih = r->emu.ih;
            return 0;
        }
    }
    const int i = eav1rec_locate_ref(r, src, plane, sx, sy);
    if (i < 0) { UNSUPPORTED(r); return -1; }
    *ref = i;
    *iw = *ih = 1 << 24; /* in-bounds by construction: clamping is the identity */
    return 0;
}

/* produce != 0: the caller writes the prep block; a write at the start of the scratch buffer
 * opens a fresh region (warp8x8t fills one region tile by tile at inner offsets) */
static uint32_t tmp_region(Eav1Walk *r, const int16_t *tmp, int produce)
{
    const int16_t *base = r->t->scratch.compinter[0];
    const ptrdiff_t off = tmp - base;
    const int idx = off >= 128 * 128;
    const ptrdiff_t inner = off - idx * 128 * 128;
    const int reg = idx ? REG_TMP1 : REG_TMP0;
    if (produce && inner == 0) return eav1rec_region_new(r, reg, (size_t)r->blk_w * r->blk_h * 2);
    return eav1rec_region(r, reg, 2) + (uint32_t)(inner * 2);
}

static void rec_mc_any(const int filter, pixel *dst8, int16_t *dst16, const pixel *src,
                       const int w, const int h, const int mx, const int my, const int dx, const int dy)
{
    REC();
    Eav1Cmd *c = eav1rec_push(r);
    c->op = EAV1_OP_MC;
    int ref, plane, sx, sy, iw, ih;
    if (resolve_src(r, src, &ref, &plane, &sx, &sy, &iw, &ih)) { c->op = EAV1_OP_NOP; return; }
    if (dst8) {
        set_dst(r, c, dst8, w, h, REG_LAP, sizeof(pixel));
    } else {
        c->w = (uint8_t)w; c->h = (uint8_t)h;
        c->flags |= EAV1_F_DST_ARENA | EAV1_F_PREP;
        c->dst_off = tmp_region(r, dst16, 1);
        c->dst_stride = (uint16_t)w;
    }
    c->plane = (c->flags & EAV1_F_DST_ARENA) ? (uint8_t)plane : c->plane;
    c->u.mc.sx = sx; c->u.mc.sy = sy; c->u.mc.iw = iw; c->u.mc.ih = ih;
    c->u.mc.mx = (uint16_t)mx; c->u.mc.my = (uint16_t)my;
    c->u.mc.dx = (uint16_t)dx; c->u.mc.dy = (uint16_t)dy;
    c->u.mc.ref = (uint8_t)ref;
    c->u.mc.filt_h = filt_h_of[filter]; c->u.mc.filt_v = filt_v_of[filter];
}

#define MC_FNS(F) \
static void rec_put_##F(pixel *dst, ptrdiff_t ds, const pixel *src, ptrdiff_t ss, int w, int h, int mx, int my HIGHBD_DECL_SUFFIX) \
{ (void)ds; (void)ss; rec_mc_any(F, dst, NULL, src, w, h, mx, my, 0, 0); } \
static void rec_prep_##F(int16_t *tmp, const pixel *src, ptrdiff_t ss, int w, int h, int mx, int my HIGHBD_DECL_SUFFIX) \
{ (void)ss; rec_mc_any(F, NULL, tmp, src, w, h, mx, my, 0, 0); } \
static void rec_put_scaled_##F(pixel *dst, ptrdiff_t ds, const pixel *src, ptrdiff_t ss, int w, int h, int mx, int my, int dx, int dy HIGHBD_DECL_SUFFIX) \
{ (void)ds; (void)ss; rec_mc_any(F, dst, NULL, src, w, h, mx, my, dx, dy); } \
static void rec_prep_scaled_##F(int16_t *tmp, const pixel *src, ptrdiff_t ss, int w, int h, int mx, int my, int dx, int dy HIGHBD_DECL_SUFFIX) \
{ (void)ss; rec_mc_any(F, NULL, tmp, src, w, h, mx, my, dx, dy); }
MC_FNS(0) MC_FNS(1) MC_FNS(2) MC_FNS(3) MC_FNS(4) MC_FNS(5) MC_FNS(6) MC_FNS(7) MC_FNS(8) MC_FNS(9)

static void rec_emu_edge(intptr_t bw, intptr_t bh, intptr_t iw, intptr_t ih, intptr_t x, intptr_t y,
                         pixel *dst, ptrdiff_t dst_stride, const pixel *src, ptrdiff_t src_stride)
{
    REC();
    (void)src_stride;
    int pl, sx, sy;
    const int ref = eav1rec_locate_ref(r, src, &pl, &sx, &sy);
    if (ref < 0) { UNSUPPORTED(r); r->emu.valid = 0; return; }
    r->emu.dst = dst; r->emu.dst_stride = dst_stride;
    r->emu.bw = (int)bw; r->emu.bh = (int)bh;
    r->emu.x = (int)x; r->emu.y = (int)y; r->emu.iw = (int)iw; r->emu.ih = (int)ih;
    r->emu.ref = ref; r->emu.plane = pl; r->emu.valid = 1;
}

static void rec_warp_any(pixel *dst8, int16_t *dst16, const ptrdiff_t dstride, const pixel *src,
                         const int16_t *abcd, const int mx, const int my)
{
    REC();
    Eav1Cmd *c = eav1rec_push(r);
    c->op = EAV1_OP_WARP;
    int ref, plane, sx, sy, iw, ih;
    if (resolve_src(r, src, &ref, &plane, &sx, &sy, &iw, &ih)) { c->op = EAV1_OP_NOP; return; }
    if (dst8) {
        set_dst(r, c, dst8, 8, 8, REG_LAP, sizeof(pixel));
    } else {
        c->w = c->h = 8;
        c->flags |= EAV1_F_DST_ARENA | EAV1_F_PREP;
        c->dst_off = tmp_region(r, dst16, 1);
        c->dst_stride = (uint16_t)dstride;
        c->plane = (uint8_t)plane;
    }
    c->u.warp.sx = sx; c->u.warp.sy = sy; c->u.warp.iw = iw; c->u.warp.ih = ih;
    c->u.warp.mx = mx; c->u.warp.my = my;
    memcpy(c->u.warp.abcd, abcd, 8);
    c->u.warp.ref = (uint8_t)ref;
}
static void rec_warp8x8(pixel *dst, ptrdiff_t ds, const pixel *src, ptrdiff_t ss, const int16_t *abcd, int mx, int my HIGHBD_DECL_SUFFIX)
{ (void)ss; rec_warp_any(dst, NULL, ds, src, abcd, mx, my); }
static void rec_warp8x8t(int16_t *tmp, const ptrdiff_t ts, const pixel *src, ptrdiff_t ss, const int16_t *abcd, int mx, int my HIGHBD_DECL_SUFFIX)
{ (void)ss; rec_warp_any(NULL, tmp, ts, src, abcd, mx, my); }

static Eav1Cmd *rec_comp(int op, pixel *dst, const int16_t *t1, const int16_t *t2, int w, int h)
{
    REC();
    Eav1Cmd *c = eav1rec_push(r);
    c->op = (uint8_t)op;
    set_dst(r, c, dst, w, h, REG_LAP, sizeof(pixel));
    c->u.comp.t1 = tmp_region(r, t1, 0);
    c->u.comp.t2 = tmp_region(r, t2, 0);
    return c;
}
static void rec_avg(pixel *dst, ptrdiff_t ds, const int16_t *t1, const int16_t *t2, int w, int h HIGHBD_DECL_SUFFIX)
{ (void)ds; rec_comp(EAV1_OP_AVG, dst, t1, t2, w, h); }
static void rec_w_avg(pixel *dst, ptrdiff_t ds, const int16_t *t1, const int16_t *t2, int w, int h, int weight HIGHBD_DECL_SUFFIX)
{ (void)ds; rec_comp(EAV1_OP_W_AVG, dst, t1, t2, w, h)->u.comp.weight = (int16_t)weight; }
static void rec_mask(pixel *dst, ptrdiff_t ds, const int16_t *t1, const int16_t *t2, int w, int h, const uint8_t *mask HIGHBD_DECL_SUFFIX)
{
    REC();
    (void)ds;
    Eav1Cmd *c = rec_comp(EAV1_OP_MASK, dst, t1, t2, w, h);
    const ptrdiff_t off = mask - (const uint8_t *)&dav1d_masks;
    if (off >= 0 && off < (ptrdiff_t)sizeof(dav1d_masks)) c->u.comp.mask_off = (uint32_t)off;
    else {
        c->flags |= EAV1_F_MASK_ARENA;
        c->u.comp.mask_off = eav1rec_region(r, REG_SEGMASK, 1) + (uint32_t)(mask - r->t->scratch.seg_mask);
    }
}
static void rec_w_mask_any(int ss_hor, int ss_ver, pixel *dst, const int16_t *t1, const int16_t *t2, int w, int h, uint8_t *mask, int sign)
{
    REC();
    Eav1Cmd *c = rec_comp(EAV1_OP_W_MASK, dst, t1, t2, w, h);
    if (mask == r->t->scratch.seg_mask) c->u.comp.mask_off = eav1rec_region_new(r, REG_SEGMASK, (size_t)r->blk_w * r->blk_h);
    else c->u.comp.mask_off = eav1rec_region(r, REG_SEGMASK, 1) + (uint32_t)(mask - r->t->scratch.seg_mask);
    c->u.comp.sign = (uint8_t)sign; c->u.comp.ss_hor = (uint8_t)ss_hor; c->u.comp.ss_ver = (uint8_t)ss_ver;
}
static void rec_w_mask_444(pixel *dst, ptrdiff_t ds, const int16_t *t1, const int16_t *t2, int w, int h, uint8_t *mask, int sign HIGHBD_DECL_SUFFIX)
{ (void)ds; rec_w_mask_any(0, 0, dst, t1, t2, w, h, mask, sign); }
static void rec_w_mask_422(pixel *dst, ptrdiff_t ds, const int16_t *t1, const int16_t *t2, int w, int h, uint8_t *mask, int sign HIGHBD_DECL_SUFFIX)
{ (void)ds; rec_w_mask_any(1, 0, dst, t1, t2, w, h, mask, sign); }
static void rec_w_mask_420(pixel *dst, ptrdiff_t ds, const int16_t *t1, const int16_t *t2, int w, int h, uint8_t *mask, int sign HIGHBD_DECL_SUFFIX)
{ (void)ds; rec_w_mask_any(1, 1, dst, t1, t2, w, h, mask, sign); }

static void rec_blend(pixel *dst, ptrdiff_t ds, const pixel *tmp, int w, int h, const uint8_t *mask)
{
    REC();
    (void)ds; (void)tmp;
    Eav1Cmd *c = eav1rec_push(r);
    c->op = EAV1_OP_BLEND;
    set_dst(r, c, dst, w, h, REG_LAP, sizeof(pixel));
    c->u.blend.tmp_off = eav1rec_region(r, REG_II, sizeof(pixel));
    c->u.blend.mask_off = (uint32_t)(mask - (const uint8_t *)&dav1d_masks);
}
static void rec_blend_dir(int op, pixel *dst, int w, int h)
{
    REC();
    Eav1Cmd *c = eav1rec_push(r);
    c->op = (uint8_t)op;
    set_dst(r, c, dst, w, h, REG_LAP, sizeof(pixel));
    c->u.blend.tmp_off = eav1rec_region(r, REG_LAP, sizeof(pixel));
}
static void rec_blend_v(pixel *dst, ptrdiff_t ds, const pixel *tmp, int w, int h) { (void)ds; (void)tmp; rec_blend_dir(EAV1_OP_BLEND_V, dst, w, h); }
static void rec_blend_h(pixel *dst, ptrdiff_t ds, const pixel *tmp, int w, int h) { (void)ds; (void)tmp; rec_blend_dir(EAV1_OP_BLEND_H, dst, w, h); }

/* mc.resize has two callers: dav1d_filter_sbrow_resize (recon_tmpl.c:2053), replaced below, and backup_lpf inside
 * dav1d_copy_lpf (lf_apply_tmpl.c:76-88), replaced by a no-op. Both upscales run on the device as whole-frame
 * launches (Eav1FrameDesc.superres), so nothing may arrive here. */
static void rec_resize(pixel *dst, ptrdiff_t ds, const pixel *src, ptrdiff_t ss, int dst_w, int h, int src_w, int dx, int mx HIGHBD_DECL_SUFFIX)
{
    REC();
    (void)dst; (void)ds; (void)src; (void)ss; (void)dst_w; (void)h; (void)src_w; (void)dx; (void)mx;
    UNSUPPORTED(r);
}

void bitfn(eav1rec_mc_dsp_init)(Dav1dMCDSPContext *const c)
{
#define MC_SET(F) c->mc[F] = rec_put_##F; c->mct[F] = rec_prep_##F; \
    c->mc_scaled[F] = rec_put_scaled_##F; c->mct_scaled[F] = rec_prep_scaled_##F;
    MC_SET(0) MC_SET(1) MC_SET(2) MC_SET(3) MC_SET(4) MC_SET(5) MC_SET(6) MC_SET(7) MC_SET(8) MC_SET(9)
    c->avg = rec_avg; c->w_avg = rec_w_avg; c->mask = rec_mask;
    c->w_mask[0] = rec_w_mask_444; c->w_mask[1] = rec_w_mask_422; c->w_mask[2] = rec_w_mask_420;
    c->blend = rec_blend; c->blend_v = rec_blend_v; c->blend_h = rec_blend_h;
    c->warp8x8 = rec_warp8x8; c->warp8x8t = rec_warp8x8t;
    c->emu_edge = rec_emu_edge; c->resize = rec_resize;
}

/* ---------------------------------------------------------------- intra */

/* Replaces dav1d_prepare_intra_edges (ipred_prepare_tmpl.c:75-204): resolves the
 * implementation mode and angle exactly as the reference does, and notes which neighbouring
 * samples exist; the samples themselves are gathered on the device (ipred.cuh build_edge). */
enum IntraPredMode
bitfn(eav1rec_prepare_intra_edges)(const int x, const int have_left, const int y, const int have_top,
                                   const int w, const int h, const enum EdgeFlags edge_flags,
                                   const pixel *const dst, const ptrdiff_t stride,
                                   const pixel *prefilter_toplevel_sb_edge,
                                   enum IntraPredMode mode, int *const angle,
                                   const int tw, const int th, const int filter_edge,
                                   pixel *const topleft_out HIGHBD_DECL_SUFFIX)
{
    REC();
    (void)stride; (void)prefilter_toplevel_sb_edge;
    static const uint8_t conv_dc[2][2] = { { DC_128_PRED, TOP_DC_PRED }, { LEFT_DC_PRED, DC_PRED } };
    static const uint8_t conv_paeth[2][2] = { { DC_128_PRED, VERT_PRED }, { HOR_PRED, PAETH_PRED } };
    static const uint8_t angle_map[8] = { 90, 180, 45, 135, 113, 157, 203, 67 };
    /* bit0 left, bit1 top, bit2 topleft, bit3 topright, bit4 bottomleft */
    static const uint8_t needs_tab[N_IMPL_INTRA_PRED_MODES] = {
        [DC_PRED] = 3, [VERT_PRED] = 2, [HOR_PRED] = 1, [LEFT_DC_PRED] = 1, [TOP_DC_PRED] = 2,
        [DC_128_PRED] = 0, [Z1_PRED] = 2 | 8 | 4, [Z2_PRED] = 7, [Z3_PRED] = 1 | 16 | 4,
        [SMOOTH_PRED] = 3, [SMOOTH_V_PRED] = 3, [SMOOTH_H_PRED] = 3, [PAETH_PRED] = 7, [FILTER_PRED] = 7,
    };
    if (mode >= VERT_PRED && mode <= VERT_LEFT_PRED) {
        *angle = angle_map[mode - VERT_PRED] + 3 * *angle;
        if (*angle <= 90) mode = *angle < 90 && have_top ? Z1_PRED : VERT_PRED;
        else if (*angle < 180) mode = Z2_PRED;
        else mode = *angle > 180 && have_left ? Z3_PRED : HOR_PRED;
    } else if (mode == DC_PRED) mode = conv_dc[have_left][have_top];
    else if (mode == PAETH_PRED) mode = conv_paeth[have_left][have_top];

    const int needs = needs_tab[mode];
    int ex, ey;
    const int pl = eav1rec_locate_cur(r, dst, &ex, &ey);
    r->edge.valid = pl >= 0;
    if (pl < 0) { UNSUPPORTED(r); return mode; }
    r->edge.topleft = topleft_out;
    r->edge.plane = pl; r->edge.ex = ex; r->edge.ey = ey;
    r->edge.mode = mode; r->edge.tw = tw; r->edge.th = th;
    int have = (have_left ? 1 : 0) | (have_top ? 2 : 0);
    r->edge.n_left = r->edge.n_bl = r->edge.n_top = r->edge.n_tr = 0;
    if (needs & 1) {
        const int sz = th << 2;
        if (have_left) r->edge.n_left = imin(sz, (h - y) << 2);
        if (needs & 16) {
            const int have_bl = (!have_left || y + th >= h) ? 0 : (edge_flags & EDGE_I444_LEFT_HAS_BOTTOM);
            if (have_bl) { have |= 8; r->edge.n_bl = imin(sz, (h - y - th) << 2); }
        }
    }
    if (needs & 2) {
        const int sz = tw << 2;
        if (have_top) r->edge.n_top = imin(sz, (w - x) << 2);
        if (needs & 8) {
            const int have_tr = (!have_top || x + tw >= w) ? 0 : (edge_flags & EDGE_I444_TOP_HAS_RIGHT);
            if (have_tr) { have |= 16; r->edge.n_tr = imin(sz, (w - x - tw) << 2); }
        }
    }
    if ((needs & 4) && mode == Z2_PRED && tw + th >= 6 && filter_edge) have |= 4;
    r->edge.have = have;
    return mode;
}

static void fill_edge(Eav1Walk *r, Eav1Cmd *c)
{
    c->u.ipred.ex = (int16_t)r->edge.ex; c->u.ipred.ey = (int16_t)r->edge.ey;
    c->u.ipred.mode = (uint8_t)r->edge.mode; c->u.ipred.have = (uint8_t)r->edge.have;
    c->u.ipred.n_left = (uint8_t)r->edge.n_left; c->u.ipred.n_bl = (uint8_t)r->edge.n_bl;
    c->u.ipred.n_top = (uint8_t)r->edge.n_top; c->u.ipred.n_tr = (uint8_t)r->edge.n_tr;
    c->u.ipred.tw = (uint8_t)r->edge.tw; c->u.ipred.th = (uint8_t)r->edge.th;
}

static void rec_ipred(const int m, pixel *dst, const pixel *topleft, int w, int h, int angle, int max_w, int max_h)
{
    REC();
    Eav1Cmd *c = eav1rec_push(r);
    c->op = EAV1_OP_IPRED;
    if (!r->edge.valid || r->edge.topleft != topleft || r->edge.mode != m) { UNSUPPORTED(r); c->op = EAV1_OP_NOP; return; }
    set_dst(r, c, dst, w, h, REG_II, sizeof(pixel));
    if (c->flags & EAV1_F_DST_ARENA) c->plane = (uint8_t)r->edge.plane;
    fill_edge(r, c);
    c->u.ipred.angle = (int16_t)angle;
    c->u.ipred.max_w = (int16_t)imin(max_w, 16383); c->u.ipred.max_h = (int16_t)imin(max_h, 16383);
}
#define IPRED_FN(M) \
static void rec_ipred_##M(pixel *dst, ptrdiff_t s, const pixel *tl, int w, int h, int a, int mw, int mh HIGHBD_DECL_SUFFIX) \
{ (void)s; rec_ipred(M, dst, tl, w, h, a, mw, mh); }
IPRED_FN(0) IPRED_FN(1) IPRED_FN(2) IPRED_FN(3) IPRED_FN(4) IPRED_FN(5) IPRED_FN(6) IPRED_FN(7)
IPRED_FN(8) IPRED_FN(9) IPRED_FN(10) IPRED_FN(11) IPRED_FN(12) IPRED_FN(13)

static void rec_cfl_ac(int16_t *ac, const pixel *ypx, ptrdiff_t stride, int w_pad, int h_pad, int cw, int ch)
{
    REC();
    (void)ac; (void)stride;
    int lx, ly;
    const int pl = eav1rec_locate_cur(r, ypx, &lx, &ly);
    r->ac.valid = pl == 0;
    if (pl != 0) { UNSUPPORTED(r); return; }
    r->ac.lx = lx; r->ac.ly = ly; r->ac.w_pad = w_pad; r->ac.h_pad = h_pad; r->ac.cw = cw; r->ac.ch = ch;
}
static void rec_cfl_pred(const int m, pixel *dst, const pixel *topleft, int w, int h, int alpha)
{
    REC();
    Eav1Cmd *c = eav1rec_push(r);
    c->op = EAV1_OP_CFL;
    if (!r->ac.valid || !r->edge.valid || r->edge.topleft != topleft || r->edge.mode != m ||
        r->ac.cw != w || r->ac.ch != h) { UNSUPPORTED(r); c->op = EAV1_OP_NOP; return; }
    set_dst(r, c, dst, w, h, REG_II, sizeof(pixel));
    fill_edge(r, c);
    c->u.ipred.lx = (int16_t)r->ac.lx; c->u.ipred.ly = (int16_t)r->ac.ly;
    c->u.ipred.alpha = (int8_t)alpha;
    c->u.ipred.w_pad = (uint8_t)r->ac.w_pad; c->u.ipred.h_pad = (uint8_t)r->ac.h_pad;
}
#define CFL_FN(M) \
static void rec_cfl_pred_##M(pixel *dst, ptrdiff_t s, const pixel *tl, int w, int h, const int16_t *ac, int alpha HIGHBD_DECL_SUFFIX) \
{ (void)s; (void)ac; rec_cfl_pred(M, dst, tl, w, h, alpha); }
CFL_FN(0) CFL_FN(3) CFL_FN(4) CFL_FN(5)

static void rec_pal_pred(pixel *dst, ptrdiff_t stride, const pixel *pal, const uint8_t *idx, int w, int h)
{
    REC();
    (void)stride;
    Eav1Cmd *c = eav1rec_push(r);
    c->op = EAV1_OP_PAL;
    set_dst(r, c, dst, w, h, REG_II, sizeof(pixel));
    c->u.pal.pal_off = eav1rec_payload(r, pal, 8 * sizeof(pixel), 16);
    c->u.pal.idx_off = eav1rec_payload(r, idx, (size_t)(w * h) >> 1, 16);
}

void bitfn(eav1rec_intra_pred_dsp_init)(Dav1dIntraPredDSPContext *const c)
{
#define IPRED_SET(M) c->intra_pred[M] = rec_ipred_##M;
    IPRED_SET(0) IPRED_SET(1) IPRED_SET(2) IPRED_SET(3) IPRED_SET(4) IPRED_SET(5) IPRED_SET(6) IPRED_SET(7)
    IPRED_SET(8) IPRED_SET(9) IPRED_SET(10) IPRED_SET(11) IPRED_SET(12) IPRED_SET(13)
    c->cfl_ac[0] = c->cfl_ac[1] = c->cfl_ac[2] = rec_cfl_ac;
    c->cfl_pred[DC_PRED] = rec_cfl_pred_0; c->cfl_pred[LEFT_DC_PRED] = rec_cfl_pred_3;
    c->cfl_pred[TOP_DC_PRED] = rec_cfl_pred_4; c->cfl_pred[DC_128_PRED] = rec_cfl_pred_5;
    c->pal_pred = rec_pal_pred;
}

/* ---------------------------------------------------------------- loop filter / restoration */

static void rec_lf_any(const int dir, pixel *dst, const uint32_t *mask, const uint8_t (*lvl)[4], const int n_masks)
{
    REC();
    int x, y;
    const int pl = eav1rec_locate_cur(r, dst, &x, &y);
    if (pl < 0) { UNSUPPORTED(r); return; }
    const uint32_t any = mask[0] | mask[1] | (n_masks > 2 ? mask[2] : 0);
    if (!any) return;
    Eav1LfRec v, *lr = &v;
    lr->x = (uint16_t)x; lr->y = (uint16_t)y; lr->plane = (uint8_t)pl; lr->dir = (uint8_t)dir; lr->pad = 0;
    lr->vmask[0] = mask[0]; lr->vmask[1] = mask[1]; lr->vmask[2] = n_masks > 2 ? mask[2] : 0;
    lr->lvl_off = (uint32_t)((const uint8_t *)lvl - (const uint8_t *)r->f->lf.level);
    eav1rec_push_lf(r, &v);
}
#define LF_FN(name, dir, n) \
static void name(pixel *dst, ptrdiff_t stride, const uint32_t *mask, const uint8_t (*lvl)[4], ptrdiff_t lvl_stride, \
                 const Av1FilterLUT *lut, int wh HIGHBD_DECL_SUFFIX) \
{ (void)stride; (void)lvl_stride; (void)lut; (void)wh; rec_lf_any(dir, dst, mask, lvl, n); }
LF_FN(rec_lf_h_y, 0, 3) LF_FN(rec_lf_v_y, 1, 3) LF_FN(rec_lf_h_uv, 0, 2) LF_FN(rec_lf_v_uv, 1, 2)

void bitfn(eav1rec_loop_filter_dsp_init)(Dav1dLoopFilterDSPContext *const c)
{
    c->loop_filter_sb[0][0] = rec_lf_h_y; c->loop_filter_sb[0][1] = rec_lf_v_y;
    c->loop_filter_sb[1][0] = rec_lf_h_uv; c->loop_filter_sb[1][1] = rec_lf_v_uv;
}

static void rec_lr_any(const int kind, pixel *dst, int w, int h, const LooprestorationParams *params, int edges)
{
    REC();
    int x, y;
    /* loop restoration works on the super-resolved picture (lr_apply_tmpl.c: f->sr_cur), which is f->cur itself
     * unless the frame uses super-resolution */
    const int pl = eav1rec_locate_sr(r, dst, &x, &y);
    if (pl < 0) { UNSUPPORTED(r); return; }
    Eav1LrRec v, *lr = &v;
    memset(&v, 0, sizeof(v));
    lr->x = (uint16_t)x; lr->y = (uint16_t)y; lr->w = (uint16_t)w; lr->h = (uint16_t)h;
    lr->plane = (uint8_t)pl; lr->edges = (uint8_t)edges; lr->kind = (uint8_t)kind;
    if (kind <= EAV1_LR_WIENER5) memcpy(lr->p.filter, params->filter, sizeof(lr->p.filter));
    else { lr->p.sgr.s0 = params->sgr.s0; lr->p.sgr.s1 = params->sgr.s1; lr->p.sgr.w0 = params->sgr.w0; lr->p.sgr.w1 = params->sgr.w1; }
    eav1rec_push_lr(r, &v);
}
#define LR_FN(name, kind) \
static void name(pixel *dst, ptrdiff_t stride, const_left_pixel_row left, const pixel *lpf, int w, int h, \
                 const LooprestorationParams *params, enum LrEdgeFlags edges HIGHBD_DECL_SUFFIX) \
{ (void)stride; (void)left; (void)lpf; rec_lr_any(kind, dst, w, h, params, edges); }
LR_FN(rec_wiener7, EAV1_LR_WIENER7) LR_FN(rec_wiener5, EAV1_LR_WIENER5)
LR_FN(rec_sgr5, EAV1_LR_SGR_5X5) LR_FN(rec_sgr3, EAV1_LR_SGR_3X3) LR_FN(rec_sgrmix, EAV1_LR_SGR_MIX)

void bitfn(eav1rec_loop_restoration_dsp_init)(Dav1dLoopRestorationDSPContext *const c, int bpc)
{
    (void)bpc;
    c->wiener[0] = rec_wiener7; c->wiener[1] = rec_wiener5;
    c->sgr[0] = rec_sgr5; c->sgr[1] = rec_sgr3; c->sgr[2] = rec_sgrmix;
}

/* CDEF is driven on the device (needs pixel-derived direction/variance): the host table
 * must never be reached because filter_sbrow below skips dav1d_filter_sbrow_cdef. */
static int rec_cdef_dir(const pixel *dst, ptrdiff_t s, unsigned *var HIGHBD_DECL_SUFFIX)
{ (void)dst; (void)s; *var = 0; eav1rec_unsupported(&eav1_walk); return 0; }
static void rec_cdef_fb(pixel *dst, ptrdiff_t stride, const_left_pixel_row_2px left, const pixel *top, const pixel *bottom,
                        int pri, int sec, int dir, int damping, enum CdefEdgeFlags edges HIGHBD_DECL_SUFFIX)
{ (void)dst; (void)stride; (void)left; (void)top; (void)bottom; (void)pri; (void)sec; (void)dir; (void)damping; (void)edges;
  eav1rec_unsupported(&eav1_walk); }
void bitfn(eav1rec_cdef_dsp_init)(Dav1dCdefDSPContext *const c)
{
    c->dir = rec_cdef_dir; c->fb[0] = c->fb[1] = c->fb[2] = rec_cdef_fb;
}

/* Film grain is applied at output time through dav1d_apply_grain (lib.c:311-330), which is
 * re-bound to the engine separately; the per-row DSP table is therefore never used. */
void bitfn(eav1rec_film_grain_dsp_init)(Dav1dFilmGrainDSPContext *const c)
{
    memset(c, 0, sizeof(*c));
}

/* ---------------------------------------------------------------- B4 wrappers */

void bitfn(eav1rec_recon_b_intra)(Dav1dTaskContext *const t, const enum BlockSize bs,
                                  const enum EdgeFlags intra_edge_flags, const Av1Block *const b)
{
    eav1rec_block_begin(eav1rec_enter(t->f, t), bs);
    bitfn(eav1ref_recon_b_intra)(t, bs, intra_edge_flags, b);
}

int bitfn(eav1rec_recon_b_inter)(Dav1dTaskContext *const t, const enum BlockSize bs, const Av1Block *const b)
{
    eav1rec_block_begin(eav1rec_enter(t->f, t), bs);
    return bitfn(eav1ref_recon_b_inter)(t, bs, b);
}

void bitfn(eav1rec_backup_ipred_edge)(Dav1dTaskContext *const t) { (void)t; }
void bitfn(eav1rec_copy_lpf)(Dav1dFrameContext *const f, pixel *const src[3], const int sby) { (void)f; (void)src; (void)sby; }

/* The post-filter stages as dav1d's task scheduler calls them one by one (thread_task.c:809-864).
 * Deblocking and loop-restoration drivers of the reference run against the recording tables;
 * CDEF and the super-resolution upscale are driven on the device as whole-frame stages. */
void bitfn(eav1rec_filter_sbrow_deblock_cols)(Dav1dFrameContext *const f, const int sby)
{
    eav1rec_enter(f, NULL);
    bitfn(eav1ref_filter_sbrow_deblock_cols)(f, sby);
}
void bitfn(eav1rec_filter_sbrow_deblock_rows)(Dav1dFrameContext *const f, const int sby)
{
    eav1rec_enter(f, NULL);
    bitfn(eav1ref_filter_sbrow_deblock_rows)(f, sby);
}
void bitfn(eav1rec_filter_sbrow_cdef)(Dav1dTaskContext *const tc, const int sby) { (void)tc; (void)sby; }
/* super-resolution is one whole-frame stage of the engine (Eav1FrameDesc.superres, filled in eav1rec_frame_end) */
void bitfn(eav1rec_filter_sbrow_resize)(Dav1dFrameContext *const f, const int sby) { (void)f; (void)sby; }
void bitfn(eav1rec_filter_sbrow_lr)(Dav1dFrameContext *const f, const int sby)
{
    eav1rec_enter(f, NULL);
    bitfn(eav1ref_filter_sbrow_lr)(f, sby);
}

/* Replaces dav1d_filter_sbrow (recon_tmpl.c:2100-2109), the single-threaded path: all stages
 * of one superblock row; after the last row the frame is handed over. */
void bitfn(eav1rec_filter_sbrow)(Dav1dFrameContext *const f, const int sby)
{
    Eav1Walk *w = eav1rec_enter(f, NULL);
    bitfn(eav1ref_filter_sbrow_deblock_cols)(f, sby);
    bitfn(eav1ref_filter_sbrow_deblock_rows)(f, sby);
    if (f->lf.restore_planes) bitfn(eav1ref_filter_sbrow_lr)(f, sby);
    if (sby == f->sbh - 1) eav1rec_frame_end(w->r, f, 0);
}
