/*
 * oracle/ref_harness.c — TEST INFRASTRUCTURE ONLY (never linked into the product).
 *
 * Thin flat-C entry points around the *reference's own* dav1d C code, compiled
 * in place from /root/reference by oracle/Makefile into oracle/_ref/libdav1d_ref.so.
 * Nothing here computes pixels: every function forwards to the reference
 *   - stream level: dav1d_open / dav1d_send_data / dav1d_get_picture
 *     (ext/dav1d/src/lib.c:144,438,458) -> decoded visible planes (the input of
 *     the per-frame MD5 of ext/dav1d/tools/output/md5.c:194-250);
 *   - operator level: the C function tables filled by dav1d_*_dsp_init_{8,16}bpc
 *     (ext/dav1d/src/internal.h:62-70), i.e. boundary B3 of SURVEY.md §8(b).
 */
#include <errno.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "config.h"
#include "dav1d/dav1d.h"
#include "src/internal.h"

/* ------------------------------------------------------------------ stream */

typedef struct RefDec {
    Dav1dContext *c;
    Dav1dData data;
    int eof_sent;
} RefDec;

typedef struct RefFrameInfo {
    int w, h, layout, bpc;
    int frame_type, show_frame, showable;
    int64_t offset;
} RefFrameInfo;

static void no_free(const uint8_t *d, void *cookie) { (void)d; (void)cookie; }
static void quiet_log(void *cookie, const char *fmt, va_list ap) { (void)cookie; (void)fmt; (void)ap; }

RefDec *ref_dec_open(const uint8_t *obu, size_t sz, int n_threads, int max_frame_delay,
                     int apply_grain, int inloop_filters, int quiet)
{
    RefDec *d = calloc(1, sizeof(*d));
    Dav1dSettings s;
    dav1d_default_settings(&s);
    s.n_threads = n_threads;
    s.max_frame_delay = max_frame_delay;
    s.apply_grain = apply_grain;
    s.inloop_filters = inloop_filters;
    if (quiet) s.logger.callback = quiet_log;
    if (dav1d_open(&d->c, &s) < 0) { free(d); return NULL; }
    if (dav1d_data_wrap(&d->data, obu, sz, no_free, NULL) < 0) {
        dav1d_close(&d->c); free(d); return NULL;
    }
    return d;
}

/* Returns 0 and fills info/planes (visible area, tightly packed Y,U,V, 16-bit LE
 * for hbd) on success; 1 at end of stream; <0 on decoder error.
 * `out` may be NULL to only query; cap is its capacity in bytes. */
int ref_dec_next(RefDec *d, RefFrameInfo *info, uint8_t *out, size_t cap, size_t *used)
{
    Dav1dPicture p;
    for (;;) {
        memset(&p, 0, sizeof(p));
        int res;
        if (d->data.sz > 0) {
            res = dav1d_send_data(d->c, &d->data);
            if (res < 0 && res != DAV1D_ERR(EAGAIN)) return res;
        }
        res = dav1d_get_picture(d->c, &p);
        if (res == 0) break;
        if (res != DAV1D_ERR(EAGAIN)) return res;
        if (d->data.sz == 0) {
            /* drain */
            res = dav1d_get_picture(d->c, &p);
            if (res == 0) break;
            return res == DAV1D_ERR(EAGAIN) ? 1 : res;
        }
    }
    info->w = p.p.w; info->h = p.p.h; info->layout = p.p.layout; info->bpc = p.p.bpc;
    info->frame_type = p.frame_hdr->frame_type;
    info->show_frame = p.frame_hdr->show_frame;
    info->showable = p.frame_hdr->showable_frame;
    info->offset = p.m.offset;
    const int hbd = p.p.bpc > 8;
    const int ss_hor = p.p.layout != DAV1D_PIXEL_LAYOUT_I444;
    const int ss_ver = p.p.layout == DAV1D_PIXEL_LAYOUT_I420;
    size_t n = 0;
    for (int pl = 0; pl < (p.p.layout == DAV1D_PIXEL_LAYOUT_I400 ? 1 : 3); pl++) {
        const int w = pl ? (p.p.w + ss_hor) >> ss_hor : p.p.w;
        const int h = pl ? (p.p.h + ss_ver) >> ss_ver : p.p.h;
        const size_t row = (size_t)w << hbd;
        const uint8_t *src = p.data[pl];
        for (int y = 0; y < h; y++) {
            if (out && n + row <= cap) memcpy(out + n, src, row);
            n += row;
            src += p.stride[!!pl];
        }
    }
    if (used) *used = n;
    dav1d_picture_unref(&p);
    return 0;
}

void ref_dec_close(RefDec *d)
{
    if (!d) return;
    dav1d_data_unref(&d->data);
    dav1d_close(&d->c);
    free(d);
}

/* --------------------------------------------------------------- operators */

static Dav1dDSPContext g_dsp[3]; /* 8, 10, 12 bpc */
static int g_dsp_init[3];

static const Dav1dDSPContext *get_dsp(int bpc)
{
    const int i = (bpc - 8) >> 1;
    if (!g_dsp_init[i]) {
        Dav1dDSPContext *d = &g_dsp[i];
        if (bpc == 8) {
            dav1d_cdef_dsp_init_8bpc(&d->cdef);
            dav1d_intra_pred_dsp_init_8bpc(&d->ipred);
            dav1d_itx_dsp_init_8bpc(&d->itx, bpc);
            dav1d_loop_filter_dsp_init_8bpc(&d->lf);
            dav1d_loop_restoration_dsp_init_8bpc(&d->lr, bpc);
            dav1d_mc_dsp_init_8bpc(&d->mc);
            dav1d_film_grain_dsp_init_8bpc(&d->fg);
        } else {
            dav1d_cdef_dsp_init_16bpc(&d->cdef);
            dav1d_intra_pred_dsp_init_16bpc(&d->ipred);
            dav1d_itx_dsp_init_16bpc(&d->itx, bpc);
            dav1d_loop_filter_dsp_init_16bpc(&d->lf);
            dav1d_loop_restoration_dsp_init_16bpc(&d->lr, bpc);
            dav1d_mc_dsp_init_16bpc(&d->mc);
            dav1d_film_grain_dsp_init_16bpc(&d->fg);
        }
        g_dsp_init[i] = 1;
    }
    return &g_dsp[i];
}

/* The 16bpc function types carry a trailing `int bitdepth_max`
 * (include/common/bitdepth.h:72-74); cast accordingly. */
#define BDMAX(bpc) ((1 << (bpc)) - 1)

void ref_itxfm_add(int bpc, int tx, int txtp, void *dst, ptrdiff_t stride, void *coef, int eob)
{
    const Dav1dDSPContext *d = get_dsp(bpc);
    if (bpc == 8)
        d->itx.itxfm_add[tx][txtp](dst, stride, coef, eob);
    else
        ((void (*)(void *, ptrdiff_t, void *, int, int))d->itx.itxfm_add[tx][txtp])(dst, stride, coef, eob, BDMAX(bpc));
}

void ref_mc_put(int bpc, int filter2d, void *dst, ptrdiff_t dst_stride, const void *src,
                ptrdiff_t src_stride, int w, int h, int mx, int my)
{
    const Dav1dDSPContext *d = get_dsp(bpc);
    if (bpc == 8) d->mc.mc[filter2d](dst, dst_stride, src, src_stride, w, h, mx, my);
    else ((void (*)(void *, ptrdiff_t, const void *, ptrdiff_t, int, int, int, int, int))
              d->mc.mc[filter2d])(dst, dst_stride, src, src_stride, w, h, mx, my, BDMAX(bpc));
}

void ref_mc_prep(int bpc, int filter2d, int16_t *tmp, const void *src, ptrdiff_t src_stride,
                 int w, int h, int mx, int my)
{
    const Dav1dDSPContext *d = get_dsp(bpc);
    if (bpc == 8) d->mc.mct[filter2d](tmp, src, src_stride, w, h, mx, my);
    else ((void (*)(int16_t *, const void *, ptrdiff_t, int, int, int, int, int))
              d->mc.mct[filter2d])(tmp, src, src_stride, w, h, mx, my, BDMAX(bpc));
}

void ref_mc_put_scaled(int bpc, int filter2d, void *dst, ptrdiff_t dst_stride, const void *src,
                       ptrdiff_t src_stride, int w, int h, int mx, int my, int dx, int dy)
{
    const Dav1dDSPContext *d = get_dsp(bpc);
    if (bpc == 8) d->mc.mc_scaled[filter2d](dst, dst_stride, src, src_stride, w, h, mx, my, dx, dy);
    else ((void (*)(void *, ptrdiff_t, const void *, ptrdiff_t, int, int, int, int, int, int, int))
              d->mc.mc_scaled[filter2d])(dst, dst_stride, src, src_stride, w, h, mx, my, dx, dy, BDMAX(bpc));
}

void ref_mc_prep_scaled(int bpc, int filter2d, int16_t *tmp, const void *src, ptrdiff_t src_stride,
                        int w, int h, int mx, int my, int dx, int dy)
{
    const Dav1dDSPContext *d = get_dsp(bpc);
    if (bpc == 8) d->mc.mct_scaled[filter2d](tmp, src, src_stride, w, h, mx, my, dx, dy);
    else ((void (*)(int16_t *, const void *, ptrdiff_t, int, int, int, int, int, int, int))
              d->mc.mct_scaled[filter2d])(tmp, src, src_stride, w, h, mx, my, dx, dy, BDMAX(bpc));
}

void ref_avg(int bpc, void *dst, ptrdiff_t stride, const int16_t *t1, const int16_t *t2, int w, int h)
{
    const Dav1dDSPContext *d = get_dsp(bpc);
    if (bpc == 8) d->mc.avg(dst, stride, t1, t2, w, h);
    else ((void (*)(void *, ptrdiff_t, const int16_t *, const int16_t *, int, int, int))
              d->mc.avg)(dst, stride, t1, t2, w, h, BDMAX(bpc));
}

void ref_w_avg(int bpc, void *dst, ptrdiff_t stride, const int16_t *t1, const int16_t *t2, int w, int h, int weight)
{
    const Dav1dDSPContext *d = get_dsp(bpc);
    if (bpc == 8) d->mc.w_avg(dst, stride, t1, t2, w, h, weight);
    else ((void (*)(void *, ptrdiff_t, const int16_t *, const int16_t *, int, int, int, int))
              d->mc.w_avg)(dst, stride, t1, t2, w, h, weight, BDMAX(bpc));
}

void ref_mask(int bpc, void *dst, ptrdiff_t stride, const int16_t *t1, const int16_t *t2, int w, int h, const uint8_t *mask)
{
    const Dav1dDSPContext *d = get_dsp(bpc);
    if (bpc == 8) d->mc.mask(dst, stride, t1, t2, w, h, mask);
    else ((void (*)(void *, ptrdiff_t, const int16_t *, const int16_t *, int, int, const uint8_t *, int))
              d->mc.mask)(dst, stride, t1, t2, w, h, mask, BDMAX(bpc));
}

void ref_w_mask(int bpc, int ss_idx /*0=444,1=422,2=420*/, void *dst, ptrdiff_t stride, const int16_t *t1,
                const int16_t *t2, int w, int h, uint8_t *mask, int sign)
{
    const Dav1dDSPContext *d = get_dsp(bpc);
    if (bpc == 8) d->mc.w_mask[ss_idx](dst, stride, t1, t2, w, h, mask, sign);
    else ((void (*)(void *, ptrdiff_t, const int16_t *, const int16_t *, int, int, uint8_t *, int, int))
              d->mc.w_mask[ss_idx])(dst, stride, t1, t2, w, h, mask, sign, BDMAX(bpc));
}

void ref_blend(int bpc, void *dst, ptrdiff_t stride, const void *tmp, int w, int h, const uint8_t *mask)
{
    get_dsp(bpc)->mc.blend(dst, stride, tmp, w, h, mask);
}
void ref_blend_v(int bpc, void *dst, ptrdiff_t stride, const void *tmp, int w, int h)
{
    get_dsp(bpc)->mc.blend_v(dst, stride, tmp, w, h);
}
void ref_blend_h(int bpc, void *dst, ptrdiff_t stride, const void *tmp, int w, int h)
{
    get_dsp(bpc)->mc.blend_h(dst, stride, tmp, w, h);
}

void ref_warp8x8(int bpc, void *dst, ptrdiff_t dst_stride, const void *src, ptrdiff_t src_stride,
                 const int16_t *abcd, int mx, int my)
{
    const Dav1dDSPContext *d = get_dsp(bpc);
    if (bpc == 8) d->mc.warp8x8(dst, dst_stride, src, src_stride, abcd, mx, my);
    else ((void (*)(void *, ptrdiff_t, const void *, ptrdiff_t, const int16_t *, int, int, int))
              d->mc.warp8x8)(dst, dst_stride, src, src_stride, abcd, mx, my, BDMAX(bpc));
}

void ref_warp8x8t(int bpc, int16_t *tmp, ptrdiff_t tmp_stride, const void *src, ptrdiff_t src_stride,
                  const int16_t *abcd, int mx, int my)
{
    const Dav1dDSPContext *d = get_dsp(bpc);
    if (bpc == 8) d->mc.warp8x8t(tmp, tmp_stride, src, src_stride, abcd, mx, my);
    else ((void (*)(int16_t *, ptrdiff_t, const void *, ptrdiff_t, const int16_t *, int, int, int))
              d->mc.warp8x8t)(tmp, tmp_stride, src, src_stride, abcd, mx, my, BDMAX(bpc));
}

void ref_emu_edge(int bpc, intptr_t bw, intptr_t bh, intptr_t iw, intptr_t ih, intptr_t x, intptr_t y,
                  void *dst, ptrdiff_t dst_stride, const void *src, ptrdiff_t src_stride)
{
    get_dsp(bpc)->mc.emu_edge(bw, bh, iw, ih, x, y, dst, dst_stride, src, src_stride);
}

void ref_resize(int bpc, void *dst, ptrdiff_t dst_stride, const void *src, ptrdiff_t src_stride,
                int dst_w, int h, int src_w, int dx, int mx)
{
    const Dav1dDSPContext *d = get_dsp(bpc);
    if (bpc == 8) d->mc.resize(dst, dst_stride, src, src_stride, dst_w, h, src_w, dx, mx);
    else ((void (*)(void *, ptrdiff_t, const void *, ptrdiff_t, int, int, int, int, int, int))
              d->mc.resize)(dst, dst_stride, src, src_stride, dst_w, h, src_w, dx, mx, BDMAX(bpc));
}

void ref_intra_pred(int bpc, int mode, void *dst, ptrdiff_t stride, const void *topleft,
                    int w, int h, int angle, int max_w, int max_h)
{
    const Dav1dDSPContext *d = get_dsp(bpc);
    if (bpc == 8) d->ipred.intra_pred[mode](dst, stride, topleft, w, h, angle, max_w, max_h);
    else ((void (*)(void *, ptrdiff_t, const void *, int, int, int, int, int, int))
              d->ipred.intra_pred[mode])(dst, stride, topleft, w, h, angle, max_w, max_h, BDMAX(bpc));
}

void ref_cfl_ac(int bpc, int layout_idx /*0=420,1=422,2=444*/, int16_t *ac, const void *y, ptrdiff_t stride,
                int w_pad, int h_pad, int cw, int ch)
{
    get_dsp(bpc)->ipred.cfl_ac[layout_idx](ac, y, stride, w_pad, h_pad, cw, ch);
}

void ref_cfl_pred(int bpc, int mode, void *dst, ptrdiff_t stride, const void *topleft, int w, int h,
                  const int16_t *ac, int alpha)
{
    const Dav1dDSPContext *d = get_dsp(bpc);
    if (bpc == 8) d->ipred.cfl_pred[mode](dst, stride, topleft, w, h, ac, alpha);
    else ((void (*)(void *, ptrdiff_t, const void *, int, int, const int16_t *, int, int))
              d->ipred.cfl_pred[mode])(dst, stride, topleft, w, h, ac, alpha, BDMAX(bpc));
}

void ref_pal_pred(int bpc, void *dst, ptrdiff_t stride, const void *pal, const uint8_t *idx, int w, int h)
{
    get_dsp(bpc)->ipred.pal_pred(dst, stride, pal, idx, w, h);
}

/* prepare_intra_edges is not in a DSP table but is part of the path (K3, D1):
 * ext/dav1d/src/ipred_prepare_tmpl.c:75-204. Returns the implementation mode. */
int ref_prepare_intra_edges(int bpc, int x, int have_left, int y, int have_top, int w, int h,
                            int edge_flags, const void *dst, ptrdiff_t stride,
                            const void *prefilter_toplevel_sb_edge, int mode, int *angle,
                            int tw, int th, int filter_edge, void *topleft_out)
{
    if (bpc == 8)
        return dav1d_prepare_intra_edges_8bpc(x, have_left, y, have_top, w, h, edge_flags, dst, stride,
                                              prefilter_toplevel_sb_edge, mode, angle, tw, th,
                                              filter_edge, topleft_out);
    return dav1d_prepare_intra_edges_16bpc(x, have_left, y, have_top, w, h, edge_flags, dst, stride,
                                           prefilter_toplevel_sb_edge, mode, angle, tw, th,
                                           filter_edge, topleft_out, BDMAX(bpc));
}

void ref_loop_filter_sb(int bpc, int plane_uv, int dir_v, void *dst, ptrdiff_t stride, const uint32_t *mask,
                        const uint8_t (*lvl)[4], ptrdiff_t lvl_stride, const Av1FilterLUT *lut, int w)
{
    const Dav1dDSPContext *d = get_dsp(bpc);
    if (bpc == 8) d->lf.loop_filter_sb[plane_uv][dir_v](dst, stride, mask, lvl, lvl_stride, lut, w);
    else ((void (*)(void *, ptrdiff_t, const uint32_t *, const uint8_t (*)[4], ptrdiff_t,
                    const Av1FilterLUT *, int, int))
              d->lf.loop_filter_sb[plane_uv][dir_v])(dst, stride, mask, lvl, lvl_stride, lut, w, BDMAX(bpc));
}

int ref_cdef_dir(int bpc, const void *dst, ptrdiff_t stride, unsigned *var)
{
    const Dav1dDSPContext *d = get_dsp(bpc);
    if (bpc == 8) return d->cdef.dir(dst, stride, var);
    return ((int (*)(const void *, ptrdiff_t, unsigned *, int))d->cdef.dir)(dst, stride, var, BDMAX(bpc));
}

void ref_cdef_fb(int bpc, int idx /*0=8x8,1=4x8,2=4x4*/, void *dst, ptrdiff_t stride, const void *left,
                 const void *top, const void *bottom, int pri, int sec, int dir, int damping, int edges)
{
    const Dav1dDSPContext *d = get_dsp(bpc);
    if (bpc == 8) d->cdef.fb[idx](dst, stride, left, top, bottom, pri, sec, dir, damping, edges);
    else ((void (*)(void *, ptrdiff_t, const void *, const void *, const void *, int, int, int, int, int, int))
              d->cdef.fb[idx])(dst, stride, left, top, bottom, pri, sec, dir, damping, edges, BDMAX(bpc));
}

void ref_lr_wiener(int bpc, int idx /*0=7tap,1=5tap*/, void *dst, ptrdiff_t stride, const void *left,
                   const void *lpf, int w, int h, const LooprestorationParams *params, int edges)
{
    const Dav1dDSPContext *d = get_dsp(bpc);
    if (bpc == 8) d->lr.wiener[idx](dst, stride, left, lpf, w, h, params, edges);
    else ((void (*)(void *, ptrdiff_t, const void *, const void *, int, int, const LooprestorationParams *, int, int))
              d->lr.wiener[idx])(dst, stride, left, lpf, w, h, params, edges, BDMAX(bpc));
}

void ref_lr_sgr(int bpc, int idx /*0=5x5,1=3x3,2=mix*/, void *dst, ptrdiff_t stride, const void *left,
                const void *lpf, int w, int h, const LooprestorationParams *params, int edges)
{
    const Dav1dDSPContext *d = get_dsp(bpc);
    if (bpc == 8) d->lr.sgr[idx](dst, stride, left, lpf, w, h, params, edges);
    else ((void (*)(void *, ptrdiff_t, const void *, const void *, int, int, const LooprestorationParams *, int, int))
              d->lr.sgr[idx])(dst, stride, left, lpf, w, h, params, edges, BDMAX(bpc));
}

/* Film grain applied to a whole picture through the reference driver
 * (ext/dav1d/src/fg_apply_tmpl.c:225-241). Planes are caller-owned. */
void ref_apply_grain(int bpc, int w, int h, int layout, int mtrx_identity, const Dav1dFilmGrainData *fg,
                     void *const out_data[3], void *const in_data[3], const ptrdiff_t stride[2])
{
    const Dav1dDSPContext *d = get_dsp(bpc);
    Dav1dPicture in, out;
    Dav1dFrameHeader fh;
    Dav1dSequenceHeader sh;
    memset(&in, 0, sizeof(in)); memset(&out, 0, sizeof(out)); memset(&fh, 0, sizeof(fh)); memset(&sh, 0, sizeof(sh));
    fh.film_grain.data = *fg;
    sh.mtrx = mtrx_identity ? DAV1D_MC_IDENTITY : DAV1D_MC_BT709;
    in.seq_hdr = out.seq_hdr = &sh;
    in.frame_hdr = out.frame_hdr = &fh;
    in.p.w = out.p.w = w; in.p.h = out.p.h = h;
    in.p.layout = out.p.layout = layout; in.p.bpc = out.p.bpc = bpc;
    for (int i = 0; i < 3; i++) { in.data[i] = in_data[i]; out.data[i] = out_data[i]; }
    for (int i = 0; i < 2; i++) in.stride[i] = out.stride[i] = stride[i];
    if (bpc == 8) dav1d_apply_grain_8bpc(&d->fg, &out, &in);
    else dav1d_apply_grain_16bpc(&d->fg, &out, &in);
}

/* Constant tables the kernels must reproduce verbatim (ext/dav1d/src/tables.c,
 * wedge.c). Exposed so tests can diff the product's device tables against them. */
const void *ref_table(const char *name, size_t *size)
{
#define T(n) if (!strcmp(name, #n)) { *size = sizeof(n); return n; }
    T(dav1d_mc_subpel_filters) T(dav1d_mc_warp_filter) T(dav1d_resize_filter)
    T(dav1d_sm_weights) T(dav1d_dr_intra_derivative) T(dav1d_filter_intra_taps)
    T(dav1d_obmc_masks) T(dav1d_gaussian_sequence) T(dav1d_sgr_x_by_x)
    T(dav1d_cdef_directions) T(dav1d_sgr_params) T(dav1d_block_dimensions)
    T(dav1d_txfm_dimensions)
#undef T
    *size = 0;
    return NULL;
}
