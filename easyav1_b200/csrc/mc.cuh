// mc.cuh — inter prediction (K2 of SURVEY.md §8a) as CTA-cooperative device code.
//
// Behaviour follows ext/dav1d/src/mc_tmpl.c: put/prep 8-tap (:130-187, :247-306) with the
// 4-tap variants for w<=4 / h<=4 (:115-122), bilinear (:434-626), scaled (:190-245, :308-358,
// :492-580), avg/w_avg/mask (:628-680), blend/blend_v/blend_h (:683-721), w_mask (:724-783),
// warp_affine_8x8(t) (:799-866). emu_edge (:868-916) is folded into the source fetch: every
// read is clamped to the [0,iw) x [0,ih) rectangle, which is exactly the replicated-edge patch
// the reference builds in scratch memory first.
#pragma once
#include "dev_common.cuh"

namespace eav1 {

template <typename pixel>
struct SrcView {
    const uint8_t *p; int stride; int iw, ih;
    EAV1_DEVM int at(int x, int y) const {
        x = iclip(x, 0, iw - 1); y = iclip(y, 0, ih - 1);
        return reinterpret_cast<const pixel *>(p + (ptrdiff_t)y * stride)[x];
    }
};

EAV1_DEV int intermediate_bits(int bitdepth_max) {
    return bitdepth_max == 0xff ? 4 : 14 - (32 - clz32((unsigned)bitdepth_max));
}

template <typename BD> struct PrepBias { static constexpr int v = BD::hbd ? 8192 : 0; };

EAV1_DEV const int8_t *subpel_filter(const Eav1Tables *tab, int kind, int frac, int len) {
    // kind: 0 regular, 1 smooth, 2 sharp; 4-tap set for short dimensions
    return tab->mc_subpel_filters[len > 4 ? kind : 3 + (kind & 1)][frac - 1];
}

template <typename pixel>
EAV1_DEV int tap8_h(const SrcView<pixel> &s, int x, int y, const int8_t *f) {
    int sum = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) sum += f[k] * s.at(x + k - 3, y);
    return sum;
}
template <typename pixel>
EAV1_DEV int tap8_v(const SrcView<pixel> &s, int x, int y, const int8_t *f) {
    int sum = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) sum += f[k] * s.at(x, y + k - 3);
    return sum;
}

// Unscaled put/prep. dst8 (pixels) xor dst16 (prep) receives a w x h block with the given
// stride (elements). `mid` is CTA scratch of at least (32+7)*128 int16.
template <typename BD>
EAV1_DEV void mc_block(const Eav1Tables *tab, const SrcView<typename BD::pixel> &src, int sx, int sy,
                       int w, int h, int mx, int my, int filt_h, int filt_v,
                       typename BD::pixel *dst8, int16_t *dst16, int dstride,
                       int bitdepth_max, int16_t *mid, int tid, int nthr)
{
    typedef typename BD::pixel pixel;
    const int ib = intermediate_bits(bitdepth_max);
    const int bias = PrepBias<BD>::v;
    const bool bilin = filt_h == 3;
    if (!mx && !my) {
        for (int i = tid; i < w * h; i += nthr) {
            const int y = i / w, x = i - y * w;
            const int v = src.at(sx + x, sy + y);
            if (dst8) dst8[y * dstride + x] = (pixel)v;
            else dst16[y * dstride + x] = (int16_t)((v << ib) - bias);
        }
        CTA_SYNC();
        return;
    }
    if (bilin) {
        if (mx && my) {
            for (int y0 = 0; y0 < h; y0 += 32) {
                const int th = imin(32, h - y0);
                for (int i = tid; i < (th + 1) * w; i += nthr) {
                    const int r = i / w, x = i - r * w;
                    const int a = src.at(sx + x, sy + y0 + r), b = src.at(sx + x + 1, sy + y0 + r);
                    mid[i] = (int16_t)((16 * a + mx * (b - a) + ((1 << (4 - ib)) >> 1)) >> (4 - ib));
                }
                CTA_SYNC();
                for (int i = tid; i < th * w; i += nthr) {
                    const int r = i / w, x = i - r * w;
                    const int a = mid[i], b = mid[i + w];
                    const int t = 16 * a + my * (b - a);
                    if (dst8) dst8[(y0 + r) * dstride + x] = (pixel)iclip((t + ((1 << (4 + ib)) >> 1)) >> (4 + ib), 0, bitdepth_max);
                    else dst16[(y0 + r) * dstride + x] = (int16_t)(((t + 8) >> 4) - bias);
                }
                CTA_SYNC();
            }
        } else {
            for (int i = tid; i < w * h; i += nthr) {
                const int y = i / w, x = i - y * w;
                const int a = src.at(sx + x, sy + y);
                const int b = mx ? src.at(sx + x + 1, sy + y) : src.at(sx + x, sy + y + 1);
                const int t = 16 * a + (mx ? mx : my) * (b - a);
                if (dst8) {
                    if (mx) {
                        const int px = (t + ((1 << (4 - ib)) >> 1)) >> (4 - ib);
                        dst8[y * dstride + x] = (pixel)iclip((px + ((1 << ib) >> 1)) >> ib, 0, bitdepth_max);
                    } else {
                        dst8[y * dstride + x] = (pixel)iclip((t + 8) >> 4, 0, bitdepth_max);
                    }
                } else {
                    dst16[y * dstride + x] = (int16_t)(((t + ((1 << (4 - ib)) >> 1)) >> (4 - ib)) - bias);
                }
            }
            CTA_SYNC();
        }
        return;
    }
    const int8_t *fh = mx ? subpel_filter(tab, filt_h, mx, w) : nullptr;
    const int8_t *fv = my ? subpel_filter(tab, filt_v, my, h) : nullptr;
    if (fh && fv) {
        const int sh1 = 6 - ib, r1 = (1 << sh1) >> 1;
        for (int y0 = 0; y0 < h; y0 += 32) {
            const int th = imin(32, h - y0);
            for (int i = tid; i < (th + 7) * w; i += nthr) {
                const int r = i / w, x = i - r * w;
                mid[i] = (int16_t)((tap8_h(src, sx + x, sy + y0 + r - 3, fh) + r1) >> sh1);
            }
            CTA_SYNC();
            for (int i = tid; i < th * w; i += nthr) {
                const int r = i / w, x = i - r * w;
                int sum = 0;
#pragma unroll
                for (int k = 0; k < 8; k++) sum += fv[k] * mid[(r + k) * w + x];
                if (dst8) dst8[(y0 + r) * dstride + x] = (pixel)iclip((sum + ((1 << (6 + ib)) >> 1)) >> (6 + ib), 0, bitdepth_max);
                else dst16[(y0 + r) * dstride + x] = (int16_t)(((sum + 32) >> 6) - bias);
            }
            CTA_SYNC();
        }
        return;
    }
    for (int i = tid; i < w * h; i += nthr) {
        const int y = i / w, x = i - y * w;
        if (fh) {
            const int sum = tap8_h(src, sx + x, sy + y, fh);
            if (dst8) dst8[y * dstride + x] = (pixel)iclip((sum + 32 + ((1 << (6 - ib)) >> 1)) >> 6, 0, bitdepth_max);
            else dst16[y * dstride + x] = (int16_t)(((sum + ((1 << (6 - ib)) >> 1)) >> (6 - ib)) - bias);
        } else {
            const int sum = tap8_v(src, sx + x, sy + y, fv);
            if (dst8) dst8[y * dstride + x] = (pixel)iclip((sum + 32) >> 6, 0, bitdepth_max);
            else dst16[y * dstride + x] = (int16_t)(((sum + ((1 << (6 - ib)) >> 1)) >> (6 - ib)) - bias);
        }
    }
    CTA_SYNC();
}

// Scaled put/prep: (mx,my) are 1/1024 positions of sample (0,0), (dx,dy) the per-sample steps.
// (sx,sy) addresses the sample the reference's `src` pointer points at.
template <typename BD>
EAV1_DEV void mc_scaled_block(const Eav1Tables *tab, const SrcView<typename BD::pixel> &src, int sx, int sy,
                              int w, int h, int mx, int my, int dx, int dy, int filt_h, int filt_v,
                              typename BD::pixel *dst8, int16_t *dst16, int dstride,
                              int bitdepth_max, int tid, int nthr)
{
    typedef typename BD::pixel pixel;
    const int ib = intermediate_bits(bitdepth_max);
    const int bias = PrepBias<BD>::v;
    const bool bilin = filt_h == 3;
    for (int i = tid; i < w * h; i += nthr) {
        const int y = i / w, x = i - y * w;
        const int px = mx + x * dx, py = my + y * dy;
        const int ox = px >> 10, oy = py >> 10;
        int out;
        if (bilin) {
            const int fx = (px & 0x3ff) >> 6, fy = (py & 0x3ff) >> 6;
            int m[2];
            for (int r = 0; r < 2; r++) {
                const int a = src.at(sx + ox, sy + oy + r), b = src.at(sx + ox + 1, sy + oy + r);
                m[r] = (int16_t)((16 * a + fx * (b - a) + ((1 << (4 - ib)) >> 1)) >> (4 - ib));
            }
            const int t = 16 * m[0] + fy * (m[1] - m[0]);
            if (dst8) out = iclip((t + ((1 << (4 + ib)) >> 1)) >> (4 + ib), 0, bitdepth_max);
            else out = ((t + 8) >> 4) - bias;
        } else {
            const int fxi = (px & 0x3ff) >> 6, fyi = (py & 0x3ff) >> 6;
            const int8_t *fh = fxi ? subpel_filter(tab, filt_h, fxi, w) : nullptr;
            const int8_t *fv = fyi ? subpel_filter(tab, filt_v, fyi, h) : nullptr;
            int m[8];
            for (int r = fv ? 0 : 3; r < (fv ? 8 : 4); r++) {
                const int yy = sy + oy + r - 3;
                m[r] = fh ? (int16_t)((tap8_h(src, sx + ox, yy, fh) + ((1 << (6 - ib)) >> 1)) >> (6 - ib))
                          : (int16_t)(src.at(sx + ox, yy) << ib);
            }
            if (fv) {
                int sum = 0;
                for (int k = 0; k < 8; k++) sum += fv[k] * m[k];
                if (dst8) out = iclip((sum + ((1 << (6 + ib)) >> 1)) >> (6 + ib), 0, bitdepth_max);
                else out = ((sum + 32) >> 6) - bias;
            } else {
                if (dst8) out = iclip((m[3] + ((1 << ib) >> 1)) >> ib, 0, bitdepth_max);
                else out = m[3] - bias;
            }
        }
        if (dst8) dst8[y * dstride + x] = (pixel)out;
        else dst16[y * dstride + x] = (int16_t)out;
    }
    CTA_SYNC();
}

// 8x8 affine warp. (sx,sy) addresses the sample the reference's `src` pointer points at.
template <typename BD>
EAV1_DEV void warp8x8_block(const Eav1Tables *tab, const SrcView<typename BD::pixel> &src, int sx, int sy,
                            const int16_t abcd[4], int mx, int my,
                            typename BD::pixel *dst8, int16_t *dst16, int dstride,
                            int bitdepth_max, int16_t *mid, int tid, int nthr)
{
    typedef typename BD::pixel pixel;
    const int ib = intermediate_bits(bitdepth_max);
    const int bias = PrepBias<BD>::v;
    for (int i = tid; i < 15 * 8; i += nthr) {
        const int y = i >> 3, x = i & 7;
        const int tmx = mx + y * abcd[1] + x * abcd[0];
        const int8_t *f = tab->mc_warp_filter[64 + ((tmx + 512) >> 10)];
        mid[i] = (int16_t)((tap8_h(src, sx + x, sy + y - 3, f) + ((1 << (7 - ib)) >> 1)) >> (7 - ib));
    }
    CTA_SYNC();
    for (int i = tid; i < 64; i += nthr) {
        const int y = i >> 3, x = i & 7;
        const int tmy = my + y * abcd[3] + x * abcd[2];
        const int8_t *f = tab->mc_warp_filter[64 + ((tmy + 512) >> 10)];
        int sum = 0;
#pragma unroll
        for (int k = 0; k < 8; k++) sum += f[k] * mid[(y + k) * 8 + x];
        if (dst8) dst8[y * dstride + x] = (pixel)iclip((sum + ((1 << (7 + ib)) >> 1)) >> (7 + ib), 0, bitdepth_max);
        else dst16[y * dstride + x] = (int16_t)(((sum + 64) >> 7) - bias);
    }
    CTA_SYNC();
}

// ---- compound / blend ------------------------------------------------------------------
template <typename BD>
EAV1_DEV void comp_avg(typename BD::pixel *dst, int dstride, const int16_t *t1, const int16_t *t2,
                       int w, int h, int bitdepth_max, int tid, int nthr)
{
    const int ib = intermediate_bits(bitdepth_max), sh = ib + 1;
    const int rnd = (1 << ib) + PrepBias<BD>::v * 2;
    for (int i = tid; i < w * h; i += nthr) {
        const int y = i / w, x = i - y * w;
        dst[y * dstride + x] = (typename BD::pixel)iclip((t1[i] + t2[i] + rnd) >> sh, 0, bitdepth_max);
    }
    CTA_SYNC();
}

template <typename BD>
EAV1_DEV void comp_w_avg(typename BD::pixel *dst, int dstride, const int16_t *t1, const int16_t *t2,
                         int w, int h, int weight, int bitdepth_max, int tid, int nthr)
{
    const int ib = intermediate_bits(bitdepth_max), sh = ib + 4;
    const int rnd = (8 << ib) + PrepBias<BD>::v * 16;
    for (int i = tid; i < w * h; i += nthr) {
        const int y = i / w, x = i - y * w;
        dst[y * dstride + x] = (typename BD::pixel)iclip((t1[i] * weight + t2[i] * (16 - weight) + rnd) >> sh, 0, bitdepth_max);
    }
    CTA_SYNC();
}

template <typename BD>
EAV1_DEV void comp_mask(typename BD::pixel *dst, int dstride, const int16_t *t1, const int16_t *t2,
                        int w, int h, const uint8_t *mask, int bitdepth_max, int tid, int nthr)
{
    const int ib = intermediate_bits(bitdepth_max), sh = ib + 6;
    const int rnd = (32 << ib) + PrepBias<BD>::v * 64;
    for (int i = tid; i < w * h; i += nthr) {
        const int y = i / w, x = i - y * w;
        const int m = mask[i];
        dst[y * dstride + x] = (typename BD::pixel)iclip((t1[i] * m + t2[i] * (64 - m) + rnd) >> sh, 0, bitdepth_max);
    }
    CTA_SYNC();
}

// w_mask: blends and writes the (optionally 2x1 / 2x2 down-sampled) mask, stride w>>ss_hor.
template <typename BD>
EAV1_DEV void comp_w_mask(typename BD::pixel *dst, int dstride, const int16_t *t1, const int16_t *t2,
                          int w, int h, uint8_t *mask, int sign, int ss_hor, int ss_ver,
                          int bitdepth_max, int tid, int nthr)
{
    const int ib = intermediate_bits(bitdepth_max), sh = ib + 6;
    const int bitdepth = 32 - clz32((unsigned)bitdepth_max);
    const int rnd = (32 << ib) + PrepBias<BD>::v * 64;
    const int mask_sh = bitdepth + ib - 4, mask_rnd = 1 << (mask_sh - 5);
    // one work item per mask sample (covers 1, 2 or 4 pixels)
    const int mw = w >> ss_hor, mh = h >> ss_ver;
    for (int i = tid; i < mw * mh; i += nthr) {
        const int myy = i / mw, mxx = i - myy * mw;
        int msum = 0;
        for (int yy = 0; yy <= ss_ver; yy++)
            for (int xx = 0; xx <= ss_hor; xx++) {
                const int x = (mxx << ss_hor) + xx, y = (myy << ss_ver) + yy;
                const int a = t1[y * w + x], b = t2[y * w + x];
                const int diff = a - b;
                const int m = imin(38 + ((iabs(diff) + mask_rnd) >> mask_sh), 64);
                dst[y * dstride + x] = (typename BD::pixel)iclip((diff * m + b * 64 + rnd) >> sh, 0, bitdepth_max);
                msum += m;
            }
        if (ss_hor && ss_ver) mask[i] = (uint8_t)((msum + 2 - sign) >> 2);
        else if (ss_hor) mask[i] = (uint8_t)((msum + 1 - sign) >> 1);
        else mask[i] = (uint8_t)msum;
    }
    CTA_SYNC();
}

template <typename pixel>
EAV1_DEV void blend_mask(pixel *dst, int dstride, const pixel *tmp, int w, int h, const uint8_t *mask, int tid, int nthr)
{
    for (int i = tid; i < w * h; i += nthr) {
        const int y = i / w, x = i - y * w;
        const int m = mask[i], a = dst[y * dstride + x], b = tmp[i];
        dst[y * dstride + x] = (pixel)((a * (64 - m) + b * m + 32) >> 6);
    }
    CTA_SYNC();
}

template <typename pixel>
EAV1_DEV void blend_v(const Eav1Tables *tab, pixel *dst, int dstride, const pixel *tmp, int w, int h, int tid, int nthr)
{
    const uint8_t *mask = &tab->obmc_masks[w];
    const int ww = (w * 3) >> 2;
    for (int i = tid; i < ww * h; i += nthr) {
        const int y = i / ww, x = i - y * ww;
        const int m = mask[x], a = dst[y * dstride + x], b = tmp[y * w + x];
        dst[y * dstride + x] = (pixel)((a * (64 - m) + b * m + 32) >> 6);
    }
    CTA_SYNC();
}

template <typename pixel>
EAV1_DEV void blend_h(const Eav1Tables *tab, pixel *dst, int dstride, const pixel *tmp, int w, int h, int tid, int nthr)
{
    const uint8_t *mask = &tab->obmc_masks[h];
    const int hh = (h * 3) >> 2;
    for (int i = tid; i < w * hh; i += nthr) {
        const int y = i / w, x = i - y * w;
        const int m = mask[y], a = dst[y * dstride + x], b = tmp[y * w + x];
        dst[y * dstride + x] = (pixel)((a * (64 - m) + b * m + 32) >> 6);
    }
    CTA_SYNC();
}

} // namespace eav1
