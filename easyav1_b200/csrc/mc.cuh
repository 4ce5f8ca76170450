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


// ---- 8-bit fast path helpers -------------------------------------------------------------
// 32 bits starting `sh` bits into the 64-bit value hi:lo (sh in 0,8,16,24)
EAV1_DEV uint32_t funnel_r(uint32_t lo, uint32_t hi, int sh) {
#if defined(__CUDA_ARCH__)
    return __funnelshift_r(lo, hi, sh);
#else
    return (uint32_t)((((uint64_t)hi << 32) | lo) >> sh);
#endif
}
// acc + sum of four (unsigned pixel byte) x (signed filter byte) products
EAV1_DEV int dot4_us(uint32_t px4, uint32_t f4, int acc) {
#if defined(__CUDA_ARCH__)
    int d;
    asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(px4), "r"(f4), "r"(acc));
    return d;
#else
    for (int k = 0; k < 4; k++) acc += (int)((px4 >> (8 * k)) & 0xff) * (int)(int8_t)((f4 >> (8 * k)) & 0xff);
    return acc;
#endif
}

// 16 pixels of row y starting at the 4-aligned column xa, as four little-endian words.
// `inside` promises the addressed window needs no edge replication.
EAV1_DEV void fetch16(const SrcView<uint8_t> &s, int xa, int y, bool inside, uint32_t W[4]) {
    if (inside) {
        const uint32_t *p = reinterpret_cast<const uint32_t *>(s.p + (ptrdiff_t)y * s.stride + xa);
        W[0] = p[0]; W[1] = p[1]; W[2] = p[2]; W[3] = p[3];
    } else {
#pragma unroll
        for (int i = 0; i < 4; i++)
            W[i] = (uint32_t)s.at(xa + 4 * i, y) | ((uint32_t)s.at(xa + 4 * i + 1, y) << 8) |
                   ((uint32_t)s.at(xa + 4 * i + 2, y) << 16) | ((uint32_t)s.at(xa + 4 * i + 3, y) << 24);
    }
}
// 4 pixels at (x..x+3, y) as one word
EAV1_DEV uint32_t fetch4(const SrcView<uint8_t> &s, int x, int y, bool inside) {
    if (inside) {
        const int a = x & 3;
        const uint32_t *p = reinterpret_cast<const uint32_t *>(s.p + (ptrdiff_t)y * s.stride + (x - a));
        const uint32_t lo = p[0];
        return a ? funnel_r(lo, p[1], 8 * a) : lo;
    }
    return (uint32_t)s.at(x, y) | ((uint32_t)s.at(x + 1, y) << 8) | ((uint32_t)s.at(x + 2, y) << 16) |
           ((uint32_t)s.at(x + 3, y) << 24);
}
// horizontal 8-tap on 4 adjacent outputs whose first tap sits at column x0 of row y
EAV1_DEV void hfilt4(const SrcView<uint8_t> &s, int x0, int y, bool inside, uint32_t f0, uint32_t f1, int sum[4]) {
    const int a = x0 & 3;
    uint32_t W[4];
    fetch16(s, x0 - a, y, inside, W);
    const int sh = 8 * a;
    const uint32_t v0 = funnel_r(W[0], W[1], sh), v1 = funnel_r(W[1], W[2], sh), v2 = funnel_r(W[2], W[3], sh);
#pragma unroll
    for (int j = 0; j < 4; j++)
        sum[j] = dot4_us(funnel_r(v0, v1, 8 * j), f0, dot4_us(funnel_r(v1, v2, 8 * j), f1, 0));
}
EAV1_DEV void store4_px(uint8_t *d, const int v[4]) {
    const uint32_t w = (uint32_t)v[0] | ((uint32_t)v[1] << 8) | ((uint32_t)v[2] << 16) | ((uint32_t)v[3] << 24);
    if (((uintptr_t)d & 3) == 0) *reinterpret_cast<uint32_t *>(d) = w;
    else { d[0] = (uint8_t)v[0]; d[1] = (uint8_t)v[1]; d[2] = (uint8_t)v[2]; d[3] = (uint8_t)v[3]; }
}
EAV1_DEV void store4_i16(int16_t *d, const int v[4]) {
    // prep destinations are 8-byte aligned (arena offsets are 16-aligned, strides multiples of 4)
    uint2 w;
    w.x = (uint32_t)(uint16_t)v[0] | ((uint32_t)(uint16_t)v[1] << 16);
    w.y = (uint32_t)(uint16_t)v[2] | ((uint32_t)(uint16_t)v[3] << 16);
    *reinterpret_cast<uint2 *>(d) = w;
}

// 8-bit put/prep for w >= 4, four horizontally adjacent samples per work item. Arithmetic is
// the reference's (mc_tmpl.c:130-187, :247-306) with intermediate_bits = 4 and no prep bias.
EAV1_DEVH void mc_block8_fast(const Eav1Tables *tab, const SrcView<uint8_t> &src, int sx, int sy,
                             int w, int h, int fw, int fh, int mx, int my, int filt_h, int filt_v,
                             uint8_t *dst8, int16_t *dst16, int dstride, int16_t *mid, int tid, int nthr)
{
    const int lq = ulog2((unsigned)w) - 2, qmask = (1 << lq) - 1;     // quads per row = 1 << lq
    const bool inside = sx >= 3 && sy >= 3 && sx + w + 12 <= src.iw && sy + h + 4 <= src.ih;
    uint32_t fh0 = 0, fh1 = 0;
    int fv[8];
    if (mx) {
        const uint32_t *f = reinterpret_cast<const uint32_t *>(subpel_filter(tab, filt_h, mx, fw));
        fh0 = f[0]; fh1 = f[1];
    }
    if (my) {
        const int8_t *f = subpel_filter(tab, filt_v, my, fh);
#pragma unroll
        for (int k = 0; k < 8; k++) fv[k] = f[k];
    }
    if (!mx && !my) {
        for (int it = tid; it < (h << lq); it += nthr) {
            const int y = it >> lq, x = (it & qmask) << 2;
            const uint32_t v = fetch4(src, sx + x, sy + y, inside);
            int o[4];
#pragma unroll
            for (int j = 0; j < 4; j++) o[j] = (int)((v >> (8 * j)) & 0xff);
            if (dst8) store4_px(dst8 + y * dstride + x, o);
            else {
#pragma unroll
                for (int j = 0; j < 4; j++) o[j] <<= 4;
                store4_i16(dst16 + y * dstride + x, o);
            }
        }
    } else if (mx && !my) {
        for (int it = tid; it < (h << lq); it += nthr) {
            const int y = it >> lq, x = (it & qmask) << 2;
            int sum[4];
            hfilt4(src, sx + x - 3, sy + y, inside, fh0, fh1, sum);
            if (dst8) {
#pragma unroll
                for (int j = 0; j < 4; j++) sum[j] = iclip((sum[j] + 34) >> 6, 0, 255);
                store4_px(dst8 + y * dstride + x, sum);
            } else {
#pragma unroll
                for (int j = 0; j < 4; j++) sum[j] = (sum[j] + 2) >> 2;
                store4_i16(dst16 + y * dstride + x, sum);
            }
        }
    } else if (!mx) {
        for (int it = tid; it < (h << lq); it += nthr) {
            const int y = it >> lq, x = (it & qmask) << 2;
            int sum[4] = { 0, 0, 0, 0 };
#pragma unroll
            for (int k = 0; k < 8; k++) {
                const uint32_t v = fetch4(src, sx + x, sy + y + k - 3, inside);
#pragma unroll
                for (int j = 0; j < 4; j++) sum[j] += fv[k] * (int)((v >> (8 * j)) & 0xff);
            }
            if (dst8) {
#pragma unroll
                for (int j = 0; j < 4; j++) sum[j] = iclip((sum[j] + 32) >> 6, 0, 255);
                store4_px(dst8 + y * dstride + x, sum);
            } else {
#pragma unroll
                for (int j = 0; j < 4; j++) sum[j] = (sum[j] + 2) >> 2;
                store4_i16(dst16 + y * dstride + x, sum);
            }
        }
    } else {
        for (int y0 = 0; y0 < h; y0 += 32) {
            const int th = imin(32, h - y0);
#pragma unroll 2
            for (int it = tid; it < ((th + 7) << lq); it += nthr) {
                const int r = it >> lq, x = (it & qmask) << 2;
                int sum[4];
                hfilt4(src, sx + x - 3, sy + y0 + r - 3, inside, fh0, fh1, sum);
#pragma unroll
                for (int j = 0; j < 4; j++) sum[j] = (sum[j] + 2) >> 2;
                store4_i16(mid + r * w + x, sum);
            }
            CTA_SYNC();
#pragma unroll 2
            for (int it = tid; it < (th << lq); it += nthr) {
                const int r = it >> lq, x = (it & qmask) << 2;
                int sum[4] = { 0, 0, 0, 0 };
#pragma unroll
                for (int k = 0; k < 8; k++) {
                    const uint2 m = *reinterpret_cast<const uint2 *>(mid + (r + k) * w + x);
                    sum[0] += fv[k] * (int)(int16_t)(m.x & 0xffff);
                    sum[1] += fv[k] * (int)(int16_t)(m.x >> 16);
                    sum[2] += fv[k] * (int)(int16_t)(m.y & 0xffff);
                    sum[3] += fv[k] * (int)(int16_t)(m.y >> 16);
                }
                if (dst8) {
#pragma unroll
                    for (int j = 0; j < 4; j++) sum[j] = iclip((sum[j] + 512) >> 10, 0, 255);
                    store4_px(dst8 + (y0 + r) * dstride + x, sum);
                } else {
#pragma unroll
                    for (int j = 0; j < 4; j++) sum[j] = (sum[j] + 32) >> 6;
                    store4_i16(dst16 + (y0 + r) * dstride + x, sum);
                }
            }
            CTA_SYNC();
        }
        return;
    }
    CTA_SYNC();
}

// 16-bit put/prep for w >= 4 whose whole 8-tap footprint lies inside the reference picture (no edge replication: the
// caller checks), four horizontally adjacent outputs per work item. Same arithmetic as the general path of mc_block
// below (mc_tmpl.c:130-187, :247-306), but the taps sit in registers, the eleven samples four outputs share are read
// once, rows come as plain loads without per-tap clamping, and items are split by shifts instead of divisions.
EAV1_DEVH void mc_block16_fast(const Eav1Tables *tab, const SrcView<uint16_t> &src, int sx, int sy,
                              int w, int h, int fw, int fh, int mx, int my, int filt_h, int filt_v,
                              uint16_t *dst8, int16_t *dst16, int dstride, int bitdepth_max, int16_t *mid,
                              int tid, int nthr)
{
    const int ib = intermediate_bits(bitdepth_max), bias = 8192;
    const int lq = ulog2((unsigned)w) - 2, qmask = (1 << lq) - 1;
    int fhc[8] = { 0, 0, 0, 0, 0, 0, 0, 0 }, fvc[8] = { 0, 0, 0, 0, 0, 0, 0, 0 };
    if (mx) {
        const int8_t *f = subpel_filter(tab, filt_h, mx, fw);
#pragma unroll
        for (int k = 0; k < 8; k++) fhc[k] = f[k];
    }
    if (my) {
        const int8_t *f = subpel_filter(tab, filt_v, my, fh);
#pragma unroll
        for (int k = 0; k < 8; k++) fvc[k] = f[k];
    }
    auto row = [&](int y) { return reinterpret_cast<const uint16_t *>(src.p + (ptrdiff_t)y * src.stride); };
    // four horizontally filtered outputs whose first tap is at column x0 of row y
    auto hquad = [&](int x0, int y, int sum[4]) {
        const uint16_t *r = row(y) + x0;
        int v[11];
#pragma unroll
        for (int k = 0; k < 11; k++) v[k] = r[k];
#pragma unroll
        for (int j = 0; j < 4; j++) {
            int a = 0;
#pragma unroll
            for (int k = 0; k < 8; k++) a += fhc[k] * v[j + k];
            sum[j] = a;
        }
    };
    auto put4 = [&](int x, int y, const int o[4]) {          // final samples
        uint16_t *d = dst8 + y * dstride + x;
        if (((uintptr_t)d & 7) == 0) {
            uint2 q; q.x = (uint32_t)o[0] | ((uint32_t)o[1] << 16); q.y = (uint32_t)o[2] | ((uint32_t)o[3] << 16);
            *reinterpret_cast<uint2 *>(d) = q;
        } else { d[0] = (uint16_t)o[0]; d[1] = (uint16_t)o[1]; d[2] = (uint16_t)o[2]; d[3] = (uint16_t)o[3]; }
    };
    const int n_items = h << lq;
    if (mx && my) {
        const int sh1 = 6 - ib, r1 = (1 << sh1) >> 1;
        for (int y0 = 0; y0 < h; y0 += 32) {
            const int th = imin(32, h - y0);
            for (int it = tid; it < ((th + 7) << lq); it += nthr) {
                const int r = it >> lq, x = (it & qmask) << 2;
                int sum[4];
                hquad(sx + x - 3, sy + y0 + r - 3, sum);
#pragma unroll
                for (int j = 0; j < 4; j++) sum[j] = (sum[j] + r1) >> sh1;
                store4_i16(mid + r * w + x, sum);
            }
            CTA_SYNC();
            for (int it = tid; it < (th << lq); it += nthr) {
                const int r = it >> lq, x = (it & qmask) << 2;
                int sum[4] = { 0, 0, 0, 0 };
#pragma unroll
                for (int k = 0; k < 8; k++) {
                    const uint2 m = *reinterpret_cast<const uint2 *>(mid + (r + k) * w + x);
                    sum[0] += fvc[k] * (int)(int16_t)(m.x & 0xffff);
                    sum[1] += fvc[k] * (int)(int16_t)(m.x >> 16);
                    sum[2] += fvc[k] * (int)(int16_t)(m.y & 0xffff);
                    sum[3] += fvc[k] * (int)(int16_t)(m.y >> 16);
                }
                if (dst8) {
#pragma unroll
                    for (int j = 0; j < 4; j++) sum[j] = iclip((sum[j] + ((1 << (6 + ib)) >> 1)) >> (6 + ib), 0, bitdepth_max);
                    put4(x, y0 + r, sum);
                } else {
#pragma unroll
                    for (int j = 0; j < 4; j++) sum[j] = ((sum[j] + 32) >> 6) - bias;
                    store4_i16(dst16 + (y0 + r) * dstride + x, sum);
                }
            }
            CTA_SYNC();
        }
        return;
    }
    for (int it = tid; it < n_items; it += nthr) {
        const int y = it >> lq, x = (it & qmask) << 2;
        int sum[4];
        if (mx) {
            hquad(sx + x - 3, sy + y, sum);
            if (dst8) {
#pragma unroll
                for (int j = 0; j < 4; j++) sum[j] = iclip((sum[j] + 32 + ((1 << (6 - ib)) >> 1)) >> 6, 0, bitdepth_max);
            } else {
#pragma unroll
                for (int j = 0; j < 4; j++) sum[j] = ((sum[j] + ((1 << (6 - ib)) >> 1)) >> (6 - ib)) - bias;
            }
        } else if (my) {
            sum[0] = sum[1] = sum[2] = sum[3] = 0;
#pragma unroll
            for (int k = 0; k < 8; k++) {
                const uint16_t *r = row(sy + y + k - 3) + sx + x;
#pragma unroll
                for (int j = 0; j < 4; j++) sum[j] += fvc[k] * (int)r[j];
            }
            if (dst8) {
#pragma unroll
                for (int j = 0; j < 4; j++) sum[j] = iclip((sum[j] + 32) >> 6, 0, bitdepth_max);
            } else {
#pragma unroll
                for (int j = 0; j < 4; j++) sum[j] = ((sum[j] + ((1 << (6 - ib)) >> 1)) >> (6 - ib)) - bias;
            }
        } else {
            const uint16_t *r = row(sy + y) + sx + x;
#pragma unroll
            for (int j = 0; j < 4; j++) sum[j] = dst8 ? (int)r[j] : ((int)r[j] << ib) - bias;
        }
        if (dst8) put4(x, y, sum);
        else store4_i16(dst16 + y * dstride + x, sum);
    }
    CTA_SYNC();
}

// Unscaled put/prep. dst8 (pixels) xor dst16 (prep) receives a w x h block with the given
// stride (elements). `mid` is CTA scratch of at least (32+7)*128 int16.
// (w,h) is the rectangle produced; (fw,fh) the dimensions of the whole prediction block, which
// select the 4-tap or 8-tap filter set -- they differ when a block is produced tile by tile.
template <typename BD>
EAV1_DEVH void mc_block(const Eav1Tables *tab, const SrcView<typename BD::pixel> &src, int sx, int sy,
                       int w, int h, int fw, int fh, int mx, int my, int filt_h, int filt_v,
                       typename BD::pixel *dst8, int16_t *dst16, int dstride,
                       int bitdepth_max, int16_t *mid, int tid, int nthr)
{
    typedef typename BD::pixel pixel;
    const int ib = intermediate_bits(bitdepth_max);
    const int bias = PrepBias<BD>::v;
    const bool bilin = filt_h == 3;
    if constexpr (!BD::hbd) {
        if (w >= 4 && !bilin && (!dst16 || (dstride & 3) == 0)) {
            mc_block8_fast(tab, src, sx, sy, w, h, fw, fh, mx, my, filt_h, filt_v, dst8, dst16, dstride, mid, tid, nthr);
            return;
        }
    } else {
        if (w >= 4 && !bilin && (!dst16 || (dstride & 3) == 0) &&
            sx >= 3 && sy >= 3 && sx + w + 4 <= src.iw && sy + h + 4 <= src.ih) {
            mc_block16_fast(tab, src, sx, sy, w, h, fw, fh, mx, my, filt_h, filt_v, dst8, dst16, dstride, bitdepth_max,
                            mid, tid, nthr);
            return;
        }
    }
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
    const int8_t *fh_ = mx ? subpel_filter(tab, filt_h, mx, fw) : nullptr;
    const int8_t *fv = my ? subpel_filter(tab, filt_v, my, fh) : nullptr;
    if (fh_ && fv) {
        const int sh1 = 6 - ib, r1 = (1 << sh1) >> 1;
        for (int y0 = 0; y0 < h; y0 += 32) {
            const int th = imin(32, h - y0);
            for (int i = tid; i < (th + 7) * w; i += nthr) {
                const int r = i / w, x = i - r * w;
                mid[i] = (int16_t)((tap8_h(src, sx + x, sy + y0 + r - 3, fh_) + r1) >> sh1);
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
        if (fh_) {
            const int sum = tap8_h(src, sx + x, sy + y, fh_);
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
EAV1_DEVH void mc_scaled_block(const Eav1Tables *tab, const SrcView<typename BD::pixel> &src, int sx, int sy,
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
EAV1_DEVH void warp8x8_block(const Eav1Tables *tab, const SrcView<typename BD::pixel> &src, int sx, int sy,
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
// Four samples of two int16 prep blocks (stride w, 8-byte aligned when w >= 4)
EAV1_DEV void load4_i16(const int16_t *p, int v[4]) {
    const uint2 m = *reinterpret_cast<const uint2 *>(p);
    v[0] = (int16_t)(m.x & 0xffff); v[1] = (int16_t)(m.x >> 16);
    v[2] = (int16_t)(m.y & 0xffff); v[3] = (int16_t)(m.y >> 16);
}
template <typename pixel>
EAV1_DEV void store4_any(pixel *d, const int v[4]) {
    if constexpr (sizeof(pixel) == 1) store4_px(reinterpret_cast<uint8_t *>(d), v);
    else { d[0] = (pixel)v[0]; d[1] = (pixel)v[1]; d[2] = (pixel)v[2]; d[3] = (pixel)v[3]; }
}

// avg / w_avg / mask share one loop: out = clip((a*wa + b*wb + rnd) >> sh) with per-sample weights
// for mask (mc_tmpl.c:628-680). mode 0 avg, 1 w_avg, 2 mask.
template <typename BD, int MODE>
EAV1_DEV void comp_blend3(typename BD::pixel *dst, int dstride, const int16_t *t1, const int16_t *t2,
                          int w, int h, int weight, const uint8_t *mask, int bitdepth_max, int tid, int nthr)
{
    const int ib = intermediate_bits(bitdepth_max);
    const int sh = MODE == 0 ? ib + 1 : MODE == 1 ? ib + 4 : ib + 6;
    const int rnd = MODE == 0 ? (1 << ib) + PrepBias<BD>::v * 2
                  : MODE == 1 ? (8 << ib) + PrepBias<BD>::v * 16 : (32 << ib) + PrepBias<BD>::v * 64;
    if (w >= 4) {
        const int lq = ulog2((unsigned)w) - 2, qmask = (1 << lq) - 1;
        for (int it = tid; it < (h << lq); it += nthr) {
            const int y = it >> lq, x = (it & qmask) << 2, i = y * w + x;
            int a[4], b[4], o[4];
            load4_i16(t1 + i, a); load4_i16(t2 + i, b);
            uint32_t m4 = 0;
            if (MODE == 2) {
                const uint8_t *mp = mask + i;
                if (((uintptr_t)mp & 3) == 0) m4 = *reinterpret_cast<const uint32_t *>(mp);
                else m4 = (uint32_t)mp[0] | ((uint32_t)mp[1] << 8) | ((uint32_t)mp[2] << 16) | ((uint32_t)mp[3] << 24);
            }
#pragma unroll
            for (int j = 0; j < 4; j++) {
                int v;
                if (MODE == 0) v = a[j] + b[j];
                else if (MODE == 1) v = a[j] * weight + b[j] * (16 - weight);
                else { const int m = (int)((m4 >> (8 * j)) & 0xff); v = a[j] * m + b[j] * (64 - m); }
                o[j] = iclip((v + rnd) >> sh, 0, bitdepth_max);
            }
            store4_any(dst + y * dstride + x, o);
        }
    } else {
        const FastDiv fd_w(w);
        for (int i = tid; i < w * h; i += nthr) {
            const int y = fd_w.div(i), x = i - y * w;
            int v;
            if (MODE == 0) v = t1[i] + t2[i];
            else if (MODE == 1) v = t1[i] * weight + t2[i] * (16 - weight);
            else { const int m = mask[i]; v = t1[i] * m + t2[i] * (64 - m); }
            dst[y * dstride + x] = (typename BD::pixel)iclip((v + rnd) >> sh, 0, bitdepth_max);
        }
    }
    CTA_SYNC();
}

template <typename BD>
EAV1_DEV void comp_avg(typename BD::pixel *dst, int dstride, const int16_t *t1, const int16_t *t2,
                       int w, int h, int bitdepth_max, int tid, int nthr)
{
    comp_blend3<BD, 0>(dst, dstride, t1, t2, w, h, 0, nullptr, bitdepth_max, tid, nthr);
}

template <typename BD>
EAV1_DEV void comp_w_avg(typename BD::pixel *dst, int dstride, const int16_t *t1, const int16_t *t2,
                         int w, int h, int weight, int bitdepth_max, int tid, int nthr)
{
    comp_blend3<BD, 1>(dst, dstride, t1, t2, w, h, weight, nullptr, bitdepth_max, tid, nthr);
}

template <typename BD>
EAV1_DEV void comp_mask(typename BD::pixel *dst, int dstride, const int16_t *t1, const int16_t *t2,
                        int w, int h, const uint8_t *mask, int bitdepth_max, int tid, int nthr)
{
    comp_blend3<BD, 2>(dst, dstride, t1, t2, w, h, 0, mask, bitdepth_max, tid, nthr);
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
    const FastDiv fd_mw(mw);
    for (int i = tid; i < mw * mh; i += nthr) {
        const int myy = fd_mw.div(i), mxx = i - myy * mw;
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
    const FastDiv fd_w(w);
    for (int i = tid; i < w * h; i += nthr) {
        const int y = fd_w.div(i), x = i - y * w;
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
    const FastDiv fd_ww(ww);
    for (int i = tid; i < ww * h; i += nthr) {
        const int y = fd_ww.div(i), x = i - y * ww;
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
    const FastDiv fd_w(w);
    for (int i = tid; i < w * hh; i += nthr) {
        const int y = fd_w.div(i), x = i - y * w;
        const int m = mask[y], a = dst[y * dstride + x], b = tmp[y * w + x];
        dst[y * dstride + x] = (pixel)((a * (64 - m) + b * m + 32) >> 6);
    }
    CTA_SYNC();
}

} // namespace eav1
