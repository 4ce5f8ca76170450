// ipred.cuh — intra prediction (K3 + the edge builder of D1) as
// CTA-cooperative device code.
//
// Behaviour follows ext/dav1d/src/ipred_tmpl.c (DC family :86-218, v/h :220-243, paeth :244,
// smooth :267-326, edge filter/upsample :327-407, z1/z2/z3 :408-617, filter-intra :618,
// cfl_ac :657, cfl_pred :72-84, pal_pred :717) and the edge gathering of
// dav1d_prepare_intra_edges (ext/dav1d/src/ipred_prepare_tmpl.c:75-204). The edge array is
// built in shared memory straight from the reconstructed current frame on the device (the
// host only says how many neighbouring samples are available), then every output sample is
// computed independently by one thread; filter-intra runs as a wavefront of its 4x2 cells.
#pragma once
#include "dev_common.cuh"

namespace eav1 {

// implementation modes (ext/dav1d/src/levels.h:94-131)
enum {
    IP_DC = 0, IP_VERT, IP_HOR, IP_LEFT_DC, IP_TOP_DC, IP_DC_128, IP_Z1, IP_Z2, IP_Z3,
    IP_SMOOTH, IP_SMOOTH_V, IP_SMOOTH_H, IP_PAETH, IP_FILTER
};

#define EDGE_C 144 // index of the top-left sample inside the shared edge arrays

// needs: bit0 left, bit1 top, bit2 topleft, bit3 topright, bit4 bottomleft
EAV1_DEV int ipred_needs(int mode) {
    // five bits per mode, packed (a local array would be built on the stack at every call):
    // { 3, 2, 1, 1, 2, 0, 2|8|4, 1|2|4, 1|16|4, 3, 3, 3, 7, 7 }
    const unsigned long long packed =
        3ull | 2ull << 5 | 1ull << 10 | 1ull << 15 | 2ull << 20 | 0ull << 25 | 14ull << 30 | 7ull << 35 |
        21ull << 40 | 3ull << 45 | 3ull << 50 | 3ull << 55;
    return mode < 12 ? (int)((packed >> (5 * mode)) & 31) : 7;
}

// Gather the edge of the block at (ex,ey) in `pl` into E[-2*sz_l .. 2*sz_t] (E = &edge[EDGE_C]).
template <typename pixel>
EAV1_DEV void build_edge(const Plane &pl, const Eav1Cmd &c, int bitdepth_max, uint16_t *edge, int tid, int nthr)
{
    uint16_t *E = edge + EDGE_C;
    const int ex = c.u.ipred.ex, ey = c.u.ipred.ey;
    const int mode = c.u.ipred.mode;
    const int needs = ipred_needs(mode);
    const int have_left = c.u.ipred.have & 1, have_top = (c.u.ipred.have >> 1) & 1;
    const int bitdepth = 32 - clz32((unsigned)bitdepth_max);
    const int szl = c.u.ipred.th << 2, szt = c.u.ipred.tw << 2;
    PROF_BEGIN();
    PROF_MARK(20);
    // One pass over every edge sample that is needed: [0,nl) left, [nl,nl+nbl) bottom-left, then top and
    // top-right. Each sample is one independent load (or a constant), so the whole gather costs one
    // trip to L2 instead of one per part.
    const int nl = (needs & 1) ? szl : 0, nbl = ((needs & 1) && (needs & 16)) ? szl : 0;
    const int nt = (needs & 2) ? szt : 0, ntr = ((needs & 2) && (needs & 8)) ? szt : 0;
    const int n_left = c.u.ipred.n_left, n_top = c.u.ipred.n_top;
    const int nb = (c.u.ipred.have & 8) ? c.u.ipred.n_bl : 0, nr = (c.u.ipred.have & 16) ? c.u.ipred.n_tr : 0;
    // The usual case — no top-right / bottom-left part, every needed neighbour exists and none has to be replicated —
    // needs none of the case analysis below: left column, top row and corner are plain loads. This code runs once per
    // transform block on the frame's critical chain, where every instruction is a few cycles of a single warp.
    if (!nbl && !ntr && (!nl || (have_left && n_left >= szl)) && (!nt || (have_top && n_top >= szt)) &&
        (!(needs & 4) || (have_left && have_top && !(c.u.ipred.have & 4)))) {
        const int nc = (needs & 4) ? 1 : 0;
        for (int j = tid; j < nl + nt + nc; j += nthr) {
            int px, py, slot;
            if (j < nl) { px = ex - 1; py = ey + j; slot = -(1 + j); }
            else if (j < nl + nt) { px = ex + (j - nl); py = ey - 1; slot = 1 + (j - nl); }
            else { px = ex - 1; py = ey - 1; slot = 0; }
            E[slot] = (uint16_t)*px_at<pixel>(pl, px, py);
        }
        PROF_MARK(22);
        CTA_SYNC();
        PROF_MARK(24);
        return;
    }
    const int fill_l = ((1 << bitdepth) >> 1) + 1, fill_t = ((1 << bitdepth) >> 1) - 1;
    const int total = nl + nbl + nt + ntr;
    // where edge sample j comes from: (px,py) in the plane, or the constant cst >= 0; slot = index into E
    auto locate = [&](int j, int &px, int &py, int &cst, int &slot) {
        cst = -1;
        if (j < nl + nbl) {   // left column, going down
            const int i = j < nl ? j : j - nl;
            slot = j < nl ? -(1 + i) : -(szl + 1 + i);
            px = ex - 1;
            if (j < nl) py = ey + imin(i, n_left - 1);
            else if (nb) py = ey + szl + imin(i, nb - 1);
            else py = ey + imin(szl - 1, n_left - 1);
            if (!have_left) { if (have_top) { px = ex; py = ey - 1; } else cst = fill_l; }
        } else {              // top row, going right
            const int jj = j - nl - nbl;
            const int i = jj < nt ? jj : jj - nt;
            slot = jj < nt ? 1 + i : 1 + szt + i;
            py = ey - 1;
            if (jj < nt) px = ex + imin(i, n_top - 1);
            else if (nr) px = ex + szt + imin(i, nr - 1);
            else px = ex + imin(szt - 1, n_top - 1);
            if (!have_top) { if (have_left) { px = ex - 1; py = ey; } else cst = fill_t; }
        }
    };
    // two samples per thread and round, both loads in flight before either is stored
    for (int j = tid; j < total; j += 2 * nthr) {
        int px0, py0, c0, s0, px1 = 0, py1 = 0, c1 = 0, s1 = 0;
        const bool two = j + nthr < total;
        locate(j, px0, py0, c0, s0);
        if (two) locate(j + nthr, px1, py1, c1, s1);
        int v0 = c0, v1 = c1;
        if (c0 < 0) v0 = *px_at<pixel>(pl, px0, py0);
        if (two && c1 < 0) v1 = *px_at<pixel>(pl, px1, py1);
        E[s0] = (uint16_t)v0;
        if (two) E[s1] = (uint16_t)v1;
    }
    PROF_MARK(22);
    // top-left sample: the last thread fetches what it needs itself, so the array is complete after one barrier
    if ((needs & 4) && tid == nthr - 1) {
        int v;
        if (have_left) v = have_top ? *px_at<pixel>(pl, ex - 1, ey - 1) : *px_at<pixel>(pl, ex - 1, ey);
        else v = have_top ? *px_at<pixel>(pl, ex, ey - 1) : (1 << bitdepth) >> 1;
        if (c.u.ipred.have & 4) {
            // corner filter of z2 (ipred_tmpl.c): 5/6/5 taps over left[0], top-left, top[0] as the edge array holds them
            int l0, t0;
            if (needs & 1) l0 = have_left ? *px_at<pixel>(pl, ex - 1, ey + imin(0, c.u.ipred.n_left - 1))
                                          : (have_top ? *px_at<pixel>(pl, ex, ey - 1) : ((1 << bitdepth) >> 1) + 1);
            else l0 = 0;
            if (needs & 2) t0 = have_top ? *px_at<pixel>(pl, ex + imin(0, c.u.ipred.n_top - 1), ey - 1)
                                         : (have_left ? *px_at<pixel>(pl, ex - 1, ey) : ((1 << bitdepth) >> 1) - 1);
            else t0 = 0;
            v = ((l0 + t0) * 5 + v * 6 + 8) >> 4;
        }
        E[0] = (uint16_t)v;
    }
    PROF_MARK(23);
    CTA_SYNC();
    PROF_MARK(24);
}

EAV1_DEV int filter_strength(int wh, int angle, int is_sm) {
    if (is_sm) {
        if (wh <= 8) { if (angle >= 64) return 2; if (angle >= 40) return 1; }
        else if (wh <= 16) { if (angle >= 48) return 2; if (angle >= 20) return 1; }
        else if (wh <= 24) { if (angle >= 4) return 3; }
        else return 3;
    } else {
        if (wh <= 8) { if (angle >= 56) return 1; }
        else if (wh <= 16) { if (angle >= 40) return 1; }
        else if (wh <= 24) { if (angle >= 32) return 3; if (angle >= 16) return 2; if (angle >= 8) return 1; }
        else if (wh <= 32) { if (angle >= 32) return 3; if (angle >= 4) return 2; return 1; }
        else return 3;
    }
    return 0;
}
EAV1_DEV int use_upsample(int wh, int angle, int is_sm) { return angle < 40 && wh <= (16 >> is_sm); }

// out[i], i in [0,sz): 5-tap smoothing of in[clip(i-2..i+2, from, to-1)] for lim_from <= i < lim_to
EAV1_DEV void edge_filter(uint16_t *out, int sz, int lim_from, int lim_to, const uint16_t *in,
                          int from, int to, int strength, int tid, int nthr)
{
    const uint8_t k[3][5] = { { 0, 4, 8, 4, 0 }, { 0, 5, 6, 5, 0 }, { 2, 4, 4, 4, 2 } };
    for (int i = tid; i < sz; i += nthr) {
        if (i < imin(sz, lim_from) || i >= imin(lim_to, sz)) {
            out[i] = in[iclip(i, from, to - 1)];
        } else {
            int s = 0;
            for (int j = 0; j < 5; j++) s += in[iclip(i - 2 + j, from, to - 1)] * k[strength - 1][j];
            out[i] = (uint16_t)((s + 8) >> 4);
        }
    }
}

// out[0 .. 2*hsz-2]
EAV1_DEV void edge_upsample(uint16_t *out, int hsz, const uint16_t *in, int from, int to,
                            int bitdepth_max, int tid, int nthr)
{
    for (int i = tid; i < hsz; i += nthr) {
        out[i * 2] = in[iclip(i, from, to - 1)];
        if (i < hsz - 1) {
            const int s = -in[iclip(i - 1, from, to - 1)] + 9 * in[iclip(i, from, to - 1)] +
                          9 * in[iclip(i + 1, from, to - 1)] - in[iclip(i + 2, from, to - 1)];
            out[i * 2 + 1] = (uint16_t)iclip((s + 8) >> 4, 0, bitdepth_max);
        }
    }
}

// four horizontally adjacent samples in one store (x is a multiple of 4, rows are 4-sample aligned)
template <typename pixel>
EAV1_DEV void store4(pixel *q, int v0, int v1, int v2, int v3) {
    if (sizeof(pixel) == 1)
        *reinterpret_cast<uint32_t *>(q) = (uint32_t)v0 | (uint32_t)v1 << 8 | (uint32_t)v2 << 16 | (uint32_t)v3 << 24;
    else {
        uint2 o; o.x = (uint32_t)v0 | (uint32_t)v1 << 16; o.y = (uint32_t)v2 | (uint32_t)v3 << 16;
        *reinterpret_cast<uint2 *>(q) = o;
    }
}

// the same with the residual of the transform block added (row-major int16, `r` points at this group)
// Residuals fetched before the edge was gathered: the thread's first two groups of four samples (i = tid and
// i = tid + nthr), whose latency is long gone when the prediction is ready. Two groups cover a 32x32 block on a
// 128-thread group and a 16x16 block on a warp; further groups are loaded where they are used.
struct ResPre { uint2 a, b; int tid, nthr; };

template <typename pixel>
EAV1_DEV void store4r(pixel *q, int v0, int v1, int v2, int v3, const int16_t *r, int bitdepth_max,
                      int i, const ResPre &pre) {
    if (r) {
        const uint2 rr = i == pre.tid ? pre.a : i == pre.tid + pre.nthr ? pre.b : *reinterpret_cast<const uint2 *>(r);
        v0 = iclip(v0 + (int16_t)(rr.x & 0xffff), 0, bitdepth_max); v1 = iclip(v1 + (int16_t)(rr.x >> 16), 0, bitdepth_max);
        v2 = iclip(v2 + (int16_t)(rr.y & 0xffff), 0, bitdepth_max); v3 = iclip(v3 + (int16_t)(rr.y >> 16), 0, bitdepth_max);
    }
    store4<pixel>(q, v0, v1, v2, v3);
}

// sum of p[0], p[step], ... (n terms), every thread gets the total: each warp reduces on its own
EAV1_DEV unsigned edge_sum(const uint16_t *p, int step, int n, int tid, int nthr) {
#if defined(__CUDA_ARCH__)
    unsigned s = 0;
    if (nthr >= 32) {
        for (int i = tid & 31; i < n; i += 32) s += p[i * step];
        for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        return s;
    }
#endif
    unsigned t = 0;
    for (int i = 0; i < n; i++) t += p[i * step];
    return t;
}

// E[1..w] + E[-1..-h] in one pass (one reduction instead of two)
EAV1_DEV unsigned edge_sum2(const uint16_t *E, int w, int h, int tid, int nthr) {
#if defined(__CUDA_ARCH__)
    if (nthr >= 32) {
        unsigned s = 0;
        for (int i = tid & 31; i < w + h; i += 32) s += i < h ? E[-1 - i] : E[1 + i - h];
        for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        return s;
    }
#endif
    unsigned t = 0;
    for (int i = 0; i < w; i++) t += E[1 + i];
    for (int i = 0; i < h; i++) t += E[-1 - i];
    return t;
}

// Predict a w x h block into dst (stride in pixels) from the edge array (edge[EDGE_C] = top-left).
// edge2 is scratch of the same size.
// With `res` (row-major int16 residual of the same block) the sum prediction + residual is written in
// one go; returns whether it did so. `tile` is shared scratch for a 32x32 block (filter-intra).
template <typename pixel>
EAV1_DEV bool ipred_block(const Eav1Tables *tab, pixel *dst, int stride, uint16_t *edge, uint16_t *edge2,
                          int mode, int w, int h, int angle_in, int max_w, int max_h,
                          int bitdepth_max, int tid, int nthr, uint16_t *tile, const int16_t *res = nullptr,
                          const ResPre &res_pre = ResPre())
{
    const uint16_t *E = edge + EDGE_C;
    // w and h are powers of two >= 4: samples are handled four to a thread, i4 -> (y, x = 4*(i4 & m4))
    const int lw = ctz32((unsigned)w), lw4 = lw - 2, m4 = (w >> 2) - 1, n4 = (w * h) >> 2;
    PROF_BEGIN();
    PROF_MARK(25);
    switch (mode) {
    case IP_DC: case IP_LEFT_DC: case IP_TOP_DC: case IP_DC_128: {
        int dc;
        if (mode == IP_DC_128) dc = (bitdepth_max + 1) >> 1;
        else if (mode == IP_TOP_DC) dc = ((w >> 1) + edge_sum(&E[1], 1, w, tid, nthr)) >> lw;
        else if (mode == IP_LEFT_DC) dc = ((h >> 1) + edge_sum(&E[-1], -1, h, tid, nthr)) >> ctz32(h);
        else {
            unsigned s = ((w + h) >> 1) + edge_sum2(E, w, h, tid, nthr);
            s >>= ctz32(w + h);
            if (w != h) {
                const bool quarter = w > h * 2 || h > w * 2;
                if (sizeof(pixel) == 1) { s *= quarter ? 0x3334 : 0x5556; s >>= 16; }
                else { s *= quarter ? 0x6667 : 0xAAAB; s >>= 17; }
            }
            dc = s;
        }
        for (int i = tid; i < n4; i += nthr) store4r<pixel>(dst + (i >> lw4) * stride + ((i & m4) << 2), dc, dc, dc, dc, res ? res + (i << 2) : nullptr, bitdepth_max, i, res_pre);
        break;
    }
    case IP_VERT:
        for (int i = tid; i < n4; i += nthr) {
            const int x = (i & m4) << 2;
            store4r<pixel>(dst + (i >> lw4) * stride + x, E[1 + x], E[2 + x], E[3 + x], E[4 + x], res ? res + (i << 2) : nullptr, bitdepth_max, i, res_pre);
        }
        break;
    case IP_HOR:
        for (int i = tid; i < n4; i += nthr) {
            const int y = i >> lw4, v = E[-(1 + y)];
            store4r<pixel>(dst + y * stride + ((i & m4) << 2), v, v, v, v, res ? res + (i << 2) : nullptr, bitdepth_max, i, res_pre);
        }
        break;
    case IP_PAETH: {
        const int tl = E[0];
        for (int i = tid; i < n4; i += nthr) {
            const int y = i >> lw4, x = (i & m4) << 2;
            const int l = E[-(1 + y)];
            int v[4];
#pragma unroll
            for (int k = 0; k < 4; k++) {
                const int t = E[1 + x + k];
                const int base = l + t - tl;
                const int ld = iabs(l - base), td = iabs(t - base), tld = iabs(tl - base);
                v[k] = ld <= td && ld <= tld ? l : td <= tld ? t : tl;
            }
            store4r<pixel>(dst + y * stride + x, v[0], v[1], v[2], v[3], res ? res + (i << 2) : nullptr, bitdepth_max, i, res_pre);
        }
        break;
    }
    case IP_SMOOTH: case IP_SMOOTH_V: case IP_SMOOTH_H: {
        const uint8_t *wh_ = &tab->sm_weights[w], *wv = &tab->sm_weights[h];
        const int right = E[w], bottom = E[-h];
        for (int i = tid; i < n4; i += nthr) {
            const int y = i >> lw4, x0 = (i & m4) << 2;
            int v[4];
#pragma unroll
            for (int k = 0; k < 4; k++) {
                const int x = x0 + k;
                if (mode == IP_SMOOTH)
                    v[k] = (wv[y] * E[1 + x] + (256 - wv[y]) * bottom + wh_[x] * E[-(1 + y)] + (256 - wh_[x]) * right + 256) >> 9;
                else if (mode == IP_SMOOTH_V)
                    v[k] = (wv[y] * E[1 + x] + (256 - wv[y]) * bottom + 128) >> 8;
                else
                    v[k] = (wh_[x] * E[-(1 + y)] + (256 - wh_[x]) * right + 128) >> 8;
            }
            store4r<pixel>(dst + y * stride + x0, v[0], v[1], v[2], v[3], res ? res + (i << 2) : nullptr, bitdepth_max, i, res_pre);
        }
        break;
    }
    case IP_Z1: {
        const int is_sm = (angle_in >> 9) & 1, filt = angle_in >> 10;
        const int angle = angle_in & 511;
        int dx = tab->dr_intra_derivative[angle >> 1];
        const int up = filt ? use_upsample(w + h, 90 - angle, is_sm) : 0;
        const uint16_t *top;
        int max_base_x;
        if (up) {
            edge_upsample(edge2, w + h, &E[1], -1, w + imin(w, h), bitdepth_max, tid, nthr);
            top = edge2; max_base_x = 2 * (w + h) - 2; dx <<= 1;
        } else {
            const int fs = filt ? filter_strength(w + h, 90 - angle, is_sm) : 0;
            if (fs) {
                edge_filter(edge2, w + h, 0, w + h, &E[1], -1, w + imin(w, h), fs, tid, nthr);
                top = edge2; max_base_x = w + h - 1;
            } else { top = &E[1]; max_base_x = w + imin(w, h) - 1; }
        }
        CTA_SYNC();
        const int inc = 1 + up;
        for (int i = tid; i < n4; i += nthr) {
            const int y = i >> lw4, x0 = (i & m4) << 2;
            const int xpos = dx * (y + 1), frac = xpos & 0x3E;
            int v[4];
#pragma unroll
            for (int k = 0; k < 4; k++) {
                const int base = (xpos >> 6) + (x0 + k) * inc;
                if (base < max_base_x) v[k] = (top[base] * (64 - frac) + top[base + 1] * frac + 32) >> 6;
                else v[k] = top[max_base_x];
            }
            store4r<pixel>(dst + y * stride + x0, v[0], v[1], v[2], v[3], res ? res + (i << 2) : nullptr, bitdepth_max, i, res_pre);
        }
        break;
    }
    case IP_Z2: {
        const int is_sm = (angle_in >> 9) & 1, filt = angle_in >> 10;
        const int angle = angle_in & 511;
        int dy = tab->dr_intra_derivative[(angle - 90) >> 1];
        int dx = tab->dr_intra_derivative[(180 - angle) >> 1];
        const int up_left = filt ? use_upsample(w + h, 180 - angle, is_sm) : 0;
        const int up_above = filt ? use_upsample(w + h, angle - 90, is_sm) : 0;
        uint16_t *T = edge2 + EDGE_C; // new top-left
        if (up_above) {
            edge_upsample(T, w + 1, E, 0, w + 1, bitdepth_max, tid, nthr);
            dx <<= 1;
        } else {
            const int fs = filt ? filter_strength(w + h, angle - 90, is_sm) : 0;
            if (fs) edge_filter(&T[1], w, 0, max_w, &E[1], -1, w, fs, tid, nthr);
            else for (int i = tid; i < w; i += nthr) T[1 + i] = E[1 + i];
        }
        if (up_left) {
            edge_upsample(&T[-h * 2], h + 1, &E[-h], 0, h + 1, bitdepth_max, tid, nthr);
            dy <<= 1;
        } else {
            const int fs = filt ? filter_strength(w + h, 180 - angle, is_sm) : 0;
            if (fs) edge_filter(&T[-h], h, h - max_h, h, &E[-h], 0, h + 1, fs, tid, nthr);
            else for (int i = tid; i < h; i += nthr) T[-h + i] = E[-h + i];
        }
        CTA_SYNC();
        if (tid == 0) T[0] = E[0];
        CTA_SYNC();
        const int inc_x = 1 + up_above;
        const uint16_t *left = &T[-(1 + up_left)];
        for (int i = tid; i < n4; i += nthr) {
            const int y = i >> lw4, x0 = (i & m4) << 2;
            const int xpos = ((1 + up_above) << 6) - dx * (y + 1);
            int o[4];
#pragma unroll
            for (int k = 0; k < 4; k++) {
                const int x = x0 + k;
                const int base_x = (xpos >> 6) + x * inc_x;
                int v;
                if (base_x >= 0) {
                    const int fx = xpos & 0x3E;
                    v = T[base_x] * (64 - fx) + T[base_x + 1] * fx;
                } else {
                    const int ypos = (y << (6 + up_left)) - dy * (x + 1);
                    const int base_y = ypos >> 6, fy = ypos & 0x3E;
                    v = left[-base_y] * (64 - fy) + left[-(base_y + 1)] * fy;
                }
                o[k] = (v + 32) >> 6;
            }
            store4r<pixel>(dst + y * stride + x0, o[0], o[1], o[2], o[3], res ? res + (i << 2) : nullptr, bitdepth_max, i, res_pre);
        }
        break;
    }
    case IP_Z3: {
        const int is_sm = (angle_in >> 9) & 1, filt = angle_in >> 10;
        const int angle = angle_in & 511;
        int dy = tab->dr_intra_derivative[(270 - angle) >> 1];
        const int up = filt ? use_upsample(w + h, angle - 180, is_sm) : 0;
        const uint16_t *left;
        int max_base_y;
        if (up) {
            edge_upsample(edge2, w + h, &E[-(w + h)], imax(w - h, 0), w + h + 1, bitdepth_max, tid, nthr);
            left = &edge2[2 * (w + h) - 2]; max_base_y = 2 * (w + h) - 2; dy <<= 1;
        } else {
            const int fs = filt ? filter_strength(w + h, angle - 180, is_sm) : 0;
            if (fs) {
                edge_filter(edge2, w + h, 0, w + h, &E[-(w + h)], imax(w - h, 0), w + h + 1, fs, tid, nthr);
                left = &edge2[w + h - 1]; max_base_y = w + h - 1;
            } else { left = &E[-1]; max_base_y = h + imin(w, h) - 1; }
        }
        CTA_SYNC();
        const int inc = 1 + up;
        for (int i = tid; i < n4; i += nthr) {
            const int y = i >> lw4, x0 = (i & m4) << 2;
            int v[4];
#pragma unroll
            for (int k = 0; k < 4; k++) {
                const int ypos = dy * (x0 + k + 1), frac = ypos & 0x3E;
                const int base = (ypos >> 6) + y * inc;
                if (base < max_base_y) v[k] = (left[-base] * (64 - frac) + left[-(base + 1)] * frac + 32) >> 6;
                else v[k] = left[-max_base_y];
            }
            store4r<pixel>(dst + y * stride + x0, v[0], v[1], v[2], v[3], res ? res + (i << 2) : nullptr, bitdepth_max, i, res_pre);
        }
        break;
    }
    default: { // IP_FILTER: 4x2 cells, cell (cx,cy) needs (cx-1,cy), (cx,cy-1), (cx-1,cy-1)
        // The cells predict from each other, so the block is built in shared memory (one
        // anti-diagonal of cells per step) and written out once, residual included.
        const int8_t (*taps)[8] = tab->filter_intra_taps[angle_in & 511];
        const int cw = w >> 2, ch = h >> 1;
        uint16_t *T = tile;   // w x h, stride w (filter-intra blocks are at most 32x32)
        // the 8 x 7 taps as sixteen words (tap 7 of every row is zero padding)
        uint32_t tw[16];
        if (tid < ch || nthr == 1) {
            const uint32_t *tp = reinterpret_cast<const uint32_t *>(&taps[0][0]);
#pragma unroll
            for (int k = 0; k < 16; k++) tw[k] = tp[k];
        }
        for (int d = 0; d < cw + ch - 1; d++) {
            for (int cy = tid; cy < ch; cy += nthr) {
                const int cx = d - cy;
                if (cx < 0 || cx >= cw) continue;
                const int x0 = cx * 4, y0 = cy * 2;
                int p[7];
                if (cy == 0) { p[0] = E[x0]; for (int k = 0; k < 4; k++) p[1 + k] = E[1 + x0 + k]; }
                else {
                    p[0] = x0 ? T[(y0 - 1) * w + x0 - 1] : E[-y0];
                    for (int k = 0; k < 4; k++) p[1 + k] = T[(y0 - 1) * w + x0 + k];
                }
                if (cx == 0) { p[5] = E[-(1 + y0)]; p[6] = E[-(2 + y0)]; }
                else { p[5] = T[y0 * w + x0 - 1]; p[6] = T[(y0 + 1) * w + x0 - 1]; }
                if (sizeof(pixel) == 1) {
                    // 8-bit samples: seven taps = two byte dot products
                    const uint32_t P0 = (uint32_t)p[0] | (uint32_t)p[1] << 8 | (uint32_t)p[2] << 16 | (uint32_t)p[3] << 24;
                    const uint32_t P1 = (uint32_t)p[4] | (uint32_t)p[5] << 8 | (uint32_t)p[6] << 16;
#pragma unroll
                    for (int k = 0; k < 8; k++) {
                        const int acc = dot4_us(P0, tw[2 * k], dot4_us(P1, tw[2 * k + 1], 0));
                        T[(y0 + (k >> 2)) * w + x0 + (k & 3)] = (uint16_t)iclip((acc + 8) >> 4, 0, bitdepth_max);
                    }
                } else {
#pragma unroll
                    for (int k = 0; k < 8; k++) {
                        int acc = 0;
#pragma unroll
                        for (int j = 0; j < 7; j++) acc += (int)(int8_t)((tw[2 * k + (j >> 2)] >> (8 * (j & 3))) & 0xff) * p[j];
                        T[(y0 + (k >> 2)) * w + x0 + (k & 3)] = (uint16_t)iclip((acc + 8) >> 4, 0, bitdepth_max);
                    }
                }
            }
            // only the first ch <= 16 threads take part in the steps: a warp-level barrier is enough
#if defined(__CUDA_ARCH__)
            if (nthr > 32) { if (tid < 32) __syncwarp(); } else CTA_SYNC();
#endif
        }
        CTA_SYNC();
        for (int i = tid; i < n4; i += nthr) {
            const uint16_t *t = T + (i << 2);
            store4r<pixel>(dst + (i >> lw4) * stride + ((i & m4) << 2), t[0], t[1], t[2], t[3],
                           res ? res + (i << 2) : nullptr, bitdepth_max, i, res_pre);
        }
        break;
    }
    }
    PROF_MARK(26);
    // no barrier here: the caller's per-command barrier follows
    return res != nullptr;
}

// cfl_ac into ac[cw*ch] (shared). ypx addresses luma sample (lx,ly) of the current frame.
template <typename pixel>
EAV1_DEV void cfl_ac_block(const Plane &luma, int lx, int ly, int w_pad, int h_pad, int cw, int ch,
                           int ss_hor, int ss_ver, int16_t *ac, int32_t *red, int tid, int nthr)
{
    const int vw = cw - 4 * w_pad, vh = ch - 4 * h_pad;
    int part = 0;
    for (int i = tid; i < cw * ch; i += nthr) {
        const int y = i / cw, x = i - y * cw;
        const int sx = imin(x, vw - 1), sy = imin(y, vh - 1);
        const pixel *p = px_at<pixel>(luma, lx + (sx << ss_hor), ly + (sy << ss_ver));
        int s = p[0];
        if (ss_hor) s += p[1];
        if (ss_ver) {
            const pixel *q = px_at<pixel>(luma, lx + (sx << ss_hor), ly + (sy << ss_ver) + 1);
            s += q[0];
            if (ss_hor) s += q[1];
        }
        const int v = s << (1 + !ss_ver + !ss_hor);
        ac[i] = (int16_t)v;
        part += v;
    }
    red[tid] = part;
    CTA_SYNC();
    if (tid == 0) {
        int tot = 0;
        for (int i = 0; i < nthr; i++) tot += red[i];
        const int log2sz = ctz32(cw) + ctz32(ch);
        red[0] = (tot + ((1 << log2sz) >> 1)) >> log2sz;
    }
    CTA_SYNC();
    const int avg = red[0];
    for (int i = tid; i < cw * ch; i += nthr) ac[i] = (int16_t)(ac[i] - avg);
    CTA_SYNC();
}

template <typename pixel>
EAV1_DEV void cfl_pred_block(pixel *dst, int stride, const uint16_t *edge, int mode, int w, int h,
                             const int16_t *ac, int alpha, int bitdepth_max, int tid, int nthr)
{
    const uint16_t *E = edge + EDGE_C;
    int dc;
    if (mode == IP_DC_128) dc = (bitdepth_max + 1) >> 1;
    else if (mode == IP_TOP_DC) { unsigned s = w >> 1; for (int i = 0; i < w; i++) s += E[1 + i]; dc = s >> ctz32(w); }
    else if (mode == IP_LEFT_DC) { unsigned s = h >> 1; for (int i = 0; i < h; i++) s += E[-(1 + i)]; dc = s >> ctz32(h); }
    else {
        unsigned s = (w + h) >> 1;
        for (int i = 0; i < w; i++) s += E[1 + i];
        for (int i = 0; i < h; i++) s += E[-(1 + i)];
        s >>= ctz32(w + h);
        if (w != h) {
            const bool quarter = w > h * 2 || h > w * 2;
            if (sizeof(pixel) == 1) { s *= quarter ? 0x3334 : 0x5556; s >>= 16; }
            else { s *= quarter ? 0x6667 : 0xAAAB; s >>= 17; }
        }
        dc = s;
    }
    for (int i = tid; i < w * h; i += nthr) {
        const int y = i / w, x = i - y * w;
        const int diff = alpha * ac[i];
        dst[y * stride + x] = (pixel)iclip(dc + apply_sign((iabs(diff) + 32) >> 6, diff), 0, bitdepth_max);
    }
    CTA_SYNC();
}

template <typename pixel>
EAV1_DEV void pal_pred_block(pixel *dst, int stride, const pixel *pal, const uint8_t *idx, int w, int h, int tid, int nthr)
{
    for (int i = tid; i < (w * h) >> 1; i += nthr) {
        const int y = (2 * i) / w, x = 2 * i - y * w;
        const int v = idx[i];
        dst[y * stride + x] = pal[v & 7];
        dst[y * stride + x + 1] = pal[v >> 4];
    }
    CTA_SYNC();
}

} // namespace eav1
