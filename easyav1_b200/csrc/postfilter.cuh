// postfilter.cuh — whole-frame in-loop filters: deblocking (K4/D3), CDEF (K5/D5) and loop
// restoration (K6/D6) of SURVEY.md §8a, as device code shared by the kernels in engine.cu.
//
// Unlike the reference, which filters in place superblock row by superblock row and therefore
// needs line/column backups (cdef_line, lr_lpf_line, lr_bak, pre_lr_border), every stage here
// reads one resident frame and writes another:
//     recon+deblock -> A        CDEF: A -> B        LR: (B, stripe borders from A) -> P
// so each output sample only ever sees the pre-stage samples the specification prescribes.
#pragma once
#include "dev_common.cuh"

namespace eav1 {

struct PfParams {
    Plane a[3];              // deblocked frame (CDEF input, LR stripe-border rows)
    Plane b[3];              // CDEF output / LR input
    Plane p[3];              // final picture
    int bitdepth_max;
    int ss_hor, ss_ver, layout;
    int w, h;                // visible luma size
    int bw4, bh4;
    // deblock
    const uint8_t *level;    // copy of f->lf.level
    int b4_stride;
    uint8_t lut_e[64], lut_i[64];
    // cdef
    const uint8_t *lfmask;   // Av1Filter array (1348 bytes each)
    int sb128w;
    int cdef_damping;
    uint8_t cdef_y[8], cdef_uv[8];
    int cdef_mirror;         // CDEF stores every filtered sample in p[] as well: loop restoration follows, and its output
                             // picture then starts as the copy of b[] it needs without a second whole-picture copy
    const Eav1Tables *tab;
};

// ------------------------------------------------------------------------------- deblock
// Behaviour: loop_filter() ext/dav1d/src/loopfilter_tmpl.c:38-170 and the four sb128 wrappers
// :172-246. One thread filters lines [line0, line1) of one 4-sample edge unit (the kernel gives
// every line its own thread: four times the parallelism, and for row edges neighbouring threads
// touch neighbouring samples; the host emulation runs all four in one call).
template <typename pixel>
EAV1_DEV void lf_edge4(pixel *dst, int E, int I, int H, ptrdiff_t stridea, ptrdiff_t strideb, int wd, int bitdepth_max,
                       int line0 = 0, int line1 = 4)
{
    const int bd8 = (32 - clz32((unsigned)bitdepth_max)) - 8;
    const int F = 1 << bd8;
    E <<= bd8; I <<= bd8; H <<= bd8;
    const int dmin = -128 * (1 << bd8), dmax = 128 * (1 << bd8) - 1;
    dst += stridea * line0;
    for (int i = line0; i < line1; i++, dst += stridea) {
        const int p1 = dst[strideb * -2], p0 = dst[strideb * -1];
        const int q0 = dst[0], q1 = dst[strideb];
        int p2 = 0, p3 = 0, q2 = 0, q3 = 0;
        bool fm = iabs(p1 - p0) <= I && iabs(q1 - q0) <= I && iabs(p0 - q0) * 2 + (iabs(p1 - q1) >> 1) <= E;
        if (wd > 4) {
            p2 = dst[strideb * -3]; q2 = dst[strideb * 2];
            fm = fm && iabs(p2 - p1) <= I && iabs(q2 - q1) <= I;
            if (wd > 6) {
                p3 = dst[strideb * -4]; q3 = dst[strideb * 3];
                fm = fm && iabs(p3 - p2) <= I && iabs(q3 - q2) <= I;
            }
        }
        if (!fm) continue;
        bool flat8out = false, flat8in = false;
        int p6 = 0, p5 = 0, p4 = 0, q4 = 0, q5 = 0, q6 = 0;
        if (wd >= 16) {
            p6 = dst[strideb * -7]; p5 = dst[strideb * -6]; p4 = dst[strideb * -5];
            q4 = dst[strideb * 4]; q5 = dst[strideb * 5]; q6 = dst[strideb * 6];
            flat8out = iabs(p6 - p0) <= F && iabs(p5 - p0) <= F && iabs(p4 - p0) <= F &&
                       iabs(q4 - q0) <= F && iabs(q5 - q0) <= F && iabs(q6 - q0) <= F;
        }
        if (wd >= 6)
            flat8in = iabs(p2 - p0) <= F && iabs(p1 - p0) <= F && iabs(q1 - q0) <= F && iabs(q2 - q0) <= F;
        if (wd >= 8)
            flat8in = flat8in && iabs(p3 - p0) <= F && iabs(q3 - q0) <= F;
        if (wd >= 16 && flat8out && flat8in) {
            // 13-tap: out[k] = (sum of a sliding 13-sample window with edge replication, centre x2 ... ) >> 4
            const int v[14] = { p6, p5, p4, p3, p2, p1, p0, q0, q1, q2, q3, q4, q5, q6 };
            int out[12];
            for (int k = 0; k < 12; k++) {
                // output sample v[k+1]; taps span v[k+1-6 .. k+1+6] clamped, plus centre and its
                // two neighbours once more (weights 1..1,2,2,2,1..1 collapse to this form)
                int s = 8;
                for (int t = -6; t <= 6; t++) s += v[iclip(k + 1 + t, 0, 13)];
                s += v[k + 1] + v[iclip(k, 0, 13)] + v[iclip(k + 2, 0, 13)];
                out[k] = s >> 4;
            }
            for (int k = 0; k < 12; k++) dst[strideb * (k - 6)] = (pixel)out[k];
        } else if (wd >= 8 && flat8in) {
            dst[strideb * -3] = (pixel)((p3 + p3 + p3 + 2 * p2 + p1 + p0 + q0 + 4) >> 3);
            dst[strideb * -2] = (pixel)((p3 + p3 + p2 + 2 * p1 + p0 + q0 + q1 + 4) >> 3);
            dst[strideb * -1] = (pixel)((p3 + p2 + p1 + 2 * p0 + q0 + q1 + q2 + 4) >> 3);
            dst[0]            = (pixel)((p2 + p1 + p0 + 2 * q0 + q1 + q2 + q3 + 4) >> 3);
            dst[strideb * 1]  = (pixel)((p1 + p0 + q0 + 2 * q1 + q2 + q3 + q3 + 4) >> 3);
            dst[strideb * 2]  = (pixel)((p0 + q0 + q1 + 2 * q2 + q3 + q3 + q3 + 4) >> 3);
        } else if (wd == 6 && flat8in) {
            dst[strideb * -2] = (pixel)((p2 + 2 * p2 + 2 * p1 + 2 * p0 + q0 + 4) >> 3);
            dst[strideb * -1] = (pixel)((p2 + 2 * p1 + 2 * p0 + 2 * q0 + q1 + 4) >> 3);
            dst[0]            = (pixel)((p1 + 2 * p0 + 2 * q0 + 2 * q1 + q2 + 4) >> 3);
            dst[strideb * 1]  = (pixel)((p0 + 2 * q0 + 2 * q1 + 2 * q2 + q2 + 4) >> 3);
        } else {
            const bool hev = iabs(p1 - p0) > H || iabs(q1 - q0) > H;
            if (hev) {
                int f = iclip(p1 - q1, dmin, dmax);
                f = iclip(3 * (q0 - p0) + f, dmin, dmax);
                const int f1 = imin(f + 4, dmax) >> 3, f2 = imin(f + 3, dmax) >> 3;
                dst[strideb * -1] = (pixel)iclip(p0 + f2, 0, bitdepth_max);
                dst[0] = (pixel)iclip(q0 - f1, 0, bitdepth_max);
            } else {
                int f = iclip(3 * (q0 - p0), dmin, dmax);
                const int f1 = imin(f + 4, dmax) >> 3, f2 = imin(f + 3, dmax) >> 3;
                dst[strideb * -1] = (pixel)iclip(p0 + f2, 0, bitdepth_max);
                dst[0] = (pixel)iclip(q0 - f1, 0, bitdepth_max);
                f = (f1 + 1) >> 1;
                dst[strideb * -2] = (pixel)iclip(p1 + f, 0, bitdepth_max);
                dst[strideb * 1] = (pixel)iclip(q1 - f, 0, bitdepth_max);
            }
        }
    }
}

// The same filter on one line held in registers: a[8 + k] = q_k, a[7 - k] = p_k. Returns how many samples on each side
// of the edge it changed (0, 1, 2, 3 or 6). Arithmetic identical to lf_edge4 (loopfilter_tmpl.c:38-170).
EAV1_DEV int lf_core(int *a, int E, int I, int H, int wd, int bitdepth_max)
{
    const int bd8 = (32 - clz32((unsigned)bitdepth_max)) - 8;
    const int F = 1 << bd8;
    E <<= bd8; I <<= bd8; H <<= bd8;
    const int dmin = -128 * (1 << bd8), dmax = 128 * (1 << bd8) - 1;
    const int p1 = a[6], p0 = a[7], q0 = a[8], q1 = a[9];
    const int p2 = a[5], q2 = a[10], p3 = a[4], q3 = a[11];
    bool fm = iabs(p1 - p0) <= I && iabs(q1 - q0) <= I && iabs(p0 - q0) * 2 + (iabs(p1 - q1) >> 1) <= E;
    if (wd > 4) {
        fm = fm && iabs(p2 - p1) <= I && iabs(q2 - q1) <= I;
        if (wd > 6) fm = fm && iabs(p3 - p2) <= I && iabs(q3 - q2) <= I;
    }
    if (!fm) return 0;
    bool flat8out = false, flat8in = false;
    if (wd >= 16)
        flat8out = iabs(a[1] - p0) <= F && iabs(a[2] - p0) <= F && iabs(a[3] - p0) <= F &&
                   iabs(a[12] - q0) <= F && iabs(a[13] - q0) <= F && iabs(a[14] - q0) <= F;
    if (wd >= 6) flat8in = iabs(p2 - p0) <= F && iabs(p1 - p0) <= F && iabs(q1 - q0) <= F && iabs(q2 - q0) <= F;
    if (wd >= 8) flat8in = flat8in && iabs(p3 - p0) <= F && iabs(q3 - q0) <= F;
    if (wd >= 16 && flat8out && flat8in) {
        int v[14], out[12];
#pragma unroll
        for (int k = 0; k < 14; k++) v[k] = a[1 + k];          // p6 .. q6
#pragma unroll
        for (int k = 0; k < 12; k++) {
            int sum = 8;
#pragma unroll
            for (int t = -6; t <= 6; t++) sum += v[iclip(k + 1 + t, 0, 13)];
            sum += v[k + 1] + v[iclip(k, 0, 13)] + v[iclip(k + 2, 0, 13)];
            out[k] = sum >> 4;
        }
#pragma unroll
        for (int k = 0; k < 12; k++) a[2 + k] = out[k];         // p5 .. q5
        return 6;
    }
    if (wd >= 8 && flat8in) {
        a[5] = (p3 + p3 + p3 + 2 * p2 + p1 + p0 + q0 + 4) >> 3;
        a[6] = (p3 + p3 + p2 + 2 * p1 + p0 + q0 + q1 + 4) >> 3;
        a[7] = (p3 + p2 + p1 + 2 * p0 + q0 + q1 + q2 + 4) >> 3;
        a[8] = (p2 + p1 + p0 + 2 * q0 + q1 + q2 + q3 + 4) >> 3;
        a[9] = (p1 + p0 + q0 + 2 * q1 + q2 + q3 + q3 + 4) >> 3;
        a[10] = (p0 + q0 + q1 + 2 * q2 + q3 + q3 + q3 + 4) >> 3;
        return 3;
    }
    if (wd == 6 && flat8in) {
        a[6] = (p2 + 2 * p2 + 2 * p1 + 2 * p0 + q0 + 4) >> 3;
        a[7] = (p2 + 2 * p1 + 2 * p0 + 2 * q0 + q1 + 4) >> 3;
        a[8] = (p1 + 2 * p0 + 2 * q0 + 2 * q1 + q2 + 4) >> 3;
        a[9] = (p0 + 2 * q0 + 2 * q1 + 2 * q2 + q2 + 4) >> 3;
        return 2;
    }
    const bool hev = iabs(p1 - p0) > H || iabs(q1 - q0) > H;
    if (hev) {
        int f = iclip(p1 - q1, dmin, dmax);
        f = iclip(3 * (q0 - p0) + f, dmin, dmax);
        const int f1 = imin(f + 4, dmax) >> 3, f2 = imin(f + 3, dmax) >> 3;
        a[7] = iclip(p0 + f2, 0, bitdepth_max);
        a[8] = iclip(q0 - f1, 0, bitdepth_max);
        return 1;
    }
    int f = iclip(3 * (q0 - p0), dmin, dmax);
    const int f1 = imin(f + 4, dmax) >> 3, f2 = imin(f + 3, dmax) >> 3;
    a[7] = iclip(p0 + f2, 0, bitdepth_max);
    a[8] = iclip(q0 - f1, 0, bitdepth_max);
    f = (f1 + 1) >> 1;
    a[6] = iclip(p1 + f, 0, bitdepth_max);
    a[9] = iclip(q1 - f, 0, bitdepth_max);
    return 2;
}

// One line of a column edge (the samples lie along the row): the 8 or 16 samples around the edge come in as aligned
// vector loads (the edge position is a multiple of 4 samples) and go back as vector stores that cover only this edge's
// own half blocks — [x-2, x+2) for the 4- and 6-tap filters, [x-4, x+4) for the 8-tap, [x-8, x+8) for the 16-tap — which
// no other edge of the pass writes. The per-sample version issued 8-16 loads and up to 12 stores per line and stalled
// on the load/store queue (ncu at 3840x2160: lg_throttle 47 per issue).
// HALF samples on each side of e into a[8 - HALF .. 8 + HALF) / back from there; all indices are compile-time
template <typename pixel, int HALF>
EAV1_DEV void lf_row_load(const pixel *e, int *a)
{
#pragma unroll
    for (int q = 0; q < HALF / 2; q++) {
        if (sizeof(pixel) == 1) {
            const uint32_t v = reinterpret_cast<const uint32_t *>(e - HALF)[q];
#pragma unroll
            for (int b = 0; b < 4; b++) a[8 - HALF + 4 * q + b] = (int)((v >> (8 * b)) & 0xff);
        } else {
            const uint2 v = reinterpret_cast<const uint2 *>(e - HALF)[q];
            a[8 - HALF + 4 * q] = (int)(v.x & 0xffff); a[8 - HALF + 4 * q + 1] = (int)(v.x >> 16);
            a[8 - HALF + 4 * q + 2] = (int)(v.y & 0xffff); a[8 - HALF + 4 * q + 3] = (int)(v.y >> 16);
        }
    }
}
template <typename pixel, int HALF>
EAV1_DEV void lf_row_store(pixel *e, const int *a)
{
#pragma unroll
    for (int q = 0; q < HALF / 2; q++) {
        const int *s = &a[8 - HALF + 4 * q];
        if (sizeof(pixel) == 1)
            reinterpret_cast<uint32_t *>(e - HALF)[q] = (uint32_t)s[0] | ((uint32_t)s[1] << 8) | ((uint32_t)s[2] << 16) | ((uint32_t)s[3] << 24);
        else {
            uint2 v; v.x = (uint32_t)s[0] | ((uint32_t)s[1] << 16); v.y = (uint32_t)s[2] | ((uint32_t)s[3] << 16);
            reinterpret_cast<uint2 *>(e - HALF)[q] = v;
        }
    }
}

template <typename pixel>
EAV1_DEV void lf_line_h(pixel *e, int E, int I, int H, int wd, int bitdepth_max)
{
    int a[16];
    if (wd >= 16) lf_row_load<pixel, 8>(e, a); else lf_row_load<pixel, 4>(e, a);
    const int n = lf_core(a, E, I, H, wd, bitdepth_max);
    if (!n) return;
    if (n <= 2) {   // p1 p0 | q0 q1: two 2-sample stores
        if (sizeof(pixel) == 1) {
            *reinterpret_cast<uint16_t *>(e - 2) = (uint16_t)(a[6] | (a[7] << 8));
            *reinterpret_cast<uint16_t *>(e) = (uint16_t)(a[8] | (a[9] << 8));
        } else {
            *reinterpret_cast<uint32_t *>(e - 2) = (uint32_t)a[6] | ((uint32_t)a[7] << 16);
            *reinterpret_cast<uint32_t *>(e) = (uint32_t)a[8] | ((uint32_t)a[9] << 16);
        }
        return;
    }
    if (n == 6) lf_row_store<pixel, 8>(e, a); else lf_row_store<pixel, 4>(e, a);
}

// One 4-sample edge unit `u` (0..31) of one recorded loop_filter_sb call.
template <typename pixel>
EAV1_DEV void lf_unit(const PfParams &pp, const Eav1LfRec &r, int u, int line0 = 0, int line1 = 4)
{
    const uint32_t bit = 1u << u;
    const uint32_t any = r.vmask[0] | r.vmask[1] | r.vmask[2];
    if (!(any & bit)) return;
    const Plane &pl = pp.a[r.plane];
    const ptrdiff_t stride = pl.stride / (ptrdiff_t)sizeof(pixel);
    const uint8_t *l = pp.level + r.lvl_off + (r.dir ? (size_t)u * 4 : (size_t)u * pp.b4_stride * 4);
    int L = l[0];
    if (!L) L = r.dir ? *(l - (ptrdiff_t)pp.b4_stride * 4) : *(l - 4);
    if (!L) return;
    const int H = L >> 4, E = pp.lut_e[L], I = pp.lut_i[L];
    int wd;
    if (r.plane == 0) wd = 4 << ((r.vmask[2] & bit) ? 2 : (r.vmask[1] & bit) ? 1 : 0);
    else wd = 4 + 2 * ((r.vmask[1] & bit) ? 1 : 0);
    if (r.dir == 0) {
        pixel *dst = px_at<pixel>(pl, r.x, r.y + 4 * u);
        if (r.x >= 8) for (int i = line0; i < line1; i++) lf_line_h<pixel>(dst + stride * i, E, I, H, wd, pp.bitdepth_max);
        else lf_edge4<pixel>(dst, E, I, H, stride, 1, wd, pp.bitdepth_max, line0, line1);
    } else {
        pixel *dst = px_at<pixel>(pl, r.x + 4 * u, r.y);
        lf_edge4<pixel>(dst, E, I, H, 1, stride, wd, pp.bitdepth_max, line0, line1);
    }
}

// ------------------------------------------------------------------------------- CDEF
// Behaviour: dav1d_cdef_brow (ext/dav1d/src/cdef_apply_tmpl.c:97-308) for the per-8x8
// decisions, cdef_find_dir_c / cdef_filter_block_c (ext/dav1d/src/cdef_tmpl.c:238-315,
// :104-215) for the arithmetic. Out of place (a -> b), one 8x8 luma block (+ its chroma)
// per thread group.

struct CdefSmem {
    alignas(4) int16_t tile[3][12 * 12];  // padded inputs, INT16_MIN marks "not available"
    int32_t part[8][16];       // per direction: sums of the samples on each of its (up to 15) lines
    int32_t cost[8];
    int32_t dir, var;
};

EAV1_DEV int cdef_constrain(int diff, int threshold, int shift) {
    const int adiff = iabs(diff);
    return apply_sign(imin(adiff, imax(0, threshold - (adiff >> shift))), diff);
}

// load the (w+4)x(h+4) neighbourhood of the block at (x0,y0) into a stride-12 tile
template <typename pixel>
EAV1_DEV void cdef_load(const Plane &pl, int x0, int y0, int w, int h, int edges, int16_t *tile, int tid, int nthr)
{
    if (sizeof(pixel) == 1 && (x0 & 3) == 0) {
        // 8-bit: aligned 32-bit loads. Row ty covers columns x0-4 .. x0+w+3 in (w+8)/4 words; the first
        // word is skipped at the left frame edge (it would lie before the row), rows outside the frame
        // are never touched.
        const int nw = (w + 8) >> 2, lnw = nw == 4 ? 2 : 0;   // 4 words (w = 8) or 3 (w = 4)
        const int items = (h + 4) * nw;
        const FastDiv fd(nw);
        for (int i = tid; i < items; i += nthr) {
            const int ty = lnw ? i >> lnw : fd.div(i), wi = i - ty * nw;
            const int y = ty - 2;
            const bool row_ok = (y >= 0 || (edges & 4)) && (y < h || (edges & 8));
            // the last word holds only right neighbours: never read past the frame's last column
            const bool word_ok = row_ok && (wi > 0 || (edges & 1)) && (wi < nw - 1 || (edges & 2));
            uint32_t v = 0;
            if (word_ok) v = *reinterpret_cast<const uint32_t *>(pl.p + (ptrdiff_t)(y0 + y) * pl.stride + x0 - 4 + 4 * wi);
#pragma unroll
            for (int b = 0; b < 4; b++) {
                const int tx = 4 * wi + b - 2, x = tx - 2;
                if (tx < 0 || tx >= w + 4) continue;
                const bool ok = row_ok && (x >= 0 || (edges & 1)) && (x < w || (edges & 2));
                tile[ty * 12 + tx] = ok ? (int16_t)((v >> (8 * b)) & 0xff) : (int16_t)-32768;
            }
        }
        return;
    }
    if (sizeof(pixel) == 2 && (x0 & 1) == 0) {
        // 16-bit: two samples per aligned 32-bit load; the 2-sample halo on either side is exactly one word
        const int nw = (w + 4) >> 1;
        const FastDiv fdw(nw);
        for (int i = tid; i < (h + 4) * nw; i += nthr) {
            const int ty = fdw.div(i), wi = i - ty * nw, y = ty - 2;
            const bool ok = (y >= 0 || (edges & 4)) && (y < h || (edges & 8)) && (wi > 0 || (edges & 1)) && (wi < nw - 1 || (edges & 2));
            uint32_t v = 0x80008000u;   // two INT16_MIN markers
            if (ok) v = *reinterpret_cast<const uint32_t *>(pl.p + (ptrdiff_t)(y0 + y) * pl.stride + (ptrdiff_t)(x0 - 2 + 2 * wi) * 2);
            *reinterpret_cast<uint32_t *>(&tile[ty * 12 + 2 * wi]) = v;
        }
        return;
    }
    const FastDiv fd(w + 4);
    for (int i = tid; i < (w + 4) * (h + 4); i += nthr) {
        const int ty = fd.div(i), tx = i - ty * (w + 4);
        const int x = tx - 2, y = ty - 2;
        const bool ok = (x >= 0 || (edges & 1)) && (x < w || (edges & 2)) && (y >= 0 || (edges & 4)) && (y < h || (edges & 8));
        tile[ty * 12 + tx] = ok ? (int16_t)*px_at<pixel>(pl, x0 + x, y0 + y) : (int16_t)-32768;
    }
}

// The three windows of one block for a 32-thread group: every load of every plane is issued before the first
// shared-memory store, so the group waits for memory once (the per-plane loops above wait once per iteration: up to
// five round trips for a 10-bit 4:2:0 block). Same tile contents as cdef_load. Needs x0 % 4 == 0 (8-bit) or
// x0 % 2 == 0 (16-bit), true for the 8x8 luma grid and its subsampled chroma positions.
#ifdef __CUDACC__
template <typename pixel>
EAV1_DEV void cdef_load_group(const Plane *pl, int lx, int ly, int cx, int cy, int cw, int chh, bool chroma, int edges,
                              int16_t (*tiles)[12 * 12], int lane)
{
    constexpr bool HBD = sizeof(pixel) == 2;
    constexpr int K = HBD ? 3 : 2;                 // words per lane and plane: 72 (16-bit) or 48 (8-bit) at most
    uint32_t v[3][K];
    int at[3][K];                                  // row * 16 + word, -1: nothing to do
    const int n_planes = chroma ? 3 : 1;
#pragma unroll
    for (int p = 0; p < 3; p++) {
        const int w = p ? cw : 8, h = p ? chh : 8, x0 = p ? cx : lx, y0 = p ? cy : ly;
        const int nw = HBD ? (w + 4) >> 1 : (w + 8) >> 2;
        const FastDiv fd(nw);
#pragma unroll
        for (int k = 0; k < K; k++) {
            const int i = lane + 32 * k;
            at[p][k] = -1;
            v[p][k] = HBD ? 0x80008000u : 0u;
            if (p >= n_planes || i >= (h + 4) * nw) continue;
            const int ty = fd.div(i), wi = i - ty * nw, y = ty - 2;
            at[p][k] = ty * 16 + wi;
            const bool row_ok = (y >= 0 || (edges & 4)) && (y < h || (edges & 8));
            const bool ok = row_ok && (wi > 0 || (edges & 1)) && (wi < nw - 1 || (edges & 2));
            if (ok) v[p][k] = *reinterpret_cast<const uint32_t *>(pl[p].p + (ptrdiff_t)(y0 + y) * pl[p].stride +
                                                                   (HBD ? (ptrdiff_t)(x0 - 2 + 2 * wi) * 2 : (ptrdiff_t)(x0 - 4 + 4 * wi)));
        }
    }
#pragma unroll
    for (int p = 0; p < 3; p++) {
        const int w = p ? cw : 8, h = p ? chh : 8;
        int16_t *tile = tiles[p];
#pragma unroll
        for (int k = 0; k < K; k++) {
            if (at[p][k] < 0) continue;
            const int ty = at[p][k] >> 4, wi = at[p][k] & 15, y = ty - 2;
            if (HBD) {
                *reinterpret_cast<uint32_t *>(&tile[ty * 12 + 2 * wi]) = v[p][k];
            } else {
                const bool row_ok = (y >= 0 || (edges & 4)) && (y < h || (edges & 8));
#pragma unroll
                for (int half = 0; half < 2; half++) {
                    const int tx = 4 * wi - 2 + 2 * half;      // even: the pair is inside the tile row or outside it
                    if (tx < 0 || tx >= w + 4) continue;
                    uint32_t pair = 0;
#pragma unroll
                    for (int b = 0; b < 2; b++) {
                        const int x = tx + b - 2;
                        const bool ok = row_ok && (x >= 0 || (edges & 1)) && (x < w || (edges & 2));
                        pair |= (ok ? (v[p][k] >> (16 * half + 8 * b)) & 0xffu : 0x8000u) << (16 * b);
                    }
                    *reinterpret_cast<uint32_t *>(&tile[ty * 12 + tx]) = pair;
                }
            }
        }
    }
}
#endif

// Filters the w x h block of one plane, or of two planes with identical parameters at once (the two
// chroma planes: 2 x 16 samples of a 4:2:0 block keep a whole warp busy instead of half of it).
template <typename pixel>
EAV1_DEV void cdef_filter(const Eav1Tables *tab, const Plane &out, int x0, int y0, int w, int h, const int16_t *tile,
                          int pri, int sec, int dir, int damping, int bitdepth_max, int tid, int nthr,
                          const Plane *out2 = nullptr, const int16_t *tile2 = nullptr,
                          const Plane *mirror = nullptr, const Plane *mirror2 = nullptr)
{
    const int bd8 = (32 - clz32((unsigned)bitdepth_max)) - 8;
    const int pri_tap = 4 - ((pri >> bd8) & 1);
    const int pri_shift = pri ? imax(0, damping - ulog2(pri)) : 0;
    const int sec_shift = sec ? damping - ulog2(sec) : 0;
    // tap offsets inside the stride-12 tile, fetched once (cdef_directions rows dir+2, dir+4, dir)
    int op[2], o2[2], o3[2];
    for (int k = 0; k < 2; k++) {
        op[k] = tab->cdef_directions[dir + 2][k];
        o2[k] = tab->cdef_directions[dir + 4][k];
        o3[k] = tab->cdef_directions[dir][k];
    }
    const int lw = ulog2((unsigned)w), wh = w * h;
    for (int ii = tid; ii < (out2 ? 2 * wh : wh); ii += nthr) {
        const bool second = ii >= wh;
        const int i = second ? ii - wh : ii;
        const int y = i >> lw, x = i & (w - 1);
        const int16_t *t = (second ? tile2 : tile) + (y + 2) * 12 + x + 2;
        const int px = t[0];
        int sum = 0, mx = px, mn = px;
#pragma unroll
        for (int k = 0; k < 2; k++) {
            if (pri) {
                const int off = op[k];
                const int p0 = t[off], p1 = t[-off];
                const int tap = k ? ((pri_tap & 3) | 2) : pri_tap;
                sum += tap * cdef_constrain(p0 - px, pri, pri_shift);
                sum += tap * cdef_constrain(p1 - px, pri, pri_shift);
                if ((unsigned)p0 < (unsigned)mn) mn = p0;
                if ((unsigned)p1 < (unsigned)mn) mn = p1;
                mx = imax(imax(p0, p1), mx);
            }
            if (sec) {
                const int s0 = t[o2[k]], s1 = t[-o2[k]], s2 = t[o3[k]], s3 = t[-o3[k]];
                const int tap = 2 - k;
                sum += tap * (cdef_constrain(s0 - px, sec, sec_shift) + cdef_constrain(s1 - px, sec, sec_shift) +
                              cdef_constrain(s2 - px, sec, sec_shift) + cdef_constrain(s3 - px, sec, sec_shift));
                if ((unsigned)s0 < (unsigned)mn) mn = s0;
                if ((unsigned)s1 < (unsigned)mn) mn = s1;
                if ((unsigned)s2 < (unsigned)mn) mn = s2;
                if ((unsigned)s3 < (unsigned)mn) mn = s3;
                mx = imax(imax(imax(s0, s1), imax(s2, s3)), mx);
            }
        }
        int v = px + ((sum - (sum < 0) + 8) >> 4);
        if (pri && sec) v = iclip(v, mn, mx);
        *px_at<pixel>(second ? *out2 : out, x0 + x, y0 + y) = (pixel)v;
        if (mirror) *px_at<pixel>(second ? *mirror2 : *mirror, x0 + x, y0 + y) = (pixel)v;
    }
}

// Direction search on the 8x8 block held in tile[0] (interior at +2,+2), cdef_find_dir
// (cdef_tmpl.c:220-315). Step 1: every thread adds its samples into the line sums of all eight
// directions (shared-memory adds; 64 samples x 8 directions spread over the group). Step 2: eight
// threads turn the line sums of one direction each into its cost.
#if defined(__CUDA_ARCH__)
#define EAV1_SMEM_ADD(p, v) atomicAdd((p), (v))
#else
#define EAV1_SMEM_ADD(p, v) (*(p) += (v))
#endif
EAV1_DEV void cdef_dir_partial(const int16_t *tile, int bd8, int32_t (*part)[16], int tid, int nthr)
{
    // pairs of horizontally adjacent samples: directions whose line index uses x >> 1 (1, 3) or
    // ignores x (2) take the pair as one add
    for (int i = tid; i < 32; i += nthr) {
        const int y = i >> 2, x = (i & 3) << 1;
        const int a = (tile[(y + 2) * 12 + x + 2] >> bd8) - 128, b = (tile[(y + 2) * 12 + x + 3] >> bd8) - 128;
        EAV1_SMEM_ADD(&part[0][y + x], a);             EAV1_SMEM_ADD(&part[0][y + x + 1], b);
        EAV1_SMEM_ADD(&part[1][y + (x >> 1)], a + b);
        EAV1_SMEM_ADD(&part[2][y], a + b);
        EAV1_SMEM_ADD(&part[3][3 + y - (x >> 1)], a + b);
        EAV1_SMEM_ADD(&part[4][7 + y - x], a);         EAV1_SMEM_ADD(&part[4][6 + y - x], b);
        EAV1_SMEM_ADD(&part[5][3 - (y >> 1) + x], a);  EAV1_SMEM_ADD(&part[5][4 - (y >> 1) + x], b);
        EAV1_SMEM_ADD(&part[6][x], a);                 EAV1_SMEM_ADD(&part[6][x + 1], b);
        EAV1_SMEM_ADD(&part[7][(y >> 1) + x], a);      EAV1_SMEM_ADD(&part[7][(y >> 1) + x + 1], b);
    }
}

EAV1_DEV void cdef_dir_cost(const int32_t *part, int n, int32_t *cost)
{
    const int div_table[7] = { 840, 420, 280, 210, 168, 140, 120 };
    unsigned c = 0;
    if (n == 2 || n == 6) {
        for (int i = 0; i < 8; i++) c += part[i] * part[i];
        c *= 105;
    } else if (n == 0 || n == 4) {
        for (int i = 0; i < 7; i++) c += (part[i] * part[i] + part[14 - i] * part[14 - i]) * div_table[i];
        c += part[7] * part[7] * 105;
    } else {
        for (int m = 0; m < 5; m++) c += part[3 + m] * part[3 + m];
        c *= 105;
        for (int m = 0; m < 3; m++) c += (part[m] * part[m] + part[10 - m] * part[10 - m]) * div_table[2 * m + 1];
    }
    cost[n] = (int32_t)c;
}

// One 8x8 luma block at 4x4-unit position (bx,by), bx and by even.
template <typename pixel>
EAV1_DEV void cdef_block(const PfParams &pp, int bx, int by, bool valid, CdefSmem &sm, int tid, int nthr)
{
    if (!valid) { CTA_SYNC(); CTA_SYNC(); CTA_SYNC(); CTA_SYNC(); return; } // keep barrier count equal across groups
    const int bd8 = (32 - clz32((unsigned)pp.bitdepth_max)) - 8;
    const int has_chroma = pp.layout != 0;
    const int cw = 8 >> pp.ss_hor, chh = 8 >> pp.ss_ver;
    const int lx = bx * 4, ly = by * 4, cx = lx >> pp.ss_hor, cy = ly >> pp.ss_ver;
    // per-64x64 strength index and per-8x8 "has coefficients" bit (lf_mask.h:51-58)
    const uint8_t *flt = pp.lfmask + (size_t)((by >> 5) * pp.sb128w + (bx >> 5)) * 1348;
    const int sb64_idx = ((by & 16) >> 3) + ((bx & 16) >> 4);
    const int cdef_idx = (int8_t)flt[1280 + sb64_idx];
    const uint16_t *noskip = reinterpret_cast<const uint16_t *>(flt + 1284) + ((by & 30) >> 1) * 2;
    const uint32_t noskip_mask = ((uint32_t)noskip[1] << 16) | noskip[0];
    int y_lvl = 0, uv_lvl = 0;
    bool active = cdef_idx != -1;
    if (active) { y_lvl = pp.cdef_y[cdef_idx]; uv_lvl = pp.cdef_uv[cdef_idx]; active = y_lvl || uv_lvl; }
    if (active) active = (noskip_mask & (3u << (bx & 30))) != 0;
    // NOTE: several blocks share one CTA, so every block must pass the same CTA_SYNC()s:
    // inactive blocks are predicated, never early-returned.
    const int edges = (bx > 0 ? 1 : 0) | (bx + 2 < pp.bw4 ? 2 : 0) | (by > 0 ? 4 : 0) | (by + 2 < pp.bh4 ? 8 : 0);
    const int y_pri = (y_lvl >> 2) << bd8;
    int y_sec = y_lvl & 3; y_sec += y_sec == 3; y_sec <<= bd8;
    const int uv_pri = (uv_lvl >> 2) << bd8;
    int uv_sec = uv_lvl & 3; uv_sec += uv_sec == 3; uv_sec <<= bd8;
    const int damping = pp.cdef_damping + bd8;

    // The output picture starts as a copy of the input (made by the caller before this stage), so only filtered
    // blocks are written. A whole group (one warp, barriers are __syncwarp) with nothing to filter leaves at once.
    if (!active && nthr <= 32) return;
    EAV1_GRID_DEP_WAIT();       // the decision above used uploaded side data only; pixels from here on
    if (active) {
        for (int i = tid; i < 8 * 16; i += nthr) (&sm.part[0][0])[i] = 0;
#ifdef __CUDACC__
        if (nthr == 32) cdef_load_group<pixel>(pp.a, lx, ly, cx, cy, cw, chh, has_chroma && uv_lvl, edges, sm.tile, tid);
        else
#endif
        {
            cdef_load<pixel>(pp.a[0], lx, ly, 8, 8, edges, sm.tile[0], tid, nthr);
            if (has_chroma && uv_lvl) {
                cdef_load<pixel>(pp.a[1], cx, cy, cw, chh, edges, sm.tile[1], tid, nthr);
                cdef_load<pixel>(pp.a[2], cx, cy, cw, chh, edges, sm.tile[2], tid, nthr);
            }
        }
    }
    CTA_SYNC();
    const bool need_dir = active && (y_pri || uv_pri);
    if (need_dir) cdef_dir_partial(sm.tile[0], bd8, sm.part, tid, nthr);
    CTA_SYNC();
    if (need_dir)
        for (int n = tid; n < 8; n += nthr) cdef_dir_cost(sm.part[n], n, sm.cost);
    CTA_SYNC();
    if (need_dir && tid == 0) {
        int best = 0;
        unsigned bc = (unsigned)sm.cost[0];
        for (int n = 1; n < 8; n++) if ((unsigned)sm.cost[n] > bc) { bc = (unsigned)sm.cost[n]; best = n; }
        sm.dir = best;
        sm.var = (int32_t)((bc - (unsigned)sm.cost[best ^ 4]) >> 10);
    }
    CTA_SYNC();
    int dir = 0;
    unsigned var = 0;
    if (need_dir) { dir = sm.dir; var = (unsigned)sm.var; }
    // luma
    int pri = 0, ydir = 0;
    bool do_y = false;
    if (active) {
        if (y_pri) {
            const int i = (var >> 6) ? imin(ulog2(var >> 6), 12) : 0;
            pri = var ? (y_pri * (4 + i) + 8) >> 4 : 0;
            ydir = dir;
            do_y = pri || y_sec;
        } else do_y = y_sec != 0;
    }
    if (do_y) cdef_filter<pixel>(pp.tab, pp.b[0], lx, ly, 8, 8, sm.tile[0], pri, y_sec, ydir, damping, pp.bitdepth_max, tid, nthr,
                                 nullptr, nullptr, pp.cdef_mirror ? &pp.p[0] : nullptr);
    if (has_chroma) {
        if (active && uv_lvl) {
            const uint8_t uvd422[8] = { 7, 0, 2, 4, 5, 6, 6, 6 };
            const int uvdir = uv_pri ? (pp.layout == 2 ? uvd422[dir] : dir) : 0;
            cdef_filter<pixel>(pp.tab, pp.b[1], cx, cy, cw, chh, sm.tile[1], uv_pri, uv_sec, uvdir, damping - 1, pp.bitdepth_max, tid, nthr,
                               &pp.b[2], sm.tile[2], pp.cdef_mirror ? &pp.p[1] : nullptr, &pp.p[2]);
        }
    }
}

// ------------------------------------------------------------------------------- super-resolution
// Behaviour: resize_c, ext/dav1d/src/mc_tmpl.c:918-955 (8-tap horizontal upscale, 64 phases). The
// reference walks a row with the recurrence mx += dx; src_x += mx >> 14; mx &= 0x3fff; its state after
// x steps is the 14-bit fixed-point position mx0 + x * dx, so every output sample is independent.
struct ResizeParams {
    Plane src[3], dst[3];
    int dst_w[3], src_w[3], h[3];
    int dx[3], mx0[3];
    int n_planes, bitdepth_max;
    const Eav1Tables *tab;
};

template <typename pixel>
EAV1_DEV void resize_sample(const ResizeParams &rp, const int8_t (*filt)[8], int pl, int x, int y)
{
    const int pos = rp.mx0[pl] + x * rp.dx[pl];
    const int src_x = -1 + (pos >> 14);
    const int8_t *const F = filt[(pos & 0x3fff) >> 8];
    const pixel *const row = px_at<pixel>(rp.src[pl], 0, y);
    const int last = rp.src_w[pl] - 1;
    int sum = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) sum += F[k] * row[iclip(src_x - 3 + k, 0, last)];
    *px_at<pixel>(rp.dst[pl], x, y) = (pixel)iclip((-sum + 64) >> 7, 0, rp.bitdepth_max);
}

// ------------------------------------------------------------------------------- loop restoration
// Behaviour: wiener_c (ext/dav1d/src/looprestoration_tmpl.c:250-387) and sgr_5x5_c / sgr_3x3_c /
// sgr_mix_c (:679-1384) with sgr_calc_row_ab (:505-523) and the finish filters (:555-626).
// The reference streams rows through ring buffers; here each CTA builds the padded input of a
// <=32-column slice of one restoration-unit stripe in shared memory and evaluates the 2-D
// definitions directly. Padding rules (lr_apply_tmpl.c:36-98, lf_apply_tmpl.c:41-102): rows
// above/below the stripe are the two nearest DEBLOCKED (pre-CDEF) rows, the third replicated;
// at the frame top/bottom and left/right the edge sample is replicated.

#define LR_TW 32
#ifndef LR_SUB_ROWS
#define LR_SUB_ROWS 32     // a stripe (<= 64 rows) is filtered as two halves by two CTAs: twice the groups in flight
#endif
struct alignas(16) LrSmem {
    uint16_t in[(LR_SUB_ROWS + 6) * (LR_TW + 6)];
    union alignas(16) {
        uint16_t hor[(LR_SUB_ROWS + 6) * LR_TW];        // Wiener: horizontally filtered rows
        struct {
            int32_t a3[(LR_SUB_ROWS + 2) * (LR_TW + 2)]; uint16_t b3[(LR_SUB_ROWS + 2) * (LR_TW + 2)];
            int32_t a5[(LR_SUB_ROWS / 2 + 2) * (LR_TW + 2)]; uint16_t b5[(LR_SUB_ROWS / 2 + 2) * (LR_TW + 2)];
        } s;
    };
};

// Rows row0 .. row0+nrows-1 of the stripe (plus 3 above and below) and columns x0-3 .. x0+tw+2 of the unit. Rows that
// lie outside the stripe follow the reference's rule (2 saved pre-CDEF rows from A, the third replicated, or the
// stripe's own edge row at the frame border); rows inside it, also beyond the sub-range, are the CDEF output B.
template <typename pixel>
EAV1_DEV void lr_load(const PfParams &pp, const Eav1LrRec &r, int x0, int tw, int row0, int nrows, uint16_t *in, int tid, int nthr)
{
    const Plane &B = pp.b[r.plane], &A = pp.a[r.plane];
    const int ss_ver = r.plane ? pp.ss_ver : 0;
    const int ph = (pp.h + ss_ver) >> ss_ver;
    const int W = tw + 6, H = nrows + 6;
    // Full-width slices whose 3-sample halo exists on both sides (everything but the first / last slice of a picture
    // row) are read as aligned 32-bit words: columns x0-4 .. x0+tw+3, of which the tile keeps x0-3 .. x0+tw+2.
    if (tw == LR_TW && (x0 > 0 || (r.edges & 1)) && (x0 + tw + 4 <= r.w || (r.edges & 2))) {
        constexpr int SPW = 4 / (int)sizeof(pixel), NW = (LR_TW + 8) / SPW;
        for (int i = tid; i < NW * H; i += nthr) {
            const int ty = i / NW, wi = i - ty * NW;
            const int ry = row0 + ty - 3;
            const Plane *src = &B;
            int sy;
            if (ry < 0) {
                if (r.edges & 4) { src = &A; sy = r.y + imax(ry, -2); } else sy = r.y;
            } else if (ry >= r.h) {
                if (r.edges & 8) { src = &A; sy = imin(r.y + r.h + imin(ry - r.h, 1), ph - 1); } else sy = r.y + r.h - 1;
            } else sy = r.y + ry;
            const uint32_t v = *reinterpret_cast<const uint32_t *>(src->p + (ptrdiff_t)sy * src->stride +
                                                                    (ptrdiff_t)(r.x + x0 - 4 + SPW * wi) * (int)sizeof(pixel));
#pragma unroll
            for (int b = 0; b < SPW; b++) {
                const int tx = SPW * wi + b - 1;
                if (tx >= 0 && tx < W) in[ty * W + tx] = (uint16_t)(sizeof(pixel) == 1 ? (v >> (8 * b)) & 0xff : (v >> (16 * b)) & 0xffff);
            }
        }
        return;
    }
    const FastDiv fdW(W);
    for (int i = tid; i < W * H; i += nthr) {
        const int ty = fdW.div(i), tx = i - ty * W;
        const int ry = row0 + ty - 3;
        int cx = x0 + tx - 3;
        if (cx < 0 && !(r.edges & 1)) cx = 0;
        if (cx >= r.w && !(r.edges & 2)) cx = r.w - 1;
        const int ax = r.x + cx;
        int v;
        if (ry < 0) {
            if (r.edges & 4) v = *px_at<pixel>(A, ax, r.y + imax(ry, -2));
            else v = *px_at<pixel>(B, ax, r.y);
        } else if (ry >= r.h) {
            if (r.edges & 8) v = *px_at<pixel>(A, ax, imin(r.y + r.h + imin(ry - r.h, 1), ph - 1));
            else v = *px_at<pixel>(B, ax, r.y + r.h - 1);
        } else v = *px_at<pixel>(B, ax, r.y + ry);
        in[i] = (uint16_t)v;
    }
}

// four finished samples of one row, x a multiple of 4: one 32-bit (8-bit samples) or 64-bit store
template <typename pixel>
EAV1_DEV void lr_store4(const Plane &P, int x, int y, const int v[4])
{
    uint8_t *d = P.p + (ptrdiff_t)y * P.stride + (ptrdiff_t)x * (int)sizeof(pixel);
    if (sizeof(pixel) == 1) *reinterpret_cast<uint32_t *>(d) = (uint32_t)v[0] | ((uint32_t)v[1] << 8) | ((uint32_t)v[2] << 16) | ((uint32_t)v[3] << 24);
    else { uint2 q; q.x = (uint32_t)v[0] | ((uint32_t)v[1] << 16); q.y = (uint32_t)v[2] | ((uint32_t)v[3] << 16); *reinterpret_cast<uint2 *>(d) = q; }
}

template <typename pixel>
EAV1_DEV void lr_wiener_tile(const PfParams &pp, const Eav1LrRec &r, int x0, int tw, int row0, int nrows, LrSmem &sm, int tid, int nthr)
{
    const int bitdepth = 32 - clz32((unsigned)pp.bitdepth_max);
    const int rb_h = 3 + (bitdepth == 12) * 2, rb_v = 11 - (bitdepth == 12) * 2;
    const int clip_limit = 1 << (bitdepth + 1 + 7 - rb_h);
    const int W = tw + 6, H = nrows + 6;
    lr_load<pixel>(pp, r, x0, tw, row0, nrows, sm.in, tid, nthr);
    CTA_SYNC();
    const FastDiv fdt(tw);
    if (tw == LR_TW) {
        // four outputs per thread from ten inputs read as five aligned pairs (the tile row pitch, 38, is even)
        int f[7];
#pragma unroll
        for (int k = 0; k < 7; k++) f[k] = r.p.filter[0][k];
        for (int i = tid; i < H * (LR_TW / 4); i += nthr) {
            const int ty = i / (LR_TW / 4), x = (i - ty * (LR_TW / 4)) * 4;
            const uint32_t *row = reinterpret_cast<const uint32_t *>(sm.in + ty * W + x);
            int sv[10];
#pragma unroll
            for (int q = 0; q < 5; q++) { const uint32_t w2 = row[q]; sv[2 * q] = (int)(w2 & 0xffff); sv[2 * q + 1] = (int)(w2 >> 16); }
            uint32_t o[4];
#pragma unroll
            for (int j = 0; j < 4; j++) {
                int sum = 1 << (bitdepth + 6);
                if (sizeof(pixel) == 1) sum += sv[j + 3] * 128;
#pragma unroll
                for (int k = 0; k < 7; k++) sum += sv[j + k] * f[k];
                o[j] = (uint32_t)iclip((sum + (1 << (rb_h - 1))) >> rb_h, 0, clip_limit - 1);
            }
            uint2 q2; q2.x = o[0] | (o[1] << 16); q2.y = o[2] | (o[3] << 16);
            *reinterpret_cast<uint2 *>(&sm.hor[ty * LR_TW + x]) = q2;
        }
    } else
    for (int i = tid; i < H * tw; i += nthr) {
        const int ty = fdt.div(i), x = i - ty * tw;
        const uint16_t *row = sm.in + ty * W + x;
        int sum = 1 << (bitdepth + 6);
        if (sizeof(pixel) == 1) sum += row[3] * 128;
        for (int k = 0; k < 7; k++) sum += row[k] * r.p.filter[0][k];
        sm.hor[ty * LR_TW + x] = (uint16_t)iclip((sum + (1 << (rb_h - 1))) >> rb_h, 0, clip_limit - 1);
    }
    CTA_SYNC();
    const Plane &P = pp.p[r.plane];
    if (tw == LR_TW) {
        for (int i = tid; i < nrows * (LR_TW / 4); i += nthr) {
            const int y = i / (LR_TW / 4), x = (i - y * (LR_TW / 4)) * 4;
            int sum[4], v[4];
#pragma unroll
            for (int j = 0; j < 4; j++) sum[j] = -(1 << (bitdepth + (rb_v - 1)));
            for (int k = 0; k < 7; k++) {
                const int f = r.p.filter[1][k];
                const uint2 h4 = *reinterpret_cast<const uint2 *>(&sm.hor[(y + k) * LR_TW + x]);
                sum[0] += (int)(h4.x & 0xffff) * f; sum[1] += (int)(h4.x >> 16) * f;
                sum[2] += (int)(h4.y & 0xffff) * f; sum[3] += (int)(h4.y >> 16) * f;
            }
#pragma unroll
            for (int j = 0; j < 4; j++) v[j] = iclip((sum[j] + (1 << (rb_v - 1))) >> rb_v, 0, pp.bitdepth_max);
            lr_store4<pixel>(P, r.x + x0 + x, r.y + row0 + y, v);
        }
        CTA_SYNC();
        return;
    }
    for (int i = tid; i < nrows * tw; i += nthr) {
        const int y = fdt.div(i), x = i - y * tw;
        int sum = -(1 << (bitdepth + (rb_v - 1)));
        for (int k = 0; k < 7; k++) sum += sm.hor[(y + k) * LR_TW + x] * r.p.filter[1][k];
        *px_at<pixel>(P, r.x + x0 + x, r.y + row0 + y) = (pixel)iclip((sum + (1 << (rb_v - 1))) >> rb_v, 0, pp.bitdepth_max);
    }
    CTA_SYNC();
}

// box statistics -> (A,B) of sgr_calc_row_ab for the window of radius rad centred at tile
// position (cx,cy) (tile coordinates: sample (0,0) of the stripe slice sits at in[3*W+3]).
EAV1_DEV void sgr_ab(const PfParams &pp, const uint16_t *in, int W, int cx, int cy, int rad, int s,
                     int32_t *A, uint16_t *B)
{
    const int bd8 = (32 - clz32((unsigned)pp.bitdepth_max)) - 8;
    unsigned sum = 0, sumsq = 0;
    for (int dy = -rad; dy <= rad; dy++)
        for (int dx = -rad; dx <= rad; dx++) {
            const unsigned v = in[(cy + 3 + dy) * W + cx + 3 + dx];
            sum += v; sumsq += v * v;
        }
    const int n = rad == 1 ? 9 : 25, one_by_x = rad == 1 ? 455 : 164;
    const int a = (int)((sumsq + ((1u << (2 * bd8)) >> 1)) >> (2 * bd8));
    const int b = (int)((sum + ((1u << bd8) >> 1)) >> bd8);
    const unsigned p = (unsigned)imax(a * n - b * b, 0);
    const unsigned z = (p * (unsigned)s + (1u << 19)) >> 20;
    const unsigned x = pp.tab->sgr_x_by_x[z < 255 ? z : 255];
    *A = (int32_t)((x * sum * (unsigned)one_by_x + (1u << 11)) >> 12);
    *B = (uint16_t)x;
}

template <typename pixel>
EAV1_DEV void lr_sgr_tile(const PfParams &pp, const Eav1LrRec &r, int x0, int tw, int row0, int nrows, LrSmem &sm, int tid, int nthr)
{
    // row0 is even, so the 5x5 filter's row pairs (computed on rows -1, 1, 3, ... of the stripe) keep their parity
    const int W = tw + 6, SW = tw + 2, h = nrows;
    const bool do5 = r.kind != EAV1_LR_SGR_3X3, do3 = r.kind != EAV1_LR_SGR_5X5;
    lr_load<pixel>(pp, r, x0, tw, row0, nrows, sm.in, tid, nthr);
    CTA_SYNC();
    const FastDiv fdS(SW), fdt(tw);
    if (do3)
        for (int i = tid; i < (h + 2) * SW; i += nthr) {
            const int ty = fdS.div(i), tx = i - ty * SW;
            sgr_ab(pp, sm.in, W, tx - 1, ty - 1, 1, (int)r.p.sgr.s1, &sm.s.a3[i], &sm.s.b3[i]);
        }
    const int n5 = (h + 1) / 2 + 1; // rows -1, 1, 3, ...
    if (do5)
        for (int i = tid; i < n5 * SW; i += nthr) {
            const int ty = fdS.div(i), tx = i - ty * SW;
            sgr_ab(pp, sm.in, W, tx - 1, 2 * ty - 1, 2, (int)r.p.sgr.s0, &sm.s.a5[i], &sm.s.b5[i]);
        }
    CTA_SYNC();
    const Plane &P = pp.p[r.plane];
    // one finished sample of the stripe slice (weighted sum of the box-filter outputs around it)
    auto sample = [&](int x, int y) -> int {
        const int src = sm.in[(y + 3) * W + x + 3];
        int t3 = 0, t5 = 0;
        if (do3) {
            const int32_t *A = sm.s.a3 + (y + 1) * SW + x + 1;
            const uint16_t *Bq = sm.s.b3 + (y + 1) * SW + x + 1;
            const int a = (Bq[0] + Bq[-1] + Bq[1] + Bq[-SW] + Bq[SW]) * 4 + (Bq[-SW - 1] + Bq[SW - 1] + Bq[-SW + 1] + Bq[SW + 1]) * 3;
            const int b = (A[0] + A[-1] + A[1] + A[-SW] + A[SW]) * 4 + (A[-SW - 1] + A[SW - 1] + A[-SW + 1] + A[SW + 1]) * 3;
            t3 = (b - a * src + (1 << 8)) >> 9;
        }
        if (do5) {
            if (!(y & 1)) {
                const int32_t *A0 = sm.s.a5 + (y >> 1) * SW + x + 1, *A1 = A0 + SW;
                const uint16_t *B0 = sm.s.b5 + (y >> 1) * SW + x + 1, *B1 = B0 + SW;
                const int a = (B0[0] + B1[0]) * 6 + (B0[-1] + B1[-1] + B0[1] + B1[1]) * 5;
                const int b = (A0[0] + A1[0]) * 6 + (A0[-1] + A1[-1] + A0[1] + A1[1]) * 5;
                t5 = (b - a * src + (1 << 8)) >> 9;
            } else {
                const int32_t *A1 = sm.s.a5 + ((y + 1) >> 1) * SW + x + 1;
                const uint16_t *B1 = sm.s.b5 + ((y + 1) >> 1) * SW + x + 1;
                const int a = B1[0] * 6 + (B1[-1] + B1[1]) * 5;
                const int b = A1[0] * 6 + (A1[-1] + A1[1]) * 5;
                t5 = (b - a * src + (1 << 7)) >> 8;
            }
        }
        int v;
        if (r.kind == EAV1_LR_SGR_MIX) v = r.p.sgr.w0 * t5 + r.p.sgr.w1 * t3;
        else if (r.kind == EAV1_LR_SGR_5X5) v = r.p.sgr.w0 * t5;
        else v = r.p.sgr.w1 * t3;
        return iclip(src + ((v + (1 << 10)) >> 11), 0, pp.bitdepth_max);
    };
    if (tw == LR_TW) {
        for (int i = tid; i < h * (LR_TW / 4); i += nthr) {
            const int y = i / (LR_TW / 4), x = (i - y * (LR_TW / 4)) * 4;
            int v4[4];
#pragma unroll
            for (int j = 0; j < 4; j++) v4[j] = sample(x + j, y);
            lr_store4<pixel>(P, r.x + x0 + x, r.y + row0 + y, v4);
        }
    } else {
        for (int i = tid; i < h * tw; i += nthr) {
            const int y = fdt.div(i), x = i - y * tw;
            *px_at<pixel>(P, r.x + x0 + x, r.y + row0 + y) = (pixel)sample(x, y);
        }
    }
    CTA_SYNC();
}

template <typename pixel>
EAV1_DEV void lr_tile(const PfParams &pp, const Eav1LrRec &r, int slice, int sub, LrSmem &sm, int tid, int nthr)
{
    const int x0 = slice * LR_TW, row0 = sub * LR_SUB_ROWS;
    if (x0 >= r.w || row0 >= r.h) return; // uniform across the CTA
    const int tw = imin(LR_TW, r.w - x0), nrows = imin(LR_SUB_ROWS, r.h - row0);
    if (r.kind <= EAV1_LR_WIENER5) lr_wiener_tile<pixel>(pp, r, x0, tw, row0, nrows, sm, tid, nthr);
    else lr_sgr_tile<pixel>(pp, r, x0, tw, row0, nrows, sm, tid, nthr);
}

} // namespace eav1
