// filmgrain.cuh — film grain synthesis on the device.
//
// Behaviour follows ext/dav1d/src/filmgrain_tmpl.c (generate_grain_y_c :50-86,
// generate_grain_uv_c :88-156, fgy_32x32xn_c :169-276, fguv_32x32xn_c :278-421) and
// fg_apply_tmpl.c (generate_scaling :41-97, prep_grain :100-159, apply_grain_row :161-223);
// the organisation is ours:
//   * fg_prepare   one CTA: scaling LUTs in closed form (one entry per thread), the per-block
//                  template offsets of every 32-row band (one LFSR per band), the three 73x82
//                  grain templates. The 16-bit LFSR is linear over GF(2), so each template row
//                  starts from a state obtained by a jump matrix and rows are drawn in parallel;
//                  the auto-regressive filter runs as a skewed wavefront over the rows in shared memory.
//   * fg_strip     one CTA per 128x32 luma strip and its chroma: 4 pixels per thread, 32-bit /
//                  64-bit accesses, the templates and LUTs come through the read-only path.
#pragma once
#include "dev_common.cuh"

namespace eav1 {

enum { FG_GW = 82, FG_GH = 73, FG_SUBW = 44, FG_SUBH = 38, FG_BLK = 32, FG_STRIP = 4 };

struct FgParams {
    Eav1GrainParams g;
    Plane src[3], dst[3];
    int w, h, layout, ss_hor, ss_ver, bitdepth_max;
    int rows, cols;                 // 32x32 luma blocks
    const Eav1Tables *tab;
    int16_t *grain;                 // [3][FG_GH][FG_GW]
    uint8_t *scaling;               // [3][4096]
    uint8_t *offsets;               // [rows][cols]
};

struct FgSmem {
    int16_t buf[3][FG_GH][FG_GW];
    uint32_t jump[2][16];
    uint32_t row_state[2 * FG_GH];
};

EAV1_DEV uint32_t fg_lfsr(uint32_t r)
{
    const uint32_t bit = (r ^ (r >> 1) ^ (r >> 3) ^ (r >> 12)) & 1;
    return (r >> 1) | (bit << 15);
}
EAV1_DEV int fg_round2(int x, int shift) { return (x + ((1 << shift) >> 1)) >> shift; }

// one entry of the 8-bit scaling function (fg_apply_tmpl.c:41-81)
EAV1_DEV int fg_scale8(const uint8_t (*pt)[2], int num, int i)
{
    if (!num) return 0;
    if (i < pt[0][0]) return pt[0][1];
    if (i >= pt[num - 1][0]) return pt[num - 1][1];
    int k = 0;
    while (i >= pt[k + 1][0]) k++;
    const int bx = pt[k][0], by = pt[k][1];
    const int dx = pt[k + 1][0] - bx, dy = pt[k + 1][1] - by;
    const int delta = dy * ((0x10000 + (dx >> 1)) / dx);
    return (by + ((0x8000 + (i - bx) * delta) >> 16)) & 0xff;
}

// entry idx of the table at bit depth 8 + sx (fg_apply_tmpl.c:83-96 fills the entries between
// two 8-bit nodes by linear interpolation inside the point range only)
EAV1_DEV int fg_scale(const uint8_t (*pt)[2], int num, int idx, int sx)
{
    const int i8 = idx >> sx, n = idx & ((1 << sx) - 1);
    const int b0 = fg_scale8(pt, num, i8);
    if (!n || !num || i8 < pt[0][0] || i8 >= pt[num - 1][0]) return b0;
    const int range = fg_scale8(pt, num, i8 + 1) - b0;
    return (b0 + ((((1 << sx) >> 1) + n * range) >> sx)) & 0xff;
}

// Draw n_planes templates of W x H Gaussian samples. Plane p starts from seed[p]; row y of a
// plane starts from the state after y*W LFSR steps, found with the W-step jump matrix.
EAV1_DEV void fg_draw(FgSmem &sm, const FgParams &P, int first_plane, int n_planes, int W, int H,
                      const uint32_t seed[2], int shift, int tid, int nthr)
{
    for (int i = tid; i < 16; i += nthr) {
        uint32_t s = 1u << i;
        for (int k = 0; k < W; k++) s = fg_lfsr(s);
        sm.jump[0][i] = s;
    }
    CTA_SYNC();
    for (int p = tid; p < n_planes; p += nthr) {
        uint32_t s = seed[p];
        for (int y = 0; y < H; y++) {
            sm.row_state[p * H + y] = s;
            uint32_t nx = 0;
            for (int i = 0; i < 16; i++) if (s >> i & 1) nx ^= sm.jump[0][i];
            s = nx;
        }
    }
    CTA_SYNC();
    for (int r = tid; r < n_planes * H; r += nthr) {
        const int p = r / H, y = r - p * H;
        uint32_t s = sm.row_state[r];
        for (int x = 0; x < W; x++) {
            s = fg_lfsr(s);
            sm.buf[first_plane + p][y][x] = (int16_t)fg_round2(P.tab->gaussian_sequence[(s >> 5) & 2047], shift);
        }
    }
    CTA_SYNC();
}

#ifdef __CUDACC__
// The wavefront below with one thread per row (every row of every plane has a thread of its own), the lag a
// compile-time constant: the 2*LAG*(LAG+1) coefficients of the thread's plane sit in registers, the tap loops are
// unrolled, and the taps are summed on four accumulators. A step is then one round of independent shared-memory loads,
// a short multiply-add tree, one store and the barrier; the generic loop spent ~1200 cycles per step on 24 dependent
// iterations (constant-bank coefficient fetch, address arithmetic, loop control), 0.33 ms per frame in all.
template <int LAG>
EAV1_DEV void fg_ar_rows(FgSmem &sm, const FgParams &P, int first_plane, int n_planes, int W, int H,
                         int subx, int suby, int tid)
{
    const Eav1GrainParams &g = P.g;
    constexpr int skew = LAG + 1, NC = 2 * LAG * (LAG + 1);
    const int bd8 = 32 - clz32((unsigned)P.bitdepth_max) - 8;
    const int gmin = -(128 << bd8), gmax = (128 << bd8) - 1;
    const int rows = H - 3, colsn = W - 6;
    const int steps = colsn + (rows - 1) * skew;
    const bool mine = tid < n_planes * rows;
    const int p = mine ? tid / rows : 0, yy = mine ? tid - p * rows : 0;
    const int pl = first_plane + p, y = yy + 3;
    const int8_t *coeff = pl ? g.ar_coeffs_uv[pl - 1] : g.ar_coeffs_y;
    int c[NC + 1];
#pragma unroll
    for (int k = 0; k < NC; k++) c[k] = coeff[k];
    c[NC] = pl ? coeff[NC] : 0;      // the luma-coupling coefficient of a chroma plane
    const bool coupled = pl && g.num_y_points;
    for (int step = 0; step < steps; step++) {
        const int xx = step - yy * skew;
        if (mine && xx >= 0 && xx < colsn) {
            const int x = xx + 3;
            int acc[4] = { 0, 0, 0, 0 };
            int k = 0;
#pragma unroll
            for (int dy = -LAG; dy <= 0; dy++)
#pragma unroll
                for (int dx = -LAG; dx <= LAG; dx++) {
                    if (dy == 0 && dx >= 0) continue;
                    acc[k & 3] += c[k] * sm.buf[pl][y + dy][x + dx];
                    k++;
                }
            int sum = (acc[0] + acc[1]) + (acc[2] + acc[3]);
            if (coupled) {
                const int lx = ((x - 3) << subx) + 3, ly = ((y - 3) << suby) + 3;
                int luma = 0;
                for (int i = 0; i <= suby; i++)
                    for (int j = 0; j <= subx; j++) luma += sm.buf[0][ly + i][lx + j];
                sum += fg_round2(luma, subx + suby) * c[NC];
            }
            sm.buf[pl][y][x] = (int16_t)iclip(sm.buf[pl][y][x] + fg_round2(sum, g.ar_coeff_shift), gmin, gmax);
        }
        __syncthreads();
    }
}
#endif

// Auto-regressive filter as a wavefront: row y works on column x at step (x-3) + (y-3)*(lag+1),
// one step after the row above finished column x+lag.
EAV1_DEV void fg_ar(FgSmem &sm, const FgParams &P, int first_plane, int n_planes, int W, int H,
                    int subx, int suby, int tid, int nthr)
{
    const Eav1GrainParams &g = P.g;
    const int lag = g.ar_coeff_lag, skew = lag + 1;
    const int bd8 = 32 - clz32((unsigned)P.bitdepth_max) - 8;
    const int gmin = -(128 << bd8), gmax = (128 << bd8) - 1;
    const int rows = H - 3, colsn = W - 6;
    const int steps = colsn + (rows - 1) * skew;
#ifdef __CUDACC__
    if (n_planes * rows <= nthr) {
        switch (lag) {
        case 0: fg_ar_rows<0>(sm, P, first_plane, n_planes, W, H, subx, suby, tid); return;
        case 1: fg_ar_rows<1>(sm, P, first_plane, n_planes, W, H, subx, suby, tid); return;
        case 2: fg_ar_rows<2>(sm, P, first_plane, n_planes, W, H, subx, suby, tid); return;
        default: fg_ar_rows<3>(sm, P, first_plane, n_planes, W, H, subx, suby, tid); return;
        }
    }
#endif
    for (int step = 0; step < steps; step++) {
        for (int r = tid; r < n_planes * rows; r += nthr) {
            const int p = r / rows, yy = r - p * rows;
            const int xx = step - yy * skew;
            if (xx < 0 || xx >= colsn) continue;
            const int x = xx + 3, y = yy + 3, pl = first_plane + p;
            const int8_t *coeff = pl ? g.ar_coeffs_uv[pl - 1] : g.ar_coeffs_y;
            int sum = 0;
            for (int dy = -lag; dy <= 0; dy++)
                for (int dx = -lag; dx <= lag; dx++) {
                    if (!dx && !dy) break;
                    sum += *(coeff++) * sm.buf[pl][y + dy][x + dx];
                }
            if (pl && g.num_y_points) {
                const int lx = ((x - 3) << subx) + 3, ly = ((y - 3) << suby) + 3;
                int luma = 0;
                for (int i = 0; i <= suby; i++)
                    for (int j = 0; j <= subx; j++) luma += sm.buf[0][ly + i][lx + j];
                sum += fg_round2(luma, subx + suby) * *coeff;
            }
            sm.buf[pl][y][x] = (int16_t)iclip(sm.buf[pl][y][x] + fg_round2(sum, g.ar_coeff_shift), gmin, gmax);
        }
        CTA_SYNC();
    }
}

EAV1_DEV void fg_prepare(const FgParams &P, FgSmem &sm, int tid, int nthr)
{
    const Eav1GrainParams &g = P.g;
    const int bd8 = 32 - clz32((unsigned)P.bitdepth_max) - 8;
    const int n_scale = 256 << bd8;
    for (int i = tid; i < 3 * n_scale; i += nthr) {
        const int pl = i / n_scale, idx = i - pl * n_scale;
        P.scaling[pl * 4096 + idx] = (uint8_t)(pl ? fg_scale(g.uv_points[pl - 1], g.num_uv_points[pl - 1], idx, bd8)
                                                  : fg_scale(g.y_points, g.num_y_points, idx, bd8));
    }
    for (int row = tid; row < P.rows; row += nthr) {
        uint32_t s = g.seed;
        s ^= (uint32_t)((row * 37 + 178) & 0xff) << 8;
        s ^= (uint32_t)((row * 173 + 105) & 0xff);
        for (int col = 0; col < P.cols; col++) {
            s = fg_lfsr(s);
            P.offsets[row * P.cols + col] = (uint8_t)(s >> 8);
        }
    }
    const int shift = 4 - bd8 + g.grain_scale_shift;
    uint32_t seed[2] = { g.seed, 0 };
    fg_draw(sm, P, 0, 1, FG_GW, FG_GH, seed, shift, tid, nthr);
    fg_ar(sm, P, 0, 1, FG_GW, FG_GH, 0, 0, tid, nthr);
    if (P.layout) {
        const int W = P.ss_hor ? FG_SUBW : FG_GW, H = P.ss_ver ? FG_SUBH : FG_GH;
        seed[0] = g.seed ^ 0xb524; seed[1] = g.seed ^ 0x49d8;
        fg_draw(sm, P, 1, 2, W, H, seed, shift, tid, nthr);
        fg_ar(sm, P, 1, 2, W, H, P.ss_hor, P.ss_ver, tid, nthr);
    }
    int16_t *flat = &sm.buf[0][0][0];
    for (int i = tid; i < 3 * FG_GH * FG_GW; i += nthr) P.grain[i] = flat[i];
}

// ---- applying the templates ------------------------------------------------------------------

template <typename pixel> struct FgVec;
template <> struct FgVec<uint8_t> {
    uint32_t v;
    EAV1_DEVM void load(const uint8_t *p) { v = EAV1_LDG(reinterpret_cast<const uint32_t *>(p)); }
    EAV1_DEVM void store(uint8_t *p) const { *reinterpret_cast<uint32_t *>(p) = v; }
    EAV1_DEVM int get(int i) const { return (v >> (8 * i)) & 0xff; }
    EAV1_DEVM void set(int i, int x) { v = (v & ~(0xffu << (8 * i))) | ((uint32_t)x << (8 * i)); }
};
template <> struct FgVec<uint16_t> {
    uint2 v;
    EAV1_DEVM void load(const uint16_t *p) { v = EAV1_LDG(reinterpret_cast<const uint2 *>(p)); }
    EAV1_DEVM void store(uint16_t *p) const { *reinterpret_cast<uint2 *>(p) = v; }
    EAV1_DEVM int get(int i) const { return ((i & 2 ? v.y : v.x) >> (16 * (i & 1))) & 0xffff; }
    EAV1_DEVM void set(int i, int x) {
        uint32_t &w = i & 2 ? v.y : v.x;
        w = (w & ~(0xffffu << (16 * (i & 1)))) | ((uint32_t)x << (16 * (i & 1)));
    }
};

struct FgClip { int gmin, gmax; };

// grain of the sample (xb, yb) of block (row, bxi): template lookup at the block's pseudo-random
// offset, blended over the 2-sample (1 when subsampled) seam with the left / upper neighbours
EAV1_DEV int fg_grain_at(const FgParams &P, const int16_t *lut, int row, int bxi, int xb, int yb,
                         int sx, int sy, FgClip c)
{
    const int bw = FG_BLK >> sx, bh = FG_BLK >> sy, mx = 2 >> sx, my = 2 >> sy;
    const uint8_t *off = P.offsets + row * P.cols + bxi;
    const int o = EAV1_LDG(off);
    const int ox = 3 + mx * (3 + (o >> 4)), oy = 3 + my * (3 + (o & 15));
    int gr = EAV1_LDG(&lut[(oy + yb) * FG_GW + ox + xb]);
    const bool xov = P.g.overlap_flag && bxi > 0 && xb < mx;
    const bool yov = P.g.overlap_flag && row > 0 && yb < my;
    const int wx0 = sx ? 23 : xb ? 17 : 27, wx1 = sx ? 22 : xb ? 27 : 17;
    const int wy0 = sy ? 23 : yb ? 17 : 27, wy1 = sy ? 22 : yb ? 27 : 17;
    if (xov) {
        const int ol = EAV1_LDG(off - 1);
        const int lx = 3 + mx * (3 + (ol >> 4)), ly = 3 + my * (3 + (ol & 15));
        const int old = EAV1_LDG(&lut[(ly + yb) * FG_GW + lx + xb + bw]);
        gr = iclip(fg_round2(old * wx0 + gr * wx1, 5), c.gmin, c.gmax);
    }
    if (yov) {
        const int ot = EAV1_LDG(off - P.cols);
        const int tx = 3 + mx * (3 + (ot >> 4)), ty = 3 + my * (3 + (ot & 15));
        int top = EAV1_LDG(&lut[(ty + yb + bh) * FG_GW + tx + xb]);
        if (xov) {
            const int oc = EAV1_LDG(off - P.cols - 1);
            const int cx = 3 + mx * (3 + (oc >> 4)), cy = 3 + my * (3 + (oc & 15));
            const int old = EAV1_LDG(&lut[(cy + yb + bh) * FG_GW + cx + xb + bw]);
            top = iclip(fg_round2(old * wx0 + top * wx1, 5), c.gmin, c.gmax);
        }
        gr = iclip(fg_round2(top * wy0 + gr * wy1, 5), c.gmin, c.gmax);
    }
    return gr;
}

// one 128x32 luma strip (4 grain blocks of band `row`) and the chroma that belongs to it
template <typename pixel>
EAV1_DEV void fg_strip(const FgParams &P, int strip, int row, int tid, int nthr)
{
    const Eav1GrainParams &g = P.g;
    const int bd8 = 32 - clz32((unsigned)P.bitdepth_max) - 8;
    const FgClip gc = { -(128 << bd8), (128 << bd8) - 1 };
    const int minv = g.clip_to_restricted_range ? 16 << bd8 : 0;
    const int maxy = g.clip_to_restricted_range ? 235 << bd8 : P.bitdepth_max;
    const int maxc = g.clip_to_restricted_range ? (g.mtrx_identity ? 235 : 240) << bd8 : P.bitdepth_max;
    const int x0 = strip * FG_STRIP * FG_BLK, y0 = row * FG_BLK;
    const int bh = imin(FG_BLK, P.h - y0);
    const pixel *const luma = reinterpret_cast<const pixel *>(P.src[0].p);
    const int lstride = P.src[0].stride / (int)sizeof(pixel);

    for (int item = tid; item < bh * 32; item += nthr) {
        const int yb = item >> 5, xs = (item & 31) * 4, x = x0 + xs;
        if (x >= P.w) continue;
        FgVec<pixel> v;
        v.load(luma + (size_t)(y0 + yb) * lstride + x);
        if (g.num_y_points)
            for (int i = 0; i < 4; i++) {
                const int s = v.get(i);
                const int gr = fg_grain_at(P, P.grain, row, (x + i) >> 5, (xs + i) & 31, yb, 0, 0, gc);
                const int noise = fg_round2(EAV1_LDG(&P.scaling[s]) * gr, g.scaling_shift);
                v.set(i, iclip(s + noise, minv, maxy));
            }
        v.store(reinterpret_cast<pixel *>(P.dst[0].p) + (size_t)(y0 + yb) * (P.dst[0].stride / (int)sizeof(pixel)) + x);
    }
    if (!P.layout) return;

    const int sx = P.ss_hor, sy = P.ss_ver;
    const int cw = (P.w + sx) >> sx, bhc = (bh + sy) >> sy;
    const int x0c = x0 >> sx, y0c = y0 >> sy, per_row = 32 >> sx, bwc = FG_BLK >> sx;
    for (int pl = 1; pl < 3; pl++) {
        const bool on = g.chroma_scaling_from_luma || g.num_uv_points[pl - 1];
        const uint8_t *const scal = P.scaling + (g.chroma_scaling_from_luma ? 0 : pl * 4096);
        const int16_t *const lut = P.grain + pl * FG_GH * FG_GW;
        const pixel *const src = reinterpret_cast<const pixel *>(P.src[pl].p);
        pixel *const dst = reinterpret_cast<pixel *>(P.dst[pl].p);
        const int sstride = P.src[pl].stride / (int)sizeof(pixel), dstride = P.dst[pl].stride / (int)sizeof(pixel);
        for (int item = tid; item < bhc * per_row; item += nthr) {
            const int yb = item / per_row, xs = (item - yb * per_row) * 4, x = x0c + xs;
            if (x >= cw) continue;
            FgVec<pixel> v;
            v.load(src + (size_t)(y0c + yb) * sstride + x);
            if (on) {
                const pixel *const lrow = luma + (size_t)(y0 + (yb << sy)) * lstride;
                for (int i = 0; i < 4; i++) {
                    if (x + i >= cw) break;
                    const int s = v.get(i);
                    const int lx = (x + i) << sx;
                    int val = EAV1_LDG(&lrow[lx]);
                    if (sx) val = (val + EAV1_LDG(&lrow[imin(lx + 1, P.w - 1)]) + 1) >> 1;
                    if (!g.chroma_scaling_from_luma) {
                        const int combined = val * g.uv_luma_mult[pl - 1] + s * g.uv_mult[pl - 1];
                        val = iclip((combined >> 6) + g.uv_offset[pl - 1] * (1 << bd8), 0, P.bitdepth_max);
                    }
                    const int gr = fg_grain_at(P, lut, row, (x + i) / bwc, (xs + i) & (bwc - 1), yb, sx, sy, gc);
                    const int noise = fg_round2(EAV1_LDG(&scal[val]) * gr, g.scaling_shift);
                    v.set(i, iclip(s + noise, minv, maxc));
                }
            }
            v.store(dst + (size_t)(y0c + yb) * dstride + x);
        }
    }
}

} // namespace eav1
