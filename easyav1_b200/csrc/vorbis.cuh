// vorbis.cuh — batched Vorbis inverse MDCT + window + overlap-add.
//
// What it reproduces (ext/minivorbis/minivorbis.h): mdct_backward :7054-7148 (y[t] = sum_k X[k]
// cos(2pi/n (t + 1/2 + n/4)(k + 1/2)), unscaled), the four long/short overlap cases of
// vorbis_synthesis_blockin :9195-9300 with the half windows of _vorbis_window_get :11992, and the
// planar -> interleaved copy of src/easyav1.c:2471-2492. How it does it is ours:
//   * the transform is a DCT-IV of the n/2 spectral lines, run as one n/4-point complex FFT
//     (Stockham radix-4 in shared memory, one butterfly per thread) between two twiddle passes;
//     the reference's split-radix butterflies + bit reversal are not followed, which is why
//     parity is a tolerance (1e-5 of full scale), not bit equality;
//   * the n time samples are never materialised: by the TDAC symmetries both halves of a block
//     are signed mirror images of the n/2 DCT-IV outputs, so lapping reads them in place;
//   * one CTA walks a run of consecutive blocks of one channel, keeping the previous block's
//     DCT-IV output in shared memory, so the overlap state never goes through HBM inside a
//     run (one block per run is transformed twice, by the run that precedes it too).
#pragma once
#include "dev_common.cuh"

namespace eav1 {

struct cpx { float x, y; };

struct VorbisParams {
    int channels, n_blocks, run_len;
    int M[2];                    // half block sizes (blocksize/2) = spectral lines per block
    const uint8_t *flags;        // [n_blocks] 0 short, 1 long
    const uint32_t *spec_off;    // [n_blocks] start of block b in `spectra`, in floats, channel 0
    const uint32_t *out_off;     // [n_blocks + 1] first output sample of the segment block b completes
    const float *spectra;        // block-major: [b][channel][M[flags[b]]]
    const cpx *tw[2];            // [M/2] exp(-i pi (8k+1) / (8M)): pre- and post-twiddle of the DCT-IV
    const cpx *fftw[2];          // [M/2] exp(-2 pi i j / (M/2))
    const float *win[2];         // [M] rising half window
    const float *state_in;       // [channels][M[1]] DCT-IV output of the block before block 0 (or null)
    int state_W;                 // its size flag, -1: none (stream start / after a seek)
    float *state_out;            // [channels][M[1]] DCT-IV output of the last block of the batch
    float *pcm;                  // planar [channel][pcm_stride] or interleaved [sample][channels]
    size_t pcm_stride;
    int interleave;
};

EAV1_DEV cpx cmul(cpx a, cpx b) { return { a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x }; }
EAV1_DEV cpx cadd(cpx a, cpx b) { return { a.x + b.x, a.y + b.y }; }
EAV1_DEV cpx csub(cpx a, cpx b) { return { a.x - b.x, a.y - b.y }; }

// DCT-IV of the M lines at X into u_out (M floats); `work` is a second buffer of M floats.
// The FFT stages ping-pong between the two; the start buffer is picked so that the result of the
// last pass lands in u_out. All threads of the group must call it.
EAV1_DEV void vorbis_dct4(const VorbisParams &P, int s, const float *X, float *u_out, float *work,
                          int tid, int nthr)
{
    const int M = P.M[s], H = M >> 1;
    const cpx *tw = P.tw[s], *W = P.fftw[s];
    int lg = 0;
    while ((1 << lg) < H) lg++;
    const int n_stages = (lg & 1) + (lg >> 1);
    // after n_stages swaps the spectrum sits in the start buffer (even) or the other one (odd);
    // the post-twiddle then writes u into the buffer that does not hold the spectrum
    cpx *x = reinterpret_cast<cpx *>(n_stages & 1 ? u_out : work);
    cpx *y = reinterpret_cast<cpx *>(n_stages & 1 ? work : u_out);
    for (int n = tid; n < H; n += nthr) {
        const cpx t = { EAV1_LDG(&X[2 * n]), EAV1_LDG(&X[M - 1 - 2 * n]) };
        x[n] = cmul(t, tw[n]);
    }
    CTA_SYNC();
    int p = 1;
    if (lg & 1) {
        const int T = H >> 1;
        for (int i = tid; i < T; i += nthr) {
            const cpx u0 = x[i], u1 = x[i + T];      // p == 1: the twiddle is 1
            y[2 * i] = cadd(u0, u1);
            y[2 * i + 1] = csub(u0, u1);
        }
        CTA_SYNC();
        cpx *t = x; x = y; y = t;
        p = 2;
    }
    while (p < H) {
        const int T = H >> 2;
        for (int i = tid; i < T; i += nthr) {
            const int k = i & (p - 1), j = ((i - k) << 2) + k;
            const int m = k * (H / (4 * p));
            const cpx u0 = x[i];
            const cpx u1 = cmul(x[i + T], W[m]);
            const cpx u2 = cmul(x[i + 2 * T], W[2 * m]);
            const cpx u3 = cmul(x[i + 3 * T], W[3 * m]);
            const cpx a0 = cadd(u0, u2), a1 = csub(u0, u2), a2 = cadd(u1, u3);
            const cpx d = csub(u1, u3);
            const cpx a3 = { d.y, -d.x };            // (u1 - u3) * -i
            y[j] = cadd(a0, a2);
            y[j + p] = cadd(a1, a3);
            y[j + 2 * p] = csub(a0, a2);
            y[j + 3 * p] = csub(a1, a3);
        }
        CTA_SYNC();
        cpx *t = x; x = y; y = t;
        p <<= 2;
    }
    // x holds the FFT; the post-twiddle writes u into the other work buffer
    float *u = reinterpret_cast<float *>(y);
    for (int k = tid; k < H; k += nthr) {
        const cpx d = cmul(x[k], tw[k]);
        u[2 * k] = d.x;
        u[M - 1 - 2 * k] = -d.y;
    }
    CTA_SYNC();
}

// time samples of a block from its DCT-IV output u (M values): second half R[i] = y[M + i],
// first half L[i] = y[i]
EAV1_DEV float vorbis_R(const float *u, int M, int i) { return i < (M >> 1) ? -u[(M >> 1) - 1 - i] : -u[i - (M >> 1)]; }
EAV1_DEV float vorbis_L(const float *u, int M, int i) { return i < (M >> 1) ? u[i + (M >> 1)] : -u[3 * (M >> 1) - 1 - i]; }

EAV1_DEV void vorbis_store(const VorbisParams &P, int ch, size_t pos, float v)
{
    if (P.interleave) P.pcm[pos * P.channels + ch] = v;
    else P.pcm[(size_t)ch * P.pcm_stride + pos] = v;
}

// lap the previous block (up, size flag sp) with the current one (uc, sc): the samples that
// vorbis_synthesis_blockin makes final, (M[sp] + M[sc]) / 2 of them, starting at `pos`
EAV1_DEV void vorbis_lap(const VorbisParams &P, int ch, size_t pos, const float *up, int sp, const float *uc, int sc,
                         int tid, int nthr)
{
    const int Mp = P.M[sp], Mc = P.M[sc];
    if (Mp == Mc) {
        const float *w = P.win[sc];
        for (int i = tid; i < Mc; i += nthr)
            vorbis_store(P, ch, pos + i, vorbis_R(up, Mp, i) * EAV1_LDG(&w[Mc - 1 - i]) + vorbis_L(uc, Mc, i) * EAV1_LDG(&w[i]));
    } else if (Mp > Mc) {           // long -> short
        const float *w = P.win[0];
        const int lead = (Mp >> 1) - (Mc >> 1);
        for (int i = tid; i < lead + Mc; i += nthr) {
            float v = vorbis_R(up, Mp, i);
            if (i >= lead) v = v * EAV1_LDG(&w[Mc - 1 - (i - lead)]) + vorbis_L(uc, Mc, i - lead) * EAV1_LDG(&w[i - lead]);
            vorbis_store(P, ch, pos + i, v);
        }
    } else {                        // short -> long
        const float *w = P.win[0];
        const int lead = (Mc >> 1) - (Mp >> 1);
        for (int i = tid; i < (Mc >> 1) + (Mp >> 1); i += nthr) {
            float v = vorbis_L(uc, Mc, lead + i);
            if (i < Mp) v = vorbis_R(up, Mp, i) * EAV1_LDG(&w[Mp - 1 - i]) + v * EAV1_LDG(&w[i]);
            vorbis_store(P, ch, pos + i, v);
        }
    }
}

// one run: blocks [run*run_len, ...) of channel ch; smem = 3 * M[1] floats
EAV1_DEV void vorbis_run(const VorbisParams &P, int run, int ch, float *smem, int tid, int nthr)
{
    const int M1 = P.M[1];
    float *buf[3] = { smem, smem + M1, smem + 2 * M1 };
    const int b0 = run * P.run_len, b1 = imin(b0 + P.run_len, P.n_blocks);
    int sp = -1;                   // size flag of the block whose DCT-IV output sits in buf[0]
    if (b0 > 0) {
        sp = P.flags[b0 - 1];
        vorbis_dct4(P, sp, P.spectra + (size_t)P.spec_off[b0 - 1] + (size_t)ch * P.M[sp], buf[0], buf[1], tid, nthr);
    } else if (P.state_W >= 0) {
        sp = P.state_W;
        for (int i = tid; i < P.M[sp]; i += nthr) buf[0][i] = P.state_in[(size_t)ch * M1 + i];
        CTA_SYNC();
    }
    for (int b = b0; b < b1; b++) {
        const int sc = P.flags[b];
        vorbis_dct4(P, sc, P.spectra + (size_t)P.spec_off[b] + (size_t)ch * P.M[sc], buf[1], buf[2], tid, nthr);
        if (sp >= 0) vorbis_lap(P, ch, P.out_off[b], buf[0], sp, buf[1], sc, tid, nthr);
        CTA_SYNC();
        float *t = buf[0]; buf[0] = buf[1]; buf[1] = t;
        sp = sc;
    }
    if (b1 == P.n_blocks && sp >= 0 && P.state_out)
        for (int i = tid; i < P.M[sp]; i += nthr) P.state_out[(size_t)ch * M1 + i] = buf[0][i];
}

} // namespace eav1
