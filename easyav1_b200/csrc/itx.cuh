// itx.cuh — AV1 inverse transforms (K1 of SURVEY.md §8a) as CTA-cooperative device code.
//
// Behaviour follows the reference's inv_txfm_add_c (ext/dav1d/src/itx_tmpl.c:43-120) and
// its 1-D cores (ext/dav1d/src/itx_1d.c:65-1017, WHT :1066): integer butterflies with a
// 12-bit rounding after every rotation, a clip after every add/sub stage, the rect2
// (x*181+128)>>8 pre-scale, the per-size inter-pass shift, the DC-only shortcut and the
// final (x+8)>>4. The formulation here is our own: every rotation goes through one
// primitive m2() evaluated in a multiply type wide enough to be exact (int32 for 8-bit
// coefficients, int64 for high bit-depth), which yields the reference's values without
// reproducing its overflow-avoidance rewrites (itx_1d.c:39-63 explains they are
// value-preserving).
//
// Layout in shared memory: element (row y, column x) of the tile lives at s[x*P + y] with an
// odd pitch P, so both passes (thread-per-row, then thread-per-column) are bank-conflict free,
// and the reference's column-major coefficient order (coeff[y + x*sh]) loads contiguously.
#pragma once
#include "dev_common.cuh"

namespace eav1 {

enum { T1D_DCT = 0, T1D_ADST = 1, T1D_IDENTITY = 2, T1D_FLIPADST = 3 };

// (a*ca + b*cb + 2048) >> 12, exact.
template <typename W> EAV1_DEV int m2(int a, int ca, int b, int cb) {
    return (int)(((W)a * ca + (W)b * cb + 2048) >> 12);
}
// ((a) * 181 + 128) >> 8  (the 1/sqrt2 rotation)
template <typename W> EAV1_DEV int r181(int a) { return (int)(((W)a * 181 + 128) >> 8); }

#define CL(v) iclip((v), mn, mx)

// ---- DCT: recursive even/odd decomposition -------------------------------------------
// idctN(e[]) consumes N inputs in natural order (in[k] = coefficient k) and produces N
// outputs in natural order. Input is passed as two arrays: the caller splits even/odd.

template <typename W>
EAV1_DEV void idct4(const int in[4], int out[4], int mn, int mx) {
    const int t0 = r181<W>(in[0] + in[2]);
    const int t1 = r181<W>(in[0] - in[2]);
    const int t2 = m2<W>(in[1], 1567, in[3], -3784);
    const int t3 = m2<W>(in[1], 3784, in[3], 1567);
    out[0] = CL(t0 + t3);
    out[1] = CL(t1 + t2);
    out[2] = CL(t1 - t2);
    out[3] = CL(t0 - t3);
}

template <typename W>
EAV1_DEV void idct8(const int in[8], int out[8], int mn, int mx) {
    int ev[4] = { in[0], in[2], in[4], in[6] }, e[4];
    idct4<W>(ev, e, mn, mx);
    const int a4 = m2<W>(in[1], 799, in[7], -4017);
    const int a7 = m2<W>(in[1], 4017, in[7], 799);
    const int a5 = m2<W>(in[5], 3406, in[3], -2276);
    const int a6 = m2<W>(in[5], 2276, in[3], 3406);
    const int t4 = CL(a4 + a5), b5 = CL(a4 - a5);
    const int t7 = CL(a7 + a6), b6 = CL(a7 - a6);
    const int t5 = r181<W>(b6 - b5);
    const int t6 = r181<W>(b6 + b5);
    out[0] = CL(e[0] + t7); out[7] = CL(e[0] - t7);
    out[1] = CL(e[1] + t6); out[6] = CL(e[1] - t6);
    out[2] = CL(e[2] + t5); out[5] = CL(e[2] - t5);
    out[3] = CL(e[3] + t4); out[4] = CL(e[3] - t4);
}

// odd half of the 16-point DCT: o[k] = in[2k+1], result t[0..7] = t8..t15
template <typename W>
EAV1_DEV void idct16_odd(const int o[8], int t[8], int mn, int mx) {
    const int a8  = m2<W>(o[0], 401,  o[7], -4076), a15 = m2<W>(o[0], 4076, o[7], 401);
    const int a9  = m2<W>(o[4], 3166, o[3], -2598), a14 = m2<W>(o[4], 2598, o[3], 3166);
    const int a10 = m2<W>(o[2], 1931, o[5], -3612), a13 = m2<W>(o[2], 3612, o[5], 1931);
    const int a11 = m2<W>(o[6], 3920, o[1], -1189), a12 = m2<W>(o[6], 1189, o[1], 3920);
    const int b8  = CL(a8 + a9),   b9  = CL(a8 - a9);
    const int b10 = CL(a11 - a10), b11 = CL(a11 + a10);
    const int b12 = CL(a12 + a13), b13 = CL(a12 - a13);
    const int b14 = CL(a15 - a14), b15 = CL(a15 + a14);
    const int c9  = m2<W>(b14, 1567, b9, -3784),   c14 = m2<W>(b14, 3784, b9, 1567);
    const int c10 = m2<W>(b13, -3784, b10, -1567), c13 = m2<W>(b13, 1567, b10, -3784);
    const int d8  = CL(b8 + b11),  d11 = CL(b8 - b11);
    const int d9  = CL(c9 + c10),  d10 = CL(c9 - c10);
    const int d12 = CL(b15 - b12), d15 = CL(b15 + b12);
    const int d13 = CL(c14 - c13), d14 = CL(c14 + c13);
    t[0] = d8;
    t[1] = d9;
    t[2] = r181<W>(d13 - d10);
    t[3] = r181<W>(d12 - d11);
    t[4] = r181<W>(d12 + d11);
    t[5] = r181<W>(d13 + d10);
    t[6] = d14;
    t[7] = d15;
}

template <typename W>
EAV1_DEV void idct16(const int in[16], int out[16], int mn, int mx) {
    int ev[8], od[8], e[8], t[8];
#pragma unroll
    for (int i = 0; i < 8; i++) { ev[i] = in[2 * i]; od[i] = in[2 * i + 1]; }
    idct8<W>(ev, e, mn, mx);
    idct16_odd<W>(od, t, mn, mx);
#pragma unroll
    for (int i = 0; i < 8; i++) {
        out[i] = CL(e[i] + t[7 - i]);
        out[15 - i] = CL(e[i] - t[7 - i]);
    }
}

// odd half of the 32-point DCT: o[k] = in[2k+1], result t[0..15] = t16..t31
template <typename W>
EAV1_DEV void idct32_odd(const int o[16], int t[16], int mn, int mx) {
    // stage 1: 8 rotations pairing in[k] with in[32-k]
    const int a16 = m2<W>(o[0], 201,  o[15], -4091), a31 = m2<W>(o[0], 4091, o[15], 201);
    const int a17 = m2<W>(o[8], 3035, o[7],  -2751), a30 = m2<W>(o[8], 2751, o[7],  3035);
    const int a18 = m2<W>(o[4], 1751, o[11], -3703), a29 = m2<W>(o[4], 3703, o[11], 1751);
    const int a19 = m2<W>(o[12], 3857, o[3], -1380), a28 = m2<W>(o[12], 1380, o[3], 3857);
    const int a20 = m2<W>(o[2], 995,  o[13], -3973), a27 = m2<W>(o[2], 3973, o[13], 995);
    const int a21 = m2<W>(o[10], 3513, o[5], -2106), a26 = m2<W>(o[10], 2106, o[5], 3513);
    const int a22 = m2<W>(o[6], 2440, o[9],  -3290), a25 = m2<W>(o[6], 3290, o[9],  2440);
    const int a23 = m2<W>(o[14], 4052, o[1], -601),  a24 = m2<W>(o[14], 601,  o[1], 4052);
    // stage 2
    const int b16 = CL(a16 + a17), b17 = CL(a16 - a17);
    const int b18 = CL(a19 - a18), b19 = CL(a19 + a18);
    const int b20 = CL(a20 + a21), b21 = CL(a20 - a21);
    const int b22 = CL(a23 - a22), b23 = CL(a23 + a22);
    const int b24 = CL(a24 + a25), b25 = CL(a24 - a25);
    const int b26 = CL(a27 - a26), b27 = CL(a27 + a26);
    const int b28 = CL(a28 + a29), b29 = CL(a28 - a29);
    const int b30 = CL(a31 - a30), b31 = CL(a31 + a30);
    // stage 3
    const int c17 = m2<W>(b30, 799,   b17, -4017), c30 = m2<W>(b30, 4017, b17, 799);
    const int c18 = m2<W>(b29, -4017, b18, -799),  c29 = m2<W>(b29, 799,  b18, -4017);
    const int c21 = m2<W>(b26, 3406,  b21, -2276), c26 = m2<W>(b26, 2276, b21, 3406);
    const int c22 = m2<W>(b25, -2276, b22, -3406), c25 = m2<W>(b25, 3406, b22, -2276);
    // stage 4
    const int d16 = CL(b16 + b19), d19 = CL(b16 - b19);
    const int d17 = CL(c17 + c18), d18 = CL(c17 - c18);
    const int d20 = CL(b23 - b20), d23 = CL(b23 + b20);
    const int d21 = CL(c22 - c21), d22 = CL(c22 + c21);
    const int d24 = CL(b24 + b27), d27 = CL(b24 - b27);
    const int d25 = CL(c25 + c26), d26 = CL(c25 - c26);
    const int d28 = CL(b31 - b28), d31 = CL(b31 + b28);
    const int d29 = CL(c30 - c29), d30 = CL(c30 + c29);
    // stage 5
    const int e18 = m2<W>(d29, 1567,  d18, -3784), e29 = m2<W>(d29, 3784, d18, 1567);
    const int e19 = m2<W>(d28, 1567,  d19, -3784), e28 = m2<W>(d28, 3784, d19, 1567);
    const int e20 = m2<W>(d27, -3784, d20, -1567), e27 = m2<W>(d27, 1567, d20, -3784);
    const int e21 = m2<W>(d26, -3784, d21, -1567), e26 = m2<W>(d26, 1567, d21, -3784);
    // stage 6
    const int f16 = CL(d16 + d23), f23 = CL(d16 - d23);
    const int f17 = CL(d17 + d22), f22 = CL(d17 - d22);
    const int f18 = CL(e18 + e21), f21 = CL(e18 - e21);
    const int f19 = CL(e19 + e20), f20 = CL(e19 - e20);
    const int f24 = CL(d31 - d24), f31 = CL(d31 + d24);
    const int f25 = CL(d30 - d25), f30 = CL(d30 + d25);
    const int f26 = CL(e29 - e26), f29 = CL(e29 + e26);
    const int f27 = CL(e28 - e27), f28 = CL(e28 + e27);
    // stage 7
    t[0] = f16; t[1] = f17; t[2] = f18; t[3] = f19;
    t[4] = r181<W>(f27 - f20);
    t[5] = r181<W>(f26 - f21);
    t[6] = r181<W>(f25 - f22);
    t[7] = r181<W>(f24 - f23);
    t[8] = r181<W>(f24 + f23);
    t[9] = r181<W>(f25 + f22);
    t[10] = r181<W>(f26 + f21);
    t[11] = r181<W>(f27 + f20);
    t[12] = f28; t[13] = f29; t[14] = f30; t[15] = f31;
}

template <typename W>
EAV1_DEV void idct32(const int in[32], int out[32], int mn, int mx) {
    int ev[16], od[16], e[16], t[16];
#pragma unroll
    for (int i = 0; i < 16; i++) { ev[i] = in[2 * i]; od[i] = in[2 * i + 1]; }
    idct16<W>(ev, e, mn, mx);
    idct32_odd<W>(od, t, mn, mx);
#pragma unroll
    for (int i = 0; i < 16; i++) {
        out[i] = CL(e[i] + t[15 - i]);
        out[31 - i] = CL(e[i] - t[15 - i]);
    }
}

template <typename W> EAV1_DEV int m1(int a, int c) { return (int)(((W)a * c + 2048) >> 12); }

// odd half of the 64-point DCT. Only the first 32 coefficients of a 64-point AV1
// transform can be non-zero, so the 16 inputs are o[k] = in[2k+1], k < 16, and each
// stage-1 rotation degenerates to a single product (itx_1d.c:451-482).
template <typename W>
EAV1_DEV void idct64_odd(const int o[16], int t[32], int mn, int mx) {
    const int a32 = m1<W>(o[0], 101),    a63 = m1<W>(o[0], 4095);
    const int a33 = m1<W>(o[15], -2824), a62 = m1<W>(o[15], 2967);
    const int a34 = m1<W>(o[8], 1660),   a61 = m1<W>(o[8], 3745);
    const int a35 = m1<W>(o[7], -1474),  a60 = m1<W>(o[7], 3822);
    const int a36 = m1<W>(o[4], 897),    a59 = m1<W>(o[4], 3996);
    const int a37 = m1<W>(o[11], -2191), a58 = m1<W>(o[11], 3461);
    const int a38 = m1<W>(o[12], 2359),  a57 = m1<W>(o[12], 3349);
    const int a39 = m1<W>(o[3], -700),   a56 = m1<W>(o[3], 4036);
    const int a40 = m1<W>(o[2], 501),    a55 = m1<W>(o[2], 4065);
    const int a41 = m1<W>(o[13], -2520), a54 = m1<W>(o[13], 3229);
    const int a42 = m1<W>(o[10], 2019),  a53 = m1<W>(o[10], 3564);
    const int a43 = m1<W>(o[5], -1092),  a52 = m1<W>(o[5], 3948);
    const int a44 = m1<W>(o[6], 1285),   a51 = m1<W>(o[6], 3889);
    const int a45 = m1<W>(o[9], -1842),  a50 = m1<W>(o[9], 3659);
    const int a46 = m1<W>(o[14], 2675),  a49 = m1<W>(o[14], 3102);
    const int a47 = m1<W>(o[1], -301),   a48 = m1<W>(o[1], 4085);

    const int b32 = CL(a32 + a33), b33 = CL(a32 - a33), b34 = CL(a35 - a34), b35 = CL(a35 + a34);
    const int b36 = CL(a36 + a37), b37 = CL(a36 - a37), b38 = CL(a39 - a38), b39 = CL(a39 + a38);
    const int b40 = CL(a40 + a41), b41 = CL(a40 - a41), b42 = CL(a43 - a42), b43 = CL(a43 + a42);
    const int b44 = CL(a44 + a45), b45 = CL(a44 - a45), b46 = CL(a47 - a46), b47 = CL(a47 + a46);
    const int b48 = CL(a48 + a49), b49 = CL(a48 - a49), b50 = CL(a51 - a50), b51 = CL(a51 + a50);
    const int b52 = CL(a52 + a53), b53 = CL(a52 - a53), b54 = CL(a55 - a54), b55 = CL(a55 + a54);
    const int b56 = CL(a56 + a57), b57 = CL(a56 - a57), b58 = CL(a59 - a58), b59 = CL(a59 + a58);
    const int b60 = CL(a60 + a61), b61 = CL(a60 - a61), b62 = CL(a63 - a62), b63 = CL(a63 + a62);

    const int c33 = m2<W>(b33, -4076, b62, 401),   c62 = m2<W>(b33, 401,   b62, 4076);
    const int c34 = m2<W>(b34, -401,  b61, -4076), c61 = m2<W>(b34, -4076, b61, 401);
    const int c37 = m2<W>(b37, -2598, b58, 3166),  c58 = m2<W>(b37, 3166,  b58, 2598);
    const int c38 = m2<W>(b38, -3166, b57, -2598), c57 = m2<W>(b38, -2598, b57, 3166);
    const int c41 = m2<W>(b41, -3612, b54, 1931),  c54 = m2<W>(b41, 1931,  b54, 3612);
    const int c42 = m2<W>(b42, -1931, b53, -3612), c53 = m2<W>(b42, -3612, b53, 1931);
    const int c45 = m2<W>(b45, -1189, b50, 3920),  c50 = m2<W>(b45, 3920,  b50, 1189);
    const int c46 = m2<W>(b46, -3920, b49, -1189), c49 = m2<W>(b46, -1189, b49, 3920);

    const int d32 = CL(b32 + b35), d35 = CL(b32 - b35), d33 = CL(c33 + c34), d34 = CL(c33 - c34);
    const int d36 = CL(b39 - b36), d39 = CL(b39 + b36), d37 = CL(c38 - c37), d38 = CL(c38 + c37);
    const int d40 = CL(b40 + b43), d43 = CL(b40 - b43), d41 = CL(c41 + c42), d42 = CL(c41 - c42);
    const int d44 = CL(b47 - b44), d47 = CL(b47 + b44), d45 = CL(c46 - c45), d46 = CL(c46 + c45);
    const int d48 = CL(b48 + b51), d51 = CL(b48 - b51), d49 = CL(c49 + c50), d50 = CL(c49 - c50);
    const int d52 = CL(b55 - b52), d55 = CL(b55 + b52), d53 = CL(c54 - c53), d54 = CL(c54 + c53);
    const int d56 = CL(b56 + b59), d59 = CL(b56 - b59), d57 = CL(c57 + c58), d58 = CL(c57 - c58);
    const int d60 = CL(b63 - b60), d63 = CL(b63 + b60), d61 = CL(c62 - c61), d62 = CL(c62 + c61);

    const int e34 = m2<W>(d34, -4017, d61, 799),   e61 = m2<W>(d34, 799,   d61, 4017);
    const int e35 = m2<W>(d35, -4017, d60, 799),   e60 = m2<W>(d35, 799,   d60, 4017);
    const int e36 = m2<W>(d36, -799,  d59, -4017), e59 = m2<W>(d36, -4017, d59, 799);
    const int e37 = m2<W>(d37, -799,  d58, -4017), e58 = m2<W>(d37, -4017, d58, 799);
    const int e42 = m2<W>(d42, -2276, d53, 3406),  e53 = m2<W>(d42, 3406,  d53, 2276);
    const int e43 = m2<W>(d43, -2276, d52, 3406),  e52 = m2<W>(d43, 3406,  d52, 2276);
    const int e44 = m2<W>(d44, -3406, d51, -2276), e51 = m2<W>(d44, -2276, d51, 3406);
    const int e45 = m2<W>(d45, -3406, d50, -2276), e50 = m2<W>(d45, -2276, d50, 3406);

    const int f32 = CL(d32 + d39), f39 = CL(d32 - d39), f33 = CL(d33 + d38), f38 = CL(d33 - d38);
    const int f34 = CL(e34 + e37), f37 = CL(e34 - e37), f35 = CL(e35 + e36), f36 = CL(e35 - e36);
    const int f40 = CL(d47 - d40), f47 = CL(d47 + d40), f41 = CL(d46 - d41), f46 = CL(d46 + d41);
    const int f42 = CL(e45 - e42), f45 = CL(e45 + e42), f43 = CL(e44 - e43), f44 = CL(e44 + e43);
    const int f48 = CL(d48 + d55), f55 = CL(d48 - d55), f49 = CL(d49 + d54), f54 = CL(d49 - d54);
    const int f50 = CL(e50 + e53), f53 = CL(e50 - e53), f51 = CL(e51 + e52), f52 = CL(e51 - e52);
    const int f56 = CL(d63 - d56), f63 = CL(d63 + d56), f57 = CL(d62 - d57), f62 = CL(d62 + d57);
    const int f58 = CL(e61 - e58), f61 = CL(e61 + e58), f59 = CL(e60 - e59), f60 = CL(e60 + e59);

    const int g36 = m2<W>(f36, -3784, f59, 1567),  g59 = m2<W>(f36, 1567,  f59, 3784);
    const int g37 = m2<W>(f37, -3784, f58, 1567),  g58 = m2<W>(f37, 1567,  f58, 3784);
    const int g38 = m2<W>(f38, -3784, f57, 1567),  g57 = m2<W>(f38, 1567,  f57, 3784);
    const int g39 = m2<W>(f39, -3784, f56, 1567),  g56 = m2<W>(f39, 1567,  f56, 3784);
    const int g40 = m2<W>(f40, -1567, f55, -3784), g55 = m2<W>(f40, -3784, f55, 1567);
    const int g41 = m2<W>(f41, -1567, f54, -3784), g54 = m2<W>(f41, -3784, f54, 1567);
    const int g42 = m2<W>(f42, -1567, f53, -3784), g53 = m2<W>(f42, -3784, f53, 1567);
    const int g43 = m2<W>(f43, -1567, f52, -3784), g52 = m2<W>(f43, -3784, f52, 1567);

    const int h32 = CL(f32 + f47), h47 = CL(f32 - f47), h33 = CL(f33 + f46), h46 = CL(f33 - f46);
    const int h34 = CL(f34 + f45), h45 = CL(f34 - f45), h35 = CL(f35 + f44), h44 = CL(f35 - f44);
    const int h36 = CL(g36 + g43), h43 = CL(g36 - g43), h37 = CL(g37 + g42), h42 = CL(g37 - g42);
    const int h38 = CL(g38 + g41), h41 = CL(g38 - g41), h39 = CL(g39 + g40), h40 = CL(g39 - g40);
    const int h48 = CL(f63 - f48), h63 = CL(f63 + f48), h49 = CL(f62 - f49), h62 = CL(f62 + f49);
    const int h50 = CL(f61 - f50), h61 = CL(f61 + f50), h51 = CL(f60 - f51), h60 = CL(f60 + f51);
    const int h52 = CL(g59 - g52), h59 = CL(g59 + g52), h53 = CL(g58 - g53), h58 = CL(g58 + g53);
    const int h54 = CL(g57 - g54), h57 = CL(g57 + g54), h55 = CL(g56 - g55), h56 = CL(g56 + g55);

    t[0] = h32; t[1] = h33; t[2] = h34; t[3] = h35; t[4] = h36; t[5] = h37; t[6] = h38; t[7] = h39;
    t[8]  = r181<W>(h55 - h40); t[23] = r181<W>(h55 + h40);
    t[9]  = r181<W>(h54 - h41); t[22] = r181<W>(h54 + h41);
    t[10] = r181<W>(h53 - h42); t[21] = r181<W>(h53 + h42);
    t[11] = r181<W>(h52 - h43); t[20] = r181<W>(h52 + h43);
    t[12] = r181<W>(h51 - h44); t[19] = r181<W>(h51 + h44);
    t[13] = r181<W>(h50 - h45); t[18] = r181<W>(h50 + h45);
    t[14] = r181<W>(h49 - h46); t[17] = r181<W>(h49 + h46);
    t[15] = r181<W>(h48 - h47); t[16] = r181<W>(h48 + h47);
    t[24] = h56; t[25] = h57; t[26] = h58; t[27] = h59; t[28] = h60; t[29] = h61; t[30] = h62; t[31] = h63;
}

// ---- ADST ---------------------------------------------------------------------------
template <typename W>
EAV1_DEV void iadst4(const int in[4], int out[4]) {
    out[0] = (int)(((W)1321 * in[0] + (W)3803 * in[2] + (W)2482 * in[3] + (W)3344 * in[1] + 2048) >> 12);
    out[1] = (int)(((W)2482 * in[0] - (W)1321 * in[2] - (W)3803 * in[3] + (W)3344 * in[1] + 2048) >> 12);
    out[2] = (int)(((W)209 * (in[0] - in[2] + in[3]) + 128) >> 8);
    out[3] = (int)(((W)3803 * in[0] + (W)2482 * in[2] - (W)1321 * in[3] - (W)3344 * in[1] + 2048) >> 12);
}

template <typename W>
EAV1_DEV void iadst8(const int in[8], int out[8], int mn, int mx) {
    const int a0 = m2<W>(in[7], 4076, in[0], 401),  a1 = m2<W>(in[7], 401,  in[0], -4076);
    const int a2 = m2<W>(in[5], 3612, in[2], 1931), a3 = m2<W>(in[5], 1931, in[2], -3612);
    const int a4 = m2<W>(in[3], 2598, in[4], 3166), a5 = m2<W>(in[3], 3166, in[4], -2598);
    const int a6 = m2<W>(in[1], 1189, in[6], 3920), a7 = m2<W>(in[1], 3920, in[6], -1189);
    const int t0 = CL(a0 + a4), t1 = CL(a1 + a5), t2 = CL(a2 + a6), t3 = CL(a3 + a7);
    const int t4 = CL(a0 - a4), t5 = CL(a1 - a5), t6 = CL(a2 - a6), t7 = CL(a3 - a7);
    const int b4 = m2<W>(t4, 3784, t5, 1567),  b5 = m2<W>(t4, 1567, t5, -3784);
    const int b6 = m2<W>(t7, 3784, t6, -1567), b7 = m2<W>(t7, 1567, t6, 3784);
    out[0] = CL(t0 + t2);
    out[7] = -CL(t1 + t3);
    const int u2 = CL(t0 - t2), u3 = CL(t1 - t3);
    out[1] = -CL(b4 + b6);
    out[6] = CL(b5 + b7);
    const int u6 = CL(b4 - b6), u7 = CL(b5 - b7);
    out[3] = -r181<W>(u2 + u3);
    out[4] = r181<W>(u2 - u3);
    out[2] = r181<W>(u6 + u7);
    out[5] = -r181<W>(u6 - u7);
}

template <typename W>
EAV1_DEV void iadst16(const int in[16], int out[16], int mn, int mx) {
    const int a0  = m2<W>(in[15], 4091, in[0], 201),   a1  = m2<W>(in[15], 201,  in[0], -4091);
    const int a2  = m2<W>(in[13], 3973, in[2], 995),   a3  = m2<W>(in[13], 995,  in[2], -3973);
    const int a4  = m2<W>(in[11], 3703, in[4], 1751),  a5  = m2<W>(in[11], 1751, in[4], -3703);
    const int a6  = m2<W>(in[9], 3290,  in[6], 2440),  a7  = m2<W>(in[9], 2440,  in[6], -3290);
    const int a8  = m2<W>(in[7], 2751,  in[8], 3035),  a9  = m2<W>(in[7], 3035,  in[8], -2751);
    const int a10 = m2<W>(in[5], 2106,  in[10], 3513), a11 = m2<W>(in[5], 3513,  in[10], -2106);
    const int a12 = m2<W>(in[3], 1380,  in[12], 3857), a13 = m2<W>(in[3], 3857,  in[12], -1380);
    const int a14 = m2<W>(in[1], 601,   in[14], 4052), a15 = m2<W>(in[1], 4052,  in[14], -601);

    const int b0 = CL(a0 + a8),  b1 = CL(a1 + a9),  b2 = CL(a2 + a10), b3 = CL(a3 + a11);
    const int b4 = CL(a4 + a12), b5 = CL(a5 + a13), b6 = CL(a6 + a14), b7 = CL(a7 + a15);
    const int b8 = CL(a0 - a8),  b9 = CL(a1 - a9),  b10 = CL(a2 - a10), b11 = CL(a3 - a11);
    const int b12 = CL(a4 - a12), b13 = CL(a5 - a13), b14 = CL(a6 - a14), b15 = CL(a7 - a15);

    const int c8  = m2<W>(b8, 4017,  b9, 799),    c9  = m2<W>(b8, 799,   b9, -4017);
    const int c10 = m2<W>(b10, 2276, b11, 3406),  c11 = m2<W>(b10, 3406, b11, -2276);
    const int c12 = m2<W>(b13, 4017, b12, -799),  c13 = m2<W>(b13, 799,  b12, 4017);
    const int c14 = m2<W>(b15, 2276, b14, -3406), c15 = m2<W>(b15, 3406, b14, 2276);

    const int d0 = CL(b0 + b4), d1 = CL(b1 + b5), d2 = CL(b2 + b6), d3 = CL(b3 + b7);
    const int d4 = CL(b0 - b4), d5 = CL(b1 - b5), d6 = CL(b2 - b6), d7 = CL(b3 - b7);
    const int d8 = CL(c8 + c12), d9 = CL(c9 + c13), d10 = CL(c10 + c14), d11 = CL(c11 + c15);
    const int d12 = CL(c8 - c12), d13 = CL(c9 - c13), d14 = CL(c10 - c14), d15 = CL(c11 - c15);

    const int e4  = m2<W>(d4, 3784, d5, 1567),    e5  = m2<W>(d4, 1567, d5, -3784);
    const int e6  = m2<W>(d7, 3784, d6, -1567),   e7  = m2<W>(d7, 1567, d6, 3784);
    const int e12 = m2<W>(d12, 3784, d13, 1567),  e13 = m2<W>(d12, 1567, d13, -3784);
    const int e14 = m2<W>(d15, 3784, d14, -1567), e15 = m2<W>(d15, 1567, d14, 3784);

    out[0]  = CL(d0 + d2);
    out[15] = -CL(d1 + d3);
    const int f2 = CL(d0 - d2), f3 = CL(d1 - d3);
    out[3]  = -CL(e4 + e6);
    out[12] = CL(e5 + e7);
    const int f6 = CL(e4 - e6), f7 = CL(e5 - e7);
    out[1]  = -CL(d8 + d10);
    out[14] = CL(d9 + d11);
    const int f10 = CL(d8 - d10), f11 = CL(d9 - d11);
    out[2]  = CL(e12 + e14);
    out[13] = -CL(e13 + e15);
    const int f14 = CL(e12 - e14), f15 = CL(e13 - e15);

    out[7]  = -r181<W>(f2 + f3);
    out[8]  = r181<W>(f2 - f3);
    out[4]  = r181<W>(f6 + f7);
    out[11] = -r181<W>(f6 - f7);
    out[6]  = r181<W>(f10 + f11);
    out[9]  = -r181<W>(f10 - f11);
    out[5]  = -r181<W>(f14 + f15);
    out[10] = r181<W>(f14 - f15);
}

#undef CL

// One 1-D transform of length n (4..64) of kind `type`, in place on c[i*stride].
// For n == 64 only c[0..31] are read (the rest is defined to be zero).
template <typename W>
EAV1_DEVN void itx_1d(int type, int n, int32_t *c, int stride, int mn, int mx) {
    if (type == T1D_IDENTITY) {
        switch (n) {
        case 4:
            for (int i = 0; i < 4; i++) { const int v = c[i * stride]; c[i * stride] = v + (int)(((W)v * 1697 + 2048) >> 12); }
            break;
        case 8:
            for (int i = 0; i < 8; i++) c[i * stride] *= 2;
            break;
        case 16:
            for (int i = 0; i < 16; i++) { const int v = c[i * stride]; c[i * stride] = 2 * v + (int)(((W)v * 1697 + 1024) >> 11); }
            break;
        default:
            for (int i = 0; i < 32; i++) c[i * stride] *= 4;
            break;
        }
        return;
    }
    if (type == T1D_DCT) {
        switch (n) {
        case 4: {
            int in[4], out[4];
            for (int i = 0; i < 4; i++) in[i] = c[i * stride];
            idct4<W>(in, out, mn, mx);
            for (int i = 0; i < 4; i++) c[i * stride] = out[i];
            break;
        }
        case 8: {
            int in[8], out[8];
            for (int i = 0; i < 8; i++) in[i] = c[i * stride];
            idct8<W>(in, out, mn, mx);
            for (int i = 0; i < 8; i++) c[i * stride] = out[i];
            break;
        }
        case 16: {
            int in[16], out[16];
            for (int i = 0; i < 16; i++) in[i] = c[i * stride];
            idct16<W>(in, out, mn, mx);
            for (int i = 0; i < 16; i++) c[i * stride] = out[i];
            break;
        }
        case 32: {
            int in[32], out[32];
            for (int i = 0; i < 32; i++) in[i] = c[i * stride];
            idct32<W>(in, out, mn, mx);
            for (int i = 0; i < 32; i++) c[i * stride] = out[i];
            break;
        }
        default: {
            int ev[32], od[16], e[32], t[32];
            for (int i = 0; i < 16; i++) { ev[i] = c[2 * i * stride]; od[i] = c[(2 * i + 1) * stride]; ev[16 + i] = 0; }
            idct32<W>(ev, e, mn, mx);
            idct64_odd<W>(od, t, mn, mx);
            for (int i = 0; i < 32; i++) {
                c[i * stride] = iclip(e[i] + t[31 - i], mn, mx);
                c[(63 - i) * stride] = iclip(e[i] - t[31 - i], mn, mx);
            }
            break;
        }
        }
        return;
    }
    // ADST / FLIPADST
    const int flip = type == T1D_FLIPADST;
    switch (n) {
    case 4: {
        int in[4], out[4];
        for (int i = 0; i < 4; i++) in[i] = c[i * stride];
        iadst4<W>(in, out);
        for (int i = 0; i < 4; i++) c[(flip ? 3 - i : i) * stride] = out[i];
        break;
    }
    case 8: {
        int in[8], out[8];
        for (int i = 0; i < 8; i++) in[i] = c[i * stride];
        iadst8<W>(in, out, mn, mx);
        for (int i = 0; i < 8; i++) c[(flip ? 7 - i : i) * stride] = out[i];
        break;
    }
    default: {
        int in[16], out[16];
        for (int i = 0; i < 16; i++) in[i] = c[i * stride];
        iadst16<W>(in, out, mn, mx);
        for (int i = 0; i < 16; i++) c[(flip ? 15 - i : i) * stride] = out[i];
        break;
    }
    }
}

EAV1_DEV void wht4(int32_t *c, int stride) {
    const int in0 = c[0], in1 = c[stride], in2 = c[2 * stride], in3 = c[3 * stride];
    const int t0 = in0 + in1, t2 = in2 - in3, t4 = (t0 - t2) >> 1;
    const int t3 = t4 - in3, t1 = t4 - in1;
    c[0] = t0 - t3; c[stride] = t3; c[2 * stride] = t1; c[3 * stride] = t2 + t1;
}

// (row type, column type) for the table index used by itxfm_add[tx][txtp]
// (itx_tmpl.c:206-260 binds index X_Y to the function whose first/row pass is Y's kind).
EAV1_DEV void txtp_1d_types(int txtp, int &row, int &col) {
    // index:            DCT_DCT ADST_DCT DCT_ADST ADST_ADST FLIP_DCT DCT_FLIP FLIP_FLIP ADST_FLIP FLIP_ADST IDTX V_DCT H_DCT V_ADST H_ADST V_FLIP H_FLIP
    const uint32_t rows = 0u | (0u << 2) | (1u << 4) | (1u << 6) | (0u << 8) | (3u << 10) | (3u << 12) | (3u << 14) |
                          (1u << 16) | (2u << 18) | (2u << 20) | (0u << 22) | (2u << 24) | (1u << 26) | (2u << 28) | (3u << 30);
    const uint32_t cols = 0u | (1u << 2) | (0u << 4) | (1u << 6) | (3u << 8) | (0u << 10) | (3u << 12) | (1u << 14) |
                          (3u << 16) | (2u << 18) | (0u << 20) | (2u << 22) | (1u << 24) | (2u << 26) | (3u << 28) | (2u << 30);
    row = (rows >> (2 * txtp)) & 3;
    col = (cols >> (2 * txtp)) & 3;
}

struct TxDim { uint8_t w, h, shift; };
EAV1_DEV TxDim tx_dim(int tx) {
    // enum RectTxfmSize order (levels.h:44-78); shift per itx_tmpl.c:160-178
    // packed tables (a local array would be rebuilt on the stack at every call): log2(w) - 2 and
    // log2(h) - 2 in three bits, the shift in two bits per transform size
    // W = { 4, 8, 16, 32, 64, 4, 8, 8, 16, 16, 32, 32, 64, 4, 16, 8, 32, 16, 64 }
    // H = { 4, 8, 16, 32, 64, 8, 4, 16, 8, 32, 16, 64, 32, 16, 4, 32, 8, 64, 16 }
    // S = { 0, 1, 2, 2, 2, 0, 0, 1, 1, 1, 1, 1, 1, 1, 1, 2, 2, 2, 2 }
    const unsigned long long LW = 0x1132846d2244688ull, LH = 0xa161389940c688ull, SS = 0x2a955542a4ull;
    TxDim d;
    d.w = (uint8_t)(4 << ((LW >> (3 * tx)) & 7)); d.h = (uint8_t)(4 << ((LH >> (3 * tx)) & 7)); d.shift = (uint8_t)((SS >> (2 * tx)) & 3);
    return d;
}

// ---- register path (both dimensions <= 32) ---------------------------------------------
// One lane owns one row (first pass) and then one column (second pass) entirely in registers:
// the coefficients are read straight from the payload (column-major, so consecutive lanes read
// consecutive addresses), the only shared-memory traffic is the transpose between the passes,
// and the second pass adds into the destination directly. One group barrier per transform.

// A residual travels as int16. The only value that does not fit is +32768: a 12-bit 32-point identity column whose
// input sits at the inter-pass clip limit (131071 * 4 + 8 >> 4). Saturating it to 32767 cannot change the
// reconstructed sample, which is clipped to bitdepth_max <= 4095 after the add (found by tests/test_ops.py).
EAV1_DEV int16_t res16(int v) { return (int16_t)(v > 32767 ? 32767 : v); }

// N-point 1-D transform of kind `type` (T1D_FLIPADST is computed as ADST; the caller mirrors
// the output index), in place on a register array.
template <typename W, int N>
EAV1_DEV void itx_1d_reg(int type, int (&v)[N], int mn, int mx) {
    if (type == T1D_IDENTITY) {
#pragma unroll
        for (int i = 0; i < N; i++) {
            if (N == 4) v[i] = v[i] + (int)(((W)v[i] * 1697 + 2048) >> 12);
            else if (N == 8) v[i] *= 2;
            else if (N == 16) v[i] = 2 * v[i] + (int)(((W)v[i] * 1697 + 1024) >> 11);
            else v[i] *= 4;
        }
        return;
    }
    int o[N];
    if (type == T1D_DCT || N == 32) {
        if constexpr (N == 4) idct4<W>(v, o, mn, mx);
        else if constexpr (N == 8) idct8<W>(v, o, mn, mx);
        else if constexpr (N == 16) idct16<W>(v, o, mn, mx);
        else idct32<W>(v, o, mn, mx);
    } else {
        if constexpr (N == 4) iadst4<W>(v, o);
        else if constexpr (N == 8) iadst8<W>(v, o, mn, mx);
        else if constexpr (N == 16) iadst16<W>(v, o, mn, mx);
    }
#pragma unroll
    for (int i = 0; i < N; i++) v[i] = o[i];
}

// first pass of row y: coefficients -> s[x*P + y] (already rounded/clipped for the second pass)
template <typename BD, int N>
EAV1_DEV void itx_row_reg(const typename BD::coef *coeff, int sh, int y, int type, int is_rect2, int shift,
                          int row_min, int row_max, int col_min, int col_max, int32_t *s, int P)
{
    typedef typename BD::wide W;
    if (y >= sh) {       // a row past the last one that was shipped: all zero in, all zero out (every 1-D transform is linear)
#pragma unroll
        for (int x = 0; x < N; x++) s[x * P + y] = 0;
        return;
    }
    int v[N];
#pragma unroll
    for (int x = 0; x < N; x++) {
        int c = coeff[y + x * sh];
        if (is_rect2) c = (int)(((W)c * 181 + 128) >> 8);
        v[x] = c;
    }
    itx_1d_reg<W, N>(type, v, row_min, row_max);
    const int rnd = (1 << shift) >> 1;
    const int flip = type == T1D_FLIPADST;
#pragma unroll
    for (int x = 0; x < N; x++)
        s[(flip ? N - 1 - x : x) * P + y] = iclip((v[x] + rnd) >> shift, col_min, col_max);
}

// second pass of column x, added into dst
template <typename BD, int N>
EAV1_DEV void itx_col_reg(const int32_t *s, int P, int x, int type, int col_min, int col_max,
                          typename BD::pixel *dst, int dst_stride, int bitdepth_max, int16_t *res, int res_stride)
{
    typedef typename BD::wide W;
    typedef typename BD::pixel pixel;
    int v[N];
#pragma unroll
    for (int y = 0; y < N; y++) v[y] = s[x * P + y];
    const int flip = type == T1D_FLIPADST;
    if (res) {   // residual only (phase 1): row-major int16 block, added to the prediction later
        itx_1d_reg<W, N>(type, v, col_min, col_max);
#pragma unroll
        for (int y = 0; y < N; y++) res[(flip ? N - 1 - y : y) * res_stride + x] = res16((v[y] + 8) >> 4);
        return;
    }
    pixel *d = dst + x;
    if (N <= 16) {
        // issue the destination loads before the butterflies so their latency hides behind them
        int px[N];
#pragma unroll
        for (int y = 0; y < N; y++) px[y] = d[(flip ? N - 1 - y : y) * dst_stride];
        itx_1d_reg<W, N>(type, v, col_min, col_max);
#pragma unroll
        for (int y = 0; y < N; y++)
            d[(flip ? N - 1 - y : y) * dst_stride] = (pixel)iclip(px[y] + ((v[y] + 8) >> 4), 0, bitdepth_max);
    } else {
        itx_1d_reg<W, N>(type, v, col_min, col_max);
#pragma unroll
        for (int y = 0; y < N; y++) {
            pixel *q = d + (flip ? N - 1 - y : y) * dst_stride;
            *q = (pixel)iclip(*q + ((v[y] + 8) >> 4), 0, bitdepth_max);
        }
    }
}

template <typename BD>
EAV1_DEVH void itx_add_reg(typename BD::pixel *dst, int dst_stride, const typename BD::coef *coeff, int w, int h,
                           int shift, int rt, int ct, int bitdepth_max, int32_t *s, int tid, int nthr, int16_t *res, int rows)
{
    const int is_rect2 = (w * 2 == h) || (h * 2 == w);
    const int P = h + 1;
    int row_min, col_min;
    if (!BD::hbd) { row_min = col_min = -32768; }
    else {
        row_min = (int)((unsigned)~bitdepth_max << 7);
        col_min = (int)((unsigned)~bitdepth_max << 5);
    }
    const int row_max = ~row_min, col_max = ~col_min;
    PROF_BEGIN();
    for (int y = tid; y < h; y += nthr) {
        switch (w) {
        case 4:  itx_row_reg<BD, 4>(coeff, rows, y, rt, is_rect2, shift, row_min, row_max, col_min, col_max, s, P); break;
        case 8:  itx_row_reg<BD, 8>(coeff, rows, y, rt, is_rect2, shift, row_min, row_max, col_min, col_max, s, P); break;
        case 16: itx_row_reg<BD, 16>(coeff, rows, y, rt, is_rect2, shift, row_min, row_max, col_min, col_max, s, P); break;
        default: itx_row_reg<BD, 32>(coeff, rows, y, rt, is_rect2, shift, row_min, row_max, col_min, col_max, s, P); break;
        }
    }
    PROF_MARK(10);
    CTA_SYNC();
    PROF_MARK(11);
    for (int x = tid; x < w; x += nthr) {
        switch (h) {
        case 4:  itx_col_reg<BD, 4>(s, P, x, ct, col_min, col_max, dst, dst_stride, bitdepth_max, res, w); break;
        case 8:  itx_col_reg<BD, 8>(s, P, x, ct, col_min, col_max, dst, dst_stride, bitdepth_max, res, w); break;
        case 16: itx_col_reg<BD, 16>(s, P, x, ct, col_min, col_max, dst, dst_stride, bitdepth_max, res, w); break;
        default: itx_col_reg<BD, 32>(s, P, x, ct, col_min, col_max, dst, dst_stride, bitdepth_max, res, w); break;
        }
    }
    PROF_MARK(12);
    CTA_SYNC();
    PROF_MARK(13);
}

// dst += inverse_transform(coef). `s` is the CTA's shared tile (>= 64*65 int32).
template <typename BD>
// `rows`: how many rows of the (column-major, min(w,32) x min(h,32)) coefficient block were shipped, which is also the
// stride between its columns; the rows behind them are zero (itx_tmpl.c:87-105 last_nonzero_col). 0 = the whole block.
EAV1_DEVH void itx_add(typename BD::pixel *dst, int dst_stride /*pixels*/, const typename BD::coef *coeff,
                      int eob, int tx, int txtp, int bitdepth_max, int32_t *s, int tid, int nthr, int16_t *res = nullptr,
                      int rows = 0)
{
    typedef typename BD::wide W;
    const TxDim d = tx_dim(tx);
    const int w = d.w, h = d.h;
    if (rows <= 0 || rows > imin(h, 32)) rows = imin(h, 32);
    if (txtp == 16) { // WHT_WHT, 4x4 only
        if (tid == 0) {
            for (int y = 0; y < 4; y++)
                for (int x = 0; x < 4; x++) s[y * 4 + x] = coeff[y + x * 4] >> 2;     /* lossless blocks always ship all four rows */
            for (int y = 0; y < 4; y++) wht4(&s[y * 4], 1);
            for (int x = 0; x < 4; x++) wht4(&s[x], 4);
            for (int y = 0; y < 4; y++)
                for (int x = 0; x < 4; x++)
                    dst[y * dst_stride + x] = (typename BD::pixel)iclip(dst[y * dst_stride + x] + s[y * 4 + x], 0, bitdepth_max);
        }
        CTA_SYNC();
        return;
    }
    const int is_rect2 = (w * 2 == h) || (h * 2 == w);
    const int rnd = (1 << d.shift) >> 1;
    if (eob < (txtp == 0)) {
        int dc = coeff[0];
        if (is_rect2) dc = (int)(((W)dc * 181 + 128) >> 8);
        dc = (int)(((W)dc * 181 + 128) >> 8);
        dc = (dc + rnd) >> d.shift;
        dc = (int)(((W)dc * 181 + 128 + 2048) >> 12);
        for (int i = tid; i < w * h; i += nthr) {
            const int y = i / w, x = i - y * w;
            dst[y * dst_stride + x] = (typename BD::pixel)iclip(dst[y * dst_stride + x] + dc, 0, bitdepth_max);
        }
        CTA_SYNC();
        return;
    }
    int rt, ct;
    txtp_1d_types(txtp, rt, ct);
#ifndef EAV1_ITX_REG
#define EAV1_ITX_REG 1
#endif
    if (EAV1_ITX_REG && w <= 32 && h <= 32) {
        itx_add_reg<BD>(dst, dst_stride, coeff, w, h, d.shift, rt, ct, bitdepth_max, s, tid, nthr, res, rows);
        return;
    }
    const int sh = imin(h, 32), sw = imin(w, 32);
    const int P = h + 1;
    int row_min, col_min;
    if (!BD::hbd) { row_min = col_min = -32768; }
    else {
        row_min = (int)((unsigned)~bitdepth_max << 7);
        col_min = (int)((unsigned)~bitdepth_max << 5);
    }
    const int row_max = ~row_min, col_max = ~col_min;

    // load (column-major source: index i = y + x*sh)
    for (int i = tid; i < sw * sh; i += nthr) {
        const int x = i / sh, y = i - x * sh;
        int v = y < rows ? coeff[y + x * rows] : 0;
        if (is_rect2) v = (int)(((W)v * 181 + 128) >> 8);
        s[x * P + y] = v;
    }
    CTA_SYNC();
    // first pass: rows y < sh, length w
    for (int y = tid; y < sh; y += nthr)
        itx_1d<W>(rt, w, &s[y], P, row_min, row_max);
    CTA_SYNC();
    // inter-pass rounding/clip; rows sh..h-1 (64-point columns) are implicitly zero
    for (int i = tid; i < w * sh; i += nthr) {
        const int x = i / sh, y = i - x * sh;
        s[x * P + y] = iclip((s[x * P + y] + rnd) >> d.shift, col_min, col_max);
    }
    CTA_SYNC();
    for (int x = tid; x < w; x += nthr)
        itx_1d<W>(ct, h, &s[x * P], 1, col_min, col_max);
    CTA_SYNC();
    if (res) {
        for (int i = tid; i < w * h; i += nthr) {
            const int y = i / w, x = i - y * w;
            res[i] = res16((s[x * P + y] + 8) >> 4);
        }
        CTA_SYNC();
        return;
    }
    for (int i = tid; i < w * h; i += nthr) {
        const int y = i / w, x = i - y * w;
        dst[y * dst_stride + x] = (typename BD::pixel)iclip(dst[y * dst_stride + x] + ((s[x * P + y] + 8) >> 4), 0, bitdepth_max);
    }
    CTA_SYNC();
}

// Is this transform block reconstructed in two steps (residual in phase 1, add in phase 2)?
// DC-only blocks and the lossless WHT stay single-step: their residual is a constant / trivial.
EAV1_DEV bool itx_two_step(int eob, int txtp) { return txtp != 16 && !(eob < (txtp == 0)); }

// dst += res (row-major int16 residual of a w x h transform block, w and h powers of two)
template <typename BD>
EAV1_DEV void res_add(typename BD::pixel *dst, int dst_stride, const int16_t *res, int w, int h,
                      int bitdepth_max, int tid, int nthr)
{
    typedef typename BD::pixel pixel;
    const int lw = ulog2((unsigned)w);
    if (w >= 4 && !BD::hbd) {
        // four samples per thread: one 64-bit residual load, one 32-bit pixel load/store
        const int n4 = (w * h) >> 2, lw4 = lw - 2;
        for (int i = tid; i < n4; i += nthr) {
            const int y = i >> lw4, x = (i & ((1 << lw4) - 1)) << 2;
            const uint2 r = *reinterpret_cast<const uint2 *>(res + (i << 2));
            uint32_t *q = reinterpret_cast<uint32_t *>(dst + y * dst_stride + x);
            const uint32_t p = *q;
            const int r0 = (int16_t)(r.x & 0xffff), r1 = (int16_t)(r.x >> 16), r2 = (int16_t)(r.y & 0xffff), r3 = (int16_t)(r.y >> 16);
            const uint32_t o = (uint32_t)iclip((int)(p & 0xff) + r0, 0, 255) | (uint32_t)iclip((int)((p >> 8) & 0xff) + r1, 0, 255) << 8 |
                               (uint32_t)iclip((int)((p >> 16) & 0xff) + r2, 0, 255) << 16 | (uint32_t)iclip((int)(p >> 24) + r3, 0, 255) << 24;
            *q = o;
        }
    } else {
        for (int i = tid; i < w * h; i += nthr) {
            const int y = i >> lw, x = i & (w - 1);
            pixel *q = dst + y * dst_stride + x;
            *q = (pixel)iclip(*q + res[i], 0, bitdepth_max);
        }
    }
}

} // namespace eav1
