// tests/emu/emu_vorbis.cpp — DEBUGGING AID, TEST INFRASTRUCTURE ONLY (see emu_engine.cpp).
// Host emulation of vorbis.cu: the device code of easyav1_b200/csrc/vorbis.cuh run with one
// "thread" per CTA behind the same eav1cu_vorbis_* entry points.
#include <math.h>
#include <string.h>
#include <vector>

#include "../../include/easyav1_cuda.h"
#include "../../easyav1_b200/csrc/vorbis.cuh"

using namespace eav1;

struct Eav1Vorbis {
    int channels, bs[2];
    std::vector<cpx> tw[2], fftw[2];
    std::vector<float> win[2], state[2];
    int state_cur = 0, state_W = -1;
    long long launches = 0;
};

extern "C" {

int eav1cu_vorbis_open(Eav1Vorbis **out, int, int channels, int bs0, int bs1, const float *w0, const float *w1)
{
    Eav1Vorbis *v = new Eav1Vorbis();
    v->channels = channels; v->bs[0] = bs0; v->bs[1] = bs1;
    const float *win_in[2] = { w0, w1 };
    for (int s = 0; s < 2; s++) {
        const int M = v->bs[s] / 2, H = M / 2;
        v->tw[s].resize(H); v->fftw[s].resize(H); v->win[s].resize(M);
        for (int k = 0; k < H; k++) {
            const double a = -M_PI * (8.0 * k + 1.0) / (8.0 * M), b = -2.0 * M_PI * k / H;
            v->tw[s][k] = { (float)cos(a), (float)sin(a) };
            v->fftw[s][k] = { (float)cos(b), (float)sin(b) };
        }
        for (int i = 0; i < M; i++) {
            const double x = sin((i + 0.5) / (2.0 * M) * M_PI);
            v->win[s][i] = win_in[s] ? win_in[s][i] : (float)sin(0.5 * M_PI * x * x);
        }
        v->state[s].assign((size_t)channels * (bs1 / 2), 0.f);
    }
    *out = v;
    return 0;
}

void eav1cu_vorbis_close(Eav1Vorbis *v) { delete v; }
int eav1cu_vorbis_restart(Eav1Vorbis *v) { v->state_W = -1; return 0; }
long long eav1cu_vorbis_launch_count(Eav1Vorbis *v) { return v->launches; }
int eav1cu_vorbis_last_error(Eav1Vorbis *, char *msg, size_t cap) { if (msg && cap) msg[0] = 0; return 0; }

int eav1cu_vorbis_imdct_ola(Eav1Vorbis *v, const uint8_t *flags, int n_blocks, const float *spectra,
                            float *pcm, size_t cap_samples, int interleave, size_t *n_samples)
{
    std::vector<uint32_t> spec_off(n_blocks), out_off(n_blocks + 1);
    size_t so = 0, oo = 0;
    int prev = v->state_W;
    for (int b = 0; b < n_blocks; b++) {
        const int W = flags[b] ? 1 : 0;
        spec_off[b] = (uint32_t)so; out_off[b] = (uint32_t)oo;
        so += (size_t)v->channels * (v->bs[W] / 2);
        if (prev >= 0) oo += v->bs[prev] / 4 + v->bs[W] / 4;
        prev = W;
    }
    out_off[n_blocks] = (uint32_t)oo;
    if (oo > cap_samples) return EAV1CU_ERR_INVAL;
    VorbisParams P;
    memset(&P, 0, sizeof(P));
    P.channels = v->channels; P.n_blocks = n_blocks; P.run_len = 8;
    for (int s = 0; s < 2; s++) {
        P.M[s] = v->bs[s] / 2;
        P.tw[s] = v->tw[s].data(); P.fftw[s] = v->fftw[s].data(); P.win[s] = v->win[s].data();
    }
    std::vector<uint8_t> fl(n_blocks);
    for (int b = 0; b < n_blocks; b++) fl[b] = flags[b] ? 1 : 0;
    P.flags = fl.data(); P.spec_off = spec_off.data(); P.out_off = out_off.data(); P.spectra = spectra;
    P.state_in = v->state[v->state_cur].data(); P.state_W = v->state_W;
    P.state_out = v->state[v->state_cur ^ 1].data();
    P.pcm = pcm; P.pcm_stride = cap_samples; P.interleave = interleave;
    std::vector<float> smem(3 * (size_t)P.M[1]);
    for (int run = 0; run < (n_blocks + P.run_len - 1) / P.run_len; run++)
        for (int ch = 0; ch < v->channels; ch++) vorbis_run(P, run, ch, smem.data(), 0, 1);
    v->launches++;
    v->state_cur ^= 1;
    v->state_W = flags[n_blocks - 1] ? 1 : 0;
    if (n_samples) *n_samples = oo;
    return 0;
}

} // extern "C"
