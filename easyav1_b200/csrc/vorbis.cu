// vorbis.cu — C ABI of the batched Vorbis IMDCT + window + overlap-add (include/easyav1_cuda.h,
// "Vorbis float slice"). Device code: vorbis.cuh. sm_100a only; no CPU fallback.
#include <cuda_runtime.h>
#include <math.h>
#include <stdio.h>
#include <string.h>
#include <string>
#include <vector>

#include "../../include/easyav1_cuda.h"
#include "vorbis.cuh"

using namespace eav1;

__global__ void __launch_bounds__(128)
vorbis_kernel(VorbisParams P)
{
    extern __shared__ float vorbis_smem[];
    vorbis_run(P, blockIdx.x, blockIdx.y, vorbis_smem, threadIdx.x, blockDim.x);
}

struct VBuf { void *d = nullptr; size_t cap = 0; };

struct Eav1Vorbis {
    int device = 0, channels = 0, bs[2] = {0, 0};
    cudaStream_t stream = nullptr;
    bool own_stream = true;
    cpx *d_tw[2] = {}, *d_fftw[2] = {};
    float *d_win[2] = {};
    float *d_state[2] = {};          // ping-pong: a batch reads one and writes the other
    int state_cur = 0, state_W = -1;
    VBuf flags, spec_off, out_off, spectra, pcm;
    void *h_pin = nullptr; size_t h_pin_cap = 0;   // pinned staging for the host entry point
    // staged batch (measurement)
    VorbisParams staged; bool have_staged = false; size_t staged_samples = 0;
    int err = 0; std::string errmsg;
    long long launches = 0, h2d = 0, d2h = 0;
};

static int vfail(Eav1Vorbis *v, int code, const char *what, cudaError_t e = cudaSuccess)
{
    if (v && !v->err) {
        v->err = code; v->errmsg = what;
        if (e != cudaSuccess) { v->errmsg += ": "; v->errmsg += cudaGetErrorString(e); }
    }
    return code;
}
#define VCK(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return vfail(v, EAV1CU_ERR_CUDA, #call, e_); } while (0)

static int vensure(Eav1Vorbis *v, VBuf &b, size_t bytes)
{
    if (bytes <= b.cap) return 0;
    if (b.d) cudaFree(b.d);
    b.d = nullptr; b.cap = 0;
    const size_t cap = bytes + bytes / 4 + 256;
    VCK(cudaMalloc(&b.d, cap));
    b.cap = cap;
    return 0;
}

static bool pow2(int n) { return n > 0 && !(n & (n - 1)); }

extern "C" {

int eav1cu_vorbis_open(Eav1Vorbis **out, int device, int channels, int blocksize0, int blocksize1,
                       const float *window0, const float *window1)
{
    if (!out) return EAV1CU_ERR_INVAL;
    *out = nullptr;
    // blocksizes: powers of two, 64..8192, short <= long (Vorbis I, minivorbis.h:8650-8653)
    if (channels < 1 || channels > 255 || !pow2(blocksize0) || !pow2(blocksize1) || blocksize0 < 64 ||
        blocksize1 > 8192 || blocksize0 > blocksize1)
        return EAV1CU_ERR_INVAL;
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || device < 0 || device >= n) return EAV1CU_ERR_NO_DEVICE;
    Eav1Vorbis *v = new Eav1Vorbis();
    v->device = device; v->channels = channels; v->bs[0] = blocksize0; v->bs[1] = blocksize1;
    if (cudaSetDevice(device) != cudaSuccess || cudaStreamCreateWithFlags(&v->stream, cudaStreamNonBlocking) != cudaSuccess) {
        delete v;
        return EAV1CU_ERR_CUDA;
    }
    const float *win_in[2] = { window0, window1 };
    for (int s = 0; s < 2; s++) {
        const int M = v->bs[s] / 2, H = M / 2;
        std::vector<cpx> tw(H), fw(H);
        for (int k = 0; k < H; k++) {
            const double a = -M_PI * (8.0 * k + 1.0) / (8.0 * M), b = -2.0 * M_PI * k / H;
            tw[k] = { (float)cos(a), (float)sin(a) };
            fw[k] = { (float)cos(b), (float)sin(b) };
        }
        std::vector<float> win(M);
        for (int i = 0; i < M; i++) {
            // the Vorbis window, sin(pi/2 sin^2((i + 1/2) pi / n)); a caller that links the reference's
            // tables (minivorbis.h:9917-11990) passes them instead
            const double x = sin((i + 0.5) / (2.0 * M) * M_PI);
            win[i] = win_in[s] ? win_in[s][i] : (float)sin(0.5 * M_PI * x * x);
        }
        if (cudaMalloc(&v->d_tw[s], H * sizeof(cpx)) != cudaSuccess || cudaMalloc(&v->d_fftw[s], H * sizeof(cpx)) != cudaSuccess ||
            cudaMalloc(&v->d_win[s], M * sizeof(float)) != cudaSuccess ||
            cudaMalloc(&v->d_state[s], (size_t)channels * (v->bs[1] / 2) * sizeof(float)) != cudaSuccess ||
            cudaMemcpy(v->d_tw[s], tw.data(), H * sizeof(cpx), cudaMemcpyHostToDevice) != cudaSuccess ||
            cudaMemcpy(v->d_fftw[s], fw.data(), H * sizeof(cpx), cudaMemcpyHostToDevice) != cudaSuccess ||
            cudaMemcpy(v->d_win[s], win.data(), M * sizeof(float), cudaMemcpyHostToDevice) != cudaSuccess) {
            eav1cu_vorbis_close(v);
            return EAV1CU_ERR_CUDA;
        }
    }
    const size_t smem = 3 * (size_t)(v->bs[1] / 2) * sizeof(float);
    if (smem > 48 * 1024)
        cudaFuncSetAttribute(vorbis_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    *out = v;
    return 0;
}

void eav1cu_vorbis_close(Eav1Vorbis *v)
{
    if (!v) return;
    cudaSetDevice(v->device);
    if (v->stream) cudaStreamSynchronize(v->stream);
    for (int s = 0; s < 2; s++) {
        if (v->d_tw[s]) cudaFree(v->d_tw[s]);
        if (v->d_fftw[s]) cudaFree(v->d_fftw[s]);
        if (v->d_win[s]) cudaFree(v->d_win[s]);
        if (v->d_state[s]) cudaFree(v->d_state[s]);
    }
    for (VBuf *b : { &v->flags, &v->spec_off, &v->out_off, &v->spectra, &v->pcm }) if (b->d) cudaFree(b->d);
    if (v->h_pin) cudaFreeHost(v->h_pin);
    if (v->own_stream && v->stream) cudaStreamDestroy(v->stream);
    delete v;
}

int eav1cu_vorbis_last_error(Eav1Vorbis *v, char *msg, size_t cap)
{
    if (!v) return EAV1CU_ERR_INVAL;
    if (msg && cap) snprintf(msg, cap, "%s", v->errmsg.c_str());
    return v->err;
}

/* vorbis_synthesis_restart (minivorbis.h:9159): forget the overlap state, e.g. after a seek */
int eav1cu_vorbis_restart(Eav1Vorbis *v)
{
    if (!v) return EAV1CU_ERR_INVAL;
    v->state_W = -1;
    return 0;
}

int eav1cu_vorbis_set_stream(Eav1Vorbis *v, void *cuda_stream)
{
    if (!v) return EAV1CU_ERR_INVAL;
    cudaSetDevice(v->device);
    cudaStreamSynchronize(v->stream);
    if (v->own_stream && v->stream) cudaStreamDestroy(v->stream);
    v->stream = (cudaStream_t)cuda_stream;
    v->own_stream = false;
    return 0;
}

long long eav1cu_vorbis_launch_count(Eav1Vorbis *v) { return v ? v->launches : 0; }

// upload one batch and fill the kernel parameters; *n_samples = samples per channel it will emit
static int vorbis_prepare(Eav1Vorbis *v, VorbisParams &P, const uint8_t *flags, int n_blocks, const float *spectra,
                          size_t cap_samples, int interleave, size_t *n_samples, bool from_pinned_copy)
{
    if (v->err) return v->err;
    if (!flags || !spectra || n_blocks <= 0) return vfail(v, EAV1CU_ERR_INVAL, "vorbis: empty batch");
    VCK(cudaSetDevice(v->device));
    std::vector<uint32_t> spec_off(n_blocks), out_off(n_blocks + 1);
    size_t so = 0, oo = 0;
    int prev = v->state_W;
    for (int b = 0; b < n_blocks; b++) {
        const int W = flags[b] ? 1 : 0;
        spec_off[b] = (uint32_t)so;
        out_off[b] = (uint32_t)oo;
        so += (size_t)v->channels * (v->bs[W] / 2);
        if (prev >= 0) oo += v->bs[prev] / 4 + v->bs[W] / 4;
        prev = W;
        if (so > 0xffffffffu) return vfail(v, EAV1CU_ERR_INVAL, "vorbis: batch too large");
    }
    out_off[n_blocks] = (uint32_t)oo;
    if (n_samples) *n_samples = oo;
    if (oo > cap_samples) return vfail(v, EAV1CU_ERR_INVAL, "vorbis: pcm buffer too small");
    int r;
    if ((r = vensure(v, v->flags, n_blocks)) || (r = vensure(v, v->spec_off, n_blocks * 4)) ||
        (r = vensure(v, v->out_off, (n_blocks + 1) * 4)) || (r = vensure(v, v->spectra, so * 4)) ||
        (r = vensure(v, v->pcm, (size_t)v->channels * (oo ? oo : 1) * 4)))
        return r;
    // one pinned staging buffer: [spectra | spec_off | out_off | flags]
    const size_t need = so * 4 + (size_t)(2 * n_blocks + 1) * 4 + n_blocks;
    if (need > v->h_pin_cap) {
        if (v->h_pin) cudaFreeHost(v->h_pin);
        v->h_pin = nullptr; v->h_pin_cap = 0;
        VCK(cudaHostAlloc(&v->h_pin, need + need / 4, cudaHostAllocDefault));
        v->h_pin_cap = need + need / 4;
    }
    uint8_t *hp = (uint8_t *)v->h_pin;
    (void)from_pinned_copy;
    memcpy(hp, spectra, so * 4);
    memcpy(hp + so * 4, spec_off.data(), n_blocks * 4);
    memcpy(hp + so * 4 + n_blocks * 4, out_off.data(), (n_blocks + 1) * 4);
    for (int b = 0; b < n_blocks; b++) hp[so * 4 + (2 * n_blocks + 1) * 4 + b] = flags[b] ? 1 : 0;
    VCK(cudaMemcpyAsync(v->spectra.d, hp, so * 4, cudaMemcpyHostToDevice, v->stream));
    VCK(cudaMemcpyAsync(v->spec_off.d, hp + so * 4, n_blocks * 4, cudaMemcpyHostToDevice, v->stream));
    VCK(cudaMemcpyAsync(v->out_off.d, hp + so * 4 + n_blocks * 4, (n_blocks + 1) * 4, cudaMemcpyHostToDevice, v->stream));
    VCK(cudaMemcpyAsync(v->flags.d, hp + so * 4 + (2 * n_blocks + 1) * 4, n_blocks, cudaMemcpyHostToDevice, v->stream));
    v->h2d += (long long)need;
    memset(&P, 0, sizeof(P));
    P.channels = v->channels; P.n_blocks = n_blocks; P.run_len = 8;
    for (int s = 0; s < 2; s++) {
        P.M[s] = v->bs[s] / 2;
        P.tw[s] = v->d_tw[s]; P.fftw[s] = v->d_fftw[s]; P.win[s] = v->d_win[s];
    }
    P.flags = (const uint8_t *)v->flags.d;
    P.spec_off = (const uint32_t *)v->spec_off.d;
    P.out_off = (const uint32_t *)v->out_off.d;
    P.spectra = (const float *)v->spectra.d;
    P.state_in = v->d_state[v->state_cur];
    P.state_W = v->state_W;
    P.state_out = v->d_state[v->state_cur ^ 1];
    P.pcm = (float *)v->pcm.d;
    P.pcm_stride = oo;
    P.interleave = interleave;
    return 0;
}

static int vorbis_launch(Eav1Vorbis *v, const VorbisParams &P)
{
    const dim3 grid((P.n_blocks + P.run_len - 1) / P.run_len, P.channels);
    const size_t smem = 3 * (size_t)P.M[1] * sizeof(float);
    vorbis_kernel<<<grid, 128, smem, v->stream>>>(P);
    v->launches++;
    VCK(cudaGetLastError());
    return 0;
}

/* Host entry point: flags[n_blocks] (0 short / 1 long), spectra block-major [b][channel][blocksize/2],
 * pcm out planar [channel][cap_samples] or interleaved [sample][channels] (src/easyav1.c:2471-2492).
 * The overlap state carries over to the next call like vorbis_dsp_state.pcm does. */
int eav1cu_vorbis_imdct_ola(Eav1Vorbis *v, const uint8_t *flags, int n_blocks, const float *spectra,
                            float *pcm, size_t cap_samples, int interleave, size_t *n_samples)
{
    if (!v || !pcm) return EAV1CU_ERR_INVAL;
    VorbisParams P;
    size_t ns = 0;
    int r = vorbis_prepare(v, P, flags, n_blocks, spectra, cap_samples, interleave, &ns, true);
    if (r) return r;
    if ((r = vorbis_launch(v, P))) return r;
    if (ns) {
        if (interleave) VCK(cudaMemcpyAsync(pcm, v->pcm.d, ns * v->channels * 4, cudaMemcpyDeviceToHost, v->stream));
        else VCK(cudaMemcpy2DAsync(pcm, cap_samples * 4, v->pcm.d, ns * 4, ns * 4, v->channels, cudaMemcpyDeviceToHost, v->stream));
        v->d2h += (long long)(ns * v->channels * 4);
    }
    VCK(cudaStreamSynchronize(v->stream));
    v->state_cur ^= 1;
    v->state_W = flags[n_blocks - 1] ? 1 : 0;
    if (n_samples) *n_samples = ns;
    return 0;
}

/* Measurement: keep a batch resident in device memory and replay the kernel on it. The overlap
 * state the batch starts from is frozen at stage time; replays do not advance it. */
int eav1cu_vorbis_stage(Eav1Vorbis *v, const uint8_t *flags, int n_blocks, const float *spectra, int interleave,
                        size_t *n_samples)
{
    if (!v) return EAV1CU_ERR_INVAL;
    size_t ns = 0;
    int r = vorbis_prepare(v, v->staged, flags, n_blocks, spectra, (size_t)-1, interleave, &ns, false);
    if (r) return r;
    VCK(cudaStreamSynchronize(v->stream));
    v->have_staged = true;
    v->staged_samples = ns;
    if (n_samples) *n_samples = ns;
    return 0;
}

int eav1cu_vorbis_run_staged(Eav1Vorbis *v)
{
    if (!v || !v->have_staged) return EAV1CU_ERR_INVAL;
    VCK(cudaSetDevice(v->device));
    return vorbis_launch(v, v->staged);
}

int eav1cu_vorbis_fetch(Eav1Vorbis *v, float *pcm, size_t cap_floats)
{
    if (!v || !v->have_staged || !pcm) return EAV1CU_ERR_INVAL;
    const size_t n = v->staged_samples * v->channels;
    if (n > cap_floats) return EAV1CU_ERR_INVAL;
    VCK(cudaSetDevice(v->device));
    VCK(cudaMemcpyAsync(pcm, v->pcm.d, n * 4, cudaMemcpyDeviceToHost, v->stream));
    VCK(cudaStreamSynchronize(v->stream));
    return 0;
}

int eav1cu_vorbis_sync(Eav1Vorbis *v)
{
    if (!v) return EAV1CU_ERR_INVAL;
    VCK(cudaSetDevice(v->device));
    VCK(cudaStreamSynchronize(v->stream));
    return v->err;
}

} // extern "C"
