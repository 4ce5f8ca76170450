// engine.cu — libeasyav1_cuda.so: context, device picture pool, command upload, level-by-level
// launch of the reconstruction kernels and the whole-frame post-filter kernels, C ABI.
// Built for sm_100a only (see easyav1_b200/csrc/Makefile). No CPU fallback.
#include <cuda_runtime.h>
#include <stdio.h>
#include <string.h>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/easyav1_cuda.h"
#include "dev_common.cuh"
#include "exec.cuh"
#include "postfilter.cuh"
#include "sched.h"

using namespace eav1;

// ------------------------------------------------------------------ kernels

template <typename BD>
__global__ void __launch_bounds__(128)
recon_level_kernel(FrameCtx fc, const Eav1Cmd *__restrict__ cmds, const Chain *__restrict__ chains)
{
    __shared__ CtaScratch sm;
    __shared__ Eav1Cmd cur;
    const Chain ch = chains[blockIdx.x];
    for (uint32_t k = 0; k < ch.count; k++) {
        // stage the command in shared memory (60 bytes = 15 words)
        if (threadIdx.x < sizeof(Eav1Cmd) / 4)
            reinterpret_cast<uint32_t *>(&cur)[threadIdx.x] =
                reinterpret_cast<const uint32_t *>(&cmds[ch.first + k])[threadIdx.x];
        __syncthreads();
        exec_cmd<BD>(fc, cur, sm, threadIdx.x, blockDim.x);
        __syncthreads();
    }
}

template <typename pixel>
__global__ void __launch_bounds__(128)
deblock_kernel(PfParams pp, const Eav1LfRec *__restrict__ recs, int n)
{
    const int gid = blockIdx.x * blockDim.x + threadIdx.x;
    const int ri = gid >> 5;
    if (ri >= n) return;
    lf_unit<pixel>(pp, recs[ri], gid & 31);
}

template <typename pixel>
__global__ void __launch_bounds__(256)
cdef_kernel(PfParams pp)
{
    __shared__ CdefSmem sm[8];
    const int g = threadIdx.x >> 5;
    const int bx = (blockIdx.x * 8 + g) * 2, by = blockIdx.y * 2;
    cdef_block<pixel>(pp, bx, by, bx < pp.bw4 && by < pp.bh4, sm[g], threadIdx.x & 31, 32);
}

template <typename pixel>
__global__ void __launch_bounds__(128)
lr_kernel(PfParams pp, const Eav1LrRec *__restrict__ recs)
{
    __shared__ LrSmem sm;
    __shared__ Eav1LrRec r;
    if (threadIdx.x < sizeof(Eav1LrRec) / 4)
        reinterpret_cast<uint32_t *>(&r)[threadIdx.x] = reinterpret_cast<const uint32_t *>(&recs[blockIdx.y])[threadIdx.x];
    __syncthreads();
    lr_tile<pixel>(pp, r, blockIdx.x, sm, threadIdx.x, blockDim.x);
}

// ------------------------------------------------------------------ context

struct Picture {
    bool used = false;
    int w = 0, h = 0, layout = 0, bpc = 0;
    uint8_t *base = nullptr;
    size_t bytes = 0;
    Plane pl[3];
};

struct DevBuf {
    uint8_t *d = nullptr; size_t cap = 0;
    uint8_t *h = nullptr; size_t hcap = 0; // pinned staging
};

struct Eav1cuCtx {
    int device = 0;
    cudaStream_t stream = nullptr;
    Eav1Tables *d_tab = nullptr;
    bool have_tab = false;
    uint8_t *d_masks = nullptr; size_t masks_bytes = 0;
    std::vector<Picture> pics;
    DevBuf cmds, chains, payload, arena, lfrec, lflvl, lrrec, lfmask;
    Picture work[2];           // post-filter ping-pong pictures
    std::vector<Eav1LfRec> lf_sorted;
    Scheduler sched;
    Schedule plan;
    int err = 0;
    std::string errmsg;
    std::mutex mu;
    long long n_launches = 0;
};

static int fail(Eav1cuCtx *c, int code, const char *what, cudaError_t e = cudaSuccess)
{
    if (c && !c->err) {
        c->err = code;
        c->errmsg = what;
        if (e != cudaSuccess) { c->errmsg += ": "; c->errmsg += cudaGetErrorString(e); }
    }
    return code;
}
#define CK(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return fail(ctx, EAV1CU_ERR_CUDA, #call, e_); } while (0)

static int ensure(Eav1cuCtx *ctx, DevBuf &b, size_t bytes, bool pinned_too)
{
    if (bytes > b.cap) {
        if (b.d) cudaFree(b.d);
        size_t cap = bytes + bytes / 2 + 4096;
        CK(cudaMalloc(&b.d, cap));
        b.cap = cap;
    }
    if (pinned_too && bytes > b.hcap) {
        if (b.h) cudaFreeHost(b.h);
        size_t cap = bytes + bytes / 2 + 4096;
        CK(cudaHostAlloc(&b.h, cap, cudaHostAllocDefault));
        b.hcap = cap;
    }
    return 0;
}

static void free_buf(DevBuf &b)
{
    if (b.d) cudaFree(b.d);
    if (b.h) cudaFreeHost(b.h);
    b = DevBuf();
}

static int upload(Eav1cuCtx *ctx, DevBuf &b, const void *src, size_t bytes)
{
    if (!bytes) return 0;
    int r = ensure(ctx, b, bytes, true);
    if (r) return r;
    memcpy(b.h, src, bytes);
    CK(cudaMemcpyAsync(b.d, b.h, bytes, cudaMemcpyHostToDevice, ctx->stream));
    return 0;
}

static int pic_setup(Eav1cuCtx *ctx, Picture &p, int w, int h, int layout, int bpc)
{
    const int hbd = bpc > 8;
    const int ss_hor = layout != 3, ss_ver = layout == 1;
    // pad to 128 luma pixels so 64x64 transforms / CDEF blocks at the frame edge stay inside
    const int aw = (w + 127) & ~127, ah = (h + 127) & ~127;
    const int cw = layout ? aw >> ss_hor : 0, chh = layout ? ah >> ss_ver : 0;
    const size_t ystride = (size_t)aw << hbd, cstride = (size_t)cw << hbd;
    const size_t ybytes = ystride * ah, cbytes = cstride * chh;
    const size_t total = ybytes + 2 * cbytes + 256;
    if (p.base && p.bytes < total) { cudaFree(p.base); p.base = nullptr; }
    if (!p.base) { CK(cudaMalloc(&p.base, total)); p.bytes = total; }
    p.w = w; p.h = h; p.layout = layout; p.bpc = bpc;
    p.pl[0].p = p.base; p.pl[0].stride = (int)ystride; p.pl[0].w = aw; p.pl[0].h = ah;
    p.pl[1].p = p.base + ybytes; p.pl[1].stride = (int)cstride; p.pl[1].w = cw; p.pl[1].h = chh;
    p.pl[2].p = p.base + ybytes + cbytes; p.pl[2].stride = (int)cstride; p.pl[2].w = cw; p.pl[2].h = chh;
    return 0;
}

// ------------------------------------------------------------------ C ABI

extern "C" {

int eav1cu_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) return 0;
    return n;
}

int eav1cu_ctx_create(Eav1cuCtx **out, int device)
{
    if (!out) return EAV1CU_ERR_INVAL;
    *out = nullptr;
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n <= 0 || device >= n) return EAV1CU_ERR_NO_DEVICE;
    Eav1cuCtx *ctx = new Eav1cuCtx();
    ctx->device = device;
    if (cudaSetDevice(device) != cudaSuccess ||
        cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess ||
        cudaMalloc(&ctx->d_tab, sizeof(Eav1Tables)) != cudaSuccess) {
        delete ctx;
        return EAV1CU_ERR_CUDA;
    }
    *out = ctx;
    return 0;
}

void eav1cu_ctx_destroy(Eav1cuCtx *ctx)
{
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    for (auto &p : ctx->pics) if (p.base) cudaFree(p.base);
    for (auto &p : ctx->work) if (p.base) cudaFree(p.base);
    free_buf(ctx->cmds); free_buf(ctx->chains); free_buf(ctx->payload); free_buf(ctx->arena);
    free_buf(ctx->lfrec); free_buf(ctx->lflvl); free_buf(ctx->lrrec); free_buf(ctx->lfmask);
    if (ctx->d_tab) cudaFree(ctx->d_tab);
    if (ctx->d_masks) cudaFree(ctx->d_masks);
    cudaStreamDestroy(ctx->stream);
    delete ctx;
}

int eav1cu_last_error(Eav1cuCtx *ctx, char *msg, size_t cap)
{
    if (!ctx) return EAV1CU_ERR_INVAL;
    if (msg && cap) snprintf(msg, cap, "%s", ctx->errmsg.c_str());
    return ctx->err;
}

int eav1cu_host_alloc(void **p, size_t bytes)
{
    if (!p) return EAV1CU_ERR_INVAL;
    return cudaHostAlloc(p, bytes, cudaHostAllocPortable) == cudaSuccess ? 0 : EAV1CU_ERR_NOMEM;
}

void eav1cu_host_free(void *p) { if (p) cudaFreeHost(p); }

long long eav1cu_launch_count(Eav1cuCtx *ctx) { return ctx ? ctx->n_launches : 0; }

int eav1cu_set_tables(Eav1cuCtx *ctx, const Eav1Tables *t)
{
    if (!ctx || !t) return EAV1CU_ERR_INVAL;
    CK(cudaSetDevice(ctx->device));
    CK(cudaMemcpyAsync(ctx->d_tab, t, sizeof(*t), cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    ctx->have_tab = true;
    return 0;
}

int eav1cu_set_masks(Eav1cuCtx *ctx, const uint8_t *blob, size_t bytes)
{
    if (!ctx || !blob) return EAV1CU_ERR_INVAL;
    CK(cudaSetDevice(ctx->device));
    if (ctx->d_masks) cudaFree(ctx->d_masks);
    CK(cudaMalloc(&ctx->d_masks, bytes));
    CK(cudaMemcpy(ctx->d_masks, blob, bytes, cudaMemcpyHostToDevice));
    ctx->masks_bytes = bytes;
    return 0;
}

int eav1cu_sync(Eav1cuCtx *ctx)
{
    if (!ctx) return EAV1CU_ERR_INVAL;
    CK(cudaStreamSynchronize(ctx->stream));
    return ctx->err;
}

int eav1cu_pic_alloc(Eav1cuCtx *ctx, int w, int h, int layout, int bpc)
{
    if (!ctx || w <= 0 || h <= 0) return EAV1CU_ERR_INVAL;
    std::lock_guard<std::mutex> g(ctx->mu);
    CK(cudaSetDevice(ctx->device));
    int id = -1;
    for (size_t i = 0; i < ctx->pics.size(); i++)
        if (!ctx->pics[i].used) { id = (int)i; break; }
    if (id < 0) { ctx->pics.emplace_back(); id = (int)ctx->pics.size() - 1; }
    Picture &p = ctx->pics[id];
    int r = pic_setup(ctx, p, w, h, layout, bpc);
    if (r) return r;
    p.used = true;
    return id;
}

int eav1cu_pic_release(Eav1cuCtx *ctx, int id)
{
    if (!ctx || id < 0 || id >= (int)ctx->pics.size()) return EAV1CU_ERR_INVAL;
    std::lock_guard<std::mutex> g(ctx->mu);
    ctx->pics[id].used = false; // memory is kept for reuse
    return 0;
}

static int pic_copy(Eav1cuCtx *ctx, int id, void *const data[3], const ptrdiff_t stride[2], bool to_device)
{
    if (!ctx || id < 0 || id >= (int)ctx->pics.size() || !ctx->pics[id].used) return EAV1CU_ERR_INVAL;
    CK(cudaSetDevice(ctx->device));
    Picture &p = ctx->pics[id];
    const int hbd = p.bpc > 8;
    const int ss_hor = p.layout != 3, ss_ver = p.layout == 1;
    for (int pl = 0; pl < (p.layout ? 3 : 1); pl++) {
        const int w = pl ? (p.w + ss_hor) >> ss_hor : p.w;
        const int h = pl ? (p.h + ss_ver) >> ss_ver : p.h;
        if (to_device)
            CK(cudaMemcpy2DAsync(p.pl[pl].p, p.pl[pl].stride, data[pl], stride[!!pl], (size_t)w << hbd, h,
                                 cudaMemcpyHostToDevice, ctx->stream));
        else
            CK(cudaMemcpy2DAsync(data[pl], stride[!!pl], p.pl[pl].p, p.pl[pl].stride, (size_t)w << hbd, h,
                                 cudaMemcpyDeviceToHost, ctx->stream));
    }
    CK(cudaStreamSynchronize(ctx->stream));
    return ctx->err;
}

int eav1cu_pic_upload(Eav1cuCtx *ctx, int id, const void *const data[3], const ptrdiff_t stride[2])
{
    return pic_copy(ctx, id, (void *const *)data, stride, true);
}

int eav1cu_pic_download(Eav1cuCtx *ctx, int id, void *const data[3], const ptrdiff_t stride[2])
{
    return pic_copy(ctx, id, data, stride, false);
}

int eav1cu_arena_upload(Eav1cuCtx *ctx, size_t off, const void *src, size_t bytes)
{
    if (!ctx) return EAV1CU_ERR_INVAL;
    CK(cudaSetDevice(ctx->device));
    if (off + bytes > ctx->arena.cap) {
        // grow, preserving nothing (tests upload before the first submit)
        int r = ensure(ctx, ctx->arena, off + bytes, false);
        if (r) return r;
    }
    CK(cudaMemcpy(ctx->arena.d + off, src, bytes, cudaMemcpyHostToDevice));
    return 0;
}

int eav1cu_arena_download(Eav1cuCtx *ctx, size_t off, void *dst, size_t bytes)
{
    if (!ctx || off + bytes > ctx->arena.cap) return EAV1CU_ERR_INVAL;
    CK(cudaSetDevice(ctx->device));
    CK(cudaStreamSynchronize(ctx->stream));
    CK(cudaMemcpy(dst, ctx->arena.d + off, bytes, cudaMemcpyDeviceToHost));
    return 0;
}

static void fill_planes(const Picture &p, Plane out[3]) { for (int i = 0; i < 3; i++) out[i] = p.pl[i]; }


static int postfilter_run(Eav1cuCtx *ctx, const Eav1FrameDesc *fd, Picture &A, Picture &cur)
{
    const bool hbd = fd->bpc > 8;
    const bool has_cdef = fd->cdef_enable != 0, has_lr = fd->n_lr_recs > 0;
    PfParams pp;
    memset(&pp, 0, sizeof(pp));
    pp.bitdepth_max = (1 << fd->bpc) - 1;
    pp.layout = fd->layout; pp.ss_hor = fd->layout != 3; pp.ss_ver = fd->layout == 1;
    pp.w = fd->w; pp.h = fd->h; pp.bw4 = fd->bw4; pp.bh4 = fd->bh4;
    pp.b4_stride = fd->b4_stride;
    memcpy(pp.lut_e, fd->lut_e, 64); memcpy(pp.lut_i, fd->lut_i, 64);
    pp.sb128w = fd->sb128w; pp.cdef_damping = fd->cdef_damping;
    memcpy(pp.cdef_y, fd->cdef_y_strength, 8); memcpy(pp.cdef_uv, fd->cdef_uv_strength, 8);
    pp.tab = ctx->d_tab;
    for (int i = 0; i < 3; i++) pp.a[i] = A.pl[i];
    int r;
    if (fd->n_lf_recs) {
        // partition the recorded calls: all column-edge filters run before all row-edge filters
        std::vector<Eav1LfRec> &v = ctx->lf_sorted;
        v.resize(fd->n_lf_recs);
        size_t n0 = 0;
        for (size_t i = 0; i < fd->n_lf_recs; i++) if (fd->lf_recs[i].dir == 0) v[n0++] = fd->lf_recs[i];
        size_t n1 = n0;
        for (size_t i = 0; i < fd->n_lf_recs; i++) if (fd->lf_recs[i].dir != 0) v[n1++] = fd->lf_recs[i];
        if ((r = upload(ctx, ctx->lfrec, v.data(), v.size() * sizeof(Eav1LfRec)))) return r;
        if ((r = upload(ctx, ctx->lflvl, fd->lf_level, fd->lf_level_bytes))) return r;
        pp.level = ctx->lflvl.d;
        const Eav1LfRec *d = reinterpret_cast<const Eav1LfRec *>(ctx->lfrec.d);
        for (int pass = 0; pass < 2; pass++) {
            const size_t off = pass ? n0 : 0, cnt = pass ? n1 - n0 : n0;
            if (!cnt) continue;
            const unsigned grid = (unsigned)((cnt * 32 + 127) / 128);
            if (hbd) deblock_kernel<uint16_t><<<grid, 128, 0, ctx->stream>>>(pp, d + off, (int)cnt);
            else     deblock_kernel<uint8_t><<<grid, 128, 0, ctx->stream>>>(pp, d + off, (int)cnt);
            ctx->n_launches++;
        }
    }
    if (has_cdef) {
        Picture &dst = has_lr ? ctx->work[1] : cur;
        for (int i = 0; i < 3; i++) pp.b[i] = dst.pl[i];
        if ((r = upload(ctx, ctx->lfmask, fd->lf_mask, (size_t)fd->sb128w * fd->sb128h * 1348))) return r;
        pp.lfmask = ctx->lfmask.d;
        dim3 grid((fd->bw4 / 2 + 8) / 8, (fd->bh4 + 1) / 2);
        if (hbd) cdef_kernel<uint16_t><<<grid, 256, 0, ctx->stream>>>(pp);
        else     cdef_kernel<uint8_t><<<grid, 256, 0, ctx->stream>>>(pp);
        ctx->n_launches++;
    }
    if (has_lr) {
        Picture &src = has_cdef ? ctx->work[1] : A;
        for (int i = 0; i < 3; i++) { pp.b[i] = src.pl[i]; pp.p[i] = cur.pl[i]; }
        CK(cudaMemcpyAsync(cur.base, src.base, std::min(cur.bytes, src.bytes), cudaMemcpyDeviceToDevice, ctx->stream));
        if ((r = upload(ctx, ctx->lrrec, fd->lr_recs, fd->n_lr_recs * sizeof(Eav1LrRec)))) return r;
        dim3 grid(384 / LR_TW, (unsigned)fd->n_lr_recs);
        const Eav1LrRec *d = reinterpret_cast<const Eav1LrRec *>(ctx->lrrec.d);
        if (hbd) lr_kernel<uint16_t><<<grid, 128, 0, ctx->stream>>>(pp, d);
        else     lr_kernel<uint8_t><<<grid, 128, 0, ctx->stream>>>(pp, d);
        ctx->n_launches++;
    }
    CK(cudaGetLastError());
    return 0;
}



int eav1cu_frame_submit(Eav1cuCtx *ctx, const Eav1FrameDesc *fd,
                        const Eav1Cmd *cmds, size_t n_cmds,
                        const void *payload, size_t payload_bytes, size_t arena_bytes)
{
    if (!ctx || !fd) return EAV1CU_ERR_INVAL;
    if (ctx->err) return ctx->err;
    if (!ctx->have_tab) return fail(ctx, EAV1CU_ERR_NOTABLES, "eav1cu_set_tables not called");
    if (fd->cur < 0 || fd->cur >= (int)ctx->pics.size() || !ctx->pics[fd->cur].used)
        return fail(ctx, EAV1CU_ERR_INVAL, "bad current picture id");
    CK(cudaSetDevice(ctx->device));
    Picture &cur = ctx->pics[fd->cur];
    const bool hbd = fd->bpc > 8;
    const bool has_cdef = fd->cdef_enable != 0, has_lr = fd->n_lr_recs > 0;
    // recon + deblock happen in A; CDEF / LR then move the frame towards `cur` out of place
    Picture *A = &cur;
    if (has_cdef || has_lr) {
        int r0 = pic_setup(ctx, ctx->work[0], cur.w, cur.h, cur.layout, cur.bpc);
        if (r0) return r0;
        if (has_cdef && has_lr && (r0 = pic_setup(ctx, ctx->work[1], cur.w, cur.h, cur.layout, cur.bpc))) return r0;
        A = &ctx->work[0];
    }

    FrameCtx fc;
    memset(&fc, 0, sizeof(fc));
    fill_planes(*A, fc.cur);
    for (int r = 0; r < 7; r++) {
        const int id = fd->ref[r];
        if (id >= 0 && id < (int)ctx->pics.size() && ctx->pics[id].base) fill_planes(ctx->pics[id], fc.ref[r]);
        else fill_planes(*A, fc.ref[r]);
    }
    fill_planes(*A, fc.ref[7]);
    fc.bitdepth_max = (1 << fd->bpc) - 1;
    fc.ss_hor = fd->layout != 3;
    fc.ss_ver = fd->layout == 1;
    fc.tab = ctx->d_tab;
    fc.masks = ctx->d_masks;

    if (n_cmds) {
        int pw[3] = { A->pl[0].w, A->pl[1].w, A->pl[2].w };
        int ph[3] = { A->pl[0].h, A->pl[1].h, A->pl[2].h };
        ctx->sched.reset(pw, ph);
        ctx->sched.build(cmds, n_cmds, ctx->plan);
        int r;
        if ((r = upload(ctx, ctx->cmds, cmds, n_cmds * sizeof(Eav1Cmd)))) return r;
        if ((r = upload(ctx, ctx->chains, ctx->plan.chains.data(), ctx->plan.chains.size() * sizeof(Chain)))) return r;
        if ((r = upload(ctx, ctx->payload, payload, payload_bytes))) return r;
        if ((r = ensure(ctx, ctx->arena, arena_bytes + 16, false))) return r;
        fc.arena = ctx->arena.d;
        fc.payload = ctx->payload.d;
        const Eav1Cmd *d_cmds = reinterpret_cast<const Eav1Cmd *>(ctx->cmds.d);
        const Chain *d_chains = reinterpret_cast<const Chain *>(ctx->chains.d);
        for (size_t l = 0; l + 1 < ctx->plan.level_start.size(); l++) {
            const uint32_t a = ctx->plan.level_start[l], b = ctx->plan.level_start[l + 1];
            if (a == b) continue;
            if (hbd) recon_level_kernel<BD16><<<b - a, 128, 0, ctx->stream>>>(fc, d_cmds, d_chains + a);
            else     recon_level_kernel<BD8><<<b - a, 128, 0, ctx->stream>>>(fc, d_cmds, d_chains + a);
            ctx->n_launches++;
        }
        CK(cudaGetLastError());
    }
    int r = postfilter_run(ctx, fd, *A, cur);
    if (r) return r;
    return ctx->err;
}

} // extern "C"
