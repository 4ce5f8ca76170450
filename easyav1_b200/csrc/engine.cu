// engine.cu — libeasyav1_cuda.so: context, device picture pool, command upload, per-frame plan
// (phase-1 items, residual items, dataflow units and their ticket order), the kernels of a frame
// (motion compensation + residuals, dataflow reconstruction, deblock, CDEF, loop restoration), C ABI.
// Built for sm_100a only (see easyav1_b200/csrc/Makefile). No CPU fallback.
#include <cuda_runtime.h>
#include <stdio.h>
#include <string.h>
#include <stdlib.h>
#include <algorithm>
#include <atomic>
#include <chrono>
#include <mutex>
#include <thread>
#include <string>
#include <vector>

#include "../../include/easyav1_cuda.h"
#include "dev_common.cuh"
#include "exec.cuh"
#include "postfilter.cuh"
#include "filmgrain.cuh"
#include "convert.cuh"
#include "sched.h"
#include "coverage.h"

using namespace eav1;

// Launch epochs are unique across every context of the process: a done-flag array that lands on
// memory a destroyed context used can then never hold a value that looks current.
static std::atomic<uint32_t> g_epoch{0};
// process-wide totals over every context (eav1cu_global_counters): launches, host->device bytes, device->host bytes, frames
static std::atomic<long long> g_total[4];

// Many decoders of one process submit frames from their own threads. The CUDA calls of one frame (a copy, about ten
// launches, a few event records) are short; issued from many threads at once they queue up inside the driver and
// every call becomes slow for everybody. One process-wide lock around each frame's burst of calls keeps them
// back to back (measured with 16 streams on one B200: see DESIGN.md); the CPU work of plan_build stays outside.
// EAV1_TIMING=1: per call-site totals of the CUDA runtime calls on the frame path, printed when the process exits
// (development aid for the multi-stream host path).
enum { TC_SETDEV, TC_H2D_PACK, TC_PIC_COPY, TC_EVREC, TC_EVSYNC_SLOT, TC_EVSYNC_FENCE, TC_WAITEV, TC_D2D, TC_ALLOC, TC_LAUNCHES, TC_N };
struct CallStats {
    std::atomic<long long> n[TC_N], ns[TC_N], mx[TC_N];
    const bool on = getenv("EAV1_TIMING") != nullptr;
    ~CallStats() {
        if (!on) return;
        static const char *nm[TC_N] = { "cudaSetDevice", "cudaMemcpyAsync (packed lists)", "picture copy", "cudaEventRecord", "cudaEventSynchronize (slot)",
                                        "cudaEventSynchronize (fence)", "cudaStreamWaitEvent", "cudaMemcpyAsync D2D", "cudaMalloc/Free/HostAlloc", "kernel launches of a frame" };
        for (int i = 0; i < TC_N; i++)
            if (n[i]) fprintf(stderr, "EAV1 TIMING: %-34s %9lld calls, mean %8.1f us, max %9.1f us, total %8.1f ms\n", nm[i], (long long)n[i],
                              1e-3 * ns[i] / n[i], 1e-3 * mx[i], 1e-6 * ns[i]);
    }
    void add(int i, long long d) {
        n[i]++; ns[i] += d;
        long long m = mx[i];
        while (d > m && !mx[i].compare_exchange_weak(m, d)) {}
    }
};
static CallStats g_calls;
struct CallTimer {
    int i; std::chrono::steady_clock::time_point t0;
    explicit CallTimer(int idx) : i(idx) { if (g_calls.on) t0 = std::chrono::steady_clock::now(); }
    ~CallTimer() { if (g_calls.on) g_calls.add(i, std::chrono::duration_cast<std::chrono::nanoseconds>(std::chrono::steady_clock::now() - t0).count()); }
};
#define TCK(idx, call) do { CallTimer ct_(idx); CK(call); } while (0)

// How host threads wait for a frame (dav1d_get_picture -> eav1cu_fence_wait, and the wait for a staging slot).
// A lone decoder spins inside the driver: waking a sleeping thread costs tens of microseconds per frame, which is a
// quarter of the time of a 360p frame in easyAV1's one-picture-per-call loop (1,730 -> 2,200 frames/s) and 7 % at
// 1080p. With many decoders per process the waiters must sleep (cudaEventBlockingSync): spinning waiters slow every
// CUDA call of the submitting threads down. The mode is chosen when a context is created or taken from the cache,
// by the number of live contexts; EAV1_SPIN_WAIT=1 / 0 forces spinning / sleeping.
static std::atomic<int> g_live_ctx{0};
static bool want_spin(int live_including_this)
{
    if (const char *e = getenv("EAV1_SPIN_WAIT")) return atoi(e) != 0;
    static const unsigned cores = std::thread::hardware_concurrency();
    return live_including_this <= 1 && cores >= 4;   // a polling waiter needs a core of its own
}

struct SubmitMutex {      // EAV1_NO_SUBMIT_LOCK=1 turns it off (A/B measurements)
    std::mutex m;
    const bool on = getenv("EAV1_NO_SUBMIT_LOCK") == nullptr;
    void lock() { if (on) m.lock(); }
    void unlock() { if (on) m.unlock(); }
};
static SubmitMutex g_submit_mu;

// Which launches carry the programmatic-dependency attribute. EAV1_PDL (A/B runs): bit 0 post-filter kernels, bit 1 the
// dataflow kernel, bit 2 motion compensation, bit 3 early release. Default: all three, and early release while the
// process has a single decoder (one 1080p stream 13.07 -> 12.16 ms per 120 frames). With several decoders on the GPU
// early release is off: CTAs parked in pdl_wait() hold the slots the other streams' kernels would run in (8 streams
// 24.1k -> 18.3k frames/s with it, 23.8k -> 24.1k with the attribute alone).
static int pdl_mask()
{
    static const int m = [] { const char *e = getenv("EAV1_PDL"); return e ? atoi(e) : 7; }();
    return m;
}
static void pdl_set_early(int device, int live, cudaStream_t st)
{
    static std::mutex mu;
    static int cur[64];          // per device: 0 unknown, 1 off, 2 on
    const char *e = getenv("EAV1_PDL");
    const int want = e ? ((atoi(e) & 8) != 0) : live <= 1;
    std::lock_guard<std::mutex> g(mu);
    if (device < 0 || device >= 64 || cur[device] == want + 1) return;
    if (cudaMemcpyToSymbolAsync(c_pdl_early, &want, sizeof(want), 0, cudaMemcpyHostToDevice, st) == cudaSuccess &&
        cudaStreamSynchronize(st) == cudaSuccess)
        cur[device] = want + 1;
}
template <typename... KArgs, typename... Args>
static cudaError_t launch_k(void (*kernel)(KArgs...), dim3 grid, dim3 block, cudaStream_t st, bool pdl, Args &&...args)
{
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = grid; cfg.blockDim = block; cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

// ------------------------------------------------------------------ kernels

// One launch reconstructs a whole frame. CTAs draw chain tickets in decode order (which is a
// topological order of the dependency graph, sched.h), wait until the listed producer chains
// have published their done flag, run the chain, publish their own flag. A CTA only ever waits
// for tickets smaller than its own, and those are held by CTAs that are already running, so the
// kernel makes progress for any grid size. Flags hold the launch epoch, so they never need
// clearing; ctrl[0] is the ticket counter, ctrl[1] counts retired CTAs (the last one re-arms both).
// What a worker needs to run one chain, fetched with a single 32-byte load once it holds a ticket.
struct alignas(32) ChainRec {
    uint32_t first, count;      // commands [first, first+count)
    uint32_t dep_off, n_deps;   // producers: deps[dep_off .. dep_off+n_deps) (chain ids)
    uint32_t id;                // this chain's id (index of its done flag)
    uint32_t pad[3];            // pad[0]: some later unit waits for this one
};

struct ReconArgs {
    const Eav1Cmd *cmds;
    const uint32_t *deps;
    const ChainRec *big;          // chains run by a whole CTA, in decode order
    const ChainRec *small;        // chains run by one warp, in decode order
    uint32_t *done;
    uint32_t *ctrl;               // [0] big tickets, [1] small tickets, [2] retired workers
    uint32_t n_big, n_small;
    uint32_t n_big_ctas, n_workers;
    uint32_t epoch;
    uint32_t debug;               // count executed units in ctrl[4] (EAV1_DEBUG_CHECK)
    unsigned long long *hist;     // development aid: per (op, size) cycle histogram behind the trace words
    unsigned long long *trace;    // development aid (eav1cu_staged_trace): 6 words per unit, else null
};

#define EAV1_TRACE_WORDS 12
__device__ __forceinline__ unsigned long long global_ns()
{
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

__device__ __forceinline__ uint32_t ld_acquire(const uint32_t *p)
{
    uint32_t v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

__device__ __forceinline__ void st_release(uint32_t *p, uint32_t v)
{
    asm volatile("st.release.gpu.global.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}

#define EAV1_CMD_STAGE 8   // commands fetched into shared memory per round trip

struct WorkerSmem {
    Eav1Cmd cmd[EAV1_CMD_STAGE];
    uint32_t ticket;
};

// One work group (a CTA of 128 threads or a single warp) draining one ticket list.
template <typename BD>
__device__ __forceinline__ void recon_worker(const FrameCtx &fc, const ReconArgs &ra, const ChainRec *recs,
                                             uint32_t n_list, uint32_t *counter, CtaScratch &sm, WorkerSmem &ws,
                                             int tid, int nthr)
{
    uint32_t next = 0;
    bool dep_waited = false;    // pdl_wait() once, before the first command touches pixels: tickets, unit records and
                                // commands are this launch's own (or uploaded) state
    if (tid == 0) next = atomicAdd(counter, 1u);
    for (;;) {
        if (tid == 0) ws.ticket = next;
        CTA_SYNC();
        const uint32_t k = ws.ticket;
        if (k >= n_list) break;
        const uint4 r0 = *reinterpret_cast<const uint4 *>(&recs[k]);
        const uint32_t id = recs[k].id, fed = recs[k].pad[0];
        // draw the following ticket now: the round trip hides behind this chain's work
        if (tid == 0) next = atomicAdd(counter, 1u);
        const uint32_t first = r0.x, count = r0.y, d0 = r0.z, nd = r0.w;
        unsigned long long t_ticket = 0, t_ready = 0;
        if (ra.trace && tid == 0) t_ticket = global_ns();
        // fetch the unit's first commands while its producers are still being polled
        {
            const uint32_t n = min(count, (uint32_t)EAV1_CMD_STAGE);
            const uint32_t *src = reinterpret_cast<const uint32_t *>(&ra.cmds[first]);
            uint32_t *dst = reinterpret_cast<uint32_t *>(ws.cmd);
            for (uint32_t i = tid; i < n * (sizeof(Eav1Cmd) / 4); i += nthr) dst[i] = src[i];
        }
        for (uint32_t d = tid; d < nd; d += nthr) {
            const uint32_t *flag = &ra.done[ra.deps[d0 + d]];
            unsigned ns = 20;
#ifndef EAV1_POLL_MAX_NS
#define EAV1_POLL_MAX_NS 80
#endif
            while (ld_acquire(flag) != ra.epoch) { if (EAV1_POLL_MAX_NS) { __nanosleep(ns); if (ns < EAV1_POLL_MAX_NS) ns <<= 1; } }
        }
        CTA_SYNC();
        if (ra.trace && tid == 0) {
            t_ready = global_ns();
            unsigned long long last = 0, who = ~0ull;
            for (uint32_t d = 0; d < nd; d++) {
                const uint32_t p = ra.deps[d0 + d];
                const unsigned long long t = *(volatile unsigned long long *)&ra.trace[EAV1_TRACE_WORDS * (size_t)p + 2];
                if (t >= last) { last = t; who = p; }
            }
            ra.trace[EAV1_TRACE_WORDS * (size_t)id + 3] = who;
        }
        if (!dep_waited) { pdl_wait(); dep_waited = true; }
        long long cyc_stage = 0, cyc_op[3] = { 0, 0, 0 }, n_op[3] = { 0, 0, 0 };
        for (uint32_t base = 0; base < count; base += EAV1_CMD_STAGE) {
            const uint32_t n = min(count - base, (uint32_t)EAV1_CMD_STAGE);
            long long c0 = ra.trace ? clock64() : 0;
            if (base) {
                const uint32_t *src = reinterpret_cast<const uint32_t *>(&ra.cmds[first + base]);
                uint32_t *dst = reinterpret_cast<uint32_t *>(ws.cmd);
                for (uint32_t i = tid; i < n * (sizeof(Eav1Cmd) / 4); i += nthr) dst[i] = src[i];
                CTA_SYNC();
            }
            if (ra.trace) { const long long c1 = clock64(); cyc_stage += c1 - c0; c0 = c1; }
            for (uint32_t c = 0; c < n;) {
                const uint32_t c_this = c;
                c += exec_cmd<BD>(fc, ws.cmd[c], sm, tid, nthr, c + 1 < n ? &ws.cmd[c + 1] : nullptr);
                CTA_SYNC();
                if (ra.trace) {
                    const long long c1 = clock64();
                    const int op = ws.cmd[c_this].op, k = op == EAV1_OP_ITX ? 1 : (op == EAV1_OP_IPRED || op == EAV1_OP_CFL) ? 0 : 2;
                    cyc_op[k] += c1 - c0; n_op[k]++;
                    if (tid == 0 && ra.hist) {   // [op][log2(w*h)] -> cycles, count
                        const int lg = 31 - __clz((int)ws.cmd[c_this].w * (int)ws.cmd[c_this].h);
                        unsigned long long *hh = ra.hist + ((op & 15) * 16 + (lg & 15)) * 2;
                        if (op == EAV1_OP_IPRED) {   // tables 15 / 14: by prediction mode, all sizes / 16x16 and larger
                            unsigned long long *hm = ra.hist + (15 * 16 + (ws.cmd[c_this].u.ipred.mode & 15)) * 2;
                            atomicAdd(hm, (unsigned long long)(c1 - c0)); atomicAdd(hm + 1, 1ull);
                            if (lg >= 8) {
                                hm = ra.hist + (14 * 16 + (ws.cmd[c_this].u.ipred.mode & 15)) * 2;
                                atomicAdd(hm, (unsigned long long)(c1 - c0)); atomicAdd(hm + 1, 1ull);
                            }
                        }
                        atomicAdd(hh, (unsigned long long)(c1 - c0)); atomicAdd(hh + 1, 1ull);
                    }
                    c0 = c1;
                }
            }
        }
        if (ra.debug == 1 && tid == 0) atomicAdd(&ra.ctrl[4], 1u);
        if (ra.trace && tid == 0) {
            unsigned long long *t = &ra.trace[EAV1_TRACE_WORDS * (size_t)id];
            t[0] = t_ticket; t[1] = t_ready; t[2] = global_ns();
            t[6] = cyc_stage; t[7] = cyc_op[0]; t[8] = cyc_op[1]; t[9] = cyc_op[2];
            t[10] = (unsigned long long)n_op[0] << 32 | (unsigned long long)n_op[1]; t[11] = n_op[2];
            t[4] = ((unsigned long long)first << 32) | count; t[5] = ((unsigned long long)k << 32) | (blockIdx.x << 8) | (nthr == 32 ? (threadIdx.x >> 5) : 255u);
        }
        if (fed && tid == 0) {   // nobody polls the flag of a unit without consumers
            // release at gpu scope: the group barrier after the unit's last command orders every
            // thread's stores before this one (cumulativity), so no separate fence is needed
#ifdef EAV1_EXTRA_FENCE
            __threadfence();
#endif
            st_release(&ra.done[id], ra.epoch);
        }
    }
    if (tid == 0) {
        const uint32_t retired = atomicAdd(&ra.ctrl[2], 1u);
        if (retired == ra.n_workers - 1) { ra.ctrl[0] = 0; ra.ctrl[1] = 0; ra.ctrl[2] = 0; __threadfence(); }
    }
}

#ifndef EAV1_RECON_MIN_CTAS
#define EAV1_RECON_MIN_CTAS 3   // 168 registers, 444 resident work groups: measured best (1: 255 regs, 4: 128 regs with spills)
#endif
template <typename BD>
__global__ void __launch_bounds__(128, EAV1_RECON_MIN_CTAS)
recon_frame_kernel(FrameCtx fc, ReconArgs ra)
{
    __shared__ alignas(16) unsigned char scratch[4 * EAV1_WARP_SCRATCH > sizeof(CtaScratch) ? 4 * EAV1_WARP_SCRATCH : sizeof(CtaScratch)];
    __shared__ WorkerSmem ws[4];
    if (blockIdx.x < ra.n_big_ctas) {
        recon_worker<BD>(fc, ra, ra.big, ra.n_big, &ra.ctrl[0], *reinterpret_cast<CtaScratch *>(scratch),
                         ws[0], threadIdx.x, 128);
    } else {
        const int w = threadIdx.x >> 5;
        recon_worker<BD>(fc, ra, ra.small, ra.n_small, &ra.ctrl[1],
                         *reinterpret_cast<CtaScratch *>(scratch + w * EAV1_WARP_SCRATCH),
                         ws[w], threadIdx.x & 31, 32);
    }
}

// Phase 2a: dataflow units without any producer inside the frame (on inter frames nearly all of them: the residual
// add on top of a finished motion-compensated block, the average / mask of two prep blocks, ...). They need no
// tickets and no flags: one warp per unit of at most 32x32 samples, one CTA per larger one, a plain grid.
// units[i] = (first command, number of commands); the large units lead the grid.
template <typename BD>
__global__ void __launch_bounds__(128, EAV1_RECON_MIN_CTAS)
flat_units_kernel(FrameCtx fc, const Eav1Cmd *__restrict__ cmds, const uint2 *__restrict__ units, uint32_t n_big, uint32_t n_small)
{
    pdl_wait();
    __shared__ alignas(16) unsigned char scratch[4 * EAV1_WARP_SCRATCH > sizeof(CtaScratch) ? 4 * EAV1_WARP_SCRATCH : sizeof(CtaScratch)];
    __shared__ Eav1Cmd cmd[4][2];
    int tid, nthr, w;
    uint2 u;
    if (blockIdx.x < n_big) {
        tid = threadIdx.x; nthr = 128; w = 0;
        u = units[blockIdx.x];
    } else {
        w = threadIdx.x >> 5; tid = threadIdx.x & 31; nthr = 32;
        const uint32_t it = (blockIdx.x - n_big) * 4 + w;
        if (it >= n_small) return;
        u = units[n_big + it];
    }
    CtaScratch &sm = *reinterpret_cast<CtaScratch *>(scratch + (nthr == 32 ? w * EAV1_WARP_SCRATCH : 0));
    for (uint32_t c = 0; c < u.y;) {
        const uint32_t n = min(u.y - c, 2u);
        for (uint32_t i = tid; i < n * (sizeof(Eav1Cmd) / 4); i += nthr)
            reinterpret_cast<uint32_t *>(&cmd[w][0])[i] = reinterpret_cast<const uint32_t *>(&cmds[u.x + c])[i];
        CTA_SYNC();
        c += exec_cmd<BD>(fc, cmd[w][0], sm, tid, nthr, n > 1 ? &cmd[w][1] : nullptr);
        CTA_SYNC();
    }
}

// Phase 1: motion compensation from reference pictures. One warp per work item (a block or a
// 32x32 tile of it); items are independent, so this is a plain grid with no flags or tickets.
template <typename BD>
__global__ void __launch_bounds__(128)
mc_kernel(FrameCtx fc, const Eav1Cmd *__restrict__ cmds, const uint32_t *__restrict__ items, uint32_t n_items)
{
    __shared__ alignas(16) unsigned char scratch[4 * EAV1_WARP_SCRATCH];
    __shared__ Eav1Cmd cmd[4];
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t it = blockIdx.x * 4 + w;
    if (it >= n_items) return;
    const uint32_t item = items[it];
    const uint32_t ci = item & 0xffffff;
    if (lane < (int)(sizeof(Eav1Cmd) / 4))
        reinterpret_cast<uint32_t *>(&cmd[w])[lane] = reinterpret_cast<const uint32_t *>(&cmds[ci])[lane];
    __syncwarp();
    pdl_wait();     // the work lists above were uploaded by the host; pixels are touched from here on
    const Eav1Cmd &c = cmd[w];
    CtaScratch &sm = *reinterpret_cast<CtaScratch *>(scratch + w * EAV1_WARP_SCRATCH);
    if (c.op == EAV1_OP_MC) {
        if (c.u.mc.dx) exec_mc_rect<BD>(fc, c, 0, 0, c.w, c.h, sm, lane, 32);
        else {
            const int x0 = (int)((item >> 24) & 3) * 32, y0 = (int)((item >> 26) & 3) * 32;
            exec_mc_rect<BD>(fc, c, x0, y0, min(32, c.w - x0), min(32, c.h - y0), sm, lane, 32);
        }
    } else if (c.op == EAV1_OP_WARP) {
        exec_warp<BD>(fc, c, sm, lane, 32);
    }
}

// Phase 1b: residuals. The inverse transform of a block depends on nothing but its coefficients, so
// every two-step ITX command of the frame is an independent item: one warp per block up to 32x32
// (register path), one CTA per block with a 64-point side. Items arrive sorted by (size, type) so
// that neighbouring warps run the same code. The residual lands in the arena; phase 2 adds it.
template <typename BD>
__global__ void __launch_bounds__(128, 4)
resid_kernel(FrameCtx fc, const Eav1Cmd *__restrict__ cmds, const uint32_t *__restrict__ items,
             uint32_t n_small, uint32_t n_small_ctas, uint32_t n_items)
{
    __shared__ alignas(16) unsigned char scratch[4 * EAV1_WARP_SCRATCH > sizeof(CtaScratch) ? 4 * EAV1_WARP_SCRATCH : sizeof(CtaScratch)];
    __shared__ Eav1Cmd cmd[4];
    const uint32_t n_big = n_items - n_small;     // the slow items lead the grid
    if (blockIdx.x >= n_big) {
        const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
        const uint32_t it = (blockIdx.x - n_big) * 4 + w;
        if (it >= n_small) return;
        const uint32_t ci = items[it];
        if (lane < (int)(sizeof(Eav1Cmd) / 4))
            reinterpret_cast<uint32_t *>(&cmd[w])[lane] = reinterpret_cast<const uint32_t *>(&cmds[ci])[lane];
        __syncwarp();
        exec_itx_residual<BD>(fc, cmd[w], *reinterpret_cast<CtaScratch *>(scratch + w * EAV1_WARP_SCRATCH), lane, 32);
    } else {
        const uint32_t it = n_small + blockIdx.x;
        if (it >= n_items) return;
        const uint32_t ci = items[it];
        if (threadIdx.x < sizeof(Eav1Cmd) / 4)
            reinterpret_cast<uint32_t *>(&cmd[0])[threadIdx.x] = reinterpret_cast<const uint32_t *>(&cmds[ci])[threadIdx.x];
        __syncthreads();
        exec_itx_residual<BD>(fc, cmd[0], *reinterpret_cast<CtaScratch *>(scratch), threadIdx.x, 128);
    }
}

// Output conversion: one thread per luma sample, rows of 128 samples per CTA row.
template <typename pixel>
__global__ void __launch_bounds__(128)
convert_kernel(ConvParams P)
{
    const int x = blockIdx.x * 128 + threadIdx.x, y = blockIdx.y;
    if (x < P.w) convert_sample<pixel>(P, x, y);
}

// debugging aid (EAV1_DEBUG_CHECK=2): stream-ordered checksum of an uploaded buffer
__global__ void checksum_kernel(const uint32_t *p, size_t n_words, unsigned long long *out)
{
    unsigned long long s = 0;
    for (size_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n_words; i += (size_t)gridDim.x * blockDim.x) s += p[i];
    atomicAdd(out, s);
}

template <typename pixel>
__global__ void __launch_bounds__(128)
deblock_kernel(PfParams pp, const Eav1LfRec *__restrict__ recs, int n)
{
    pdl_wait();
    // one thread per line of a 4-sample edge unit: 128 threads = the 32 units of one recorded call
    const int ri = blockIdx.x;
    if (ri >= n) return;
    const int line = threadIdx.x & 3;
    lf_unit<pixel>(pp, recs[ri], threadIdx.x >> 2, line, line + 1);
}

// Whole-picture duplicate ahead of CDEF / loop restoration (they write only what they filter). A copy-engine node
// between two kernels of a stream costs several microseconds of hand-over in either direction; 3 MB that sit in L2
// take a kernel two.
__global__ void __launch_bounds__(256)
picture_copy_kernel(uint4 *__restrict__ dst, uint4 *__restrict__ dst2, const uint4 *__restrict__ src, size_t n16)
{
    pdl_wait();
    for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < n16; i += (size_t)gridDim.x * 256) {
        const uint4 v = src[i];
        dst[i] = v;
        if (dst2) dst2[i] = v;
    }
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
    pdl_wait();
    lr_tile<pixel>(pp, r, blockIdx.x, blockIdx.z, sm, threadIdx.x, blockDim.x);
}

// Super-resolution: one thread per output sample, blockIdx.z = plane, the 64-phase filter table in shared memory.
template <typename pixel>
__global__ void __launch_bounds__(256)
resize_kernel(ResizeParams rp)
{
    __shared__ int8_t filt[64][8];
    for (int i = threadIdx.x; i < 64 * 8 / 4; i += blockDim.x)
        reinterpret_cast<uint32_t *>(&filt[0][0])[i] = reinterpret_cast<const uint32_t *>(&rp.tab->resize_filter[0][0])[i];
    __syncthreads();
    const int pl = blockIdx.z, y = blockIdx.y, x = blockIdx.x * blockDim.x + threadIdx.x;
    if (y >= rp.h[pl] || x >= rp.dst_w[pl]) return;
    resize_sample<pixel>(rp, filt, pl, x, y);
}

__global__ void __launch_bounds__(256)
fg_prepare_kernel(FgParams P)
{
    __shared__ FgSmem sm;
    fg_prepare(P, sm, threadIdx.x, blockDim.x);
}

template <typename pixel>
__global__ void __launch_bounds__(256)
fg_apply_kernel(FgParams P)
{
    fg_strip<pixel>(P, blockIdx.x, blockIdx.y, threadIdx.x, blockDim.x);
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

// A frame ready to launch: every list it needs is resident in device memory. The streaming
// path points these at the context's reusable buffers; eav1cu_frame_stage() gives a frame
// its own copies so it can be replayed with no host work (bench.py's device-resident leg).
struct FramePlan {
    Eav1FrameDesc fd;
    const Eav1Cmd *d_cmds = nullptr;
    const uint32_t *d_deps = nullptr, *d_items = nullptr, *d_ritems = nullptr;
    uint32_t n_ritems = 0, n_ritems_small = 0;   // two-step transform blocks (phase 1b)
    long long resid_px = 0;
    const ChainRec *d_recs = nullptr;
    const uint2 *d_flat = nullptr;               // producer-free units: (first command, count), large ones first
    uint32_t n_flat_big = 0, n_flat_small = 0;
    uint32_t n_items = 0;
    long long mc_px = 0, unit_px = 0;
    uint32_t n_chains = 0, n_deps = 0, n_levels = 0, n_big = 0, n_small = 0;
    double big_share = 0;
    const uint8_t *d_payload = nullptr;
    size_t arena_bytes = 0, payload_bytes = 0, n_cmds = 0;
    const Eav1LfRec *d_lf = nullptr; size_t n_lf0 = 0, n_lf1 = 0;
    const uint8_t *d_lflvl = nullptr;
    const uint8_t *d_lfmask = nullptr;
    const Eav1LrRec *d_lr = nullptr; size_t n_lr = 0;
    std::vector<void *> owned;   // device allocations owned by a staged plan
    long long launches = 0;
};

// Per-frame upload buffers exist twice: while the kernels of frame k still read slot k&1, the
// host already fills the other slot for frame k+1 (the `done` event guards reuse two frames later).
struct Slot {
    DevBuf stage;   // every list of the frame, packed (one host->device copy)
    cudaEvent_t done = nullptr;
    bool used = false;
};

struct Eav1cuCtx {
    int device = 0;
    cudaStream_t stream = nullptr;
    Eav1Tables *d_tab = nullptr;
    bool have_tab = false;
    uint8_t *d_masks = nullptr; size_t masks_bytes = 0;
    std::vector<Picture> pics;
    DevBuf arena;
    Slot slot[2];
    int cur_slot = 0;
    Slot &S() { return slot[cur_slot]; }
    // fences: events the host side can wait on (frame finished, picture downloaded)
    static const int N_FENCE = 64;
    cudaEvent_t fence_ev[N_FENCE] = {};
    unsigned long long fence_next = 1;
    std::mutex fence_mu;
    Picture work[4];           // post-filter ping-pong pictures; [2], [3]: super-resolved CDEF output / deblocked frame
    std::vector<Eav1LfRec> lf_sorted;
    Scheduler sched;
    DepGraph graph;
    Schedule plan;
    UnitGraph units, units2;      // all phase-2 units / those that wait for or feed another unit
    std::vector<uint2> flat_tmp;
    std::vector<uint8_t> stage_tmp;
    std::vector<uint32_t> remap_tmp;
    std::vector<ChainRec> rec_tmp;
    std::vector<uint8_t> fed_tmp;
    uint32_t *d_done = nullptr; size_t done_cap = 0;
    uint32_t *d_ctrl = nullptr;
    uint8_t *d_fg = nullptr;   // film grain: templates, scaling LUTs, block offsets
    uint32_t epoch = 0;
    int recon_grid = 148 * 4;  // the widest persistent grid: every SM filled to the kernel's occupancy
    int recon_sms = 148;
    bool recon_grid_fixed = false;
    int debug_check = 0;
    double t_build = 0, t_launch = 0, t_wait = 0;   // host seconds spent in frame_submit (EAV1_TIMING)
    double t_sched = 0, t_pack = 0, t_lock = 0, t_copycall = 0;   // inside plan_build: dependency graph, memcpy into staging, wait for the submit lock, the copy call
    unsigned long long *d_sums = nullptr;   // [frame & 63][2]: device checksums of payload / cmds
    unsigned long long h_sums[64][2] = {};  // host checksums of the same uploads
    long long sum_frame[64] = {};
    std::vector<uint32_t> est_tmp, order_tmp, bucket_tmp, ritem_tmp, tail_tmp, order2_tmp;
    std::vector<Eav1Cmd> cmd_tmp;
    std::vector<uint32_t> ritem_tmp2;
    unsigned long long *d_trace = nullptr;   // set only inside eav1cu_staged_trace
    int est_weight = 8;           // ticket key = est_weight * est - tail (EAV1_EST_WEIGHT; any weight >= 1 keeps the order topological)
    bool no_flat = true;          // unless EAV1_FLAT is set every phase-2 unit goes through the dataflow kernel: the separate flat launch
                                  // for producer-free units measured slower on one stream (recon 8.7 -> 10.4 ms per 120 frames, one more
                                  // launch per frame, same critical chain) and no faster with several streams
    bool serial_phase1 = false;   // EAV1_SERIAL_PHASE1: residuals on the main stream (profiling, debugging)
    int warp_mode = 0;   // 0: every chain on a CTA, 1: independent small chains on warps, 2: all small chains on warps
    int err = 0;
    std::string errmsg;
    std::mutex mu, err_mu;      // mu: picture pool; err_mu: err / errmsg (fail() is reached from any thread)
    long long n_launches = 0, h2d_bytes = 0, d2h_bytes = 0, n_frames = 0, staged_graph_kernels = 0;
    DevBuf conv;                        // output of eav1cu_pic_convert before it is copied to the host
    bool own_stream = true;
    cudaStream_t stream2 = nullptr;            // phase 1b (residuals) runs beside phase 1 (motion compensation)
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
    cudaEvent_t ev_fg_ready = nullptr, ev_fg_free = nullptr;   // film grain: templates built on stream2, applied on the main stream
    cudaEvent_t ev_fg_cap[2] = { nullptr, nullptr };          // the same two roles inside a stream capture: an event last
                                                               // recorded in a capture cannot be waited for outside it
    cudaStream_t stream_fg = nullptr;   // film-grain template builds (eav1cu_apply_grain)
    unsigned event_flags = cudaEventDisableTiming | cudaEventBlockingSync;   // of the events host threads wait on
    bool fg_busy = false;
    // optional per-stage device timing (bench.py's roofline leg)
    bool stage_timing = false;
    struct Span { cudaEvent_t a, b; int stage; int launches; };
    std::vector<Span> spans;
    std::vector<cudaEvent_t> ev_pool;
    double stage_ms[EAV1CU_N_STAGES] = {};
    long long stage_launches[EAV1CU_N_STAGES] = {}, stage_calls[EAV1CU_N_STAGES] = {};
    bool capture = false;
    cudaGraph_t staged_graph = nullptr;          // every staged frame in submit order, captured once (eav1cu_staged_graph_build)
    cudaGraphExec_t staged_exec = nullptr;
    FramePlan live;
    std::vector<FramePlan *> staged;
};

static int fail(Eav1cuCtx *c, int code, const char *what, cudaError_t e = cudaSuccess)
{
    std::unique_lock<std::mutex> g;
    if (c) g = std::unique_lock<std::mutex>(c->err_mu);
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
    if (bytes <= b.cap && (!pinned_too || bytes <= b.hcap)) return 0;
    CallTimer ct_(TC_ALLOC);
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
    if (p.base && p.bytes < total) { CallTimer ct_(TC_ALLOC); cudaFree(p.base); p.base = nullptr; }
    if (!p.base) { CallTimer ct_(TC_ALLOC); CK(cudaMalloc(&p.base, total)); p.bytes = total; }
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

// Contexts are recycled within the process: opening a decoder right after another one was
// closed (the usual pattern of a player or a transcoder working through a playlist) finds its
// stream, staging buffers and device pictures already allocated.
static std::mutex g_cache_mu;
static std::vector<Eav1cuCtx *> g_cache;
static const size_t MAX_CACHED_CTX = 96;   // a 64-stream server re-opening its decoders finds all of them (bench config C5)

// Called with the context idle (fresh or from the cache): events of the other kind are dropped and made again on demand.
static void ctx_set_wait_mode(Eav1cuCtx *ctx, bool spin)
{
    const unsigned flags = cudaEventDisableTiming | (spin ? 0u : (unsigned)cudaEventBlockingSync);
    if (flags == ctx->event_flags) return;
    ctx->event_flags = flags;
    for (Slot &sl : ctx->slot) if (sl.done) { cudaEventDestroy(sl.done); sl.done = nullptr; sl.used = false; }
    for (cudaEvent_t &e : ctx->fence_ev) if (e) { cudaEventDestroy(e); e = nullptr; }
}

int eav1cu_ctx_create(Eav1cuCtx **out, int device)
{
    if (!out) return EAV1CU_ERR_INVAL;
    *out = nullptr;
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n <= 0 || device >= n) return EAV1CU_ERR_NO_DEVICE;
    {
        std::lock_guard<std::mutex> g(g_cache_mu);
        for (size_t i = 0; i < g_cache.size(); i++)
            if (g_cache[i]->device == device) {
                *out = g_cache[i];
                g_cache.erase(g_cache.begin() + i);
                cudaSetDevice(device);
                const int live = ++g_live_ctx;
                ctx_set_wait_mode(*out, want_spin(live));
                pdl_set_early(device, live, (*out)->stream);
                return 0;
            }
    }
    Eav1cuCtx *ctx = new Eav1cuCtx();
    ctx->device = device;
    ctx->pics.reserve(512);   // never reallocates: other threads index it while frames are in flight
    if (cudaSetDevice(device) != cudaSuccess ||
        cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess ||
        cudaMalloc(&ctx->d_tab, sizeof(Eav1Tables)) != cudaSuccess ||
        cudaMalloc(&ctx->d_ctrl, 64) != cudaSuccess ||
        cudaMemsetAsync(ctx->d_ctrl, 0, 64, ctx->stream) != cudaSuccess ||
        cudaStreamSynchronize(ctx->stream) != cudaSuccess) {
        delete ctx;
        return EAV1CU_ERR_CUDA;
    }
    {
        // persistent grid: every SM filled to the occupancy the kernel reaches (multiple of the SM count)
        int sms = 148, per8 = 4, per16 = 4;
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per8, recon_frame_kernel<BD8>, 128, 0);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per16, recon_frame_kernel<BD16>, 128, 0);
        int per = std::max(1, std::min(per8, per16));
        if (const char *e = getenv("EAV1_RECON_CTAS_PER_SM")) per = std::max(1, std::min(per, atoi(e)));
        ctx->recon_grid = sms * per;
        ctx->recon_sms = sms;
        ctx->recon_grid_fixed = getenv("EAV1_RECON_CTAS_PER_SM") != nullptr;
    }
    if (const char *e = getenv("EAV1_WARP_MODE")) ctx->warp_mode = atoi(e);
    ctx->serial_phase1 = getenv("EAV1_SERIAL_PHASE1") != nullptr;
    ctx->no_flat = getenv("EAV1_FLAT") == nullptr;
    if (const char *e = getenv("EAV1_EST_WEIGHT")) ctx->est_weight = std::max(1, std::min(16, atoi(e)));
    if (!ctx->stream2) {
        if (cudaStreamCreateWithFlags(&ctx->stream2, cudaStreamNonBlocking) != cudaSuccess ||
            cudaEventCreateWithFlags(&ctx->ev_fork, cudaEventDisableTiming) != cudaSuccess ||
            cudaEventCreateWithFlags(&ctx->ev_join, cudaEventDisableTiming) != cudaSuccess) {
            if (ctx->stream2) { cudaStreamDestroy(ctx->stream2); ctx->stream2 = nullptr; }   // run phase 1b on the main stream
        }
    }
    if (const char *e = getenv("EAV1_DEBUG_CHECK")) ctx->debug_check = atoi(e) ? atoi(e) : 1;
    if (ctx->debug_check == 2) { cudaMalloc(&ctx->d_sums, 64 * 2 * 8); cudaMemset(ctx->d_sums, 0, 64 * 2 * 8); cudaDeviceSynchronize(); }
    const int live = ++g_live_ctx;
    ctx_set_wait_mode(ctx, want_spin(live));
    pdl_set_early(device, live, ctx->stream);
    *out = ctx;
    return 0;
}

static void ctx_free(Eav1cuCtx *ctx)
{
    if (!ctx) return;
    if (getenv("EAV1_TIMING") && ctx->n_frames)
        fprintf(stderr, "EAV1 TIMING: %lld frames; per frame on the submitting thread: slot wait %.3f ms, plan build + uploads %.3f ms, launches %.3f ms\n",
                ctx->n_frames, 1e3 * ctx->t_wait / ctx->n_frames, 1e3 * ctx->t_build / ctx->n_frames, 1e3 * ctx->t_launch / ctx->n_frames);
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    for (auto &p : ctx->pics) if (p.base) cudaFree(p.base);
    for (auto &p : ctx->work) if (p.base) cudaFree(p.base);
    free_buf(ctx->arena);
    for (Slot &sl : ctx->slot) {
        free_buf(sl.stage);
        if (sl.done) cudaEventDestroy(sl.done);
    }
    for (cudaEvent_t e : ctx->fence_ev) if (e) cudaEventDestroy(e);
    if (ctx->d_done) cudaFree(ctx->d_done);
    if (ctx->conv.d) cudaFree(ctx->conv.d);
    if (ctx->d_ctrl) cudaFree(ctx->d_ctrl);
    if (ctx->d_fg) cudaFree(ctx->d_fg);
    if (ctx->d_tab) cudaFree(ctx->d_tab);
    if (ctx->d_masks) cudaFree(ctx->d_masks);
    for (auto &sp : ctx->spans) { cudaEventDestroy(sp.a); cudaEventDestroy(sp.b); }
    for (cudaEvent_t e : ctx->ev_pool) cudaEventDestroy(e);
    for (FramePlan *sp : ctx->staged) { for (void *d : sp->owned) cudaFree(d); delete sp; }
    if (ctx->staged_exec) cudaGraphExecDestroy(ctx->staged_exec);
    if (ctx->staged_graph) cudaGraphDestroy(ctx->staged_graph);
    if (ctx->own_stream) cudaStreamDestroy(ctx->stream);
    if (ctx->stream2) cudaStreamDestroy(ctx->stream2);
    if (ctx->ev_fork) cudaEventDestroy(ctx->ev_fork);
    if (ctx->ev_join) cudaEventDestroy(ctx->ev_join);
    if (ctx->stream_fg) cudaStreamDestroy(ctx->stream_fg);
    if (ctx->ev_fg_ready) cudaEventDestroy(ctx->ev_fg_ready);
    if (ctx->ev_fg_free) cudaEventDestroy(ctx->ev_fg_free);
    for (cudaEvent_t e : ctx->ev_fg_cap) if (e) cudaEventDestroy(e);
    delete ctx;
}

void eav1cu_ctx_destroy(Eav1cuCtx *ctx)
{
    if (!ctx) return;
    const int live = --g_live_ctx;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    pdl_set_early(ctx->device, live, ctx->stream);
    bool keep = ctx->own_stream && !ctx->err && !getenv("EAV1_NO_CTX_CACHE");
    if (keep) {
        // back to a just-created state, keeping every allocation
        if (getenv("EAV1_TIMING") && ctx->n_frames)
            fprintf(stderr, "EAV1 TIMING: %lld frames; per frame on the submitting thread: slot wait %.3f ms, plan build + uploads %.3f ms (graph %.3f, pack %.3f, lock wait %.3f, copy call %.3f), launches %.3f ms\n",
                    ctx->n_frames, 1e3 * ctx->t_wait / ctx->n_frames, 1e3 * ctx->t_build / ctx->n_frames, 1e3 * ctx->t_sched / ctx->n_frames,
                    1e3 * ctx->t_pack / ctx->n_frames, 1e3 * ctx->t_lock / ctx->n_frames, 1e3 * ctx->t_copycall / ctx->n_frames, 1e3 * ctx->t_launch / ctx->n_frames);
        for (FramePlan *sp : ctx->staged) { for (void *d : sp->owned) cudaFree(d); delete sp; }
        ctx->staged.clear();
        if (ctx->staged_exec) { cudaGraphExecDestroy(ctx->staged_exec); ctx->staged_exec = nullptr; }
        if (ctx->staged_graph) { cudaGraphDestroy(ctx->staged_graph); ctx->staged_graph = nullptr; }
        for (auto &p : ctx->pics) p.used = false;
        ctx->have_tab = false;
        ctx->capture = false; ctx->stage_timing = false;
        for (auto &sp : ctx->spans) { ctx->ev_pool.push_back(sp.a); ctx->ev_pool.push_back(sp.b); }
        ctx->spans.clear();
        for (int i = 0; i < EAV1CU_N_STAGES; i++) { ctx->stage_ms[i] = 0; ctx->stage_launches[i] = 0; ctx->stage_calls[i] = 0; }
        ctx->n_launches = ctx->h2d_bytes = ctx->d2h_bytes = ctx->n_frames = 0;
        ctx->t_build = ctx->t_launch = ctx->t_wait = 0;
        ctx->t_sched = ctx->t_pack = ctx->t_lock = ctx->t_copycall = 0;
        std::lock_guard<std::mutex> g(g_cache_mu);
        if (g_cache.size() < MAX_CACHED_CTX) { g_cache.push_back(ctx); return; }
    }
    ctx_free(ctx);
}

int eav1cu_last_error(Eav1cuCtx *ctx, char *msg, size_t cap)
{
    if (!ctx) return EAV1CU_ERR_INVAL;
    std::lock_guard<std::mutex> g(ctx->err_mu);
    if (msg && cap) snprintf(msg, cap, "%s", ctx->errmsg.c_str());
    return ctx->err;
}

int eav1cu_clear_error(Eav1cuCtx *ctx)
{
    if (!ctx) return EAV1CU_ERR_INVAL;
    std::lock_guard<std::mutex> g(ctx->err_mu);
    if (ctx->err == EAV1CU_ERR_INVAL) { ctx->err = 0; ctx->errmsg.clear(); }
    return ctx->err;
}

int eav1cu_host_alloc(void **p, size_t bytes)
{
    if (!p) return EAV1CU_ERR_INVAL;
    CallTimer ct_(TC_ALLOC);
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
    if (ctx->d_masks && ctx->masks_bytes != bytes) { cudaFree(ctx->d_masks); ctx->d_masks = nullptr; }
    if (!ctx->d_masks) CK(cudaMalloc(&ctx->d_masks, bytes));
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
    if (id < 0) {
        if (ctx->pics.size() >= 512) return fail(ctx, EAV1CU_ERR_NOMEM, "device picture pool exhausted");
        ctx->pics.emplace_back(); id = (int)ctx->pics.size() - 1;
    }
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

static int pic_copy(Eav1cuCtx *ctx, int id, void *const data[3], const ptrdiff_t stride[2], bool to_device, bool wait = true)
{
    if (!ctx || id < 0 || id >= (int)ctx->pics.size() || !ctx->pics[id].base) return EAV1CU_ERR_INVAL;
    CK(cudaSetDevice(ctx->device));
    Picture &p = ctx->pics[id];
    const int hbd = p.bpc > 8;
    const int ss_hor = p.layout != 3, ss_ver = p.layout == 1;
    std::unique_lock<SubmitMutex> g(g_submit_mu);
    // The host picture usually has the device picture's geometry (same strides, planes the same distance apart:
    // both follow dav1d's allocation rule): then the whole picture is one contiguous copy instead of three 2-D ones.
    const int np = p.layout ? 3 : 1;
    bool same = stride[0] == p.pl[0].stride && (np == 1 || stride[1] == p.pl[1].stride);
    for (int pl = 1; pl < np && same; pl++)
        same = (const uint8_t *)data[pl] - (const uint8_t *)data[0] == p.pl[pl].p - p.pl[0].p;
    if (same) {
        const int lw = np == 1 ? p.w : (p.w + ss_hor) >> ss_hor, lh = np == 1 ? p.h : (p.h + ss_ver) >> ss_ver;
        const size_t bytes = (size_t)(p.pl[np - 1].p - p.pl[0].p) + (size_t)p.pl[np - 1].stride * (lh - 1) + ((size_t)lw << hbd);
        if (to_device) { ctx->h2d_bytes += (long long)bytes; TCK(TC_PIC_COPY, cudaMemcpyAsync(p.pl[0].p, data[0], bytes, cudaMemcpyHostToDevice, ctx->stream)); }
        else { ctx->d2h_bytes += (long long)bytes; g_total[2] += (long long)bytes; TCK(TC_PIC_COPY, cudaMemcpyAsync(data[0], p.pl[0].p, bytes, cudaMemcpyDeviceToHost, ctx->stream)); }
    } else
    for (int pl = 0; pl < np; pl++) {
        const int w = pl ? (p.w + ss_hor) >> ss_hor : p.w;
        const int h = pl ? (p.h + ss_ver) >> ss_ver : p.h;
        if (to_device) ctx->h2d_bytes += ((long long)w << hbd) * h; else { ctx->d2h_bytes += ((long long)w << hbd) * h; g_total[2] += ((long long)w << hbd) * h; }
        if (to_device)
            CK(cudaMemcpy2DAsync(p.pl[pl].p, p.pl[pl].stride, data[pl], stride[!!pl], (size_t)w << hbd, h,
                                 cudaMemcpyHostToDevice, ctx->stream));
        else
            CK(cudaMemcpy2DAsync(data[pl], stride[!!pl], p.pl[pl].p, p.pl[pl].stride, (size_t)w << hbd, h,
                                 cudaMemcpyDeviceToHost, ctx->stream));
    }
    g.unlock();
    if (wait) CK(cudaStreamSynchronize(ctx->stream));
    return ctx->err;
}

int eav1cu_pic_convert(Eav1cuCtx *ctx, int id, const Eav1Convert *cv, void *dst, size_t dst_stride, size_t dst_bytes)
{
    if (!ctx || !cv || !dst || id < 0 || id >= (int)ctx->pics.size() || !ctx->pics[id].base) return EAV1CU_ERR_INVAL;
    if (ctx->err) return ctx->err;
    const Picture &p = ctx->pics[id];
    const size_t need = conv_needed(cv->format, p.layout, p.bpc, p.w, p.h, dst_stride);
    if (!need || need > dst_bytes) return EAV1CU_ERR_INVAL;
    CK(cudaSetDevice(ctx->device));
    if (need > ctx->conv.cap) {
        if (ctx->conv.d) cudaFree(ctx->conv.d);
        ctx->conv.d = nullptr; ctx->conv.cap = 0;
        CK(cudaMalloc((void **)&ctx->conv.d, need));
        ctx->conv.cap = need;
    }
    ConvParams P;
    memset(&P, 0, sizeof(P));
    for (int i = 0; i < 3; i++) P.src[i] = p.pl[i];
    P.w = p.w; P.h = p.h; P.mono = p.layout == 0; P.ss_hor = p.layout != 3; P.ss_ver = p.layout == 1;
    P.bpc = p.bpc; P.format = cv->format;
    conv_coefficients(P, cv->matrix, cv->full_range);
    P.dst = ctx->conv.d; P.dst_stride = (long long)dst_stride;
    {
        std::lock_guard<SubmitMutex> g(g_submit_mu);
        const dim3 grid((unsigned)((p.w + 127) / 128), (unsigned)p.h);
        if (p.bpc > 8) convert_kernel<uint16_t><<<grid, 128, 0, ctx->stream>>>(P);
        else convert_kernel<uint8_t><<<grid, 128, 0, ctx->stream>>>(P);
        CK(cudaGetLastError());
        ctx->n_launches++; g_total[0]++;
        CK(cudaMemcpyAsync(dst, ctx->conv.d, need, cudaMemcpyDeviceToHost, ctx->stream));
        ctx->d2h_bytes += (long long)need; g_total[2] += (long long)need;
    }
    CK(cudaStreamSynchronize(ctx->stream));
    return ctx->err;
}

int eav1cu_pic_download_async(Eav1cuCtx *ctx, int id, void *const data[3], const ptrdiff_t stride[2])
{
    return pic_copy(ctx, id, data, stride, false, false);
}

int eav1cu_fence(Eav1cuCtx *ctx, unsigned long long *ticket)
{
    if (!ctx || !ticket) return EAV1CU_ERR_INVAL;
    CK(cudaSetDevice(ctx->device));
    std::lock_guard<std::mutex> g(ctx->fence_mu);
    const unsigned long long t = ctx->fence_next++;
    cudaEvent_t &e = ctx->fence_ev[t % Eav1cuCtx::N_FENCE];
    if (!e) CK(cudaEventCreateWithFlags(&e, ctx->event_flags));
    else TCK(TC_EVSYNC_FENCE, cudaEventSynchronize(e));   // the ticket issued N_FENCE fences ago
    {
        std::lock_guard<SubmitMutex> gs(g_submit_mu);
        TCK(TC_EVREC, cudaEventRecord(e, ctx->stream));
    }
    *ticket = t;
    return 0;
}

int eav1cu_fence_wait(Eav1cuCtx *ctx, unsigned long long ticket)
{
    if (!ctx || !ticket) return EAV1CU_ERR_INVAL;
    CK(cudaSetDevice(ctx->device));
    cudaEvent_t e = nullptr;
    {
        std::lock_guard<std::mutex> g(ctx->fence_mu);
        if (ticket >= ctx->fence_next) return EAV1CU_ERR_INVAL;
        if (ticket + Eav1cuCtx::N_FENCE <= ctx->fence_next) return ctx->err;   // recycled: it completed long ago
        e = ctx->fence_ev[ticket % Eav1cuCtx::N_FENCE];
    }
    TCK(TC_EVSYNC_FENCE, cudaEventSynchronize(e));
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
        // grow, keeping what earlier uploads put there
        CK(cudaStreamSynchronize(ctx->stream));
        DevBuf bigger;
        int r = ensure(ctx, bigger, (off + bytes) * 2, false);
        if (r) return r;
        if (ctx->arena.d) { CK(cudaMemcpy(bigger.d, ctx->arena.d, ctx->arena.cap, cudaMemcpyDeviceToDevice)); cudaFree(ctx->arena.d); }
        ctx->arena.d = bigger.d; ctx->arena.cap = bigger.cap;
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

} // extern "C"

static cudaEvent_t ev_get(Eav1cuCtx *ctx)
{
    if (!ctx->ev_pool.empty()) { cudaEvent_t e = ctx->ev_pool.back(); ctx->ev_pool.pop_back(); return e; }
    cudaEvent_t e = nullptr;
    cudaEventCreate(&e);
    return e;
}

struct StageTimer {
    Eav1cuCtx *ctx; int stage; long long l0; cudaEvent_t a = nullptr; cudaStream_t on;
    StageTimer(Eav1cuCtx *c, int st, cudaStream_t s = nullptr) : ctx(c), stage(st), l0(c->n_launches), on(s ? s : c->stream) {
        if (ctx->stage_timing) { a = ev_get(ctx); cudaEventRecord(a, on); }
    }
    ~StageTimer() {
        if (!a) return;
        cudaEvent_t b = ev_get(ctx);
        cudaEventRecord(b, on);
        ctx->spans.push_back({a, b, stage, (int)(ctx->n_launches - l0)});
    }
};

// Everything a frame's kernels read from the host travels in ONE packed buffer and one host->device copy:
// put() only notes (source, size, where the device pointer goes); flush() lays the pieces out 256 bytes apart
// in the slot's pinned staging buffer (streaming) or in a private device allocation (staged plan), fills in the
// device pointers and queues the single copy. The sources (the context's scratch vectors, the caller's lists)
// stay untouched until then.
struct Uploader {
    Eav1cuCtx *ctx; FramePlan *own;
    struct Piece { const void *src; size_t bytes; const void **out; };
    std::vector<Piece> pieces;
    void put(const void *src, size_t bytes, const void **out) {
        *out = nullptr;
        if (bytes) pieces.push_back({src, bytes, out});
    }
    template <typename T> void put(const void *src, size_t bytes, const T **out) { put(src, bytes, reinterpret_cast<const void **>(out)); }
    int flush(DevBuf &stage) {
        size_t total = 0;
        for (const Piece &p : pieces) total += (p.bytes + 255) & ~(size_t)255;
        if (!total) return 0;
        uint8_t *dev = nullptr, *hst = nullptr;
        if (!own) {
            int r = ensure(ctx, stage, total, true);
            if (r) return r;
            dev = stage.d; hst = stage.h;
        } else {
            void *d = nullptr;
            CK(cudaMalloc(&d, total));
            own->owned.push_back(d);
            dev = (uint8_t *)d;
            ctx->stage_tmp.resize(total);
            hst = ctx->stage_tmp.data();
        }
        size_t off = 0;
        auto ta = std::chrono::steady_clock::now();
        for (const Piece &p : pieces) {
            memcpy(hst + off, p.src, p.bytes);
            *p.out = dev + off;
            off += (p.bytes + 255) & ~(size_t)255;
        }
        auto tb = std::chrono::steady_clock::now();
        {
            std::lock_guard<SubmitMutex> g(g_submit_mu);
            auto tc = std::chrono::steady_clock::now();
            TCK(TC_H2D_PACK, cudaMemcpyAsync(dev, hst, total, cudaMemcpyHostToDevice, ctx->stream));
            auto td = std::chrono::steady_clock::now();
            ctx->t_lock += std::chrono::duration<double>(tc - tb).count();
            ctx->t_copycall += std::chrono::duration<double>(td - tc).count();
        }
        ctx->t_pack += std::chrono::duration<double>(tb - ta).count();
        if (own) CK(cudaStreamSynchronize(ctx->stream));
        else { ctx->h2d_bytes += (long long)total; g_total[1] += (long long)total; }
        pieces.clear();
        return 0;
    }
};

static int plan_build(Eav1cuCtx *ctx, FramePlan &pl, bool staged, const Eav1FrameDesc *fd,
                      const Eav1Cmd *cmds, size_t n_cmds,
                      const void *payload, size_t payload_bytes, size_t arena_bytes)
{
    if (fd->cur < 0 || fd->cur >= (int)ctx->pics.size() || !ctx->pics[fd->cur].base)
        return fail(ctx, EAV1CU_ERR_INVAL, "bad current picture id");
    Picture &cur = ctx->pics[fd->cur];
    pl.fd = *fd;
    pl.fd.lf_recs = nullptr; pl.fd.lf_level = nullptr; pl.fd.lf_mask = nullptr; pl.fd.lr_recs = nullptr;
    const bool has_cdef = fd->cdef_enable != 0, has_lr = fd->n_lr_recs > 0;
    int r;
    if (fd->superres) {
        if (fd->sr_pic < 0 || fd->sr_pic >= (int)ctx->pics.size() || !ctx->pics[fd->sr_pic].base || fd->sr_pic == fd->cur)
            return fail(ctx, EAV1CU_ERR_INVAL, "bad super-resolution picture id");
        const Picture &sr = ctx->pics[fd->sr_pic];
        if (sr.w != fd->sr_w || sr.h != cur.h || sr.layout != cur.layout || sr.bpc != cur.bpc)
            return fail(ctx, EAV1CU_ERR_INVAL, "super-resolution picture geometry mismatch");
        // reconstruction + deblock happen in `cur` itself; CDEF needs one more picture of the coded size,
        // loop restoration the upscaled CDEF output and (for its stripe borders) the upscaled deblocked frame
        if (has_cdef && (r = pic_setup(ctx, ctx->work[1], cur.w, cur.h, cur.layout, cur.bpc))) return r;
        if (has_lr) {
            if ((r = pic_setup(ctx, ctx->work[2], sr.w, sr.h, sr.layout, sr.bpc))) return r;
            if (has_cdef && (r = pic_setup(ctx, ctx->work[3], sr.w, sr.h, sr.layout, sr.bpc))) return r;
        }
    } else if (has_cdef || has_lr) {
        if ((r = pic_setup(ctx, ctx->work[0], cur.w, cur.h, cur.layout, cur.bpc))) return r;
        if (has_cdef && has_lr && (r = pic_setup(ctx, ctx->work[1], cur.w, cur.h, cur.layout, cur.bpc))) return r;
    }
    Uploader up{ctx, staged ? &pl : nullptr};
    // `pl` may be the context's reused live plan: nothing of the previous frame may survive (a submission without
    // commands skips the block below)
    pl.n_chains = pl.n_deps = pl.n_levels = 0;
    pl.n_items = pl.n_ritems = pl.n_ritems_small = pl.n_flat_big = pl.n_flat_small = pl.n_big = pl.n_small = 0;
    pl.d_cmds = nullptr; pl.d_deps = nullptr; pl.d_items = nullptr; pl.d_ritems = nullptr; pl.d_recs = nullptr; pl.d_flat = nullptr;
    pl.d_payload = nullptr; pl.resid_px = 0;
    pl.arena_bytes = arena_bytes; pl.payload_bytes = payload_bytes; pl.n_cmds = n_cmds;
    if (n_cmds) {
        int pw[3] = { cur.pl[0].w, cur.pl[1].w, cur.pl[2].w };
        int ph[3] = { cur.pl[0].h, cur.pl[1].h, cur.pl[2].h };
        auto ts0 = std::chrono::steady_clock::now();
        ctx->sched.reset(pw, ph);
        ctx->sched.build_units(cmds, n_cmds, ctx->units);
        ctx->t_sched += std::chrono::duration<double>(std::chrono::steady_clock::now() - ts0).count();
        const UnitGraph &gall = ctx->units;
        // Two-step transform blocks: give each a residual block behind the arena and list it for
        // phase 1b, sorted by (transform size, type). The engine works on its own copy of the list.
        {
            std::vector<Eav1Cmd> &cc = ctx->cmd_tmp;
            cc.assign(cmds, cmds + n_cmds);
            std::vector<uint32_t> &ri = ctx->ritem_tmp, &bk = ctx->bucket_tmp;
            size_t off = (arena_bytes + 31) & ~(size_t)31;
            // small blocks (one warp each) largest first, same (size, type) adjacent; blocks with a
            // 64-point side (one CTA each) behind them in the list, but first in the grid
            const int NK = 19 * 17;
            bk.assign(8 * NK + 1, 0);
            auto key = [NK](const Eav1Cmd &c) {
                if (c.w > 32 || c.h > 32) return 7 * NK + c.u.itx.tx * 17 + c.u.itx.txtp;
                int lg = 0; while ((2 << lg) <= (int)c.w * c.h) lg++;      // 4 (4x4) .. 10 (32x32)
                return (10 - lg) * NK + c.u.itx.tx * 17 + c.u.itx.txtp;
            };
            uint32_t n_two = 0;
            pl.resid_px = 0;
            for (size_t i = 0; i < n_cmds; i++) {
                Eav1Cmd &c = cc[i];
                if (c.op != EAV1_OP_ITX || (c.flags & EAV1_F_DST_ARENA) || c.u.itx.tx >= 19 || c.u.itx.txtp >= 16 ||
                    c.u.itx.eob < (c.u.itx.txtp == 0))
                    continue;
                c.flags |= EAV1_F_RESID;
                c.dst_off = (uint32_t)off;
                off += (size_t)c.w * c.h * 2;
                pl.resid_px += (long long)c.w * c.h;
                bk[key(c) + 1]++;
                n_two++;
            }
            for (size_t b = 1; b < bk.size(); b++) bk[b] += bk[b - 1];
            pl.n_ritems_small = bk[7 * NK];
            ri.resize(n_two);
            for (size_t i = 0; i < n_cmds; i++)
                if (cc[i].flags & EAV1_F_RESID && cc[i].op == EAV1_OP_ITX) ri[bk[key(cc[i])]++] = (uint32_t)i;
            pl.n_ritems = n_two;
            arena_bytes = off;
            pl.arena_bytes = arena_bytes;
            cmds = cc.data();
            if (n_two) {
                up.put(ri.data(), ri.size() * 4, &pl.d_ritems);
            }
        }
        up.put(cmds, n_cmds * sizeof(Eav1Cmd), &pl.d_cmds);
        pl.n_items = (uint32_t)gall.mc_items.size();
        if (staged) {
            pl.mc_px = pl.unit_px = 0;
            for (uint32_t it : gall.mc_items) {
                const Eav1Cmd &c = cmds[it & 0xffffff];
                const int x0 = (int)((it >> 24) & 3) * 32, y0 = (int)((it >> 26) & 3) * 32;
                pl.mc_px += (c.op == EAV1_OP_MC && !c.u.mc.dx) ? (long long)std::min(32, c.w - x0) * std::min(32, c.h - y0) : (long long)c.w * c.h;
            }
            for (size_t u = 0; u < gall.unit_cmd.size(); u++)
                for (uint32_t k = 0; k < gall.unit_cnt[u]; k++)
                    pl.unit_px += (long long)cmds[gall.unit_cmd[u] + k].w * cmds[gall.unit_cmd[u] + k].h;
        }
        if (pl.n_items) {
            // Largest items first (they are independent, any order is legal): the four warps of a CTA
            // get items of similar size, and the kernel's tail is made of small items, not of a 32x32
            // tile that happened to be recorded last.
            std::vector<uint32_t> &sorted = ctx->ritem_tmp2, &bk = ctx->bucket_tmp;
            auto weight = [&](uint32_t it) {
                const Eav1Cmd &c = cmds[it & 0xffffff];
                const int x0 = (int)((it >> 24) & 3) * 32, y0 = (int)((it >> 26) & 3) * 32;
                const int px = (c.op == EAV1_OP_MC && !c.u.mc.dx) ? std::min(32, c.w - x0) * std::min(32, c.h - y0) : (int)c.w * c.h;
                int lg = 0; while ((2 << lg) <= px) lg++;
                return 15 - std::min(lg, 15);   // bucket 0 = largest
            };
            bk.assign(17, 0);
            for (uint32_t it : gall.mc_items) bk[weight(it) + 1]++;
            for (size_t b = 1; b < bk.size(); b++) bk[b] += bk[b - 1];
            sorted.resize(gall.mc_items.size());
            for (uint32_t it : gall.mc_items) sorted[bk[weight(it)]++] = it;
            up.put(sorted.data(), sorted.size() * 4, &pl.d_items);
        }
        // Units that wait for nothing inside the frame run in a flat launch of their own (phase 2a); only the
        // units with a producer among the phase-2 units go through the dataflow kernel, and a producer that
        // already ran in phase 2a is no longer a dependency. On inter frames that leaves a few dozen units.
        {
            const UnitGraph &ga = ctx->units;
            UnitGraph &g2 = ctx->units2;
            g2.mc_items.clear(); g2.unit_cmd.clear(); g2.unit_cnt.clear(); g2.dep_start.clear(); g2.deps.clear();
            std::vector<uint2> &flat = ctx->flat_tmp;
            std::vector<uint32_t> &remap = ctx->remap_tmp;
            const size_t nu = ga.unit_cmd.size();
            remap.assign(nu, UINT32_MAX);
            flat.clear();
            size_t n_flat_big = 0;
            for (int pass = 0; pass < 2; pass++)        // large units lead the list
                for (size_t u = 0; u < nu; u++) {
                    if (ga.dep_start[u + 1] != ga.dep_start[u] || ctx->no_flat) continue;
                    const Eav1Cmd &c = cmds[ga.unit_cmd[u]];
                    const bool big = c.w > 32 || c.h > 32;
                    if (big != (pass == 0)) continue;
                    flat.push_back(make_uint2(ga.unit_cmd[u], ga.unit_cnt[u]));
                    n_flat_big += big;
                }
            for (size_t u = 0; u < nu; u++) {
                if (ga.dep_start[u + 1] == ga.dep_start[u] && !ctx->no_flat) continue;
                remap[u] = (uint32_t)g2.unit_cmd.size();
                g2.unit_cmd.push_back(ga.unit_cmd[u]);
                g2.unit_cnt.push_back(ga.unit_cnt[u]);
                g2.dep_start.push_back((uint32_t)g2.deps.size());
                for (uint32_t k = ga.dep_start[u]; k < ga.dep_start[u + 1]; k++)
                    if (remap[ga.deps[k]] != UINT32_MAX) g2.deps.push_back(remap[ga.deps[k]]);
            }
            g2.dep_start.push_back((uint32_t)g2.deps.size());
            pl.n_flat_big = (uint32_t)n_flat_big; pl.n_flat_small = (uint32_t)(flat.size() - n_flat_big);
            if (!flat.empty()) {
                up.put(flat.data(), flat.size() * sizeof(uint2), &pl.d_flat);
            }
        }
        const UnitGraph &g = ctx->units2;
        // never hand the kernel a null list, even when nothing has a producer
        static const uint32_t zero = 0;
        up.put(g.deps.empty() ? &zero : g.deps.data(), std::max<size_t>(g.deps.size(), 1) * 4, &pl.d_deps);
        pl.n_chains = (uint32_t)g.unit_cmd.size(); pl.n_deps = (uint32_t)g.deps.size();
        if (staged) pl.n_levels = Scheduler::unit_levels(g);
        if (pl.n_chains) {
            // Two ticket lists, each in decode order. A unit runs on a single warp when it is at
            // most 32x32 (scratch size) and neither waits for nor feeds another unit (only
            // throughput matters); everything else gets a whole CTA (latency matters).
            const size_t nc = g.unit_cmd.size();
            std::vector<uint8_t> &fed = ctx->fed_tmp;
            fed.assign(nc, 0);
            for (uint32_t p : g.deps) fed[p] = 1;
            std::vector<ChainRec> &rt = ctx->rec_tmp;
            rt.resize(nc);
            size_t nb = 0;
            double px_big = 0, px_small = 0;
            std::vector<uint32_t> &order = ctx->order_tmp;
            Scheduler::ticket_order(g, (uint32_t)ctx->est_weight, order, ctx->est_tmp, ctx->tail_tmp, ctx->bucket_tmp);
            for (int pass = 0; pass < 2; pass++) {
                size_t pos = pass ? nb : 0;
                for (size_t oi = 0; oi < nc; oi++) {
                    const size_t u = order[oi];
                    const Eav1Cmd &c = cmds[g.unit_cmd[u]];
                    const bool linked = fed[u] || g.dep_start[u + 1] != g.dep_start[u];
                    bool big = c.w > 32 || c.h > 32;
                    if (ctx->warp_mode == 0) big = true;
                    else if (ctx->warp_mode == 1) big |= linked;
                    if (big != (pass == 0)) continue;
                    ChainRec &rc = rt[pos++];
                    rc.first = g.unit_cmd[u]; rc.count = g.unit_cnt[u];
                    rc.dep_off = g.dep_start[u]; rc.n_deps = g.dep_start[u + 1] - g.dep_start[u];
                    rc.id = (uint32_t)u; rc.pad[0] = fed[u]; rc.pad[1] = rc.pad[2] = 0;
                    if (big) px_big += (double)c.w * c.h; else px_small += (double)c.w * c.h;
                }
                if (!pass) nb = pos;
            }
            pl.n_big = (uint32_t)nb; pl.n_small = (uint32_t)(nc - nb);
            pl.big_share = px_big + px_small > 0 ? px_big / (px_big + px_small) : 0;
            up.put(rt.data(), rt.size() * sizeof(ChainRec), &pl.d_recs);
        } else { pl.n_big = pl.n_small = 0; }
        up.put(payload, payload_bytes, &pl.d_payload);
        if (arena_bytes + 16 > ctx->arena.cap) {
            // grow; the old contents are kept (operator-level callers upload prep blocks before they submit)
            CK(cudaStreamSynchronize(ctx->stream));
            DevBuf bigger;
            if ((r = ensure(ctx, bigger, arena_bytes + 16, false))) return r;
            if (ctx->arena.d) { CK(cudaMemcpy(bigger.d, ctx->arena.d, ctx->arena.cap, cudaMemcpyDeviceToDevice)); cudaFree(ctx->arena.d); }
            ctx->arena.d = bigger.d; ctx->arena.cap = bigger.cap;
        }
        if (g.unit_cmd.size() > ctx->done_cap) {
            if (ctx->d_done) { CK(cudaStreamSynchronize(ctx->stream)); cudaFree(ctx->d_done); ctx->d_done = nullptr; }
            const size_t cap = g.unit_cmd.size() * 2 + 4096;
            CK(cudaMalloc(&ctx->d_done, cap * 4));
            CK(cudaMemsetAsync(ctx->d_done, 0, cap * 4, ctx->stream));
            ctx->done_cap = cap;
        }
    }
    if (fd->n_lf_recs) {
        // partition the recorded calls: all column-edge filters run before all row-edge filters
        std::vector<Eav1LfRec> &v = ctx->lf_sorted;
        v.resize(fd->n_lf_recs);
        size_t n0 = 0;
        for (size_t i = 0; i < fd->n_lf_recs; i++) if (fd->lf_recs[i].dir == 0) v[n0++] = fd->lf_recs[i];
        size_t n1 = n0;
        for (size_t i = 0; i < fd->n_lf_recs; i++) if (fd->lf_recs[i].dir != 0) v[n1++] = fd->lf_recs[i];
        up.put(v.data(), v.size() * sizeof(Eav1LfRec), &pl.d_lf);
        pl.n_lf0 = n0; pl.n_lf1 = n1 - n0;
        up.put(fd->lf_level, fd->lf_level_bytes, &pl.d_lflvl);
    } else { pl.n_lf0 = pl.n_lf1 = 0; }
    if (has_cdef)
        up.put(fd->lf_mask, (size_t)fd->sb128w * fd->sb128h * 1348, &pl.d_lfmask);
    pl.n_lr = fd->n_lr_recs;
    if (has_lr) up.put(fd->lr_recs, fd->n_lr_recs * sizeof(Eav1LrRec), &pl.d_lr);
    return up.flush(ctx->S().stage);
}

// device-to-device duplicate of a picture on stream st (EAV1_COPY_ENGINE=1: through the copy engine, for A/B runs)
// dst2 (optional): a second destination of the same size
static cudaError_t picture_copy(Eav1cuCtx *ctx, void *dst, const void *src, size_t bytes, cudaStream_t st, void *dst2 = nullptr)
{
    static const bool engine = getenv("EAV1_COPY_ENGINE") != nullptr;
    if (engine || (bytes & 15) || ((uintptr_t)dst & 15) || ((uintptr_t)src & 15) || ((uintptr_t)dst2 & 15)) {
        cudaError_t e = cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToDevice, st);
        if (e == cudaSuccess && dst2) e = cudaMemcpyAsync(dst2, src, bytes, cudaMemcpyDeviceToDevice, st);
        return e;
    }
    const size_t n16 = bytes / 16;
    const unsigned grid = (unsigned)std::min<size_t>((n16 + 255) / 256, (size_t)ctx->recon_sms * 8);
    ctx->n_launches++;
    return launch_k(picture_copy_kernel, dim3(grid), dim3(256), st, (pdl_mask() & 1) != 0, static_cast<uint4 *>(dst), static_cast<uint4 *>(dst2), static_cast<const uint4 *>(src), n16);
}

static int plan_launch(Eav1cuCtx *ctx, FramePlan &pl)
{
    const Eav1FrameDesc *fd = &pl.fd;
    Picture &cur = ctx->pics[fd->cur];
    const bool hbd = fd->bpc > 8;
    const bool has_cdef = fd->cdef_enable != 0, has_lr = pl.n_lr > 0;
    cudaStream_t st = ctx->stream;
    long long n0 = ctx->n_launches;
    // recon + deblock happen in A; CDEF / LR then move the frame towards `cur` out of place.
    // With super-resolution A is `cur` itself (coded width) and the frame ends up in fd->sr_pic.
    const bool superres = fd->superres != 0;
    Picture &A = superres ? cur : (has_cdef || has_lr) ? ctx->work[0] : cur;

    FrameCtx fc;
    memset(&fc, 0, sizeof(fc));
    fill_planes(A, fc.cur);
    for (int r = 0; r < 7; r++) {
        const int id = fd->ref[r];
        if (id >= 0 && id < (int)ctx->pics.size() && ctx->pics[id].base) fill_planes(ctx->pics[id], fc.ref[r]);
        else fill_planes(A, fc.ref[r]);
    }
    fill_planes(A, fc.ref[7]);
    fc.bitdepth_max = (1 << fd->bpc) - 1;
    fc.ss_hor = fd->layout != 3;
    fc.ss_ver = fd->layout == 1;
    fc.tab = ctx->d_tab;
    fc.masks = ctx->d_masks;
    fc.arena = ctx->arena.d;
    fc.payload = pl.d_payload;
    // A submission without commands is a post-filter call on its own (operator-level use): fd->cur then holds the
    // reconstructed, unfiltered frame, which the filters expect in A.
    if (!pl.n_cmds && A.base != cur.base)
        CK(cudaMemcpyAsync(A.base, cur.base, std::min(A.bytes, cur.bytes), cudaMemcpyDeviceToDevice, st));
    // Phase 1: motion compensation on the main stream, the residuals of the same frame beside it
    // on a second stream (they share nothing); the dataflow kernel waits for both.
    const bool fork = pl.n_ritems && ctx->stream2 && !ctx->serial_phase1;
    cudaStream_t rs = fork ? ctx->stream2 : st;
    if (fork) { TCK(TC_EVREC, cudaEventRecord(ctx->ev_fork, st)); TCK(TC_WAITEV, cudaStreamWaitEvent(rs, ctx->ev_fork, 0)); }
    if (pl.n_ritems) {
        StageTimer tm(ctx, 6, rs);
        const unsigned small_ctas = (pl.n_ritems_small + 3) / 4;
        const unsigned grid = small_ctas + (pl.n_ritems - pl.n_ritems_small);
        if (hbd) resid_kernel<BD16><<<grid, 128, 0, rs>>>(fc, pl.d_cmds, pl.d_ritems, pl.n_ritems_small, small_ctas, pl.n_ritems);
        else     resid_kernel<BD8><<<grid, 128, 0, rs>>>(fc, pl.d_cmds, pl.d_ritems, pl.n_ritems_small, small_ctas, pl.n_ritems);
        ctx->n_launches++;
    }
    if (fork) TCK(TC_EVREC, cudaEventRecord(ctx->ev_join, rs));
    if (pl.n_items) {
        StageTimer tm(ctx, 4);
        const unsigned grid = (pl.n_items + 3) / 4;
        const bool pdl = (pdl_mask() & 4) != 0;
        if (hbd) CK(launch_k(mc_kernel<BD16>, dim3(grid), dim3(128), st, pdl, fc, pl.d_cmds, pl.d_items, pl.n_items));
        else     CK(launch_k(mc_kernel<BD8>, dim3(grid), dim3(128), st, pdl, fc, pl.d_cmds, pl.d_items, pl.n_items));
        ctx->n_launches++;
    }
    if (fork) TCK(TC_WAITEV, cudaStreamWaitEvent(st, ctx->ev_join, 0));
    if (pl.n_flat_big + pl.n_flat_small) {
        StageTimer tm(ctx, 0);
        const unsigned grid = pl.n_flat_big + (pl.n_flat_small + 3) / 4;
        if (hbd) flat_units_kernel<BD16><<<grid, 128, 0, st>>>(fc, pl.d_cmds, pl.d_flat, pl.n_flat_big, pl.n_flat_small);
        else     flat_units_kernel<BD8><<<grid, 128, 0, st>>>(fc, pl.d_cmds, pl.d_flat, pl.n_flat_big, pl.n_flat_small);
        ctx->n_launches++;
    }
    if (pl.n_chains) {
        StageTimer tm(ctx, 0);
        if (pl.n_chains > ctx->done_cap) return fail(ctx, EAV1CU_ERR_INVAL, "done-flag array smaller than the staged frame");
        ReconArgs ra;
        ra.cmds = pl.d_cmds; ra.deps = pl.d_deps;
        uint32_t ep = ++g_epoch;
        if (ep == 0 || ep < ctx->epoch) {
            // the 32-bit launch epoch wrapped (about a week of launches): never-written flags hold 0 and stale flags
            // could match again, so clear this context's flags (stream-ordered) and never hand out epoch 0
            CK(cudaMemsetAsync(ctx->d_done, 0, ctx->done_cap * 4, st));
            while (ep == 0) ep = ++g_epoch;
        }
        ra.done = ctx->d_done; ra.ctrl = ctx->d_ctrl; ra.epoch = ctx->epoch = ep;
        ra.big = pl.d_recs; ra.small = pl.d_recs + pl.n_big;
        ra.n_big = pl.n_big; ra.n_small = pl.n_small;
        // CTAs are split between the two kinds in proportion to their pixel work
        // The dataflow kernel is bound by the frame's dependency chains, not by its worker count (one CTA per SM
        // instead of three costs one stream 9 %), but at three CTAs per SM its 168 registers per thread fill the register
        // file and nothing of another stream runs beside it. With several decoders on the GPU a narrower grid wins:
        // 8 streams 20.7k -> 23.5k frames/s at one CTA per SM, 4 streams 19.9k -> 21.3k.
        unsigned grid = (unsigned)ctx->recon_grid;
        if (!ctx->recon_grid_fixed) {
            const int live = g_live_ctx.load(std::memory_order_relaxed);
            const unsigned per = live >= 4 ? 1 : live >= 2 ? 2 : 3;
            grid = std::min(grid, (unsigned)ctx->recon_sms * per);
        }
        unsigned big_ctas = pl.n_big ? std::max(1u, (unsigned)(grid * pl.big_share + 0.5)) : 0;
        big_ctas = std::min(big_ctas, pl.n_big);
        unsigned small_ctas = pl.n_small ? std::min<unsigned>((pl.n_small + 3) / 4, std::max(1u, grid - std::min(grid - 1, big_ctas))) : 0;
        if (!pl.n_small) big_ctas = std::min<unsigned>(pl.n_big, grid);   // never more CTAs than units: the rest of the GPU stays free for other streams
        grid = big_ctas + small_ctas;
        ra.n_big_ctas = big_ctas; ra.n_workers = big_ctas + 4 * small_ctas;
        ra.debug = (uint32_t)ctx->debug_check;
        ra.trace = ctx->d_trace;
        ra.hist = ctx->d_trace ? ctx->d_trace + (size_t)pl.n_chains * EAV1_TRACE_WORDS : nullptr;
        if (ctx->debug_check == 2) {
            // verify what the kernels will read: checksum payload and commands on the device, in stream order
            const int slot = (int)(ctx->n_frames & 63);
            unsigned long long old[2];
            if (ctx->sum_frame[slot]) {   // the entry written 64 frames ago has long completed: compare it
                cudaMemcpy(old, ctx->d_sums + slot * 2, 16, cudaMemcpyDeviceToHost);
                if (old[0] != ctx->h_sums[slot][0] || old[1] != ctx->h_sums[slot][1])
                    fprintf(stderr, "EAV1 DEBUG: frame %lld upload mismatch payload %llx/%llx cmds %llx/%llx\n", ctx->sum_frame[slot] - 1, old[0], ctx->h_sums[slot][0], old[1], ctx->h_sums[slot][1]);
            }
            cudaMemsetAsync(ctx->d_sums + slot * 2, 0, 16, st);
            unsigned long long hs = 0;
            const uint32_t *hp = reinterpret_cast<const uint32_t *>(ctx->S().stage.h + ((const uint8_t *)pl.d_payload - ctx->S().stage.d));
            for (size_t i = 0; i < pl.payload_bytes / 4; i++) hs += hp[i];
            ctx->h_sums[slot][0] = hs;
            hs = 0;
            hp = reinterpret_cast<const uint32_t *>(ctx->S().stage.h + ((const uint8_t *)pl.d_cmds - ctx->S().stage.d));
            for (size_t i = 0; i < pl.n_cmds * sizeof(Eav1Cmd) / 4; i++) hs += hp[i];
            ctx->h_sums[slot][1] = hs;
            ctx->sum_frame[slot] = ctx->n_frames + 1;
            if (pl.payload_bytes) checksum_kernel<<<64, 256, 0, st>>>(reinterpret_cast<const uint32_t *>(pl.d_payload), pl.payload_bytes / 4, ctx->d_sums + slot * 2);
            checksum_kernel<<<64, 256, 0, st>>>(reinterpret_cast<const uint32_t *>(pl.d_cmds), pl.n_cmds * sizeof(Eav1Cmd) / 4, ctx->d_sums + slot * 2 + 1);
        }
        if (ctx->debug_check == 1) {
            uint32_t c[8];
            CK(cudaStreamSynchronize(st));
            CK(cudaMemcpy(c, ctx->d_ctrl, sizeof(c), cudaMemcpyDeviceToHost));
            if (c[0] | c[1] | c[2]) fprintf(stderr, "EAV1 DEBUG: ctrl not armed before launch: %u %u %u (frame %lld)\n", c[0], c[1], c[2], ctx->n_frames);
            CK(cudaMemset(ctx->d_ctrl + 4, 0, 4));
            CK(cudaDeviceSynchronize());
        }
        const bool pdl = (pdl_mask() & 2) != 0;
        if (hbd) CK(launch_k(recon_frame_kernel<BD16>, dim3(grid), dim3(128), st, pdl, fc, ra));
        else     CK(launch_k(recon_frame_kernel<BD8>, dim3(grid), dim3(128), st, pdl, fc, ra));
        ctx->n_launches++;
        if (ctx->debug_check == 1) {
            uint32_t c[8];
            CK(cudaStreamSynchronize(st));
            CK(cudaMemcpy(c, ctx->d_ctrl, sizeof(c), cudaMemcpyDeviceToHost));
            if (c[4] != pl.n_chains || (c[0] | c[1] | c[2]))
                fprintf(stderr, "EAV1 DEBUG: frame %lld executed %u of %u units, ctrl after = %u %u %u, grid %u workers %u\n",
                        ctx->n_frames, c[4], pl.n_chains, c[0], c[1], c[2], grid, ra.n_workers);
        }
    }

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
    pp.level = pl.d_lflvl;
    if (pl.n_lf0 + pl.n_lf1) {
    StageTimer tm(ctx, 1);
    for (int pass = 0; pass < 2; pass++) {
        const size_t off = pass ? pl.n_lf0 : 0, cnt = pass ? pl.n_lf1 : pl.n_lf0;
        if (!cnt) continue;
        const unsigned grid = (unsigned)cnt;
        const bool pdl = (pdl_mask() & 1) != 0;
        if (hbd) CK(launch_k(deblock_kernel<uint16_t>, dim3(grid), dim3(128), st, pdl, pp, pl.d_lf + off, (int)cnt));
        else     CK(launch_k(deblock_kernel<uint8_t>, dim3(grid), dim3(128), st, pdl, pp, pl.d_lf + off, (int)cnt));
        ctx->n_launches++;
    }
    }
    bool cdef_mirrored = false;
    if (has_cdef) {
        StageTimer tm(ctx, 2);
        Picture &dst = (has_lr || superres) ? ctx->work[1] : cur;
        for (int i = 0; i < 3; i++) pp.b[i] = dst.pl[i];
        pp.lfmask = pl.d_lfmask;
        // Most 8x8 blocks of a typical frame are left alone by CDEF (skipped, strength 0, cdef_idx -1): the whole
        // picture is copied by the copy engine first (L2-resident, a few microseconds) and the kernel only writes
        // the blocks it filters; a warp copying its own 96 bytes spent its life waiting for them.
        // With loop restoration next (and no super-resolution in between) the final picture receives the same copy
        // and the same filtered samples: LR's own whole-picture copy is not needed then.
        cdef_mirrored = has_lr && !superres && &dst != &cur && cur.bytes == dst.bytes;
        pp.cdef_mirror = cdef_mirrored;
        if (cdef_mirrored) for (int i = 0; i < 3; i++) pp.p[i] = cur.pl[i];
        TCK(TC_D2D, picture_copy(ctx, dst.base, A.base, std::min(dst.bytes, A.bytes), st, cdef_mirrored ? cur.base : nullptr));
        dim3 grid((fd->bw4 / 2 + 8) / 8, (fd->bh4 + 1) / 2);
        const bool pdl = (pdl_mask() & 1) != 0;
        if (hbd) CK(launch_k(cdef_kernel<uint16_t>, grid, dim3(256), st, pdl, pp));
        else     CK(launch_k(cdef_kernel<uint8_t>, grid, dim3(256), st, pdl, pp));
        ctx->n_launches++;
    }
    Picture *lr_src = has_cdef ? &ctx->work[1] : &A, *lr_out = &cur;
    if (superres) {
        // mc.resize over the whole frame (recon_tmpl.c:2053-2086): the CDEF output becomes the pre-LR picture at
        // the upscaled width; with CDEF on, the deblocked frame is upscaled as well because the rows above and
        // below a restoration stripe are deblocked, not CDEF-filtered, samples (lf_apply_tmpl.c:76-88)
        StageTimer tm(ctx, 3);
        Picture &out = ctx->pics[fd->sr_pic];
        auto resize = [&](const Picture &from, const Picture &to) {
            ResizeParams rp;
            memset(&rp, 0, sizeof(rp));
            rp.n_planes = fd->layout ? 3 : 1;
            int max_w = 0, max_h = 0;
            for (int i = 0; i < rp.n_planes; i++) {
                const int sh = i && fd->layout != 3, sv = i && fd->layout == 1;
                rp.src[i] = from.pl[i]; rp.dst[i] = to.pl[i];
                rp.dst_w[i] = (fd->sr_w + sh) >> sh; rp.src_w[i] = (4 * fd->bw4 + sh) >> sh; rp.h[i] = (fd->h + sv) >> sv;
                rp.dx[i] = fd->resize_step[!!i]; rp.mx0[i] = fd->resize_start[!!i];
                max_w = std::max(max_w, rp.dst_w[i]); max_h = std::max(max_h, rp.h[i]);
            }
            rp.bitdepth_max = (1 << fd->bpc) - 1;
            rp.tab = ctx->d_tab;
            const dim3 grid((max_w + 255) / 256, max_h, rp.n_planes);
            if (hbd) resize_kernel<uint16_t><<<grid, 256, 0, st>>>(rp);
            else     resize_kernel<uint8_t><<<grid, 256, 0, st>>>(rp);
            ctx->n_launches++;
        };
        if (has_lr) {
            resize(*lr_src, ctx->work[2]);
            if (has_cdef) resize(A, ctx->work[3]);
            for (int i = 0; i < 3; i++) pp.a[i] = has_cdef ? ctx->work[3].pl[i] : ctx->work[2].pl[i];
            lr_src = &ctx->work[2];
            lr_out = &out;
        } else resize(*lr_src, out);
    }
    if (has_lr) {
        StageTimer tm(ctx, 3);
        Picture &src = *lr_src, &cur = *lr_out;
        for (int i = 0; i < 3; i++) { pp.b[i] = src.pl[i]; pp.p[i] = cur.pl[i]; }
        if (!cdef_mirrored) TCK(TC_D2D, picture_copy(ctx, cur.base, src.base, std::min(cur.bytes, src.bytes), st));
        dim3 grid(384 / LR_TW, (unsigned)pl.n_lr, 64 / LR_SUB_ROWS);
        const bool pdl = (pdl_mask() & 1) != 0;
        if (hbd) CK(launch_k(lr_kernel<uint16_t>, grid, dim3(128), st, pdl, pp, pl.d_lr));
        else     CK(launch_k(lr_kernel<uint8_t>, grid, dim3(128), st, pdl, pp, pl.d_lr));
        ctx->n_launches++;
    }
    CK(cudaGetLastError());
    pl.launches = ctx->n_launches - n0;
    return 0;
}

extern "C" {

int eav1cu_frame_submit(Eav1cuCtx *ctx, const Eav1FrameDesc *fd,
                        const Eav1Cmd *cmds, size_t n_cmds,
                        const void *payload, size_t payload_bytes, size_t arena_bytes)
{
    if (!ctx || !fd) return EAV1CU_ERR_INVAL;
    if (ctx->err) return ctx->err;
    if (!ctx->have_tab) return fail(ctx, EAV1CU_ERR_NOTABLES, "eav1cu_set_tables not called");
    coverage_count(fd, cmds, n_cmds);
    CK(cudaSetDevice(ctx->device));
    ctx->cur_slot ^= 1;
    Slot &sl = ctx->S();
    if (!sl.done) CK(cudaEventCreateWithFlags(&sl.done, ctx->event_flags));
    auto t0 = std::chrono::steady_clock::now();
    if (sl.used) TCK(TC_EVSYNC_SLOT, cudaEventSynchronize(sl.done));   // the frame two submits ago has finished with this slot
    auto t1 = std::chrono::steady_clock::now();
    int r = plan_build(ctx, ctx->live, false, fd, cmds, n_cmds, payload, payload_bytes, arena_bytes);
    if (r) return r;
    auto t2 = std::chrono::steady_clock::now();
    {
        std::lock_guard<SubmitMutex> g(g_submit_mu);
        { CallTimer ct_(TC_LAUNCHES); if ((r = plan_launch(ctx, ctx->live))) return r; }
        TCK(TC_EVREC, cudaEventRecord(sl.done, ctx->stream));
    }
    auto t3 = std::chrono::steady_clock::now();
    ctx->t_wait += std::chrono::duration<double>(t1 - t0).count();
    ctx->t_build += std::chrono::duration<double>(t2 - t1).count();
    ctx->t_launch += std::chrono::duration<double>(t3 - t2).count();
    sl.used = true;
    ctx->n_frames++;
    g_total[0] += ctx->live.launches; g_total[3]++;
    if (ctx->capture) {
        // keep a replayable copy of this frame (bench.py's device-resident leg)
        FramePlan *sp = new FramePlan();
        if ((r = plan_build(ctx, *sp, true, fd, cmds, n_cmds, payload, payload_bytes, arena_bytes))) { delete sp; return r; }
        ctx->staged.push_back(sp);
    }
    return ctx->err;
}

int eav1cu_coverage(Eav1Coverage *out, int reset) { return coverage_read(out, reset); }

size_t eav1cu_abi_sizeof(int which)
{
    switch (which) {
    case 0: return sizeof(Eav1Cmd);
    case 1: return sizeof(Eav1LfRec);
    case 2: return sizeof(Eav1LrRec);
    case 3: return sizeof(Eav1FrameDesc);
    case 4: return sizeof(Eav1StagedInfo);
    case 5: return sizeof(Eav1Tables);
    case 6: return sizeof(Eav1Coverage);
    default: return 0;
    }
}

int eav1cu_plan_levels(const Eav1Cmd *cmds, size_t n_cmds, const int plane_w[3], const int plane_h[3],
                       uint32_t *chain_first, uint32_t *chain_level, size_t cap)
{
    if (!cmds || !plane_w || !plane_h) return EAV1CU_ERR_INVAL;
    Scheduler sc;
    Schedule plan;
    sc.reset(plane_w, plane_h);
    sc.build(cmds, n_cmds, plan);
    size_t k = 0;
    for (size_t l = 0; l + 1 < plan.level_start.size(); l++)
        for (uint32_t c = plan.level_start[l]; c < plan.level_start[l + 1]; c++, k++)
            if (k < cap) {
                if (chain_first) chain_first[k] = plan.chains[c].first;
                if (chain_level) chain_level[k] = (uint32_t)l;
            }
    return (int)plan.chains.size();
}

int eav1cu_plan_units(const Eav1Cmd *cmds, size_t n_cmds, const int plane_w[3], const int plane_h[3],
                      uint32_t *unit_first, uint32_t *unit_count, uint32_t *unit_level, size_t cap_units,
                      uint32_t *mc_items, size_t cap_items, size_t *n_items)
{
    if (!cmds || !plane_w || !plane_h) return EAV1CU_ERR_INVAL;
    Scheduler sc;
    UnitGraph g;
    sc.reset(plane_w, plane_h);
    sc.build_units(cmds, n_cmds, g);
    std::vector<uint32_t> lvl;
    Scheduler::unit_levels(g, &lvl);
    for (size_t u = 0; u < g.unit_cmd.size() && u < cap_units; u++) {
        if (unit_first) unit_first[u] = g.unit_cmd[u];
        if (unit_count) unit_count[u] = g.unit_cnt[u];
        if (unit_level) unit_level[u] = lvl[u];
    }
    for (size_t i = 0; i < g.mc_items.size() && i < cap_items; i++) if (mc_items) mc_items[i] = g.mc_items[i];
    if (n_items) *n_items = g.mc_items.size();
    return (int)g.unit_cmd.size();
}

/* The order in which the dataflow kernel hands out the units of eav1cu_plan_units (ticket t runs
 * unit order[t]) and the producer lists it waits on (CSR: dep_start has n+1 entries). No GPU needed. */
int eav1cu_plan_tickets(const Eav1Cmd *cmds, size_t n_cmds, const int plane_w[3], const int plane_h[3],
                        int est_weight, uint32_t *order, size_t cap_units, uint32_t *dep_start,
                        uint32_t *deps, size_t cap_deps, size_t *n_deps)
{
    if (!cmds || !plane_w || !plane_h || est_weight < 1) return EAV1CU_ERR_INVAL;
    Scheduler sc;
    UnitGraph g;
    sc.reset(plane_w, plane_h);
    sc.build_units(cmds, n_cmds, g);
    std::vector<uint32_t> ord, est, tail, bucket;
    Scheduler::ticket_order(g, (uint32_t)est_weight, ord, est, tail, bucket);
    const size_t n = g.unit_cmd.size();
    if (n > cap_units || g.deps.size() > cap_deps) return EAV1CU_ERR_INVAL;
    for (size_t u = 0; u < n; u++) { if (order) order[u] = ord[u]; if (dep_start) dep_start[u] = g.dep_start[u]; }
    if (dep_start) dep_start[n] = g.dep_start[n];
    for (size_t i = 0; i < g.deps.size(); i++) if (deps) deps[i] = g.deps[i];
    if (n_deps) *n_deps = g.deps.size();
    return (int)n;
}

int eav1cu_global_counters(long long out[4])
{
    if (!out) return EAV1CU_ERR_INVAL;
    for (int i = 0; i < 4; i++) out[i] = g_total[i];
    return 0;
}

int eav1cu_counters(Eav1cuCtx *ctx, long long out[4])
{
    if (!ctx || !out) return EAV1CU_ERR_INVAL;
    out[0] = ctx->n_launches; out[1] = ctx->h2d_bytes; out[2] = ctx->d2h_bytes; out[3] = ctx->n_frames;
    return 0;
}

int eav1cu_set_stage_timing(Eav1cuCtx *ctx, int on)
{
    if (!ctx) return EAV1CU_ERR_INVAL;
    ctx->stage_timing = on != 0;
    return 0;
}

int eav1cu_stage_times(Eav1cuCtx *ctx, double *ms, long long *launches, long long *calls, int reset)
{
    if (!ctx) return EAV1CU_ERR_INVAL;
    CK(cudaSetDevice(ctx->device));
    CK(cudaStreamSynchronize(ctx->stream));
    for (auto &sp : ctx->spans) {
        float t = 0.f;
        if (cudaEventElapsedTime(&t, sp.a, sp.b) == cudaSuccess) {
            ctx->stage_ms[sp.stage] += t;
            ctx->stage_launches[sp.stage] += sp.launches;
            ctx->stage_calls[sp.stage]++;
        }
        ctx->ev_pool.push_back(sp.a); ctx->ev_pool.push_back(sp.b);
    }
    ctx->spans.clear();
    for (int i = 0; i < EAV1CU_N_STAGES; i++) {
        if (ms) ms[i] = ctx->stage_ms[i];
        if (launches) launches[i] = ctx->stage_launches[i];
        if (calls) calls[i] = ctx->stage_calls[i];
        if (reset) { ctx->stage_ms[i] = 0; ctx->stage_launches[i] = 0; ctx->stage_calls[i] = 0; }
    }
    return 0;
}

int eav1cu_set_capture(Eav1cuCtx *ctx, int on)
{
    if (!ctx) return EAV1CU_ERR_INVAL;
    ctx->capture = on != 0;
    return 0;
}

int eav1cu_staged_count(Eav1cuCtx *ctx) { return ctx ? (int)ctx->staged.size() : 0; }

int eav1cu_staged_replay(Eav1cuCtx *ctx, int idx)
{
    if (!ctx || idx < 0 || idx >= (int)ctx->staged.size()) return EAV1CU_ERR_INVAL;
    if (ctx->err) return ctx->err;
    std::lock_guard<SubmitMutex> g(g_submit_mu);
    return plan_launch(ctx, *ctx->staged[idx]);
}

/* Replays every staged frame in submit order, `passes` times, from one C loop (bench.py runs one host thread per
 * staged decoder so that several streams keep the GPU busy at once). Returns 0 or the first error. */
int eav1cu_staged_replay_all(Eav1cuCtx *ctx, int passes)
{
    if (!ctx || passes < 0) return EAV1CU_ERR_INVAL;
    if (ctx->err) return ctx->err;
    CK(cudaSetDevice(ctx->device));
    for (int p = 0; p < passes; p++)
        for (FramePlan *pl : ctx->staged) {
            std::lock_guard<SubmitMutex> g(g_submit_mu);
            const int r = plan_launch(ctx, *pl);
            if (r) return r;
        }
    return 0;
}

/* The staged frames of a context as ONE CUDA graph: the launches of all frames (both streams of phase 1, the
 * device-to-device copy of loop restoration) are captured once, a replay is a single cudaGraphLaunch. The dataflow
 * kernel's done flags carry launch epochs that are baked into the captured arguments, so the graph starts by clearing
 * this context's flags (an epoch can then never be met again from an earlier replay). Launch-bound replays (several
 * staged streams sharing the GPU) then run at the device's pace instead of the host's. */
struct FgLaunch;
static int fg_fill(Eav1cuCtx *ctx, int src_pic, int dst_pic, const Eav1GrainParams *g, FgLaunch &L);
static int fg_side_stream(Eav1cuCtx *ctx);
static cudaError_t fg_launch_prepare(const FgLaunch &L, cudaStream_t st);
static cudaError_t fg_launch_apply(const FgLaunch &L, cudaStream_t st);
static FgLaunch *fg_launch_new();
static void fg_launch_delete(FgLaunch *);

int eav1cu_staged_graph_build_grain(Eav1cuCtx *ctx, int grain_dst_pic, const Eav1GrainParams *grain)
{
    if (!ctx || ctx->staged.empty()) return EAV1CU_ERR_INVAL;
    if (ctx->err) return ctx->err;
    CK(cudaSetDevice(ctx->device));
    if (grain) {        // everything the grain launches need must exist before the capture begins
        FgLaunch *probe = fg_launch_new();
        int r = fg_fill(ctx, ctx->staged[0]->fd.cur, grain_dst_pic, grain, *probe);
        fg_launch_delete(probe);
        if (!r) r = fg_side_stream(ctx);
        if (r) return r;
    }
    CK(cudaStreamSynchronize(ctx->stream));
    if (ctx->staged_exec) { cudaGraphExecDestroy(ctx->staged_exec); ctx->staged_exec = nullptr; }
    if (ctx->staged_graph) { cudaGraphDestroy(ctx->staged_graph); ctx->staged_graph = nullptr; }
    const bool timing = ctx->stage_timing;
    ctx->stage_timing = false;
    std::lock_guard<SubmitMutex> g(g_submit_mu);
    CK(cudaStreamBeginCapture(ctx->stream, cudaStreamCaptureModeThreadLocal));
    int r = 0;
    const long long n0 = ctx->n_launches;
    if (ctx->d_done && cudaMemsetAsync(ctx->d_done, 0, ctx->done_cap * 4, ctx->stream) != cudaSuccess) r = EAV1CU_ERR_CUDA;
    FgLaunch *fg = grain ? fg_launch_new() : nullptr;
    // Film grain of every frame runs on a stream of its own, template build and application both: the application
    // only reads the finished frame, so the next frame's kernels need not queue behind it. The grain stream joins the
    // capture here, waits for each frame's last kernel before that frame's application, and is joined at the end. The
    // main stream waits for it only before a frame that reuses the picture of an application still outstanding.
    std::vector<int> fg_reading;
    if (grain && (cudaEventRecord(ctx->ev_fg_cap[0], ctx->stream) != cudaSuccess ||
                  cudaStreamWaitEvent(ctx->stream_fg, ctx->ev_fg_cap[0], 0) != cudaSuccess)) r = EAV1CU_ERR_CUDA;
    for (FramePlan *pl : ctx->staged) {
        if (r) break;
        if (grain) {
            r = fg_fill(ctx, pl->fd.cur, grain_dst_pic, grain, *fg);
            if (r) break;
            if (std::find(fg_reading.begin(), fg_reading.end(), pl->fd.cur) != fg_reading.end() ||
                (pl->fd.sr_pic >= 0 && std::find(fg_reading.begin(), fg_reading.end(), pl->fd.sr_pic) != fg_reading.end())) {
                if (cudaEventRecord(ctx->ev_fg_cap[1], ctx->stream_fg) != cudaSuccess ||
                    cudaStreamWaitEvent(ctx->stream, ctx->ev_fg_cap[1], 0) != cudaSuccess) { r = EAV1CU_ERR_CUDA; break; }
                fg_reading.clear();
            }
            if (fg_launch_prepare(*fg, ctx->stream_fg) != cudaSuccess) { r = EAV1CU_ERR_CUDA; break; }
        }
        r = plan_launch(ctx, *pl);
        if (grain && !r) {
            if (cudaEventRecord(ctx->ev_fg_cap[0], ctx->stream) != cudaSuccess ||
                cudaStreamWaitEvent(ctx->stream_fg, ctx->ev_fg_cap[0], 0) != cudaSuccess ||
                fg_launch_apply(*fg, ctx->stream_fg) != cudaSuccess) r = EAV1CU_ERR_CUDA;
            fg_reading.push_back(pl->fd.cur);
            ctx->n_launches += 2;
        }
    }
    if (grain && !r && (cudaEventRecord(ctx->ev_fg_cap[1], ctx->stream_fg) != cudaSuccess ||
                        cudaStreamWaitEvent(ctx->stream, ctx->ev_fg_cap[1], 0) != cudaSuccess)) r = EAV1CU_ERR_CUDA;
    if (fg) fg_launch_delete(fg);
    ctx->staged_graph_kernels = ctx->n_launches - n0;   // kernel nodes of the graph
    ctx->n_launches = n0;                               // capturing launched nothing
    cudaGraph_t graph = nullptr;
    const cudaError_t e = cudaStreamEndCapture(ctx->stream, &graph);
    ctx->stage_timing = timing;
    if (r || e != cudaSuccess) {
        if (graph) cudaGraphDestroy(graph);
        cudaGetLastError();
        return r ? r : fail(ctx, EAV1CU_ERR_CUDA, "cudaStreamEndCapture", e);
    }
    ctx->staged_graph = graph;
    CK(cudaGraphInstantiate(&ctx->staged_exec, graph, 0));
    return 0;
}

int eav1cu_staged_graph_build(Eav1cuCtx *ctx) { return eav1cu_staged_graph_build_grain(ctx, -1, nullptr); }

int eav1cu_staged_graph_launch(Eav1cuCtx *ctx, int passes)
{
    if (!ctx || !ctx->staged_exec || passes < 0) return EAV1CU_ERR_INVAL;
    if (ctx->err) return ctx->err;
    CK(cudaSetDevice(ctx->device));
    for (int p = 0; p < passes; p++) {
        std::lock_guard<SubmitMutex> g(g_submit_mu);
        CK(cudaGraphLaunch(ctx->staged_exec, ctx->stream));
        ctx->n_launches += ctx->staged_graph_kernels;         // the kernels the graph runs
    }
    return 0;
}

/* Development aid: replay staged frame idx once with per-unit timestamps. out receives 6 words per
 * dataflow unit, indexed by unit id (decode order): ticket drawn, producers ready, done (global
 * timer, ns), id of the producer that finished last (~0 = none), first command << 32 | count,
 * ticket << 32 | CTA << 8 | warp (255 = whole CTA). Returns the number of units, < 0 on error. */
int eav1cu_staged_trace(Eav1cuCtx *ctx, int idx, unsigned long long *out, size_t cap_units)
{
    if (!ctx || !out || idx < 0 || idx >= (int)ctx->staged.size()) return EAV1CU_ERR_INVAL;
    if (ctx->err) return ctx->err;
    CK(cudaSetDevice(ctx->device));
    FramePlan &pl = *ctx->staged[idx];
    const size_t n = pl.n_chains;
    if (n + 43 > cap_units) return EAV1CU_ERR_INVAL;   /* 12 words per unit + a 512-word (op, size) histogram */
    if (!n) return 0;
    const size_t words = n * EAV1_TRACE_WORDS + 512;
    CK(cudaMalloc(&ctx->d_trace, words * sizeof(unsigned long long)));
    CK(cudaMemsetAsync(ctx->d_trace, 0, words * sizeof(unsigned long long), ctx->stream));
    int r = plan_launch(ctx, pl);
    cudaError_t e = cudaStreamSynchronize(ctx->stream);
    if (!r && e == cudaSuccess) e = cudaMemcpy(out, ctx->d_trace, words * sizeof(unsigned long long), cudaMemcpyDeviceToHost);
    cudaFree(ctx->d_trace);
    ctx->d_trace = nullptr;
    if (r) return r;
    if (e != cudaSuccess) return fail(ctx, EAV1CU_ERR_CUDA, "staged trace", e);
    return (int)n;
}

/* Development aid: cycle counters of -DEAV1_PROF=1 builds (zeros otherwise); reading resets them. */
int eav1cu_prof_read(unsigned long long out[64])
{
#if defined(EAV1_PROF) && EAV1_PROF
    static const unsigned long long zero[64] = { 0 };
    if (cudaMemcpyFromSymbol(out, g_prof, sizeof(zero)) != cudaSuccess) return EAV1CU_ERR_CUDA;
    if (cudaMemcpyToSymbol(g_prof, zero, sizeof(zero)) != cudaSuccess) return EAV1CU_ERR_CUDA;
    return 1;
#else
    memset(out, 0, 64 * sizeof(out[0]));
    return 0;
#endif
}

int eav1cu_staged_info(Eav1cuCtx *ctx, int idx, Eav1StagedInfo *out)
{
    if (!ctx || !out || idx < 0 || idx >= (int)ctx->staged.size()) return EAV1CU_ERR_INVAL;
    const FramePlan &p = *ctx->staged[idx];
    memset(out, 0, sizeof(*out));
    out->cur = p.fd.cur; out->w = p.fd.w; out->h = p.fd.h; out->bpc = p.fd.bpc; out->layout = p.fd.layout;
    out->n_levels = (int)p.n_levels;
    out->n_chains = (int)p.n_chains;
    out->n_big = (int)p.n_big; out->n_small = (int)p.n_small;
    out->n_mc_items = (int)p.n_items; out->mc_px = p.mc_px; out->unit_px = p.unit_px;
    out->n_resid_items = (int)p.n_ritems; out->resid_px = p.resid_px;
    out->launches = (int)p.launches;
    out->n_lf = (int)(p.n_lf0 + p.n_lf1); out->n_lr = (int)p.n_lr;
    out->cdef = p.fd.cdef_enable;
    out->n_cmds = (long long)p.n_cmds; out->payload_bytes = (long long)p.payload_bytes;
    for (int i = 0; i < 7; i++) if (p.fd.ref[i] >= 0) out->inter = 1;
    return 0;
}

enum { FG_GRAIN_BYTES = 36864, FG_SCALE_BYTES = 3 * 4096, FG_OFFSET_BYTES = 65536 };

} // extern "C" (helpers)

struct FgLaunch { FgParams P; int bpc; dim3 grid; };
static FgLaunch *fg_launch_new() { return new FgLaunch(); }
static void fg_launch_delete(FgLaunch *L) { delete L; }

// validation + kernel parameters of one film-grain application (allocates the scratch on first use)
static int fg_fill(Eav1cuCtx *ctx, int src_pic, int dst_pic, const Eav1GrainParams *g, FgLaunch &L)
{
    const int np = (int)ctx->pics.size();
    if (src_pic < 0 || src_pic >= np || dst_pic < 0 || dst_pic >= np || src_pic == dst_pic ||
        !ctx->pics[src_pic].base || !ctx->pics[dst_pic].base)
        return fail(ctx, EAV1CU_ERR_INVAL, "apply_grain: bad picture id");
    const Picture &src = ctx->pics[src_pic], &dst = ctx->pics[dst_pic];
    if (src.w != dst.w || src.h != dst.h || src.layout != dst.layout || src.bpc != dst.bpc)
        return fail(ctx, EAV1CU_ERR_INVAL, "apply_grain: geometry mismatch");
    if (!ctx->have_tab) return fail(ctx, EAV1CU_ERR_INVAL, "apply_grain: tables not set");
    if (g->ar_coeff_lag < 0 || g->ar_coeff_lag > 3 || g->num_y_points > 14 || g->num_uv_points[0] > 10 ||
        g->num_uv_points[1] > 10)
        return fail(ctx, EAV1CU_ERR_INVAL, "apply_grain: bad parameters");
    coverage_count_grain(src.bpc, g);
    CK(cudaSetDevice(ctx->device));
    if (!ctx->d_fg) CK(cudaMalloc(&ctx->d_fg, FG_GRAIN_BYTES + FG_SCALE_BYTES + FG_OFFSET_BYTES));
    FgParams &P = L.P;
    memset(&P, 0, sizeof(P));
    P.g = *g;
    for (int i = 0; i < 3; i++) { P.src[i] = src.pl[i]; P.dst[i] = dst.pl[i]; }
    P.w = src.w; P.h = src.h; P.layout = src.layout;
    P.ss_hor = src.layout != 3; P.ss_ver = src.layout == 1;
    P.bitdepth_max = (1 << src.bpc) - 1;
    P.rows = (src.h + 31) / 32; P.cols = (src.w + 31) / 32;
    if ((size_t)P.rows * P.cols > FG_OFFSET_BYTES) return fail(ctx, EAV1CU_ERR_INVAL, "apply_grain: picture too large");
    P.tab = ctx->d_tab;
    P.grain = (int16_t *)ctx->d_fg;
    P.scaling = ctx->d_fg + FG_GRAIN_BYTES;
    P.offsets = ctx->d_fg + FG_GRAIN_BYTES + FG_SCALE_BYTES;
    L.bpc = src.bpc;
    L.grid = dim3((P.cols + FG_STRIP - 1) / FG_STRIP, P.rows);
    return 0;
}
static int fg_side_stream(Eav1cuCtx *ctx)
{
    if (ctx->ev_fg_ready) return 0;
    // Highest priority: the template build is one CTA of dependent steps beside kernels of thousands of CTAs. At
    // equal priority its CTA queued behind every pending CTA of the kernels launched before it and the build ran
    // after the frame instead of beside it.
    int prio_lo = 0, prio_hi = 0;
    cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi);
    if (cudaStreamCreateWithPriority(&ctx->stream_fg, cudaStreamNonBlocking, prio_hi) != cudaSuccess ||
        cudaEventCreateWithFlags(&ctx->ev_fg_ready, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&ctx->ev_fg_free, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&ctx->ev_fg_cap[0], cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&ctx->ev_fg_cap[1], cudaEventDisableTiming) != cudaSuccess)
        return fail(ctx, EAV1CU_ERR_CUDA, "film grain stream / events");
    return 0;
}
static cudaError_t fg_launch_prepare(const FgLaunch &L, cudaStream_t st)
{
    fg_prepare_kernel<<<1, 256, 0, st>>>(L.P);
    return cudaGetLastError();
}
static cudaError_t fg_launch_apply(const FgLaunch &L, cudaStream_t st)
{
    if (L.bpc > 8) fg_apply_kernel<uint16_t><<<L.grid, 256, 0, st>>>(L.P);
    else fg_apply_kernel<uint8_t><<<L.grid, 256, 0, st>>>(L.P);
    return cudaGetLastError();
}

extern "C" {

int eav1cu_apply_grain(Eav1cuCtx *ctx, int src_pic, int dst_pic, const Eav1GrainParams *g)
{
    if (!ctx || !g) return EAV1CU_ERR_INVAL;
    if (ctx->err) return ctx->err;
    FgLaunch L;
    if (int r = fg_fill(ctx, src_pic, dst_pic, g, L)) return r;
    std::lock_guard<SubmitMutex> gs(g_submit_mu);
    StageTimer tm(ctx, 5);
    // The templates, LUTs and offsets depend on the parameters only (one CTA, ~0.1 ms of dependent steps): they are
    // built on the second stream while the kernels of the frame itself still run on the main one; the apply kernel
    // waits for them, and the next call's build waits until this apply has finished with the shared scratch.
    // A stream of its own, not the one the residual kernels use: there the 0.2 ms template build of frame N sat in
    // front of the residuals of frame N+1 and, through them, in front of its reconstruction.
    const bool side = ctx->stream2 && !ctx->serial_phase1 && !ctx->stage_timing;
    if (side) if (int r = fg_side_stream(ctx)) return r;
    cudaStream_t ps = side ? ctx->stream_fg : ctx->stream;
    if (side && ctx->fg_busy) CK(cudaStreamWaitEvent(ps, ctx->ev_fg_free, 0));
    CK(fg_launch_prepare(L, ps));
    if (side) { CK(cudaEventRecord(ctx->ev_fg_ready, ps)); CK(cudaStreamWaitEvent(ctx->stream, ctx->ev_fg_ready, 0)); }
    CK(fg_launch_apply(L, ctx->stream));
    if (side) { CK(cudaEventRecord(ctx->ev_fg_free, ctx->stream)); ctx->fg_busy = true; }
    ctx->n_launches += 2;
    return 0;
}

} // extern "C"

// ---- film grain without any allocation (eav1cu_apply_grain_noalloc) -----------------------------
// Same kernels as eav1cu_apply_grain; the scratch they need (grain templates, scaling LUTs, block
// offsets) is this file-scope device array instead of a cudaMalloc'd buffer. Every GPU that loads the
// library holds these bytes until the process exits. In host code `g_fg_static` is only a stand-in for
// the symbol, not a device address: the address each device gave it is looked up once per device with
// cudaGetSymbolAddress (that device current) and it is that pointer the kernels receive.
enum { FG_STATIC_OFFSET_BYTES = EAV1CU_NOALLOC_MAX_GRAIN_BLOCKS,
       FG_STATIC_BYTES = FG_GRAIN_BYTES + FG_SCALE_BYTES + FG_STATIC_OFFSET_BYTES };
static __device__ __align__(256) unsigned char g_fg_static[FG_STATIC_BYTES];

enum { FG_STATIC_MAX_DEVICES = 64 };
static std::mutex g_fg_static_mu;
static uint8_t *g_fg_static_ptr[FG_STATIC_MAX_DEVICES];

// device address of g_fg_static on `device` (which must be current); looked up on first use per device
static cudaError_t fg_static_scratch(int device, uint8_t **out)
{
    std::lock_guard<std::mutex> g(g_fg_static_mu);
    if (!g_fg_static_ptr[device]) {
        void *p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_fg_static);
        if (e != cudaSuccess) return e;
        if (!p) return cudaErrorInvalidSymbol;
        g_fg_static_ptr[device] = (uint8_t *)p;
    }
    *out = g_fg_static_ptr[device];
    return cudaSuccess;
}

extern "C" {

int eav1cu_apply_grain_noalloc(Eav1cuCtx *ctx, int src_pic, int dst_pic, const Eav1GrainParams *g)
{
    if (!ctx || !g) return EAV1CU_ERR_INVAL;
    if (ctx->err) return ctx->err;
    const int np = (int)ctx->pics.size();
    if (src_pic < 0 || src_pic >= np || dst_pic < 0 || dst_pic >= np || src_pic == dst_pic ||
        !ctx->pics[src_pic].base || !ctx->pics[dst_pic].base)
        return fail(ctx, EAV1CU_ERR_INVAL, "apply_grain_noalloc: bad picture id");
    const Picture &src = ctx->pics[src_pic], &dst = ctx->pics[dst_pic];
    if (src.w != dst.w || src.h != dst.h || src.layout != dst.layout || src.bpc != dst.bpc)
        return fail(ctx, EAV1CU_ERR_INVAL, "apply_grain_noalloc: geometry mismatch");
    if (!ctx->have_tab) return fail(ctx, EAV1CU_ERR_INVAL, "apply_grain_noalloc: tables not set");
    if (g->ar_coeff_lag < 0 || g->ar_coeff_lag > 3 || g->num_y_points > 14 || g->num_uv_points[0] > 10 ||
        g->num_uv_points[1] > 10)
        return fail(ctx, EAV1CU_ERR_INVAL, "apply_grain_noalloc: bad parameters");
    if (ctx->device < 0 || ctx->device >= FG_STATIC_MAX_DEVICES)
        return fail(ctx, EAV1CU_ERR_INVAL, "apply_grain_noalloc: device index beyond the per-device address table");
    FgParams P;
    memset(&P, 0, sizeof(P));
    P.rows = (src.h + 31) / 32; P.cols = (src.w + 31) / 32;
    if ((size_t)P.rows * P.cols > (size_t)FG_STATIC_OFFSET_BYTES)
        return fail(ctx, EAV1CU_ERR_INVAL,
                    "apply_grain_noalloc: picture has more than EAV1CU_NOALLOC_MAX_GRAIN_BLOCKS 32x32 luma blocks "
                    "(static scratch limit); use eav1cu_apply_grain");
    CK(cudaSetDevice(ctx->device));
    uint8_t *scratch = nullptr;
    const cudaError_t se = fg_static_scratch(ctx->device, &scratch);
    if (se != cudaSuccess) return fail(ctx, EAV1CU_ERR_CUDA, "cudaGetSymbolAddress(g_fg_static)", se);
    P.g = *g;
    for (int i = 0; i < 3; i++) { P.src[i] = src.pl[i]; P.dst[i] = dst.pl[i]; }
    P.w = src.w; P.h = src.h; P.layout = src.layout;
    P.ss_hor = src.layout != 3; P.ss_ver = src.layout == 1;
    P.bitdepth_max = (1 << src.bpc) - 1;
    P.tab = ctx->d_tab;
    // fg_prepare_kernel rewrites every byte the apply kernel reads (all three templates, the LUT range
    // of this bit depth, rows x cols offsets), so nothing of the previous call's contents has to be cleared.
    P.grain = (int16_t *)scratch;
    P.scaling = scratch + FG_GRAIN_BYTES;
    P.offsets = scratch + FG_GRAIN_BYTES + FG_SCALE_BYTES;
    fg_prepare_kernel<<<1, 256, 0, ctx->stream>>>(P);
    const dim3 grid((P.cols + FG_STRIP - 1) / FG_STRIP, P.rows);
    if (src.bpc > 8) fg_apply_kernel<uint16_t><<<grid, 256, 0, ctx->stream>>>(P);
    else fg_apply_kernel<uint8_t><<<grid, 256, 0, ctx->stream>>>(P);
    ctx->n_launches += 2;
    CK(cudaGetLastError());
    return 0;
}

int eav1cu_set_stream(Eav1cuCtx *ctx, void *cuda_stream)
{
    if (!ctx) return EAV1CU_ERR_INVAL;
    CK(cudaSetDevice(ctx->device));
    CK(cudaStreamSynchronize(ctx->stream));
    if (ctx->own_stream) cudaStreamDestroy(ctx->stream);
    ctx->stream = (cudaStream_t)cuda_stream;
    ctx->own_stream = false;
    return 0;
}

} // extern "C"
