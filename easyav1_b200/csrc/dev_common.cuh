// dev_common.cuh — shared device-side definitions for the sm_100a reconstruction kernels.
//
// Every per-command routine is written as a CTA-cooperative function
//     fn(..., int tid, int nthr)
// whose threads stride over independent output elements and meet at CTA_SYNC().
// (tests/emu builds the same code for the host with one "thread" per CTA, purely
// as a debugging aid for arithmetic; it is never part of the product.)
#pragma once
#include <stdint.h>
#include <stddef.h>
#include "../../include/easyav1_cuda.h"

#if defined(__CUDACC__)
#define EAV1_DEV __device__ __forceinline__
#define EAV1_DEVN __device__ __noinline__
// helpers on the per-item hot path of the flat kernels: inlined unless EAV1_HOT_NOINLINE is set
#if defined(EAV1_HOT_NOINLINE) && EAV1_HOT_NOINLINE
#define EAV1_DEVH __device__ __noinline__
#else
#define EAV1_DEVH __device__ __forceinline__
#endif
#define EAV1_DEVM __device__ __forceinline__
// group barrier: a work group is either one warp (nthr == 32) or the whole CTA; `nthr` must be in scope
#define CTA_SYNC() do { if (nthr <= 32) __syncwarp(); else __syncthreads(); } while (0)
#define EAV1_LDG(p) __ldg(p)
// Programmatic dependent launch (sm_90+): a kernel launched with the attribute may be scheduled while its predecessor
// in the stream drains. It may read what the host uploaded (work lists) but must not touch anything a kernel wrote
// before pdl_wait() returns (a no-op for a kernel launched the ordinary way). Every kernel of the frame pipeline calls
// it exactly once per thread. c_pdl_early != 0: the kernel then also releases its own successor, whose CTAs move into
// slots as they fall free and have their prologue behind them when the last CTA of this kernel leaves. Wait before
// release, always: a successor's prologue then implies that everything before this kernel has completed.
__constant__ int c_pdl_early;
static __device__ __forceinline__ void pdl_wait()
{
    asm volatile("griddepcontrol.wait;" ::: "memory");
    if (c_pdl_early) asm volatile("griddepcontrol.launch_dependents;" :::);
}
#define EAV1_GRID_DEP_WAIT() pdl_wait()
#else
#define EAV1_GRID_DEP_WAIT() ((void)0)
#define EAV1_DEV static inline
#define EAV1_DEVN static
#define EAV1_DEVH static inline
#define EAV1_DEVM inline
#define CTA_SYNC() ((void)0)
#define EAV1_LDG(p) (*(p))
struct alignas(8) uint2 { uint32_t x, y; };
struct alignas(16) uint4 { uint32_t x, y, z, w; };
#endif

// Development aid (-DEAV1_PROF=1 builds only): cycle accounting between marks inside one command.
#if defined(__CUDACC__) && defined(EAV1_PROF) && EAV1_PROF
__device__ unsigned long long g_prof[64];
__device__ __forceinline__ void prof_mark(long long &t, int slot, int tid) {
    if (tid == 0) { const long long n = clock64(); atomicAdd(&g_prof[2 * slot], (unsigned long long)(n - t)); atomicAdd(&g_prof[2 * slot + 1], 1ull); t = n; }
}
#define PROF_BEGIN() long long prof_t = clock64()
#define PROF_MARK(slot) prof_mark(prof_t, slot, tid)
#else
#define PROF_BEGIN() ((void)0)
#define PROF_MARK(slot) ((void)0)
#endif

namespace eav1 {

struct Plane {
    uint8_t *p;   // base pointer (bytes)
    int stride;   // bytes
    int w, h;     // allocated (padded) size in pixels
};

struct FrameCtx {
    Plane cur[3];
    Plane ref[8][3];          // [7] aliases cur (intrabc)
    int bitdepth_max;
    int ss_hor, ss_ver;
    uint8_t *arena;           // per-frame scratch (prep blocks, lap, masks, ...)
    const uint8_t *payload;   // per-frame coefficients / palettes
    const Eav1Tables *tab;
    const uint8_t *masks;     // wedge / inter-intra mask blob
};

// Per-CTA shared scratch: the 64x64 int32 transform tile (odd pitch), or the MC
// intermediate rows, or the intra edge/ac buffers -- never live at the same time.
// A single warp working on commands of at most 32x32 pixels only touches the first
// EAV1_WARP_SCRATCH bytes (32x33 int32 tile, (32+7)x32 int16 rows, the whole `ip` part), so four
// warps of a CTA can each own such a slice.
#define EAV1_WARP_SCRATCH 4224
struct alignas(16) CtaScratch {
    union {
        int32_t itx[64 * 65];
        int16_t mc_mid[(32 + 7) * 128];
        struct {
            uint16_t edge[288];      // intra edge array, top-left sample at [EDGE_C]
            uint16_t edge2[288];     // filtered / upsampled edge
            int16_t  ac[32 * 32];    // cfl ac
            int32_t  red[256];       // CTA reduction scratch
        } ip;
    };
};

EAV1_DEV int iclip(int v, int lo, int hi) { return v < lo ? lo : v > hi ? hi : v; }
EAV1_DEV int imin(int a, int b) { return a < b ? a : b; }
EAV1_DEV int imax(int a, int b) { return a > b ? a : b; }
EAV1_DEV int iabs(int a) { return a < 0 ? -a : a; }
EAV1_DEV int apply_sign(int v, int s) { return s < 0 ? -v : v; }
EAV1_DEV int ctz32(unsigned v) {
#if defined(__CUDA_ARCH__)
    return __ffs(v) - 1;
#else
    return __builtin_ctz(v);
#endif
}
EAV1_DEV int clz32(unsigned v) {
#if defined(__CUDA_ARCH__)
    return __clz(v);
#else
    return __builtin_clz(v);
#endif
}
EAV1_DEV int ulog2(unsigned v) { return 31 - clz32(v); }

// i / d for many i and one small runtime divisor: one real division per loop (the magic number),
// a multiply-high per element. Exact for d <= 1024 and i < 2^22 (error term i * (magic*d - 2^32) < 2^32).
struct FastDiv {
    unsigned magic, d;
    EAV1_DEVM explicit FastDiv(int dd) : magic((unsigned)((0x100000000ull + (unsigned)dd - 1) / (unsigned)dd)), d((unsigned)dd) {}
    EAV1_DEVM int div(int i) const {
        if (d == 1) return i;
#if defined(__CUDA_ARCH__)
        return (int)__umulhi((unsigned)i, magic);
#else
        return (int)(((unsigned long long)(unsigned)i * magic) >> 32);
#endif
    }
};

// Bit-depth traits: pixel / coefficient types and the multiply width the inverse
// transforms need (see itx.cuh).
struct BD8 {
    typedef uint8_t pixel;
    typedef int16_t coef;
    typedef int32_t wide;
    static constexpr bool hbd = false;
};
struct BD16 {
    typedef uint16_t pixel;
    typedef int32_t coef;
    typedef int64_t wide;
    static constexpr bool hbd = true;
};

template <typename pixel>
EAV1_DEV pixel *px_at(const Plane &pl, int x, int y) {
    return reinterpret_cast<pixel *>(pl.p + (ptrdiff_t)y * pl.stride) + x;
}

} // namespace eav1
