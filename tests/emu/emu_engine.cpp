// tests/emu/emu_engine.cpp — DEBUGGING AID, TEST INFRASTRUCTURE ONLY.
//
// Compiles the device code of easyav1_b200/csrc/*.cuh for the host (one "thread" per CTA,
// commands executed strictly in recorded order) behind the same eav1cu_* C ABI, so the
// arithmetic of the kernels and the host-side recorder can be diffed against the reference
// decoder in the build container, which has no GPU. It is never built into, loaded by or
// shipped with the product: easyav1_b200 only ever loads libeasyav1_cuda.so and fails
// loudly without a CUDA device.
#include <stdio.h>
#include <time.h>
#include <stdlib.h>
#include <string.h>
#include <algorithm>
#include <vector>
#include <string>

#include "../../include/easyav1_cuda.h"
#include "../../easyav1_b200/csrc/dev_common.cuh"
#include "../../easyav1_b200/csrc/exec.cuh"
#include "../../easyav1_b200/csrc/sched.h"
#include "../../easyav1_b200/csrc/postfilter.cuh"
#include "../../easyav1_b200/csrc/filmgrain.cuh"
#include "../../easyav1_b200/csrc/convert.cuh"
#include "../../easyav1_b200/csrc/coverage.h"

using namespace eav1;

struct Picture {
    bool used = false;
    int w = 0, h = 0, layout = 0, bpc = 0;
    std::vector<uint8_t> mem;
    Plane pl[3];
};

struct Eav1cuCtx {
    Eav1Tables tab;
    bool have_tab = false;
    std::vector<uint8_t> masks;
    std::vector<Picture> pics;
    std::vector<uint8_t> arena;
    Picture work[4];
    long long launches = 0;
    int err = 0;
};

static void pic_setup(Picture &p, int w, int h, int layout, int bpc)
{
    const int hbd = bpc > 8;
    const int ss_hor = layout != 3, ss_ver = layout == 1;
    const int aw = (w + 127) & ~127, ah = (h + 127) & ~127;
    const int cw = layout ? aw >> ss_hor : 0, chh = layout ? ah >> ss_ver : 0;
    const size_t ystride = (size_t)aw << hbd, cstride = (size_t)cw << hbd;
    const size_t ybytes = ystride * ah, cbytes = cstride * chh;
    p.mem.assign(ybytes + 2 * cbytes + 256, 0);
    p.w = w; p.h = h; p.layout = layout; p.bpc = bpc;
    uint8_t *b = p.mem.data();
    p.pl[0] = { b, (int)ystride, aw, ah };
    p.pl[1] = { b + ybytes, (int)cstride, cw, chh };
    p.pl[2] = { b + ybytes + cbytes, (int)cstride, cw, chh };
}

#include "emu_postfilter.inc"

#include <execinfo.h>
#include <signal.h>
#include <unistd.h>
// EAV1_EMU_BT=1: print a native backtrace on SIGSEGV (there is no debugger in the build container)
#include <ucontext.h>
static void segv_bt(int, siginfo_t *si, void *uc_)
{
    ucontext_t *uc = (ucontext_t *)uc_;
    void *fr[48];
    int n = backtrace(fr, 48);
    backtrace_symbols_fd(fr, n, 2);
    // a call through a bad pointer leaves no frame: the caller is on top of the stack
    void **sp = (void **)uc->uc_mcontext.gregs[REG_RSP];
    void *cand[8];
    for (int i = 0; i < 8; i++) cand[i] = sp[i];
    fprintf(stderr, "fault addr %p rip %p; stack words:\n", si->si_addr, (void *)uc->uc_mcontext.gregs[REG_RIP]);
    backtrace_symbols_fd(cand, 8, 2);
    _exit(139);
}

extern "C" {

int eav1cu_device_count(void) { return 1; }
int eav1cu_ctx_create(Eav1cuCtx **out, int) { if (getenv("EAV1_EMU_BT")) { struct sigaction sa; memset(&sa, 0, sizeof(sa)); sa.sa_sigaction = segv_bt; sa.sa_flags = SA_SIGINFO; sigaction(SIGSEGV, &sa, NULL); } *out = new Eav1cuCtx(); (*out)->pics.reserve(512); return 0; }
void eav1cu_ctx_destroy(Eav1cuCtx *c) { delete c; }
int eav1cu_last_error(Eav1cuCtx *c, char *msg, size_t cap) { if (msg && cap) msg[0] = 0; return c->err; }
int eav1cu_clear_error(Eav1cuCtx *c) { if (c->err == EAV1CU_ERR_INVAL) c->err = 0; return c->err; }
int eav1cu_set_tables(Eav1cuCtx *c, const Eav1Tables *t) { c->tab = *t; c->have_tab = true; return 0; }
int eav1cu_set_masks(Eav1cuCtx *c, const uint8_t *b, size_t n) { c->masks.assign(b, b + n); return 0; }
int eav1cu_sync(Eav1cuCtx *c) { return c->err; }
int eav1cu_host_alloc(void **p, size_t n) { *p = malloc(n); return *p ? 0 : EAV1CU_ERR_NOMEM; }
void eav1cu_host_free(void *p) { free(p); }
long long eav1cu_launch_count(Eav1cuCtx *c) { return c->launches; }

int eav1cu_pic_alloc(Eav1cuCtx *c, int w, int h, int layout, int bpc)
{
    int id = -1;
    for (size_t i = 0; i < c->pics.size(); i++) if (!c->pics[i].used) { id = (int)i; break; }
    if (id < 0) { c->pics.emplace_back(); id = (int)c->pics.size() - 1; }
    pic_setup(c->pics[id], w, h, layout, bpc);
    c->pics[id].used = true;
    return id;
}
int eav1cu_pic_release(Eav1cuCtx *c, int id) { c->pics[id].used = false; return 0; }

static int pic_copy(Eav1cuCtx *c, int id, void *const data[3], const ptrdiff_t stride[2], bool up)
{
    static const bool null_engine = getenv("EAV1_EMU_NULL") && atoi(getenv("EAV1_EMU_NULL")) > 1;
    if (null_engine) return 0;   // host-cost profiling, level 2: no picture traffic either
    Picture &p = c->pics[id];
    const int hbd = p.bpc > 8, ss_hor = p.layout != 3, ss_ver = p.layout == 1;
    for (int pl = 0; pl < (p.layout ? 3 : 1); pl++) {
        const int w = pl ? (p.w + ss_hor) >> ss_hor : p.w, h = pl ? (p.h + ss_ver) >> ss_ver : p.h;
        for (int y = 0; y < h; y++) {
            uint8_t *d = p.pl[pl].p + (size_t)y * p.pl[pl].stride;
            uint8_t *s = (uint8_t *)data[pl] + (ptrdiff_t)y * stride[!!pl];
            if (up) memcpy(d, s, (size_t)w << hbd); else memcpy(s, d, (size_t)w << hbd);
        }
    }
    return 0;
}
int eav1cu_pic_upload(Eav1cuCtx *c, int id, const void *const data[3], const ptrdiff_t stride[2]) { return pic_copy(c, id, (void *const *)data, stride, true); }
int eav1cu_pic_download(Eav1cuCtx *c, int id, void *const data[3], const ptrdiff_t stride[2]) { return pic_copy(c, id, data, stride, false); }
int eav1cu_arena_upload(Eav1cuCtx *c, size_t off, const void *src, size_t n) { if (c->arena.size() < off + n) c->arena.resize(off + n); memcpy(c->arena.data() + off, src, n); return 0; }
int eav1cu_arena_download(Eav1cuCtx *c, size_t off, void *dst, size_t n) { if (c->arena.size() < off + n) return EAV1CU_ERR_INVAL; memcpy(dst, c->arena.data() + off, n); return 0; }

int eav1cu_pic_download_async(Eav1cuCtx *c, int id, void *const data[3], const ptrdiff_t stride[2])
{ return eav1cu_pic_download(c, id, data, stride); }

int eav1cu_pic_convert(Eav1cuCtx *c, int id, const Eav1Convert *cv, void *dst, size_t dst_stride, size_t dst_bytes)
{
    if (!c || !cv || !dst || id < 0 || id >= (int)c->pics.size()) return EAV1CU_ERR_INVAL;
    const Picture &p = c->pics[id];
    const size_t need = conv_needed(cv->format, p.layout, p.bpc, p.w, p.h, dst_stride);
    if (!need || need > dst_bytes) return EAV1CU_ERR_INVAL;
    ConvParams P;
    memset(&P, 0, sizeof(P));
    for (int i = 0; i < 3; i++) P.src[i] = p.pl[i];
    P.w = p.w; P.h = p.h; P.mono = p.layout == 0; P.ss_hor = p.layout != 3; P.ss_ver = p.layout == 1;
    P.bpc = p.bpc; P.format = cv->format;
    conv_coefficients(P, cv->matrix, cv->full_range);
    P.dst = (uint8_t *)dst; P.dst_stride = (long long)dst_stride;
    for (int y = 0; y < p.h; y++)
        for (int x = 0; x < p.w; x++) {
            if (p.bpc > 8) convert_sample<uint16_t>(P, x, y); else convert_sample<uint8_t>(P, x, y);
        }
    return 0;
}
int eav1cu_fence(Eav1cuCtx *, unsigned long long *ticket) { static unsigned long long n; *ticket = ++n; return 0; }
int eav1cu_fence_wait(Eav1cuCtx *, unsigned long long) { return 0; }

int eav1cu_frame_submit(Eav1cuCtx *c, const Eav1FrameDesc *fd, const Eav1Cmd *cmds, size_t n,
                        const void *payload, size_t, size_t arena_bytes)
{
    if (!c->have_tab) return EAV1CU_ERR_NOTABLES;
    coverage_count(fd, cmds, n);
    if (getenv("EAV1_EMU_NULL")) {  // host-cost profiling: record only
        static size_t tot_cmds, tot_lf, tot_lr, frames;
        tot_cmds += n; tot_lf += fd->n_lf_recs; tot_lr += fd->n_lr_recs; frames++;
        if (getenv("EAV1_EMU_STATS")) {
            size_t hist[16] = {0}, px[16] = {0};
            for (size_t i = 0; i < n; i++) { hist[cmds[i].op & 15]++; px[cmds[i].op & 15] += (size_t)cmds[i].w * cmds[i].h; }
            fprintf(stderr, "frame %zu cmds %zu lf %zu lr %zu cdef %d |", frames, n, (size_t)fd->n_lf_recs, (size_t)fd->n_lr_recs, fd->cdef_enable);
            static const char *nm[] = {"nop","mc","warp","avg","wavg","mask","wmask","blend","blendv","blendh","ipred","cfl","pal","itx"};
            for (int o = 1; o < 14; o++) if (hist[o]) fprintf(stderr, " %s %zu/%zuk", nm[o], hist[o], px[o] >> 10);
            fprintf(stderr, "\n");
        }
        if (getenv("EAV1_EMU_PLAN")) {   // planner analysis: critical path of the dataflow graph in commands
            static eav1::Scheduler sch; static eav1::UnitGraph g;
            const Picture &cp = c->pics[fd->cur];
            int pw[3] = { cp.pl[0].w, cp.pl[1].w, cp.pl[2].w }, ph[3] = { cp.pl[0].h, cp.pl[1].h, cp.pl[2].h };
            struct timespec ta, tb; clock_gettime(CLOCK_MONOTONIC, &ta);
            sch.reset(pw, ph);
            sch.build_units(cmds, n, g);
            clock_gettime(CLOCK_MONOTONIC, &tb);
            static double tot_ms; tot_ms += 1e3 * (tb.tv_sec - ta.tv_sec) + 1e-6 * (tb.tv_nsec - ta.tv_nsec);
            fprintf(stderr, "planner: frame %zu build_units %.3f ms (running mean %.3f ms)\n", frames,
                    1e3 * (tb.tv_sec - ta.tv_sec) + 1e-6 * (tb.tv_nsec - ta.tv_nsec), tot_ms / frames);
            const size_t nu = g.unit_cmd.size();
            std::vector<uint32_t> cp_cmds(nu, 0), cp_units(nu, 0), pred(nu, UINT32_MAX);
            size_t best = 0; size_t tot = 0, hist[8] = {0};
            for (size_t u = 0; u < nu; u++) {
                uint32_t m = 0, mu = 0, pr = UINT32_MAX;
                for (uint32_t d = g.dep_start[u]; d < g.dep_start[u + 1]; d++)
                    if (cp_cmds[g.deps[d]] > m) { m = cp_cmds[g.deps[d]]; mu = cp_units[g.deps[d]]; pr = g.deps[d]; }
                cp_cmds[u] = m + g.unit_cnt[u]; cp_units[u] = mu + 1; pred[u] = pr;
                if (cp_cmds[u] > cp_cmds[best]) best = u;
                tot += g.unit_cnt[u];
                int b = 0; while ((1u << (b + 1)) <= g.unit_cnt[u] && b < 7) b++;
                hist[b]++;
            }
            fprintf(stderr, "plan frame %zu: units %zu cmds-in-units %zu mc-items %zu | critical path %u cmds over %u units | unit size hist(log2):",
                    frames, nu, tot, g.mc_items.size(), nu ? cp_cmds[best] : 0, nu ? cp_units[best] : 0);
            for (int b = 0; b < 8; b++) fprintf(stderr, " %zu", hist[b]);
            fprintf(stderr, "\n");
            if (atoi(getenv("EAV1_EMU_PLAN")) > 1 && nu) {
                for (uint32_t u = (uint32_t)best; u != UINT32_MAX; u = pred[u]) {
                    const Eav1Cmd &k = cmds[g.unit_cmd[u]];
                    fprintf(stderr, "   unit %u: %u cmds, first op %d plane %d %dx%d at (%d,%d)\n", u, g.unit_cnt[u], k.op, k.plane, k.w, k.h, k.x, k.y);
                }
            }
        }
        return 0;
    }
    Picture &cur = c->pics[fd->cur];
    const bool has_cdef = fd->cdef_enable != 0, has_lr = fd->n_lr_recs > 0;
    Picture *A = &cur;
    if (fd->superres) {
        const Picture &sr = c->pics[fd->sr_pic];
        if (has_cdef) pic_setup(c->work[1], cur.w, cur.h, cur.layout, cur.bpc);
        if (has_lr) { pic_setup(c->work[2], sr.w, sr.h, sr.layout, sr.bpc); if (has_cdef) pic_setup(c->work[3], sr.w, sr.h, sr.layout, sr.bpc); }
    } else if (has_cdef || has_lr) {
        pic_setup(c->work[0], cur.w, cur.h, cur.layout, cur.bpc);
        if (has_cdef && has_lr) pic_setup(c->work[1], cur.w, cur.h, cur.layout, cur.bpc);
        A = &c->work[0];
    }
    if (!n && A != &cur) memcpy(A->mem.data(), cur.mem.data(), std::min(A->mem.size(), cur.mem.size()));   // post-filter call on its own
    FrameCtx fc;
    memset(&fc, 0, sizeof(fc));
    for (int i = 0; i < 3; i++) fc.cur[i] = A->pl[i];
    for (int r = 0; r < 7; r++) {
        const Picture &rp = fd->ref[r] >= 0 ? c->pics[fd->ref[r]] : *A;
        for (int i = 0; i < 3; i++) fc.ref[r][i] = rp.pl[i];
    }
    for (int i = 0; i < 3; i++) fc.ref[7][i] = A->pl[i];
    fc.bitdepth_max = (1 << fd->bpc) - 1;
    fc.ss_hor = fd->layout != 3; fc.ss_ver = fd->layout == 1;
    fc.tab = &c->tab;
    fc.masks = c->masks.data();
    // same two-step split as the engine's planner: residual blocks behind the arena (phase 1b),
    // the ITX command itself then only adds (EAV1_F_RESID)
    static thread_local std::vector<Eav1Cmd> patched;
    patched.assign(cmds, cmds + n);
    size_t off = (arena_bytes + 31) & ~(size_t)31;
    for (size_t i = 0; i < n; i++) {
        Eav1Cmd &k = patched[i];
        if (k.op != EAV1_OP_ITX || (k.flags & EAV1_F_DST_ARENA) || k.u.itx.tx >= 19 || k.u.itx.txtp >= 16 ||
            k.u.itx.eob < (k.u.itx.txtp == 0))
            continue;
        k.flags |= EAV1_F_RESID;
        k.dst_off = (uint32_t)off;
        off += (size_t)k.w * k.h * 2;
    }
    arena_bytes = off;
    cmds = patched.data();
    if (c->arena.size() < arena_bytes + 16) c->arena.resize(arena_bytes + 16);
    fc.arena = c->arena.data();
    fc.payload = (const uint8_t *)payload;
    static thread_local CtaScratch sm;
    for (size_t i = 0; i < n; i++)
        if (cmds[i].op == EAV1_OP_ITX && (cmds[i].flags & EAV1_F_RESID)) {
            if (fd->bpc > 8) exec_itx_residual<BD16>(fc, cmds[i], sm, 0, 1);
            else exec_itx_residual<BD8>(fc, cmds[i], sm, 0, 1);
        }
    for (size_t i = 0; i < n; i++) {
        if (fd->bpc > 8) exec_cmd<BD16>(fc, cmds[i], sm, 0, 1);
        else exec_cmd<BD8>(fc, cmds[i], sm, 0, 1);
    }
    c->launches++;
    return emu_postfilter(c, fd, *A, cur);
}

int eav1cu_coverage(Eav1Coverage *out, int reset) { return coverage_read(out, reset); }
size_t eav1cu_abi_sizeof(int which) { return which == 6 ? sizeof(Eav1Coverage) : which == 3 ? sizeof(Eav1FrameDesc) : 0; }

int eav1cu_apply_grain(Eav1cuCtx *c, int src_pic, int dst_pic, const Eav1GrainParams *g)
{
    const Picture &src = c->pics[src_pic], &dst = c->pics[dst_pic];
    coverage_count_grain(src.bpc, g);
    static thread_local std::vector<uint8_t> buf(36864 + 3 * 4096 + 65536);
    static thread_local FgSmem sm;
    FgParams P;
    memset(&P, 0, sizeof(P));
    P.g = *g;
    for (int i = 0; i < 3; i++) { P.src[i] = src.pl[i]; P.dst[i] = dst.pl[i]; }
    P.w = src.w; P.h = src.h; P.layout = src.layout;
    P.ss_hor = src.layout != 3; P.ss_ver = src.layout == 1;
    P.bitdepth_max = (1 << src.bpc) - 1;
    P.rows = (src.h + 31) / 32; P.cols = (src.w + 31) / 32;
    P.tab = &c->tab;
    P.grain = (int16_t *)buf.data();
    P.scaling = buf.data() + 36864;
    P.offsets = buf.data() + 36864 + 3 * 4096;
    if (getenv("EAV1_EMU_STATS"))
        fprintf(stderr, "grain: seed %u y-points %d uv-points %d %d cfl %d lag %d overlap %d clip %d shift %d\n", g->seed, g->num_y_points,
                g->num_uv_points[0], g->num_uv_points[1], g->chroma_scaling_from_luma, g->ar_coeff_lag, g->overlap_flag,
                g->clip_to_restricted_range, g->scaling_shift);
    fg_prepare(P, sm, 0, 1);
    for (int row = 0; row < P.rows; row++)
        for (int s = 0; s < (P.cols + FG_STRIP - 1) / FG_STRIP; s++) {
            if (src.bpc > 8) fg_strip<uint16_t>(P, s, row, 0, 1);
            else fg_strip<uint8_t>(P, s, row, 0, 1);
        }
    c->launches += 2;
    return 0;
}

} // extern "C"

// ---- self checks of small device helpers, callable from the CPU tests ------------------------
extern "C" long long eav1emu_fastdiv_mismatches(int d_max, int i_max)
{
    long long bad = 0;
    for (int d = 1; d <= d_max; d++) {
        const eav1::FastDiv fd(d);
        for (int i = 0; i < i_max; i += (i < 4096 ? 1 : 61)) bad += fd.div(i) != i / d;
        bad += fd.div(i_max - 1) != (i_max - 1) / d;
    }
    return bad;
}
