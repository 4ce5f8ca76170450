// coverage.h — tool-coverage histogram over the frames handed to eav1cu_frame_submit (host code).
// Which (operator, transform size / type, prediction mode, filter, scaled?, bit depth) combinations the
// test streams actually execute; off unless EAV1_COVERAGE is set in the environment. Shared by the CUDA
// engine and the host emulation so that the CPU tests see the same numbers. tests/test_coverage.py pins
// the set of combinations the fixtures must keep reaching.
#pragma once
#include <stdlib.h>
#include <string.h>
#include <mutex>
#include "../../include/easyav1_cuda.h"

namespace eav1 {

struct CoverageState {
    std::mutex mu;
    Eav1Coverage c;
    bool on;
    CoverageState() : on(getenv("EAV1_COVERAGE") != nullptr) { memset(&c, 0, sizeof(c)); }
};

inline CoverageState &coverage_state() { static CoverageState s; return s; }

inline int cov_log2(int v) { int l = 0; while ((2 << l) <= v) l++; return l; }

inline void coverage_count(const Eav1FrameDesc *fd, const Eav1Cmd *cmds, size_t n)
{
    CoverageState &s = coverage_state();
    if (!s.on) return;
    std::lock_guard<std::mutex> g(s.mu);
    Eav1Coverage &c = s.c;
    const int bd = fd->bpc == 8 ? 0 : fd->bpc == 10 ? 1 : 2;
    c.frames[bd]++;
    c.layout[fd->layout & 3]++;
    for (size_t i = 0; i < n; i++) {
        const Eav1Cmd &k = cmds[i];
        const int lw = cov_log2(k.w) & 7, lh = cov_log2(k.h) & 7;
        switch (k.op) {
        case EAV1_OP_ITX:
            if (k.u.itx.tx < 19 && k.u.itx.txtp < 17) {
                c.itx[bd][k.u.itx.tx][k.u.itx.txtp]++;
                c.itx_eob[bd][k.u.itx.eob < (k.u.itx.txtp == 0) ? 0 : k.u.itx.eob < 10 ? 1 : 2]++;
            }
            break;
        case EAV1_OP_MC: {
            const int prep = (k.flags & EAV1_F_PREP) != 0, scaled = k.u.mc.dx != 0;
            const int sub = scaled ? 3 : (k.u.mc.mx ? 1 : 0) | (k.u.mc.my ? 2 : 0);
            c.mc[bd > 0][scaled][prep][k.u.mc.filt_h & 3][k.u.mc.filt_v & 3][sub]++;
            c.mc_wh[scaled][lw][lh]++;
            if (k.u.mc.ref == 7) c.intrabc++;
            break;
        }
        case EAV1_OP_WARP: c.warp[bd > 0][(k.flags & EAV1_F_PREP) != 0]++; break;
        case EAV1_OP_AVG: c.comp[bd > 0][0]++; break;
        case EAV1_OP_W_AVG: c.comp[bd > 0][1]++; break;
        case EAV1_OP_MASK: c.comp[bd > 0][2]++; break;
        case EAV1_OP_W_MASK: c.comp[bd > 0][3]++; c.w_mask_ss[k.u.comp.ss_hor + k.u.comp.ss_ver]++; break;
        case EAV1_OP_BLEND: c.blend[bd > 0][0]++; break;
        case EAV1_OP_BLEND_V: c.blend[bd > 0][1]++; break;
        case EAV1_OP_BLEND_H: c.blend[bd > 0][2]++; break;
        case EAV1_OP_IPRED:
            if (k.u.ipred.mode < 14) { c.ipred[bd][k.u.ipred.mode]++; c.ipred_wh[lw][lh]++; }
            break;
        case EAV1_OP_CFL:
            if (k.u.ipred.mode < 14) c.cfl[bd > 0][k.u.ipred.mode]++;
            break;
        case EAV1_OP_PAL: c.pal[bd > 0]++; break;
        default: break;
        }
    }
    for (size_t i = 0; i < fd->n_lf_recs; i++) {
        const Eav1LfRec &r = fd->lf_recs[i];
        for (int sz = 0; sz < 3; sz++)
            if (r.vmask[sz]) c.lf[bd > 0][r.plane > 0][r.dir & 1][sz] += (uint64_t)__builtin_popcount(r.vmask[sz]);
    }
    for (size_t i = 0; i < fd->n_lr_recs; i++)
        if (fd->lr_recs[i].kind < 5) c.lr[bd > 0][fd->lr_recs[i].kind]++;
    if (fd->cdef_enable) c.cdef_frames[bd > 0]++;
    if (fd->superres) c.superres_frames[bd > 0]++;
}

inline void coverage_count_grain(int bpc, const Eav1GrainParams *g)
{
    CoverageState &s = coverage_state();
    if (!s.on) return;
    std::lock_guard<std::mutex> lk(s.mu);
    s.c.grain[bpc > 8][g->chroma_scaling_from_luma ? 1 : 0]++;
}

inline int coverage_read(Eav1Coverage *out, int reset)
{
    CoverageState &s = coverage_state();
    std::lock_guard<std::mutex> g(s.mu);
    if (out) *out = s.c;
    if (reset) memset(&s.c, 0, sizeof(s.c));
    return s.on ? 1 : 0;
}

} // namespace eav1
