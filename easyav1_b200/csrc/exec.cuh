// exec.cuh — executes one recorded reconstruction command (struct Eav1Cmd) on a CTA.
// One CTA owns one coding block ("chain") and runs its commands in host-walk order, so
// intra-block dependencies (prediction -> residual -> next transform block's edge, prep ->
// avg, w_mask -> chroma mask, mc -> OBMC blend) are ordinary program order + CTA_SYNC().
#pragma once
#include "dev_common.cuh"
#include "itx.cuh"
#include "mc.cuh"
#include "ipred.cuh"

namespace eav1 {

// Motion compensation of the sub-rectangle (x0,y0,w,h) of command c (whole block, or one
// 32x32 tile of it when the block is spread over several warps).
template <typename BD>
EAV1_DEV void exec_mc_rect(const FrameCtx &fc, const Eav1Cmd &c, int x0, int y0, int w, int h,
                           CtaScratch &sm, int tid, int nthr)
{
    typedef typename BD::pixel pixel;
    pixel *dst8 = nullptr;
    int16_t *dst16 = nullptr;
    int dstride;
    if (c.flags & EAV1_F_DST_ARENA) {
        dstride = c.dst_stride;
        if (c.flags & EAV1_F_PREP) dst16 = reinterpret_cast<int16_t *>(fc.arena + c.dst_off) + y0 * dstride + x0;
        else dst8 = reinterpret_cast<pixel *>(fc.arena + c.dst_off) + y0 * dstride + x0;
    } else {
        const Plane &pl = fc.cur[c.plane];
        dst8 = px_at<pixel>(pl, c.x + x0, c.y + y0);
        dstride = pl.stride / (int)sizeof(pixel);
    }
    const Plane &rp = fc.ref[c.u.mc.ref][c.plane];
    SrcView<pixel> src;
    src.p = rp.p; src.stride = rp.stride;
    src.iw = c.u.mc.iw; src.ih = c.u.mc.ih;
    if (c.u.mc.dx)   // scaled references are never tiled (x0 == y0 == 0)
        mc_scaled_block<BD>(fc.tab, src, c.u.mc.sx, c.u.mc.sy, w, h, c.u.mc.mx, c.u.mc.my,
                            c.u.mc.dx, c.u.mc.dy, c.u.mc.filt_h, c.u.mc.filt_v, dst8, dst16, dstride,
                            fc.bitdepth_max, tid, nthr);
    else
        mc_block<BD>(fc.tab, src, c.u.mc.sx + x0, c.u.mc.sy + y0, w, h, c.w, c.h, c.u.mc.mx, c.u.mc.my,
                     c.u.mc.filt_h, c.u.mc.filt_v, dst8, dst16, dstride, fc.bitdepth_max, sm.mc_mid, tid, nthr);
}

template <typename BD>
EAV1_DEV void exec_warp(const FrameCtx &fc, const Eav1Cmd &c, CtaScratch &sm, int tid, int nthr)
{
    typedef typename BD::pixel pixel;
    pixel *dst8 = nullptr;
    int16_t *dst16 = nullptr;
    int dstride;
    if (c.flags & EAV1_F_DST_ARENA) {
        dstride = c.dst_stride;
        if (c.flags & EAV1_F_PREP) dst16 = reinterpret_cast<int16_t *>(fc.arena + c.dst_off);
        else dst8 = reinterpret_cast<pixel *>(fc.arena + c.dst_off);
    } else {
        const Plane &pl = fc.cur[c.plane];
        dst8 = px_at<pixel>(pl, c.x, c.y);
        dstride = pl.stride / (int)sizeof(pixel);
    }
    const Plane &rp = fc.ref[c.u.warp.ref][c.plane];
    SrcView<pixel> src;
    src.p = rp.p; src.stride = rp.stride;
    src.iw = c.u.warp.iw; src.ih = c.u.warp.ih;
    warp8x8_block<BD>(fc.tab, src, c.u.warp.sx, c.u.warp.sy, c.u.warp.abcd, c.u.warp.mx, c.u.warp.my,
                      dst8, dst16, dstride, fc.bitdepth_max, sm.mc_mid, tid, nthr);
}

// Phase 1 of a two-step transform block: residual of ITX command c into the arena (c.dst_off).
template <typename BD>
EAV1_DEV void exec_itx_residual(const FrameCtx &fc, const Eav1Cmd &c, CtaScratch &sm, int tid, int nthr)
{
    itx_add<BD>(nullptr, 0, (const typename BD::coef *)(fc.payload + c.u.itx.cf_off), c.u.itx.eob, c.u.itx.tx,
                c.u.itx.txtp, fc.bitdepth_max, sm.itx, tid, nthr, reinterpret_cast<int16_t *>(fc.arena + c.dst_off), c.u.itx.rows);
}

// Inlined into the worker: as a separate function it saves and restores its callee-saved registers
// through local memory on every command, which costs a fifth of the dataflow kernel's time
// (EAV1_DEDUPE=1 builds keep one shared copy instead; measured 11.6 vs 9.3 ms per 120 frames).
#ifndef EAV1_DEDUPE
#define EAV1_DEDUPE 0
#endif
#if EAV1_DEDUPE
#define EAV1_EXEC_LINK EAV1_DEVN
#else
#define EAV1_EXEC_LINK EAV1_DEV
#endif
// Returns how many commands it consumed: an intra prediction directly followed by the residual
// add of the same transform block (`next`) finishes the block in one go, without a second trip
// through the dispatch.
template <typename BD>
EAV1_EXEC_LINK int exec_cmd(const FrameCtx &fc, const Eav1Cmd &c, CtaScratch &sm, int tid, int nthr,
                            const Eav1Cmd *next = nullptr)
{
    typedef typename BD::pixel pixel;
    const int bdmax = fc.bitdepth_max;
    PROF_BEGIN();
    // destination
    pixel *dst8 = nullptr;
    int16_t *dst16 = nullptr;
    int dstride;
    if (c.flags & EAV1_F_DST_ARENA) {
        dstride = c.dst_stride;
        if (c.flags & EAV1_F_PREP) dst16 = reinterpret_cast<int16_t *>(fc.arena + c.dst_off);
        else dst8 = reinterpret_cast<pixel *>(fc.arena + c.dst_off);
    } else {
        const Plane &pl = fc.cur[c.plane];
        dst8 = px_at<pixel>(pl, c.x, c.y);
        dstride = pl.stride / (int)sizeof(pixel);
    }
    PROF_MARK(0);
    switch (c.op) {
    case EAV1_OP_MC:
        exec_mc_rect<BD>(fc, c, 0, 0, c.w, c.h, sm, tid, nthr);
        break;
    case EAV1_OP_WARP:
        exec_warp<BD>(fc, c, sm, tid, nthr);
        break;
    case EAV1_OP_AVG:
        PROF_MARK(6);
        comp_avg<BD>(dst8, dstride, (const int16_t *)(fc.arena + c.u.comp.t1), (const int16_t *)(fc.arena + c.u.comp.t2),
                     c.w, c.h, bdmax, tid, nthr);
        PROF_MARK(7);
        break;
    case EAV1_OP_W_AVG:
        comp_w_avg<BD>(dst8, dstride, (const int16_t *)(fc.arena + c.u.comp.t1), (const int16_t *)(fc.arena + c.u.comp.t2),
                       c.w, c.h, c.u.comp.weight, bdmax, tid, nthr);
        break;
    case EAV1_OP_MASK: {
        const uint8_t *mask = (c.flags & EAV1_F_MASK_ARENA) ? fc.arena + c.u.comp.mask_off : fc.masks + c.u.comp.mask_off;
        comp_mask<BD>(dst8, dstride, (const int16_t *)(fc.arena + c.u.comp.t1), (const int16_t *)(fc.arena + c.u.comp.t2),
                      c.w, c.h, mask, bdmax, tid, nthr);
        break;
    }
    case EAV1_OP_W_MASK:
        comp_w_mask<BD>(dst8, dstride, (const int16_t *)(fc.arena + c.u.comp.t1), (const int16_t *)(fc.arena + c.u.comp.t2),
                        c.w, c.h, fc.arena + c.u.comp.mask_off, c.u.comp.sign, c.u.comp.ss_hor, c.u.comp.ss_ver,
                        bdmax, tid, nthr);
        break;
    case EAV1_OP_BLEND:
        blend_mask<pixel>(dst8, dstride, (const pixel *)(fc.arena + c.u.blend.tmp_off), c.w, c.h,
                          fc.masks + c.u.blend.mask_off, tid, nthr);
        break;
    case EAV1_OP_BLEND_V:
        blend_v<pixel>(fc.tab, dst8, dstride, (const pixel *)(fc.arena + c.u.blend.tmp_off), c.w, c.h, tid, nthr);
        break;
    case EAV1_OP_BLEND_H:
        blend_h<pixel>(fc.tab, dst8, dstride, (const pixel *)(fc.arena + c.u.blend.tmp_off), c.w, c.h, tid, nthr);
        break;
    case EAV1_OP_IPRED:
        PROF_MARK(1);
        {
            // the residual add of the same transform block follows: finish the block here
            const bool fused = next && next->op == EAV1_OP_ITX && (next->flags & EAV1_F_RESID) && next->plane == c.plane &&
                               next->x == c.x && next->y == c.y && next->w == c.w && next->h == c.h &&
                               !(c.flags & EAV1_F_DST_ARENA);
            const int16_t *res = fused ? (const int16_t *)(fc.arena + next->dst_off) : nullptr;
            ResPre res_pre = ResPre();
            res_pre.tid = tid; res_pre.nthr = nthr;
            if (fused && tid < ((c.w * c.h) >> 2)) res_pre.a = *reinterpret_cast<const uint2 *>(res + (tid << 2));
            if (fused && tid + nthr < ((c.w * c.h) >> 2)) res_pre.b = *reinterpret_cast<const uint2 *>(res + ((tid + nthr) << 2));
            build_edge<pixel>(fc.cur[c.plane], c, bdmax, sm.ip.edge, tid, nthr);
            PROF_MARK(2);
            const bool added = ipred_block<pixel>(fc.tab, dst8, dstride, sm.ip.edge, sm.ip.edge2, c.u.ipred.mode, c.w, c.h,
                                                  c.u.ipred.angle, c.u.ipred.max_w, c.u.ipred.max_h, bdmax, tid, nthr,
                                                  reinterpret_cast<uint16_t *>(sm.ip.ac), res, res_pre);
            if (fused && !added) { CTA_SYNC(); res_add<BD>(dst8, dstride, res, c.w, c.h, bdmax, tid, nthr); }
            PROF_MARK(3);
            if (fused) return 2;
        }
        break;
    case EAV1_OP_CFL:
        cfl_ac_block<pixel>(fc.cur[0], c.u.ipred.lx, c.u.ipred.ly, c.u.ipred.w_pad, c.u.ipred.h_pad, c.w, c.h,
                            fc.ss_hor, fc.ss_ver, sm.ip.ac, sm.ip.red, tid, nthr);
        build_edge<pixel>(fc.cur[c.plane], c, bdmax, sm.ip.edge, tid, nthr);
        cfl_pred_block<pixel>(dst8, dstride, sm.ip.edge, c.u.ipred.mode, c.w, c.h, sm.ip.ac, c.u.ipred.alpha,
                              bdmax, tid, nthr);
        break;
    case EAV1_OP_PAL:
        pal_pred_block<pixel>(dst8, dstride, (const pixel *)(fc.payload + c.u.pal.pal_off),
                              fc.payload + c.u.pal.idx_off, c.w, c.h, tid, nthr);
        break;
    case EAV1_OP_ITX:
        PROF_MARK(4);
        if (c.flags & EAV1_F_RESID)
            res_add<BD>(dst8, dstride, (const int16_t *)(fc.arena + c.dst_off), c.w, c.h, bdmax, tid, nthr);
        else
            itx_add<BD>(dst8, dstride, (const typename BD::coef *)(fc.payload + c.u.itx.cf_off), c.u.itx.eob,
                        c.u.itx.tx, c.u.itx.txtp, bdmax, sm.itx, tid, nthr, nullptr, c.u.itx.rows);
        PROF_MARK(5);
        break;
    default:
        break;
    }
    return 1;
}

} // namespace eav1
