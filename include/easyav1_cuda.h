/*
 * easyav1_cuda.h — C ABI of libeasyav1_cuda.so, the B200 (sm_100a) AV1 reconstruction engine.
 *
 * Plain pointers and sizes only (no torch / C++ types). This is the boundary "B5" of
 * SURVEY.md §8(b); it is what the dav1d host side (entropy decode + block walk, kept in C)
 * binds instead of dav1d's per-bitdepth DSP objects:
 *
 *   reference interface replaced                         entry point here
 *   ---------------------------------------------------  ---------------------------------
 *   Dav1dDSPContext tables, ext/dav1d/src/internal.h:62   eav1cu_frame_submit (recorded DSP
 *     itx.itxfm_add  (src/itx.h:37-72)                      commands, struct Eav1Cmd below:
 *     mc.*           (src/mc.h:38-162)                      one record per DSP call the host
 *     ipred.*        (src/ipred.h:43-90) + prepare_intra_   walk would have made)
 *       edges (src/ipred_prepare.h:80-92)
 *   f->bd_fn.filter_sbrow* (src/recon.h:57-69;            eav1cu_frame_submit post-filter
 *     lf_apply_tmpl.c:313-466, cdef_apply_tmpl.c:97,        stage (Eav1FrameDesc.lf/cdef/lr
 *     lr_apply_tmpl.c:168)                                  side data, whole frame)
 *   dav1d_apply_grain_{8,16}bpc (src/fg_apply.h)          eav1cu_apply_grain
 *   Dav1dPicAllocator (include/dav1d/picture.h:133-145)   eav1cu_pic_alloc / _release / _download
 *   mdct_backward + vorbis_synthesis_blockin              eav1cu_vorbis_open / _imdct_ola
 *     (ext/minivorbis/minivorbis.h:7054, 9195)
 *
 * Every function returns 0 on success or a negative EAV1CU_ERR_* code; CUDA errors are
 * latched in the context (eav1cu_last_error) because the DSP surface it replaces is void.
 * There is NO CPU fallback: without a CUDA device every entry point fails with
 * EAV1CU_ERR_NO_DEVICE.
 */
#ifndef EASYAV1_CUDA_H
#define EASYAV1_CUDA_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define EAV1CU_ERR_NO_DEVICE (-1)
#define EAV1CU_ERR_CUDA      (-2)
#define EAV1CU_ERR_INVAL     (-3)
#define EAV1CU_ERR_NOMEM     (-4)
#define EAV1CU_ERR_NOTABLES  (-5)

typedef struct Eav1cuCtx Eav1cuCtx;

/* ---- constant tables (values come from ext/dav1d/src/tables.c and wedge.c; the host
 * side uploads its own copies once, the engine never regenerates them) ---------------- */
typedef struct Eav1Tables {
    int8_t   mc_subpel_filters[6][15][8]; /* tables.c:445 */
    int8_t   mc_warp_filter[193][8];      /* tables.c:550 */
    int8_t   resize_filter[64][8];        /* tables.c:655 */
    uint8_t  sm_weights[128];             /* tables.c:690 */
    uint16_t dr_intra_derivative[44];     /* tables.c:719 */
    int8_t   filter_intra_taps[5][8][8];  /* tables.c:765, canonical [filter][pixel yy*4+xx][tap p0..p6] */
    uint8_t  obmc_masks[64];              /* tables.c:814 */
    int16_t  gaussian_sequence[2048];     /* tables.c:832 */
    uint8_t  sgr_x_by_x[256];             /* tables.c:423 */
    int8_t   cdef_directions[12][2];      /* tables.c:400, rows 2+dir-2 .. 2+dir+2, in units of a 12-px stride */
    uint16_t sgr_params[16][2];           /* tables.c:415 */
} Eav1Tables;

/* ---- recorded reconstruction commands ---------------------------------------------- */
enum Eav1Op {
    EAV1_OP_NOP = 0,
    EAV1_OP_MC,      /* mc.mc / mc.mct / mc_scaled / mct_scaled (emu_edge folded into clamping) */
    EAV1_OP_WARP,    /* mc.warp8x8 / warp8x8t */
    EAV1_OP_AVG,     /* mc.avg */
    EAV1_OP_W_AVG,   /* mc.w_avg */
    EAV1_OP_MASK,    /* mc.mask */
    EAV1_OP_W_MASK,  /* mc.w_mask[ss] */
    EAV1_OP_BLEND,   /* mc.blend */
    EAV1_OP_BLEND_V, /* mc.blend_v */
    EAV1_OP_BLEND_H, /* mc.blend_h */
    EAV1_OP_IPRED,   /* prepare_intra_edges + ipred.intra_pred[m] */
    EAV1_OP_CFL,     /* ipred.cfl_ac + prepare_intra_edges + ipred.cfl_pred[m] */
    EAV1_OP_PAL,     /* ipred.pal_pred */
    EAV1_OP_ITX,     /* itx.itxfm_add[tx][txtp] */
    EAV1_N_OPS
};

/* Eav1Cmd.flags */
#define EAV1_F_DST_ARENA  1  /* destination is the per-frame scratch arena (dst_off, dst_stride) */
#define EAV1_F_PREP       2  /* MC/WARP: int16 "prep" output instead of pixels */
#define EAV1_F_BLOCK_HEAD 4  /* first command of a coding block (= scheduling unit) */
#define EAV1_F_MASK_ARENA 8  /* MASK/BLEND: mask lives in the arena (w_mask output), else in the mask table blob */
#define EAV1_F_RESID      16 /* ITX, set by the engine's planner only: the residual was computed in phase 1 and lies
                                at arena offset dst_off (row-major int16, stride w); phase 2 just adds it */

/* mask_off for table masks indexes the blob passed to eav1cu_set_masks (wedge + ii masks). */

typedef struct Eav1Cmd {
    uint8_t  op, plane, w, h;   /* w,h in pixels, 1..128 */
    int16_t  x, y;              /* destination position in the plane (pixels) when !DST_ARENA */
    uint32_t dst_off;           /* arena byte offset when DST_ARENA */
    uint16_t dst_stride;        /* arena stride in elements when DST_ARENA */
    uint8_t  flags, pad0;
    union {
        struct {                /* EAV1_OP_MC */
            int32_t  sx, sy;    /* integer source position of the block's (0,0) sample (may be out of frame) */
            int32_t  iw, ih;    /* source plane clamp bounds (emu_edge iw, ih) */
            uint16_t mx, my;    /* 1/16 subpel (0..15), or 1/1024 fraction for scaled */
            uint16_t dx, dy;    /* scaled step (0 = unscaled) */
            uint8_t  ref;       /* 0..6 reference slot, 7 = current frame (intrabc) */
            uint8_t  filt_h, filt_v; /* 0 regular,1 smooth,2 sharp (8-tap) ; 3 = bilinear */
            uint8_t  pad;
        } mc;
        struct {                /* EAV1_OP_WARP (8x8) */
            int32_t  sx, sy, iw, ih;
            int32_t  mx, my;
            int16_t  abcd[4];
            uint8_t  ref, pad[3];
        } warp;
        struct {                /* AVG / W_AVG / MASK / W_MASK */
            uint32_t t1, t2;    /* arena byte offsets of the two int16 prep blocks (stride = w) */
            uint32_t mask_off;  /* MASK: mask source; W_MASK: arena offset the mask is written to */
            int16_t  weight;    /* W_AVG */
            uint8_t  sign;      /* W_MASK */
            uint8_t  ss_hor, ss_ver; /* W_MASK mask subsampling */
        } comp;
        struct {                /* BLEND / BLEND_V / BLEND_H */
            uint32_t tmp_off;   /* arena byte offset of the pixel tmp block (stride = w) */
            uint32_t mask_off;  /* BLEND: table-blob offset of the ii mask (stride = w) */
        } blend;
        struct {                /* IPRED and the edge part of CFL */
            int16_t  ex, ey;    /* position (pixels, in `plane` of the CURRENT frame) whose neighbours form the edge */
            int16_t  angle;     /* final angle | flags as passed to intra_pred (or filter idx) */
            uint8_t  mode;      /* final implementation mode (after prepare_intra_edges) */
            uint8_t  have;      /* bit0 have_left, bit1 have_top, bit2 z2 corner filter */
            uint8_t  n_left, n_bl, n_top, n_tr; /* available px counts; 255 = "not needed"; n_bl/n_tr 0 = replicate */
            uint8_t  tw, th;    /* edge sizes in 4px units */
            int16_t  max_w, max_h; /* clamped to 16383 */
            /* CFL extras */
            int16_t  lx, ly;    /* luma position for cfl_ac */
            int8_t   alpha;
            uint8_t  w_pad, h_pad;
            uint8_t  pad;
        } ipred;
        struct {                /* PAL */
            uint32_t pal_off;   /* payload byte offset: 8 pixels */
            uint32_t idx_off;   /* payload byte offset: w*h/2 packed indices */
        } pal;
        struct {                /* ITX */
            uint32_t cf_off;    /* payload byte offset of the coefficients: min(w,32) columns of `rows` entries each */
            int16_t  eob;
            uint8_t  tx, txtp;  /* enum RectTxfmSize, enum TxfmType (table index as used by itxfm_add[tx][txtp]) */
            uint8_t  rows;      /* rows of the column-major min(w,32) x min(h,32) block that are shipped (= the stride
                                   between columns); the rows behind them are zero: last_nonzero_col + 1 of
                                   itx_tmpl.c:87-105, derived from eob. 0 = the whole block (stride min(h,32)) */
        } itx;
        uint8_t raw[44];
    } u;
} Eav1Cmd; /* 60 bytes */

/* ---- post-filter side data ------------------------------------------------------------
 * Deblocking and loop restoration are recorded at the DSP-call level as well: the host's
 * sbrow drivers (lf_apply_tmpl.c:313-466, lr_apply_tmpl.c:107-202) run unchanged against a
 * recording Dav1dLoopFilterDSPContext / Dav1dLoopRestorationDSPContext, so tile-edge mask
 * fix-ups, frame-edge skipping and restoration-unit geometry arrive already resolved.
 * CDEF needs pixel-derived decisions (direction, variance) and is driven entirely on the
 * device from dav1d's Av1Filter array (cdef_idx, noskip_mask). */
typedef struct Eav1LfRec {      /* one lf.loop_filter_sb[plane>0][dir] call (loopfilter.h:39-43) */
    uint16_t x, y;              /* pixel position of `dst` in the plane */
    uint8_t  plane, dir;        /* dir: 0 = column edges (filter across x), 1 = row edges */
    uint16_t pad;
    uint32_t vmask[3];          /* which 4px units have an edge, by filter size class */
    uint32_t lvl_off;           /* byte offset of `lvl` inside the uploaded level array */
} Eav1LfRec;                    /* 24 bytes */

enum { EAV1_LR_WIENER7 = 0, EAV1_LR_WIENER5, EAV1_LR_SGR_5X5, EAV1_LR_SGR_3X3, EAV1_LR_SGR_MIX };
typedef struct Eav1LrRec {      /* one lr.wiener[]/lr.sgr[] call (looprestoration.h:64-70) */
    uint16_t x, y, w, h;        /* unit position/size in the plane; h = stripe height */
    uint8_t  plane, edges, kind, pad;
    union {
        int16_t filter[2][8];   /* Wiener taps as in LooprestorationParams.filter */
        struct { uint32_t s0, s1; int16_t w0, w1; } sgr;
    } p;
} Eav1LrRec;                    /* 44 bytes */

typedef struct Eav1FrameDesc {
    int w, h;                 /* visible frame size (luma) */
    int bpc;                  /* 8, 10, 12 */
    int layout;               /* Dav1dPixelLayout: 0=I400 1=I420 2=I422 3=I444 */
    int bw4, bh4;             /* frame size in 4x4 units (f->bw, f->bh) */
    int sb128;                /* seq_hdr->sb128 */
    int cur;                  /* picture id being reconstructed (receives the final frame) */
    int ref[7];               /* picture ids of the 7 references (-1 = none) */
    /* deblock */
    const Eav1LfRec *lf_recs; size_t n_lf_recs;
    const uint8_t *lf_level;  size_t lf_level_bytes; /* copy of f->lf.level */
    int b4_stride;
    uint8_t lut_e[64], lut_i[64];  /* Av1FilterLUT.e / .i (lf_mask.h:60-65) */
    /* cdef */
    int cdef_enable;
    int cdef_damping;
    uint8_t cdef_y_strength[8], cdef_uv_strength[8];
    const void *lf_mask;      /* Av1Filter[sb128h][sb128w] (1348 bytes each, lf_mask.h:51-58) */
    int sb128w, sb128h;
    /* loop restoration */
    const Eav1LrRec *lr_recs; size_t n_lr_recs;
    /* super-resolution (frame_hdr->super_res.enabled, width[0] != width[1]): reconstruction, deblocking and
     * CDEF run at the coded width in picture `cur`; mc.resize (mc_tmpl.c:918-955, driven by
     * dav1d_filter_sbrow_resize, recon_tmpl.c:2053-2086) then upscales every row into picture `sr_pic`
     * (sr_w luma samples wide), where loop restoration runs (its records are in sr_pic coordinates; the
     * stripe-border rows are the resized deblocked rows, lf_apply_tmpl.c:76-88). sr_pic receives the final
     * frame and is what later frames reference. superres == 0: everything happens in `cur`. */
    int superres, sr_pic, sr_w;
    int resize_step[2], resize_start[2];   /* f->resize_step / f->resize_start (decode.c:3532-3539): luma, chroma */
} Eav1FrameDesc;

/* ---- lifecycle ------------------------------------------------------------------------ */
int  eav1cu_device_count(void);
int  eav1cu_ctx_create(Eav1cuCtx **out, int device);
void eav1cu_ctx_destroy(Eav1cuCtx *ctx);
int  eav1cu_last_error(Eav1cuCtx *ctx, char *msg, size_t cap);
/* Forget a latched EAV1CU_ERR_INVAL (a bad argument of some earlier call); CUDA failures stay latched because the
 * device context is unusable after them. Returns 0 if the context is usable afterwards. Called by the host side
 * from dav1d_flush, so that a decoder can restart at the next key frame (ext/dav1d/src/lib.c:526). */
int  eav1cu_clear_error(Eav1cuCtx *ctx);
int  eav1cu_set_tables(Eav1cuCtx *ctx, const Eav1Tables *t);
int  eav1cu_set_masks(Eav1cuCtx *ctx, const uint8_t *blob, size_t bytes);
int  eav1cu_sync(Eav1cuCtx *ctx);

/* ---- pictures (device-resident reference pool) ---------------------------------------- */
int  eav1cu_pic_alloc(Eav1cuCtx *ctx, int w, int h, int layout, int bpc);           /* returns id >= 0 */
int  eav1cu_pic_release(Eav1cuCtx *ctx, int id);
int  eav1cu_pic_upload(Eav1cuCtx *ctx, int id, const void *const data[3], const ptrdiff_t stride[2]);
int  eav1cu_pic_download(Eav1cuCtx *ctx, int id, void *const data[3], const ptrdiff_t stride[2]);

/* ---- output conversion -------------------------------------------------------
 * An ADDITIONAL getter: the planar picture easyav1_get_video_frame / dav1d_get_picture hand out is unchanged.
 * Converts a finished picture on the device and copies the result to `dst` (host memory; pinned memory makes the
 * copy faster), so that the format every consumer of the reference converts to anyway (the SDL player uploads
 * Y/U/V textures and lets the GPU convert, tools/easyav1_player.c:1338; colour metadata mapping
 * src/easyav1.c:3287-3507) crosses PCIe once. No reference code computes these formats, so parity is against a
 * numpy restatement of the integer formulas below (tests/test_convert.py): "parity unpinned".
 *   NV12   8-bit 4:2:0 only: Y rows, then (h+1)/2 rows of interleaved U,V; row pitch dst_stride bytes
 *   P010   10/12-bit 4:2:0 only: the same with 16-bit samples, value << (16 - bpc)
 *   RGBA8  any layout and bit depth: R,G,B,255 per sample, chroma taken from the co-sited (x >> ss_hor, y >> ss_ver)
 *          sample. With y' = Y - y_off, u' = U - c_off, v' = V - c_off:
 *            R = clip8((ky*y' + krv*v' + 4096) >> 13),  G = clip8((ky*y' - kgu*u' - kgv*v' + 4096) >> 13),
 *            B = clip8((ky*y' + kbu*u' + 4096) >> 13)
 *          ky = rint(8192 * 255 / (yr << (bpc-8))), k.. = rint(8192 * 255 * m / (cr << (bpc-8))) with
 *          yr, cr = 219, 224 (limited range) or 255, 255 (full), m = 2(1-Kr), 2(1-Kb)Kb/Kg, 2(1-Kr)Kr/Kg, 2(1-Kb),
 *          (Kr, Kb) = (0.2126, 0.0722) for matrix 1 (BT.709), (0.2627, 0.0593) for 9 (BT.2020 NCL), else (0.299, 0.114);
 *          y_off = 16 << (bpc-8) (limited) or 0, c_off = 128 << (bpc-8). Monochrome pictures use u' = v' = 0.
 * Returns 0, EAV1CU_ERR_INVAL for a format the picture's layout / bit depth does not allow or a dst that is too
 * small (needed bytes: dst_stride * (h + (h+1)/2) for NV12 / P010, dst_stride * h for RGBA8). Synchronous. */
enum { EAV1CU_FMT_NV12 = 1, EAV1CU_FMT_P010 = 2, EAV1CU_FMT_RGBA8 = 3 };
typedef struct Eav1Convert {
    int32_t format;        /* EAV1CU_FMT_* */
    int32_t matrix;        /* enum Dav1dMatrixCoefficients (include/dav1d/headers.h), RGBA8 only */
    int32_t full_range;    /* Dav1dSequenceHeader.color_range, RGBA8 only */
} Eav1Convert;
int  eav1cu_pic_convert(Eav1cuCtx *ctx, int id, const Eav1Convert *cv, void *dst, size_t dst_stride, size_t dst_bytes);

/* Pinned host memory for output pictures (backs the Dav1dPicAllocator replacement). */
int  eav1cu_host_alloc(void **p, size_t bytes);
void eav1cu_host_free(void *p);

/* ---- frame reconstruction --------------------------------------------------------------- */
/* Asynchronous variant of eav1cu_pic_download: queues the copies on the context's stream and
 * returns; pair it with a fence. `data` must be pinned (eav1cu_host_alloc) to really overlap. */
int  eav1cu_pic_download_async(Eav1cuCtx *ctx, int id, void *const data[3], const ptrdiff_t stride[2]);
/* A fence marks everything queued so far on the context's stream; eav1cu_fence_wait blocks the
 * calling host thread until that work has finished (callable from any thread). */
int  eav1cu_fence(Eav1cuCtx *ctx, unsigned long long *ticket);
int  eav1cu_fence_wait(Eav1cuCtx *ctx, unsigned long long ticket);

/* Reconstruct one frame from its recorded commands (in host walk order) and run the in-loop
 * post filters; the result stays resident as picture fd->cur. Asynchronous on the context's
 * stream; eav1cu_pic_download / eav1cu_sync order after it. */
/* Per-call (operator-level) use, as tests/test_ops.py drives it: a command list of independent commands with no
 * post-filter side data runs exactly those DSP operators on picture fd->cur; a submission with n_cmds == 0 runs only
 * the post filters described by fd on the samples fd->cur already holds. */
int  eav1cu_frame_submit(Eav1cuCtx *ctx, const Eav1FrameDesc *fd,
                         const Eav1Cmd *cmds, size_t n_cmds,
                         const void *payload, size_t payload_bytes, size_t arena_bytes);

/* Scratch arena access (prep blocks / masks), for operator-level parity tests. */
int  eav1cu_arena_upload(Eav1cuCtx *ctx, size_t off, const void *src, size_t bytes);
int  eav1cu_arena_download(Eav1cuCtx *ctx, size_t off, void *dst, size_t bytes);
/* Number of kernel launches issued so far by this context (bench.py's gpu_launches). */
long long eav1cu_launch_count(Eav1cuCtx *ctx);

/* Measurement support. With capture on, every frame passed to eav1cu_frame_submit also keeps
 * a private device-resident copy of its work lists, so the frame can be re-launched with no
 * host work and no host->device traffic (bench.py's `value` leg: inputs resident in HBM).
 * Replaying the staged frames in submit order reproduces the original pictures, because each
 * frame reads the same reference slots it read the first time. */
typedef struct Eav1StagedInfo {
    int cur, w, h, bpc, layout;
    int n_levels, n_chains, launches;
    int n_lf, n_lr, cdef, inter;
    int n_big, n_small;      /* dataflow units run by a whole CTA / by a single warp */
    int n_mc_items, n_resid_items; /* phase-1 work items: motion compensation (<= 32x32 each) / transform blocks */
    long long n_cmds, payload_bytes;
    long long mc_px, unit_px; /* samples produced by phase 1 / touched by phase-2 units */
    long long resid_px;       /* samples of the transform blocks whose residual phase 1b computes */
} Eav1StagedInfo;
/* sizeof() of the ABI structs as compiled (0 Eav1Cmd, 1 Eav1LfRec, 2 Eav1LrRec, 3 Eav1FrameDesc,
 * 4 Eav1StagedInfo, 5 Eav1Tables), so bindings can verify their mirrors. */
size_t eav1cu_abi_sizeof(int which);
/* Host-only: the dependency levels eav1cu_frame_submit would give these commands. Chains
 * (coding blocks) come back sorted by level; returns the number of chains. Needs no device. */
int  eav1cu_plan_levels(const Eav1Cmd *cmds, size_t n_cmds, const int plane_w[3], const int plane_h[3],
                        uint32_t *chain_first, uint32_t *chain_level, size_t cap);
/* Host-only: the two-phase plan eav1cu_frame_submit makes. Phase-1 motion-compensation items
 * (command index | tile_x << 24 | tile_y << 26) and the dataflow units (first command, number of
 * fused commands, dependency level) in decode order. Returns the number of units. */
int  eav1cu_plan_units(const Eav1Cmd *cmds, size_t n_cmds, const int plane_w[3], const int plane_h[3],
                       uint32_t *unit_first, uint32_t *unit_count, uint32_t *unit_level, size_t cap_units,
                       uint32_t *mc_items, size_t cap_items, size_t *n_items);
/* Ticket order of the dataflow kernel for the units of eav1cu_plan_units (ticket t runs unit order[t]) and
 * the producer lists (CSR, dep_start has n + 1 entries); est_weight as in csrc/sched.h. Needs no GPU. */
int  eav1cu_plan_tickets(const Eav1Cmd *cmds, size_t n_cmds, const int plane_w[3], const int plane_h[3],
                         int est_weight, uint32_t *order, size_t cap_units, uint32_t *dep_start,
                         uint32_t *deps, size_t cap_deps, size_t *n_deps);
/* out[0] kernel launches, out[1] host->device bytes, out[2] device->host bytes, out[3] frames */
int  eav1cu_counters(Eav1cuCtx *ctx, long long out[4]);
/* the same four totals over every context of the process since it started (frames submitted through
 * eav1cu_frame_submit only; staged replays are not counted) */
int  eav1cu_global_counters(long long out[4]);
int  eav1cu_set_capture(Eav1cuCtx *ctx, int on);
/* Test support: which coding tools the frames submitted by this process exercised. Counting is on only when
 * EAV1_COVERAGE is set in the environment when the library is first used (eav1cu_coverage then returns 1,
 * else 0 and zeros). Indices: bit depth 0/1/2 = 8/10/12 bits, `hbd` 0/1 = 8-bit / high bit depth;
 * tx = enum RectTxfmSize, txtp = enum TxfmType (16 = WHT_WHT), mode = the implementation intra mode
 * (ext/dav1d/src/levels.h); w / h as floor(log2). */
typedef struct Eav1Coverage {
    uint64_t frames[3], layout[4];
    uint64_t itx[3][19][17];            /* [bit depth][tx][txtp] */
    uint64_t itx_eob[3][3];             /* [bit depth][dc-only shortcut / eob < 10 / larger] */
    uint64_t mc[2][2][2][4][4][4];      /* [hbd][scaled][prep][filt_h][filt_v][bit0 mx != 0, bit1 my != 0; scaled: 3] */
    uint64_t mc_wh[2][8][8];            /* [scaled][w][h] */
    uint64_t intrabc;
    uint64_t warp[2][2];                /* [hbd][put / prep] */
    uint64_t comp[2][4];                /* [hbd][avg, w_avg, mask, w_mask] */
    uint64_t w_mask_ss[3];              /* 4:4:4, 4:2:2, 4:2:0 */
    uint64_t blend[2][3];               /* [hbd][blend, blend_v, blend_h] */
    uint64_t ipred[3][14];              /* [bit depth][mode] */
    uint64_t ipred_wh[8][8];
    uint64_t cfl[2][14];                /* [hbd][mode] */
    uint64_t pal[2];
    uint64_t lf[2][2][2][3];            /* [hbd][chroma][dir][filter size class]: 4-sample edge units */
    uint64_t lr[2][5];                  /* [hbd][EAV1_LR_*] recorded calls */
    uint64_t cdef_frames[2], superres_frames[2];
    uint64_t grain[2][2];               /* [hbd][chroma scaled from luma] eav1cu_apply_grain calls */
} Eav1Coverage;
int  eav1cu_coverage(Eav1Coverage *out, int reset);
/* Per-stage device time (CUDA events on the context's stream around each stage of every
 * frame): index 0 dataflow reconstruction (intra, residual, compound/blend), 1 deblock, 2 CDEF,
 * 3 loop restoration, 4 motion compensation, 5 film grain, 6 inverse-transform residuals. Each array has EAV1CU_N_STAGES entries: accumulated
 * milliseconds, kernel launches and stage invocations since the last reset. */
#define EAV1CU_N_STAGES 7
int  eav1cu_set_stage_timing(Eav1cuCtx *ctx, int on);
int  eav1cu_stage_times(Eav1cuCtx *ctx, double *ms, long long *launches, long long *calls, int reset);
int  eav1cu_staged_count(Eav1cuCtx *ctx);
int  eav1cu_staged_replay(Eav1cuCtx *ctx, int idx);
int  eav1cu_staged_info(Eav1cuCtx *ctx, int idx, Eav1StagedInfo *out);
/* all staged frames in submit order, `passes` times, queued from one C loop (one host thread per context lets
 * several staged streams share the GPU: bench.py's multi-stream device-resident leg) */
int  eav1cu_staged_replay_all(Eav1cuCtx *ctx, int passes);
/* The same replay captured once as a CUDA graph (every kernel of every staged frame, both phase-1 streams, the
 * flag reset in front): _build captures and instantiates, _launch queues `passes` replays, one cudaGraphLaunch each.
 * The stream passed to eav1cu_set_stream (or the context's own) is the capture and launch stream. */
int  eav1cu_staged_graph_build(Eav1cuCtx *ctx);
int  eav1cu_staged_graph_launch(Eav1cuCtx *ctx, int passes);
/* Development aid: replay staged frame idx once with per-unit device timestamps. out gets 6 words
 * per dataflow unit (index = unit id): ticket drawn, producers ready, done (ns), id of the producer
 * that finished last (~0: none), first command << 32 | count, ticket << 32 | CTA << 8 | warp. */
/* Development aid: cycle counters of -DEAV1_PROF=1 builds (all zero otherwise); reading resets them. */
int  eav1cu_prof_read(unsigned long long out[64]);
int  eav1cu_staged_trace(Eav1cuCtx *ctx, int idx, unsigned long long *out, size_t cap_units);
/* Run on a caller-owned cudaStream_t (passed as void*) so the caller can record its own
 * events around the launches; the context never destroys a stream it did not create. */
int  eav1cu_set_stream(Eav1cuCtx *ctx, void *cuda_stream);

/* ---- film grain synthesis ----------------------------------------------
 * Replaces dav1d_apply_grain_{8,16}bpc (ext/dav1d/src/fg_apply_tmpl.c:225-241: prep_grain :100,
 * apply_grain_row :161) and the Dav1dFilmGrainDSPContext entries behind it (filmgrain_tmpl.c:
 * generate_grain_y :50, generate_grain_uv :88, fgy_32x32xn :169, fguv_32x32xn :278).
 * Eav1GrainParams is Dav1dFilmGrainData (include/dav1d/headers.h:315-333) with fixed-width
 * members plus the one sequence-header bit the chroma clip depends on. */
typedef struct Eav1GrainParams {
    uint32_t seed;
    int32_t  num_y_points;
    uint8_t  y_points[14][2];
    int32_t  chroma_scaling_from_luma;
    int32_t  num_uv_points[2];
    uint8_t  uv_points[2][10][2];
    int32_t  scaling_shift, ar_coeff_lag;
    int8_t   ar_coeffs_y[24];
    int8_t   ar_coeffs_uv[2][28];
    int32_t  ar_coeff_shift, grain_scale_shift;
    int32_t  uv_mult[2], uv_luma_mult[2], uv_offset[2];
    int32_t  overlap_flag, clip_to_restricted_range;
    int32_t  mtrx_identity;             /* seq_hdr->mtrx == DAV1D_MC_IDENTITY (fg_apply_tmpl.c:172) */
} Eav1GrainParams;
/* dst = src + grain, both device pictures of the same geometry; stream-ordered after every frame
 * submitted so far. Three launches: grain templates + scaling LUTs + block offsets, then all planes. */
int  eav1cu_apply_grain(Eav1cuCtx *ctx, int src_pic, int dst_pic, const Eav1GrainParams *g);
/* eav1cu_staged_graph_build with film grain (parameters g) applied to every staged frame into picture dst_pic inside
 * the same graph: the template build of a frame forks onto a stream of its own at the start of the frame and the
 * apply joins after the frame's last kernel, as eav1cu_apply_grain arranges it for direct launches. */
int  eav1cu_staged_graph_build_grain(Eav1cuCtx *ctx, int dst_pic, const Eav1GrainParams *g);

/* eav1cu_apply_grain_noalloc has NO counterpart in the reference; it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid it). Same arguments, same kernels and bit-identical output as
 * eav1cu_apply_grain, but the scratch those kernels need (grain templates, scaling LUTs, per-block
 * template offsets; 83,968 bytes) is a file-scope static __device__ array that every GPU loading the
 * library holds until the process exits, so the call performs no cudaMalloc / cudaFree / allocating
 * library call. The caller still passes input and output: src_pic / dst_pic are pictures it allocated
 * earlier with eav1cu_pic_alloc on a context whose tables are set.
 *   limit      pictures of at most EAV1CU_NOALLOC_MAX_GRAIN_BLOCKS 32x32 luma blocks,
 *              ceil(w/32) * ceil(h/32) <= 34,816 (e.g. 8192x4352); anything larger is rejected with
 *              EAV1CU_ERR_INVAL and a message naming the limit (use eav1cu_apply_grain then).
 *   sharing    every call on a device uses the same array, which keeps the previous call's contents;
 *              each call rewrites everything it later reads, so nothing needs clearing, but two calls
 *              whose kernels overlap on the GPU (different streams / contexts of one device) RACE on
 *              it. Serialise them (one stream, or an event between them).
 *   set-up     the array's device address is looked up once per device (cudaGetSymbolAddress, on the
 *              first call for that device); callers under graph capture make one call outside capture
 *              first. */
#define EAV1CU_NOALLOC_MAX_GRAIN_BLOCKS 34816
int  eav1cu_apply_grain_noalloc(Eav1cuCtx *ctx, int src_pic, int dst_pic, const Eav1GrainParams *g);

/* ---- Vorbis float slice ----------------------------------------------------
 * Replaces, for a batch of decoded packets, mdct_backward (ext/minivorbis/minivorbis.h:7054-7148,
 * called from mapping0_inverse :17825), the window + overlap-add of vorbis_synthesis_blockin
 * (:9195-9300) and the planar -> interleaved copy of src/easyav1.c:2471-2492. The bit-serial part of
 * vorbis_synthesis (floor1, residue, codebooks) stays on the host and produces the spectra.
 * One object per audio track; the overlap state carries over between calls like
 * vorbis_dsp_state.pcm does, eav1cu_vorbis_restart() is vorbis_synthesis_restart (:9159).
 *   flags[n_blocks]    0 short / 1 long block (vorbis_block.W)
 *   spectra            block-major: [block][channel][blocksize/2] floats, as mapping0_inverse holds
 *                      them in vb->pcm[channel] just before the transform
 *   pcm                planar [channel][cap_samples] or interleaved [sample][channels]
 *   *n_samples         samples per channel made final by this batch: sum over consecutive block
 *                      pairs of blocksize[lW]/4 + blocksize[W]/4 (the very first block of a stream
 *                      only primes the overlap, minivorbis.h:9300-9308)
 * window0 / window1: the half windows (blocksize/2 floats) the host's decoder uses
 * (_vorbis_window_get, minivorbis.h:11992), or NULL to have them generated from the Vorbis window
 * formula. Results match the reference within 1e-5 of full scale (different butterfly order). */
typedef struct Eav1Vorbis Eav1Vorbis;
int  eav1cu_vorbis_open(Eav1Vorbis **out, int device, int channels, int blocksize0, int blocksize1,
                        const float *window0, const float *window1);
void eav1cu_vorbis_close(Eav1Vorbis *v);
int  eav1cu_vorbis_last_error(Eav1Vorbis *v, char *msg, size_t cap);
int  eav1cu_vorbis_restart(Eav1Vorbis *v);
int  eav1cu_vorbis_imdct_ola(Eav1Vorbis *v, const uint8_t *flags, int n_blocks, const float *spectra,
                             float *pcm, size_t cap_samples, int interleave, size_t *n_samples);
/* measurement: keep one batch resident in device memory, replay the kernel on it, read it back */
int  eav1cu_vorbis_stage(Eav1Vorbis *v, const uint8_t *flags, int n_blocks, const float *spectra,
                         int interleave, size_t *n_samples);
int  eav1cu_vorbis_run_staged(Eav1Vorbis *v);
int  eav1cu_vorbis_fetch(Eav1Vorbis *v, float *pcm, size_t cap_floats);
int  eav1cu_vorbis_sync(Eav1Vorbis *v);
int  eav1cu_vorbis_set_stream(Eav1Vorbis *v, void *cuda_stream);
long long eav1cu_vorbis_launch_count(Eav1Vorbis *v);

#ifdef __cplusplus
}
#endif
#endif /* EASYAV1_CUDA_H */
