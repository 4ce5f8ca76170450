/*
 * rec.h — host-side command recorder that stands in for dav1d's Dav1dDSPContext tables.
 *
 * The reference's host code (OBU parse, msac entropy decode, decode_b / recon_b_* block walk,
 * lf_apply / lr_apply sbrow drivers) is compiled IN PLACE from /root/reference and linked with
 * this recorder instead of the reference's *_tmpl.c DSP objects. Every DSP call the walk
 * makes (ext/dav1d/src/internal.h:62-70) appends one Eav1Cmd / Eav1LfRec / Eav1LrRec; nothing
 * here touches pixels. At the last filter_sbrow of a frame the lists go to libeasyav1_cuda.
 */
#ifndef EAV1_REC_H
#define EAV1_REC_H

#include <stddef.h>
#include <stdint.h>

#include "config.h"
#include "dav1d/dav1d.h"
#include "src/internal.h"

#include "easyav1_cuda.h"

/* what the picture allocator hangs on Dav1dPicture.allocator_data */
typedef struct Eav1PicData {
    void *host_base;
    size_t host_bytes;
    int dev_id;
    struct Eav1Rec *rec;
} Eav1PicData;

enum { REG_TMP0, REG_TMP1, REG_LAP, REG_II, REG_SEGMASK, N_REGIONS };

typedef struct Eav1Rec {
    const Dav1dContext *c;
    Eav1cuCtx *cu;
    /* growable lists */
    Eav1Cmd *cmds; size_t n_cmds, cap_cmds;
    uint8_t *payload; size_t n_payload, cap_payload;
    Eav1LfRec *lf; size_t n_lf, cap_lf;
    Eav1LrRec *lr; size_t n_lr, cap_lr;
    size_t arena_used;
    /* current frame / block */
    const Dav1dFrameContext *f;
    const Dav1dTaskContext *t;
    int frame_open;
    int block_head;               /* next command starts a new coding block */
    int blk_w, blk_h;             /* luma size of the current block */
    uint32_t region_off[N_REGIONS];
    uint8_t region_live[N_REGIONS];
    /* pending emu_edge */
    struct {
        const void *dst; ptrdiff_t dst_stride; int bw, bh;
        int x, y, iw, ih, ref, plane, valid;
    } emu;
    /* pending prepare_intra_edges */
    struct {
        const void *topleft; int valid;
        int ex, ey, plane, mode, have, n_left, n_bl, n_top, n_tr, tw, th;
    } edge;
    /* pending cfl_ac */
    struct { int valid, lx, ly, w_pad, h_pad, cw, ch; } ac;
    void *alloc_state;
    int unsupported;              /* sticky: a tool we do not run yet was seen */
    int error;
    long long frames_submitted;
} Eav1Rec;

extern __thread Eav1Rec *eav1_tls_rec;

Eav1Rec *eav1rec_for_ctx(const Dav1dContext *c);
Eav1Cmd *eav1rec_push(Eav1Rec *r);
uint32_t eav1rec_payload(Eav1Rec *r, const void *src, size_t bytes, size_t align);
uint32_t eav1rec_region(Eav1Rec *r, int reg, size_t elem_bytes);
uint32_t eav1rec_region_new(Eav1Rec *r, int reg, size_t bytes);
Eav1LfRec *eav1rec_push_lf(Eav1Rec *r);
Eav1LrRec *eav1rec_push_lr(Eav1Rec *r);
void eav1rec_block_begin(Eav1Rec *r, const Dav1dTaskContext *t, int bs);
int  eav1rec_frame_end(Eav1Rec *r, const Dav1dFrameContext *f);

/* locate a pixel pointer inside the current picture: returns plane or -1 */
int eav1rec_locate_cur(const Eav1Rec *r, const void *p, int *x, int *y);
/* locate a pixel pointer inside a reference (0..6) or the current picture (7) */
int eav1rec_locate_ref(const Eav1Rec *r, const void *p, int *plane, int *x, int *y);

#endif
