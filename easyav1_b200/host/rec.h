/*
 * rec.h — host-side command recorder that stands in for dav1d's Dav1dDSPContext tables.
 *
 * The reference's HOST code (OBU parse, msac entropy decode, decode_b / recon_b_* block walk,
 * lf_apply / lr_apply sbrow drivers, the frame/tile task scheduler of thread_task.c) is compiled
 * IN PLACE from /root/reference and linked with this recorder instead of the reference's
 * *_tmpl.c DSP objects. Every DSP call the walk makes (ext/dav1d/src/internal.h:62-70) appends
 * one Eav1Cmd / Eav1LfRec / Eav1LrRec; nothing here touches pixels.
 *
 * Threading follows dav1d's own model (thread_task.c): many frames and tiles are walked
 * concurrently by the worker pool. Recording state is therefore split three ways:
 *   Eav1Walk   thread-local: the coding block being recorded right now
 *   Eav1Frame  one per Dav1dFrameContext slot: command lists (one per tile), coefficient
 *              payload, deblock / restoration records of the frame in that slot
 *   Eav1Rec    one per Dav1dContext: engine handle, job queue, submit thread
 * When a frame's last task has finished (dav1d_decode_frame_exit) its lists are detached into
 * an Eav1Job; one submit thread per decoder hands jobs to libeasyav1_cuda in decode order and
 * brings the planes back asynchronously; dav1d_get_picture waits for the picture's fence.
 */
#ifndef EAV1_REC_H
#define EAV1_REC_H

#include <pthread.h>
#include <stdatomic.h>
#include <stddef.h>
#include <stdint.h>

#include "config.h"
#include "dav1d/dav1d.h"
#include "src/internal.h"

#include "easyav1_cuda.h"

struct Eav1Rec;

/* what the picture allocator hangs on Dav1dPicture.allocator_data */
typedef struct Eav1PicData {
    void *host_base;
    size_t host_bytes;
    int dev_id;
    struct Eav1Rec *rec;
    atomic_int refs;                    /* the Dav1dPicture itself + every queued job using it */
    atomic_ullong fence;                /* 0: nothing pending, else engine fence of its download */
    atomic_int failed;
    void *alloc_state;
} Eav1PicData;

enum { REG_TMP0, REG_TMP1, REG_LAP, REG_II, REG_SEGMASK, N_REGIONS };

typedef struct Eav1List { Eav1Cmd *cmds; size_t n, cap; } Eav1List;

#define EAV1_MAX_TILES 4096             /* 64 x 64, ext/dav1d/include/dav1d/headers.h */

/* everything recorded for the frame currently decoded in one Dav1dFrameContext slot */
typedef struct Eav1Frame {
    int open;
    Eav1List *tiles;                    /* EAV1_MAX_TILES lists; only [0, n_tiles) are used */
    int n_tiles;
    uint8_t *payload; size_t cap_payload; atomic_size_t n_payload;
    atomic_size_t arena_used;
    pthread_mutex_t mu;                 /* guards lf / lr (filter tasks of different rows race) */
    Eav1LfRec *lf; size_t n_lf, cap_lf;
    Eav1LrRec *lr; size_t n_lr, cap_lr;
    atomic_int unsupported;
    unsigned long long seq_q[4]; int seq_n;   /* decode-order tickets of frames routed to this slot */
} Eav1Frame;

/* a finished frame on its way to the engine (rec_common.c) */
typedef struct Eav1Job Eav1Job;

typedef struct Eav1Rec {
    const Dav1dContext *c;
    Eav1cuCtx *cu;
    Eav1Frame *frames; int n_frames;
    int threaded;                       /* dav1d runs frame threads: frames end in the exit hook */
    int inline_submit;                  /* single-threaded walk: the walking thread hands frames to the engine itself */
    /* job queue (decode order) */
    pthread_mutex_t q_mu; pthread_cond_t q_cv, q_idle;
    pthread_mutex_t cu_mu;              /* engine calls: the submit thread vs. film grain at output time */
    Eav1Job *q_head, *job_free;
    unsigned long long seq_assign, seq_submit;
    int q_stop, q_busy;
    pthread_t submit_thread; int have_thread;
    void *alloc_state;
    atomic_int error;                   /* engine failure (latched until dav1d_flush / the next shown key frame) */
    atomic_llong frames_submitted;
} Eav1Rec;

/* thread-local: the block being walked */
typedef struct Eav1Walk {
    Eav1Rec *r;
    Eav1Frame *fr;
    Eav1List *list;
    const Dav1dFrameContext *f;
    const Dav1dTaskContext *t;
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
} Eav1Walk;

extern __thread Eav1Walk eav1_walk;

Eav1Rec *eav1rec_for_ctx(const Dav1dContext *c);
/* bind the calling thread to frame f (and tile state ts, may be NULL for filter tasks) */
Eav1Walk *eav1rec_enter(const Dav1dFrameContext *f, const Dav1dTaskContext *t);
void eav1rec_unsupported(Eav1Walk *w);   /* the frame being walked cannot be recorded: it alone fails */
Eav1Cmd *eav1rec_push(Eav1Walk *w);
uint32_t eav1rec_payload(Eav1Walk *w, const void *src, size_t bytes, size_t align);
/* room for `bytes` in the frame's payload; *dst receives where to write them (NULL when the payload is full) */
uint32_t eav1rec_payload_reserve(Eav1Walk *w, size_t bytes, size_t align, void **dst);
uint32_t eav1rec_region(Eav1Walk *w, int reg, size_t elem_bytes);
uint32_t eav1rec_region_new(Eav1Walk *w, int reg, size_t bytes);
Eav1LfRec *eav1rec_push_lf(Eav1Walk *w, const Eav1LfRec *v);
Eav1LrRec *eav1rec_push_lr(Eav1Walk *w, const Eav1LrRec *v);
void eav1rec_block_begin(Eav1Walk *w, int bs);
/* frame finished (retval as passed to dav1d_decode_frame_exit): detach + queue, or run inline */
void eav1rec_frame_end(Eav1Rec *r, const Dav1dFrameContext *f, int retval);

/* locate a pixel pointer inside the current picture: returns plane or -1 */
int eav1rec_locate_cur(const Eav1Walk *w, const void *p, int *x, int *y);
/* same inside the super-resolved picture f->sr_cur (== f->cur without super-resolution) */
int eav1rec_locate_sr(const Eav1Walk *w, const void *p, int *x, int *y);
/* locate a pixel pointer inside a reference (0..6) or the current picture (7) */
int eav1rec_locate_ref(const Eav1Walk *w, const void *p, int *plane, int *x, int *y);

int eav1rec_picture_convert(const Dav1dPicture *p, int format, int matrix, int full_range, void *dst, size_t dst_stride, size_t dst_bytes);

#endif
