/*
 * rec_common.c — bit-depth independent part of the recorder: decoder open/close wrappers,
 * pinned picture allocator tied to the device picture pool, constant-table upload, per-frame
 * list management, the hooks into dav1d's frame life cycle and the submit thread. See rec.h.
 */
#include <errno.h>
#include <pthread.h>
#include <stdio.h>
#include <time.h>
#include <stdlib.h>
#include <string.h>

#include "rec.h"
#include "common/frame.h"
#include "src/tables.h"
#include "src/wedge.h"
#include "src/decode.h"
#include "src/thread_task.h"

/* the reference's own entry points, renamed at compile time (host/Makefile) */
int  eav1ref_dav1d_open(Dav1dContext **c_out, const Dav1dSettings *s);
void eav1ref_dav1d_close(Dav1dContext **c_out);
int  eav1ref_dav1d_get_picture(Dav1dContext *c, Dav1dPicture *out);
void eav1ref_dav1d_flush(Dav1dContext *c);

/* pinned host memory comes from the engine so this library needs no CUDA headers */
int  eav1cu_host_alloc(void **p, size_t bytes);
void eav1cu_host_free(void *p);

#define FENCE_PENDING (~0ULL)

__thread Eav1Walk eav1_walk;

#define MAX_RECS 256
static pthread_mutex_t g_mu = PTHREAD_MUTEX_INITIALIZER;
static Eav1Rec *g_recs[MAX_RECS];

/* Every thread caches the recorder it used last. A decoder can be torn down by another thread (the last
 * dav1d_picture_unref), so the cached pointer is only dereferenced while no recorder has been destroyed since it was
 * cached: destroy_rec bumps the generation. */
static atomic_uint g_rec_generation;
static __thread unsigned t_rec_generation;

Eav1Rec *eav1rec_for_ctx(const Dav1dContext *c)
{
    Eav1Rec *r = eav1_walk.r;
    const unsigned gen = atomic_load_explicit(&g_rec_generation, memory_order_acquire);
    if (r && t_rec_generation == gen && r->c == c) return r;
    t_rec_generation = gen;
    pthread_mutex_lock(&g_mu);
    r = NULL;
    for (int i = 0; i < MAX_RECS; i++)
        if (g_recs[i] && g_recs[i]->c == c) { r = g_recs[i]; break; }
    pthread_mutex_unlock(&g_mu);
    eav1_walk.r = r;
    return r;
}

/* ------------------------------------------------------------------ lists */

static void *grow(void *p, size_t *cap, size_t need, size_t elem)
{
    if (need <= *cap) return p;
    size_t n = *cap ? *cap * 2 : 4096;
    while (n < need) n *= 2;
    p = realloc(p, n * elem);
    if (!p) abort();
    *cap = n;
    return p;
}

void eav1rec_unsupported(Eav1Walk *w)
{
    if (w && w->fr) atomic_store(&w->fr->unsupported, 1);
}

Eav1Cmd *eav1rec_push(Eav1Walk *w)
{
    Eav1List *l = w->list;
    l->cmds = grow(l->cmds, &l->cap, l->n + 1, sizeof(Eav1Cmd));
    Eav1Cmd *c = &l->cmds[l->n++];
    memset(c, 0, sizeof(*c));
    if (w->block_head) { c->flags |= EAV1_F_BLOCK_HEAD; w->block_head = 0; }
    return c;
}

/* The payload buffer is sized for the worst case when the frame opens (one coefficient per
 * sample plus palette data), so concurrent tiles only need an atomic bump. */
uint32_t eav1rec_payload(Eav1Walk *w, const void *src, size_t bytes, size_t align)
{
    Eav1Frame *fr = w->fr;
    const size_t take = (bytes + align - 1) & ~(align - 1);
    const size_t off = atomic_fetch_add(&fr->n_payload, take);
    if (off + bytes > fr->cap_payload) { atomic_store(&fr->unsupported, 1); return 0; }
    memcpy(fr->payload + off, src, bytes);
    return (uint32_t)off;
}

uint32_t eav1rec_payload_reserve(Eav1Walk *w, size_t bytes, size_t align, void **dst)
{
    Eav1Frame *fr = w->fr;
    const size_t take = (bytes + align - 1) & ~(align - 1);
    const size_t off = atomic_fetch_add(&fr->n_payload, take);
    if (off + bytes > fr->cap_payload) { atomic_store(&fr->unsupported, 1); *dst = NULL; return 0; }
    *dst = fr->payload + off;
    return (uint32_t)off;
}

/* A producer starts a fresh region every time (single assignment): no command ever overwrites
 * scratch that an earlier command of the frame still has to read, so the engine is free to run
 * all motion compensation of a frame before anything else. */
uint32_t eav1rec_region_new(Eav1Walk *w, int reg, size_t bytes)
{
    const size_t take = (bytes + 15) & ~(size_t)15;
    const size_t off = atomic_fetch_add(&w->fr->arena_used, take);
    w->region_off[reg] = (uint32_t)off;
    w->region_live[reg] = 1;
    return (uint32_t)off;
}

uint32_t eav1rec_region(Eav1Walk *w, int reg, size_t elem_bytes)
{
    if (!w->region_live[reg]) eav1rec_region_new(w, reg, (size_t)w->blk_w * w->blk_h * elem_bytes);
    return w->region_off[reg];
}

Eav1LfRec *eav1rec_push_lf(Eav1Walk *w, const Eav1LfRec *v)
{
    Eav1Frame *fr = w->fr;
    pthread_mutex_lock(&fr->mu);
    fr->lf = grow(fr->lf, &fr->cap_lf, fr->n_lf + 1, sizeof(Eav1LfRec));
    Eav1LfRec *x = &fr->lf[fr->n_lf++];
    *x = *v;
    pthread_mutex_unlock(&fr->mu);
    return x;
}

Eav1LrRec *eav1rec_push_lr(Eav1Walk *w, const Eav1LrRec *v)
{
    Eav1Frame *fr = w->fr;
    pthread_mutex_lock(&fr->mu);
    fr->lr = grow(fr->lr, &fr->cap_lr, fr->n_lr + 1, sizeof(Eav1LrRec));
    Eav1LrRec *x = &fr->lr[fr->n_lr++];
    *x = *v;
    pthread_mutex_unlock(&fr->mu);
    return x;
}

/* ------------------------------------------------------------------ frame open */

static size_t payload_worst_case(const Dav1dFrameContext *f)
{
    const size_t luma = (size_t)f->bw * 4 * f->bh * 4;
    const int layout = f->cur.p.layout;
    const size_t samples = layout == DAV1D_PIXEL_LAYOUT_I400 ? luma :
                           layout == DAV1D_PIXEL_LAYOUT_I420 ? luma * 3 / 2 :
                           layout == DAV1D_PIXEL_LAYOUT_I422 ? luma * 2 : luma * 3;
    const size_t coef = f->cur.p.bpc > 8 ? 4 : 2;
    /* one coefficient per sample + palette indices (<= 1 byte/sample incl. the 16-byte rounding) */
    return samples * (coef + 1) + (1 << 16);
}

static void frame_begin(Eav1Frame *fr, const Dav1dFrameContext *f)
{
    pthread_mutex_lock(&fr->mu);
    if (!fr->open) {
        const size_t need = payload_worst_case(f);
        if (need > fr->cap_payload) {
            free(fr->payload);
            fr->payload = malloc(need);
            if (!fr->payload) abort();
            fr->cap_payload = need;
        }
        for (int i = 0; i < fr->n_tiles; i++) fr->tiles[i].n = 0;
        fr->n_tiles = f->frame_hdr->tiling.cols * f->frame_hdr->tiling.rows;
        if (fr->n_tiles > EAV1_MAX_TILES) fr->n_tiles = EAV1_MAX_TILES;
        for (int i = 0; i < fr->n_tiles; i++) fr->tiles[i].n = 0;
        atomic_store(&fr->n_payload, 0);
        atomic_store(&fr->arena_used, 0);
        fr->n_lf = fr->n_lr = 0;
        atomic_store(&fr->unsupported, 0);
        fr->open = 1;
    }
    pthread_mutex_unlock(&fr->mu);
}

Eav1Walk *eav1rec_enter(const Dav1dFrameContext *f, const Dav1dTaskContext *t)
{
    Eav1Walk *w = &eav1_walk;
    Eav1Rec *r = eav1rec_for_ctx(f->c);
    w->r = r;
    w->f = f;
    w->t = t;
    w->fr = &r->frames[f - f->c->fc];
    if (!w->fr->open) frame_begin(w->fr, f);
    w->list = NULL;
    if (t && t->ts) {
        const ptrdiff_t ti = t->ts - f->ts;
        w->list = &w->fr->tiles[ti >= 0 && ti < w->fr->n_tiles ? ti : 0];
    }
    return w;
}

void eav1rec_block_begin(Eav1Walk *w, int bs)
{
    w->block_head = 1;
    w->blk_w = dav1d_block_dimensions[bs][0] * 4;
    w->blk_h = dav1d_block_dimensions[bs][1] * 4;
    memset(w->region_live, 0, sizeof(w->region_live));
    w->emu.valid = 0;
    w->edge.valid = 0;
    w->ac.valid = 0;
}

/* ------------------------------------------------------------------ pointer provenance */

static int locate(const Dav1dPicture *p, const void *ptr, int *plane, int *x, int *y)
{
    const int hbd = p->p.bpc > 8;
    const int ss_ver = p->p.layout == DAV1D_PIXEL_LAYOUT_I420;
    const int ah = (p->p.h + 127) & ~127;
    for (int pl = 0; pl < (p->p.layout == DAV1D_PIXEL_LAYOUT_I400 ? 1 : 3); pl++) {
        const uint8_t *base = p->data[pl];
        if (!base) continue;
        const ptrdiff_t stride = p->stride[!!pl];
        const int h = pl ? ah >> ss_ver : ah;
        const ptrdiff_t off = (const uint8_t *)ptr - base;
        if (off < 0 || off >= stride * h) continue;
        *plane = pl;
        *y = (int)(off / stride);
        *x = (int)(off % stride) >> hbd;
        return 1;
    }
    return 0;
}

int eav1rec_locate_cur(const Eav1Walk *w, const void *p, int *x, int *y)
{
    int pl;
    if (w->f && locate(&w->f->cur, p, &pl, x, y)) return pl;
    return -1;
}

int eav1rec_locate_sr(const Eav1Walk *w, const void *p, int *x, int *y)
{
    int pl;
    if (w->f && locate(&w->f->sr_cur.p, p, &pl, x, y)) return pl;
    return -1;
}

int eav1rec_locate_ref(const Eav1Walk *w, const void *p, int *plane, int *x, int *y)
{
    const Dav1dFrameContext *f = w->f;
    if (!f) return -1;
    for (int i = 0; i < 7; i++)
        if (f->refp[i].p.data[0] && locate(&f->refp[i].p, p, plane, x, y)) return i;
    if (locate(&f->cur, p, plane, x, y)) return 7;
    return -1;
}

/* ------------------------------------------------------------------ picture allocator
 * Replaces dav1d_default_picture_alloc (ext/dav1d/src/picture.c:46-80): same geometry
 * (128-aligned, planes contiguous Y,U,V, strides never a multiple of 1024), but the host
 * buffer is pinned and paired with a device-resident picture of the engine's pool. A picture
 * stays alive (host buffer and device slot) while a queued job still reads or writes it. */

typedef struct HostBuf { void *p; size_t bytes; } HostBuf;
typedef struct AllocState {
    Eav1Rec *rec;
    pthread_mutex_t mu;
    HostBuf free_list[32];
    int outstanding, closing;
} AllocState;

static void destroy_rec(Eav1Rec *r, AllocState *as);

/* pinned buffers outlive a decoder: the next one opened in this process takes them over */
#define PIN_POOL 2048   /* 64 concurrent streams x (8 references + frames in flight + output queue) */
static HostBuf g_pin_pool[PIN_POOL];

static int pin_pool_get(size_t total, HostBuf *out)
{
    int ok = 0;
    pthread_mutex_lock(&g_mu);
    for (int i = 0; i < PIN_POOL; i++)
        if (g_pin_pool[i].p && g_pin_pool[i].bytes >= total && g_pin_pool[i].bytes <= total * 2) {
            *out = g_pin_pool[i];
            g_pin_pool[i].p = NULL;
            ok = 1;
            break;
        }
    pthread_mutex_unlock(&g_mu);
    return ok;
}

static int pin_pool_put(HostBuf b)
{
    int ok = 0;
    pthread_mutex_lock(&g_mu);
    for (int i = 0; i < PIN_POOL; i++)
        if (!g_pin_pool[i].p) { g_pin_pool[i] = b; ok = 1; break; }
    pthread_mutex_unlock(&g_mu);
    return ok;
}

static int pic_alloc_cb(Dav1dPicture *pic, void *cookie)
{
    AllocState *as = cookie;
    Eav1Rec *r = as->rec;
    const int hbd = pic->p.bpc > 8;
    const int aligned_w = (pic->p.w + 127) & ~127;
    const int aligned_h = (pic->p.h + 127) & ~127;
    const int has_chroma = pic->p.layout != DAV1D_PIXEL_LAYOUT_I400;
    const int ss_ver = pic->p.layout == DAV1D_PIXEL_LAYOUT_I420;
    const int ss_hor = pic->p.layout != DAV1D_PIXEL_LAYOUT_I444;
    ptrdiff_t y_stride = (ptrdiff_t)aligned_w << hbd;
    ptrdiff_t uv_stride = has_chroma ? y_stride >> ss_hor : 0;
    if (!(y_stride & 1023)) y_stride += 64;
    if (!(uv_stride & 1023) && has_chroma) uv_stride += 64;
    pic->stride[0] = y_stride;
    pic->stride[1] = uv_stride;
    const size_t y_sz = (size_t)y_stride * aligned_h;
    const size_t uv_sz = (size_t)uv_stride * (aligned_h >> ss_ver);
    const size_t total = y_sz + 2 * uv_sz + 64;

    Eav1PicData *pd = calloc(1, sizeof(*pd));
    if (!pd) return DAV1D_ERR(ENOMEM);
    pthread_mutex_lock(&as->mu);
    for (int i = 0; i < 32; i++)
        if (as->free_list[i].p && as->free_list[i].bytes >= total && as->free_list[i].bytes <= total * 2) {
            pd->host_base = as->free_list[i].p;
            pd->host_bytes = as->free_list[i].bytes;
            as->free_list[i].p = NULL;
            break;
        }
    pthread_mutex_unlock(&as->mu);
    if (!pd->host_base) {
        HostBuf hb;
        if (pin_pool_get(total, &hb)) { pd->host_base = hb.p; pd->host_bytes = hb.bytes; }
    }
    if (!pd->host_base) {
        if (eav1cu_host_alloc(&pd->host_base, total)) { free(pd); return DAV1D_ERR(ENOMEM); }
        pd->host_bytes = total;
    }
    pd->dev_id = eav1cu_pic_alloc(r->cu, pic->p.w, pic->p.h, pic->p.layout, pic->p.bpc);
    if (pd->dev_id < 0) { eav1cu_host_free(pd->host_base); free(pd); return DAV1D_ERR(ENOMEM); }
    pd->rec = r;
    pd->alloc_state = as;
    atomic_init(&pd->refs, 1);
    atomic_init(&pd->fence, 0);
    atomic_init(&pd->failed, 0);
    uint8_t *d = pd->host_base;
    pic->data[0] = d;
    pic->data[1] = has_chroma ? d + y_sz : NULL;
    pic->data[2] = has_chroma ? d + y_sz + uv_sz : NULL;
    pic->allocator_data = pd;
    pthread_mutex_lock(&as->mu);
    as->outstanding++;
    pthread_mutex_unlock(&as->mu);
    return 0;
}

static void pd_unref(Eav1PicData *pd)
{
    if (atomic_fetch_sub(&pd->refs, 1) != 1) return;
    AllocState *as = pd->alloc_state;
    eav1cu_pic_release(pd->rec->cu, pd->dev_id);
    int kept = 0;
    pthread_mutex_lock(&as->mu);
    for (int i = 0; i < 32; i++)
        if (!as->free_list[i].p) {
            as->free_list[i].p = pd->host_base;
            as->free_list[i].bytes = pd->host_bytes;
            kept = 1;
            break;
        }
    const int last = --as->outstanding == 0 && as->closing;
    pthread_mutex_unlock(&as->mu);
    if (!kept) eav1cu_host_free(pd->host_base);
    Eav1Rec *r = pd->rec;
    free(pd);
    if (last) destroy_rec(r, as);
}

static void pic_release_cb(Dav1dPicture *pic, void *cookie)
{
    (void)cookie;
    if (pic->allocator_data) pd_unref(pic->allocator_data);
}

/* ------------------------------------------------------------------ tables */

static void fill_tables(Eav1Tables *t)
{
    memset(t, 0, sizeof(*t));
    memcpy(t->mc_subpel_filters, dav1d_mc_subpel_filters, sizeof(t->mc_subpel_filters));
    memcpy(t->mc_warp_filter, dav1d_mc_warp_filter, sizeof(t->mc_warp_filter));
    memcpy(t->resize_filter, dav1d_resize_filter, sizeof(t->resize_filter));
    memcpy(t->sm_weights, dav1d_sm_weights, sizeof(t->sm_weights));
    memcpy(t->dr_intra_derivative, dav1d_dr_intra_derivative, sizeof(t->dr_intra_derivative));
    memcpy(t->obmc_masks, dav1d_obmc_masks, sizeof(t->obmc_masks));
    memcpy(t->gaussian_sequence, dav1d_gaussian_sequence, sizeof(t->gaussian_sequence));
    memcpy(t->sgr_x_by_x, dav1d_sgr_x_by_x, sizeof(t->sgr_x_by_x));
    memcpy(t->cdef_directions, dav1d_cdef_directions, sizeof(t->cdef_directions));
    memcpy(t->sgr_params, dav1d_sgr_params, sizeof(t->sgr_params));
    /* filter-intra taps: undo the per-arch interleave of tables.c:747-763 */
    for (int f = 0; f < 5; f++)
        for (int k = 0; k < 8; k++) {
            const int8_t *p = dav1d_filter_intra_taps[f];
#if ARCH_X86
            const int o[7] = { 0, 1, 16, 17, 32, 33, 48 };
            for (int j = 0; j < 7; j++) t->filter_intra_taps[f][k][j] = p[2 * k + o[j]];
#else
            for (int j = 0; j < 7; j++) t->filter_intra_taps[f][k][j] = p[k + 8 * j];
#endif
        }
}

/* the tables exactly as the product uploads them (operator-level tests drive the engine directly) */
void eav1rec_fill_tables(Eav1Tables *t) { fill_tables(t); }
size_t eav1rec_masks(const void **blob) { *blob = &dav1d_masks; return sizeof(dav1d_masks); }

/* ------------------------------------------------------------------ jobs */

struct Eav1Job {
    struct Eav1Job *next;
    unsigned long long seq;
    int skip;                           /* failed / aborted frame: only keeps the order going */
    int no_download;                    /* never shown as is (film grain makes a new picture) */
    Eav1FrameDesc fd;
    Eav1Cmd *cmds; size_t n_cmds, cap_cmds;
    uint8_t *payload; size_t n_payload, cap_payload;
    size_t arena_bytes;
    Eav1LfRec *lf; size_t n_lf, cap_lf;
    Eav1LrRec *lr; size_t n_lr, cap_lr;
    uint8_t *lf_level; size_t cap_level;
    uint8_t *lf_mask; size_t cap_mask;
    Eav1PicData *cur, *ref[7];
    Eav1PicData *coded;                 /* super-resolution: the coded-width picture (cur is the upscaled one) */
    void *cur_data[3]; ptrdiff_t cur_stride[2];
};

static void job_submit(Eav1Rec *r, Eav1Job *j, int digest, int sync_each);
static void job_release(Eav1Rec *r, Eav1Job *j);

static Eav1Job *job_get(Eav1Rec *r)
{
    pthread_mutex_lock(&r->q_mu);
    Eav1Job *j = r->job_free;
    if (j) r->job_free = j->next;
    pthread_mutex_unlock(&r->q_mu);
    if (!j) { j = calloc(1, sizeof(*j)); if (!j) abort(); }
    j->next = NULL; j->skip = 0; j->cur = NULL; j->coded = NULL;
    memset(j->ref, 0, sizeof(j->ref));
    j->n_cmds = j->n_payload = j->n_lf = j->n_lr = 0;
    return j;
}

static void job_free_all(Eav1Job *j)
{
    while (j) {
        Eav1Job *n = j->next;
        free(j->cmds); free(j->payload); free(j->lf); free(j->lr); free(j->lf_level); free(j->lf_mask);
        free(j);
        j = n;
    }
}

static void enqueue(Eav1Rec *r, Eav1Job *j)
{
    pthread_mutex_lock(&r->q_mu);
    Eav1Job **pp = &r->q_head;
    while (*pp && (*pp)->seq < j->seq) pp = &(*pp)->next;
    j->next = *pp;
    *pp = j;
    pthread_cond_broadcast(&r->q_cv);
    pthread_mutex_unlock(&r->q_mu);
}

static Eav1PicData *pd_of(const Dav1dPicture *p)
{
    return p->data[0] ? p->allocator_data : NULL;
}

/* ext/dav1d/src/lib.c:303-309 */
static int has_grain(const Dav1dFrameHeader *hdr)
{
    const Dav1dFilmGrainData *fg = &hdr->film_grain.data;
    return fg->num_y_points || fg->num_uv_points[0] || fg->num_uv_points[1] ||
           (fg->clip_to_restricted_range && fg->chroma_scaling_from_luma);
}

/* Detach everything recorded for the frame in slot f into a job. Runs on whichever thread
 * finished the frame (under dav1d's task lock in threaded mode), so it only moves pointers and
 * copies the two side tables dav1d is about to reuse (lf.level, lf.mask). */
void eav1rec_frame_end(Eav1Rec *r, const Dav1dFrameContext *f, int retval)
{
    Eav1Frame *fr = &r->frames[f - f->c->fc];
    unsigned long long seq = 0;
    int have_seq = 0;
    pthread_mutex_lock(&r->q_mu);
    if (fr->seq_n > 0) {
        seq = fr->seq_q[0];
        memmove(&fr->seq_q[0], &fr->seq_q[1], sizeof(fr->seq_q[0]) * (size_t)(fr->seq_n - 1));
        fr->seq_n--;
        have_seq = 1;
    }
    pthread_mutex_unlock(&r->q_mu);
    if (!have_seq) { fr->open = 0; return; }

    Eav1Job *j = job_get(r);
    j->seq = seq;
    const Dav1dFrameHeader *const hdr = f->frame_hdr;
    /* `cur` is the picture that is shown and referenced: f->sr_cur, which is f->cur itself unless the frame is
     * super-resolved (decode.c:3524-3529); `coded` is where reconstruction, deblocking and CDEF happen */
    const int superres = hdr && hdr->width[0] != hdr->width[1];
    Eav1PicData *coded = pd_of(&f->cur);
    Eav1PicData *cur = superres ? pd_of(&f->sr_cur.p) : coded;
    if (!coded) cur = NULL;
    /* A frame fails on its own: a construct the recorder could not express, or (test hook) EAV1_TEST_FAIL_FRAME.
     * A frame that references a failed picture fails too, as in dav1d (FRAME_ERROR progress, thread_task.c), so the
     * damage ends at the next key frame instead of being handed out as good pictures. */
    int unsupported = atomic_load(&fr->unsupported);
    {
        static int fail_at = -2;
        if (fail_at == -2) { const char *e = getenv("EAV1_TEST_FAIL_FRAME"); fail_at = e ? atoi(e) : -1; }
        if (fail_at >= 0 && (long long)seq == fail_at) unsupported = 1;
    }
    int bad_ref = 0;
    if (hdr && IS_INTER_OR_SWITCH(hdr))
        for (int i = 0; i < 7; i++) {
            Eav1PicData *rp = pd_of(&f->refp[i].p);
            if (rp && atomic_load(&rp->failed)) bad_ref = 1;
        }
    j->skip = retval != 0 || !cur || !hdr || unsupported || bad_ref;
    if (!retval && (unsupported || bad_ref) && cur) {
        if (unsupported)
            fprintf(stderr, "easyav1_b200: frame %llu uses a construct the recorder cannot express (no CPU fallback): dropped\n", seq);
        atomic_store(&cur->failed, 1);
    }
    if (!j->skip) {
        if (!fr->open) frame_begin(fr, f);
        Eav1FrameDesc *fd = &j->fd;
        memset(fd, 0, sizeof(*fd));
        fd->w = f->cur.p.w; fd->h = f->cur.p.h;
        fd->bpc = f->cur.p.bpc;
        fd->layout = f->cur.p.layout;
        fd->bw4 = f->bw; fd->bh4 = f->bh;
        fd->sb128 = f->seq_hdr->sb128;
        fd->cur = coded->dev_id;
        j->cur = cur;
        atomic_fetch_add(&cur->refs, 1);
        atomic_store(&cur->fence, FENCE_PENDING);
        if (superres) {
            j->coded = coded;
            atomic_fetch_add(&coded->refs, 1);
            fd->superres = 1;
            fd->sr_pic = cur->dev_id;
            fd->sr_w = f->sr_cur.p.p.w;
            for (int i = 0; i < 2; i++) { fd->resize_step[i] = f->resize_step[i]; fd->resize_start[i] = f->resize_start[i]; }
        }
        for (int i = 0; i < 7; i++) {
            Eav1PicData *rp = IS_INTER_OR_SWITCH(hdr) ? pd_of(&f->refp[i].p) : NULL;
            fd->ref[i] = rp ? rp->dev_id : -1;
            j->ref[i] = rp;
            if (rp) atomic_fetch_add(&rp->refs, 1);
        }
        for (int i = 0; i < 3; i++) j->cur_data[i] = f->sr_cur.p.data[i];
        j->cur_stride[0] = f->sr_cur.p.stride[0]; j->cur_stride[1] = f->sr_cur.p.stride[1];
        /* output_image (lib.c:311-330) hands out a grain-synthesised copy, never this picture */
        j->no_download = f->c->apply_grain && has_grain(hdr);
        /* command lists: tiles are independent in reconstruction, so any tile order is legal */
        size_t total = 0;
        for (int i = 0; i < fr->n_tiles; i++) total += fr->tiles[i].n;
        if (fr->n_tiles == 1) {
            /* one tile (the usual case): its list becomes the job's by exchanging the two buffers; this runs under
             * dav1d's task lock, where copying a few hundred KB would stall every worker */
            Eav1List *t0 = &fr->tiles[0];
            t0->cmds = grow(t0->cmds, &t0->cap, total + 1, sizeof(Eav1Cmd));
            Eav1Cmd *p = j->cmds; const size_t c = j->cap_cmds;
            j->cmds = t0->cmds; j->cap_cmds = t0->cap;
            t0->cmds = p; t0->cap = c; t0->n = 0;
        } else {
            j->cmds = grow(j->cmds, &j->cap_cmds, total + 1, sizeof(Eav1Cmd));
            size_t pos = 0;
            for (int i = 0; i < fr->n_tiles; i++) {
                memcpy(j->cmds + pos, fr->tiles[i].cmds, fr->tiles[i].n * sizeof(Eav1Cmd));
                pos += fr->tiles[i].n;
            }
        }
        j->n_cmds = total;
        /* payload / lf / lr: swap buffers with the frame slot */
        { uint8_t *p = j->payload; size_t c = j->cap_payload;
          j->payload = fr->payload; j->cap_payload = fr->cap_payload; j->n_payload = atomic_load(&fr->n_payload);
          fr->payload = p; fr->cap_payload = c; }
        { Eav1LfRec *p = j->lf; size_t c = j->cap_lf;
          j->lf = fr->lf; j->cap_lf = fr->cap_lf; j->n_lf = fr->n_lf; fr->lf = p; fr->cap_lf = c; }
        { Eav1LrRec *p = j->lr; size_t c = j->cap_lr;
          j->lr = fr->lr; j->cap_lr = fr->cap_lr; j->n_lr = fr->n_lr; fr->lr = p; fr->cap_lr = c; }
        j->arena_bytes = atomic_load(&fr->arena_used);
        fd->lf_recs = j->lf; fd->n_lf_recs = j->n_lf;
        fd->lr_recs = j->lr; fd->n_lr_recs = j->n_lr;
        fd->b4_stride = (int)f->b4_stride;
        fd->lf_level_bytes = (size_t)f->b4_stride * 32 * f->sb128h * 4;
        if (j->n_lf) {
            j->lf_level = grow(j->lf_level, &j->cap_level, fd->lf_level_bytes, 1);
            memcpy(j->lf_level, f->lf.level, fd->lf_level_bytes);
        }
        fd->lf_level = j->lf_level;
        memcpy(fd->lut_e, f->lf.lim_lut.e, 64);
        memcpy(fd->lut_i, f->lf.lim_lut.i, 64);
        fd->cdef_enable = f->seq_hdr->cdef && (f->c->inloop_filters & DAV1D_INLOOPFILTER_CDEF);
        fd->cdef_damping = hdr->cdef.damping;
        for (int i = 0; i < 8; i++) {
            fd->cdef_y_strength[i] = (uint8_t)hdr->cdef.y_strength[i];
            fd->cdef_uv_strength[i] = (uint8_t)hdr->cdef.uv_strength[i];
        }
        fd->sb128w = f->sb128w; fd->sb128h = f->sb128h;
        if (fd->cdef_enable) {
            const size_t mb = (size_t)f->sb128w * f->sb128h * sizeof(Av1Filter);
            j->lf_mask = grow(j->lf_mask, &j->cap_mask, mb, 1);
            memcpy(j->lf_mask, f->lf.mask, mb);
        }
        fd->lf_mask = j->lf_mask;
    }
    fr->open = 0;
    if (r->inline_submit) {
        /* single-threaded walk: frames end here in decode order on the caller's thread, so the frame goes to the
         * engine right away (no hand-over to the submit thread, no wake-up); everything stays asynchronous on the
         * GPU side and dav1d_get_picture / eav1rec_picture_wait wait for the fence as usual */
        job_submit(r, j, 0, 0);
        pthread_mutex_lock(&r->q_mu);
        r->seq_submit = j->seq + 1;
        pthread_cond_broadcast(&r->q_idle);
        pthread_mutex_unlock(&r->q_mu);
        job_release(r, j);
        pthread_mutex_lock(&r->q_mu);
        j->next = r->job_free;
        r->job_free = j;
        pthread_mutex_unlock(&r->q_mu);
        return;
    }
    enqueue(r, j);
}

/* The submit thread: takes jobs strictly in decode order, hands them to the engine, queues the
 * device->host copy of the finished planes and publishes the fence get_picture waits on. */
static uint64_t fnv(const void *p, size_t n, uint64_t h)
{
    const uint8_t *b = p;
    for (size_t i = 0; i < n; i++) { h ^= b[i]; h *= 1099511628211ULL; }
    return h;
}

/* EAV1_JOB_HASH=1: print a digest of every job (debugging aid for nondeterminism; the filter
 * records are summed because their order legitimately depends on thread timing) */
static void job_digest(const Eav1Job *j)
{
    if (atoi(getenv("EAV1_JOB_HASH")) == 2) {
        if (j->seq < 3) fprintf(stderr, "JOB %llu cmds %zu payload %zu lf %zu lr %zu arena %zu cur %d w %d h %d bw4 %d first-coef-bytes %02x%02x%02x%02x\n", j->seq, j->n_cmds, j->n_payload, j->n_lf, j->n_lr, j->arena_bytes, j->fd.cur, j->fd.w, j->fd.h, j->fd.bw4, j->payload[0], j->payload[1], j->payload[2], j->payload[3]);
        return;
    }
    uint64_t hc = fnv(j->cmds, j->n_cmds * sizeof(Eav1Cmd), 14695981039346656037ULL);
    uint64_t hp = fnv(j->payload, j->n_payload, 14695981039346656037ULL);
    uint64_t hl = 0, hr = 0;
    for (size_t i = 0; i < j->n_lf; i++) hl += fnv(&j->lf[i], sizeof(Eav1LfRec), 14695981039346656037ULL);
    for (size_t i = 0; i < j->n_lr; i++) hr += fnv(&j->lr[i], sizeof(Eav1LrRec), 14695981039346656037ULL);
    uint64_t hv = j->n_lf ? fnv(j->lf_level, j->fd.lf_level_bytes, 1) : 0;
    uint64_t hm = j->fd.cdef_enable ? fnv(j->lf_mask, (size_t)j->fd.sb128w * j->fd.sb128h * 1348, 1) : 0;
    fprintf(stderr, "JOB %llu cmds %zu %016llx payload %zu %016llx lf %zu %016llx lr %zu %016llx lvl %016llx mask %016llx cur %d refs %d %d %d %d %d %d %d arena %zu\n",
            j->seq, j->n_cmds, (unsigned long long)hc, j->n_payload, (unsigned long long)hp, j->n_lf, (unsigned long long)hl,
            j->n_lr, (unsigned long long)hr, (unsigned long long)hv, (unsigned long long)hm, j->fd.cur,
            j->fd.ref[0], j->fd.ref[1], j->fd.ref[2], j->fd.ref[3], j->fd.ref[4], j->fd.ref[5], j->fd.ref[6], j->arena_bytes);
}

/* hand one finished frame to the engine: reconstruction, download of the planes, fence */
static void job_submit(Eav1Rec *r, Eav1Job *j, int digest, int sync_each)
{
    if (j->skip) return;
    if (digest) job_digest(j);
    int res = atomic_load(&r->error);
    pthread_mutex_lock(&r->cu_mu);
    if (!res)
        res = eav1cu_frame_submit(r->cu, &j->fd, j->cmds, j->n_cmds, j->payload, j->n_payload, j->arena_bytes);
    unsigned long long ticket = 0;
    if (!res && !j->no_download)
        res = eav1cu_pic_download_async(r->cu, j->fd.superres ? j->fd.sr_pic : j->fd.cur, j->cur_data, j->cur_stride);
    if (!res) res = eav1cu_fence(r->cu, &ticket);
    if (!res && sync_each) res = eav1cu_sync(r->cu);
    pthread_mutex_unlock(&r->cu_mu);
    if (res) {
        char msg[256] = "";
        eav1cu_last_error(r->cu, msg, sizeof(msg));
        fprintf(stderr, "easyav1_b200: frame reconstruction failed (%d) %s\n", res, msg);
        atomic_store(&r->error, res);
        atomic_store(&j->cur->failed, 1);
    }
    atomic_store(&j->cur->fence, ticket);
    atomic_fetch_add(&r->frames_submitted, 1);
}

/* The pictures may go once the engine has the launches queued: device slots are reused in stream order, and the
 * host buffer of `cur` stays referenced by its Dav1dPicture or, for a frame nobody keeps, is only recycled after
 * its fence. */
static void job_release(Eav1Rec *r, Eav1Job *j)
{
    if (j->skip) return;
    Eav1PicData *cur = j->cur;
    const unsigned long long t = atomic_load(&cur->fence);
    if (atomic_load(&cur->refs) == 1 && t) eav1cu_fence_wait(r->cu, t);
    pd_unref(cur);
    if (j->coded) pd_unref(j->coded);
    for (int i = 0; i < 7; i++) if (j->ref[i]) pd_unref(j->ref[i]);
}

static void *submit_main(void *arg)
{
    Eav1Rec *r = arg;
    const int digest = getenv("EAV1_JOB_HASH") != NULL;
    const int sync_each = getenv("EAV1_SYNC_EACH") != NULL;
    const int timing = getenv("EAV1_TIMING") != NULL;
    double t_idle = 0, t_busy = 0; long n_jobs = 0, n_waited = 0;
    struct timespec ts0, ts1;
    pthread_mutex_lock(&r->q_mu);
    for (;;) {
        if (timing) clock_gettime(CLOCK_MONOTONIC, &ts0);
        int waited = 0;
        while (!r->q_stop && !(r->q_head && r->q_head->seq == r->seq_submit)) {
            waited = 1;
            pthread_cond_wait(&r->q_cv, &r->q_mu);
        }
        if (timing) {
            clock_gettime(CLOCK_MONOTONIC, &ts1);
            t_idle += (ts1.tv_sec - ts0.tv_sec) + 1e-9 * (ts1.tv_nsec - ts0.tv_nsec);
            n_waited += waited;
        }
        if (!(r->q_head && r->q_head->seq == r->seq_submit)) break;   /* stop, nothing runnable */
        Eav1Job *j = r->q_head;
        r->q_head = j->next;
        r->q_busy = 1;
        pthread_mutex_unlock(&r->q_mu);

        job_submit(r, j, digest, sync_each);
        pthread_mutex_lock(&r->q_mu);
        r->seq_submit++;
        r->q_busy = 0;
        pthread_cond_broadcast(&r->q_idle);
        pthread_mutex_unlock(&r->q_mu);
        job_release(r, j);
        if (timing) {
            clock_gettime(CLOCK_MONOTONIC, &ts0);
            t_busy += (ts0.tv_sec - ts1.tv_sec) + 1e-9 * (ts0.tv_nsec - ts1.tv_nsec);
            n_jobs++;
        }
        pthread_mutex_lock(&r->q_mu);
        j->next = r->job_free;
        r->job_free = j;
    }
    pthread_mutex_unlock(&r->q_mu);
    if (timing && n_jobs)
        fprintf(stderr, "EAV1 TIMING: submit thread: %ld jobs, busy %.3f ms/job, idle %.3f ms/job, found the queue empty %ld times\n",
                n_jobs, 1e3 * t_busy / n_jobs, 1e3 * t_idle / n_jobs, n_waited);
    return NULL;
}

/* block until the picture's planes are in host memory; returns 0 or DAV1D_ERR(EIO) */
static int pd_wait(Eav1Rec *r, Eav1PicData *pd)
{
    if (!pd) return 0;
    if (atomic_load(&pd->fence) == FENCE_PENDING) {
        pthread_mutex_lock(&r->q_mu);
        while (atomic_load(&pd->fence) == FENCE_PENDING && !r->q_stop)
            pthread_cond_wait(&r->q_idle, &r->q_mu);
        pthread_mutex_unlock(&r->q_mu);
    }
    const unsigned long long t = atomic_load(&pd->fence);
    if (t && t != FENCE_PENDING && eav1cu_fence_wait(r->cu, t)) return DAV1D_ERR(EIO);
    return atomic_load(&pd->failed) ? DAV1D_ERR(EIO) : 0;
}

/* ------------------------------------------------------------------ film grain
 * host/Makefile renames, in lib.c, the calls dav1d_apply_grain() makes once it has allocated the
 * output picture (lib.c:485-524): dav1d_apply_grain_{8,16}bpc (single-threaded) and
 * dav1d_task_delayed_fg (worker pool) both arrive here. `in` is device-resident once its frame
 * has been handed to the engine; the grain kernels run behind it on the same stream and the new
 * picture comes back like any other (fence on `out`, waited for in dav1d_get_picture). */
static void apply_grain_gpu(Dav1dPicture *const out, const Dav1dPicture *const in)
{
    Eav1PicData *po = pd_of(out), *pi = pd_of(in);
    if (!po || !pi) return;
    Eav1Rec *r = po->rec;
    /* wait until the frame that reconstructs `in` has been queued on the stream */
    if (atomic_load(&pi->fence) == FENCE_PENDING) {
        pthread_mutex_lock(&r->q_mu);
        while (atomic_load(&pi->fence) == FENCE_PENDING && !r->q_stop)
            pthread_cond_wait(&r->q_idle, &r->q_mu);
        pthread_mutex_unlock(&r->q_mu);
    }
    if (atomic_load(&pi->failed) || atomic_load(&r->error)) { atomic_store(&po->failed, 1); return; }
    const Dav1dFilmGrainData *d = &out->frame_hdr->film_grain.data;
    Eav1GrainParams g;
    memset(&g, 0, sizeof(g));
    g.seed = d->seed;
    g.num_y_points = d->num_y_points;
    memcpy(g.y_points, d->y_points, sizeof(g.y_points));
    g.chroma_scaling_from_luma = d->chroma_scaling_from_luma;
    for (int i = 0; i < 2; i++) {
        g.num_uv_points[i] = d->num_uv_points[i];
        g.uv_mult[i] = d->uv_mult[i];
        g.uv_luma_mult[i] = d->uv_luma_mult[i];
        g.uv_offset[i] = d->uv_offset[i];
    }
    memcpy(g.uv_points, d->uv_points, sizeof(g.uv_points));
    g.scaling_shift = d->scaling_shift;
    g.ar_coeff_lag = d->ar_coeff_lag;
    memcpy(g.ar_coeffs_y, d->ar_coeffs_y, sizeof(g.ar_coeffs_y));
    memcpy(g.ar_coeffs_uv, d->ar_coeffs_uv, sizeof(g.ar_coeffs_uv));
    g.ar_coeff_shift = (int32_t)d->ar_coeff_shift;
    g.grain_scale_shift = d->grain_scale_shift;
    g.overlap_flag = d->overlap_flag;
    g.clip_to_restricted_range = d->clip_to_restricted_range;
    g.mtrx_identity = out->seq_hdr->mtrx == DAV1D_MC_IDENTITY;
    unsigned long long ticket = 0;
    pthread_mutex_lock(&r->cu_mu);
    int res = eav1cu_apply_grain(r->cu, pi->dev_id, po->dev_id, &g);
    if (!res) res = eav1cu_pic_download_async(r->cu, po->dev_id, out->data, out->stride);
    if (!res) res = eav1cu_fence(r->cu, &ticket);
    pthread_mutex_unlock(&r->cu_mu);
    if (res) {
        fprintf(stderr, "easyav1_b200: film grain synthesis failed (%d)\n", res);
        atomic_store(&r->error, res);
        atomic_store(&po->failed, 1);
    }
    atomic_store(&po->fence, ticket);
}

void eav1hook_apply_grain_8bpc(const Dav1dFilmGrainDSPContext *const dsp, Dav1dPicture *const out,
                               const Dav1dPicture *const in)
{ (void)dsp; apply_grain_gpu(out, in); }
void eav1hook_apply_grain_16bpc(const Dav1dFilmGrainDSPContext *const dsp, Dav1dPicture *const out,
                                const Dav1dPicture *const in)
{ (void)dsp; apply_grain_gpu(out, in); }
void eav1hook_task_delayed_fg(Dav1dContext *const c, Dav1dPicture *const out, const Dav1dPicture *const in)
{ (void)c; apply_grain_gpu(out, in); }

/* ------------------------------------------------------------------ hooks into the frame life cycle
 * host/Makefile renames the CALLS (not the definitions) of three reference functions:
 *   obu.c:          dav1d_submit_frame       -> eav1hook_submit_frame        (decode.c:3321)
 *   thread_task.c:  dav1d_decode_frame_init  -> eav1hook_decode_frame_init   (decode.c:2750)
 *   thread_task.c,
 *   lib.c:          dav1d_decode_frame_exit  -> eav1hook_decode_frame_exit   (decode.c:3242)
 */

/* Frames enter the decoder here one at a time, in decode order: hand out the order ticket. */
int eav1hook_submit_frame(Dav1dContext *const c)
{
    Eav1Rec *r = eav1rec_for_ctx(c);
    if (!r) return dav1d_submit_frame(c);
    const unsigned idx = c->n_fc > 1 ? c->frame_thread.next : 0;
    Eav1Frame *fr = &r->frames[idx];
    pthread_mutex_lock(&r->q_mu);
    const unsigned long long seq = r->seq_assign++;
    if (fr->seq_n < 4) fr->seq_q[fr->seq_n++] = seq;
    pthread_mutex_unlock(&r->q_mu);
    const int res = dav1d_submit_frame(c);
    /* Single-threaded decoding has finished the frame by now; if it failed before reaching the
     * last filter row (or the threaded path refused the frame) the ticket is still queued. */
    int orphan = 0;
    pthread_mutex_lock(&r->q_mu);
    if (c->n_fc == 1 || res < 0)
        for (int i = 0; i < fr->seq_n; i++)
            if (fr->seq_q[i] == seq) {
                memmove(&fr->seq_q[i], &fr->seq_q[i + 1], sizeof(fr->seq_q[0]) * (size_t)(fr->seq_n - 1 - i));
                fr->seq_n--;
                orphan = 1;
                break;
            }
    pthread_mutex_unlock(&r->q_mu);
    if (orphan) {
        fr->open = 0;
        if (r->inline_submit) {
            pthread_mutex_lock(&r->q_mu);
            r->seq_submit = seq + 1;
            pthread_cond_broadcast(&r->q_idle);
            pthread_mutex_unlock(&r->q_mu);
        } else {
            Eav1Job *j = job_get(r);
            j->seq = seq;
            j->skip = 1;
            enqueue(r, j);
        }
    }
    return res;
}

int eav1hook_decode_frame_init(Dav1dFrameContext *const f)
{
    const int res = dav1d_decode_frame_init(f);
    Eav1Rec *r = eav1rec_for_ctx(f->c);
    if (r && !res) {
        Eav1Frame *fr = &r->frames[f - f->c->fc];
        fr->open = 0;
        frame_begin(fr, f);
    }
    return res;
}

void eav1hook_decode_frame_exit(Dav1dFrameContext *const f, const int retval)
{
    Eav1Rec *r = eav1rec_for_ctx(f->c);
    if (r) eav1rec_frame_end(r, f, retval);
    dav1d_decode_frame_exit(f, retval);
}

/* ------------------------------------------------------------------ open / close */

static void destroy_rec(Eav1Rec *r, AllocState *as)
{
    if (!r) return;
    pthread_mutex_lock(&g_mu);
    for (int i = 0; i < MAX_RECS; i++) if (g_recs[i] == r) g_recs[i] = NULL;
    atomic_fetch_add_explicit(&g_rec_generation, 1, memory_order_release);
    pthread_mutex_unlock(&g_mu);
    if (eav1_walk.r == r) eav1_walk.r = NULL;
    if (as) {
        for (int i = 0; i < 32; i++)
            if (as->free_list[i].p && !pin_pool_put(as->free_list[i])) eav1cu_host_free(as->free_list[i].p);
        pthread_mutex_destroy(&as->mu);
        free(as);
    }
    if (r->cu) eav1cu_ctx_destroy(r->cu);
    for (int i = 0; i < r->n_frames; i++) {
        Eav1Frame *fr = &r->frames[i];
        if (fr->tiles) for (int t = 0; t < EAV1_MAX_TILES; t++) free(fr->tiles[t].cmds);
        free(fr->tiles); free(fr->payload); free(fr->lf); free(fr->lr);
        pthread_mutex_destroy(&fr->mu);
    }
    free(r->frames);
    job_free_all(r->q_head);
    job_free_all(r->job_free);
    pthread_mutex_destroy(&r->q_mu);
    pthread_cond_destroy(&r->q_cv);
    pthread_cond_destroy(&r->q_idle);
    free(r);
}

static void stop_submit_thread(Eav1Rec *r)
{
    if (!r->have_thread) return;
    pthread_mutex_lock(&r->q_mu);
    r->q_stop = 1;
    pthread_cond_broadcast(&r->q_cv);
    pthread_cond_broadcast(&r->q_idle);
    pthread_mutex_unlock(&r->q_mu);
    pthread_join(r->submit_thread, NULL);
    r->have_thread = 0;
}

/* Drop-in for dav1d_open (ext/dav1d/src/lib.c:144): same signature and error codes. The GPU
 * ordinal comes from EASYAV1_CUDA_DEVICE (default 0) so easyav1.h stays unchanged. Threading is
 * dav1d's own: n_threads / max_frame_delay keep their meaning (lib.c:100-142), the workers walk
 * frames and tiles concurrently and only record; EASYAV1_THREADS overrides n_threads. */
int dav1d_open(Dav1dContext **c_out, const Dav1dSettings *s)
{
    if (!c_out || !s) return DAV1D_ERR(EINVAL);
    Eav1Rec *r = calloc(1, sizeof(*r));
    AllocState *as = calloc(1, sizeof(*as));
    if (!r || !as) { free(r); free(as); return DAV1D_ERR(ENOMEM); }
    pthread_mutex_init(&as->mu, NULL);
    pthread_mutex_init(&r->q_mu, NULL);
    pthread_mutex_init(&r->cu_mu, NULL);
    pthread_cond_init(&r->q_cv, NULL);
    pthread_cond_init(&r->q_idle, NULL);
    as->rec = r;
    const char *dev = getenv("EASYAV1_CUDA_DEVICE");
    if (eav1cu_ctx_create(&r->cu, dev ? atoi(dev) : 0)) {
        fprintf(stderr, "easyav1_b200: no usable CUDA device (there is no CPU fallback)\n");
        destroy_rec(r, as);
        return DAV1D_ERR(ENODEV);
    }
    Dav1dSettings ss = *s;
    const char *thr = getenv("EASYAV1_THREADS");
    if (thr) ss.n_threads = atoi(thr);
    const char *fdl = getenv("EASYAV1_FRAME_DELAY");
    if (fdl) ss.max_frame_delay = atoi(fdl);
    /* tile threads without frame threads would end frames inside decode.c where no hook
     * reaches; a frame delay of 1 therefore means the plain single-threaded walk */
    if (ss.max_frame_delay == 1) ss.n_threads = 1;
    ss.allocator.cookie = as;
    ss.allocator.alloc_picture_callback = pic_alloc_cb;
    ss.allocator.release_picture_callback = pic_release_cb;
    /* the recorder must exist before the workers start: register under a placeholder first */
    const int res = eav1ref_dav1d_open(c_out, &ss);
    if (res < 0) { destroy_rec(r, as); return res; }
    Dav1dContext *c = *c_out;
    if (c->n_fc == 1 && c->n_tc > 1) {   /* cannot happen with dav1d's n_fc rule; be safe */
        eav1ref_dav1d_close(c_out);
        destroy_rec(r, as);
        return DAV1D_ERR(EINVAL);
    }
    r->c = c;
    r->threaded = c->n_fc > 1;
    r->inline_submit = !r->threaded && !getenv("EASYAV1_NO_INLINE_SUBMIT");
    r->n_frames = (int)c->n_fc;
    r->frames = calloc((size_t)r->n_frames, sizeof(Eav1Frame));
    if (!r->frames) { eav1ref_dav1d_close(c_out); destroy_rec(r, as); return DAV1D_ERR(ENOMEM); }
    for (int i = 0; i < r->n_frames; i++) {
        pthread_mutex_init(&r->frames[i].mu, NULL);
        r->frames[i].tiles = calloc(EAV1_MAX_TILES, sizeof(Eav1List));
        if (!r->frames[i].tiles) { eav1ref_dav1d_close(c_out); destroy_rec(r, as); return DAV1D_ERR(ENOMEM); }
    }
    Eav1Tables tab;
    fill_tables(&tab);
    if (eav1cu_set_tables(r->cu, &tab) ||
        eav1cu_set_masks(r->cu, (const uint8_t *)&dav1d_masks, sizeof(dav1d_masks))) {
        eav1ref_dav1d_close(c_out);
        destroy_rec(r, as);
        return DAV1D_ERR(EIO);
    }
    pthread_mutex_lock(&g_mu);
    int ok = 0;
    for (int i = 0; i < MAX_RECS; i++) if (!g_recs[i]) { g_recs[i] = r; ok = 1; break; }
    pthread_mutex_unlock(&g_mu);
    if (ok && pthread_create(&r->submit_thread, NULL, submit_main, r) == 0) r->have_thread = 1;
    if (!ok || !r->have_thread) { eav1ref_dav1d_close(c_out); destroy_rec(r, as); return DAV1D_ERR(ENOMEM); }
    /* remember the allocator state for close */
    r->alloc_state = as;
    return 0;
}

/* Drop-in for dav1d_close (ext/dav1d/src/lib.c:610). Pictures still held by the caller keep
 * the allocator state and the device context alive until they are released. */
void dav1d_close(Dav1dContext **c_out)
{
    if (!c_out || !*c_out) return;
    Eav1Rec *r = eav1rec_for_ctx(*c_out);
    eav1ref_dav1d_close(c_out);     /* flushes: frames in flight end through the exit hook */
    if (!r) return;
    stop_submit_thread(r);
    AllocState *as = r->alloc_state;
    pthread_mutex_lock(&as->mu);
    as->closing = 1;
    const int busy = as->outstanding > 0;
    pthread_mutex_unlock(&as->mu);
    if (!busy) destroy_rec(r, as);
}

/* Drop-in for dav1d_flush (ext/dav1d/src/lib.c:526; easyAV1 calls it on every seek, src/easyav1.c:3059). The
 * reference's flush ends the frames in flight (they arrive as skipped jobs); afterwards the queue is drained and an
 * engine error that was not a CUDA failure is forgotten, so that decoding can restart at the next key frame. */
void dav1d_flush(Dav1dContext *c)
{
    eav1ref_dav1d_flush(c);
    Eav1Rec *r = eav1rec_for_ctx(c);
    if (!r) return;
    pthread_mutex_lock(&r->q_mu);
    while ((r->q_head || r->q_busy) && !r->q_stop) pthread_cond_wait(&r->q_idle, &r->q_mu);
    pthread_mutex_unlock(&r->q_mu);
    if (atomic_load(&r->error) == EAV1CU_ERR_INVAL && eav1cu_clear_error(r->cu) == 0) atomic_store(&r->error, 0);
}

/* for tools / tests: how many frames went through the engine and whether it failed */
long long eav1rec_frames_submitted(const Dav1dContext *c) { Eav1Rec *r = eav1rec_for_ctx(c); return r ? atomic_load(&r->frames_submitted) : -1; }
int eav1rec_error(const Dav1dContext *c) { Eav1Rec *r = eav1rec_for_ctx(c); return r ? atomic_load(&r->error) : -1; }
/* the engine context behind a Dav1dContext, for measurement (bench.py) */
Eav1cuCtx *eav1rec_engine(const Dav1dContext *c) { Eav1Rec *r = eav1rec_for_ctx(c); return r ? r->cu : NULL; }
long long eav1rec_launches(const Dav1dContext *c) { Eav1Rec *r = eav1rec_for_ctx(c); return r ? eav1cu_launch_count(r->cu) : -1; }
/* wait until every frame handed over so far has been given to the engine */
void eav1rec_drain(const Dav1dContext *c)
{
    Eav1Rec *r = eav1rec_for_ctx(c);
    if (!r) return;
    pthread_mutex_lock(&r->q_mu);
    while ((r->q_head || r->q_busy) && !r->q_stop) pthread_cond_wait(&r->q_idle, &r->q_mu);
    pthread_mutex_unlock(&r->q_mu);
    eav1cu_sync(r->cu);
}

/* For a caller that multiplexes many decoders on few threads (host/pump.c): dav1d_get_picture without the wait for
 * the planes, and the wait as a call of its own. The picture may be passed on, referenced and released as usual;
 * only its sample data must not be read before eav1rec_picture_wait returned 0. */
int eav1rec_get_picture_async(Dav1dContext *c, Dav1dPicture *out)
{
    const int res = eav1ref_dav1d_get_picture(c, out);
    Eav1Rec *r = eav1rec_for_ctx(c);
    if (r && res == 0 && atomic_load(&r->error)) { dav1d_picture_unref(out); return DAV1D_ERR(EIO); }
    return res;
}

int eav1rec_picture_wait(Dav1dPicture *p)
{
    Eav1PicData *pd = p && p->data[0] ? p->allocator_data : NULL;
    if (!pd) return 0;
    return pd_wait(pd->rec, pd);
}

/* SURVEY.md N4 — an ADDITIONAL getter beside dav1d_get_picture: the picture `p` (as handed out by dav1d_get_picture,
 * still referenced by the caller) converted on the device to NV12 / P010 / RGBA8 (include/easyav1_cuda.h,
 * eav1cu_pic_convert) and copied to dst. matrix < 0 / full_range < 0 take the values of the picture's sequence header
 * (what src/easyav1.c:3287-3507 reports to the application). Returns 0 or a negative DAV1D_ERR. */
int eav1rec_picture_convert(const Dav1dPicture *p, int format, int matrix, int full_range, void *dst, size_t dst_stride, size_t dst_bytes)
{
    Eav1PicData *pd = p && p->data[0] ? p->allocator_data : NULL;
    if (!pd || !dst) return DAV1D_ERR(EINVAL);
    Eav1Rec *r = pd->rec;
    const int w = pd_wait(r, pd);               /* its frame has been reconstructed (and it did not fail) */
    if (w) return w;
    Eav1Convert cv;
    cv.format = format;
    cv.matrix = matrix >= 0 ? matrix : (p->seq_hdr ? (int)p->seq_hdr->mtrx : 2);
    cv.full_range = full_range >= 0 ? full_range : (p->seq_hdr ? p->seq_hdr->color_range : 0);
    pthread_mutex_lock(&r->cu_mu);
    const int res = eav1cu_pic_convert(r->cu, pd->dev_id, &cv, dst, dst_stride, dst_bytes);
    pthread_mutex_unlock(&r->cu_mu);
    return res == 0 ? 0 : res == EAV1CU_ERR_INVAL ? DAV1D_ERR(EINVAL) : DAV1D_ERR(EIO);
}

/* Drop-in for dav1d_get_picture (ext/dav1d/src/lib.c:458): the picture dav1d hands out was
 * reconstructed asynchronously; wait for its planes to arrive in (pinned) host memory. The DSP
 * surface we replace is void, so engine failures are latched and surface here as DAV1D_ERR(EIO);
 * easyAV1 maps that to EASYAV1_STATUS_DECODER_ERROR (src/easyav1.c:2399-2412). No CPU fallback. */
int dav1d_get_picture(Dav1dContext *c, Dav1dPicture *out)
{
    const int res = eav1ref_dav1d_get_picture(c, out);
    Eav1Rec *r = eav1rec_for_ctx(c);
    if (!r) return res;
    if (res == 0 && out->data[0]) {
        const int w = pd_wait(r, out->allocator_data);
        if (w) { dav1d_picture_unref(out); return w; }
    }
    if (res == 0 && atomic_load(&r->error)) { dav1d_picture_unref(out); return DAV1D_ERR(EIO); }
    return res;
}
