/*
 * rec_common.c — bit-depth independent part of the recorder: decoder open/close wrappers,
 * pinned picture allocator tied to the device picture pool, constant-table upload, list
 * management and the per-frame submit. See rec.h.
 */
#include <errno.h>
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "rec.h"
#include "common/frame.h"
#include "src/tables.h"
#include "src/wedge.h"

/* the reference's own entry points, renamed at compile time (host/Makefile) */
int  eav1ref_dav1d_open(Dav1dContext **c_out, const Dav1dSettings *s);
void eav1ref_dav1d_close(Dav1dContext **c_out);

/* pinned host memory comes from the engine so this library needs no CUDA headers */
int  eav1cu_host_alloc(void **p, size_t bytes);
void eav1cu_host_free(void *p);

__thread Eav1Rec *eav1_tls_rec;

#define MAX_RECS 256
static pthread_mutex_t g_mu = PTHREAD_MUTEX_INITIALIZER;
static Eav1Rec *g_recs[MAX_RECS];

Eav1Rec *eav1rec_for_ctx(const Dav1dContext *c)
{
    Eav1Rec *r = eav1_tls_rec;
    if (r && r->c == c) return r;
    pthread_mutex_lock(&g_mu);
    r = NULL;
    for (int i = 0; i < MAX_RECS; i++)
        if (g_recs[i] && g_recs[i]->c == c) { r = g_recs[i]; break; }
    pthread_mutex_unlock(&g_mu);
    eav1_tls_rec = r;
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

Eav1Cmd *eav1rec_push(Eav1Rec *r)
{
    r->cmds = grow(r->cmds, &r->cap_cmds, r->n_cmds + 1, sizeof(Eav1Cmd));
    Eav1Cmd *c = &r->cmds[r->n_cmds++];
    memset(c, 0, sizeof(*c));
    if (r->block_head) { c->flags |= EAV1_F_BLOCK_HEAD; r->block_head = 0; }
    return c;
}

uint32_t eav1rec_payload(Eav1Rec *r, const void *src, size_t bytes, size_t align)
{
    size_t off = (r->n_payload + align - 1) & ~(align - 1);
    r->payload = grow(r->payload, &r->cap_payload, off + bytes, 1);
    memcpy(r->payload + off, src, bytes);
    r->n_payload = off + bytes;
    return (uint32_t)off;
}

uint32_t eav1rec_region(Eav1Rec *r, int reg, size_t elem_bytes)
{
    if (!r->region_live[reg]) {
        size_t off = (r->arena_used + 15) & ~(size_t)15;
        r->region_off[reg] = (uint32_t)off;
        r->arena_used = off + (size_t)r->blk_w * r->blk_h * elem_bytes;
        r->region_live[reg] = 1;
    }
    return r->region_off[reg];
}

/* A producer starts a fresh region every time (single assignment): no command ever overwrites
 * scratch that an earlier command of the frame still has to read, so the engine is free to run
 * all motion compensation of a frame before anything else. */
uint32_t eav1rec_region_new(Eav1Rec *r, int reg, size_t bytes)
{
    size_t off = (r->arena_used + 15) & ~(size_t)15;
    r->region_off[reg] = (uint32_t)off;
    r->arena_used = off + bytes;
    r->region_live[reg] = 1;
    return (uint32_t)off;
}

Eav1LfRec *eav1rec_push_lf(Eav1Rec *r)
{
    r->lf = grow(r->lf, &r->cap_lf, r->n_lf + 1, sizeof(Eav1LfRec));
    return &r->lf[r->n_lf++];
}

Eav1LrRec *eav1rec_push_lr(Eav1Rec *r)
{
    r->lr = grow(r->lr, &r->cap_lr, r->n_lr + 1, sizeof(Eav1LrRec));
    Eav1LrRec *x = &r->lr[r->n_lr++];
    memset(x, 0, sizeof(*x));
    return x;
}

void eav1rec_block_begin(Eav1Rec *r, const Dav1dTaskContext *t, int bs)
{
    r->f = t->f;
    r->t = t;
    r->frame_open = 1;
    r->block_head = 1;
    r->blk_w = dav1d_block_dimensions[bs][0] * 4;
    r->blk_h = dav1d_block_dimensions[bs][1] * 4;
    memset(r->region_live, 0, sizeof(r->region_live));
    r->emu.valid = 0;
    r->edge.valid = 0;
    r->ac.valid = 0;
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

int eav1rec_locate_cur(const Eav1Rec *r, const void *p, int *x, int *y)
{
    int pl;
    if (r->f && locate(&r->f->cur, p, &pl, x, y)) return pl;
    return -1;
}

int eav1rec_locate_ref(const Eav1Rec *r, const void *p, int *plane, int *x, int *y)
{
    const Dav1dFrameContext *f = r->f;
    if (!f) return -1;
    for (int i = 0; i < 7; i++)
        if (f->refp[i].p.data[0] && locate(&f->refp[i].p, p, plane, x, y)) return i;
    if (locate(&f->cur, p, plane, x, y)) return 7;
    return -1;
}

/* ------------------------------------------------------------------ picture allocator
 * Replaces dav1d_default_picture_alloc (ext/dav1d/src/picture.c:46-80): same geometry
 * (128-aligned, planes contiguous Y,U,V, strides never a multiple of 1024), but the host
 * buffer is pinned and paired with a device-resident picture of the engine's pool. */

typedef struct HostBuf { void *p; size_t bytes; } HostBuf;
typedef struct AllocState {
    Eav1Rec *rec;
    pthread_mutex_t mu;
    HostBuf free_list[32];
    int outstanding, closing;
} AllocState;

static void destroy_rec(Eav1Rec *r, AllocState *as);

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
        if (eav1cu_host_alloc(&pd->host_base, total)) { free(pd); return DAV1D_ERR(ENOMEM); }
        pd->host_bytes = total;
    }
    pd->dev_id = eav1cu_pic_alloc(r->cu, pic->p.w, pic->p.h, pic->p.layout, pic->p.bpc);
    if (pd->dev_id < 0) { eav1cu_host_free(pd->host_base); free(pd); return DAV1D_ERR(ENOMEM); }
    pd->rec = r;
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

static void pic_release_cb(Dav1dPicture *pic, void *cookie)
{
    AllocState *as = cookie;
    Eav1PicData *pd = pic->allocator_data;
    if (!pd) return;
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

/* ------------------------------------------------------------------ open / close */

static void destroy_rec(Eav1Rec *r, AllocState *as)
{
    if (!r) return;
    pthread_mutex_lock(&g_mu);
    for (int i = 0; i < MAX_RECS; i++) if (g_recs[i] == r) g_recs[i] = NULL;
    pthread_mutex_unlock(&g_mu);
    if (eav1_tls_rec == r) eav1_tls_rec = NULL;
    if (as) {
        for (int i = 0; i < 32; i++) if (as->free_list[i].p) eav1cu_host_free(as->free_list[i].p);
        pthread_mutex_destroy(&as->mu);
        free(as);
    }
    if (r->cu) eav1cu_ctx_destroy(r->cu);
    free(r->cmds); free(r->payload); free(r->lf); free(r->lr);
    free(r);
}

/* Drop-in for dav1d_open (ext/dav1d/src/lib.c:144): same signature and error codes. The GPU
 * ordinal comes from EASYAV1_CUDA_DEVICE (default 0) so easyav1.h stays unchanged. */
int dav1d_open(Dav1dContext **c_out, const Dav1dSettings *s)
{
    if (!c_out || !s) return DAV1D_ERR(EINVAL);
    Eav1Rec *r = calloc(1, sizeof(*r));
    AllocState *as = calloc(1, sizeof(*as));
    if (!r || !as) { free(r); free(as); return DAV1D_ERR(ENOMEM); }
    pthread_mutex_init(&as->mu, NULL);
    as->rec = r;
    const char *dev = getenv("EASYAV1_CUDA_DEVICE");
    if (eav1cu_ctx_create(&r->cu, dev ? atoi(dev) : 0)) {
        fprintf(stderr, "easyav1_b200: no usable CUDA device (there is no CPU fallback)\n");
        destroy_rec(r, as);
        return DAV1D_ERR(ENODEV);
    }
    Dav1dSettings ss = *s;
    ss.n_threads = 1;       /* host walk is serial per stream; streams scale across threads/GPUs */
    ss.max_frame_delay = 1;
    ss.allocator.cookie = as;
    ss.allocator.alloc_picture_callback = pic_alloc_cb;
    ss.allocator.release_picture_callback = pic_release_cb;
    const int res = eav1ref_dav1d_open(c_out, &ss);
    if (res < 0) { destroy_rec(r, as); return res; }
    r->c = *c_out;
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
    if (!ok) { eav1ref_dav1d_close(c_out); destroy_rec(r, as); return DAV1D_ERR(ENOMEM); }
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
    eav1ref_dav1d_close(c_out);
    if (!r) return;
    AllocState *as = r->alloc_state;
    pthread_mutex_lock(&as->mu);
    as->closing = 1;
    const int busy = as->outstanding > 0;
    pthread_mutex_unlock(&as->mu);
    if (!busy) destroy_rec(r, as);
}

/* ------------------------------------------------------------------ frame submit */

static int dev_id_of(const Dav1dPicture *p)
{
    const Eav1PicData *pd = p->allocator_data;
    return pd && p->data[0] ? pd->dev_id : -1;
}

/* Called after the host walk and the lf/lr sbrow drivers of the last superblock row ran.
 * Ships the frame's lists to the engine and brings the finished planes back into the
 * Dav1dPicture the application will see (ext/dav1d/include/dav1d/picture.h). */
int eav1rec_frame_end(Eav1Rec *r, const Dav1dFrameContext *f)
{
    const Dav1dFrameHeader *const hdr = f->frame_hdr;
    Eav1FrameDesc fd;
    memset(&fd, 0, sizeof(fd));
    fd.w = f->cur.p.w; fd.h = f->cur.p.h;
    fd.bpc = f->cur.p.bpc;
    fd.layout = f->cur.p.layout;
    fd.bw4 = f->bw; fd.bh4 = f->bh;
    fd.sb128 = f->seq_hdr->sb128;
    fd.cur = dev_id_of(&f->cur);
    for (int i = 0; i < 7; i++)
        fd.ref[i] = IS_INTER_OR_SWITCH(hdr) ? dev_id_of(&f->refp[i].p) : -1;
    fd.lf_recs = r->lf; fd.n_lf_recs = r->n_lf;
    fd.lf_level = (const uint8_t *)f->lf.level;
    fd.b4_stride = (int)f->b4_stride;
    fd.lf_level_bytes = (size_t)f->b4_stride * 32 * f->sb128h * 4;
    memcpy(fd.lut_e, f->lf.lim_lut.e, 64);
    memcpy(fd.lut_i, f->lf.lim_lut.i, 64);
    fd.cdef_enable = f->seq_hdr->cdef && (f->c->inloop_filters & DAV1D_INLOOPFILTER_CDEF);
    fd.cdef_damping = hdr->cdef.damping;
    for (int i = 0; i < 8; i++) {
        fd.cdef_y_strength[i] = (uint8_t)hdr->cdef.y_strength[i];
        fd.cdef_uv_strength[i] = (uint8_t)hdr->cdef.uv_strength[i];
    }
    fd.lf_mask = f->lf.mask;
    fd.sb128w = f->sb128w; fd.sb128h = f->sb128h;
    fd.lr_recs = r->lr; fd.n_lr_recs = r->n_lr;

    int res = 0;
    if (hdr->width[0] != hdr->width[1]) { r->unsupported = 1; res = EAV1CU_ERR_INVAL; } /* super-res: not yet */
    if (!res && r->unsupported) res = EAV1CU_ERR_INVAL;
    if (!res)
        res = eav1cu_frame_submit(r->cu, &fd, r->cmds, r->n_cmds, r->payload, r->n_payload, r->arena_used);
    if (!res) {
        void *data[3] = { f->cur.data[0], f->cur.data[1], f->cur.data[2] };
        res = eav1cu_pic_download(r->cu, fd.cur, data, f->cur.stride);
    }
    if (res) {
        char msg[256] = "";
        eav1cu_last_error(r->cu, msg, sizeof(msg));
        fprintf(stderr, "easyav1_b200: frame reconstruction failed (%d) %s%s\n", res, msg,
                r->unsupported ? " [unsupported coding tool in stream]" : "");
        r->error = res;
    }
    r->n_cmds = r->n_payload = r->n_lf = r->n_lr = 0;
    r->arena_used = 0;
    r->frame_open = 0;
    r->frames_submitted++;
    return res;
}

/* for tools / tests: how many frames went through the engine and whether it failed */
long long eav1rec_frames_submitted(const Dav1dContext *c) { Eav1Rec *r = eav1rec_for_ctx(c); return r ? r->frames_submitted : -1; }
int eav1rec_error(const Dav1dContext *c) { Eav1Rec *r = eav1rec_for_ctx(c); return r ? r->error : -1; }
/* the engine context behind a Dav1dContext, for measurement (bench.py) */
Eav1cuCtx *eav1rec_engine(const Dav1dContext *c) { Eav1Rec *r = eav1rec_for_ctx(c); return r ? r->cu : NULL; }
long long eav1rec_launches(const Dav1dContext *c) { Eav1Rec *r = eav1rec_for_ctx(c); return r ? eav1cu_launch_count(r->cu) : -1; }

/* Drop-in for dav1d_get_picture (ext/dav1d/src/lib.c:458): the DSP surface we replace is void,
 * so engine failures are latched by the recorder and surfaced here as DAV1D_ERR(EIO); easyAV1
 * maps that to EASYAV1_STATUS_DECODER_ERROR (src/easyav1.c:2399-2412). No CPU fallback. */
int eav1ref_dav1d_get_picture(Dav1dContext *c, Dav1dPicture *out);
int dav1d_get_picture(Dav1dContext *c, Dav1dPicture *out)
{
    const int res = eav1ref_dav1d_get_picture(c, out);
    Eav1Rec *r = eav1rec_for_ctx(c);
    if (r && r->error) {
        if (res == 0) dav1d_picture_unref(out);
        return DAV1D_ERR(EIO);
    }
    return res;
}
