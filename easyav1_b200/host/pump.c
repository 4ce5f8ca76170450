/*
 * pump.c — many streams on few threads (the host pipeline around the path; the packet fan-out
 * of src/easyav1.c:1904-2132 / 2271-2333 for S decoders at once).
 *
 * A server that decodes S streams on one GPU must not run S x n threads: with the reconstruction on the device a
 * stream's host thread spends a good part of every frame waiting for the GPU, and 64 such threads on 16 cores
 * mostly preempt each other while holding locks. The pump runs W worker threads (about one per core). Every
 * stream is a single-threaded dav1d decoder (n_threads = 1: the entropy decode and the block walk of a frame run
 * on the worker that feeds it, and that worker hands the finished frame to the engine itself, rec_common.c
 * inline_submit). A worker takes the next stream that is not being served, and either
 *   - feeds it until one more picture comes out (dav1d_send_data / eav1rec_get_picture_async: the picture is
 *     returned while the GPU still works on it), or
 *   - retires its oldest pending picture (eav1rec_picture_wait, copy the visible planes out, unref)
 * when the stream already has DEPTH pictures in flight or has no input left. The GPU latency of one stream is
 * thereby hidden behind the entropy decode of the others, and no thread ever spins.
 *
 * Output: for every stream the visible Y|U|V planes of each picture are packed into that stream's buffer (each
 * picture overwrites the previous one, like eav1_stream_next), optionally passed to a callback first.
 */
#include <errno.h>
#include <pthread.h>
#include <stdatomic.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "config.h"
#include "dav1d/dav1d.h"

int eav1rec_get_picture_async(Dav1dContext *c, Dav1dPicture *out);
int eav1rec_picture_wait(Dav1dPicture *p);

#define PUMP_DEPTH 3

typedef void (*eav1_pump_cb)(void *user, int stream, int frame, const Dav1dPicture *pic);

typedef struct PumpStream {
    Dav1dContext *c;
    Dav1dData data;
    Dav1dPicture pend[PUMP_DEPTH];
    int n_pend, head;
    int drained, finished, error;
    int frames;
    uint8_t *out; size_t cap, used;
    pthread_mutex_t mu;
} PumpStream;

typedef struct Eav1Pump {
    PumpStream *st; int n_streams;
    atomic_int next, n_finished;
    eav1_pump_cb cb; void *user;
} Eav1Pump;

static void no_free(const uint8_t *d, void *cookie) { (void)d; (void)cookie; }
static void quiet_log(void *cookie, const char *fmt, va_list ap) { (void)cookie; (void)fmt; (void)ap; }

static void pack_out(PumpStream *s, const Dav1dPicture *p)
{
    const int hbd = p->p.bpc > 8;
    const int ss_hor = p->p.layout != DAV1D_PIXEL_LAYOUT_I444;
    const int ss_ver = p->p.layout == DAV1D_PIXEL_LAYOUT_I420;
    size_t n = 0;
    for (int pl = 0; pl < (p->p.layout == DAV1D_PIXEL_LAYOUT_I400 ? 1 : 3); pl++) {
        const int w = pl ? (p->p.w + ss_hor) >> ss_hor : p->p.w;
        const int h = pl ? (p->p.h + ss_ver) >> ss_ver : p->p.h;
        const size_t row = (size_t)w << hbd;
        const uint8_t *src = p->data[pl];
        for (int y = 0; y < h; y++) {
            if (s->out && n + row <= s->cap) memcpy(s->out + n, src, row);
            n += row;
            src += p->stride[!!pl];
        }
    }
    s->used = n;
}

static void retire_one(Eav1Pump *pm, int si)
{
    PumpStream *s = &pm->st[si];
    Dav1dPicture *p = &s->pend[s->head];
    const int res = eav1rec_picture_wait(p);
    if (res < 0) s->error = res;
    else {
        if (pm->cb) pm->cb(pm->user, si, s->frames, p);
        pack_out(s, p);
        s->frames++;
    }
    dav1d_picture_unref(p);
    s->head = (s->head + 1) % PUMP_DEPTH;
    s->n_pend--;
}

/* one unit of work for stream si; the caller holds its lock */
static void step(Eav1Pump *pm, int si)
{
    PumpStream *s = &pm->st[si];
    if (s->n_pend == PUMP_DEPTH || ((s->drained || s->error) && s->n_pend)) { retire_one(pm, si); return; }
    if (s->drained || s->error) {
        s->finished = 1;
        atomic_fetch_add(&pm->n_finished, 1);
        return;
    }
    Dav1dPicture *p = &s->pend[(s->head + s->n_pend) % PUMP_DEPTH];
    for (;;) {
        memset(p, 0, sizeof(*p));
        int res;
        if (s->data.sz > 0) {
            res = dav1d_send_data(s->c, &s->data);
            if (res < 0 && res != DAV1D_ERR(EAGAIN)) { s->error = res; return; }
        }
        res = eav1rec_get_picture_async(s->c, p);
        if (res == 0) { s->n_pend++; return; }
        if (res != DAV1D_ERR(EAGAIN)) { s->error = res; return; }
        if (s->data.sz == 0) {
            res = eav1rec_get_picture_async(s->c, p);
            if (res == 0) { s->n_pend++; return; }
            if (res != DAV1D_ERR(EAGAIN)) s->error = res;
            s->drained = 1;
            return;
        }
    }
}

static void *worker(void *arg)
{
    Eav1Pump *pm = arg;
    int idle = 0;
    while (atomic_load(&pm->n_finished) < pm->n_streams) {
        const int si = (int)((unsigned)atomic_fetch_add(&pm->next, 1) % (unsigned)pm->n_streams);
        PumpStream *s = &pm->st[si];
        if (s->finished || pthread_mutex_trylock(&s->mu)) {
            if (++idle >= pm->n_streams) { sched_yield(); idle = 0; }
            continue;
        }
        idle = 0;
        if (!s->finished) step(pm, si);
        pthread_mutex_unlock(&s->mu);
    }
    return NULL;
}

/* Decode n_streams OBU buffers to the end on n_workers threads. out[i] (cap bytes each, may be NULL) receives the
 * packed planes of stream i's pictures, frames[i] their number, used[i] the size of the last one. Returns 0 or
 * the first decoder error. apply_grain / inloop_filters as in Dav1dSettings. */
int eav1_pump_run(const uint8_t *const *obu, const size_t *sz, int n_streams, int n_workers, int apply_grain,
                  int inloop_filters, uint8_t *const *out, size_t cap, int *frames, size_t *used,
                  eav1_pump_cb cb, void *user)
{
    if (n_streams <= 0 || n_workers <= 0) return DAV1D_ERR(EINVAL);
    Eav1Pump pm;
    memset(&pm, 0, sizeof(pm));
    pm.st = calloc((size_t)n_streams, sizeof(PumpStream));
    if (!pm.st) return DAV1D_ERR(ENOMEM);
    pm.n_streams = n_streams; pm.cb = cb; pm.user = user;
    int err = 0, opened = 0;
    for (int i = 0; i < n_streams && !err; i++, opened++) {
        PumpStream *s = &pm.st[i];
        pthread_mutex_init(&s->mu, NULL);
        Dav1dSettings st;
        dav1d_default_settings(&st);
        st.n_threads = 1;                /* the pump's workers are the thread pool */
        st.max_frame_delay = 1;
        st.apply_grain = apply_grain;
        st.inloop_filters = inloop_filters;
        st.logger.callback = quiet_log;
        s->out = out ? out[i] : NULL; s->cap = cap;
        if ((err = dav1d_open(&s->c, &st)) < 0) break;
        if ((err = dav1d_data_wrap(&s->data, obu[i], sz[i], no_free, NULL)) < 0) { dav1d_close(&s->c); break; }
    }
    if (!err) {
        pthread_t *th = calloc((size_t)n_workers, sizeof(*th));
        int started = 0;
        for (int i = 0; i < n_workers; i++) if (pthread_create(&th[i], NULL, worker, &pm) == 0) started++; else break;
        if (!started) err = DAV1D_ERR(EAGAIN);
        for (int i = 0; i < started; i++) pthread_join(th[i], NULL);
        free(th);
    }
    for (int i = 0; i < opened; i++) {
        PumpStream *s = &pm.st[i];
        if (!err && s->error) err = s->error;
        if (frames) frames[i] = s->frames;
        if (used) used[i] = s->used;
        while (s->n_pend) { dav1d_picture_unref(&s->pend[s->head]); s->head = (s->head + 1) % PUMP_DEPTH; s->n_pend--; }
        dav1d_data_unref(&s->data);
        dav1d_close(&s->c);
        pthread_mutex_destroy(&s->mu);
    }
    free(pm.st);
    return err;
}
