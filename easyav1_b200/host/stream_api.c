/*
 * stream_api.c — flat C helpers over the (unchanged) dav1d public API exported by this
 * library, for bindings that cannot describe Dav1dSettings / Dav1dPicture (ctypes):
 * feed a buffer of OBUs, pull decoded frames as tightly packed visible Y|U|V planes.
 * Mirrors what src/easyav1.c:2360-2432 (decode_video) does with dav1d_send_data /
 * dav1d_get_picture.
 */
#include <errno.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "config.h"
#include "dav1d/dav1d.h"

typedef struct Eav1Stream {
    Dav1dContext *c;
    Dav1dData data;
} Eav1Stream;

typedef struct Eav1StreamFrame {
    int w, h, layout, bpc;
    int frame_type, show_frame, showable;
    int64_t offset;
} Eav1StreamFrame;

static void no_free(const uint8_t *d, void *cookie) { (void)d; (void)cookie; }
static void quiet_log(void *cookie, const char *fmt, va_list ap) { (void)cookie; (void)fmt; (void)ap; }

Eav1Stream *eav1_stream_open(const uint8_t *obu, size_t sz, int apply_grain, int inloop_filters, int quiet)
{
    Eav1Stream *s = calloc(1, sizeof(*s));
    if (!s) return NULL;
    Dav1dSettings st;
    dav1d_default_settings(&st);
    st.apply_grain = apply_grain;
    st.inloop_filters = inloop_filters;
    if (quiet) st.logger.callback = quiet_log;
    if (dav1d_open(&s->c, &st) < 0) { free(s); return NULL; }
    if (dav1d_data_wrap(&s->data, obu, sz, no_free, NULL) < 0) { dav1d_close(&s->c); free(s); return NULL; }
    return s;
}

Dav1dContext *eav1_stream_ctx(Eav1Stream *s) { return s->c; }

/* 0 = frame returned, 1 = end of stream, <0 = DAV1D_ERR */
int eav1_stream_next(Eav1Stream *s, Eav1StreamFrame *info, uint8_t *out, size_t cap, size_t *used)
{
    Dav1dPicture p;
    for (;;) {
        memset(&p, 0, sizeof(p));
        int res;
        if (s->data.sz > 0) {
            res = dav1d_send_data(s->c, &s->data);
            if (res < 0 && res != DAV1D_ERR(EAGAIN)) return res;
        }
        res = dav1d_get_picture(s->c, &p);
        if (res == 0) break;
        if (res != DAV1D_ERR(EAGAIN)) return res;
        if (s->data.sz == 0) {
            res = dav1d_get_picture(s->c, &p);
            if (res == 0) break;
            return res == DAV1D_ERR(EAGAIN) ? 1 : res;
        }
    }
    info->w = p.p.w; info->h = p.p.h; info->layout = p.p.layout; info->bpc = p.p.bpc;
    info->frame_type = p.frame_hdr->frame_type;
    info->show_frame = p.frame_hdr->show_frame;
    info->showable = p.frame_hdr->showable_frame;
    info->offset = p.m.offset;
    const int hbd = p.p.bpc > 8;
    const int ss_hor = p.p.layout != DAV1D_PIXEL_LAYOUT_I444;
    const int ss_ver = p.p.layout == DAV1D_PIXEL_LAYOUT_I420;
    size_t n = 0;
    for (int pl = 0; pl < (p.p.layout == DAV1D_PIXEL_LAYOUT_I400 ? 1 : 3); pl++) {
        const int w = pl ? (p.p.w + ss_hor) >> ss_hor : p.p.w;
        const int h = pl ? (p.p.h + ss_ver) >> ss_ver : p.p.h;
        const size_t row = (size_t)w << hbd;
        const uint8_t *src = p.data[pl];
        for (int y = 0; y < h; y++) {
            if (out && n + row <= cap) memcpy(out + n, src, row);
            n += row;
            src += p.stride[!!pl];
        }
    }
    if (used) *used = n;
    dav1d_picture_unref(&p);
    return 0;
}

void eav1_stream_close(Eav1Stream *s)
{
    if (!s) return;
    dav1d_data_unref(&s->data);
    dav1d_close(&s->c);
    free(s);
}
