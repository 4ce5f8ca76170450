/* vorbis_shim.c — puts the CUDA IMDCT + window + overlap-add (eav1cu_vorbis_imdct_ola,
 * csrc/vorbis.cu) behind easyav1_get_audio_frame without touching
 * src/easyav1.c: the three functions below replace minivorbis' weak-marked ones at link time
 * (easyav1_unit.c).
 *
 *   mdct_backward            only parks the spectrum of one channel: the transform itself runs on
 *                            the GPU together with the window and the overlap-add, which need the
 *                            block type of the neighbours that only blockin knows
 *   vorbis_synthesis_blockin hands the parked spectra of the block to the GPU object of this
 *                            decoder, receives the samples the block makes final and publishes
 *                            them through v->pcm / pcm_returned / pcm_current, which is where
 *                            vorbis_synthesis_pcmout (:9411) and easyav1.c's interleave loop
 *                            (src/easyav1.c:2468-2492) pick them up
 *   vorbis_dsp_clear         releases the GPU object, then runs the reference's own clean-up
 *
 * The overlap state (the second half of the previous block) lives on the device; v->pcm only
 * ever holds finished samples here. There is no CPU fallback: without a usable GPU blockin fails
 * with OV_EFAULT and easyav1 reports a decoder error. */
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "minivorbis/minivorbis.h"      /* declarations only: this file does not define *_IMPL */
#include "easyav1_cuda.h"

extern void eav1ref_vorbis_dsp_clear(vorbis_dsp_state *v);
extern int eav1ref_mdct_size(const void *lookup);
extern ogg_int64_t *eav1ref_vorbis_sample_count(vorbis_dsp_state *v);
extern int eav1ref_vorbis_halfrate(vorbis_dsp_state *v);
extern const float *_vorbis_window_get(int n);   /* minivorbis.h:11992 */

/* ---- spectra parked by mdct_backward until their block is submitted -------------------------- */

typedef struct { float *where; int n; size_t off; } Parked;
static __thread Parked *t_parked;
static __thread int t_n_parked, t_cap_parked;
static __thread float *t_spec;
static __thread size_t t_spec_used, t_spec_cap;
static __thread float *t_plane;            /* finished samples of one block, planar */
static __thread size_t t_plane_cap;

void mdct_backward(void *lookup, float *in, float *out)
{
    const int n = eav1ref_mdct_size(lookup);
    const size_t half = (size_t)n / 2;
    if (t_n_parked == t_cap_parked) {
        t_cap_parked = t_cap_parked ? 2 * t_cap_parked : 8;
        t_parked = realloc(t_parked, (size_t)t_cap_parked * sizeof(*t_parked));
    }
    if (t_spec_used + half > t_spec_cap) {
        t_spec_cap = 2 * (t_spec_used + half);
        t_spec = realloc(t_spec, t_spec_cap * sizeof(float));
    }
    if (!t_parked || !t_spec) { t_n_parked = t_cap_parked = 0; t_spec_used = t_spec_cap = 0; return; }
    memcpy(t_spec + t_spec_used, in, half * sizeof(float));
    t_parked[t_n_parked++] = (Parked){ out, n, t_spec_used };
    t_spec_used += half;
}

/* ---- one GPU object per vorbis_dsp_state ------------------------------------------------------ */

typedef struct { vorbis_dsp_state *key; Eav1Vorbis *gpu; int channels, bs[2]; } Track;
static pthread_mutex_t g_mu = PTHREAD_MUTEX_INITIALIZER;
static Track *g_tracks;
static int g_n_tracks, g_cap_tracks;

static int log2i(int v) { int n = 0; while (v > 1) { v >>= 1; n++; } return n; }

static Eav1Vorbis *track_of(vorbis_dsp_state *v)
{
    const int ch = v->vi->channels;
    const int bs0 = (int)vorbis_info_blocksize(v->vi, 0), bs1 = (int)vorbis_info_blocksize(v->vi, 1);
    Eav1Vorbis *gpu = NULL;
    pthread_mutex_lock(&g_mu);
    int slot = -1;
    for (int i = 0; i < g_n_tracks; i++)
        if (g_tracks[i].key == v) { slot = i; break; }
    if (slot >= 0 && (g_tracks[slot].channels != ch || g_tracks[slot].bs[0] != bs0 || g_tracks[slot].bs[1] != bs1)) {
        eav1cu_vorbis_close(g_tracks[slot].gpu);     /* the address was reused by another stream */
        g_tracks[slot] = g_tracks[--g_n_tracks];
        slot = -1;
    }
    if (slot < 0) {
        if (g_n_tracks == g_cap_tracks) {
            const int cap = g_cap_tracks ? 2 * g_cap_tracks : 8;
            Track *t = realloc(g_tracks, (size_t)cap * sizeof(*t));
            if (!t) { pthread_mutex_unlock(&g_mu); return NULL; }
            g_tracks = t; g_cap_tracks = cap;
        }
        const char *dev = getenv("EASYAV1_CUDA_DEVICE");
        /* the reference's own window tables (:9917-11990), so the samples differ from the
         * reference's only by the order of the butterflies */
        const int rc = eav1cu_vorbis_open(&gpu, dev ? atoi(dev) : 0, ch, bs0, bs1,
                                          _vorbis_window_get(log2i(bs0) - 6), _vorbis_window_get(log2i(bs1) - 6));
        if (rc) {
            fprintf(stderr, "easyav1_b200: the Vorbis transform needs a CUDA device (error %d; there is no CPU fallback)\n", rc);
            pthread_mutex_unlock(&g_mu);
            return NULL;
        }
        g_tracks[g_n_tracks++] = (Track){ v, gpu, ch, { bs0, bs1 } };
    } else {
        gpu = g_tracks[slot].gpu;
    }
    pthread_mutex_unlock(&g_mu);
    return gpu;
}

void vorbis_dsp_clear(vorbis_dsp_state *v)
{
    if (v) {
        pthread_mutex_lock(&g_mu);
        for (int i = 0; i < g_n_tracks; i++)
            if (g_tracks[i].key == v) {
                eav1cu_vorbis_close(g_tracks[i].gpu);
                g_tracks[i] = g_tracks[--g_n_tracks];
                break;
            }
        pthread_mutex_unlock(&g_mu);
    }
    eav1ref_vorbis_dsp_clear(v);
}

/* ---- vorbis_synthesis_blockin ------------------------------------------------------------------ */

/* Samples that leave the end (end-of-stream page) or the start (first page) of what is buffered when
 * the stream's granule position disagrees with the running count, as :9322-9400 define it. */
static void trim(vorbis_dsp_state *v, long extra, int at_end)
{
    const long have = v->pcm_current - v->pcm_returned;
    if (extra < 0) extra = 0;
    if (at_end) {
        v->pcm_current -= (int)(extra > have ? have : extra);
    } else {
        v->pcm_returned += (int)extra;
        if (v->pcm_returned > v->pcm_current) v->pcm_returned = v->pcm_current;
    }
}

int vorbis_synthesis_blockin(vorbis_dsp_state *v, vorbis_block *vb)
{
    Parked *parked = t_parked;
    const int n_parked = t_n_parked;
    t_n_parked = 0; t_spec_used = 0;                 /* whatever was parked belongs to this call */

    if (!vb) return OV_EINVAL;
    if (v->pcm_returned != -1 && v->pcm_current > v->pcm_returned) return OV_EINVAL;   /* unread samples (:9203) */
    if (eav1ref_vorbis_halfrate(v)) return OV_EIMPL;   /* vorbis_synthesis_halfrate: not used by easyAV1 */

    const long bs[2] = { vorbis_info_blocksize(v->vi, 0), vorbis_info_blocksize(v->vi, 1) };
    ogg_int64_t *sample_count = eav1ref_vorbis_sample_count(v);
    const int fresh = v->pcm_returned == -1;         /* after vorbis_synthesis_init/_restart (:9159) */

    v->lW = v->W;
    v->W = vb->W;
    v->nW = -1;
    if (v->sequence == -1 || v->sequence + 1 != vb->sequence) {
        v->granulepos = -1;                          /* a gap: positions are unknown until the next page says */
        *sample_count = -1;
    }
    v->sequence = vb->sequence;
    const long advance = bs[v->lW] / 4 + bs[v->W] / 4;

    if (vb->pcm) {                                   /* NULL after vorbis_synthesis_trackonly (seeking) */
        const int ch = v->vi->channels;
        Eav1Vorbis *gpu = track_of(v);
        if (!gpu) return OV_EFAULT;
        if (n_parked != ch) return OV_EFAULT;
        for (int j = 0; j < ch; j++)
            if (parked[j].where != vb->pcm[j] || parked[j].n != bs[v->W] || parked[j].off != (size_t)j * (bs[v->W] / 2))
                return OV_EFAULT;
        v->glue_bits += vb->glue_bits; v->time_bits += vb->time_bits;
        v->floor_bits += vb->floor_bits; v->res_bits += vb->res_bits;

        if (fresh && eav1cu_vorbis_restart(gpu)) return OV_EFAULT;
        /* finished samples land at the start of v->pcm[]: planar, one row of pcm_storage floats per channel */
        const size_t need = (size_t)ch * (size_t)v->pcm_storage;
        if (need > t_plane_cap) {
            free(t_plane);
            t_plane = malloc(need * sizeof(float));
            t_plane_cap = t_plane ? need : 0;
        }
        float *plane = t_plane;
        if (!plane) return OV_EFAULT;
        const uint8_t flag = (uint8_t)(v->W != 0);
        size_t made = 0;
        const int rc = eav1cu_vorbis_imdct_ola(gpu, &flag, 1, t_spec, plane, (size_t)v->pcm_storage, 0, &made);
        if (rc || (fresh ? made != 0 : made != (size_t)advance)) {
            char msg[256] = "";
            eav1cu_vorbis_last_error(gpu, msg, sizeof(msg));
            fprintf(stderr, "easyav1_b200: Vorbis block failed on the GPU (%d) %s\n", rc, msg);
            return OV_EFAULT;
        }
        for (int j = 0; j < ch; j++) memcpy(v->pcm[j], plane + (size_t)j * v->pcm_storage, made * sizeof(float));
        v->centerW = v->centerW ? 0 : (int)(bs[1] / 2);   /* kept in step for vorbis_synthesis_lapout users */
        v->pcm_returned = 0;
        v->pcm_current = (int)made;
    }

    /* running sample count and granule position (:9311-9400) */
    if (*sample_count == -1) *sample_count = 0;
    else *sample_count += advance;

    if (v->granulepos == -1) {
        if (vb->granulepos != -1) {
            v->granulepos = vb->granulepos;
            if (*sample_count > v->granulepos)       /* a short first (or only) page */
                trim(v, (long)(*sample_count - vb->granulepos), vb->eofflag);
        }
    } else {
        v->granulepos += advance;
        if (vb->granulepos != -1 && v->granulepos != vb->granulepos) {
            if (v->granulepos > vb->granulepos && vb->eofflag)     /* a partial last block */
                trim(v, (long)(v->granulepos - vb->granulepos), 1);
            v->granulepos = vb->granulepos;
        }
    }
    if (vb->eofflag) v->eofflag = 1;
    return 0;
}
