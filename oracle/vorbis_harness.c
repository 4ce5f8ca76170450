/* oracle/vorbis_harness.c — TEST INFRASTRUCTURE ONLY.
 *
 * Compiles the reference's Vorbis decoder (ext/minivorbis/minivorbis.h, single header, the same
 * OGG_IMPL / VORBIS_IMPL switches as src/easyav1.c:3-8) as plain C where it lies and exposes the
 * float slice of SURVEY.md rows V1-V3 through flat wrappers:
 *   vref_mdct_backward   mdct_backward            (minivorbis.h:7054) with mdct_init tables (:6709)
 *   vref_window          _vorbis_window_get       (minivorbis.h:11992)
 *   vref_open/_block     vorbis_synthesis_blockin (minivorbis.h:9195) + pcmout/read (:9411-9435)
 *                        driven with hand-made blocks whose pcm[] is the IMDCT of given spectra,
 *                        exactly as mapping0_inverse leaves them (minivorbis.h:17825)
 * Nothing here is shipped or called by the product. */
#define OGG_IMPL
#define VORBIS_IMPL
#include "minivorbis.h"

#include <stdlib.h>
#include <string.h>

static mdct_lookup g_mdct[16];

static mdct_lookup *lookup_for(int n)
{
    int l = 0;
    while ((1 << l) < n) l++;
    if (l >= 16 || (1 << l) != n) return NULL;
    if (!g_mdct[l].n) mdct_init(&g_mdct[l], n);
    return &g_mdct[l];
}

/* in: n/2 spectral lines, out: n time samples (not windowed) */
int vref_mdct_backward(int n, const float *in, float *out)
{
    mdct_lookup *m = lookup_for(n);
    if (!m) return -1;
    float *tmp = malloc(sizeof(float) * (size_t)n);
    memcpy(tmp, in, sizeof(float) * (size_t)(n / 2));
    mdct_backward(m, tmp, tmp);          /* in place, like mapping0_inverse */
    memcpy(out, tmp, sizeof(float) * (size_t)n);
    free(tmp);
    return 0;
}

/* half window (n/2 entries) of block size n */
const float *vref_window(int n)
{
    return _vorbis_window_get(ov_ilog(n) - 7);
}

typedef struct VRef {
    vorbis_info vi;
    codec_setup_info *ci;
    vorbis_dsp_state v;
    vorbis_block vb;
    float **pcm;
    long seq;
} VRef;

VRef *vref_open(int channels, int bs0, int bs1)
{
    VRef *r = calloc(1, sizeof(*r));
    vorbis_info_init(&r->vi);
    r->ci = r->vi.codec_setup;
    r->vi.channels = channels;
    r->vi.rate = 48000;
    r->ci->blocksizes[0] = bs0;
    r->ci->blocksizes[1] = bs1;
    r->ci->modes = 1;
    r->ci->mode_param[0] = _ogg_calloc(1, sizeof(*r->ci->mode_param[0]));
    if (vorbis_synthesis_init(&r->v, &r->vi)) { free(r); return NULL; }
    memset(&r->vb, 0, sizeof(r->vb));
    r->vb.vd = &r->v;
    r->pcm = calloc((size_t)channels, sizeof(float *));
    for (int i = 0; i < channels; i++) r->pcm[i] = calloc((size_t)bs1, sizeof(float));
    r->vb.pcm = r->pcm;
    return r;
}

/* One block (W = 0 short / 1 long) for all channels; spectra = [channel][n/2]. Finished samples are
 * appended planar to out[channel * cap + pos .. ). Returns the number of samples produced. */
int vref_block(VRef *r, int W, const float *spectra, float *out, long cap, long pos)
{
    const int n = (int)r->ci->blocksizes[W];
    mdct_lookup *m = r->v.backend_state ? ((private_state *)r->v.backend_state)->transform[W][0] : NULL;
    if (!m) return -1;
    for (int c = 0; c < r->vi.channels; c++) {
        memcpy(r->pcm[c], spectra + (size_t)c * (n / 2), sizeof(float) * (size_t)(n / 2));
        mdct_backward(m, r->pcm[c], r->pcm[c]);
    }
    r->vb.W = W;
    r->vb.lW = r->v.W;
    r->vb.nW = 0;
    r->vb.pcmend = n;
    r->vb.sequence = r->seq++;
    r->vb.granulepos = -1;
    r->vb.eofflag = 0;
    if (vorbis_synthesis_blockin(&r->v, &r->vb)) return -2;
    float **pcm;
    const int got = vorbis_synthesis_pcmout(&r->v, &pcm);
    if (got > 0) {
        if (pos + got > cap) return -3;
        for (int c = 0; c < r->vi.channels; c++)
            memcpy(out + (size_t)c * cap + pos, pcm[c], sizeof(float) * (size_t)got);
        vorbis_synthesis_read(&r->v, got);
    }
    return got;
}

void vref_restart(VRef *r) { vorbis_synthesis_restart(&r->v); r->seq = 0; }

void vref_close(VRef *r)
{
    if (!r) return;
    for (int i = 0; i < r->vi.channels; i++) free(r->pcm[i]);
    free(r->pcm);
    vorbis_dsp_clear(&r->v);
    vorbis_info_clear(&r->vi);
    free(r);
}
