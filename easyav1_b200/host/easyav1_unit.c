/* easyav1_unit.c — the translation unit libeasyav1_b200.so is built from: the reference's
 * src/easyav1.c (which carries minivorbis through its own OGG_IMPL/VORBIS_IMPL include) compiled
 * UNMODIFIED where it lies, with three of minivorbis' functions marked weak so that the strong
 * definitions of vorbis_shim.c take their place at link time:
 *
 *   mdct_backward             ext/minivorbis/minivorbis.h:7054  (called by mapping0_inverse :17825)
 *   vorbis_synthesis_blockin  :9195                             (called by src/easyav1.c:2448,2459)
 *   vorbis_dsp_clear          :8770                             (called by src/easyav1.c:3977)
 *
 * Everything else of the audio path (Ogg packet handling, codebooks, floor1, residue, channel
 * coupling: the bit-serial part, SURVEY.md 8(a) "(host)") stays the reference's own code. What a
 * maintainer would change in the reference instead of this link-time trick is in INTEGRATION.md. */
#pragma weak mdct_backward
#pragma weak vorbis_synthesis_blockin
#pragma weak vorbis_dsp_clear

#include "easyav1.c"

/* ---- what vorbis_shim.c needs from minivorbis' private half ---------------------------------- */

/* the reference's own vorbis_dsp_clear (frees the decoder state); the exported name is the shim's */
extern __typeof__(vorbis_dsp_clear) eav1ref_vorbis_dsp_clear __attribute__((alias("vorbis_dsp_clear")));

int eav1ref_mdct_size(const void *lookup) { return ((const mdct_lookup *)lookup)->n; }

/* private_state.sample_count (:5312), the only private field the block bookkeeping touches */
ogg_int64_t *eav1ref_vorbis_sample_count(vorbis_dsp_state *v)
{
    return &((private_state *)v->backend_state)->sample_count;
}

int eav1ref_vorbis_halfrate(vorbis_dsp_state *v)
{
    return ((codec_setup_info *)v->vi->codec_setup)->halfrate_flag;
}
