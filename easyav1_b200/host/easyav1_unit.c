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

/* ---- an additional getter beside easyav1_get_video_frame (src/easyav1.h:738), which stays as it is ----
 * The frame easyav1_get_video_frame last returned (easyav1->video.picture, src/easyav1.c:3520-3568), converted on the
 * GPU to format 1 = NV12, 2 = P010, 3 = RGBA8 (include/easyav1_cuda.h) with the stream's own matrix / range, written
 * to dst. Returns EASYAV1_STATUS_OK or EASYAV1_STATUS_ERROR (no frame yet, format not possible for this stream,
 * dst too small). */
extern int eav1rec_picture_convert(const Dav1dPicture *p, int format, int matrix, int full_range, void *dst,
                                   size_t dst_stride, size_t dst_bytes);

easyav1_status easyav1_b200_get_video_frame_converted(easyav1_t *easyav1, int format, void *dst, size_t dst_stride, size_t dst_bytes)
{
    if (!easyav1 || !dst || easyav1->video.active != EASYAV1_TRUE || !easyav1->video.picture.data[0])
        return EASYAV1_STATUS_ERROR;
    return eav1rec_picture_convert(&easyav1->video.picture, format, -1, -1, dst, dst_stride, dst_bytes) == 0 ?
           EASYAV1_STATUS_OK : EASYAV1_STATUS_ERROR;
}
