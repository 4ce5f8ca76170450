/* placeholder: filled in when the Vorbis slice (V1-V3) is built */
int vorbis_ref_placeholder(void) { return 0; }
