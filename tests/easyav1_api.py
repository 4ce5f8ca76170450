"""ctypes view of the reference's public C99 API (src/easyav1.h) as exported by
libeasyav1_b200.so = the UNMODIFIED src/easyav1.c + nestegg compiled where they lie and linked
against libdav1d_b200.so (easyav1_b200/host/Makefile). TEST INFRASTRUCTURE."""
import ctypes as C
import hashlib


class Settings(C.Structure):   # easyav1_settings, src/easyav1.h:405-428
    _fields_ = [("enable_video", C.c_int), ("enable_audio", C.c_int), ("skip_unprocessed_frames", C.c_int),
                ("interlace_audio", C.c_int), ("close_handle_on_destroy", C.c_int),
                ("cb_video", C.c_void_p), ("cb_audio", C.c_void_p), ("cb_userdata", C.c_void_p),
                ("video_track", C.c_uint), ("audio_track", C.c_uint), ("use_fast_seeking", C.c_int),
                ("audio_offset_time", C.c_int64), ("log_level", C.c_int)]


class VideoFrame(C.Structure):   # easyav1_video_frame, src/easyav1.h:288-303
    _fields_ = [("pixel_layout", C.c_int), ("bits_per_color", C.c_int), ("color_space", C.c_int),
                ("color_primaries", C.c_int), ("transfer_characteristics", C.c_int), ("matrix_coefficients", C.c_int),
                ("chroma_sample_position", C.c_int), ("width", C.c_uint), ("height", C.c_uint),
                ("timestamp", C.c_uint64), ("data", C.c_void_p * 3), ("stride", C.c_size_t * 3)]


def bind(path):
    L = C.CDLL(path)
    L.easyav1_default_settings.restype = Settings
    L.easyav1_init_from_memory.restype = C.c_void_p
    L.easyav1_init_from_memory.argtypes = [C.c_void_p, C.c_size_t, C.POINTER(Settings)]
    L.easyav1_init_from_filename.restype = C.c_void_p
    L.easyav1_init_from_filename.argtypes = [C.c_char_p, C.POINTER(Settings)]
    L.easyav1_decode_next.argtypes = [C.c_void_p]
    L.easyav1_has_video_frame.argtypes = [C.c_void_p]
    L.easyav1_get_video_frame.restype = C.POINTER(VideoFrame)
    L.easyav1_get_video_frame.argtypes = [C.c_void_p]
    L.easyav1_is_finished.argtypes = [C.c_void_p]
    L.easyav1_destroy.argtypes = [C.POINTER(C.c_void_p)]
    L.easyav1_get_total_video_frames_processed.restype = C.c_uint64
    L.easyav1_get_total_video_frames_processed.argtypes = [C.c_void_p]
    return L


def frame_md5(fr, bpc, layout):
    """MD5 of the visible Y|U|V samples (same packing as the golden manifest)"""
    bps = 2 if bpc > 8 else 1
    w, h = fr.width, fr.height
    sx, sy = int(layout != 3), int(layout == 1)
    dims = [(w, h)] if layout == 0 else [(w, h), ((w + sx) >> sx, (h + sy) >> sy), ((w + sx) >> sx, (h + sy) >> sy)]
    md = hashlib.md5()
    for pl, (pw, ph) in enumerate(dims):
        base, stride = fr.data[pl], fr.stride[pl]
        for y in range(ph):
            md.update(C.string_at(base + y * stride, pw * bps))
    return md.hexdigest()


def decode_webm(L, webm, bpc=8, layout=1):
    """easyav1_init_from_memory -> easyav1_decode_next loop (the loop of easyav1.h:35-53); returns MD5s"""
    keep = C.create_string_buffer(webm, len(webm))
    st = L.easyav1_default_settings()
    st.enable_audio = 0
    st.log_level = 1
    h = L.easyav1_init_from_memory(keep, len(webm), C.byref(st))
    if not h:
        raise RuntimeError("easyav1_init_from_memory failed")
    out = []
    try:
        while L.easyav1_decode_next(h) == 1:
            if L.easyav1_has_video_frame(h) == 1:
                fr = L.easyav1_get_video_frame(h)
                out.append(frame_md5(fr.contents, bpc, layout))
        finished = L.easyav1_is_finished(h) == 1
    finally:
        hp = C.c_void_p(h)
        L.easyav1_destroy(C.byref(hp))
    return out, finished


VIDEO_CB = C.CFUNCTYPE(None, C.POINTER(VideoFrame), C.c_void_p)


def play_file(L, path, bpc=8, layout=1, timeout_s=30.0):
    """easyav1_init_from_filename -> easyav1_play: the library's own playback thread decodes in real time and hands
    every frame to the video callback (src/easyav1.h:531-551). Returns (MD5s in callback order, finished)."""
    import time
    L.easyav1_play.argtypes = [C.c_void_p]
    L.easyav1_stop.argtypes = [C.c_void_p]
    L.easyav1_stop.restype = None
    out = []

    def on_video(frame, userdata):
        out.append(frame_md5(frame.contents, bpc, layout))
    cb = VIDEO_CB(on_video)
    st = L.easyav1_default_settings()
    st.enable_audio = 0
    st.skip_unprocessed_frames = 0
    st.log_level = 1
    st.cb_video = C.cast(cb, C.c_void_p)
    h = L.easyav1_init_from_filename(path.encode(), C.byref(st))
    if not h:
        raise RuntimeError("easyav1_init_from_filename failed")
    try:
        assert L.easyav1_play(h) == 1, "easyav1_play failed"
        t0 = time.time()
        while L.easyav1_is_finished(h) != 1 and time.time() - t0 < timeout_s:
            time.sleep(0.02)
        finished = L.easyav1_is_finished(h) == 1
        L.easyav1_stop(h)
    finally:
        hp = C.c_void_p(h)
        L.easyav1_destroy(C.byref(hp))
    return out, finished


class AudioFrame(C.Structure):   # easyav1_audio_frame, src/easyav1.h:309-319
    _fields_ = [("channels", C.c_uint), ("samples", C.c_uint), ("timestamp", C.c_uint64), ("bytes", C.c_size_t),
                ("pcm", C.c_void_p)]


def decode_audio(L, webm, interlace=True, seek_after=None, seek_to_ms=0, enable_video=True):
    """The loop of easyav1.h:35-53 with audio enabled. Returns ([(timestamp, samples, float32 bytes per frame)], number
    of video frames). With seek_after=k, easyav1_seek_to_timestamp(seek_to_ms) is called once after k audio frames."""
    keep = C.create_string_buffer(webm, len(webm))
    L.easyav1_get_audio_frame.restype = C.POINTER(AudioFrame)
    L.easyav1_get_audio_frame.argtypes = [C.c_void_p]
    L.easyav1_seek_to_timestamp.argtypes = [C.c_void_p, C.c_uint64]
    L.easyav1_get_audio_channels.argtypes = [C.c_void_p]
    st = L.easyav1_default_settings()
    st.enable_audio = 1
    st.enable_video = 1 if enable_video else 0
    st.interlace_audio = 1 if interlace else 0
    st.log_level = 1
    h = L.easyav1_init_from_memory(keep, len(webm), C.byref(st))
    if not h:
        raise RuntimeError("easyav1_init_from_memory failed")
    out, n_video, sought = [], 0, False
    try:
        while L.easyav1_decode_next(h) == 1:
            if L.easyav1_has_video_frame(h) == 1:
                L.easyav1_get_video_frame(h)
                n_video += 1
            afr = L.easyav1_get_audio_frame(h)     # NULL when no new samples (src/easyav1.h:769)
            if afr:
                fr = afr.contents
                if interlace:
                    data = C.string_at(fr.pcm, fr.samples * fr.channels * 4)
                else:
                    ptrs = C.cast(fr.pcm, C.POINTER(C.c_void_p))
                    data = b"".join(C.string_at(ptrs[c], fr.samples * 4) for c in range(fr.channels))
                out.append((int(fr.timestamp), int(fr.samples), data))
                if seek_after is not None and not sought and len(out) >= seek_after:
                    sought = True
                    if L.easyav1_seek_to_timestamp(h, seek_to_ms) != 1:
                        raise RuntimeError("easyav1_seek_to_timestamp failed")
    finally:
        hp = C.c_void_p(h)
        L.easyav1_destroy(C.byref(hp))
    return out, n_video
