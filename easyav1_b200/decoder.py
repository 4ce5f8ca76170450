"""Stream-level binding: OBU bytes in, decoded planes out, through the unchanged dav1d API
(dav1d_open / dav1d_send_data / dav1d_get_picture) of libdav1d_b200.so — the call sequence
of src/easyav1.c:2360-2432."""
import ctypes as C

from .lib import load_host, StreamFrame, EngineError


class Decoder:
    def __init__(self, obu: bytes, apply_grain=1, inloop_filters=7, host=None, max_frame_bytes=None):
        self.L = host or load_host()
        self._keep = C.create_string_buffer(obu, len(obu))
        self.h = self.L.eav1_stream_open(self._keep, len(obu), apply_grain, inloop_filters, 1)
        if not self.h:
            raise EngineError("dav1d_open failed (no CUDA device? there is no CPU fallback)")
        self.cap = max_frame_bytes or (1 << 20)
        self.buf = C.create_string_buffer(self.cap)

    def __iter__(self):
        return self

    def __next__(self):
        info = StreamFrame()
        used = C.c_size_t(0)
        r = self.L.eav1_stream_next(self.h, C.byref(info), self.buf, self.cap, C.byref(used))
        if r == 1:
            raise StopIteration
        if r < 0:
            raise EngineError("decoder error %d" % r)
        if used.value > self.cap:
            raise EngineError("frame of %d bytes exceeds max_frame_bytes=%d" % (used.value, self.cap))
        return info, self.buf.raw[:used.value]

    @property
    def engine(self):
        """Eav1cuCtx* behind this stream (measurement hooks of include/easyav1_cuda.h)."""
        return self.L.eav1rec_engine(self.L.eav1_stream_ctx(self.h))

    @property
    def launches(self):
        return self.L.eav1rec_launches(self.L.eav1_stream_ctx(self.h))

    @property
    def frames_submitted(self):
        return self.L.eav1rec_frames_submitted(self.L.eav1_stream_ctx(self.h))

    def close(self):
        if self.h:
            self.L.eav1_stream_close(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def decode_obu(obu: bytes, **kw):
    d = Decoder(obu, **kw)
    try:
        return list(d)
    finally:
        d.close()
