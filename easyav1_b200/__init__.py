"""easyav1_b200 — B200 (sm_100a) AV1 reconstruction engine behind dav1d's DSP surface.

Only what the hot path needs: csrc/ (CUDA kernels + C ABI, libeasyav1_cuda.so), host/ (the
recorder that replaces dav1d's Dav1dDSPContext tables, libdav1d_b200.so) and thin ctypes
bindings. There is no CPU fallback: loading fails loudly when the CUDA libraries are missing.
"""
from .lib import load_engine, load_host, EngineError  # noqa: F401
from .decoder import Decoder, decode_obu  # noqa: F401
