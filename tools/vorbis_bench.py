#!/usr/bin/env python3
"""Config C3: 4096 blocks x 6 channels, block sizes 256/2048, through the batched
Vorbis IMDCT + window + overlap-add kernel. Prints device-resident and host-to-host throughput and
the HBM roofline fraction (algorithmic bytes = 8 * blocksize per (block, channel))."""
import ctypes as C
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tests import vorbis_ops as vo   # batch generator only; the oracle is not timed here


def main():
    E = C.CDLL(os.path.join(ROOT, "easyav1_b200", "libeasyav1_cuda.so"))
    channels, bs, n_blocks = 6, (256, 2048), int(sys.argv[1]) if len(sys.argv) > 1 else 4096
    flags, spectra = vo.make_batch(3, channels, bs, n_blocks)
    E.eav1cu_vorbis_open.argtypes = [C.POINTER(C.c_void_p), C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
    E.eav1cu_vorbis_stage.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.POINTER(C.c_size_t)]
    E.eav1cu_vorbis_run_staged.argtypes = [C.c_void_p]
    E.eav1cu_vorbis_set_stream.argtypes = [C.c_void_p, C.c_void_p]
    E.eav1cu_vorbis_sync.argtypes = [C.c_void_p]
    E.eav1cu_vorbis_imdct_ola.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_size_t, C.c_int,
                                          C.POINTER(C.c_size_t)]
    h = C.c_void_p()
    assert E.eav1cu_vorbis_open(C.byref(h), 0, channels, bs[0], bs[1], None, None) == 0
    stream = torch.cuda.Stream()
    E.eav1cu_vorbis_set_stream(h, C.c_void_p(stream.cuda_stream))
    n = C.c_size_t(0)
    assert E.eav1cu_vorbis_stage(h, flags.ctypes.data, n_blocks, spectra.ctypes.data, 0, C.byref(n)) == 0
    alg_bytes = float(sum(8 * bs[int(f)] for f in flags)) * channels
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    times = []
    with torch.cuda.stream(stream):
        for it in range(13):
            flush.fill_(it & 255)
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(stream)
            assert E.eav1cu_vorbis_run_staged(h) == 0
            b.record(stream)
            stream.synchronize()
            if it >= 3:
                times.append(a.elapsed_time(b))
    ms = float(np.median(times))
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))).get("hbm_gbs", 6457.4) if os.path.exists(
        os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6457.4
    # host to host
    cap = int(n.value) + bs[1]
    out = np.zeros((channels, cap), np.float32)
    E.eav1cu_vorbis_restart = E.eav1cu_vorbis_restart
    t_host = []
    for it in range(6):
        E.eav1cu_vorbis_restart(h)
        t0 = time.perf_counter()
        assert E.eav1cu_vorbis_imdct_ola(h, flags.ctypes.data, n_blocks, spectra.ctypes.data, out.ctypes.data, cap, 0,
                                         C.byref(n)) == 0
        t_host.append(time.perf_counter() - t0)
    res = {"workload": "C3: %d blocks x %d ch, blocksizes %s" % (n_blocks, channels, list(bs)),
           "samples_per_channel": int(n.value), "kernel_ms": ms, "alg_bytes": alg_bytes,
           "achieved_gbs": alg_bytes / ms / 1e6, "peak_gbs": peak, "frac": alg_bytes / ms / 1e6 / peak,
           "msamples_per_s_device": n.value * channels / ms / 1e3,
           "msamples_per_s_host_to_host": n.value * channels / float(np.median(t_host[1:])) / 1e6}
    print(json.dumps(res))


if __name__ == "__main__":
    main()
