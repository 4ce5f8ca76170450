#!/usr/bin/env python3
"""Development aid: how fast is the HOST side on this box?  Reference dav1d C at several thread
counts, and our recorder with the null emulation engine (parse + entropy + recording only)."""
import ctypes as C, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import load_clip, ref_decode_pass
man, obu = load_clip()
cap = man["w"] * man["h"] * 4
try:
    print([l for l in open("/proc/cpuinfo") if "model name" in l][0].strip(), "| cpus", os.cpu_count(),
          "| affinity", len(os.sched_getaffinity(0)))
except Exception as e:
    print(e)
for thr in (1, 2, 4, 8, 16, 32):
    if thr > 2 * (os.cpu_count() or 1):
        break
    n, dt, _ = ref_decode_pass(obu, cap, thr)
    n, dt, _ = ref_decode_pass(obu, cap, thr)
    print("reference n_threads=%d: %.1f fps" % (thr, n / dt))
emu = os.path.join(ROOT, "tests/emu/libdav1d_b200_emu.so")
if os.path.exists(emu):
    os.environ.setdefault("EAV1_EMU_NULL", "1")   # 2: no picture copies either
    from easyav1_b200.lib import _bind_host, StreamFrame
    host = _bind_host(C.CDLL(emu, mode=C.RTLD_GLOBAL))
    buf = C.create_string_buffer(cap)
    for rep in range(2):
        keep = C.create_string_buffer(obu, len(obu))
        t0 = time.perf_counter()
        h = host.eav1_stream_open(keep, len(obu), 1, 7, 1)
        info, used, n = StreamFrame(), C.c_size_t(0), 0
        while host.eav1_stream_next(h, C.byref(info), buf, cap, C.byref(used)) == 0:
            n += 1
        host.eav1_stream_close(h)
        dt = time.perf_counter() - t0
        print("host-only (record, null engine): %.1f fps" % (n / dt))
