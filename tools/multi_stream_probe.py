#!/usr/bin/env python3
"""Development aid: S concurrent decoders end to end on one GPU; prints frames/s, host CPU utilisation and
(EAV1_TIMING=1) the submit threads' per-frame time split.
usage: multi_stream_probe.py S [passes] [threads_each] [active]      one Python thread per stream
       multi_stream_probe.py S passes pump W                        host/pump.c with W workers"""
import os, resource, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench

def main():
    S = int(sys.argv[1]); passes = int(sys.argv[2]) if len(sys.argv) > 2 else 2
    from easyav1_b200.lib import load_engine, load_host
    load_engine(); host = load_host()
    man, obu = bench.load_clip()
    cores = bench.host_threads()
    if len(sys.argv) > 3 and sys.argv[3] == "pump":
        import ctypes as C, hashlib
        from easyav1_b200.lib import pump_run
        W = int(sys.argv[4])
        cap = man["w"] * man["h"] * 4
        keep = C.create_string_buffer(obu, len(obu))
        bufs = [C.create_string_buffer(cap) for _ in range(S)]
        class K:
            def __len__(self): return len(obu)
        obus = [obu] * S
        pump_run(host, obus, W, cap, bufs=bufs)
        r0 = resource.getrusage(resource.RUSAGE_SELF); t0 = time.perf_counter(); n = 0
        for _ in range(passes):
            frames, last = pump_run(host, obus, W, cap, bufs=bufs)
            n += sum(frames)
        wall = time.perf_counter() - t0; r1 = resource.getrusage(resource.RUSAGE_SELF)
        assert all(hashlib.md5(x).hexdigest() == man["frames"][-1]["md5"] for x in last)
        cpu = (r1.ru_utime - r0.ru_utime) + (r1.ru_stime - r0.ru_stime)
        print("S=%d pump workers=%d: %.0f frames/s; %.1f of %d cores busy; %.2f cpu-ms per frame" % (S, W, n / wall, cpu / wall, cores, 1e3 * cpu / n))
        return
    thr = int(sys.argv[3]) if len(sys.argv) > 3 else max(1, cores // S)
    active = int(sys.argv[4]) if len(sys.argv) > 4 else None
    bench.multi_stream_e2e(host, man, obu, S, 1, thr, active)
    r0 = resource.getrusage(resource.RUSAGE_SELF); t0 = time.perf_counter()
    last, n, dt, stats = bench.multi_stream_e2e(host, man, obu, S, passes, thr, active)
    r1 = resource.getrusage(resource.RUSAGE_SELF); wall = time.perf_counter() - t0
    cpu = (r1.ru_utime - r0.ru_utime) + (r1.ru_stime - r0.ru_stime)
    print("S=%d active=%s threads_each=%d: %.0f frames/s; cpu %.1f s over %.2f s wall = %.1f of %d cores busy (user %.1f sys %.1f); %.2f cpu-ms per frame"
          % (S, active, thr, n / dt, cpu, wall, cpu / wall, cores, r1.ru_utime - r0.ru_utime, r1.ru_stime - r0.ru_stime, 1e3 * cpu / n))

main()
