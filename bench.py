#!/usr/bin/env python3
"""bench.py — AV1 decoded frames/s on the BASELINE.json configs[1] workload.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

Workload: tests/golden/c2_1080p_bench.ivf — 1920x1080 8-bit 4:2:0, 120 frames (1 intra + 119
inter), CDEF + loop restoration on. One "step" = one pass of the
reconstruction path over the whole 120-frame stream.

  value         frames/s with every work list (commands, coefficients, filter side data) already
                resident in HBM: the staged frames are re-launched in decode order, timed with
                CUDA events on the launching stream, L2 flushed between steps.
  e2e           frames/s through the reference-facing API (dav1d_open / dav1d_send_data /
                dav1d_get_picture of libdav1d_b200.so) from HOST OBU bytes to HOST planes:
                entropy decode, host->device copies of the work lists, kernels, device->host
                copy of every picture are all inside the timed region.
  roofline      the stage with the largest share of device time, algorithmic bytes of
                SURVEY.md §8(d) / its average duration (CUDA events around the stage).
  cpu_baseline  the reference's own dav1d C build (oracle/_ref) on the host cores, same stream.

N > 1 (torchrun, one process per GPU): every rank decodes its own copy of the stream (streams
shard, nothing is exchanged: SURVEY.md §8(e)); value = N * frames / max-over-ranks time.
"""
import argparse
import ctypes as C
import hashlib
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

CLIP = "c2_1080p_bench"
METRIC = "AV1 decoded frames/s (1080p)"
UNIT = "frames/s"


def load_clip(name=None):
    from tests.golden.gen_fixtures import read_ivf
    man = json.load(open(os.path.join(ROOT, "tests", "golden", "manifest.json")))[name or CLIP]
    obu = b"".join(read_ivf(os.path.join(ROOT, "tests", "golden", man["file"])))
    return man, obu


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs"
        except Exception:
            pass
    return 6550.0, "B200_PROFILING.md fallback"


def ncu_traffic(kernel_name):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of `kernel_name`, from the newest committed
    `ncu --set full` summary under profiles/ (tools/profile_round.sh writes it); None if there is none.
    It is a profiler figure (cold cache, serialised launches), reported beside the live numbers."""
    import glob
    best = None
    for path in sorted(glob.glob(os.path.join(ROOT, "profiles", "r*_full_summary.json"))):
        try:
            d = json.load(open(path))
        except Exception:
            continue
        for k, v in d.items():
            if kernel_name in k:
                rd = [x for n, x in v.items() if n.startswith("dram__bytes_read.sum")]
                wr = [x for n, x in v.items() if n.startswith("dram__bytes_write.sum")]
                unit = [n for n in v if n.startswith("dram__bytes_read.sum")]
                if rd and wr and unit:
                    scale = 1e6 if "Mbyte" in unit[0] else 1e3 if "Kbyte" in unit[0] else 1e9 if "Gbyte" in unit[0] else 1.0
                    best = {"bytes_per_launch": int((rd[0] + wr[0]) * scale), "source": os.path.basename(path)}
    return best


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]

    def __init__(self, index):
        self.index, self.rows, self.p = index, [], None

    def start(self):
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                       "--format=csv,noheader,nounits", "-lms", "100"],
                                      stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.p = None

    def _read(self):
        for line in self.p.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if not self.p:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.t.join(timeout=2)
        sm, mx, reasons = [], None, set()
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx = float(r[1])
                for i, n in enumerate(self.NAMES):
                    if r[2 + i].lower().startswith("active"):
                        reasons.add(n)
            except Exception:
                continue
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


# ----------------------------------------------------------------------------- reference arm

def ref_decode_pass(obu, cap, n_threads):
    """One pass of the reference's dav1d C build (oracle/_ref) over the stream; returns
    (frames, seconds, md5 of last frame). Planes are packed into a host buffer like our e2e leg."""
    from tests import refdec
    L = refdec.lib()
    keep = C.create_string_buffer(obu, len(obu))
    buf = C.create_string_buffer(cap)
    t0 = time.perf_counter()
    d = L.ref_dec_open(keep, len(obu), n_threads, 0, 1, 7, 1)
    if not d:
        raise RuntimeError("ref_dec_open failed")
    n = 0
    info, used = refdec.RefFrameInfo(), C.c_size_t(0)
    while True:
        r = L.ref_dec_next(d, C.byref(info), buf, cap, C.byref(used))
        if r == 1:
            break
        if r < 0:
            raise RuntimeError("reference decoder error %d" % r)
        n += 1
    L.ref_dec_close(d)
    dt = time.perf_counter() - t0
    return n, dt, hashlib.md5(buf.raw[:used.value]).hexdigest()


def host_threads():
    try:
        n = len(os.sched_getaffinity(0))
    except Exception:
        n = os.cpu_count() or 1
    return max(1, min(n, 256))   # DAV1D_MAX_THREADS (ext/dav1d/src/internal.h)


def run_reference(args, man, obu):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cap = man["w"] * man["h"] * 4
    thr = host_threads()
    for _ in range(args.warmup):
        ref_decode_pass(obu, cap, thr)
    total_t, total_n, md5 = 0.0, 0, None
    for _ in range(args.steps):
        n, dt, md5 = ref_decode_pass(obu, cap, thr)
        total_t += dt
        total_n += n
    assert md5 == man["frames"][-1]["md5"], "reference decode does not match its own golden MD5"
    fps = total_n / total_t
    sample = "%d passes of the %d-frame %s stream, dav1d C (no asm), n_threads=%d, frame delay auto" % (
        args.steps, len(man["frames"]), CLIP, thr)
    multi = None
    if args.streams > 1:
        t_each = max(2, thr // args.streams)
        concurrent_passes(lambda: ref_decode_pass(obu, cap, t_each)[0], args.streams, 1)
        n_multi, t_multi = concurrent_passes(lambda: ref_decode_pass(obu, cap, t_each)[0], args.streams, args.steps)
        multi = {"streams": args.streams, "threads_per_stream": t_each, "value": n_multi / t_multi, "unit": UNIT}
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": fps, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total_t / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8",
        "data": "synthetic",
        "config": {"workload": "C2: 1920x1080 8-bit 4:2:0, %d frames, CDEF+LR, single stream (%s.ivf)"
                               % (len(man["frames"]), CLIP)},
        "cpu_baseline": {"value": fps, "unit": UNIT, "cores": thr, "kind": "reference", "sample": sample},
        "e2e": {"value": fps, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        **({"e2e_multi_stream": multi} if multi else {}),
    }))


# ----------------------------------------------------------------------------- our arm

def e2e_pass(host, obu, cap, buf):
    """HOST OBU bytes -> HOST planes through the dav1d API of libdav1d_b200.so."""
    from easyav1_b200.lib import StreamFrame
    keep = C.create_string_buffer(obu, len(obu))
    t0 = time.perf_counter()
    h = host.eav1_stream_open(keep, len(obu), 1, 7, 1)
    if not h:
        raise RuntimeError("dav1d_open failed: no CUDA device (there is no CPU fallback)")
    eng = host.eav1rec_engine(host.eav1_stream_ctx(h))
    info, used, n = StreamFrame(), C.c_size_t(0), 0
    while True:
        r = host.eav1_stream_next(h, C.byref(info), buf, cap, C.byref(used))
        if r == 1:
            break
        if r < 0:
            raise RuntimeError("decoder error %d" % r)
        n += 1
    from easyav1_b200.lib import load_engine
    cnt = (C.c_longlong * 4)()
    load_engine().eav1cu_counters(eng, cnt)
    host.eav1_stream_close(h)
    dt = time.perf_counter() - t0
    return n, dt, hashlib.md5(buf.raw[:used.value]).hexdigest(), list(cnt)


def concurrent_passes(fn, n_streams, steps):
    """n_streams threads, each running `steps` passes of fn() (ctypes releases the GIL inside the
    decoders); returns (frames, wall seconds)."""
    frames = [0] * n_streams
    errs = []

    def work(i):
        try:
            for _ in range(steps):
                frames[i] += fn()
        except Exception as e:  # surfaced by the caller
            errs.append(e)

    ts = [threading.Thread(target=work, args=(i,)) for i in range(n_streams)]
    t0 = time.perf_counter()
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    dt = time.perf_counter() - t0
    if errs:
        raise errs[0]
    return sum(frames), dt


def stage_bytes(info, P):
    """Algorithmic HBM bytes per frame and stage (SURVEY.md §8(d), 8-bit samples). The
    "residual + prediction" row (read ref P + coefs + write P) is split over our two kernels:
    motion compensation reads one reference sample and writes one sample per predicted sample;
    the dataflow kernel reads the coefficient/palette payload actually shipped and reads +
    writes every sample its commands touch. Each post filter = read P + write P."""
    mc = 2 * info.mc_px
    # two-step transform blocks: phase 1b reads the coefficient payload and writes an int16 residual
    # per sample; the dataflow kernel reads that residual and reads + writes every sample it touches
    resid = info.payload_bytes + 2 * info.resid_px
    recon = 2 * info.resid_px + 2 * info.unit_px
    return [recon, 2 * P if info.n_lf else 0, 2 * P if info.cdef else 0, 2 * P if info.n_lr else 0, mc, 0, resid]


def run_ours(args, man, obu):
    import torch
    import torch.distributed as dist
    from easyav1_b200.lib import load_engine, load_host, StagedInfo, StreamFrame

    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the reconstruction path has no CPU fallback")
    torch.cuda.set_device(local)
    os.environ["EASYAV1_CUDA_DEVICE"] = str(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    E, host = load_engine(), load_host()
    n_frames = len(man["frames"])
    cap = man["w"] * man["h"] * 4
    buf = C.create_string_buffer(cap)
    P = man["w"] * man["h"] * 3 // 2

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- stage the stream once: decode with capture on, check parity of every frame
    keep = C.create_string_buffer(obu, len(obu))
    h = host.eav1_stream_open(keep, len(obu), 1, 7, 1)
    if not h:
        raise SystemExit("bench.py: dav1d_open failed (no CUDA device?)")
    eng = host.eav1rec_engine(host.eav1_stream_ctx(h))
    E.eav1cu_set_capture(eng, 1)
    info, used = StreamFrame(), C.c_size_t(0)
    for i in range(n_frames):
        r = host.eav1_stream_next(h, C.byref(info), buf, cap, C.byref(used))
        assert r == 0, "decode failed at frame %d (%d)" % (i, r)
        assert hashlib.md5(buf.raw[:used.value]).hexdigest() == man["frames"][i]["md5"], "parity lost at frame %d" % i
    E.eav1cu_set_capture(eng, 0)
    n_staged = E.eav1cu_staged_count(eng)
    infos = []
    for i in range(n_staged):
        si = StagedInfo()
        E.eav1cu_staged_info(eng, i, C.byref(si))
        infos.append(si)

    stream = torch.cuda.Stream()
    E.eav1cu_set_stream(eng, C.c_void_p(stream.cuda_stream))
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")   # > 126 MB L2

    def replay_step():
        for i in range(n_staged):
            r = E.eav1cu_staged_replay(eng, i)
            if r:
                raise RuntimeError("replay failed: %d" % r)

    def last_frame_md5():
        si = infos[-1]
        w, hh = si.w, si.h
        cw, ch = (w + 1) // 2, (hh + 1) // 2
        y = C.create_string_buffer(w * hh); u = C.create_string_buffer(cw * ch); v = C.create_string_buffer(cw * ch)
        data = (C.c_void_p * 3)(C.addressof(y), C.addressof(u), C.addressof(v))
        stride = (C.c_ssize_t * 2)(w, cw)
        assert E.eav1cu_pic_download(eng, si.cur, data, stride) == 0
        return hashlib.md5(y.raw + u.raw + v.raw).hexdigest()

    with torch.cuda.stream(stream):
        for _ in range(max(args.warmup, 3)):
            replay_step()
        torch.cuda.synchronize()
        assert last_frame_md5() == man["frames"][-1]["md5"], "replayed frames differ from the golden MD5"
        E.eav1cu_stage_times(eng, None, None, None, 1)
        E.eav1cu_set_stage_timing(eng, 1)
        cnt0 = (C.c_longlong * 4)(); E.eav1cu_counters(eng, cnt0)
        clocks = ClockSampler(local)
        clocks.start()
        barrier()
        prof = os.environ.get("EAV1_BENCH_PROFILE")     # ncu --profile-from-start off: first timed step only

        def timed_steps(profile_first):
            total = 0.0
            for step in range(args.steps):
                flush.fill_(1)                      # evict L2 between timed steps
                torch.cuda.synchronize()
                if profile_first and step == 0:
                    torch.cuda.profiler.start()
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record(stream)
                replay_step()
                b.record(stream)
                b.synchronize()
                if profile_first and step == 0:
                    torch.cuda.profiler.stop()
                total += a.elapsed_time(b)
            return total

        # `value`: K steps with nothing but the kernels on the stream. The per-stage breakdown (and
        # the roofline of the dominant kernel) comes from K more steps of the same replay with a CUDA
        # event pair around every stage; those ~14 extra event records per frame cost a few percent,
        # which is why they are kept out of the headline number (both totals are reported).
        E.eav1cu_set_stage_timing(eng, 0)
        dev_ms = timed_steps(bool(prof))
        cnt_mid = (C.c_longlong * 4)(); E.eav1cu_counters(eng, cnt_mid)
        E.eav1cu_stage_times(eng, None, None, None, 1)
        E.eav1cu_set_stage_timing(eng, 1)
        dev_ms_staged = timed_steps(False)
        barrier()
        clk = clocks.stop()
        cnt1 = (C.c_longlong * 4)(); E.eav1cu_counters(eng, cnt1)
        ms = (C.c_double * 7)(); ln = (C.c_longlong * 7)(); calls = (C.c_longlong * 7)()
        E.eav1cu_stage_times(eng, ms, ln, calls, 1)
        E.eav1cu_set_stage_timing(eng, 0)
        assert last_frame_md5() == man["frames"][-1]["md5"], "timed replay differs from the golden MD5"
    dev_ms = max_over_ranks(dev_ms)
    launches = int(cnt_mid[0] - cnt0[0])     # kernels launched inside the K timed steps of `value`
    torch.cuda.synchronize()
    host.eav1_stream_close(h)

    # ---- end to end through the dav1d API: host OBUs in, host planes out
    for _ in range(max(args.warmup, 3)):
        e2e_pass(host, obu, cap, buf)
    barrier()
    e2e_t, e2e_n, h2d, d2h = 0.0, 0, 0, 0
    t_all0 = time.perf_counter()
    for _ in range(args.steps):
        n, dt, md5, cnt = e2e_pass(host, obu, cap, buf)
        assert md5 == man["frames"][-1]["md5"], "e2e decode differs from the golden MD5"
        e2e_t += dt; e2e_n += n; h2d += cnt[1]; d2h += cnt[2]
    barrier()
    e2e_t = max_over_ranks(time.perf_counter() - t_all0)

    # ---- optional: S independent streams per GPU, end to end (streams shard, nothing is exchanged)
    multi = None
    if args.streams > 1:
        thr = max(2, host_threads() // args.streams)
        old = os.environ.get("EASYAV1_THREADS")
        os.environ["EASYAV1_THREADS"] = str(thr)
        bufs = [C.create_string_buffer(cap) for _ in range(args.streams)]
        idx = iter(range(1 << 30))
        lock = threading.Lock()

        def one_pass():
            with lock:
                b = bufs[next(idx) % args.streams]
            n, _, md5, _ = e2e_pass(host, obu, cap, b)
            assert md5 == man["frames"][-1]["md5"], "multi-stream decode differs from the golden MD5"
            return n

        concurrent_passes(one_pass, args.streams, 1)     # warm
        barrier()
        n_multi, t_multi = concurrent_passes(one_pass, args.streams, args.steps)
        t_multi = max_over_ranks(t_multi)
        multi = {"streams_per_gpu": args.streams, "threads_per_stream": thr, "value": world * n_multi / t_multi, "unit": UNIT}
        if old is None:
            os.environ.pop("EASYAV1_THREADS", None)
        else:
            os.environ["EASYAV1_THREADS"] = old

    # ---- roofline of the stage with the largest share of device time
    KS = [0, 1, 2, 3, 4, 6]
    tot_bytes = [0] * 7
    for si in infos:
        sb = stage_bytes(si, P)
        for k in KS:
            tot_bytes[k] += sb[k]
    names = ["recon_frame_kernel (dataflow: intra prediction, residual add, compound, blend)", "deblock_kernel (cols+rows)",
             "cdef_kernel", "lr_kernel (+ plane copy)", "mc_kernel (motion compensation)", "fg (unused here)",
             "resid_kernel (inverse transforms -> residual blocks)"]
    stage_ms = [ms[k] / args.steps for k in range(7)]          # per step (= per 120 frames)
    dom = max(KS, key=lambda k: stage_ms[k])
    peak, peak_src = measured_peak()
    achieved = (tot_bytes[dom] / 1e9) / (stage_ms[dom] / 1e3) if stage_ms[dom] > 0 else 0.0
    stages = {names[k].split(" ")[0]: {"ms_per_step": round(stage_ms[k], 4),
                                       "launches_per_step": int(ln[k] // args.steps),
                                       "alg_bytes_per_step": int(tot_bytes[k]),
                                       "gbs": round((tot_bytes[k] / 1e9) / (stage_ms[k] / 1e3), 2) if stage_ms[k] > 0 else None}
              for k in KS}
    all_bytes = sum(tot_bytes)
    out = {
        "metric": METRIC, "value": world * n_frames * args.steps / (dev_ms / 1e3), "unit": UNIT,
        "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": dev_ms / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": "C2: 1920x1080 8-bit 4:2:0, %d frames, CDEF+LR, single stream per GPU (%s.ivf)"
                               % (n_frames, CLIP),
                   "step": "one pass over the %d-frame stream" % n_frames,
                   "l2": "256 MiB flush write between timed steps", "streams_per_gpu": 1,
                   "ms_per_step_with_stage_events": round(dev_ms_staged / args.steps, 3)},
        "roofline": {"bound": "hbm", "kernel": names[dom], "achieved": round(achieved, 2), "peak": peak,
                     "unit": "GB/s", "frac": round(achieved / peak, 5),
                     "traffic": (ncu_traffic(names[dom].split(" ")[0]) or {}).get("bytes_per_launch"),
                     "traffic_source": (ncu_traffic(names[dom].split(" ")[0]) or {}).get("source"),
                     "peak_source": peak_src,
                     "alg_bytes_per_launch": int(tot_bytes[dom] / max(1, calls[dom] // args.steps)),
                     "whole_frame_gbs": round((all_bytes / 1e9) / (dev_ms / args.steps / 1e3), 2),
                     "stages": stages},
        "e2e": {"value": world * e2e_n / e2e_t, "unit": UNIT, "h2d_bytes_per_step": int(h2d // args.steps),
                "d2h_bytes_per_step": int(d2h // args.steps)},
        "gpu_launches": launches,
        "clocks": clk,
    }
    if multi:
        out["e2e_multi_stream"] = multi
    if rank == 0:
        thr = host_threads()
        t0 = time.perf_counter()
        n, dt, md5 = ref_decode_pass(obu, cap, thr)     # warm
        passes, tt, nn = 0, 0.0, 0
        while passes < 1 or (tt < 10.0 and passes < 20):
            n, dt, md5 = ref_decode_pass(obu, cap, thr)
            tt += dt; nn += n; passes += 1
        out["cpu_baseline"] = {"value": nn / tt, "unit": UNIT, "cores": thr, "kind": "reference",
                               "sample": "%d passes of the same %d-frame stream through oracle/_ref "
                                         "(reference dav1d C build, n_threads=%d)" % (passes, n_frames, thr)}
        print(json.dumps(out))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--streams", type=int, default=1,
                    help="also measure S concurrent decoders per GPU end to end (BASELINE.json configs[4]); "
                         "adds an e2e_multi_stream object, the headline keys stay single-stream")
    args = ap.parse_args()
    man, obu = load_clip()
    if args.impl == "reference":
        run_reference(args, man, obu)
    else:
        run_ours(args, man, obu)


if __name__ == "__main__":
    main()
