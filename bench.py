#!/usr/bin/env python3
"""bench.py — AV1 decoded frames/s on the BASELINE.json configs, one JSON line.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--only c2,c4,c3,c1,c5]

N = 1. Top-level keys = configs[1] (C2): tests/golden/c2_1080p_bench.ivf, 1920x1080 8-bit 4:2:0, 120 frames,
CDEF + loop restoration, one stream; one "step" = one pass of the reconstruction path over the stream.
  value         frames/s with every work list already resident in HBM: the staged frames are re-launched in
                decode order, CUDA events on the launching stream, L2 flushed between steps.
  e2e           frames/s through the reference-facing API (dav1d_open / dav1d_send_data / dav1d_get_picture of
                libdav1d_b200.so) from HOST OBU bytes to HOST planes; entropy decode, H2D, kernels, D2H all timed.
  roofline      the stage with the largest share of device time (algorithmic bytes of SURVEY.md §8(d)).
  cpu_baseline  the reference's own dav1d C build (oracle/_ref) on the host cores, same stream.
  configs       the other BASELINE.json configs, each with value / e2e / roofline / cpu_baseline / parity:
                C4 3840x2160 10-bit (+ film grain on every shown frame), C3 Vorbis IMDCT / window / overlap-add,
                C1 the easyAV1 C99 API on a 640x360 WebM (the reference's own benchmark loop), C5 64 concurrent
                1080p streams (device-resident replay of S = 1 / 4 / 8 / 16 staged streams, end to end with 64).
N > 1 (torchrun, one process per GPU, gloo for the barrier and the MAX over ranks; nothing is exchanged on the
data path): the top-level keys are configs[4] (C5): 64 streams dealt round-robin to the ranks
(easyav1_b200.shard.assign), 64 / N decoders per rank sharing that rank's share of the host cores.
--impl reference times the reference's CPU implementation the same way: one stream on all host threads at
N = 1, 64 concurrent reference decoders sharing the host cores at N > 1 (rank 0 only).
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
CLIP4 = "c4_2160p_10bit_bench"
CLIP1 = "c1_360p_bench"
METRIC = "AV1 decoded frames/s (1080p)"
UNIT = "frames/s"
N_STREAMS_C5 = 64


def workload_c2(n_frames):
    return "C2: 1920x1080 8-bit 4:2:0, %d frames, CDEF+LR, single stream (%s.ivf)" % (n_frames, CLIP)


def workload_c5(n):
    return "C5: %d concurrent 1920x1080 8-bit streams (copies of %s.ivf), sharded by stream over %d GPU(s)" % (
        N_STREAMS_C5, CLIP, n)


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
    `ncu --set full` summary under profiles/; None if there is none. A profiler figure (cold cache, serialised
    launches), reported beside the live numbers."""
    import glob
    best = None
    for path in sorted(glob.glob(os.path.join(ROOT, "profiles", "r*_full_summary.json"))):
        try:
            d = json.load(open(path))
        except Exception:
            continue
        for k, v in d.items():
            if kernel_name in k:
                def in_bytes(prefix):     # ncu picks a unit per metric ("[Mbyte]", "[Kbyte]", ...)
                    for n, x in v.items():
                        if n.startswith(prefix):
                            return x * (1e6 if "Mbyte" in n else 1e3 if "Kbyte" in n else 1e9 if "Gbyte" in n else 1.0)
                    return None
                rd, wr = in_bytes("dram__bytes_read.sum"), in_bytes("dram__bytes_write.sum")
                if rd is not None and wr is not None:
                    best = {"bytes_per_launch": int(rd + wr), "source": os.path.basename(path),
                            "launches_captured": v.get("launches_captured")}
    return best


class ClockSampler:
    """nvidia-smi clocks / throttle reasons / utilisation during the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap,utilization.gpu")
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
        sm, mx, reasons, util = [], None, set(), []
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx = float(r[1])
                for i, n in enumerate(self.NAMES):
                    if r[2 + i].lower().startswith("active"):
                        reasons.add(n)
                util.append(float(r[6]))
            except Exception:
                continue
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm), "gpu_util_mean": round(sum(util) / len(util), 1) if util else None}


def host_threads():
    try:
        n = len(os.sched_getaffinity(0))
    except Exception:
        n = os.cpu_count() or 1
    return max(1, min(n, 256))   # DAV1D_MAX_THREADS (ext/dav1d/src/internal.h)


def concurrent_passes(fn, n_streams, steps):
    """n_streams threads, each running `steps` passes of fn(i) (ctypes releases the GIL inside the decoders);
    returns (per-stream results of the last pass, frames, wall seconds)."""
    frames = [0] * n_streams
    last = [None] * n_streams
    errs = []

    def work(i):
        try:
            for _ in range(steps):
                n, res = fn(i)
                frames[i] += n
                last[i] = res
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
    return last, sum(frames), dt


# ----------------------------------------------------------------------------- reference (CPU) legs

def ref_decode_pass(obu, cap, n_threads, buf=None, keep=None):
    """One pass of the reference's dav1d C build (oracle/_ref) over the stream: host OBUs -> host planes packed
    into a caller buffer, exactly what our e2e leg does. Returns (frames, seconds, md5 of the last frame)."""
    from tests import refdec
    L = refdec.lib()
    keep = keep or C.create_string_buffer(obu, len(obu))
    buf = buf or C.create_string_buffer(cap)
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


def ref_bounded(obu, cap, thr, man, budget_s=10.0, max_passes=20):
    """cpu_baseline sample: whole passes of the stream for about budget_s seconds"""
    keep, buf = C.create_string_buffer(obu, len(obu)), C.create_string_buffer(cap)
    ref_decode_pass(obu, cap, thr, buf, keep)     # warm
    passes, tt, nn, md5 = 0, 0.0, 0, None
    while passes < 1 or (tt < budget_s and passes < max_passes):
        n, dt, md5 = ref_decode_pass(obu, cap, thr, buf, keep)
        tt += dt; nn += n; passes += 1
    assert md5 == man["frames"][-1]["md5"], "reference decode does not match its own golden MD5"
    return nn / tt, passes


def ref_multi_stream(obu, cap, n_streams, passes, man):
    """n_streams concurrent reference decoders sharing the host cores"""
    thr = max(1, host_threads() // n_streams)
    keep = C.create_string_buffer(obu, len(obu))
    bufs = [C.create_string_buffer(cap) for _ in range(n_streams)]

    def one(i):
        n, _, md5 = ref_decode_pass(obu, cap, thr, bufs[i], keep)
        return n, md5
    last, n, dt = concurrent_passes(one, n_streams, passes)
    assert all(m == man["frames"][-1]["md5"] for m in last), "reference multi-stream decode differs from the golden MD5"
    return n / dt, thr


def ref_vorbis(flags, spectra, channels, bs):
    """minivorbis (oracle/_ref/libvorbis_ref.so): mdct_backward + vorbis_synthesis_blockin over the batch"""
    from tests import vorbis_ops as vo
    vo.ref_decode(flags[:64], spectra, channels, bs)       # warm (tables)
    t0 = time.perf_counter()
    pcm = vo.ref_decode(flags, spectra, channels, bs)
    return pcm, time.perf_counter() - t0


def easyav1_api_pass(L, webm_keep, webm_len, want_md5=None):
    """The loop of the reference's own benchmark tool (tools/easyav1_benchmark.c:124-185): init from memory,
    easyav1_decode_next until it stops, easyav1_get_video_frame for every frame; ns clock instead of ms."""
    from tests import easyav1_api as api
    st = L.easyav1_default_settings()
    st.enable_audio = 0
    st.log_level = 1
    t0 = time.perf_counter()
    h = L.easyav1_init_from_memory(webm_keep, webm_len, C.byref(st))
    if not h:
        raise RuntimeError("easyav1_init_from_memory failed")
    n, last = 0, None
    while L.easyav1_decode_next(h) == 1:
        if L.easyav1_has_video_frame(h) == 1:
            last = L.easyav1_get_video_frame(h)
            n += 1
    md5 = api.frame_md5(last.contents, 8, 1) if (want_md5 and last) else None
    hp = C.c_void_p(h)
    L.easyav1_destroy(C.byref(hp))
    dt = time.perf_counter() - t0
    if want_md5:
        assert md5 == want_md5, "easyAV1 API: last frame differs from the golden MD5"
    return n, dt


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if rank != 0:
        return
    man, obu = load_clip()
    cap = man["w"] * man["h"] * 4
    thr = host_threads()
    n_frames = len(man["frames"])
    out = {"impl": "reference", "metric": METRIC, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
           "warmup": args.warmup, "higher_is_better": True, "vs_baseline": None, "dtype": "u8", "data": "synthetic",
           "gpu_launches": 0}
    if world > 1 or args.gpus > 1:
        # configs[4]: the same 64 streams our arm decodes, as 64 concurrent reference decoders on the host cores
        ref_multi_stream(obu, cap, N_STREAMS_C5, 1, man)     # warm
        total_n, total_t = 0, 0.0
        for _ in range(max(1, min(args.steps, 3))):
            t0 = time.perf_counter()
            fps, t_each = ref_multi_stream(obu, cap, N_STREAMS_C5, 1, man)
            total_t += time.perf_counter() - t0
            total_n += N_STREAMS_C5 * n_frames
        fps = total_n / total_t
        out.update({"value": fps, "ms_per_step": 1e3 * total_t / max(1, min(args.steps, 3)), "scaling": "strong",
                    "config": {"workload": workload_c5(args.gpus), "step": "one pass of all 64 streams",
                               "threads_per_stream": t_each, "host_threads": thr},
                    "cpu_baseline": {"value": fps, "unit": UNIT, "cores": thr, "kind": "reference",
                                     "sample": "%d passes of 64 concurrent %d-frame streams, dav1d C (no asm), %d thread(s) each"
                                               % (max(1, min(args.steps, 3)), n_frames, t_each)},
                    "e2e": {"value": fps, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}})
        print(json.dumps(out))
        return
    keep, buf = C.create_string_buffer(obu, len(obu)), C.create_string_buffer(cap)
    for _ in range(args.warmup):
        ref_decode_pass(obu, cap, thr, buf, keep)
    total_t, total_n, md5 = 0.0, 0, None
    for _ in range(args.steps):
        n, dt, md5 = ref_decode_pass(obu, cap, thr, buf, keep)
        total_t += dt
        total_n += n
    assert md5 == man["frames"][-1]["md5"], "reference decode does not match its own golden MD5"
    fps = total_n / total_t
    out.update({"value": fps, "ms_per_step": 1e3 * total_t / args.steps, "scaling": "weak",
                "config": {"workload": workload_c2(n_frames), "step": "one pass over the %d-frame stream" % n_frames},
                "cpu_baseline": {"value": fps, "unit": UNIT, "cores": thr, "kind": "reference",
                                 "sample": "%d passes of the %d-frame %s stream, dav1d C (no asm), n_threads=%d, frame delay auto"
                                           % (args.steps, n_frames, CLIP, thr)},
                "e2e": {"value": fps, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}})
    print(json.dumps(out))


# ----------------------------------------------------------------------------- our arm

def e2e_pass(host, obu, cap, buf, keep=None, gate=None):
    """HOST OBU bytes -> HOST planes through the dav1d API of libdav1d_b200.so. Returns (frames, seconds,
    md5 of the last frame, engine counters); the md5 is computed outside the timed span."""
    from easyav1_b200.lib import StreamFrame, load_engine
    keep = keep or C.create_string_buffer(obu, len(obu))
    t0 = time.perf_counter()
    h = host.eav1_stream_open(keep, len(obu), 1, 7, 1)
    if not h:
        raise RuntimeError("dav1d_open failed: no CUDA device (there is no CPU fallback)")
    eng = host.eav1rec_engine(host.eav1_stream_ctx(h))
    info, used, n = StreamFrame(), C.c_size_t(0), 0
    while True:
        if gate is not None:
            with gate:      # at most `gate` streams work on a frame at the same time (the others are parked, not spinning)
                r = host.eav1_stream_next(h, C.byref(info), buf, cap, C.byref(used))
        else:
            r = host.eav1_stream_next(h, C.byref(info), buf, cap, C.byref(used))
        if r == 1:
            break
        if r < 0:
            raise RuntimeError("decoder error %d" % r)
        n += 1
    cnt = (C.c_longlong * 4)()
    load_engine().eav1cu_counters(eng, cnt)
    host.eav1_stream_close(h)
    dt = time.perf_counter() - t0
    return n, dt, hashlib.md5(buf.raw[:used.value]).hexdigest(), list(cnt)


def stage_bytes(info, P):
    """Algorithmic HBM bytes per frame and stage; P = bytes of all planes. Motion compensation
    reads one reference sample and writes one sample per predicted sample; phase 1b reads the coefficient payload
    and writes an int16 residual per sample; the dataflow kernel reads that residual and reads + writes every
    sample its commands touch; each post filter = read P + write P."""
    bps = 2 if info.bpc > 8 else 1
    mc = 2 * bps * info.mc_px
    resid = info.payload_bytes + 2 * info.resid_px
    recon = 2 * info.resid_px + 2 * bps * info.unit_px
    return [recon, 2 * P if info.n_lf else 0, 2 * P if info.cdef else 0, 2 * P if info.n_lr else 0, mc, 0, resid]


STAGE_NAMES = ["recon_frame_kernel (dataflow: intra prediction, residual add, compound, blend)", "deblock_kernel (cols+rows)",
               "cdef_kernel", "lr_kernel (+ plane copy)", "mc_kernel (motion compensation)", "fg_prepare_kernel + fg_apply_kernel (film grain)",
               "resid_kernel (inverse transforms -> residual blocks)"]


class Staged:
    """One decoder whose frames are resident on the device and can be replayed with no host work."""

    def __init__(self, E, host, man, obu, check=True):
        from easyav1_b200.lib import StagedInfo, StreamFrame
        self.E, self.host, self.man = E, host, man
        self.keep = C.create_string_buffer(obu, len(obu))
        cap = man["w"] * man["h"] * 4
        buf = C.create_string_buffer(cap)
        self.h = host.eav1_stream_open(self.keep, len(obu), 1, 7, 1)
        if not self.h:
            raise SystemExit("bench.py: dav1d_open failed (no CUDA device?)")
        self.eng = host.eav1rec_engine(host.eav1_stream_ctx(self.h))
        E.eav1cu_set_capture(self.eng, 1)
        info, used = StreamFrame(), C.c_size_t(0)
        for i in range(len(man["frames"])):
            r = host.eav1_stream_next(self.h, C.byref(info), buf, cap, C.byref(used))
            assert r == 0, "decode failed at frame %d (%d)" % (i, r)
            if check:
                assert hashlib.md5(buf.raw[:used.value]).hexdigest() == man["frames"][i]["md5"], "parity lost at frame %d" % i
        host.eav1rec_drain(host.eav1_stream_ctx(self.h))
        E.eav1cu_set_capture(self.eng, 0)
        self.n_staged = E.eav1cu_staged_count(self.eng)
        self.infos = []
        for i in range(self.n_staged):
            si = StagedInfo()
            E.eav1cu_staged_info(self.eng, i, C.byref(si))
            self.infos.append(si)

    def set_stream(self, torch_stream):
        self.E.eav1cu_set_stream(self.eng, C.c_void_p(torch_stream.cuda_stream))

    def replay(self, passes=1):
        r = self.E.eav1cu_staged_replay_all(self.eng, passes)
        if r:
            raise RuntimeError("replay failed: %d" % r)

    def graph_build(self, grain_pic=None, grain=None):
        if grain is None:
            r = self.E.eav1cu_staged_graph_build(self.eng)
        else:       # film grain applied to every staged frame inside the same graph
            r = self.E.eav1cu_staged_graph_build_grain(self.eng, grain_pic, C.byref(grain))
        if r:
            raise RuntimeError("graph capture failed: %d" % r)

    def graph_replay(self, passes=1):
        r = self.E.eav1cu_staged_graph_launch(self.eng, passes)
        if r:
            raise RuntimeError("graph replay failed: %d" % r)

    def last_frame_md5(self, pic=None):
        """MD5 of the planes of the last staged frame's picture (or of picture id `pic` of the same geometry)"""
        si = self.infos[-1]
        w, hh, bps = si.w, si.h, 2 if si.bpc > 8 else 1
        cw, ch = (w + 1) // 2, (hh + 1) // 2
        y = C.create_string_buffer(w * hh * bps); u = C.create_string_buffer(cw * ch * bps); v = C.create_string_buffer(cw * ch * bps)
        data = (C.c_void_p * 3)(C.addressof(y), C.addressof(u), C.addressof(v))
        stride = (C.c_ssize_t * 2)(w * bps, cw * bps)
        assert self.E.eav1cu_pic_download(self.eng, si.cur if pic is None else pic, data, stride) == 0
        return hashlib.md5(y.raw + u.raw + v.raw).hexdigest()

    def close(self):
        if self.h:
            self.host.eav1_stream_close(self.h)
            self.h = None


def device_leg(E, host, man, obu, steps, warmup, torch, grain=None, prof=False):
    """Device-resident replay of one staged stream: (ms per step, per-stage ms/launches/calls, stage bytes, infos,
    launches per step, ms per step with stage events). grain: apply film grain to every shown frame as well."""
    st = Staged(E, host, man, obu)
    stream = torch.cuda.Stream()
    st.set_stream(stream)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")   # > 126 MB L2
    n_staged = st.n_staged
    shown = [i for i in range(n_staged)]
    gp = None
    if grain is not None:
        from tests import ops
        E.eav1cu_pic_alloc.restype = C.c_int
        E.eav1cu_pic_alloc.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int]
        E.eav1cu_apply_grain.argtypes = [C.c_void_p, C.c_int, C.c_int, C.POINTER(ops.GrainParams)]
        si = st.infos[0]
        busy = {x.cur for x in st.infos}       # the decoder has released these ids, the staged frames still use them
        gp = -1
        for _ in range(64):
            gp = E.eav1cu_pic_alloc(st.eng, si.w, si.h, si.layout, si.bpc)
            assert gp >= 0
            if gp not in busy:
                break
        assert gp not in busy

    use_graph = [False]

    def replay_step():
        if use_graph[0]:
            st.graph_replay(1)     # all frames of the step (and their film grain) as one CUDA graph, built once outside the timed region
            return
        if grain is None:
            st.replay(1)
            return
        for i in shown:
            r = E.eav1cu_staged_replay(st.eng, i)
            if r:
                raise RuntimeError("replay failed: %d" % r)
            r = E.eav1cu_apply_grain(st.eng, st.infos[i].cur, gp, C.byref(grain))
            if r:
                raise RuntimeError("apply_grain failed: %d" % r)

    def timed(n_steps, profile_first=False):
        total = 0.0
        for step in range(n_steps):
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

    with torch.cuda.stream(stream):
        for _ in range(max(warmup, 3)):
            replay_step()
        torch.cuda.synchronize()
        assert st.last_frame_md5() == man["frames"][-1]["md5"], "replayed frames differ from the golden MD5"
        E.eav1cu_set_stage_timing(st.eng, 0)
        if not os.environ.get("EAV1_BENCH_NO_GRAPH"):
            grained = st.last_frame_md5(gp) if grain is not None else None      # the last frame with grain, direct launches
            st.graph_build(gp, grain) if grain is not None else st.graph_build()
            use_graph[0] = True
            for _ in range(2):
                replay_step()
            torch.cuda.synchronize()
            assert st.last_frame_md5() == man["frames"][-1]["md5"], "graph replay differs from the golden MD5"
            if grain is not None:
                assert st.last_frame_md5(gp) == grained, "film grain inside the graph differs from the direct launches"
        cnt0 = (C.c_longlong * 4)(); E.eav1cu_counters(st.eng, cnt0)
        dev_ms = timed(steps, prof)
        graph_used = use_graph[0]
        use_graph[0] = False    # the instrumented pass below needs direct launches (event pairs around the stages)
        cnt1 = (C.c_longlong * 4)(); E.eav1cu_counters(st.eng, cnt1)
        # per-stage breakdown: the same replay with a CUDA event pair around every stage (kept out of `value`)
        E.eav1cu_stage_times(st.eng, None, None, None, 1)
        E.eav1cu_set_stage_timing(st.eng, 1)
        dev_ms_staged = timed(steps)
        ms = (C.c_double * 7)(); ln = (C.c_longlong * 7)(); calls = (C.c_longlong * 7)()
        E.eav1cu_stage_times(st.eng, ms, ln, calls, 1)
        E.eav1cu_set_stage_timing(st.eng, 0)
        assert st.last_frame_md5() == man["frames"][-1]["md5"], "timed replay differs from the golden MD5"
    torch.cuda.synchronize()
    infos = st.infos
    bps = 2 if infos[0].bpc > 8 else 1
    P = man["w"] * man["h"] * 3 // 2 * bps
    tot_bytes = [0] * 7
    for si in infos:
        sb = stage_bytes(si, P)
        for k in range(7):
            tot_bytes[k] += sb[k]
    if grain is not None:
        tot_bytes[5] = 2 * P * len(shown)
    res = dict(ms_per_step=dev_ms / steps, ms_per_step_events=dev_ms_staged / steps,
               stage_ms=[ms[k] / steps for k in range(7)], stage_launches=[int(ln[k] // steps) for k in range(7)],
               stage_calls=[int(calls[k] // steps) for k in range(7)], stage_bytes=tot_bytes,
               launches=int(cnt1[0] - cnt0[0]), replay="one CUDA graph per step" if graph_used else "direct launches", n_frames=len(man["frames"]), n_staged=n_staged,
               coded_fraction=sum(si.resid_px for si in infos) / float(sum(1 for _ in infos) * man["w"] * man["h"] * 3 // 2),
               payload_bytes=int(sum(si.payload_bytes for si in infos)))
    st.close()
    del flush
    return res


def roofline_of(dev, names=STAGE_NAMES):
    peak, peak_src = measured_peak()
    KS = [k for k in range(7) if dev["stage_ms"][k] > 0]
    dom = max(KS, key=lambda k: dev["stage_ms"][k])
    gbs = lambda k: (dev["stage_bytes"][k] / 1e9) / (dev["stage_ms"][k] / 1e3) if dev["stage_ms"][k] > 0 else None
    stages = {names[k].split(" ")[0]: {"ms_per_step": round(dev["stage_ms"][k], 4), "launches_per_step": dev["stage_launches"][k],
                                       "alg_bytes_per_step": int(dev["stage_bytes"][k]), "gbs": round(gbs(k), 2)} for k in KS}
    achieved = gbs(dom)
    tr = ncu_traffic(names[dom].split(" ")[0]) or {}
    return {"bound": "hbm", "kernel": names[dom], "achieved": round(achieved, 2), "peak": peak, "unit": "GB/s",
            "frac": round(achieved / peak, 5), "traffic": tr.get("bytes_per_launch"), "traffic_source": tr.get("source"),
            "peak_source": peak_src,
            "alg_bytes_per_launch": int(dev["stage_bytes"][dom] / max(1, dev["stage_calls"][dom])),
            "whole_frame_gbs": round((sum(dev["stage_bytes"]) / 1e9) / (dev["ms_per_step"] / 1e3), 2),
            "whole_frame_frac": round((sum(dev["stage_bytes"]) / 1e9) / (dev["ms_per_step"] / 1e3) / peak, 5),
            "stages": stages}


def multi_stream_device(E, host, man, obu, n_streams, passes, torch, graph=True, sampler=None):
    """S staged decoders, each on its own CUDA stream and host thread, replayed together: aggregate frames/s and
    GB/s of the device alone (inputs resident in HBM). One start event, one end event per stream; the elapsed time
    is the latest end."""
    sts = [Staged(E, host, man, obu, check=(i == 0)) for i in range(n_streams)]
    streams = [torch.cuda.Stream() for _ in range(n_streams)]
    for st, s in zip(sts, streams):
        st.set_stream(s)
        if graph:
            st.graph_build()

    def run(p):
        if graph:               # one cudaGraphLaunch per stream and pass: a single host thread keeps all streams fed
            for _ in range(p):
                for st in sts:
                    st.graph_replay(1)
            return
        ts = [threading.Thread(target=st.replay, args=(p,)) for st in sts]
        for t in ts:
            t.start()
        for t in ts:
            t.join()

    run(1)                                   # warm
    torch.cuda.synchronize()
    for st in sts:
        assert st.last_frame_md5() == man["frames"][-1]["md5"], "multi-stream replay differs from the golden MD5"
    cnt0 = [(C.c_longlong * 4)() for _ in sts]
    for st, c in zip(sts, cnt0):
        E.eav1cu_counters(st.eng, c)
    a = torch.cuda.Event(enable_timing=True)
    ends = [torch.cuda.Event(enable_timing=True) for _ in sts]
    torch.cuda.synchronize()
    if sampler:
        sampler.start()
    a.record(streams[0])
    run(passes)
    for e, s in zip(ends, streams):
        e.record(s)
    torch.cuda.synchronize()
    clk = sampler.stop() if sampler else None
    ms = max(a.elapsed_time(e) for e in ends)
    launches = 0
    for st, c in zip(sts, cnt0):
        c1 = (C.c_longlong * 4)(); E.eav1cu_counters(st.eng, c1)
        launches += int(c1[0] - c[0])
    for st in sts:
        assert st.last_frame_md5() == man["frames"][-1]["md5"], "multi-stream replay differs from the golden MD5"
    bps = 2 if sts[0].infos[0].bpc > 8 else 1
    P = man["w"] * man["h"] * 3 // 2 * bps
    alg = sum(sum(stage_bytes(si, P)) for si in sts[0].infos) * n_streams * passes
    for st in sts:
        st.close()
    frames = n_streams * passes * len(man["frames"])
    return {"streams": n_streams, "passes": passes, "value": frames / (ms / 1e3), "unit": UNIT, "ms": round(ms, 3),
            "whole_frame_gbs": round(alg / 1e9 / (ms / 1e3), 1), "gpu_launches": launches,
            "replay": "one CUDA graph per stream (all frames captured once)" if graph else "direct launches, one host thread per stream",
            **({"gpu_util_mean": clk.get("gpu_util_mean"), "sm_mhz": clk.get("sm_mhz")} if clk else {})}


def pump_e2e(host, man, obu, n_streams, passes, workers):
    """n_streams decoders multiplexed on `workers` threads by host/pump.c (host OBUs -> host planes): every decoder
    is single-threaded, a worker feeds whichever stream is free and retires its pictures a few frames later.
    Returns (md5 of each stream's last picture, frames, seconds, engine byte counters are not split per stream)."""
    from easyav1_b200.lib import pump_run
    cap = man["w"] * man["h"] * 4
    bufs = [C.create_string_buffer(cap) for _ in range(n_streams)]
    obus = [obu] * n_streams
    from easyav1_b200.lib import load_engine
    E = load_engine()
    c0 = (C.c_longlong * 4)(); E.eav1cu_global_counters(c0)
    n, t0, last = 0, time.perf_counter(), None
    for _ in range(passes):
        frames, last = pump_run(host, obus, workers, cap, bufs=bufs)
        n += sum(frames)
    dt = time.perf_counter() - t0
    c1 = (C.c_longlong * 4)(); E.eav1cu_global_counters(c1)
    stats = {"launches": int(c1[0] - c0[0]), "h2d": int(c1[1] - c0[1]), "d2h": int(c1[2] - c0[2])}
    return [hashlib.md5(x).hexdigest() for x in last], n, dt, stats


def multi_stream_e2e(host, man, obu, n_streams, passes, threads_each, active=None):
    """n_streams concurrent decoders (host OBUs -> host planes), one Python thread each"""
    cap = man["w"] * man["h"] * 4
    keep = C.create_string_buffer(obu, len(obu))
    bufs = [C.create_string_buffer(cap) for _ in range(n_streams)]
    old = os.environ.get("EASYAV1_THREADS")
    os.environ["EASYAV1_THREADS"] = str(threads_each)
    stats = {"h2d": 0, "d2h": 0}
    lock = threading.Lock()

    gate = threading.BoundedSemaphore(active) if active and active < n_streams else None

    def one(i):
        n, _, md5, cnt = e2e_pass(host, obu, cap, bufs[i], keep, gate)
        with lock:
            stats["h2d"] += cnt[1]; stats["d2h"] += cnt[2]
        return n, md5
    try:
        last, n, dt = concurrent_passes(one, n_streams, passes)
    finally:
        if old is None:
            os.environ.pop("EASYAV1_THREADS", None)
        else:
            os.environ["EASYAV1_THREADS"] = old
    return last, n, dt, stats


def vorbis_leg(E, torch):
    """configs[2] (C3): 4096 blocks x 6 channels, blocksizes 256 / 2048, IMDCT + window + overlap-add"""
    import numpy as np
    from tests import vorbis_ops as vo
    channels, bs, n_blocks = 6, (256, 2048), 4096
    flags, spectra = vo.make_batch(3, channels, bs, n_blocks)
    E.eav1cu_vorbis_open.argtypes = [C.POINTER(C.c_void_p), C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
    E.eav1cu_vorbis_stage.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.POINTER(C.c_size_t)]
    E.eav1cu_vorbis_run_staged.argtypes = [C.c_void_p]
    E.eav1cu_vorbis_set_stream.argtypes = [C.c_void_p, C.c_void_p]
    E.eav1cu_vorbis_restart.argtypes = [C.c_void_p]
    E.eav1cu_vorbis_close.argtypes = [C.c_void_p]
    E.eav1cu_vorbis_launch_count.restype = C.c_longlong
    E.eav1cu_vorbis_launch_count.argtypes = [C.c_void_p]
    E.eav1cu_vorbis_imdct_ola.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_size_t, C.c_int,
                                          C.POINTER(C.c_size_t)]
    h = C.c_void_p()
    assert E.eav1cu_vorbis_open(C.byref(h), torch.cuda.current_device(), channels, bs[0], bs[1], None, None) == 0
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
    peak, peak_src = measured_peak()
    cap = int(n.value) + bs[1]
    out = np.zeros((channels, cap), np.float32)
    t_host = []
    for it in range(6):
        E.eav1cu_vorbis_restart(h)
        t0 = time.perf_counter()
        assert E.eav1cu_vorbis_imdct_ola(h, flags.ctypes.data, n_blocks, spectra.ctypes.data, out.ctypes.data, cap, 0,
                                         C.byref(n)) == 0
        t_host.append(time.perf_counter() - t0)
    t_e2e = float(np.median(t_host[1:]))
    ref_pcm, t_ref = ref_vorbis(flags, spectra, channels, bs)
    m = min(ref_pcm.shape[1], int(n.value))
    err = float(np.max(np.abs(out[:, :m] - ref_pcm[:, :m])))
    E.eav1cu_vorbis_close(h)
    samples = int(n.value) * channels
    return {"workload": "C3: %d blocks x %d channels, blocksizes %s, IMDCT + window + overlap-add" % (n_blocks, channels, list(bs)),
            "metric": "Vorbis PCM samples/s", "unit": "Msamples/s", "dtype": "f32",
            "value": samples / ms / 1e3, "ms_per_step": ms,
            "e2e": {"value": samples / t_e2e / 1e6, "unit": "Msamples/s", "h2d_bytes_per_step": int(spectra.nbytes + flags.nbytes),
                    "d2h_bytes_per_step": int(samples * 4)},
            "roofline": {"bound": "hbm", "kernel": "vorbis_kernel", "achieved": round(alg_bytes / ms / 1e6, 1), "peak": peak,
                         "unit": "GB/s", "frac": round(alg_bytes / ms / 1e6 / peak, 4), "traffic": None, "peak_source": peak_src,
                         "alg_bytes_per_launch": int(alg_bytes)},
            "cpu_baseline": {"value": samples / t_ref / 1e6, "unit": "Msamples/s", "cores": 1, "kind": "reference",
                             "sample": "the whole batch once through minivorbis mdct_backward + vorbis_synthesis_blockin (oracle/_ref/libvorbis_ref.so)"},
            "parity": {"max_abs_err": err, "tolerance": 1e-5, "ok": bool(err <= 1e-5)}}


def random_grain():
    """seeded Dav1dFilmGrainData for configs[3] (rng(4), overlap on)"""
    import numpy as np
    from tests.test_filmgrain import random_grain as rg
    g = rg(np.random.default_rng(4), 10, 1, cfl=0, overlap=1, clip=1)
    return g


def run_ours(args):
    import torch
    import torch.distributed as dist
    from easyav1_b200.lib import load_engine, load_host
    from easyav1_b200 import shard

    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the reconstruction path has no CPU fallback")
    torch.cuda.set_device(local)
    os.environ["EASYAV1_CUDA_DEVICE"] = str(local)
    if world > 1:
        dist.init_process_group("gloo")      # barrier + MAX only: nothing is exchanged on the data path
    E, host = load_engine(), load_host()
    man, obu = load_clip()
    n_frames = len(man["frames"])
    cap = man["w"] * man["h"] * 4
    only = set(args.only.split(",")) if args.only else None
    want = lambda k: only is None or k in only

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    cores = host_threads()
    if world > 1:
        # ---- configs[4]: 64 streams dealt to the ranks; S decoders per rank share cores / world host threads
        mine = shard.assign(N_STREAMS_C5, world, rank)
        S = len(mine)
        thr = max(1, cores // (world * max(S, 1)))
        clocks = ClockSampler(local)
        clocks.start()
        # device-resident: this rank's S staged streams replayed together
        barrier()
        dev = multi_stream_device(E, host, man, obu, S, max(1, args.steps), torch)
        dev_ms = max_over_ranks(dev["ms"])
        launches = int(sum_over_ranks(dev["gpu_launches"]))
        gbs = sum_over_ranks(dev["whole_frame_gbs"] * dev["ms"]) / dev_ms
        # end to end
        workers = max(2, (cores + cores // 4) // world)
        pump_e2e(host, man, obu, S, 1, workers)          # warm
        barrier()
        t0 = time.perf_counter()
        passes = max(1, min(args.steps, 3))
        last, n_e2e, _, stats = pump_e2e(host, man, obu, S, passes, workers)
        barrier()
        e2e_t = max_over_ranks(time.perf_counter() - t0)
        clk = clocks.stop()
        ok = all(m == man["frames"][-1]["md5"] for m in last)
        ok = sum_over_ranks(0 if ok else 1) == 0
        assert ok, "a stream's last frame differs from the golden MD5 (= the 1-GPU result)"
        peak, peak_src = measured_peak()
        total_frames = N_STREAMS_C5 * n_frames
        out = {"metric": METRIC, "value": total_frames * max(1, args.steps) / (dev_ms / 1e3), "unit": UNIT, "n_gpus": world,
               "steps": args.steps, "warmup": max(args.warmup, 1), "ms_per_step": dev_ms / max(1, args.steps),
               "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
               "config": {"workload": workload_c5(world), "step": "one pass of all 64 streams", "streams_per_gpu": S,
                          "host_pipeline": "host/pump.c: %d single-threaded decoders on %d worker threads per rank" % (S, workers),
                          "host_threads": cores,
                          "l2": "working set of %d streams per GPU exceeds L2; no flush" % S,
                          "parity": "last frame of every stream on every rank equals the golden MD5"},
               "roofline": {"bound": "hbm", "kernel": "whole frame, %d streams per GPU" % S, "achieved": round(gbs / world, 1),
                            "peak": peak, "unit": "GB/s", "frac": round(gbs / world / peak, 5), "traffic": None,
                            "peak_source": peak_src, "aggregate_gbs": round(gbs, 1)},
               "e2e": {"value": total_frames * passes / e2e_t, "unit": UNIT,
                       "h2d_bytes_per_step": int(sum_over_ranks(stats["h2d"]) // passes),
                       "d2h_bytes_per_step": int(sum_over_ranks(stats["d2h"]) // passes)},
               "gpu_launches": launches, "clocks": clk}
        if rank == 0:
            fps, t_each = ref_multi_stream(obu, cap, N_STREAMS_C5, 1, man)
            out["cpu_baseline"] = {"value": fps, "unit": UNIT, "cores": cores, "kind": "reference",
                                   "sample": "one pass of 64 concurrent reference decoders (oracle/_ref), %d thread(s) each" % t_each}
            print(json.dumps(out))
        dist.barrier()
        dist.destroy_process_group()
        return

    # ---- N = 1: configs[1] (C2) is the headline
    buf = C.create_string_buffer(cap)
    clocks = ClockSampler(local)
    clocks.start()
    dev = device_leg(E, host, man, obu, args.steps, args.warmup, torch, prof=bool(os.environ.get("EAV1_BENCH_PROFILE")))
    clk = clocks.stop()
    keep = C.create_string_buffer(obu, len(obu))
    for _ in range(max(args.warmup, 3)):
        e2e_pass(host, obu, cap, buf, keep)
    e2e_t, e2e_n, h2d, d2h = 0.0, 0, 0, 0
    for _ in range(args.steps):
        n, dt, md5, cnt = e2e_pass(host, obu, cap, buf, keep)
        assert md5 == man["frames"][-1]["md5"], "e2e decode differs from the golden MD5"
        e2e_t += dt; e2e_n += n; h2d += cnt[1]; d2h += cnt[2]
    out = {
        "metric": METRIC, "value": n_frames / (dev["ms_per_step"] / 1e3), "unit": UNIT,
        "n_gpus": 1, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": dev["ms_per_step"], "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": workload_c2(n_frames), "step": "one pass over the %d-frame stream" % n_frames,
                   "l2": "256 MiB flush write between timed steps", "streams_per_gpu": 1,
                   "replay": dev["replay"],
                   "coded_fraction_c": round(dev["coded_fraction"], 4),
                   "ms_per_step_with_stage_events": round(dev["ms_per_step_events"], 3)},
        "roofline": roofline_of(dev),
        "e2e": {"value": e2e_n / e2e_t, "unit": UNIT, "h2d_bytes_per_step": int(h2d // args.steps),
                "d2h_bytes_per_step": int(d2h // args.steps)},
        "gpu_launches": dev["launches"],
        "clocks": clk,
    }
    fps, passes = ref_bounded(obu, cap, cores, man)
    out["cpu_baseline"] = {"value": fps, "unit": UNIT, "cores": cores, "kind": "reference",
                           "sample": "%d passes of the same %d-frame stream through oracle/_ref (reference dav1d C build, "
                                     "n_threads=%d)" % (passes, n_frames, cores)}
    configs = {}
    # ---- configs[3] (C4): 3840x2160 10-bit, film grain applied to every shown frame in the device leg
    if want("c4"):
        try:
            man4, obu4 = load_clip(CLIP4)
            cap4 = man4["w"] * man4["h"] * 4
            g = random_grain()
            d4 = device_leg(E, host, man4, obu4, max(2, args.steps // 2), 3, torch, grain=g)
            buf4, keep4 = C.create_string_buffer(cap4), C.create_string_buffer(obu4, len(obu4))
            e2e_pass(host, obu4, cap4, buf4, keep4)
            t4, n4, h4, dd4 = 0.0, 0, 0, 0
            for _ in range(2):
                n, dt, md5, cnt = e2e_pass(host, obu4, cap4, buf4, keep4)
                assert md5 == man4["frames"][-1]["md5"], "C4 e2e decode differs from the golden MD5"
                t4 += dt; n4 += n; h4 += cnt[1]; dd4 += cnt[2]
            fps4, p4 = ref_bounded(obu4, cap4, cores, man4, budget_s=6.0, max_passes=6)
            configs["C4"] = {
                "workload": "C4: 3840x2160 10-bit 4:2:0, %d frames (%s.ivf); device leg adds eav1cu_apply_grain (seeded "
                            "Dav1dFilmGrainData, overlap on) on every frame" % (len(man4["frames"]), CLIP4),
                "metric": "AV1 decoded frames/s (4K 10-bit)", "unit": UNIT, "dtype": "u16",
                "value": len(man4["frames"]) / (d4["ms_per_step"] / 1e3), "ms_per_step": d4["ms_per_step"],
                "coded_fraction_c": round(d4["coded_fraction"], 4),
                "roofline": roofline_of(d4),
                "e2e": {"value": n4 / t4, "unit": UNIT, "h2d_bytes_per_step": int(h4 // 2), "d2h_bytes_per_step": int(dd4 // 2),
                        "note": "the stream carries no film grain, so the end-to-end leg is decode only"},
                "gpu_launches": d4["launches"],
                "cpu_baseline": {"value": fps4, "unit": UNIT, "cores": cores, "kind": "reference",
                                 "sample": "%d passes of the same stream through oracle/_ref, n_threads=%d" % (p4, cores)},
                "parity": "every frame MD5-exact while staging; last frame re-checked after the timed replays and every e2e pass"}
        except KeyError:
            configs["C4"] = {"unavailable": "fixture %s missing from tests/golden/manifest.json" % CLIP4}
    # ---- configs[2] (C3): Vorbis float slice
    if want("c3"):
        configs["C3"] = vorbis_leg(E, torch)
    # ---- configs[0] (C1): the easyAV1 C99 API on a 640x360 WebM, our library vs the reference's own
    if want("c1"):
        try:
            from tests import easyav1_api as api
            man1 = json.load(open(os.path.join(ROOT, "tests", "golden", "manifest.json")))[CLIP1]
            webm = open(os.path.join(ROOT, "tests", "golden", CLIP1 + ".webm"), "rb").read()
            keep1 = C.create_string_buffer(webm, len(webm))
            L = api.bind(os.path.join(ROOT, "easyav1_b200", "libeasyav1_b200.so"))
            R = api.bind(os.path.join(ROOT, "oracle", "_ref", "libeasyav1_ref.so"))
            want_md5 = man1["frames"][-1]["md5"]
            easyav1_api_pass(L, keep1, len(webm), want_md5)
            t1, n1 = 0.0, 0
            for _ in range(max(3, args.steps)):
                n, dt = easyav1_api_pass(L, keep1, len(webm))
                t1 += dt; n1 += n
            easyav1_api_pass(R, keep1, len(webm), want_md5)
            tr, nr = 0.0, 0
            for _ in range(3):
                n, dt = easyav1_api_pass(R, keep1, len(webm))
                tr += dt; nr += n
            configs["C1"] = {
                "workload": "C1: easyav1_benchmark loop (easyav1_init_from_memory / easyav1_decode_next / easyav1_get_video_frame) "
                            "on a 640x360 8-bit 4:2:0 WebM, %d frames (%s.webm, video only: the benchmark disables audio)"
                            % (len(man1["frames"]), CLIP1),
                "metric": "AV1 decoded frames/s (360p, easyAV1 API)", "unit": UNIT, "dtype": "u8",
                "value": None, "e2e": {"value": n1 / t1, "unit": UNIT, "frames_per_pass": n1 // max(3, args.steps)},
                "cpu_baseline": {"value": nr / tr, "unit": UNIT, "cores": cores, "kind": "reference",
                                 "sample": "3 passes through oracle/_ref/libeasyav1_ref.so (src/easyav1.c + nestegg + dav1d C, unmodified)"},
                "parity": "last frame MD5-exact through both libraries"}
        except (KeyError, OSError) as e:
            configs["C1"] = {"unavailable": str(e)}
    # ---- configs[4] (C5) on one GPU: device-resident S-stream replay, 64 streams end to end
    if want("c5"):
        peak, _ = measured_peak()
        devs = [multi_stream_device(E, host, man, obu, s, max(8, 96 // s), torch, sampler=ClockSampler(local)) for s in (1, 4, 8, 16)]
        devs.append(multi_stream_device(E, host, man, obu, 4, 24, torch, graph=False))
        clk5 = {"gpu_util_mean": max((d.get("gpu_util_mean") or 0) for d in devs)}
        workers = cores + cores // 4
        pump_e2e(host, man, obu, N_STREAMS_C5, 1, workers)          # warm
        last, n5, t5, st5 = pump_e2e(host, man, obu, N_STREAMS_C5, 1, workers)
        assert all(m == man["frames"][-1]["md5"] for m in last), "a stream's last frame differs from the golden MD5"
        fps5, t_each = ref_multi_stream(obu, cap, N_STREAMS_C5, 1, man)
        best = max(devs, key=lambda d: d["value"])
        configs["C5"] = {
            "workload": workload_c5(1), "metric": METRIC, "unit": UNIT, "dtype": "u8",
            "value": best["value"], "device_resident": devs,
            "roofline": {"bound": "hbm", "kernel": "whole frame, %d staged streams replayed together" % best["streams"],
                         "achieved": best["whole_frame_gbs"], "peak": peak, "unit": "GB/s",
                         "frac": round(best["whole_frame_gbs"] / peak, 5), "traffic": None},
            "gpu_util_mean_best_device_leg": clk5.get("gpu_util_mean"),
            "e2e": {"value": n5 / t5, "unit": UNIT, "streams": N_STREAMS_C5,
                    "host_pipeline": "host/pump.c: 64 single-threaded decoders on %d worker threads" % workers,
                    "h2d_bytes_per_step": st5["h2d"], "d2h_bytes_per_step": st5["d2h"], "gpu_launches": st5["launches"]},
            "cpu_baseline": {"value": fps5, "unit": UNIT, "cores": cores, "kind": "reference",
                             "sample": "one pass of 64 concurrent reference decoders (oracle/_ref), %d thread(s) each" % t_each},
            "parity": "last frame of every stream MD5-exact"}
    if configs:
        out["configs"] = configs
    print(json.dumps(out))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--only", default="", help="comma list of the extra config legs to run at N=1 (c4,c3,c1,c5); default all")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
