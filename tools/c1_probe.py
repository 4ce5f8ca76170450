#!/usr/bin/env python3
"""Development aid: where does a pass through the easyAV1 API go (init / decode loop / destroy), ours vs the reference,
at several EASYAV1_THREADS.   python tools/c1_probe.py [clip]"""
import ctypes as C, json, os, subprocess, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
CHILD = """
import ctypes as C, sys, time, json
sys.path.insert(0, %r)
from tests import easyav1_api as api
L = api.bind(sys.argv[1]); webm = open(sys.argv[2], 'rb').read(); keep = C.create_string_buffer(webm, len(webm))
res = []
for rep in range(4):
    st = L.easyav1_default_settings(); st.enable_audio = 0; st.log_level = 1
    t0 = time.perf_counter(); h = L.easyav1_init_from_memory(keep, len(webm), C.byref(st)); t1 = time.perf_counter()
    n = 0
    while L.easyav1_decode_next(h) == 1:
        if L.easyav1_has_video_frame(h) == 1:
            L.easyav1_get_video_frame(h); n += 1
    t2 = time.perf_counter(); hp = C.c_void_p(h); L.easyav1_destroy(C.byref(hp)); t3 = time.perf_counter()
    res.append((n, round((t1 - t0) * 1e3, 2), round((t2 - t1) * 1e3, 2), round((t3 - t2) * 1e3, 2), round(n / (t3 - t0), 1), round(n / (t2 - t1), 1)))
print(json.dumps(res))
""" % ROOT
clip = sys.argv[1] if len(sys.argv) > 1 else "c1_360p_bench"
path = os.path.join(ROOT, "tests", "golden", clip + ".webm")
for name, lib, envs in (("reference", "oracle/_ref/libeasyav1_ref.so", [{}]),
                        ("ours", "easyav1_b200/libeasyav1_b200.so",
                         [{}] + [{"EASYAV1_THREADS": t} for t in os.environ.get("C1_THREADS", "1,4,8").split(",")] +
                         [{"EASYAV1_THREADS": t, "EASYAV1_FRAME_DELAY": d} for t, d in
                          [x.split(":") for x in os.environ.get("C1_COMBOS", "").split(",") if x]] +
                         ([{"EAV1_TIMING": "1"}] if os.environ.get("C1_TIMING") else []))):
    for env in envs:
        e = dict(os.environ); e.update(env)
        out = subprocess.run([sys.executable, "-c", CHILD, os.path.join(ROOT, lib), path], capture_output=True, text=True, env=e, timeout=300)
        print(name, env, "(frames, init ms, loop ms, destroy ms, fps whole, fps loop):", out.stdout.strip().splitlines()[-1] if out.stdout.strip() else out.stderr[-500:])
        if "EAV1_TIMING" in env:
            print(out.stderr[-3000:])
