#!/usr/bin/env python3
"""Development aid: end-to-end frames/s of one stream through libdav1d_b200.so against EASYAV1_THREADS / EASYAV1_FRAME_DELAY.
    python tools/thread_sweep.py c2_1080p_bench 4,8,12,16,24,32"""
import os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CHILD = """
import sys, os, ctypes as C, json
sys.path.insert(0, %r)
import bench
from easyav1_b200.lib import load_host
man, obu = bench.load_clip(sys.argv[1])
host = load_host(); cap = man['w'] * man['h'] * 4 * (2 if man.get('bpc', 8) > 8 else 1); buf = C.create_string_buffer(cap)
best = 0
for rep in range(int(sys.argv[2])):
    n, dt, md5, cnt = bench.e2e_pass(host, obu, cap, buf)
    if rep: best = max(best, n / dt)
print(json.dumps(round(best, 1)))
""" % ROOT
clip = sys.argv[1]
reps = "3" if "2160" in clip else "5"
for thr in sys.argv[2].split(","):
    for fd in ([""] + (sys.argv[3].split(",") if len(sys.argv) > 3 else [])):
        e = dict(os.environ, EASYAV1_THREADS=thr)
        if fd: e["EASYAV1_FRAME_DELAY"] = fd
        out = subprocess.run([sys.executable, "-c", CHILD, clip, reps], capture_output=True, text=True, env=e, timeout=600)
        print(clip, "threads", thr, "frame_delay", fd or "auto", "best fps", out.stdout.strip().splitlines()[-1] if out.stdout.strip() else out.stderr[-300:], flush=True)
