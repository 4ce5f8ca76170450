#!/usr/bin/env python3
"""Development aid: device-resident step time and per-stage GPU time of the bench clip (C2, or EAV1_CLIP).
usage: stage_probe.py [steps]   -> one line: ms per step, then ms per step of every stage"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench, torch
from easyav1_b200.lib import load_engine, load_host
E, host = load_engine(), load_host()
man, obu = bench.load_clip(os.environ.get("EAV1_CLIP"))
steps = int(sys.argv[1]) if len(sys.argv) > 1 else 10
r = bench.device_leg(E, host, man, obu, steps, 3, torch)
names = ["recon", "deblock", "cdef", "lr", "mc", "grain", "resid"]
print("step %.3f ms (%s; %.3f with stage events) | " % (r["ms_per_step"], r["replay"], r["ms_per_step_events"]) +
      " ".join("%s %.3f" % (n, v) for n, v in zip(names, r["stage_ms"])))
