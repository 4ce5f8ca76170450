#!/usr/bin/env python3
"""Development aid: device-resident replay of S staged streams as CUDA graphs. usage: multi_stream_device_probe.py S [passes]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench, torch
from easyav1_b200.lib import load_engine, load_host
E, host = load_engine(), load_host()
man, obu = bench.load_clip()
for S in [int(x) for x in sys.argv[1].split(",")]:
    r = bench.multi_stream_device(E, host, man, obu, S, int(sys.argv[2]) if len(sys.argv) > 2 else max(4, 48 // S), torch, sampler=bench.ClockSampler(0))
    print("S=%d: %.0f frames/s, %.0f GB/s, util %s" % (S, r["value"], r["whole_frame_gbs"], r.get("gpu_util_mean")))
