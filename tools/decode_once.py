#!/usr/bin/env python3
"""Development aid: decode one fixture once through libdav1d_b200.so (a target for ncu / compute-sanitizer).
    python tools/decode_once.py c4_2160p_10bit_bench"""
import ctypes as C, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
from easyav1_b200.lib import load_host
man, obu = bench.load_clip(sys.argv[1] if len(sys.argv) > 1 else None)
host = load_host()
cap = man["w"] * man["h"] * 4 * (2 if "10bit" in man["file"] or "12" in man["file"] else 1)
buf = C.create_string_buffer(cap)
n, dt, md5, cnt = bench.e2e_pass(host, obu, cap, buf)
print(n, "frames", round(n / dt, 1), "frames/s", md5, md5 == man["frames"][-1]["md5"])
