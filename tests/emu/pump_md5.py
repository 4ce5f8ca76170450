#!/usr/bin/env python3
"""Run host/pump.c over several fixtures at once on the HOST emulation engine (own process) and print, per stream,
the frame count and the MD5 of the last picture as JSON. usage: pump_md5.py n_workers name [name ...]"""
import hashlib, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from easyav1_b200.lib import load_host, pump_run
from tests.golden.gen_fixtures import read_ivf
from tests.golden import synth_streams

man = json.load(open(os.path.join(ROOT, "tests/golden/manifest.json")))
obus = []
for name in sys.argv[2:]:
    if name.startswith("syn_"):
        obus.append(synth_streams.build(name)[0])
    else:
        obus.append(b"".join(read_ivf(os.path.join(ROOT, "tests/golden", man[name]["file"]))))
host = load_host(os.path.join(ROOT, "tests/emu/libdav1d_b200_emu.so"))
frames, last = pump_run(host, obus, int(sys.argv[1]), 1 << 22)
print(json.dumps({"frames": frames, "md5": [hashlib.md5(x).hexdigest() for x in last]}))
