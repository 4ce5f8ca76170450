#!/usr/bin/env python3
"""Decode a fixture through the recorder + HOST emulation engine in a process of its own (so
the emulation library never shares a symbol namespace with the CUDA engine) and print the
per-frame MD5s as JSON.  usage: decode_md5.py <fixture> [truncate_fraction]"""
import hashlib, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from easyav1_b200.lib import load_host, EngineError
from easyav1_b200.decoder import Decoder
from tests.golden.gen_fixtures import read_ivf

if sys.argv[1].startswith("syn_"):     # synthetic stream, rebuilt from its seed (tests/golden/synth_streams.py)
    from tests.golden import synth_streams
    obu, _ = synth_streams.build(sys.argv[1])
    man = {"w": 3840, "h": 2160}
else:
    man = json.load(open(os.path.join(ROOT, "tests/golden/manifest.json")))[sys.argv[1]]
    obu = b"".join(read_ivf(os.path.join(ROOT, "tests/golden", man["file"])))
if len(sys.argv) > 2:
    obu = obu[: int(len(obu) * float(sys.argv[2]))]
host = load_host(os.path.join(ROOT, "tests/emu/libdav1d_b200_emu.so"))
out, err = [], None
for rep in range(int(os.environ.get("EMU_REPEAT", "1"))):
    d = Decoder(obu, host=host, max_frame_bytes=man["w"] * man["h"] * 4)
    cur = []
    try:
        for _, p in d:
            cur.append(hashlib.md5(p).hexdigest())
    except EngineError as e:
        err = str(e)
    d.close()
    out.append(cur)
print(json.dumps({"passes": out, "error": err}))
