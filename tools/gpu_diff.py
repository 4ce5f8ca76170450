#!/usr/bin/env python3
"""Debug helper: decode a fixture through the recorder + HOST EMULATION of the kernels and
diff against the golden MD5s / the reference frames. Not a test of the product."""
import hashlib, json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import ctypes as C
from easyav1_b200.lib import load_host
from easyav1_b200.decoder import Decoder
from tests.golden.gen_fixtures import read_ivf
from tests import refdec

def main():
    name = sys.argv[1]
    filt = int(sys.argv[2]) if len(sys.argv) > 2 else 7
    man = json.load(open(os.path.join(ROOT, "tests/golden/manifest.json")))[name]
    obu = b"".join(read_ivf(os.path.join(ROOT, "tests/golden", man["file"])))
    host = load_host() if os.environ.get("EAV1_GPU") else load_host(os.path.join(ROOT, "tests/emu/libdav1d_b200_emu.so"))
    cap = man["w"] * man["h"] * 4
    ref = list(refdec.decode_all_big(obu, cap, inloop_filters=filt))
    d = Decoder(obu, host=host, max_frame_bytes=cap, inloop_filters=filt)
    ok = True
    for i, (info, planes) in enumerate(d):
        rinfo, rplanes = ref[i]
        same = planes == rplanes
        ok &= same
        if not same:
            a = np.frombuffer(planes, np.uint8); b = np.frombuffer(rplanes, np.uint8)
            bad = np.nonzero(a != b)[0]
            ysz = info.w * info.h
            first = int(bad[0])
            if first < ysz: where = ("Y", first % info.w, first // info.w)
            else:
                cw = (info.w + 1) // 2; csz = cw * ((info.h + 1) // 2)
                o = first - ysz; pl = "U" if o < csz else "V"; o %= csz
                where = (pl, o % cw, o // cw)
            print("frame", i, "type", info.frame_type, "MISMATCH", len(bad), "bytes; first at", where,
                  "got", a[first], "want", b[first])
        else:
            print("frame", i, "type", info.frame_type, "ok", hashlib.md5(planes).hexdigest())
    print("ALL OK" if ok else "FAILED")

main()
