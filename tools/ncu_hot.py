#!/usr/bin/env python3
"""Development aid: the hottest SASS instructions (warp stall samples) per kernel from `ncu --page source --csv`.
    ncu -i X.ncu-rep --page source --csv | python tools/ncu_hot.py [N] > hot.txt"""
import collections, csv, sys
top = int(sys.argv[1]) if len(sys.argv) > 1 else 30
csv.field_size_limit(1 << 30)
agg, inst, src, tot = {}, {}, {}, collections.Counter()
kern, hdr = None, None
for r in csv.reader(sys.stdin):
    if not r:
        continue
    if r[0] == "Kernel Name":
        kern = r[1].split("(")[0]
        agg.setdefault(kern, collections.Counter()); inst.setdefault(kern, collections.Counter()); src.setdefault(kern, {})
        continue
    if r[0] == "Address":
        hdr = r
        si, ws, ie = hdr.index("Source"), hdr.index("Warp Stall Sampling (All Samples)"), hdr.index("Instructions Executed")
        continue
    if hdr is None or kern is None or len(r) != len(hdr):
        continue
    try:
        v, e = int(r[ws]), int(r[ie])
    except ValueError:
        continue
    agg[kern][r[0]] += v; inst[kern][r[0]] += e; src[kern][r[0]] = r[si]; tot[kern] += v
for k in agg:
    print("== %s: %d stall samples" % (k, tot[k]))
    for a, v in agg[k].most_common(top):
        print("%5.1f%%  executed %9d  %s" % (100.0 * v / max(tot[k], 1), inst[k][a], src[k][a][:120]))
