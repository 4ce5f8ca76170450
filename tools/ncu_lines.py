#!/usr/bin/env python3
"""Development aid: aggregate an ncu report's source page by source line (stall samples + instructions).
   ncu -i X.ncu-rep --page source --csv --print-source cuda,sass > src.csv ; tools/ncu_lines.py src.csv [N]"""
import csv, collections, sys
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
cur = None; hdr = None
agg = collections.Counter(); inst = collections.Counter()
for r in rows:
    if not r: continue
    if r[0] == "File Path": cur = r[1].split('/')[-1]; continue
    if r[0] == "Function Name": continue
    if r[0] == "Line No": hdr = r; si = hdr.index("# Samples"); ii = hdr.index("Instructions Executed"); continue
    if hdr is None or cur is None or r[0] == "": continue
    try: line = int(r[0]); s = int(r[si]); i = int(r[ii])
    except Exception: continue
    k = (cur, line, r[1].strip()[:100])
    agg[k] += s; inst[k] += i
tot = sum(agg.values()) or 1; ti = sum(inst.values()) or 1
print("total samples", tot, "warp instructions", ti)
byfile = collections.Counter(); bi = collections.Counter()
for k, s in agg.items(): byfile[k[0]] += s
for k, s in inst.items(): bi[k[0]] += s
print("samples by file:", byfile.most_common()); print("instructions by file:", bi.most_common())
for k, s in agg.most_common(top):
    print(f"{s:6d} {s/tot:5.1%} i={inst[k]/ti:5.1%} {k[0]}:{k[1]}  {k[2]}")
