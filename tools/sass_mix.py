#!/usr/bin/env python3
"""Development aid: instruction mix of every kernel of libeasyav1_cuda.so from `cuobjdump -sass` — load / store widths,
shared-memory traffic, integer dot products, packed 16x2 integer ops, TMA / async-copy mnemonics, barriers.
    python tools/sass_mix.py > profiles/<tag>_sass_mix.txt"""
import collections, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "easyav1_b200", "libeasyav1_cuda.so")
out = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
GROUPS = [("LDG.8", r"^LDG\.E\.(U8|S8)"), ("LDG.16", r"^LDG\.E\.(U16|S16)"), ("LDG.32", r"^LDG\.E(\.(CONSTANT|STRONG\.\w+|LTC\w+))*$"),
          ("LDG.64", r"^LDG\.E\.64"), ("LDG.128", r"^LDG\.E\.128"), ("STG.8", r"^STG\.E\.U8"), ("STG.16", r"^STG\.E\.U16"),
          ("STG.32", r"^STG\.E(\.STRONG\.\w+)?$"), ("STG.64", r"^STG\.E\.64"), ("STG.128", r"^STG\.E\.128"),
          ("LDS", r"^LDS"), ("STS", r"^STS"), ("ATOMS", r"^ATOMS"), ("ATOM/RED", r"^(ATOMG|RED)"), ("IDP (dp4a/dp2a)", r"^IDP"),
          ("IMAD", r"^IMAD"), ("16x2 int", r"\.(S|U)?16x2"), ("SHFL", r"^SHFL"), ("BAR", r"^BAR"), ("WARPSYNC", r"^WARPSYNC"),
          ("LDGSTS", r"^LDGSTS"), ("UTMA*", r"^UTMA"), ("UBLKCP", r"^UBLKCP"), ("NANOSLEEP", r"^NANOSLEEP")]
kern, mix, total = None, collections.OrderedDict(), collections.Counter()
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        kern = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip().split("(")[0]
        mix[kern] = collections.Counter()
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4,6}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]+)", line)
    if m and kern:
        op = m.group(1)
        total[kern] += 1
        for g, pat in GROUPS:
            if re.search(pat, op):
                mix[kern][g] += 1
print("instruction mix per kernel of %s (cuobjdump -sass, sm_100a)\n" % os.path.basename(so))
names = [g for g, _ in GROUPS]
for k, c in mix.items():
    print("%s: %d instructions" % (k, total[k]))
    print("    " + ", ".join("%s %d" % (g, c[g]) for g in names if c[g]))
