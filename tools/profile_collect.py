#!/usr/bin/env python3
"""Development aid: copy what tools/profile_round.sh <tag> brought back (gpurun_out/) into profiles/ and
derive the per-kernel summary and the kernel shares.   python tools/profile_collect.py r1k"""
import collections, csv, json, os, shutil, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag = sys.argv[1]
G, P = os.path.join(ROOT, "gpurun_out"), os.path.join(ROOT, "profiles")
pairs = [("%s_launches.csv", "%s_launches_bench_timed_step.csv"), ("%s_bench.json", "%s_bench.json"),
         ("%s_vorbis.json", "%s_vorbis_c3.json"), ("%s_full_raw.csv", "%s_full_raw_recon_resid_mc_cdef_lr_deblock.csv"),
         ("%s_vorbis_full_raw.csv", "%s_vorbis_full_raw.csv"), ("%s_grain_full_raw.csv", "%s_grain_full_raw.csv")]
for a, b in pairs:
    src = os.path.join(G, a % tag)
    if os.path.exists(src) and os.path.getsize(src):
        shutil.copy(src, os.path.join(P, b % tag))
    else:
        print("missing:", src)
raw = os.path.join(P, "%s_full_raw_recon_resid_mc_cdef_lr_deblock.csv" % tag)
if os.path.exists(raw):
    subprocess.run([sys.executable, os.path.join(ROOT, "tools", "ncu_summary.py"), raw,
                    os.path.join(P, "%s_full_summary.json" % tag)], stdout=subprocess.DEVNULL)
rows = list(csv.reader(open(os.path.join(P, "%s_launches_bench_timed_step.csv" % tag))))
for i, r in enumerate(rows):
    if "Kernel Name" in r:
        hdr, start = r, i + 1
        break
agg, cnt = collections.Counter(), collections.Counter()
for r in rows[start:]:
    if len(r) < len(hdr):
        continue
    try:
        v = float(r[hdr.index("Metric Value")].replace(",", ""))
    except ValueError:
        continue
    k = r[hdr.index("Kernel Name")].split("(")[0]
    agg[k] += v
    cnt[k] += 1
tot = sum(agg.values())
with open(os.path.join(P, "%s_kernel_share.txt" % tag), "w") as f:
    f.write("ncu --metrics gpu__time_duration.sum --clock-control none, first timed step of `python bench.py --steps 2 --warmup 3` (120 frames)\n")
    f.write("(cold-cache, serialised launches: compare the SHARES with bench.py's stage events, not the absolutes)\n")
    for k, v in agg.most_common():
        f.write("%-40s %6d launches %10.3f ms  %5.1f%%\n" % (k, cnt[k], v / 1e6, 100 * v / tot))
    d = json.load(open(os.path.join(P, "%s_bench.json" % tag)))
    st = d["roofline"]["stages"]
    t = sum(x["ms_per_step"] for k, x in st.items() if k != "resid_kernel")
    f.write("\nbench.py stage events (instrumented pass, ms per step; resid_kernel runs beside mc_kernel on a second stream):\n")
    for k, x in st.items():
        f.write("%-40s %10.3f ms  %5.1f%% of the main-stream sum\n" % (k, x["ms_per_step"], 100 * x["ms_per_step"] / t))
    f.write("\nvalue %.1f frames/s (%.2f ms per step without stage events), e2e %.1f frames/s, cpu_baseline %.1f frames/s\n"
            % (d["value"], d["ms_per_step"], d["e2e"]["value"], d["cpu_baseline"]["value"]))
print(open(os.path.join(P, "%s_kernel_share.txt" % tag)).read())
