"""Which coding tools the parity streams actually execute. The engine histograms every command it is
handed (EAV1_COVERAGE=1, struct Eav1Coverage in include/easyav1_cuda.h); tools/coverage_run.py decodes a
list of streams in one process, checks every frame against its golden MD5 and prints the histogram. The
assertions below are the floor the fixtures must keep reaching, so that a changed fixture cannot silently
drop a tool: every entry of the reference's itxfm_add[19][17] table that exists (itx_tmpl.c:225-290), every
2-D filter pair for put / prep, unscaled and scaled, all 14 intra modes, CfL, palette, intra block copy,
warp, the four compound and three blend operators, every deblock filter size, every restoration kind,
CDEF, super-resolution and film grain - for 8-bit and for high bit depth. tests/golden/coverage.json is the
histogram of the GPU run the fixtures were last changed with."""
import json
import os
import subprocess
import sys

import pytest

from tests.golden import synth_streams

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CLIPS = ["intra_64x64_nofilt", "cif_intra", "cif_inter", "cif_tools_s4", "cif_tools_s2", "c1_360p", "c2_1080p",
         "cif_10bit", "c4_720p_10bit", "cif_grain_t1", "cif_grain_t15", "cif_grain_est", "cif_444", "cif_422",
         "cif_400", "qcif_lossless", "cif_screen"]

T16 = list(range(16))
VALID_ITX = {}
for _tx in (0, 5, 13, 6, 1, 7, 14, 8):          # 4x4 4x8 4x16 8x4 8x8 8x16 16x4 16x8: all 16 types
    VALID_ITX[_tx] = T16
VALID_ITX[0] = T16 + [16]                       # + WHT_WHT (lossless)
VALID_ITX[2] = list(range(12))                  # 16x16: no identity x adst
for _tx in (15, 9, 16, 10, 3):                  # 8x32 16x32 32x8 32x16 32x32: DCT_DCT, IDTX
    VALID_ITX[_tx] = [0, 9]
for _tx in (17, 11, 18, 12, 4):                 # 16x64 32x64 64x16 64x32 64x64: DCT_DCT
    VALID_ITX[_tx] = [0]


def run(engine, names):
    env = dict(os.environ, EAV1_COVERAGE="1", EASYAV1_THREADS="3")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "coverage_run.py"), engine] + names, env=env,
                       stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=1500)
    assert r.returncode == 0, r.stderr[-3000:]
    d = json.loads(r.stdout.strip().splitlines()[-1])
    assert d["_on"] == 1
    return d


def missing(d, full):
    """list of required histogram entries that are zero"""
    out = []
    itx = d["itx"]
    for tx, types in VALID_ITX.items():
        for tp in types:
            if not itx[0][tx][tp]:
                out.append(("itx 8-bit", tx, tp))
            if full and not (itx[1][tx][tp] or itx[2][tx][tp]):
                out.append(("itx hbd", tx, tp))
    for bd in range(3 if full else 1):
        for cls in range(3):
            if not d["itx_eob"][bd][cls]:
                out.append(("itx eob class", bd, cls))
        for mode in range(14):
            if not d["ipred"][bd][mode]:
                out.append(("ipred", bd, mode))
    pairs = [(a, b) for a in range(3) for b in range(3)] + [(3, 3)]
    for hbd in range(2 if full else 1):
        for scaled in range(2):
            for prep in range(2):
                for fh, fv in pairs:
                    n = sum(d["mc"][hbd][scaled][prep][fh][fv])
                    if scaled and hbd:       # rare combination: put or prep is enough per filter pair
                        n += sum(d["mc"][hbd][scaled][1 - prep][fh][fv])
                    if not n and (full or not scaled):
                        out.append(("mc", hbd, scaled, prep, fh, fv))
                if not sum(sum(sum(x) for x in fh) for fh in d["mc"][hbd][scaled][prep]):
                    out.append(("mc any filter", hbd, scaled, prep))
        for prep in range(2):
            for sub in range(4):                # integer / horizontal / vertical / 2-D sub-pel positions
                if not sum(d["mc"][hbd][0][prep][f][f][sub] for f in range(4)):
                    out.append(("mc subpel", hbd, prep, sub))
        for name, n in (("warp", 2), ("comp", 4), ("blend", 3)):
            for i in range(n):
                if not d[name][hbd][i]:
                    out.append((name, hbd, i))
        for mode in (0, 3, 4, 5):               # cfl_pred: DC, LEFT_DC, TOP_DC, DC_128
            if not d["cfl"][hbd][mode]:
                out.append(("cfl", hbd, mode))
        for chroma in range(2):
            for direction in range(2):
                for size in range(2 if chroma else 3):
                    if not d["lf"][hbd][chroma][direction][size]:
                        out.append(("lf", hbd, chroma, direction, size))
        for kind in range(5):
            if not d["lr"][hbd][kind]:
                out.append(("lr", hbd, kind))
        for name in ("cdef_frames", "superres_frames", "pal"):
            if not d[name][hbd]:
                out.append((name, hbd))
        if not (d["grain"][hbd][0] and d["grain"][hbd][1]):
            out.append(("grain", hbd))
    for i in range(3):
        if not d["w_mask_ss"][i]:
            out.append(("w_mask subsampling", i))
    if not d["intrabc"]:
        out.append(("intrabc",))
    for i in range(4):
        if not d["layout"][i]:
            out.append(("layout", i))
    # block sizes: every put/prep and intra size class from 4 to 64 (and 128 for motion compensation)
    for l in range(2, 8 if full else 7):
        if not (sum(d["mc_wh"][0][l]) and sum(row[l] for row in d["mc_wh"][0])):
            out.append(("mc size", l))
    for l in range(2, 7):
        if not (sum(d["ipred_wh"][l]) and sum(row[l] for row in d["ipred_wh"])):
            out.append(("ipred size", l))
    return out


def test_coverage_on_host_emulation():
    """the quick streams through the host emulation: the histogram works and already spans the 8-bit tools"""
    if not os.path.exists(os.path.join(ROOT, "tests", "emu", "libdav1d_b200_emu.so")):
        pytest.skip("tests/emu not built (needs the reference sources)")
    names = ["cif_tools_s2", "cif_tools_s4", "cif_screen", "cif_grain_t15", "cif_grain_est", "cif_422", "cif_400",
             "qcif_lossless", "syn_420_8", "syn_scaled_8", "syn_gmv_8", "syn_sr_8", "syn_intrabc", "syn_444_8",
             "syn_lossless", "syn_grain_8_444"]
    d = run("emu", names)
    assert d["_frames_checked"] == sum(d["frames"]) or d["_frames_checked"] > 0
    assert missing(d, full=False) == []


@pytest.mark.gpu
def test_coverage_cuda():
    """every fixture and every synthetic stream through the CUDA engine: bit-exact (checked by the runner)
    and together they reach every required tool at 8-bit and at high bit depth"""
    names = CLIPS + sorted(synth_streams.STREAMS)
    d = run("cuda", names)
    out = os.path.join(ROOT, "gpurun_out")
    if os.path.isdir(out):
        json.dump(d, open(os.path.join(out, "coverage_cuda.json"), "w"))
    assert missing(d, full=True) == []
