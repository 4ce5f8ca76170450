#!/usr/bin/env python3
"""Generate the AV1 bitstream fixtures + golden per-frame MD5s committed under tests/golden/.

Run in the BUILD container only (needs Pillow's AVIF plugin = libaom encoder, and
oracle/_ref/libdav1d_ref.so = the reference's own dav1d C build, see oracle/Makefile):

    python tests/golden/gen_fixtures.py

Method: no media is vendored in the reference and no
stand-alone encoder exists offline, so clips are synthesised with numpy, encoded
through Pillow's animated-AVIF writer (aom), the av01 samples are demuxed from the
ISOBMFF container and written as IVF.  The golden answer of every clip is the
per-frame MD5 of the visible Y,U,V planes as decoded by the reference C decoder
(method of ext/dav1d/tools/output/md5.c:194-250, reset per frame).
"""
import hashlib
import io
import json
import os
import struct
import sys

import numpy as np
from PIL import Image

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))


def synth_frames(w, h, n, seed, noise=3.0, pan=(3, 2)):
    """Low-pass noise field panning + two moving shapes + per-frame noise (C1/C2 recipe)."""
    rng = np.random.default_rng(seed)
    big_w, big_h = w + abs(pan[0]) * n + 16, h + abs(pan[1]) * n + 16
    lo = rng.integers(0, 256, size=((big_h + 7) // 8 + 2, (big_w + 7) // 8 + 2, 3), dtype=np.uint8)
    field = np.asarray(Image.fromarray(lo).resize((big_w, big_h), Image.BICUBIC), dtype=np.float32)
    frames = []
    for i in range(n):
        ox, oy = pan[0] * i, pan[1] * i
        f = field[oy:oy + h, ox:ox + w].copy()
        # moving shapes
        cx = int((0.2 + 0.6 * i / max(n - 1, 1)) * w)
        cy = int(0.3 * h)
        r = max(h // 10, 4)
        f[max(cy - r, 0):cy + r, max(cx - r, 0):cx + r] = (220, 40, 40)
        cy2 = int((0.8 - 0.5 * i / max(n - 1, 1)) * h)
        cx2 = int(0.7 * w)
        yy, xx = np.ogrid[:h, :w]
        f[(yy - cy2) ** 2 + (xx - cx2) ** 2 < r * r] = (30, 200, 90)
        f += rng.normal(0.0, noise, size=f.shape).astype(np.float32)
        frames.append(Image.fromarray(np.clip(f, 0, 255).astype(np.uint8)))
    return frames


def synth_zoom_frames(w, h, n, seed, noise=0.0):
    """10-bit frames (uint16 RGB): a low-pass noise field that drifts, rotates and zooms slowly (so that a
    translational model alone does not predict it), one moving square, optional per-frame noise. Config C4's
    content recipe."""
    import math
    rng = np.random.default_rng(seed)
    bw, bh = w + 256, h + 256
    lo = rng.integers(0, 256, size=((bh + 15) // 16 + 2, (bw + 15) // 16 + 2, 3), dtype=np.uint8)
    field = Image.fromarray(lo).resize((bw, bh), Image.BICUBIC)
    frames = []
    for i in range(n):
        ang = math.radians(0.15 * i)
        z = 1.0 + 0.003 * i
        c, s = math.cos(ang) / z, math.sin(ang) / z
        cx, cy = bw / 2 + 2.0 * i, bh / 2 + 1.0 * i
        a, b, cc = c, -s, cx - c * w / 2 + s * h / 2
        d, e, f = s, c, cy - s * w / 2 - c * h / 2
        img = field.transform((w, h), Image.AFFINE, (a, b, cc, d, e, f), resample=Image.BILINEAR)
        arr = np.asarray(img, dtype=np.float32).copy()
        r = h // 10
        x0 = int((0.2 + 0.6 * i / max(n - 1, 1)) * w)
        y0 = int(0.3 * h)
        arr[y0 - r:y0 + r, x0 - r:x0 + r] = (220, 40, 40)
        arr += rng.normal(0.0, noise, size=arr.shape).astype(np.float32)
        frames.append(np.clip(arr * 4.0, 0, 1023).astype(np.uint16))
    return frames


def synth_zoom_frames_float(w, h, n, seed, grid=32):
    """The synth_zoom_frames recipe carried out in float until the final 10-bit rounding (field interpolation and the
    per-frame drift / rotation / zoom warp are bicubic on float32): the 8-bit staircase of the integer recipe, which
    costs an encoder 250 KB per 4K inter frame, is gone and the same quality setting lands at a streaming-like
    ~100 KB per frame (24 Mbit/s at 30 frames/s)."""
    import math
    import cv2
    rng = np.random.default_rng(seed)
    bw, bh = w + 256, h + 256
    lo = rng.uniform(0, 255, size=((bh + grid - 1) // grid + 2, (bw + grid - 1) // grid + 2, 3)).astype(np.float32)
    field = cv2.resize(lo, (bw, bh), interpolation=cv2.INTER_CUBIC)
    frames = []
    for i in range(n):
        ang = math.radians(0.15 * i)
        z = 1.0 + 0.003 * i
        c, s = math.cos(ang) / z, math.sin(ang) / z
        cx, cy = bw / 2 + 2.0 * i, bh / 2 + 1.0 * i
        m = np.array([[c, -s, cx - c * w / 2 + s * h / 2], [s, c, cy - s * w / 2 - c * h / 2]], dtype=np.float64)
        arr = cv2.warpAffine(field, m, (w, h), flags=cv2.INTER_CUBIC | cv2.WARP_INVERSE_MAP)
        r = h // 10
        x0 = int((0.2 + 0.6 * i / max(n - 1, 1)) * w)
        y0 = int(0.3 * h)
        arr[y0 - r:y0 + r, x0 - r:x0 + r] = (220, 40, 40)
        frames.append(np.clip(np.rint(arr * 4.0), 0, 1023).astype(np.uint16))
    return frames


def encode_cv2_10bit(frames16, quality, speed):
    """uint16 10-bit RGB frames through OpenCV's animated-AVIF writer (libavif/aom)"""
    import cv2
    anim = cv2.Animation()
    anim.frames = [cv2.cvtColor((a << 6).astype(np.uint16), cv2.COLOR_RGB2BGR) for a in frames16]
    anim.durations = [33] * len(frames16)
    path = "/tmp/_eav1_fixture.avif"
    assert cv2.imwriteanimation(path, anim, [cv2.IMWRITE_AVIF_DEPTH, 10, cv2.IMWRITE_AVIF_QUALITY, quality,
                                             cv2.IMWRITE_AVIF_SPEED, speed])
    buf = open(path, "rb").read()
    os.remove(path)
    return buf


def boxes(buf, start, end):
    pos = start
    while pos + 8 <= end:
        size, typ = struct.unpack(">I4s", buf[pos:pos + 8])
        hdr = 8
        if size == 1:
            size = struct.unpack(">Q", buf[pos + 8:pos + 16])[0]
            hdr = 16
        elif size == 0:
            size = end - pos
        yield typ.decode("latin1"), pos + hdr, pos + size
        pos += size


def find(buf, start, end, path):
    for typ, s, e in boxes(buf, start, end):
        if typ == path[0]:
            if len(path) == 1:
                return s, e
            r = find(buf, s, e, path[1:])
            if r:
                return r
    return None


def avif_samples(buf):
    """Return the list of av01 samples (bytes) of the first video track."""
    stbl = find(buf, 0, len(buf), ["moov", "trak", "mdia", "minf", "stbl"])
    assert stbl, "no moov/trak (not an image sequence?)"
    s, e = stbl
    sizes, chunk_offs, stsc = [], [], []
    for typ, bs, be in boxes(buf, s, e):
        if typ == "stsz":
            _, fixed, cnt = struct.unpack(">III", buf[bs:bs + 12])
            sizes = [fixed] * cnt if fixed else list(struct.unpack(">%dI" % cnt, buf[bs + 12:bs + 12 + 4 * cnt]))
        elif typ == "stco":
            _, cnt = struct.unpack(">II", buf[bs:bs + 8])
            chunk_offs = list(struct.unpack(">%dI" % cnt, buf[bs + 8:bs + 8 + 4 * cnt]))
        elif typ == "co64":
            _, cnt = struct.unpack(">II", buf[bs:bs + 8])
            chunk_offs = list(struct.unpack(">%dQ" % cnt, buf[bs + 8:bs + 8 + 8 * cnt]))
        elif typ == "stsc":
            _, cnt = struct.unpack(">II", buf[bs:bs + 8])
            stsc = [struct.unpack(">III", buf[bs + 8 + 12 * i:bs + 20 + 12 * i]) for i in range(cnt)]
    samples, si = [], 0
    for ci, off in enumerate(chunk_offs):
        per = 1
        for first, n, _ in stsc:
            if ci + 1 >= first:
                per = n
        pos = off
        for _ in range(per):
            if si >= len(sizes):
                break
            samples.append(bytes(buf[pos:pos + sizes[si]]))
            pos += sizes[si]
            si += 1
    assert si == len(sizes), (si, len(sizes))
    return samples


def write_ivf(path, samples, w, h, fps=30):
    with open(path, "wb") as f:
        f.write(b"DKIF" + struct.pack("<HH4sHHIIII", 0, 32, b"AV01", w, h, fps, 1, len(samples), 0))
        for i, s in enumerate(samples):
            f.write(struct.pack("<IQ", len(s), i))
            f.write(s)


def read_ivf(path):
    buf = open(path, "rb").read()
    assert buf[:4] == b"DKIF"
    hdr = struct.unpack("<H", buf[6:8])[0]
    pos, out = hdr, []
    while pos + 12 <= len(buf):
        sz = struct.unpack("<I", buf[pos:pos + 4])[0]
        out.append(buf[pos + 12:pos + 12 + sz])
        pos += 12 + sz
    return out


def encode(frames, advanced, quality=55, speed=6):
    bio = io.BytesIO()
    advanced = dict(advanced)
    advanced.pop("container", None)
    subsampling = advanced.pop("subsampling", "4:2:0")
    frames[0].save(bio, format="AVIF", save_all=True, append_images=frames[1:], codec="aom",
                   quality=quality, speed=speed, subsampling=subsampling, duration=33,
                   advanced=advanced)
    return bio.getvalue()


def encode_cv2(frames, depth, quality, speed):
    """High-bit-depth clips: OpenCV's animated-AVIF writer (libavif/aom) takes 16-bit input."""
    import cv2
    arrs = []
    for i, f in enumerate(frames):
        a = np.asarray(f).astype(np.uint16)
        # use the extra precision: a low-amplitude ramp below the 8-bit LSB
        a = (a << (depth - 8)) + ((np.arange(a.shape[1])[None, :, None] + i) % (1 << (depth - 8))).astype(np.uint16)
        arrs.append(cv2.cvtColor((a << (16 - depth)).astype(np.uint16), cv2.COLOR_RGB2BGR))
    anim = cv2.Animation()
    anim.frames = arrs
    anim.durations = [33] * len(arrs)
    path = "/tmp/_eav1_fixture.avif"
    ok = cv2.imwriteanimation(path, anim, [cv2.IMWRITE_AVIF_DEPTH, depth, cv2.IMWRITE_AVIF_QUALITY, quality,
                                          cv2.IMWRITE_AVIF_SPEED, speed])
    assert ok
    buf = open(path, "rb").read()
    os.remove(path)
    return buf


def golden_md5(obu):
    from tests.refdec import decode_all_big as _dec
    decode_all = lambda o: _dec(o, 3840 * 2160 * 4)
    out = []
    for info, planes in decode_all(obu):
        out.append({"md5": hashlib.md5(planes).hexdigest(), "w": info.w, "h": info.h,
                    "bpc": info.bpc, "layout": info.layout, "frame_type": info.frame_type})
    return out


CLIPS = {
    # name: (w, h, n_frames, seed, advanced-options, quality, speed)
    # intra-only, no post filters: recon parity in isolation
    "intra_64x64_nofilt": (64, 64, 2, 11, {"enable-cdef": "0", "enable-restoration": "0", "enable-palette": "0",
                                            "enable-intrabc": "0"}, 60, 6),
    "cif_intra": (352, 288, 2, 12, {"enable-cdef": "1", "enable-restoration": "1"}, 55, 6),
    "cif_inter": (352, 288, 8, 13, {"enable-cdef": "1", "enable-restoration": "1"}, 55, 6),
    "c1_360p": (640, 360, 12, 1, {"enable-cdef": "1", "enable-restoration": "1"}, 55, 6),
    "c2_1080p": (1920, 1080, 6, 7, {"enable-cdef": "1", "enable-restoration": "1"}, 60, 4, 6.0),
    # slower encoder presets switch on the tools speed 6 never picks: loop restoration
    # (Wiener + SGR), OBMC (blend_v/blend_h), wedge / inter-intra / diff-weighted masks
    "cif_tools_s4": (352, 288, 7, 21, {"enable-cdef": "1", "enable-restoration": "1"}, 60, 4, 8.0),
    "cif_tools_s2": (352, 288, 7, 22, {"enable-cdef": "1", "enable-restoration": "1"}, 50, 2, 10.0),
    # 10-bit 4:2:0 (the BITDEPTH=16 template of every kernel); encoded through OpenCV (depth key)
    "cif_10bit": (352, 288, 6, 31, {"depth": 10}, 60, 4, 8.0),
    "c4_720p_10bit": (1280, 720, 4, 32, {"depth": 10}, 60, 6, 6.0),
    # config C4's frame size (BASELINE.json configs[3]): 3840x2160 10-bit 4:2:0, a few frames, low noise
    # so the file stays small (the OpenCV writer exposes neither film grain nor tool switches)
    "c4_2160p_10bit": (3840, 2160, 3, 33, {"depth": 10}, 30, 8, 0.7),
    # film grain (aom's built-in parameter sets 1 and 15 = many scaling points without block overlap /
    # chroma scaled from luma; and a grain table the encoder estimated itself from the noisy source)
    "cif_grain_t1": (352, 288, 4, 41, {"film-grain-test": "1"}, 60, 6, 6.0),
    "cif_grain_t15": (352, 288, 4, 42, {"film-grain-test": "15"}, 60, 6, 6.0),
    "cif_grain_est": (352, 288, 5, 43, {"denoise-noise-level": "20"}, 60, 6, 8.0),
    # other chroma layouts, lossless (Walsh-Hadamard), screen content (palette, intra block copy)
    "cif_444": (352, 288, 5, 44, {"subsampling": "4:4:4"}, 60, 4, 6.0),
    "cif_422": (352, 288, 5, 45, {"subsampling": "4:2:2"}, 60, 4, 6.0),
    "cif_400": (352, 288, 4, 46, {"subsampling": "4:0:0"}, 60, 6, 6.0),
    "qcif_lossless": (176, 144, 3, 47, {"subsampling": "4:4:4"}, 100, 6, 3.0),
    "cif_screen": (352, 288, 5, 48, {"subsampling": "4:4:4", "tune-content": "screen"}, 60, 4, 0.0),
    # config C1 (easyav1_benchmark's input): 640x360 in a WebM container for the unchanged easyAV1 API.
    # chroma-sample-position is set because easyav1_get_video_frame refuses streams that leave it
    # unknown (src/easyav1.c:3493-3503)
    "c1_360p_webm": (640, 360, 48, 2, {"enable-cdef": "1", "enable-restoration": "1", "chroma-sample-position": "1",
                                       "container": "webm"}, 55, 6),
    # config C1 at its full length (BASELINE.json configs[0]: 240 frames), the bench's easyAV1-API leg
    "c1_360p_bench": (640, 360, 240, 3, {"enable-cdef": "1", "enable-restoration": "1", "chroma-sample-position": "1",
                                        "container": "webm"}, 55, 6),
    # config C4 for the bench (BASELINE.json configs[3]): 3840x2160 10-bit, 30 frames of drifting / rotating /
    # zooming content, generated in float (synth_zoom_frames_float): 2.9 MB = 98 KB per frame, a streaming-like
    # 24 Mbit/s at 30 frames/s (the integer recipe costs 7.7 MB at the same quality, 1 MB per frame at quality 30-40)
    "c4_2160p_10bit_bench": (3840, 2160, 30, 34, {"depth": 10, "content": "zoom_float"}, 12, 8, 0.0),
    # the bench workload (BASELINE.json configs[1]): same recipe, 120 frames. Speed 4 because
    # speed 6 never signals loop restoration on this content (checked from the recorded frames)
    "c2_1080p_bench": (1920, 1080, 120, 7, {"enable-cdef": "1", "enable-restoration": "1"}, 60, 4, 6.0),
}


def main():
    only = sys.argv[1:]
    manifest_path = os.path.join(HERE, "manifest.json")
    manifest = json.load(open(manifest_path)) if os.path.exists(manifest_path) else {}
    for name, spec in CLIPS.items():
        w, h, n, seed, adv, q, spd = spec[:7]
        noise = spec[7] if len(spec) > 7 else 3.0
        if only and name not in only:
            continue
        if adv.get("content") == "zoom_float":
            avif = encode_cv2_10bit(synth_zoom_frames_float(w, h, n, seed), q, spd)
        elif adv.get("content") == "zoom":
            avif = encode_cv2_10bit(synth_zoom_frames(w, h, n, seed, noise=noise), q, spd)
        else:
            frames = synth_frames(w, h, n, seed, noise=noise)
            avif = encode_cv2(frames, adv["depth"], q, spd) if "depth" in adv else encode(frames, adv, q, spd)
        samples = avif_samples(avif)
        ivf = os.path.join(HERE, name + ".ivf")
        write_ivf(ivf, samples, w, h)
        if adv.get("container") == "webm":
            from tests.webm import mux_av1
            open(os.path.join(HERE, name + ".webm"), "wb").write(mux_av1(samples, w, h))
        obu = b"".join(samples)
        g = golden_md5(obu)
        manifest[name] = {"file": name + ".ivf", "w": w, "h": h, "frames": g, "bytes": len(obu),
                          "encoder": "OpenCV AVIF/aom" if "depth" in adv else "Pillow AVIF/aom", "advanced": adv, "quality": q, "speed": spd,
                          "seed": seed, "noise": noise}
        print(name, len(samples), "samples", len(obu), "bytes ->", len(g), "decoded frames",
              [x["frame_type"] for x in g])
    json.dump(manifest, open(manifest_path, "w"), indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
