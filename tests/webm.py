"""Minimal WebM (Matroska) writer for the fixtures: one V_AV1 video track, SimpleBlocks in
clusters, Segment Info with TimecodeScale + Duration — what easyAV1's demuxer (nestegg) asks for
(src/easyav1.c:1281-1400, 1611-1620). TEST INFRASTRUCTURE."""
import struct


def _vint(n, length=None):
    if length is None:
        length = 1
        while n >= (1 << (7 * length)) - 1:
            length += 1
    return ((1 << (7 * length)) | n).to_bytes(length, "big")


def _el(eid, payload):
    return bytes.fromhex(eid) + _vint(len(payload)) + payload


def _uint(eid, v):
    n = max(1, (v.bit_length() + 7) // 8)
    return _el(eid, v.to_bytes(n, "big"))


def mux_av1(samples, w, h, fps=30, keyframes=None):
    """samples: list of AV1 temporal units (bytes). Returns the WebM file as bytes."""
    frame_ns = 1000000000 // fps
    head = _el("1A45DFA3", _uint("4286", 1) + _uint("42F7", 1) + _uint("42F2", 4) + _uint("42F3", 8) +
               _el("4282", b"webm") + _uint("4287", 2) + _uint("4285", 2))
    dur_ms = len(samples) * 1000.0 / fps
    info = _el("1549A966", _uint("2AD7B1", 1000000) + _el("4489", struct.pack(">d", dur_ms)) +
               _el("4D80", b"easyav1_b200 tests") + _el("5741", b"easyav1_b200 tests"))
    video = _el("E0", _uint("B0", w) + _uint("BA", h))
    track = _el("AE", _uint("D7", 1) + _uint("73C5", 1) + _uint("83", 1) + _uint("9C", 0) + _el("86", b"V_AV1") +
                _uint("23E383", frame_ns) + video)
    tracks = _el("1654AE6B", track)
    clusters = b""
    per_cluster = 30
    for c0 in range(0, len(samples), per_cluster):
        t0 = c0 * 1000 // fps
        body = _uint("E7", t0)
        for i in range(c0, min(c0 + per_cluster, len(samples))):
            rel = i * 1000 // fps - t0
            key = (i == 0) if keyframes is None else (i in keyframes)
            body += _el("A3", b"\x81" + struct.pack(">h", rel) + (b"\x80" if key else b"\x00") + samples[i])
        clusters += _el("1F43B675", body)
    segment = _el("18538067", info + tracks + clusters)
    return head + segment
