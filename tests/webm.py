"""Minimal WebM (Matroska) writer for the fixtures: one V_AV1 video track and optionally one
A_VORBIS audio track, SimpleBlocks in clusters, Segment Info with TimecodeScale + Duration, Cues for
the cluster starts — what easyAV1's demuxer (nestegg) asks for (src/easyav1.c:1281-1400, 1428-1500,
1611-1620, 2929-2959). TEST INFRASTRUCTURE."""
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


def _fixed_uint(eid, v, n=8):
    return _el(eid, v.to_bytes(n, "big"))


def mux_av1_vorbis(samples, w, h, fps, vorbis_private, audio, rate, channels, per_cluster=30):
    """samples: AV1 temporal units (the first one a key frame); audio: [(packet bytes, start sample)].
    Track 1 = video, track 2 = audio; blocks of a cluster are interleaved by time; one cue per cluster."""
    frame_ns = 1000000000 // fps
    head = _el("1A45DFA3", _uint("4286", 1) + _uint("42F7", 1) + _uint("42F2", 4) + _uint("42F3", 8) +
               _el("4282", b"webm") + _uint("4287", 2) + _uint("4285", 2))
    dur_ms = len(samples) * 1000.0 / fps
    info = _el("1549A966", _uint("2AD7B1", 1000000) + _el("4489", struct.pack(">d", dur_ms)) +
               _el("4D80", b"easyav1_b200 tests") + _el("5741", b"easyav1_b200 tests"))
    vtrack = _el("AE", _uint("D7", 1) + _uint("73C5", 1) + _uint("83", 1) + _uint("9C", 0) + _el("86", b"V_AV1") +
                 _uint("23E383", frame_ns) + _el("E0", _uint("B0", w) + _uint("BA", h)))
    atrack = _el("AE", _uint("D7", 2) + _uint("73C5", 2) + _uint("83", 2) + _uint("9C", 0) + _el("86", b"A_VORBIS") +
                 _el("63A2", vorbis_private) +
                 _el("E1", _el("B5", struct.pack(">d", float(rate))) + _uint("9F", channels)))
    tracks = _el("1654AE6B", vtrack + atrack)
    blocks = [(i * 1000 // fps, 0, 1, i == 0, samples[i]) for i in range(len(samples))]
    blocks += [(start * 1000 // rate, 1, 2, True, pkt) for pkt, start in audio]
    blocks.sort(key=lambda b: (b[0], b[1]))
    clusters, times = [], []
    span = per_cluster * 1000 // fps
    t0 = 0
    while t0 * fps < len(samples) * 1000 or any(b[0] >= t0 for b in blocks):
        mine = [b for b in blocks if t0 <= b[0] < t0 + span]
        if not mine and not any(b[0] >= t0 for b in blocks):
            break
        body = _uint("E7", t0)
        for t, _, track, key, data in mine:
            body += _el("A3", bytes([0x80 | track]) + struct.pack(">h", t - t0) + (b"\x80" if key else b"\x00") + data)
        clusters.append(_el("1F43B675", body))
        times.append(t0)
        t0 += span

    def cues_for(first_cluster_pos):
        pos, pts = first_cluster_pos, b""
        for t, c in zip(times, clusters):
            pts += _el("BB", _fixed_uint("B3", t) + _el("B7", _fixed_uint("F7", 1, 1) + _fixed_uint("F1", pos)))
            pos += len(c)
        return _el("1C53BB6B", pts)

    cues_len = len(cues_for(0))
    cues = cues_for(len(info) + len(tracks) + cues_len)
    assert len(cues) == cues_len
    segment = _el("18538067", info + tracks + cues + b"".join(clusters))
    return head + segment
