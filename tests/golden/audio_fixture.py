"""Builds the WebM fixture with an audio track: the AV1 samples of c1_360p_webm.ivf
(committed) + a synthetic Vorbis track from vorbis_synth.py, muxed by tests/webm.py. The file is small and
fully determined by the seeds, so it is built on demand instead of being committed. TEST INFRASTRUCTURE."""
import os

from tests import webm
from tests.golden import vorbis_synth as vs
from tests.golden.gen_fixtures import read_ivf

HERE = os.path.dirname(os.path.abspath(__file__))
RATE, CHANNELS, BS0, BS1 = 48000, 2, 256, 2048


def build(seed=7, seconds=1.6, channels=CHANNELS, bs0=BS0, bs1=BS1):
    samples = read_ivf(os.path.join(HERE, "c1_360p_webm.ivf"))
    hdrs = vs.headers(channels, RATE, bs0, bs1)
    audio, start, n = [], 0, 0
    pk = vs.packets(400, seed, bs0, bs1)
    fin = vs.samples_after([f for _, f in pk], bs0, bs1)
    while n < len(pk) and start < seconds * RATE:
        audio.append((pk[n][0], start))
        start += fin[n]
        n += 1
    data = webm.mux_av1_vorbis(samples, 640, 360, 30, vs.xiph_private(hdrs), audio, RATE, channels)
    return data, [f for _, f in pk[:n]]
