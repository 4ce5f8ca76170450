"""Minimal Vorbis I stream writer for the audio fixtures: the three header packets
(identification, comment, setup) written bit by bit after the Vorbis I specification as
minivorbis parses it (vorbis_synthesis_headerin -> _vorbis_unpack_info/_comment/_books,
ext/minivorbis/minivorbis.h), and audio packets whose payload is seeded random bits.

All codebooks are complete Huffman trees (every entry has length log2(entries)), so ANY bit
sequence decodes to valid floor posts, residue classifications and VQ entries; a packet that
runs out of bits ends the way the specification says (remaining values are zero). The spectra the
reference's host code hands to mdct_backward are therefore dense, bounded and reproducible,
and both block sizes, all four window transitions, channel coupling and residue types 1 and 2
occur. TEST INFRASTRUCTURE: nothing here is used by the product.
"""
import random
import struct


class BitWriter:
    """LSB-first bit packer (oggpack_write)"""

    def __init__(self):
        self.acc = 0
        self.n = 0

    def put(self, value, bits):
        assert 0 <= value < (1 << bits) or bits == 0, (value, bits)
        self.acc |= value << self.n
        self.n += bits

    def raw(self, data):
        for b in data:
            self.put(b, 8)

    def bytes(self):
        return self.acc.to_bytes((self.n + 7) // 8, "little")


def ilog(v):
    return v.bit_length()


def float32_pack(x):
    """Vorbis packed float (inverse of _float32_unpack): 21-bit mantissa, exponent biased by 788"""
    if x == 0:
        return 0
    sign = 0x80000000 if x < 0 else 0
    x = abs(x)
    exp = 0
    # mantissa as an integer < 2^21
    while x != int(x) or x >= (1 << 21):
        if x >= (1 << 21):
            x /= 2
            exp += 1
        else:
            x *= 2
            exp -= 1
    return sign | ((exp + 788) << 21) | int(x)


def codebook(w, dim, entries, lookup=None):
    """lookup: None or (minimum, delta, value_bits, multiplicands) for lookup type 1"""
    length = entries.bit_length() - 1
    assert 1 << length == entries
    w.put(0x564342, 24)
    w.put(dim, 16)
    w.put(entries, 24)
    w.put(0, 1)            # not ordered
    w.put(0, 1)            # not sparse
    for _ in range(entries):
        w.put(length - 1, 5)
    if lookup is None:
        w.put(0, 4)
        return
    minimum, delta, value_bits, mults = lookup
    w.put(1, 4)
    w.put(float32_pack(minimum), 32)
    w.put(float32_pack(delta), 32)
    w.put(value_bits - 1, 4)
    w.put(0, 1)            # sequence_p
    quantvals = 1
    while (quantvals + 1) ** dim <= entries:
        quantvals += 1
    assert len(mults) == quantvals, (len(mults), quantvals)
    for m in mults:
        w.put(m, value_bits)


def headers(channels=2, rate=48000, bs0=256, bs1=2048):
    ident = BitWriter()
    ident.raw(b"\x01vorbis")
    ident.put(0, 32)
    ident.put(channels, 8)
    ident.put(rate, 32)
    ident.put(0, 32); ident.put(0, 32); ident.put(0, 32)
    ident.put(bs0.bit_length() - 1, 4)
    ident.put(bs1.bit_length() - 1, 4)
    ident.put(1, 1)

    vendor = b"easyav1_b200 tests"
    comment = BitWriter()
    comment.raw(b"\x03vorbis")
    comment.put(len(vendor), 32)
    comment.raw(vendor)
    comment.put(0, 32)
    comment.put(1, 1)

    s = BitWriter()
    s.raw(b"\x05vorbis")
    # -- codebooks
    s.put(5 - 1, 8)
    codebook(s, 1, 16)                                             # 0: floor post deviations
    codebook(s, 2, 4)                                              # 1: residue classification words (2 classes ^ 2)
    codebook(s, 2, 16, (-2.0, 1.0, 2, [0, 1, 2, 3]))                # 2: coarse VQ
    codebook(s, 4, 16, (-0.375, 0.25, 2, [0, 3]))                   # 3: fine VQ
    codebook(s, 1, 8, (-1.0, 0.25, 4, [0, 1, 3, 4, 5, 7, 8, 12]))   # 4: scalar
    # -- time domain transforms (placeholder)
    s.put(0, 6); s.put(0, 16)
    # -- floors: two floor1 set-ups
    s.put(2 - 1, 6)
    for posts, mult in (([4, 12, 24, 40, 56, 72, 96, 112], 2), ([3, 9, 20, 33, 50, 64, 90, 120, 150, 200, 230, 250], 1)):
        rangebits = 7 if max(posts) < 128 else 8
        dim = 4
        s.put(1, 16)
        s.put(len(posts) // dim, 5)
        for _ in range(len(posts) // dim):
            s.put(0, 4)                # every partition uses class 0
        s.put(dim - 1, 3); s.put(0, 2)  # class 0: 4 posts, no subclasses
        s.put(0 + 1, 8)                # its single book: 0
        s.put(mult - 1, 2)
        s.put(rangebits, 4)
        for x in posts:
            s.put(x, rangebits)
    # -- residues
    s.put(2 - 1, 6)
    for rtype, end, grouping, books in ((1, 96, 16, [2, 4]), (2, 1536, 32, [2, 3])):
        s.put(rtype, 16)
        s.put(0, 24); s.put(end, 24); s.put(grouping - 1, 24)
        s.put(2 - 1, 6)                # two classifications
        s.put(1, 8)                    # classification book
        s.put(0, 3); s.put(0, 1)       # class 0: no stages (zeros)
        s.put(3, 3); s.put(0, 1)       # class 1: two stages
        for b in books:
            s.put(b, 8)
    # -- mappings
    s.put(2 - 1, 6)
    for floor, residue, coupled in ((0, 0, False), (1, 1, channels >= 2)):
        s.put(0, 16)
        s.put(0, 1)                    # one submap
        if coupled:
            s.put(1, 1); s.put(0, 8)
            s.put(0, ilog(channels - 1)); s.put(1, ilog(channels - 1))
        else:
            s.put(0, 1)
        s.put(0, 2)
        s.put(0, 8); s.put(floor, 8); s.put(residue, 8)
    # -- modes: 0 = short blocks, 1 = long blocks
    s.put(2 - 1, 6)
    for blockflag, mapping in ((0, 0), (1, 1)):
        s.put(blockflag, 1); s.put(0, 16); s.put(0, 16); s.put(mapping, 8)
    s.put(1, 1)
    return ident.bytes(), comment.bytes(), s.bytes()


def packets(n, seed, bs0=256, bs1=2048, payload=(40, 260), long_ratio=0.6):
    """n audio packets; returns [(bytes, blockflag)]. The first bits are written explicitly (packet
    type 0, the mode, and for long blocks the two window flags), the rest is random."""
    rnd = random.Random(seed)
    out = []
    for i in range(n):
        long_block = rnd.random() < long_ratio
        w = BitWriter()
        w.put(0, 1)
        w.put(1 if long_block else 0, 1)
        if long_block:
            w.put(rnd.getrandbits(1), 1); w.put(rnd.getrandbits(1), 1)
        nbytes = rnd.randint(*payload) if long_block else rnd.randint(payload[0] // 4, payload[1] // 4)
        while w.n < nbytes * 8:
            w.put(rnd.getrandbits(8), 8)
        out.append((w.bytes()[:nbytes], int(long_block)))
    return out


def samples_after(flags, bs0=256, bs1=2048):
    """samples made final after each packet (vorbis_synthesis_blockin: blocksize[lW]/4 + blocksize[W]/4,
    nothing for the first block)"""
    bs = (bs0, bs1)
    out, prev = [], None
    for f in flags:
        out.append(0 if prev is None else bs[prev] // 4 + bs[f] // 4)
        prev = f
    return out


def xiph_private(hdrs):
    """Matroska CodecPrivate of A_VORBIS: Xiph lacing of the three headers"""
    out = bytes([len(hdrs) - 1])
    for h in hdrs[:-1]:
        n = len(h)
        out += b"\xff" * (n // 255) + bytes([n % 255])
    return out + b"".join(hdrs)


if __name__ == "__main__":
    i, c, s = headers()
    print(len(i), len(c), len(s), struct.unpack("<I", i[12:16]))
