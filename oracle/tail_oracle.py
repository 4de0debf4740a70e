"""TEST INFRASTRUCTURE — CPU restatement of the reference's tail-noise model with the draws of the B200 path.

Follows /root/reference/py/tksm_badread.py:
  KDE_noise_generator.load / __init__ / noise_seq   :886-959
  CustomDist.__init__ / __call__                    :970-985
  Custom2Dist.__call__                              :1014-1024
  call site (tail between molecule and right flank) :335-341

The reference draws from Python's unseeded `random`; the product keys Philox4x32-10 by (seed, global read
ordinal) instead (tksm_b200/csrc/tail.hpp).  This module restates the reference's arithmetic in plain
Python floats with those draws, so that `Sequencer.tail_draw` must agree with it byte for byte, and
tests/golden/ref_tail.json.gz (statistics of the unmodified reference, oracle/make_goldens.py tail) pins
the distribution.  Only tests/, smoke() and bench.py's CPU legs may import this file.
"""
import bisect
import gzip
import json

TAG_READ, TAG_TAIL = 0, 4
M32 = 0xFFFFFFFF


def philox4x32_10(c, k):
    c0, c1, c2, c3 = c
    k0, k1 = k
    for _ in range(10):
        p0 = 0xD2511F53 * c0
        p1 = 0xCD9E8D57 * c2
        c0, c1, c2, c3 = ((p1 >> 32) ^ c1 ^ k0) & M32, p1 & M32, ((p0 >> 32) ^ c3 ^ k1) & M32, p0 & M32
        k0 = (k0 + 0x9E3779B9) & M32
        k1 = (k1 + 0xBB67AE85) & M32
    return [c0, c1, c2, c3]


def unit(w):
    return (w + 0.5) / 4294967296.0


class TailModel:
    def __init__(self, dc):
        self.lx = [float(x) for x in dc["lx"]]
        self.ly = [float(y) for y in dc["ly"]]
        self.cdf = []
        for i in range(len(self.ly)):                       # CustomDist.__init__ :977-980
            pdf = [float(v) for v in dc["grid"][i]]
            total = 0.0
            for v in pdf:
                total += v
            acc, row = 0.0, []
            for v in pdf:
                acc = v / total + acc
                row.append(acc)
            self.cdf.append(row)
        self.trans = [[float(v) for v in row] for row in dc["trans"]]
        self.ratio = float(dc["ratio"])
        self.bases = list(dc["bases"])

    @classmethod
    def load(cls, path):
        with (gzip.open(path, "rt") if path.endswith(".gz") else open(path)) as f:
            return cls(json.load(f))

    def noise_seq(self, seed, ordinal, frag_len):
        """bytes: the tail of read `ordinal` (noise_seq(len(fragment)), :910-927)."""
        key = (seed & M32, seed >> 32)
        olo, ohi = ordinal & M32, ordinal >> 32
        w = philox4x32_10((3, TAG_READ, olo, ohi), key)
        if unit(w[0]) > self.ratio:                         # :911
            return b""
        y = float(frag_len)                                 # Custom2Dist.__call__ :1014-1024
        pos = bisect.bisect_left(self.ly, y)
        if pos < len(self.ly) - 1 and abs(self.ly[pos] - y) > abs(self.ly[pos + 1] - y):
            pos += 1
        mult = 1.0
        if pos >= len(self.ly):
            mult = pos / self.ly[-1]
            pos = len(self.ly) - 1
        q = bisect.bisect_left(self.cdf[pos], unit(w[1]))    # CustomDist.__call__ :982-985
        q = min(q, len(self.lx) - 1)
        x = int(self.lx[q] * mult)
        if x <= 0:
            return b""
        index = w[2] & 3                                    # random.choices(arange(4)) :918
        out = bytearray()
        for i in range(x):
            if i & 3 == 0:
                d = philox4x32_10((i >> 2, TAG_TAIL, olo, ohi), key)
            t = self.trans[index]
            c0 = t[0]
            c1 = c0 + t[1]
            c2 = c1 + t[2]
            c3 = c2 + t[3]
            v = unit(d[i & 3]) * c3                         # random.choices(..., weights): bisect_right, capped
            index = 0 if v < c0 else 1 if v < c1 else 2 if v < c2 else 3
            out += self.bases[index].encode()
        return bytes(out)
