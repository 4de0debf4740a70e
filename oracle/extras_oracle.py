"""TEST INFRASTRUCTURE — CPU restatement of the Badread extras (junk / random reads, adapters, chimeras, glitches)
with the draws of the B200 path.

The reference (/root/reference/py/tksm_badread.py) does not have these features: only the two unused constants
CHIMERA_START_ADAPTER_CHANCE / CHIMERA_END_ADAPTER_CHANCE (:61-63) hint at them, and the upstream Badread source is
not part of /root/reference.  PARITY UNPINNED: this module restates the statistical definition of DESIGN.md 3.10
(after the semantics `badread simulate` documents for --junk_reads, --random_reads, --chimeras, --start_adapter,
--end_adapter, --glitches) on plain strings, with the Philox draws of tksm_b200/csrc/extras.hpp, so that the product's
segment-level implementation (csrc/extras.cpp) must agree with it byte for byte.  Only tests/ may import this file.
"""
import math

from oracle.tail_oracle import philox4x32_10, unit

TAG_EXTRA = 5
START_ADAPTER = "AATGTACTTCGTTCAGTTACGTATTGCT"
END_ADAPTER = "GCAATACGTAACTGAACGAAGT"
CHIMERA_END_ADAPTER_CHANCE = 0.25
CHIMERA_START_ADAPTER_CHANCE = 0.25


def words(seed, ordinal, index, kind, g=0):
    return philox4x32_10((index, TAG_EXTRA | (kind << 8) | (g << 12), ordinal & 0xFFFFFFFF, ordinal >> 32),
                         (seed & 0xFFFFFFFF, seed >> 32))


def random_bases(seed, ordinal, kind, g, n):
    out = []
    w = None
    for i in range(n):
        if i & 63 == 0:
            w = words(seed, ordinal, i >> 6, kind, g)
        out.append("ACGT"[(w[(i >> 4) & 3] >> (2 * (i & 15))) & 3])
    return "".join(out)


def geometric(mean, u):
    """numpy's geometric(1 / mean) by inversion (support 1, 2, ...)."""
    if not mean > 1.0:
        return 1
    k = math.ceil(math.log(1.0 - u) / math.log(1.0 - 1.0 / mean))
    return max(1, min(int(k), 4000000000))


def geometric0(mean, u):
    return geometric(mean, u) if mean > 0.0 else 0


def adapter_bases(length, amount, u):
    """How many bases of an adapter are present: all of it for amount >= 1, else length * x, x ~ Beta(a / (1 - a), 1)
    (mean a: 'the average fraction of the adapter present'), rounded half up."""
    if amount >= 1.0:
        return length
    if not amount > 0.0:
        return 0
    return min(length, int(length * math.pow(u, (1.0 - amount) / amount) + 0.5))


def add_glitches(frag, seed, ordinal, rate, size, skip):
    """Badread's add_glitches: pieces of geometric length kept, then `size` random bases put in and `skip` bases lost."""
    out, i, g = [], 0, 0
    while True:
        w = words(seed, ordinal, 0, 3, g & 0xFFFFF)
        dist = geometric(rate, unit(w[0]))
        out.append(frag[i:i + dist])
        i += dist
        if i >= len(frag):
            break
        out.append(random_bases(seed, ordinal, 4, g & 0xFFFFF, geometric0(size, unit(w[1]))))
        i += geometric0(skip, unit(w[2]))
        g += 1
    return "".join(out)


def apply(seqs, seed, ordinal_base=0, junk_reads=0.0, random_reads=0.0, chimeras=0.0, start_adapter=(0.0, 0.0),
          end_adapter=(0.0, 0.0), glitches=(0.0, 0.0, 0.0), start_adapter_seq=START_ADAPTER, end_adapter_seq=END_ADAPTER,
          tails=None):
    """seqs: the error-free molecule of every read, in read order (upper case; str or bytes).  Returns the fragment
    every read is sequenced from, as str.  Percentages as on the command line (0..100).  tails: optional list of the tail of every read
    (oracle/tail_oracle.py), which goes between the molecule and what follows it."""
    junk, rnd, chim = junk_reads / 100.0, random_reads / 100.0, chimeras / 100.0
    s_rate, s_amount = start_adapter[0] / 100.0, start_adapter[1] / 100.0
    e_rate, e_amount = end_adapter[0] / 100.0, end_adapter[1] / 100.0
    start_adapter_seq, end_adapter_seq = (start_adapter_seq or "").upper(), (end_adapter_seq or "").upper()
    out = []
    seqs = [m.decode() if isinstance(m, (bytes, bytearray)) else m for m in seqs]
    for r, mol in enumerate(seqs):
        o = ordinal_base + r
        w = words(seed, o, 0, 0)
        wa = words(seed, o, 0, 2)
        parts = []
        if s_rate > 0 and start_adapter_seq and unit(wa[0]) < s_rate:
            k = adapter_bases(len(start_adapter_seq), s_amount, unit(wa[1]))
            parts.append(start_adapter_seq[len(start_adapter_seq) - k:])
        u0 = unit(w[0])
        if junk > 0 and u0 < junk:
            ul = 1 + w[1] % 5
            parts.append("".join("ACGT"[(w[2] >> (2 * (i % ul))) & 3] for i in range(len(mol))))
        elif rnd > 0 and u0 < junk + rnd:
            parts.append(random_bases(seed, o, 1, 0, len(mol)))
        else:
            parts.append(mol)
        if tails is not None:
            parts.append(tails[r])
        if chim > 0 and r > 0 and unit(w[3]) < chim:      # (the first read of a call has no read before it)
            wc = words(seed, o, 0, 5)
            if end_adapter_seq and unit(wc[0]) < CHIMERA_END_ADAPTER_CHANCE:
                parts.append(end_adapter_seq)
            if start_adapter_seq and unit(wc[1]) < CHIMERA_START_ADAPTER_CHANCE:
                parts.append(start_adapter_seq)
            parts.append(seqs[r - 1])
        if e_rate > 0 and end_adapter_seq and unit(wa[2]) < e_rate:
            parts.append(end_adapter_seq[:adapter_bases(len(end_adapter_seq), e_amount, unit(wa[3]))])
        frag = "".join(parts)
        if glitches[0] > 0:
            frag = add_glitches(frag, seed, o, glitches[0], glitches[1], glitches[2])
        out.append(frag)
    return out


# ---------------------------------------------------------------- fused front-end steps
# Restated on interval lists, following the modules of the reference (draws: Philox as in csrc/extras.hpp, since the
# modules' std::mt19937 streams are not reproduced):
#   plA  /root/reference/src/polyA.cpp:133-148      Flp  src/strand_man.cpp:47-53, src/interval.h:908-920
#   Tag  src/tag.cpp:84-114, src/util.h:53-92       SCB  src/scb.cpp:73-81
TAG_FUSE = 6
IUPAC = {"A": "A", "G": "G", "T": "T", "C": "C", "U": "U", "R": "GA", "Y": "TC", "K": "GT", "M": "AC", "S": "GC", "W": "AT",
         "B": "GTC", "D": "GAT", "H": "ACT", "V": "GCA", "N": "AGCT"}          # src/util.h:61-76


def fuse_words(seed, ordinal, step, index):
    return philox4x32_10((index, TAG_FUSE | (step << 8), ordinal & 0xFFFFFFFF, ordinal >> 32), (seed & 0xFFFFFFFF, seed >> 32))


def parse_records(text):
    """[(id, depth, comment, [(chrom, start, end, strand, mods), ...])] of the records with depth > 0."""
    if isinstance(text, (bytes, bytearray)):
        text = text.decode()
    out, cur = [], None
    for ln in text.split("\n"):
        if not ln:
            continue
        f = ln.split("\t")
        if f[0][0] == "+":
            cur = (f[0][1:], int(f[1]), f[2] if len(f) > 2 else "", [])
            if cur[1] > 0:
                out.append(cur)
        elif cur is not None:
            f = f + [""] if len(f) == 4 else f
            cur[3].append((f[0], int(f[1]), int(f[2]), f[3], f[4]))
    return out


def polya_length(dist, params, lo, hi, seed, ordinal, step):
    w = fuse_words(seed, ordinal, step, 0)
    two_pi = 6.283185307179586
    if dist == "normal":
        x = params[0] + params[1] * (math.sqrt(-2.0 * math.log(unit(w[0]))) * math.cos(two_pi * unit(w[1])))
    elif dist == "weibull":
        x = params[1] * math.pow(-math.log(1.0 - unit(w[0])), 1.0 / params[0])
    elif dist == "poisson":
        u, pk, k = unit(w[0]), math.exp(-params[0]), 0
        c = pk
        while u > c and k < 100000:
            k += 1
            pk *= params[0] / k
            c += pk
        x = float(k)
    else:   # gamma: shape, scale (Marsaglia-Tsang)
        a = params[0] + 1.0 if params[0] < 1.0 else params[0]
        d = a - 1.0 / 3.0
        c = 1.0 / math.sqrt(9.0 * d)
        x = 0.0
        for it in range(1, 1000):
            v = fuse_words(seed, ordinal, step, it)
            z = math.sqrt(-2.0 * math.log(unit(v[0]))) * math.cos(two_pi * unit(v[1]))
            t = 1.0 + c * z
            if t <= 0.0:
                continue
            t3 = t * t * t
            if math.log(unit(v[2])) < 0.5 * z * z + d - d * t3 + d * math.log(t3):
                x = d * t3
                break
        if params[0] < 1.0:
            x *= math.pow(unit(w[3]), 1.0 / params[0])
        x *= params[1]
    return max(lo, min(hi, int(x)))       # int(): toward zero, like the module's `int poly_a_len = dist(rand_gen)`


def tag_sequence(fmt, seed, ordinal, step, base):
    if fmt and fmt[0].isdigit():
        fmt = "N" * int(fmt)
    out, w = [], None
    for j, ch in enumerate(fmt):
        if j & 3 == 0:
            w = fuse_words(seed, ordinal, step, base + (j >> 2))
        s = IUPAC.get(ch, "")
        if s:
            out.append(s[(w[j & 3] * len(s)) >> 32])
    return "".join(out)


def front_end(records, seed, ordinal_base, spec):
    """records: parse_records(); spec as tksm_seq_set_front_end takes it.  Returns (id, intervals) per READ (depth
    expanded), i.e. what oracle/seq_oracle.parse_mdf would give for the MDF the module pipeline would have written."""
    steps = [s for s in spec.split(";") if s]
    out, ordinal = [], ordinal_base
    for rid, depth, comment, ivs in records:
        ivs = list(ivs)
        for k, st in enumerate(steps):
            name, _, arg = st.partition(":")
            if name == "plA":
                dist, _, nums = arg.partition("=")
                v = [float(x) for x in nums.split(",")]
                npar = 1 if dist == "poisson" else 2
                lo, hi = (int(v[npar]), int(v[npar + 1])) if len(v) == npar + 2 else (0, 5000)
                n = polya_length(dist, v[:npar], lo, hi, seed, ordinal, k)
                if n > 0:
                    ivs.append(("A" * n, 0, n, "+", ""))
            elif name == "Flp":
                if unit(fuse_words(seed, ordinal, k, 0)[0]) < float(arg):
                    ivs = [(c, s, e, "-" if sd == "+" else "+", m) for c, s, e, sd, m in reversed(ivs)]
            elif name == "Tag":
                f5, _, f3 = arg.partition(",")
                t5, t3 = tag_sequence(f5, seed, ordinal, k, 0), tag_sequence(f3, seed, ordinal, k, 1 << 16)
                if t5:
                    ivs.insert(0, (t5, 0, len(t5), "+", ""))
                if t3:
                    ivs.append((t3, 0, len(t3), "+", ""))
            elif name == "SCB":
                cb = None
                for item in comment.split(";"):
                    if item.startswith("CB="):
                        cb = item[3:].split(",")[0]
                        break
                if not cb:
                    raise KeyError("CB")
                if cb != ".":
                    ivs.append((cb, 0, len(cb), "+", ""))
            else:
                raise ValueError(name)
        out.extend((rid, ivs) for _ in range(depth))
        ordinal += depth
    return out
