"""oracle/bulk_inputs.py — synthetic inputs of BASELINE configs[1] without the product.  TEST INFRASTRUCTURE.

bench.py's CPU legs (`--impl reference`, `cpu_baseline`) time the reference's algorithm on molecules of the README
bulk pipeline shape Tsb -> Trc -> plA -> Flp -> Tag (README.md:100-120) and must not depend on the library under
test for their inputs: this module builds a small random genome (FASTA text) and the MDF of such molecules with
numpy, in the emitter format of src/interval.h:898-905, and oracle/seq_oracle.py materialises the sequences
(py/sequence.py:303-313).  Same distributions as tksm_b200/csrc/synth.cpp shape 1: 1 + Geom(1/8) exons of
log-normal length (median 150) and introns (median 1500), 3'-anchored truncation to a log-normal target (median
900), polyA N(15, 7.5), flip with p = 0.5, 28-nt and 22-nt tags.
"""
import numpy as np

TAG5, TAG3 = "AATGTACTTCGTTCAGTTACGTATTGCT", "GCAATACGTAACTGAACGAAGT"


def genome(n_bases=8_000_000, n_contigs=4, seed=42):
    rng = np.random.default_rng(seed)
    sizes = [n_bases // n_contigs] * n_contigs
    out, names = [], []
    for i, n in enumerate(sizes):
        g = rng.choice(np.frombuffer(b"ACGT", np.uint8), size=n)
        p = 0
        while p < n:                       # 1 % N in runs
            if rng.random() < 0.01:
                run = int(rng.integers(1, 2000))
                g[p:p + run] = ord("N")
                p += run
            p += int(rng.integers(1000, 50000))
        s = g.tobytes().decode()
        names.append((f"chr{i + 1}", n))
        out.append(f">chr{i + 1}\n" + "\n".join(s[j:j + 60] for j in range(0, n, 60)) + "\n")
    return "".join(out), names


def bulk_mdf(contigs, n_mol, seed=1000):
    rng = np.random.default_rng(seed)
    tot = sum(n for _, n in contigs)
    lines = []
    m = 0
    while m < n_mol:
        name, clen = contigs[int(np.searchsorted(np.cumsum([n for _, n in contigs]), rng.random() * tot))]
        nex = 1
        while nex < 40 and rng.random() > 1 / 8:
            nex += 1
        el = np.clip(rng.lognormal(np.log(150), 0.7, nex), 20, 5000).astype(np.int64)
        il = np.clip(rng.lognormal(np.log(1500), 1.0, nex), 50, 100000).astype(np.int64)
        il[0] = 0
        span = int(el.sum() + il.sum())
        if span + 2 > clen:
            continue
        pos = int(rng.random() * (clen - span))
        segs = []
        for e in range(nex):
            pos += int(il[e])
            segs.append([name, pos, pos + int(el[e]), True])
            pos += int(el[e])
        plus = rng.random() < 0.5
        if not plus:
            segs.reverse()
        for s in segs:
            s[3] = plus
        target = int(min(max(rng.lognormal(np.log(900), 0.6), 100), 20000))       # Trc: keep the 3' end
        total = sum(s[2] - s[1] for s in segs)
        while total > target and segs:
            s = segs[0]
            ln = s[2] - s[1]
            cut = min(ln, total - target)
            if cut == ln:
                segs.pop(0)
            elif s[3]:
                s[1] += cut
            else:
                s[2] -= cut
            total -= cut
        pa = int(round(15 + 7.5 * rng.normal()))                                  # plA
        if pa > 0:
            segs.append(["A" * pa, 0, pa, True])
        if rng.random() < 0.5:                                                    # Flp
            segs.reverse()
            for s in segs:
                s[3] = not s[3]
        segs.insert(0, [TAG5, 0, len(TAG5), True])                                # Tag
        segs.append([TAG3, 0, len(TAG3), True])
        lines.append(f"+M{m}\t1\ttid=T{m};")
        for c, a, b, p in segs:
            lines.append(f"{c}\t{a}\t{b}\t{'+' if p else '-'}\t")
        m += 1
    return "\n".join(lines) + "\n"


def molecules(n_mol, seed=1000, genome_bases=8_000_000):
    """(error-free sequences, molecule ids) of n_mol bulk-shape molecules."""
    from oracle import seq_oracle as so
    fasta, contigs = genome(genome_bases)
    refs = so.load_fasta(fasta)
    mols = so.parse_mdf(bulk_mdf(contigs, n_mol, seed))
    seqs = [so.molecule_sequence(refs, iv) for _, iv in mols]
    return [s.encode() if isinstance(s, str) else s for s in seqs], [m for m, _ in mols]
