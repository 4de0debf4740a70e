"""Badread extras on the host (no GPU): csrc/extras.cpp builds the molecule of every read as a segment list; a small
native harness (tests/native/extras_emul.cpp) materialises those segments from the host copy of the genome, and the
result must equal oracle/extras_oracle.py's restatement on plain strings byte for byte -- the same comparison
tests/test_extras.py makes through the kernels."""
import os
import subprocess

import numpy as np
import pytest

from oracle import extras_oracle as xo
from oracle import seq_oracle as so
from test_gpu_parity import handmade_mdf, make_fasta

HERE = os.path.dirname(os.path.abspath(__file__))
SEED = 4242
CONFIGS = [
    dict(junk_reads=20.0, random_reads=15.0),
    dict(start_adapter=(90.0, 60.0), end_adapter=(50.0, 20.0)),
    dict(start_adapter=(100.0, 100.0), end_adapter=(100.0, 100.0), start_adapter_seq="ACGTTGCA", end_adapter_seq="ttgacc"),
    dict(glitches=(150.0, 25.0, 25.0)),
    dict(glitches=(40.0, 0.0, 7.0)),
    dict(chimeras=30.0, start_adapter=(40.0, 50.0), end_adapter=(40.0, 50.0)),
    dict(junk_reads=5.0, random_reads=5.0, chimeras=10.0, start_adapter=(90.0, 60.0), end_adapter=(50.0, 20.0),
         glitches=(300.0, 25.0, 25.0)),
]


@pytest.fixture(scope="module")
def emul(built, tmp_path_factory):
    from tksm_b200.sequence import Sequencer
    csrc = os.path.join(HERE, "..", "tksm_b200", "csrc")
    bdir = os.path.join(HERE, "native", "_build")
    os.makedirs(bdir, exist_ok=True)
    exe = os.path.join(bdir, "extras_emul")
    subprocess.check_call(["g++", "-O1", "-std=c++20", "-I" + csrc, "-o", exe, os.path.join(HERE, "native", "extras_emul.cpp"),
                           os.path.join(csrc, "host.cpp"), os.path.join(csrc, "extras.cpp"), os.path.join(csrc, "tail.cpp")])
    tmp = tmp_path_factory.mktemp("extras")
    fasta = make_fasta(np.random.default_rng(77), [("chr1", 200000), ("chr2", 120000)])
    s = Sequencer(host_only=True, seed=SEED)
    s.add_reference_text(fasta)
    s.load_models("random", "random")
    mdf = s.synth_mdf(3, 150, seed=5) + handmade_mdf().encode() + s.synth_mdf(1, 250, seed=6, min_len=300, max_len=9000)
    (tmp / "g.fa").write_text(fasta)
    (tmp / "in.mdf").write_bytes(mdf)
    refs = so.load_fasta(fasta)
    seqs = [so.molecule_sequence(refs, iv).decode() for _, iv in so.parse_mdf(mdf)]
    return exe, str(tmp / "g.fa"), str(tmp / "in.mdf"), seqs


@pytest.mark.parametrize("cfg", CONFIGS)
@pytest.mark.parametrize("ordinal_base", [0, 1000])
def test_segment_level_extras_equal_the_string_restatement(emul, cfg, ordinal_base):
    exe, fa, mdf, seqs = emul
    args = [cfg.get("junk_reads", 0), cfg.get("random_reads", 0), cfg.get("chimeras", 0), *cfg.get("start_adapter", (0, 0)),
            *cfg.get("end_adapter", (0, 0)), *cfg.get("glitches", (0, 0, 0)), cfg.get("start_adapter_seq", xo.START_ADAPTER),
            cfg.get("end_adapter_seq", xo.END_ADAPTER)]
    out = subprocess.run([exe, fa, mdf, str(SEED), str(ordinal_base)] + [str(a) for a in args], capture_output=True, text=True, check=True)
    got = out.stdout.split("\n")[:-1]
    want = xo.apply(seqs, SEED, ordinal_base, **cfg)
    assert len(got) == len(want) == len(seqs)
    bad = [i for i, (g, w) in enumerate(zip(got, want)) if g != w]
    assert not bad, (bad[:5], got[bad[0]][:80], want[bad[0]][:80])
    assert sum(g != m for g, m in zip(got, seqs)) > 20      # the flags did something


def test_all_off_leaves_the_molecules_alone(emul):
    exe, fa, mdf, seqs = emul
    out = subprocess.run([exe, fa, mdf, str(SEED), "0"] + ["0"] * 10 + [xo.START_ADAPTER, xo.END_ADAPTER], capture_output=True, text=True, check=True)
    assert out.stdout.split("\n")[:-1] == seqs


# ---------------------------------------------------------------- fused front-end steps
FUSE_SPECS = [
    "plA:normal=15,7.5;Flp:0.5;Tag:AATGTACTTCGTTCAGTTACGTATTGCT,GCAATACGTAACTGAACGAAGT",     # README bulk pipeline
    "plA:poisson=20;Tag:,10;SCB;Tag:,AGATCGGAAGAGCGTCGTGTAG",                                  # README single-cell head
    "plA:gamma=2,8,5,40;Flp:1",
    "plA:gamma=0.6,30;Flp:0.3;plA:weibull=1.5,30,0,25;Tag:NNRYBU,4",
]


def with_cb(mdf, seed):
    """Every header gets a CB= comment: a barcode, now and then '.', now and then two values."""
    rng = np.random.default_rng(seed)
    out = []
    for ln in mdf.decode().split("\n"):
        if ln.startswith("+"):
            f = ln.split("\t")
            r = rng.random()
            cb = "." if r < 0.1 else "".join(rng.choice(list("ACGT"), 16))
            if r > 0.9:
                cb += ",TTTT"
            f[2] = f"CB={cb};" + (f[2] if len(f) > 2 else "")
            ln = "\t".join(f[:3])
        out.append(ln)
    return "\n".join(out).encode()


@pytest.fixture(scope="module")
def fuse_world(emul, tmp_path_factory):
    from tksm_b200.sequence import Sequencer
    exe = emul[0]
    tmp = tmp_path_factory.mktemp("fuse")
    fasta = make_fasta(np.random.default_rng(78), [("chr1", 150000), ("chr2", 90000)])
    s = Sequencer(host_only=True, seed=SEED)
    s.add_reference_text(fasta)
    s.load_models("random", "random")
    mdf = with_cb(s.synth_mdf(0, 200, seed=4, min_len=200, max_len=4000) + handmade_mdf().encode() + s.synth_mdf(3, 100, seed=5), 3)
    (tmp / "g.fa").write_text(fasta)
    (tmp / "in.mdf").write_bytes(mdf)
    return exe, str(tmp / "g.fa"), str(tmp / "in.mdf"), so.load_fasta(fasta), mdf


def run_emul(exe, fa, mdf_path, ordinal_base, spec, cfg=None):
    cfg = cfg or {}
    args = [cfg.get("junk_reads", 0), cfg.get("random_reads", 0), cfg.get("chimeras", 0), *cfg.get("start_adapter", (0, 0)),
            *cfg.get("end_adapter", (0, 0)), *cfg.get("glitches", (0, 0, 0)), cfg.get("start_adapter_seq", xo.START_ADAPTER),
            cfg.get("end_adapter_seq", xo.END_ADAPTER), spec]
    out = subprocess.run([exe, fa, mdf_path, str(SEED), str(ordinal_base)] + [str(a) for a in args], capture_output=True, text=True, check=True)
    return out.stdout.split("\n")[:-1]


@pytest.mark.parametrize("spec", FUSE_SPECS)
@pytest.mark.parametrize("ordinal_base", [0, 77])
def test_fused_steps_equal_the_interval_restatement(fuse_world, spec, ordinal_base):
    exe, fa, mdf_path, refs, mdf = fuse_world
    reads = xo.front_end(xo.parse_records(mdf), SEED, ordinal_base, spec)
    want = [so.molecule_sequence(refs, iv).decode() for _, iv in reads]
    got = run_emul(exe, fa, mdf_path, ordinal_base, spec)
    assert len(got) == len(want) == len(so.parse_mdf(mdf))
    bad = [i for i, (g, w) in enumerate(zip(got, want)) if g != w]
    assert not bad, (bad[:5], got[bad[0]][-60:], want[bad[0]][-60:])
    # ... and the per-read extras act on the molecules the steps made (chimeras take the fused molecule of the read before)
    cfg = dict(chimeras=30.0, start_adapter=(50.0, 50.0), glitches=(400.0, 10.0, 10.0), junk_reads=4.0)
    assert run_emul(exe, fa, mdf_path, ordinal_base, spec, cfg) == xo.apply(want, SEED, ordinal_base, **cfg)


def test_fused_steps_follow_their_modules_in_distribution(fuse_world, tmp_path):
    """Five thousand records (10 000 reads) through plA / Flp / Tag: lengths, rates and letters as the standalone modules define them."""
    from tksm_b200.sequence import Sequencer
    exe, fa, _, refs, _ = fuse_world
    s = Sequencer(host_only=True, seed=SEED)
    s.add_reference_text(open(fa).read())
    s.load_models("random", "random")
    mdf = s.synth_mdf(0, 10000, seed=8, min_len=200, max_len=1500)
    recs = xo.parse_records(mdf)
    first = np.concatenate(([0], np.cumsum([d for _, d, _, _ in recs])[:-1]))     # first read of every record
    all_mols = [so.molecule_sequence(refs, iv).decode() for _, iv in so.parse_mdf(mdf)]
    mols = [all_mols[i] for i in first]
    assert len(mols) >= 5000
    p = tmp_path / "big.mdf"
    p.write_bytes(mdf)
    emul_all = globals()["run_emul"]

    def run_emul(*a):           # one read per record: the depth copies of a record share the result
        out = emul_all(*a)
        assert len(out) == len(all_mols)
        assert all(out[i] == out[i - 1] for i in range(1, len(out)) if i not in first_set and all_mols[i] == all_mols[i - 1])
        return [out[i] for i in first]
    first_set = set(first.tolist())
    comp = str.maketrans("ACGT", "TGCA")
    # polyA: normal(15, 7.5) truncated toward zero and clamped at 0 (src/polyA.cpp:133-148)
    got = run_emul(exe, fa, str(p), 0, "plA:normal=15,7.5")
    tails = np.array([len(g) - len(m) for g, m in zip(got, mols)])
    assert all(g == m + "A" * t for g, m, t in zip(got, mols, tails))
    x = np.random.default_rng(1).normal(15, 7.5, 400000)
    expect = np.clip(np.trunc(x), 0, 5000)
    assert abs(tails.mean() - expect.mean()) < 0.3 and abs(tails.std() - expect.std()) < 0.3
    assert abs((tails == 0).mean() - (expect == 0).mean()) < 0.01
    # poisson(20) and weibull(1.5, 30): means
    t = np.array([len(g) - len(m) for g, m in zip(run_emul(exe, fa, str(p), 0, "plA:poisson=20"), mols)])
    assert abs(t.mean() - 20) < 0.3 and abs(t.var() - 20) < 1.5
    t = np.array([len(g) - len(m) for g, m in zip(run_emul(exe, fa, str(p), 0, "plA:weibull=1.5,30"), mols)])
    assert abs(t.mean() - np.trunc(30 * np.random.default_rng(2).weibull(1.5, 400000)).mean()) < 0.6
    t = np.array([len(g) - len(m) for g, m in zip(run_emul(exe, fa, str(p), 0, "plA:gamma=2,8"), mols)])
    assert abs(t.mean() - np.trunc(np.random.default_rng(3).gamma(2, 8, 400000)).mean()) < 0.5
    # flip: half of the molecules reverse-complemented (src/strand_man.cpp:47-53)
    got = run_emul(exe, fa, str(p), 0, "Flp:0.5")
    flipped = sum(g == m.translate(comp)[::-1] and g != m for g, m in zip(got, mols))
    assert flipped + sum(g == m for g, m in zip(got, mols)) >= len(mols) - 2 and abs(flipped / len(mols) - 0.5) < 0.025
    # tag: 10 random bases in front (uniform letters), the fixed adapter behind (src/tag.cpp:84-114)
    got = run_emul(exe, fa, str(p), 0, "Tag:10,GCAATACGTAACTGAACGAAGT")
    assert all(g[10:] == m + "GCAATACGTAACTGAACGAAGT" for g, m in zip(got, mols))
    letters = np.array([list(g[:10]) for g in got]).ravel()
    assert all(abs((letters == b).mean() - 0.25) < 0.01 for b in "ACGT")


def test_host_parser_token_and_integer_corner_cases(emul, tmp_path):
    """csrc/host.cpp: MdfParser compares contig names through their first 16 bytes and reads plain digit runs eight at a
    time; everything else takes python's int() (py/sequence.py:207, 216-217).  Names of 1 to 33 bytes, literals of the
    same lengths (one byte off a contig name), integers with leading zeros, signs, blanks, underscores and nine to
    thirteen digits, modification lists and a last line without a line end: the molecules must be the oracle's."""
    import random
    exe = emul[0]
    rng = random.Random(6)
    names = ["c", "chr12345", "chr_fifteen_abc", "chr_sixteen_abcd", "chr_seventeen_abc", "chr_seventeen_abd",
             "a_very_long_contig_name_of_thirty", "ACGTACGTACGTACGTA"]
    fasta = "".join(f">{n}\n{''.join(rng.choice('ACGTN') for _ in range(5000))}\n" for n in names)
    lits = ["A", "ACGTACGT", "ACGTACGTACGTACG", "ACGTACGTACGTACGT", "ACGTACGTACGTACGTC", "chr_seventeen_abX",
            "chr_sixteen_abcX", "acgtnnRYacgtacgtacgtacgtacgtac", "a_very_long_contig_name_of_thirtY"]

    def fmt(v):
        r = rng.random()
        if v < 0 or r > 0.45:
            return str(v)
        if r < 0.15:
            return f"{v:09d}"
        if r < 0.25:
            return f"{v:016d}"
        if r < 0.30:
            return f" {v} "
        if r < 0.35:
            return f"+{v}"
        if r < 0.40 and v >= 1000:
            return f"{v // 1000}_{v % 1000:03d}"
        return f"{v:017d}"

    lines = []
    for m in range(1500):
        lines.append(f"+m{m}\t{fmt(rng.choice([0, 1, 1, 1, 2]))}\tCB=AAAC,x;")
        for _ in range(rng.randint(1, 8)):
            if rng.random() < 0.6:
                a = rng.randint(0, 4000)
                b = a + rng.randint(1, 900)
                if rng.random() < 0.1:
                    b = rng.choice([123456789, 1234567890123, 99999999, 100000000, 4999, 5000, 5001])
                mods = ""
                if rng.random() < 0.3 and b <= 5000:
                    mods = ",".join(f"{fmt(rng.randint(0, b - a - 1)).strip()}{rng.choice('ACGTacgtN')}" for _ in range(rng.randint(1, 3)))
                lines.append(f"{rng.choice(names)}\t{fmt(a)}\t{fmt(b)}\t{rng.choice('+-')}\t{mods}")
            else:
                lines.append(f"{rng.choice(lits)}\t{fmt(rng.randint(-3, 5))}\t{fmt(rng.randint(-3, 40))}\t{rng.choice('+-')}\t")
    mdf = "\n".join(lines)                      # the last token of the text has nothing readable behind it
    (tmp_path / "odd.fa").write_text(fasta)
    (tmp_path / "odd.mdf").write_text(mdf)
    refs = so.load_fasta(fasta)
    want = [so.molecule_sequence(refs, iv).decode() for _, iv in so.parse_mdf(mdf)]
    assert len(want) > 1000 and any(len(w) > 3000 for w in want)
    out = subprocess.run([exe, str(tmp_path / "odd.fa"), str(tmp_path / "odd.mdf"), str(SEED), "0"] + ["0"] * 10 +
                         [xo.START_ADAPTER, xo.END_ADAPTER], capture_output=True, text=True, check=True)
    assert out.stdout.split("\n")[:-1] == want
