"""Badread extras (junk / random reads, adapters, chimeras, glitches): tksm_seq_set_extras, csrc/extras.cpp.

The reference has none of them (parity unpinned), so the checks are: (1) with every rate at zero the
output is byte-identical to a run without the call; (2) the fragment every read is sequenced from -- which the
error-free output shows -- equals oracle/extras_oracle.py's restatement on plain strings byte for byte, and the
errored FASTQ equals the CPU twin run on those fragments; (3) on 10 000 reads the rates and lengths are the ones
the flags ask for; (4) the streaming pipeline gives the bytes of a single device batch (chimeras reach across
device batches).
"""
import os

import numpy as np
import pytest

from oracle import extras_oracle as xo
from oracle import seq_oracle as so
from test_gpu_parity import first_diff, handmade_mdf, make_fasta

pytestmark = pytest.mark.gpu

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
def world(built):
    from tksm_b200.sequence import Sequencer
    rng = np.random.default_rng(77)
    fasta = make_fasta(rng, [("chr1", 200000), ("chr2", 120000)])
    s = Sequencer(seed=SEED)
    s.add_reference_text(fasta)
    s.load_models("random", "random")
    return {"s": s, "refs": so.load_fasta(fasta), "fasta": fasta}


def molecules(world, mdf):
    mols = so.parse_mdf(mdf)
    return [so.molecule_sequence(world["refs"], iv).decode() for _, iv in mols], [m for m, _ in mols]


def records(out, fastq):
    lines = out.decode().split("\n")
    per = 4 if fastq else 2
    return [(lines[i], lines[i + 1]) for i in range(0, len(lines) - 1, per)]


def test_all_rates_zero_is_byte_identical(world):
    from tksm_b200.sequence import BADREAD, PERFECT
    s = world["s"]
    mdf = s.synth_mdf(1, 300, seed=3) + handmade_mdf().encode()
    s.set_extras()
    want, want_perf = s.simulate(mdf, BADREAD, True), s.simulate(mdf, PERFECT, False)
    s.set_extras(junk_reads=0.0, start_adapter=(0.0, 60.0), glitches=(0.0, 25.0, 25.0))
    assert s.simulate(mdf, BADREAD, True) == want and s.simulate(mdf, PERFECT, False) == want_perf
    s.set_extras()


@pytest.mark.parametrize("cfg", CONFIGS)
def test_fragments_equal_the_string_restatement_and_reads_equal_the_twin(world, cfg):
    from tksm_b200.sequence import BADREAD, PERFECT
    s = world["s"]
    mdf = s.synth_mdf(3, 150, seed=5) + handmade_mdf().encode() + s.synth_mdf(1, 250, seed=6, min_len=300, max_len=9000)
    seqs, ids = molecules(world, mdf)
    try:
        s.set_extras(**cfg)
        want = xo.apply(seqs, SEED, 0, **cfg)
        got = records(s.simulate(mdf, PERFECT, False), False)
        assert len(got) == len(want)
        bad = [(i, len(g[1]), len(w)) for i, (g, w) in enumerate(zip(got, want)) if g[1] != w]
        assert not bad, bad[:5]
        # the same through a device-resident batch at another ordinal base (no read before the first one: the
        # restatement is given the same view)
        b = s.batch(mdf, ordinal_base=1000)
        b.run(PERFECT, 0)
        got2 = records(b.fetch(PERFECT), False)
        want2 = xo.apply(seqs, SEED, 1000, **cfg)
        assert [g[1] for g in got2] == want2
        # errors, alignment and qscores on those fragments: the CPU twin
        fq = s.simulate(mdf, BADREAD, True)
        twin = so.twin_simulate(s.tables(), [w.encode() for w in want], ids, ordinal_base=0, badread=True, fastq=True)
        d = first_diff(fq, twin, True)
        assert d is None, d
    finally:
        s.set_extras()


def test_streaming_batches_carry_the_chimera_partner(world, monkeypatch):
    from tksm_b200.sequence import PERFECT, Sequencer
    s = world["s"]
    mdf = s.synth_mdf(1, 2000, seed=9)
    cfg = dict(chimeras=25.0, glitches=(500.0, 10.0, 10.0), junk_reads=3.0)
    try:
        s.set_extras(**cfg)
        want = s.simulate(mdf, PERFECT, False)
    finally:
        s.set_extras()
    seqs, _ = molecules(world, mdf)
    assert [r[1] for r in records(want, False)] == xo.apply(seqs, SEED, 0, **cfg)
    monkeypatch.setenv("TKSM_SEQ_CHUNK_BYTES", "20000")
    s2 = Sequencer(seed=SEED, batch_reads=53)
    s2.add_reference_text(make_fasta(np.random.default_rng(77), [("chr1", 200000), ("chr2", 120000)]))
    s2.load_models("random", "random")
    s2.set_extras(**cfg)
    assert s2.simulate(mdf, PERFECT, False) == want


def test_rates_and_lengths_on_ten_thousand_reads(world):
    from tksm_b200.sequence import PERFECT
    s = world["s"]
    mdf = s.synth_mdf(1, 10000, seed=12, min_len=300, max_len=3000)
    seqs, _ = molecules(world, mdf)
    n = len(seqs)

    def fragments(**cfg):
        try:
            s.set_extras(**cfg)
            return [r[1] for r in records(s.simulate(mdf, PERFECT, False), False)]
        finally:
            s.set_extras()

    def periodic(x):
        return any(x == (x[:p] * (len(x) // p + 1))[:len(x)] for p in range(1, 6))

    def near(count, p, total=n):   # binomial, 4.5 sigma
        return abs(count - p * total) <= 4.5 * (p * (1 - p) * total) ** 0.5 + 1

    # junk 7 %, random 4 %: same length as the molecule, periodic / not
    fr = fragments(junk_reads=7.0, random_reads=4.0)
    changed = [(f, m) for f, m in zip(fr, seqs) if f != m]
    assert all(len(f) == len(m) for f, m in changed)
    assert near(sum(periodic(f) for f, _ in changed), 0.07) and near(sum(not periodic(f) for f, _ in changed), 0.04)
    # adapters: rate = share of reads that carry one, amount = mean share of it present
    fr = fragments(start_adapter=(90.0, 60.0), end_adapter=(50.0, 20.0))
    ks = np.array([f.index(m) for f, m in zip(fr, seqs)])
    ke = np.array([len(f) - len(m) - k for f, m, k in zip(fr, seqs, ks)])
    assert all(f[:k] == xo.START_ADAPTER[len(xo.START_ADAPTER) - k:] for f, k in zip(fr, ks))
    assert all(f[len(f) - k:] == xo.END_ADAPTER[:k] for f, k in zip(fr, ke) if k)
    # (a present adapter can round to zero bases: the restatement says how often at these settings)
    u = (np.arange(200000) + 0.5) / 200000
    exp_s = np.floor(28 * u ** ((1 - 0.6) / 0.6) + 0.5)
    exp_e = np.floor(22 * u ** ((1 - 0.2) / 0.2) + 0.5)
    assert near((ks > 0).sum(), 0.9 * (exp_s > 0).mean()) and near((ke > 0).sum(), 0.5 * (exp_e > 0).mean())
    assert abs(ks[ks > 0].mean() - exp_s[exp_s > 0].mean()) < 0.5 and abs(ke[ke > 0].mean() - exp_e[exp_e > 0].mean()) < 0.5
    assert abs(exp_s.mean() / 28 - 0.6) < 0.02 and abs(exp_e.mean() / 22 - 0.2) < 0.02
    # glitches every 1000 bases on average, 25 bases in, none lost: the reads grow by 25 per glitch
    fr = fragments(glitches=(1000.0, 25.0, 0.0))
    total = sum(len(m) for m in seqs)
    grown = sum(len(f) for f in fr) - total
    expect = total / 1000.0 * 25.0
    assert abs(grown - expect) < 0.06 * expect
    # ... and 25 bases lost per glitch, nothing put in: they shrink (a glitch near the end loses less)
    fr = fragments(glitches=(1000.0, 0.0, 25.0))
    lost = total - sum(len(f) for f in fr)
    assert 0.85 * expect < lost < 1.03 * expect
    # chimeras: 12 % of the reads run on into the molecule before them
    fr = fragments(chimeras=12.0)
    chim = sum(1 for i in range(1, n) if fr[i] != seqs[i] and fr[i].startswith(seqs[i]) and fr[i].endswith(seqs[i - 1]))
    assert near(chim, 0.12) and sum(f != m for f, m in zip(fr, seqs)) == chim


# ---------------------------------------------------------------- fused front-end steps
@pytest.mark.parametrize("spec", [
    "plA:normal=15,7.5;Flp:0.5;Tag:AATGTACTTCGTTCAGTTACGTATTGCT,GCAATACGTAACTGAACGAAGT",
    "plA:poisson=20;Tag:,10;SCB;Tag:,AGATCGGAAGAGCGTCGTGTAG",
])
def test_fused_front_end_through_the_kernels(world, spec, tmp_path):
    """tksm_seq_set_front_end / --fuse: the molecules the steps make (oracle/extras_oracle.front_end on interval lists)
    are what the error-free output shows and what the CPU twin must be given to reproduce the errored reads; the
    command line flag gives the same bytes."""
    import subprocess
    from test_extras_host import with_cb
    from tksm_b200 import _lib
    from tksm_b200.sequence import BADREAD, PERFECT
    s = world["s"]
    mdf = with_cb(s.synth_mdf(0, 300, seed=4, min_len=200, max_len=6000) + handmade_mdf().encode() + s.synth_mdf(3, 100, seed=5), 3)
    reads = xo.front_end(xo.parse_records(mdf), SEED, 0, spec)
    want = [so.molecule_sequence(world["refs"], iv).decode() for _, iv in reads]
    ids = [i for i, _ in reads]
    try:
        s.set_front_end(spec)
        perfect = s.simulate(mdf, PERFECT, False)
        assert [r[1] for r in records(perfect, False)] == want
        fq = s.simulate(mdf, BADREAD, True)
        d = first_diff(fq, so.twin_simulate(s.tables(), [w.encode() for w in want], ids, badread=True, fastq=True), True)
        assert d is None, d
        # with the per-read extras on top
        cfg = dict(chimeras=20.0, end_adapter=(60.0, 40.0), glitches=(600.0, 8.0, 8.0))
        s.set_extras(**cfg)
        assert [r[1] for r in records(s.simulate(mdf, PERFECT, False), False)] == xo.apply(want, SEED, 0, **cfg)
    finally:
        s.set_extras()
        s.set_front_end(None)
    assert s.simulate(mdf, PERFECT, False) != perfect
    # the front-end binary
    (tmp_path / "g.fa").write_text(world["fasta"])
    (tmp_path / "in.mdf").write_bytes(mdf)
    exe = os.path.join(os.path.dirname(_lib.LIB_PATH), "tksm")
    subprocess.check_call([exe, "sequence", "-i", str(tmp_path / "in.mdf"), "-r", str(tmp_path / "g.fa"), "--perfect",
                           str(tmp_path / "p.fasta"), "-s", str(SEED), "--fuse", spec], stderr=subprocess.DEVNULL)
    assert (tmp_path / "p.fasta").read_bytes() == perfect
