"""Tail-noise model (py/tksm_badread.py:886-1036).

CPU: the product's draws (tksm_seq_tail_draw) equal the oracle restatement byte for byte, both follow the
statistics of the unmodified reference (tests/golden/ref_tail.json.gz, oracle/make_goldens.py tail), and the
loader rejects what the reference rejects.  GPU: reads with tails through the kernels equal the CPU twin
fed with the oracle's tails; the perfect output and error_free_length leave the tail out."""
import gzip
import json
import os

import numpy as np
import pytest

from oracle import seq_oracle as so
from oracle import tail_oracle
from tksm_b200.sequence import BADREAD, PERFECT, SeqError, Sequencer

GOLD = os.path.join(os.path.dirname(__file__), "golden")
MODEL = os.path.join(GOLD, "tail_model.json")


def golden():
    with gzip.open(os.path.join(GOLD, "ref_tail.json.gz"), "rt") as f:
        return json.load(f)


def host_sequencer(seed=42):
    s = Sequencer(seed=seed, host_only=True)
    s.load_models("random", "random", MODEL)
    return s


def test_draws_equal_the_oracle():
    tm = tail_oracle.TailModel.load(MODEL)
    for seed in (42, (7 << 32) | 123):
        s = host_sequencer(seed)
        n_tail = 0
        for i, frag_len in enumerate([1, 150, 199, 200, 201, 340, 360, 500, 999, 1400, 2000, 2001, 70000]):
            for ordinal in list(range(400 * i, 400 * i + 400)) + [(1 << 32) + i, (5 << 40) + 17 * i]:
                want = tm.noise_seq(seed, ordinal, frag_len)
                got = s.tail_draw(ordinal, frag_len)
                assert got == want, (seed, ordinal, frag_len)
                n_tail += bool(got)
        assert n_tail > 500


def test_no_noise_draws_nothing():
    s = Sequencer(host_only=True)
    s.load_models("random", "random")
    assert all(s.tail_draw(o, 500) == b"" for o in range(50))
    s.load_models("random", "random", "no_noise")
    assert s.tail_draw(3, 500) == b""


def test_distribution_follows_the_reference():
    g = golden()
    s = host_sequencer(1234)
    n = 40000
    for frag_len, ref in g["frag_lens"].items():
        frag_len = int(frag_len)
        lens, pairs, first = {}, np.zeros((4, 4)), np.zeros(4)
        for o in range(n):
            t = s.tail_draw(o + 100000 * (frag_len % 97), frag_len)
            lens[len(t)] = lens.get(len(t), 0) + 1
            if t:
                idx = ["AGTC".index(chr(c)) for c in t]
                first[idx[0]] += 1
                np.add.at(pairs, (idx[:-1], idx[1:]), 1)
        ref_lens = {int(k): v for k, v in ref["lens"].items()}
        assert set(lens) <= set(ref_lens) | {0}, (frag_len, sorted(lens))
        # total variation between the two length histograms (both 40 k draws; sampling noise ~0.01)
        keys = set(lens) | set(ref_lens)
        tv = 0.5 * sum(abs(lens.get(k, 0) / n - ref_lens.get(k, 0) / g["n"]) for k in keys)
        assert tv < 0.02, (frag_len, tv)
        rp = np.array(ref["pairs"], float)
        if rp.sum() == 0:
            assert pairs.sum() == 0
            continue
        # transition rows of the Markov chain
        for a in range(4):
            if rp[a].sum() > 2000:
                assert np.abs(pairs[a] / pairs[a].sum() - rp[a] / rp[a].sum()).max() < 0.02, (frag_len, a)
        rf = np.array(ref["first"], float)
        assert np.abs(first / first.sum() - rf / rf.sum()).max() < 0.03, frag_len


def test_label_quirks_follow_the_reference():
    """Custom2Dist.__call__ (py/tksm_badread.py:1014-1024): the row is the first label >= y unless the NEXT label is
    nearer (340 and 360 both use the 500 row), and past the last label the multiplier is the label COUNT over the
    last label (4 / 2000), which empties every tail."""
    s = host_sequencer(9)
    assert [s.tail_draw(o, 340) for o in range(3000)] == [s.tail_draw(o, 500) for o in range(3000)]
    assert all(s.tail_draw(o, 2001) == b"" for o in range(3000))
    assert any(s.tail_draw(o, 2000) for o in range(3000))


def test_loader_errors(tmp_path):
    s = Sequencer(host_only=True)
    with pytest.raises(SeqError):
        s.load_models("random", "random", str(tmp_path / "missing.json"))
    bad = tmp_path / "bad.json"
    dc = json.load(open(MODEL))
    del dc["ratio"]
    bad.write_text(json.dumps(dc))
    with pytest.raises(SeqError, match="ratio"):      # KeyError: 'ratio' in the reference
        s.load_models("random", "random", str(bad))
    bad.write_text("{not json")
    with pytest.raises(SeqError):
        s.load_models("random", "random", str(bad))
    gz = tmp_path / "m.json.gz"
    with gzip.open(gz, "wt") as f:
        json.dump(json.load(open(MODEL)), f)
    s.load_models("random", "random", str(gz))          # .gz like gzip.open(model_name, "rt")
    assert any(s.tail_draw(o, 500) for o in range(200))


def test_model_name_lookup(tmp_path, monkeypatch):
    """*.tail.gz files under $TKSM_MODELS/badread are found by name (py/sequence.py:17-31)."""
    d = tmp_path / "models" / "badread"
    d.mkdir(parents=True)
    with gzip.open(d / "mytail.tail.gz", "wt") as f:
        json.dump(json.load(open(MODEL)), f)
    monkeypatch.setenv("TKSM_MODELS", str(tmp_path / "models"))
    s = Sequencer(host_only=True)
    s.load_models("random", "random", "mytail")
    assert any(s.tail_draw(o, 500) for o in range(200))


# --------------------------------------------------------------------------------------------- GPU

def world():
    rng = np.random.default_rng(21)
    g = rng.choice(np.frombuffer(b"ACGT", np.uint8), size=80_000).tobytes().decode()
    fasta = ">c1\n" + g + "\n"
    lines = []
    for i in range(700):
        st = int(rng.integers(0, 78000))
        ln = int(rng.choice([150, 340, 500, 1400, 1999, 2600]))
        depth = int(rng.choice([1, 1, 1, 3]))
        lines.append(f"+m{i}\t{depth}\t\nc1\t{st}\t{st + ln}\t{'+-'[i & 1]}\t\n")
        if i % 50 == 0:   # a literal segment and a modification in front of the tail
            lines[-1] += f"ACGTTTGA\t0\t8\t+\t\n"
        if i % 70 == 0:
            lines[-1] = f"+m{i}\t{depth}\t\nc1\t{st}\t{st + ln}\t-\t3G,7T\n"
    return fasta, "".join(lines)


def expected(s, fasta, mdf, seed, ordinal_base, badread, fastq):
    tm = tail_oracle.TailModel.load(MODEL)
    refs = so.load_fasta(fasta)
    mols = so.parse_mdf(mdf)
    seqs = [so.molecule_sequence(refs, iv) for _, iv in mols]
    ids = [m for m, _ in mols]
    tails = [tm.noise_seq(seed, ordinal_base + r, len(q)) for r, q in enumerate(seqs)]
    assert sum(map(bool, tails)) > 50
    return so.twin_simulate(s.tables(), seqs, ids, ordinal_base=ordinal_base, badread=badread, fastq=fastq, tails=tails)


@pytest.mark.gpu
@pytest.mark.parametrize("models", [("random", "random"), ("nanopore2020", "nanopore2020")])
def test_gpu_reads_with_tails_equal_the_twin(models):
    fasta, mdf = world()
    s = Sequencer(seed=77)
    s.add_reference_text(fasta)
    s.load_models(models[0], models[1], MODEL)
    want = expected(s, fasta, mdf, 77, 1000, True, True)
    got = s.simulate(mdf, ordinal_base=1000)
    assert got == want
    # device-resident batch path
    b = s.batch(mdf, ordinal_base=1000)
    b.run()
    assert b.fetch() == want
    # perfect output: no tails, error_free_length without the tail
    wantp = expected(s, fasta, mdf, 77, 1000, False, False)
    gotp = s.simulate(mdf, which=PERFECT, fastq=False, ordinal_base=1000)
    assert gotp == wantp
    s2 = Sequencer(seed=77)
    s2.add_reference_text(fasta)
    s2.load_models(models[0], models[1])
    assert s2.simulate(mdf, which=PERFECT, fastq=False, ordinal_base=1000) == wantp


@pytest.mark.gpu
def test_gpu_tails_do_not_depend_on_chunking():
    fasta, mdf = world()
    outs = []
    for chunk, reads in ((1 << 20, 0), (4096, 64), (700, 5)):
        os.environ["TKSM_SEQ_CHUNK_BYTES"] = str(chunk)
        try:
            s = Sequencer(seed=3, batch_reads=reads) if reads else Sequencer(seed=3)
            s.add_reference_text(fasta)
            s.load_models("random", "random", MODEL)
            outs.append(s.simulate(mdf, ordinal_base=12345))
        finally:
            del os.environ["TKSM_SEQ_CHUNK_BYTES"]
    assert outs[0] == outs[1] == outs[2]


@pytest.mark.gpu
def test_gpu_end_to_end_extra_length_follows_the_reference():
    """600 molecules of 500 bases at 99.5 % identity through the unmodified sequence.py with the tail model:
    length - error_free_length is the tail (plus a little error noise)."""
    ref = np.array(golden()["e2e"]["extra"])
    rng = np.random.default_rng(3)
    g = rng.choice(np.frombuffer(b"ACGT", np.uint8), size=60_000).tobytes().decode()
    mdf = "".join(f"+m{i}\t1\t\nc1\t{st}\t{st + 500}\t+\t\n" for i, st in enumerate(rng.integers(0, 59000, 6000)))
    s = Sequencer(identity=(99.5, 99.9, 0.1), seed=8)
    s.add_reference_text(">c1\n" + g + "\n")
    s.load_models("random", "random", MODEL)
    recs = so.parse_records(s.simulate(mdf, fastq=False), fastq=False)
    extra = []
    for h, _, _ in recs:
        kv = dict(t.split("=") for t in h.split()[1:])
        assert kv["error_free_length"] == "500"
        extra.append(int(kv["length"]) - 500)
    extra = np.array(extra)
    for thr in (2, 5, 20, 80, 150):
        a, b = np.mean(extra > thr), np.mean(ref > thr)
        assert abs(a - b) < 4 * np.sqrt(b * (1 - b) / len(ref) + 1e-4) + 0.01, (thr, a, b)


@pytest.mark.gpu
def test_gpu_cli_tail_model_flag(tmp_path):
    """`tksm sequence ... --badread-tail-model <path>` (py/sequence.py:108-114, 343-345) writes the bytes of the
    library call; both outputs at once: the perfect file holds no tails."""
    import subprocess
    exe = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tksm_b200", "tksm")
    fasta, mdf = world()
    s = Sequencer(seed=7)
    s.add_reference_text(fasta)
    s.load_models("random", "random", MODEL)
    want, want_perf = s.simulate(mdf, BADREAD, True), s.simulate(mdf, PERFECT, False)
    (tmp_path / "g.fa").write_text(fasta)
    (tmp_path / "in.mdf").write_text(mdf)
    out, perf = tmp_path / "o.fastq", tmp_path / "p.fasta"
    p = subprocess.run([exe, "sequence", "-i", str(tmp_path / "in.mdf"), "-r", str(tmp_path / "g.fa"), "-o", str(out),
                        "--perfect", str(perf), "--badread-error-model", "random", "--badread-qscore-model", "random",
                        "--badread-tail-model", MODEL, "-s", "7", "-t", "2"], capture_output=True, text=True, timeout=120)
    assert p.returncode == 0, p.stderr
    assert out.read_bytes() == want and perf.read_bytes() == want_perf
    p = subprocess.run([exe, "sequence", "-i", str(tmp_path / "in.mdf"), "-r", str(tmp_path / "g.fa"), "-o", str(out),
                        "--badread-tail-model", str(tmp_path / "nope.json")], capture_output=True, text=True, timeout=120)
    assert p.returncode == 1 and "Error" in p.stderr
