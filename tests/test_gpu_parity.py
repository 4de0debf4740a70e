"""GPU parity: the CUDA path, called through the C ABI, against the oracles on the same inputs.

* perfect output (interval fetch, mods, reverse-complement, literal segments, ordering, record text)
  must equal the Python restatement of the reference byte for byte;
* badread output must equal the CPU twin of the PB1 model bit for bit (oracle/twin.c).
"""
import os

import numpy as np
import pytest

from oracle import seq_oracle as so

pytestmark = pytest.mark.gpu

HAVE_MODELS = os.path.exists(os.path.join(os.environ["TKSM_MODELS"], "badread", "nanopore2020.error.gz"))


def make_fasta(rng, contigs, n_rate=0.01, lower=True):
    out = []
    for name, length in contigs:
        seq = rng.choice(np.frombuffer(b"ACGT", np.uint8), size=length)
        # N runs, a few IUPAC codes, soft-masked stretches
        pos = 0
        while pos < length:
            if rng.random() < n_rate:
                run = int(rng.integers(1, 400))
                seq[pos:pos + run] = ord("N")
                pos += run
            pos += int(rng.integers(50, 3000))
        for p in rng.integers(0, length, size=5):
            seq[p] = rng.choice(np.frombuffer(b"RYKMSW", np.uint8))
        s = seq.tobytes().decode()
        if lower:
            a = int(rng.integers(0, max(1, length - 500)))
            s = s[:a] + s[a:a + 400].lower() + s[a + 400:]
        out.append(f">{name} synthetic contig\n" + "\n".join(s[i:i + 60] for i in range(0, len(s), 60)))
    return "\n".join(out) + "\n"


def first_diff(got, want, fastq=True):
    if got == want:
        return None
    try:
        g, w = so.parse_records(got, fastq), so.parse_records(want, fastq)
    except AssertionError as e:
        return f"unparseable output ({e}); sizes {len(got)} vs {len(want)}"
    if len(g) != len(w):
        return f"record count {len(g)} vs {len(w)}"
    for i, (a, b) in enumerate(zip(g, w)):
        if a != b:
            msg = [f"record {i}:"]
            if a[0] != b[0]:
                msg.append(f" header\n   got  {a[0]}\n   want {b[0]}")
            if a[1] != b[1]:
                j = next((x for x in range(min(len(a[1]), len(b[1]))) if a[1][x] != b[1][x]), min(len(a[1]), len(b[1])))
                msg.append(f" seq differs at {j} (len {len(a[1])} vs {len(b[1])}): "
                           f"got {a[1][max(0, j - 10):j + 10]} want {b[1][max(0, j - 10):j + 10]}")
            if fastq and a[2] != b[2]:
                j = next((x for x in range(min(len(a[2]), len(b[2]))) if a[2][x] != b[2][x]), min(len(a[2]), len(b[2])))
                nd = sum(1 for x, y in zip(a[2], b[2]) if x != y)
                msg.append(f" qual differs at {j} ({nd} positions; len {len(a[2])} vs {len(b[2])}): "
                           f"got {a[2][max(0, j - 10):j + 10]} want {b[2][max(0, j - 10):j + 10]}")
            return "".join(msg)
    return "records equal but bytes differ"


@pytest.fixture(scope="module")
def world(built):
    from tksm_b200.sequence import Sequencer
    rng = np.random.default_rng(2024)
    fasta = make_fasta(rng, [("chr1", 300000), ("chr2", 120000), ("scaffold_3", 30000)])
    return {"fasta": fasta, "refs": so.load_fasta(fasta), "Sequencer": Sequencer}


def oracle_inputs(world, mdf):
    mols = so.parse_mdf(mdf)
    seqs = [so.molecule_sequence(world["refs"], iv) for _, iv in mols]
    return seqs, [m for m, _ in mols]


def handmade_mdf():
    """Edge cases of the deterministic stage (SURVEY.md §4 vectors and §8a footnotes)."""
    return (
        "+m1\t1\tinfo=x;\n"
        "chr1\t10\t14\t+\t2C,3T\n"
        "chr1\t16\t19\t+\t1G\n"
        "+lit\t1\t\n"
        "TT\t0\t2\t+\t\n"
        "chr1\t0\t4\t+\t\n"
        "+minus\t2\tdepth2;\n"
        "chr1\t100\t160\t-\t\n"
        "+clamp\t1\t\n"
        "chr2\t119990\t999999\t+\t\n"
        "+lower_lit\t1\t\n"
        "acgtn\t0\t5\t-\t0r\n"
        "+nomods4\t1\t\n"
        "chr2\t5\t45\t-\n"
        "+multi\t3\tCB=ACGT;tid=T1;\n"
        "AATGTACTTCGTTCAGTTACGTATTGCT\t0\t28\t+\t\n"
        "chr2\t5000\t5400\t-\t5A,17c,399T\n"
        "chr2\t3000\t3200\t-\t\n"
        "AAAAAAAAAAAAAAAAAAA\t0\t19\t+\t\n"
        "+zero_depth\t0\t\n"
        "chr1\t0\t10\t+\t\n"
        "+single\t1\t\n"
        "chr1\t77\t78\t+\t\n"
        "+empty_slice\t1\t\n"
        "chr1\t50\t50\t+\t\n"
        "chr1\t200\t260\t+\t\n"
        "+negative\t1\t\n"
        "chr1\t-20\t-5\t+\t-1G\n"
    )


def test_perfect_handmade_matches_python_oracle(world):
    s = world["Sequencer"](seed=5)
    s.add_reference_text(world["fasta"])
    s.load_models("random", "random")
    from tksm_b200.sequence import PERFECT
    mdf = handmade_mdf()
    seqs, ids = oracle_inputs(world, mdf)
    for fastq in (False, True):
        got = so.parse_records(s.simulate(mdf, PERFECT, fastq=fastq), fastq)
        assert len(got) == len(seqs)
        for (hdr, seq, qual), want, mid in zip(got, seqs, ids):
            uuid = hdr.split(" ")[0][1:]
            assert len(uuid) == 36
            assert (hdr + "\n").encode() + seq + b"\n" + (b"+\n" + qual + b"\n" if fastq else b"") == \
                so.perfect_record(uuid, mid, want, fastq)


@pytest.mark.parametrize("shape,n", [(0, 400), (1, 1500), (2, 60), (3, 600)])
def test_perfect_synthetic_matches_python_oracle(world, shape, n):
    from tksm_b200.sequence import PERFECT
    s = world["Sequencer"](seed=9)
    s.add_reference_text(world["fasta"])
    s.load_models("random", "random")
    mdf = s.synth_mdf(shape, n, seed=shape + 1, min_len=200, max_len=60000)
    seqs, ids = oracle_inputs(world, mdf)
    got = s.simulate(mdf, PERFECT, fastq=False)
    want = so.twin_simulate(s.tables(), seqs, ids, badread=False, fastq=False)
    assert first_diff(got, want, False) is None


CASES = [
    # error model, qscore model, identity, skip_qual, shape, n
    ("random", "random", (84.0, 99.0, 5.5), False, 1, 300),
    ("random", "ideal", (90.0, 98.0, 3.0), False, 1, 200),
    ("random", "random", (84.0, 99.0, 5.5), False, 3, 300),     # BASELINE configs[2] shape: scRNA Tsb->SCB->PCR
]
if HAVE_MODELS:
    CASES += [
        ("nanopore2020", "nanopore2020", (84.0, 99.0, 5.5), False, 1, 1200),
        ("nanopore2020", "nanopore2020", (84.0, 99.0, 5.5), True, 0, 300),
        ("nanopore2018", "nanopore2018", (87.0, 97.0, 4.0), False, 1, 300),
        ("pacbio2016", "pacbio2016", (99.0, 99.9, 0.5), False, 1, 300),
        ("nanopore2020", "nanopore2020", (95.0, 95.0, 0.0), False, 0, 200),
        ("nanopore2020", "nanopore2020", (60.0, 80.0, 8.0), False, 1, 200),
        ("nanopore2020", "nanopore2020", (84.0, 99.0, 5.5), False, 2, 80),
        ("nanopore2020", "nanopore2020", (84.0, 99.0, 5.5), False, 3, 500),
        ("pacbio2016", "pacbio2016", (99.0, 99.9, 0.5), False, 3, 300),   # configs[3]: high identity, qscore-dominated
    ]


@pytest.mark.parametrize("em,qm,ident,skip,shape,n", CASES)
def test_badread_matches_cpu_twin(world, em, qm, ident, skip, shape, n):
    from tksm_b200.sequence import BADREAD
    s = world["Sequencer"](identity=ident, seed=1234, skip_qual_compute=skip)
    s.add_reference_text(world["fasta"])
    s.load_models(em, qm)
    mdf = s.synth_mdf(shape, n, seed=17 + shape, min_len=300, max_len=70000)
    if shape == 1:
        mdf += handmade_mdf().encode()
    seqs, ids = oracle_inputs(world, mdf)
    for fastq in (True, False):
        got = s.simulate(mdf, BADREAD, fastq=fastq, ordinal_base=1000)
        want = so.twin_simulate(s.tables(), seqs, ids, ordinal_base=1000, badread=True, fastq=fastq)
        d = first_diff(got, want, fastq)
        assert d is None, d
    assert s.last_stats["status_or"] & 3 == 0


def test_device_batch_rerun_is_deterministic_and_ordinal_keyed(world):
    from tksm_b200.sequence import BADREAD, FASTQ
    s = world["Sequencer"](seed=77)
    s.add_reference_text(world["fasta"])
    s.load_models("random", "random")
    mdf = s.synth_mdf(1, 500, seed=5)
    b = s.batch(mdf)
    st1 = b.run(BADREAD, FASTQ)
    out1 = b.fetch(BADREAD)
    st2 = b.run(BADREAD, FASTQ)
    out2 = b.fetch(BADREAD)
    assert out1 == out2 and st1["out_bytes"] == st2["out_bytes"]
    # splitting the stream in two batches gives the same bytes: results depend on the read
    # ordinal only, not on batch or GPU boundaries
    recs = mdf.split(b"\n+M")          # record boundaries (a strand field is never at line start)
    half = len(recs) // 2
    a = b"\n+M".join(recs[:half]) + b"\n"
    c = b"+M" + b"\n+M".join(recs[half:])
    na = len(so.parse_mdf(a))
    out_split = s.simulate(a, BADREAD, True, 0) + s.simulate(c, BADREAD, True, na)
    assert out_split == out1


def test_streaming_pipeline_is_chunking_invariant(world, tmp_path, monkeypatch):
    """The pipelined path (host parser threads -> upload -> kernels -> download, three device slots)
    must produce the bytes of a single device batch whatever the text chunk size, the device chunk
    size (batch_reads) or the sink (host buffer, plain file, .gz file)."""
    import gzip
    from tksm_b200.sequence import BADREAD, FASTQ, PERFECT
    s = world["Sequencer"](seed=99)
    s.add_reference_text(world["fasta"])
    s.load_models("random", "random")
    mdf = s.synth_mdf(1, 3000, seed=21) + handmade_mdf().encode()
    b = s.batch(mdf)
    b.run(BADREAD | PERFECT, FASTQ)
    want, want_perf = b.fetch(BADREAD), b.fetch(PERFECT)
    assert s.simulate(mdf, BADREAD, True) == want
    for chunk_bytes, batch_reads in ((4096, 0), (50_000, 37), (1, 0)):
        monkeypatch.setenv("TKSM_SEQ_CHUNK_BYTES", str(chunk_bytes))
        s2 = world["Sequencer"](seed=99, batch_reads=batch_reads)
        s2.add_reference_text(world["fasta"])
        s2.load_models("random", "random")
        assert s2.simulate(mdf, BADREAD, True) == want
        assert s2.simulate(mdf, PERFECT, True) == want_perf
        with s2.simulate_view(mdf, BADREAD, True) as rb:
            assert rb.tobytes() == want
        assert s2.last_stats["reads"] == len(so.parse_mdf(mdf))
        src = tmp_path / "in.mdf"
        src.write_bytes(mdf)
        s2.run_files(str(src), badread=str(tmp_path / "o.fastq"), perfect=str(tmp_path / "p.fq.gz"))
        assert (tmp_path / "o.fastq").read_bytes() == want
        assert gzip.open(tmp_path / "p.fq.gz").read() == want_perf
    monkeypatch.delenv("TKSM_SEQ_CHUNK_BYTES")
    # an input error surfaces as an error of the call, after which the handle keeps working
    with pytest.raises(Exception):
        s.simulate(mdf + b"chr1\t0\t5\n", BADREAD, True)
    assert s.simulate(mdf, BADREAD, True) == want


def test_shard_plan_through_the_kernels(world):
    """every rank simulates its shard with the global ordinal of its first read; the
    rank outputs concatenate to the single-GPU bytes (here the ranks run one after the other)."""
    from tksm_b200.sequence import BADREAD
    s = world["Sequencer"](seed=31)
    s.add_reference_text(world["fasta"])
    s.load_models("random", "random")
    mdf = b"+d3\t3\tx;\nchr1\t100\t400\t+\t\n" + s.synth_mdf(1, 700, seed=8) + handmade_mdf().encode()
    whole = s.simulate(mdf, BADREAD, True)
    for n_ranks in (2, 3, 8):
        parts = [s.simulate(mdf[b0:b1], BADREAD, True, ordinal_base=ob) for b0, b1, ob, _ in s.shard_plan(mdf, n_ranks)]
        assert b"".join(parts) == whole


def test_cli_front_end_files_and_fifo(world, tmp_path):
    """`tksm sequence` (src/tksm.cpp:164-166 -> Sequencer_module -> tksm_sequence_main): the command
    rule `sequence` of the Snakefile runs (Snakefile:283-305), on files and on a FIFO input
    (enable_piping, Snakefile:247-280), against the library call on the same text."""
    import subprocess
    import threading
    from tksm_b200.sequence import BADREAD, PERFECT
    exe = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tksm_b200", "tksm")
    s = world["Sequencer"](seed=7)
    s.add_reference_text(world["fasta"])
    s.load_models("random", "random")
    mdf = s.synth_mdf(1, 400, seed=4) + handmade_mdf().encode()
    want = s.simulate(mdf, BADREAD, True)
    want_perf = s.simulate(mdf, PERFECT, False)
    fa, src = tmp_path / "g.fa", tmp_path / "in.mdf"
    fa.write_text(world["fasta"])
    src.write_bytes(mdf)
    common = ["-r", str(fa), "--badread-error-model", "random", "--badread-qscore-model", "random", "-s", "7", "-t", "3"]
    out, perf = tmp_path / "o.fastq", tmp_path / "p.fasta"
    p = subprocess.run([exe, "sequence", "-i", str(src), "-o", str(out), "--perfect", str(perf)] + common,
                       capture_output=True, text=True, timeout=120)
    assert p.returncode == 0, p.stderr
    assert out.read_bytes() == want and perf.read_bytes() == want_perf
    fifo = tmp_path / "in.fifo"
    os.mkfifo(fifo)

    def feed():
        with open(fifo, "wb") as f:
            for i in range(0, len(mdf), 7919):   # odd-sized writes: records straddle them
                f.write(mdf[i:i + 7919])
    th = threading.Thread(target=feed)
    th.start()
    out2 = tmp_path / "o2.fq"
    p = subprocess.run([exe, "sequence", "-i", str(fifo), "-o", str(out2)] + common, capture_output=True, text=True, timeout=120)
    th.join()
    assert p.returncode == 0, p.stderr
    assert out2.read_bytes() == want
    # a broken record makes the module exit 1 like the reference's uncaught exception does
    bad = tmp_path / "bad.mdf"
    bad.write_bytes(mdf + b"chr1\t0\n")
    p = subprocess.run([exe, "sequence", "-i", str(bad), "-o", str(tmp_path / "o3.fq")] + common, capture_output=True, text=True, timeout=120)
    assert p.returncode == 1 and "not enough values" in p.stderr


def test_fused_and_split_kernels_agree(world, monkeypatch):
    """The simulate stage runs as two kernels per length class by default (error injection | expansion +
    qscores, state handed over through global memory); TKSM_SIM_SPLIT=0 runs the fused kernel."""
    from tksm_b200.sequence import BADREAD, PERFECT
    s = world["Sequencer"](seed=5)
    s.add_reference_text(world["fasta"])
    s.load_models("nanopore2020" if HAVE_MODELS else "random", "nanopore2020" if HAVE_MODELS else "random")
    mdf = s.synth_mdf(1, 600, seed=12) + s.synth_mdf(2, 40, seed=13, min_len=300, max_len=60000) + handmade_mdf().encode()
    out = {}
    for mode in ("1", "0"):
        monkeypatch.setenv("TKSM_SIM_SPLIT", mode)
        out[mode] = (s.simulate(mdf, BADREAD, True), s.simulate(mdf, PERFECT, False), s.simulate(mdf, BADREAD, False))
    assert out["0"] == out["1"]


@pytest.mark.gpu
def test_cli_gpus_flag_shards_in_one_process(world, tmp_path):
    """`tksm sequence --gpus N`: one process, one handle and one thread per shard of the plan,
    files written in shard order hold the bytes of the one-GPU run.  On a one-GPU box $TKSM_GPU_LIST places every
    shard on device 0; with more devices visible the default placement (device i for shard i) runs as well."""
    import subprocess
    import torch
    from tksm_b200.sequence import BADREAD, PERFECT
    exe = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tksm_b200", "tksm")
    s = world["Sequencer"](seed=11)
    s.add_reference_text(world["fasta"])
    s.load_models("random", "random")
    mdf = s.synth_mdf(1, 900, seed=6) + handmade_mdf().encode()
    want, want_perf = s.simulate(mdf, BADREAD, True), s.simulate(mdf, PERFECT, False)
    fa, src = tmp_path / "g.fa", tmp_path / "in.mdf"
    fa.write_text(world["fasta"])
    src.write_bytes(mdf)
    common = ["-r", str(fa), "--badread-error-model", "random", "--badread-qscore-model", "random", "-s", "11", "-t", "4"]
    runs = [(3, {"TKSM_GPU_LIST": "0,0,0"}), (2, {"TKSM_GPU_LIST": "0"})]
    if torch.cuda.device_count() >= 2:
        runs.append((2, {}))
    for n, env in runs:
        out, perf = tmp_path / f"o{n}.fastq", tmp_path / f"p{n}.fasta"
        p = subprocess.run([exe, "sequence", "-i", str(src), "-o", str(out), "--perfect", str(perf), "--gpus", str(n)] + common,
                           capture_output=True, text=True, timeout=180, env={**os.environ, **env})
        assert p.returncode == 0, p.stderr
        assert out.read_bytes() == want and perf.read_bytes() == want_perf, (n, env)
    # more shards than records, and an empty input
    tiny = tmp_path / "tiny.mdf"
    tiny.write_bytes(b"+a\t1\t\nchr1\t10\t300\t+\t\n")
    out = tmp_path / "tiny.fastq"
    p = subprocess.run([exe, "sequence", "-i", str(tiny), "-o", str(out), "--gpus", "4"] + common, capture_output=True,
                       text=True, timeout=180, env={**os.environ, "TKSM_GPU_LIST": "0"})
    assert p.returncode == 0, p.stderr
    assert out.read_bytes() == s.simulate(tiny.read_bytes(), BADREAD, True)
    empty = tmp_path / "empty.mdf"
    empty.write_bytes(b"")
    p = subprocess.run([exe, "sequence", "-i", str(empty), "-o", str(out), "--gpus", "2"] + common, capture_output=True,
                       text=True, timeout=180, env={**os.environ, "TKSM_GPU_LIST": "0"})
    assert p.returncode == 0 and out.read_bytes() == b""
    # a broken record in the second shard fails the run like the one-GPU run does
    bad = tmp_path / "bad.mdf"
    bad.write_bytes(mdf + b"chr1\t0\n")
    p = subprocess.run([exe, "sequence", "-i", str(bad), "-o", str(out), "--gpus", "2"] + common, capture_output=True,
                       text=True, timeout=180, env={**os.environ, "TKSM_GPU_LIST": "0"})
    assert p.returncode == 1 and "Error" in p.stderr
    # a stream that cannot be read twice is dealt to the devices piece by piece, not slurped: same bytes
    import threading
    for deal_bytes in ("7000", "100000000"):
        fifo = tmp_path / f"in{deal_bytes}.fifo"
        os.mkfifo(fifo)
        feeder = threading.Thread(target=lambda: open(fifo, "wb").write(mdf))
        feeder.start()
        out, perf = tmp_path / "of.fastq", tmp_path / "pf.fasta"
        p = subprocess.run([exe, "sequence", "-i", str(fifo), "-o", str(out), "--perfect", str(perf), "--gpus", "3"] + common,
                           capture_output=True, text=True, timeout=180,
                           env={**os.environ, "TKSM_GPU_LIST": "0,0,0", "TKSM_SEQ_DEAL_BYTES": deal_bytes})
        feeder.join()
        assert p.returncode == 0, p.stderr
        assert out.read_bytes() == want and perf.read_bytes() == want_perf, deal_bytes
    p = subprocess.run([exe, "sequence", "-i", str(bad), "-o", str(out), "--gpus", "2"] + common, capture_output=True, text=True,
                       timeout=180, env={**os.environ, "TKSM_GPU_LIST": "0", "TKSM_SEQ_DEAL": "1", "TKSM_SEQ_DEAL_BYTES": "9000"})
    assert p.returncode == 1 and "Error" in p.stderr
    # a regular file above the size a shard plan is made for is dealt as well
    p = subprocess.run([exe, "sequence", "-i", str(src), "-o", str(out), "--perfect", str(perf), "--gpus", "2"] + common,
                       capture_output=True, text=True, timeout=180,
                       env={**os.environ, "TKSM_GPU_LIST": "0,0", "TKSM_SEQ_PLAN_MAX_BYTES": "1000", "TKSM_SEQ_DEAL_BYTES": "30000",
                            "TKSM_SEQ_TRACE": "1"})
    assert p.returncode == 0, p.stderr
    assert "[trace] dealt" in p.stderr and "dealt 1 pieces" not in p.stderr
    assert out.read_bytes() == want and perf.read_bytes() == want_perf


@pytest.mark.gpu
def test_gz_sink_blocks(world, tmp_path, monkeypatch):
    """`.gz` outputs (get_output_file, py/sequence.py:291-296) are written as independently deflated blocks, one
    gzip member each, by helper threads: Python's gzip and zcat read them back as the plain bytes, whatever the
    block size; an empty MDF gives a valid empty .gz."""
    import gzip
    import subprocess
    from tksm_b200.sequence import BADREAD, PERFECT
    s = world["Sequencer"](seed=3)
    s.add_reference_text(world["fasta"])
    s.load_models("random", "random")
    mdf = s.synth_mdf(1, 300, seed=8)
    want, want_perf = s.simulate(mdf, BADREAD, True), s.simulate(mdf, PERFECT, False)
    src = tmp_path / "in.mdf"
    src.write_bytes(mdf)
    for block in ("777", "65536", None):
        if block:
            monkeypatch.setenv("TKSM_GZ_BLOCK_BYTES", block)
        else:
            monkeypatch.delenv("TKSM_GZ_BLOCK_BYTES")
        o, p = tmp_path / f"o{block}.fastq.gz", tmp_path / f"p{block}.fa.gz"
        s.run_files(str(src), badread=str(o), perfect=str(p))
        assert gzip.open(o).read() == want and gzip.open(p).read() == want_perf
        assert subprocess.run(["zcat", str(o)], capture_output=True, timeout=60).stdout == want
        if block == "777":
            assert o.read_bytes().count(b"\x1f\x8b\x08") >= len(want) // 777   # one member per block
    empty = tmp_path / "empty.mdf"
    empty.write_bytes(b"")
    s.run_files(str(empty), badread=str(tmp_path / "e.fq.gz"))
    assert (tmp_path / "e.fq.gz").stat().st_size > 0 and gzip.open(tmp_path / "e.fq.gz").read() == b""


# ---------------------------------------------------------------- BASELINE configs[0] at its size
def _twin_span(args):
    """One span of the molecules through the CPU twin (worker process)."""
    tables_kw, fasta, mdf_part, lo, fastq = args
    from tksm_b200.sequence import Sequencer
    s = Sequencer(host_only=True, **tables_kw)
    s.add_reference_text(fasta)
    s.load_models("nanopore2020", "nanopore2020")
    refs = so.load_fasta(fasta)
    mols = so.parse_mdf(mdf_part)
    seqs = [so.molecule_sequence(refs, iv) for _, iv in mols]
    return so.twin_simulate(s.tables(), seqs, [m for m, _ in mols], ordinal_base=lo, badread=True, fastq=fastq)


@pytest.mark.skipif(not HAVE_MODELS, reason="Badread model files not present")
def test_c1_sized_run_matches_cpu_twin(built):
    """BASELINE configs[0]: 10 k Tsb-shape molecules (multi-exon, both strands, depth > 1) from a 1 Mbp genome,
    nanopore2020 models, FASTQ with qscores: the CUDA path against the CPU twin, byte for byte.  The twin runs in
    worker processes over spans of the molecules (results depend on the global read ordinal only)."""
    import multiprocessing as mp
    from tksm_b200.sequence import BADREAD, Sequencer
    rng = np.random.default_rng(31)
    fasta = make_fasta(rng, [("chr1", 600000), ("chr2", 400000)])
    s = Sequencer(seed=4321)
    s.add_reference_text(fasta)
    s.load_models("nanopore2020", "nanopore2020")
    mdf = s.synth_mdf(0, 10000, seed=3, min_len=300, max_len=20000)
    got = s.simulate(mdf, BADREAD, fastq=True)
    assert s.last_stats["reads"] >= 10000 and s.last_stats["status_or"] & 3 == 0
    # the twin needs the effective kappa the device measured
    kw = dict(seed=4321, corr=(float(s.tables().calib_overshoot), float(s.tables().calib_kappa)))
    text = mdf.decode()
    recs = text.split("\n+")
    recs = [recs[0]] + ["+" + r for r in recs[1:]]
    depth = [int(r.split("\n", 1)[0].split("\t")[1]) for r in recs]
    procs = min(os.cpu_count() or 1, 16)
    per = (len(recs) + procs - 1) // procs
    jobs, lo = [], 0
    for i in range(0, len(recs), per):
        part = "\n".join(x.rstrip("\n") for x in recs[i:i + per]) + "\n"
        jobs.append((kw, fasta, part, lo, True))
        lo += sum(depth[i:i + per])
    with mp.get_context("fork").Pool(len(jobs)) as pool:
        want = b"".join(pool.map(_twin_span, jobs))
    d = first_diff(got, want, True)
    assert d is None, d


@pytest.mark.skipif(not HAVE_MODELS, reason="Badread model files not present")
def test_model_given_by_path_under_another_name(built, tmp_path):
    """--badread-error-model / --badread-qscore-model take a name or a path (py/tksm_badread.py:84-89).  Nothing may
    hang on the file name: a copy under another name gives the same measured calibration and the same reads."""
    import shutil
    from tksm_b200.sequence import BADREAD, Sequencer
    src = os.path.join(os.environ["TKSM_MODELS"], "badread")
    em, qm = str(tmp_path / "my_flowcell.error.gz"), str(tmp_path / "my_flowcell.qscore.gz")
    shutil.copy(os.path.join(src, "nanopore2018.error.gz"), em)
    shutil.copy(os.path.join(src, "nanopore2018.qscore.gz"), qm)
    rng = np.random.default_rng(5)
    fasta = make_fasta(rng, [("chr1", 80000)])
    outs, kappas = [], []
    for e, q in (("nanopore2018", "nanopore2018"), (em, qm)):
        s = Sequencer(seed=99)
        s.add_reference_text(fasta)
        s.load_models(e, q)
        kappas.append((float(s.tables().calib_kappa_measured), float(s.tables().calib_kappa)))
        outs.append(s.simulate(s.synth_mdf(1, 300, seed=8), BADREAD, fastq=True))
    assert kappas[0] == kappas[1] and 0.2 < kappas[0][0] < 0.9
    assert outs[0] == outs[1]
    # and the device's measurement is the twin's (oracle/fit_calibration.py -> tests/golden/calibration.json)
    import json
    cal = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "calibration.json")))
    assert kappas[0][0] == pytest.approx(cal["nanopore2018"], abs=2e-3)


def test_insertion_heavy_reads_rerun_in_larger_slots(built, tmp_path):
    """A read that outgrows twice its fragment does not fit the slot of the first pass (ST_OPS_OVERFLOW /
    ST_SCRATCH_OVERFLOW): the chunk runs again in slots no edit script can outgrow (engine.cu: slot_launch(grow))
    and comes out as the CPU twin, which knows no slots, writes it.  The model here turns every 3-mer into a 6-mer."""
    import itertools
    from tksm_b200.sequence import BADREAD, Sequencer
    em = tmp_path / "insertions.error"
    with open(em, "w") as f:
        for km in itertools.product("ACGT", repeat=3):
            km = "".join(km)
            ins = "".join("ACGT"[("ACGT".index(km[1]) + 1 + i) % 4] for i in range(3))
            f.write(f"{km},0.02;{km[0]}{km[1]}{ins}{km[2]},0.9;\n")
    rng = np.random.default_rng(11)
    fasta = make_fasta(rng, [("chr1", 120000)], n_rate=0.0, lower=False)
    s = Sequencer(identity=(40.0, 40.0, 0.0), seed=5)
    s.add_reference_text(fasta)
    s.load_models(str(em), "random")
    mdf = s.synth_mdf(1, 60, seed=3, min_len=40, max_len=12000)
    refs = so.load_fasta(fasta)
    mols = so.parse_mdf(mdf)
    seqs = [so.molecule_sequence(refs, iv) for _, iv in mols]
    ids = [m for m, _ in mols]
    for fastq in (True, False):
        got = s.simulate(mdf, BADREAD, fastq=fastq)
        assert s.last_stats["status_or"] & 3 == 0
        assert s.last_stats["read_bases"] > 2.2 * s.last_stats["raw_bases"]     # the first pass cannot have held them
        want = so.twin_simulate(s.tables(), seqs, ids, badread=True, fastq=fastq)
        d = first_diff(got, want, fastq)
        assert d is None, d
    # a device-resident batch takes the same second pass
    from tksm_b200.sequence import FASTQ
    b = s.batch(mdf)
    st = b.run(BADREAD, FASTQ)
    assert st["status_or"] & 3 == 0 and b.fetch(BADREAD) == s.simulate(mdf, BADREAD, fastq=True)


@pytest.mark.parametrize("models", [("random", "random")] + ([("nanopore2020", "nanopore2020")] if HAVE_MODELS else []))
def test_device_mdf_tokeniser_matches_the_host_parser(world, monkeypatch, models):
    """with TKSM_SEQ_DEVICE_PARSE=1 the streaming pipeline uploads the MDF text and builds the molecule
    batch on the device (csrc/mdf_device.cuh).  The records must be the bytes the host parser's batch gives, for every
    MDF shape, whatever the chunking; text outside the emitter's format (negative indices, CR line ends) and every
    input error go through the host parser and behave as before."""
    from tksm_b200.sequence import BADREAD, FASTQ, PERFECT
    s = world["Sequencer"](seed=31)
    s.add_reference_text(world["fasta"])
    s.load_models(*models)
    clean = b"".join(s.synth_mdf(shape, n, seed=40 + shape, min_len=300, max_len=30000)
                     for shape, n in ((0, 400), (1, 1500), (2, 60), (3, 700)))
    clean += b"+zero_depth\t0\t\nchr1\t0\t10\t+\t\n+lit\t2\tx=1;\nTTacgN\t1\t5\t-\t2g\nchr2\t119990\t999999\t+\t\n+four\t1\t\nchr2\t5\t45\t-\n"
    monkeypatch.delenv("TKSM_SEQ_DEVICE_PARSE", raising=False)
    b = s.batch(clean)
    b.run(BADREAD | PERFECT, FASTQ)
    want, want_perf = b.fetch(BADREAD), b.fetch(PERFECT)
    host_h2d = None
    for chunk_bytes in (0, 30000):
        if chunk_bytes:
            monkeypatch.setenv("TKSM_SEQ_CHUNK_BYTES", str(chunk_bytes))
        monkeypatch.delenv("TKSM_SEQ_DEVICE_PARSE", raising=False)
        assert s.simulate(clean, BADREAD, True) == want
        host_h2d = s.last_stats["h2d_bytes"]
        monkeypatch.setenv("TKSM_SEQ_DEVICE_PARSE", "1")
        assert s.simulate(clean, BADREAD, True) == want
        assert s.last_stats["h2d_bytes"] == len(clean) != host_h2d          # the text went up, not a host-built blob
        assert s.last_stats["reads"] == len(so.parse_mdf(clean))
        assert s.simulate(clean, PERFECT, True) == want_perf
    monkeypatch.delenv("TKSM_SEQ_CHUNK_BYTES")
    # the handmade records hold negative indices and modification positions: that batch takes the host parser
    odd = s.synth_mdf(1, 200, seed=3) + handmade_mdf().encode()
    monkeypatch.delenv("TKSM_SEQ_DEVICE_PARSE")
    want_odd = s.simulate(odd, BADREAD, True)
    monkeypatch.setenv("TKSM_SEQ_DEVICE_PARSE", "1")
    assert s.simulate(odd, BADREAD, True) == want_odd
    assert s.last_stats["h2d_bytes"] != len(odd)
    crlf = clean.replace(b"\n", b"\r\n")
    assert s.simulate(crlf, BADREAD, True) == want
    # errors surface as they do from the host parser, and the handle keeps working
    for bad in (b"chr1\t0\t5\n", b"chr1\t0\tx\t+\t\n", b"chr1\t0\t5\t+\t9A\n", b"\n", b"\t0\t5\t+\t\n"):
        with pytest.raises(Exception):
            s.simulate(clean + bad, BADREAD, True)
    assert s.simulate(clean, BADREAD, True) == want
