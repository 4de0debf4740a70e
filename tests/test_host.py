"""CPU tests of the host side of libtksm_seq.so: the C ABI surface, loud failure without a device,
model packing (csrc/model.cpp) against the oracle restatement, FASTA store, MDF parser error
behaviour and the `tksm sequence` command line.  No compute call is made (no GPU here)."""
import ctypes as C
import os
import re
import subprocess
import sys

import numpy as np
import pytest

from oracle import model_oracle as mo
from oracle import seq_oracle as so

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
MODELS = os.path.join(os.environ["TKSM_MODELS"], "badread")
HAVE_MODELS = os.path.exists(os.path.join(MODELS, "nanopore2020.error.gz"))
needs_models = pytest.mark.skipif(not HAVE_MODELS, reason="model files not staged by build()")


@pytest.fixture(scope="module")
def Sequencer(built):
    from tksm_b200.sequence import Sequencer
    return Sequencer


def test_c_abi_exports_every_declared_symbol(built):
    from tksm_b200 import _lib
    hdr = open(os.path.join(ROOT, "include", "tksm_seq.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(tksm_seq(?:uence)?_[a-z_]+)\s*\(", hdr))
    assert len(declared) >= 20
    assert declared == set(_lib.EXPORTS)
    lib = C.CDLL(_lib.LIB_PATH)
    for name in declared:
        assert hasattr(lib, name), name
    # the reference-shaped C++ class is exported too (src/tksm.cpp:164-166 links against it)
    syms = subprocess.run(["nm", "-DC", _lib.LIB_PATH], capture_output=True, text=True).stdout
    assert "Sequencer_module::run()" in syms and "Sequencer_module::Sequencer_module(int, char**)" in syms


def test_product_never_imports_the_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, "tksm_b200")):
        for f in files:
            if f.endswith((".py", ".cpp", ".cu", ".cuh", ".hpp", ".h")):
                text = open(os.path.join(dirpath, f), errors="replace").read()
                assert not re.search(r"^\s*(from|import)\s+oracle|#include\s+\"[^\"]*oracle/|dlopen\(.*oracle", text, re.M), f


def test_create_fails_loudly_without_a_device(Sequencer):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a device is present")
    from tksm_b200.sequence import SeqError
    with pytest.raises(SeqError, match="no CPU fallback"):
        Sequencer()
    s = Sequencer(host_only=True)
    s.add_reference_text(">c\nACGT\n")
    s.load_models("random", "random")
    with pytest.raises(SeqError, match="host_only"):
        s.simulate("+m\t1\t\nc\t0\t4\t+\t\n")


def test_genome_store_keeps_every_byte(Sequencer):
    fasta = ">chr1 some description\nACGTNNNNacgtnRYKM\nacgtACGT\r\n>chr2\nTTTT\n>chr1x\nGG\n"
    s = Sequencer(host_only=True)
    s.add_reference_text(fasta)
    refs = so.load_fasta(fasta)
    for name, seq in refs.items():
        assert s.fetch_reference(name, 0, len(seq)) == seq.upper().encode("latin-1"), name
    from tksm_b200.sequence import SeqError
    with pytest.raises(SeqError):
        s.add_reference_text(">a\nAC\n\nGT\n")     # IndexError in the reference (py/sequence.py:177)


def test_large_reference_packed_by_several_threads(Sequencer):
    """Above 8 MiB of text the contigs are packed side by side by host threads (no two contigs share a word).  Contig
    lengths around multiples of 64 put a contig's last word next to its neighbour's first; a line that raises in the
    middle leaves the contigs before it and none after it."""
    from tksm_b200.sequence import SeqError
    rng = np.random.default_rng(21)
    alphabet = np.frombuffer(b"ACGTacgtNn", np.uint8)
    probs = np.array([0.22] * 4 + [0.02] * 4 + [0.03, 0.01])
    probs = probs / probs.sum()
    lengths = [64, 128, 6400, 63, 65, 1, 16, 640000, 1_000_000, 64000] * 6
    parts = []
    for i, n in enumerate(lengths):
        sq = rng.choice(alphabet, size=n, p=probs).tobytes().decode()
        if n >= 6400:
            a = int(rng.integers(0, n - 3000))
            sq = sq[:a] + "N" * 2500 + sq[a + 2500:]
        parts.append(f">s{i} x\n" + "\n".join(sq[j:j + 64] for j in range(0, n, 64)) + "\n")
    fasta = "".join(parts)
    assert len(fasta) > (8 << 20)
    refs = so.load_fasta(fasta)
    s = Sequencer(host_only=True)
    s.add_reference_text(fasta)
    for name, sq in refs.items():
        assert s.fetch_reference(name, 0, len(sq)) == sq.upper().encode(), name
    cut = fasta.index(">s37 ")                                  # 640 000 bases
    mid = fasta.index("\n", cut) + 1
    broken = fasta[:mid + 650] + "\n" + fasta[mid + 650:]       # an empty line after ten lines of s37
    s2 = Sequencer(host_only=True)
    with pytest.raises(SeqError, match="empty line"):
        s2.add_reference_text(broken)
    assert s2.fetch_reference("s36", 0, len(refs["s36"])) == refs["s36"].upper().encode()
    for gone in ("s37", "s38", "s59"):
        with pytest.raises(SeqError):
            s2.fetch_reference(gone, 0, 1)
    s2.add_reference_text(">again\nACGTNACGT\n")               # the handle keeps working
    assert s2.fetch_reference("again", 0, 9) == b"ACGTNACGT" and s2.fetch_reference("s7", 0, 16) == refs["s7"][:16].upper().encode()


def test_reference_files_plain_gz_and_later_files_override(Sequencer, tmp_path):
    """py/sequence.py:188-194: every -r file is read (gzip by its magic bytes, py/tksm_badread.py:259-290) and a name
    given by a later file replaces the earlier one (dict.update).  A plain file is packed from its mapping, a .gz one
    from memory; an empty file gives the reference's '' contig and a missing one an error."""
    import gzip
    from tksm_b200.sequence import SeqError
    rng = np.random.default_rng(9)
    body = lambda n: "".join(rng.choice(list("ACGTNacgt"), size=n))
    a = f">chr1 first\n{body(5000)}\n>chr2\n{body(777)}\n"
    b = f">chr2\n{body(100)}\n>chr3\n{body(64)}\n"
    (tmp_path / "a.fa").write_text(a)
    with gzip.open(tmp_path / "b.fa.gz", "wt") as f:
        f.write(b)
    (tmp_path / "c.fasta").write_text(b)          # the same records, plain
    (tmp_path / "empty.fa").write_text("")
    refs = so.load_fasta(a)
    so.load_fasta(b, refs)
    for second in ("b.fa.gz", "c.fasta"):
        s = Sequencer(host_only=True)
        s.load_references([str(tmp_path / "a.fa"), str(tmp_path / second), str(tmp_path / "empty.fa")])
        for name, sq in refs.items():
            assert s.fetch_reference(name, 0, len(sq)) == sq.upper().encode(), (second, name)
    with pytest.raises(SeqError):
        Sequencer(host_only=True).load_references([str(tmp_path / "missing.fa")])


@pytest.mark.parametrize("width,eol", [(60, "\n"), (70, "\r\n"), (17, "\n"), (61, "\r"), (0, "\n")])
def test_genome_store_packs_lines_where_they_lie(Sequencer, width, eol):
    """csrc/host.cpp: GenomeStore packs sequence lines 16 and 8 bases at a time into the 2-bit words, whatever the line
    width leaves misaligned, and keeps every byte outside ACGT in runs (stretches of one byte 16 at a time).  Contigs
    of 1 to 200 000 bases with N stretches across the 16 / 32 byte boundaries, mixed-case stretches, lower case and
    IUPAC codes, a contig name given twice (the later record wins, py/sequence.py:184): every byte the oracle's
    load_fasta holds comes back, whole contigs and ranges."""
    import random
    rng = random.Random(width * 7 + len(eol))

    def seq(n, pn, plow, piupac):
        out, i = [], 0
        while i < n:
            if rng.random() < pn:
                k = rng.choice([1, 2, 15, 16, 17, 31, 32, 33, 100, 1000])
                out.append("".join(rng.choice("Nn") for _ in range(k)) if rng.random() < 0.2 else rng.choice("NnNNRY-*") * k)
            else:
                k = rng.choice([1, 3, 16, 40, 200, 1000])
                piece = "".join(rng.choice("ACGT" if rng.random() > plow else "acgtACGT") for _ in range(k))
                if rng.random() < piupac:
                    piece = piece[:k // 2] + rng.choice("RYKMSWBDHVU.") + piece[k // 2 + 1:]
                out.append(piece)
            i += k
        return "".join(out)[:n]

    contigs = [("chr1 some words", seq(200000, 0.02, 0.3, 0.05)), ("chr2", seq(70001, 0.05, 0.5, 0.1)), ("c63", seq(63, 0, 0, 0)),
               ("c64", seq(64, 0, 0, 0)), ("c65", seq(65, 0.1, 0.5, 0)), ("c1", seq(1, 0, 0, 0)), ("chr2", seq(999, 0.3, 0.5, 0.2))]
    lines = []
    for name, sq in contigs:
        lines.append(">" + name)
        lines.extend([sq] if width == 0 else [sq[i:i + width] for i in range(0, len(sq), width)])
    fasta = eol.join(lines) + (eol if width != 17 else "")
    refs = so.load_fasta(fasta)
    assert len(refs["chr2"]) == 999
    s = Sequencer(host_only=True)
    s.add_reference_text(fasta)
    for name, sq in refs.items():
        want = sq.upper().encode("latin-1")
        assert s.fetch_reference(name, 0, len(sq)) == want, name
        for _ in range(50):
            a = rng.randrange(len(sq))
            b = min(len(sq), a + rng.choice([1, 15, 16, 17, 64, 1000]))
            assert s.fetch_reference(name, a, b - a) == want[a:b], (name, a, b)
    # a second text on the same handle (a later -r file): earlier contigs stay, names given again are replaced
    s.add_reference_text(">chr2\nACGTN\n>other\nGGGG" + "N" * 40 + "\n")
    assert s.fetch_reference("chr2", 0, 5) == b"ACGTN" and s.fetch_reference("other", 0, 44) == b"GGGG" + b"N" * 40
    assert s.fetch_reference("c65", 0, 65) == refs["c65"].upper().encode("latin-1")


@needs_models
@pytest.mark.parametrize("model", ["nanopore2020", "nanopore2018", "pacbio2016"])
def test_packed_error_table_matches_oracle_distribution(Sequencer, model):
    s = Sequencer(host_only=True)
    s.load_models(model, "random")
    tn = s.tables_numpy()
    assert tn["k"] == 7 and tn["has_err_table"] == 1
    rows = list(mo.iter_error_model(os.path.join(MODELS, f"{model}.error.gz")))
    assert len(rows) == 4 ** 7
    qcls, qrep = tn["qcls"], tn["qrep"]
    class_dev = []
    for kmer, alts in rows:          # every row of the model: the CUDA == twin tests take these tables as given
        pc, dist, prand = mo.trial_distribution(kmer, alts)
        # arrivals: the k-mer's trials arrive at the mean change probability of its class (255 classes
        # of equal width over the 16 384 rows), which is its own to within a class width
        class_dev.append(abs(qrep[qcls[mo.kmer_code(kmer)]] - pc))
        assert class_dev[-1] <= 0.004, kmer
        q, pdist, prandp = mo.decode_error_row(tn, kmer)
        assert q == pytest.approx(pc, abs=2e-5), kmer
        assert prandp == pytest.approx(prand, abs=2e-5), kmer
        assert set(pdist) == set(dist), kmer
        for slots, p in dist.items():
            assert pdist[slots] == pytest.approx(p, abs=2e-5), (kmer, slots)
    assert np.mean(class_dev) < 1.5e-3


@needs_models
@pytest.mark.parametrize("model", ["nanopore2020", "pacbio2016", "ideal", "random"])
def test_packed_qscore_rows_match_oracle(Sequencer, model):
    s = Sequencer(host_only=True)
    s.load_models("random", model)
    tn = s.tables_numpy()
    m = mo.builtin_qscore_model(model) if model in ("ideal", "random") else \
        mo.load_qscore_model(os.path.join(MODELS, f"{model}.qscore.gz"))
    assert tn["kq"] == mo.qscore_kmer_size(m)
    import random
    rng = random.Random(5)
    cigars = list(m)                 # every cigar of the model, then windows the model does not list
    for _ in range(400):
        w = min(rng.choice([1, 3, 5, 7, 9]), tn["kq"])
        ops = [rng.choices("=XI", weights=[0.8, 0.1, 0.1])[0] for _ in range(w)]
        cig = ops[0]
        for o in ops[1:]:
            if rng.random() < 0.15:
                cig += "D" * rng.choice([1, 1, 2, 3, 6])
            cig += o
        cigars.append(cig)
    for cig in cigars:
        if cig[0] == "D" or cig[-1] == "D":
            continue
        scores, probs = m[mo.resolve_cigar(m, cig)]
        want = {q: p for q, p in zip(scores, probs) if p > 0}
        tot = sum(want.values())
        got = mo.decode_qrow(tn, mo.packed_row_for(tn, cig))
        assert set(got) <= set(want), cig
        for q, p in want.items():
            # one 16-bit uniform per read base: alias thresholds carry 16 - lg fraction bits
            assert got.get(q, 0.0) == pytest.approx(p / tot, abs=1e-4), (cig, q)


def test_identity_table_follows_the_beta_quantiles(Sequencer):
    from scipy import stats
    s = Sequencer(identity=(84.0, 99.0, 5.5), host_only=True)
    s.load_models("random", "random")
    tab = s.tables_numpy()["ident_tab"]
    a, b = mo.beta_parameters(84.0, 5.5, 99.0)
    qs = np.linspace(0, 1, len(tab))
    want = 0.99 * stats.beta.ppf(qs[1:-1], a, b)
    assert np.abs(tab[1:-1] - want).max() < 2e-4
    s2 = Sequencer(identity=(90.0, 90.0, 3.0), host_only=True)
    s2.load_models("random", "random")
    t2 = s2.tables_numpy()
    assert t2["ident_const"] == 1 and t2["ident_mean"] == pytest.approx(0.9)


def test_mdf_parser_raises_where_the_reference_raises(Sequencer):
    from tksm_b200.sequence import SeqError
    s = Sequencer(host_only=True)
    s.add_reference_text(">c\nACGTACGTAC\n")
    s.load_models("random", "random")
    good = "+m\t2\tk=v;\nc\t0\t4\t+\t1G\nAAAA\t0\t4\t-\t\n"
    s.batch(good).close()
    for bad in ("+m\t1\t\nc\t0\t4\n",                  # not enough values to unpack
                "+m\t1\t\nc\t0\t4\t+\t\textra\n",       # too many values
                "+m\t1\t\nc\tx\t4\t+\t\n",             # int()
                "+m\tone\t\nc\t0\t4\t+\t\n",           # depth
                "+m\t1\t\nc\t0\t4\t+\t9G\n",           # modification out of range
                "+m\t1\t\n\nc\t0\t4\t+\t\n",           # empty line
                "+m\t1\t\n\t0\t4\t+\t\n"):           # empty first field: line[0][0]
        with pytest.raises(SeqError):
            s.batch(bad)


def run_cli(args, env=None):
    exe = os.path.join(ROOT, "tksm_b200", "tksm")
    e = dict(os.environ)
    e.update(env or {})
    return subprocess.run([exe, "sequence"] + args, capture_output=True, text=True, env=e)


def test_cli_flags_and_exit_codes(built, tmp_path):
    # py/sequence.py:34-165: usage errors exit 2, validation errors exit 1 with the same messages
    p = run_cli(["--list"])
    assert p.returncode == 0
    dests = p.stdout.split()
    for d in ["input", "references", "badread", "perfect", "skip_qual_compute", "output_format", "threads",
              "badread_identity", "badread_error_model", "badread_qscore_model", "badread_tail_model", "list"]:
        assert d in dests
    assert run_cli([]).returncode == 2
    assert run_cli(["-i", "x.mdf", "-r", "g.fa"]).returncode == 2            # neither -o nor --perfect
    assert run_cli(["-i", "x.mdf", "-r", "g.fa", "-o", "o.fq", "--bogus"]).returncode == 2
    p = run_cli(["-i", "x.mdf", "-r", "g.fa", "-o", "o.fq", "--badread-identity", "101,99,5"])
    assert p.returncode == 1 and "mean read identity cannot be more than 100" in p.stderr
    p = run_cli(["-i", "x.mdf", "-r", "g.fa", "-o", "o.fq", "--badread-identity", "90,85,5"])
    assert p.returncode == 1 and "cannot be larger than max identity" in p.stderr
    p = run_cli(["-i", "x.mdf", "-r", "g.fa", "-o", "o.fq", "-O", "bam"])
    assert p.returncode == 2 and "invalid choice" in p.stderr
    p = run_cli(["-h"])
    assert p.returncode == 0 and "TKSM_MODELS" in p.stdout
    # a valid command line gets as far as needing the device
    import torch
    if not torch.cuda.is_available():
        fa = tmp_path / "g.fa"
        fa.write_text(">c\nACGT\n")
        mdf = tmp_path / "x.mdf"
        mdf.write_text("+m\t1\t\nc\t0\t4\t+\t\n")
        p = run_cli(["-i", str(mdf), "-r", str(fa), "--perfect", str(tmp_path / "o.fasta")])
        assert p.returncode == 1 and "no CPU fallback" in p.stderr


def test_model_lookup_first_directory_wins(Sequencer, tmp_path):
    # py/sequence.py:21-31: directories are scanned in reverse, so the first entry wins on a clash
    for d, line in (("a", "=;1;5:1.0,\nX;1;5:1.0,\nI;1;5:1.0,\n"), ("b", "=;1;7:1.0,\nX;1;7:1.0,\nI;1;7:1.0,\n")):
        os.makedirs(tmp_path / d / "badread")
        import gzip
        with gzip.open(tmp_path / d / "badread" / "toy.qscore.gz", "wt") as f:
            f.write("overall;3;1:1.0,\n" + line)
    old = os.environ["TKSM_MODELS"]
    try:
        os.environ["TKSM_MODELS"] = f"{tmp_path / 'a'}:{tmp_path / 'b'}"
        s = Sequencer(host_only=True)
        s.load_models("random", "toy")
        tn = s.tables_numpy()
        assert set(mo.decode_qrow(tn, mo.packed_row_for(tn, "="))) == {5}
    finally:
        os.environ["TKSM_MODELS"] = old


@pytest.mark.parametrize("shape", [0, 1, 2, 3])
def test_synthetic_mdf_shapes_parse_like_the_reference_reader(Sequencer, shape):
    """tksm_seq_synth_mdf writes the emitter format of src/interval.h:898-905 for the BASELINE shapes (Tsb, README
    bulk, length skew, scRNA with PCR mods): the oracle's restatement of mdf_generator / mdf_to_seq reads it (mods
    in range, `.cycle` ids kept) and the shard plan counts the same reads."""
    rng = np.random.default_rng(1)
    g = rng.choice(np.frombuffer(b"ACGT", np.uint8), size=400000).tobytes().decode()
    fasta = ">chr1\n" + g[:300000] + "\n>chr2\n" + g[300000:] + "\n"
    s = Sequencer(host_only=True)
    s.add_reference_text(fasta)
    n = 300
    mdf = s.synth_mdf(shape, n, seed=4, min_len=300, max_len=50000).decode()
    mols = so.parse_mdf(mdf)
    assert len(mols) == n                       # depth expanded
    refs = so.load_fasta(fasta)
    seqs = [so.molecule_sequence(refs, iv) for _, iv in mols]
    assert all(len(q) > 0 for q in seqs)
    plan = s.shard_plan(mdf, 3)
    assert sum(p[3] for p in plan) == n
    if shape == 3:
        assert any("." in m for m, _ in mols)
        with_mods = [l for l in mdf.split("\n") if l and l[0] != "+" and l.split("\t")[4]]
        assert len(with_mods) > 50
        assert any(l[0] in "ACGT" and l.split("\t")[0] not in ("chr1", "chr2") for l in with_mods)   # a literal with a mod
