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
