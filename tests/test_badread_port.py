"""oracle/badread_port.c (literal C port of the reference's sequential algorithm, the timed CPU
baseline) against the statistics of reads simulated by the unmodified Python reference."""
import os

import pytest

from oracle import seq_oracle as so
from oracle import stats as st
from oracle.make_goldens import dist_inputs
from test_distribution import compare, load_ref, needs_models


def test_align_kmers_docstring_examples(built):
    # py/tksm_badread.py:151-153
    import numpy as np
    from oracle import badread_port as bp
    buf = np.zeros((4, 16), np.uint8)
    for kmer, alt, want in (("ACGT", "ACGTT", ["A", "C", "GT", "T"]), ("ACGT", "ACT", ["A", "C", "", "T"])):
        assert bp.lib().port_align_kmers(kmer.encode(), 4, alt.encode(), len(alt), buf.ctypes.data) == 0
        got = [bytes(buf[j, 1:1 + buf[j, 0]]).decode() for j in range(4)]
        assert got == want


@needs_models
@pytest.mark.parametrize("name,n_reads", [("nanopore2020", 500), ("random", 400)])
def test_port_matches_reference_distributions(built, name, n_reads):
    from oracle import badread_port as bp
    ref = load_ref(name)
    c = ref["config"]
    ident = tuple(float(x) for x in c["ident"].split(","))
    fasta, mdf = dist_inputs(**ref["inputs"])
    refs = so.load_fasta(fasta)
    mols = so.parse_mdf(mdf)[:n_reads]
    seqs = [so.molecule_sequence(refs, iv) for _, iv in mols]
    seqs = [s.encode() if isinstance(s, str) else s for s in seqs]
    ids = [m for m, _ in mols]
    model = bp.get_model(c["em"], c["qm"], ident)
    reads, quals, idn = bp.simulate(model, seqs, seed=3, threads=os.cpu_count() or 1)
    # thread-count independence (one RNG stream per read)
    r1, _, _ = bp.simulate(model, seqs[:20], seed=3, threads=1)
    assert r1 == reads[:20]
    fq, pf = bp.to_fastq(seqs, ids, reads, quals, idn)
    compare(name, st.summarise(fq, pf), ref)
