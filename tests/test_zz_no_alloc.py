"""tksm_seq_record_offsets_static: the record-offset scan with its scratch in a static device array (no allocation).

Collected last on purpose.  The entry point is checked against the existing one that
contains the same scan, tksm_seq_simulate_text: the offsets it computes from the record lengths must be the byte offsets
at which that call put the records, exactly (integer work: no tolerance), on two calls in a row that share the scratch.
"""
import ctypes as C

import numpy as np
import pytest


def _records(out, fastq):
    lines = out.split(b"\n")
    assert lines[-1] == b""
    per = 4 if fastq else 2
    recs = [b"\n".join(lines[i:i + per]) + b"\n" for i in range(0, len(lines) - 1, per)]
    assert b"".join(recs) == out
    return recs


def _offsets_static(L, torch, lengths):
    n = len(lengths)
    d_len = torch.from_numpy(np.asarray(lengths, dtype=np.int64).astype(np.int32)).cuda()   # u32 values below 2^31
    d_off = torch.full((max(n, 1),), -1, dtype=torch.int64, device="cuda")
    d_tot = torch.full((1,), -1, dtype=torch.int64, device="cuda")
    torch.cuda.synchronize()
    rc = L.tksm_seq_record_offsets_static(0, d_len.data_ptr(), n, d_off.data_ptr(), d_tot.data_ptr(), None)
    assert rc == 0, L.tksm_seq_last_error(None)
    torch.cuda.synchronize()
    return d_off.cpu().numpy()[:n], int(d_tot.item())


@pytest.mark.gpu
def test_static_record_offsets_match_the_output_of_simulate_text(built):
    import torch
    from tksm_b200 import _lib
    from tksm_b200.sequence import BADREAD, Sequencer
    L = _lib.lib()
    s = Sequencer(seed=3)
    rng = np.random.default_rng(8)
    s.add_reference_text(">chr1\n" + "".join(rng.choice(list("ACGT"), 200000)) + "\n")
    s.load_models("random", "random")
    # two calls in a row: a large FASTQ run (several scan blocks), then a smaller FASTA run that leaves most of the
    # scratch as the first call wrote it, then the first one again
    big = s.simulate(s.synth_mdf(1, 9000, seed=1, min_len=200, max_len=900), BADREAD, fastq=True)
    small = s.simulate(s.synth_mdf(0, 700, seed=2, min_len=200, max_len=900), BADREAD, fastq=False)
    for out, fastq in ((big, True), (small, False), (big, True)):
        recs = _records(out, fastq)
        lengths = [len(r) for r in recs]
        assert len(recs) >= 700 and (len(recs) > 2 * 4096 or not fastq)
        off, total = _offsets_static(L, torch, lengths)
        want = np.concatenate(([0], np.cumsum(lengths)[:-1]))
        assert total == len(out)
        assert np.array_equal(off, want)
        first = b"@" if fastq else b">"
        for i in (0, 1, len(recs) // 2, len(recs) - 1):
            assert out[off[i]:off[i] + lengths[i]] == recs[i] and recs[i][:1] == first
    # set-up form and the limit
    off, total = _offsets_static(L, torch, [])
    assert total == 0
    dummy = torch.zeros(1, dtype=torch.int64, device="cuda")
    rc = L.tksm_seq_record_offsets_static(0, dummy.data_ptr(), (1 << 24) + 1, dummy.data_ptr(), dummy.data_ptr(), None)
    assert rc == 2 and b"exceed the static scratch" in L.tksm_seq_last_error(None)
