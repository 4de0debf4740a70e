"""Alignment of the simulated read back to its fragment (py/tksm_badread.py:613).

Three statements of the same banded alignment are compared on the reads the CPU twin simulates:
  * oracle/nw.c        the full-matrix stand-in for edlib the reference fixtures were made with,
  * oracle/twin.c      twin_bnw, the banded integer DP that defines what the kernel computes,
  * tksm_b200/csrc/align.cuh  the product's bit-vector chunk functions, compiled for the host by
                       tests/native/align_emul.cpp and driven the way the CUDA team drives them
                       (assumed chunk-edge states, redone until the neighbours agree).
The GPU suite then checks kernel == twin byte for byte (tests/test_gpu_parity.py).
"""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

from oracle import seq_oracle as so
from oracle import stats as st
from oracle.make_goldens import dist_inputs

HERE = os.path.dirname(os.path.abspath(__file__))
HAVE_MODELS = os.path.exists(os.path.join(os.environ["TKSM_MODELS"], "badread", "nanopore2020.error.gz"))
needs_models = pytest.mark.skipif(not HAVE_MODELS, reason="Badread model files not present")

CB = C.CFUNCTYPE(None, C.POINTER(C.c_uint8), C.c_uint32, C.POINTER(C.c_uint8), C.c_uint32, C.POINTER(C.c_uint8),
                 C.POINTER(C.c_uint8), C.c_int, C.POINTER(C.c_uint32))


@pytest.fixture(scope="module")
def emul():
    src = os.path.join(HERE, "native", "align_emul.cpp")
    out = os.path.join(HERE, "native", "_build", "libalign_emul.so")
    os.makedirs(os.path.dirname(out), exist_ok=True)
    hdr = os.path.join(HERE, "..", "tksm_b200", "csrc", "align.cuh")
    if not os.path.exists(out) or os.path.getmtime(out) < max(os.path.getmtime(src), os.path.getmtime(hdr)):
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-Wall", "-o", out, src])
    lib = C.CDLL(out)
    lib.align_emul.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_int,
                               C.c_void_p, C.c_void_p, C.c_void_p]
    return lib


def twin_alignments(em, qm, ident, n_reads, callback):
    """Runs the twin over the distribution-test molecules; `callback` sees every alignment."""
    from tksm_b200.sequence import Sequencer
    fasta, mdf = dist_inputs(n_mol=n_reads)
    s = Sequencer(identity=ident, seed=42, host_only=True)
    s.add_reference_text(fasta)
    s.load_models(em, qm)
    refs = so.load_fasta(fasta)
    mols = so.parse_mdf(mdf)
    seqs = [so.molecule_sequence(refs, iv) for _, iv in mols]
    ids = [m for m, _ in mols]
    lib = so.twin_lib()
    cb = CB(callback)
    lib.twin_set_align_cb.argtypes = [CB]
    lib.twin_set_align_cb(cb)
    try:
        so.twin_simulate(s.tables(), seqs, ids, badread=True, fastq=True)
    finally:
        lib.twin_set_align_cb(C.cast(None, CB))


CODE = np.arange(256, dtype=np.uint8)
for _i, _ch in enumerate(b"ACGT"):
    CODE[_ch] = _i


@needs_models
@pytest.mark.parametrize("em,qm,ident,n_reads,threads", [
    ("nanopore2020", "nanopore2020", (84.0, 99.0, 5.5), 250, 32),
    ("pacbio2016", "pacbio2016", (84.0, 99.0, 5.5), 120, 64),
    ("nanopore2018", "nanopore2018", (70.0, 85.0, 6.0), 120, 32),    # low identity: some reads need the 64-bit band
    ("pacbio2016", "pacbio2016", (99.0, 99.9, 0.5), 120, 256),
])
def test_bitvector_chunks_equal_the_integer_band(built, emul, em, qm, ident, n_reads, threads):
    seen = {"n": 0, "wide": 0, "bad": []}

    def cb(fa, n, sq, m, ops, dc, status, soff):
        seen["n"] += 1
        frag = CODE[np.frombuffer(C.string_at(fa, n), np.uint8)]
        read = CODE[np.frombuffer(C.string_at(sq, m), np.uint8)]
        slen = np.diff(np.ctypeslib.as_array(soff, (n + 1,)).astype(np.int64)).astype(np.uint8)
        eq, cols, rounds = C.c_uint32(), C.c_uint32(), (C.c_uint32 * 2)()
        rb = read.copy()
        r = emul.align_emul(frag.ctypes.data, slen.ctypes.data, rb.ctypes.data, n, m, threads, 0, C.byref(eq),
                            C.byref(cols), rounds)
        if r == 1:
            seen["wide"] += 1
            rb = read.copy()
            r = emul.align_emul(frag.ctypes.data, slen.ctypes.data, rb.ctypes.data, n, m, threads, 1, C.byref(eq),
                                C.byref(cols), rounds)
        if status:                       # the twin gave up on both bands: so must the product
            if r != 1:
                seen["bad"].append(("twin failed, product did not", n, m))
            return
        want = np.frombuffer(C.string_at(ops, m), np.uint8)
        wdc = np.frombuffer(C.string_at(dc, m), np.uint8)
        wop = (np.where(want == ord("="), 0, np.where(want == ord("X"), 1, 2)) | (wdc.astype(np.int64) << 2)).astype(np.uint8)
        if r != 0 or not np.array_equal(rb, wop):
            seen["bad"].append((r, n, m))
        elif eq.value != int((want == ord("=")).sum()):
            seen["bad"].append(("matches", eq.value))

    twin_alignments(em, qm, ident, n_reads, cb)
    assert seen["n"] == n_reads
    assert not seen["bad"], seen["bad"][:5]


@needs_models
def test_band_around_the_edit_script_finds_the_full_matrix_alignment(built):
    """twin_bnw (band of 32 / 64 rows around the edit script) against oracle/nw.c on whole reads."""
    seen = {"n": 0, "diff": 0, "fail": 0}

    def cb(fa, n, sq, m, ops, dc, status, soff):
        seen["n"] += 1
        if status:
            seen["fail"] += 1
            return
        frag, read = C.string_at(fa, n), C.string_at(sq, m)
        o, d = C.string_at(ops, m), C.string_at(dc, m)
        full = st.align_ops(read, frag).tobytes()
        mine = b"".join((b"D" * d[i] if i else b"") + o[i:i + 1] for i in range(m))
        lead = len(full) - len(full.lstrip(b"D"))
        body = re.sub(b"D{16,}", b"D" * 15, full.lstrip(b"D").rstrip(b"D"))   # runs are recorded up to 15
        if mine != body or min(lead, 15) != d[0]:
            seen["diff"] += 1

    twin_alignments("nanopore2020", "nanopore2020", (84.0, 99.0, 5.5), 400, cb)
    assert seen["n"] == 400 and seen["fail"] == 0 and seen["diff"] == 0, seen


@pytest.fixture(scope="module")
def emul_one():
    """Host build of align_kernel's per-thread routine: the lines of kernels.cuh from `struct ByteStream` to the end
    of align_one, with the few CUDA names stubbed (tests/native/align_one_emul.cpp)."""
    src = os.path.join(HERE, "native", "align_one_emul.cpp")
    kern = os.path.join(HERE, "..", "tksm_b200", "csrc", "kernels.cuh")
    bdir = os.path.join(HERE, "native", "_build")
    os.makedirs(bdir, exist_ok=True)
    lines = open(kern).read().split("\n")
    a = next(i for i, l in enumerate(lines) if l.startswith("struct ByteStream {"))
    b = next(i for i, l in enumerate(lines) if l.startswith("// -------------------------------------------------------------- align_kernel"))
    open(os.path.join(bdir, "align_one_body.inc"), "w").write("\n".join(lines[a:b]) + "\n")
    out = os.path.join(bdir, "libalign_one_emul.so")
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-o", out, src])
    lib = C.CDLL(out)
    lib.align_one_emul.argtypes = [C.c_void_p] * 4 + [C.c_uint32, C.c_uint32, C.c_int, C.c_void_p, C.c_void_p]
    return lib


@needs_models
@pytest.mark.parametrize("em,ident,n_reads", [("nanopore2020", (84.0, 99.0, 5.5), 250), ("nanopore2018", (70.0, 85.0, 6.0), 100)])
def test_thread_per_read_alignment_equals_the_integer_band(built, emul_one, em, ident, n_reads):
    """align_kernel (one thread per read: streamed forward pass, traceback through register windows) against twin_bnw, with
    the kernel's band rule: 32-bit band up to ALIGN_WIDE64_FROM slots, retried with 64 bits; 64 bits at once above."""
    seen = {"n": 0, "wide": 0, "bad": []}

    def cb(fa, n, sq, m, ops, dc, status, soff):
        seen["n"] += 1
        frag = CODE[np.frombuffer(C.string_at(fa, n), np.uint8)]
        read = CODE[np.frombuffer(C.string_at(sq, m), np.uint8)]
        slen = np.diff(np.ctypeslib.as_array(soff, (n + 1,)).astype(np.int64)).astype(np.uint8)
        out = np.zeros(m, np.uint8)
        eq, cols = C.c_uint32(), C.c_uint32()
        wide = 1 if n > 2048 else 0
        r = emul_one.align_one_emul(frag.ctypes.data, slen.ctypes.data, read.ctypes.data, out.ctypes.data, n, m, wide,
                                    C.byref(eq), C.byref(cols))
        if r == 1 and not wide:
            wide = 1
            r = emul_one.align_one_emul(frag.ctypes.data, slen.ctypes.data, read.ctypes.data, out.ctypes.data, n, m, 1,
                                        C.byref(eq), C.byref(cols))
        seen["wide"] += wide
        if status:
            if r != 1:
                seen["bad"].append(("twin failed, kernel routine did not", n, m))
            return
        want = np.frombuffer(C.string_at(ops, m), np.uint8)
        wdc = np.frombuffer(C.string_at(dc, m), np.uint8)
        wop = (np.where(want == ord("="), 0, np.where(want == ord("X"), 1, 2)) | (wdc.astype(np.int64) << 2)).astype(np.uint8)
        if r != 0 or not np.array_equal(out, wop) or eq.value != int((want == ord("=")).sum()):
            seen["bad"].append((r, wide, n, m))

    twin_alignments(em, em, ident, n_reads, cb)
    assert seen["n"] == n_reads and seen["wide"] > 0
    assert not seen["bad"], seen["bad"][:5]
