"""oracle/badread_port.py — driver of oracle/badread_port.c.  TEST INFRASTRUCTURE ONLY.

The C file is a literal port of the reference's sequential Badread algorithm
(py/tksm_badread.py:119-213, 324-451, 584-655, 741-745); this module parses the model files the way
ErrorModel.load_from_file (py/tksm_badread.py:91-117) and QScoreModel.load_from_file (:546-582) do and
hands flat arrays to it.  Used by tests/test_badread_port.py (statistics against the reference's own
reads) and by bench.py as the timed CPU baseline.  Never imported by tksm_b200/.
"""
import ctypes as C
import os
import subprocess

import numpy as np

from . import model_oracle as mo

_HERE = os.path.dirname(os.path.abspath(__file__))
_lib = None
SLOT_CAP = 15


class PortModel(C.Structure):
    _fields_ = [
        ("k", C.c_int), ("has_table", C.c_int),
        ("alt_begin", C.c_void_p), ("alt_prob", C.c_void_p), ("alt_slots", C.c_void_p),
        ("kq", C.c_int), ("n_cigars", C.c_int),
        ("cigar_blob", C.c_void_p), ("cigar_off", C.c_void_p), ("q_begin", C.c_void_p),
        ("q_score", C.c_void_p), ("q_prob", C.c_void_p), ("hash", C.c_void_p), ("hash_mask", C.c_uint32),
        ("ident_const", C.c_int),
        ("ident_mean", C.c_double), ("beta_a", C.c_double), ("beta_b", C.c_double), ("ident_max", C.c_double),
    ]


def lib():
    global _lib
    if _lib is None:
        so = os.path.join(_HERE, "_build", "libbadread_port.so")
        if not os.path.exists(so):
            subprocess.check_call(["make", "-C", _HERE, "_build/libbadread_port.so"])
        _lib = C.CDLL(so)
        _lib.port_simulate.restype = C.c_uint64
        _lib.port_simulate.argtypes = [C.POINTER(PortModel), C.c_uint64, C.c_void_p, C.c_void_p, C.c_uint64, C.c_int,
                                       C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p,
                                       C.c_void_p]
        _lib.port_align_kmers.restype = C.c_int
        _lib.port_align_kmers.argtypes = [C.c_char_p, C.c_int, C.c_char_p, C.c_int, C.c_void_p]
        _lib.port_hash.restype = C.c_uint32
        _lib.port_hash.argtypes = [C.c_char_p, C.c_int]
    return _lib


def _models_dir():
    return os.path.join(os.environ.get("TKSM_MODELS", os.path.join(os.path.dirname(_HERE), "tksm_b200", "models")),
                        "badread")


def _resolve(name, kind):
    if os.path.exists(name):
        return name
    p = os.path.join(_models_dir(), f"{name}.{kind}.gz")
    if os.path.exists(p):
        return p
    raise FileNotFoundError(f"{kind} model {name!r}")


class Model:
    """Holds the numpy buffers the C struct points into."""

    def __init__(self, error_model="nanopore2020", qscore_model="nanopore2020", identity=(84.0, 99.0, 5.5)):
        L = lib()
        self.keep = []
        m = PortModel()
        # ---- error model (py/tksm_badread.py:70-117)
        if error_model == "random":
            m.k, m.has_table = 1, 0
        else:
            k = None
            begin, probs, slots = [0], [], []
            rows = {}
            for kmer, alts in mo.iter_error_model(_resolve(error_model, "error")):
                k = k or len(kmer)
                assert len(kmer) == k and alts[0][0] == kmer
                rows[kmer] = alts
            n_kmers = 4 ** k
            buf = np.zeros((k, 1 + SLOT_CAP), np.uint8)
            for code in range(n_kmers):
                kmer = "".join("ACGT"[(code >> (2 * (k - 1 - j))) & 3] for j in range(k))
                for alt, p in rows.get(kmer, []):
                    rc = L.port_align_kmers(kmer.encode(), k, alt.encode(), len(alt), buf.ctypes.data)
                    assert rc == 0, (kmer, alt)
                    slots.append(buf.copy())
                    probs.append(p)
                begin.append(len(probs))
            m.k, m.has_table = k, 1
            a = np.array(begin, np.uint32)
            b = np.array(probs, np.float64)
            c = np.ascontiguousarray(np.stack(slots)) if slots else np.zeros((0, k, 1 + SLOT_CAP), np.uint8)
            self.keep += [a, b, c]
            m.alt_begin, m.alt_prob, m.alt_slots = a.ctypes.data, b.ctypes.data, c.ctypes.data
        # ---- qscore model (py/tksm_badread.py:464-582)
        qm = mo.builtin_qscore_model(qscore_model) if qscore_model in ("random", "ideal") else \
            mo.load_qscore_model(_resolve(qscore_model, "qscore"))
        cigars = list(qm)
        blob = "".join(cigars).encode()
        off = np.zeros(len(cigars) + 1, np.uint32)
        off[1:] = np.cumsum([len(c) for c in cigars])
        qb = np.zeros(len(cigars) + 1, np.uint32)
        qb[1:] = np.cumsum([len(qm[c][0]) for c in cigars])
        qs = np.array([q for c in cigars for q in qm[c][0]], np.uint8)
        qp = np.array([p for c in cigars for p in qm[c][1]], np.float64)
        size = 1
        while size < 4 * len(cigars):
            size *= 2
        table = np.full(size, -1, np.int32)
        for i, cg in enumerate(cigars):
            h = L.port_hash(cg.encode(), len(cg)) & (size - 1)
            while table[h] >= 0:
                h = (h + 1) & (size - 1)
            table[h] = i
        blob_arr = np.frombuffer(blob + b"\0", np.uint8).copy()
        self.keep += [blob_arr, off, qb, qs, qp, table]
        m.kq, m.n_cigars = mo.qscore_kmer_size(qm), len(cigars)
        m.cigar_blob, m.cigar_off, m.q_begin = blob_arr.ctypes.data, off.ctypes.data, qb.ctypes.data
        m.q_score, m.q_prob, m.hash, m.hash_mask = qs.ctypes.data, qp.ctypes.data, table.ctypes.data, size - 1
        # ---- identities (py/tksm_badread.py:702-757)
        mean, mx, sd = identity
        if mean == mx or sd == 0.0:
            m.ident_const, m.ident_mean = 1, mean / 100.0
        else:
            m.ident_const = 0
            m.beta_a, m.beta_b = mo.beta_parameters(mean, sd, mx)
            m.ident_max = mx / 100.0
        self.c = m


_cache = {}


def get_model(error_model="nanopore2020", qscore_model="nanopore2020", identity=(84.0, 99.0, 5.5)):
    key = (error_model, qscore_model, tuple(identity))
    if key not in _cache:
        _cache[key] = Model(error_model, qscore_model, identity)
    return _cache[key]


def simulate(model, seqs, seed=1, compute_qscores=True, threads=1, want_reads=True):
    """seqs: list of bytes.  Returns (reads, quals, identities) or just the base count."""
    n = len(seqs)
    blob = np.frombuffer(b"".join(seqs) + b"\0", np.uint8)
    off = np.zeros(n + 1, np.uint64)
    off[1:] = np.cumsum([len(s) for s in seqs], dtype=np.uint64)
    lens = np.zeros(max(n, 1), np.uint32)
    ident = np.zeros(max(n, 1), np.float64)
    if not want_reads:
        return int(lib().port_simulate(C.byref(model.c), n, blob.ctypes.data, off.ctypes.data, seed,
                                       int(compute_qscores), threads, None, None, None, 0, lens.ctypes.data,
                                       ident.ctypes.data))
    cap = int(off[-1]) * 2 + 64 * n + 64
    oseq, oqual = np.zeros(cap, np.uint8), np.zeros(cap, np.uint8)
    ooff = np.zeros(n + 1, np.uint64)
    lib().port_simulate(C.byref(model.c), n, blob.ctypes.data, off.ctypes.data, seed, int(compute_qscores), threads,
                        oseq.ctypes.data, oqual.ctypes.data, ooff.ctypes.data, cap, lens.ctypes.data,
                        ident.ctypes.data)
    assert int(ooff[n]) <= cap
    reads = [oseq[int(ooff[i]):int(ooff[i + 1])].tobytes() for i in range(n)]
    quals = [oqual[int(ooff[i]):int(ooff[i + 1])].tobytes() for i in range(n)]
    return reads, quals, ident[:n].copy()


def simulate_count(seqs, threads=1, error_model="nanopore2020", qscore_model="nanopore2020",
                   identity=(84.0, 99.0, 5.5), seed=1):
    """bench.py's CPU leg: simulated bases for `seqs` (errors + qscores), model load excluded by caching."""
    return simulate(get_model(error_model, qscore_model, identity), seqs, seed, True, threads, want_reads=False)


def to_fastq(seqs, ids, reads, quals, identities):
    """FASTQ + perfect FASTA bytes in the layout oracle/stats.summarise() expects."""
    fq, pf = [], []
    for mid, raw, rd, ql, idn in zip(ids, seqs, reads, quals, identities):
        mid = mid if isinstance(mid, str) else mid.decode()
        fq.append(f"@r length={len(rd)} error_free_length={len(raw)} read_identity={idn * 100.0:.2f}% "
                  f"molecule_id={mid}\n".encode() + rd + b"\n+\n" + ql + b"\n")
        pf.append(f">r length={len(raw)} error_free_length={len(raw)} read_identity=100.00% "
                  f"molecule_id={mid}\n".encode() + raw + b"\n")
    return b"".join(fq), b"".join(pf)
