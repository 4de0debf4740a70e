"""oracle/seq_oracle.py — CPU restatement of the Seq path.  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import this module; the
product (tksm_b200/) never does.

Deterministic stage (byte-exact oracle), restated from the reference with its quirks:
    load_fasta           <- generate_fasta / get_reference_seqs   py/sequence.py:168-194
    parse_mdf            <- mdf_generator                         py/sequence.py:197-221
    molecule_sequence    <- mdf_to_seq + apply_modifications
                            + reverse_complement                  py/sequence.py:224-239, 303-313
    perfect_record       <- perfect() + formatters                py/sequence.py:261-288
It is pinned against the unmodified reference by tests/golden/deterministic_*.json
(oracle/make_goldens.py runs /root/reference/py/sequence.py --perfect to produce them) and the
README / docstring vectors of SURVEY.md §4.

Stochastic stage: `twin_simulate` drives oracle/twin.c (the CPU twin of the PB1 model, bit-exact
oracle for the CUDA kernels) on the sequences produced by the deterministic stage above.
"""
import ctypes as C
import gzip
import io
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_BUILD = os.path.join(_HERE, "_build")

_COMP = bytes.maketrans(b"ACGTN", b"TGCAN")  # every other byte maps to itself (py/sequence.py:225-226)


def build(targets=("libtwin.so", "libnw.so", "libbadread_port.so")):
    subprocess.check_call(["make", "-C", _HERE] + [f"_build/{t}" for t in targets])


def _text_lines(text):
    """Lines as Python's text-mode iteration yields them after `.rstrip("\\n")`: universal newlines."""
    if isinstance(text, bytes):
        text = text.decode("latin-1")
    return io.StringIO(text, newline=None).read().split("\n")


def load_fasta(text, into=None):
    """name -> sequence (str, as stored: not upper-cased).  Later records override earlier names.
    An empty line raises IndexError like the reference does (`l[0]` on an empty string)."""
    refs = {} if into is None else into
    lines = _text_lines(text)
    if lines and lines[-1] == "":
        lines.pop()  # the terminator of the last line, not an empty line
    name, chunks = "", []
    for ln in lines:
        if ln[0] == ">":
            if chunks:
                refs[name] = "".join(chunks)
                chunks = []
            name = ln[1:].split(" ")[0]
        else:
            chunks.append(ln)
    refs[name] = "".join(chunks)
    return refs


def load_fasta_file(path, into=None):
    opener = gzip.open if path.endswith(".gz") else open
    with opener(path, "rb") as f:
        return load_fasta(f.read(), into)


def parse_mdf(text):
    """[(molecule_id, [(chrom, start, end, strand, mods), ...]), ...] with `depth` copies per record."""
    out = []
    lines = _text_lines(text)
    if lines and lines[-1] == "":
        lines.pop()
    cur_id, cur_depth, cur_iv = None, 0, []
    for ln in lines:
        f = ln.split("\t")
        if f[0][0] == "+":
            if cur_id is not None:
                out.extend((cur_id, cur_iv) for _ in range(cur_depth))
            cur_id, cur_depth, cur_iv = f[0][1:], int(f[1]), []
        else:
            if len(f) == 4:  # the C++ reader's tolerance (src/mdf.h:88-90); the Python reference needs 5
                f = f + [""]
            chrom, start, end, strand, mods = f
            cur_iv.append((chrom, int(start), int(end), strand, mods))
    if cur_id is not None:
        out.extend((cur_id, cur_iv) for _ in range(cur_depth))
    return out


def molecule_sequence(refs, intervals):
    """Error-free molecule as bytes."""
    parts = []
    for chrom, start, end, strand, mods in intervals:
        seg = bytearray(refs.get(chrom, chrom)[start:end].upper().encode("latin-1"))
        if mods:
            for item in mods.split(","):
                seg[int(item[:-1])] = ord(item[-1])   # IndexError when out of range, as in the reference
        seg = bytes(seg)
        parts.append(seg if strand == "+" else seg.translate(_COMP)[::-1])
    return b"".join(parts)


def perfect_record(read_id, molecule_id, seq, fastq):
    head = (f"{'@' if fastq else '>'}{read_id} length={len(seq)} error_free_length={len(seq)} "
            f"read_identity=100.00% molecule_id={molecule_id}").encode()
    if fastq:
        return head + b"\n" + seq + b"\n+\n" + b"K" * len(seq) + b"\n"
    return head + b"\n" + seq + b"\n"


# ----------------------------------------------------------------- CPU twin

_twin = None


def twin_lib():
    global _twin
    if _twin is None:
        so = os.path.join(_BUILD, "libtwin.so")
        if not os.path.exists(so):
            build(("libtwin.so",))
        _twin = C.CDLL(so)
        _twin.twin_simulate.restype = C.c_size_t
        _twin.twin_simulate.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                        C.c_uint64, C.c_int, C.c_int, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p]
        _twin.twin_set_tail_lens.restype = None
        _twin.twin_set_tail_lens.argtypes = [C.c_void_p]
        _twin.twin_expneg.restype = C.c_float
        _twin.twin_expneg.argtypes = [C.c_float]
    return _twin


def philox4x32_10(counter, key):
    c = (C.c_uint32 * 4)(*counter)
    k = (C.c_uint32 * 2)(*key)
    o = (C.c_uint32 * 4)()
    twin_lib().twin_philox(c, k, o)
    return list(o)


def twin_simulate(tables, seqs, ids, ordinal_base=0, badread=True, fastq=True, want_lengths=False, tails=None):
    """tables: ctypes struct with the layout of tksm_seq_tables (include/tksm_seq.h);
    seqs: list of bytes (error-free molecules); ids: list of str/bytes;
    tails: optional list of bytes, the tail-noise bases of each read (oracle/tail_oracle.py)."""
    lib = twin_lib()
    n = len(seqs)
    if tails is not None:
        assert len(tails) == n
        tl = np.array([len(t) for t in tails] + [0], np.uint32)
        lib.twin_set_tail_lens(tl.ctypes.data)
        try:
            return twin_simulate(tables, [s + t for s, t in zip(seqs, tails)], ids, ordinal_base, badread, fastq,
                                 want_lengths)
        finally:
            lib.twin_set_tail_lens(None)
    seq_off = np.zeros(n + 1, np.uint64)
    seq_off[1:] = np.cumsum([len(s) for s in seqs], dtype=np.uint64)
    idb = [i.encode() if isinstance(i, str) else i for i in ids]
    id_off = np.zeros(n + 1, np.uint64)
    id_off[1:] = np.cumsum([len(s) for s in idb], dtype=np.uint64)
    sblob = np.frombuffer(b"".join(seqs) + b"\0", np.uint8)
    iblob = np.frombuffer(b"".join(idb) + b"\0", np.uint8)
    lens = np.zeros(max(n, 1), np.uint32)
    ident = np.zeros(2 * max(n, 1), np.uint32)
    cap = int(seq_off[-1]) * 5 + 400 * n + int(id_off[-1]) + 64
    out = np.zeros(cap, np.uint8)
    need = lib.twin_simulate(C.byref(tables), n, sblob.ctypes.data, seq_off.ctypes.data, iblob.ctypes.data,
                             id_off.ctypes.data, ordinal_base, 1 if badread else 0, 1 if fastq else 0,
                             out.ctypes.data, cap, lens.ctypes.data, ident.ctypes.data)
    if need > cap:
        out = np.zeros(need, np.uint8)
        lib.twin_simulate(C.byref(tables), n, sblob.ctypes.data, seq_off.ctypes.data, iblob.ctypes.data,
                          id_off.ctypes.data, ordinal_base, 1 if badread else 0, 1 if fastq else 0,
                          out.ctypes.data, need, lens.ctypes.data, ident.ctypes.data)
    data = out[:need].tobytes()
    if want_lengths:
        return data, lens[:n].copy(), ident[:2 * n].reshape(-1, 2).copy()
    return data


def parse_records(data, fastq=True):
    """[(header:str, seq:bytes, qual:bytes|None)] from FASTQ/FASTA bytes (single-line records)."""
    lines = data.split(b"\n")
    assert lines[-1] == b""
    lines.pop()
    step = 4 if fastq else 2
    assert len(lines) % step == 0, "truncated output"
    recs = []
    for i in range(0, len(lines), step):
        if fastq:
            assert lines[i + 2] == b"+"
            recs.append((lines[i].decode(), lines[i + 1], lines[i + 3]))
        else:
            recs.append((lines[i].decode(), lines[i + 1], None))
    return recs
