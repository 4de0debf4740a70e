"""Host-side mirror of the reference's Seq module interface (py/sequence.py) over the C ABI.

The reference exposes a script (`tksm sequence ...`) whose pieces are `get_reference_seqs`,
`mdf_generator`, `mdf_to_seq`, `badread`/`perfect` and the two formatters.  Here the same steps are
methods of `Sequencer`; the arithmetic runs in libtksm_seq.so on the GPU, never in Python.
"""
import ctypes as C
import os
import sys

import numpy as np

from . import _lib
from ._lib import BADREAD, FASTA, FASTQ, FASTQ_BADREAD, FASTQ_PERFECT, PERFECT, Config, Stats, Tables


class SeqError(RuntimeError):
    pass


def output_is_fastq(path):
    """get_output_file's rule (py/sequence.py:291-300): extension after stripping .gz."""
    p = path[:-3] if path.endswith(".gz") else path
    return p.endswith(".fastq") or p.endswith(".fq")


class Sequencer:
    def __init__(self, identity=(84.0, 99.0, 5.5), seed=42, skip_qual_compute=False, device=0, host_only=False,
                 corr=None, batch_reads=0, batch_bases=0, threads=0):
        self._L = _lib.lib()
        cfg = Config()
        self._L.tksm_seq_default_config(C.byref(cfg))
        cfg.device = device
        cfg.seed = seed
        cfg.identity_mean, cfg.identity_max, cfg.identity_stdev = identity
        cfg.skip_qual_compute = int(skip_qual_compute)
        cfg.host_only = int(host_only)
        cfg.batch_reads = batch_reads
        cfg.batch_bases = batch_bases
        cfg.threads = threads   # host MDF parser threads; 0 = all hardware threads (max 32)
        if corr is not None:
            cfg.calib_overshoot, cfg.calib_kappa = corr
        self._h = self._L.tksm_seq_create(C.byref(cfg))
        if not self._h:
            raise SeqError(self._L.tksm_seq_last_error(None).decode())
        self.last_stats = None

    def close(self):
        if getattr(self, "_h", None):
            self._L.tksm_seq_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc):
        if rc:
            raise SeqError(self._L.tksm_seq_last_error(self._h).decode())

    # ---- Badread extras (include/tksm_seq.h: tksm_seq_extras); no reference counterpart
    def set_extras(self, **kw):
        """junk_reads, random_reads, chimeras (percent); start_adapter=(rate, amount), end_adapter=(rate, amount);
        glitches=(rate, size, skip); start_adapter_seq, end_adapter_seq.  No arguments: all off."""
        if not kw:
            self._ck(self._L.tksm_seq_set_extras(self._h, None))
            return
        x = _lib.Extras()
        x.junk_reads = kw.pop("junk_reads", 0.0)
        x.random_reads = kw.pop("random_reads", 0.0)
        x.chimeras = kw.pop("chimeras", 0.0)
        x.start_adapter_rate, x.start_adapter_amount = kw.pop("start_adapter", (0.0, 0.0))
        x.end_adapter_rate, x.end_adapter_amount = kw.pop("end_adapter", (0.0, 0.0))
        x.glitch_rate, x.glitch_size, x.glitch_skip = kw.pop("glitches", (0.0, 0.0, 0.0))
        for name in ("start_adapter_seq", "end_adapter_seq"):
            v = kw.pop(name, None)
            setattr(x, name, v.encode() if isinstance(v, str) else v)
        if kw:
            raise TypeError(f"unknown extras: {sorted(kw)}")
        self._ck(self._L.tksm_seq_set_extras(self._h, C.byref(x)))

    def set_front_end(self, spec=None):
        """Fused front-end steps, e.g. "plA:normal=15,7.5;Flp:0.5;Tag:FMT5,FMT3;SCB" (include/tksm_seq.h); None: none."""
        self._ck(self._L.tksm_seq_set_front_end(self._h, spec.encode() if spec else None))

    # ---- get_reference_seqs (py/sequence.py:189-194)
    def load_references(self, paths):
        arr = (C.c_char_p * len(paths))(*[p.encode() for p in paths])
        self._ck(self._L.tksm_seq_load_references(self._h, arr, len(paths)))

    def add_reference_text(self, fasta_text):
        b = fasta_text.encode() if isinstance(fasta_text, str) else bytes(fasta_text)
        self._ck(self._L.tksm_seq_add_reference_text(self._h, b, len(b)))

    def synth_reference(self, names, lengths, n_frac=0.01, seed=42):
        n = len(names)
        a = (C.c_char_p * n)(*[s.encode() for s in names])
        l = (C.c_uint64 * n)(*lengths)
        self._ck(self._L.tksm_seq_synth_reference(self._h, a, l, n, n_frac, seed))

    def fetch_reference(self, contig, start, length):
        buf = C.create_string_buffer(length + 1)
        self._ck(self._L.tksm_seq_fetch_reference(self._h, contig.encode(), start, length, buf))
        return buf.raw[:length]

    # ---- ErrorModel / QScoreModel / tail model (py/sequence.py:337-345)
    def load_models(self, error_model=None, qscore_model=None, tail_model=None):
        enc = lambda s: None if s is None else s.encode()
        self._ck(self._L.tksm_seq_load_models(self._h, enc(error_model), enc(qscore_model), enc(tail_model)))

    def tail_draw(self, ordinal, frag_len):
        """tail_model.noise_seq(frag_len) for the read with this global ordinal (py/tksm_badread.py:910-927)."""
        n = C.c_uint32()
        buf = C.create_string_buffer(4096)
        self._ck(self._L.tksm_seq_tail_draw(self._h, ordinal, frag_len, buf, 4096, C.byref(n)))
        if n.value > 4096:
            buf = C.create_string_buffer(n.value)
            self._ck(self._L.tksm_seq_tail_draw(self._h, ordinal, frag_len, buf, n.value, C.byref(n)))
        return buf.raw[:n.value]

    def tables(self):
        t = Tables()
        self._ck(self._L.tksm_seq_get_tables(self._h, C.byref(t)))
        return t

    def tables_numpy(self):
        t = self.tables()
        view = lambda p, n, dt: np.ctypeslib.as_array(p, shape=(n,)).view(dt).copy() if n else np.zeros(0, dt)
        return {
            "k": t.k, "has_err_table": t.has_err_table,
            "q16": view(t.q16, t.n_q16, np.uint16),
            "err_alias": view(t.err_alias, t.n_err_alias, np.uint32),
            "err_alts": view(t.err_alts, t.n_err_alts, np.uint64),
            "ident_const": t.ident_const, "ident_mean": t.ident_mean,
            "ident_tab": view(t.ident_tab, t.n_ident_tab, np.float32),
            "delta_tab": view(t.delta_tab, t.n_delta_tab, np.float32),
            "calib_overshoot": t.calib_overshoot, "calib_kappa": t.calib_kappa, "kq": t.kq,
            "q_direct": view(t.q_direct, t.n_q_direct, np.uint32),
            "q_direct_off": list(t.q_direct_off),
            "q_hash": view(t.q_hash, 2 * t.n_q_hash, np.uint64),
            "q_hash_shift": t.q_hash_shift, "q_hash_mask": t.q_hash_mask,
            "q_rows": view(t.q_rows, 2 * t.n_q_rows, np.uint32),
            "q_alias": view(t.q_alias, t.n_q_alias, np.uint32),
            "seed": t.seed, "compute_qual": t.compute_qual,
            "qcls": view(t.qcls, t.n_qcls, np.uint8),
            "qrep": view(t.qrep, t.n_qrep, np.float64),
            "thr": view(t.thr, t.n_thr, np.uint16),
        }

    # ---- mdf_generator + mdf_to_seq + formatter on host buffers (py/sequence.py:197-320)
    def simulate(self, mdf_text, which=BADREAD, fastq=True, ordinal_base=0):
        b = mdf_text.encode() if isinstance(mdf_text, str) else bytes(mdf_text)
        out, n, st = C.c_void_p(), C.c_size_t(), Stats()
        fmt = FASTQ if fastq else FASTA
        self._ck(self._L.tksm_seq_simulate_text(self._h, b, len(b), which, fmt, ordinal_base, C.byref(out), C.byref(n),
                                                C.byref(st)))
        try:
            data = C.string_at(out, n.value)
        finally:
            self._L.tksm_seq_free(out)
        self.last_stats = st.as_dict()
        return data

    def simulate_view(self, mdf_text, which=BADREAD, fastq=True, ordinal_base=0):
        """Same call without the copy into a Python bytes object: returns a RecordBuffer that views the
        library-owned (page-locked) record buffer; release it with .close() or a `with` block."""
        b = mdf_text.encode() if isinstance(mdf_text, str) else mdf_text
        out, n, st = C.c_void_p(), C.c_size_t(), Stats()
        fmt = FASTQ if fastq else FASTA
        self._ck(self._L.tksm_seq_simulate_text(self._h, b, len(b), which, fmt, ordinal_base, C.byref(out), C.byref(n),
                                                C.byref(st)))
        self.last_stats = st.as_dict()
        return RecordBuffer(self._L, out, n.value)

    # ---- the __main__ loop (py/sequence.py:347-376)
    def run_files(self, mdf_path, badread=None, perfect=None):
        st = Stats()
        enc = lambda s: None if s is None else s.encode()
        self._ck(self._L.tksm_seq_run_files(self._h, mdf_path.encode(), enc(badread), enc(perfect), C.byref(st)))
        self.last_stats = st.as_dict()
        return self.last_stats

    # ---- multi-GPU sharding: contiguous shards of about equal simulated bases
    def shard_plan(self, mdf_text, world):
        """[(byte_begin, byte_end, ordinal_base, n_reads)] per rank; rank i simulates
        mdf_text[byte_begin:byte_end] with ordinal_base and the rank outputs concatenate, in rank
        order, to the bytes of a single-GPU run."""
        b = mdf_text.encode() if isinstance(mdf_text, str) else mdf_text
        bb = (C.c_uint64 * (world + 1))()
        ob = (C.c_uint64 * (world + 1))()
        self._ck(self._L.tksm_seq_shard_plan(self._h, b, len(b), world, bb, ob))
        return [(bb[i], bb[i + 1], ob[i], ob[i + 1] - ob[i]) for i in range(world)]

    # ---- device-resident batches (bench)
    def batch(self, mdf_text, ordinal_base=0):
        return Batch(self, mdf_text, ordinal_base)

    def synth_mdf(self, shape, n_molecules, seed=42, min_len=500, max_len=100000):
        out, n = C.c_void_p(), C.c_size_t()
        rc = self._L.tksm_seq_synth_mdf(self._h, shape, n_molecules, seed, min_len, max_len, C.byref(out), C.byref(n))
        if rc:
            raise SeqError("synth_mdf failed (no reference loaded?)")
        try:
            return C.string_at(out, n.value)
        finally:
            self._L.tksm_seq_free(out)


class RecordBuffer:
    """FASTQ/FASTA bytes owned by the library (tksm_seq_free releases them)."""

    def __init__(self, L, ptr, n):
        self._L, self._ptr, self.nbytes = L, ptr, n

    def array(self):
        return np.ctypeslib.as_array(C.cast(self._ptr, C.POINTER(C.c_uint8)), shape=(max(self.nbytes, 1),))[:self.nbytes]

    def tobytes(self):
        return C.string_at(self._ptr, self.nbytes)

    def close(self):
        if self._ptr:
            self._L.tksm_seq_free(self._ptr)
            self._ptr = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Batch:
    def __init__(self, seq, mdf_text, ordinal_base=0):
        self._s = seq
        b = mdf_text.encode() if isinstance(mdf_text, str) else bytes(mdf_text)
        self._b = seq._L.tksm_seq_batch_create(seq._h, b, len(b), ordinal_base)
        if not self._b:
            raise SeqError(seq._L.tksm_seq_last_error(seq._h).decode())

    def run(self, which=BADREAD, fmt=FASTQ):
        st = Stats()
        self._s._ck(self._s._L.tksm_seq_batch_run(self._s._h, self._b, which, fmt, C.byref(st)))
        return st.as_dict()

    def fetch(self, which=BADREAD):
        out, n = C.c_void_p(), C.c_size_t()
        self._s._ck(self._s._L.tksm_seq_batch_fetch(self._s._h, self._b, which, C.byref(out), C.byref(n)))
        try:
            return C.string_at(out, n.value)
        finally:
            self._s._L.tksm_seq_free(out)

    def close(self):
        if self._b:
            self._s._L.tksm_seq_batch_destroy(self._s._h, self._b)
            self._b = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def main(argv=None):
    """`tksm sequence` entry point: argv[0] == "sequence" (src/tksm.cpp:164-166)."""
    argv = list(sys.argv if argv is None else argv)
    if not argv or os.path.basename(argv[0]) != "sequence":
        argv = ["sequence"] + argv[1:]
    L = _lib.lib()
    arr = (C.c_char_p * len(argv))(*[a.encode() for a in argv])
    return L.tksm_sequence_main(len(argv), arr)


if __name__ == "__main__":
    sys.exit(main())
