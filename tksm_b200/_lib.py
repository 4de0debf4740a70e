"""ctypes binding of libtksm_seq.so (include/tksm_seq.h).  No fallback: a missing or unloadable
library raises, and a missing CUDA device makes tksm_seq_create() fail unless host_only is set."""
import ctypes as C
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libtksm_seq.so")


class Config(C.Structure):
    _fields_ = [
        ("device", C.c_int),
        ("seed", C.c_uint64),
        ("identity_mean", C.c_double),
        ("identity_max", C.c_double),
        ("identity_stdev", C.c_double),
        ("skip_qual_compute", C.c_int),
        ("calib_overshoot", C.c_double),
        ("calib_kappa", C.c_double),
        ("batch_reads", C.c_uint64),
        ("batch_bases", C.c_uint64),
        ("host_only", C.c_int),
        ("threads", C.c_int),
    ]


class Extras(C.Structure):
    """tksm_seq_extras (include/tksm_seq.h): Badread extras, percentages 0..100, all off by default."""
    _fields_ = [
        ("junk_reads", C.c_double), ("random_reads", C.c_double), ("chimeras", C.c_double),
        ("start_adapter_rate", C.c_double), ("start_adapter_amount", C.c_double),
        ("end_adapter_rate", C.c_double), ("end_adapter_amount", C.c_double),
        ("glitch_rate", C.c_double), ("glitch_size", C.c_double), ("glitch_skip", C.c_double),
        ("start_adapter_seq", C.c_char_p), ("end_adapter_seq", C.c_char_p),
    ]


class Stats(C.Structure):
    _fields_ = [
        ("reads", C.c_uint64),
        ("raw_bases", C.c_uint64),
        ("read_bases", C.c_uint64),
        ("out_bytes", C.c_uint64 * 2),
        ("algorithmic_bytes", C.c_uint64),
        ("h2d_bytes", C.c_uint64),
        ("d2h_bytes", C.c_uint64),
        ("kernel_ms", C.c_double),
        ("simulate_ms", C.c_double),
        ("kernel_launches", C.c_uint32),
        ("status_or", C.c_uint32),
    ]

    def as_dict(self):
        d = {}
        for name, _ in self._fields_:
            v = getattr(self, name)
            d[name] = list(v) if name == "out_bytes" else v
        return d


class Tables(C.Structure):
    _fields_ = [
        ("k", C.c_int), ("has_err_table", C.c_int),
        ("q16", C.POINTER(C.c_uint16)), ("n_q16", C.c_size_t),
        ("err_alias", C.POINTER(C.c_uint32)), ("n_err_alias", C.c_size_t),
        ("err_alts", C.POINTER(C.c_uint64)), ("n_err_alts", C.c_size_t),
        ("ident_const", C.c_int), ("ident_mean", C.c_float),
        ("ident_tab", C.POINTER(C.c_float)), ("n_ident_tab", C.c_size_t),
        ("delta_tab", C.POINTER(C.c_float)), ("n_delta_tab", C.c_size_t),
        ("calib_overshoot", C.c_float), ("calib_kappa", C.c_float),
        ("kq", C.c_int),
        ("q_direct", C.POINTER(C.c_uint32)), ("n_q_direct", C.c_size_t), ("q_direct_off", C.c_uint32 * 5),
        ("q_hash", C.POINTER(C.c_uint64)), ("n_q_hash", C.c_size_t),
        ("q_hash_shift", C.c_uint32), ("q_hash_mask", C.c_uint32),
        ("q_rows", C.POINTER(C.c_uint32)), ("n_q_rows", C.c_size_t),
        ("q_alias", C.POINTER(C.c_uint32)), ("n_q_alias", C.c_size_t),
        ("seed", C.c_uint64),
        ("compute_qual", C.c_int),
        ("qcls", C.POINTER(C.c_uint8)), ("n_qcls", C.c_size_t),
        ("qrep", C.POINTER(C.c_double)), ("n_qrep", C.c_size_t),
        ("thr", C.POINTER(C.c_uint16)), ("n_thr", C.c_size_t),
        ("calib_kappa_measured", C.c_float),
    ]


BADREAD, PERFECT = 1, 2
FASTA, FASTQ_BADREAD, FASTQ_PERFECT, FASTQ = 0, 1, 2, 3

_lib = None

EXPORTS = [
    "tksm_seq_default_config", "tksm_seq_create", "tksm_seq_destroy", "tksm_seq_last_error", "tksm_seq_free",
    "tksm_seq_load_references", "tksm_seq_add_reference_text", "tksm_seq_synth_reference",
    "tksm_seq_fetch_reference", "tksm_seq_commit_reference", "tksm_seq_load_models", "tksm_seq_get_tables",
    "tksm_seq_run_files", "tksm_seq_simulate_text", "tksm_seq_batch_create", "tksm_seq_batch_run",
    "tksm_seq_batch_fetch", "tksm_seq_batch_destroy", "tksm_seq_synth_mdf", "tksm_seq_shard_plan", "tksm_seq_tail_draw", "tksm_seq_run_files_multi",
    "tksm_sequence_main", "tksm_seq_record_offsets_static", "tksm_seq_set_extras", "tksm_seq_set_front_end",
]


def build(force=False):
    """Compiles the library in-tree (nvcc, sm_100a).  Works without a GPU."""
    if force or not os.path.exists(LIB_PATH):
        subprocess.check_call(["make", "-C", os.path.join(_HERE, "csrc"), "-j8"])
    return LIB_PATH


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(there is no CPU fallback for the Seq hot path)")
    L = C.CDLL(LIB_PATH)
    vp, cp, u64, sz = C.c_void_p, C.c_char_p, C.c_uint64, C.c_size_t
    L.tksm_seq_default_config.argtypes = [C.POINTER(Config)]
    L.tksm_seq_default_config.restype = None
    L.tksm_seq_create.argtypes = [C.POINTER(Config)]
    L.tksm_seq_create.restype = vp
    L.tksm_seq_destroy.argtypes = [vp]
    L.tksm_seq_destroy.restype = None
    L.tksm_seq_last_error.argtypes = [vp]
    L.tksm_seq_last_error.restype = cp
    L.tksm_seq_free.argtypes = [vp]
    L.tksm_seq_free.restype = None
    L.tksm_seq_load_references.argtypes = [vp, C.POINTER(cp), C.c_int]
    L.tksm_seq_add_reference_text.argtypes = [vp, cp, sz]
    L.tksm_seq_synth_reference.argtypes = [vp, C.POINTER(cp), C.POINTER(u64), C.c_int, C.c_double, u64]
    L.tksm_seq_fetch_reference.argtypes = [vp, cp, u64, u64, C.c_char_p]
    L.tksm_seq_commit_reference.argtypes = [vp]
    L.tksm_seq_load_models.argtypes = [vp, cp, cp, cp]
    L.tksm_seq_get_tables.argtypes = [vp, C.POINTER(Tables)]
    L.tksm_seq_run_files.argtypes = [vp, cp, cp, cp, C.POINTER(Stats)]
    L.tksm_seq_simulate_text.argtypes = [vp, cp, sz, C.c_int, C.c_int, u64, C.POINTER(vp), C.POINTER(sz),
                                         C.POINTER(Stats)]
    L.tksm_seq_batch_create.argtypes = [vp, cp, sz, u64]
    L.tksm_seq_batch_create.restype = vp
    L.tksm_seq_batch_run.argtypes = [vp, vp, C.c_int, C.c_int, C.POINTER(Stats)]
    L.tksm_seq_batch_fetch.argtypes = [vp, vp, C.c_int, C.POINTER(vp), C.POINTER(sz)]
    L.tksm_seq_batch_destroy.argtypes = [vp, vp]
    L.tksm_seq_batch_destroy.restype = None
    L.tksm_seq_shard_plan.argtypes = [vp, cp, sz, C.c_int, C.POINTER(u64), C.POINTER(u64)]
    L.tksm_seq_run_files_multi.argtypes = [C.POINTER(vp), C.c_int, cp, cp, cp, C.POINTER(Stats)]
    L.tksm_seq_tail_draw.argtypes = [vp, u64, C.c_uint32, vp, sz, C.POINTER(C.c_uint32)]
    L.tksm_seq_synth_mdf.argtypes = [vp, C.c_int, u64, u64, C.c_uint32, C.c_uint32, C.POINTER(vp), C.POINTER(sz)]
    L.tksm_sequence_main.argtypes = [C.c_int, C.POINTER(cp)]
    L.tksm_seq_set_extras.argtypes = [vp, C.POINTER(Extras)]
    L.tksm_seq_set_front_end.argtypes = [vp, cp]
    L.tksm_seq_record_offsets_static.argtypes = [C.c_int, vp, C.c_uint32, vp, vp, vp]
    _lib = L
    return L
