/* tksm_seq.h — C ABI of the B200-native `tksm sequence` (Seq exit module) hot path.
 *
 * Drop-in boundary.  Every entry point names the reference interface it replaces
 * (paths relative to the vpc-ccg/tksm tree):
 *
 *   tksm_sequence_main          Sequencer_module{argc-1, argv+1}.run()   src/tksm.cpp:164-166,
 *                               src/sequence.cpp:30-54, src/python_runner.h:44-73 (argv[0] == "sequence")
 *   tksm_seq_create/_destroy    parse_args() state + Identities(...)      py/sequence.py:34-165, 331-336
 *   tksm_seq_load_references    get_reference_seqs()                      py/sequence.py:189-194
 *   tksm_seq_load_models        ErrorModel(...), QScoreModel(...),
 *                               KDE_noise_generator.load(...)             py/sequence.py:337-345
 *   tksm_seq_tail_draw          tail_model.noise_seq(len(fragment))       py/tksm_badread.py:335, 910-927
 *   tksm_seq_run_files          the __main__ loop                         py/sequence.py:347-376
 *   tksm_seq_simulate_text      mdf_generator + mdf_to_seq + formatter    py/sequence.py:197-320
 *
 * Conventions: plain pointers and sizes only; the caller owns every buffer it passes in;
 * buffers returned by the library are released with tksm_seq_free(); no exception crosses
 * the boundary; functions return 0 on success, 1 on a runtime/input error, 2 on a usage
 * error (same codes `tksm sequence` exits with) and tksm_seq_last_error() describes the
 * failure.  There is no CPU fallback: without a CUDA device tksm_seq_create() fails.
 */
#ifndef TKSM_SEQ_H
#define TKSM_SEQ_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct tksm_seq tksm_seq;
typedef struct tksm_seq_batch tksm_seq_batch;

typedef struct tksm_seq_config {
    int device;               /* CUDA device ordinal */
    uint64_t seed;            /* -s/--seed (src/module.h:88-91); default 42 */
    double identity_mean;     /* --badread-identity mean,max,stdev (percent); default 84,99,5.5 */
    double identity_max;
    double identity_stdev;
    int skip_qual_compute;    /* --skip-qual-compute */
    double calib_overshoot;  /* identity calibration: reference overshoot in errors per read, */
    double calib_kappa;      /* alignment explain-away coefficient; NaN => measured on the device when the models are loaded */
    uint64_t batch_reads;     /* streaming batch limits, 0 => defaults */
    uint64_t batch_bases;
    int host_only;            /* 1 => no device (table packing / parsing only; run calls fail) */
    int threads;              /* -t/--threads: host threads parsing MDF chunks; 0 => all hardware threads (max 32) */
} tksm_seq_config;

/* which outputs a run produces */
enum { TKSM_SEQ_BADREAD = 1, TKSM_SEQ_PERFECT = 2 };
/* record formats (py/sequence.py:291-300: chosen per output file by its extension);
 * bit 0 = badread output is FASTQ, bit 1 = perfect output is FASTQ */
enum { TKSM_SEQ_FASTA = 0, TKSM_SEQ_FASTQ_BADREAD = 1, TKSM_SEQ_FASTQ_PERFECT = 2, TKSM_SEQ_FASTQ = 3 };

typedef struct tksm_seq_stats {
    uint64_t reads;
    uint64_t raw_bases;        /* sum of error-free lengths */
    uint64_t read_bases;       /* sum of simulated read lengths (badread output) */
    uint64_t out_bytes[2];     /* [0] badread, [1] perfect */
    uint64_t algorithmic_bytes;/* inputs read + records written */
    uint64_t h2d_bytes, d2h_bytes;
    double kernel_ms;          /* CUDA-event time of the kernel sequence */
    double simulate_ms;        /* of which: simulate kernels */
    uint32_t kernel_launches;
    uint32_t status_or;        /* OR of per-read status words (0 = clean) */
} tksm_seq_stats;

/* host copies of the packed model tables, for the parity tests and the CPU twin */
typedef struct tksm_seq_tables {
    int k, has_err_table;
    const uint16_t* q16; size_t n_q16;
    const uint32_t* err_alias; size_t n_err_alias;
    const uint64_t* err_alts; size_t n_err_alts;
    int ident_const; float ident_mean;
    const float* ident_tab; size_t n_ident_tab;
    const float* delta_tab; size_t n_delta_tab;
    float calib_overshoot, calib_kappa;
    int kq;
    const uint32_t* q_direct; size_t n_q_direct; uint32_t q_direct_off[5];
    const uint64_t* q_hash; size_t n_q_hash;      /* entries of 2 x u64: key, row */
    uint32_t q_hash_shift, q_hash_mask;
    const uint32_t* q_rows; size_t n_q_rows;      /* pairs: alias offset, log2 columns */
    const uint32_t* q_alias; size_t n_q_alias;
    uint64_t seed;
    int compute_qual;
    const uint8_t* qcls; size_t n_qcls;           /* 4^k change-probability classes (0 entries without a table) */
    const double* qrep; size_t n_qrep;            /* 256 class representatives */
    const uint16_t* thr; size_t n_thr;            /* [64 delta bins][256 classes][8]: P(N >= m) * 65536 */
    float calib_kappa_measured;                   /* explain-away coefficient measured on the device when the models were
                                                     loaded (NaN: host-only handle, or calib_kappa given in the config) */
} tksm_seq_tables;

void tksm_seq_default_config(tksm_seq_config* cfg);
tksm_seq* tksm_seq_create(const tksm_seq_config* cfg);   /* NULL on failure, see tksm_seq_last_error(NULL) */
void tksm_seq_destroy(tksm_seq* h);
const char* tksm_seq_last_error(const tksm_seq* h);
void tksm_seq_free(void* p);

int tksm_seq_load_references(tksm_seq* h, const char* const* fasta_paths, int n_paths);
int tksm_seq_add_reference_text(tksm_seq* h, const char* fasta_text, size_t len);
/* synthetic reference generated on the device: contig i is named names[i] and holds lengths[i]
 * uniform ACGT bases with n_frac of them replaced by 'N' runs (bench helper) */
int tksm_seq_synth_reference(tksm_seq* h, const char* const* names, const uint64_t* lengths, int n_contigs,
                             double n_frac, uint64_t seed);
/* copies the upper-cased bytes of contig[start, start+len) into out (tests) */
int tksm_seq_fetch_reference(tksm_seq* h, const char* contig, uint64_t start, uint64_t len, char* out);
int tksm_seq_commit_reference(tksm_seq* h);   /* upload; implicit on first run */

/* names are looked up in $TKSM_MODELS (py/sequence.py:17-31) unless they are paths;
 * NULL selects the reference defaults (nanopore2020 if found else random; no_noise).
 * tail_model: "no_noise", a *.tail.gz name found in $TKSM_MODELS, or the path of a JSON file (plain or .gz)
 * as KDE_noise_generator.save writes it (py/tksm_badread.py:929-959) */
int tksm_seq_load_models(tksm_seq* h, const char* error_model, const char* qscore_model, const char* tail_model);
/* the tail-noise draw of one read (tail_model.noise_seq(len(fragment)), py/tksm_badread.py:335,910-927): the
 * bases appended to a molecule of frag_len bases for the read with this global ordinal; *len = 0 without a
 * tail; out is filled when *len <= cap.  Host only (tests, diagnostics). */
int tksm_seq_tail_draw(tksm_seq* h, uint64_t ordinal, uint32_t frag_len, char* out, size_t cap, uint32_t* len);
int tksm_seq_get_tables(tksm_seq* h, tksm_seq_tables* out);

/* Badread extras (north_star names them; the reference does not have them -- only the two unused constants of
 * py/tksm_badread.py:61-63 -- so there is no reference interface to cite and parity is unpinned; semantics after
 * `badread simulate`, defined in DESIGN.md 3.10).  All off by default and after tksm_seq_set_extras(h, NULL); with
 * every rate at zero the output is byte-identical to a run without the call.  Percentages are 0..100. */
typedef struct tksm_seq_extras {
    double junk_reads;             /* --junk_reads: percent of reads that are low-complexity junk of the molecule's length */
    double random_reads;           /* --random_reads: percent of reads that are random sequence of the molecule's length */
    double chimeras;               /* --chimeras: percent of reads that run on into the molecule of the read before them */
    double start_adapter_rate, start_adapter_amount;   /* --start_adapter rate,amount: percent of reads with the adapter; mean percent of it present */
    double end_adapter_rate, end_adapter_amount;       /* --end_adapter rate,amount */
    double glitch_rate, glitch_size, glitch_skip;      /* --glitches rate,size,skip: mean distance between glitches in bases (0: none), mean random bases inserted, mean bases lost */
    const char* start_adapter_seq; /* NULL: AATGTACTTCGTTCAGTTACGTATTGCT */
    const char* end_adapter_seq;   /* NULL: GCAATACGTAACTGAACGAAGT */
} tksm_seq_extras;
int tksm_seq_set_extras(tksm_seq* h, const tksm_seq_extras* x);

/* Fused front-end steps: the modules that usually run between the transcript sampler and Seq as
 * in-process edits of every MDF record, in the order given, before the record is sequenced:
 *   "plA:normal=15,7.5"  src/polyA.cpp:133-148 (normal | poisson | gamma | weibull = parameters[,min,max])
 *   "Flp:0.5"            src/strand_man.cpp:47-53, src/interval.h:908-920
 *   "Tag:FMT5,FMT3"      src/tag.cpp:84-114, src/util.h:53-92 (IUPAC formats, a number n = n times N, either may be empty)
 *   "SCB"                src/scb.cpp:73-81 (appends the first value of the record's CB= comment unless it is ".")
 * joined by ';', e.g. the README bulk pipeline "plA:normal=15,7.5;Flp:0.5;Tag:AATGTACTTCGTTCAGTTACGTATTGCT,GCAATACGTAACTGAACGAAGT".
 * The modules draw from std::mt19937; the fused steps draw from Philox keyed by (seed, ordinal of the record's first
 * read), so a fused run matches the module pipeline in distribution, not byte for byte.  NULL or "": none (default). */
int tksm_seq_set_front_end(tksm_seq* h, const char* spec);

/* MDF file (or FIFO) -> output files; NULL path => that output is not produced */
int tksm_seq_run_files(tksm_seq* h, const char* mdf_path, const char* badread_path, const char* perfect_path,
                       tksm_seq_stats* stats);

/* MDF text in host memory -> records in host memory (library-owned *out) */
int tksm_seq_simulate_text(tksm_seq* h, const char* mdf, size_t mdf_len, int which, int format, uint64_t ordinal_base,
                           char** out, size_t* out_len, tksm_seq_stats* stats);

/* device-resident path: parse + upload once, run the kernel sequence any number of times */
tksm_seq_batch* tksm_seq_batch_create(tksm_seq* h, const char* mdf, size_t mdf_len, uint64_t ordinal_base);
int tksm_seq_batch_run(tksm_seq* h, tksm_seq_batch* b, int which, int format, tksm_seq_stats* stats);
int tksm_seq_batch_fetch(tksm_seq* h, tksm_seq_batch* b, int which, char** out, size_t* out_len);
void tksm_seq_batch_destroy(tksm_seq* h, tksm_seq_batch* b);

/* synthetic MDF text in the emitter format of src/interval.h:898-905 (bench / tests);
 * shape: 0 = Tsb only (depth >= 1), 1 = Tsb->Trc->plA->Flp->Tag (README bulk), 2 = single long
 * intervals with log-uniform lengths in [min_len, max_len], 3 = scRNA Tsb->SCB->PCR (barcode, UMI and adapter
 * literals, ".cycle" ids, base mods on every interval kind) */
int tksm_seq_synth_mdf(tksm_seq* h, int shape, uint64_t n_molecules, uint64_t seed, uint32_t min_len,
                       uint32_t max_len, char** out, size_t* out_len);

/* Multi-GPU sharding (no reference counterpart: the reference forks --threads workers over
 * one stream, py/sequence.py:358-366).  Cuts the MDF text at record boundaries into `world`
 * contiguous shards of about equal simulated bases (sum over records of depth x interval lengths).
 * byte_begin[i] .. byte_begin[i+1] is the text of shard i and ordinal_base[i] the global ordinal of
 * its first read (after depth expansion), so that every rank keys Philox exactly as a single-GPU
 * run would; both arrays have world + 1 entries (the last ones: text length, total reads).
 * Host only: works on a host_only handle. */
int tksm_seq_shard_plan(tksm_seq* h, const char* mdf, size_t mdf_len, int world, uint64_t* byte_begin,
                        uint64_t* ordinal_base);

/* One process, several GPUs (`tksm sequence --gpus N`; no reference counterpart): handles[i] is
 * a handle on device i with the same references, models, seed and identity settings loaded.  The MDF (file or
 * FIFO, read to its end) is cut by the shard plan above, shard i runs on handles[i] in its own thread, and the
 * records are written in shard order, so the files hold the bytes of a one-GPU run. */
int tksm_seq_run_files_multi(tksm_seq* const* handles, int n_handles, const char* mdf_path, const char* badread_path,
                             const char* perfect_path, tksm_seq_stats* stats);

/* Record offsets without allocation.  This entry point has no counterpart in the reference; it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid it).  It is the record-offset scan of the formatting stage on its own:
 * d_rec_off[i] = d_rec_len[0] + ... + d_rec_len[i-1] (byte offset of record i in the output file) and
 * *d_total = the sum of all lengths.  d_rec_len (n_records x u32), d_rec_off (n_records x u64) and d_total (one u64)
 * are DEVICE pointers on `device`, owned by the caller; cuda_stream is a cudaStream_t (NULL: the default stream) and
 * the kernels are only enqueued on it: synchronise the stream before reading the results.  The call allocates
 * nothing (no cudaMalloc / cudaFree, no allocating library routine): the per-block sums it needs live in a
 * file-scope device array of 32 784 bytes that every GPU loading the library holds until the process exits, which
 * limits a call to TKSM_SEQ_STATIC_SCAN_MAX_RECORDS records (more: return code 2, tksm_seq_last_error(NULL) says
 * so).  All calls on a device share that array: calls that may overlap on the GPU (different streams, different
 * threads) race and must be serialised by the caller; calls in stream order are safe and nothing has to be cleared
 * between them.  The first call on a device looks up the array's address (cudaGetSymbolAddress); a call with
 * n_records = 0 does only that and writes *d_total = 0, so it can serve as set-up before a graph capture. */
#define TKSM_SEQ_STATIC_SCAN_MAX_RECORDS (1u << 24)
int tksm_seq_record_offsets_static(int device, const uint32_t* d_rec_len, uint32_t n_records, uint64_t* d_rec_off,
                                   uint64_t* d_total, void* cuda_stream);

/* the module entry point: argv[0] == "sequence" */
int tksm_sequence_main(int argc, char** argv);

#ifdef __cplusplus
}
#endif
#endif
