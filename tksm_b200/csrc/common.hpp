// Shared host/device definitions for the tksm Seq hot path (B200 build).
//
// Everything in here is plain data: packed model tables, the molecule batch
// (structure-of-arrays) and the constants of the "PB1" stochastic model
// (Poissonised Badread).  The CPU twin in oracle/twin.c
// restates the same constants independently; tests assert both agree.
#pragma once
#include <cstdint>

namespace tk {

// ---------------------------------------------------------------- Philox tags
// counter = (c0 = index, c1 = tag word, c2 = read ordinal lo, c3 = read ordinal hi)
// key     = (seed lo, seed hi)
enum : uint32_t {
    TAG_READ = 0,    // c0 = 0: identity draw (w0); c0 = 1: uuid (w0..w3); c0 = 2: flanks (w0 left, w1 right)
    TAG_ARRIVE = 1,  // | epoch << 8 ; c0 = start position >> 3 ; 16-bit lane (position & 7) of the 128 bits
    TAG_TRIAL = 2,   // | epoch << 8 | trial << 20 ; c0 = start position
    TAG_QUAL = 3,    // c0 = untrimmed read base index >> 3 ; 16-bit lane (index & 7) of the 128 bits
    TAG_TAIL = 4,    // host only (tail.cpp): c0 = tail base index >> 2 ; word (index & 3). TAG_READ c0 = 3: tail yes/no, length, start state
};

constexpr int MAX_K = 7;             // error-model k-mer size supported (4^7 rows)
constexpr int ERR_ALIAS_COLS = 32;   // alias columns per k-mer row
constexpr int ERR_ALT_STRIDE = 32;   // alt words per k-mer row
constexpr uint32_t ERR_OUTCOME_RANDOM = 31;
constexpr int MAX_TRIALS_PER_EPOCH = 8;   // Poisson count cap per start position and epoch
constexpr int MAX_SLOT_LEN = 5;      // bases one fragment slot can hold after an edit
constexpr int IDENT_TAB_BINS = 4096;
constexpr int DELTA_TAB_BINS = 64;
constexpr int Q_CLASSES = 256;       // change-probability classes of the k-mers; class 255 = "not in the table" (q = 1)
constexpr uint32_t Q_CLASS_RANDOM = 255;
constexpr int THR_PER_CLASS = 8;     // P(N >= m), m = 1..8, as u16 thresholds (== MAX_TRIALS_PER_EPOCH)
constexpr float TIME_CAP = 100.0f;   // py/tksm_badread.py:354 (loop_count > 100 * frag_len)
constexpr uint32_t KEY_UNTOUCHED = 0xFFFFFFFFu;

// content16 (low half of a slot key)
//   [15:13] len (0..5)   [12:11] kind   [10:1] bases, 2 bits each, first base lowest   [0] zero
// kind 0: literal bases; kind 1: original byte followed by one inserted base;
// kind 2: one inserted base followed by the original byte.
constexpr uint32_t CONTENT_LEN_SHIFT = 13;
constexpr uint32_t CONTENT_KIND_SHIFT = 11;

// cigar symbols used by the qscore stage
enum : uint32_t { OP_EQ = 0, OP_X = 1, OP_I = 2, OP_D = 3 };

// ------------------------------------------------------------ qscore tables
struct QRow {
    uint32_t off;   // first alias column
    uint32_t lg;    // log2 of the column count
};

struct QHashEntry {
    uint64_t key;   // 0 = empty; else sentinel-1-prefixed 2-bit symbol string
    uint32_t row;
    uint32_t pad;
};

// ------------------------------------------------------------- device views
struct DevTables {
    // genome
    const uint32_t* genome;       // 2-bit codes, 16 bases per word, A=0 C=1 G=2 T=3
    const uint64_t* exc_start;    // sorted run starts (global base coordinate)
    const uint32_t* exc_len;
    const uint8_t* exc_chr;       // upper-cased byte of the run
    uint32_t n_exc;
    // error model
    int k;                        // k-mer size (1 for the "random" model)
    int has_err_table;            // 0 => every trial is a random change
    const uint8_t* qcls;          // 4^k: change-probability class of the k-mer
    const uint16_t* thr;          // [DELTA_TAB_BINS][Q_CLASSES][THR_PER_CLASS]: P(Poisson(q delta) >= m) * 65536
    const uint32_t* err_alias;    // 4^k * 32: thr22 | a << 22 | b << 27
    const uint64_t* err_alts;     // 4^k * 32 edit words
    // identity
    int ident_const;              // 1 => every read gets ident_mean
    float ident_mean;
    const float* ident_tab;       // IDENT_TAB_BINS + 1 quantiles (already times max identity)
    const float* delta_tab;       // DELTA_TAB_BINS epoch lengths over target in [0.5, 1]
    float calib_overshoot, calib_kappa;           // effective target: 1 - g, g - kappa g^2 = (1 - target) + overshoot / n
    // qscore model
    int kq;                       // odd, <= 9
    const uint32_t* q_direct;     // concatenated direct maps for widths 1,3,5,7,9 (3^w entries each)
    uint32_t q_direct_off[5];
    const QHashEntry* q_hash;
    uint32_t q_hash_shift;        // 64 - log2(size)
    uint32_t q_hash_mask;
    const QRow* q_rows;
    // fast path for 9-wide windows (kq == 9), entries = alias offset << 3 | log2 columns; 0 pointers when kq != 9
    const uint32_t* q_t9;         // 3^9: D-free windows, base-3 code of the ops (first op lowest)
    const uint32_t* q_t9d1;       // 3^9 * 24: one D run of 1..3 at gap g: code + 19683 * (3 g + dc - 1)
    const uint32_t* q_t9d2;       // 3^9 * 28: single D columns at gaps g1 < g2: code + 19683 * (g1 (15 - g1) / 2 + g2 - g1 - 1)
    const uint32_t* q_alias;      // thr16 | qa << 16 | qb << 23
    uint32_t seed_lo, seed_hi;
};

// seg_src layout
constexpr uint64_t SEG_LITERAL = 1ull << 63;
constexpr uint64_t SEG_MINUS = 1ull << 62;
constexpr uint64_t SEG_HAS_EXC = 1ull << 61;
constexpr uint64_t SEG_OFF_MASK = (1ull << 56) - 1;

// what the expansion of a read tells the alignment and qscore kernels
struct alignas(16) RInfo {
    uint32_t m;            // read length with flanks
    uint32_t start_trim, end_trim;
    uint32_t flags;        // RINFO_*
};

struct DevBatch {
    uint32_t n_reads;
    uint64_t ordinal_base;
    const uint32_t* read_mol;      // molecule index per read (depth expansion)
    const uint32_t* mol_seg_begin; // M + 1
    const uint32_t* mol_raw_len;   // M
    const uint32_t* mol_id_begin;  // M + 1, into id_pool
    const uint64_t* seg_src;
    const uint32_t* seg_len;
    const uint32_t* seg_mod_begin; // S + 1 or nullptr when the batch has no mods
    const uint32_t* mod_pos;
    const uint8_t* mod_chr;
    const uint8_t* lit_pool;       // literal segments as fragment codes
    const uint8_t* id_pool;
    // per-read scratch produced by the simulate kernel
    const uint64_t* scr_off;       // R + 1; read r owns [scr_off[r], scr_off[r+1]) in each arena
    uint8_t* scr_seq;
    uint8_t* scr_qual;
    uint8_t* scr_ops;              // read codes with flanks (expansion kernels); nullptr without qscores
    uint8_t* scr_cig;              // op | dcount << 2 per read base (align_kernel); nullptr without qscores
    RInfo* rinfo;                  // per read, written by the expansion kernels; nullptr without qscores
    uint8_t* scr_raw;              // error-free sequence (perfect output) or nullptr
    const uint64_t* raw_off;       // R + 1 when scr_raw != nullptr
    uint32_t* read_len;            // L'
    uint32_t* ident_num;           // matches
    uint32_t* ident_den;           // alignment columns
    uint32_t* status;              // 0 ok, else error bits
};

enum : uint32_t {
    ST_OK = 0,
    ST_OPS_OVERFLOW = 1,      // the read outgrew twice its fragment
    ST_SCRATCH_OVERFLOW = 2,  // the read outgrew its scratch slot
    ST_EPOCH_CAP = 4,         // time cap of the error loop (py/tksm_badread.py:354)
    ST_ALIGN_OOB = 16,        // internal error: the alignment would have left its trace array
    ST_BAND = 8,              // the alignment left its widest band: qscores follow the edit script itself
};
// ident_den carries this flag when (num, den) are the matches / columns of the read's alignment (printed as they
// are); without it they are the edit-script counts the calibration of DESIGN.md 3.5 maps to an aligned identity
constexpr uint32_t IDENT_MEASURED = 0x80000000u;

// run flags
enum : uint32_t {
    RUN_ERRORS = 1,     // inject errors (badread output requested)
    RUN_QUAL = 2,       // compute qscores
    RUN_RAW = 4,        // also store the error-free sequence
    RUN_NO_ALIGN = 8,   // experiments ($TKSM_SEQ_DEBUG bit 0): qscores from the edit script's own cigar, no alignment
    RUN_NO_WIDE = 16,   // experiments ($TKSM_SEQ_DEBUG bit 1): no retry with the 64-bit band
};

}  // namespace tk
