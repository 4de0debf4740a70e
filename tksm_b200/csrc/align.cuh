// Banded global alignment of a simulated read back to its fragment, around the edit script.
//
// The reference takes the CIGAR its qscore windows are cut from, and the read_identity= header
// token, from a global alignment of the finished read against the fragment
// (py/tksm_badread.py:613 edlib.align(seq, frag, task="path"); :244-257 identity_from_edlib_cigar).
// The edit script PB1 applied is one valid alignment; the optimal one (unit costs; ties broken as
// the edlib stand-in the fixtures were made with does: trace back from the end, diagonal first,
// then I, then D) is looked for in a band of BW rows around it:
//
//   DP cell (i, j): i read bases and j fragment slots consumed.  Column j holds the rows
//   lo_j .. lo_j + BW - 1, lo_j = max(1, o_j - BW / 2), o_j = read bases the script produced from
//   the first j slots.  Row 0 and column 0 are the usual boundaries.  Cells of the previous column
//   outside its band count as its edge cell plus one per row of distance; the cell above the band
//   top of the current column is no predecessor.  The traceback must stay ALIGN_MARGIN cells
//   inside the band, otherwise the alignment fails and the caller retries with 64-bit words.
//
// Columns are computed with Myers' bit-vector recurrence (vertical deltas Pv / Mv of the band in
// one machine word, horizontal input +1 at the band top); per column the word pair
// (Dg, Pv) is kept: Dg = "the diagonal move is optimal", Pv = "the move from above is optimal",
// which is all the traceback needs.  oracle/twin.c: twin_bnw is the plain integer form of the
// same definition; tests compare them cell path by cell path.
//
// Parallelism: the columns of a read are cut into one chunk per thread of the team.  A chunk
// needs the band state at its left edge, which only the chunk before it can tell, so every thread
// first assumes the state of an undisturbed diagonal, and chunks whose assumption turns out wrong
// (it is compared with what the neighbour actually produced) are redone until nothing changes:
// lane 0 starts from the true boundary, so after r rounds the first r chunks are exact and the
// fixpoint is the sequential result.  The traceback is parallelised the same way from the other
// end (entry cell assumed on the script path, corrected by the neighbour's exit cell).
//
// The per-thread functions are plain C++ (no intrinsics) so that tests/ can compile them for the
// host and run the same chunking sequentially against the twin.
#pragma once
#include <cstdint>

#ifndef TK_HD
#ifdef __CUDACC__
#define TK_HD __host__ __device__ __forceinline__
#else
#define TK_HD inline
#endif
#endif

namespace tk {

constexpr uint32_t ALIGN_MARGIN = 4;      // cells the traceback keeps away from the band edges
constexpr uint32_t ALIGN_MIN_CHUNK = 16;  // columns per thread at least
constexpr uint32_t ALIGN_DC_CAP = 15;     // D columns recorded in front of one read base
constexpr uint32_t ALIGN_WARMUP = 12;     // columns a thread runs ahead of its chunk to guess the state at its edge

enum : uint32_t { MV_DIAG = 0, MV_UP = 1, MV_LEFT = 2, MV_OUT = 3 };

template <typename BV>
struct AlignCtx {
    const uint8_t* frag;   // n fragment codes (0..3, or the raw byte)
    const uint8_t* slen;   // n slot lengths of the edit script
    uint8_t* rb;           // m read codes; the traceback replaces them by op | dcount << 2
    BV* trace;             // 2 words per column: [(within * stride + owner) * 2 + {0: Dg, 1: Pv}]
    uint32_t n, m;
    uint32_t chunk;        // columns per thread
    uint32_t stride;       // threads of the team
    uint32_t magic;        // floor(2^32 / chunk) + 1: (x * magic) >> 32 == x / chunk for x * chunk < 2^32
};

template <typename BV>
TK_HD uint32_t align_lo(uint32_t o) {
    constexpr uint32_t H = 4 * sizeof(BV);
    return o > H + 1 ? o - H : 1u;
}

template <typename BV>
TK_HD size_t align_trace_index(const AlignCtx<BV>& A, uint32_t j) {
    const uint32_t jj = j - 1;
    const uint32_t owner = (uint32_t)(((uint64_t)jj * A.magic) >> 32);
    const uint32_t within = jj - owner * A.chunk;
    return ((size_t)within * A.stride + owner) * 2;
}

// State of an undisturbed diagonal through (o, .): values grow by one per row away from it.
template <typename BV>
TK_HD void align_ideal_state(uint32_t o, BV& Pv, BV& Mv) {
    constexpr uint32_t BW = 8 * sizeof(BV);
    const uint32_t lo = align_lo<BV>(o);
    const uint32_t nm = o >= lo ? o - lo + 1 : 0;   // rows lo .. o step down towards the diagonal
    Mv = nm >= BW ? ~BV(0) : (BV(1) << nm) - 1;
    Pv = ~Mv;
}

// Columns [c0, c1) of the band.  In: (Pv, Mv) of column c0 - 1 (o_prev read bases in front of it);
// out: the state of column c1 - 1.  STORE: writes the (Dg, Pv) words of its columns.
template <typename BV, bool STORE = true>
TK_HD void align_forward_chunk(const AlignCtx<BV>& A, uint32_t c0, uint32_t c1, uint32_t o_prev, BV& Pv, BV& Mv) {
    constexpr uint32_t BW = 8 * sizeof(BV);
    uint32_t lo_prev = align_lo<BV>(o_prev);
    // match masks of the band rows for the four bases: bit b <-> row lo + b <-> read base lo + b - 1
    BV M0 = 0, M1 = 0, M2 = 0, M3 = 0;
    for (uint32_t b = 0; b < BW; b++) {
        const uint32_t row = lo_prev + b;
        if (row > A.m) break;
        const uint32_t c = A.rb[row - 1];
        const BV bit = BV(1) << b;
        M0 |= c == 0 ? bit : BV(0);
        M1 |= c == 1 ? bit : BV(0);
        M2 |= c == 2 ? bit : BV(0);
        M3 |= c == 3 ? bit : BV(0);
    }
    size_t ti = align_trace_index(A, c0);
    for (uint32_t j = c0; j < c1; j++) {
        const uint32_t o = o_prev + A.slen[j - 1];
        const uint32_t lo = align_lo<BV>(o);
        const uint32_t s = lo - lo_prev;
        if (s) {
            // the band moves down s rows: rows that enter at the bottom were outside the previous column
            const BV fill = ~BV(0) << (BW - s);
            Pv = (Pv >> s) | fill;
            Mv >>= s;
            M0 >>= s;
            M1 >>= s;
            M2 >>= s;
            M3 >>= s;
            for (uint32_t t = 0; t < s; t++) {
                const uint32_t row = lo_prev + BW + t;
                if (row <= A.m) {
                    const uint32_t c = A.rb[row - 1];
                    const BV bit = BV(1) << (BW - s + t);
                    M0 |= c == 0 ? bit : BV(0);
                    M1 |= c == 1 ? bit : BV(0);
                    M2 |= c == 2 ? bit : BV(0);
                    M3 |= c == 3 ? bit : BV(0);
                }
            }
        } else if (lo > 1) {
            // same top row as the previous column: the cell above it was outside that column's band
            Pv &= ~BV(1);
            Mv |= BV(1);
        }
        const uint32_t f = A.frag[j - 1];
        BV Eq;
        if (f < 4) {
            Eq = f == 0 ? M0 : f == 1 ? M1 : f == 2 ? M2 : M3;
        } else {
            Eq = 0;
            for (uint32_t b = 0; b < BW; b++) {
                const uint32_t row = lo + b;
                if (row > A.m) break;
                if (A.rb[row - 1] == f) Eq |= BV(1) << b;
            }
        }
        // Myers' column step with horizontal input +1 at the band top
        const BV Xv = Eq | Mv;
        const BV Xh = (((Eq & Pv) + Pv) ^ Pv) | Eq;
        BV Ph = Mv | ~(Xh | Pv);
        BV Mh = Pv & Xh;
        const BV D0 = Xh | Mv;
        Ph = (Ph << 1) | BV(1);
        Mh <<= 1;
        Pv = Mh | ~(Xv | Ph);
        Mv = Ph & Xv;
        if (STORE) {
            A.trace[ti] = Eq | ~D0;
            A.trace[ti + 1] = Pv;
            ti += (size_t)A.stride * 2;
        }
        o_prev = o;
        lo_prev = lo;
    }
}

template <typename BV>
TK_HD uint32_t align_decide(const AlignCtx<BV>& A, uint32_t i, uint32_t j, uint32_t oj) {
    constexpr uint32_t BW = 8 * sizeof(BV);
    if (i == 0) return MV_LEFT;
    if (j == 0) return MV_UP;
    const uint32_t lo = align_lo<BV>(oj);
    if (i < lo + (lo > 1 ? ALIGN_MARGIN : 0u) || i + ALIGN_MARGIN >= lo + BW) return MV_OUT;
    const uint32_t b = i - lo;
    const size_t ti = align_trace_index(A, j);
    if ((A.trace[ti] >> b) & 1) return MV_DIAG;
    if ((A.trace[ti + 1] >> b) & 1) return MV_UP;
    return MV_LEFT;
}

// Walks the path back from cell (i, j) (oj = o_j) while it is in columns >= cmin (to the origin when
// `bottom`).  `top`: (i, j) is the end cell (m, n), whose trailing D run is walked first.  A step
// consumes one read base and then the D run in front of it, so the cell handed to the next thread
// always continues with a read base.  WRITE: store op | dcount << 2 over the read codes and count
// matches; without it only the exit cell is computed.  Returns false when the band was left.
template <typename BV, bool WRITE>
TK_HD bool align_trace_chunk(const AlignCtx<BV>& A, uint32_t cmin, bool bottom, bool top, uint32_t& i, uint32_t& j,
                             uint32_t& oj, uint32_t& eq, uint32_t& cols) {
    if (top) {
        while (j > 0) {
            const uint32_t d = align_decide(A, i, j, oj);
            if (d == MV_OUT) return false;
            if (d != MV_LEFT) break;
            oj -= A.slen[j - 1];
            j--;
            cols++;
        }
    }
    while (i > 0 && (j >= cmin || bottom)) {
        const uint32_t d = align_decide(A, i, j, oj);
        if (d == MV_OUT) return false;
        if (d == MV_LEFT) {   // only from a mistaken entry cell of a round that will be redone
            oj -= A.slen[j - 1];
            j--;
            continue;
        }
        const uint32_t bi = i - 1;
        uint32_t op = 2u;   // OP_I
        if (d == MV_DIAG) {
            if (WRITE) op = A.rb[bi] == A.frag[j - 1] ? 0u : 1u;   // OP_EQ : OP_X
            oj -= A.slen[j - 1];
            j--;
        }
        i--;
        uint32_t run = 0;
        while (j > 0) {
            const uint32_t d2 = align_decide(A, i, j, oj);
            if (d2 == MV_OUT) return false;
            if (d2 != MV_LEFT) break;
            oj -= A.slen[j - 1];
            j--;
            run++;
        }
        if (WRITE) {
            A.rb[bi] = (uint8_t)(op | ((run > ALIGN_DC_CAP ? ALIGN_DC_CAP : run) << 2));
            eq += op == 0u;
        }
        cols += 1 + run;
    }
    return true;
}

// chunking of n columns over `threads` threads
TK_HD void align_chunking(uint32_t n, uint32_t threads, uint32_t& chunk, uint32_t& lanes, uint32_t& magic) {
    chunk = (n + threads - 1) / threads;
    if (chunk < ALIGN_MIN_CHUNK) chunk = ALIGN_MIN_CHUNK;
    lanes = (n + chunk - 1) / chunk;
    magic = (uint32_t)(0x100000000ull / chunk) + 1u;
}

}  // namespace tk
