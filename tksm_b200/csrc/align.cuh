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
// one machine word, horizontal input +1 at the band top).  Per column a cell {Dg, Pv, Mv, arr} is
// kept: Dg = "the diagonal move is optimal", Pv = "the move from above is optimal" (all the
// traceback needs), Mv to recognise a recomputed column that has come back to the known state,
// arr = row at which the last traceback first started a read base in this column.
// oracle/twin.c: twin_bnw is the plain integer form of the same definition; tests compare them.
//
// Parallelism: the columns of a read are cut into one chunk per thread of the team.  A chunk
// needs the band state at its left edge, which only the chunk before it can tell, so every thread
// first assumes the state of an undisturbed diagonal; chunks whose assumption turns out wrong (it
// is compared with what the neighbour actually produced) are redone, only as far as their state
// differs from what they had computed before, until nothing changes.  Lane 0 starts from the true
// boundary, so after r rounds the first r chunks are exact and the fixpoint is the sequential
// result.  The traceback is parallelised the same way from the other end (entry cell assumed on
// the script path, corrected by the neighbour's exit cell, redone until it meets its old path).
//
// The per-thread functions are plain C++ (no intrinsics) so that tests/ can compile them for the
// host and run the same chunking sequentially against the twin.
#pragma once
#include <cstddef>
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
constexpr uint32_t ALIGN_CELL_BYTES = 32; // trace bytes per column the caller provides (a cell of 64-bit words)
constexpr uint32_t ALIGN_WIDE64_FROM = 2048; // fragments with more slots start with the 64-bit band (no 32-bit attempt)
constexpr uint32_t ALIGN_WIDE_FROM = 8192; // ... and above this the team-parallel form of it (FIN_FULL kernels) does them
constexpr uint32_t ALIGN_ROW_BITS = 18;   // arr = walk tag << 18 | row; rows stay below 2^18 - 1 (reads of up to 2 x 131 070 bases)

enum : uint32_t { MV_DIAG = 0, MV_UP = 1, MV_LEFT = 2, MV_OUT = 3 };

template <typename BV>
struct alignas(4 * sizeof(BV)) AlignCell {
    BV dg, pv, mv, arr;
};

// the cell of a pass that needs neither rounds nor redone walks (align_kernel: one thread per read)
template <typename BV>
struct alignas(2 * sizeof(BV)) AlignCell2 {
    BV dg, pv;
};

template <typename BV, typename CELL = AlignCell<BV>>
struct AlignCtx {
    const uint8_t* frag;   // n fragment codes (0..3, or the raw byte)
    const uint8_t* slen;   // n slot lengths of the edit script
    const uint8_t* rb;     // m read codes
    uint8_t* out;          // op | dcount << 2 per read base (may be the array of the read codes)
    CELL* trace;           // cell of column j at [within * stride + owner], (owner, within) = divmod(j - 1, chunk)
    uint32_t n, m;
    uint32_t chunk;        // columns per thread
    uint32_t stride;       // threads of the team
    uint32_t cells;        // capacity of the trace array (cells)
    uint32_t* oob;         // set when an access would have left the trace array (never expected; the access is skipped)
};

// TK_ALIGN_PARANOID: every index is checked; the first violation is recorded in A.oob[1..4] (site, index, limit, j)
// and the access is redirected to element 0.
#ifdef TK_ALIGN_PARANOID
#define TK_AIDX(A, site, idx, lim, jj) \
    (((uint32_t)(idx) < (uint32_t)(lim)) ? (uint32_t)(idx) : ((A).oob[0] ? 0u : ((A).oob[0] = 1, (A).oob[1] = (site), (A).oob[2] = (uint32_t)(idx), (A).oob[3] = (uint32_t)(lim), (A).oob[4] = (uint32_t)(jj), 0u)))
#else
#define TK_AIDX(A, site, idx, lim, jj) (idx)
#endif

template <typename BV>
TK_HD uint32_t align_lo(uint32_t o) {
    constexpr uint32_t H = 4 * sizeof(BV);
    return o > H + 1 ? o - H : 1u;
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

// The read bases of the band rows as bit planes: bit b <-> row lo + b <-> read base lo + b - 1.
template <typename BV>
struct AlignPlanes {
    BV b0, b1, v, np;   // code bit 0, code bit 1, "is A/C/G/T", "is N"
    TK_HD void put(uint32_t c, uint32_t pos) {
        b0 |= BV(c & 1u) << pos;
        b1 |= BV((c >> 1) & 1u) << pos;
        v |= BV(c < 4u) << pos;
        np |= BV(c == 78u) << pos;
    }
};

// One thread's share of the forward pass: columns [c0, c1), ob read bases in front of column c0 - 1's slot.
template <typename BV>
struct AlignLane {
    uint32_t c0, c1, ob;
    AlignPlanes<BV> pl;   // planes of the band of column c0 - 1 (they depend on the read only, not on the round)
};

template <typename BV, typename CTX>
TK_HD void align_lane_init(const CTX& A, AlignLane<BV>& Ln) {
    constexpr uint32_t BW = 8 * sizeof(BV);
    const uint32_t lo = align_lo<BV>(Ln.ob);
    Ln.pl.b0 = Ln.pl.b1 = Ln.pl.v = Ln.pl.np = 0;
    for (uint32_t b = 0; b < BW; b++) {
        const uint32_t row = lo + b;
        if (row > A.m) break;
        Ln.pl.put(A.rb[TK_AIDX(A, 1, row - 1, A.m, Ln.c0)], b);
    }
}

// Columns [c0, c1) of the band.  In: (Pv, Mv) of column c0 - 1; out: the state of column c1 - 1.
// RECHECK: the columns have been computed before from another input state; returns true (and leaves
// Pv / Mv alone) as soon as a column ends in the state stored for it, since everything after it is
// then what it was.
template <typename BV, bool RECHECK, typename CTX>
TK_HD bool align_forward_chunk(const CTX& A, const AlignLane<BV>& Ln, BV& Pv_io, BV& Mv_io) {
    constexpr uint32_t BW = 8 * sizeof(BV);
    BV Pv = Pv_io, Mv = Mv_io;
    AlignPlanes<BV> pl = Ln.pl;
    uint32_t o_prev = Ln.ob;
    uint32_t lo_prev = align_lo<BV>(o_prev);
    auto* cell = A.trace + (Ln.c0 - 1) / A.chunk;   // first column of the chunk: within = 0, owner = the thread
    for (uint32_t j = Ln.c0; j < Ln.c1; j++, cell += A.stride) {
        if ((size_t)(cell - A.trace) >= A.cells) {
            *A.oob = 1;
            break;
        }
        BV old_pv = 0, old_mv = 0;
        if (RECHECK) {
            old_pv = cell->pv;
            old_mv = cell->mv;
        }
        const uint32_t o = o_prev + A.slen[TK_AIDX(A, 2, j - 1, A.n, j)];
        const uint32_t lo = align_lo<BV>(o);
        const uint32_t s = lo - lo_prev;
        if (s) {
            // the band moves down s rows: rows that enter at the bottom were outside the previous column
            Pv = (Pv >> s) | (~BV(0) << (BW - s));
            Mv >>= s;
            pl.b0 >>= s;
            pl.b1 >>= s;
            pl.v >>= s;
            pl.np >>= s;
            for (uint32_t t = 0; t < s; t++) {
                const uint32_t row = lo_prev + BW + t;
                if (row <= A.m) pl.put(A.rb[TK_AIDX(A, 3, row - 1, A.m, j)], BW - s + t);
            }
        } else if (lo > 1) {
            // same top row as the previous column: the cell above it was outside that column's band
            Pv &= ~BV(1);
            Mv |= BV(1);
        }
        const uint32_t f = A.frag[TK_AIDX(A, 4, j - 1, A.n, j)];
        BV Eq;
        if (f < 4) {
            const BV f0 = BV(0) - BV(f & 1u), f1 = BV(0) - BV(f >> 1);
            Eq = pl.v & ~((pl.b0 ^ f0) | (pl.b1 ^ f1));
        } else if (f == 78u) {
            Eq = pl.np;
        } else {
            Eq = 0;
            for (uint32_t b = 0; b < BW; b++) {
                const uint32_t row = lo + b;
                if (row > A.m) break;
                if (A.rb[TK_AIDX(A, 5, row - 1, A.m, j)] == f) Eq |= BV(1) << b;
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
        if (RECHECK) {
            cell->dg = Eq | ~D0;
            cell->pv = Pv;
            cell->mv = Mv;
            if (Pv == old_pv && Mv == old_mv) return true;
        } else {
            AlignCell<BV> c;
            c.dg = Eq | ~D0;
            c.pv = Pv;
            c.mv = Mv;
            c.arr = ~BV(0);   // no walk has been here (the trace array is reused from read to read)
            *cell = c;
        }
        o_prev = o;
        lo_prev = lo;
    }
    Pv_io = Pv;
    Mv_io = Mv;
    return false;
}

// Position of the traceback inside the trace array: column j, its cell index and (owner, within).
struct AlignCursor {
    uint32_t owner = 0, within = 0;
    size_t idx = 0;
    template <typename CTX>
    TK_HD void seek(const CTX& A, uint32_t j) {
        const uint32_t jj = j - 1;
        owner = jj / A.chunk;
        within = jj - owner * A.chunk;
        idx = (size_t)within * A.stride + owner;
    }
    template <typename CTX>
    TK_HD void left(const CTX& A) {
        if (within) {
            within--;
            idx -= A.stride;
        } else {
            owner--;
            within = A.chunk - 1;
            idx = (size_t)within * A.stride + owner;
        }
    }
};

enum : int { TRACE_FIRST = 0, TRACE_REDO = 1, TRACE_WRITE = 2 };

// Walks the path back from cell (i, j) (oj = o_j), which must be one the path leaves by consuming a read
// base (or the end cell (m, n) when `top`), while it is in columns >= cmin (to the origin when `bottom`),
// and returns the cell it hands to the thread below in (i, j, oj).
//   TRACE_FIRST  records in every column the row at which a read base was first started, tagged `tag`;
//   TRACE_REDO   the same from a corrected entry cell; stops with `merged` as soon as it starts a base where
//                a walk tagged >= emin did (the caller keeps emin = the oldest walk that ended in the exit
//                cell it holds): from there on it is that walk, and (i, j, oj) are left alone;
//   TRACE_WRITE  stores op | dcount << 2 over the read codes and counts matches and columns.
// Returns false when the band was left.
template <typename BV, int MODE, typename CTX>
TK_HD bool align_trace_chunk(const CTX& A, uint32_t cmin, bool bottom, bool top, uint32_t& i_io, uint32_t& j_io,
                             uint32_t& oj_io, uint32_t& eq, uint32_t& cols, bool& merged, uint32_t tag = 0,
                             uint32_t emin = 0) {
    constexpr uint32_t BW = 8 * sizeof(BV);
    uint32_t i = i_io, j = j_io, oj = oj_io;
    merged = false;
    (void)top;   // the end cell needs no special case: its D run is walked before the first read base starts
    AlignCursor cur;
    BV dg = 0, pv = 0;
    uint32_t lo = 1, arr_old = 0;
    bool newcol = true;
    auto enter = [&]() {   // (j, oj) changed: fetch the column's cell
        if (j > 0 && cur.idx >= A.cells) {
            *A.oob = 1;
            j = 0;
            i = 0;
        }
        if (j > 0) {
            const auto& c = A.trace[cur.idx];
            dg = c.dg;
            pv = c.pv;
            if constexpr (MODE == TRACE_REDO) arr_old = (uint32_t)c.arr;
            lo = align_lo<BV>(oj);
        }
        newcol = true;
    };
    if (j > 0) cur.seek(A, j);
    enter();
    bool have = false;
    uint32_t bi = 0, op = 0, run = 0;
    while (true) {
        uint32_t d;
        if (i == 0) {
            if (j == 0) break;
            d = MV_LEFT;
        } else if (j == 0) {
            d = MV_UP;
        } else {
            if (i < lo + (lo > 1 ? ALIGN_MARGIN : 0u) || i + ALIGN_MARGIN >= lo + BW) return false;
            const uint32_t b = i - lo;
            d = ((dg >> b) & 1) ? MV_DIAG : ((pv >> b) & 1) ? MV_UP : MV_LEFT;
        }
        if (d == MV_LEFT) {
            run++;
            oj -= A.slen[TK_AIDX(A, 6, j - 1, A.n, i)];
            j--;
            if (j > 0) cur.left(A);
            enter();
            continue;
        }
        // a read base starts here: the one before it is complete (its D run is known now)
        if (have) {
            if (MODE == TRACE_WRITE) {
                A.out[TK_AIDX(A, 7, bi, A.m, j)] = (uint8_t)(op | ((run > ALIGN_DC_CAP ? ALIGN_DC_CAP : run) << 2));
                eq += op == 0u;
            }
            cols += 1 + run;
        } else {
            cols += run;   // D run after the last read base (top), nothing otherwise
        }
        if (!(j >= cmin || bottom)) {
            i_io = i;
            j_io = j;
            oj_io = oj;
            return true;
        }
        if constexpr (MODE != TRACE_WRITE) if (newcol && j > 0) {
            if (MODE == TRACE_REDO && (arr_old & ((1u << ALIGN_ROW_BITS) - 1)) == i && (arr_old >> ALIGN_ROW_BITS) >= emin) {
                merged = true;
                return true;
            }
            A.trace[cur.idx].arr = (BV)((tag << ALIGN_ROW_BITS) | i);
            newcol = false;
        }
        bi = i - 1;
        have = true;
        run = 0;
        op = 2u;   // OP_I
        i--;
        if (d == MV_DIAG) {
            if (MODE == TRACE_WRITE) op = A.rb[TK_AIDX(A, 8, bi, A.m, j)] == A.frag[TK_AIDX(A, 9, j - 1, A.n, i)] ? 0u : 1u;   // OP_EQ : OP_X
            oj -= A.slen[TK_AIDX(A, 10, j - 1, A.n, i)];
            j--;
            if (j > 0) cur.left(A);
            enter();
        }
    }
    if (have) {
        if (MODE == TRACE_WRITE) {
            A.out[bi] = (uint8_t)(op | ((run > ALIGN_DC_CAP ? ALIGN_DC_CAP : run) << 2));
            eq += op == 0u;
        }
        cols += 1 + run;
    } else {
        cols += run;
    }
    i_io = 0;
    j_io = 0;
    oj_io = 0;
    return true;
}

// chunking of n columns over `threads` threads.  The state of a band of BW rows remembers about BW columns, and a chunk
// shorter than that rarely ends in a state that does not depend on the one it started from (every such chunk costs
// a round): chunks are at least 4 BW columns long.
TK_HD void align_chunking(uint32_t n, uint32_t threads, uint32_t band_bits, uint32_t& chunk, uint32_t& lanes) {
    chunk = (n + threads - 1) / threads;
    const uint32_t min_chunk = band_bits >= 64 ? 4 * band_bits : ALIGN_MIN_CHUNK;
    if (chunk < min_chunk) chunk = min_chunk;
    lanes = (n + chunk - 1) / chunk;
}

}  // namespace tk
