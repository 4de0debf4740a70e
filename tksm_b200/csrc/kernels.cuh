// sm_100a kernels of the Seq hot path.
//
//   simulate_kernel<W>   one team of W warps per read: materialise the molecule from the 2-bit
//                        genome (interval fetch, base mods, reverse-complement, literal segments;
//                        py/sequence.py:303-313), inject Badread-model errors with the PB1
//                        Philox-keyed restatement of sequence_fragment (py/tksm_badread.py:324-451),
//                        scan slot lengths into read offsets, sample CIGAR-window qscores
//                        (py/tksm_badread.py:584-655) and store read + quals in scratch.
//   reclen_kernel        FASTQ/FASTA record length per read (py/sequence.py:242-288).
//   scan kernels         device-wide exclusive scan of record lengths -> output offsets.
//   format_kernel        header text + sequence + quals into the contiguous output buffer.
//
// No tensor cores: nothing here is a contraction.  The kernels are byte/integer work staged in
// shared memory; DESIGN.md §4 has the per-kernel byte accounting.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>
#include <type_traits>

#include "align.cuh"
#include "common.hpp"

namespace tk {

// ------------------------------------------------------------------ Philox
__device__ __forceinline__ uint4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0,
                                               uint32_t k1) {
#pragma unroll
    for (int r = 0; r < 10; r++) {
        uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
        uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
        uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
        c0 = n0;
        c1 = lo1;
        c2 = n2;
        c3 = lo0;
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    return make_uint4(c0, c1, c2, c3);
}

__device__ __forceinline__ uint32_t word_of(const uint4& v, uint32_t i) {
    return i == 0 ? v.x : i == 1 ? v.y : i == 2 ? v.z : v.w;
}

// exp(-x) for x >= 0 from IEEE fma/mul only, so the CPU twin reproduces it bit for bit.
__device__ __forceinline__ float expneg(float x) {
    float y = __fmul_rn(x, 1.44269504f);
    if (y >= 120.0f) return 0.0f;
    int i = (int)y;
    float f = __fsub_rn(y, (float)i);
    // 2^-f = exp(-f ln2), Taylor to degree 9 on [0, 1)
    float t = __fmul_rn(f, -0.693147181f);
    float p = 2.75573192e-6f;
    p = __fmaf_rn(p, t, 2.48015873e-5f);
    p = __fmaf_rn(p, t, 1.98412698e-4f);
    p = __fmaf_rn(p, t, 1.38888889e-3f);
    p = __fmaf_rn(p, t, 8.33333333e-3f);
    p = __fmaf_rn(p, t, 4.16666667e-2f);
    p = __fmaf_rn(p, t, 1.66666667e-1f);
    p = __fmaf_rn(p, t, 0.5f);
    p = __fmaf_rn(p, t, 1.0f);
    p = __fmaf_rn(p, t, 1.0f);
    return __fmul_rn(p, __uint_as_float((uint32_t)(127 - i) << 23));
}

// ------------------------------------------------------------------- teams
template <int W>
struct Team {
    static constexpr int SIZE = W * 32;
    int rank;  // thread rank inside the team
    int id;    // team index inside the CTA
    int bar;   // named barrier of the team (W > 1): id + 1, or id when the CTA holds 16 teams
    __device__ __forceinline__ void sync() const {
        if (W == 1)
            __syncwarp();
        else
            asm volatile("bar.sync %0, %1;" ::"r"(bar), "n"(W * 32) : "memory");
    }
};

__device__ __forceinline__ uint32_t warp_incl_scan(uint32_t v) {
    const int lane = threadIdx.x & 31;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        uint32_t o = __shfl_up_sync(0xffffffffu, v, d);
        if (lane >= d) v += o;
    }
    return v;
}
__device__ __forceinline__ uint32_t warp_sum(uint32_t v) {
#pragma unroll
    for (int d = 16; d; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
    return v;
}

// Exclusive scan of one value per thread over the team; `total` gets the team sum.
// `wsum` is W + 1 words of team-private shared memory.
template <int W>
__device__ __forceinline__ uint32_t team_excl_scan(const Team<W>& tm, uint32_t v, uint32_t* wsum, uint32_t& total) {
    uint32_t inc = warp_incl_scan(v);
    if (W == 1) {
        total = __shfl_sync(0xffffffffu, inc, 31);
        return inc - v;
    }
    const int w = tm.rank >> 5;
    if ((tm.rank & 31) == 31) wsum[w] = inc;
    tm.sync();
    uint32_t base = 0, tot = 0;
#pragma unroll
    for (int i = 0; i < W; i++) {
        uint32_t s = wsum[i];
        if (i < w) base += s;
        tot += s;
    }
    tm.sync();
    total = tot;
    return base + inc - v;
}

template <int W>
__device__ __forceinline__ uint32_t team_sum(const Team<W>& tm, uint32_t v, uint32_t* wsum) {
    uint32_t s = warp_sum(v);
    if (W == 1) return s;
    if ((tm.rank & 31) == 0) wsum[tm.rank >> 5] = s;
    tm.sync();
    uint32_t tot = 0;
#pragma unroll
    for (int i = 0; i < W; i++) tot += wsum[i];
    tm.sync();
    return tot;
}

// --------------------------------------------------------------- content16
__device__ __forceinline__ uint32_t content_len(uint32_t c) { return (c >> CONTENT_LEN_SHIFT) & 7u; }
__device__ __forceinline__ uint32_t content_kind(uint32_t c) { return (c >> CONTENT_KIND_SHIFT) & 3u; }
__device__ __forceinline__ uint32_t content_base(uint32_t c, uint32_t i) { return (c >> (1 + 2 * i)) & 3u; }

// (match lost, inserted bases) of a changed slot whose original fragment code is b
__device__ __forceinline__ void slot_weights(uint32_t c, uint32_t b, uint32_t& loss, uint32_t& ins) {
    uint32_t len = content_len(c);
    if (len <= 1) {
        loss = 1;
        ins = 0;
        return;
    }
    ins = len - 1;
    if (content_kind(c) != 0) {
        loss = 0;
        return;
    }
    loss = !(b < 4 && (content_base(c, 0) == b || content_base(c, len - 1) == b));
}

__device__ __forceinline__ uint8_t code_ascii(uint32_t c) { return c < 4 ? (uint8_t)(0x54474341u >> (8 * c)) : (uint8_t)c; }
__device__ __forceinline__ uint32_t ascii_code(uint32_t ch) {
    return ch == 'A' ? 0u : ch == 'C' ? 1u : ch == 'G' ? 2u : ch == 'T' ? 3u : ch;
}

// crossing test: estimated identity (n - loss) / (n + ins) <= effective target (tq / 2^24)
__device__ __forceinline__ bool crossed(uint32_t n, uint32_t loss, uint32_t ins, uint32_t tq) {
    return ((uint64_t)(n - loss) << 24) <= (uint64_t)tq * (uint64_t)(n + ins);
}

constexpr int SIM_THREADS = 768;     // 24 warps per CTA, one CTA per SM
constexpr int SIM_WARPS = SIM_THREADS / 32;
constexpr uint32_t SIM_FIXED_SMEM = 64 + 512;  // CTA-wide: lut3[64] | alias row of the all-'=' window (128 cells)
constexpr uint32_t MISC_WORDS = 64;  // per-team scratch words
constexpr uint32_t MISC_WSUM = 8;    // W + 1 words of scan partials
constexpr uint32_t MISC_LISTN = 40;  // touched-list length (teams of more than one warp)
constexpr uint32_t MISC_ALIGN_OOB = 44;   // (+ 4 words of detail in TK_ALIGN_PARANOID builds)
// dummy to keep the numbering visible:   // set when the alignment would have left its trace array
constexpr uint32_t MISC_ALIGN_FLAG = 42;  // two words: "some thread's chunk-edge assumption was wrong" of the alignment rounds
constexpr uint32_t QUEUE_BYTES = 64 * 4 + 64 * 4;  // per-warp compaction queue: 64 x (u32, u32)

struct SimParams {
    DevTables T;
    DevBatch B;
    const uint32_t* work;   // read indices of this length class
    uint32_t n_work;
    uint32_t* counter;      // dynamic work counter
    uint32_t flags;         // RUN_*
    uint32_t ncap;          // fragment capacity of one team workspace (multiple of 16)
    uint32_t cap_list;      // touched-list capacity
    uint32_t slab_bytes;    // bytes of one team workspace
    uint8_t* gws;           // global workspaces (one per CTA) when GWS
    uint8_t* mid;           // PHASE 1 -> 2 records, 3 * scratch bytes + 64 per read
    uint8_t* trace;         // PHASE 2: alignment trace words, trace_bytes per team (global, L2-resident)
    uint64_t trace_bytes;
    uint32_t* dbg;          // TK_ALIGN_PARANOID builds: first recorded violation
    uint32_t* redo;         // PHASE 2: reads to be finished by the PHASE 3 kernel of the class (64-bit band)
    uint32_t* redo_count;
    const uint32_t* n_work_dev;  // launches on a device-side list: number of entries (device memory); nullptr: n_work
    uint32_t rb_mul;        // read codes a read may produce per fragment slot: 2, or GROW_RB_MUL in the larger slots of a second pass
    uint32_t* redo64;       // FIN_EXPAND: list of the 64-bit alignment instance (nullptr: that instance runs on all reads)
    uint32_t* redo64_count;
};

// Team workspace layout, shared by the host (sizing) and the kernel:
//   T1[256] u16 | hist[256] HistT | misc[64] u32 | W queues | list[cap_list] u32 | key[ncap] u32 | frag[ncap + 32] u8
struct SlabLayout {
    uint32_t off_hist, off_misc, off_queue, off_list, off_key, off_frag, bytes;
};
__host__ __device__ inline SlabLayout slab_layout(int W, uint32_t hist_elem_bytes, uint32_t ncap, uint32_t cap_list) {
    SlabLayout L;
    L.off_hist = 512u;
    L.off_misc = L.off_hist + 256 * hist_elem_bytes;
    L.off_queue = L.off_misc + MISC_WORDS * 4;
    L.off_list = L.off_queue + (uint32_t)W * QUEUE_BYTES;
    L.off_key = (L.off_list + 4 * cap_list + 15u) & ~15u;
    L.off_frag = L.off_key + 4 * ncap;
    L.bytes = (L.off_frag + ncap + 32 + 15u) & ~15u;
    return L;
}

// PHASE 2 (expansion, alignment, qscores) team workspace:
//   small part: misc[64] u32 | W x 512 bytes (warp compaction queues; the chunk-edge exchange of the alignment)
//   big part:   frag[ncap + 32] u8 | slen[ncap] u8 (slot lengths) | rb[2 ncap + 32] u8 (read codes, then cigar ops)
// The big part follows the small one in shared memory, or lives in global memory for the GWS class.
__host__ __device__ inline uint32_t slab2_small_bytes(int W) { return MISC_WORDS * 4 + (uint32_t)W * QUEUE_BYTES; }
__host__ __device__ inline uint32_t slab2_big_bytes(uint32_t ncap) { return (4 * ncap + 64 + 15u) & ~15u; }
// columns of alignment trace one team needs room for (a chunk per thread, at least ALIGN_MIN_CHUNK columns each)
__host__ __device__ inline uint64_t trace_team_bytes(int W, uint32_t ncap) {
    const uint64_t cols = (uint64_t)(ncap / (32u * W) + 1 > 256u ? ncap / (32u * W) + 1 : 256u) * 32u * W;   // align_chunking, 64-bit band
    return cols * ALIGN_CELL_BYTES;
}

// shared-memory bytes of the GWS class for the part of the workspace in front of the candidate list
__host__ __device__ inline uint32_t gws_small_bytes(int W) {
    return (slab_layout(W, 8, 16, 16).off_list + 15u) & ~15u;
}

// Finds the first of 256 bins at which the cumulative (loss, ins) crosses the target.
// Executed by the first warp of the team; results broadcast through misc[].
//   misc[0] = bin (256 when no crossing), misc[1] = cum loss before bin, misc[2] = cum ins before bin,
//   misc[3] = total loss, misc[4] = total ins
template <typename HistT>
__device__ __forceinline__ void hist_scan_warp(const HistT* hist, uint32_t* misc, uint32_t n, uint32_t loss0,
                                               uint32_t ins0, uint32_t tq) {
    const int lane = threadIdx.x & 31;
    uint32_t sa = 0, si = 0;
    uint32_t va[8], vi[8];
#pragma unroll
    for (int b = 0; b < 8; b++) {
        HistT h = hist[lane * 8 + b];
        if (sizeof(HistT) == 4) {
            va[b] = (uint32_t)h & 0xFFFFu;
            vi[b] = (uint32_t)h >> 16;
        } else {
            va[b] = (uint32_t)((uint64_t)h & 0xFFFFFFFFull);
            vi[b] = (uint32_t)((uint64_t)h >> 32);
        }
        sa += va[b];
        si += vi[b];
    }
    const uint32_t ea = warp_incl_scan(sa) - sa, ei = warp_incl_scan(si) - si;  // exclusive over lanes
    int first = 8;
    uint32_t ca = ea, ci = ei, fa = 0, fi = 0;
#pragma unroll
    for (int b = 0; b < 8; b++) {
        const uint32_t na = ca + va[b], ni = ci + vi[b];
        if (first == 8 && crossed(n, loss0 + na, ins0 + ni, tq)) {
            first = b;
            fa = ca;
            fi = ci;
        }
        ca = na;
        ci = ni;
    }
    const uint32_t ball = __ballot_sync(0xffffffffu, first < 8);
    const uint32_t tot_a = __shfl_sync(0xffffffffu, ea + sa, 31), tot_i = __shfl_sync(0xffffffffu, ei + si, 31);
    if (ball == 0) {
        if (lane == 0) {
            misc[0] = 256;
            misc[1] = misc[2] = 0;
            misc[3] = tot_a;
            misc[4] = tot_i;
        }
    } else {
        const int src = __ffs(ball) - 1;
        if (lane == src) {
            misc[0] = (uint32_t)(lane * 8 + first);
            misc[1] = fa;
            misc[2] = fi;
            misc[3] = tot_a;
            misc[4] = tot_i;
        }
    }
}

template <typename HistT>
__device__ __forceinline__ void hist_add(HistT* hist, uint32_t bin, uint32_t loss, uint32_t ins) {
    if (sizeof(HistT) == 4)
        atomicAdd((uint32_t*)&hist[bin], loss | (ins << 16));
    else
        atomicAdd((unsigned long long*)&hist[bin], (unsigned long long)loss | ((unsigned long long)ins << 32));
}

// ---------------------------------------------------------- qscore lookup
// ops byte: op | dcount << 2 ; stored in the upper halves of the slot keys
__device__ __forceinline__ uint32_t ops_at(const uint8_t* keyb, uint32_t j) { return keyb[4 * (j >> 1) + 2 + (j & 1)]; }
__device__ __forceinline__ void ops_put(uint8_t* keyb, uint32_t j, uint32_t v) { keyb[4 * (j >> 1) + 2 + (j & 1)] = (uint8_t)v; }

// py/tksm_badread.py:584-598 for the 9-column window of one read base held in registers:
//   ops18: op of column i at bits 2i; gaps: D count in front of column i (i >= 1) at bits 4(i-1);
//   t0: first trim level (the window is already shrunk at the read ends).
// Level t keeps columns t..8-t; a window without D goes to the direct table of its width, one with D
// to the hash under the packed key of model.cpp: qkey_from_cigar; a miss trims one column per side
// (leading/trailing D runs go with them, as cigar[1:-1].strip("D") does).
__device__ __forceinline__ uint32_t qscore_row(const DevTables& T, const uint8_t* lut3, uint32_t ops18, uint32_t gaps,
                                               uint32_t t0) {
    for (uint32_t t = t0;; t++) {
        const uint32_t w = 9u - 2u * t;
        const uint32_t o = (ops18 >> (2u * t)) & ((1u << (2u * w)) - 1u);
        const uint32_t gmask = w == 9u ? 0xFFFFFFFFu : ((1u << (4u * (w - 1u))) - 1u);
        const uint32_t g = (uint32_t)(((uint64_t)gaps >> (4u * t)) & gmask);
        if (g == 0u) {
            const uint32_t code3 = lut3[o & 63u] + 27u * lut3[(o >> 6) & 63u] + 729u * lut3[o >> 12];
            return T.q_direct[T.q_direct_off[w >> 1] + code3];
        }
        const uint64_t key = ((uint64_t)(0x10u | w) << 56) | ((uint64_t)o << 32) | g;
        uint32_t h = (uint32_t)((key * 0x9E3779B97F4A7C15ull) >> T.q_hash_shift);
        while (true) {
            const QHashEntry e = T.q_hash[h];
            if (e.key == key) return e.row;
            if (e.key == 0) break;
            h = (h + 1) & T.q_hash_mask;
        }
    }
}

// ------------------------------------------------------- warp compaction
// Rare, divergent work (a trial at a start position; a qscore window that misses the direct
// tables) is not done where it is found: the lanes that found some push it to a 64-entry ring in
// shared memory and the warp drains 32 entries at a time with every lane busy.
struct WarpQueue {
    uint32_t* a;
    uint32_t* b;
    uint32_t head, tail;  // warp-uniform
    __device__ __forceinline__ void push(bool pred, uint32_t va, uint32_t vb) {
        const uint32_t m = __ballot_sync(0xffffffffu, pred);
        if (pred) {
            const uint32_t slot = (tail + __popc(m & ((1u << (threadIdx.x & 31)) - 1))) & 63u;
            a[slot] = va;
            b[slot] = vb;
        }
        tail += __popc(m);
    }
    // number of entries the next drain handles: 32 when at least that many are queued, or the
    // remainder when `flush`
    __device__ __forceinline__ uint32_t ready(bool flush) const {
        const uint32_t cnt = tail - head;
        return cnt >= 32 ? 32u : (flush ? cnt : 0u);
    }
    __device__ __forceinline__ bool take(uint32_t cnt, uint32_t& va, uint32_t& vb) {
        const uint32_t lane = threadIdx.x & 31;
        const bool mine = lane < cnt;
        __syncwarp();
        if (mine) {
            va = a[(head + lane) & 63u];
            vb = b[(head + lane) & 63u];
        }
        __syncwarp();
        head += cnt;
        return mine;
    }
};

// four fragment bytes -> their 2-bit codes packed in 8 bits / a 4-bit mask of the bytes that are not ACGT codes
__device__ __forceinline__ uint32_t pack4(uint32_t x) {
    x &= 0x03030303u;
    x = (x | (x >> 6)) & 0x000F000Fu;
    return (x | (x >> 12)) & 0xFFu;
}
__device__ __forceinline__ uint32_t nz4(uint32_t x, uint32_t mask) {
    return ((__vcmpne4(x & mask, 0u) & 0x08040201u) * 0x01010101u) >> 24;
}

// ------------------------------------------------------------ simulate one
// PHASE 1 (inject_read): materialise + error injection, then slot keys and fragment go to the read's
// `mid` record in global memory.  PHASE 2 (finish_read): expansion, alignment, qscores.  Two kernels
// halve the hot code an SM has to keep in its instruction caches.
template <int W, bool GWS>
__device__ __forceinline__ void inject_read(const Team<W>& tm, const SimParams& P, uint32_t r, uint8_t* ws, uint8_t* wsl,
                                            const uint8_t* qcls_s) {
    // ws: team workspace (shared memory; global memory for the GWS class).  wsl: where its small, hot part lives
    // (threshold row, histogram, scratch words, compaction rings): == ws, or shared memory for the GWS class.
    constexpr int PHASE = 1;
    using HistT = typename std::conditional<(GWS || W >= 8), uint64_t, uint32_t>::type;
    constexpr int SIZE = Team<W>::SIZE;
    const DevTables& T = P.T;
    const DevBatch& B = P.B;
    const SlabLayout LY = slab_layout(W, (uint32_t)sizeof(HistT), P.ncap, P.cap_list);

    uint16_t* T1 = (uint16_t*)wsl;
    HistT* hist = (HistT*)(wsl + LY.off_hist);
    uint32_t* misc = (uint32_t*)(wsl + LY.off_misc);
    uint32_t* wsum = misc + MISC_WSUM;
    uint32_t* wq_a = (uint32_t*)(wsl + LY.off_queue + (uint32_t)(tm.rank >> 5) * QUEUE_BYTES);
    uint32_t* wq_b = wq_a + 64;
    uint32_t* list = (uint32_t*)(ws + LY.off_list);
    uint32_t* key = (uint32_t*)(ws + LY.off_key);
    uint8_t* keyb = (uint8_t*)key;
    uint8_t* frag = ws + LY.off_frag;
    const uint32_t* fragw = (const uint32_t*)frag;
    const int lane = tm.rank & 31;

    const uint32_t mol = B.read_mol[r];
    const uint32_t L = B.mol_raw_len[mol];
    const uint32_t k = (uint32_t)T.k;
    const uint32_t n = L + 2 * k;
    const uint64_t ordinal = B.ordinal_base + r;
    const uint32_t olo = (uint32_t)ordinal, ohi = (uint32_t)(ordinal >> 32);
    const uint32_t k0 = T.seed_lo, k1 = T.seed_hi;

    const uint64_t sbase = B.scr_off[r];
    const uint32_t scap = (uint32_t)(B.scr_off[r + 1] - sbase);
    uint32_t loss_c = 0, ins_c = 0;  // committed totals
    uint32_t status = ST_OK;
    // mid record of the read (PHASE 1 -> 2): keys (n rounded up to 4 slots) | fragment (n + 32 bytes);
    // 3 * scratch bytes + 64 per read always holds it (scratch is 2 L + 32 per read, rounded up to 16)
    uint4* mid4 = (uint4*)(P.mid + 3 * sbase + 64 * (uint64_t)r);
    const uint32_t key_q = (n + 3) >> 2, frag_q = (n + 32 + 15) >> 4;  // 16-byte units
    if (PHASE != 2) {
    // ---- 1. materialise flank + molecule + flank as fragment codes
    {
        uint4 fl = philox4x32_10(2, TAG_READ, olo, ohi, k0, k1);
        if (tm.rank < (int)k) {
            frag[tm.rank] = (uint8_t)((fl.x >> (2 * tm.rank)) & 3u);
            frag[k + L + tm.rank] = (uint8_t)((fl.y >> (2 * tm.rank)) & 3u);
        }
        if (tm.rank < 32) frag[n + tm.rank] = 0;
        const uint32_t s0 = B.mol_seg_begin[mol], s1 = B.mol_seg_begin[mol + 1];
        uint32_t off = k;
        for (uint32_t s = s0; s < s1; s++) {
            const uint64_t src = B.seg_src[s];
            const uint32_t len = B.seg_len[s];
            const bool minus = src & SEG_MINUS;
            const uint64_t so = src & SEG_OFF_MASK;
            if (src & SEG_LITERAL) {
                for (uint32_t x = tm.rank; x < len; x += SIZE) {
                    uint32_t c = B.lit_pool[so + (minus ? len - 1 - x : x)];
                    if (minus && c < 4) c ^= 3u;
                    frag[off + x] = (uint8_t)c;
                }
            } else {
                const bool exc = src & SEG_HAS_EXC;
                for (uint32_t x = tm.rank; x < len; x += SIZE) {
                    uint64_t g = so + (minus ? len - 1 - x : x);
                    uint32_t c = (__ldg(&T.genome[g >> 4]) >> (2 * (uint32_t)(g & 15))) & 3u;
                    if (minus) c ^= 3u;
                    if (exc) {
                        // last run starting at or before g
                        int lo = 0, hi = (int)T.n_exc;
                        while (lo < hi) {
                            int mid = (lo + hi) >> 1;
                            if (T.exc_start[mid] <= g) lo = mid + 1; else hi = mid;
                        }
                        if (lo > 0 && g < T.exc_start[lo - 1] + T.exc_len[lo - 1]) c = T.exc_chr[lo - 1];
                    }
                    frag[off + x] = (uint8_t)c;
                }
            }
            if (B.seg_mod_begin) {
                const uint32_t m0 = B.seg_mod_begin[s], m1 = B.seg_mod_begin[s + 1];
                if (m1 > m0) {
                    tm.sync();
                    for (uint32_t m = m0 + tm.rank; m < m1; m += SIZE) {
                        uint32_t pos = B.mod_pos[m];
                        uint32_t c = ascii_code(B.mod_chr[m]);
                        if (minus && c < 4) c ^= 3u;
                        frag[off + (minus ? len - 1 - pos : pos)] = (uint8_t)c;
                    }
                }
            }
            off += len;
        }
    }
    // key init (16 bytes per store; the array is a multiple of 16 slots long)
    {
        uint4* k4 = (uint4*)key;
        const uint32_t n4 = (n + 3) >> 2;
        for (uint32_t p = tm.rank; p < n4; p += SIZE) k4[p] = make_uint4(KEY_UNTOUCHED, KEY_UNTOUCHED, KEY_UNTOUCHED, KEY_UNTOUCHED);
    }
    tm.sync();

    if ((P.flags & RUN_RAW) && B.scr_raw) {
        uint8_t* raw = B.scr_raw + B.raw_off[r];
        for (uint32_t x = tm.rank; x < L; x += SIZE) raw[x] = code_ascii(frag[k + x]);
    }
    if (!(P.flags & RUN_ERRORS)) {
        if (tm.rank == 0) {
            B.read_len[r] = L;
            B.ident_num[r] = 1;
            B.ident_den[r] = 1;
            B.status[r] = ST_OK;
        }
        return;
    }

    // ---- 2. per-read target identity, effective target and epoch length class
    float target;
    if (T.ident_const) {
        target = T.ident_mean;
    } else {
        uint32_t u = philox4x32_10(0, TAG_READ, olo, ohi, k0, k1).x;
        uint32_t idx = u >> 20;
        float f = __fmul_rn((float)(u & 0xFFFFFu), 9.5367431640625e-07f);  // 2^-20
        float a = T.ident_tab[idx], b = T.ident_tab[idx + 1];
        target = __fmaf_rn(f, __fsub_rn(b, a), a);
    }
    // generative error g with g - kappa g^2 = (1 - target) + overshoot / n
    float teff;
    {
        float a = __fsub_rn(1.0f, target);
        if (n <= 1000) {  // py/tksm_badread.py:408; the overshoot grows with the error level up to 10 %
            float s = __fmul_rn(a, 10.0f);
            if (s > 1.0f) s = 1.0f;
            a = __fadd_rn(a, __fdiv_rn(__fmul_rn(T.calib_overshoot, s), (float)n));
        }
        if (T.calib_kappa > 0.0f) {
            float x = __fsub_rn(1.0f, __fmul_rn(__fmul_rn(4.0f, T.calib_kappa), a));
            if (!(x > 0.0f)) x = 0.0f;
            teff = __fsub_rn(1.0f, __fdiv_rn(__fsub_rn(1.0f, __fsqrt_rn(x)), __fmul_rn(2.0f, T.calib_kappa)));
        } else
            teff = __fsub_rn(1.0f, a);
    }
    uint32_t tq;
    if (!(teff > 0.0f))
        tq = 0;
    else if (teff >= 1.0f)
        tq = 1u << 24;
    else
        tq = (uint32_t)__fmul_rn(teff, 16777216.0f);
    int bin = (int)__fmul_rn(__fsub_rn(teff, 0.5f), 128.0f);
    bin = bin < 0 ? 0 : bin > DELTA_TAB_BINS - 1 ? DELTA_TAB_BINS - 1 : bin;
    const float delta = T.delta_tab[bin];
    const uint16_t* thr_bin = T.thr + (size_t)bin * Q_CLASSES * THR_PER_CLASS;
    for (uint32_t c = tm.rank; c < Q_CLASSES; c += SIZE) T1[c] = thr_bin[c * THR_PER_CLASS];
    if (tm.rank == 0) {
        misc[MISC_LISTN] = 0;
    }

    // ---- 3. error injection epochs
    uint32_t changed_c = 0;
    uint32_t list_n = 0;  // W == 1: the touched-list length lives in a (warp-uniform) register
    const uint32_t npos = n - k;
    const uint32_t vm0 = k >= 4 ? 0xFCFCFCFCu : (0xFCFCFCFCu >> (8 * (4 - k)));
    const uint32_t vm1 = k <= 4 ? 0u : (0xFCFCFCFCu >> (8 * (8 - k)));
    const uint32_t codemask = (1u << (2 * k)) - 1;
    const uint32_t kbits = (1u << k) - 1;
    const bool has_table = T.has_err_table != 0;

    // runs `f(p)` for every touched slot (every slot when the list overflowed)
    auto for_touched = [&](uint32_t cnt, auto&& f) {
        if (cnt <= P.cap_list) {
            for (uint32_t e = tm.rank; e < cnt; e += SIZE) f(list[e]);
        } else {
            for (uint32_t p = tm.rank; p < n; p += SIZE)
                if (key[p] != KEY_UNTOUCHED) f(p);
        }
    };

    if (!crossed(n, 0, 0, tq)) {
        for (uint32_t epoch = 0;; epoch++) {
            for (uint32_t b = tm.rank; b < 256; b += SIZE) hist[b] = 0;
            tm.sync();
            // --- arrivals: N ~ Poisson(q * delta) trials per start position, found ones are queued
            WarpQueue tq_{wq_a, wq_b, 0, 0};
            auto run_trials = [&](uint32_t cnt) {
                // an entry is ONE trial of a start position: (i, u16 | trial index << 16).  A position that drew more
                // than one trial puts the next one back into the queue (a loop up to the warp's largest count ran its
                // later rounds with one or two lanes)
                uint32_t i = 0, uw = 0;
                const bool mine = tq_.take(cnt, i, uw);
                const uint32_t u16 = uw & 0xFFFFu;
                const int tr = (int)(uw >> 16);
                uint32_t x0 = 0, x1 = 0, code = 0;
                bool valid = false;
                int N = 0;
                if (mine) {
                    const uint32_t qw = i >> 2, sh = 8 * (i & 3u);
                    x0 = __funnelshift_r(fragw[qw], fragw[qw + 1], sh);
                    x1 = __funnelshift_r(fragw[qw + 1], fragw[qw + 2], sh);
                    valid = has_table && (((x0 & vm0) | (x1 & vm1)) == 0);
                    uint32_t cls = Q_CLASS_RANDOM;
                    if (valid) {
                        code = (pack4(x0) | (pack4(x1) << 8)) & codemask;
                        cls = qcls_s[code];
                    }
                    const uint4 th = __ldg((const uint4*)(thr_bin + cls * THR_PER_CLASS));
                    const uint32_t uu = u16 | (u16 << 16);
                    // halves of th hold P(N >= m): count those above the uniform
                    N = __popc(__vcmpgtu2(th.x, uu) & 0x00010001u) + __popc(__vcmpgtu2(th.y, uu) & 0x00010001u) +
                        __popc(__vcmpgtu2(th.z, uu) & 0x00010001u) + __popc(__vcmpgtu2(th.w, uu) & 0x00010001u);
                }
                {
                    // per lane: up to k slot updates (cm = mask of slots to update)
                    uint32_t cm = 0, tm16 = 0, rc_content = 0;
                    uint64_t aw = 0;
                    bool table_alt = false;
                    if (tr < N) {
                        const uint4 wr = philox4x32_10(i, TAG_TRIAL | (epoch << 8) | ((uint32_t)tr << 20), olo, ohi, k0, k1);
                        tm16 = wr.x >> 16;
                        tm16 += tm16 == 0;
                        uint32_t outcome = ERR_OUTCOME_RANDOM;
                        if (valid) {
                            uint32_t e = __ldg(&T.err_alias[code * ERR_ALIAS_COLS + (wr.y >> 27)]);
                            uint32_t fr = (wr.y >> 5) & 0x3FFFFFu;
                            outcome = fr < (e & 0x3FFFFFu) ? ((e >> 22) & 31u) : (e >> 27);
                        }
                        if (outcome != ERR_OUTCOME_RANDOM) {
                            aw = __ldg(&T.err_alts[code * ERR_ALT_STRIDE + outcome]);
                            cm = (uint32_t)(aw >> 21) & 0x7Fu;
                            table_alt = true;
                        } else {
                            // py/tksm_badread.py:199-213
                            uint32_t type = __umulhi(wr.z, 3u);
                            uint32_t pos = __umulhi(wr.w, k);
                            uint32_t r2 = wr.z & 0xFFFFu;
                            uint32_t ob = (pos < 4 ? (x0 >> (8 * pos)) : (x1 >> (8 * (pos - 4)))) & 0xFFu;
                            if (type == 0) {
                                uint32_t nb = ob < 4 ? ((ob + 1 + ((r2 * 3u) >> 16)) & 3u) : (r2 & 3u);
                                rc_content = (1u << CONTENT_LEN_SHIFT) | (nb << 1);
                            } else if (type == 1) {
                                uint32_t nb = r2 & 3u;
                                uint32_t before = (r2 >> 2) & 1u;
                                rc_content = (2u << CONTENT_LEN_SHIFT) | ((before ? 2u : 1u) << CONTENT_KIND_SHIFT) | (nb << 1);
                            } else {
                                rc_content = 0;
                            }
                            cm = 1u << pos;
                        }
                    }
                    while (__any_sync(0xffffffffu, cm != 0)) {
                        bool fresh = false;
                        uint32_t p = 0;
                        if (cm) {
                            const uint32_t j = __ffs(cm) - 1;
                            cm &= cm - 1;
                            uint32_t content = rc_content;
                            if (table_alt) {
                                const uint32_t lens = (uint32_t)aw & 0x1FFFFFu;
                                const uint32_t len = (lens >> (3 * j)) & 7u;
                                // bases before slot j: sum of the 3-bit length fields below it
                                uint32_t below = lens & ((1u << (3 * j)) - 1);
                                uint32_t t2 = (below & 0x1C71C7u) + ((below >> 3) & 0x1C71C7u);
                                const uint32_t nb = ((t2 * 0x41041u) >> 18) & 0x3Fu;
                                const uint32_t bases = (uint32_t)(aw >> (28 + 2 * nb)) & ((1u << (2 * len)) - 1);
                                content = (len << CONTENT_LEN_SHIFT) | (bases << 1);
                            }
                            p = i + j;
                            const uint32_t old = atomicMin(&key[p], (tm16 << 16) | content);
                            fresh = old == KEY_UNTOUCHED;
                        }
                        // first touch of a slot: remember it
                        const uint32_t fm = __ballot_sync(0xffffffffu, fresh);
                        if (fm) {
                            uint32_t base;
                            if (W == 1) {
                                base = list_n;
                                list_n += __popc(fm);
                            } else {
                                if (lane == 0) base = atomicAdd(&misc[MISC_LISTN], (uint32_t)__popc(fm));
                                base = __shfl_sync(0xffffffffu, base, 0);
                            }
                            if (fresh) {
                                const uint32_t e = base + __popc(fm & ((1u << lane) - 1));
                                if (e < P.cap_list) list[e] = p;
                            }
                        }
                    }
                }
                tq_.push(mine && tr + 1 < N, i, u16 | ((uint32_t)(tr + 1) << 16));
            };
            const uint32_t nblk = (npos + 7) >> 3;
            const uint32_t nblk_r = (nblk + 31u) & ~31u;  // whole warps iterate together
            for (uint32_t blk = tm.rank; blk < nblk_r; blk += SIZE) {
                const bool live = blk < nblk;
                uint32_t codes = 0, bad = 0xFFFFu;
                uint4 ur = make_uint4(0, 0, 0, 0);
                if (live) {
                    const uint2* fp = (const uint2*)(frag + 8 * blk);
                    const uint2 fa = fp[0], fb = fp[1];
                    if (has_table) {
                        codes = pack4(fa.x) | (pack4(fa.y) << 8) | (pack4(fb.x) << 16) | (pack4(fb.y) << 24);
                        bad = nz4(fa.x, 0xFCFCFCFCu) | (nz4(fa.y, 0xFCFCFCFCu) << 4) | (nz4(fb.x, 0xFCFCFCFCu) << 8) |
                              (nz4(fb.y, 0xFCFCFCFCu) << 12);
                    }
                    ur = philox4x32_10(blk, TAG_ARRIVE | (epoch << 8), olo, ohi, k0, k1);
                }
#pragma unroll 1
                for (uint32_t t = 0; t < 8; t++) {
                    const uint32_t i = 8 * blk + t;
                    const uint32_t wsel = t < 4 ? (t < 2 ? ur.x : ur.y) : (t < 6 ? ur.z : ur.w);
                    const uint32_t u16 = (wsel >> (16 * (t & 1u))) & 0xFFFFu;
                    bool hit = false;
                    if (live && i < npos) {
                        const uint32_t cls = ((bad >> t) & kbits) ? Q_CLASS_RANDOM : qcls_s[(codes >> (2 * t)) & codemask];
                        hit = u16 < T1[cls];
                    }
                    tq_.push(hit, i, u16);
                    // (a drain may put up to 32 further trials back: drain until fewer than 32 wait, so that the next
                    // 32 arrivals still fit the ring of 64)
                    for (uint32_t cnt; (cnt = tq_.ready(false)) != 0;) run_trials(cnt);
                }
            }
            for (uint32_t cnt; (cnt = tq_.ready(true)) != 0;) run_trials(cnt);
            tm.sync();
            const uint32_t ncand = W == 1 ? list_n : misc[MISC_LISTN];
            // histogram of candidate change times, high byte
            for_touched(ncand, [&](uint32_t p) {
                const uint32_t kv = key[p];
                if ((kv >> 16) != 0) {
                    uint32_t a, in;
                    slot_weights(kv & 0xFFFFu, frag[p], a, in);
                    hist_add<HistT>(hist, kv >> 24, a, in);
                }
            });
            tm.sync();
            if (tm.rank < 32) hist_scan_warp<HistT>(hist, misc, n, loss_c, ins_c, tq);
            tm.sync();
            const uint32_t hbin = misc[0], pa = misc[1], pi = misc[2], ta = misc[3], ti = misc[4];
            if (hbin < 256) {
                // refine inside the crossing bin on the low time byte
                tm.sync();
                for (uint32_t b = tm.rank; b < 256; b += SIZE) hist[b] = 0;
                tm.sync();
                for_touched(ncand, [&](uint32_t p) {
                    const uint32_t kv = key[p];
                    if ((kv >> 24) == hbin && (kv >> 16) != 0) {
                        uint32_t a, in;
                        slot_weights(kv & 0xFFFFu, frag[p], a, in);
                        hist_add<HistT>(hist, (kv >> 16) & 255u, a, in);
                    }
                });
                tm.sync();
                if (tm.rank < 32) hist_scan_warp<HistT>(hist, misc, n, loss_c + pa, ins_c + pi, tq);
                tm.sync();
                const uint32_t lo = misc[0];
                const uint32_t tstar = (hbin << 8) | (lo < 256 ? lo : 255u);
                tm.sync();
                // drop later candidates, total up the kept ones
                uint32_t la = 0, li = 0;
                for_touched(ncand, [&](uint32_t p) {
                    const uint32_t kv = key[p];
                    if ((kv >> 16) != 0) {
                        if ((kv >> 16) > tstar) {
                            key[p] = KEY_UNTOUCHED;
                        } else {
                            uint32_t a, in;
                            slot_weights(kv & 0xFFFFu, frag[p], a, in);
                            la += a;
                            li += in;
                        }
                    }
                });
                loss_c += team_sum<W>(tm, la, wsum);
                ins_c += team_sum<W>(tm, li, wsum);
                break;
            }
            // no crossing in this epoch: commit every candidate and go on
            uint32_t cnt = 0;
            for_touched(ncand, [&](uint32_t p) {
                const uint32_t kv = key[p];
                if ((kv >> 16) != 0) {
                    key[p] = kv & 0xFFFFu;
                    cnt++;
                }
            });
            loss_c += ta;
            ins_c += ti;
            changed_c += team_sum<W>(tm, cnt, wsum);
            tm.sync();
            // py/tksm_badread.py:354-361
            if ((uint64_t)changed_c * 10 > (uint64_t)n * 9) break;
            if (__fmul_rn((float)(epoch + 1), delta) > TIME_CAP) {
                status |= ST_EPOCH_CAP;
                break;
            }
        }
    }
    tm.sync();
    }  // PHASE != 2
    if (PHASE == 1) {
        const uint4* k4 = (const uint4*)key;
        const uint4* f4 = (const uint4*)frag;
        for (uint32_t x = tm.rank; x < key_q; x += SIZE) mid4[x] = k4[x];
        for (uint32_t x = tm.rank; x < frag_q; x += SIZE) mid4[key_q + x] = f4[x];
        if (tm.rank == 0) {
            B.ident_num[r] = n - loss_c;
            B.ident_den[r] = n + ins_c;
            B.status[r] = status;
        }
        tm.sync();
        return;
    }
}

// ---- team-parallel alignment (align.cuh has the per-thread pieces and the definition)
// One chunk of columns per thread; chunk-edge states (forward) and entry cells (traceback) are
// assumed, compared with what the neighbour produced and redone until they agree.  xch: one 16-byte
// cell per thread.  Returns false when the traceback left the band (rb is then untouched).
template <int W, typename BV>
__device__ __forceinline__ bool align_team_impl(const Team<W>& tm, AlignCtx<BV>& A, uint4* xch, uint32_t* misc, uint32_t* wsum,
                                                uint32_t& eq_out, uint32_t& cols_out) {
    uint32_t lanes;
    align_chunking(A.n, (uint32_t)Team<W>::SIZE, 8u * (uint32_t)sizeof(BV), A.chunk, lanes);
    A.stride = (uint32_t)Team<W>::SIZE;
    const uint32_t l = (uint32_t)tm.rank;
    const bool active = l < lanes;
    AlignLane<BV> Ln;
    Ln.c0 = 1 + l * A.chunk;
    Ln.c1 = min(A.n + 1, Ln.c0 + A.chunk);
    uint32_t mine = 0;
    if (active)
        for (uint32_t j = Ln.c0; j < Ln.c1; j++) mine += A.slen[j - 1];
    uint32_t tot;
    Ln.ob = team_excl_scan<W>(tm, mine, wsum, tot);   // read bases in front of the chunk
    if (active) align_lane_init(A, Ln);
    // agreement flags of the rounds (teams of more than one warp): two words, used alternately
    uint32_t* flag = misc + MISC_ALIGN_FLAG;
    if (W > 1) {
        if (tm.rank == 0) flag[0] = flag[1] = 0;
        tm.sync();
    }
    uint32_t round = 0;
    auto team_any = [&](bool pred) -> bool {
        bool any;
        if (W == 1) {
            any = __any_sync(0xffffffffu, pred);
        } else {
            if (pred) flag[round & 1u] = 1;
            tm.sync();
            any = flag[round & 1u] != 0;
            if (tm.rank == 0) flag[(round + 1) & 1u] = 0;
        }
        round++;
        return any;
    };
    // ---- forward
    {
        BV P, M, oP = 0, oM = 0;
        if (l == 0) {
            P = ~BV(0);
            M = 0;
        } else {
            align_ideal_state<BV>(Ln.ob, P, M);
        }
        if (active) {
            oP = P;
            oM = M;
            align_forward_chunk<BV, false>(A, Ln, oP, oM);
        }
        while (true) {
            xch[l] = make_uint4((uint32_t)oP, (uint32_t)((uint64_t)oP >> 32), (uint32_t)oM, (uint32_t)((uint64_t)oM >> 32));
            tm.sync();
            bool ch = false;
            if (active && l > 0) {
                const uint4 v = xch[l - 1];
                const BV nP = (BV)(((uint64_t)v.y << 32) | v.x), nM = (BV)(((uint64_t)v.w << 32) | v.z);
                ch = nP != P || nM != M;
                P = nP;
                M = nM;
            }
            if (!team_any(ch)) break;
            if (ch) {
                BV tP = P, tM = M;
                if (!align_forward_chunk<BV, true>(A, Ln, tP, tM)) {
                    oP = tP;
                    oM = tM;
                }
            }
        }
    }
    __threadfence_block();
    tm.sync();
    // ---- traceback: exit cells until they agree, then the writing pass
    uint32_t ei, ej, eo;
    const bool top = l + 1 == lanes;
    if (top) {
        ei = A.m;
        ej = A.n;
        eo = A.m;
    } else {
        ei = Ln.ob + mine;
        ej = Ln.c1 - 1;
        eo = Ln.ob + mine;
    }
    bool ok = true;
    {
        uint32_t xi = ei, xj = ej, xo = eo, tag = 1, emin = 1;
        if (active) {
            uint32_t e = 0, c = 0;
            bool merged;
            ok = align_trace_chunk<BV, TRACE_FIRST>(A, Ln.c0, l == 0, top, xi, xj, xo, e, c, merged, 1u, 1u);
        }
        while (true) {
            xch[l] = make_uint4(xi, xj, xo, 0u);
            tm.sync();
            bool ch = false;
            if (active && !top) {
                const uint4 v = xch[l + 1];
                ch = v.x != ei || v.y != ej || v.z != eo;
                ei = v.x;
                ej = v.y;
                eo = v.z;
            }
            if (!team_any(ch)) break;
            if (ch) {
                uint32_t ti = ei, tj = ej, to = eo, e = 0, c = 0;
                bool merged;
                tag++;
                const bool r = align_trace_chunk<BV, TRACE_REDO>(A, Ln.c0, l == 0, top, ti, tj, to, e, c, merged, tag, emin);
                if (!merged) {
                    ok = r;
                    xi = ti;
                    xj = tj;
                    xo = to;
                    emin = tag;
                }
            }
        }
    }
    if (team_any(active && !ok)) return false;
    uint32_t eq = 0, cols = 0;
    if (active) {
        bool merged;
        align_trace_chunk<BV, TRACE_WRITE>(A, Ln.c0, l == 0, top, ei, ej, eo, eq, cols, merged);
    }
    eq_out = team_sum<W>(tm, eq, wsum);
    cols_out = team_sum<W>(tm, cols, wsum);
    return true;
}

// ------------------------------------------------------------- PHASE 2
// Expansion of the slot contents into the read, alignment of the read back to the fragment
// (py/tksm_badread.py:613), qscores from the alignment's window cigars (:584-655).
//   wsl: misc | compaction queues (shared memory);  ws: frag | slen | rb (shared, or global for the GWS class)
// MODE (one kernel each; reads of up to ALIGN_WIDE_FROM slots go EXPAND -> align_kernel -> QUAL, the others FULL):
//   FIN_EXPAND  expansion only: read codes to the scr_ops arena, slot lengths behind the fragment in the `mid` record,
//               (read length with flanks, trims) to rinfo; nothing but the small workspace part in shared memory;
//   FIN_QUAL    qscores from the ops align_kernel left in scr_ops;
//   FIN_FULL    expansion, team-parallel alignment with the 64-bit band, qscores, all on the workspace arrays:
//               long reads and the redo lists (reads align_kernel gave up on).
enum : int { FIN_EXPAND = 2, FIN_FULL = 3, FIN_QUAL = 4 };
constexpr uint32_t RINFO_SKIP = 1;     // the qscore kernel of the class leaves the read alone: the 64-bit alignment instance /
                                       // the FIN_FULL kernel finish it
constexpr uint32_t RINFO_BAND_LEFT = 4; // the 64-bit instance gave up on the read: the FIN_FULL kernel need not try that band again
constexpr uint32_t RINFO_WIDE_OK = 2;  // aligned by the 64-bit instance: its qscores come from the FIN_QUAL launch on that list

template <int W, bool GWS, int MODE>
__device__ __forceinline__ void finish_read(const Team<W>& tm, const SimParams& P, uint32_t r, uint8_t* ws, uint8_t* wsl,
                                            const uint8_t* lut3, uint8_t* trace) {
    constexpr int SIZE = Team<W>::SIZE;
    const DevTables& T = P.T;
    const DevBatch& B = P.B;
    uint32_t* misc = (uint32_t*)wsl;
    uint32_t* wsum = misc + MISC_WSUM;
    uint8_t* qbase = wsl + MISC_WORDS * 4;
    uint32_t* wq_a = (uint32_t*)(qbase + (uint32_t)(tm.rank >> 5) * QUEUE_BYTES);
    uint32_t* wq_b = wq_a + 64;
    const uint32_t mol = B.read_mol[r];
    const uint32_t L = B.mol_raw_len[mol];
    const uint32_t k = (uint32_t)T.k;
    const uint32_t n = L + 2 * k;
    const uint64_t ordinal = B.ordinal_base + r;
    const uint32_t olo = (uint32_t)ordinal, ohi = (uint32_t)(ordinal >> 32);
    const uint32_t k0 = T.seed_lo, k1 = T.seed_hi;
    const uint64_t sbase = B.scr_off[r];
    const uint32_t scap = (uint32_t)(B.scr_off[r + 1] - sbase);
    const uint4* mid4 = (const uint4*)(P.mid + 3 * sbase + 64 * (uint64_t)r);
    const uint32_t key_q = (n + 3) >> 2, frag_q = (n + 32 + 15) >> 4;  // 16-byte units
    const uint32_t* gkey = (const uint32_t*)mid4;   // slot keys stay in global memory (each is read once or twice)
    // fragment, slot lengths, read codes: workspace arrays (FIN_FULL), else the `mid` record and the scr_ops arena
    uint8_t* frag = MODE == FIN_FULL ? ws : (uint8_t*)(mid4 + key_q);
    uint8_t* slen = MODE == FIN_FULL ? ws + P.ncap + 32 : (uint8_t*)(mid4 + key_q + frag_q);
    uint8_t* rb = MODE == FIN_FULL ? slen + P.ncap : (MODE == FIN_QUAL ? B.scr_cig : B.scr_ops) + sbase;
    const uint32_t* fragw = (const uint32_t*)frag;
    if (MODE == FIN_FULL) {
        uint4* f4 = (uint4*)frag;
        for (uint32_t x = tm.rank; x < frag_q; x += SIZE) f4[x] = mid4[key_q + x];
    }
    uint32_t status = B.status[r];
    if (tm.rank == 0) misc[MISC_ALIGN_OOB] = 0;
    auto key_len = [&](uint32_t p) -> uint32_t {
        const uint32_t c = __ldg(gkey + p) & 0xFFFFu;
        return c == 0xFFFFu ? 1u : content_len(c);
    };
    uint32_t start_trim = 0, end_trim = 0;
    if (MODE != FIN_QUAL) {
        for (uint32_t j = 0; j < k; j++) {
            start_trim += key_len(j);
            end_trim += key_len(n - k + j);
        }
    }
    tm.sync();

    // ---- slot lengths -> read offsets; read bases into rb (codes) and the scratch arena (text);
    //      `ops_mode`: instead, the cigar ops of the edit script itself (used when the alignment fails)
    uint8_t* oseq = B.scr_seq + sbase;
    const uint32_t rbcap = (MODE == FIN_FULL || B.scr_ops) ? P.rb_mul * n : 0u;   // no read codes without qscores
    uint32_t Ttot = 0;
    auto expand_pass = [&](const bool ops_mode) {
        uint32_t carry = 0;
        WarpQueue xq{wq_a, wq_b, 0, 0};
        auto expand_changed = [&](uint32_t cnt) {
            uint32_t pv = 0, o = 0;
            if (!xq.take(cnt, pv, o)) return;
            const uint32_t p = pv & 0x0FFFFFFFu, dc = pv >> 28;
            const uint32_t c = __ldg(gkey + p) & 0xFFFFu, b = frag[p];
            const uint32_t len = content_len(c), kind = content_kind(c);
            const bool body = p >= k && p < n - k;
            uint32_t eqpos = 0xFFu;  // index inside the slot that aligns to the original base
            bool first_x = false;
            if (kind == 1u) eqpos = 0;
            else if (kind == 2u) eqpos = 1;
            else if (len >= 2 && b < 4 && content_base(c, 0) == b) eqpos = 0;
            else if (len >= 2 && b < 4 && content_base(c, len - 1) == b) eqpos = len - 1;
            else first_x = true;
            for (uint32_t u = 0; u < len; u++) {
                const uint32_t j = o + u;
                uint32_t code;
                if (kind == 0u) code = content_base(c, u);
                else if (kind == 1u) code = u == 0 ? b : content_base(c, 0);
                else code = u == 0 ? content_base(c, 0) : b;
                if (ops_mode) {
                    const uint32_t op = u == eqpos ? OP_EQ : (first_x && u == 0) ? OP_X : OP_I;
                    if (j < rbcap) rb[j] = (uint8_t)(op | ((u == 0 ? dc : 0u) << 2));
                } else {
                    if (j < rbcap) rb[j] = (uint8_t)code;
                    if (body) {
                        const uint32_t d = j - start_trim;
                        if (d < scap) oseq[d] = code_ascii(code);
                    }
                }
            }
        };
        for (uint32_t base = 0; base < n; base += 4 * SIZE) {
            const uint32_t p0 = base + 4 * tm.rank;
            uint32_t c[4], len[4], s = 0;
            uint32_t fb4 = 0;
            if (p0 < n) {
                const uint4 kq4 = __ldg((const uint4*)(gkey + p0));  // the key array is padded to a multiple of 4 slots
                c[0] = kq4.x & 0xFFFFu;
                c[1] = kq4.y & 0xFFFFu;
                c[2] = kq4.z & 0xFFFFu;
                c[3] = kq4.w & 0xFFFFu;
                fb4 = fragw[p0 >> 2];
            } else {
                c[0] = c[1] = c[2] = c[3] = 0;
            }
#pragma unroll
            for (int t = 0; t < 4; t++) {
                const uint32_t p = p0 + t;
                len[t] = p < n ? (c[t] == 0xFFFFu ? 1u : content_len(c[t])) : 0u;
                s += len[t];
            }
            if (!ops_mode && p0 < n) *(uint32_t*)(slen + p0) = len[0] | (len[1] << 8) | (len[2] << 16) | (len[3] << 24);
            uint32_t tot;
            uint32_t o = carry + team_excl_scan<W>(tm, s, wsum, tot);
            carry += tot;
            // deleted slots right before this lane's first slot (only the edit-script cigar needs them)
            uint32_t pend = 0;
            if (ops_mode && p0 < n) {
                for (uint32_t q = p0; q > 0 && pend < 15; q--) {
                    if ((__ldg(gkey + q - 1) & 0xFFFFu) >> CONTENT_LEN_SHIFT != 0) break;  // 0xFFFF and len >= 1 both stop
                    pend++;
                }
            }
#pragma unroll
            for (int t = 0; t < 4; t++) {
                const uint32_t p = p0 + t;
                bool changed = false;
                uint32_t qv = 0, qo = 0;
                if (p < n) {
                    if (len[t] == 0) {
                        pend = pend < 15 ? pend + 1 : 15;
                    } else {
                        const uint32_t dc = pend;
                        pend = 0;
                        if (c[t] == 0xFFFFu) {
                            // unchanged slot: one '=' column
                            const uint32_t code = (fb4 >> (8 * t)) & 0xFFu;
                            if (ops_mode) {
                                if (o < rbcap) rb[o] = (uint8_t)(OP_EQ | (dc << 2));
                            } else {
                                if (o < rbcap) rb[o] = (uint8_t)code;
                                if (p >= k && p < n - k) {
                                    const uint32_t d = o - start_trim;
                                    if (d < scap) oseq[d] = code_ascii(code);
                                }
                            }
                        } else {
                            // changed slot (about one in six): queued, expanded 32 at a time with every lane busy
                            changed = true;
                            qv = p | (dc << 28);
                            qo = o;
                        }
                        o += len[t];
                    }
                }
                xq.push(changed, qv, qo);
                const uint32_t cnt = xq.ready(false);
                if (cnt) expand_changed(cnt);
            }
        }
        for (uint32_t cnt; (cnt = xq.ready(true)) != 0;) expand_changed(cnt);
        Ttot = carry;
        tm.sync();
    };
    if (MODE != FIN_QUAL) {
        expand_pass(false);
        const uint32_t Lp = Ttot - start_trim - end_trim;
        if (Ttot > P.rb_mul * n) status |= ST_OPS_OVERFLOW;
        if (Lp > scap) status |= ST_SCRATCH_OVERFLOW;
        if (tm.rank == 0) {
            B.read_len[r] = Lp;
            B.status[r] = status;
        }
    }
    if (MODE == FIN_EXPAND) {
        // reads above ALIGN_WIDE_FROM slots take the 64-bit band at once: the FIN_FULL kernel of the class does them
        const bool aligned = (P.flags & RUN_QUAL) && !(status & (ST_OPS_OVERFLOW | ST_SCRATCH_OVERFLOW));
        const bool full = aligned && n > ALIGN_WIDE_FROM;
        // above ALIGN_WIDE64_FROM slots the 64-bit band at once: in a class that also holds shorter reads, through the
        // list of the 64-bit alignment instance (P.redo64; the other classes run that instance on all their reads)
        const bool wide64 = aligned && !full && n > ALIGN_WIDE64_FROM && P.redo64;
        if (tm.rank == 0) {
            if (B.rinfo) B.rinfo[r] = RInfo{Ttot, start_trim, end_trim, (full || wide64) ? RINFO_SKIP : 0u};
            if (full) P.redo[atomicAdd(P.redo_count, 1u)] = r;
            if (wide64) P.redo64[atomicAdd(P.redo64_count, 1u)] = r;
        }
        tm.sync();
        return;
    }
    if (MODE == FIN_QUAL) {
        const RInfo ri = B.rinfo[r];
        // the launch on the class's work list does the reads the 32-bit band aligned, the one on the 64-bit instance's
        // list (P.n_work_dev set) those that instance aligned
        if (P.n_work_dev ? !(ri.flags & RINFO_WIDE_OK) : (ri.flags & RINFO_SKIP) != 0) return;
        Ttot = ri.m;
        start_trim = ri.start_trim;
        end_trim = ri.end_trim;
    }

    // ---- qscores for the read bases that survive trimming
    if ((P.flags & RUN_QUAL) && !(status & (ST_OPS_OVERFLOW | ST_SCRATCH_OVERFLOW))) {
        // the read is aligned back to the fragment; the ops of the alignment replace the read codes in rb
        if (MODE == FIN_FULL) {
            uint32_t eq = 0, cols = 0;
            bool ok = false;
            // (a read the thread-per-read 64-bit instance gave up on would leave the same band here: same definition)
            const bool band_left = P.n_work_dev && B.rinfo && (B.rinfo[r].flags & RINFO_BAND_LEFT);
            if (!(P.flags & RUN_NO_ALIGN) && !band_left) {
                using BV = uint64_t;
                AlignCtx<BV> A;
                A.frag = frag;
                A.slen = slen;
                A.rb = rb;
                A.out = rb;
                A.trace = (AlignCell<BV>*)trace;
                A.cells = (uint32_t)(P.trace_bytes / sizeof(AlignCell<BV>));
                A.oob = misc + MISC_ALIGN_OOB;
                A.n = n;
                A.m = Ttot;
                ok = align_team_impl<W, BV>(tm, A, (uint4*)qbase, misc, wsum, eq, cols);
            }
            tm.sync();
            if (misc[MISC_ALIGN_OOB]) {
                ok = false;
                status |= ST_ALIGN_OOB;
#ifdef TK_ALIGN_PARANOID
                if (tm.rank == 0 && P.dbg && atomicAdd(P.dbg, 1u) == 0) {
                    P.dbg[1] = r;
                    for (int q = 0; q < 5; q++) P.dbg[2 + q] = misc[MISC_ALIGN_OOB + q];
                    P.dbg[7] = n;
                    P.dbg[8] = Ttot;
                }
#endif
            }
            if (ok) {
                if (tm.rank == 0) {
                    B.ident_num[r] = eq;
                    B.ident_den[r] = cols | IDENT_MEASURED;
                }
            } else {
                // never seen at the shipped models' usual identities: the edit script itself is the cigar then
                status |= ST_BAND;
                if (tm.rank == 0) B.status[r] = status;
                expand_pass(true);
            }
        }
        uint8_t* oq = B.scr_qual + sbase;
        const int margin = (T.kq - 1) >> 1;
        const uint32_t jb = start_trim, je = Ttot - end_trim;
        WarpQueue sq{wq_a, wq_b, 0, 0};
        const uint32_t* q0row = (const uint32_t*)(lut3 + 64);
        const uint32_t ent0 = T.q_t9 ? __ldg(&T.q_t9[0]) : 0u;   // the all-match window (its alias row is q0row)
        auto sample = [&](uint32_t off, uint32_t lg, uint32_t u, uint32_t j) {
            const uint32_t col = lg ? (u >> (16 - lg)) : 0u;
            const uint32_t e = __ldg(&T.q_alias[off + col]);
            const uint32_t fr = (u << lg) & 0xFFFFu;
            const uint32_t q = fr < (e & 0xFFFFu) ? ((e >> 16) & 127u) : ((e >> 23) & 127u);
            oq[j - jb] = (uint8_t)(q + 33);
        };
        // windows the direct tables do not cover: generic trim-and-retry lookup, 32 at a time
        auto run_slow = [&](uint32_t cnt) {
            uint32_t j = 0, u = 0;
            if (!sq.take(cnt, j, u)) return;
            // columns j-4 .. j+4, shrunk symmetrically at the read ends (py/tksm_badread.py:632-639)
            int t0 = 4 - margin;
            t0 = max(t0, max(4 - (int)j, (int)j + 4 - (int)(Ttot - 1)));
            uint32_t ops18 = 0, gaps = 0;
            for (int i = t0; i <= 8 - t0; i++) {
                const uint32_t o = rb[j + (uint32_t)i - 4u];
                ops18 |= (o & 3u) << (2 * i);
                if (i > t0) gaps |= (o >> 2) << (4 * (i - 1));
            }
            const QRow rw = T.q_rows[qscore_row(T, lut3, ops18, gaps, (uint32_t)t0)];
            sample(rw.off, rw.lg, u, j);
        };
        const uint32_t g0 = jb >> 3, g1 = (je + 7) >> 3;  // groups of 8 read bases [g0, g1)
        const uint32_t ng_r = (g1 - g0 + 31u) & ~31u;
        for (uint32_t it = tm.rank; it < ng_r; it += SIZE) {
            const uint32_t g = g0 + it, j0 = 8 * g;
            const bool live = g < g1;
            // 9-wide windows of j0..j0+7 read the ops of j0-4 .. j0+11
            const bool fast = live && T.q_t9 && j0 >= 4 && j0 + 11 < Ttot;
            uint4 ur = make_uint4(0, 0, 0, 0);
            uint32_t x0 = 0, x1 = 0, x2 = 0, x3 = 0, opsP = 0, dmask = 0;
            if (live) ur = philox4x32_10(g, TAG_QUAL, olo, ohi, k0, k1);
            if (fast) {
                const uint32_t* kp = (const uint32_t*)(rb + (j0 - 4));   // j0 is a multiple of 8
                x0 = kp[0];
                x1 = kp[1];
                x2 = kp[2];
                x3 = kp[3];
                opsP = pack4(x0) | (pack4(x1) << 8) | (pack4(x2) << 16) | (pack4(x3) << 24);
                dmask = nz4(x0, 0x3C3C3C3Cu) | (nz4(x1, 0x3C3C3C3Cu) << 4) | (nz4(x2, 0x3C3C3C3Cu) << 8) |
                        (nz4(x3, 0x3C3C3C3Cu) << 12);
            }
#pragma unroll 1
            for (uint32_t t = 0; t < 8; t++) {
                const uint32_t j = j0 + t;
                const bool want = live && j >= jb && j < je;
                const uint32_t wsel = t < 4 ? (t < 2 ? ur.x : ur.y) : (t < 6 ? ur.z : ur.w);
                const uint32_t u16 = (wsel >> (16 * (t & 1u))) & 0xFFFFu;
                bool slow = want;
                if (want && fast) {
                    // One path for every window the closure tables cover (no deletion; one D run of up to 3; two single
                    // D columns): which table and which entry is worked out with selects, so that the lanes of a warp
                    // stay together (each kind used to be a branch of its own: 14 of 32 lanes active on average).
                    const uint32_t ops18 = (opsP >> (2 * t)) & 0x3FFFFu;
                    const uint32_t gm = (dmask >> (t + 1)) & 0xFFu;
                    const uint32_t code3 = lut3[ops18 & 63u] + 27u * lut3[(ops18 >> 6) & 63u] + 729u * lut3[ops18 >> 12];
                    // D count in front of column bi of the 16-column strip (bi = t + 1 + gap)
                    auto dc_of = [&](uint32_t bi) {
                        const uint32_t w = bi < 4 ? x0 : bi < 8 ? x1 : bi < 12 ? x2 : x3;
                        return (w >> (8 * (bi & 3u) + 2)) & 15u;
                    };
                    const uint32_t rest = gm & (gm - 1);
                    const uint32_t g1 = gm ? __ffs(gm) - 1 : 0u, g2 = rest ? __ffs(rest) - 1 : 0u;
                    const uint32_t d1 = dc_of(t + 1 + g1), d2 = dc_of(t + 1 + g2);
                    const bool none = gm == 0;
                    const bool one = gm != 0 && rest == 0 && d1 <= 3;
                    const bool two = rest != 0 && (rest & (rest - 1)) == 0 && d1 == 1 && d2 == 1;
                    const bool allmatch = (gm | ops18) == 0;   // nine matches: entry and alias row are at hand
                    const uint32_t* tbl = one ? T.q_t9d1 : two ? T.q_t9d2 : T.q_t9;
                    const uint32_t idx = code3 + 19683u * (one ? 3 * g1 + d1 - 1 : two ? g1 * (15 - g1) / 2 + g2 - g1 - 1 : 0u);
                    if (none || one || two) {
                        slow = false;
                        const uint32_t ent = allmatch ? ent0 : __ldg(tbl + idx);
                        const uint32_t off = ent >> 3, lg = ent & 7u;
                        const uint32_t col = lg ? (u16 >> (16 - lg)) : 0u;
                        const uint32_t e = allmatch ? q0row[col] : __ldg(&T.q_alias[off + col]);
                        const uint32_t fr = (u16 << lg) & 0xFFFFu;
                        const uint32_t q = fr < (e & 0xFFFFu) ? ((e >> 16) & 127u) : ((e >> 23) & 127u);
                        oq[j - jb] = (uint8_t)(q + 33);
                    }
                }
                sq.push(slow, j, u16);
                const uint32_t cnt = sq.ready(false);
                if (cnt) run_slow(cnt);
            }
        }
        for (uint32_t cnt; (cnt = sq.ready(true)) != 0;) run_slow(cnt);
    }
    tm.sync();
}

template <int W, bool GWS, int PHASE>
__global__ void __launch_bounds__(PHASE >= 2 ? 1024 : SIM_THREADS, 1) simulate_kernel(const SimParams P) {
    extern __shared__ __align__(16) uint8_t smem[];
    // 24 warps per CTA; the PHASE 2 kernels need fewer registers and less shared memory per team and
    // run 32 warps per CTA where the team workspaces fit (the host decides: blockDim.x)
    const int TEAMS = (int)(blockDim.x >> 5) / W;
    Team<W> tm;
    tm.rank = threadIdx.x % (W * 32);
    tm.id = threadIdx.x / (W * 32);
    tm.bar = TEAMS > 15 ? tm.id : tm.id + 1;
    // k-mer change-probability classes staged in shared memory for the whole CTA, then a 64-entry
    // table turning three packed cigar ops into their base-3 code
    const uint32_t nq = (P.T.has_err_table && PHASE == 1) ? (1u << (2 * P.T.k)) : 0u;  // arrivals only
    uint8_t* qcls_s = smem;
    for (uint32_t i = threadIdx.x; i < nq; i += blockDim.x) qcls_s[i] = P.T.qcls[i];
    const uint32_t qcls_bytes = (nq + 15) & ~15u;
    uint8_t* lut3 = smem + qcls_bytes;
    if (threadIdx.x < 64) lut3[threadIdx.x] = (uint8_t)((threadIdx.x & 3) + 3 * ((threadIdx.x >> 2) & 3) + 9 * (threadIdx.x >> 4));
    // the window of nine matches is by far the most frequent one: its alias row lives in shared memory
    uint32_t* q0row = (uint32_t*)(lut3 + 64);
    if ((PHASE == FIN_FULL || PHASE == FIN_QUAL) && P.T.q_t9 && threadIdx.x < 128) {
        const uint32_t e0 = P.T.q_t9[0];
        if (threadIdx.x < (1u << (e0 & 7u))) q0row[threadIdx.x] = P.T.q_alias[(e0 >> 3) + threadIdx.x];
    }
    __syncthreads();
    // GWS class: the large arrays of the (long) read in global memory, the small hot part of the workspace in
    // shared memory.  P.slab_bytes: bytes of one team workspace (shared classes) / of one global workspace (GWS).
    uint8_t* pool = smem + qcls_bytes + SIM_FIXED_SMEM;
    uint8_t *ws, *wsl;
    size_t pool_used;
    if (PHASE == FIN_EXPAND || PHASE == FIN_QUAL) {
        // no large arrays: P.slab_bytes is the small part
        wsl = pool + (size_t)tm.id * P.slab_bytes;
        ws = nullptr;
        pool_used = (size_t)TEAMS * P.slab_bytes;
    } else if (PHASE == FIN_FULL) {
        const uint32_t small = slab2_small_bytes(W);
        if (GWS) {
            wsl = pool;
            ws = P.gws + (size_t)blockIdx.x * P.slab_bytes;
            pool_used = small;
        } else {
            wsl = pool + (size_t)tm.id * P.slab_bytes;
            ws = wsl + small;
            pool_used = (size_t)TEAMS * P.slab_bytes;
        }
    } else {
        if (GWS) {
            wsl = pool;
            ws = P.gws + (size_t)blockIdx.x * P.slab_bytes;
            pool_used = gws_small_bytes(W);
        } else {
            ws = wsl = pool + (size_t)tm.id * P.slab_bytes;
            pool_used = (size_t)TEAMS * P.slab_bytes;
        }
    }
    uint32_t* ticket = (uint32_t*)(pool + pool_used) + tm.id;
    uint8_t* trace = PHASE == FIN_FULL ? P.trace + ((size_t)blockIdx.x * TEAMS + tm.id) * P.trace_bytes : nullptr;
    const uint32_t n_work = P.n_work_dev ? *P.n_work_dev : P.n_work;
    while (true) {
        if (tm.rank == 0) *ticket = atomicAdd(P.counter, 1u);
        tm.sync();
        const uint32_t wi = *ticket;
        tm.sync();
        if (wi >= n_work) break;
        if (PHASE >= 2)
            finish_read<W, GWS, PHASE>(tm, P, P.work[wi], ws, wsl, lut3, trace);
        else
            inject_read<W, GWS>(tm, P, P.work[wi], ws, wsl, qcls_s);
    }
}

// ------------------------------------------------------- align_kernel
// Alignment of reads of up to ALIGN_WIDE_FROM slots: ONE THREAD PER READ, a warp takes 32 consecutive reads of the
// sorted work list (near-equal lengths).  The forward pass is align.cuh's column step run over the whole read
// (one chunk, true initial state: no assumptions, no rounds); fragment codes, slot lengths and read codes stream
// through 8-byte windows fetched one window ahead; the (Dg, Pv) cells of column j of the 32 reads of a warp are
// neighbours in the trace array ([(j - 1) * 32 + lane]: one coalesced store per column).  The traceback is the walk
// of align_trace_chunk<TRACE_WRITE> over the whole read with the same windows running backwards and the next
// column's cell fetched ahead; it writes op | dcount << 2 per read base to the scr_cig arena (the read codes stay
// intact).  A read whose walk leaves the 32-bit band goes to a list the 64-bit instance of the kernel works off; one
// that leaves that band too goes to the redo list of its class, where the FIN_FULL kernel gives it the edit script's
// own cigar.
struct AlignKParams {
    DevBatch B;
    const uint32_t* work;
    uint32_t n_work;
    uint32_t* counter;      // work ticket
    const uint8_t* mid;
    uint8_t* trace;         // trace_warp_bytes per warp of the grid
    uint64_t trace_warp_bytes;
    uint32_t k;
    uint32_t flags;
    uint32_t* redo;         // reads this kernel gives up on: for its 64-bit instance (32-bit instance) / the FIN_FULL kernel
    uint32_t* redo_count;
    const uint32_t* n_work_dev;   // 64-bit instance: number of entries of `work` (device memory)
    const uint32_t* work_begin_dev;   // ... and the entry it starts at (an earlier launch did those before it); nullptr: 0
};

// sequential bytes of a global array through 8-byte windows, the next window in flight (base 8-byte aligned; reads up
// to 15 bytes past the last byte asked for)
struct ByteStream {
    const uint2* p;
    uint64_t win, nxt;
    uint32_t left;
    __device__ __forceinline__ static uint64_t ld(const uint2* q) {
        const uint2 v = __ldg(q);
        return ((uint64_t)v.y << 32) | v.x;
    }
    __device__ __forceinline__ void open(const uint8_t* base) {
        p = (const uint2*)base;
        win = ld(p++);
        nxt = ld(p++);
        left = 8;
    }
    __device__ __forceinline__ uint32_t next() {
        if (left == 0) {
            win = nxt;
            nxt = ld(p++);
            left = 8;
        }
        const uint32_t c = (uint32_t)win & 0xFFu;
        win >>= 8;
        left--;
        return c;
    }
};
// the same running backwards: byte(idx) for non-increasing idx
struct ByteStreamBack {
    const uint2* base;
    uint64_t win, nxt;
    uint32_t q;   // window holds bytes [8 q, 8 q + 8)
    __device__ __forceinline__ void open(const uint8_t* b, uint32_t idx) {
        base = (const uint2*)b;
        q = idx >> 3;
        win = ByteStream::ld(base + q);
        nxt = q ? ByteStream::ld(base + q - 1) : 0ull;
    }
    __device__ __forceinline__ uint32_t at(uint32_t idx) {
        while ((idx >> 3) != q) {   // one window down as a rule; more after a long run of D or I moves
            q--;
            win = nxt;
            if (q) nxt = ByteStream::ld(base + q - 1);
        }
        return (uint32_t)(win >> (8u * (idx & 7u))) & 0xFFu;
    }
};

// 2 BW rows of one bit plane of the read, bit 0 of `cur` = row lo of the band: the band itself and the rows the next
// columns will shift into it
template <typename BV>
struct AlignAhead {
    BV cur, nxt;
    __device__ __forceinline__ void shift(uint32_t s) {   // 1 <= s < BW
        constexpr uint32_t BW = 8 * sizeof(BV);
        cur = (cur >> s) | (nxt << (BW - s));
        nxt >>= s;
    }
    __device__ __forceinline__ void insert(uint32_t bits8, uint32_t p) {   // eight rows at p <= 2 BW - 8
        constexpr uint32_t BW = 8 * sizeof(BV);
        if (p < BW) {
            cur |= BV(bits8) << p;
            if (p + 8 > BW) nxt |= BV(bits8) >> (BW - p);
        } else {
            nxt |= BV(bits8) << (p - BW);
        }
    }
};
constexpr uint32_t ALIGN_PREFETCH_COLS = 12;
__device__ __forceinline__ void prefetch_l2(const void* p) {
#ifdef __CUDA_ARCH__
    asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
#else
    (void)p;
#endif
}
// bits 0, 8, 16, 24 of x -> bits 0..3
__device__ __forceinline__ uint32_t gather4(uint32_t x) { return (x * 0x10204080u) >> 28; }

// Eq of a column whose fragment byte is not A/C/G/T (an N as a rule): the rows of the band that hold the same byte
template <typename BV>
__device__ inline BV align_eq_rare(const uint8_t* rb, uint32_t lo, uint32_t m, uint32_t f) {
    constexpr uint32_t BW = 8 * sizeof(BV);
    BV Eq = 0;
#pragma unroll 1
    for (uint32_t b = 0; b < BW; b++) {
        const uint32_t row = lo + b;
        if (row > m) break;
        if (rb[row - 1] == f) Eq |= BV(1) << b;
    }
    return Eq;
}

template <typename BV>
__device__ __forceinline__ bool align_one(const uint8_t* frag, const uint8_t* slen, const uint8_t* rb, uint8_t* out, uint32_t n,
                                          uint32_t m, AlignCell2<BV>* cells, uint32_t off, uint32_t nmax, bool valid,
                                          uint32_t& eq_out, uint32_t& cols_out) {
    // every thread of the warp comes here (`valid`: it has a read); n = m = 0 otherwise; nmax = largest n of the warp
    constexpr uint32_t BW = 8 * sizeof(BV);
    static_assert(ALIGN_MARGIN == 4, "the cells carry four bits of column data where the margin rows would be");
    constexpr uint32_t CELL_SH = BW - ALIGN_MARGIN;
    constexpr BV CELL_LOW = (BV(1) << CELL_SH) - 1;
    const uint2* l8 = (const uint2*)slen;
    const uint2* f8 = (const uint2*)frag;
    const uint2* r8 = (const uint2*)rb;
    // ---- forward: the column step of align_forward_chunk over columns 1 .. n from the true boundary, the 32 threads
    //      of the warp in step (same instructions for every thread: the per-thread branches of the chunked form cost
    //      a warp the sum of their paths).  The read enters as bit planes of 2 BW rows, refilled eight rows at a time
    //      for the whole warp at once when one of its threads runs low.
    {
        AlignAhead<BV> a0{0, 0}, a1{0, 0}, ax{0, 0}, an{0, 0};   // code bit 0, code bit 1, "not A/C/G/T, or beyond the read", "N"
        uint32_t avail = 0, gnext = 0;   // rows the planes hold from row lo on; next eight-byte group of the read
        uint64_t w_ahead = m ? ByteStream::ld(r8) : 0ull;   // that group, fetched one refill ahead
        auto refill = [&]() {
            const uint64_t w = w_ahead;
            w_ahead = 8 * (gnext + 1) < m ? ByteStream::ld(r8 + gnext + 1) : 0ull;
            const uint32_t x0 = (uint32_t)w, x1 = (uint32_t)(w >> 32);
            const uint32_t t0 = x0 & 0xFCFCFCFCu, t1 = x1 & 0xFCFCFCFCu;
            uint32_t inv = 0;
            if (t0 | t1) {   // (rare: an N of the fragment that came through, or another IUPAC letter)
                inv = gather4((((t0 & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | t0) >> 7 & 0x01010101u) |
                      gather4((((t1 & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | t1) >> 7 & 0x01010101u) << 4;
                const uint32_t y0 = x0 ^ 0x4E4E4E4Eu, y1 = x1 ^ 0x4E4E4E4Eu;   // zero bytes: N (78)
                const uint32_t z0 = ~(((y0 & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | y0) >> 7 & 0x01010101u;
                const uint32_t z1 = ~(((y1 & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | y1) >> 7 & 0x01010101u;
                uint32_t isn = gather4(z0) | gather4(z1) << 4;
                const uint32_t have_n = m > 8 * gnext ? m - 8 * gnext : 0u;
                if (have_n < 8) isn &= (1u << have_n) - 1u;
                an.insert(isn, avail);
            }
            const uint32_t have = m > 8 * gnext ? m - 8 * gnext : 0u;   // read bases of this group
            if (have < 8) inv |= 0xFFu << have;
            a0.insert(gather4(x0 & 0x01010101u) | gather4(x1 & 0x01010101u) << 4, avail);
            a1.insert(gather4((x0 >> 1) & 0x01010101u) | gather4((x1 >> 1) & 0x01010101u) << 4, avail);
            ax.insert(inv & 0xFFu, avail);
            avail += 8;
            gnext++;
        };
#pragma unroll 1
        for (uint32_t t = 0; t < 2 * BW / 8; t++) refill();
        BV Pv = ~BV(0), Mv = 0;
        uint32_t o_prev = 0, lo_prev = 1;
        AlignCell2<BV>* cell = cells + (size_t)off * 32;
        uint64_t lw_n = 0, fw_n = 0;
        if (n) {
            lw_n = ByteStream::ld(l8);
            fw_n = ByteStream::ld(f8);
        }
#pragma unroll 1
        for (uint32_t j0 = 1; j0 <= nmax; j0 += 8) {
            const uint64_t lw = lw_n, fw = fw_n;
            if (j0 + 8 <= n) {   // (the group after the last one is never fetched)
                lw_n = ByteStream::ld(l8 + (j0 >> 3) + 1);
                fw_n = ByteStream::ld(f8 + (j0 >> 3) + 1);
            }
#pragma unroll
            for (uint32_t t8 = 0; t8 < 8; t8++) {
                const bool live = j0 + t8 <= n;
                if (__any_sync(0xffffffffu, live && avail < BW + 13)) {
                    if (avail + 8 <= 2 * BW) refill();
                }
                if (live) {
                    const uint32_t len = (uint32_t)(lw >> (8 * t8)) & 0xFFu;
                    const uint32_t o = o_prev + len;
                    const uint32_t lo = align_lo<BV>(o);
                    const uint32_t s = lo - lo_prev;
                    if (s) {
                        Pv = (Pv >> s) | (~BV(0) << (BW - s));
                        Mv >>= s;
                        a0.shift(s);
                        a1.shift(s);
                        ax.shift(s);
                        an.shift(s);
                        avail -= s;
                    } else if (lo > 1) {
                        Pv &= ~BV(1);
                        Mv |= BV(1);
                    }
                    const uint32_t f = (uint32_t)(fw >> (8 * t8)) & 0xFFu;
                    BV Eq;
                    if (f < 4) {
                        const BV f0 = BV(0) - BV(f & 1u), f1 = BV(0) - BV(f >> 1);
                        Eq = ~(ax.cur | (a0.cur ^ f0) | (a1.cur ^ f1));
                    } else if (f == 78u) {
                        Eq = an.cur;
                    } else {
                        Eq = align_eq_rare<BV>(rb, lo, m, f);
                    }
                    const BV Xv = Eq | Mv;
                    const BV Xh = (((Eq & Pv) + Pv) ^ Pv) | Eq;
                    BV Ph = Mv | ~(Xh | Pv);
                    BV Mh = Pv & Xh;
                    const BV D0 = Xh | Mv;
                    Ph = (Ph << 1) | BV(1);
                    Mh <<= 1;
                    Pv = Mh | ~(Xv | Ph);
                    Mv = Ph & Xv;
                    // the walk never looks at the top ALIGN_MARGIN rows of a band: their bits carry the slot length and the
                    // fragment code of the column, so that it needs neither array again
                    AlignCell2<BV> c;
                    c.dg = ((Eq | ~D0) & CELL_LOW) | (BV(len | (f < 4 ? 0u : 8u)) << CELL_SH);
                    c.pv = (Pv & CELL_LOW) | (BV(f & 3u) << CELL_SH);
                    cell[(size_t)t8 * 32] = c;
                    o_prev = o;
                    lo_prev = lo;
                }
            }
            cell += 8 * 32;
        }
    }
    // ---- traceback over the whole read (align_trace_chunk<TRACE_WRITE> with cmin = 1, bottom, top), the 32 threads of
    // the warp one move at a time.  Every move depends on the one before it, so what the walk reads is fetched ahead
    // into registers: the cells of the next two columns (they carry the slot length and the fragment code of their
    // column) and the read codes eight at a time with the group after them in flight.  The rows of the trace array are shifted per thread (`off`) so that the
    // walks of a warp start in the same row and, moving at about the same pace, keep reading the same 256-byte rows
    // (measured against cells contiguous per thread: 7.3 ms instead of 9.8 ms for the largest class of the bulk shape).
    const AlignCell2<BV>* crow = cells + (size_t)off * 32;   // cell of column x at crow[(x - 1) * 32]
    uint32_t i = m, j = n, oj = m;
    uint64_t rwin = 0, rnx = 0;
    uint32_t gi = 0;   // eight-byte group of the read the window holds: (i - 1) >> 3
    AlignCell2<BV> c0{}, c1{}, c2{};   // cells of columns j, j - 1, j - 2
    if (valid && n) {
        c0 = crow[(size_t)(n - 1) * 32];
        if (n >= 2) c1 = crow[(size_t)(n - 2) * 32];
        if (n >= 3) c2 = crow[(size_t)(n - 3) * 32];
    }
    if (valid && m) {
        gi = (m - 1) >> 3;
        rwin = ByteStream::ld(r8 + gi);
        if (gi) rnx = ByteStream::ld(r8 + gi - 1);
    }
    auto byte_of = [](uint64_t w, uint32_t idx) { return (uint32_t)(w >> (8u * (idx & 7u))) & 0xFFu; };
    bool have = false, failed = false;
    uint32_t bi = 0, op = 0, run = 0, eq = 0, cols = 0, acc = 0;
    auto flush = [&]() {
        if (have) {
            acc |= (op | ((run > ALIGN_DC_CAP ? ALIGN_DC_CAP : run) << 2)) << (8u * (bi & 3u));
            if ((bi & 3u) == 0) {   // four ops per store (the array is 16-byte aligned and padded)
                *(uint32_t*)(out + bi) = acc;
                acc = 0;
            }
            eq += op == 0u;
            cols += 1 + run;
        } else {
            cols += run;   // D run after the last read base
        }
    };
    bool done = !valid || (i | j) == 0;
    while (__any_sync(0xffffffffu, !done)) {
        // the move of every thread first, then the threads that start a read base together, then those that change
        // column together (left to itself the compiler gives each kind of move its own copy of the shared parts and
        // the warp runs them one after the other)
        uint32_t d = MV_OUT;   // MV_OUT: no move (done)
        if (!done) {
            d = MV_LEFT;
            if (i != 0) {
                d = MV_UP;
                if (j != 0) {
                    const uint32_t lo = align_lo<BV>(oj);
                    const uint32_t b = (i - lo) & (BW - 1);
                    d = ((c0.dg >> b) & 1) ? MV_DIAG : ((c0.pv >> b) & 1) ? MV_UP : MV_LEFT;
                    if (i < lo + (lo > 1 ? ALIGN_MARGIN : 0u) || i + ALIGN_MARGIN >= lo + BW) {
                        failed = true;
                        done = true;
                        d = MV_OUT;
                    }
                }
            }
        }
        __syncwarp();
        run += d == MV_LEFT;
        if (d == MV_DIAG || d == MV_UP) {
            // a read base starts here: the one before it is complete (its D run is known now)
            flush();
            bi = i - 1;
            have = true;
            run = 0;
            op = 2u;   // OP_I
            if (d == MV_DIAG) {   // OP_EQ : OP_X
                const uint32_t fc = ((uint32_t)(c0.dg >> (CELL_SH + 3)) & 1u) ? (uint32_t)frag[j - 1] : (uint32_t)(c0.pv >> CELL_SH) & 3u;
                op = byte_of(rwin, bi) == fc ? 0u : 1u;
            }
            i--;
            if (i != 0 && ((i - 1) & 7u) == 7u) {   // the next base is the last of the group below
                rwin = rnx;
                gi--;
                if (gi) rnx = ByteStream::ld(r8 + gi - 1);
            }
        }
        __syncwarp();
        if (d == MV_DIAG || d == MV_LEFT) {   // to column j - 1
            oj -= (uint32_t)(c0.dg >> CELL_SH) & 7u;
            j--;
            c0 = c1;
            c1 = c2;
            if (j >= 3) c2 = crow[(size_t)(j - 3) * 32];
            // (the cells were written a forward pass ago: they come from DRAM, so the row the walk reaches in
            // ALIGN_PREFETCH_COLS moves is asked for now)
            if (j >= 3 + ALIGN_PREFETCH_COLS) prefetch_l2(crow + (size_t)(j - 3 - ALIGN_PREFETCH_COLS) * 32);
        }
        __syncwarp();
        if (d != MV_OUT) done = (i | j) == 0;
    }
    if (failed || !valid) return false;
    flush();
    eq_out = eq;
    cols_out = cols;
    return true;
}

// -------------------------------------------------------------- align_kernel
// One thread per read, a warp takes 32 consecutive reads of the work list (similar lengths: the list is sorted).
constexpr int ALIGN_K_THREADS = 128;
template <typename BV>
__global__ void __launch_bounds__(ALIGN_K_THREADS) align_kernel(const AlignKParams K) {
    const uint32_t n_work = K.n_work_dev ? *K.n_work_dev : K.n_work;
    const uint32_t begin = K.work_begin_dev ? *K.work_begin_dev : 0u;
    const DevBatch& B = K.B;
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    uint8_t* region = K.trace + (size_t)warp * K.trace_warp_bytes;
    while (true) {
        uint32_t base = 0;
        if (lane == 0) base = begin + atomicAdd(K.counter, 32u);
        base = __shfl_sync(0xffffffffu, base, 0);
        if (base >= n_work) break;
        const uint32_t wi = base + lane;
        bool valid = false;
        uint32_t r = 0, n = 0, m = 0, rflags = 0;
        const uint8_t *frag = nullptr, *slen = nullptr, *rb = nullptr;
        uint8_t* out = nullptr;
        if (wi < n_work) {
            r = K.work[wi];
            const RInfo ri = B.rinfo[r];
            const uint32_t st = B.status[r];
            rflags = ri.flags;
            // on its list the 64-bit instance takes what the 32-bit one (or the expansion) left for it: RINFO_SKIP is set
            // there; on a class's own work list it leaves the reads above ALIGN_WIDE_FROM slots to the FIN_FULL kernel
            if ((!(ri.flags & RINFO_SKIP) || K.n_work_dev) && !(st & (ST_OPS_OVERFLOW | ST_SCRATCH_OVERFLOW))) {
                valid = true;
                n = B.mol_raw_len[B.read_mol[r]] + 2 * K.k;
                m = ri.m;
                const uint64_t sbase = B.scr_off[r];
                const uint4* mid4 = (const uint4*)(K.mid + 3 * sbase + 64 * (uint64_t)r);
                const uint32_t key_q = (n + 3) >> 2, frag_q = (n + 32 + 15) >> 4;
                frag = (const uint8_t*)(mid4 + key_q);
                slen = (const uint8_t*)(mid4 + key_q + frag_q);
                rb = B.scr_ops + sbase;
                out = B.scr_cig + sbase;
            }
        }
        uint32_t eq = 0, cols = 0;
        const bool run_it = valid && !(K.flags & RUN_NO_ALIGN) && !(sizeof(BV) == 8 && (K.flags & RUN_NO_WIDE));
        // the longest fragment of the warp: the forward pass runs the warp in step up to it, and it sets the row in
        // which every walk of the warp starts
        uint32_t nmax = run_it ? n : 0u;
        for (int dd = 16; dd; dd >>= 1) nmax = max(nmax, __shfl_xor_sync(0xffffffffu, nmax, dd));
        const bool ok = align_one<BV>(frag, slen, rb, out, run_it ? n : 0u, run_it ? m : 0u, (AlignCell2<BV>*)region + lane,
                                      run_it ? nmax - n : 0u, nmax, run_it, eq, cols);
        if (valid) {
            if (ok) {
                B.ident_num[r] = eq;
                B.ident_den[r] = cols | IDENT_MEASURED;
                if (sizeof(BV) == 8) B.rinfo[r].flags = rflags | RINFO_WIDE_OK;
            } else {
                // the band was left: the read goes on to the 64-bit instance, from there to the FIN_FULL kernel of the
                // class (which gives it the edit script's own cigar); the qscore kernel leaves it alone meanwhile
                B.rinfo[r].flags = rflags | RINFO_SKIP | ((sizeof(BV) == 8 && run_it) ? RINFO_BAND_LEFT : 0u);
                K.redo[atomicAdd(K.redo_count, 1u)] = r;
            }
        }
        __syncwarp();
    }
}

// -------------------------------------------------------------- larger slots
// A read may not outgrow twice its fragment in the first pass of a chunk (scratch slot of 2 L + 32 bytes, 2 n read
// codes).  A chunk in which one did (ST_OPS_OVERFLOW / ST_SCRATCH_OVERFLOW) runs again with slots GROW_SLOT_MUL times
// as large and GROW_RB_MUL n read codes (MAX_SLOT_LEN bases per slot: the most an edit script can produce), in the
// length class that holds grow_len(n) slots, since the workspaces of a class hold twice its capacity in read codes.
constexpr uint32_t GROW_SLOT_MUL = 3, GROW_RB_MUL = MAX_SLOT_LEN;
__host__ __device__ inline uint32_t grow_len(uint32_t n) { return (GROW_RB_MUL * n + 1) / 2; }

__global__ void scale_offsets_kernel(const uint64_t* in, uint32_t count, uint32_t mul, uint64_t* out) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < count) out[i] = in[i] * mul;
}

// -------------------------------------------------------------- work order
// Reads are handed to the simulate kernels class by class, inside a class by ascending target
// identity (32 buckets: the error level sets the number of trials) and then by descending length,
// so that the teams of a CTA that start reads together (lockstep) finish them together.  One
// counting sort over a 17-bit key: class(3) | identity bucket(5) | quantised length(9).
constexpr uint32_t WORK_KEY_BITS = 17;
constexpr uint32_t WORK_BINS = 1u << WORK_KEY_BITS;

struct WorkSortParams {
    uint32_t n_reads;
    uint64_t ordinal_base;
    uint32_t seed_lo, seed_hi;
    int ident_const;
    uint32_t k;
    const uint32_t* read_mol;
    const uint32_t* mol_raw_len;
    uint32_t ncap[5];       // fragment capacity of the shared-memory classes
    uint32_t len_shift[6];  // quantisation of (capacity - n) per class
    uint32_t gws_cap;       // capacity used for the global-workspace class
    uint32_t grow;          // classes by grow_len(n): second pass of a chunk with an outgrown slot (engine.cu: slot_launch)
    uint32_t* keys;         // R
    uint32_t* bins;         // WORK_BINS (zeroed), turned into cursors by work_scan_kernel
    uint32_t* work;         // R: sorted read indices
};

__global__ void work_key_kernel(const WorkSortParams W) {
    const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= W.n_reads) return;
    uint32_t n = W.mol_raw_len[W.read_mol[r]] + 2 * W.k;
    if (W.grow) n = grow_len(n);
    uint32_t c = 0;
    while (c < 5 && n > W.ncap[c]) c++;
    const uint32_t cap = c < 5 ? W.ncap[c] : W.gws_cap;
    uint32_t lq = (cap - min(n, cap)) >> W.len_shift[c];
    lq = min(lq, 511u);
    uint32_t bucket = 0;
    if (!W.ident_const) {
        const uint64_t ordinal = W.ordinal_base + r;
        bucket = philox4x32_10(0, TAG_READ, (uint32_t)ordinal, (uint32_t)(ordinal >> 32), W.seed_lo, W.seed_hi).x >> 27;
    }
    const uint32_t key = (c << 14) | (bucket << 9) | lq;
    W.keys[r] = key;
    atomicAdd(&W.bins[key], 1u);
}

// exclusive scan of the WORK_BINS counters, one block of 1024 threads: every warp scans its own
// contiguous 1/32 of the bins with coalesced row loads, then adds the total of the warps before it
__global__ void __launch_bounds__(1024) work_scan_kernel(uint32_t* bins) {
    __shared__ uint32_t wtot[32];
    constexpr uint32_t PER_WARP = WORK_BINS / 32, ROWS = PER_WARP / 32;
    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint32_t* mine = bins + warp * PER_WARP;
    uint32_t carry = 0;
    for (uint32_t i = 0; i < ROWS; i++) {
        const uint32_t v = mine[i * 32 + lane];
        const uint32_t inc = warp_incl_scan(v);
        mine[i * 32 + lane] = carry + inc - v;
        carry += __shfl_sync(0xffffffffu, inc, 31);
    }
    if (lane == 0) wtot[warp] = carry;
    __syncthreads();
    if (threadIdx.x < 32) {
        const uint32_t v = wtot[threadIdx.x];
        wtot[threadIdx.x] = warp_incl_scan(v) - v;
    }
    __syncthreads();
    const uint32_t base = wtot[warp];
    if (base)
        for (uint32_t i = 0; i < ROWS; i++) mine[i * 32 + lane] += base;
}

__global__ void work_scatter_kernel(const WorkSortParams W) {
    const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= W.n_reads) return;
    W.work[atomicAdd(&W.bins[W.keys[r]], 1u)] = r;
}

// -------------------------------------------------------------- formatting
struct FmtParams {
    uint32_t n_reads;
    uint64_t ordinal_base;
    uint32_t seed_lo, seed_hi;
    const uint32_t* read_mol;
    const uint32_t* mol_raw_len;
    const uint32_t* mol_tail_len; // nullptr unless a tail-noise model is loaded: bases of the raw length that are tail
    const uint32_t* mol_id_begin;
    const uint8_t* id_pool;
    const uint32_t* read_len;     // nullptr => perfect output (length = error-free length)
    const uint32_t* ident_num;
    const uint32_t* ident_den;
    const uint8_t* seq;           // scratch arena
    const uint64_t* seq_off;
    const uint8_t* qual;          // nullptr => 'K' * len
    uint32_t fastq;               // 1 FASTQ, 0 FASTA
    double kappa;                 // identity calibration (header text)
    uint32_t* rec_len;
    const uint64_t* rec_off;
    uint8_t* out;
    // batch totals accumulated by reclen_kernel (badread output only; nullptr otherwise)
    const uint32_t* status;
    unsigned long long* sum_read_len;
    uint32_t* status_or;
};

__device__ __forceinline__ uint32_t dec_digits(uint32_t v) {
    uint32_t d = 1;
    while (v >= 10) {
        v /= 10;
        d++;
    }
    return d;
}

// round-half-even of (num / den * 100.0) * 100 computed on the exact binary value of the double,
// i.e. the digits Python's f"{x * 100.0:.2f}" prints (py/sequence.py:255)
// The identity printed is the aligned identity the generative one corresponds to:
// a = g - kappa g^2 with g = 1 - num / den.
__device__ __forceinline__ uint32_t pct_hundredths(uint32_t num, uint32_t den, double kappa) {
    const double g = __dsub_rn(1.0, __ddiv_rn((double)num, (double)den));
    const double a = __dsub_rn(g, __dmul_rn(__dmul_rn(kappa, g), g));
    double y = __dmul_rn(__dsub_rn(1.0, a), 100.0);
    uint64_t bits = (uint64_t)__double_as_longlong(y);
    int ex = (int)((bits >> 52) & 0x7FF);
    if (ex == 0) return 0;
    uint64_t M = (bits & 0xFFFFFFFFFFFFFull) | (1ull << 52);
    int sh = 1075 - ex;  // y = M * 2^-sh
    if (sh <= 0) return 10000;  // cannot happen for y <= 100
    uint64_t v = M * 100ull;   // < 2^60
    if (sh >= 64) return 0;
    uint64_t q = v >> sh, rem = v & ((1ull << sh) - 1), half = 1ull << (sh - 1);
    if (rem > half || (rem == half && (q & 1))) q++;
    return (uint32_t)q;
}

__device__ __forceinline__ uint32_t pct_len(uint32_t h) { return h >= 10000 ? 6u : h >= 1000 ? 5u : 4u; }

// read_identity of read r in hundredths of a percent: the identity of its alignment when it was aligned
// (py/tksm_badread.py:613-614), else the calibrated edit-script estimate (:443-444)
template <typename FP>
__device__ __forceinline__ uint32_t read_pct(const FP& F, uint32_t r) {
    const uint32_t den = F.ident_den[r];
    return pct_hundredths(F.ident_num[r], den & ~IDENT_MEASURED, (den & IDENT_MEASURED) ? 0.0 : F.kappa);
}

__device__ __forceinline__ uint32_t record_len(const FmtParams& F, uint32_t r) {
    const uint32_t mol = F.read_mol[r];
    // error_free_length = len(raw_seq), the molecule without its tail (py/sequence.py:253)
    const uint32_t L = F.mol_raw_len[mol] - (F.mol_tail_len ? F.mol_tail_len[mol] : 0u);
    // (a read that outgrew its scratch slot is flagged and reported by the host; its record is cut to the slot)
    const uint32_t Lp = F.read_len ? min(F.read_len[r], (uint32_t)(F.seq_off[r + 1] - F.seq_off[r])) : L;
    const uint32_t h = F.read_len ? read_pct(F, r) : 10000u;
    const uint32_t idl = F.mol_id_begin[mol + 1] - F.mol_id_begin[mol];
    // "@" uuid " length=" Lp " error_free_length=" L " read_identity=" pct "% molecule_id=" id "\n" seq "\n"
    uint32_t len = 1 + 36 + 8 + dec_digits(Lp) + 19 + dec_digits(L) + 15 + pct_len(h) + 14 + idl + 1 + Lp + 1;
    if (F.fastq) len += 2 + Lp + 1;  // "+\n" qual "\n"
    return len;
}

__global__ void reclen_kernel(const FmtParams F) {
    uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t lp = 0, stt = 0;
    if (r < F.n_reads) {
        F.rec_len[r] = record_len(F, r);
        if (F.sum_read_len) {
            lp = F.read_len[r];
            stt = F.status[r];
        }
    }
    if (F.sum_read_len) {
        lp = warp_sum(lp);
        stt = __reduce_or_sync(0xffffffffu, stt);
        if ((threadIdx.x & 31) == 0) {
            atomicAdd(F.sum_read_len, (unsigned long long)lp);
            if (stt) atomicOr(F.status_or, stt);
        }
    }
}

// exclusive scan of u32 -> u64, three small kernels (block sums, scan of sums, final)
constexpr int SCAN_BLOCK = 1024;
constexpr int SCAN_ITEMS = 4;
__global__ void scan_block_sums(const uint32_t* in, uint32_t n, uint64_t* sums) {
    __shared__ uint64_t sh[32];
    uint64_t s = 0;
    uint32_t base = blockIdx.x * SCAN_BLOCK * SCAN_ITEMS;
    for (int t = 0; t < SCAN_ITEMS; t++) {
        uint32_t i = base + t * SCAN_BLOCK + threadIdx.x;
        if (i < n) s += in[i];
    }
    for (int d = 16; d; d >>= 1) s += __shfl_xor_sync(0xffffffffu, s, d);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x < 32) {
        s = sh[threadIdx.x];
        for (int d = 16; d; d >>= 1) s += __shfl_xor_sync(0xffffffffu, s, d);
        if (threadIdx.x == 0) sums[blockIdx.x] = s;
    }
}
__global__ void scan_sums(uint64_t* sums, uint32_t nb, uint64_t* total) {
    // single block; serial over chunks of 1024
    __shared__ uint64_t sh[32];
    __shared__ uint64_t carry_s;
    if (threadIdx.x == 0) carry_s = 0;
    __syncthreads();
    for (uint32_t base = 0; base < nb; base += 1024) {
        uint32_t i = base + threadIdx.x;
        uint64_t v = i < nb ? sums[i] : 0, inc = v;
        const int lane = threadIdx.x & 31;
        for (int d = 1; d < 32; d <<= 1) {
            uint64_t o = __shfl_up_sync(0xffffffffu, inc, d);
            if (lane >= d) inc += o;
        }
        if (lane == 31) sh[threadIdx.x >> 5] = inc;
        __syncthreads();
        uint64_t wb = 0;
        for (int w = 0; w < (int)(threadIdx.x >> 5); w++) wb += sh[w];
        uint64_t c = carry_s;
        if (i < nb) sums[i] = c + wb + inc - v;
        __syncthreads();
        if (threadIdx.x == 1023) carry_s = c + wb + inc;
        __syncthreads();
    }
    if (threadIdx.x == 0) *total = carry_s;
}
__global__ void scan_final(const uint32_t* in, uint32_t n, const uint64_t* sums, uint64_t* out) {
    __shared__ uint64_t sh[32];
    __shared__ uint64_t carry_s;
    if (threadIdx.x == 0) carry_s = sums[blockIdx.x];
    __syncthreads();
    uint32_t base = blockIdx.x * SCAN_BLOCK * SCAN_ITEMS;
    for (int t = 0; t < SCAN_ITEMS; t++) {
        uint32_t i = base + t * SCAN_BLOCK + threadIdx.x;
        uint64_t v = i < n ? in[i] : 0, inc = v;
        const int lane = threadIdx.x & 31;
        for (int d = 1; d < 32; d <<= 1) {
            uint64_t o = __shfl_up_sync(0xffffffffu, inc, d);
            if (lane >= d) inc += o;
        }
        if (lane == 31) sh[threadIdx.x >> 5] = inc;
        __syncthreads();
        uint64_t wb = 0;
        for (int w = 0; w < (int)(threadIdx.x >> 5); w++) wb += sh[w];
        uint64_t c = carry_s;
        if (i < n) out[i] = c + wb + inc - v;
        __syncthreads();
        if (threadIdx.x == SCAN_BLOCK - 1) carry_s = c + wb + inc;
        __syncthreads();
    }
}

__device__ __forceinline__ char hexdig(uint32_t v) { return (char)(v < 10 ? '0' + v : 'a' + (v - 10)); }

// Warp-wide byte copy with 16-byte stores: a few head bytes bring the destination to a 16-byte
// boundary, then every lane moves 16 bytes per step (the source is read as aligned words and
// funnel-shifted into place), then the tail.  May read up to 7 bytes past src + n inside the same
// allocation (scratch arenas and the header buffer are padded).
__device__ __forceinline__ void warp_copy(uint8_t* dst, const uint8_t* src, uint32_t n, uint32_t lane) {
    uint32_t head = (uint32_t)((16u - (uint32_t)((uintptr_t)dst & 15u)) & 15u);
    head = min(head, n);
    if (lane < head) dst[lane] = src[lane];
    dst += head;
    src += head;
    n -= head;
    const uint32_t nv = n >> 4;
    const uint32_t sh = (uint32_t)((uintptr_t)src & 3u) * 8u;
    const uint32_t* sw = (const uint32_t*)((uintptr_t)src & ~(uintptr_t)3);
    uint4* dv = (uint4*)dst;
    for (uint32_t i = lane; i < nv; i += 32) {
        const uint32_t* p = sw + 4 * i;
        const uint32_t a = p[0], b = p[1], c = p[2], d = p[3], e = sh ? p[4] : 0u;
        dv[i] = make_uint4(__funnelshift_r(a, b, sh), __funnelshift_r(b, c, sh), __funnelshift_r(c, d, sh),
                           __funnelshift_r(d, e, sh));
    }
    const uint32_t done = nv << 4, tail = n - done;
    if (lane < tail) dst[done + lane] = src[done + lane];
}

// one warp per read
__global__ void __launch_bounds__(256) format_kernel(const FmtParams F) {
    __shared__ __align__(16) uint8_t hdr_s[8][160];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint8_t* hdr = hdr_s[warp];
    for (uint32_t r = blockIdx.x * 8 + warp; r < F.n_reads; r += gridDim.x * 8) {
        const uint32_t mol = F.read_mol[r];
        const uint32_t L = F.mol_raw_len[mol] - (F.mol_tail_len ? F.mol_tail_len[mol] : 0u);
        const uint32_t Lp = F.read_len ? min(F.read_len[r], (uint32_t)(F.seq_off[r + 1] - F.seq_off[r])) : L;
        const uint32_t id0 = F.mol_id_begin[mol], idl = F.mol_id_begin[mol + 1] - id0;
        uint32_t hl = 0;
        if (lane == 0) {
            const uint64_t ordinal = F.ordinal_base + r;
            uint4 u = philox4x32_10(1, TAG_READ, (uint32_t)ordinal, (uint32_t)(ordinal >> 32), F.seed_lo, F.seed_hi);
            uint32_t w[4] = {u.x, u.y, u.z, u.w};
            uint32_t p = 0;
            hdr[p++] = F.fastq ? '@' : '>';
            for (int d = 0; d < 32; d++) {
                if (d == 8 || d == 12 || d == 16 || d == 20) hdr[p++] = '-';
                hdr[p++] = (uint8_t)hexdig((w[d >> 3] >> (28 - 4 * (d & 7))) & 15u);
            }
            auto puts_ = [&](const char* s) {
                while (*s) hdr[p++] = (uint8_t)*s++;
            };
            auto putu = [&](uint32_t v) {
                uint32_t nd = dec_digits(v);
                for (uint32_t d = 0; d < nd; d++) {
                    hdr[p + nd - 1 - d] = (uint8_t)('0' + v % 10);
                    v /= 10;
                }
                p += nd;
            };
            puts_(" length=");
            putu(Lp);
            puts_(" error_free_length=");
            putu(L);
            puts_(" read_identity=");
            uint32_t h = F.read_len ? read_pct(F, r) : 10000u;
            putu(h / 100);
            hdr[p++] = '.';
            hdr[p++] = (uint8_t)('0' + (h / 10) % 10);
            hdr[p++] = (uint8_t)('0' + h % 10);
            puts_("% molecule_id=");
            hl = p;
        }
        hl = __shfl_sync(0xffffffffu, hl, 0);
        __syncwarp();
        uint8_t* o = F.out + F.rec_off[r];
        warp_copy(o, hdr, hl, (uint32_t)lane);
        o += hl;
        for (uint32_t x = lane; x < idl; x += 32) o[x] = F.id_pool[id0 + x];
        o += idl;
        if (lane == 0) o[0] = '\n';
        o += 1;
        const uint8_t* s = F.seq + F.seq_off[r];
        warp_copy(o, s, Lp, (uint32_t)lane);
        o += Lp;
        if (lane == 0) o[0] = '\n';
        o += 1;
        if (F.fastq) {
            if (lane == 0) {
                o[0] = '+';
                o[1] = '\n';
            }
            o += 2;
            if (F.qual) {
                const uint8_t* q = F.qual + F.seq_off[r];
                warp_copy(o, q, Lp, (uint32_t)lane);
            } else {
                for (uint32_t x = lane; x < Lp; x += 32) o[x] = 'K';
            }
            o += Lp;
            if (lane == 0) o[0] = '\n';
        }
        __syncwarp();
    }
}

}  // namespace tk
