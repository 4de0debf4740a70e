// Host emulation of the team-parallel alignment of tksm_b200/csrc/align.cuh (TEST CODE).
//
// The per-thread chunk functions are the product's own (compiled here for the host); this file only
// replays what the CUDA team does with them: every "thread" of a team of `threads` lanes assumes
// the state at its chunk edge, chunks are redone until the assumptions agree with the neighbours'
// results, then the traceback does the same from the other end.  tests/test_align.py checks the
// outcome against the integer DP of oracle/twin.c for every read the twin aligns.
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "../../tksm_b200/csrc/align.cuh"

using namespace tk;

template <typename BV>
static int run(const uint8_t* frag, const uint8_t* slen, uint8_t* rb, uint32_t n, uint32_t m, uint32_t threads,
               uint32_t* eq_out, uint32_t* cols_out, uint32_t* rounds) {
    AlignCtx<BV> A;
    uint32_t lanes;
    align_chunking(n, threads, 8 * (uint32_t)sizeof(BV), A.chunk, lanes);
    A.frag = frag;
    A.slen = slen;
    A.rb = rb;
    A.out = rb;
    A.n = n;
    A.m = m;
    A.stride = threads;
    // the trace array is reused from read to read on the device: start from garbage
    std::vector<AlignCell<BV>> trace((size_t)A.chunk * threads + 1);
    memset((void*)trace.data(), 0x5A, trace.size() * sizeof(AlignCell<BV>));
    A.trace = trace.data();
    A.cells = (uint32_t)trace.size();
    uint32_t oob = 0;
    A.oob = &oob;
    std::vector<AlignLane<BV>> ln(lanes);
    uint32_t o = 0;
    for (uint32_t l = 0; l < lanes; l++) {
        ln[l].c0 = 1 + l * A.chunk;
        ln[l].c1 = ln[l].c0 + A.chunk < n + 1 ? ln[l].c0 + A.chunk : n + 1;
        ln[l].ob = o;
        for (uint32_t j = ln[l].c0; j < ln[l].c1; j++) o += slen[j - 1];
    }
    if (o != m) return -1;
    for (uint32_t l = 0; l < lanes; l++) align_lane_init(A, ln[l]);
    // forward fixpoint
    std::vector<BV> inP(lanes), inM(lanes), outP(lanes), outM(lanes);
    std::vector<char> need(lanes, 1);
    for (uint32_t l = 0; l < lanes; l++) {
        if (l == 0) { inP[l] = ~BV(0); inM[l] = 0; }
        else align_ideal_state<BV>(ln[l].ob, inP[l], inM[l]);
    }
    rounds[0] = 0;
    for (;;) {
        for (uint32_t l = 0; l < lanes; l++)
            if (need[l]) {
                BV P = inP[l], M = inM[l];
                const bool conv = rounds[0] == 0 ? align_forward_chunk<BV, false>(A, ln[l], P, M)
                                                 : align_forward_chunk<BV, true>(A, ln[l], P, M);
                if (!conv) { outP[l] = P; outM[l] = M; }
            }
        rounds[0]++;
        bool any = false;
        for (uint32_t l = lanes; l-- > 1;) {
            need[l] = outP[l - 1] != inP[l] || outM[l - 1] != inM[l];
            if (need[l]) { inP[l] = outP[l - 1]; inM[l] = outM[l - 1]; any = true; }
        }
        need[0] = 0;
        if (!any) break;
    }
    // traceback fixpoint (no writes), then the writing pass
    struct Cell { uint32_t i, j, o; };
    std::vector<Cell> ent(lanes), ext(lanes);
    std::vector<char> ok(lanes, 1);
    std::vector<uint32_t> tag(lanes, 1), emin(lanes, 1);
    for (uint32_t l = 0; l < lanes; l++) {
        if (l == lanes - 1) ent[l] = {m, n, m};
        else ent[l] = {ln[l + 1].ob, ln[l].c1 - 1, ln[l + 1].ob};
        need[l] = 1;
    }
    rounds[1] = 0;
    for (;;) {
        for (uint32_t l = 0; l < lanes; l++)
            if (need[l]) {
                Cell c = ent[l];
                uint32_t e = 0, k = 0;
                bool merged = false;
                if (rounds[1] == 0) {
                    ok[l] = align_trace_chunk<BV, TRACE_FIRST>(A, ln[l].c0, l == 0, l == lanes - 1, c.i, c.j, c.o, e, k, merged, 1, 1);
                    ext[l] = c;
                } else {
                    tag[l]++;
                    const bool r = align_trace_chunk<BV, TRACE_REDO>(A, ln[l].c0, l == 0, l == lanes - 1, c.i, c.j, c.o, e, k, merged,
                                                                     tag[l], emin[l]);
                    if (!merged) { ok[l] = r; ext[l] = c; emin[l] = tag[l]; }
                }
            }
        rounds[1]++;
        bool any = false;
        for (uint32_t l = 0; l + 1 < lanes; l++) {
            const Cell& x = ext[l + 1];
            need[l] = x.i != ent[l].i || x.j != ent[l].j || x.o != ent[l].o;
            if (need[l]) { ent[l] = x; any = true; }
        }
        need[lanes - 1] = 0;
        if (!any) break;
    }
    for (uint32_t l = 0; l < lanes; l++)
        if (!ok[l]) return 1;
    uint32_t eq = 0, cols = 0;
    for (uint32_t l = 0; l < lanes; l++) {
        Cell c = ent[l];
        bool merged;
        if (!align_trace_chunk<BV, TRACE_WRITE>(A, ln[l].c0, l == 0, l == lanes - 1, c.i, c.j, c.o, eq, cols, merged)) return 1;
    }
    *eq_out = eq;
    *cols_out = cols;
    return oob ? -2 : 0;
}

// wide: 0 = 32-bit band, 1 = 64-bit band.  Returns 0 ok, 1 band left, -1 inconsistent input.
extern "C" int align_emul(const uint8_t* frag, const uint8_t* slen, uint8_t* rb, uint32_t n, uint32_t m, uint32_t threads,
                          int wide, uint32_t* eq, uint32_t* cols, uint32_t* rounds) {
    return wide ? run<uint64_t>(frag, slen, rb, n, m, threads, eq, cols, rounds)
                : run<uint32_t>(frag, slen, rb, n, m, threads, eq, cols, rounds);
}
