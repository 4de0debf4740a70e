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
    const uint32_t warm = getenv("ALIGN_WARM") ? (uint32_t)atoi(getenv("ALIGN_WARM")) : ALIGN_WARMUP;
    AlignCtx<BV> A;
    uint32_t lanes;
    align_chunking(n, threads, A.chunk, lanes, A.magic);
    A.frag = frag;
    A.slen = slen;
    A.rb = rb;
    A.n = n;
    A.m = m;
    A.stride = threads;
    std::vector<BV> trace((size_t)A.chunk * threads * 2 + 2, 0);
    A.trace = trace.data();
    std::vector<uint32_t> c0(lanes), c1(lanes), ob(lanes + 1);
    uint32_t o = 0;
    for (uint32_t l = 0; l < lanes; l++) {
        c0[l] = 1 + l * A.chunk;
        c1[l] = c0[l] + A.chunk < n + 1 ? c0[l] + A.chunk : n + 1;
        ob[l] = o;
        for (uint32_t j = c0[l]; j < c1[l]; j++) o += slen[j - 1];
    }
    ob[lanes] = o;
    if (o != m) return -1;
    // forward fixpoint
    std::vector<BV> inP(lanes), inM(lanes), outP(lanes), outM(lanes);
    std::vector<char> need(lanes, 1);
    for (uint32_t l = 0; l < lanes; l++) {
        if (l == 0) { inP[l] = ~BV(0); inM[l] = 0; }
        else {
            // warm-up: the state at the chunk edge is guessed by running a few columns up to it from an
            // undisturbed diagonal
            const uint32_t wu = c0[l] - 1 < warm ? c0[l] - 1 : warm;
            uint32_t ow = ob[l];
            for (uint32_t j = c0[l] - wu; j < c0[l]; j++) ow -= slen[j - 1];
            BV P, M;
            if (c0[l] - wu == 1) { P = ~BV(0); M = 0; }
            else align_ideal_state<BV>(ow, P, M);
            if (wu) align_forward_chunk<BV, false>(A, c0[l] - wu, c0[l], ow, P, M);
            inP[l] = P; inM[l] = M;
        }
    }
    rounds[0] = 0;
    for (;;) {
        rounds[0]++;
        for (uint32_t l = 0; l < lanes; l++)
            if (need[l]) {
                BV P = inP[l], M = inM[l];
                align_forward_chunk(A, c0[l], c1[l], ob[l], P, M);
                outP[l] = P; outM[l] = M;
            }
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
    for (uint32_t l = 0; l < lanes; l++) {
        if (l == lanes - 1) ent[l] = {m, n, m};
        else ent[l] = {ob[l + 1], c1[l] - 1, ob[l + 1]};
        need[l] = 1;
    }
    rounds[1] = 0;
    for (;;) {
        rounds[1]++;
        for (uint32_t l = 0; l < lanes; l++)
            if (need[l]) {
                Cell c = ent[l];
                uint32_t e = 0, k = 0;
                ok[l] = align_trace_chunk<BV, false>(A, c0[l], l == 0, l == lanes - 1, c.i, c.j, c.o, e, k);
                ext[l] = c;
            }
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
        if (!align_trace_chunk<BV, true>(A, c0[l], l == 0, l == lanes - 1, c.i, c.j, c.o, eq, cols)) return 1;
    }
    *eq_out = eq;
    *cols_out = cols;
    return 0;
}

// wide: 0 = 32-bit band, 1 = 64-bit band.  Returns 0 ok, 1 band left, -1 inconsistent input.
extern "C" int align_emul(const uint8_t* frag, const uint8_t* slen, uint8_t* rb, uint32_t n, uint32_t m, uint32_t threads,
                          int wide, uint32_t* eq, uint32_t* cols, uint32_t* rounds) {
    return wide ? run<uint64_t>(frag, slen, rb, n, m, threads, eq, cols, rounds)
                : run<uint32_t>(frag, slen, rb, n, m, threads, eq, cols, rounds);
}
