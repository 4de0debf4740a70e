// Tail-noise model of the Seq hot path (host side): extra bases between the molecule and the right flank.
//
// Replaces (reference, /root/reference):
//   KDE_noise_generator.load / noise_seq     py/tksm_badread.py:886-959
//   CustomDist, Custom2Dist.__call__         py/tksm_badread.py:970-1036
//   the call site                            py/tksm_badread.py:335-341
//
// The reference draws from Python's `random`; here the draws of a read come from Philox keyed by
// (seed, global read ordinal), like every other draw of the path:
//   counter (3, TAG_READ, ordinal)  w0: tail or no tail (ratio)  w1: length (inverse CDF)  w2 & 3: start state
//   counter (i >> 2, TAG_TAIL, ordinal) word i & 3: base i of the Markov chain
// The tail travels to the device as one more literal segment of the read's molecule, so the kernels
// see it as part of the fragment, exactly where the reference puts it.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

#include "host.hpp"

namespace tk {

struct TailModel {
    bool loaded = false;              // false: no_noise (Mock_noise_generator)
    std::vector<double> lx, ly;       // tail lengths; fragment-length labels
    std::vector<std::vector<double>> cdf;   // one cumulative row per ly entry (CustomDist.cdf)
    double trans[4][4] = {};          // transition weights, row = previous state
    double ratio = 0;                 // fraction of reads with a tail
    uint8_t bases[4] = {'A', 'G', 'T', 'C'};

    void load_json(const std::string& text);                      // throws std::runtime_error
    // length of the tail of read `ordinal` whose molecule has frag_len bases (0: none); start = first chain state
    uint32_t draw_length(uint64_t seed, uint64_t ordinal, uint32_t frag_len, uint32_t* start) const;
    // the tail bases as fragment codes (host.hpp: frag_code)
    void draw_bases(uint64_t seed, uint64_t ordinal, uint32_t start, uint32_t n, uint8_t* out) const;
};

void philox4x32_10_host(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1, uint32_t out[4]);

// (the tail joins the molecule of its read in extras.cpp: apply_read_transforms)

}  // namespace tk
