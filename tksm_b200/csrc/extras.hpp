// Badread extras of the Seq hot path (host side): junk reads, random reads, adapters, glitches, chimeras.
//
// north_star names them as part of the path; the reference does not have them (py/tksm_badread.py holds only the two
// unused constants CHIMERA_START_ADAPTER_CHANCE / CHIMERA_END_ADAPTER_CHANCE, :61-63, and the upstream Badread source
// is not in /root/reference), so there is nothing to pin them against: PARITY UNPINNED.  They follow the semantics
// Badread publishes for `badread simulate` (--junk_reads, --random_reads, --chimeras, --start_adapter, --end_adapter,
// --glitches), restated as the statistical definition in DESIGN.md 3.10, and are OFF by default: with every rate at
// zero the output is byte-identical to a build without them.
//
// Like the tail-noise model they only change what the fragment of a read is made of, so they act on the molecule's
// segment list on the host (literal segments added, intervals cut) and the kernels see an ordinary molecule.  Every
// draw comes from Philox keyed by (seed, global read ordinal):
//   counter (0, TAG_EXTRA | 0 << 8, ordinal)  w0: junk / random / real   w1: junk unit length   w2: junk unit bases
//                                             w3: chimera or not
//   counter (i >> 6, TAG_EXTRA | 1 << 8, ordinal)  word (i >> 4) & 3, bits 2 (i & 15): base i of a random read
//   counter (0, TAG_EXTRA | 2 << 8, ordinal)  w0: start adapter or not  w1: its amount  w2: end adapter or not
//                                             w3: its amount
//   counter (0, TAG_EXTRA | 3 << 8 | g << 12, ordinal)  glitch g: w0 distance, w1 size, w2 skip
//   counter (i >> 6, TAG_EXTRA | 4 << 8 | g << 12, ordinal)  base i of the sequence glitch g inserts
//   counter (0, TAG_EXTRA | 5 << 8, ordinal)  chimera junction: w0: end adapter of the first part or not
//                                             w1: start adapter of the second part or not
// oracle/extras_oracle.py restates the same on plain strings; tests compare the two byte for byte.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

#include "host.hpp"
#include "tail.hpp"

namespace tk {

constexpr uint32_t TAG_EXTRA = 5;   // (common.hpp: TAG_READ .. TAG_TAIL = 0 .. 4)

struct ExtrasConfig {
    double junk = 0, random = 0, chimera = 0;              // fractions of reads (0 .. 1)
    double start_rate = 0, start_amount = 0;               // fraction of reads with a start adapter; mean fraction of it present
    double end_rate = 0, end_amount = 0;
    double glitch_rate = 0, glitch_size = 0, glitch_skip = 0;   // mean distance between glitches (0: none), mean bases inserted / lost
    std::vector<uint8_t> start_seq, end_seq;               // adapters as fragment codes
    bool any() const {
        return junk > 0 || random > 0 || chimera > 0 || (start_rate > 0 && !start_seq.empty()) ||
               (end_rate > 0 && !end_seq.empty()) || glitch_rate > 0;
    }
};

// ---------------------------------------------------------------- fused front-end steps
// The modules that usually sit between the transcript sampler and Seq only edit the segment list of a molecule
// record; as optional in-process steps they save the MDF round trips of a pipeline.  Each follows its module:
//   plA  src/polyA.cpp:133-148   a literal 'A' run of dist() bases, clamped to [min, max], appended (plus strand)
//   Flp  src/strand_man.cpp:47-53, src/interval.h:908-920   with probability p the segments are reversed and every
//        strand is toggled
//   Tag  src/tag.cpp:84-114, src/util.h:53-92 (fmt2seq)   a sequence drawn from the 5' format (IUPAC letters: one of
//        the letter's bases each; a number n: n times N) prepended, one from the 3' format appended
//   SCB  src/scb.cpp:73-81   the first value of the record's CB= comment appended, unless it is "."
// Steps run in the order given, once per MDF record (the `depth` copies of a record share the result), before the
// per-read extras.  The modules draw from std::mt19937 streams; here the draws of a record come from Philox keyed
// by the ordinal of the record's first read: counter (i, TAG_FUSE | step << 8, ordinal), so a fused run matches a
// pipeline of the standalone modules in distribution, not byte for byte.
constexpr uint32_t TAG_FUSE = 6;

struct FrontStep {
    enum Kind { POLYA, FLIP, TAG, SCB } kind = POLYA;
    enum Dist { NORMAL, POISSON, GAMMA, WEIBULL } dist = NORMAL;
    double p1 = 0, p2 = 0;             // distribution parameters (plA); p1 = flip probability (Flp)
    int min_len = 0, max_len = 5000;   // src/polyA.cpp defaults
    std::string fmt5, fmt3;            // Tag formats (numbers already expanded to N runs)
};
struct FrontEnd {
    std::vector<FrontStep> steps;
    bool any() const { return !steps.empty(); }
    bool needs_cb() const {
        for (const FrontStep& s : steps)
            if (s.kind == FrontStep::SCB) return true;
        return false;
    }
};
// "plA:normal=15,7.5[,min,max];Flp:0.5;Tag:FMT5,FMT3;SCB" (an empty format is allowed: "Tag:,10"); throws on a bad spec
FrontEnd parse_front_end(const std::string& spec);

// the molecule a chimeric read continues into: the molecule of the read before it (in global read order), after the
// fused front-end steps
struct ChimeraCarry {
    bool have = false;
    std::vector<uint64_t> seg_src;      // literal offsets relative to `lit`
    std::vector<uint32_t> seg_len;
    std::vector<uint32_t> seg_mod_begin{0};
    std::vector<uint32_t> mod_pos;
    std::vector<uint8_t> mod_chr;
    std::vector<uint8_t> lit;
};

// One molecule per read with the extras and the tail of that read applied; mol_tail_len filled when a tail model
// is loaded.  `ordinal_base` is the global ordinal of the first read of the first part.  `carry` (may be null: no
// chimeras across calls) holds the molecule of the read before the first one and is left holding the last one.
HostBatch apply_read_transforms(const TailModel& T, const ExtrasConfig& X, const FrontEnd& F, uint64_t seed,
                                uint64_t ordinal_base, const std::vector<const HostBatch*>& parts, ChimeraCarry* carry);

}  // namespace tk
