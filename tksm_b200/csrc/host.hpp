// Host-side data model of the Seq hot path: reference genome store, MDF parser, molecule batch.
//
// Replaces (reference, /root/reference):
//   generate_fasta / get_reference_seqs   py/sequence.py:168-194
//   mdf_generator                         py/sequence.py:197-221
//   the host part of mdf_to_seq           py/sequence.py:303-313 (contig lookup, slice clamping,
//                                         modification parsing; the byte work runs on the GPU)
#pragma once
#include <cstdint>
#include <string>
#include <unordered_map>
#include <vector>

#include "common.hpp"

namespace tk {

struct Contig {
    uint64_t offset;  // global base coordinate of base 0
    uint64_t length;
};

// Whole reference as 2-bit codes plus a sorted run list for every byte that is not ACGT
// after upper-casing (N runs, IUPAC codes, anything else the FASTA holds).
struct GenomeStore {
    std::unordered_map<std::string, Contig> contigs;
    // bit min(len, 63) set for every contig-name length: most literal segment names (barcodes, UMIs, adapters)
    // are ruled out as contigs without hashing them
    uint64_t name_len_mask = 0;
    void note_name(size_t len) { name_len_mask |= 1ull << (len < 63 ? len : 63); }
    bool name_len_possible(size_t len) const { return (name_len_mask >> (len < 63 ? len : 63)) & 1ull; }
    std::vector<uint32_t> words;        // 16 bases per word
    uint64_t n_bases = 0;               // global coordinate space used so far
    std::vector<uint64_t> exc_start;
    std::vector<uint32_t> exc_len;
    std::vector<uint8_t> exc_chr;
    // one bit per 2^EXC_BLOCK_SHIFT bases: set when an exception run touches the block (built lazily by
    // index_exceptions(); lets overlaps_exception() answer "no" without a search for most intervals)
    static constexpr int EXC_BLOCK_SHIFT = 14;
    std::vector<uint64_t> exc_block_bits;
    size_t exc_indexed = 0;   // number of runs covered by exc_block_bits
    void index_exceptions();

    void add_contig(const std::string& name, const char* seq, size_t len);
    // streaming form of add_contig (load_fasta_text packs the sequence lines where they lie)
    struct ExcRun {
        uint64_t start = 0;
        uint32_t len = 0;
        uint8_t chr = 0;   // the run's byte, upper-cased
        uint8_t raw = 0;   // its first byte as written
        bool one_raw = false;   // every byte of the run is `raw` (no mixed case)
    };
    void begin_contig();
    void append_bases(const char* seq, size_t len, ExcRun& run);
    void end_contig(const std::string& name, uint64_t off, ExcRun& run);
    void load_fasta(const std::string& path);              // plain or .gz
    void load_fasta_text(const char* text, size_t len);
    bool overlaps_exception(uint64_t start, uint64_t len) const;
    // test / oracle helper: the upper-cased bytes of [start, start + len)
    std::string fetch(uint64_t start, uint64_t len) const;
};

struct HostBatch {
    uint64_t ordinal_base = 0;
    std::vector<uint32_t> read_mol;
    std::vector<uint32_t> mol_seg_begin{0};
    std::vector<uint32_t> mol_raw_len;
    std::vector<uint32_t> mol_tail_len;   // empty unless a tail-noise model appended tails (tail.cpp): bases of mol_raw_len that are tail
    std::vector<uint32_t> mol_id_begin{0};
    std::vector<uint64_t> seg_src;
    std::vector<uint32_t> seg_len;
    std::vector<uint32_t> seg_mod_begin{0};
    std::vector<uint32_t> mod_pos;
    std::vector<uint8_t> mod_chr;
    std::vector<uint8_t> lit_pool;
    std::vector<uint8_t> id_pool;
    // first value of the CB= comment of every molecule (MdfParser::capture_cb; the fused SCB step appends it)
    std::vector<uint32_t> mol_cb_begin{0};
    std::vector<uint8_t> cb_pool;
    uint64_t total_raw_bases = 0;   // over reads (depth expanded)
    uint64_t genomic_bases = 0, literal_bases = 0;  // over reads, for the algorithmic-bytes figure
    // literal de-duplication: direct-mapped on a hash of the encoded bytes (repeated tags and poly(A) runs share
    // pool bytes; unique barcodes / UMIs overwrite a slot and cost no allocation)
    struct LitSlot {
        uint64_t hash = 0;
        uint32_t off = 0, len = 0;   // len == 0: empty
    };
    std::vector<LitSlot> lit_index;

    size_t n_reads() const { return read_mol.size(); }
    size_t n_mols() const { return mol_raw_len.size(); }
    void clear();
    void append(const HostBatch& other);  // concatenates another batch parsed from the next text range
};

// Streaming MDF parser.  feed() consumes complete records only and returns the number of bytes
// consumed; the caller keeps the tail and calls again with more data (FIFO-safe, never seeks).
// With `final` set the remaining bytes are parsed as the last record(s).
// Throws std::runtime_error where the reference would raise (py/sequence.py:205-218, 229-239).
// One deliberate leniency: an interval line of four fields (no modifications column) is taken as one with an empty
// fifth field, as the reference's C++ reader does (src/mdf.h:88-90: the modules before Seq write such lines);
// py/sequence.py:215 would raise ValueError.  Tested in tests/test_gpu_parity.py (handmade record `nomods4`).
struct MdfParser {
    const GenomeStore* genome;
    bool capture_cb = false;   // keep the first value of the header's CB= comment per molecule (fused SCB step)
    explicit MdfParser(const GenomeStore* g) : genome(g) {}
    size_t feed(const char* text, size_t len, bool final, HostBatch& out, size_t max_reads = SIZE_MAX,
                uint64_t max_bases = UINT64_MAX);
};

// Contiguous shards of about equal simulated bases, cut at record boundaries (multi-GPU runs).
// byte_begin / ordinal_base get world + 1 entries.
void plan_shards(const char* text, size_t len, int world, std::vector<uint64_t>& byte_begin,
                 std::vector<uint64_t>& ordinal_base);

// Fragment code of one byte: ACGT -> 0..3, anything else keeps its (ASCII) value.
inline uint8_t frag_code(uint8_t c) {
    switch (c) {
        case 'A': return 0;
        case 'C': return 1;
        case 'G': return 2;
        case 'T': return 3;
        default: return c;
    }
}
inline uint8_t ascii_upper(uint8_t c) { return (c >= 'a' && c <= 'z') ? (uint8_t)(c - 32) : c; }

}  // namespace tk
