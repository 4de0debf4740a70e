// MDF tokeniser on the device: the text of a batch of complete MDF records -> the molecule batch
// (structure of arrays, DevBatch) the simulate kernels read, without the host parser.
//
// Replaces, for input in the emitter's format (src/interval.h:898-905): mdf_generator and the host part of mdf_to_seq
// (py/sequence.py:197-221, 303-313), i.e. what host.cpp: MdfParser + engine.cu: prepare_host do on the host threads.
// Anything outside that format -- CR line ends, signed / underscored / spaced integers, malformed or empty lines,
// negative modification positions, more than 16 of them on a line, a depth expansion beyond the buffers -- raises a flag instead of
// being interpreted here, and the batch goes through the host parser, which stays the definition of the reference's
// error behaviour and of its corner cases (python slices with negative bounds, int() syntax).
//
// Passes (all sizes from the host's count of lines and header lines while it copies the text to page-locked memory):
//   mark_lines / line_starts     newline positions -> start offset of every line
//   parse_lines                  one thread per line: fields, integers, contig lookup (open-addressing hash of the
//                                reference's names), slice clamping, segment source word, modification count
//   header_records / keep_lines  record index of every line, depth > 0 filter
//   fill_segments / fill_records the arrays of DevBatch at the offsets prefix sums over lines / records give:
//                                segments, modifications, literal codes, ids, depth expansion, scratch offsets,
//                                length-class counts
#pragma once
#include <cstdint>

#include "common.hpp"

namespace tk {

constexpr uint32_t MDF_TILE = 256;   // bytes per thread of the newline passes

struct ContigTable {     // device copy of GenomeStore::contigs
    const uint64_t* slot_hash;   // 0 = empty
    const uint32_t* slot_contig;
    uint32_t mask;               // slots - 1
    const uint32_t* name_off;    // per contig, n + 1
    const uint8_t* names;
    const uint64_t* offset;
    const uint64_t* length;
    uint64_t name_len_mask;      // GenomeStore::name_len_mask
};

__host__ __device__ inline uint64_t mdf_name_hash(const uint8_t* p, uint32_t n) {
    uint64_t h = 0xCBF29CE484222325ull;
    for (uint32_t i = 0; i < n; i++) h = (h ^ p[i]) * 0x100000001B3ull;
    return h | 1ull;
}

struct MdfTotals {       // device -> host, once per batch
    uint64_t reads, mols, segs, mods, lit_bytes, id_bytes;
    uint64_t scr_total, raw_total, total_raw_bases, gen_bases, lit_bases, per_read_segs, per_read_mods, idsum;
    uint32_t n_work[8], n_work_grow[8];
    uint32_t max_n_gws, max_n_gws_grow;
    uint32_t bad;        // != 0: the batch must go through the host parser (first reason seen)
    uint32_t pad;
};

struct MdfParams {
    const uint8_t* text;
    uint64_t n_bytes;
    uint32_t n_lines, n_records;   // counted by the host (records: header lines, whatever their depth)
    uint32_t read_cap;             // capacity of the per-read arrays
    ContigTable ct;
    const uint64_t* exc_start;     // DevTables: exception runs of the genome (sorted)
    const uint32_t* exc_len;
    uint32_t n_exc;
    uint32_t k;                    // error-model k-mer size (fragment = molecule + 2 k flank slots)
    uint32_t ncap[5];              // length-class capacities
    // per line
    uint64_t* line_start;          // n_lines + 1
    uint8_t* kind;                 // 0 header, 1 interval
    uint32_t* a32;                 // header: depth; interval: segment length
    uint32_t* b32;                 // header: id length; interval: modification count
    uint32_t* c32;                 // interval: literal bytes (0 for a genomic segment)
    uint64_t* src;                 // interval: segment source word (literal offset still missing)
    uint32_t* is_hdr;              // scan input
    uint64_t* hdr_excl;            // n_lines + 1: headers before the line
    uint32_t* keep_seg;            // scan inputs per line ...
    uint32_t* nm_k;
    uint32_t* lit_k;
    uint32_t* len_k;
    uint64_t* seg_idx;             // ... and their exclusive prefix sums (n_lines + 1 each)
    uint64_t* mod_off;
    uint64_t* lit_off;
    uint64_t* len_prefix;
    // per record
    uint32_t* rec_line;
    uint32_t* rec_keep;
    uint32_t* rec_depth;
    uint32_t* rec_idlen;
    uint32_t* rec_slot;            // depth x scratch slot, depth x raw slot
    uint32_t* rec_rawslot;
    uint64_t* mol_of_rec;          // n_records + 1 each
    uint64_t* read_of_rec;
    uint64_t* id_of_rec;
    uint64_t* so_of_rec;
    uint64_t* ro_of_rec;
    // outputs (DevBatch arrays)
    uint32_t* read_mol;
    uint32_t* mol_seg_begin;
    uint32_t* mol_raw_len;
    uint32_t* mol_id_begin;
    uint64_t* seg_src;
    uint32_t* seg_len;
    uint32_t* seg_mod_begin;
    uint32_t* mod_pos;
    uint8_t* mod_chr;
    uint8_t* lit_pool;
    uint8_t* id_pool;
    uint64_t* scr_off;
    uint64_t* raw_off;
    MdfTotals* totals;
};

enum : uint32_t {
    MDF_BAD_CR = 1, MDF_BAD_EMPTY = 2, MDF_BAD_FIELDS = 3, MDF_BAD_INT = 4, MDF_BAD_MOD = 5, MDF_BAD_RANGE = 6,
    MDF_BAD_READS = 7, MDF_BAD_HEADER = 8,
};

__device__ inline void mdf_flag(const MdfParams& P, uint32_t why) { atomicCAS(&P.totals->bad, 0u, why); }

// ---- newline passes
__global__ void mdf_mark_lines(const uint8_t* text, uint64_t n, uint32_t* tile_count) {
    const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const uint64_t b0 = t * MDF_TILE;
    if (b0 >= n) return;
    const uint64_t b1 = b0 + MDF_TILE < n ? b0 + MDF_TILE : n;
    uint32_t c = 0;
    for (uint64_t i = b0; i < b1; i++) c += text[i] == '\n';
    tile_count[t] = c;
}
// line 0 starts at byte 0; every newline starts the next line (the text ends with a newline: the host sees to it)
__global__ void mdf_line_starts(const uint8_t* text, uint64_t n, const uint64_t* tile_off, uint64_t* line_start, uint32_t n_lines) {
    const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const uint64_t b0 = t * MDF_TILE;
    if (t == 0) line_start[0] = 0;
    if (b0 >= n) return;
    const uint64_t b1 = b0 + MDF_TILE < n ? b0 + MDF_TILE : n;
    uint64_t li = tile_off[t] + 1;
    for (uint64_t i = b0; i < b1; i++)
        if (text[i] == '\n') {
            if (li <= n_lines) line_start[li] = i + 1;
            li++;
        }
}

// ---- one thread per line
__device__ inline bool mdf_digits(const uint8_t* p, const uint8_t* e, uint64_t& v) {   // 1..18 decimal digits, nothing else
    if (p >= e || e - p > 18) return false;
    uint64_t x = 0;
    for (; p < e; p++) {
        const uint32_t d = (uint32_t)*p - '0';
        if (d > 9) return false;
        x = x * 10 + d;
    }
    v = x;
    return true;
}

__device__ inline bool mdf_overlaps_exception(const MdfParams& P, uint64_t start, uint64_t len) {
    if (P.n_exc == 0 || len == 0) return false;
    uint32_t lo = 0, hi = P.n_exc;   // first run with start >= `start`
    while (lo < hi) {
        const uint32_t mid = (lo + hi) >> 1;
        if (P.exc_start[mid] < start) lo = mid + 1; else hi = mid;
    }
    if (lo < P.n_exc && P.exc_start[lo] < start + len) return true;
    if (lo > 0 && P.exc_start[lo - 1] + P.exc_len[lo - 1] > start) return true;
    return false;
}

// modification list "12C,40T": count (a position listed twice counts once), validity (digits + one character,
// position inside the slice, at most 16 positions);
// with out_pos != nullptr the entries are written too
__device__ inline int mdf_mods(const uint8_t* p, const uint8_t* e, uint32_t slen, uint32_t* out_pos, uint8_t* out_chr) {
    uint32_t seen[16];
    int n = 0;
    while (p <= e) {
        const uint8_t* c = p;
        while (c < e && *c != ',') c++;
        if (c - p < 2) return -1;
        uint64_t pos;
        if (!mdf_digits(p, c - 1, pos) || pos >= slen) return -1;
        if (n >= 16) return -1;
        int at = n;
        for (int x = 0; x < n; x++)
            if (seen[x] == (uint32_t)pos) at = x;   // listed order: the last entry of a position wins (host.cpp: MdfParser)
        if (out_pos) {
            out_pos[at] = (uint32_t)pos;
            out_chr[at] = c[-1];
        }
        if (at == n) {
            seen[n] = (uint32_t)pos;
            n++;
        }
        if (c >= e) break;
        p = c + 1;
    }
    return n;
}

__global__ void mdf_parse_lines(const MdfParams P) {
    const uint32_t li = blockIdx.x * blockDim.x + threadIdx.x;
    if (li >= P.n_lines) return;
    const uint8_t* b = P.text + P.line_start[li];
    const uint8_t* e = P.text + P.line_start[li + 1] - 1;   // the newline
    P.kind[li] = 2;
    P.is_hdr[li] = 0;
    P.a32[li] = P.b32[li] = P.c32[li] = 0;
    P.src[li] = 0;
    if (e <= b || *b == '\t') {   // empty line, empty first field: line[0][0] raises (py/sequence.py:207)
        mdf_flag(P, MDF_BAD_EMPTY);
        return;
    }
    const uint8_t* tab[5];
    int ntab = 0;
    for (const uint8_t* p = b; p < e; p++) {
        if (*p == '\t') {
            if (ntab < 5) tab[ntab] = p;
            ntab++;
        } else if (*p == '\r') {
            mdf_flag(P, MDF_BAD_CR);
            return;
        }
    }
    if (*b == '+') {
        uint64_t depth;
        if (ntab == 0 || !mdf_digits(tab[0] + 1, ntab > 1 ? tab[1] : e, depth)) {
            mdf_flag(P, MDF_BAD_HEADER);
            return;
        }
        if (depth > 0x3FFFFFFFull) {
            mdf_flag(P, MDF_BAD_READS);
            return;
        }
        P.kind[li] = 0;
        P.is_hdr[li] = 1;
        P.a32[li] = (uint32_t)depth;
        P.b32[li] = (uint32_t)(tab[0] - (b + 1));
        return;
    }
    const int nf = ntab + 1;
    if (nf < 4 || nf > 5) {
        mdf_flag(P, MDF_BAD_FIELDS);
        return;
    }
    uint64_t start, end;
    if (!mdf_digits(tab[0] + 1, tab[1], start) || !mdf_digits(tab[1] + 1, tab[2], end)) {
        mdf_flag(P, MDF_BAD_INT);
        return;
    }
    const uint8_t* f3e = nf == 5 ? tab[3] : e;
    const bool plus = (f3e - (tab[2] + 1) == 1) && tab[2][1] == '+';
    const uint32_t clen = (uint32_t)(tab[0] - b);
    // contig or "the name is the sequence" (py/sequence.py:307)
    uint64_t n = clen, base = 0;
    bool literal = true;
    if ((P.ct.name_len_mask >> (clen < 63 ? clen : 63)) & 1ull) {
        const uint64_t h = mdf_name_hash(b, clen);
        for (uint32_t s = (uint32_t)(h >> 17) & P.ct.mask;; s = (s + 1) & P.ct.mask) {
            const uint64_t sh = P.ct.slot_hash[s];
            if (sh == 0) break;
            if (sh == h) {
                const uint32_t c = P.ct.slot_contig[s];
                const uint32_t o = P.ct.name_off[c];
                bool same = P.ct.name_off[c + 1] - o == clen;
                for (uint32_t i = 0; same && i < clen; i++) same = P.ct.names[o + i] == b[i];
                if (same) {
                    literal = false;
                    n = P.ct.length[c];
                    base = P.ct.offset[c];
                    break;
                }
            }
        }
    }
    const uint64_t s = start < n ? start : n;
    uint64_t t = end < n ? end : n;
    if (t < s) t = s;
    if (t - s > 0xFFFFFF00ull) {
        mdf_flag(P, MDF_BAD_RANGE);
        return;
    }
    const uint32_t slen = (uint32_t)(t - s);
    int nm = 0;
    if (nf == 5 && e > tab[3] + 1) {
        nm = mdf_mods(tab[3] + 1, e, slen, nullptr, nullptr);
        if (nm < 0) {
            mdf_flag(P, MDF_BAD_MOD);
            return;
        }
    }
    uint64_t src;
    if (literal) {
        src = SEG_LITERAL | (uint64_t)(P.line_start[li] + s);   // text offset for now; fill_segments puts the pool offset in
    } else {
        src = base + s;
        if (mdf_overlaps_exception(P, base + s, slen)) src |= SEG_HAS_EXC;
    }
    if (!plus) src |= SEG_MINUS;
    P.kind[li] = 1;
    P.a32[li] = slen;
    P.b32[li] = (uint32_t)nm;
    P.c32[li] = literal ? slen : 0u;
    P.src[li] = src;
}

// ---- records: header line li is record hdr_excl[li]
__global__ void mdf_header_records(const MdfParams P) {
    const uint32_t li = blockIdx.x * blockDim.x + threadIdx.x;
    if (li >= P.n_lines || P.kind[li] != 0) return;
    const uint32_t rec = (uint32_t)P.hdr_excl[li];
    if (rec >= P.n_records) return;   // (cannot happen: the host counted the same header lines)
    const uint32_t depth = P.a32[li];
    P.rec_line[rec] = li;
    P.rec_keep[rec] = depth > 0;
    P.rec_depth[rec] = depth;
    P.rec_idlen[rec] = depth > 0 ? P.b32[li] : 0u;
}

// interval lines of records with depth > 0 are segments (py/sequence.py:205-218: a record is yielded depth times;
// lines before the first header are parsed and dropped)
__global__ void mdf_keep_lines(const MdfParams P) {
    const uint32_t li = blockIdx.x * blockDim.x + threadIdx.x;
    if (li >= P.n_lines) return;
    bool keep = false;
    if (P.kind[li] == 1) {
        const uint64_t h = P.hdr_excl[li];
        keep = h > 0 && P.rec_keep[h - 1];
    }
    P.keep_seg[li] = keep;
    P.nm_k[li] = keep ? P.b32[li] : 0u;
    P.lit_k[li] = keep ? P.c32[li] : 0u;
    P.len_k[li] = keep ? P.a32[li] : 0u;
}

__device__ inline uint8_t mdf_frag_code(uint8_t c) {
    if (c >= 'a' && c <= 'z') c = (uint8_t)(c - 32);
    switch (c) {
        case 'A': return 0;
        case 'C': return 1;
        case 'G': return 2;
        case 'T': return 3;
        default: return c;
    }
}

__global__ void mdf_fill_segments(const MdfParams P) {
    const uint32_t li = blockIdx.x * blockDim.x + threadIdx.x;
    if (li >= P.n_lines) return;
    if (li == 0) P.seg_mod_begin[0] = 0;
    if (!P.keep_seg[li]) return;
    const uint64_t si = P.seg_idx[li];
    uint64_t src = P.src[li];
    const uint32_t slen = P.a32[li];
    if (src & SEG_LITERAL) {
        const uint8_t* from = P.text + (src & SEG_OFF_MASK);
        const uint64_t off = P.lit_off[li];
        for (uint32_t i = 0; i < slen; i++) P.lit_pool[off + i] = mdf_frag_code(from[i]);
        src = (src & ~SEG_OFF_MASK) | off;
    }
    P.seg_src[si] = src;
    P.seg_len[si] = slen;
    const uint32_t nm = P.b32[li];
    const uint64_t m0 = P.mod_off[li];
    P.seg_mod_begin[si + 1] = (uint32_t)(m0 + nm);
    if (nm) {
        // the fifth field again, this time for its entries
        const uint8_t* e = P.text + P.line_start[li + 1] - 1;
        const uint8_t* p = e;
        int tabs = 0;
        for (const uint8_t* q = P.text + P.line_start[li]; q < e; q++)
            if (*q == '\t' && ++tabs == 4) {
                p = q + 1;
                break;
            }
        mdf_mods(p, e, slen, P.mod_pos + m0, P.mod_chr + m0);
    }
}

__host__ __device__ inline uint32_t mdf_align16(uint32_t v) { return (v + 15u) & ~15u; }

// scratch / raw slots of a record's reads (inputs of two more prefix sums) once its molecule length is known
__global__ void mdf_record_slots(const MdfParams P) {
    const uint32_t rec = blockIdx.x * blockDim.x + threadIdx.x;
    if (rec >= P.n_records) return;
    P.rec_slot[rec] = P.rec_rawslot[rec] = 0;
    if (!P.rec_keep[rec]) return;
    const uint32_t li = P.rec_line[rec];
    const uint32_t end_line = rec + 1 < P.n_records ? P.rec_line[rec + 1] : P.n_lines;
    const uint64_t L = P.len_prefix[end_line] - P.len_prefix[li];
    const uint64_t depth = P.rec_depth[rec];
    if (L > 0x7FFFFF00ull || depth * (2 * L + 48) > 0xFFFFFF00ull) {
        mdf_flag(P, MDF_BAD_RANGE);
        return;
    }
    P.rec_slot[rec] = (uint32_t)(depth * mdf_align16((uint32_t)(2 * L + 32)));
    P.rec_rawslot[rec] = (uint32_t)(depth * mdf_align16((uint32_t)L));
}

__global__ void mdf_fill_records(const MdfParams P) {
    const uint32_t rec = blockIdx.x * blockDim.x + threadIdx.x;
    if (rec >= P.n_records || !P.rec_keep[rec]) return;
    const uint32_t li = P.rec_line[rec];
    const uint32_t end_line = rec + 1 < P.n_records ? P.rec_line[rec + 1] : P.n_lines;
    const uint32_t mol = (uint32_t)P.mol_of_rec[rec];
    const uint32_t L = (uint32_t)(P.len_prefix[end_line] - P.len_prefix[li]);
    const uint32_t depth = P.rec_depth[rec];
    const uint64_t r0 = P.read_of_rec[rec];
    if (r0 + depth > P.read_cap) {
        mdf_flag(P, MDF_BAD_READS);
        return;
    }
    const uint32_t s0 = (uint32_t)P.seg_idx[li], s1 = (uint32_t)P.seg_idx[end_line];
    P.mol_seg_begin[mol] = s0;
    P.mol_raw_len[mol] = L;
    const uint64_t id0 = P.id_of_rec[rec];
    P.mol_id_begin[mol] = (uint32_t)id0;
    const uint32_t idlen = P.rec_idlen[rec];
    const uint8_t* idp = P.text + P.line_start[li] + 1;
    for (uint32_t i = 0; i < idlen; i++) P.id_pool[id0 + i] = idp[i];
    const uint32_t slot = mdf_align16(2 * L + 32), rslot = mdf_align16(L);
    const uint64_t so = P.so_of_rec[rec], ro = P.ro_of_rec[rec];
    for (uint32_t d = 0; d < depth; d++) {
        P.read_mol[r0 + d] = mol;
        P.scr_off[r0 + d] = so + (uint64_t)d * slot;
        P.raw_off[r0 + d] = ro + (uint64_t)d * rslot;
    }
    // statistics and length classes (engine.cu: prepare_host)
    uint64_t gen = 0, lit = 0;
    for (uint32_t s = s0; s < s1; s++) {
        if (P.seg_src[s] & SEG_LITERAL) lit += P.seg_len[s]; else gen += P.seg_len[s];
    }
    const uint32_t nmods = P.seg_mod_begin[s1] - P.seg_mod_begin[s0];
    MdfTotals* T = P.totals;
    atomicAdd((unsigned long long*)&T->total_raw_bases, (unsigned long long)L * depth);
    atomicAdd((unsigned long long*)&T->gen_bases, (unsigned long long)gen * depth);
    atomicAdd((unsigned long long*)&T->lit_bases, (unsigned long long)lit * depth);
    atomicAdd((unsigned long long*)&T->per_read_segs, (unsigned long long)(s1 - s0) * depth);
    atomicAdd((unsigned long long*)&T->per_read_mods, (unsigned long long)nmods * depth);
    atomicAdd((unsigned long long*)&T->idsum, (unsigned long long)idlen * depth);
    const uint32_t n = L + 2 * P.k;
    uint32_t c = 0;
    while (c < 5 && n > P.ncap[c]) c++;
    atomicAdd(&T->n_work[c], depth);
    if (c == 5) atomicMax(&T->max_n_gws, n);
    const uint32_t ng = (5u * n + 1u) / 2u;   // kernels.cuh: grow_len
    while (c < 5 && ng > P.ncap[c]) c++;
    atomicAdd(&T->n_work_grow[c], depth);
    if (c == 5) atomicMax(&T->max_n_gws_grow, ng);
}

// closing entries of the offset arrays and the totals the host sizes the launch with
__global__ void mdf_finalize(const MdfParams P) {
    if (blockIdx.x || threadIdx.x) return;
    MdfTotals* T = P.totals;
    T->reads = P.read_of_rec[P.n_records];
    T->mols = P.mol_of_rec[P.n_records];
    T->segs = P.seg_idx[P.n_lines];
    T->mods = P.mod_off[P.n_lines];
    T->lit_bytes = P.lit_off[P.n_lines];
    T->id_bytes = P.id_of_rec[P.n_records];
    T->scr_total = P.so_of_rec[P.n_records];
    T->raw_total = P.ro_of_rec[P.n_records];
    if (T->reads > P.read_cap) {
        atomicCAS(&T->bad, 0u, MDF_BAD_READS);
        return;
    }
    P.mol_seg_begin[T->mols] = (uint32_t)T->segs;
    P.mol_id_begin[T->mols] = (uint32_t)T->id_bytes;
    P.scr_off[T->reads] = T->scr_total;
    P.raw_off[T->reads] = T->raw_total;
}

}  // namespace tk
