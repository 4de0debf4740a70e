// Reference genome store, MDF parser and molecule batch (see host.hpp for the reference mapping).
#include "host.hpp"

#include <emmintrin.h>
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <atomic>
#include <cstring>
#include <thread>
#include <stdexcept>
#include <string_view>

#include "model.hpp"  // read_text_maybe_gz

namespace tk {

// ------------------------------------------------------------------ genome

namespace {

// 8 bytes of A/C/G/T (either case) -> 16 bits, base j at bits 2j: A=0 C=1 G=2 T=3 is ((c >> 1) ^ (c >> 2)) & 3
inline uint32_t pack8(uint64_t c) {
    uint64_t x = ((c >> 1) ^ (c >> 2)) & 0x0303030303030303ull;
    x = (x | (x >> 6)) & 0x000F000F000F000Full;
    x = (x | (x >> 12)) & 0x000000FF000000FFull;
    return (uint32_t)((x | (x >> 24)) & 0xFFFFu);
}

}  // namespace

// Appends `len` bytes of sequence at global base n_bases.  Blocks of 16 bytes that are all A/C/G/T (either case) are
// packed with three shift-and-mask steps per 8 bases and OR-ed into the (zeroed) words through one unaligned 64-bit
// access; any other block goes byte by byte, its bytes outside ACGT extending or starting exception runs.  `run` is
// the open run of the contig (run_len == 0: none).
namespace {

struct ExcOut {
    std::vector<uint64_t>* start;
    std::vector<uint32_t>* len;
    std::vector<uint8_t>* chr;
};

// `W`: the words (large enough for [g, g + len) plus two words), `g`: global base of seq[0], advanced.
void pack_bases(uint32_t* const W, uint64_t& g_io, const char* seq, size_t len, GenomeStore::ExcRun& run, const ExcOut& exc) {
    auto flush = [&]() {
        if (run.len) {
            exc.start->push_back(run.start);
            exc.len->push_back(run.len);
            exc.chr->push_back(run.chr);
            run.len = 0;
        }
    };
    uint64_t g = g_io;
    size_t i = 0;
    const __m128i lower = _mm_set1_epi8(0x20), va = _mm_set1_epi8('a'), vc = _mm_set1_epi8('c'), vg = _mm_set1_epi8('g'),
                  vt = _mm_set1_epi8('t');
    while (i < len) {
        if (len - i >= 16) {
            const __m128i v = _mm_loadu_si128((const __m128i*)(seq + i));
            const __m128i lo = _mm_or_si128(v, lower);
            const __m128i ok = _mm_or_si128(_mm_or_si128(_mm_cmpeq_epi8(lo, va), _mm_cmpeq_epi8(lo, vc)),
                                            _mm_or_si128(_mm_cmpeq_epi8(lo, vg), _mm_cmpeq_epi8(lo, vt)));
            if (_mm_movemask_epi8(ok) == 0xFFFF) {
                flush();
                uint64_t c0, c1;
                memcpy(&c0, seq + i, 8);
                memcpy(&c1, seq + i + 8, 8);
                // (the second word is touched only when bases of this block lie in it: the word after a contig's
                // last one may belong to the next contig, which another thread is packing)
                const uint32_t grp32 = pack8(c0) | (pack8(c1) << 16);
                const unsigned sh = 2 * (unsigned)(g & 15);
                if (sh == 0) {
                    W[g >> 4] |= grp32;
                } else {
                    uint64_t w;
                    memcpy(&w, W + (g >> 4), 8);
                    w |= (uint64_t)grp32 << sh;
                    memcpy(W + (g >> 4), &w, 8);
                }
                g += 16;
                i += 16;
                continue;
            }
            if (run.len && run.one_raw && run.start + run.len == g && run.len < 0xFFFFFFE0u &&
                _mm_movemask_epi8(_mm_cmpeq_epi8(v, _mm_set1_epi8((char)run.raw))) == 0xFFFF) {
                run.len += 16;   // inside a run of one byte (the N stretches of an assembly)
                g += 16;
                i += 16;
                continue;
            }
        }
        if (len - i >= 8) {   // the same for 8 bytes (what a 60- or 70-column line leaves after its 16-byte blocks)
            const __m128i v = _mm_loadl_epi64((const __m128i*)(seq + i));
            const __m128i lo = _mm_or_si128(v, lower);
            const __m128i ok = _mm_or_si128(_mm_or_si128(_mm_cmpeq_epi8(lo, va), _mm_cmpeq_epi8(lo, vc)),
                                            _mm_or_si128(_mm_cmpeq_epi8(lo, vg), _mm_cmpeq_epi8(lo, vt)));
            if ((_mm_movemask_epi8(ok) & 0xFF) == 0xFF) {
                flush();
                uint64_t c0;
                memcpy(&c0, seq + i, 8);
                const unsigned sh = 2 * (unsigned)(g & 15);
                if (sh <= 16) {
                    W[g >> 4] |= pack8(c0) << sh;
                } else {
                    uint64_t w;
                    memcpy(&w, W + (g >> 4), 8);
                    w |= (uint64_t)pack8(c0) << sh;
                    memcpy(W + (g >> 4), &w, 8);
                }
                g += 8;
                i += 8;
                continue;
            }
        }
        const size_t stop = std::min(len, i + 8);
        for (; i < stop; i++, g++) {
            const uint8_t raw = (uint8_t)seq[i];
            const uint8_t c = ascii_upper(raw);
            const uint8_t code = frag_code(c);
            if (code < 4) {
                W[(size_t)(g >> 4)] |= (uint32_t)code << (2 * (g & 15));
                flush();
            } else if (run.len && run.chr == c && run.start + run.len == g && run.len < 0xFFFFFFFFu) {
                run.len++;
                if (run.raw != raw) run.one_raw = false;   // mixed case: the 16-byte shortcut compares raw bytes
            } else {
                flush();
                run.start = g;
                run.len = 1;
                run.chr = c;
                run.raw = raw;
                run.one_raw = true;
            }
        }
    }
    g_io = g;
}

}  // namespace

void GenomeStore::append_bases(const char* seq, size_t len, ExcRun& run) {
    const size_t need = (size_t)((n_bases + len + 63) / 64) * 4 + 4;
    if (words.size() < need) {
        if (words.capacity() < need) words.reserve(std::max(need, words.capacity() + words.capacity() / 2));
        words.resize(need, 0);
    }
    pack_bases(words.data(), n_bases, seq, len, run, ExcOut{&exc_start, &exc_len, &exc_chr});
}

void GenomeStore::begin_contig() {
    // every contig starts on a 64-base boundary so 128-bit loads of a contig prefix stay aligned
    n_bases = (n_bases + 63) & ~uint64_t(63);
}

void GenomeStore::end_contig(const std::string& name, uint64_t off, ExcRun& run) {
    if (run.len) {
        exc_start.push_back(run.start);
        exc_len.push_back(run.len);
        exc_chr.push_back(run.chr);
        run.len = 0;
    }
    const uint64_t len = n_bases - off;
    words.resize((size_t)((off + len + 63) / 64) * 4 + 4, 0);
    contigs[name] = Contig{off, len};  // later files override earlier names (dict.update)
    note_name(name.size());
}

void GenomeStore::add_contig(const std::string& name, const char* seq, size_t len) {
    begin_contig();
    const uint64_t off = n_bases;
    ExcRun run;
    append_bases(seq, len, run);
    end_contig(name, off, run);
}

namespace {

// number of '\n' and '\r' bytes in [p, p + n)
size_t count_line_ends(const char* p, size_t n) {
    size_t c = 0, i = 0;
    const __m128i vn = _mm_set1_epi8('\n'), vr = _mm_set1_epi8('\r');
    for (; i + 16 <= n; i += 16) {
        const __m128i v = _mm_loadu_si128((const __m128i*)(p + i));
        c += (size_t)__builtin_popcount((unsigned)_mm_movemask_epi8(_mm_or_si128(_mm_cmpeq_epi8(v, vn), _mm_cmpeq_epi8(v, vr))));
    }
    for (; i < n; i++) c += p[i] == '\n' || p[i] == '\r';
    return c;
}

// end of the line that starts at `pos` (universal newlines): `eol` = its last byte + 1, returns the next line's start
size_t line_end(const char* text, size_t len, size_t pos, size_t& eol) {
    const char* nl = (const char*)memchr(text + pos, '\n', len - pos);
    eol = nl ? (size_t)(nl - text) : len;
    size_t next = nl ? eol + 1 : len;
    const char* cr = (const char*)memchr(text + pos, '\r', eol - pos);
    if (cr) {
        const size_t r = (size_t)(cr - text);
        next = (r + 1 == eol && nl) ? eol + 1 : r + 1;
        eol = r;
    }
    return next;
}

}  // namespace

void GenomeStore::load_fasta_text(const char* text, size_t len) {
    // py/sequence.py:168-186 in text mode: universal newlines, name = header up to the first space,
    // a header that follows a header without sequence replaces it, an empty line raises.
    //
    // Every contig starts on a 64-base boundary, so no two contigs share a word: the records are found by their '>'
    // lines (memchr; a '>' inside a line is sequence), their bases counted (bytes that end no line), their places
    // fixed in text order, and then packed side by side by host threads, each record where its lines lie (no copy).
    // Exception runs never cross a contig and are appended in text order.  A record that raises leaves the records
    // before it in place and nothing of itself or of the later ones.
    struct Rec {
        std::string name;
        size_t body = 0, body_end = 0;   // its sequence lines (body == body_end: none)
        uint64_t bases = 0, off = 0;
        std::vector<uint64_t> exc_start;
        std::vector<uint32_t> exc_len;
        std::vector<uint8_t> exc_chr;
        std::string error;
    };
    std::vector<Rec> recs;
    {
        std::string name;              // sequence before any header belongs to the contig ""
        bool pending = true;           // no sequence since the last header (or since the start)
        auto header_at = [&](size_t p) { return text[p] == '>' && (p == 0 || text[p - 1] == '\n' || text[p - 1] == '\r'); };
        size_t pos = 0;
        while (pos < len) {
            if (header_at(pos)) {
                size_t eol;
                const size_t next = line_end(text, len, pos, eol);
                std::string_view hdr(text + pos + 1, eol - pos - 1);
                name = std::string(hdr.substr(0, hdr.find(' ')));
                pending = true;
                pos = next;
                continue;
            }
            // sequence lines up to the next header line
            size_t q = pos;
            for (;;) {
                const char* gt = (const char*)memchr(text + q, '>', len - q);
                if (!gt) {
                    q = len;
                    break;
                }
                q = (size_t)(gt - text);
                if (header_at(q)) break;
                q++;
            }
            Rec r;
            r.name = name;
            r.body = pos;
            r.body_end = q;
            recs.push_back(std::move(r));
            pending = false;
            pos = q;
        }
        if (pending) {   // the reference's final `refs[name] = "".join(chunks)` with no chunks: an empty contig
            Rec r;
            r.name = name;
            r.body = r.body_end = len;
            recs.push_back(std::move(r));
        }
    }
    // bases and places
    const size_t n_threads_max = std::max(1u, std::min(16u, std::thread::hardware_concurrency()));
    const size_t n_threads = len < ((size_t)8 << 20) ? 1 : std::min(n_threads_max, recs.size());
    auto for_each_rec = [&](auto&& fn) {
        if (n_threads <= 1) {
            for (size_t i = 0; i < recs.size(); i++) fn(i);
            return;
        }
        // largest first, so that the longest chromosome does not start last
        std::vector<size_t> order(recs.size());
        for (size_t i = 0; i < order.size(); i++) order[i] = i;
        std::sort(order.begin(), order.end(), [&](size_t a, size_t b) {
            return recs[a].body_end - recs[a].body > recs[b].body_end - recs[b].body;
        });
        std::atomic<size_t> next{0};
        std::vector<std::thread> th;
        for (size_t t = 0; t < n_threads; t++)
            th.emplace_back([&]() {
                for (size_t k; (k = next.fetch_add(1)) < order.size();) fn(order[k]);
            });
        for (auto& t : th) t.join();
    };
    for_each_rec([&](size_t i) {
        Rec& r = recs[i];
        r.bases = (uint64_t)(r.body_end - r.body) - count_line_ends(text + r.body, r.body_end - r.body);
    });
    const uint64_t keep_bases = n_bases;
    const size_t keep_words = words.size();
    uint64_t at = n_bases;
    for (Rec& r : recs) {
        at = (at + 63) & ~uint64_t(63);   // every contig starts on a 64-base boundary so 128-bit loads of a contig prefix stay aligned
        r.off = at;
        at += r.bases;
    }
    const Rec& last = recs.back();
    words.resize(std::max(words.size(), (size_t)((last.off + last.bases + 63) / 64) * 4 + 4), 0);
    uint32_t* const W = words.data();
    for_each_rec([&](size_t i) {
        Rec& r = recs[i];
        ExcRun run;
        uint64_t g = r.off;
        const ExcOut exc{&r.exc_start, &r.exc_len, &r.exc_chr};
        size_t pos = r.body;
        while (pos < r.body_end) {
            size_t eol;
            const size_t next = line_end(text, r.body_end, pos, eol);
            if (eol == pos) {
                r.error = "IndexError: empty line in FASTA (py/sequence.py:177)";
                return;
            }
            pack_bases(W, g, text + pos, eol - pos, run, exc);
            pos = next;
        }
        if (run.len) {
            r.exc_start.push_back(run.start);
            r.exc_len.push_back(run.len);
            r.exc_chr.push_back(run.chr);
        }
        if (g != r.off + r.bases) r.error = "internal error: FASTA base count";
    });
    // in text order
    for (size_t i = 0; i < recs.size(); i++) {
        Rec& r = recs[i];
        if (!r.error.empty()) {
            // nothing of this record or the later ones: their words go back to zero
            const uint64_t end = i ? recs[i - 1].off + recs[i - 1].bases : keep_bases;
            const size_t good_words = i ? (size_t)((end + 63) / 64) * 4 + 4 : keep_words;
            words.resize(good_words);
            for (size_t w = (size_t)((end + 63) / 64) * 4; w < words.size(); w++) words[w] = 0;   // the padding words it may have touched
            n_bases = end;
            throw std::runtime_error(r.error);
        }
        exc_start.insert(exc_start.end(), r.exc_start.begin(), r.exc_start.end());
        exc_len.insert(exc_len.end(), r.exc_len.begin(), r.exc_len.end());
        exc_chr.insert(exc_chr.end(), r.exc_chr.begin(), r.exc_chr.end());
        contigs[r.name] = Contig{r.off, r.bases};  // later files override earlier names (dict.update)
        note_name(r.name.size());
        n_bases = r.off + r.bases;
    }
    words.resize((size_t)((last.off + last.bases + 63) / 64) * 4 + 4, 0);
}

void GenomeStore::load_fasta(const std::string& path) {
    // a plain regular file is packed straight from its mapping (no copy of a 3 GB reference); anything else
    // (.gz, a pipe) is read into memory first
    {
        const int fd = open(path.c_str(), O_RDONLY);
        if (fd >= 0) {
            struct stat sb{};
            unsigned char magic[2] = {0, 0};
            const bool regular = fstat(fd, &sb) == 0 && S_ISREG(sb.st_mode) && sb.st_size > 0;
            const bool gz = regular && pread(fd, magic, 2, 0) == 2 && magic[0] == 0x1f && magic[1] == 0x8b;
            void* map = regular && !gz ? mmap(nullptr, (size_t)sb.st_size, PROT_READ, MAP_PRIVATE, fd, 0) : MAP_FAILED;
            close(fd);
            if (map != MAP_FAILED) {
                madvise(map, (size_t)sb.st_size, MADV_SEQUENTIAL);
                struct Unmap {
                    void* p;
                    size_t n;
                    ~Unmap() { munmap(p, n); }
                } guard{map, (size_t)sb.st_size};
                load_fasta_text((const char*)map, (size_t)sb.st_size);
                return;
            }
        }
    }
    std::string text = read_text_maybe_gz(path);
    load_fasta_text(text.data(), text.size());
}

void GenomeStore::index_exceptions() {
    const uint64_t nblk = (n_bases >> EXC_BLOCK_SHIFT) + 2;
    exc_block_bits.assign((size_t)((nblk + 63) / 64), 0);
    for (size_t i = 0; i < exc_start.size(); i++) {
        const uint64_t b0 = exc_start[i] >> EXC_BLOCK_SHIFT, b1 = (exc_start[i] + exc_len[i] - 1) >> EXC_BLOCK_SHIFT;
        for (uint64_t b = b0; b <= b1; b++) exc_block_bits[(size_t)(b >> 6)] |= 1ull << (b & 63);
    }
    exc_indexed = exc_start.size();
}

bool GenomeStore::overlaps_exception(uint64_t start, uint64_t len) const {
    if (exc_start.empty() || len == 0) return false;
    if (exc_indexed == exc_start.size() && !exc_block_bits.empty()) {
        const uint64_t b0 = start >> EXC_BLOCK_SHIFT, b1 = (start + len - 1) >> EXC_BLOCK_SHIFT;
        if (b1 - b0 < 8 && (b1 >> 6) < exc_block_bits.size()) {
            bool any = false;
            for (uint64_t b = b0; b <= b1; b++) any |= (exc_block_bits[(size_t)(b >> 6)] >> (b & 63)) & 1;
            if (!any) return false;
        }
    }
    // first run with start >= `start`; the run before it may still reach into the range
    size_t i = (size_t)(std::lower_bound(exc_start.begin(), exc_start.end(), start) - exc_start.begin());
    if (i < exc_start.size() && exc_start[i] < start + len) return true;
    if (i > 0 && exc_start[i - 1] + exc_len[i - 1] > start) return true;
    return false;
}

std::string GenomeStore::fetch(uint64_t start, uint64_t len) const {
    std::string s((size_t)len, '?');
    for (uint64_t i = 0; i < len; i++) {
        uint64_t g = start + i;
        s[(size_t)i] = "ACGT"[(words[(size_t)(g >> 4)] >> (2 * (g & 15))) & 3];
    }
    size_t r = (size_t)(std::lower_bound(exc_start.begin(), exc_start.end(), start) - exc_start.begin());
    if (r > 0) r--;
    for (; r < exc_start.size() && exc_start[r] < start + len; r++) {
        uint64_t a = std::max(exc_start[r], start), b = std::min(exc_start[r] + exc_len[r], start + len);
        for (uint64_t g = a; g < b; g++) s[(size_t)(g - start)] = (char)exc_chr[r];
    }
    return s;
}

// ------------------------------------------------------------------- batch

void HostBatch::clear() {
    read_mol.clear();
    mol_seg_begin.assign(1, 0);
    mol_raw_len.clear();
    mol_id_begin.assign(1, 0);
    mol_tail_len.clear();
    seg_src.clear();
    seg_len.clear();
    seg_mod_begin.assign(1, 0);
    mod_pos.clear();
    mod_chr.clear();
    lit_pool.clear();
    id_pool.clear();
    mol_cb_begin.assign(1, 0);
    cb_pool.clear();
    lit_index.clear();
    total_raw_bases = genomic_bases = literal_bases = 0;
}

void HostBatch::append(const HostBatch& o) {
    const uint32_t mol0 = (uint32_t)n_mols(), seg0 = (uint32_t)seg_src.size(), mod0 = (uint32_t)mod_pos.size();
    const uint64_t lit0 = lit_pool.size();
    const uint32_t id0 = (uint32_t)id_pool.size();
    auto grow = [](auto& v, size_t extra) {
        if (v.capacity() < v.size() + extra) v.reserve(std::max(v.capacity() * 2, v.size() + extra));
    };
    grow(read_mol, o.read_mol.size());
    grow(mol_seg_begin, o.mol_seg_begin.size());
    grow(mol_id_begin, o.mol_id_begin.size());
    grow(seg_src, o.seg_src.size());
    grow(seg_mod_begin, o.seg_mod_begin.size());
    for (uint32_t m : o.read_mol) read_mol.push_back(m + mol0);
    for (size_t i = 1; i < o.mol_seg_begin.size(); i++) mol_seg_begin.push_back(o.mol_seg_begin[i] + seg0);
    if (!o.mol_tail_len.empty() || !mol_tail_len.empty()) {
        mol_tail_len.resize(mol0, 0);
        mol_tail_len.insert(mol_tail_len.end(), o.mol_tail_len.begin(), o.mol_tail_len.end());
        mol_tail_len.resize(mol0 + o.n_mols(), 0);
    }
    mol_raw_len.insert(mol_raw_len.end(), o.mol_raw_len.begin(), o.mol_raw_len.end());
    for (size_t i = 1; i < o.mol_id_begin.size(); i++) mol_id_begin.push_back(o.mol_id_begin[i] + id0);
    for (uint64_t s : o.seg_src) seg_src.push_back((s & SEG_LITERAL) ? s + lit0 : s);
    seg_len.insert(seg_len.end(), o.seg_len.begin(), o.seg_len.end());
    for (size_t i = 1; i < o.seg_mod_begin.size(); i++) seg_mod_begin.push_back(o.seg_mod_begin[i] + mod0);
    mod_pos.insert(mod_pos.end(), o.mod_pos.begin(), o.mod_pos.end());
    mod_chr.insert(mod_chr.end(), o.mod_chr.begin(), o.mod_chr.end());
    lit_pool.insert(lit_pool.end(), o.lit_pool.begin(), o.lit_pool.end());
    id_pool.insert(id_pool.end(), o.id_pool.begin(), o.id_pool.end());
    total_raw_bases += o.total_raw_bases;
    genomic_bases += o.genomic_bases;
    literal_bases += o.literal_bases;
}

// -------------------------------------------------------------------- MDF

namespace {

struct Line {
    const char* b;
    const char* e;
};

// universal-newline line splitter; returns false at end of input
inline bool next_line(const char* text, size_t len, size_t& pos, Line& out, bool& terminated) {
    if (pos >= len) return false;
    const char* p = text + pos;
    const char* nl = (const char*)memchr(p, '\n', len - pos);
    size_t eol = nl ? (size_t)(nl - text) : len;
    size_t next = nl ? eol + 1 : len;
    terminated = nl != nullptr;
    const char* cr = (const char*)memchr(p, '\r', eol - pos);
    if (cr) {
        size_t r = (size_t)(cr - text);
        if (r + 1 == eol && nl) {
            next = eol + 1;
        } else {
            next = r + 1;
            terminated = true;
        }
        eol = r;
    }
    out.b = p;
    out.e = text + eol;
    pos = next;
    return true;
}

// One pass over a line: its end (universal newlines, as next_line) and the positions of its first
// tabs.  16 bytes at a time where the buffer allows.
struct LineScan {
    const char* b;
    const char* e;
    const char* tab[5];
    int ntab;        // tabs seen, capped at 6
    size_t next;
    bool terminated;
};
inline bool scan_line(const char* text, size_t len, size_t pos, LineScan& L) {
    if (pos >= len) return false;
    const char* p = text + pos;
    const char* const end = text + len;
    L.b = p;
    L.ntab = 0;
    const __m128i vt = _mm_set1_epi8('\t'), vn = _mm_set1_epi8('\n'), vr = _mm_set1_epi8('\r');
    const char* stop = nullptr;
    // 32 bytes at a time (most interval lines end inside the first block): tabs and line ends as two bit masks, the
    // tabs before the line end taken off their mask one by one
    while (end - p >= 32) {
        const __m128i v0 = _mm_loadu_si128((const __m128i*)p), v1 = _mm_loadu_si128((const __m128i*)(p + 16));
        unsigned mt = (unsigned)_mm_movemask_epi8(_mm_cmpeq_epi8(v0, vt)) |
                      ((unsigned)_mm_movemask_epi8(_mm_cmpeq_epi8(v1, vt)) << 16);
        const unsigned mn = (unsigned)_mm_movemask_epi8(_mm_or_si128(_mm_cmpeq_epi8(v0, vn), _mm_cmpeq_epi8(v0, vr))) |
                            ((unsigned)_mm_movemask_epi8(_mm_or_si128(_mm_cmpeq_epi8(v1, vn), _mm_cmpeq_epi8(v1, vr))) << 16);
        if (mn) {
            const int at = __builtin_ctz(mn);
            mt &= (1u << at) - 1u;
            stop = p + at;
        }
        while (mt) {
            const int i = __builtin_ctz(mt);
            mt &= mt - 1;
            if (L.ntab < 5) L.tab[L.ntab] = p + i;
            if (L.ntab < 6) L.ntab++;
        }
        if (stop) break;
        p += 32;
    }
    while (!stop) {
        if (end - p >= 16) {
            const __m128i v = _mm_loadu_si128((const __m128i*)p);
            unsigned m = (unsigned)_mm_movemask_epi8(
                _mm_or_si128(_mm_or_si128(_mm_cmpeq_epi8(v, vt), _mm_cmpeq_epi8(v, vn)), _mm_cmpeq_epi8(v, vr)));
            while (m) {
                const int i = __builtin_ctz(m);
                m &= m - 1;
                if (p[i] == '\t') {
                    if (L.ntab < 5) L.tab[L.ntab] = p + i;
                    if (L.ntab < 6) L.ntab++;
                } else {
                    stop = p + i;
                    break;
                }
            }
            if (!stop) p += 16;
        } else {
            for (; p < end; p++) {
                const char c = *p;
                if (c == '\t') {
                    if (L.ntab < 5) L.tab[L.ntab] = p;
                    if (L.ntab < 6) L.ntab++;
                } else if (c == '\n' || c == '\r') {
                    stop = p;
                    break;
                }
            }
            if (!stop) break;
        }
    }
    if (!stop) {
        L.e = end;
        L.next = len;
        L.terminated = false;
        return true;
    }
    L.e = stop;
    L.terminated = true;
    if (*stop == '\r' && stop + 1 < end && stop[1] == '\n')
        L.next = (size_t)(stop - text) + 2;
    else
        L.next = (size_t)(stop - text) + 1;
    return true;
}

// Python int(): optional surrounding whitespace, sign, digits with single underscores
long long py_int(const char* b, const char* e) {
    auto fail = [&]() -> long long {
        throw std::runtime_error("ValueError: invalid literal for int() with base 10: '" + std::string(b, e) + "'");
    };
    while (b < e && isspace((unsigned char)*b)) b++;
    while (e > b && isspace((unsigned char)e[-1])) e--;
    if (b == e) return fail();
    bool neg = false;
    if (*b == '+' || *b == '-') {
        neg = *b == '-';
        b++;
    }
    if (b == e || !isdigit((unsigned char)*b)) return fail();
    long long v = 0;
    bool prev_us = false;
    for (; b < e; b++) {
        if (*b == '_') {
            if (prev_us) return fail();
            prev_us = true;
            continue;
        }
        if (!isdigit((unsigned char)*b)) return fail();
        prev_us = false;
        if (v > (long long)4e17) throw std::runtime_error("MDF integer out of range");
        v = v * 10 + (*b - '0');
    }
    if (prev_us) return fail();
    return neg ? -v : v;
}

// n <= 8 decimal digits starting at p, all eight bytes at p readable: the bytes moved to the top of a word, '0' below
// them, checked and summed in three multiplications instead of a dependent multiply-add per digit
inline bool digits8(const char* p, size_t n, uint64_t& out) {
    uint64_t v;
    memcpy(&v, p, 8);
    if (n < 8) v = (v << (8 * (8 - n))) | (0x3030303030303030ull >> (8 * n));
    if (((v + 0x4646464646464646ull) | (v - 0x3030303030303030ull)) & 0x8080808080808080ull) return false;
    v -= 0x3030303030303030ull;
    v = (v * 10) + (v >> 8);
    out = (((v & 0x000000FF000000FFull) * 0x000F424000000064ull) + (((v >> 16) & 0x000000FF000000FFull) * 0x0000271000000001ull)) >> 32;
    return true;
}

// plain decimal digits (the emitter's format); anything else goes through py_int.  `readable_end`: end of the buffer
// the field lies in (bytes up to it may be read), or null.
inline long long mdf_int(const char* b, const char* e, const char* readable_end = nullptr) {
    const size_t n = (size_t)(e - b);
    if (b < e && n <= 16 && readable_end && b + 8 <= readable_end) {
        uint64_t lo, hi = 0;
        if (n <= 8 ? digits8(b, n, lo) : (digits8(b, n - 8, hi) && digits8(e - 8, 8, lo)))
            return (long long)(hi * 100000000ull + lo);
    } else if (b < e && n <= 18) {
        long long v = 0;
        const char* p = b;
        for (; p < e; p++) {
            unsigned d = (unsigned)(*p - '0');
            if (d > 9) break;
            v = v * 10 + d;
        }
        if (p == e) return v;
    }
    return py_int(b, e);
}

inline void py_slice(long long start, long long end, long long n, long long& s, long long& e) {
    auto norm = [n](long long i) {
        if (i < 0) {
            i += n;
            if (i < 0) i = 0;
        } else if (i > n) {
            i = n;
        }
        return i;
    };
    s = norm(start);
    e = norm(end);
    if (e < s) e = s;
}

struct SvHash {
    using is_transparent = void;
    size_t operator()(std::string_view s) const { return std::hash<std::string_view>{}(s); }
};

}  // namespace

void plan_shards(const char* text, size_t len, int world, std::vector<uint64_t>& byte_begin,
                 std::vector<uint64_t>& ordinal_base) {
    if (world < 1) throw std::runtime_error("shard plan: world size must be at least 1");
    // one light pass: per record its byte offset, depth and depth x sum of (end - start)
    struct Rec {
        uint64_t off, reads, bases;
    };
    // The pass is cut into ranges that start at header lines ("\n+") and run on host threads; the records of the
    // ranges, put together in text order, are those of one pass over the whole text.
    auto scan_range = [&](size_t begin, size_t end, std::vector<Rec>& recs) {
        LineScan ln;
        size_t pos = begin;
        long long depth = 0;
        bool have = false;
        while (scan_line(text, end, pos, ln)) {
            const size_t line_start = pos;
            pos = ln.next;
            if (ln.b == ln.e) continue;  // the parser reports it
            if (*ln.b == '+') {
                depth = 0;
                if (ln.ntab >= 1) {
                    try {
                        depth = mdf_int(ln.tab[0] + 1, ln.ntab > 1 ? ln.tab[1] : ln.e, text + len);
                    } catch (const std::exception&) {
                        depth = 0;
                    }
                }
                if (depth < 0) depth = 0;
                recs.push_back({(uint64_t)line_start, (uint64_t)depth, 0});
                have = true;
            } else if (have && ln.ntab >= 2) {
                try {
                    const long long a = mdf_int(ln.tab[0] + 1, ln.tab[1], text + len);
                    const long long b = mdf_int(ln.tab[1] + 1, ln.ntab > 2 ? ln.tab[2] : ln.e, text + len);
                    if (b > a && a >= 0) recs.back().bases += (uint64_t)(b - a) * recs.back().reads;
                } catch (const std::exception&) {
                }
            }
        }
    };
    std::vector<Rec> recs;
    {
        const size_t want = len < ((size_t)8 << 20) ? 1 : std::max(1u, std::min(16u, std::thread::hardware_concurrency()));
        std::vector<size_t> cuts{0};
        for (size_t t = 1; t < want; t++) {
            size_t from = std::max(cuts.back(), len / want * t);
            const char* hit = nullptr;
            while (from + 1 < len) {
                const char* nl = (const char*)memchr(text + from, '\n', len - 1 - from);
                if (!nl) break;
                if (nl[1] == '+') {
                    hit = nl;
                    break;
                }
                from = (size_t)(nl - text) + 1;
            }
            if (!hit) break;
            const size_t at = (size_t)(hit - text) + 1;
            if (at > cuts.back()) cuts.push_back(at);
        }
        cuts.push_back(len);
        const size_t n_ranges = cuts.size() - 1;
        if (n_ranges == 1) {
            scan_range(0, len, recs);
        } else {
            std::vector<std::vector<Rec>> part(n_ranges);
            std::vector<std::thread> th;
            for (size_t i = 0; i < n_ranges; i++) th.emplace_back([&, i]() { scan_range(cuts[i], cuts[i + 1], part[i]); });
            for (auto& t : th) t.join();
            size_t n = 0;
            for (auto& v : part) n += v.size();
            recs.reserve(n);
            for (auto& v : part) recs.insert(recs.end(), v.begin(), v.end());
        }
    }
    uint64_t total = 0;
    for (auto& r : recs) total += r.bases + r.reads;  // + reads: records of empty molecules still cost a read
    byte_begin.assign((size_t)world + 1, len);
    ordinal_base.assign((size_t)world + 1, 0);
    byte_begin[0] = 0;
    uint64_t acc = 0, reads = 0;
    int shard = 1;
    for (size_t i = 0; i < recs.size(); i++) {
        // record i starts shard `shard` once the records before it hold shard/world of the weight
        while (shard < world && acc * (uint64_t)world >= total * (uint64_t)shard && i > 0) {
            byte_begin[(size_t)shard] = recs[i].off;
            ordinal_base[(size_t)shard] = reads;
            shard++;
        }
        acc += recs[i].bases + recs[i].reads;
        reads += recs[i].reads;
    }
    for (; shard < world; shard++) {  // fewer records than ranks: the rest are empty shards at the end
        byte_begin[(size_t)shard] = len;
        ordinal_base[(size_t)shard] = reads;
    }
    byte_begin[(size_t)world] = len;
    ordinal_base[(size_t)world] = reads;
}

size_t MdfParser::feed(const char* text, size_t len, bool final, HostBatch& out, size_t max_reads,
                       uint64_t max_bases) {
    size_t pos = 0, consumed = 0;
    // Room for the lines of this text up front (about 25 text bytes per interval line and 8 lines per molecule in the
    // emitter's format; the streaming pipeline reserves the same before its first call, engine.cu: parse_chunk): the
    // arrays of a 1 MiB chunk would otherwise be reallocated a dozen times each, every step a fresh mapping whose
    // pages fault in, on all parser threads at once.  Pages never written cost nothing.
    {
        const size_t lines = len / 22 + 16;
        auto room = [](auto& v, size_t extra) {
            if (v.capacity() < v.size() + extra) v.reserve(v.size() + extra);
        };
        room(out.seg_src, lines);
        room(out.seg_len, lines);
        room(out.seg_mod_begin, lines);
        room(out.read_mol, lines / 4);
        room(out.mol_raw_len, lines / 4);
        room(out.mol_seg_begin, lines / 4);
        room(out.mol_id_begin, lines / 4);
        room(out.id_pool, len / 16);
    }
    // state of the record being assembled
    bool have = false;
    long long depth = 0;
    uint32_t raw_len = 0;
    uint64_t gen_b = 0, lit_b = 0;
    // roll-back marks
    size_t m_seg = 0, m_mod = 0, m_lit = 0, m_id = 0, m_cb = 0;

    // The first field of recent interval lines, resolved once: contig (or "the name is the sequence",
    // py/sequence.py:307) and, for literals, the pool offset of the last slice taken from it.  Tags,
    // adapters and the current chromosome alternate inside every molecule, so a few entries avoid
    // nearly all string hashing.
    struct TokEntry {
        std::string tok;
        uint64_t k0 = 0, k1 = 0;   // the first 16 bytes of tok, zero-padded
        bool valid = false;
        const Contig* contig = nullptr;
        long long lit_s = -1, lit_e = -1;
        size_t lit_epoch = 0;
        uint32_t lit_off = 0;
    };
    constexpr int NTOK = 64;       // direct-mapped on a hash of the first 16 bytes and the length
    TokEntry toks[NTOK];
    size_t lit_epoch = 1;  // bumped when a rollback may have dropped pool bytes
    TokEntry* last_tok = nullptr;  // exon lines of one molecule repeat the chromosome
    TokEntry scratch_tok;
    // Literal lines that repeat from molecule to molecule (tags, adapters, poly(A) runs of common lengths): the raw
    // field text and the slice taken from it, direct-mapped on its length and first / last bytes, so that a repeat
    // costs one memcmp instead of encoding and hashing the bases again.
    struct LitText {
        std::string text;
        uint64_t k0 = 0, k1 = 0;   // its first 16 bytes, as in TokEntry
        long long s = 0, e = 0;
        size_t epoch = 0;   // 0: no text stored
        uint32_t off = 0;
        uint64_t seen = 0;  // fingerprint of the last text that fell here: the text is stored when it comes a second time
    };
    constexpr size_t LIT_TEXTS = 256;
    std::vector<LitText> lit_texts(LIT_TEXTS);
    uint64_t lit_fp = 0;   // fingerprint of the text lit_text_slot() was last asked about
    uint64_t k0 = 0, k1 = 0;   // first 16 bytes of the first field of the current line, zero-padded (key_of)
    auto key_hash = [&](size_t n) {
        return (k0 * 0x9E3779B97F4A7C15ull) ^ (k1 * 0xC2B2AE3D27D4EB4Full) ^ (n * 0x165667B19E3779F9ull);
    };
    auto lit_text_slot = [&](std::string_view t) -> LitText& {
        const uint64_t h = key_hash(t.size());
        lit_fp = h | 1;
        return lit_texts[(size_t)(h >> 56) & (LIT_TEXTS - 1)];
    };
    // the text behind a key that already matched: nothing for 16 bytes or fewer, one overlapping 16-byte compare up
    // to 32 bytes (both sides hold n >= 16 bytes), libc beyond
    auto same_rest = [](const char* a, const char* b, size_t n) {
        if (n <= 16) return true;
        if (n <= 32)
            return _mm_movemask_epi8(_mm_cmpeq_epi8(_mm_loadu_si128((const __m128i*)(a + n - 16)),
                                                    _mm_loadu_si128((const __m128i*)(b + n - 16)))) == 0xFFFF;
        return memcmp(a + 16, b + 16, n - 16) == 0;
    };
    std::string lit;               // encoded literal of the current line (buffer reused)
    constexpr size_t LIT_SLOTS = 4096;
    const char* const text_end = text + len;
    auto key_of = [&](std::string_view chrom) {
        const size_t n = chrom.size();
        if (chrom.data() + 16 <= text_end) {
            memcpy(&k0, chrom.data(), 8);
            memcpy(&k1, chrom.data() + 8, 8);
        } else {
            char pad[16] = {0};
            memcpy(pad, chrom.data(), n < 16 ? n : 16);
            memcpy(&k0, pad, 8);
            memcpy(&k1, pad + 8, 8);
        }
        if (n < 8) {
            k0 &= (1ull << (8 * n)) - 1;
            k1 = 0;
        } else if (n < 16) {
            k1 &= (1ull << (8 * (n - 8))) - 1;
        }
    };
    // needs key_of(chrom); a cache entry is compared through the key words, without a call into libc
    auto resolve = [&](std::string_view chrom) -> TokEntry& {
        const size_t n = chrom.size();
        if (!genome->name_len_possible(n)) {
            // no contig name has this length: a literal (barcode, UMI, adapter, poly(A) run); it stays out of the
            // token cache, which the chromosome names keep
            scratch_tok.contig = nullptr;
            scratch_tok.lit_s = scratch_tok.lit_e = -1;
            return scratch_tok;
        }
        auto same = [&](const TokEntry& t) {
            return t.valid && t.k0 == k0 && t.k1 == k1 && t.tok.size() == n && same_rest(t.tok.data(), chrom.data(), n);
        };
        if (last_tok && same(*last_tok)) return *last_tok;
        const uint64_t hsh = key_hash(n);
        TokEntry& t = toks[(size_t)(hsh >> 58) & (NTOK - 1)];
        last_tok = &t;
        if (same(t)) return t;
        t.tok.assign(chrom);
        t.k0 = k0;
        t.k1 = k1;
        t.valid = true;
        auto it = genome->contigs.find(t.tok);
        t.contig = it == genome->contigs.end() ? nullptr : &it->second;
        t.lit_s = t.lit_e = -1;
        t.lit_epoch = lit_epoch;
        return t;
    };

    auto mark = [&]() {
        m_seg = out.seg_src.size();
        m_mod = out.mod_pos.size();
        m_lit = out.lit_pool.size();
        m_id = out.id_pool.size();
        m_cb = out.cb_pool.size();
    };
    auto rollback = [&]() {
        out.seg_src.resize(m_seg);
        out.seg_len.resize(m_seg);
        out.seg_mod_begin.resize(m_seg + 1);
        out.mod_pos.resize(m_mod);
        out.mod_chr.resize(m_mod);
        // literals appended by the rolled-back record may be indexed: drop the index entries lazily
        if (out.lit_pool.size() != m_lit) {
            for (auto& sl : out.lit_index)
                if (sl.len && (size_t)sl.off + sl.len > m_lit) sl = HostBatch::LitSlot{};
            out.lit_pool.resize(m_lit);
            lit_epoch++;
        }
        out.id_pool.resize(m_id);
        out.cb_pool.resize(m_cb);
    };
    auto commit = [&]() {
        // py/sequence.py:208-210: the record is yielded `depth` times
        if (depth > 0) {
            uint32_t mol = (uint32_t)out.n_mols();
            out.mol_raw_len.push_back(raw_len);
            out.mol_seg_begin.push_back((uint32_t)out.seg_src.size());
            out.mol_id_begin.push_back((uint32_t)out.id_pool.size());
            if (capture_cb) out.mol_cb_begin.push_back((uint32_t)out.cb_pool.size());
            for (long long d = 0; d < depth; d++) out.read_mol.push_back(mol);
            out.total_raw_bases += (uint64_t)raw_len * (uint64_t)depth;
            out.genomic_bases += gen_b * (uint64_t)depth;
            out.literal_bases += lit_b * (uint64_t)depth;
        } else {
            rollback();
        }
        have = false;
    };

    LineScan ln;
    while (true) {
        size_t line_start = pos;
        if (!scan_line(text, len, pos, ln)) break;
        if (!final && !ln.terminated) break;  // partial last line: wait for more data
        pos = ln.next;
        if (ln.b == ln.e) throw std::runtime_error("IndexError: empty line in MDF (py/sequence.py:207)");
        if (*ln.b == '\t') throw std::runtime_error("IndexError: MDF line with an empty first field (py/sequence.py:207)");
        if (*ln.b == '+') {
            if (have) {
                commit();
                consumed = line_start;
                if (out.n_reads() >= max_reads || out.total_raw_bases >= max_bases) return consumed;
            } else {
                consumed = line_start;
            }
            if (ln.ntab == 0) throw std::runtime_error("IndexError: MDF header without depth field");
            const char* t1 = ln.tab[0];
            depth = mdf_int(t1 + 1, ln.ntab > 1 ? ln.tab[1] : ln.e);
            mark();
            out.id_pool.insert(out.id_pool.end(), (const uint8_t*)ln.b + 1, (const uint8_t*)t1);
            if (capture_cb && ln.ntab > 1) {
                // comment field "key=v1,v2;key=...;" (src/interval.h: dump_comment): first value of CB
                const char* c = ln.tab[1] + 1;
                const char* ce = ln.ntab > 2 ? ln.tab[2] : ln.e;
                while (c < ce) {
                    const char* semi = (const char*)memchr(c, ';', (size_t)(ce - c));
                    const char* ie = semi ? semi : ce;
                    if (ie - c >= 3 && c[0] == 'C' && c[1] == 'B' && c[2] == '=') {
                        const char* v = c + 3;
                        const char* comma = (const char*)memchr(v, ',', (size_t)(ie - v));
                        out.cb_pool.insert(out.cb_pool.end(), (const uint8_t*)v, (const uint8_t*)(comma ? comma : ie));
                        break;
                    }
                    c = ie + 1;
                }
            }
            have = true;
            raw_len = 0;
            gen_b = lit_b = 0;
            continue;
        }
        // interval line: chr, start, end, strand[, mods]
        const char* f[6];
        f[0] = ln.b;
        for (int i = 0; i < ln.ntab && i < 5; i++) f[i + 1] = ln.tab[i] + 1;
        const int nf = ln.ntab + 1;  // number of fields seen (capped at 7)
        if (nf < 4) throw std::runtime_error("ValueError: not enough values to unpack (MDF interval line)");
        if (nf > 5) throw std::runtime_error("ValueError: too many values to unpack (MDF interval line)");
        auto fend = [&](int i) { return i + 1 < nf ? f[i + 1] - 1 : ln.e; };
        long long start = mdf_int(f[1], fend(1), text + len), end = mdf_int(f[2], fend(2), text + len);
        bool plus = (fend(3) - f[3] == 1) && *f[3] == '+';
        if (!have) continue;  // intervals before the first header are parsed and dropped (py:213)

        std::string_view chrom(f[0], (size_t)(fend(0) - f[0]));
        key_of(chrom);
        TokEntry& te = resolve(chrom);
        long long s, e;
        uint64_t src;
        uint32_t slen;
        if (te.contig) {
            py_slice(start, end, (long long)te.contig->length, s, e);
            slen = (uint32_t)(e - s);
            uint64_t g = te.contig->offset + (uint64_t)s;
            src = g;
            if (genome->overlaps_exception(g, slen)) src |= SEG_HAS_EXC;
            gen_b += slen;
        } else {
            // unknown contig: the name itself is the sequence (py/sequence.py:307)
            py_slice(start, end, (long long)chrom.size(), s, e);
            slen = (uint32_t)(e - s);
            uint32_t off;
            LitText* lt = nullptr;
            if (te.lit_s == s && te.lit_e == e && te.lit_epoch == lit_epoch) {
                off = te.lit_off;
            } else if (!chrom.empty() && (lt = &lit_text_slot(chrom))->epoch == lit_epoch && lt->s == s && lt->e == e &&
                       lt->text.size() == chrom.size() && lt->k0 == k0 && lt->k1 == k1 && same_rest(lt->text.data(), chrom.data(), chrom.size())) {
                off = lt->off;
            } else {
                lit.resize((size_t)slen);
                uint64_t hsh = 0xCBF29CE484222325ull;
                for (uint32_t i = 0; i < slen; i++) {
                    const uint8_t c = frag_code(ascii_upper((uint8_t)chrom[(size_t)s + i]));
                    lit[i] = (char)c;
                    hsh = (hsh ^ c) * 0x100000001B3ull;
                }
                if (out.lit_index.empty()) out.lit_index.resize(LIT_SLOTS);
                HostBatch::LitSlot& sl = out.lit_index[(size_t)((hsh ^ (hsh >> 29)) & (LIT_SLOTS - 1))];
                if (sl.len == slen && sl.hash == hsh && slen && memcmp(out.lit_pool.data() + sl.off, lit.data(), slen) == 0) {
                    off = sl.off;
                } else {
                    off = (uint32_t)out.lit_pool.size();
                    out.lit_pool.insert(out.lit_pool.end(), lit.begin(), lit.end());
                    sl.hash = hsh;
                    sl.off = off;
                    sl.len = slen;
                }
                te.lit_s = s;
                te.lit_e = e;
                te.lit_off = off;
                te.lit_epoch = lit_epoch;
                if (lt) {
                    if (lt->seen == lit_fp) {   // (unique barcodes and UMIs never get here)
                        lt->text.assign(chrom);
                        lt->k0 = k0;
                        lt->k1 = k1;
                        lt->s = s;
                        lt->e = e;
                        lt->off = off;
                        lt->epoch = lit_epoch;
                    } else {
                        lt->seen = lit_fp;
                    }
                }
            }
            src = SEG_LITERAL | off;
            lit_b += slen;
        }
        if (!plus) src |= SEG_MINUS;
        // modifications: applied to the forward slice before the strand flip (py/sequence.py:229-239, 308)
        if (nf == 5 && fend(4) > f[4]) {
            size_t m0 = out.mod_pos.size();
            for (const char* p = f[4]; p <= ln.e;) {
                const char* c = (const char*)memchr(p, ',', (size_t)(ln.e - p));
                const char* ie = c ? c : ln.e;
                if (ie == p) throw std::runtime_error("IndexError: empty modification entry");
                long long mp = mdf_int(p, ie - 1, text + len);   // plain digits in the emitter's format, else python's int()
                if (mp < 0) mp += slen;
                if (mp < 0 || mp >= (long long)slen)
                    throw std::runtime_error("IndexError: list assignment index out of range (modification position)");
                uint8_t ch = (uint8_t)ie[-1];
                bool dup = false;
                for (size_t x = m0; x < out.mod_pos.size(); x++)
                    if (out.mod_pos[x] == (uint32_t)mp) {
                        out.mod_chr[x] = ch;  // listed order: the last one wins
                        dup = true;
                    }
                if (!dup) {
                    out.mod_pos.push_back((uint32_t)mp);
                    out.mod_chr.push_back(ch);
                }
                if (!c) break;
                p = c + 1;
            }
        }
        out.seg_src.push_back(src);
        out.seg_len.push_back(slen);
        out.seg_mod_begin.push_back((uint32_t)out.mod_pos.size());
        raw_len += slen;
    }
    if (final) {
        if (have) commit();
        consumed = len;
    } else if (have) {
        // incomplete record: undo its partial contents, the caller will feed it again
        rollback();
    }
    return consumed;
}

}  // namespace tk
