// Badread extras and tail noise applied to the molecules of a batch, one molecule per read (see extras.hpp).
#include "extras.hpp"

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <stdexcept>

namespace tk {

namespace {

constexpr double CHIMERA_END_ADAPTER_CHANCE = 0.25;     // py/tksm_badread.py:61-63 (the only trace of the extras
constexpr double CHIMERA_START_ADAPTER_CHANCE = 0.25;   // in the reference: the two constants, unused there)

inline double unit(uint32_t w) { return ((double)w + 0.5) / 4294967296.0; }

struct Draw {
    uint64_t seed, ordinal;
    void words(uint32_t index, uint32_t kind, uint32_t g, uint32_t out[4]) const {
        philox4x32_10_host(index, TAG_EXTRA | (kind << 8) | (g << 12), (uint32_t)ordinal, (uint32_t)(ordinal >> 32), (uint32_t)seed,
                           (uint32_t)(seed >> 32), out);
    }
    // n uniform bases (fragment codes 0..3): 16 per word, 64 per Philox call
    void bases(uint32_t kind, uint32_t g, uint32_t n, uint8_t* out) const {
        uint32_t w[4];
        for (uint32_t i = 0; i < n; i++) {
            if ((i & 63u) == 0) words(i >> 6, kind, g, w);
            out[i] = (uint8_t)((w[(i >> 4) & 3u] >> (2 * (i & 15u))) & 3u);
        }
    }
};

// numpy's geometric(1 / mean) by inversion, support 1, 2, ...; mean <= 1: always 1
inline uint32_t geometric(double mean, double u) {
    if (!(mean > 1.0)) return 1;
    const double k = std::ceil(std::log(1.0 - u) / std::log(1.0 - 1.0 / mean));
    if (!(k >= 1.0)) return 1;
    return k > 4.0e9 ? 4000000000u : (uint32_t)k;
}
// the same with "mean 0 means never" (glitch size and skip)
inline uint32_t geometric0(double mean, double u) { return mean > 0.0 ? geometric(mean, u) : 0u; }

// bases of an adapter of `len` that are present: the whole of it for amount >= 1, else len x with x from
// Beta(a / (1 - a), 1) (mean a, quantile u^((1 - a) / a)), rounded half up
inline uint32_t adapter_bases(uint32_t len, double amount, double u) {
    if (amount >= 1.0) return len;
    if (!(amount > 0.0)) return 0;
    const double x = std::pow(u, (1.0 - amount) / amount);
    const uint32_t k = (uint32_t)((double)len * x + 0.5);
    return k > len ? len : k;
}

// A molecule under construction: segments in fragment order, their modifications in one array.
struct Builder {
    std::vector<uint64_t> src;
    std::vector<uint32_t> len;
    std::vector<uint32_t> mod_begin{0};
    std::vector<uint32_t> mod_pos;
    std::vector<uint8_t> mod_chr;
    uint64_t total = 0;
    void clear() {
        src.clear();
        len.clear();
        mod_begin.assign(1, 0);
        mod_pos.clear();
        mod_chr.clear();
        total = 0;
    }
    void end_seg(uint64_t s, uint32_t l) {
        src.push_back(s);
        len.push_back(l);
        mod_begin.push_back((uint32_t)mod_pos.size());
        total += l;
    }
    // fragment positions [a, b) of segment (s, l) whose modifications are mp[m0 .. m1) / mc (forward coordinates)
    void add_part(uint64_t s, uint32_t l, uint32_t a, uint32_t b, const uint32_t* mp, const uint8_t* mc, uint32_t m0, uint32_t m1) {
        if (b <= a) return;
        const bool minus = s & SEG_MINUS;
        const uint32_t f0 = minus ? l - b : a;   // forward range [f0, f0 + (b - a))
        for (uint32_t x = m0; x < m1; x++)
            if (mp[x] >= f0 && mp[x] < f0 + (b - a)) {
                mod_pos.push_back(mp[x] - f0);
                mod_chr.push_back(mc[x]);
            }
        end_seg((s & ~SEG_OFF_MASK) | ((s & SEG_OFF_MASK) + f0), b - a);
    }
    void add_literal(std::vector<uint8_t>& pool, const uint8_t* codes, uint32_t n) {
        if (!n) return;
        const uint64_t off = pool.size();
        if (off + n > SEG_OFF_MASK) throw std::runtime_error("literal pool too large");
        pool.insert(pool.end(), codes, codes + n);
        end_seg(SEG_LITERAL | off, n);
    }
};

// a molecule somewhere: arrays of a HostBatch or of the carry, literal offsets to be shifted by lit_shift
struct MolView {
    const uint64_t* src;
    const uint32_t* len;
    const uint32_t* mod_begin;   // nullptr: no modifications
    const uint32_t* mod_pos;
    const uint8_t* mod_chr;
    uint32_t n_seg;
    uint64_t lit_shift;
};

void add_molecule(Builder& B, const MolView& M) {
    for (uint32_t s = 0; s < M.n_seg; s++) {
        const uint64_t v = M.src[s];
        const uint32_t m0 = M.mod_begin ? M.mod_begin[s] : 0, m1 = M.mod_begin ? M.mod_begin[s + 1] : 0;
        B.add_part((v & SEG_LITERAL) ? v + M.lit_shift : v, M.len[s], 0, M.len[s], M.mod_pos, M.mod_chr, m0, m1);
    }
}

// Badread's add_glitches on the segment list: pieces of geometric length kept, then `size` random bases put in and
// `skip` bases of the fragment lost, until the fragment is used up
void add_glitches(const Builder& in, Builder& out, std::vector<uint8_t>& pool, const ExtrasConfig& X, const Draw& D,
                  std::vector<uint8_t>& scratch) {
    out.clear();
    size_t seg = 0;
    uint32_t off = 0;   // position inside segment `seg`
    uint64_t i = 0;
    const uint64_t n = in.total;
    auto advance = [&](uint64_t count, bool keep) {
        while (count && seg < in.src.size()) {
            const uint32_t l = in.len[seg];
            const uint32_t take = (uint32_t)std::min<uint64_t>(count, l - off);
            if (keep)
                out.add_part(in.src[seg], l, off, off + take, in.mod_pos.data(), in.mod_chr.data(), in.mod_begin[seg], in.mod_begin[seg + 1]);
            off += take;
            count -= take;
            if (off == l) {
                seg++;
                off = 0;
            }
        }
    };
    for (uint32_t g = 0;; g++) {
        uint32_t w[4];
        D.words(0, 3, g & 0xFFFFFu, w);
        const uint32_t dist = geometric(X.glitch_rate, unit(w[0]));
        advance(dist, true);
        i += dist;
        if (i >= n) break;
        const uint32_t size = geometric0(X.glitch_size, unit(w[1]));
        if (size) {
            scratch.resize(size);
            D.bases(4, g & 0xFFFFFu, size, scratch.data());
            out.add_literal(pool, scratch.data(), size);
        }
        const uint32_t skip = geometric0(X.glitch_skip, unit(w[2]));
        advance(skip, false);
        i += skip;
    }
}

// ---- fused front-end steps: one record's molecule `in` -> `out`
struct FuseDraw {
    uint64_t seed, ordinal;
    uint32_t step;
    void words(uint32_t index, uint32_t out[4]) const {
        philox4x32_10_host(index, TAG_FUSE | (step << 8), (uint32_t)ordinal, (uint32_t)(ordinal >> 32), (uint32_t)seed,
                           (uint32_t)(seed >> 32), out);
    }
};

// the length the polyA module's `int poly_a_len = dist(rand_gen)` would hold, by inversion / Box-Muller /
// Marsaglia-Tsang on Philox uniforms (the truncation toward zero is the module's)
double polya_draw(const FrontStep& S, const FuseDraw& D) {
    uint32_t w[4];
    D.words(0, w);
    const double TWO_PI = 6.283185307179586;
    switch (S.dist) {
        case FrontStep::NORMAL:
            return S.p1 + S.p2 * (std::sqrt(-2.0 * std::log(unit(w[0]))) * std::cos(TWO_PI * unit(w[1])));
        case FrontStep::WEIBULL:
            return S.p2 * std::pow(-std::log(1.0 - unit(w[0])), 1.0 / S.p1);
        case FrontStep::POISSON: {
            const double u = unit(w[0]);
            double pk = std::exp(-S.p1), c = pk;
            int k = 0;
            while (u > c && k < 100000) {
                k++;
                pk *= S.p1 / k;
                c += pk;
            }
            return (double)k;
        }
        case FrontStep::GAMMA: {   // shape p1, scale p2
            const double a = S.p1 < 1.0 ? S.p1 + 1.0 : S.p1;
            const double d = a - 1.0 / 3.0, c = 1.0 / std::sqrt(9.0 * d);
            double x = 0;
            for (uint32_t it = 1; it < 1000; it++) {
                uint32_t v[4];
                D.words(it, v);
                const double z = std::sqrt(-2.0 * std::log(unit(v[0]))) * std::cos(TWO_PI * unit(v[1]));
                const double t = 1.0 + c * z;
                if (t <= 0.0) continue;
                const double t3 = t * t * t;
                if (std::log(unit(v[2])) < 0.5 * z * z + d - d * t3 + d * std::log(t3)) {
                    x = d * t3;
                    break;
                }
            }
            if (S.p1 < 1.0) x *= std::pow(unit(w[3]), 1.0 / S.p1);
            return x * S.p2;
        }
    }
    return 0;
}

// src/util.h:53-92: the bases an IUPAC letter of a tag format stands for
const char* iupac_set(char c) {
    switch (c) {
        case 'A': return "A"; case 'G': return "G"; case 'T': return "T"; case 'C': return "C"; case 'U': return "U";
        case 'R': return "GA"; case 'Y': return "TC"; case 'K': return "GT"; case 'M': return "AC"; case 'S': return "GC";
        case 'W': return "AT"; case 'B': return "GTC"; case 'D': return "GAT"; case 'H': return "ACT"; case 'V': return "GCA";
        case 'N': return "AGCT";
        default: return "";
    }
}
// one sequence of a tag format: letter j takes word j & 3 of counter (base + (j >> 2))
void tag_draw(const std::string& fmt, const FuseDraw& D, uint32_t base, std::vector<uint8_t>& out) {
    out.clear();
    uint32_t w[4] = {0, 0, 0, 0};
    for (size_t j = 0; j < fmt.size(); j++) {
        if ((j & 3u) == 0) D.words(base + (uint32_t)(j >> 2), w);
        const char* set = iupac_set(fmt[j]);
        const size_t k = strlen(set);
        if (!k) continue;   // a letter fmt2seq has no entry for yields nothing
        out.push_back(frag_code((uint8_t)set[((uint64_t)w[j & 3u] * k) >> 32]));
    }
}

void apply_front_end(const FrontEnd& F, uint64_t seed, uint64_t ordinal, const MolView& self, const uint8_t* cb, uint32_t cb_len,
                     bool have_cb, std::vector<uint8_t>& pool, Builder& cur, Builder& tmp, std::vector<uint8_t>& scratch) {
    cur.clear();
    add_molecule(cur, self);
    for (size_t k = 0; k < F.steps.size(); k++) {
        const FrontStep& S = F.steps[k];
        const FuseDraw D{seed, ordinal, (uint32_t)k};
        if (S.kind == FrontStep::POLYA) {
            const double x = polya_draw(S, D);
            long len = x >= 2.0e9 ? 2000000000L : x <= -2.0e9 ? -2000000000L : (long)x;   // (int) conversion: toward zero
            if (len < S.min_len) len = S.min_len;
            if (len > S.max_len) len = S.max_len;
            if (len > 0) {
                scratch.assign((size_t)len, frag_code('A'));
                cur.add_literal(pool, scratch.data(), (uint32_t)len);
            }
        } else if (S.kind == FrontStep::FLIP) {
            uint32_t w[4];
            D.words(0, w);
            if (unit(w[0]) < S.p1) {
                tmp.clear();
                for (size_t s = cur.src.size(); s-- > 0;) {
                    const uint32_t m0 = cur.mod_begin[s], m1 = cur.mod_begin[s + 1];
                    for (uint32_t x = m0; x < m1; x++) {
                        tmp.mod_pos.push_back(cur.mod_pos[x]);
                        tmp.mod_chr.push_back(cur.mod_chr[x]);
                    }
                    tmp.end_seg(cur.src[s] ^ SEG_MINUS, cur.len[s]);
                }
                std::swap(cur, tmp);
            }
        } else if (S.kind == FrontStep::TAG) {
            tag_draw(S.fmt5, D, 0, scratch);
            if (!scratch.empty()) {
                tmp.clear();
                tmp.add_literal(pool, scratch.data(), (uint32_t)scratch.size());
                const MolView rest{cur.src.data(), cur.len.data(), cur.mod_begin.data(), cur.mod_pos.data(), cur.mod_chr.data(),
                                   (uint32_t)cur.src.size(), 0};
                add_molecule(tmp, rest);
                std::swap(cur, tmp);
            }
            tag_draw(S.fmt3, D, 1u << 16, scratch);
            if (!scratch.empty()) cur.add_literal(pool, scratch.data(), (uint32_t)scratch.size());
        } else {   // SCB
            if (!have_cb) throw std::runtime_error("SCB step: the MDF record has no CB comment");
            if (!(cb_len == 1 && cb[0] == '.') && cb_len) {
                scratch.resize(cb_len);
                for (uint32_t i = 0; i < cb_len; i++) scratch[i] = frag_code(ascii_upper(cb[i]));
                cur.add_literal(pool, scratch.data(), cb_len);
            }
        }
    }
}

}  // namespace

FrontEnd parse_front_end(const std::string& spec) {
    FrontEnd F;
    auto bad = [&](const std::string& why) { throw std::runtime_error("Error: --fuse: " + why + " in '" + spec + "'"); };
    auto numbers = [&](const std::string& t) {
        std::vector<double> v;
        size_t pos = 0;
        while (pos <= t.size()) {
            size_t c = t.find(',', pos);
            const std::string tok = t.substr(pos, c == std::string::npos ? std::string::npos : c - pos);
            char* e = nullptr;
            const double x = strtod(tok.c_str(), &e);
            if (tok.empty() || *e) bad("not a number: '" + tok + "'");
            v.push_back(x);
            if (c == std::string::npos) break;
            pos = c + 1;
        }
        return v;
    };
    auto expand = [](std::string f) {   // src/tag.cpp:84-91: a format that starts with a digit is that many N
        if (!f.empty() && isdigit((unsigned char)f[0])) f = std::string((size_t)std::stoi(f), 'N');
        return f;
    };
    size_t pos = 0;
    while (pos < spec.size()) {
        size_t semi = spec.find(';', pos);
        const std::string item = spec.substr(pos, semi == std::string::npos ? std::string::npos : semi - pos);
        pos = semi == std::string::npos ? spec.size() : semi + 1;
        if (item.empty()) continue;
        const size_t colon = item.find(':');
        const std::string name = item.substr(0, colon), arg = colon == std::string::npos ? "" : item.substr(colon + 1);
        FrontStep S;
        if (name == "plA") {
            const size_t eq = arg.find('=');
            if (eq == std::string::npos) bad("plA needs DIST=PARAMS");
            const std::string dn = arg.substr(0, eq);
            const std::vector<double> v = numbers(arg.substr(eq + 1));
            S.kind = FrontStep::POLYA;
            size_t np = 2;
            if (dn == "normal") S.dist = FrontStep::NORMAL;
            else if (dn == "gamma") S.dist = FrontStep::GAMMA;
            else if (dn == "weibull") S.dist = FrontStep::WEIBULL;
            else if (dn == "poisson") { S.dist = FrontStep::POISSON; np = 1; }
            else bad("unknown polyA distribution '" + dn + "'");
            if (v.size() != np && v.size() != np + 2) bad("plA:" + dn + " takes " + std::to_string(np) + " parameters (and optionally min,max)");
            S.p1 = v[0];
            if (np == 2) S.p2 = v[1];
            if (v.size() == np + 2) {
                S.min_len = (int)v[np];
                S.max_len = (int)v[np + 1];
            }
            if (S.min_len < 0 || S.max_len < 0 || S.min_len > S.max_len) bad("polyA length limits");   // src/polyA.cpp:106-117
            if ((S.dist == FrontStep::GAMMA || S.dist == FrontStep::WEIBULL) && !(S.p1 > 0 && S.p2 > 0)) bad("distribution parameters");
            if (S.dist == FrontStep::POISSON && !(S.p1 > 0 && S.p1 < 700)) bad("poisson mean");
        } else if (name == "Flp") {
            const std::vector<double> v = numbers(arg);
            if (v.size() != 1 || v[0] < 0.0 || v[0] > 1.0) bad("Flp takes a probability");   // src/strand_man.cpp:78
            S.kind = FrontStep::FLIP;
            S.p1 = v[0];
        } else if (name == "Tag") {
            const size_t comma = arg.find(',');
            S.kind = FrontStep::TAG;
            S.fmt5 = expand(arg.substr(0, comma));
            S.fmt3 = expand(comma == std::string::npos ? "" : arg.substr(comma + 1));
        } else if (name == "SCB") {
            S.kind = FrontStep::SCB;
        } else {
            bad("unknown step '" + name + "'");
        }
        if (F.steps.size() >= 200) bad("too many steps");
        F.steps.push_back(S);
    }
    return F;
}

HostBatch apply_read_transforms(const TailModel& T, const ExtrasConfig& X, const FrontEnd& F, uint64_t seed,
                                uint64_t ordinal_base, const std::vector<const HostBatch*>& parts, ChimeraCarry* carry) {
    HostBatch out;
    out.ordinal_base = ordinal_base;
    size_t R = 0;
    for (const HostBatch* hb : parts) R += hb->n_reads();
    out.read_mol.reserve(R);
    out.mol_raw_len.reserve(R);
    if (T.loaded) out.mol_tail_len.reserve(R);
    out.mol_seg_begin.reserve(R + 1);
    out.mol_id_begin.reserve(R + 1);
    const bool extras = X.any();
    Builder A, G, Rec, RecPrev, Tmp;   // Rec: the molecule of the current MDF record after the fused front-end steps
    auto view = [](const Builder& b) {
        return MolView{b.src.data(), b.len.data(), b.mod_begin.data(), b.mod_pos.data(), b.mod_chr.data(), (uint32_t)b.src.size(), 0};
    };
    bool have_rec = false;
    std::vector<uint8_t> scratch;
    uint64_t ord = ordinal_base;
    // the molecule of the read before the current one (chimera partner)
    bool have_prev = carry && carry->have;
    MolView prev{};
    uint64_t carry_lit0 = 0;
    if (have_prev && X.chimera > 0) {
        carry_lit0 = out.lit_pool.size();
        out.lit_pool.insert(out.lit_pool.end(), carry->lit.begin(), carry->lit.end());
        prev = MolView{carry->seg_src.data(), carry->seg_len.data(), carry->mod_pos.empty() ? nullptr : carry->seg_mod_begin.data(),
                       carry->mod_pos.data(), carry->mod_chr.data(), (uint32_t)carry->seg_src.size(), carry_lit0};
    } else {
        have_prev = false;
    }
    for (const HostBatch* hbp : parts) {
        const HostBatch& hb = *hbp;
        const uint64_t lit0 = out.lit_pool.size();
        out.lit_pool.insert(out.lit_pool.end(), hb.lit_pool.begin(), hb.lit_pool.end());
        const bool has_mods = !hb.mod_pos.empty();
        const bool has_cb = hb.mol_cb_begin.size() == hb.n_mols() + 1;
        uint32_t cur_m = 0xFFFFFFFFu;
        for (size_t r = 0; r < hb.n_reads(); r++, ord++) {
            const uint32_t m = hb.read_mol[r];
            if (m != cur_m) {
                // first read of an MDF record: its molecule, through the fused front-end steps (keyed by this ordinal)
                cur_m = m;
                if (have_rec && X.chimera > 0) {
                    std::swap(RecPrev, Rec);
                    prev = view(RecPrev);
                    have_prev = true;
                }
                const uint32_t s0 = hb.mol_seg_begin[m], s1 = hb.mol_seg_begin[m + 1];
                const MolView self{hb.seg_src.data() + s0, hb.seg_len.data() + s0, has_mods ? hb.seg_mod_begin.data() + s0 : nullptr,
                                   hb.mod_pos.data(), hb.mod_chr.data(), s1 - s0, lit0};
                const uint8_t* cb = has_cb ? hb.cb_pool.data() + hb.mol_cb_begin[m] : nullptr;
                const uint32_t cb_len = has_cb ? hb.mol_cb_begin[m + 1] - hb.mol_cb_begin[m] : 0;
                apply_front_end(F, seed, ord, self, cb, cb_len, has_cb && cb_len > 0, out.lit_pool, Rec, Tmp, scratch);
                if (Rec.total > 0xFFFFFF00ull) throw std::runtime_error("molecule too long after the fused front-end steps");
                have_rec = true;
            } else if (X.chimera > 0) {
                prev = view(Rec);   // the read before this one comes from the same record
                have_prev = true;
            }
            const uint32_t L = (uint32_t)Rec.total;
            const MolView self = view(Rec);
            const Draw D{seed, ord};
            A.clear();
            uint32_t w[4] = {0, 0, 0, 0}, wa[4] = {0, 0, 0, 0};
            if (extras) {
                D.words(0, 0, 0, w);
                D.words(0, 2, 0, wa);
            }
            // start adapter: its END is what is next to the fragment
            if (X.start_rate > 0 && !X.start_seq.empty() && unit(wa[0]) < X.start_rate) {
                const uint32_t len = (uint32_t)X.start_seq.size(), k = adapter_bases(len, X.start_amount, unit(wa[1]));
                A.add_literal(out.lit_pool, X.start_seq.data() + (len - k), k);
            }
            // the body: junk, random, or the molecule
            const double u0 = unit(w[0]);
            if (X.junk > 0 && u0 < X.junk) {
                const uint32_t ul = 1 + w[1] % 5;
                scratch.resize(L);
                for (uint32_t i = 0; i < L; i++) scratch[i] = (uint8_t)((w[2] >> (2 * (i % ul))) & 3u);
                A.add_literal(out.lit_pool, scratch.data(), L);
            } else if (X.random > 0 && u0 < X.junk + X.random) {
                scratch.resize(L);
                D.bases(1, 0, L, scratch.data());
                A.add_literal(out.lit_pool, scratch.data(), L);
            } else {
                add_molecule(A, self);
            }
            // tail noise (py/tksm_badread.py:335-341): between the molecule and whatever follows it
            uint32_t tl = 0;
            if (T.loaded) {
                uint32_t start = 0;
                tl = T.draw_length(seed, ord, L, &start);
                if ((uint64_t)L + tl > 0xFFFFFF00ull) tl = 0;
                if (tl) {
                    scratch.resize(tl);
                    T.draw_bases(seed, ord, start, tl, scratch.data());
                    A.add_literal(out.lit_pool, scratch.data(), tl);
                }
            }
            // chimera: the read runs on into the molecule of the read before it
            if (X.chimera > 0 && have_prev && unit(w[3]) < X.chimera) {
                uint32_t wc[4];
                D.words(0, 5, 0, wc);
                if (!X.end_seq.empty() && unit(wc[0]) < CHIMERA_END_ADAPTER_CHANCE)
                    A.add_literal(out.lit_pool, X.end_seq.data(), (uint32_t)X.end_seq.size());
                if (!X.start_seq.empty() && unit(wc[1]) < CHIMERA_START_ADAPTER_CHANCE)
                    A.add_literal(out.lit_pool, X.start_seq.data(), (uint32_t)X.start_seq.size());
                add_molecule(A, prev);
            }
            // end adapter: its START is what is next to the fragment
            if (X.end_rate > 0 && !X.end_seq.empty() && unit(wa[2]) < X.end_rate) {
                const uint32_t len = (uint32_t)X.end_seq.size(), k = adapter_bases(len, X.end_amount, unit(wa[3]));
                A.add_literal(out.lit_pool, X.end_seq.data(), k);
            }
            const Builder* F = &A;
            if (X.glitch_rate > 0) {
                add_glitches(A, G, out.lit_pool, X, D, scratch);
                F = &G;
            }
            if (F->total > 0xFFFFFF00ull) throw std::runtime_error("read too long after the Badread extras");
            // the finished molecule of this read
            out.read_mol.push_back((uint32_t)out.mol_raw_len.size());
            const uint32_t mods0 = (uint32_t)out.mod_pos.size();
            for (size_t s = 0; s < F->src.size(); s++) {
                out.seg_src.push_back(F->src[s]);
                out.seg_len.push_back(F->len[s]);
                out.seg_mod_begin.push_back(mods0 + F->mod_begin[s + 1]);
                if (F->src[s] & SEG_LITERAL) out.literal_bases += F->len[s];
                else out.genomic_bases += F->len[s];
            }
            out.mod_pos.insert(out.mod_pos.end(), F->mod_pos.begin(), F->mod_pos.end());
            out.mod_chr.insert(out.mod_chr.end(), F->mod_chr.begin(), F->mod_chr.end());
            out.mol_seg_begin.push_back((uint32_t)out.seg_src.size());
            out.mol_raw_len.push_back((uint32_t)F->total);
            if (T.loaded) out.mol_tail_len.push_back(tl);
            out.id_pool.insert(out.id_pool.end(), hb.id_pool.begin() + hb.mol_id_begin[m], hb.id_pool.begin() + hb.mol_id_begin[m + 1]);
            out.mol_id_begin.push_back((uint32_t)out.id_pool.size());
            out.total_raw_bases += F->total;
        }
    }
    if (carry && X.chimera > 0 && have_rec) {
        // keep the molecule of the last read for the first read of the next batch
        ChimeraCarry c;
        c.have = true;
        for (size_t s = 0; s < Rec.src.size(); s++) {
            uint64_t v = Rec.src[s];
            if (v & SEG_LITERAL) {
                const uint64_t o = v & SEG_OFF_MASK;
                const uint64_t no = c.lit.size();
                c.lit.insert(c.lit.end(), out.lit_pool.begin() + (long)o, out.lit_pool.begin() + (long)(o + Rec.len[s]));
                v = (v & ~SEG_OFF_MASK) | no;
            }
            c.seg_src.push_back(v);
            c.seg_len.push_back(Rec.len[s]);
            for (uint32_t x = Rec.mod_begin[s]; x < Rec.mod_begin[s + 1]; x++) {
                c.mod_pos.push_back(Rec.mod_pos[x]);
                c.mod_chr.push_back(Rec.mod_chr[x]);
            }
            c.seg_mod_begin.push_back((uint32_t)c.mod_pos.size());
        }
        *carry = std::move(c);
    }
    return out;
}

}  // namespace tk
