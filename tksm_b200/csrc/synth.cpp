// Synthetic MDF generator for benchmarks and tests.  Emits the exact text the
// upstream modules write (src/interval.h:898-905: header "+id\tdepth\tcomment", interval lines
// "chr\tstart\tend\tstrand\tmods" with the trailing tab), shaped like the output of
// Tsb (shape 0), Tsb->Trc->plA->Flp->Tag (shape 1, README bulk pipeline), a length-skew
// sweep of single-interval molecules (shape 2) or the scRNA pipeline Tsb->SCB->PCR (shape 3:
// cell barcode + UMI + adapter literals, PCR copies with ".cycle" ids and base mods, src/pcr.cpp:40-64).
#include <algorithm>
#include <charconv>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

#include "engine.hpp"

namespace {

struct Rng {
    uint64_t s;
    explicit Rng(uint64_t seed) : s(seed * 0x9E3779B97F4A7C15ull + 0xD1B54A32D192ED03ull) {}
    uint64_t next() {  // splitmix64
        uint64_t z = (s += 0x9E3779B97F4A7C15ull);
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
        z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
        return z ^ (z >> 31);
    }
    double uni() { return (double)(next() >> 11) * (1.0 / 9007199254740992.0); }
    double normal() {
        double u1 = std::max(uni(), 1e-300), u2 = uni();
        return std::sqrt(-2.0 * std::log(u1)) * std::cos(6.283185307179586 * u2);
    }
    double lognormal(double median, double sigma) { return median * std::exp(sigma * normal()); }
};

struct Seg {
    int contig;          // -1 => literal
    std::string lit;
    uint64_t start, end;
    bool plus;
    std::string mods = {};   // "pos<char>,..." relative to the interval (src/interval.h:754)
};

void put_u64(std::string& o, uint64_t v) {
    char b[24];
    auto r = std::to_chars(b, b + 24, v);
    o.append(b, r.ptr);
}

}  // namespace

extern "C" int tksm_seq_synth_mdf(tksm_seq* h, int shape, uint64_t n_molecules, uint64_t seed, uint32_t min_len,
                                  uint32_t max_len, char** out, size_t* out_len) {
    try {
        const tk::GenomeStore* g = tk::engine_genome(h);
        std::vector<std::pair<std::string, tk::Contig>> contigs(g->contigs.begin(), g->contigs.end());
        if (contigs.empty()) throw std::runtime_error("no reference loaded");
        std::sort(contigs.begin(), contigs.end(), [](auto& a, auto& b) { return a.second.offset < b.second.offset; });
        std::vector<double> cum;
        double tot = 0;
        for (auto& c : contigs) {
            tot += (double)c.second.length;
            cum.push_back(tot);
        }
        Rng rng(seed);
        std::string o;
        o.reserve((size_t)std::min<uint64_t>(n_molecules, 1u << 22) * 300);
        const std::string tag5 = "AATGTACTTCGTTCAGTTACGTATTGCT";  // 28 nt
        const std::string tag3 = "GCAATACGTAACTGAACGAAGT";        // 22 nt
        uint64_t emitted = 0, rec = 0;
        std::vector<Seg> segs;
        std::string pcr_suffix;
        while (emitted < n_molecules) {
            segs.clear();
            int ci = (int)(std::lower_bound(cum.begin(), cum.end(), rng.uni() * tot) - cum.begin());
            ci = std::min<int>(ci, (int)contigs.size() - 1);
            const uint64_t clen = contigs[(size_t)ci].second.length;
            uint64_t depth = 1;
            if (shape == 2) {
                double lo = std::log((double)std::max<uint32_t>(min_len, 1)), hi = std::log((double)std::max(max_len, min_len));
                uint64_t len = (uint64_t)std::exp(lo + (hi - lo) * rng.uni());
                len = std::min<uint64_t>(std::max<uint64_t>(len, 1), clen);
                uint64_t st = (uint64_t)(rng.uni() * (double)(clen - len + 1));
                segs.push_back(Seg{ci, "", st, st + len, rng.uni() < 0.5});
            } else {
                // transcript: 1 + Geom(mean 8) exons, log-normal exon (median 150) and intron (median 1500) sizes
                int nex = 1;
                while (nex < 40 && rng.uni() > 1.0 / 8.0) nex++;
                std::vector<std::pair<uint64_t, uint64_t>> exons;
                uint64_t span = 0;
                std::vector<uint64_t> el((size_t)nex), il((size_t)nex, 0);
                for (int e = 0; e < nex; e++) {
                    el[(size_t)e] = (uint64_t)std::min(std::max(rng.lognormal(150, 0.7), 20.0), 5000.0);
                    if (e) il[(size_t)e] = (uint64_t)std::min(std::max(rng.lognormal(1500, 1.0), 50.0), 100000.0);
                    span += el[(size_t)e] + il[(size_t)e];
                }
                if (span + 2 > clen) continue;
                uint64_t pos = (uint64_t)(rng.uni() * (double)(clen - span));
                for (int e = 0; e < nex; e++) {
                    pos += il[(size_t)e];
                    exons.emplace_back(pos, pos + el[(size_t)e]);
                    pos += el[(size_t)e];
                }
                const bool plus = rng.uni() < 0.5;
                if (!plus) std::reverse(exons.begin(), exons.end());  // transcription order
                for (auto& ex : exons) segs.push_back(Seg{ci, "", ex.first, ex.second, plus});
                if (shape == 3) {
                    // SCB: 16-nt cell barcode + 10-nt UMI literals and the 22-nt adapter in front, polyA behind
                    auto random_lit = [&](int n) {
                        std::string t((size_t)n, 'A');
                        for (auto& c : t) c = "ACGT"[rng.next() & 3];
                        return t;
                    };
                    const std::string cb = random_lit(16), umi = random_lit(10);
                    int pa = (int)std::lround(15 + 7.5 * rng.normal());
                    if (pa > 0) segs.push_back(Seg{-1, std::string((size_t)pa, 'A'), 0, (uint64_t)pa, true});
                    segs.insert(segs.begin(), Seg{-1, umi, 0, umi.size(), true});
                    segs.insert(segs.begin(), Seg{-1, cb, 0, cb.size(), true});
                    segs.insert(segs.begin(), Seg{-1, tag3, 0, tag3.size(), true});
                    // PCR: the copy of cycle path c1.c2...; every cycle mutates bases at rate 4/3 x 2.67e-4
                    // (Taq-setting1, src/pcr.cpp:36) -- here x 5 so that small test inputs carry mods
                    const int cycles = 1 + (int)(rng.next() % 4);
                    for (auto& sg : segs) {
                        const uint64_t len = sg.end - sg.start;
                        const double expect = (double)cycles * 5.0 * (4.0 / 3.0) * 2.67e-4 * (double)len;
                        int nm = (int)expect;
                        nm += rng.uni() < expect - nm;
                        for (int m = 0; m < nm && len > 0; m++) {
                            if (!sg.mods.empty()) sg.mods.push_back(',');
                            put_u64(sg.mods, (uint64_t)(rng.uni() * (double)len));
                            sg.mods.push_back("ACGT"[rng.next() & 3]);
                        }
                    }
                    pcr_suffix.clear();
                    for (int c = 0; c < cycles; c++) {
                        pcr_suffix.push_back('.');
                        put_u64(pcr_suffix, (uint64_t)c + (rng.next() & 1));
                    }
                } else if (shape == 0) {
                    depth = 1;
                    while (depth < 50 && rng.uni() > 0.5) depth++;
                    depth = std::min<uint64_t>(depth, n_molecules - emitted);
                } else {
                    // Trc: keep the 3' end, log-normal target length (median 900)
                    uint64_t target = (uint64_t)std::min(std::max(rng.lognormal(900, 0.6), 100.0), 20000.0);
                    uint64_t total = 0;
                    for (auto& s : segs) total += s.end - s.start;
                    while (total > target && !segs.empty()) {
                        Seg& s = segs.front();
                        uint64_t l = s.end - s.start, cut = std::min(l, total - target);
                        if (cut == l) {
                            segs.erase(segs.begin());
                        } else if (s.plus) {
                            s.start += cut;
                        } else {
                            s.end -= cut;
                        }
                        total -= cut;
                    }
                    // plA: literal A run, N(15, 7.5)
                    int pa = (int)std::lround(15 + 7.5 * rng.normal());
                    if (pa > 0) segs.push_back(Seg{-1, std::string((size_t)pa, 'A'), 0, (uint64_t)pa, true});
                    // Flp: p = 0.5 (src/interval.h:908-920)
                    if (rng.uni() < 0.5) {
                        std::reverse(segs.begin(), segs.end());
                        for (auto& s : segs) s.plus = !s.plus;
                    }
                    // Tag: 5' and 3' literal adapters
                    segs.insert(segs.begin(), Seg{-1, tag5, 0, tag5.size(), true});
                    segs.push_back(Seg{-1, tag3, 0, tag3.size(), true});
                }
            }
            o.push_back('+');
            o.push_back('M');
            put_u64(o, rec);
            o.append(pcr_suffix);
            o.push_back('\t');
            put_u64(o, depth);
            o.append("\ttid=T");
            put_u64(o, rec);
            o.append(";\n");
            for (auto& s : segs) {
                if (s.contig >= 0)
                    o.append(contigs[(size_t)s.contig].first);
                else
                    o.append(s.lit);
                o.push_back('\t');
                put_u64(o, s.start);
                o.push_back('\t');
                put_u64(o, s.end);
                o.append(s.plus ? "\t+\t" : "\t-\t");
                o.append(s.mods);
                o.push_back('\n');
            }
            emitted += depth;
            rec++;
        }
        *out_len = o.size();
        *out = (char*)malloc(std::max<size_t>(o.size(), 1));
        if (!*out) throw std::runtime_error("out of host memory");
        memcpy(*out, o.data(), o.size());
    } catch (const std::exception& e) {
        (void)e;
        return 1;
    }
    return 0;
}
