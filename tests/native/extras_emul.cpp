// TEST CODE: host materialisation of the per-read molecules csrc/extras.cpp builds (apply_read_transforms), so that the
// segment-level Badread extras can be compared with oracle/extras_oracle.py without a GPU (tests/test_extras_host.py).
// usage: extras_emul FASTA MDF SEED ORDINAL_BASE JUNK RANDOM CHIMERAS S_RATE S_AMOUNT E_RATE E_AMOUNT G_RATE G_SIZE G_SKIP S_SEQ E_SEQ [FUSE_SPEC]
// prints the fragment of every read, one per line (what the error-free output of the product shows).
#include <cstdio>
#include <cstring>
#include <fstream>
#include <sstream>
#include "extras.hpp"
using namespace tk;
namespace tk { std::string read_text_maybe_gz(const std::string&) { return ""; } }
static std::string slurp(const char* p) { std::ifstream f(p, std::ios::binary); std::stringstream ss; ss << f.rdbuf(); return ss.str(); }
static char comp(char c) { switch (c) { case 'A': return 'T'; case 'C': return 'G'; case 'G': return 'C'; case 'T': return 'A'; default: return c; } }
int main(int argc, char** argv) {
    GenomeStore g;
    { std::string fa = slurp(argv[1]); g.load_fasta_text(fa.data(), fa.size()); }
    g.index_exceptions();
    std::string mdf = slurp(argv[2]);
    ExtrasConfig X;
    uint64_t seed = strtoull(argv[3], 0, 10), ob = strtoull(argv[4], 0, 10);
    X.junk = atof(argv[5]) / 100; X.random = atof(argv[6]) / 100; X.chimera = atof(argv[7]) / 100;
    X.start_rate = atof(argv[8]) / 100; X.start_amount = atof(argv[9]) / 100; X.end_rate = atof(argv[10]) / 100; X.end_amount = atof(argv[11]) / 100;
    X.glitch_rate = atof(argv[12]); X.glitch_size = atof(argv[13]); X.glitch_skip = atof(argv[14]);
    for (const char* p = argv[15]; *p; p++) X.start_seq.push_back(frag_code(ascii_upper(*p)));
    for (const char* p = argv[16]; *p; p++) X.end_seq.push_back(frag_code(ascii_upper(*p)));
    FrontEnd F;
    if (argc > 17) F = parse_front_end(argv[17]);
    MdfParser p(&g);
    p.capture_cb = F.needs_cb();
    HostBatch hb;
    p.feed(mdf.data(), mdf.size(), true, hb);
    TailModel T;
    HostBatch out = apply_read_transforms(T, X, F, seed, ob, {&hb}, nullptr);
    const bool mods = !out.mod_pos.empty();
    for (size_t r = 0; r < out.n_reads(); r++) {
        uint32_t m = out.read_mol[r];
        std::string frag;
        for (uint32_t s = out.mol_seg_begin[m]; s < out.mol_seg_begin[m + 1]; s++) {
            uint64_t v = out.seg_src[s]; uint32_t len = out.seg_len[s]; bool minus = v & SEG_MINUS; uint64_t off = v & SEG_OFF_MASK;
            std::string fwd;
            if (v & SEG_LITERAL) for (uint32_t i = 0; i < len; i++) { uint8_t c = out.lit_pool[off + i]; fwd.push_back(c < 4 ? "ACGT"[c] : (char)c); }
            else fwd = g.fetch(off, len);
            if (mods) for (uint32_t x = out.seg_mod_begin[s]; x < out.seg_mod_begin[s + 1]; x++) fwd[out.mod_pos[x]] = (char)out.mod_chr[x];
            if (minus) { std::string rc(fwd.rbegin(), fwd.rend()); for (auto& c : rc) c = comp(c); fwd = rc; }
            frag += fwd;
        }
        if (frag.size() != out.mol_raw_len[m]) { fprintf(stderr, "length mismatch read %zu\n", r); return 1; }
        printf("%s\n", frag.c_str());
    }
}
