// TOOL (not product, not test): one side of the A/B parser timing of tools_parser_ab/run.sh; compiled twice, once per revision.
#include <chrono>
#include <cstdio>
#include <string>
#include "host.hpp"
namespace tk { std::string read_text_maybe_gz(const std::string&) { return ""; }
struct Ctx { GenomeStore g; };
void* RUN_MAKE(const std::string& fa) { auto* c = new Ctx; c->g.load_fasta_text(fa.data(), fa.size()); c->g.index_exceptions(); return c; }
double RUN_NAME(void* ctx, const std::string& mdf, unsigned long long* hh) {
    GenomeStore& g = ((Ctx*)ctx)->g;
    auto t0 = std::chrono::steady_clock::now();
    size_t pos = 0; uint64_t h1 = 0;
    while (pos < mdf.size()) {
        MdfParser p(&g);
        HostBatch hb;
        size_t want = std::min<size_t>(mdf.size() - pos, 1u << 20);
        bool fin = pos + want == mdf.size();
        size_t used = p.feed(mdf.data() + pos, want, fin, hb);
        if (!used) { fprintf(stderr, "stuck\n"); return 1; }
        if (hh) for (size_t i = 0; i < hb.seg_src.size(); i++) { uint64_t v = hb.seg_src[i]; h1 = h1 * 1000003 + hb.seg_len[i]; if (v & SEG_LITERAL) { uint64_t off = v & SEG_OFF_MASK; h1 = h1 * 1000003 + (v & ~SEG_OFF_MASK); for (uint32_t j = 0; j < hb.seg_len[i]; j++) h1 = h1 * 1000003 + hb.lit_pool[off + j]; } else h1 = h1 * 1000003 + v; }
        if (hh) { for (auto v : hb.mod_pos) h1 = h1 * 1000003 + v; for (auto v : hb.mod_chr) h1 = h1 * 1000003 + v; for (auto v : hb.read_mol) h1 = h1 * 1000003 + v; for (auto v : hb.id_pool) h1 = h1 * 1000003 + v; for (auto v : hb.seg_mod_begin) h1 = h1 * 1000003 + v; for (auto v : hb.mol_seg_begin) h1 = h1 * 1000003 + v; for (auto v : hb.mol_raw_len) h1 = h1 * 1000003 + v; for (auto v : hb.mol_id_begin) h1 = h1 * 1000003 + v; h1 += hb.total_raw_bases * 7 + hb.genomic_bases * 13 + hb.literal_bases * 17; }
        pos += used;
    }
    if (hh) *hh = h1;
    return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() * 1e3;
}
}
