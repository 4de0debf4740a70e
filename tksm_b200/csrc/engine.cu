// Device engine + C ABI of the Seq hot path (include/tksm_seq.h).
#include <cuda_runtime.h>

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/tksm_seq.h"
#include "common.hpp"
#include "engine.hpp"
#include "host.hpp"
#include "kernels.cuh"
#include "model.hpp"

using namespace tk;

static thread_local std::string g_create_error;

#define CK(call)                                                                                       \
    do {                                                                                               \
        cudaError_t e_ = (call);                                                                       \
        if (e_ != cudaSuccess)                                                                         \
            throw std::runtime_error(std::string("CUDA error: ") + cudaGetErrorString(e_) + " at " +   \
                                     __FILE__ + ":" + std::to_string(__LINE__));                       \
    } while (0)

namespace {

struct DBuf {
    void* p = nullptr;
    size_t cap = 0;
    void ensure(size_t bytes) {
        if (bytes <= cap) return;
        if (p) CK(cudaFree(p));
        p = nullptr;
        cap = 0;
        size_t want = bytes + bytes / 8 + 256;
        CK(cudaMalloc(&p, want));
        cap = want;
    }
    void release() {
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
    }
    template <typename T>
    T* as() const { return (T*)p; }
};

struct HBuf {  // pinned host staging
    void* p = nullptr;
    size_t cap = 0;
    void ensure(size_t bytes) {
        if (bytes <= cap) return;
        if (p) CK(cudaFreeHost(p));
        p = nullptr;
        cap = 0;
        size_t want = bytes + bytes / 8 + 256;
        CK(cudaMallocHost(&p, want));
        cap = want;
    }
    void release() {
        if (p) cudaFreeHost(p);
        p = nullptr;
        cap = 0;
    }
};

__global__ void synth_genome_kernel(uint32_t* words, uint64_t nwords, uint32_t s0, uint32_t s1) {
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    uint64_t quads = (nwords + 3) / 4;
    if (i >= quads) return;
    uint4 v = philox4x32_10((uint32_t)i, (uint32_t)(i >> 32), 0x47454E4Fu, 0, s0, s1);
    uint32_t w[4] = {v.x, v.y, v.z, v.w};
    for (int t = 0; t < 4; t++)
        if (4 * i + t < nwords) words[4 * i + t] = w[t];
}

constexpr size_t SMEM_TOTAL = 232448;  // 227 KB opt-in maximum per CTA on sm_100
constexpr int N_CLASSES = 6;           // teams of 1, 2, 4, 8, 24 warps in shared memory; 24 warps with a global workspace
constexpr int GWS_CLASS = N_CLASSES - 1;
const int CLASS_W[N_CLASSES] = {1, 2, 4, 8, 24, 24};
inline uint32_t list_capacity(uint32_t ncap) { return ncap / 5 * 2 + 64; }

}  // namespace

struct tksm_seq_batch {
    HostBatch hb;
    // device copies
    DBuf blob;
    DevBatch db{};
    std::vector<uint64_t> scr_off, raw_off;
    std::vector<uint32_t> work[N_CLASSES];
    uint32_t* d_work[N_CLASSES] = {};
    uint32_t max_n_gws = 0;
    DBuf scr_seq, scr_qual, scr_raw, per_read, out[2], rec_len[2], rec_off[2], sums, gws, counters;
    uint64_t out_bytes[2] = {0, 0};
    uint64_t out_cap[2] = {0, 0};
    uint64_t alg_in_bytes = 0;
    bool uploaded = false;
    size_t h2d_bytes = 0;
    ~tksm_seq_batch() {
        for (DBuf* d : {&blob, &scr_seq, &scr_qual, &scr_raw, &per_read, &out[0], &out[1], &rec_len[0], &rec_len[1],
                        &rec_off[0], &rec_off[1], &sums, &gws, &counters})
            d->release();
    }
};

struct tksm_seq {
    tksm_seq_config cfg;
    std::string err;
    GenomeStore genome;
    bool genome_dirty = true;
    bool genome_synthetic = false;
    ErrorModelHost em;
    QScoreModelHost qm;
    IdentityHost idh;
    std::vector<float> delta;
    float calib_overshoot = 0, calib_kappa = 0;
    bool models_loaded = false;
    // device
    cudaStream_t stream = nullptr;
    cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
    int num_sms = 148;
    DBuf d_genome, d_exc_start, d_exc_len, d_exc_chr;
    DBuf d_qcls, d_thr, d_alias, d_alts, d_ident, d_delta, d_qdirect, d_qhash, d_qrows, d_qalias, d_t9, d_t9d1;
    DevTables dt{};
    HBuf stage, stage_out;
    uint32_t ncap[N_CLASSES] = {};
    uint32_t cap_list[N_CLASSES] = {};
    uint32_t slab[N_CLASSES] = {};
    size_t smem_bytes[N_CLASSES] = {};
    std::vector<uint16_t> thr;
    bool attrs_set = false;

    ~tksm_seq() {
        for (DBuf* d : {&d_genome, &d_exc_start, &d_exc_len, &d_exc_chr, &d_qcls, &d_thr, &d_alias, &d_alts, &d_ident, &d_delta,
                        &d_qdirect, &d_qhash, &d_qrows, &d_qalias, &d_t9, &d_t9d1})
            d->release();
        stage.release();
        stage_out.release();
        for (auto& e : ev)
            if (e) cudaEventDestroy(e);
        if (stream) cudaStreamDestroy(stream);
    }
};

namespace {

template <typename T>
void upload_vec(DBuf& d, const std::vector<T>& v, cudaStream_t s) {
    d.ensure(std::max<size_t>(v.size() * sizeof(T), 16));
    if (!v.empty()) CK(cudaMemcpyAsync(d.p, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice, s));
}

void need_device(tksm_seq* h) {
    if (h->cfg.host_only) throw std::runtime_error("handle was created host_only: no device work possible");
}

void commit_reference(tksm_seq* h) {
    need_device(h);
    if (!h->genome_dirty) return;
    CK(cudaSetDevice(h->cfg.device));
    if (!h->genome_synthetic) {
        upload_vec(h->d_genome, h->genome.words, h->stream);
    }
    upload_vec(h->d_exc_start, h->genome.exc_start, h->stream);
    upload_vec(h->d_exc_len, h->genome.exc_len, h->stream);
    upload_vec(h->d_exc_chr, h->genome.exc_chr, h->stream);
    CK(cudaStreamSynchronize(h->stream));
    h->dt.genome = h->d_genome.as<uint32_t>();
    h->dt.exc_start = h->d_exc_start.as<uint64_t>();
    h->dt.exc_len = h->d_exc_len.as<uint32_t>();
    h->dt.exc_chr = h->d_exc_chr.as<uint8_t>();
    h->dt.n_exc = (uint32_t)h->genome.exc_start.size();
    h->genome_dirty = false;
}

void compute_classes(tksm_seq* h) {
    const size_t nq = h->em.has_table ? ((size_t)1 << (2 * h->em.k)) : 0;
    const size_t qcls_bytes = (nq + 15) & ~size_t(15);
    const size_t pool = SMEM_TOTAL - qcls_bytes - 64 - 128;
    for (int c = 0; c < GWS_CLASS; c++) {
        const int W = CLASS_W[c], teams = SIM_WARPS / W;
        const uint32_t hist_elem = W >= 8 ? 8 : 4;
        const size_t budget = (pool / teams) & ~size_t(15);
        uint32_t ncap = 16;
        while (slab_layout(W, hist_elem, ncap + 16, list_capacity(ncap + 16)).bytes <= budget) ncap += 16;
        h->ncap[c] = ncap;
        h->cap_list[c] = list_capacity(ncap);
        h->slab[c] = slab_layout(W, hist_elem, ncap, h->cap_list[c]).bytes;
        h->smem_bytes[c] = qcls_bytes + 64 + (size_t)h->slab[c] * teams + 128;
    }
    h->smem_bytes[GWS_CLASS] = qcls_bytes + 64 + 128;
}

void upload_models(tksm_seq* h) {
    if (h->cfg.host_only) return;
    CK(cudaSetDevice(h->cfg.device));
    upload_vec(h->d_qcls, h->em.qcls, h->stream);
    upload_vec(h->d_thr, h->thr, h->stream);
    upload_vec(h->d_alias, h->em.alias, h->stream);
    upload_vec(h->d_alts, h->em.alts, h->stream);
    upload_vec(h->d_ident, h->idh.tab, h->stream);
    upload_vec(h->d_delta, h->delta, h->stream);
    upload_vec(h->d_qdirect, h->qm.direct, h->stream);
    upload_vec(h->d_qhash, h->qm.hash, h->stream);
    upload_vec(h->d_qrows, h->qm.rows, h->stream);
    upload_vec(h->d_qalias, h->qm.alias, h->stream);
    upload_vec(h->d_t9, h->qm.t9, h->stream);
    upload_vec(h->d_t9d1, h->qm.t9d1, h->stream);
    CK(cudaStreamSynchronize(h->stream));
    DevTables& T = h->dt;
    T.k = h->em.k;
    T.has_err_table = h->em.has_table;
    T.qcls = h->d_qcls.as<uint8_t>();
    T.thr = h->d_thr.as<uint16_t>();
    T.err_alias = h->d_alias.as<uint32_t>();
    T.err_alts = h->d_alts.as<uint64_t>();
    T.ident_const = h->idh.constant;
    T.ident_mean = h->idh.mean;
    T.ident_tab = h->d_ident.as<float>();
    T.delta_tab = h->d_delta.as<float>();
    T.calib_overshoot = h->calib_overshoot;
    T.calib_kappa = h->calib_kappa;
    T.kq = h->qm.kq;
    T.q_direct = h->d_qdirect.as<uint32_t>();
    for (int i = 0; i < 5; i++) T.q_direct_off[i] = h->qm.direct_off[i];
    T.q_hash = h->d_qhash.as<QHashEntry>();
    T.q_hash_shift = h->qm.hash_shift;
    T.q_hash_mask = h->qm.hash_mask;
    T.q_rows = h->d_qrows.as<QRow>();
    T.q_alias = h->d_qalias.as<uint32_t>();
    T.q_t9 = h->qm.t9.empty() ? nullptr : h->d_t9.as<uint32_t>();
    T.q_t9d1 = h->qm.t9d1.empty() ? nullptr : h->d_t9d1.as<uint32_t>();
    T.seed_lo = (uint32_t)h->cfg.seed;
    T.seed_hi = (uint32_t)(h->cfg.seed >> 32);
}

template <int W, bool GWS>
void launch_sim(tksm_seq* h, const SimParams& P, size_t smem, int grid) {
    auto kern = simulate_kernel<W, GWS>;
    CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<grid, SIM_THREADS, smem, h->stream>>>(P);
    CK(cudaGetLastError());
}

inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// builds the upload blob and per-read scratch layout; uploads once
void batch_upload(tksm_seq* h, tksm_seq_batch* b) {
    if (b->uploaded) return;
    need_device(h);
    if (!h->models_loaded) throw std::runtime_error("models not loaded");
    commit_reference(h);
    HostBatch& hb = b->hb;
    const size_t R = hb.n_reads(), M = hb.n_mols(), S = hb.seg_src.size();
    const uint32_t k = (uint32_t)h->em.k;
    b->scr_off.resize(R + 1);
    b->raw_off.resize(R + 1);
    uint64_t so = 0, ro = 0;
    for (int c = 0; c < N_CLASSES; c++) b->work[c].clear();
    b->max_n_gws = 0;
    for (size_t r = 0; r < R; r++) {
        uint32_t L = hb.mol_raw_len[hb.read_mol[r]];
        b->scr_off[r] = so;
        b->raw_off[r] = ro;
        so += align_up((size_t)2 * L + 32, 16);
        ro += align_up((size_t)L, 16);
        uint32_t n = L + 2 * k;
        int c = 0;
        while (c < GWS_CLASS && n > h->ncap[c]) c++;
        b->work[c].push_back((uint32_t)r);
        if (c == GWS_CLASS) b->max_n_gws = std::max(b->max_n_gws, n);
    }
    b->scr_off[R] = so;
    b->raw_off[R] = ro;
    // longest first inside a class: the tail of a dynamic schedule is then made of short reads
    for (int c = 0; c < N_CLASSES; c++)
        std::stable_sort(b->work[c].begin(), b->work[c].end(), [&](uint32_t a, uint32_t d) {
            return hb.mol_raw_len[hb.read_mol[a]] > hb.mol_raw_len[hb.read_mol[d]];
        });

    const bool mods = !hb.mod_pos.empty();
    struct Part {
        const void* src;
        size_t bytes;
        size_t off;
    };
    std::vector<Part> parts;
    size_t total = 0;
    auto add = [&](const void* p, size_t bytes) {
        size_t off = total;
        parts.push_back({p, bytes, off});
        total = align_up(total + std::max<size_t>(bytes, 1), 256);
        return off;
    };
    size_t o_read_mol = add(hb.read_mol.data(), R * 4);
    size_t o_msb = add(hb.mol_seg_begin.data(), (M + 1) * 4);
    size_t o_mrl = add(hb.mol_raw_len.data(), M * 4);
    size_t o_mib = add(hb.mol_id_begin.data(), (M + 1) * 4);
    size_t o_src = add(hb.seg_src.data(), S * 8);
    size_t o_len = add(hb.seg_len.data(), S * 4);
    size_t o_smb = add(hb.seg_mod_begin.data(), mods ? (S + 1) * 4 : 0);
    size_t o_mp = add(hb.mod_pos.data(), hb.mod_pos.size() * 4);
    size_t o_mc = add(hb.mod_chr.data(), hb.mod_chr.size());
    size_t o_lit = add(hb.lit_pool.data(), hb.lit_pool.size());
    size_t o_id = add(hb.id_pool.data(), hb.id_pool.size());
    size_t o_so = add(b->scr_off.data(), (R + 1) * 8);
    size_t o_ro = add(b->raw_off.data(), (R + 1) * 8);
    size_t o_work[N_CLASSES];
    for (int c = 0; c < N_CLASSES; c++) o_work[c] = add(b->work[c].data(), b->work[c].size() * 4);

    CK(cudaSetDevice(h->cfg.device));
    h->stage.ensure(total);
    for (auto& p : parts)
        if (p.bytes) memcpy((char*)h->stage.p + p.off, p.src, p.bytes);
    b->blob.ensure(total);
    CK(cudaMemcpyAsync(b->blob.p, h->stage.p, total, cudaMemcpyHostToDevice, h->stream));
    b->h2d_bytes = total;
    char* base = (char*)b->blob.p;
    DevBatch& D = b->db;
    D.n_reads = (uint32_t)R;
    D.ordinal_base = hb.ordinal_base;
    D.read_mol = (const uint32_t*)(base + o_read_mol);
    D.mol_seg_begin = (const uint32_t*)(base + o_msb);
    D.mol_raw_len = (const uint32_t*)(base + o_mrl);
    D.mol_id_begin = (const uint32_t*)(base + o_mib);
    D.seg_src = (const uint64_t*)(base + o_src);
    D.seg_len = (const uint32_t*)(base + o_len);
    D.seg_mod_begin = mods ? (const uint32_t*)(base + o_smb) : nullptr;
    D.mod_pos = (const uint32_t*)(base + o_mp);
    D.mod_chr = (const uint8_t*)(base + o_mc);
    D.lit_pool = (const uint8_t*)(base + o_lit);
    D.id_pool = (const uint8_t*)(base + o_id);
    D.scr_off = (const uint64_t*)(base + o_so);
    D.raw_off = (const uint64_t*)(base + o_ro);
    for (int c = 0; c < N_CLASSES; c++) b->d_work[c] = (uint32_t*)(base + o_work[c]);

    b->per_read.ensure(std::max<size_t>(R, 1) * 16);
    D.read_len = b->per_read.as<uint32_t>();
    D.ident_num = D.read_len + R;
    D.ident_den = D.ident_num + R;
    D.status = D.ident_den + R;
    b->counters.ensure(64);
    // ceil(L_genomic / 4) + L_literal + 24 per interval + 5 per mod
    uint64_t per_read_segs = 0, per_read_mods = 0;
    for (size_t r = 0; r < R; r++) {
        uint32_t m = hb.read_mol[r];
        per_read_segs += hb.mol_seg_begin[m + 1] - hb.mol_seg_begin[m];
        if (mods) per_read_mods += hb.seg_mod_begin[hb.mol_seg_begin[m + 1]] - hb.seg_mod_begin[hb.mol_seg_begin[m]];
    }
    b->alg_in_bytes = (hb.genomic_bases + 3) / 4 + hb.literal_bases + 24 * per_read_segs + 5 * per_read_mods;
    CK(cudaStreamSynchronize(h->stream));
    b->uploaded = true;
}

void run_scan(tksm_seq* h, tksm_seq_batch* b, int o, uint32_t R, uint64_t* d_total) {
    const uint32_t per = SCAN_BLOCK * SCAN_ITEMS;
    const uint32_t nb = (R + per - 1) / per;
    b->sums.ensure((size_t)(nb + 2) * 8);
    scan_block_sums<<<nb, SCAN_BLOCK, 0, h->stream>>>(b->rec_len[o].as<uint32_t>(), R, b->sums.as<uint64_t>());
    scan_sums<<<1, 1024, 0, h->stream>>>(b->sums.as<uint64_t>(), nb, d_total);
    scan_final<<<nb, SCAN_BLOCK, 0, h->stream>>>(b->rec_len[o].as<uint32_t>(), R, b->sums.as<uint64_t>(),
                                                 b->rec_off[o].as<uint64_t>());
    CK(cudaGetLastError());
}

void batch_run(tksm_seq* h, tksm_seq_batch* b, int which, int format, tksm_seq_stats* st) {
    batch_upload(h, b);
    CK(cudaSetDevice(h->cfg.device));
    HostBatch& hb = b->hb;
    const uint32_t R = (uint32_t)hb.n_reads();
    tksm_seq_stats s{};
    s.reads = R;
    s.raw_bases = hb.total_raw_bases;
    s.h2d_bytes = b->h2d_bytes;
    b->out_bytes[0] = b->out_bytes[1] = 0;
    if (R == 0) {
        if (st) *st = s;
        return;
    }
    const bool want_bad = which & TKSM_SEQ_BADREAD, want_perf = which & TKSM_SEQ_PERFECT;
    const bool fastq_o[2] = {(format & TKSM_SEQ_FASTQ_BADREAD) != 0, (format & TKSM_SEQ_FASTQ_PERFECT) != 0};
    const bool fastq = fastq_o[0];
    DevBatch& D = b->db;
    uint32_t flags = 0;
    if (want_bad) flags |= RUN_ERRORS;
    if (want_bad && fastq && !h->cfg.skip_qual_compute) flags |= RUN_QUAL;  // py/sequence.py:349-353
    if (want_perf) flags |= RUN_RAW;
    if (want_bad) {
        b->scr_seq.ensure(b->scr_off[R]);
        if (flags & RUN_QUAL) b->scr_qual.ensure(b->scr_off[R]);
    }
    if (want_perf) b->scr_raw.ensure(std::max<uint64_t>(b->raw_off[R], 16));
    D.scr_seq = b->scr_seq.as<uint8_t>();
    D.scr_qual = b->scr_qual.as<uint8_t>();
    D.scr_raw = want_perf ? b->scr_raw.as<uint8_t>() : nullptr;
    // output upper bounds
    uint64_t idsum = 0;
    for (uint32_t r = 0; r < R; r++) {
        uint32_t m = hb.read_mol[r];
        idsum += hb.mol_id_begin[m + 1] - hb.mol_id_begin[m];
    }
    uint64_t cap_bad = (uint64_t)R * 128 + idsum + b->scr_off[R] * (fastq ? 2 : 1);
    uint64_t cap_perf = (uint64_t)R * 128 + idsum + hb.total_raw_bases * (fastq_o[1] ? 2 : 1);
    if (want_bad) b->out[0].ensure(cap_bad);
    if (want_perf) b->out[1].ensure(cap_perf);
    for (int o = 0; o < 2; o++)
        if ((o == 0 && want_bad) || (o == 1 && want_perf)) {
            b->rec_len[o].ensure((size_t)R * 4);
            b->rec_off[o].ensure((size_t)R * 8);
        }
    if (b->max_n_gws) {
        const uint32_t ncap = (uint32_t)align_up(b->max_n_gws + 16, 16);
        h->ncap[GWS_CLASS] = ncap;
        h->cap_list[GWS_CLASS] = ncap / 2 + 64;
        h->slab[GWS_CLASS] = slab_layout(CLASS_W[GWS_CLASS], 8, ncap, h->cap_list[GWS_CLASS]).bytes;
        b->gws.ensure((size_t)h->slab[GWS_CLASS] * h->num_sms);
    }
    uint64_t* d_totals = (uint64_t*)((char*)b->counters.p + 32);
    CK(cudaMemsetAsync(b->counters.p, 0, 64, h->stream));

    CK(cudaEventRecord(h->ev[0], h->stream));
    uint32_t launches = 0;
    for (int c = 0; c < N_CLASSES; c++) {
        if (b->work[c].empty()) continue;
        SimParams P;
        P.T = h->dt;
        P.B = D;
        P.work = b->d_work[c];
        P.n_work = (uint32_t)b->work[c].size();
        P.counter = b->counters.as<uint32_t>() + c;
        P.flags = flags;
        P.ncap = h->ncap[c];
        P.cap_list = h->cap_list[c];
        P.slab_bytes = h->slab[c];
        P.gws = b->gws.as<uint8_t>();
        const int teams = c == GWS_CLASS ? 1 : SIM_WARPS / CLASS_W[c];
        int grid = (int)std::min<size_t>((size_t)h->num_sms, (P.n_work + teams - 1) / teams);
        switch (c) {
            case 0: launch_sim<1, false>(h, P, h->smem_bytes[c], grid); break;
            case 1: launch_sim<2, false>(h, P, h->smem_bytes[c], grid); break;
            case 2: launch_sim<4, false>(h, P, h->smem_bytes[c], grid); break;
            case 3: launch_sim<8, false>(h, P, h->smem_bytes[c], grid); break;
            case 4: launch_sim<24, false>(h, P, h->smem_bytes[c], grid); break;
            default: launch_sim<24, true>(h, P, h->smem_bytes[c], grid); break;
        }
        launches++;
    }
    CK(cudaEventRecord(h->ev[1], h->stream));
    for (int o = 0; o < 2; o++) {
        if (!((o == 0 && want_bad) || (o == 1 && want_perf))) continue;
        FmtParams F{};
        F.n_reads = R;
        F.ordinal_base = hb.ordinal_base;
        F.seed_lo = h->dt.seed_lo;
        F.seed_hi = h->dt.seed_hi;
        F.read_mol = D.read_mol;
        F.mol_raw_len = D.mol_raw_len;
        F.mol_id_begin = D.mol_id_begin;
        F.id_pool = D.id_pool;
        if (o == 0) {
            F.read_len = D.read_len;
            F.ident_num = D.ident_num;
            F.ident_den = D.ident_den;
            F.seq = D.scr_seq;
            F.seq_off = D.scr_off;
            F.qual = (flags & RUN_QUAL) ? D.scr_qual : nullptr;
        } else {
            F.read_len = nullptr;
            F.seq = D.scr_raw;
            F.seq_off = D.raw_off;
            F.qual = nullptr;
        }
        F.fastq = fastq_o[o];
        F.kappa = (double)h->calib_kappa;
        F.rec_len = b->rec_len[o].as<uint32_t>();
        F.rec_off = b->rec_off[o].as<uint64_t>();
        F.out = b->out[o].as<uint8_t>();
        reclen_kernel<<<(R + 255) / 256, 256, 0, h->stream>>>(F);
        run_scan(h, b, o, R, d_totals + o);
        int grid = (int)std::min<uint32_t>((R + 7) / 8, (uint32_t)h->num_sms * 8);
        format_kernel<<<grid, 256, 0, h->stream>>>(F);
        CK(cudaGetLastError());
        launches += 5;
    }
    CK(cudaEventRecord(h->ev[2], h->stream));
    uint64_t totals[2] = {0, 0};
    CK(cudaMemcpyAsync(totals, d_totals, 16, cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    float ms_all = 0, ms_sim = 0;
    CK(cudaEventElapsedTime(&ms_all, h->ev[0], h->ev[2]));
    CK(cudaEventElapsedTime(&ms_sim, h->ev[0], h->ev[1]));
    b->out_bytes[0] = want_bad ? totals[0] : 0;
    b->out_bytes[1] = want_perf ? totals[1] : 0;
    if (b->out_bytes[0] > cap_bad || b->out_bytes[1] > cap_perf) throw std::runtime_error("output exceeds its bound");
    s.out_bytes[0] = b->out_bytes[0];
    s.out_bytes[1] = b->out_bytes[1];
    s.kernel_ms = ms_all;
    s.simulate_ms = ms_sim;
    s.kernel_launches = launches;
    s.algorithmic_bytes = b->alg_in_bytes + b->out_bytes[0] + b->out_bytes[1];
    // read_bases and status: small reductions on the host side of the per-read arrays
    if (want_bad) {
        std::vector<uint32_t> tmp((size_t)R * 4);
        CK(cudaMemcpy(tmp.data(), D.read_len, (size_t)R * 16, cudaMemcpyDeviceToHost));
        s.d2h_bytes += (size_t)R * 16;
        uint64_t rb = 0;
        uint32_t stor = 0;
        for (uint32_t r = 0; r < R; r++) {
            rb += tmp[r];
            stor |= tmp[(size_t)3 * R + r];
        }
        s.read_bases = rb;
        s.status_or = stor;
        if (stor & (ST_OPS_OVERFLOW | ST_SCRATCH_OVERFLOW))
            throw std::runtime_error("a read outgrew its scratch slot (insertion-heavy read); rerun unsupported");
    } else {
        s.read_bases = hb.total_raw_bases;
    }
    if (st) *st = s;
}

void batch_fetch(tksm_seq* h, tksm_seq_batch* b, int o, char* dst) {
    if (b->out_bytes[o] == 0) return;
    CK(cudaSetDevice(h->cfg.device));
    CK(cudaMemcpyAsync(dst, b->out[o].p, b->out_bytes[o], cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
}

}  // namespace

// ============================================================== C ABI ====

extern "C" {

void tksm_seq_default_config(tksm_seq_config* c) {
    memset(c, 0, sizeof(*c));
    c->device = 0;
    c->seed = 42;
    c->identity_mean = 84.0;
    c->identity_max = 99.0;
    c->identity_stdev = 5.5;
    c->skip_qual_compute = 0;
    c->calib_overshoot = NAN;
    c->calib_kappa = NAN;
    c->batch_reads = 0;
    c->batch_bases = 0;
    c->host_only = 0;
}

#define API_BEGIN try {
#define API_END(h)                                  \
    }                                               \
    catch (const std::exception& e) {               \
        if (h) (h)->err = e.what();                 \
        else g_create_error = e.what();             \
        return 1;                                   \
    }                                               \
    return 0;

tksm_seq* tksm_seq_create(const tksm_seq_config* cfg) {
    std::unique_ptr<tksm_seq> h(new tksm_seq());
    try {
        if (cfg) h->cfg = *cfg; else tksm_seq_default_config(&h->cfg);
        h->idh = make_identities(h->cfg.identity_mean, h->cfg.identity_stdev, h->cfg.identity_max);
        if (!h->cfg.host_only) {
            int ndev = 0;
            cudaError_t e = cudaGetDeviceCount(&ndev);
            if (e != cudaSuccess || ndev == 0)
                throw std::runtime_error(std::string("no CUDA device available (") + cudaGetErrorString(e) +
                                         "); the Seq hot path has no CPU fallback");
            if (h->cfg.device < 0 || h->cfg.device >= ndev) throw std::runtime_error("invalid device ordinal");
            CK(cudaSetDevice(h->cfg.device));
            cudaDeviceProp prop;
            CK(cudaGetDeviceProperties(&prop, h->cfg.device));
            if (prop.major < 10)
                throw std::runtime_error(std::string("device ") + prop.name + " is not sm_100; this build targets B200 only");
            h->num_sms = prop.multiProcessorCount;
            CK(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
            for (auto& e2 : h->ev) CK(cudaEventCreate(&e2));
        }
    } catch (const std::exception& e) {
        g_create_error = e.what();
        return nullptr;
    }
    return h.release();
}

void tksm_seq_destroy(tksm_seq* h) { delete h; }

const char* tksm_seq_last_error(const tksm_seq* h) { return h ? h->err.c_str() : g_create_error.c_str(); }

void tksm_seq_free(void* p) { free(p); }

int tksm_seq_load_references(tksm_seq* h, const char* const* paths, int n) {
    API_BEGIN
    for (int i = 0; i < n; i++) {
        printf("Loading reference %s...\n", paths[i]);  // py/sequence.py:192
        h->genome.load_fasta(paths[i]);
    }
    h->genome_dirty = true;
    API_END(h)
}

int tksm_seq_add_reference_text(tksm_seq* h, const char* text, size_t len) {
    API_BEGIN
    h->genome.load_fasta_text(text, len);
    h->genome_dirty = true;
    API_END(h)
}

int tksm_seq_synth_reference(tksm_seq* h, const char* const* names, const uint64_t* lengths, int n_contigs,
                             double n_frac, uint64_t seed) {
    API_BEGIN
    need_device(h);
    CK(cudaSetDevice(h->cfg.device));
    GenomeStore& g = h->genome;
    g = GenomeStore();
    uint64_t rng = seed * 0x9E3779B97F4A7C15ull + 12345;
    auto next = [&]() {
        rng ^= rng << 13;
        rng ^= rng >> 7;
        rng ^= rng << 17;
        return rng;
    };
    for (int i = 0; i < n_contigs; i++) {
        g.n_bases = (g.n_bases + 63) & ~uint64_t(63);
        uint64_t off = g.n_bases, len = lengths[i];
        g.contigs[names[i]] = Contig{off, len};
        // 'N' runs of 1..50 kbp until n_frac of the contig is covered
        uint64_t want = (uint64_t)(n_frac * (double)len), have = 0;
        std::vector<std::pair<uint64_t, uint32_t>> runs;
        while (have < want && len > 100) {
            uint32_t rl = (uint32_t)(1 + next() % 50000);
            rl = (uint32_t)std::min<uint64_t>(rl, std::min<uint64_t>(want - have, len / 4 + 1));
            uint64_t st = next() % (len - rl + 1);
            runs.emplace_back(off + st, rl);
            have += rl;
        }
        std::sort(runs.begin(), runs.end());
        uint64_t last_end = 0;
        for (auto& r : runs) {
            if (r.first < last_end) continue;  // drop overlaps
            g.exc_start.push_back(r.first);
            g.exc_len.push_back(r.second);
            g.exc_chr.push_back('N');
            last_end = r.first + r.second;
        }
        g.n_bases = off + len;
    }
    uint64_t nwords = ((g.n_bases + 63) / 64) * 4 + 4;
    h->d_genome.ensure(nwords * 4);
    uint64_t quads = (nwords + 3) / 4;
    synth_genome_kernel<<<(unsigned)((quads + 255) / 256), 256, 0, h->stream>>>(h->d_genome.as<uint32_t>(), nwords,
                                                                                (uint32_t)seed, (uint32_t)(seed >> 32));
    CK(cudaGetLastError());
    CK(cudaStreamSynchronize(h->stream));
    h->genome_synthetic = true;
    h->genome_dirty = true;
    API_END(h)
}

int tksm_seq_fetch_reference(tksm_seq* h, const char* contig, uint64_t start, uint64_t len, char* out) {
    API_BEGIN
    auto it = h->genome.contigs.find(contig);
    if (it == h->genome.contigs.end()) throw std::runtime_error(std::string("unknown contig ") + contig);
    if (start + len > it->second.length) throw std::runtime_error("fetch beyond contig end");
    uint64_t g0 = it->second.offset + start;
    if (h->genome_synthetic) {
        need_device(h);
        CK(cudaSetDevice(h->cfg.device));
        uint64_t w0 = g0 >> 4, w1 = (g0 + len + 15) >> 4;
        h->genome.words.assign((size_t)(w1 + 1), 0);
        if (w1 > w0)
            CK(cudaMemcpy(h->genome.words.data() + w0, h->d_genome.as<uint32_t>() + w0, (w1 - w0) * 4,
                          cudaMemcpyDeviceToHost));
    }
    std::string s = h->genome.fetch(g0, len);
    memcpy(out, s.data(), len);
    if (h->genome_synthetic) std::vector<uint32_t>().swap(h->genome.words);
    API_END(h)
}

int tksm_seq_commit_reference(tksm_seq* h) {
    API_BEGIN
    commit_reference(h);
    API_END(h)
}

int tksm_seq_load_models(tksm_seq* h, const char* error_model, const char* qscore_model, const char* tail_model) {
    API_BEGIN
    std::string env;
#ifdef TKSM_MODELS_PATH
    env = TKSM_MODELS_PATH;  // compiled-in directory goes first (src/sequence.cpp:38-52)
#endif
    if (const char* e = getenv("TKSM_MODELS")) {
        if (!env.empty()) env += ":";
        env += e;
    }
    ModelNames names = scan_model_dirs(env.empty() ? nullptr : env.c_str());
    std::string em = error_model ? error_model : (names.error.count("nanopore2020") ? "nanopore2020" : "random");
    std::string qm = qscore_model ? qscore_model : (names.qscore.count("nanopore2020") ? "nanopore2020" : "random");
    std::string tail = tail_model ? tail_model : "no_noise";
    if (tail != "no_noise")
        throw std::runtime_error("tail-noise models are not implemented in this build (none is shipped with tksm); use no_noise");
    fprintf(stderr, "\nLoading error model %s\n", em.c_str());
    h->em = load_error_model(em, names.error);
    fprintf(stderr, "Loading qscore model %s\n", qm.c_str());
    h->qm = load_qscore_model(qm, names.qscore);
    h->delta = make_delta_table(h->em);
    h->thr = make_threshold_table(h->em, h->delta);
    // alignment explain-away coefficient of the PB1 edit statistics per error model
    // (oracle/fit_calibration.py measures it with the CPU twin) and the reference's mean overshoot
    // in errors for fragments it re-aligns whole (py/tksm_badread.py:408-413)
    {
        std::string base = em.substr(em.find_last_of('/') == std::string::npos ? 0 : em.find_last_of('/') + 1);
        base = base.substr(0, base.find('.'));
        double kappa = 0.45;
        if (base == "nanopore2020") kappa = 0.43;
        else if (base == "nanopore2018") kappa = 0.40;
        else if (base == "pacbio2016") kappa = 0.58;
        else if (base == "random") kappa = 0.44;
        h->calib_overshoot = std::isnan(h->cfg.calib_overshoot) ? 0.8f : (float)h->cfg.calib_overshoot;
        h->calib_kappa = std::isnan(h->cfg.calib_kappa) ? (float)kappa : (float)h->cfg.calib_kappa;
    }
    compute_classes(h);
    upload_models(h);
    h->models_loaded = true;
    API_END(h)
}

int tksm_seq_get_tables(tksm_seq* h, tksm_seq_tables* t) {
    API_BEGIN
    if (!h->models_loaded) throw std::runtime_error("models not loaded");
    memset(t, 0, sizeof(*t));
    t->k = h->em.k;
    t->has_err_table = h->em.has_table;
    t->q16 = h->em.q16.data();
    t->n_q16 = h->em.q16.size();
    t->err_alias = h->em.alias.data();
    t->n_err_alias = h->em.alias.size();
    t->err_alts = h->em.alts.data();
    t->n_err_alts = h->em.alts.size();
    t->ident_const = h->idh.constant;
    t->ident_mean = h->idh.mean;
    t->ident_tab = h->idh.tab.data();
    t->n_ident_tab = h->idh.tab.size();
    t->delta_tab = h->delta.data();
    t->n_delta_tab = h->delta.size();
    t->calib_overshoot = h->calib_overshoot;
    t->calib_kappa = h->calib_kappa;
    t->kq = h->qm.kq;
    t->q_direct = h->qm.direct.data();
    t->n_q_direct = h->qm.direct.size();
    for (int i = 0; i < 5; i++) t->q_direct_off[i] = h->qm.direct_off[i];
    t->q_hash = (const uint64_t*)h->qm.hash.data();
    t->n_q_hash = h->qm.hash.size();
    t->q_hash_shift = h->qm.hash_shift;
    t->q_hash_mask = h->qm.hash_mask;
    t->q_rows = h->qm.rows.data();
    t->n_q_rows = h->qm.rows.size() / 2;
    t->q_alias = h->qm.alias.data();
    t->n_q_alias = h->qm.alias.size();
    t->seed = h->cfg.seed;
    t->compute_qual = !h->cfg.skip_qual_compute;
    t->qcls = h->em.qcls.data();
    t->n_qcls = h->em.qcls.size();
    t->qrep = h->em.qrep.data();
    t->n_qrep = h->em.qrep.size();
    t->thr = h->thr.data();
    t->n_thr = h->thr.size();
    API_END(h)
}

tksm_seq_batch* tksm_seq_batch_create(tksm_seq* h, const char* mdf, size_t mdf_len, uint64_t ordinal_base) {
    std::unique_ptr<tksm_seq_batch> b(new tksm_seq_batch());
    try {
        MdfParser parser(&h->genome);
        b->hb.ordinal_base = ordinal_base;
        parser.feed(mdf, mdf_len, true, b->hb);
        if (!h->cfg.host_only) batch_upload(h, b.get());
    } catch (const std::exception& e) {
        h->err = e.what();
        return nullptr;
    }
    return b.release();
}

int tksm_seq_batch_run(tksm_seq* h, tksm_seq_batch* b, int which, int format, tksm_seq_stats* stats) {
    API_BEGIN
    batch_run(h, b, which, format, stats);
    API_END(h)
}

int tksm_seq_batch_fetch(tksm_seq* h, tksm_seq_batch* b, int which, char** out, size_t* out_len) {
    API_BEGIN
    int o = which == TKSM_SEQ_PERFECT ? 1 : 0;
    *out_len = b->out_bytes[o];
    *out = (char*)malloc(std::max<size_t>(*out_len, 1));
    if (!*out) throw std::runtime_error("out of host memory");
    batch_fetch(h, b, o, *out);
    API_END(h)
}

void tksm_seq_batch_destroy(tksm_seq*, tksm_seq_batch* b) { delete b; }

int tksm_seq_simulate_text(tksm_seq* h, const char* mdf, size_t mdf_len, int which, int format, uint64_t ordinal_base,
                           char** out, size_t* out_len, tksm_seq_stats* stats) {
    API_BEGIN
    if (which != TKSM_SEQ_BADREAD && which != TKSM_SEQ_PERFECT)
        throw std::runtime_error("simulate_text produces one output: TKSM_SEQ_BADREAD or TKSM_SEQ_PERFECT");
    tksm_seq_batch b;
    MdfParser parser(&h->genome);
    b.hb.ordinal_base = ordinal_base;
    parser.feed(mdf, mdf_len, true, b.hb);
    tksm_seq_stats s{};
    batch_run(h, &b, which, format, &s);
    int o = which == TKSM_SEQ_PERFECT ? 1 : 0;
    *out_len = b.out_bytes[o];
    *out = (char*)malloc(std::max<size_t>(*out_len, 1));
    if (!*out) throw std::runtime_error("out of host memory");
    // pinned staging keeps the device->host copy at full PCIe rate
    h->stage_out.ensure(std::max<size_t>(*out_len, 1));
    batch_fetch(h, &b, o, (char*)h->stage_out.p);
    memcpy(*out, h->stage_out.p, *out_len);
    s.d2h_bytes += *out_len;
    if (stats) *stats = s;
    API_END(h)
}

int tksm_seq_run_files(tksm_seq* h, const char* mdf_path, const char* badread_path, const char* perfect_path,
                       tksm_seq_stats* stats) {
    API_BEGIN
    if (!badread_path && !perfect_path) throw std::runtime_error("Must specify either --output or --perfect.");
    tk::run_files(h, mdf_path, badread_path, perfect_path, stats);
    API_END(h)
}

}  // extern "C"

// ---------------------------------------------------------------- engine.hpp glue
namespace tk {

void engine_run_batch(tksm_seq* h, HostBatch& hb, int which, int format, std::vector<char>* out_bad,
                      std::vector<char>* out_perf, tksm_seq_stats* st) {
    tksm_seq_batch b;
    b.hb = std::move(hb);
    batch_run(h, &b, which, format, st);
    if (out_bad && (which & TKSM_SEQ_BADREAD)) {
        out_bad->resize(b.out_bytes[0]);
        batch_fetch(h, &b, 0, out_bad->data());
        if (st) st->d2h_bytes += b.out_bytes[0];
    }
    if (out_perf && (which & TKSM_SEQ_PERFECT)) {
        out_perf->resize(b.out_bytes[1]);
        batch_fetch(h, &b, 1, out_perf->data());
        if (st) st->d2h_bytes += b.out_bytes[1];
    }
    hb = std::move(b.hb);
}

const GenomeStore* engine_genome(const tksm_seq* h) { return &h->genome; }
const tksm_seq_config* engine_config(const tksm_seq* h) { return &h->cfg; }

}  // namespace tk
