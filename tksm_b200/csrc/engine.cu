// Device engine + C ABI of the Seq hot path (include/tksm_seq.h).
#include <cuda_runtime.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <condition_variable>
#include <deque>
#include <functional>
#include <future>
#include <mutex>
#include <thread>
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
#include "tail.hpp"
#include "extras.hpp"
#include "kernels.cuh"
#include "mdf_device.cuh"
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
        if (want > ((size_t)1 << 30) && getenv("TKSM_SEQ_TRACE")) fprintf(stderr, "[trace] device buffer of %zu MiB\n", want >> 20);
        const cudaError_t e = cudaMalloc(&p, want);
        if (e != cudaSuccess) {
            p = nullptr;
            size_t fr = 0, tot = 0;
            cudaMemGetInfo(&fr, &tot);
            throw std::runtime_error(std::string("CUDA error: ") + cudaGetErrorString(e) + " allocating " + std::to_string(want >> 20) +
                                     " MiB of device memory (" + std::to_string(fr >> 20) + " of " + std::to_string(tot >> 20) + " MiB free)");
        }
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

constexpr size_t SMEM_TOTAL = 232448;  // 227 KB opt-in maximum per CTA on sm_100 (smaller carve-outs for more L1 measured slower: 195 KB +5 %, 163 KB +19 %)
constexpr int N_CLASSES = 6;           // teams of 1, 2, 4, 8, 24 warps in shared memory; 24 warps with a global workspace
constexpr int GWS_CLASS = N_CLASSES - 1;
constexpr int GWS_W2 = 8;    // warps per read of the FIN_FULL kernel of the global-workspace class: the chunks of its alignment keep
                             // a few hundred threads busy at most, so several small CTAs per SM (GWS_FULL_CTAS) beat one of 32 warps
constexpr int GWS_FULL_CTAS = 4;   // (length-skew shape: 8 warps x 4 CTAs 171 ms, 4 x 8 189 ms, 16 x 2 205 ms, 32 x 1 236 ms)
// CTAs of the FIN_FULL launch on a class's redo list: a few reads as a rule, but in the class of 8 warps every read above
// ALIGN_WIDE_FROM slots
inline int redo_grid_cap(int c, int num_sms) { return c >= 3 ? num_sms : 32; }
const int CLASS_W[N_CLASSES] = {1, 2, 4, 8, 24, 24};
inline uint32_t list_capacity(uint32_t ncap) { return ncap / 5 * 2 + 64; }

}  // namespace

// Page-locked upload buffers shared by the host threads that pack batches and the thread that
// issues the uploads: a packer blocks in acquire() while all buffers are in flight.
class StagePool {
  public:
    StagePool(int max_buffers, bool pinned) : max_(max_buffers), pinned_(pinned) {}
    ~StagePool() {
        for (auto& b : bufs_) {
            if (pinned_) b.mem.release();
            else free(b.mem.p);
        }
    }
    // returns the index of a free buffer of at least `bytes`, or -1 when all buffers are in flight
    // (never blocks: one packer may hold many batches before the issuing thread sees the first).
    // Best fit first, so that a small batch does not take the buffer the next large one needs; when nothing fits, a
    // new buffer (else the smallest free one) is sized to the largest request seen so far: page-locking a buffer
    // takes 10-20 ms and freeing one waits for the device, so the pool has to stop growing after a call or two
    // (with "largest free buffer, grown to the request" a small buffer was re-pinned in one call out of five).
    int acquire(size_t bytes, int device) {
        std::unique_lock<std::mutex> g(mu_);
        high_water_ = std::max(high_water_, bytes);
        int idx = -1, smallest = -1;
        for (size_t i = 0; i < bufs_.size(); i++) {
            if (bufs_[i].used) continue;
            const size_t cap = bufs_[i].mem.cap;
            if (cap >= bytes && (idx < 0 || cap < bufs_[(size_t)idx].mem.cap)) idx = (int)i;
            if (smallest < 0 || cap < bufs_[(size_t)smallest].mem.cap) smallest = (int)i;
        }
        if (idx < 0 && (int)bufs_.size() < max_) {
            bufs_.emplace_back();
            idx = (int)bufs_.size() - 1;
        }
        if (idx < 0) idx = smallest;
        if (idx < 0) return -1;
        const size_t want = high_water_;
        bufs_[(size_t)idx].used = true;
        Buf& mine = bufs_[(size_t)idx];
        g.unlock();
        try {  // only this thread touches the buffer while it is `used`
            if (mine.mem.cap >= bytes) return idx;
            if (pinned_) {
                CK(cudaSetDevice(device));
                mine.mem.ensure(want);
            } else {  // host_only handles: no CUDA call
                free(mine.mem.p);
                mine.mem.cap = 0;
                mine.mem.p = malloc(want + want / 8 + 256);
                if (!mine.mem.p) throw std::runtime_error("out of host memory");
                mine.mem.cap = want + want / 8 + 256;
            }
        } catch (...) {
            release(idx);
            throw;
        }
        return idx;
    }
    char* ptr(int idx) {
        std::lock_guard<std::mutex> g(mu_);
        return (char*)bufs_[(size_t)idx].mem.p;
    }
    void release(int idx) {
        if (idx < 0) return;
        std::lock_guard<std::mutex> g(mu_);
        bufs_[(size_t)idx].used = false;
    }

  private:
    struct Buf {
        HBuf mem;
        bool used = false;
    };
    std::mutex mu_;
    std::deque<Buf> bufs_;   // deque: growing never moves the buffers other threads are filling
    size_t high_water_ = 0;
    int max_;
    bool pinned_;
};

// move-only ownership of one StagePool buffer
struct StageLease {
    StagePool* pool = nullptr;
    int idx = -1;
    StageLease() = default;
    StageLease(StagePool* p, int i) : pool(p), idx(i) {}
    StageLease(const StageLease&) = delete;
    StageLease& operator=(const StageLease&) = delete;
    StageLease(StageLease&& o) noexcept : pool(o.pool), idx(o.idx) { o.idx = -1; }
    StageLease& operator=(StageLease&& o) noexcept {
        if (this != &o) {
            reset();
            pool = o.pool;
            idx = o.idx;
            o.idx = -1;
        }
        return *this;
    }
    ~StageLease() { reset(); }
    void reset() {
        if (pool && idx >= 0) pool->release(idx);
        idx = -1;
    }
};

// One parsed MDF chunk with everything the device needs, built by host threads only
// (no CUDA call): scratch layout, length-class work lists and the upload blob.
struct Prepared {
    uint64_t n_reads = 0, n_mols = 0;
    uint64_t ordinal_base = 0;
    uint64_t total_raw_bases = 0;
    uint64_t scr_total = 0, raw_total = 0;   // bytes of the per-read scratch arenas
    uint32_t n_work[N_CLASSES] = {};
    uint32_t max_n_gws = 0;
    // the same for the second pass of a batch in which a read outgrew its slot (slot_launch: grow): classes chosen
    // by grow_len(n), so that a read may reach MAX_SLOT_LEN bases per slot
    uint32_t n_work_grow[N_CLASSES] = {};
    uint32_t max_n_gws_grow = 0;
    uint64_t alg_in_bytes = 0, idsum = 0;
    bool mods = false;
    StageLease stage;         // page-locked buffer of the handle's StagePool holding the blob,
    std::unique_ptr<char[]> heap_blob;  // or pageable memory when the pool is exhausted
    char* blob = nullptr;
    size_t blob_bytes = 0;
    size_t o_read_mol = 0, o_msb = 0, o_mrl = 0, o_mib = 0, o_src = 0, o_len = 0, o_smb = 0, o_mp = 0, o_mc = 0, o_lit = 0,
           o_id = 0, o_so = 0, o_ro = 0;
    size_t o_tail = SIZE_MAX;  // per-molecule tail lengths (tail-noise model loaded), else absent
    // device MDF tokeniser (mdf_device.cuh): `blob` is the TEXT of the batch (complete records, newline at the end);
    // the arrays are built on the device and the counts above are filled in by slot_bind_text
    bool dev_text = false;
    uint32_t text_lines = 0, text_records = 0;
    // tail-noise model loaded: the tails depend on the global read ordinal, which is known when the batch is
    // issued; the molecule batches are kept until then and the blob is built by finish_deferred()
    std::vector<const HostBatch*> deferred_parts;
    std::vector<std::shared_ptr<std::vector<std::unique_ptr<HostBatch>>>> deferred_keep;
};

// A device slot: the buffers of one chunk in flight.  The same object backs the device-resident
// batch of the C ABI (tksm_seq_batch_create / _run / _fetch).
struct tksm_seq_batch {
    Prepared prep;
    DBuf blob;
    DevBatch db{};
    uint32_t* d_work[N_CLASSES] = {};
    DBuf gws4;            // global workspaces of the FIN_FULL kernel of the class of 24 warps
    DBuf text, mdf_ws;    // device MDF tokeniser: the text of the batch and its per-line / per-record work arrays
    HBuf h_mdf;           // pinned: MdfTotals
    DBuf scr_seq, scr_qual, scr_raw, per_read, out[2], rec_len[2], rec_off[2], sums, gws, counters, work_keys, work_sorted, work_bins, mid, trace, redo, redo64, scr_ops, scr_cig, rinfo, atrace, scr_off_grow;
    HBuf stage;           // pinned upload staging
    HBuf host_out[2];     // pinned download staging (streaming to files)
    HBuf h_tot;           // pinned: record bytes per output, read bases, status
    cudaEvent_t ev[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};  // kernels begin, simulate end, kernels end, downloads end, totals on host
    cudaEvent_t ev_sorted = nullptr, ev_class[4] = {nullptr, nullptr, nullptr, nullptr};  // work order ready; class stream done
    cudaEvent_t ev_fork[4] = {nullptr, nullptr, nullptr, nullptr}, ev_side[4] = {nullptr, nullptr, nullptr, nullptr};  // side streams
    cudaEvent_t ev_early[4] = {nullptr, nullptr, nullptr, nullptr};   // expansion of a class done (its side stream starts on the long reads)
    uint64_t out_bytes[2] = {0, 0};
    uint64_t out_cap[2] = {0, 0};
    uint32_t launches = 0;
    int which = 0, format = 0;
    bool grown = false;       // the kernels ran in the larger slots of slot_launch(grow)
    int lane = 0;             // which pair of streams the batch runs on
    bool bound = false;
    bool d2h_pending = false;
    size_t h2d_bytes = 0;
    void ensure_events() {
        for (auto& e : ev)
            if (!e) CK(cudaEventCreate(&e));
        if (!ev_sorted) CK(cudaEventCreateWithFlags(&ev_sorted, cudaEventDisableTiming));
        for (auto& e : ev_class)
            if (!e) CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        for (auto& e : ev_fork)
            if (!e) CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        for (auto& e : ev_early)
            if (!e) CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        for (auto& e : ev_side)
            if (!e) CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    }
    ~tksm_seq_batch() {
        for (DBuf* d : {&blob, &scr_seq, &scr_qual, &scr_raw, &per_read, &out[0], &out[1], &rec_len[0], &rec_len[1],
                        &rec_off[0], &rec_off[1], &sums, &gws, &counters, &work_keys, &work_sorted, &work_bins, &mid, &trace, &redo,
                        &redo64, &scr_ops, &scr_cig, &rinfo, &atrace, &scr_off_grow, &text, &mdf_ws, &gws4})
            d->release();
        stage.release();
        h_mdf.release();
        host_out[0].release();
        host_out[1].release();
        h_tot.release();
        for (auto& e : ev)
            if (e) cudaEventDestroy(e);
        if (ev_sorted) cudaEventDestroy(ev_sorted);
        for (auto& e : ev_class)
            if (e) cudaEventDestroy(e);
        for (auto& e : ev_fork)
            if (e) cudaEventDestroy(e);
        for (auto& e : ev_early)
            if (e) cudaEventDestroy(e);
        for (auto& e : ev_side)
            if (e) cudaEventDestroy(e);
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
    TailModel tail;           // loaded == false: no_noise
    ExtrasConfig extras;      // Badread extras (all off by default)
    FrontEnd front;           // fused polyA / flip / tag / scb steps (none by default)
    bool per_read_molecules() const { return tail.loaded || extras.any() || front.any(); }   // the blob is built at issue time, one molecule per read
    IdentityHost idh;
    std::vector<float> delta;
    float calib_overshoot = 0, calib_kappa = 0, calib_kappa_measured = NAN;
    bool models_loaded = false;
    // device
    cudaStream_t stream = nullptr;       // uploads + kernels
    cudaStream_t copy_stream = nullptr;  // record downloads, overlapped with the next chunk's kernels
    // Every device slot of the streaming pipeline has its own lane (upload/kernel stream, class streams, side
    // streams): the kernel sequence of a batch has long single-warp chains (one thread per read in the alignment
    // kernels), which the kernels of the neighbouring batches fill.
    cudaStream_t lane_stream[3] = {nullptr, nullptr, nullptr};   // lane 0 is `stream`
    cudaStream_t class_stream[3][4] = {};  // simulate kernels of one batch, one stream per length class
    cudaStream_t side_stream[3][4] = {};   // the few reads of a class that need the 64-bit alignment band or a redo
    std::vector<std::unique_ptr<tksm_seq_batch>> slots;  // streaming pipeline (engine_stream)
    std::unique_ptr<StagePool> stage_pool;
    int num_sms = 148;
    DBuf d_genome, d_exc_start, d_exc_len, d_exc_chr;
    DBuf d_ct_hash, d_ct_contig, d_ct_nameoff, d_ct_names, d_ct_offset, d_ct_length;   // contig names for the device MDF tokeniser
    ContigTable ct{};
    DBuf d_qcls, d_thr, d_alias, d_alts, d_ident, d_delta, d_qdirect, d_qhash, d_qrows, d_qalias, d_t9, d_t9d1, d_t9d2;
    DevTables dt{};
    uint32_t ncap[N_CLASSES] = {};
    uint32_t cap_list[N_CLASSES] = {};
    uint32_t slab[N_CLASSES] = {};
    size_t smem_bytes[N_CLASSES] = {};
    // PHASE 2 kernels (split mode): lean team workspaces, 32 warps per CTA where they fit
    uint32_t slab2[N_CLASSES] = {}, threads2[N_CLASSES] = {};
    size_t smem_bytes2[N_CLASSES] = {};
    std::vector<uint16_t> thr;
    bool attrs_set = false;

    ~tksm_seq() {
        for (DBuf* d : {&d_ct_hash, &d_ct_contig, &d_ct_nameoff, &d_ct_names, &d_ct_offset, &d_ct_length}) d->release();
        for (DBuf* d : {&d_genome, &d_exc_start, &d_exc_len, &d_exc_chr, &d_qcls, &d_thr, &d_alias, &d_alts, &d_ident, &d_delta,
                        &d_qdirect, &d_qhash, &d_qrows, &d_qalias, &d_t9, &d_t9d1, &d_t9d2})
            d->release();
        slots.clear();
        if (copy_stream) cudaStreamDestroy(copy_stream);
        for (auto& lane : class_stream)
            for (auto& cs : lane)
                if (cs) cudaStreamDestroy(cs);
        for (auto& lane : side_stream)
            for (auto& cs : lane)
                if (cs) cudaStreamDestroy(cs);
        if (lane_stream[1]) cudaStreamDestroy(lane_stream[1]);
        if (lane_stream[2]) cudaStreamDestroy(lane_stream[2]);
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
    h->genome.index_exceptions();
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
    // the reference's names for the device MDF tokeniser (mdf_device.cuh): open addressing on a hash of the name
    {
        const auto& C = h->genome.contigs;
        uint32_t slots = 16;
        while (slots < 2 * C.size() + 2) slots <<= 1;
        std::vector<uint64_t> sh(slots, 0), off, len;
        std::vector<uint32_t> sc(slots, 0), noff{0};
        std::vector<uint8_t> names;
        for (const auto& kv : C) {
            const uint32_t c = (uint32_t)off.size();
            names.insert(names.end(), kv.first.begin(), kv.first.end());
            noff.push_back((uint32_t)names.size());
            off.push_back(kv.second.offset);
            len.push_back(kv.second.length);
            const uint64_t hh = mdf_name_hash((const uint8_t*)kv.first.data(), (uint32_t)kv.first.size());
            uint32_t x = (uint32_t)(hh >> 17) & (slots - 1);
            while (sh[x]) x = (x + 1) & (slots - 1);
            sh[x] = hh;
            sc[x] = c;
        }
        if (names.empty()) names.push_back(0);
        if (off.empty()) { off.push_back(0); len.push_back(0); }
        upload_vec(h->d_ct_hash, sh, h->stream);
        upload_vec(h->d_ct_contig, sc, h->stream);
        upload_vec(h->d_ct_nameoff, noff, h->stream);
        upload_vec(h->d_ct_names, names, h->stream);
        upload_vec(h->d_ct_offset, off, h->stream);
        upload_vec(h->d_ct_length, len, h->stream);
        CK(cudaStreamSynchronize(h->stream));
        h->ct = ContigTable{h->d_ct_hash.as<uint64_t>(), h->d_ct_contig.as<uint32_t>(), slots - 1, h->d_ct_nameoff.as<uint32_t>(),
                            h->d_ct_names.as<uint8_t>(), h->d_ct_offset.as<uint64_t>(), h->d_ct_length.as<uint64_t>(),
                            h->genome.name_len_mask};
    }
    h->genome_dirty = false;
}

void compute_classes(tksm_seq* h) {
    const size_t nq = h->em.has_table ? ((size_t)1 << (2 * h->em.k)) : 0;
    const size_t qcls_bytes = (nq + 15) & ~size_t(15);
    const size_t pool = SMEM_TOTAL - qcls_bytes - SIM_FIXED_SMEM - 128;
    for (int c = 0; c < GWS_CLASS; c++) {
        const int W = CLASS_W[c], teams = SIM_WARPS / W;
        const uint32_t hist_elem = W >= 8 ? 8 : 4;
        const size_t budget = (pool / teams) & ~size_t(15);
        uint32_t ncap = 16;
        while (slab_layout(W, hist_elem, ncap + 16, list_capacity(ncap + 16)).bytes <= budget) ncap += 16;
        h->ncap[c] = ncap;
        h->cap_list[c] = list_capacity(ncap);
        h->slab[c] = slab_layout(W, hist_elem, ncap, h->cap_list[c]).bytes;
        h->smem_bytes[c] = qcls_bytes + SIM_FIXED_SMEM + (size_t)h->slab[c] * teams + 128;
        // FIN_FULL kernels (expansion, team-parallel alignment, qscores: long reads and the redo lists): fragment, slot
        // lengths and read in shared memory; no k-mer classes, threshold row or histogram, so 32 warps per CTA where the
        // workspaces fit.  FIN_EXPAND / FIN_QUAL kernels: only the small workspace part.
        h->slab2[c] = slab2_small_bytes(W) + slab2_big_bytes(ncap);
        h->threads2[c] = SIM_THREADS;
        if (32 % W == 0 && SIM_FIXED_SMEM + (size_t)h->slab2[c] * (32 / W) + 128 + 64 <= SMEM_TOTAL) h->threads2[c] = 1024;
        h->smem_bytes2[c] = SIM_FIXED_SMEM + (size_t)h->slab2[c] * (h->threads2[c] / 32 / W) + 128 + 64;
    }
    // the classes of 24 warps take the 64-bit alignment band at once, which is the rule for fragments above
    // ALIGN_WIDE_FROM slots: the class below them must reach that far
    if (h->ncap[3] < ALIGN_WIDE_FROM) throw std::runtime_error("length class of 8 warps ends below ALIGN_WIDE_FROM");
    h->smem_bytes[GWS_CLASS] = qcls_bytes + SIM_FIXED_SMEM + gws_small_bytes(CLASS_W[GWS_CLASS]) + 128;
    h->threads2[GWS_CLASS] = GWS_W2 * 32;
    h->smem_bytes2[GWS_CLASS] = SIM_FIXED_SMEM + slab2_small_bytes(GWS_W2) + 128 + 64;
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
    upload_vec(h->d_t9d2, h->qm.t9d2, h->stream);
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
    T.q_t9d2 = h->qm.t9d2.empty() ? nullptr : h->d_t9d2.as<uint32_t>();
    T.seed_lo = (uint32_t)h->cfg.seed;
    T.seed_hi = (uint32_t)(h->cfg.seed >> 32);
}

template <int W, bool GWS, int PHASE>
void launch_sim_phase(cudaStream_t stream, const SimParams& P, size_t smem, int grid, int threads = SIM_THREADS) {
    auto kern = simulate_kernel<W, GWS, PHASE>;
    CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<grid, threads, smem, stream>>>(P);
    CK(cudaGetLastError());
}

// error injection and expansion + alignment + qscores as two kernels per class with their own ticket
// counters (P.counter, P.counter + 16)
struct Phase2Launch {
    size_t smem;
    int grid, threads;
    uint32_t slab_bytes;
    uint8_t* gws = nullptr;   // workspace region of the FIN_FULL kernel when it is a global-workspace instance of its own
};
// W2: warps per read of the PHASE 2 kernel (the global-workspace class is not bound by a shared-memory slab and
// puts all 32 warps of a CTA on its read)
struct AlignLaunch {
    AlignKParams K;   // 32-bit instance: its give-up list is (redo64, redo64_count)
    int grid;
    uint32_t* redo64;         // give-up list of the 32-bit instance = work list of the 64-bit one
    uint32_t* redo64_count;
    uint8_t* trace64;         // trace cells of the 64-bit instance
    uint64_t trace64_warp_bytes;
    bool wide_class;          // every read of the class is above ALIGN_WIDE64_FROM slots: 64-bit instance on all of them
    int grid64;
    cudaStream_t side;        // stream of the 64-bit instance and the redo list
    cudaEvent_t ev_fork, ev_side, ev_early;
    int redo_grid;            // CTAs of the FIN_FULL launch on the redo list
    bool early64;             // the class holds reads above ALIGN_WIDE64_FROM slots: the 64-bit instance starts on them beside the 32-bit one
};
// GWS2: the FIN_FULL kernel works on a workspace in global memory (the class of 24 warps does, with teams of GWS_W2
// warps, although its error-injection kernel has its read in shared memory)
template <int W, bool GWS, int W2 = W, bool GWS2 = GWS>
void launch_sim(cudaStream_t stream, SimParams P, size_t smem, int grid, const Phase2Launch& p2, const AlignLaunch& al) {
    launch_sim_phase<W, GWS, 1>(stream, P, smem, grid);
    if (!(P.flags & RUN_ERRORS)) return;
    uint32_t* const counters = P.counter;
    if (W >= 24) {
        // every read of these classes is above ALIGN_WIDE_FROM slots: team-parallel alignment with the 64-bit band
        P.slab_bytes = p2.slab_bytes;
        P.counter = counters + 16;
        P.n_work_dev = nullptr;
        if (p2.gws) P.gws = p2.gws;
        launch_sim_phase<W2, GWS2, FIN_FULL>(stream, P, p2.smem, p2.grid, p2.threads);
        return;
    }
    // expansion | one thread per read: alignment | qscores; then the redo list of the class
    constexpr int TEAMS = 32 / W;
    const size_t smem_small = SIM_FIXED_SMEM + (size_t)slab2_small_bytes(W) * TEAMS + 128 + 64;
    const int grid_small = (int)std::min<size_t>((size_t)grid * 2, (P.n_work + TEAMS - 1) / TEAMS);
    SimParams Q = P;
    Q.slab_bytes = slab2_small_bytes(W);
    Q.counter = counters + 16;
    launch_sim_phase<W, GWS, FIN_EXPAND>(stream, Q, smem_small, grid_small, 1024);
    if (!(P.flags & RUN_QUAL)) return;
    constexpr int AW = ALIGN_K_THREADS / 32;
    // (per device: a process may drive several GPUs)
    cudaStream_t side = al.side;
    if (al.wide_class) {
        // 64-bit band at once, on every read of the class; what leaves it goes to the FIN_FULL redo list
        AlignKParams K2 = al.K;
        K2.trace = al.trace64;
        K2.trace_warp_bytes = al.trace64_warp_bytes;
        K2.redo = P.redo;
        K2.redo_count = P.redo_count;
        align_kernel<uint64_t><<<al.grid64, ALIGN_K_THREADS, 0, stream>>>(K2);
        CK(cudaGetLastError());
        CK(cudaEventRecord(al.ev_fork, stream));
        Q.counter = counters + 24;
        launch_sim_phase<W, GWS, FIN_QUAL>(stream, Q, smem_small, grid_small, 1024);
        CK(cudaStreamWaitEvent(side, al.ev_fork, 0));
    } else {
        // side stream, at once: the reads above ALIGN_WIDE64_FROM slots (the expansion listed them; `split` = how many)
        uint32_t* const split = counters + 104;
        if (al.early64) {
            CK(cudaMemcpyAsync(split, al.redo64_count, 4, cudaMemcpyDeviceToDevice, stream));
            CK(cudaEventRecord(al.ev_early, stream));
            CK(cudaStreamWaitEvent(side, al.ev_early, 0));
            AlignKParams K1 = al.K;
            K1.work = al.redo64;
            K1.n_work_dev = split;
            K1.counter = counters + 96;
            K1.trace = al.trace64;
            K1.trace_warp_bytes = al.trace64_warp_bytes;
            K1.redo = P.redo;
            K1.redo_count = P.redo_count;
            align_kernel<uint64_t><<<al.grid64, ALIGN_K_THREADS, 0, side>>>(K1);
            CK(cudaGetLastError());
        }
        align_kernel<uint32_t><<<al.grid, ALIGN_K_THREADS, 0, stream>>>(al.K);
        CK(cudaGetLastError());
        CK(cudaEventRecord(al.ev_fork, stream));
        // main stream: qscores of the reads the 32-bit band did
        Q.counter = counters + 24;
        launch_sim_phase<W, GWS, FIN_QUAL>(stream, Q, smem_small, grid_small, 1024);
        // side stream: the reads that left the 32-bit band (about one in a thousand) and those above ALIGN_WIDE64_FROM
        // slots take the 64-bit instance and their qscores.  Long chains on few reads: beside the main stream they
        // cost nothing.
        CK(cudaStreamWaitEvent(side, al.ev_fork, 0));
        AlignKParams K2 = al.K;
        K2.work = al.redo64;
        K2.n_work_dev = al.redo64_count;
        K2.work_begin_dev = al.early64 ? split : nullptr;   // (zero without the early launch: the counters are cleared per batch)
        K2.counter = al.K.counter + 8;
        K2.trace = al.trace64;
        K2.trace_warp_bytes = al.trace64_warp_bytes;
        K2.redo = P.redo;
        K2.redo_count = P.redo_count;
        align_kernel<uint64_t><<<al.grid64, ALIGN_K_THREADS, 0, side>>>(K2);
        CK(cudaGetLastError());
        SimParams Q2 = Q;
        Q2.counter = counters + 88;
        Q2.work = al.redo64;
        Q2.n_work_dev = al.redo64_count;
        launch_sim_phase<W, GWS, FIN_QUAL>(side, Q2, smem_small, std::min(grid_small, 32), 1024);
    }
    // what left the 64-bit band too, and the reads above ALIGN_WIDE_FROM slots: the FIN_FULL kernel
    P.slab_bytes = p2.slab_bytes;
    P.counter = counters + 56;
    P.work = P.redo;
    P.n_work_dev = P.redo_count;
    if (p2.gws) P.gws = p2.gws;
    launch_sim_phase<W2, GWS2, FIN_FULL>(side, P, p2.smem, std::min(p2.grid, al.redo_grid), p2.threads);
    CK(cudaEventRecord(al.ev_side, side));
    CK(cudaStreamWaitEvent(stream, al.ev_side, 0));
}

inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }
std::atomic<uint64_t> g_ns_prep_alloc{0}, g_ns_prep_reads{0}, g_ns_prep_copy{0};
inline uint64_t now_ns() { return (uint64_t)std::chrono::duration_cast<std::chrono::nanoseconds>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

// Host-only preparation of one device batch from the molecule batches of consecutive text chunks:
// per-read scratch layout, length classes and the upload blob, written in one pass (the parts are
// concatenated straight into the blob with their indices rebased).  Thread-safe with respect to
// the handle (reads the class capacities and k only).
void prepare_host(const tksm_seq* h, const std::vector<const HostBatch*>& parts, Prepared& p) {
    const uint32_t k = (uint32_t)h->em.k;
    size_t R = 0, M = 0, S = 0, NM = 0, LIT = 0, ID = 0;
    uint64_t raw = 0, gen = 0, litb = 0;
    for (const HostBatch* hb : parts) {
        R += hb->n_reads();
        M += hb->n_mols();
        S += hb->seg_src.size();
        NM += hb->mod_pos.size();
        LIT += hb->lit_pool.size();
        ID += hb->id_pool.size();
        raw += hb->total_raw_bases;
        gen += hb->genomic_bases;
        litb += hb->literal_bases;
    }
    if (R > 0xFFFFFFF0ull || S > 0xFFFFFFF0ull || ID > 0xFFFFFFF0ull) throw std::runtime_error("device batch too large");
    p.n_reads = R;
    p.n_mols = M;
    p.total_raw_bases = raw;
    p.mods = NM > 0;
    size_t total = 0;
    auto add = [&](size_t bytes) {
        size_t off = total;
        total = align_up(total + std::max<size_t>(bytes, 1), 256);
        return off;
    };
    p.o_read_mol = add(R * 4);
    p.o_msb = add((M + 1) * 4);
    p.o_mrl = add(M * 4);
    p.o_mib = add((M + 1) * 4);
    p.o_src = add(S * 8);
    p.o_len = add(S * 4);
    p.o_smb = add(p.mods ? (S + 1) * 4 : 0);
    p.o_mp = add(NM * 4);
    p.o_mc = add(NM);
    p.o_lit = add(LIT);
    p.o_id = add(ID);
    p.o_so = add((R + 1) * 8);
    p.o_ro = add((R + 1) * 8);
    bool tails = false;
    for (const HostBatch* hb : parts) tails = tails || !hb->mol_tail_len.empty();
    p.o_tail = tails ? add(M * 4) : SIZE_MAX;
    const uint64_t t_a0 = now_ns();
    p.stage = StageLease(h->stage_pool.get(), h->stage_pool->acquire(total, h->cfg.device));
    if (p.stage.idx >= 0) {
        p.blob = h->stage_pool->ptr(p.stage.idx);
    } else {
        p.heap_blob.reset(new char[total]);
        p.blob = p.heap_blob.get();
    }
    p.blob_bytes = total;
    g_ns_prep_alloc += now_ns() - t_a0;
    char* B = p.blob;
    uint32_t* read_mol = (uint32_t*)(B + p.o_read_mol);
    uint32_t* msb = (uint32_t*)(B + p.o_msb);
    uint32_t* mrl = (uint32_t*)(B + p.o_mrl);
    uint32_t* mib = (uint32_t*)(B + p.o_mib);
    uint64_t* src = (uint64_t*)(B + p.o_src);
    uint32_t* slen = (uint32_t*)(B + p.o_len);
    uint32_t* smb = (uint32_t*)(B + p.o_smb);
    uint32_t* mp = (uint32_t*)(B + p.o_mp);
    uint8_t* mc = (uint8_t*)(B + p.o_mc);
    uint8_t* lit = (uint8_t*)(B + p.o_lit);
    uint8_t* idp = (uint8_t*)(B + p.o_id);
    uint64_t* scr_off = (uint64_t*)(B + p.o_so);
    uint64_t* raw_off = (uint64_t*)(B + p.o_ro);

    for (int c = 0; c < N_CLASSES; c++) p.n_work[c] = p.n_work_grow[c] = 0;
    uint64_t so = 0, ro = 0, per_read_segs = 0, per_read_mods = 0, idsum = 0;
    p.max_n_gws = p.max_n_gws_grow = 0;
    size_t r0 = 0, m0 = 0, s0 = 0, nm0 = 0, lit0 = 0, id0 = 0;
    msb[0] = 0;
    mib[0] = 0;
    if (p.mods) smb[0] = 0;
    for (const HostBatch* hbp : parts) {
        const HostBatch& hb = *hbp;
        const size_t r_n = hb.n_reads(), m_n = hb.n_mols(), s_n = hb.seg_src.size(), nm_n = hb.mod_pos.size();
        const uint64_t t_r0 = now_ns();
        for (size_t r = 0; r < r_n; r++) {
            const uint32_t m = hb.read_mol[r];
            const uint32_t L = hb.mol_raw_len[m];
            read_mol[r0 + r] = (uint32_t)(m0 + m);
            scr_off[r0 + r] = so;
            raw_off[r0 + r] = ro;
            so += align_up((size_t)2 * L + 32, 16);
            ro += align_up((size_t)L, 16);
            const uint32_t n = L + 2 * k;
            int c = 0;
            while (c < GWS_CLASS && n > h->ncap[c]) c++;
            p.n_work[c]++;  // the device sorts the reads of each class (work_key / scan / scatter kernels)
            if (c == GWS_CLASS) p.max_n_gws = std::max(p.max_n_gws, n);
            const uint32_t ng = grow_len(n);
            while (c < GWS_CLASS && ng > h->ncap[c]) c++;
            p.n_work_grow[c]++;
            if (c == GWS_CLASS) p.max_n_gws_grow = std::max(p.max_n_gws_grow, ng);
            per_read_segs += hb.mol_seg_begin[m + 1] - hb.mol_seg_begin[m];
            if (nm_n) per_read_mods += hb.seg_mod_begin[hb.mol_seg_begin[m + 1]] - hb.seg_mod_begin[hb.mol_seg_begin[m]];
            idsum += hb.mol_id_begin[m + 1] - hb.mol_id_begin[m];
        }
        const uint64_t t_r1 = now_ns();
        g_ns_prep_reads += t_r1 - t_r0;
        for (size_t m = 0; m < m_n; m++) {
            msb[m0 + m + 1] = (uint32_t)(s0 + hb.mol_seg_begin[m + 1]);
            mib[m0 + m + 1] = (uint32_t)(id0 + hb.mol_id_begin[m + 1]);
        }
        if (m_n) memcpy(mrl + m0, hb.mol_raw_len.data(), m_n * 4);
        if (tails) {
            uint32_t* tl = (uint32_t*)(B + p.o_tail) + m0;
            for (size_t m = 0; m < m_n; m++) tl[m] = m < hb.mol_tail_len.size() ? hb.mol_tail_len[m] : 0u;
        }
        if (lit0 == 0) {
            if (s_n) memcpy(src + s0, hb.seg_src.data(), s_n * 8);
        } else {
            for (size_t i = 0; i < s_n; i++) {
                const uint64_t v = hb.seg_src[i];
                src[s0 + i] = (v & SEG_LITERAL) ? v + lit0 : v;
            }
        }
        if (s_n) memcpy(slen + s0, hb.seg_len.data(), s_n * 4);
        if (p.mods)
            for (size_t i = 0; i < s_n; i++) smb[s0 + i + 1] = (uint32_t)(nm0 + hb.seg_mod_begin[i + 1]);
        if (nm_n) {
            memcpy(mp + nm0, hb.mod_pos.data(), nm_n * 4);
            memcpy(mc + nm0, hb.mod_chr.data(), nm_n);
        }
        if (!hb.lit_pool.empty()) memcpy(lit + lit0, hb.lit_pool.data(), hb.lit_pool.size());
        if (!hb.id_pool.empty()) memcpy(idp + id0, hb.id_pool.data(), hb.id_pool.size());
        g_ns_prep_copy += now_ns() - t_r1;
        r0 += r_n;
        m0 += m_n;
        s0 += s_n;
        nm0 += nm_n;
        lit0 += hb.lit_pool.size();
        id0 += hb.id_pool.size();
    }
    scr_off[R] = so;
    raw_off[R] = ro;
    p.scr_total = so;
    p.raw_total = ro;
    p.idsum = idsum;
    // ceil(L_genomic / 4) + L_literal + 24 per interval + 5 per mod
    p.alg_in_bytes = (gen + 3) / 4 + litb + 24 * per_read_segs + 5 * per_read_mods;
}

// Uploads the prepared chunk of slot `b` (asynchronously on the compute stream).
void slot_bind(tksm_seq* h, tksm_seq_batch* b) {
    cudaStream_t S = h->lane_stream[b->lane];
    need_device(h);
    if (!h->models_loaded) throw std::runtime_error("models not loaded");
    commit_reference(h);
    CK(cudaSetDevice(h->cfg.device));
    Prepared& p = b->prep;
    const size_t R = p.n_reads;
    b->blob.ensure(p.blob_bytes);
    CK(cudaMemcpyAsync(b->blob.p, p.blob, p.blob_bytes, cudaMemcpyHostToDevice, S));
    b->h2d_bytes = p.blob_bytes;
    char* base = (char*)b->blob.p;
    DevBatch& D = b->db;
    D.n_reads = (uint32_t)R;
    D.ordinal_base = p.ordinal_base;
    D.read_mol = (const uint32_t*)(base + p.o_read_mol);
    D.mol_seg_begin = (const uint32_t*)(base + p.o_msb);
    D.mol_raw_len = (const uint32_t*)(base + p.o_mrl);
    D.mol_id_begin = (const uint32_t*)(base + p.o_mib);
    D.seg_src = (const uint64_t*)(base + p.o_src);
    D.seg_len = (const uint32_t*)(base + p.o_len);
    D.seg_mod_begin = p.mods ? (const uint32_t*)(base + p.o_smb) : nullptr;
    D.mod_pos = (const uint32_t*)(base + p.o_mp);
    D.mod_chr = (const uint8_t*)(base + p.o_mc);
    D.lit_pool = (const uint8_t*)(base + p.o_lit);
    D.id_pool = (const uint8_t*)(base + p.o_id);
    D.scr_off = (const uint64_t*)(base + p.o_so);
    D.raw_off = (const uint64_t*)(base + p.o_ro);
    b->per_read.ensure(std::max<size_t>(R, 1) * 16);
    D.read_len = b->per_read.as<uint32_t>();
    D.ident_num = D.read_len + R;
    D.ident_den = D.ident_num + R;
    D.status = D.ident_den + R;
    b->counters.ensure(512);
    b->h_tot.ensure(256);
    b->ensure_events();
    b->bound = true;
}

// exclusive prefix sums of n u32 values into out[0 .. n], out[n] = the total (scan kernels of the formatting stage)
void dev_scan(tksm_seq_batch* b, cudaStream_t S, const uint32_t* in, uint32_t n, uint64_t* out) {
    const uint32_t per = SCAN_BLOCK * SCAN_ITEMS;
    const uint32_t nb = (n + per - 1) / per;
    if (nb) scan_block_sums<<<nb, SCAN_BLOCK, 0, S>>>(in, n, b->sums.as<uint64_t>());
    scan_sums<<<1, 1024, 0, S>>>(b->sums.as<uint64_t>(), nb, out + n);
    if (nb) scan_final<<<nb, SCAN_BLOCK, 0, S>>>(in, n, b->sums.as<uint64_t>(), out);
}

// Device MDF tokeniser (mdf_device.cuh): uploads the text of the batch and builds the molecule batch from it on the
// device.  One synchronisation, for the counts the launch configuration needs.  Returns false when the text holds
// something the device passes do not interpret: the caller parses it on the host instead.
bool slot_bind_text(tksm_seq* h, tksm_seq_batch* b) {
    cudaStream_t S = h->lane_stream[b->lane];
    need_device(h);
    if (!h->models_loaded) throw std::runtime_error("models not loaded");
    commit_reference(h);
    CK(cudaSetDevice(h->cfg.device));
    Prepared& p = b->prep;
    const uint64_t nbytes = p.blob_bytes;
    const uint32_t NL = p.text_lines, NR = p.text_records;
    if (NL == 0 || nbytes == 0) return false;
    const uint32_t read_cap = 2 * NR + 4096;
    b->text.ensure(nbytes + 64);
    CK(cudaMemcpyAsync(b->text.p, p.blob, nbytes, cudaMemcpyHostToDevice, S));
    b->h2d_bytes = nbytes;
    // work arrays
    size_t total = 0;
    auto add = [&](size_t bytes) {
        const size_t off = total;
        total = align_up(total + std::max<size_t>(bytes, 8), 256);
        return off;
    };
    const uint32_t tiles = (uint32_t)((nbytes + MDF_TILE - 1) / MDF_TILE);
    const size_t L1 = (size_t)NL + 1, R1 = (size_t)NR + 1;
    const size_t w_tot = add(sizeof(MdfTotals)), w_tc = add((size_t)tiles * 4), w_to = add(((size_t)tiles + 1) * 8), w_ls = add(L1 * 8),
                 w_kind = add(NL), w_a = add((size_t)NL * 4), w_b = add((size_t)NL * 4), w_c = add((size_t)NL * 4), w_src = add((size_t)NL * 8),
                 w_ih = add((size_t)NL * 4), w_he = add(L1 * 8), w_ks = add((size_t)NL * 4), w_nm = add((size_t)NL * 4),
                 w_lk = add((size_t)NL * 4), w_ln = add((size_t)NL * 4), w_si = add(L1 * 8), w_mo = add(L1 * 8), w_lo = add(L1 * 8),
                 w_lp = add(L1 * 8), w_rl = add(R1 * 4), w_rk = add(R1 * 4), w_rd = add(R1 * 4), w_ri = add(R1 * 4), w_rs = add(R1 * 4),
                 w_rr = add(R1 * 4), w_mr = add(R1 * 8), w_rr8 = add(R1 * 8), w_ir = add(R1 * 8), w_so = add(R1 * 8), w_ro = add(R1 * 8);
    b->mdf_ws.ensure(total);
    char* W = (char*)b->mdf_ws.p;
    // outputs: the blob of the batch at bounds every input satisfies
    const size_t S_cap = NL - std::min(NL, NR), NM_cap = nbytes / 2 + 1;
    total = 0;
    p.o_read_mol = add((size_t)read_cap * 4);
    p.o_msb = add(R1 * 4);
    p.o_mrl = add(R1 * 4);
    p.o_mib = add(R1 * 4);
    p.o_src = add(S_cap * 8);
    p.o_len = add(S_cap * 4);
    p.o_smb = add((S_cap + 1) * 4);
    p.o_mp = add(NM_cap * 4);
    p.o_mc = add(NM_cap);
    p.o_lit = add(nbytes);
    p.o_id = add(nbytes);
    p.o_so = add(((size_t)read_cap + 1) * 8);
    p.o_ro = add(((size_t)read_cap + 1) * 8);
    p.o_tail = SIZE_MAX;
    b->blob.ensure(total);
    char* base = (char*)b->blob.p;
    b->sums.ensure(((size_t)std::max(tiles, NL) / (SCAN_BLOCK * SCAN_ITEMS) + 4) * 8);
    b->h_mdf.ensure(sizeof(MdfTotals));
    MdfParams P{};
    P.text = b->text.as<uint8_t>();
    P.n_bytes = nbytes;
    P.n_lines = NL;
    P.n_records = NR;
    P.read_cap = read_cap;
    P.ct = h->ct;
    P.exc_start = h->dt.exc_start;
    P.exc_len = h->dt.exc_len;
    P.n_exc = h->dt.n_exc;
    P.k = (uint32_t)h->em.k;
    for (int c = 0; c < 5; c++) P.ncap[c] = h->ncap[c];
    P.totals = (MdfTotals*)(W + w_tot);
    P.line_start = (uint64_t*)(W + w_ls);
    P.kind = (uint8_t*)(W + w_kind);
    P.a32 = (uint32_t*)(W + w_a);
    P.b32 = (uint32_t*)(W + w_b);
    P.c32 = (uint32_t*)(W + w_c);
    P.src = (uint64_t*)(W + w_src);
    P.is_hdr = (uint32_t*)(W + w_ih);
    P.hdr_excl = (uint64_t*)(W + w_he);
    P.keep_seg = (uint32_t*)(W + w_ks);
    P.nm_k = (uint32_t*)(W + w_nm);
    P.lit_k = (uint32_t*)(W + w_lk);
    P.len_k = (uint32_t*)(W + w_ln);
    P.seg_idx = (uint64_t*)(W + w_si);
    P.mod_off = (uint64_t*)(W + w_mo);
    P.lit_off = (uint64_t*)(W + w_lo);
    P.len_prefix = (uint64_t*)(W + w_lp);
    P.rec_line = (uint32_t*)(W + w_rl);
    P.rec_keep = (uint32_t*)(W + w_rk);
    P.rec_depth = (uint32_t*)(W + w_rd);
    P.rec_idlen = (uint32_t*)(W + w_ri);
    P.rec_slot = (uint32_t*)(W + w_rs);
    P.rec_rawslot = (uint32_t*)(W + w_rr);
    P.mol_of_rec = (uint64_t*)(W + w_mr);
    P.read_of_rec = (uint64_t*)(W + w_rr8);
    P.id_of_rec = (uint64_t*)(W + w_ir);
    P.so_of_rec = (uint64_t*)(W + w_so);
    P.ro_of_rec = (uint64_t*)(W + w_ro);
    P.read_mol = (uint32_t*)(base + p.o_read_mol);
    P.mol_seg_begin = (uint32_t*)(base + p.o_msb);
    P.mol_raw_len = (uint32_t*)(base + p.o_mrl);
    P.mol_id_begin = (uint32_t*)(base + p.o_mib);
    P.seg_src = (uint64_t*)(base + p.o_src);
    P.seg_len = (uint32_t*)(base + p.o_len);
    P.seg_mod_begin = (uint32_t*)(base + p.o_smb);
    P.mod_pos = (uint32_t*)(base + p.o_mp);
    P.mod_chr = (uint8_t*)(base + p.o_mc);
    P.lit_pool = (uint8_t*)(base + p.o_lit);
    P.id_pool = (uint8_t*)(base + p.o_id);
    P.scr_off = (uint64_t*)(base + p.o_so);
    P.raw_off = (uint64_t*)(base + p.o_ro);
    uint32_t* tile_count = (uint32_t*)(W + w_tc);
    uint64_t* tile_off = (uint64_t*)(W + w_to);
    CK(cudaMemsetAsync(P.totals, 0, sizeof(MdfTotals), S));
    const uint32_t gl = (NL + 255) / 256, gr = (NR + 255) / 256, gt = (tiles + 255) / 256;
    mdf_mark_lines<<<gt, 256, 0, S>>>(P.text, nbytes, tile_count);
    dev_scan(b, S, tile_count, tiles, tile_off);
    mdf_line_starts<<<gt, 256, 0, S>>>(P.text, nbytes, tile_off, P.line_start, NL);
    mdf_parse_lines<<<gl, 256, 0, S>>>(P);
    dev_scan(b, S, P.is_hdr, NL, P.hdr_excl);
    if (NR) mdf_header_records<<<gl, 256, 0, S>>>(P);
    dev_scan(b, S, P.rec_keep, NR, P.mol_of_rec);
    dev_scan(b, S, P.rec_depth, NR, P.read_of_rec);   // (records of depth 0 add nothing)
    dev_scan(b, S, P.rec_idlen, NR, P.id_of_rec);
    mdf_keep_lines<<<gl, 256, 0, S>>>(P);
    dev_scan(b, S, P.keep_seg, NL, P.seg_idx);
    dev_scan(b, S, P.nm_k, NL, P.mod_off);
    dev_scan(b, S, P.lit_k, NL, P.lit_off);
    dev_scan(b, S, P.len_k, NL, P.len_prefix);
    mdf_fill_segments<<<gl, 256, 0, S>>>(P);
    if (NR) mdf_record_slots<<<gr, 256, 0, S>>>(P);
    dev_scan(b, S, P.rec_slot, NR, P.so_of_rec);
    dev_scan(b, S, P.rec_rawslot, NR, P.ro_of_rec);
    if (NR) mdf_fill_records<<<gr, 256, 0, S>>>(P);
    mdf_finalize<<<1, 32, 0, S>>>(P);
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(b->h_mdf.p, P.totals, sizeof(MdfTotals), cudaMemcpyDeviceToHost, S));
    CK(cudaStreamSynchronize(S));
    const MdfTotals T = *(const MdfTotals*)b->h_mdf.p;
    // the header lines the host counted must be the ones the device saw (a '+' after a CR-only line end is not)
    if (T.bad) {
        if (getenv("TKSM_SEQ_TRACE")) fprintf(stderr, "[trace] device MDF tokeniser hands a batch of %u lines to the host parser (reason %u)\n", NL, T.bad);
        return false;
    }
    p.n_reads = T.reads;
    p.n_mols = T.mols;
    p.total_raw_bases = T.total_raw_bases;
    p.scr_total = T.scr_total;
    p.raw_total = T.raw_total;
    p.idsum = T.idsum;
    p.mods = T.mods > 0;
    p.alg_in_bytes = (T.gen_bases + 3) / 4 + T.lit_bases + 24 * T.per_read_segs + 5 * T.per_read_mods;
    for (int c = 0; c < N_CLASSES; c++) {
        p.n_work[c] = T.n_work[c];
        p.n_work_grow[c] = T.n_work_grow[c];
    }
    p.max_n_gws = T.max_n_gws;
    p.max_n_gws_grow = T.max_n_gws_grow;
    if (T.reads > 0xFFFFFFF0ull || T.segs > 0xFFFFFFF0ull || T.id_bytes > 0xFFFFFFF0ull) return false;
    const size_t R = p.n_reads;
    DevBatch& D = b->db;
    D.n_reads = (uint32_t)R;
    D.ordinal_base = p.ordinal_base;
    D.read_mol = P.read_mol;
    D.mol_seg_begin = P.mol_seg_begin;
    D.mol_raw_len = P.mol_raw_len;
    D.mol_id_begin = P.mol_id_begin;
    D.seg_src = P.seg_src;
    D.seg_len = P.seg_len;
    D.seg_mod_begin = p.mods ? P.seg_mod_begin : nullptr;
    D.mod_pos = P.mod_pos;
    D.mod_chr = P.mod_chr;
    D.lit_pool = P.lit_pool;
    D.id_pool = P.id_pool;
    D.scr_off = P.scr_off;
    D.raw_off = P.raw_off;
    b->per_read.ensure(std::max<size_t>(R, 1) * 16);
    D.read_len = b->per_read.as<uint32_t>();
    D.ident_num = D.read_len + R;
    D.ident_den = D.ident_num + R;
    D.status = D.ident_den + R;
    b->counters.ensure(512);
    b->h_tot.ensure(256);
    b->ensure_events();
    b->bound = true;
    return true;
}

void run_scan(tksm_seq* h, tksm_seq_batch* b, int o, uint32_t R, uint64_t* d_total) {
    cudaStream_t S = h->lane_stream[b->lane];
    const uint32_t per = SCAN_BLOCK * SCAN_ITEMS;
    const uint32_t nb = (R + per - 1) / per;
    b->sums.ensure((size_t)(nb + 2) * 8);
    scan_block_sums<<<nb, SCAN_BLOCK, 0, S>>>(b->rec_len[o].as<uint32_t>(), R, b->sums.as<uint64_t>());
    scan_sums<<<1, 1024, 0, S>>>(b->sums.as<uint64_t>(), nb, d_total);
    scan_final<<<nb, SCAN_BLOCK, 0, S>>>(b->rec_len[o].as<uint32_t>(), R, b->sums.as<uint64_t>(),
                                                 b->rec_off[o].as<uint64_t>());
    CK(cudaGetLastError());
}

// counters buffer (64 bytes): u32 work tickets [0..5] | +32: u64 record bytes [2] | +48: u64 read bases | +56: u32 status OR
constexpr size_t CNT_TOTALS = 32, CNT_READ_BASES = 48, CNT_STATUS = 56;

// Enqueues the kernel sequence of the bound chunk on the compute stream; no host synchronisation.
// grow: second pass of a chunk in which a read outgrew its slot (slot_wait).  Every read gets GROW_SLOT_MUL times
// its scratch slot and room for MAX_SLOT_LEN bases per fragment slot, which no edit script can exceed; the length
// classes are chosen by grow_len(n) so that the workspaces of a class hold such a read.  Results depend on the read
// ordinal only, so the reads that had fitted come out as before.
void slot_launch(tksm_seq* h, tksm_seq_batch* b, int which, int format, bool grow = false) {
    cudaStream_t S = h->lane_stream[b->lane];
    if (!b->bound) throw std::runtime_error("batch not uploaded");
    CK(cudaSetDevice(h->cfg.device));
    Prepared& p = b->prep;
    const uint32_t R = (uint32_t)p.n_reads;
    b->which = which;
    b->format = format;
    b->grown = grow;
    const uint64_t scr_total = grow ? p.scr_total * GROW_SLOT_MUL : p.scr_total;
    const uint32_t* n_work = grow ? p.n_work_grow : p.n_work;
    const uint32_t max_n_gws = grow ? p.max_n_gws_grow : p.max_n_gws;
    b->out_bytes[0] = b->out_bytes[1] = 0;
    b->launches = 0;
    if (R == 0) return;
    const bool want_bad = which & TKSM_SEQ_BADREAD, want_perf = which & TKSM_SEQ_PERFECT;
    const bool fastq_o[2] = {(format & TKSM_SEQ_FASTQ_BADREAD) != 0, (format & TKSM_SEQ_FASTQ_PERFECT) != 0};
    const bool fastq = fastq_o[0];
    DevBatch& D = b->db;
    D.scr_off = (const uint64_t*)((const char*)b->blob.p + p.o_so);
    if (grow) {
        b->scr_off_grow.ensure(((size_t)R + 1) * 8);
        scale_offsets_kernel<<<(R + 256) / 256, 256, 0, S>>>(D.scr_off, R + 1, GROW_SLOT_MUL, b->scr_off_grow.as<uint64_t>());
        CK(cudaGetLastError());
        D.scr_off = b->scr_off_grow.as<uint64_t>();
    }
    uint32_t flags = 0;
    if (want_bad) flags |= RUN_ERRORS;
    if (want_bad && fastq && !h->cfg.skip_qual_compute) flags |= RUN_QUAL;  // py/sequence.py:349-353
    if (want_perf) flags |= RUN_RAW;
    {
        static const int dbg = getenv("TKSM_SEQ_DEBUG") ? atoi(getenv("TKSM_SEQ_DEBUG")) : 0;
        if (dbg & 1) flags |= RUN_NO_ALIGN;
        if (dbg & 2) flags |= RUN_NO_WIDE;
    }
    if (want_bad) {
        b->scr_seq.ensure(scr_total);
        if (flags & RUN_QUAL) {
            b->scr_qual.ensure(scr_total);
            b->scr_ops.ensure(scr_total + 64);
            b->scr_cig.ensure(scr_total + 64);
            b->rinfo.ensure((size_t)R * sizeof(RInfo));
        }
    }
    if (want_perf) b->scr_raw.ensure(std::max<uint64_t>(p.raw_total, 16));
    D.scr_seq = b->scr_seq.as<uint8_t>();
    D.scr_qual = b->scr_qual.as<uint8_t>();
    D.scr_ops = (flags & RUN_QUAL) ? b->scr_ops.as<uint8_t>() : nullptr;
    D.scr_cig = (flags & RUN_QUAL) ? b->scr_cig.as<uint8_t>() : nullptr;
    D.rinfo = (flags & RUN_QUAL) ? b->rinfo.as<RInfo>() : nullptr;
    D.scr_raw = want_perf ? b->scr_raw.as<uint8_t>() : nullptr;
    // output upper bounds
    b->out_cap[0] = want_bad ? (uint64_t)R * 128 + p.idsum + scr_total * (fastq ? 2 : 1) : 0;
    b->out_cap[1] = want_perf ? (uint64_t)R * 128 + p.idsum + p.total_raw_bases * (fastq_o[1] ? 2 : 1) : 0;
    if (want_bad) b->out[0].ensure(b->out_cap[0]);
    if (want_perf) b->out[1].ensure(b->out_cap[1]);
    for (int o = 0; o < 2; o++)
        if ((o == 0 && want_bad) || (o == 1 && want_perf)) {
            b->rec_len[o].ensure((size_t)R * 4);
            b->rec_off[o].ensure((size_t)R * 8);
        }
    uint32_t ncap_c[N_CLASSES], cap_list_c[N_CLASSES], slab_c[N_CLASSES];
    for (int c = 0; c < N_CLASSES; c++) {
        ncap_c[c] = h->ncap[c];
        cap_list_c[c] = h->cap_list[c];
        slab_c[c] = h->slab[c];
    }
    if (max_n_gws) {
        const uint32_t ncap = (uint32_t)align_up(max_n_gws + 16, 16);
        ncap_c[GWS_CLASS] = ncap;
        cap_list_c[GWS_CLASS] = ncap / 2 + 64;
        slab_c[GWS_CLASS] = std::max(slab_layout(CLASS_W[GWS_CLASS], 8, ncap, cap_list_c[GWS_CLASS]).bytes,
                                     slab2_big_bytes(ncap));
        b->gws.ensure((size_t)slab_c[GWS_CLASS] * h->num_sms * GWS_FULL_CTAS);
    }
    // the class of 24 warps: its FIN_FULL kernel runs as a global-workspace instance too (fragment, slot lengths, read)
    const size_t gws4_slab = align_up((size_t)slab2_big_bytes(ncap_c[4]) + 256, 256);
    if (n_work[4] && (flags & RUN_ERRORS)) b->gws4.ensure(gws4_slab * h->num_sms * GWS_FULL_CTAS);
    uint64_t* d_totals = (uint64_t*)((char*)b->counters.p + CNT_TOTALS);
    // u32 words: [0..5] error-injection tickets | [8..15] totals (u64 x 4) | [16..21] expansion tickets | [24..29] qscore
    // tickets | [32..40] debug | [48..53] redo counts | [56..61] redo tickets | [64..69] alignment tickets | [72..77] tickets of its 64-bit instance | [80..85] its work counts | [88..93] qscore tickets of those reads
    CK(cudaMemsetAsync(b->counters.p, 0, 512, S));
    if (want_bad) b->mid.ensure(3 * scr_total + 64 * (uint64_t)R + 64);
    // alignment trace cells.  FIN_FULL kernels: one region per team of the grid (the classes run concurrently; the redo
    // launch of a class uses the regions of its first CTAs).  align_kernel: one region per warp, (Dg, Pv) of column j of
    // the warp's 32 reads side by side.
    uint64_t trace_off[N_CLASSES] = {}, trace_team[N_CLASSES] = {}, trace_total = 0;
    uint64_t atrace_off[N_CLASSES] = {}, atrace_warp[N_CLASSES] = {}, atrace_total = 0;
    uint64_t atrace64_off[N_CLASSES] = {}, atrace64_warp[N_CLASSES] = {};
    bool wide_class[N_CLASSES] = {};
    int agrid64[N_CLASSES] = {};
    int agrid[N_CLASSES] = {};
    if (flags & RUN_QUAL)
        for (int c = 0; c < N_CLASSES; c++) {
            if (n_work[c] == 0) continue;
            const bool gws_full = c == GWS_CLASS || c == 4;   // FIN_FULL as a global-workspace instance: one team per CTA
            const int W2 = gws_full ? GWS_W2 : CLASS_W[c];
            const int teams2 = gws_full ? 1 : (int)h->threads2[c] / 32 / CLASS_W[c];
            size_t grid2 = std::min<size_t>((size_t)h->num_sms * (gws_full ? GWS_FULL_CTAS : 1), (n_work[c] + teams2 - 1) / teams2);
            // (a class of long fragments with few reads is a few long chains for align_kernel, 7 ms at 8192 slots whatever
            // the batch; measured: the team-parallel FIN_FULL kernel on such a class is no faster)
            if (CLASS_W[c] < 24) grid2 = std::min<size_t>(grid2, (size_t)redo_grid_cap(c, h->num_sms));   // redo lists only
            trace_team[c] = trace_team_bytes(W2, ncap_c[c]);
            trace_off[c] = trace_total;
            trace_total += trace_team[c] * teams2 * grid2;
            if (CLASS_W[c] < 24) {
                const uint32_t cols = std::min<uint32_t>(ncap_c[c], ALIGN_WIDE_FROM) + 2;
                constexpr int AW = ALIGN_K_THREADS / 32;
                // 32-bit instance: enough warps to fill the SMs (20 resident per SM: shared memory), bounded by the reads
                // there are and by 3 GB of cells; the 64-bit instance (about one read in a thousand) gets 16 CTAs
                atrace_warp[c] = (uint64_t)cols * 32 * sizeof(AlignCell2<uint32_t>);
                static const int aw32 = getenv("TKSM_SEQ_AW32") ? atoi(getenv("TKSM_SEQ_AW32")) : 36;   // warps per SM (56 registers)
                static const int aw64 = getenv("TKSM_SEQ_AW64") ? atoi(getenv("TKSM_SEQ_AW64")) : 16;
                const uint64_t warps_mem = std::max<uint64_t>((6144ull << 20) / atrace_warp[c], 64);
                const uint64_t warps = std::min<uint64_t>({(uint64_t)h->num_sms * aw32, (n_work[c] + 31) / 32, warps_mem});
                agrid[c] = (int)((warps + AW - 1) / AW);
                // (grow: a class also holds reads well below the capacity of the class before it)
                wide_class[c] = !grow && c > 0 && ncap_c[c - 1] >= ALIGN_WIDE64_FROM;
                atrace_off[c] = atrace_total;
                if (!wide_class[c]) atrace_total += atrace_warp[c] * AW * agrid[c];
                atrace64_warp[c] = (uint64_t)cols * 32 * sizeof(AlignCell2<uint64_t>);
                if (wide_class[c]) {
                    const uint64_t wm = std::max<uint64_t>((6144ull << 20) / atrace64_warp[c], 64);
                    const uint64_t w64 = std::min<uint64_t>({(uint64_t)h->num_sms * aw64, (n_work[c] + 31) / 32, wm});
                    agrid64[c] = (int)((w64 + AW - 1) / AW);
                } else {
                    // (a class that also holds reads above ALIGN_WIDE64_FROM slots gives this instance real work)
                    agrid64[c] = std::min(agrid[c], ncap_c[c] > ALIGN_WIDE64_FROM ? h->num_sms : 16);
                }
                atrace64_off[c] = atrace_total;
                atrace_total += atrace64_warp[c] * AW * agrid64[c];
            }
        }
    b->trace.ensure(std::max<uint64_t>(trace_total, 16));
    b->atrace.ensure(std::max<uint64_t>(atrace_total, 16));
    CK(cudaEventRecord(b->ev[0], S));
    uint32_t launches = 0;
    {
        // device-side work order (kernels.cuh: class | identity bucket | length)
        b->work_keys.ensure((size_t)R * 4);
        b->work_sorted.ensure((size_t)R * 4);
        b->redo.ensure((size_t)R * 4);
        b->redo64.ensure((size_t)R * 4);
        b->work_bins.ensure((size_t)WORK_BINS * 4);
        CK(cudaMemsetAsync(b->work_bins.p, 0, (size_t)WORK_BINS * 4, S));
        WorkSortParams WS{};
        WS.n_reads = R;
        WS.ordinal_base = p.ordinal_base;
        WS.seed_lo = h->dt.seed_lo;
        WS.seed_hi = h->dt.seed_hi;
        WS.ident_const = h->dt.ident_const;
        WS.k = (uint32_t)h->em.k;
        WS.grow = grow ? 1u : 0u;
        WS.read_mol = D.read_mol;
        WS.mol_raw_len = D.mol_raw_len;
        for (int c = 0; c < 5; c++) WS.ncap[c] = ncap_c[c];
        WS.gws_cap = ncap_c[GWS_CLASS];
        for (int c = 0; c < N_CLASSES; c++) {
            uint32_t sh = 0;
            while ((ncap_c[c] >> sh) > 511u) sh++;
            WS.len_shift[c] = sh;
        }
        WS.keys = b->work_keys.as<uint32_t>();
        WS.bins = b->work_bins.as<uint32_t>();
        WS.work = b->work_sorted.as<uint32_t>();
        work_key_kernel<<<(R + 255) / 256, 256, 0, S>>>(WS);
        work_scan_kernel<<<1, 1024, 0, S>>>(WS.bins);
        work_scatter_kernel<<<(R + 255) / 256, 256, 0, S>>>(WS);
        CK(cudaGetLastError());
        launches += 3;
        uint32_t off = 0;
        for (int c = 0; c < N_CLASSES; c++) {
            b->d_work[c] = b->work_sorted.as<uint32_t>() + off;
            off += n_work[c];
        }
    }
    // The classes are independent (own reads, own ticket counter): each gets its own stream so that
    // the CTAs of the next class fill the SMs the tail of the previous one leaves idle.
    CK(cudaEventRecord(b->ev_sorted, S));
    uint32_t used_streams = 0;
    for (int c = 0; c < N_CLASSES; c++) {
        if (n_work[c] == 0) continue;
        SimParams P;
        P.T = h->dt;
        P.B = D;
        P.work = b->d_work[c];
        P.n_work = n_work[c];
        P.counter = b->counters.as<uint32_t>() + c;
        P.flags = flags;
        P.ncap = ncap_c[c];
        P.cap_list = cap_list_c[c];
        P.slab_bytes = slab_c[c];
        P.gws = b->gws.as<uint8_t>();
        P.mid = b->mid.as<uint8_t>();
        P.trace = b->trace.as<uint8_t>() + trace_off[c];
        P.trace_bytes = trace_team[c];
        P.dbg = (uint32_t*)((char*)b->counters.p + 128);
        P.redo = b->redo.as<uint32_t>() + (b->d_work[c] - b->work_sorted.as<uint32_t>());
        P.redo_count = b->counters.as<uint32_t>() + 48 + c;
        P.n_work_dev = nullptr;
        P.rb_mul = grow ? GROW_RB_MUL : 2u;
        AlignLaunch al{};
        al.grid = agrid[c];
        al.K.B = D;
        al.K.work = P.work;
        al.K.n_work = P.n_work;
        al.K.counter = b->counters.as<uint32_t>() + 64 + c;
        al.K.mid = P.mid;
        al.K.trace = b->atrace.as<uint8_t>() + atrace_off[c];
        al.K.trace_warp_bytes = atrace_warp[c];
        al.side = h->side_stream[b->lane][used_streams % 4];
        al.ev_fork = b->ev_fork[used_streams % 4];
        al.ev_side = b->ev_side[used_streams % 4];
        al.ev_early = b->ev_early[used_streams % 4];
        al.trace64 = b->atrace.as<uint8_t>() + atrace64_off[c];
        al.trace64_warp_bytes = atrace64_warp[c];
        al.K.k = (uint32_t)h->em.k;
        al.K.flags = flags;
        // the 32-bit instance hands its failures to the 64-bit one, which hands its own to the FIN_FULL redo list
        al.redo64 = b->redo64.as<uint32_t>() + (b->d_work[c] - b->work_sorted.as<uint32_t>());
        al.redo64_count = b->counters.as<uint32_t>() + 80 + c;
        al.K.n_work_dev = nullptr;
        al.wide_class = wide_class[c];
        al.redo_grid = redo_grid_cap(c, h->num_sms);
        al.grid64 = agrid64[c];
        P.redo64 = wide_class[c] ? nullptr : al.redo64;
        al.early64 = !wide_class[c] && ncap_c[c] > ALIGN_WIDE64_FROM;
        P.redo64_count = al.redo64_count;
        al.K.redo = al.redo64;
        al.K.redo_count = al.redo64_count;
        const int teams = c == GWS_CLASS ? 1 : SIM_WARPS / CLASS_W[c];
        int grid = (int)std::min<size_t>((size_t)h->num_sms, (P.n_work + teams - 1) / teams);
        Phase2Launch p2;
        {
            const bool gws_full = c == GWS_CLASS || c == 4;
            const int teams2 = gws_full ? 1 : (int)h->threads2[c] / 32 / CLASS_W[c];
            p2.smem = h->smem_bytes2[gws_full ? GWS_CLASS : c];
            p2.grid = (int)std::min<size_t>((size_t)h->num_sms * (gws_full ? GWS_FULL_CTAS : 1), (P.n_work + teams2 - 1) / teams2);   // (as the trace regions above)
            p2.threads = (int)h->threads2[gws_full ? GWS_CLASS : c];
            p2.slab_bytes = c == GWS_CLASS ? slab_c[c] : c == 4 ? (uint32_t)gws4_slab : h->slab2[c];   // the global workspaces keep their size
            if (c == 4) p2.gws = b->gws4.as<uint8_t>();
        }
        const uint32_t si = used_streams % 4;
        cudaStream_t st = h->class_stream[b->lane][si];
        if (used_streams < 4) CK(cudaStreamWaitEvent(st, b->ev_sorted, 0));
        used_streams++;
        switch (c) {
            case 0: launch_sim<1, false>(st, P, h->smem_bytes[c], grid, p2, al); break;
            case 1: launch_sim<2, false>(st, P, h->smem_bytes[c], grid, p2, al); break;
            case 2: launch_sim<4, false>(st, P, h->smem_bytes[c], grid, p2, al); break;
            case 3: launch_sim<8, false>(st, P, h->smem_bytes[c], grid, p2, al); break;
            case 4: launch_sim<24, false, GWS_W2, true>(st, P, h->smem_bytes[c], grid, p2, al); break;
            default: launch_sim<24, true, GWS_W2>(st, P, h->smem_bytes[c], grid, p2, al); break;
        }
        // inject | FIN_FULL, or inject | expand, or inject | expand | align | qscores | [64-bit align early] | 64-bit align | its
        // qscores | FIN_FULL on the redo list (a class of the 64-bit band only: inject | expand | align | qscores | FIN_FULL)
        launches += !(flags & RUN_ERRORS) ? 1 : CLASS_W[c] >= 24 ? 2 : !(flags & RUN_QUAL) ? 2 : wide_class[c] ? 5 : al.early64 ? 8 : 7;
        if (getenv("TKSM_SEQ_DEBUG") && (atoi(getenv("TKSM_SEQ_DEBUG")) & 8)) {
            const cudaError_t e = cudaStreamSynchronize(st);
            fprintf(stderr, "[debug] class %d (W=%d, %u reads, ncap %u): %s\n", c, CLASS_W[c], P.n_work, P.ncap, cudaGetErrorString(e));
        }
    }
    for (uint32_t i = 0; i < std::min<uint32_t>(used_streams, 4); i++) {
        CK(cudaEventRecord(b->ev_class[i], h->class_stream[b->lane][i]));
        CK(cudaStreamWaitEvent(S, b->ev_class[i], 0));
    }
    CK(cudaEventRecord(b->ev[1], S));
    for (int o = 0; o < 2; o++) {
        if (!((o == 0 && want_bad) || (o == 1 && want_perf))) continue;
        FmtParams F{};
        F.n_reads = R;
        F.ordinal_base = p.ordinal_base;
        F.seed_lo = h->dt.seed_lo;
        F.seed_hi = h->dt.seed_hi;
        F.read_mol = D.read_mol;
        F.mol_raw_len = D.mol_raw_len;
        F.mol_tail_len = p.o_tail == SIZE_MAX ? nullptr : (const uint32_t*)((const char*)b->blob.p + p.o_tail);
        F.mol_id_begin = D.mol_id_begin;
        F.id_pool = D.id_pool;
        if (o == 0) {
            F.read_len = D.read_len;
            F.ident_num = D.ident_num;
            F.ident_den = D.ident_den;
            F.seq = D.scr_seq;
            F.seq_off = D.scr_off;
            F.qual = (flags & RUN_QUAL) ? D.scr_qual : nullptr;
            F.status = D.status;
            F.sum_read_len = (unsigned long long*)((char*)b->counters.p + CNT_READ_BASES);
            F.status_or = (uint32_t*)((char*)b->counters.p + CNT_STATUS);
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
        reclen_kernel<<<(R + 255) / 256, 256, 0, S>>>(F);
        run_scan(h, b, o, R, d_totals + o);
        int grid = (int)std::min<uint32_t>((R + 7) / 8, (uint32_t)h->num_sms * 8);
        format_kernel<<<grid, 256, 0, S>>>(F);
        CK(cudaGetLastError());
        launches += 5;
    }
    CK(cudaEventRecord(b->ev[2], S));
    CK(cudaMemcpyAsync(b->h_tot.p, d_totals, 224, cudaMemcpyDeviceToHost, S));
    CK(cudaEventRecord(b->ev[4], S));
    b->launches = launches;
}

// Waits for the kernels of slot `b` and fills the statistics of the chunk.
void slot_wait(tksm_seq* h, tksm_seq_batch* b, tksm_seq_stats* st) {
    cudaStream_t S = h->lane_stream[b->lane];
    Prepared& p = b->prep;
    tksm_seq_stats s{};
    s.reads = p.n_reads;
    s.raw_bases = p.total_raw_bases;
    s.h2d_bytes = b->h2d_bytes;
    if (s.reads == 0) {
        CK(cudaStreamSynchronize(S));
        p.stage.reset();
        p.heap_blob.reset();
        p.blob = nullptr;
        if (st) *st = s;
        return;
    }
    CK(cudaEventSynchronize(b->ev[4]));
    p.stage.reset();  // the upload is long done
    p.heap_blob.reset();
    p.blob = nullptr;
    const bool want_bad = b->which & TKSM_SEQ_BADREAD, want_perf = b->which & TKSM_SEQ_PERFECT;
    const uint64_t* tot = (const uint64_t*)b->h_tot.p;
    if (const uint32_t* dbg = (const uint32_t*)((const char*)b->h_tot.p + 96); dbg[0])
        fprintf(stderr, "[align paranoid] %u violations; first: read %u site %u index %u limit %u j %u (n %u, m %u)\n", dbg[0], dbg[1],
                dbg[3], dbg[4], dbg[5], dbg[6], dbg[7], dbg[8]);
    float ms_all = 0, ms_sim = 0;
    CK(cudaEventElapsedTime(&ms_all, b->ev[0], b->ev[2]));
    CK(cudaEventElapsedTime(&ms_sim, b->ev[0], b->ev[1]));
    b->out_bytes[0] = want_bad ? tot[0] : 0;
    b->out_bytes[1] = want_perf ? tot[1] : 0;
    if (b->out_bytes[0] > b->out_cap[0] || b->out_bytes[1] > b->out_cap[1]) throw std::runtime_error("output exceeds its bound");
    s.out_bytes[0] = b->out_bytes[0];
    s.out_bytes[1] = b->out_bytes[1];
    s.kernel_ms = ms_all;
    s.simulate_ms = ms_sim;
    s.kernel_launches = b->launches;
    s.algorithmic_bytes = p.alg_in_bytes + b->out_bytes[0] + b->out_bytes[1];
    if (want_bad) {
        s.read_bases = tot[2];
        s.status_or = (uint32_t)tot[3];
        if ((s.status_or & (ST_OPS_OVERFLOW | ST_SCRATCH_OVERFLOW)) && !b->grown) {
            // a read outgrew its slot (more than twice its fragment): the chunk runs again in slots no read can outgrow
            slot_launch(h, b, b->which, b->format, true);
            slot_wait(h, b, st);
            return;
        }
    } else {
        s.read_bases = p.total_raw_bases;
    }
    if (st) *st = s;
}

void batch_fetch(tksm_seq* h, tksm_seq_batch* b, int o, char* dst) {
    cudaStream_t S = h->lane_stream[b->lane];
    if (b->out_bytes[o] == 0) return;
    CK(cudaSetDevice(h->cfg.device));
    CK(cudaMemcpyAsync(dst, b->out[o].p, b->out_bytes[o], cudaMemcpyDeviceToHost, S));
    CK(cudaStreamSynchronize(S));
}

void add_stats(tksm_seq_stats& total, const tksm_seq_stats& s) {
    total.reads += s.reads;
    total.raw_bases += s.raw_bases;
    total.read_bases += s.read_bases;
    total.out_bytes[0] += s.out_bytes[0];
    total.out_bytes[1] += s.out_bytes[1];
    total.algorithmic_bytes += s.algorithmic_bytes;
    total.h2d_bytes += s.h2d_bytes;
    total.d2h_bytes += s.d2h_bytes;
    total.kernel_ms += s.kernel_ms;
    total.simulate_ms += s.simulate_ms;
    total.kernel_launches += s.kernel_launches;
    total.status_or |= s.status_or;
}

// ---------------------------------------------------------------- pinned arena pool
// Record buffers handed to the caller by tksm_seq_simulate_text are page-locked so the device
// writes them directly; tksm_seq_free() returns them here and the next call reuses them.
struct ArenaPool {
    struct Block {
        void* p;
        size_t cap;
        bool used;
    };
    std::mutex mu;
    std::vector<Block> blocks;
    // a block of at least `need` bytes: the largest free one there is (the estimate `bytes` of a call's output moves a
    // little from call to call with the grouping of its batches; page-locking a new 1.6 GB block for that costs most of
    // a second), else a new one of `bytes`
    void* acquire(size_t bytes, size_t need = 0) {
        std::lock_guard<std::mutex> g(mu);
        if (need == 0 || need > bytes) need = bytes;
        int best = -1;
        for (size_t i = 0; i < blocks.size(); i++)
            if (!blocks[i].used && blocks[i].cap >= need && (best < 0 || blocks[i].cap > blocks[(size_t)best].cap)) best = (int)i;
        if (best >= 0) {
            blocks[(size_t)best].used = true;
            return blocks[(size_t)best].p;
        }
        // drop free blocks that are too small before growing
        for (size_t i = 0; i < blocks.size();)
            if (!blocks[i].used) {
                cudaFreeHost(blocks[i].p);
                blocks.erase(blocks.begin() + (long)i);
            } else
                i++;
        void* p = nullptr;
        size_t want = bytes + bytes / 8 + 4096;
        CK(cudaMallocHost(&p, want));
        blocks.push_back({p, want, true});
        return p;
    }
    size_t capacity(void* p) {
        std::lock_guard<std::mutex> g(mu);
        for (auto& b : blocks)
            if (b.p == p) return b.cap;
        return 0;
    }
    bool release(void* p) {
        std::lock_guard<std::mutex> g(mu);
        for (auto& b : blocks)
            if (b.p == p) {
                b.used = false;
                return true;
            }
        return false;
    }
};
ArenaPool& arena_pool() {
    static ArenaPool* pool = new ArenaPool();  // never destroyed: outlives the CUDA context teardown order
    return *pool;
}

// ---------------------------------------------------------------- host worker pool
class WorkerPool {
  public:
    explicit WorkerPool(int n) {
        for (int i = 0; i < n; i++) th_.emplace_back([this] { loop(); });
    }
    ~WorkerPool() {
        {
            std::lock_guard<std::mutex> g(mu_);
            stop_ = true;
        }
        cv_.notify_all();
        for (auto& t : th_) t.join();
    }
    template <typename F>
    auto submit(F&& f) -> std::future<decltype(f())> {
        auto task = std::make_shared<std::packaged_task<decltype(f())()>>(std::forward<F>(f));
        auto fut = task->get_future();
        {
            std::lock_guard<std::mutex> g(mu_);
            q_.emplace_back([task] { (*task)(); });
        }
        cv_.notify_one();
        return fut;
    }

  private:
    void loop() {
        while (true) {
            std::function<void()> job;
            {
                std::unique_lock<std::mutex> g(mu_);
                cv_.wait(g, [this] { return stop_ || !q_.empty(); });
                if (q_.empty()) return;
                job = std::move(q_.front());
                q_.pop_front();
            }
            job();
        }
    }
    std::mutex mu_;
    std::condition_variable cv_;
    std::deque<std::function<void()>> q_;
    std::vector<std::thread> th_;
    bool stop_ = false;
};

int host_threads(const tksm_seq* h) {
    if (h->cfg.threads > 0) return h->cfg.threads;
    unsigned hc = std::thread::hardware_concurrency();
    return (int)std::min<unsigned>(std::max<unsigned>(hc, 1), 32);
}

std::atomic<uint64_t> g_ns_parse{0}, g_ns_append{0}, g_ns_prepare{0};
struct ScopedNs {
    std::atomic<uint64_t>& acc;
    std::chrono::steady_clock::time_point t0 = std::chrono::steady_clock::now();
    explicit ScopedNs(std::atomic<uint64_t>& a) : acc(a) {}
    ~ScopedNs() { acc += (uint64_t)std::chrono::duration_cast<std::chrono::nanoseconds>(std::chrono::steady_clock::now() - t0).count(); }
};

using PreparedList = std::vector<std::unique_ptr<Prepared>>;
using HostBatchList = std::vector<std::unique_ptr<HostBatch>>;

inline size_t limit_reads(const tksm_seq* h) {
    static const size_t env = getenv("TKSM_SEQ_BATCH_READS") ? (size_t)atoll(getenv("TKSM_SEQ_BATCH_READS")) : 0;
    return h->cfg.batch_reads ? (size_t)h->cfg.batch_reads : env ? env : ((size_t)1 << 18);
}
inline uint64_t limit_bases(const tksm_seq* h) { return h->cfg.batch_bases ? h->cfg.batch_bases : (1ull << 28); }

// stage 1 (host threads): one text chunk of complete records -> molecule batches
HostBatchList parse_chunk(const tksm_seq* h, const tk::TextChunk& tc) {
    ScopedNs timer(g_ns_parse);
    HostBatchList out;
    MdfParser parser(&h->genome);
    parser.capture_cb = h->front.needs_cb();
    size_t pos = 0;
    while (true) {
        std::unique_ptr<HostBatch> hb(new HostBatch());
        if (pos == 0) {  // typical shapes: ~25 text bytes per interval line, ~8 lines per molecule
            const size_t lines = tc.n / 22 + 16;
            hb->seg_src.reserve(lines);
            hb->seg_len.reserve(lines);
            hb->seg_mod_begin.reserve(lines + 1);
            hb->read_mol.reserve(lines / 4);
            hb->mol_raw_len.reserve(lines / 4);
            hb->mol_seg_begin.reserve(lines / 4 + 1);
            hb->mol_id_begin.reserve(lines / 4 + 1);
            hb->id_pool.reserve(tc.n / 16);
        }
        size_t used = parser.feed(tc.p + pos, tc.n - pos, true, *hb, limit_reads(h), limit_bases(h));
        pos += used;
        if (hb->n_reads()) out.push_back(std::move(hb));
        if (pos >= tc.n || used == 0) break;
    }
    return out;
}

// Tail-noise model loaded: append the tails (keyed by the read ordinals, known now) and build the blob.
void finish_deferred(const tksm_seq* h, Prepared& p, uint64_t ordinal, ChimeraCarry* carry) {
    ScopedNs timer(g_ns_prepare);
    const std::vector<const HostBatch*> parts = std::move(p.deferred_parts);
    const auto keep = std::move(p.deferred_keep);
    p.deferred_parts.clear();
    p.deferred_keep.clear();
    const HostBatch per_read = apply_read_transforms(h->tail, h->extras, h->front, h->cfg.seed, ordinal, parts, carry);
    prepare_host(h, {&per_read}, p);
}

// stage 2 (host threads): the batches of consecutive text chunks, concatenated up to the device
// batch limits, with scratch layout, work lists and upload blob
PreparedList assemble_group(const tksm_seq* h, const std::vector<std::shared_ptr<HostBatchList>>& parts) {
    PreparedList out;
    std::vector<std::shared_ptr<HostBatchList>> lists;
    std::vector<const HostBatch*> cur;
    uint64_t cur_reads = 0, cur_bases = 0;
    auto flush = [&]() {
        if (!cur.empty()) {
            ScopedNs timer(g_ns_prepare);
            std::unique_ptr<Prepared> p(new Prepared());
            if (h->per_read_molecules()) {
                p->deferred_parts = cur;
                p->deferred_keep = lists;
            } else {
                prepare_host(h, cur, *p);
            }
            out.push_back(std::move(p));
        }
        cur.clear();
        cur_reads = cur_bases = 0;
    };
    for (auto& f : parts) {
        lists.push_back(f);
        for (auto& hb : *lists.back()) {
            if (!cur.empty() && (cur_reads + hb->n_reads() > limit_reads(h) || cur_bases + hb->total_raw_bases > limit_bases(h)))
                flush();
            cur.push_back(hb.get());
            cur_reads += hb->n_reads();
            cur_bases += hb->total_raw_bases;
        }
    }
    flush();
    return out;
}

// The chunks of one device batch while the host threads parse them.  No thread waits for another: the
// thread that finishes the last chunk of a closed group assembles the batch.
// device MDF tokeniser: what a host thread notes about a text chunk (the text itself is copied when the group closes)
struct TextNote {
    tk::TextChunk tc;
    uint32_t lines = 0, records = 0;
};
TextNote note_chunk(const tk::TextChunk& tc) {
    ScopedNs timer(g_ns_parse);
    TextNote t;
    t.tc = tc;
    const char* p = tc.p;
    const char* const e = tc.p + tc.n;
    if (tc.n && *p == '+') t.records++;
    while (p < e) {
        const char* nl = (const char*)memchr(p, '\n', (size_t)(e - p));
        if (!nl) break;
        t.lines++;
        if (nl + 1 < e && nl[1] == '+') t.records++;
        p = nl + 1;
    }
    if (tc.n && e[-1] != '\n') t.lines++;   // the copy gets the missing newline
    return t;
}
// the chunks of a group as one text in a page-locked buffer, to be tokenised on the device
PreparedList assemble_text_group(const tksm_seq* h, const std::vector<TextNote>& notes) {
    ScopedNs timer(g_ns_prepare);
    PreparedList out;
    size_t total = 0;
    uint64_t lines = 0, records = 0;
    for (const TextNote& t : notes) {
        total += t.tc.n + 1;
        lines += t.lines;
        records += t.records;
    }
    if (lines == 0) return out;
    if (lines > 0x7FFFFFF0ull) throw std::runtime_error("device batch too large");
    std::unique_ptr<Prepared> p(new Prepared());
    p->stage = StageLease(h->stage_pool.get(), h->stage_pool->acquire(total + 64, h->cfg.device));
    if (p->stage.idx >= 0) {
        p->blob = h->stage_pool->ptr(p->stage.idx);
    } else {
        p->heap_blob.reset(new char[total + 64]);
        p->blob = p->heap_blob.get();
    }
    size_t pos = 0;
    for (const TextNote& t : notes) {
        if (!t.tc.n) continue;
        memcpy(p->blob + pos, t.tc.p, t.tc.n);
        pos += t.tc.n;
        if (p->blob[pos - 1] != '\n') p->blob[pos++] = '\n';
    }
    p->blob_bytes = pos;
    p->dev_text = true;
    p->text_lines = (uint32_t)lines;
    p->text_records = (uint32_t)records;
    out.push_back(std::move(p));
    return out;
}

struct GroupState {
    std::atomic<int> remaining{1};   // chunks still being parsed, + 1 until the group is closed
    std::mutex mu;
    std::vector<TextNote> notes;     // device tokeniser: by chunk index
    void store_note(size_t idx, TextNote t) {
        std::lock_guard<std::mutex> g(mu);
        if (notes.size() <= idx) notes.resize(idx + 1);
        notes[idx] = std::move(t);
    }
    std::vector<std::shared_ptr<HostBatchList>> lists;   // by chunk index
    std::exception_ptr error;
    std::promise<PreparedList> done;
    void store(size_t idx, std::shared_ptr<HostBatchList> r) {
        std::lock_guard<std::mutex> g(mu);
        if (lists.size() <= idx) lists.resize(idx + 1);
        lists[idx] = std::move(r);
    }
    void fail(std::exception_ptr e) {
        std::lock_guard<std::mutex> g(mu);
        if (!error) error = e;
    }
    void complete(const tksm_seq* h) {
        try {
            if (error) std::rethrow_exception(error);
            done.set_value(notes.empty() ? assemble_group(h, lists) : assemble_text_group(h, notes));
        } catch (...) {
            done.set_exception(std::current_exception());
        }
    }
};

}  // namespace

namespace tk {

// Streaming pipeline: text chunks are parsed and packed by host threads while the device works on
// earlier chunks; record downloads run on their own stream behind the kernels of the next chunk.
//   arena mode (sink == nullptr): the records of `which` (one output) are downloaded straight into a
//     page-locked buffer that is handed to the caller (*arena_out, released with tksm_seq_free).
//   sink mode: records go through the slot's pinned staging to sink->write(), in MDF order.
void engine_stream(tksm_seq* h, ChunkSource& src, int which, int format, uint64_t ordinal_base, OutputSink* sink,
                   char** arena_out, size_t* arena_len, tksm_seq_stats* stats) {
    need_device(h);
    if (!h->models_loaded) throw std::runtime_error("models not loaded");
    commit_reference(h);
    CK(cudaSetDevice(h->cfg.device));
    // device batches in flight (their kernels queued or running) and one more slot being filled
    static const int max_inflight = getenv("TKSM_SEQ_INFLIGHT") ? std::min(std::max(atoi(getenv("TKSM_SEQ_INFLIGHT")), 1), 6) : 2;
    const int NS = max_inflight + 1;
    while ((int)h->slots.size() < NS) h->slots.emplace_back(new tksm_seq_batch());
    const int nthreads = host_threads(h);
    tksm_seq_stats total{};
    uint64_t ordinal = ordinal_base;
    const int o_arena = which == TKSM_SEQ_PERFECT ? 1 : 0;
    char* arena = nullptr;
    size_t arena_cap = 0, arena_pos = 0;
    uint64_t text_total = 0, text_done = 0, cap_done = 0;  // for the arena size estimate

    // MDF tokenisation on the device (mdf_device.cuh) unless the molecules are rebuilt per read on the host anyway
    const int dev_parse_env = getenv("TKSM_SEQ_DEVICE_PARSE") ? atoi(getenv("TKSM_SEQ_DEVICE_PARSE")) : 0;
    const bool dev_parse = dev_parse_env != 0 && !h->per_read_molecules();
    ChimeraCarry chimera_carry;   // the molecule of the last read of the batch before (Badread extras: chimeras)
    struct Pending {
        std::future<PreparedList> fut;
        size_t text_bytes;
    };
    std::deque<Pending> pending;
    bool src_done = false;
    struct InFlight {
        tksm_seq_batch* b;
    };
    std::deque<InFlight> inflight;  // launched, not yet finished (at most 2)
    size_t issued = 0;

    auto release_arena = [&]() {
        if (arena) arena_pool().release(arena);
        arena = nullptr;
    };
    auto slot_reusable = [&](tksm_seq_batch* b) {
        if (b->d2h_pending) {
            CK(cudaEventSynchronize(b->ev[3]));
            b->d2h_pending = false;
            if (sink)
                for (int o = 0; o < 2; o++)
                    if (b->out_bytes[o]) sink->write(o, (const char*)b->host_out[o].p, b->out_bytes[o]);
        }
    };
    auto finish = [&](tksm_seq_batch* b) {
        tksm_seq_stats s{};
        slot_wait(h, b, &s);
        if (sink) {
            for (int o = 0; o < 2; o++) {
                if (!b->out_bytes[o]) continue;
                b->host_out[o].ensure(b->out_bytes[o]);
                CK(cudaMemcpyAsync(b->host_out[o].p, b->out[o].p, b->out_bytes[o], cudaMemcpyDeviceToHost, h->copy_stream));
                s.d2h_bytes += b->out_bytes[o];
            }
        } else {
            const size_t nbytes = b->out_bytes[o_arena];
            if (arena_pos + nbytes > arena_cap) {
                // estimate the whole output from the bound of the chunks seen so far
                size_t want = arena_pos + nbytes;
                if (text_done && text_total > text_done)
                    want = std::max<size_t>(want, (size_t)((double)cap_done / (double)text_done * (double)text_total * 1.05));
                char* bigger = (char*)arena_pool().acquire(want, arena_pos + nbytes);
                if (arena) {
                    CK(cudaStreamSynchronize(h->copy_stream));
                    memcpy(bigger, arena, arena_pos);
                    arena_pool().release(arena);
                }
                arena = bigger;
                arena_cap = arena_pool().capacity(arena);
            }
            static const bool dbg_no_d2h = getenv("TKSM_DBG_NO_D2H") != nullptr;
            if (nbytes && !dbg_no_d2h)
                CK(cudaMemcpyAsync(arena + arena_pos, b->out[o_arena].p, nbytes, cudaMemcpyDeviceToHost, h->copy_stream));
            arena_pos += nbytes;
            s.d2h_bytes += nbytes;
        }
        CK(cudaEventRecord(b->ev[3], h->copy_stream));
        b->d2h_pending = true;
        add_stats(total, s);
    };

    const bool trace = getenv("TKSM_SEQ_TRACE") != nullptr;
    const auto t_begin = std::chrono::steady_clock::now();
    auto now_ms = [&]() { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_begin).count(); };
    double w_parse = 0, w_bind = 0, w_launch = 0, w_finish = 0, w_reuse = 0, dev_ms = 0;
    try {
        WorkerPool pool(nthreads);
        std::shared_ptr<GroupState> group;
        size_t group_chunks = 0;
        // every batch pays the latency of its kernel chain (single-warp stretches in the alignment kernels), so batches
        // start at min_group chunks rather than one
        const size_t min_group = getenv("TKSM_SEQ_MIN_GROUP") ? std::max(1, atoi(getenv("TKSM_SEQ_MIN_GROUP"))) : 8;
        size_t group_text = 0, group_target = min_group, groups_made = 0, ramp = min_group;
        uint64_t text_submitted = 0;
        // text chunks (1 MiB each) per device batch in steady state
        const size_t max_group = getenv("TKSM_SEQ_MAX_GROUP") ? std::max(1, atoi(getenv("TKSM_SEQ_MAX_GROUP"))) : 64;
        auto close_group = [&]() {
            if (!group) return;
            const tksm_seq* hc = h;
            std::shared_ptr<GroupState> gs = std::move(group);
            group.reset();
            pending.push_back({gs->done.get_future(), group_text});
            // every chunk already parsed: the batch is assembled by a host thread, never by this one
            if (gs->remaining.fetch_sub(1) == 1) pool.submit([hc, gs]() { gs->complete(hc); });
            group_chunks = 0;
            group_text = 0;
            groups_made++;
            // small batches first (the device starts early) and, when the input size is known, small
            // batches last (short last download and kernel tail): a batch takes at most half of what is left
            ramp = std::min(ramp * 2, max_group);
            group_target = ramp;
            // (halving the batches towards the end of the input shortened the last download when a batch took a
            // millisecond; with the alignment chains every batch has a floor of 6 ms and full batches to the end are
            // 5 % faster: TKSM_SEQ_TAIL_DIV=2 brings the halving back)
            static const int tail_div = getenv("TKSM_SEQ_TAIL_DIV") ? std::max(1, atoi(getenv("TKSM_SEQ_TAIL_DIV"))) : 1;
            if (text_total > text_submitted) {
                const size_t left = (size_t)((text_total - text_submitted) >> 20);  // 1 MiB chunks left, about
                group_target = std::min(ramp, std::max<size_t>(min_group, left / (size_t)tail_div));
            }
        };
        while (true) {
            // keep the parser threads fed
            while (!src_done && pending.size() < 8) {  // look-ahead: the host threads keep working while this thread waits on the device
                TextChunk tc;
                if (!src.next(tc)) {
                    src_done = true;
                    close_group();
                    break;
                }
                text_total = std::max<uint64_t>(text_total, src.total_bytes_hint());
                const tksm_seq* hc = h;
                if (!group) group = std::make_shared<GroupState>();
                {
                    std::shared_ptr<GroupState> gs = group;
                    const size_t idx = group_chunks++;
                    gs->remaining.fetch_add(1);
                    pool.submit([hc, tc, gs, idx, dev_parse]() {
                        try {
                            if (dev_parse) gs->store_note(idx, note_chunk(tc));
                            else gs->store(idx, std::make_shared<HostBatchList>(parse_chunk(hc, tc)));
                        } catch (...) {
                            gs->fail(std::current_exception());
                        }
                        if (gs->remaining.fetch_sub(1) == 1) gs->complete(hc);
                    });
                }
                group_text += tc.n;
                text_submitted += tc.n;
                if (group_chunks >= group_target) close_group();
            }
            if (pending.empty()) break;
            double t0 = now_ms();
            PreparedList list = pending.front().fut.get();
            w_parse += now_ms() - t0;
            text_done += pending.front().text_bytes;
            pending.pop_front();
            for (auto& pp : list) {
                tksm_seq_batch* b = h->slots[issued % NS].get();
                t0 = now_ms();
                slot_reusable(b);
                // one lane per slot ($TKSM_SEQ_LANES=1: everything on one lane, as before the alignment kernels)
                static const int lanes = getenv("TKSM_SEQ_LANES") ? std::min(std::max(atoi(getenv("TKSM_SEQ_LANES")), 1), 3) : 3;
                b->lane = (int)(issued % (size_t)lanes);
                w_reuse += now_ms() - t0;
                b->prep = std::move(*pp);
                if (!b->prep.deferred_parts.empty()) finish_deferred(h, b->prep, ordinal, &chimera_carry);
                b->prep.ordinal_base = ordinal;
                t0 = now_ms();
                if (b->prep.dev_text) {
                    // the molecule batch is built from the text on the device; text it does not interpret (and every
                    // error the reference would raise) goes through the host parser
                    if (!slot_bind_text(h, b)) {
                        Prepared text = std::move(b->prep);
                        HostBatch hb;
                        MdfParser parser(&h->genome);
                        parser.feed(text.blob, text.blob_bytes, true, hb);
                        b->prep = Prepared();
                        prepare_host(h, {&hb}, b->prep);
                        b->prep.ordinal_base = ordinal;
                        slot_bind(h, b);
                    }
                } else {
                    slot_bind(h, b);
                }
                ordinal += b->prep.n_reads;
                const uint64_t R = b->prep.n_reads;
                cap_done += R * 128 + b->prep.idsum +
                            (o_arena == 0 ? b->prep.scr_total * ((format & TKSM_SEQ_FASTQ_BADREAD) ? 2 : 1)
                                          : b->prep.total_raw_bases * ((format & TKSM_SEQ_FASTQ_PERFECT) ? 2 : 1));
                double t1 = now_ms();
                slot_launch(h, b, which, format);
                double t2 = now_ms();
                w_bind += t1 - t0;
                w_launch += t2 - t1;
                issued++;
                inflight.push_back({b});
                if (inflight.size() >= (size_t)max_inflight) {
                    finish(inflight.front().b);
                    inflight.pop_front();
                }
                w_finish += now_ms() - t2;
                if (trace) fprintf(stderr, "[trace] chunk %zu reads %zu issued at %.2f ms\n", issued - 1, (size_t)b->prep.n_reads, t0);
            }
        }
        while (!inflight.empty()) {
            finish(inflight.front().b);
            inflight.pop_front();
        }
        if (trace)
            fprintf(stderr, "[trace] drained at %.2f ms: wait parse %.2f, slot reuse %.2f, bind %.2f, launch %.2f, finish %.2f, kernels %.2f ms\n",
                    now_ms(), w_parse, w_reuse, w_bind, w_launch, w_finish, total.kernel_ms);
        // flush the last downloads in chunk order
        for (size_t j = issued > (size_t)NS ? issued - NS : 0; j < issued; j++) slot_reusable(h->slots[j % NS].get());
        CK(cudaStreamSynchronize(h->copy_stream));
        if (trace)
            fprintf(stderr, "[trace] done at %.2f ms; host thread time: parse %.1f ms, append %.1f ms, prepare %.1f ms (cumulative)\n",
                    now_ms(), g_ns_parse.load() / 1e6, g_ns_append.load() / 1e6, g_ns_prepare.load() / 1e6);
        if (trace)
            fprintf(stderr, "[trace] prepare: stage buffer %.1f ms, per-read loop %.1f ms, copies %.1f ms (cumulative)\n",
                    g_ns_prep_alloc.load() / 1e6, g_ns_prep_reads.load() / 1e6, g_ns_prep_copy.load() / 1e6);
        (void)dev_ms;
    } catch (...) {
        cudaDeviceSynchronize();
        for (auto& sl : h->slots) sl->d2h_pending = false;
        release_arena();
        throw;
    }
    if (!sink) {
        if (!arena) arena = (char*)arena_pool().acquire(1);
        *arena_out = arena;
        *arena_len = arena_pos;
    }
    if (stats) *stats = total;
}

}  // namespace tk

namespace {

// single-buffer chunker over text in memory: cuts at record boundaries ("\n+"), small chunks first
// so the device starts early
struct MemorySource : tk::ChunkSource {
    const char* text;
    size_t len, pos = 0, target;
    size_t max_target;
    MemorySource(const char* t, size_t n, size_t first = (size_t)1 << 20, size_t max_t = (size_t)8 << 20)
        : text(t), len(n), target(first), max_target(max_t) {}
    bool next(tk::TextChunk& tc) override {
        if (pos >= len) return false;
        size_t end = pos + target;
        if (end >= len) {
            end = len;
        } else {
            const char* q = end > 0 ? (const char*)memmem(text + end - 1, len - end + 1, "\n+", 2) : nullptr;
            end = q ? (size_t)(q - text) + 1 : len;
        }
        tc.p = text + pos;
        tc.n = end - pos;
        tc.keep.reset();
        pos = end;
        target = std::min(target * 2, max_target);
        return true;
    }
    uint64_t total_bytes_hint() const override { return len; }
};

}  // namespace

// ============================================================== C ABI ====

namespace {

// PB1 stops a read on the identity of its own edit script (error g); what the reference stops on and
// what the header prints is the identity of an optimal alignment (error a), which explains part of the edits away:
// a = g - kappa g^2.  kappa depends on the error model's mix of substitutions, insertions and deletions, so it is
// MEASURED when a model is loaded: 512 random fragments of 1000 bases (fixed seed) are simulated at four constant
// target identities, once without and once with the alignment stage, and kappa is the least-squares coefficient of
// g - a on g^2.  (oracle/fit_calibration.py measures the same with the CPU twin; the GPU suite compares the two.)
constexpr int CALIB_MOLS = 512, CALIB_LEN = 1000;
// effective kappa = measured kappa + offset: real sequence explains less away than the random fragments of the
// measurement and the reference overshoots its target a little.  One constant for error models with a k-mer table, one
// for the table-less `random` model; oracle/fit_calibration.py fit (reference runs of tests/golden/ref_calib.json, which
// the distribution tests do not use)
constexpr float CALIB_OFFSET_TABLE = -0.072f, CALIB_OFFSET_RANDOM = 0.026f;
constexpr float CALIB_KAPPA_UNMEASURED = 0.48f;            // host-only handles without calib_kappa in the config
constexpr uint64_t CALIB_SEED = 0x5EEDCA11Bull;

double self_calibrate(tksm_seq* h) {
    std::string mdf;
    mdf.reserve((size_t)CALIB_MOLS * (CALIB_LEN + 40));
    uint64_t state = 12345;
    for (int m = 0; m < CALIB_MOLS; m++) {
        mdf += "+c" + std::to_string(m) + "\t1\t\n";
        for (int i = 0; i < CALIB_LEN; i += 32) {
            state += 0x9E3779B97F4A7C15ull;   // splitmix64
            uint64_t z = state;
            z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
            z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
            z ^= z >> 31;
            for (int t = 0; t < 32 && i + t < CALIB_LEN; t++) mdf += "ACGT"[(z >> (2 * t)) & 3];
        }
        mdf += "\t0\t" + std::to_string(CALIB_LEN) + "\t+\t\n";   // unknown contig name: the name is the sequence (py/sequence.py:307)
    }
    // the settings of the run are put aside: constant targets, no calibration, no tails, the calibration seed
    const DevTables saved = h->dt;
    const TailModel saved_tail = h->tail;
    const auto saved_skip = h->cfg.skip_qual_compute;
    h->tail = TailModel();
    double sx = 0, sy = 0;
    {
        tksm_seq_batch b;
        GenomeStore no_genome;
        MdfParser parser(&no_genome);
        HostBatch hb;
        parser.feed(mdf.data(), mdf.size(), true, hb);
        prepare_host(h, {&hb}, b.prep);
        b.prep.ordinal_base = 0;
        slot_bind(h, &b);
        const size_t R = b.prep.n_reads;
        std::vector<uint32_t> gen(4 * R), ali(4 * R);
        const double levels[4] = {0.78, 0.84, 0.90, 0.95};
        for (double level : levels) {
            h->dt.ident_const = 1;
            h->dt.ident_mean = (float)level;
            h->dt.calib_kappa = 0.0f;
            h->dt.calib_overshoot = 0.0f;
            h->dt.seed_lo = (uint32_t)CALIB_SEED;
            h->dt.seed_hi = (uint32_t)(CALIB_SEED >> 32);
            for (int pass = 0; pass < 2; pass++) {
                h->cfg.skip_qual_compute = pass == 0;
                slot_launch(h, &b, TKSM_SEQ_BADREAD, TKSM_SEQ_FASTQ_BADREAD);
                tksm_seq_stats st;
                slot_wait(h, &b, &st);
                CK(cudaMemcpy((pass ? ali : gen).data(), b.per_read.p, 16 * R, cudaMemcpyDeviceToHost));
            }
            for (size_t r = 0; r < R; r++) {
                if (!(ali[2 * R + r] & IDENT_MEASURED)) continue;   // not aligned (never expected here)
                const double g = 1.0 - (double)gen[R + r] / (double)gen[2 * R + r];
                const double a = 1.0 - (double)ali[R + r] / (double)(ali[2 * R + r] & ~IDENT_MEASURED);
                sx += g * g * (g - a);
                sy += g * g * g * g;
            }
        }
    }
    // (slot_bind may have committed the reference meanwhile: only the fields touched here are put back)
    h->dt.ident_const = saved.ident_const;
    h->dt.ident_mean = saved.ident_mean;
    h->dt.calib_kappa = saved.calib_kappa;
    h->dt.calib_overshoot = saved.calib_overshoot;
    h->dt.seed_lo = saved.seed_lo;
    h->dt.seed_hi = saved.seed_hi;
    h->tail = saved_tail;
    h->cfg.skip_qual_compute = saved_skip;
    if (!(sy > 0)) throw std::runtime_error("identity calibration: no read was aligned");
    return sx / sy;
}

}  // namespace

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
    c->threads = 0;
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
        h->stage_pool.reset(new StagePool(14, !h->cfg.host_only));
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
            CK(cudaStreamCreateWithFlags(&h->copy_stream, cudaStreamNonBlocking));
            for (auto& lane : h->class_stream)
                for (auto& cs : lane) CK(cudaStreamCreateWithFlags(&cs, cudaStreamNonBlocking));
            for (auto& lane : h->side_stream)
                for (auto& cs : lane) CK(cudaStreamCreateWithFlags(&cs, cudaStreamNonBlocking));
            h->lane_stream[0] = h->stream;
            CK(cudaStreamCreateWithFlags(&h->lane_stream[1], cudaStreamNonBlocking));
            CK(cudaStreamCreateWithFlags(&h->lane_stream[2], cudaStreamNonBlocking));
        }
    } catch (const std::exception& e) {
        g_create_error = e.what();
        return nullptr;
    }
    return h.release();
}

void tksm_seq_destroy(tksm_seq* h) { delete h; }

const char* tksm_seq_last_error(const tksm_seq* h) { return h ? h->err.c_str() : g_create_error.c_str(); }

void tksm_seq_free(void* p) {
    if (p && !arena_pool().release(p)) free(p);
}

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
        g.note_name(strlen(names[i]));
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
        // device-generated genome: only the words of the range come back (the exception runs are on the host)
        need_device(h);
        CK(cudaSetDevice(h->cfg.device));
        const uint64_t w0 = g0 >> 4, w1 = (g0 + len + 15) >> 4;
        std::vector<uint32_t> w((size_t)(w1 - w0 + 1), 0);
        if (w1 > w0) CK(cudaMemcpy(w.data(), h->d_genome.as<uint32_t>() + w0, (w1 - w0) * 4, cudaMemcpyDeviceToHost));
        for (uint64_t i = 0; i < len; i++) {
            const uint64_t g = g0 + i;
            out[i] = "ACGT"[(w[(size_t)((g >> 4) - w0)] >> (2 * (g & 15))) & 3];
        }
        const auto& G = h->genome;
        size_t r = (size_t)(std::lower_bound(G.exc_start.begin(), G.exc_start.end(), g0) - G.exc_start.begin());
        if (r > 0) r--;
        for (; r < G.exc_start.size() && G.exc_start[r] < g0 + len; r++) {
            const uint64_t a = std::max(G.exc_start[r], g0), b = std::min(G.exc_start[r] + G.exc_len[r], g0 + len);
            for (uint64_t g = a; g < b; g++) out[g - g0] = (char)G.exc_chr[r];
        }
    } else {
        std::string s = h->genome.fetch(g0, len);
        memcpy(out, s.data(), len);
    }
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
    // KDE_noise_generator.load (py/tksm_badread.py:937-959): "no_noise", a name found in $TKSM_MODELS, or a path
    std::string tail = tail_model ? tail_model : "no_noise";
    h->tail = TailModel();
    if (tail != "no_noise") {
        auto it = names.tail.find(tail);
        const std::string path = it != names.tail.end() ? it->second : tail;
        fprintf(stderr, "Loading tail model %s\n", path.c_str());
        h->tail.load_json(read_text_maybe_gz(path));
    }
    fprintf(stderr, "\nLoading error model %s\n", em.c_str());
    // (the two files are independent: the qscore model is read and packed while the error model's k-mers are aligned)
    std::future<QScoreModelHost> qm_loading = std::async(std::launch::async, [&]() { return load_qscore_model(qm, names.qscore); });
    try {
        h->em = load_error_model(em, names.error);
    } catch (...) {
        qm_loading.wait();
        throw;
    }
    fprintf(stderr, "Loading qscore model %s\n", qm.c_str());
    h->qm = qm_loading.get();
    h->delta = make_delta_table(h->em);
    h->thr = make_threshold_table(h->em, h->delta);
    // the reference's mean overshoot in errors for fragments it re-aligns whole
    // (py/tksm_badread.py:408-413); the explain-away coefficient of the error model is measured below
    h->calib_overshoot = std::isnan(h->cfg.calib_overshoot) ? 0.8f : (float)h->cfg.calib_overshoot;
    h->calib_kappa = std::isnan(h->cfg.calib_kappa) ? CALIB_KAPPA_UNMEASURED : (float)h->cfg.calib_kappa;
    h->calib_kappa_measured = NAN;
    compute_classes(h);
    upload_models(h);
    h->models_loaded = true;
    if (std::isnan(h->cfg.calib_kappa) && !h->cfg.host_only) {
        // kappa of THIS error model, measured by simulating random fragments with and without the alignment stage
        h->calib_kappa_measured = (float)self_calibrate(h);
        h->calib_kappa = h->calib_kappa_measured + (h->em.has_table ? CALIB_OFFSET_TABLE : CALIB_OFFSET_RANDOM);
        h->dt.calib_kappa = h->calib_kappa;
    }
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
    t->calib_kappa_measured = h->calib_kappa_measured;
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
        parser.capture_cb = h->front.needs_cb();
        HostBatch hb;
        parser.feed(mdf, mdf_len, true, hb);
        if (h->per_read_molecules()) {
            const HostBatch per_read = apply_read_transforms(h->tail, h->extras, h->front, h->cfg.seed, ordinal_base, {&hb}, nullptr);
            prepare_host(h, {&per_read}, b->prep);
        } else {
            prepare_host(h, {&hb}, b->prep);
        }
        b->prep.ordinal_base = ordinal_base;
        if (!h->cfg.host_only) {
            slot_bind(h, b.get());
            CK(cudaStreamSynchronize(h->stream));
        }
        b->prep.stage.reset();  // uploaded (or never needed): the batch keeps no host buffer
        b->prep.heap_blob.reset();
        b->prep.blob = nullptr;
    } catch (const std::exception& e) {
        h->err = e.what();
        return nullptr;
    }
    return b.release();
}

int tksm_seq_batch_run(tksm_seq* h, tksm_seq_batch* b, int which, int format, tksm_seq_stats* stats) {
    API_BEGIN
    need_device(h);
    slot_launch(h, b, which, format);
    slot_wait(h, b, stats);
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
    // text is parsed in 1 MiB chunks (TKSM_SEQ_CHUNK_BYTES overrides, tests); consecutive chunks are
    // grouped into device batches by the pipeline
    size_t first = (size_t)1 << 20, max_t = (size_t)1 << 20;   // measured: smaller first chunks change nothing (run-to-run spread of the call is +-4 ms)
    if (const char* e = getenv("TKSM_SEQ_CHUNK_BYTES")) first = max_t = std::max<size_t>((size_t)atoll(e), 1);
    MemorySource src(mdf, mdf_len, first, max_t);
    tk::engine_stream(h, src, which, format, ordinal_base, nullptr, out, out_len, stats);
    API_END(h)
}

int tksm_seq_shard_plan(tksm_seq* h, const char* mdf, size_t mdf_len, int world, uint64_t* byte_begin,
                        uint64_t* ordinal_base) {
    API_BEGIN
    std::vector<uint64_t> bb, ob;
    plan_shards(mdf, mdf_len, world, bb, ob);
    memcpy(byte_begin, bb.data(), bb.size() * 8);
    memcpy(ordinal_base, ob.data(), ob.size() * 8);
    API_END(h)
}

int tksm_seq_set_extras(tksm_seq* h, const tksm_seq_extras* x) {
    API_BEGIN
    ExtrasConfig e;
    if (x) {
        auto pct = [](double v, const char* what) {
            if (!(v >= 0.0 && v <= 100.0)) throw std::runtime_error(std::string("Error: ") + what + " must be between 0 and 100");
            return v / 100.0;
        };
        e.junk = pct(x->junk_reads, "--badread-junk-reads");
        e.random = pct(x->random_reads, "--badread-random-reads");
        e.chimera = pct(x->chimeras, "--badread-chimeras");
        if (e.junk + e.random > 1.0) throw std::runtime_error("Error: junk and random reads together exceed 100 percent");
        e.start_rate = pct(x->start_adapter_rate, "--badread-start-adapter rate");
        e.start_amount = pct(x->start_adapter_amount, "--badread-start-adapter amount");
        e.end_rate = pct(x->end_adapter_rate, "--badread-end-adapter rate");
        e.end_amount = pct(x->end_adapter_amount, "--badread-end-adapter amount");
        if (x->glitch_rate < 0 || x->glitch_size < 0 || x->glitch_skip < 0)
            throw std::runtime_error("Error: --badread-glitches values must not be negative");
        e.glitch_rate = x->glitch_rate;
        e.glitch_size = x->glitch_size;
        e.glitch_skip = x->glitch_skip;
        auto codes = [](const char* seq, const char* dflt) {
            std::vector<uint8_t> v;
            for (const char* p = seq ? seq : dflt; *p; p++) v.push_back(frag_code(ascii_upper((uint8_t)*p)));
            return v;
        };
        e.start_seq = codes(x->start_adapter_seq, "AATGTACTTCGTTCAGTTACGTATTGCT");
        e.end_seq = codes(x->end_adapter_seq, "GCAATACGTAACTGAACGAAGT");
    }
    h->extras = std::move(e);
    API_END(h)
}

int tksm_seq_set_front_end(tksm_seq* h, const char* spec) {
    API_BEGIN
    h->front = spec && *spec ? parse_front_end(spec) : FrontEnd();
    API_END(h)
}

int tksm_seq_tail_draw(tksm_seq* h, uint64_t ordinal, uint32_t frag_len, char* out, size_t cap, uint32_t* len) {
    API_BEGIN
    *len = 0;
    if (h->tail.loaded) {
        uint32_t start = 0;
        const uint32_t n = h->tail.draw_length(h->cfg.seed, ordinal, frag_len, &start);
        *len = n;
        if (n && n <= cap) {
            std::vector<uint8_t> codes(n);
            h->tail.draw_bases(h->cfg.seed, ordinal, start, n, codes.data());
            for (uint32_t i = 0; i < n; i++) out[i] = codes[i] < 4 ? "ACGT"[codes[i]] : (char)codes[i];
        }
    }
    API_END(h)
}

int tksm_seq_run_files(tksm_seq* h, const char* mdf_path, const char* badread_path, const char* perfect_path,
                       tksm_seq_stats* stats) {
    API_BEGIN
    if (!badread_path && !perfect_path) throw std::runtime_error("Must specify either --output or --perfect.");
    tk::run_files(h, mdf_path, badread_path, perfect_path, stats);
    API_END(h)
}

int tksm_seq_run_files_multi(tksm_seq* const* handles, int n_handles, const char* mdf_path, const char* badread_path,
                             const char* perfect_path, tksm_seq_stats* stats) {
    tksm_seq* h = n_handles > 0 ? handles[0] : nullptr;
    API_BEGIN
    if (n_handles < 1) throw std::runtime_error("no handles");
    if (!badread_path && !perfect_path) throw std::runtime_error("Must specify either --output or --perfect.");
    if (n_handles == 1) tk::run_files(h, mdf_path, badread_path, perfect_path, stats);
    else tk::run_files_multi(handles, n_handles, mdf_path, badread_path, perfect_path, stats);
    API_END(h)
}

}  // extern "C"

// ---------------------------------------------------------------- engine.hpp glue
namespace tk {

std::unique_ptr<ChunkSource> make_memory_source(const char* text, size_t len) {
    return std::unique_ptr<ChunkSource>(new MemorySource(text, len));
}
const GenomeStore* engine_genome(const tksm_seq* h) { return &h->genome; }
const tksm_seq_config* engine_config(const tksm_seq* h) { return &h->cfg; }

}  // namespace tk

// ---------------------------------------------------------------- record offsets without allocation
// tksm_seq_record_offsets_static (include/tksm_seq.h): the record-offset scan of the formatting stage (run_scan
// above: scan_block_sums -> scan_sums -> scan_final) for callers that may not allocate.  Its scratch, the per-block
// sums, lives in a file-scope device array instead of a DBuf; every device that loads the library holds the
// STATIC_SCAN_SCRATCH_BYTES of it until the process exits.
namespace {

constexpr uint32_t STATIC_SCAN_MAX_RECORDS = TKSM_SEQ_STATIC_SCAN_MAX_RECORDS;
constexpr uint32_t STATIC_SCAN_PER_BLOCK = tk::SCAN_BLOCK * tk::SCAN_ITEMS;
constexpr uint32_t STATIC_SCAN_MAX_BLOCKS = (STATIC_SCAN_MAX_RECORDS + STATIC_SCAN_PER_BLOCK - 1) / STATIC_SCAN_PER_BLOCK;
constexpr size_t STATIC_SCAN_SCRATCH_BYTES = (size_t)(STATIC_SCAN_MAX_BLOCKS + 2) * 8;   // 32 784 bytes
static_assert(STATIC_SCAN_SCRATCH_BYTES <= 64 * 1024, "the static scan scratch is meant to stay small");

// In host code this name is only a handle for cudaGetSymbolAddress, never a device address.
static __device__ __align__(256) uint64_t g_static_scan_sums[STATIC_SCAN_MAX_BLOCKS + 2];

constexpr int STATIC_SCAN_MAX_DEVICES = 64;
std::mutex g_static_scan_mu;
uint64_t* g_static_scan_ptr[STATIC_SCAN_MAX_DEVICES] = {};   // device address of the array on each device (set once)

// device address of the scratch on `device` (which must be current); looked up once per device
cudaError_t static_scan_scratch(int device, uint64_t** out) {
    std::lock_guard<std::mutex> g(g_static_scan_mu);
    if (!g_static_scan_ptr[device]) {
        uint64_t* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress((void**)&p, g_static_scan_sums);
        if (e != cudaSuccess) return e;
        g_static_scan_ptr[device] = p;
    }
    *out = g_static_scan_ptr[device];
    return cudaSuccess;
}

}  // namespace

extern "C" int tksm_seq_record_offsets_static(int device, const uint32_t* d_rec_len, uint32_t n_records, uint64_t* d_rec_off,
                                              uint64_t* d_total, void* cuda_stream) {
    auto fail = [](int code, const std::string& msg) {
        g_create_error = "tksm_seq_record_offsets_static: " + msg;
        return code;
    };
    if (n_records > STATIC_SCAN_MAX_RECORDS)
        return fail(2, std::to_string(n_records) + " records exceed the static scratch (at most " +
                           std::to_string(STATIC_SCAN_MAX_RECORDS) + " per call); split the call or use a batch");
    if (device < 0 || device >= STATIC_SCAN_MAX_DEVICES) return fail(2, "invalid device ordinal");
    if (!d_total || (n_records && (!d_rec_len || !d_rec_off))) return fail(2, "null pointer");
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) return fail(1, std::string("cudaSetDevice: ") + cudaGetErrorString(e));
    uint64_t* sums = nullptr;
    e = static_scan_scratch(device, &sums);
    if (e != cudaSuccess) return fail(1, std::string("cudaGetSymbolAddress: ") + cudaGetErrorString(e));
    cudaStream_t S = (cudaStream_t)cuda_stream;
    // Nothing in the scratch has to start at zero: the first kernel overwrites sums[0 .. nb) and the second one reads
    // exactly those entries (n_records == 0: none), whatever an earlier call left behind.
    const uint32_t nb = (n_records + STATIC_SCAN_PER_BLOCK - 1) / STATIC_SCAN_PER_BLOCK;
    if (nb) tk::scan_block_sums<<<nb, tk::SCAN_BLOCK, 0, S>>>(d_rec_len, n_records, sums);
    tk::scan_sums<<<1, 1024, 0, S>>>(sums, nb, d_total);
    if (nb) tk::scan_final<<<nb, tk::SCAN_BLOCK, 0, S>>>(d_rec_len, n_records, sums, d_rec_off);
    e = cudaGetLastError();
    if (e != cudaSuccess) return fail(1, std::string("kernel launch: ") + cudaGetErrorString(e));
    return 0;
}
