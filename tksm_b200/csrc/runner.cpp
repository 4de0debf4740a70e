// Streaming runner: MDF file (or FIFO) -> batches -> kernels -> FASTQ/FASTA files.
// Replaces the __main__ loop of the reference (py/sequence.py:347-376) and its output-file
// handling (get_output_file, py/sequence.py:291-300).
#include <atomic>
#include <zlib.h>

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <future>
#include <condition_variable>
#include <map>
#include <memory>
#include <mutex>
#include <stdexcept>
#include <string>
#include <thread>
#include <vector>

#include <sys/stat.h>

#include "engine.hpp"
#include "host.hpp"

namespace tk {

namespace {

bool ends_with(const std::string& s, const char* suf) {
    size_t n = strlen(suf);
    return s.size() >= n && s.compare(s.size() - n, n, suf) == 0;
}

// One gzip member per block, compressed by helper threads and written in order (a file of concatenated
// members is a valid gzip file: `zcat`, Python's gzip and htslib read it as one stream).  Replaces the
// gzip.open(path, "wt") sink of the reference (py/sequence.py:291-296), which deflates on one core.
struct GzBlocks {
    FILE* f = nullptr;
    size_t block = (size_t)4 << 20;
    size_t max_inflight = 4;
    std::string pending;
    std::deque<std::future<std::string>> inflight;
    bool wrote_any = false;

    static std::string compress(std::string raw) {
        z_stream zs{};
        if (deflateInit2(&zs, Z_DEFAULT_COMPRESSION, Z_DEFLATED, 15 + 16, 8, Z_DEFAULT_STRATEGY) != Z_OK)
            throw std::runtime_error("gzip init failed");
        std::string out(deflateBound(&zs, (uLong)raw.size()) + 64, 0);
        zs.next_in = (Bytef*)raw.data();
        zs.avail_in = (uInt)raw.size();
        zs.next_out = (Bytef*)out.data();
        zs.avail_out = (uInt)out.size();
        const int rc = deflate(&zs, Z_FINISH);
        const size_t n = out.size() - zs.avail_out;
        deflateEnd(&zs);
        if (rc != Z_STREAM_END) throw std::runtime_error("gzip write failed");
        out.resize(n);
        return out;
    }
    void open(const std::string& path) {
        f = fopen(path.c_str(), "wb");
        if (!f) throw std::runtime_error("cannot open " + path + " for writing");
        if (const char* e = getenv("TKSM_GZ_BLOCK_BYTES")) block = std::max<size_t>((size_t)atoll(e), 1);
        block = std::min<size_t>(block, (size_t)1 << 30);
        max_inflight = std::max<size_t>(std::thread::hardware_concurrency(), 2);
    }
    void put(const std::string& member) {
        if (!member.empty() && fwrite(member.data(), 1, member.size(), f) != member.size())
            throw std::runtime_error("write failed");
        wrote_any = true;
    }
    void submit(std::string raw) {
        while (inflight.size() >= max_inflight) {
            put(inflight.front().get());
            inflight.pop_front();
        }
        inflight.push_back(std::async(std::launch::async, compress, std::move(raw)));
    }
    void write(const char* p, size_t n) {
        while (n) {
            const size_t take = std::min(n, block - pending.size());
            pending.append(p, take);
            p += take;
            n -= take;
            if (pending.size() >= block) {
                std::string raw;
                raw.swap(pending);
                submit(std::move(raw));
            }
        }
    }
    void finish() {
        if (!f) return;
        if (!pending.empty()) {
            std::string raw;
            raw.swap(pending);
            submit(std::move(raw));
        }
        while (!inflight.empty()) {
            put(inflight.front().get());
            inflight.pop_front();
        }
        if (!wrote_any) put(compress(std::string()));  // an empty input still gives a valid .gz
        FILE* g = f;
        f = nullptr;
        if (fclose(g) != 0) throw std::runtime_error("write failed");
    }
    ~GzBlocks() {
        for (auto& fu : inflight)
            if (fu.valid()) fu.wait();
        if (f) fclose(f);
    }
};

struct Sink {
    FILE* f = nullptr;
    GzBlocks gz;
    bool is_gz = false;
    bool fastq = false;
    void open(const std::string& path_in) {
        std::string path = path_in;
        if (ends_with(path, ".gz")) {
            gz.open(path_in);
            is_gz = true;
            path.resize(path.size() - 3);
        } else {
            f = fopen(path_in.c_str(), "wb");
            if (!f) throw std::runtime_error("cannot open " + path_in + " for writing");
        }
        fastq = ends_with(path, ".fastq") || ends_with(path, ".fq");
    }
    void write(const char* p, size_t n) {
        if (is_gz) {
            gz.write(p, n);
        } else if (f) {
            if (n && fwrite(p, 1, n, f) != n) throw std::runtime_error("write failed");
        }
    }
    void close() {
        if (is_gz) gz.finish();
        FILE* g = f;
        f = nullptr;
        if (g && fclose(g) != 0) throw std::runtime_error("write failed");   // e.g. ENOSPC at the final flush
    }
    ~Sink() {
        if (f) fclose(f);
    }
};

}  // namespace

bool path_is_fastq(const char* path) {
    std::string p = path;
    if (ends_with(p, ".gz")) p.resize(p.size() - 3);
    return ends_with(p, ".fastq") || ends_with(p, ".fq");
}

namespace {

// Reads the MDF (file or FIFO, never seeks) in blocks and hands out chunks cut at record
// boundaries.  A record larger than the block size grows the chunk until its end is seen.
struct FileSource : ChunkSource {
    FILE* in;
    std::string carry;   // bytes after the last record boundary of the previous block
    bool eof = false;
    size_t target = (size_t)1 << 20, max_target = (size_t)1 << 20;
    explicit FileSource(FILE* f) : in(f) {
        if (const char* e = getenv("TKSM_SEQ_CHUNK_BYTES")) target = max_target = std::max<size_t>((size_t)atoll(e), 1);
    }
    bool next(TextChunk& tc) override {
        while (true) {
            if (eof && carry.empty()) return false;
            auto buf = std::make_shared<std::string>();
            buf->swap(carry);
            while (!eof) {
                size_t have = buf->size();
                buf->resize(have + target);
                size_t n = fread(&(*buf)[have], 1, target, in);
                buf->resize(have + n);
                if (n == 0) {
                    eof = true;
                    break;
                }
                // last record boundary inside the buffer ("\n+"); everything before it is complete
                size_t cut = buf->rfind("\n+");
                if (cut != std::string::npos && cut > 0) {
                    carry.assign(*buf, cut + 1, std::string::npos);
                    buf->resize(cut + 1);
                    break;
                }
            }
            target = std::min(target * 2, max_target);
            if (buf->empty()) continue;
            tc.p = buf->data();
            tc.n = buf->size();
            tc.keep = buf;
            return true;
        }
    }
};

struct FileSinks : OutputSink {
    Sink* s[2];
    void write(int o, const char* p, size_t n) override {
        if (s[o]) s[o]->write(p, n);
    }
};

}  // namespace

void run_files(tksm_seq* h, const char* mdf_path, const char* badread_path, const char* perfect_path,
               tksm_seq_stats* stats) {
    Sink bad, perf;
    // the output files exist even when the MDF is empty (py/sequence.py:347, 356)
    if (badread_path) bad.open(badread_path);
    if (perfect_path) perf.open(perfect_path);
    int which = (badread_path ? TKSM_SEQ_BADREAD : 0) | (perfect_path ? TKSM_SEQ_PERFECT : 0);
    int format = (bad.fastq ? TKSM_SEQ_FASTQ_BADREAD : 0) | (perf.fastq ? TKSM_SEQ_FASTQ_PERFECT : 0);

    FILE* in = fopen(mdf_path, "rb");
    if (!in) throw std::runtime_error(std::string("cannot open ") + mdf_path);
    struct Closer {
        FILE* f;
        ~Closer() { fclose(f); }
    } closer{in};

    FileSource src(in);
    FileSinks sinks;
    sinks.s[0] = badread_path ? &bad : nullptr;
    sinks.s[1] = perfect_path ? &perf : nullptr;
    tksm_seq_stats total{};
    engine_stream(h, src, which, format, 0, &sinks, nullptr, nullptr, &total);
    bad.close();
    perf.close();
    if (stats) *stats = total;
}

namespace {

// records of one shard held in memory until the shards before it are on disk
struct MemorySinks : OutputSink {
    std::vector<std::string> parts[2];
    void write(int o, const char* p, size_t n) override { parts[o].emplace_back(p, n); }
};

}  // namespace

namespace {

// --gpus N on a stream: pieces of the MDF (complete records, TKSM_SEQ_DEAL_BYTES each, 64 MiB by default) go to the
// devices round-robin; piece k carries the global ordinal of its first read (a light pass over its header lines
// counts the reads), so the records are those of a one-GPU run, and they are written in piece order.
void deal_stream(tksm_seq* const* handles, int n, const char* mdf_path, int which, int format, OutputSink& out, tksm_seq_stats* stats) {
    struct Piece {
        std::string text;
        uint64_t ordinal = 0;
        MemorySinks sink;
        tksm_seq_stats st{};
        std::string error;
        bool done = false;
    };
    const size_t piece_bytes = getenv("TKSM_SEQ_DEAL_BYTES") ? std::max<size_t>((size_t)atoll(getenv("TKSM_SEQ_DEAL_BYTES")), 1) : ((size_t)64 << 20);
    std::mutex mu;
    std::condition_variable cv;
    std::vector<std::deque<std::shared_ptr<Piece>>> queue((size_t)n);   // per device: pieces waiting or running
    std::map<size_t, std::shared_ptr<Piece>> finished;                  // by piece index
    std::vector<int> pending((size_t)n, 0);
    bool closed = false;
    std::atomic<bool> failed{false};   // read outside the lock by the reader loop and the workers
    std::string first_error;

    std::vector<std::thread> workers;
    for (int i = 0; i < n; i++)
        workers.emplace_back([&, i]() {
            while (true) {
                std::shared_ptr<Piece> p;
                {
                    std::unique_lock<std::mutex> g(mu);
                    cv.wait(g, [&] { return !queue[(size_t)i].empty() || closed; });
                    if (queue[(size_t)i].empty()) return;
                    p = queue[(size_t)i].front();
                    queue[(size_t)i].pop_front();
                }
                try {
                    if (!failed) {
                        auto src = make_memory_source(p->text.data(), p->text.size());
                        engine_stream(handles[i], *src, which, format, p->ordinal, &p->sink, nullptr, nullptr, &p->st);
                    }
                } catch (const std::exception& e) {
                    p->error = e.what()[0] ? e.what() : "error";
                }
                p->text.clear();
                p->text.shrink_to_fit();
                {
                    std::lock_guard<std::mutex> g(mu);
                    p->done = true;
                    pending[(size_t)i]--;
                    if (!p->error.empty()) failed = true;
                }
                cv.notify_all();
            }
        });

    tksm_seq_stats total{};
    size_t dealt = 0, written = 0;
    std::map<size_t, std::shared_ptr<Piece>> inflight;
    auto drain = [&](bool all) {   // write finished pieces in order; with `all`, wait for every piece
        while (written < dealt) {
            std::shared_ptr<Piece> p = inflight[written];
            {
                std::unique_lock<std::mutex> g(mu);
                if (!all && !p->done) return;
                cv.wait(g, [&] { return p->done; });
            }
            if (!p->error.empty() && first_error.empty()) first_error = p->error;
            if (first_error.empty()) {
                for (int o = 0; o < 2; o++)
                    for (const std::string& part : p->sink.parts[o]) out.write(o, part.data(), part.size());
                const tksm_seq_stats& s = p->st;
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
            inflight.erase(written);
            written++;
        }
    };

    FILE* in = strcmp(mdf_path, "-") == 0 ? stdin : fopen(mdf_path, "rb");
    if (!in) {
        {
            std::lock_guard<std::mutex> g(mu);
            closed = true;
        }
        cv.notify_all();
        for (auto& t : workers) t.join();
        throw std::runtime_error(std::string("cannot open ") + mdf_path);
    }
    std::string carry;
    std::vector<char> buf((size_t)4 << 20);
    uint64_t ordinal = 0;
    bool eof = false;
    size_t want = piece_bytes;   // grows while one record is larger than a piece
    while (!eof && !failed) {
        // fill up to a piece, cut at the last record boundary
        while (carry.size() < want) {
            const size_t got = fread(buf.data(), 1, buf.size(), in);
            if (got == 0) {
                eof = true;
                break;
            }
            carry.append(buf.data(), got);
        }
        std::string piece;
        if (eof) {
            piece.swap(carry);
        } else {
            size_t cut = carry.rfind("\n+");
            if (cut == std::string::npos || cut == 0) {   // one record larger than a piece: read another piece's worth
                want = carry.size() + piece_bytes;
                continue;
            }
            piece.assign(carry, 0, cut + 1);
            carry.erase(0, cut + 1);
            want = piece_bytes;
        }
        if (piece.empty()) continue;
        std::vector<uint64_t> bb, ob;
        plan_shards(piece.data(), piece.size(), 1, bb, ob);
        auto p = std::make_shared<Piece>();
        p->text.swap(piece);
        p->ordinal = ordinal;
        ordinal += ob[1];
        const size_t dev = dealt % (size_t)n;
        {
            std::unique_lock<std::mutex> g(mu);
            cv.wait(g, [&] { return pending[dev] < 2 || failed; });
            queue[dev].push_back(p);
            pending[dev]++;
        }
        inflight[dealt++] = p;
        cv.notify_all();
        drain(false);
    }
    if (in != stdin) fclose(in);
    {
        std::lock_guard<std::mutex> g(mu);
        closed = true;
    }
    cv.notify_all();
    drain(true);
    for (auto& t : workers) t.join();
    if (getenv("TKSM_SEQ_TRACE")) fprintf(stderr, "[trace] dealt %zu pieces to %d devices\n", dealt, n);
    if (!first_error.empty()) throw std::runtime_error(first_error);
    if (stats) *stats = total;
}

}  // namespace

void run_files_multi(tksm_seq* const* handles, int n, const char* mdf_path, const char* badread_path,
                     const char* perfect_path, tksm_seq_stats* stats) {
    Sink bad, perf;
    if (badread_path) bad.open(badread_path);
    if (perfect_path) perf.open(perfect_path);
    const int which = (badread_path ? TKSM_SEQ_BADREAD : 0) | (perfect_path ? TKSM_SEQ_PERFECT : 0);
    const int format = (bad.fastq ? TKSM_SEQ_FASTQ_BADREAD : 0) | (perf.fastq ? TKSM_SEQ_FASTQ_PERFECT : 0);

    // A stream that cannot be read twice (FIFO, stdin; $TKSM_SEQ_DEAL=1 forces it for files) is not slurped: pieces
    // of complete records are dealt to the devices round-robin as they arrive, each with the global ordinal of its
    // first read, and written back in order (at most two pieces per device in flight).  So is a regular file above
    // $TKSM_SEQ_PLAN_MAX_BYTES (2 GiB): the shard plan needs the whole text in memory and every shard but the first
    // keeps its records there until its turn, which a 100 M-molecule input cannot afford; with that many pieces
    // round-robin balances as well as the plan does.  $TKSM_SEQ_DEAL=0 keeps the plan.
    {
        struct stat sb{};
        const char* deal_env = getenv("TKSM_SEQ_DEAL");
        const bool known = stat(mdf_path, &sb) == 0;
        const bool fifo = known && !S_ISREG(sb.st_mode);
        const uint64_t plan_max = getenv("TKSM_SEQ_PLAN_MAX_BYTES") ? (uint64_t)atoll(getenv("TKSM_SEQ_PLAN_MAX_BYTES")) : (2ull << 30);
        const bool big = known && S_ISREG(sb.st_mode) && (uint64_t)sb.st_size > plan_max;
        if (((fifo || big) && !(deal_env && atoi(deal_env) == 0)) || (deal_env && atoi(deal_env) == 1)) {
            FileSinks direct;
            direct.s[0] = badread_path ? &bad : nullptr;
            direct.s[1] = perfect_path ? &perf : nullptr;
            deal_stream(handles, n, mdf_path, which, format, direct, stats);
            bad.close();
            perf.close();
            return;
        }
    }
    // the shard plan needs the whole text
    std::string text;
    {
        FILE* in = fopen(mdf_path, "rb");
        if (!in) throw std::runtime_error(std::string("cannot open ") + mdf_path);
        // straight into a string of the file's size (a regular file here: streams are dealt above); whatever a
        // growing file adds is appended
        struct stat sb{};
        if (fstat(fileno(in), &sb) == 0 && S_ISREG(sb.st_mode) && sb.st_size > 0) {
            text.resize((size_t)sb.st_size);
            size_t have = 0, got;
            while (have < text.size() && (got = fread(&text[have], 1, text.size() - have, in)) > 0) have += got;
            text.resize(have);
        }
        std::vector<char> buf((size_t)4 << 20);
        size_t got;
        while ((got = fread(buf.data(), 1, buf.size(), in)) > 0) text.append(buf.data(), got);
        fclose(in);
    }
    std::vector<uint64_t> bb, ob;
    plan_shards(text.data(), text.size(), n, bb, ob);

    FileSinks direct;
    direct.s[0] = badread_path ? &bad : nullptr;
    direct.s[1] = perfect_path ? &perf : nullptr;
    std::vector<MemorySinks> held((size_t)n);
    std::vector<tksm_seq_stats> st((size_t)n);
    std::vector<std::string> errors((size_t)n);
    std::vector<std::thread> threads;
    for (int i = 0; i < n; i++) {
        threads.emplace_back([&, i]() {
            try {
                st[(size_t)i] = tksm_seq_stats{};
                if (bb[(size_t)i + 1] == bb[(size_t)i]) return;
                auto src = make_memory_source(text.data() + bb[(size_t)i], (size_t)(bb[(size_t)i + 1] - bb[(size_t)i]));
                OutputSink* sink = i == 0 ? (OutputSink*)&direct : (OutputSink*)&held[(size_t)i];  // shard 0 streams to the files
                engine_stream(handles[i], *src, which, format, ob[(size_t)i], sink, nullptr, nullptr, &st[(size_t)i]);
            } catch (const std::exception& e) {
                errors[(size_t)i] = e.what()[0] ? e.what() : "error";
            }
        });
    }
    tksm_seq_stats total{};
    std::string first_error;
    for (int i = 0; i < n; i++) {
        threads[(size_t)i].join();
        if (!errors[(size_t)i].empty() && first_error.empty()) first_error = errors[(size_t)i];
        if (!first_error.empty()) continue;  // keep joining; nothing more is written
        for (int o = 0; o < 2; o++) {
            for (const std::string& part : held[(size_t)i].parts[o]) direct.write(o, part.data(), part.size());
            held[(size_t)i].parts[o].clear();
            held[(size_t)i].parts[o].shrink_to_fit();
        }
        const tksm_seq_stats& s = st[(size_t)i];
        total.reads += s.reads;
        total.raw_bases += s.raw_bases;
        total.read_bases += s.read_bases;
        total.out_bytes[0] += s.out_bytes[0];
        total.out_bytes[1] += s.out_bytes[1];
        total.algorithmic_bytes += s.algorithmic_bytes;
        total.h2d_bytes += s.h2d_bytes;
        total.d2h_bytes += s.d2h_bytes;
        total.kernel_ms = std::max(total.kernel_ms, s.kernel_ms);      // the devices run side by side
        total.simulate_ms = std::max(total.simulate_ms, s.simulate_ms);
        total.kernel_launches += s.kernel_launches;
        total.status_or |= s.status_or;
    }
    if (!first_error.empty()) throw std::runtime_error(first_error);
    bad.close();
    perf.close();
    if (stats) *stats = total;
}

}  // namespace tk
