// Streaming runner: MDF file (or FIFO) -> batches -> kernels -> FASTQ/FASTA files.
// Replaces the __main__ loop of the reference (py/sequence.py:347-376) and its output-file
// handling (get_output_file, py/sequence.py:291-300).
#include <zlib.h>

#include <cstdio>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

#include "engine.hpp"

namespace tk {

namespace {

bool ends_with(const std::string& s, const char* suf) {
    size_t n = strlen(suf);
    return s.size() >= n && s.compare(s.size() - n, n, suf) == 0;
}

struct Sink {
    FILE* f = nullptr;
    gzFile gz = nullptr;
    bool fastq = false;
    void open(const std::string& path_in) {
        std::string path = path_in;
        if (ends_with(path, ".gz")) {
            gz = gzopen(path_in.c_str(), "wb");
            if (!gz) throw std::runtime_error("cannot open " + path_in + " for writing");
            path.resize(path.size() - 3);
        } else {
            f = fopen(path_in.c_str(), "wb");
            if (!f) throw std::runtime_error("cannot open " + path_in + " for writing");
        }
        fastq = ends_with(path, ".fastq") || ends_with(path, ".fq");
    }
    void write(const char* p, size_t n) {
        if (gz) {
            while (n) {
                unsigned chunk = (unsigned)std::min<size_t>(n, 1u << 30);
                if (gzwrite(gz, p, chunk) != (int)chunk) throw std::runtime_error("gzip write failed");
                p += chunk;
                n -= chunk;
            }
        } else if (f) {
            if (n && fwrite(p, 1, n, f) != n) throw std::runtime_error("write failed");
        }
    }
    void close() {
        if (gz) gzclose(gz);
        if (f) fclose(f);
        gz = nullptr;
        f = nullptr;
    }
    ~Sink() { close(); }
};

}  // namespace

bool path_is_fastq(const char* path) {
    std::string p = path;
    if (ends_with(p, ".gz")) p.resize(p.size() - 3);
    return ends_with(p, ".fastq") || ends_with(p, ".fq");
}

void run_files(tksm_seq* h, const char* mdf_path, const char* badread_path, const char* perfect_path,
               tksm_seq_stats* stats) {
    const tksm_seq_config* cfg = engine_config(h);
    const size_t max_reads = cfg->batch_reads ? (size_t)cfg->batch_reads : ((size_t)1 << 22);
    const uint64_t max_bases = cfg->batch_bases ? cfg->batch_bases : (1ull << 30);
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

    MdfParser parser(engine_genome(h));
    HostBatch hb;
    std::vector<char> buf, out_bad, out_perf;
    const size_t CHUNK = (size_t)64 << 20;
    size_t have = 0;
    bool eof = false;
    uint64_t ordinal = 0;
    tksm_seq_stats total{};
    auto flush = [&]() {
        if (hb.n_reads() == 0) {
            hb.clear();
            return;
        }
        hb.ordinal_base = ordinal;
        tksm_seq_stats s{};
        engine_run_batch(h, hb, which, format, badread_path ? &out_bad : nullptr, perfect_path ? &out_perf : nullptr, &s);
        ordinal += s.reads;
        bad.write(out_bad.data(), badread_path ? out_bad.size() : 0);
        perf.write(out_perf.data(), perfect_path ? out_perf.size() : 0);
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
        hb.clear();
    };
    while (true) {
        if (!eof) {
            buf.resize(have + CHUNK);
            size_t n = fread(buf.data() + have, 1, CHUNK, in);
            have += n;
            if (n == 0) eof = true;
        }
        size_t pos = 0;
        while (true) {
            size_t used = parser.feed(buf.data() + pos, have - pos, eof, hb, max_reads, max_bases);
            pos += used;
            if (hb.n_reads() >= max_reads || hb.total_raw_bases >= max_bases)
                flush();
            else
                break;
            if (used == 0) break;
        }
        memmove(buf.data(), buf.data() + pos, have - pos);
        have -= pos;
        if (eof) break;
    }
    flush();
    bad.close();
    perf.close();
    if (stats) *stats = total;
}

}  // namespace tk
