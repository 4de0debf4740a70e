// Streaming runner: MDF file (or FIFO) -> batches -> kernels -> FASTQ/FASTA files.
// Replaces the __main__ loop of the reference (py/sequence.py:347-376) and its output-file
// handling (get_output_file, py/sequence.py:291-300).
#include <zlib.h>

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
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

}  // namespace tk
