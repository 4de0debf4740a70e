// Internal glue between the device engine (engine.cu) and the host runner (runner.cpp).
#pragma once
#include <memory>
#include <vector>

#include "../../include/tksm_seq.h"
#include "host.hpp"

namespace tk {

// A run of complete MDF records (starts at a '+' header line or at the beginning of the input).
struct TextChunk {
    const char* p = nullptr;
    size_t n = 0;
    std::shared_ptr<void> keep;  // owner of the bytes, when the source does not outlive the pipeline
};

struct ChunkSource {
    virtual ~ChunkSource() = default;
    virtual bool next(TextChunk& out) = 0;                  // chunks in MDF order; false at end of input
    virtual uint64_t total_bytes_hint() const { return 0; }  // whole input size when known (arena estimate)
};

struct OutputSink {
    virtual ~OutputSink() = default;
    virtual void write(int output /* 0 badread, 1 perfect */, const char* p, size_t n) = 0;  // called in MDF order
};

// engine.cu: the streaming pipeline (host parser threads -> upload -> kernels -> download)
void engine_stream(tksm_seq* h, ChunkSource& src, int which, int format, uint64_t ordinal_base, OutputSink* sink,
                   char** arena_out, size_t* arena_len, tksm_seq_stats* stats);
// chunker over MDF text in memory (the caller keeps the text alive)
std::unique_ptr<ChunkSource> make_memory_source(const char* text, size_t len);
const GenomeStore* engine_genome(const tksm_seq* h);
const tksm_seq_config* engine_config(const tksm_seq* h);

// runner.cpp: streaming MDF file -> output files (py/sequence.py:347-376)
void run_files(tksm_seq* h, const char* mdf_path, const char* badread_path, const char* perfect_path,
               tksm_seq_stats* stats);
// one process, several devices: contiguous shards of the MDF, one handle and one thread each,
// outputs written in shard order == MDF order
void run_files_multi(tksm_seq* const* handles, int n, const char* mdf_path, const char* badread_path,
                     const char* perfect_path, tksm_seq_stats* stats);

}  // namespace tk
