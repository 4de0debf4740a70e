// Internal glue between the device engine (engine.cu) and the host runner (runner.cpp).
#pragma once
#include <vector>

#include "../../include/tksm_seq.h"
#include "host.hpp"

namespace tk {

// Runs one molecule batch through the kernels and copies the requested outputs to the host.
// `hb` is moved in and handed back so its buffers can be reused.
void engine_run_batch(tksm_seq* h, HostBatch& hb, int which, int format, std::vector<char>* out_badread,
                      std::vector<char>* out_perfect, tksm_seq_stats* stats);
const GenomeStore* engine_genome(const tksm_seq* h);
const tksm_seq_config* engine_config(const tksm_seq* h);

// runner.cpp: streaming MDF file -> output files (py/sequence.py:347-376)
void run_files(tksm_seq* h, const char* mdf_path, const char* badread_path, const char* perfect_path,
               tksm_seq_stats* stats);

}  // namespace tk
