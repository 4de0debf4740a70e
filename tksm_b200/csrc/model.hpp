// Host-side model loading and packing for the Seq hot path.
//
// Replaces (reference, /root/reference):
//   ErrorModel.load_from_file / align_kmers        py/tksm_badread.py:91-117, 146-197
//   QScoreModel ctor / load_from_file / builtins   py/tksm_badread.py:464-582, 600-605
//   Identities / beta_parameters                   py/tksm_badread.py:702-757
//   set_tksm_models_dicts                          py/sequence.py:17-31
#pragma once
#include <cstdint>
#include <map>
#include <string>
#include <vector>

#include "common.hpp"

namespace tk {

struct ErrorModelHost {
    int k = 1;
    bool has_table = false;
    std::vector<uint16_t> q16;       // 4^k exact change probabilities * 65536 (host side: tests, class building)
    std::vector<uint8_t> qcls;       // 4^k change-probability class (device table)
    std::vector<double> qrep;        // Q_CLASSES representative probabilities; [255] = 1 (k-mer not in the table)
    std::vector<uint32_t> alias;
    std::vector<uint64_t> alts;
    // model averages per trial over uniformly random k-mers (drive the epoch-length table)
    double rate_change = 1.0;   // P(trial changes something)
    double rate_loss = 2.0 / 3; // E[# slots whose original base is lost]
    double rate_ins = 1.0 / 3;  // E[# inserted bases]
    double rate_slots = 1.0;    // E[# changed slots]
};

struct QScoreModelHost {
    int kq = 1;
    std::vector<uint32_t> direct;
    uint32_t direct_off[5] = {0, 0, 0, 0, 0};
    std::vector<QHashEntry> hash;
    uint32_t hash_shift = 63, hash_mask = 1;
    std::vector<uint32_t> rows;  // pairs (alias offset, log2 columns)
    std::vector<uint32_t> alias;
    std::vector<uint32_t> t9, t9d1, t9d2;  // closure tables of the 9-wide fast path (empty when kq != 9)
    std::vector<std::string> row_cigar;  // debug / tests
};

struct IdentityHost {
    bool constant = false;
    float mean = 0.84f;
    double beta_a = 0, beta_b = 0, max_identity = 1;
    std::vector<float> tab;  // IDENT_TAB_BINS + 1
};

// Throwing std::runtime_error on malformed input; callers translate to return codes.
std::string read_text_maybe_gz(const std::string& path);

// NW (unit cost) alignment with path; cigar ops over "=XID", I consumes query, D consumes target.
// Traceback from the end preferring diagonal, then I, then D (pinned rule).
std::string nw_cigar(const std::string& query, const std::string& target);

// py/tksm_badread.py:146-197
std::vector<std::string> align_kmers(const std::string& kmer, const std::string& alt);

ErrorModelHost load_error_model(const std::string& name_or_path,
                                const std::map<std::string, std::string>& names);
QScoreModelHost load_qscore_model(const std::string& name_or_path,
                                  const std::map<std::string, std::string>& names);
QScoreModelHost pack_qscore_model(const std::vector<std::pair<std::string, std::vector<std::pair<int, double>>>>& dists);
IdentityHost make_identities(double mean_pct, double stdev_pct, double max_pct);
std::vector<float> make_delta_table(const ErrorModelHost& em);
// [DELTA_TAB_BINS][Q_CLASSES][THR_PER_CLASS] thresholds: P(Poisson(qrep[c] * delta[b]) >= m) * 65536, m = 1..8
std::vector<uint16_t> make_threshold_table(const ErrorModelHost& em, const std::vector<float>& delta);

// py/sequence.py:17-31 ($TKSM_MODELS lookup; first directory in the list wins on a clash)
struct ModelNames {
    std::map<std::string, std::string> error, qscore, tail;
};
ModelNames scan_model_dirs(const char* env_value);

uint64_t qkey_from_cigar(const std::string& cigar);  // 0 when not encodable (> 31 symbols)

}  // namespace tk
