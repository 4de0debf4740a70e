// `tksm sequence` command line: same flags, defaults, messages and exit codes as the argparse
// front-end of the reference (py/sequence.py:34-165), in front of the C ABI.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "../../include/tksm_seq.h"
#include "sequence.h"

namespace {

struct Opt {
    const char* dest;     // argparse dest, printed by --list
    const char* long_;    // long flag without dashes
    char short_;          // 0 if none
    int nargs;            // 0 flag, 1 one value, -1 one or more
};

const Opt OPTS[] = {
    {"help", "help", 'h', 0},
    {"input", "input", 'i', 1},
    {"references", "references", 'r', -1},
    {"badread", "badread", 'o', 1},
    {"perfect", "perfect", 0, 1},
    {"skip_qual_compute", "skip-qual-compute", 0, 0},
    {"output_format", "output-format", 'O', 1},
    {"threads", "threads", 't', 1},
    {"badread_identity", "badread-identity", 0, 1},
    {"badread_error_model", "badread-error-model", 0, 1},
    {"badread_qscore_model", "badread-qscore-model", 0, 1},
    {"badread_tail_model", "badread-tail-model", 0, 1},
    {"list", "list", 0, 0},
    // extensions without a reference counterpart
    {"seed", "seed", 's', 1},
    {"gpus", "gpus", 0, 1},
    {"device", "device", 0, 1},
    // Badread extras (include/tksm_seq.h: tksm_seq_extras), all off unless given
    {"badread_junk_reads", "badread-junk-reads", 0, 1},
    {"badread_random_reads", "badread-random-reads", 0, 1},
    {"badread_chimeras", "badread-chimeras", 0, 1},
    {"badread_start_adapter", "badread-start-adapter", 0, 1},
    {"badread_end_adapter", "badread-end-adapter", 0, 1},
    {"badread_start_adapter_seq", "badread-start-adapter-seq", 0, 1},
    {"badread_end_adapter_seq", "badread-end-adapter-seq", 0, 1},
    {"badread_glitches", "badread-glitches", 0, 1},
    // fused front-end steps (include/tksm_seq.h: tksm_seq_set_front_end)
    {"fuse", "fuse", 0, 1},
};
const int N_OPTS = sizeof(OPTS) / sizeof(OPTS[0]);

void usage(FILE* f) {
    fprintf(f,
            "usage: sequence [-h] -i INPUT [-r REFERENCES [REFERENCES ...]] [-o BADREAD] [--perfect PERFECT]\n"
            "                [--skip-qual-compute] [-O {fastq,fasta}] [-t THREADS]\n"
            "                [--badread-identity BADREAD_IDENTITY] [--badread-error-model BADREAD_ERROR_MODEL]\n"
            "                [--badread-qscore-model BADREAD_QSCORE_MODEL] [--badread-tail-model BADREAD_TAIL_MODEL]\n"
            "                [--list] [-s SEED] [--gpus GPUS] [--device DEVICE]\n"
            "                [--badread-junk-reads PCT] [--badread-random-reads PCT] [--badread-chimeras PCT]\n"
            "                [--badread-start-adapter RATE,AMOUNT] [--badread-end-adapter RATE,AMOUNT]\n"
            "                [--badread-start-adapter-seq SEQ] [--badread-end-adapter-seq SEQ]\n"
            "                [--badread-glitches RATE,SIZE,SKIP] [--fuse 'plA:normal=15,7.5;Flp:0.5;Tag:FMT5,FMT3;SCB']\n");
}

// comma-separated numbers of one extras flag into the fields they set; false: not such a flag or not numbers
bool extras_arg(const std::string& d, const std::string& val, tksm_seq_extras& x, std::string& start_seq, std::string& end_seq,
                bool& have) {
    auto numbers = [&](size_t count, double* out) {
        size_t pos = 0;
        for (size_t i = 0; i < count; i++) {
            size_t comma = val.find(',', pos);
            if ((i + 1 < count) != (comma != std::string::npos)) return false;
            const std::string tok = val.substr(pos, comma == std::string::npos ? std::string::npos : comma - pos);
            char* e = nullptr;
            out[i] = strtod(tok.c_str(), &e);
            if (tok.empty() || *e) return false;
            pos = comma + 1;
        }
        return true;
    };
    double v[3];
    bool ok = false;
    if (d == "badread_junk_reads") { if ((ok = numbers(1, v))) x.junk_reads = v[0]; }
    else if (d == "badread_random_reads") { if ((ok = numbers(1, v))) x.random_reads = v[0]; }
    else if (d == "badread_chimeras") { if ((ok = numbers(1, v))) x.chimeras = v[0]; }
    else if (d == "badread_start_adapter") { if ((ok = numbers(2, v))) { x.start_adapter_rate = v[0]; x.start_adapter_amount = v[1]; } }
    else if (d == "badread_end_adapter") { if ((ok = numbers(2, v))) { x.end_adapter_rate = v[0]; x.end_adapter_amount = v[1]; } }
    else if (d == "badread_glitches") { if ((ok = numbers(3, v))) { x.glitch_rate = v[0]; x.glitch_size = v[1]; x.glitch_skip = v[2]; } }
    else if (d == "badread_start_adapter_seq") { start_seq = val; ok = true; }
    else if (d == "badread_end_adapter_seq") { end_seq = val; ok = true; }
    have = have || ok;
    return ok;
}

int usage_error(const std::string& msg) {
    usage(stderr);
    fprintf(stderr, "sequence: error: %s\n", msg.c_str());
    return 2;
}

}  // namespace

extern "C" int tksm_sequence_main(int argc, char** argv) {
    std::string input, badread, perfect, identity = "84.0,99.0,5.5", err_model, qs_model, tail_model, out_fmt;
    tksm_seq_extras extras{};
    std::string start_seq, end_seq;
    bool have_extras = false;
    std::string fuse_spec;
    std::vector<std::string> refs;
    bool have_err = false, have_qs = false, have_tail = false, skip_qual = false, have_input = false;
    long threads = 1;
    bool threads_given = false;
    long long seed = 42;
    int device = 0;
    long gpus = getenv("TKSM_GPUS") ? atol(getenv("TKSM_GPUS")) : 1;   // extension: $TKSM_GPUS / --gpus
    for (int a = 1; a < argc; a++) {
        std::string arg = argv[a];
        const Opt* opt = nullptr;
        std::string inline_val;
        bool has_inline = false;
        if (arg.size() >= 3 && arg[0] == '-' && arg[1] == '-') {
            std::string name = arg.substr(2);
            size_t eq = name.find('=');
            if (eq != std::string::npos) {
                inline_val = name.substr(eq + 1);
                name = name.substr(0, eq);
                has_inline = true;
            }
            // exact match, else unique prefix (argparse allow_abbrev)
            int hits = 0;
            for (int i = 0; i < N_OPTS; i++)
                if (name == OPTS[i].long_) {
                    opt = &OPTS[i];
                    hits = 1;
                    break;
                }
            if (!opt)
                for (int i = 0; i < N_OPTS; i++)
                    if (strncmp(OPTS[i].long_, name.c_str(), name.size()) == 0) {
                        opt = &OPTS[i];
                        hits++;
                    }
            if (hits > 1) return usage_error("ambiguous option: --" + name);
            if (!opt) return usage_error("unrecognized arguments: " + arg);
        } else if (arg.size() >= 2 && arg[0] == '-' && arg[1] != '-') {
            for (int i = 0; i < N_OPTS; i++)
                if (OPTS[i].short_ && OPTS[i].short_ == arg[1]) opt = &OPTS[i];
            if (!opt) return usage_error("unrecognized arguments: " + arg);
            if (arg.size() > 2) {
                inline_val = arg.substr(2);
                has_inline = true;
            }
        } else {
            return usage_error("unrecognized arguments: " + arg);
        }
        std::vector<std::string> vals;
        if (opt->nargs == 0) {
            if (has_inline) return usage_error(std::string("ignored explicit argument '") + inline_val + "'");
        } else if (has_inline) {
            vals.push_back(inline_val);
        } else if (opt->nargs == 1) {
            if (a + 1 >= argc) return usage_error(std::string("argument --") + opt->long_ + ": expected one argument");
            vals.push_back(argv[++a]);
        } else {
            while (a + 1 < argc && !(argv[a + 1][0] == '-' && argv[a + 1][1] != 0)) vals.push_back(argv[++a]);
            if (vals.empty())
                return usage_error(std::string("argument --") + opt->long_ + ": expected at least one argument");
        }
        std::string d = opt->dest;
        if (d == "help") {
            usage(stdout);
            printf(
                "\nSequencer module of tksm. Generates FASTQ/A file from MDF (molecule description file) and "
                "reference FASTA files. Note: tksm looks for sequencing model names in colon sperated paths in "
                "$TKSM_MODELS environment variable. Current $TKSM_MODELS value: %s\n",
                getenv("TKSM_MODELS") ? getenv("TKSM_MODELS") : "None");
            return 0;
        } else if (d == "list") {
            for (int i = 0; i < N_OPTS; i++) printf("%s\n", OPTS[i].dest);  // py/sequence.py:116-122
            return 0;
        } else if (d == "input") {
            input = vals[0];
            have_input = true;
        } else if (d == "references") {
            refs = vals;
        } else if (d == "badread") {
            badread = vals[0];
        } else if (d == "perfect") {
            perfect = vals[0];
        } else if (d == "skip_qual_compute") {
            skip_qual = true;
        } else if (d == "output_format") {
            if (vals[0] != "fastq" && vals[0] != "fasta")
                return usage_error("argument -O/--output-format: invalid choice: '" + vals[0] +
                                   "' (choose from 'fastq', 'fasta')");
            out_fmt = vals[0];  // parsed and never used, like the reference
        } else if (d == "threads") {
            char* e = nullptr;
            threads = strtol(vals[0].c_str(), &e, 10);
            threads_given = true;
            if (*e) return usage_error("argument -t/--threads: invalid int value: '" + vals[0] + "'");
        } else if (d == "badread_identity") {
            identity = vals[0];
        } else if (d == "badread_error_model") {
            err_model = vals[0];
            have_err = true;
        } else if (d == "badread_qscore_model") {
            qs_model = vals[0];
            have_qs = true;
        } else if (d == "badread_tail_model") {
            tail_model = vals[0];
            have_tail = true;
        } else if (d == "fuse") {
            fuse_spec = vals[0];
        } else if (d.rfind("badread_", 0) == 0 && extras_arg(d, vals[0], extras, start_seq, end_seq, have_extras)) {
            // (parsed by extras_arg)
        } else if (d.rfind("badread_", 0) == 0 && (d == "badread_junk_reads" || d == "badread_random_reads" || d == "badread_chimeras" ||
                                                   d == "badread_start_adapter" || d == "badread_end_adapter" || d == "badread_glitches")) {
            return usage_error("argument --" + std::string(d).replace(0, 8, "badread-") + ": invalid value: '" + vals[0] + "'");
        } else if (d == "seed") {
            char* e = nullptr;
            seed = strtoll(vals[0].c_str(), &e, 10);
            if (*e) return usage_error("argument -s/--seed: invalid int value: '" + vals[0] + "'");
        } else if (d == "gpus") {
            char* e = nullptr;
            gpus = strtol(vals[0].c_str(), &e, 10);
            if (*e) return usage_error("argument --gpus: invalid int value: '" + vals[0] + "'");
        } else if (d == "device") {
            device = atoi(vals[0].c_str());
        }
    }
    if (!have_input) return usage_error("the following arguments are required: -i/--input");

    // py/sequence.py:134-164
    double idp[3];
    {
        int n = 0;
        size_t pos = 0;
        bool bad = false;
        while (true) {
            size_t c = identity.find(',', pos);
            std::string item = identity.substr(pos, c == std::string::npos ? std::string::npos : c - pos);
            char* e = nullptr;
            double v = strtod(item.c_str(), &e);
            while (e && *e == ' ') e++;
            if (item.empty() || (e && *e)) bad = true;
            if (n < 3) idp[n] = v;
            n++;
            if (c == std::string::npos) break;
            pos = c + 1;
        }
        if (bad) {
            fprintf(stderr, "Error: could not parse --identity values\n");
            return 1;
        }
        if (n != 3) {
            fprintf(stderr, "AssertionError: Must specify 3 values for --badread-identity\n");
            return 1;
        }
    }
    const double mean = idp[0], mx = idp[1], sd = idp[2];
    auto die = [](const std::string& m) {
        fprintf(stderr, "%s\n", m.c_str());
        return 1;
    };
    if (mean > 100.0) return die("Error: mean read identity cannot be more than 100");
    if (mx > 100.0) return die("Error: max read identity cannot be more than 100");
    if (mean <= 50) return die("Error: mean read identity must be at least 50");
    if (mx <= 50) return die("Error: max read identity must be at least 50");
    if (mean > mx) {
        char b[256];
        snprintf(b, sizeof b, "Error: mean identity (%g) cannot be larger than max identity (%g)", mean, mx);
        return die(b);
    }
    if (sd < 0.0) return die("Error: read identity stdev cannot be negative");
    if (badread.empty() && perfect.empty()) return usage_error("Must specify either --output or --perfect.");
    if (refs.empty()) return die("TypeError: 'NoneType' object is not iterable (no -r/--references given)");

    // --gpus N: devices device .. device + N - 1, or the comma list in $TKSM_GPU_LIST (one entry per shard;
    // an ordinal may repeat, which runs several shards on one device)
    std::vector<int> devices;
    if (const char* lst = getenv("TKSM_GPU_LIST")) {
        for (const char* p = lst; *p;) {
            char* e = nullptr;
            long v = strtol(p, &e, 10);
            if (e == p) return die("Error: could not parse $TKSM_GPU_LIST");
            devices.push_back((int)v);
            p = *e == ',' ? e + 1 : e;
        }
    }
    if (gpus < 1 || gpus > 64) return die("Error: --gpus must be between 1 and 64");
    if (devices.empty())
        for (long g = 0; g < gpus; g++) devices.push_back(device + (int)g);
    else if (gpus > 1)
        devices.resize((size_t)gpus, devices.back());
    const int n = (int)devices.size();
    // -t sizes the host parser pool here (the reference forks that many workers, default 1): without -t every hardware
    // thread is used (0), split between the GPUs
    const long host_threads = !threads_given ? 0 : threads > 0 ? std::min<long>(threads, 256) : 1;

    std::vector<const char*> paths;
    for (auto& r : refs) paths.push_back(r.c_str());
    std::vector<tksm_seq*> handles((size_t)n, nullptr);
    std::vector<std::string> errors((size_t)n);
    auto set_up = [&](int i) {
        tksm_seq_config cfg;
        tksm_seq_default_config(&cfg);
        cfg.device = devices[(size_t)i];
        cfg.seed = (uint64_t)seed;
        cfg.identity_mean = mean;
        cfg.identity_max = mx;
        cfg.identity_stdev = sd;
        cfg.skip_qual_compute = skip_qual;
        const long hw = std::min<long>(std::max<long>((long)std::thread::hardware_concurrency(), 1), 32L * n);
        cfg.threads = (int)std::max<long>((host_threads == 0 ? hw : host_threads) / n, 1);
        tksm_seq* h = tksm_seq_create(&cfg);
        if (!h) {
            errors[(size_t)i] = tksm_seq_last_error(nullptr);
            return;
        }
        handles[(size_t)i] = h;
        int rc = tksm_seq_load_references(h, paths.data(), (int)paths.size());
        if (!rc && !badread.empty())
            rc = tksm_seq_load_models(h, have_err ? err_model.c_str() : nullptr, have_qs ? qs_model.c_str() : nullptr,
                                      have_tail ? tail_model.c_str() : nullptr);
        else if (!rc)
            rc = tksm_seq_load_models(h, "random", "random", nullptr);  // perfect only: models are not consulted
        if (!rc && have_extras) {
            tksm_seq_extras x = extras;
            x.start_adapter_seq = start_seq.empty() ? nullptr : start_seq.c_str();
            x.end_adapter_seq = end_seq.empty() ? nullptr : end_seq.c_str();
            rc = tksm_seq_set_extras(h, &x);
        }
        if (!rc && !fuse_spec.empty()) rc = tksm_seq_set_front_end(h, fuse_spec.c_str());
        if (rc) errors[(size_t)i] = tksm_seq_last_error(h);
    };
    if (n == 1) {
        set_up(0);
    } else {
        std::vector<std::thread> th;
        for (int i = 0; i < n; i++) th.emplace_back(set_up, i);
        for (auto& t : th) t.join();
    }
    int rc = 0;
    for (int i = 0; i < n && !rc; i++)
        if (!errors[(size_t)i].empty()) {
            fprintf(stderr, "Error: %s\n", errors[(size_t)i].c_str());
            rc = 1;
        }
    tksm_seq_stats st;
    if (!rc) {
        rc = tksm_seq_run_files_multi(handles.data(), n, input.c_str(), badread.empty() ? nullptr : badread.c_str(),
                                      perfect.empty() ? nullptr : perfect.c_str(), &st);
        if (rc) {
            fprintf(stderr, "Error: %s\n", tksm_seq_last_error(handles[0]));
        } else {
            fprintf(stderr, "Sequencing: %llu reads, %llu bases, %.1f ms on device%s\n", (unsigned long long)st.reads,
                    (unsigned long long)st.read_bases, st.kernel_ms, n > 1 ? " (slowest of the devices)" : "");
        }
    }
    for (tksm_seq* h : handles)
        if (h) tksm_seq_destroy(h);
    return rc ? 1 : 0;
}

class Sequencer_module::impl {
public:
    int argc;
    char** argv;
    impl(int c, char** v) : argc(c), argv(v) {}
};
Sequencer_module::Sequencer_module(int argc, char** argv) : pimpl{std::make_unique<impl>(argc, argv)} {}
Sequencer_module::~Sequencer_module() = default;
int Sequencer_module::run() { return tksm_sequence_main(pimpl->argc, pimpl->argv); }
