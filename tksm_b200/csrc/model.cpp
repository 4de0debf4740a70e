// Host-side model loading and packing (see model.hpp for the reference mapping).
#include "model.hpp"

#include <sys/stat.h>
#include <dirent.h>
#include <zlib.h>

#include <algorithm>
#include <exception>
#include <atomic>
#include <mutex>
#include <thread>
#include <cmath>
#include <cstring>
#include <stdexcept>

namespace tk {

std::string read_text_maybe_gz(const std::string& path) {
    // gzopen reads plain files transparently, which mirrors get_open_func's magic-byte
    // sniffing (py/tksm_badread.py:259-290).
    // A plain regular file (a reference FASTA of gigabytes) is read straight into a string of its size.
    {
        struct stat sb{};
        FILE* plain = fopen(path.c_str(), "rb");
        if (!plain) throw std::runtime_error("cannot open " + path);
        unsigned char magic[2] = {0, 0};
        const size_t got = fread(magic, 1, 2, plain);
        const bool gz = got == 2 && magic[0] == 0x1f && magic[1] == 0x8b;
        if (!gz && fstat(fileno(plain), &sb) == 0 && S_ISREG(sb.st_mode)) {
            std::string out((size_t)sb.st_size, '\0');
            rewind(plain);
            size_t have = 0;
            while (have < out.size()) {
                const size_t n = fread(out.data() + have, 1, out.size() - have, plain);
                if (n == 0) break;
                have += n;
            }
            const bool bad = ferror(plain) != 0;
            // a file that grew meanwhile: the rest
            char extra[4096];
            size_t n;
            while (!bad && have == out.size() && (n = fread(extra, 1, sizeof extra, plain)) > 0) {
                out.append(extra, n);
                have += n;
            }
            fclose(plain);
            if (bad) throw std::runtime_error("read error in " + path);
            out.resize(have);
            return out;
        }
        fclose(plain);
    }
    gzFile f = gzopen(path.c_str(), "rb");
    if (!f) throw std::runtime_error("cannot open " + path);
    std::string out;
    std::vector<char> buf(1 << 20);
    for (;;) {
        int n = gzread(f, buf.data(), (unsigned)buf.size());
        if (n < 0) {
            gzclose(f);
            throw std::runtime_error("read error in " + path);
        }
        if (n == 0) break;
        out.append(buf.data(), (size_t)n);
    }
    gzclose(f);
    return out;
}

std::string nw_cigar(const std::string& q, const std::string& t) {
    const int n = (int)q.size(), m = (int)t.size();
    std::vector<int> d((size_t)(n + 1) * (m + 1));
    auto D = [&](int i, int j) -> int& { return d[(size_t)i * (m + 1) + j]; };
    for (int i = 0; i <= n; i++) D(i, 0) = i;
    for (int j = 0; j <= m; j++) D(0, j) = j;
    for (int i = 1; i <= n; i++)
        for (int j = 1; j <= m; j++) {
            int c = D(i - 1, j - 1) + (q[i - 1] != t[j - 1]);
            c = std::min(c, D(i - 1, j) + 1);
            c = std::min(c, D(i, j - 1) + 1);
            D(i, j) = c;
        }
    std::string ops;
    int i = n, j = m;
    while (i > 0 || j > 0) {
        if (i > 0 && j > 0 && D(i, j) == D(i - 1, j - 1) + (q[i - 1] != t[j - 1])) {
            ops.push_back(q[i - 1] == t[j - 1] ? '=' : 'X');
            i--, j--;
        } else if (i > 0 && D(i, j) == D(i - 1, j) + 1) {
            ops.push_back('I');
            i--;
        } else {
            ops.push_back('D');
            j--;
        }
    }
    std::reverse(ops.begin(), ops.end());
    std::string cigar;
    for (size_t a = 0; a < ops.size();) {
        size_t b = a;
        while (b < ops.size() && ops[b] == ops[a]) b++;
        cigar += std::to_string(b - a);
        cigar.push_back(ops[a]);
        a = b;
    }
    return cigar;
}

std::vector<std::string> align_kmers(const std::string& kmer_in, const std::string& alt_in) {
    if (kmer_in.size() <= 2 || alt_in.size() <= 1) throw std::runtime_error("align_kmers: sizes");
    if (kmer_in.front() != alt_in.front() || kmer_in.back() != alt_in.back())
        throw std::runtime_error("align_kmers: alternative does not keep the end bases of " + kmer_in);
    const size_t k = kmer_in.size();
    std::vector<std::string> result(k);
    result[0] = std::string(1, kmer_in.front());
    result[k - 1] = std::string(1, kmer_in.back());
    std::string kmer = kmer_in.substr(1, k - 2), alt = alt_in.substr(1, alt_in.size() - 2);
    std::string cigar = alt.empty() ? std::to_string(kmer.size()) + "D" : nw_cigar(alt, kmer);
    size_t kp = 0, ap = 0;
    for (size_t c = 0; c < cigar.size();) {
        size_t e = c;
        while (isdigit((unsigned char)cigar[e])) e++;
        int sz = atoi(cigar.substr(c, e - c).c_str());
        char ty = cigar[e];
        c = e + 1;
        if (ty == '=' || ty == 'X') {
            for (int x = 0; x < sz; x++) result[kp++ + 1] = std::string(1, alt[ap++]);
        } else if (ty == 'D') {
            for (int x = 0; x < sz; x++) result[kp++ + 1] = "";
        } else {
            result[kp] += alt.substr(ap, (size_t)sz);
            ap += (size_t)sz;
        }
    }
    if (result[0].size() == 2) {
        char ins = result[0][1];
        result[0].resize(1);
        result[1] = std::string(1, ins) + result[1];
    }
    return result;
}

namespace {

inline int base_code(char c) {
    switch (c) {
        case 'A': return 0;
        case 'C': return 1;
        case 'G': return 2;
        case 'T': return 3;
        default: return -1;
    }
}

struct AliasCol {
    uint32_t thr, a, b;
};

// Vose alias tables over `ncols` columns with `bits`-bit thresholds.
std::vector<AliasCol> build_alias(const std::vector<double>& p_in, int ncols, int bits) {
    const size_t n = p_in.size();
    double tot = 0;
    for (double v : p_in) tot += v;
    std::vector<double> s((size_t)ncols, 0.0);
    for (size_t i = 0; i < n; i++) s[i] = p_in[i] / tot * ncols;
    std::vector<AliasCol> cols((size_t)ncols);
    std::vector<int> small, large;
    for (int i = ncols - 1; i >= 0; i--) (s[(size_t)i] < 1.0 ? small : large).push_back(i);
    const double full = (double)(1u << bits);
    while (!small.empty() && !large.empty()) {
        int l = small.back();
        small.pop_back();
        int g = large.back();
        cols[(size_t)l] = {(uint32_t)std::min(full, std::floor(s[(size_t)l] * full + 0.5)), (uint32_t)l, (uint32_t)g};
        s[(size_t)g] = (s[(size_t)g] + s[(size_t)l]) - 1.0;
        if (s[(size_t)g] < 1.0) {
            large.pop_back();
            small.push_back(g);
        }
    }
    for (int g : large) cols[(size_t)g] = {(uint32_t)full, (uint32_t)g, (uint32_t)g};
    for (int l : small) cols[(size_t)l] = {(uint32_t)full, (uint32_t)l, (uint32_t)l};
    for (auto& c : cols) {
        // a threshold of 2^bits does not fit the field: fold it into a == b
        if (c.thr >= (1u << bits)) {
            c.thr = 0;
            c.b = c.a;
        }
        if (c.thr == 0) c.a = c.b;
    }
    return cols;
}

}  // namespace

// Groups the 4^k change probabilities into Q_CLASSES - 1 classes of equal width between the smallest and the largest
// one (representative = class mean, so the total trial rate is preserved and a k-mer's arrival rate is its own to
// within half a class width, about 0.1 % of a trial) so that the arrival test of the kernels is two shared-memory
// look-ups: class of the k-mer, threshold of the class.  Class 255 is reserved for k-mers that are not in the table
// (q = 1).
static void build_q_classes(ErrorModelHost& em) {
    em.qrep.assign(Q_CLASSES, 1.0);
    if (!em.has_table) return;
    const size_t rows = em.q16.size();
    std::vector<std::pair<uint16_t, uint32_t>> order(rows);
    for (size_t c = 0; c < rows; c++) order[c] = {em.q16[c], (uint32_t)c};
    std::sort(order.begin(), order.end());
    const int K = Q_CLASSES - 1;
    const double vmin = order.front().first, span = (double)order.back().first - vmin + 1.0;
    std::vector<int> assign(rows, 0);
    for (size_t i = 0; i < rows; i++) assign[i] = std::min(K - 1, (int)(((double)order[i].first - vmin) / span * K));
    em.qcls.assign(rows, 0);
    std::vector<double> sum((size_t)K, 0.0);
    std::vector<size_t> cnt((size_t)K, 0);
    for (size_t i = 0; i < rows; i++) {
        em.qcls[order[i].second] = (uint8_t)assign[i];
        sum[(size_t)assign[i]] += order[i].first;
        cnt[(size_t)assign[i]]++;
    }
    for (int j = 0; j < K; j++) em.qrep[(size_t)j] = cnt[(size_t)j] ? sum[(size_t)j] / (double)cnt[(size_t)j] / 65536.0 : 0.0;
}

ErrorModelHost load_error_model(const std::string& name_or_path, const std::map<std::string, std::string>& names) {
    ErrorModelHost em;
    if (name_or_path == "random") {  // k = 1, every trial is a random change
        build_q_classes(em);
        return em;
    }
    std::string path = name_or_path;
    auto it = names.find(name_or_path);
    if (it != names.end()) path = it->second;
    std::string text = read_text_maybe_gz(path);

    int k = 0;
    size_t rows = 0;
    double s_change = 0, s_loss = 0, s_ins = 0, s_slots = 0;
    size_t pos = 0;
    while (pos < text.size()) {
        size_t eol = text.find('\n', pos);
        if (eol == std::string::npos) eol = text.size();
        std::string line = text.substr(pos, eol - pos);
        pos = eol + 1;
        while (!line.empty() && isspace((unsigned char)line.back())) line.pop_back();
        size_t ls = 0;
        while (ls < line.size() && isspace((unsigned char)line[ls])) ls++;
        line = line.substr(ls);
        if (line.empty()) continue;
        std::vector<std::pair<std::string, double>> alts;
        for (size_t a = 0; a < line.size();) {
            size_t b = line.find(';', a);
            if (b == std::string::npos) b = line.size();
            if (b > a) {
                std::string item = line.substr(a, b - a);
                size_t comma = item.find(',');
                if (comma == std::string::npos) throw std::runtime_error("bad error model entry: " + item);
                alts.emplace_back(item.substr(0, comma), atof(item.c_str() + comma + 1));
            }
            a = b + 1;
        }
        if (alts.empty()) continue;
        const std::string& kmer = alts[0].first;
        if (k == 0) {
            k = (int)kmer.size();
            if (k < 3 || k > MAX_K)
                throw std::runtime_error("error model k-mer size " + std::to_string(k) + " unsupported (3.." +
                                         std::to_string(MAX_K) + ")");
            rows = (size_t)1 << (2 * k);
            em.k = k;
            em.has_table = true;
            em.q16.assign(rows, 65535);
            em.alias.assign(rows * ERR_ALIAS_COLS,
                            (uint32_t)(ERR_OUTCOME_RANDOM << 22) | (uint32_t)(ERR_OUTCOME_RANDOM << 27));
            em.alts.assign(rows * ERR_ALT_STRIDE, 0);
        } else if ((int)kmer.size() != k) {
            throw std::runtime_error("error model: k-mers of different sizes");
        }
        uint32_t code = 0;
        bool valid = true;
        for (int j = 0; j < k; j++) {
            int c = base_code(kmer[(size_t)j]);
            if (c < 0) valid = false;
            code |= (uint32_t)(c & 3) << (2 * j);
        }
        if (!valid) continue;  // cannot be looked up by an ACGT fragment either

        double W = 0;
        for (auto& a : alts) W += a.second;
        double r = 1.0 - W;
        if (!(r > 0.0)) r = 0.0;
        double total = W + r, p_nc = 0;
        std::vector<double> probs;     // changing outcomes, then random
        std::vector<uint64_t> words;
        double e_loss = 0, e_ins = 0, e_slots = 0;
        for (auto& a : alts) {
            std::vector<std::string> slots = align_kmers(kmer, a.first);
            uint64_t w = 0;
            int nb = 0, nloss = 0, nins = 0, nchanged = 0;
            for (int j = 0; j < k; j++) {
                const std::string& s = slots[(size_t)j];
                if ((int)s.size() > MAX_SLOT_LEN)
                    throw std::runtime_error("error model: alternative " + a.first + " of " + kmer +
                                             " puts more than 5 bases in one slot");
                w |= (uint64_t)s.size() << (3 * j);
                bool changed = !(s.size() == 1 && s[0] == kmer[(size_t)j]);
                if (changed) {
                    w |= 1ull << (21 + j);
                    nchanged++;
                    bool kept = s.size() >= 2 && (s.front() == kmer[(size_t)j] || s.back() == kmer[(size_t)j]);
                    if (!kept) nloss++;
                    if (s.size() >= 2) nins += (int)s.size() - 1;
                }
                for (char ch : s) {
                    int c = base_code(ch);
                    if (c < 0) throw std::runtime_error("error model: non-ACGT alternative for " + kmer);
                    if (nb >= 18) throw std::runtime_error("error model: alternative too long for " + kmer);
                    w |= (uint64_t)c << (28 + 2 * nb);
                    nb++;
                }
            }
            if (nchanged == 0) {
                p_nc += a.second;
            } else {
                probs.push_back(a.second);
                words.push_back(w);
                e_loss += a.second * nloss;
                e_ins += a.second * nins;
                e_slots += a.second * nchanged;
            }
        }
        if (words.size() > ERR_OUTCOME_RANDOM) throw std::runtime_error("error model: too many alternatives for " + kmer);
        double q = 1.0 - p_nc / total;
        // random change: 1/3 sub (loss), 1/3 ins (1 base), 1/3 del (loss)
        e_loss += r * (2.0 / 3.0);
        e_ins += r * (1.0 / 3.0);
        e_slots += r;
        s_change += q;
        s_loss += e_loss / total;
        s_ins += e_ins / total;
        s_slots += e_slots / total;
        em.q16[code] = (uint16_t)std::min(65535.0, std::floor(q * 65536.0 + 0.5));
        std::vector<double> cond = probs;
        std::vector<uint32_t> outcome((size_t)probs.size());
        for (size_t x = 0; x < outcome.size(); x++) outcome[x] = (uint32_t)x;
        if (r > 0.0 || cond.empty()) {
            cond.push_back(r > 0.0 ? r : 1.0);
            outcome.push_back(ERR_OUTCOME_RANDOM);
        }
        std::vector<AliasCol> cols = build_alias(cond, ERR_ALIAS_COLS, 22);
        for (int c = 0; c < ERR_ALIAS_COLS; c++) {
            const AliasCol& ac = cols[(size_t)c];
            uint32_t oa = ac.a < outcome.size() ? outcome[ac.a] : outcome[ac.b];
            uint32_t ob = ac.b < outcome.size() ? outcome[ac.b] : oa;
            em.alias[(size_t)code * ERR_ALIAS_COLS + (size_t)c] = ac.thr | (oa << 22) | (ob << 27);
        }
        for (size_t x = 0; x < words.size(); x++) em.alts[(size_t)code * ERR_ALT_STRIDE + x] = words[x];
    }
    if (k == 0) throw std::runtime_error("error model file is empty: " + path);
    // k-mers absent from the file keep q = 1 / random change (py/tksm_badread.py:127-128);
    // they count as such in the averages.
    size_t listed = 0;
    for (size_t c = 0; c < rows; c++) listed += em.q16[c] != 65535 || (em.alias[c * ERR_ALIAS_COLS] >> 27) != ERR_OUTCOME_RANDOM;
    double missing = (double)(rows - std::min(rows, listed));
    em.rate_change = (s_change + missing) / (double)rows;
    em.rate_loss = (s_loss + missing * 2.0 / 3.0) / (double)rows;
    em.rate_ins = (s_ins + missing / 3.0) / (double)rows;
    em.rate_slots = (s_slots + missing) / (double)rows;
    build_q_classes(em);
    return em;
}

// Packed key of a D-bearing window cigar: ops o_0..o_{w-1} ('=' 0, 'X' 1, 'I' 2; w <= 9) and the
// lengths d_1..d_{w-1} of the D runs in front of ops 1..w-1 (<= 15):
//   key = (0x10 | w) << 56 | (sum o_i << 2 i) << 32 | sum d_i << 4 (i - 1)
// 0 when the cigar cannot occur as a window (leading/trailing D, more than 9 ops, a run over 15).
uint64_t qkey_from_cigar(const std::string& cigar) {
    uint64_t ops = 0, gaps = 0;
    uint32_t w = 0, run = 0;
    for (char c : cigar) {
        uint64_t o;
        switch (c) {
            case '=': o = OP_EQ; break;
            case 'X': o = OP_X; break;
            case 'I': o = OP_I; break;
            case 'D':
                if (w == 0) return 0;
                run++;
                continue;
            default: return 0;
        }
        if (w >= 9 || run > 15) return 0;
        if (w > 0) gaps |= (uint64_t)run << (4 * (w - 1));
        ops |= o << (2 * w);
        w++;
        run = 0;
    }
    if (w == 0 || run) return 0;
    return ((uint64_t)(0x10u | w) << 56) | (ops << 32) | gaps;
}

QScoreModelHost pack_qscore_model(
    const std::vector<std::pair<std::string, std::vector<std::pair<int, double>>>>& dists) {
    QScoreModelHost qm;
    std::map<std::string, uint32_t> row_of;
    for (auto& d : dists) {
        const std::string& cigar = d.first;
        int kk = 0;
        for (char c : cigar) kk += c != 'D';
        if (kk > qm.kq) qm.kq = kk;
        std::vector<double> p;
        std::vector<uint32_t> qs;
        for (auto& sp : d.second) {
            if (sp.first < 0 || sp.first > 93) throw std::runtime_error("qscore model: score out of range in " + cigar);
            if (sp.second > 0) {
                qs.push_back((uint32_t)sp.first);
                p.push_back(sp.second);
            }
        }
        if (p.empty()) {
            // random.choices raises on all-zero weights; keep the row unusable in the same way
            throw std::runtime_error("qscore model: no positive weight for " + cigar);
        }
        int lg = 0;
        while ((1u << lg) < p.size()) lg++;
        if (lg > 7) throw std::runtime_error("qscore model: more than 128 scores for " + cigar);
        // one 16-bit uniform per read base: lg bits pick the column, 16 - lg bits are the fraction;
        // thresholds are stored left-aligned in 16 bits
        std::vector<AliasCol> cols = build_alias(p, 1 << lg, 16 - lg);
        uint32_t off = (uint32_t)qm.alias.size();
        for (auto& ac : cols) {
            uint32_t qa = ac.a < qs.size() ? qs[ac.a] : qs[ac.b];
            uint32_t qb = ac.b < qs.size() ? qs[ac.b] : qa;
            qm.alias.push_back((ac.thr << lg) | (qa << 16) | (qb << 23));
        }
        // later duplicates override earlier ones, as dict assignment does
        auto it = row_of.find(cigar);
        uint32_t row;
        if (it == row_of.end()) {
            row = (uint32_t)qm.row_cigar.size();
            row_of[cigar] = row;
            qm.row_cigar.push_back(cigar);
            qm.rows.push_back(off);
            qm.rows.push_back((uint32_t)lg);
        } else {
            row = it->second;
            qm.rows[2 * row] = off;
            qm.rows[2 * row + 1] = (uint32_t)lg;
        }
    }
    for (const char* need : {"=", "X", "I"})
        if (!row_of.count(need)) throw std::runtime_error(std::string("qscore model lacks the '") + need + "' cigar");
    if (qm.kq > 9 || qm.kq % 2 == 0)
        throw std::runtime_error("qscore model k-mer size " + std::to_string(qm.kq) + " unsupported (odd, <= 9)");

    // direct maps for D-free windows, with the trim-by-2 fallback already resolved
    // (py/tksm_badread.py:584-598)
    const int widths[5] = {1, 3, 5, 7, 9};
    const char sym[3] = {'=', 'X', 'I'};
    for (int wi = 0; wi < 5; wi++) {
        int w = widths[wi];
        qm.direct_off[wi] = (uint32_t)qm.direct.size();
        if (w > qm.kq) continue;
        uint32_t count = 1;
        for (int x = 0; x < w; x++) count *= 3;
        for (uint32_t code = 0; code < count; code++) {
            std::string s((size_t)w, '=');
            uint32_t c = code;
            for (int x = 0; x < w; x++) {
                s[(size_t)x] = sym[c % 3];
                c /= 3;
            }
            while (!row_of.count(s)) s = s.substr(1, s.size() - 2);
            qm.direct.push_back(row_of[s]);
        }
    }
    // fast-path closures for 9-wide windows: no D, or a single D run of 1..3 columns
    if (qm.kq == 9) {
        auto ent = [&](const std::string& start) {
            std::string s = start;
            while (!row_of.count(s)) {
                s = s.substr(1, s.size() - 2);
                size_t a = s.find_first_not_of('D'), b = s.find_last_not_of('D');
                s = s.substr(a, b - a + 1);
            }
            const uint32_t row = row_of.find(s)->second;   // (read-only: the closures are filled by several threads)
            return (qm.rows[2 * row] << 3) | qm.rows[2 * row + 1];
        };
        // a million closure lookups through the cigar map, about a second on one thread: the 19 683 windows are
        // independent and write their own entries
        auto for_each_code = [](auto&& fn) {
            const unsigned nt = std::min(16u, std::max(1u, std::thread::hardware_concurrency()));
            std::atomic<uint32_t> next{0};
            std::exception_ptr err;
            std::mutex mu;
            auto work = [&]() {
                try {
                    for (uint32_t c0; (c0 = next.fetch_add(128)) < 19683u;)
                        for (uint32_t c = c0; c < std::min(19683u, c0 + 128); c++) fn(c);
                } catch (...) {
                    std::lock_guard<std::mutex> g(mu);
                    if (!err) err = std::current_exception();
                }
            };
            std::vector<std::thread> th;
            for (unsigned t = 1; t < nt; t++) th.emplace_back(work);
            work();
            for (auto& t : th) t.join();
            if (err) std::rethrow_exception(err);
        };
        qm.t9.resize(19683);
        qm.t9d1.assign((size_t)19683 * 24, 0);
        for_each_code([&](uint32_t code) {
            std::string s(9, '=');
            uint32_t c = code;
            for (int x = 0; x < 9; x++) {
                s[(size_t)x] = sym[c % 3];
                c /= 3;
            }
            qm.t9[code] = ent(s);
            for (int g = 0; g < 8; g++)
                for (int dc = 1; dc <= 3; dc++) {
                    std::string d = s.substr(0, (size_t)g + 1) + std::string((size_t)dc, 'D') + s.substr((size_t)g + 1);
                    qm.t9d1[code + (size_t)19683 * (3 * g + dc - 1)] = ent(d);
                }
        });
        // two single-column deletions (the most frequent windows the one-run table misses)
        qm.t9d2.assign((size_t)19683 * 28, 0);
        for_each_code([&](uint32_t code) {
            std::string s(9, '=');
            uint32_t c = code;
            for (int x = 0; x < 9; x++) {
                s[(size_t)x] = sym[c % 3];
                c /= 3;
            }
            for (int g1 = 0; g1 < 8; g1++)
                for (int g2 = g1 + 1; g2 < 8; g2++) {
                    std::string d = s.substr(0, (size_t)g1 + 1) + "D" + s.substr((size_t)g1 + 1, (size_t)(g2 - g1)) + "D" +
                                    s.substr((size_t)g2 + 1);
                    qm.t9d2[code + (size_t)19683 * (size_t)(g1 * (15 - g1) / 2 + g2 - g1 - 1)] = ent(d);
                }
        });
    }
    // hash of D-bearing cigars
    size_t nD = 0;
    for (auto& kv : row_of) nD += kv.first.find('D') != std::string::npos;
    uint32_t lg = 4;
    while (((size_t)1 << lg) < 2 * nD + 8) lg++;
    qm.hash.assign((size_t)1 << lg, QHashEntry{0, 0, 0});
    qm.hash_shift = 64 - lg;
    qm.hash_mask = (1u << lg) - 1;
    for (auto& kv : row_of) {
        if (kv.first.find('D') == std::string::npos) continue;
        if (kv.first.front() == 'D' || kv.first.back() == 'D') continue;  // unreachable (py:643-644)
        uint64_t key = qkey_from_cigar(kv.first);
        if (!key) continue;
        uint32_t h = (uint32_t)((key * 0x9E3779B97F4A7C15ull) >> qm.hash_shift);
        while (qm.hash[h].key) h = (h + 1) & qm.hash_mask;
        qm.hash[h] = QHashEntry{key, kv.second, 0};
    }
    return qm;
}

QScoreModelHost load_qscore_model(const std::string& name_or_path, const std::map<std::string, std::string>& names) {
    std::vector<std::pair<std::string, std::vector<std::pair<int, double>>>> dists;
    auto uniform = [](int lo, int hi) {
        std::vector<std::pair<int, double>> v;
        for (int q = lo; q <= hi; q++) v.emplace_back(q, 1.0 / (hi - lo + 1));
        return v;
    };
    if (name_or_path == "random") {  // py/tksm_badread.py:487-497
        for (const char* c : {"=", "X", "I"}) dists.emplace_back(c, uniform(1, 20));
        return pack_qscore_model(dists);
    }
    if (name_or_path == "ideal") {  // py/tksm_badread.py:499-544
        dists.emplace_back("X", uniform(1, 3));
        dists.emplace_back("I", uniform(1, 3));
        dists.emplace_back("=", uniform(4, 7));
        dists.emplace_back("===", uniform(8, 20));
        dists.emplace_back("=====", uniform(21, 30));
        dists.emplace_back("=======", uniform(31, 40));
        dists.emplace_back("=========", uniform(41, 50));
        return pack_qscore_model(dists);
    }
    std::string path = name_or_path;
    auto it = names.find(name_or_path);
    if (it != names.end()) path = it->second;
    std::string text = read_text_maybe_gz(path);
    size_t pos = 0;
    while (pos < text.size()) {
        size_t eol = text.find('\n', pos);
        if (eol == std::string::npos) eol = text.size();
        std::string line = text.substr(pos, eol - pos);
        pos = eol + 1;
        while (!line.empty() && isspace((unsigned char)line.back())) line.pop_back();
        size_t ls = 0;
        while (ls < line.size() && isspace((unsigned char)line[ls])) ls++;
        line = line.substr(ls);
        std::vector<std::string> parts;
        for (size_t a = 0;;) {
            size_t b = line.find(';', a);
            if (b == std::string::npos) {
                parts.push_back(line.substr(a));
                break;
            }
            parts.push_back(line.substr(a, b - a));
            a = b + 1;
        }
        if (parts[0] == "overall") continue;
        if (parts.size() < 3) throw std::runtime_error(path + " does not seem to be a valid qscore model file");
        std::vector<std::pair<int, double>> sp;
        const std::string& body = parts[2];
        for (size_t a = 0; a < body.size();) {
            size_t b = body.find(',', a);
            if (b == std::string::npos) b = body.size();
            if (b > a) {
                std::string item = body.substr(a, b - a);
                size_t colon = item.find(':');
                if (colon == std::string::npos)
                    throw std::runtime_error(path + " does not seem to be a valid qscore model file");
                char* end = nullptr;
                long q = strtol(item.c_str(), &end, 10);
                if (end != item.c_str() + colon)
                    throw std::runtime_error(path + " does not seem to be a valid qscore model file");
                sp.emplace_back((int)q, atof(item.c_str() + colon + 1));
            }
            a = b + 1;
        }
        dists.emplace_back(parts[0], sp);
    }
    return pack_qscore_model(dists);
}

IdentityHost make_identities(double mean_pct, double stdev_pct, double max_pct) {
    IdentityHost id;
    double mean = mean_pct / 100.0, sd = stdev_pct / 100.0, mx = max_pct / 100.0;
    id.max_identity = mx;
    id.mean = (float)mean;
    if (mean == mx || sd == 0.0) {  // py/tksm_badread.py:711-720, 741-743
        id.constant = true;
        id.tab.assign(IDENT_TAB_BINS + 1, (float)mean);
        return id;
    }
    // py/tksm_badread.py:747-757 (called with the percent values; the formula is scale free)
    double u = mean_pct, s = stdev_pct, m = max_pct;
    double a = (((1 - (u / m)) / ((s / m) * (s / m))) - (m / u)) * ((u / m) * (u / m));
    double b = a * ((m / u) - 1);
    if (a < 0.0 || b < 0.0)
        throw std::runtime_error(
            "invalid beta parameters for identity distribution - trying increasing the maximum identity or "
            "reducing the standard deviation");
    id.beta_a = a;
    id.beta_b = b;
    // Quantile table of Beta(a, b) by numeric integration of the density on a fine grid.
    const int N = 1 << 21;
    std::vector<double> cdf((size_t)N + 1, 0.0);
    auto logpdf = [&](double x) { return (a - 1) * std::log(x) + (b - 1) * std::log1p(-x); };
    double mode = (a > 1 && b > 1) ? (a - 1) / (a + b - 2) : 0.5;
    double lmax = logpdf(std::min(std::max(mode, 1e-9), 1 - 1e-9));
    auto pdf = [&](double x) {
        if (x <= 0.0 || x >= 1.0) return 0.0;
        return std::exp(logpdf(x) - lmax);
    };
    // midpoint rule: robust to integrable end-point singularities
    for (int i = 0; i < N; i++) cdf[(size_t)i + 1] = cdf[(size_t)i] + pdf((i + 0.5) / N);
    double tot = cdf[(size_t)N];
    id.tab.resize(IDENT_TAB_BINS + 1);
    size_t j = 0;
    for (int t = 0; t <= IDENT_TAB_BINS; t++) {
        double target = tot * t / IDENT_TAB_BINS;
        while (j < (size_t)N && cdf[j + 1] < target) j++;
        double x;
        if (j >= (size_t)N) {
            x = 1.0;
        } else {
            double lo = cdf[j], hi = cdf[j + 1];
            double f = hi > lo ? (target - lo) / (hi - lo) : 0.0;
            x = (j + f) / N;
        }
        id.tab[(size_t)t] = (float)(mx * x);
    }
    return id;
}

std::vector<float> make_delta_table(const ErrorModelHost& em) {
    // Expected time (in trials per start position) at which the generative identity of a
    // uniformly random fragment falls to the lower edge of each target bin, times a
    // head-room factor.  Only has to be a sensible upper estimate: a read that has not
    // crossed its target at the end of an epoch simply runs another epoch.
    std::vector<float> tab(DELTA_TAB_BINS);
    const double rho = std::max(em.rate_slots, 1e-6);  // rate at which a given slot is hit
    // head-room over the expected stopping time: larger => fewer reads need a second epoch (they
    // hold up the warps that start reads together), more candidate trials per epoch
    double head_room = 1.3;
    if (const char* e = getenv("TKSM_DELTA_HEADROOM")) head_room = std::max(1.0, atof(e));
    const double f_loss = em.rate_loss / rho, f_ins = em.rate_ins / rho;
    for (int b = 0; b < DELTA_TAB_BINS; b++) {
        double target = 0.5 + b / 128.0;
        // identity(c) = (1 - f_loss c) / (1 + f_ins c), c = fraction of changed slots
        double c = (1.0 - target) / (f_loss + target * f_ins);
        c = std::min(c, 0.9);
        double t = -std::log1p(-c) / rho;
        double d = head_room * t + 0.02;
        tab[(size_t)b] = (float)std::min(d, 50.0);
    }
    return tab;
}

std::vector<uint16_t> make_threshold_table(const ErrorModelHost& em, const std::vector<float>& delta) {
    std::vector<uint16_t> tab((size_t)DELTA_TAB_BINS * Q_CLASSES * THR_PER_CLASS);
    for (int b = 0; b < DELTA_TAB_BINS; b++)
        for (int c = 0; c < Q_CLASSES; c++) {
            const double x = em.qrep[(size_t)c] * (double)delta[(size_t)b];
            double term = std::exp(-x), cum = 0.0;  // cum = P(N < m)
            for (int m = 1; m <= THR_PER_CLASS; m++) {
                cum += term;                         // P(N <= m - 1)
                term *= x / (double)m;
                double p = std::max(0.0, 1.0 - cum); // P(N >= m)
                tab[((size_t)b * Q_CLASSES + (size_t)c) * THR_PER_CLASS + (size_t)(m - 1)] =
                    (uint16_t)std::min(65535.0, std::floor(p * 65536.0 + 0.5));
            }
        }
    return tab;
}

ModelNames scan_model_dirs(const char* env_value) {
    ModelNames out;
    if (!env_value) return out;
    std::vector<std::string> dirs;
    std::string v = env_value;
    for (size_t a = 0;;) {
        size_t b = v.find(':', a);
        dirs.push_back(v.substr(a, b == std::string::npos ? std::string::npos : b - a));
        if (b == std::string::npos) break;
        a = b + 1;
    }
    // reversed iteration + dict overwrite == first directory wins (py/sequence.py:21-31)
    for (auto it = dirs.rbegin(); it != dirs.rend(); ++it) {
        std::string dir = *it + "/badread";
        DIR* d = opendir(dir.c_str());
        if (!d) continue;
        while (dirent* e = readdir(d)) {
            std::string name = e->d_name;
            auto take = [&](const char* kind, std::map<std::string, std::string>& m) {
                std::string suf = std::string(".") + kind + ".gz";
                if (name.size() > suf.size() && name.compare(name.size() - suf.size(), suf.size(), suf) == 0)
                    m[name.substr(0, name.size() - suf.size())] = dir + "/" + name;
            };
            take("error", out.error);
            take("qscore", out.qscore);
            take("tail", out.tail);
        }
        closedir(d);
    }
    return out;
}

}  // namespace tk
