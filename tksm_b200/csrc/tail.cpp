// Tail-noise model: JSON loader, per-read draws, and the batch transform that appends the tails.
// Reference behaviour: py/tksm_badread.py:886-1036 (see tail.hpp).
#include "tail.hpp"

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <map>
#include <memory>
#include <stdexcept>

namespace tk {

void philox4x32_10_host(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1, uint32_t out[4]) {
    for (int r = 0; r < 10; r++) {
        const uint64_t p0 = 0xD2511F53ull * c0, p1 = 0xCD9E8D57ull * c2;
        const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0, n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
        c0 = n0;
        c1 = (uint32_t)p1;
        c2 = n2;
        c3 = (uint32_t)p0;
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    out[0] = c0;
    out[1] = c1;
    out[2] = c2;
    out[3] = c3;
}

namespace {

// ---- a JSON reader for what json.dump writes (py/tksm_badread.py:929-935)
struct JVal {
    enum Kind { NUL, NUM, STR, ARR, OBJ, BOOL } kind = NUL;
    double num = 0;
    std::string str;
    std::vector<JVal> arr;
    std::map<std::string, JVal> obj;
};

struct JParser {
    const char* p;
    const char* e;
    [[noreturn]] void fail(const char* what) const { throw std::runtime_error(std::string("tail model: malformed JSON (") + what + ")"); }
    void ws() {
        while (p < e && (*p == ' ' || *p == '\n' || *p == '\t' || *p == '\r')) p++;
    }
    bool lit(const char* s) {
        const size_t n = strlen(s);
        if ((size_t)(e - p) >= n && memcmp(p, s, n) == 0) {
            p += n;
            return true;
        }
        return false;
    }
    std::string string_() {
        std::string s;
        p++;  // opening quote
        while (p < e && *p != '"') {
            if (*p == '\\') {
                if (++p >= e) fail("escape");
                switch (*p) {
                    case 'n': s += '\n'; break;
                    case 't': s += '\t'; break;
                    case 'r': s += '\r'; break;
                    case 'b': s += '\b'; break;
                    case 'f': s += '\f'; break;
                    case 'u': {
                        if (e - p < 5) fail("\\u escape");
                        unsigned v = (unsigned)strtoul(std::string(p + 1, p + 5).c_str(), nullptr, 16);
                        s += (char)(v < 128 ? v : '?');
                        p += 4;
                        break;
                    }
                    default: s += *p;
                }
                p++;
            } else {
                s += *p++;
            }
        }
        if (p >= e) fail("unterminated string");
        p++;
        return s;
    }
    JVal value(int depth = 0) {
        if (depth > 64) fail("nesting");
        ws();
        if (p >= e) fail("unexpected end");
        JVal v;
        if (*p == '{') {
            v.kind = JVal::OBJ;
            p++;
            ws();
            if (p < e && *p == '}') {
                p++;
                return v;
            }
            while (true) {
                ws();
                if (p >= e || *p != '"') fail("object key");
                std::string k = string_();
                ws();
                if (p >= e || *p != ':') fail("':'");
                p++;
                v.obj[k] = value(depth + 1);
                ws();
                if (p < e && *p == ',') {
                    p++;
                    continue;
                }
                if (p < e && *p == '}') {
                    p++;
                    return v;
                }
                fail("object");
            }
        }
        if (*p == '[') {
            v.kind = JVal::ARR;
            p++;
            ws();
            if (p < e && *p == ']') {
                p++;
                return v;
            }
            while (true) {
                v.arr.push_back(value(depth + 1));
                ws();
                if (p < e && *p == ',') {
                    p++;
                    continue;
                }
                if (p < e && *p == ']') {
                    p++;
                    return v;
                }
                fail("array");
            }
        }
        if (*p == '"') {
            v.kind = JVal::STR;
            v.str = string_();
            return v;
        }
        if (lit("true") || lit("false")) {
            v.kind = JVal::BOOL;
            v.num = p[-2] == 'u';
            return v;
        }
        if (lit("null")) return v;
        if (lit("NaN")) {
            v.kind = JVal::NUM;
            v.num = NAN;
            return v;
        }
        if (lit("Infinity")) {
            v.kind = JVal::NUM;
            v.num = INFINITY;
            return v;
        }
        if (lit("-Infinity")) {
            v.kind = JVal::NUM;
            v.num = -INFINITY;
            return v;
        }
        char* end = nullptr;
        std::string tok(p, std::min<size_t>((size_t)(e - p), 64));
        v.num = strtod(tok.c_str(), &end);
        if (end == tok.c_str()) fail("value");
        v.kind = JVal::NUM;
        p += end - tok.c_str();
        return v;
    }
};

const JVal& need(const JVal& o, const char* key) {
    auto it = o.obj.find(key);
    if (o.kind != JVal::OBJ || it == o.obj.end()) throw std::runtime_error(std::string("KeyError: '") + key + "' (tail model)");
    return it->second;
}

std::vector<double> numbers(const JVal& a, const char* what) {
    if (a.kind != JVal::ARR) throw std::runtime_error(std::string("tail model: '") + what + "' is not a list");
    std::vector<double> v;
    for (const JVal& x : a.arr) {
        if (x.kind == JVal::ARR && x.arr.size() == 1 && x.arr[0].kind == JVal::NUM) {
            v.push_back(x.arr[0].num);
            continue;
        }
        if (x.kind != JVal::NUM) throw std::runtime_error(std::string("tail model: '") + what + "' holds a non-number");
        v.push_back(x.num);
    }
    return v;
}

inline double unit(uint32_t w) { return ((double)w + 0.5) / 4294967296.0; }

}  // namespace

void TailModel::load_json(const std::string& text) {
    JParser jp{text.data(), text.data() + text.size()};
    const JVal root = jp.value();
    lx = numbers(need(root, "lx"), "lx");
    ly = numbers(need(root, "ly"), "ly");
    const JVal& g = need(root, "grid");
    if (lx.empty() || ly.empty()) throw std::runtime_error("tail model: empty lx / ly");
    if (g.kind != JVal::ARR || g.arr.size() < ly.size()) throw std::runtime_error("IndexError: tail model grid has fewer rows than ly");
    cdf.clear();
    for (size_t i = 0; i < ly.size(); i++) {
        // CustomDist.__init__ (py/tksm_badread.py:977-980): cdf[i] = pdf[i] / sum + cdf[i - 1]
        std::vector<double> pdf = numbers(g.arr[i], "grid row");
        if (pdf.size() != lx.size()) throw std::runtime_error("tail model: grid row length differs from lx");
        double sum = 0;
        for (double v : pdf) sum += v;
        std::vector<double> c(pdf.size());
        double acc = 0;
        for (size_t j = 0; j < pdf.size(); j++) {
            acc = pdf[j] / sum + acc;
            c[j] = acc;
        }
        cdf.push_back(std::move(c));
    }
    const JVal& tr = need(root, "trans");
    if (tr.kind != JVal::ARR || tr.arr.size() != 4) throw std::runtime_error("tail model: 'trans' must be 4 x 4");
    for (int a = 0; a < 4; a++) {
        std::vector<double> row = numbers(tr.arr[a], "trans row");
        if (row.size() != 4) throw std::runtime_error("tail model: 'trans' must be 4 x 4");
        for (int b = 0; b < 4; b++) trans[a][b] = row[b];
    }
    need(root, "begin");  // stored by the reference, never consulted (noise_seq starts from a uniform state)
    const JVal& r = need(root, "ratio");
    if (r.kind != JVal::NUM) throw std::runtime_error("tail model: 'ratio' is not a number");
    ratio = r.num;
    const JVal& bs = need(root, "bases");
    if (bs.kind != JVal::ARR || bs.arr.size() < 4) throw std::runtime_error("IndexError: tail model 'bases' needs 4 entries");
    for (int a = 0; a < 4; a++) {
        if (bs.arr[a].kind != JVal::STR || bs.arr[a].str.size() != 1)
            throw std::runtime_error("tail model: 'bases' entries must be single characters");
        bases[a] = (uint8_t)bs.arr[a].str[0];
    }
    loaded = true;
}

uint32_t TailModel::draw_length(uint64_t seed, uint64_t ordinal, uint32_t frag_len, uint32_t* start) const {
    uint32_t w[4];
    philox4x32_10_host(3, TAG_READ, (uint32_t)ordinal, (uint32_t)(ordinal >> 32), (uint32_t)seed, (uint32_t)(seed >> 32), w);
    *start = w[2] & 3u;
    if (unit(w[0]) > ratio) return 0;  // py/tksm_badread.py:911
    // Custom2Dist.__call__ (py/tksm_badread.py:1014-1024)
    const double y = (double)frag_len;
    size_t pos = (size_t)(std::lower_bound(ly.begin(), ly.end(), y) - ly.begin());
    if (pos + 1 < ly.size() && std::fabs(ly[pos] - y) > std::fabs(ly[pos + 1] - y)) pos++;
    double mult = 1.0;
    if (pos >= ly.size()) {
        mult = (double)pos / ly.back();  // the label count over the last label, as written there
        pos = ly.size() - 1;
    }
    // CustomDist.__call__ (py/tksm_badread.py:982-985); an index past the end (rounding of the last
    // cumulative value) takes the last entry where the reference raises IndexError
    const std::vector<double>& c = cdf[pos];
    size_t q = (size_t)(std::lower_bound(c.begin(), c.end(), unit(w[1])) - c.begin());
    if (q >= c.size()) q = c.size() - 1;
    const double x = lx[q] * mult;
    if (!(x >= 1.0)) return 0;
    return (uint32_t)std::min(x, 16777216.0);  // int(): truncation
}

void TailModel::draw_bases(uint64_t seed, uint64_t ordinal, uint32_t start, uint32_t n, uint8_t* out) const {
    uint32_t state = start, w[4] = {0, 0, 0, 0};
    for (uint32_t i = 0; i < n; i++) {
        if ((i & 3u) == 0)
            philox4x32_10_host(i >> 2, TAG_TAIL, (uint32_t)ordinal, (uint32_t)(ordinal >> 32), (uint32_t)seed,
                               (uint32_t)(seed >> 32), w);
        // random.choices(range(4), weights): bisect_right of u * total in the cumulative weights, capped at 3
        const double* t = trans[state];
        const double c0 = t[0], c1 = c0 + t[1], c2 = c1 + t[2], c3 = c2 + t[3];
        const double v = unit(w[i & 3u]) * c3;
        state = v < c0 ? 0u : v < c1 ? 1u : v < c2 ? 2u : 3u;
        out[i] = frag_code(bases[state]);
    }
}

}  // namespace tk
