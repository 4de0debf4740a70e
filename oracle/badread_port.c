/* oracle/badread_port.c — literal C port of the reference's sequential Badread path.
 * TEST INFRASTRUCTURE ONLY: loaded by tests/ and by bench.py's CPU legs (cpu_baseline,
 * --impl reference); the product (tksm_b200/) never links or loads it.
 *
 * What it restates, statement by statement (vpc-ccg/tksm):
 *   ERROR_MODEL_PY.align_kmers            py/tksm_badread.py:146-197
 *   ErrorModel.add_errors_to_kmer         py/tksm_badread.py:119-144
 *   ERROR_MODEL_PY.add_one_random_change  py/tksm_badread.py:199-213
 *   MISC_PY.get_random_* / identity_from_edlib_cigar   py/tksm_badread.py:223-257
 *   SIMULATE_PY.sequence_fragment         py/tksm_badread.py:324-451
 *   QScoreModel.get_qscore                py/tksm_badread.py:584-598
 *   QSCOREMODEL_PY.get_qscores            py/tksm_badread.py:607-655
 *   Identities.get_identity               py/tksm_badread.py:741-745
 * Unlike oracle/twin.c this is the reference's own algorithm: one trial at a time, a running
 * identity estimate re-anchored by real alignments every 25 changes, qscores from a re-alignment
 * of read to fragment.  It is the CPU baseline bench.py times, and tests/test_badread_port.py
 * checks its reads against the statistics of the unmodified Python reference
 * (tests/golden/ref_dist_*.json.gz).  edlib (absent third-party dependency) is
 * replaced by oracle/nw.c.  Python's unseeded `random` / numpy.random become one xoshiro256**
 * stream per read, so the port is reproducible and independent of the thread count.
 */
#include <math.h>
#include <pthread.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "nw.c"

#define MAX_K 9
#define SLOT_CAP 15 /* bases one fragment slot can hold (alternatives are at most 11 long) */

typedef struct {
    uint8_t len;
    char b[SLOT_CAP];
} slot_t;

typedef struct {
    int k;         /* error-model k-mer size; 1 for the "random" model */
    int has_table; /* 0: every trial is add_one_random_change */
    const uint32_t* alt_begin; /* 4^k + 1 offsets into the alternative arrays */
    const double* alt_prob;
    const slot_t* alt_slots;   /* k slots per alternative (filled by port_align_kmers) */
    int kq;                    /* qscore k-mer size */
    int n_cigars;
    const char* cigar_blob;    /* concatenated cigar strings */
    const uint32_t* cigar_off; /* n_cigars + 1 */
    const uint32_t* q_begin;   /* n_cigars + 1 offsets into q_score / q_prob */
    const uint8_t* q_score;
    const double* q_prob;
    const int32_t* hash;       /* open addressing over cigar strings, -1 = empty */
    uint32_t hash_mask;
    int ident_const;
    double ident_mean, beta_a, beta_b, ident_max;
} port_model;

/* ------------------------------------------------------------------ RNG */
typedef struct { uint64_t s[4]; } rng_t;
static uint64_t splitmix(uint64_t* x) {
    uint64_t z = (*x += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
static void rng_seed(rng_t* r, uint64_t seed, uint64_t stream) {
    uint64_t x = seed * 0xD1342543DE82EF95ull + stream;
    for (int i = 0; i < 4; i++) r->s[i] = splitmix(&x);
}
static inline uint64_t rotl(uint64_t x, int k) { return (x << k) | (x >> (64 - k)); }
static inline uint64_t rng_next(rng_t* r) {
    uint64_t* s = r->s;
    uint64_t res = rotl(s[1] * 5, 7) * 9, t = s[1] << 17;
    s[2] ^= s[0]; s[3] ^= s[1]; s[1] ^= s[2]; s[0] ^= s[3]; s[2] ^= t; s[3] = rotl(s[3], 45);
    return res;
}
static inline double rng_double(rng_t* r) { return (double)(rng_next(r) >> 11) * (1.0 / 9007199254740992.0); }
/* random.randint(a, b), both ends included */
static inline int64_t rng_randint(rng_t* r, int64_t a, int64_t b) {
    uint64_t span = (uint64_t)(b - a) + 1;
    return a + (int64_t)(((__uint128_t)rng_next(r) * span) >> 64);
}
static double rng_normal(rng_t* r) {
    double u1 = rng_double(r), u2 = rng_double(r);
    if (u1 < 1e-300) u1 = 1e-300;
    return sqrt(-2.0 * log(u1)) * cos(6.283185307179586 * u2);
}
static double rng_gamma(rng_t* r, double a) { /* Marsaglia-Tsang */
    if (a < 1.0) {
        double u = rng_double(r);
        if (u < 1e-300) u = 1e-300;
        return rng_gamma(r, a + 1.0) * pow(u, 1.0 / a);
    }
    double d = a - 1.0 / 3.0, c = 1.0 / sqrt(9.0 * d);
    for (;;) {
        double x = rng_normal(r), v = 1.0 + c * x;
        if (v <= 0) continue;
        v = v * v * v;
        double u = rng_double(r);
        if (u < 1.0 - 0.0331 * x * x * x * x) return d * v;
        if (log(u) < 0.5 * x * x + d * (1.0 - v + log(v))) return d * v;
    }
}
static double rng_beta(rng_t* r, double a, double b) {
    double x = rng_gamma(r, a), y = rng_gamma(r, b);
    return x / (x + y);
}

static const char BASES[4] = {'A', 'C', 'G', 'T'};
static inline char random_base(rng_t* r) { return BASES[rng_randint(r, 0, 3)]; } /* :223-229 */
static inline char random_different_base(rng_t* r, char b) {                     /* :231-236 */
    char x = random_base(r);
    while (x == b) x = random_base(r);
    return x;
}

/* ------------------------------------------------------- align_kmers (:146-197) */
/* out: k slots.  Returns 0 on success. */
int port_align_kmers(const char* kmer, int k, const char* alt, int alt_len, slot_t* out) {
    if (k <= 2 || alt_len <= 1 || kmer[0] != alt[0] || kmer[k - 1] != alt[alt_len - 1]) return -1;
    for (int j = 0; j < k; j++) out[j].len = 0;
    out[0].len = 1; out[0].b[0] = kmer[0];
    out[k - 1].len = 1; out[k - 1].b[0] = kmer[k - 1];
    const char* km = kmer + 1; int kn = k - 2;
    const char* al = alt + 1;  int an = alt_len - 2;
    char ops[64];
    int nops;
    if (an == 0) {
        nops = kn;
        memset(ops, 'D', (size_t)kn);
    } else {
        nops = nw_align(al, an, km, kn, ops, NULL); /* edlib.align(alt, kmer, task="path") */
        if (nops < 0) return -1;
    }
    int kp = 0, ap = 0;
    /* middle slots start as None in the reference and are assigned before being appended to */
    for (int c = 0; c < nops; c++) {
        if (ops[c] == '=' || ops[c] == 'X') {
            out[kp + 1].len = 1; out[kp + 1].b[0] = al[ap];
            ap++; kp++;
        } else if (ops[c] == 'D') {
            out[kp + 1].len = 0;
            kp++;
        } else { /* I: result[kmer_pos] += alt[alt_pos] */
            slot_t* s = &out[kp];
            if (s->len < SLOT_CAP) s->b[s->len++] = al[ap];
            ap++;
        }
    }
    if (out[0].len == 2) { /* :191-196 */
        char ins = out[0].b[1];
        out[0].len = 1;
        slot_t* s = &out[1];
        if (s->len < SLOT_CAP) {
            memmove(s->b + 1, s->b, s->len);
            s->b[0] = ins;
            s->len++;
        }
    }
    return 0;
}

/* ------------------------------------------- add_errors_to_kmer (:119-144, 199-213) */
static void one_random_change(rng_t* r, const char* kmer, int k, slot_t* out) {
    for (int j = 0; j < k; j++) { out[j].len = 1; out[j].b[0] = kmer[j]; }
    int type = (int)rng_randint(r, 0, 2);            /* random.choice(["s", "i", "d"]) */
    int pos = (int)rng_randint(r, 0, k - 1);
    if (type == 0) {
        out[pos].b[0] = random_different_base(r, out[pos].b[0]);
    } else if (type == 1) {
        if (rng_double(r) < 0.5) { out[pos].b[1] = random_base(r); out[pos].len = 2; }
        else { out[pos].b[1] = out[pos].b[0]; out[pos].b[0] = random_base(r); out[pos].len = 2; }
    } else {
        out[pos].len = 0;
    }
}

static int kmer_code(const char* kmer, int k) {
    int code = 0;
    for (int j = 0; j < k; j++) {
        int c = kmer[j] == 'A' ? 0 : kmer[j] == 'C' ? 1 : kmer[j] == 'G' ? 2 : kmer[j] == 'T' ? 3 : -1;
        if (c < 0) return -1;
        code = code * 4 + c;   /* first base most significant: the order the model file lists k-mers in */
    }
    return code;
}

static void add_errors_to_kmer(const port_model* M, rng_t* r, const char* kmer, slot_t* out) {
    const int k = M->k;
    int code = M->has_table ? kmer_code(kmer, k) : -1;
    if (code < 0 || M->alt_begin[code] == M->alt_begin[code + 1]) { /* random model / k-mer not in the table */
        one_random_change(r, kmer, k, out);
        return;
    }
    const uint32_t a0 = M->alt_begin[code], a1 = M->alt_begin[code + 1];
    double sum = 0;
    for (uint32_t a = a0; a < a1; a++) sum += M->alt_prob[a];
    double rest = 1.0 - sum;
    if (rest < 0.0) rest = 0.0;
    /* random.choices(alts + [None], weights = probs + [rest]) */
    double x = rng_double(r) * (sum + rest), c = 0;
    for (uint32_t a = a0; a < a1; a++) {
        c += M->alt_prob[a];
        if (x < c) {
            memcpy(out, M->alt_slots + (size_t)a * (size_t)k, sizeof(slot_t) * (size_t)k);
            return;
        }
    }
    one_random_change(r, kmer, k, out);
}

/* identity_from_edlib_cigar (:244-257) on an op string */
static double ops_identity(const char* ops, int n) {
    if (n <= 0) return 0.0;
    int m = 0;
    for (int i = 0; i < n; i++) m += ops[i] == '=';
    return (double)m / (double)n;
}

static size_t join_slots(const slot_t* s, size_t a, size_t b, char* out) {
    size_t n = 0;
    for (size_t p = a; p < b; p++) { memcpy(out + n, s[p].b, s[p].len); n += s[p].len; }
    return n;
}

/* ------------------------------------------------ get_qscore (:584-598) */
static uint32_t fnv(const char* s, int n) {
    uint32_t h = 2166136261u;
    for (int i = 0; i < n; i++) h = (h ^ (uint8_t)s[i]) * 16777619u;
    return h;
}
uint32_t port_hash(const char* s, int n) { return fnv(s, n); }

static int find_cigar(const port_model* M, const char* s, int n) {
    uint32_t h = fnv(s, n) & M->hash_mask;
    for (;;) {
        int32_t e = M->hash[h];
        if (e < 0) return -1;
        uint32_t o = M->cigar_off[e];
        if ((int)(M->cigar_off[e + 1] - o) == n && memcmp(M->cigar_blob + o, s, (size_t)n) == 0) return e;
        h = (h + 1) & M->hash_mask;
    }
}

static char get_qscore(const port_model* M, rng_t* r, const char* cigar, int n) {
    for (;;) {
        int e = find_cigar(M, cigar, n);
        if (e >= 0) {
            uint32_t q0 = M->q_begin[e], q1 = M->q_begin[e + 1];
            double sum = 0;
            for (uint32_t q = q0; q < q1; q++) sum += M->q_prob[q];
            double x = rng_double(r) * sum, c = 0;
            uint8_t val = M->q_score[q1 - 1];
            for (uint32_t q = q0; q < q1; q++) {
                c += M->q_prob[q];
                if (x < c) { val = M->q_score[q]; break; }
            }
            return (char)(val + 33);
        }
        /* cigar = cigar[1:-1].strip("D") */
        cigar++; n -= 2;
        while (n > 0 && cigar[0] == 'D') { cigar++; n--; }
        while (n > 0 && cigar[n - 1] == 'D') n--;
        if (n <= 0) return '!';   /* cannot happen: '=', 'X', 'I' are always in the model */
    }
}

/* get_qscores (:607-655).  Returns the alignment identity; qual gets seq_len chars. */
static double get_qscores(const port_model* M, rng_t* r, const char* seq, int seq_len, const char* frag, int frag_len,
                          char* qual) {
    char* full = (char*)malloc((size_t)seq_len + (size_t)frag_len + 1);
    int ncol = nw_align(seq, seq_len, frag, frag_len, full, NULL);   /* edlib.align(seq, frag, task="path") */
    if (ncol < 0) { free(full); memset(qual, '!', (size_t)seq_len); return 0.0; }
    double identity = ops_identity(full, ncol);
    int* pos2col = (int*)malloc(sizeof(int) * ((size_t)seq_len + 1));
    int i = 0;
    for (int j = 0; j < ncol; j++)
        if (full[j] != 'D') pos2col[i++] = j;
    const int margins = (M->kq - 1) / 2;
    for (i = 0; i < seq_len; i++) {
        int start = i - margins, end = i + margins;
        while (start < 0 || end >= seq_len) { start++; end--; }
        int cs = pos2col[start], ce = pos2col[end];
        qual[i] = get_qscore(M, r, full + cs, ce - cs + 1);
    }
    free(pos2col);
    free(full);
    return identity;
}

/* ------------------------------------------------ sequence_fragment (:324-451) */
enum { ALIGNMENT_INTERVAL = 25, ALIGNMENT_SIZE = 1000 }; /* SETTINGS_PY, :45-46 */

typedef struct {
    char* seq; char* qual; int len; double identity;
} read_t;

static int sequence_fragment(const port_model* M, rng_t* r, const char* mol, int mol_len, double target,
                             int compute_qscores, read_t* out) {
    const int k = M->k;
    const int frag_len = mol_len + 2 * k;   /* tail noise is "" (no_noise) */
    char* fragment = (char*)malloc((size_t)frag_len + 1);
    for (int j = 0; j < k; j++) fragment[j] = random_base(r);
    memcpy(fragment + k, mol, (size_t)mol_len);
    for (int j = 0; j < k; j++) fragment[k + mol_len + j] = random_base(r);
    slot_t* nb = (slot_t*)malloc(sizeof(slot_t) * (size_t)frag_len);
    for (int p = 0; p < frag_len; p++) { nb[p].len = 1; nb[p].b[0] = fragment[p]; }
    char* joined = (char*)malloc((size_t)frag_len * SLOT_CAP + 1);
    char* ops = (char*)malloc((size_t)frag_len * (SLOT_CAP + 1) + 2);

    double errors = 0.0;
    long change_count = 0, loop_count = 0;
    const int max_kmer_index = frag_len - 1 - k;
    slot_t nk[MAX_K];
    for (;;) {
        loop_count++;
        if (loop_count > 100L * frag_len) break;
        if ((double)change_count > 0.9 * frag_len) break;
        double estimated_identity = 1.0 - errors / frag_len;
        if (estimated_identity <= target) break;
        int i = (int)rng_randint(r, 0, max_kmer_index);
        const char* kmer = fragment + i;
        add_errors_to_kmer(M, r, kmer, nk);
        int same = 1;
        for (int j = 0; j < k && same; j++) same = nk[j].len == 1 && nk[j].b[0] == kmer[j];
        if (same) continue;
        for (int j = 0; j < k; j++) {
            const char fb = fragment[i + j];
            const int differs = !(nk[j].len == 1 && nk[j].b[0] == fb);
            slot_t* cur = &nb[i + j];
            if (differs && cur->len == 1 && cur->b[0] == fb) {
                *cur = nk[j];
                change_count++;
                int new_errors = nk[j].len < 2 ? 1 : nk[j].len - 1;
                errors += new_errors * pow(estimated_identity, 1.5);
                if (change_count % ALIGNMENT_INTERVAL == 0) {
                    if (frag_len <= ALIGNMENT_SIZE) {
                        int jl = (int)join_slots(nb, 0, (size_t)frag_len, joined);
                        int nc = nw_align(fragment, frag_len, joined, jl, ops, NULL);
                        errors = (1.0 - ops_identity(ops, nc)) * frag_len;
                    } else {
                        int pos = (int)rng_randint(r, 0, frag_len - ALIGNMENT_SIZE);
                        int pos2 = pos + ALIGNMENT_SIZE;
                        int jl = (int)join_slots(nb, (size_t)pos, (size_t)pos2, joined);
                        int nc = nw_align(fragment + pos, ALIGNMENT_SIZE, joined, jl, ops, NULL);
                        double est = (1.0 - ops_identity(ops, nc)) * frag_len;
                        double w = (double)ALIGNMENT_SIZE / frag_len;
                        errors = est * w + errors * (1 - w);
                    }
                }
            }
        }
    }
    int start_trim = (int)join_slots(nb, 0, (size_t)k, joined);
    int end_trim = (int)join_slots(nb, (size_t)(frag_len - k), (size_t)frag_len, joined);
    int seq_len = (int)join_slots(nb, 0, (size_t)frag_len, joined);
    char* qual = (char*)malloc((size_t)seq_len + 1);
    double identity;
    if (compute_qscores) {
        identity = get_qscores(M, r, joined, seq_len, fragment, frag_len, qual);
    } else {
        memset(qual, 'K', (size_t)seq_len);
        identity = 1.0 - errors / frag_len;
    }
    int L = seq_len - start_trim - end_trim;   /* seq[start_trim:-end_trim], end_trim >= 1 */
    if (L < 0) L = 0;
    out->seq = (char*)malloc((size_t)L + 1);
    out->qual = (char*)malloc((size_t)L + 1);
    memcpy(out->seq, joined + start_trim, (size_t)L);
    memcpy(out->qual, qual + start_trim, (size_t)L);
    out->len = L;
    out->identity = identity;
    free(qual); free(ops); free(joined); free(nb); free(fragment);
    return 0;
}

/* the Pool(threads).imap_unordered(..., chunksize=10) of py/sequence.py:363-368 as a thread pool */
typedef struct {
    const port_model* M; uint64_t n_reads; const char* seqs; const uint64_t* seq_off; uint64_t seed;
    int compute_qscores; read_t* res; uint64_t next;
} job_t;

static void* worker(void* arg) {
    job_t* J = (job_t*)arg;
    for (;;) {
        uint64_t a = __atomic_fetch_add(&J->next, 10, __ATOMIC_RELAXED);
        if (a >= J->n_reads) break;
        uint64_t b = a + 10 < J->n_reads ? a + 10 : J->n_reads;
        for (uint64_t r = a; r < b; r++) {
            rng_t rng;
            rng_seed(&rng, J->seed, r);
            const port_model* M = J->M;
            double target = M->ident_const ? M->ident_mean : M->ident_max * rng_beta(&rng, M->beta_a, M->beta_b); /* :741-745 */
            sequence_fragment(M, &rng, J->seqs + J->seq_off[r], (int)(J->seq_off[r + 1] - J->seq_off[r]), target,
                              J->compute_qscores, &J->res[r]);
        }
    }
    return NULL;
}

/* Simulates n_reads molecules.  out_seq / out_qual (may be NULL) receive the reads back to back at
 * out_off[r] (capacity out_cap bytes each; reads that do not fit are dropped from the buffers but still
 * counted).  Returns the total number of simulated bases. */
uint64_t port_simulate(const port_model* M, uint64_t n_reads, const char* seqs, const uint64_t* seq_off, uint64_t seed,
                       int compute_qscores, int threads, char* out_seq, char* out_qual, uint64_t* out_off,
                       uint64_t out_cap, uint32_t* read_len, double* identity) {
    uint64_t total = 0;
    read_t* res = (read_t*)calloc(n_reads ? n_reads : 1, sizeof(read_t));
    job_t job = {M, n_reads, seqs, seq_off, seed, compute_qscores, res, 0};
    if (threads < 1) threads = 1;
    if (threads > 256) threads = 256;
    pthread_t th[256];
    for (int t = 1; t < threads; t++) pthread_create(&th[t], NULL, worker, &job);
    worker(&job);
    for (int t = 1; t < threads; t++) pthread_join(th[t], NULL);
    for (uint64_t r = 0; r < n_reads; r++) total += (uint64_t)res[r].len;
    uint64_t off = 0;
    for (uint64_t r = 0; r < n_reads; r++) {
        if (read_len) read_len[r] = (uint32_t)res[r].len;
        if (identity) identity[r] = res[r].identity;
        if (out_off) out_off[r] = off;
        if (out_seq && off + (uint64_t)res[r].len <= out_cap) {
            memcpy(out_seq + off, res[r].seq, (size_t)res[r].len);
            if (out_qual) memcpy(out_qual + off, res[r].qual, (size_t)res[r].len);
        }
        off += (uint64_t)res[r].len;
        free(res[r].seq); free(res[r].qual);
    }
    if (out_off) out_off[n_reads] = off;
    free(res);
    return total;
}
