/* oracle/nw.c — global (Needleman-Wunsch, unit cost) alignment with path.  TEST INFRASTRUCTURE.
 *
 * Stand-in for the one third-party dependency of the reference's Seq path that is absent from
 * /root/reference: python-edlib (conda `python-edlib`, unpinned: env.yaml:8, env-dev.yaml:10).
 * Call sites it serves: py/tksm_badread.py:170 (align_kmers), :409-411 and :422-426 (identity
 * re-estimation), :613 (get_qscores).  Semantics: edlib.align(query, target, mode="NW",
 * task="path")["cigar"] — extended CIGAR over "=XID", I consumes query, D consumes target.
 *
 * Published algorithm restated: unit-cost edit distance by dynamic programming, here banded
 * with band doubling (Ukkonen) instead of edlib's bit-vectors; both return an optimal path.
 * Which optimal path edlib returns on ties is not pinned by anything in the reference
 * ("parity unpinned" for tie-breaking).  Rule fixed here: trace back from the
 * end preferring the diagonal, then I, then D.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

/* Writes the op string (one char per alignment column, not NUL terminated) into ops, which
 * must hold n + m bytes.  Returns the number of columns, or -1 when out of memory.
 * *dist_out (may be NULL) receives the edit distance. */
int nw_align(const char* q, int n, const char* t, int m, char* ops, int* dist_out) {
    if (n == 0 || m == 0) {
        int c = 0;
        for (int i = 0; i < n; i++) ops[c++] = 'I';
        for (int j = 0; j < m; j++) ops[c++] = 'D';
        if (dist_out) *dist_out = c;
        return c;
    }
    int diff = n > m ? n - m : m - n;
    int w = diff + 16;
    for (;;) {
        if (w > (n > m ? n : m)) w = n > m ? n : m;
        /* cell (i, j) lives at row i, column j - i + w, for |j - i| <= w */
        const int width = 2 * w + 1;
        const int INF = 1 << 29;
        int* prev = (int*)malloc(sizeof(int) * (size_t)width);
        int* cur = (int*)malloc(sizeof(int) * (size_t)width);
        if (!prev || !cur) {
            free(prev); free(cur);
            return -1;
        }
        for (int c = 0; c < width; c++) {
            int j = c - w;
            prev[c] = (j >= 0 && j <= m) ? j : INF;
        }
        for (int i = 1; i <= n; i++) {
            for (int c = 0; c < width; c++) {
                int j = i + c - w;
                if (j < 0 || j > m) {
                    cur[c] = INF;
                    continue;
                }
                if (j == 0) {
                    cur[c] = i;
                    continue;
                }
                int best = prev[c] + (q[i - 1] != t[j - 1]);          /* (i-1, j-1): same column */
                if (c + 1 < width && prev[c + 1] + 1 < best) best = prev[c + 1] + 1; /* (i-1, j): I */
                if (c > 0 && cur[c - 1] + 1 < best) best = cur[c - 1] + 1;           /* (i, j-1): D */
                cur[c] = best;
            }
            int* sw = prev; prev = cur; cur = sw;
        }
        int dist = prev[m - n + w];
        free(prev); free(cur);
        if (dist <= w || w >= (n > m ? n : m)) {
            /* exact inside this band: redo with the full band matrix kept for the traceback */
            int* d = (int*)malloc(sizeof(int) * (size_t)(n + 1) * (size_t)width);
            if (!d) return -1;
#define D(i, j) d[(size_t)(i) * (size_t)width + (size_t)((j) - (i) + w)]
#define IN(i, j) ((j) >= 0 && (j) <= m && (j) - (i) + w >= 0 && (j) - (i) + w < width)
            for (int i = 0; i <= n; i++)
                for (int c = 0; c < width; c++) {
                    int j = i + c - w;
                    int v = INF;
                    if (j >= 0 && j <= m) {
                        if (i == 0) v = j;
                        else if (j == 0) v = i;
                        else {
                            v = D(i - 1, j - 1) + (q[i - 1] != t[j - 1]);
                            if (IN(i - 1, j) && D(i - 1, j) + 1 < v) v = D(i - 1, j) + 1;
                            if (IN(i, j - 1) && D(i, j - 1) + 1 < v) v = D(i, j - 1) + 1;
                        }
                    }
                    d[(size_t)i * (size_t)width + (size_t)c] = v;
                }
            int i = n, j = m, c = 0;
            char* rev = (char*)malloc((size_t)(n + m) + 1);
            while (i > 0 || j > 0) {
                if (i > 0 && j > 0 && D(i, j) == D(i - 1, j - 1) + (q[i - 1] != t[j - 1])) {
                    rev[c++] = q[i - 1] == t[j - 1] ? '=' : 'X';
                    i--; j--;
                } else if (i > 0 && IN(i - 1, j) && D(i, j) == D(i - 1, j) + 1) {
                    rev[c++] = 'I';
                    i--;
                } else {
                    rev[c++] = 'D';
                    j--;
                }
            }
            for (int x = 0; x < c; x++) ops[x] = rev[c - 1 - x];
            free(rev);
            free(d);
#undef D
#undef IN
            if (dist_out) *dist_out = dist;
            return c;
        }
        w *= 2;
    }
}
