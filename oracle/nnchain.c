/* CPU oracle: C restatement of fastcluster's NN-chain linkage -- TEST INFRASTRUCTURE ONLY.
 *
 * The reference calls fastcluster.linkage(y, method) at bam2clusters.py:385,403,411,430.
 * fastcluster (PyPI, 1.1.x, unpinned: README.rst:17) is NOT vendored under the reference
 * tree and is not installable here, so this file restates its published algorithm
 * (Muellner 2013, "fastcluster": NN_chain_core + generate_SciPy_dendrogram) as written
 * down in SURVEY.md appendix A.  Pinned by scipy's known-answer vectors and by
 * scipy.cluster.hierarchy.linkage on tie-free inputs (tests/test_oracle_linkage.py);
 * merge ORDER under exact ties is "parity unpinned" (library specific).
 *
 * Build: make -C oracle libnnchain.so   (gcc -O2 -ffp-contract=off: no FMA contraction, the
 * stock x86-64 fastcluster wheels have none either).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

enum { M_SINGLE = 0, M_COMPLETE = 1, M_AVERAGE = 2, M_WEIGHTED = 3, M_WARD = 4 };

typedef struct { int64_t a, b; double d; int64_t seq; } rec_t;

static int cmp_rec(const void *x, const void *y) {
    const rec_t *p = (const rec_t *)x, *q = (const rec_t *)y;
    if (p->d < q->d) return -1;
    if (p->d > q->d) return 1;
    return (p->seq > q->seq) - (p->seq < q->seq); /* stable */
}

static int64_t uf_find(int64_t *parent, int64_t x) {
    int64_t root = x;
    while (parent[root] != root) root = parent[root];
    while (parent[x] != root) { int64_t nx = parent[x]; parent[x] = root; x = nx; }
    return root;
}

/* packed strict upper triangle, row-major: D(i,j), i<j */
#define DIDX(i, j) ((((2 * N - 3 - (i)) * (i)) >> 1) + (j) - 1)
#define D_(i, j) D[DIDX((int64_t)(i), (int64_t)(j))]

/* stats[0] = NN scans, stats[1] = sum of active counts over scans,
 * stats[2] = merges,  stats[3] = sum of active counts over Lance-Williams updates */
int fc_linkage(int64_t N, double *D, int method, double *Z, int64_t *stats) {
    int64_t NN = N * (N - 1) / 2;
    for (int64_t k = 0; k < NN; ++k)
        if (isnan(D[k])) return -1;
    if (method == M_WARD)
        for (int64_t k = 0; k < NN; ++k) D[k] *= D[k];

    int64_t *succ = malloc((N + 1) * sizeof *succ), *pred = malloc((N + 1) * sizeof *pred);
    int64_t *chain = malloc(N * sizeof *chain), *members = malloc(N * sizeof *members);
    rec_t *rec = malloc((N - 1) * sizeof *rec);
    if (!succ || !pred || !chain || !members || !rec) return -2;
    for (int64_t i = 0; i <= N; ++i) { succ[i] = i + 1; pred[i] = i - 1; }
    for (int64_t i = 0; i < N; ++i) members[i] = 1;
    int64_t start = 0, tip = 0, idx1 = 0, idx2 = 0, i, nactive = N;
    double min = 0, size1 = 0, size2 = 0;
    int64_t nscan = 0, scan_sum = 0, upd_sum = 0;

    for (int64_t j = 0; j < N - 1; ++j) {
        if (tip <= 3) {
            chain[0] = idx1 = start;
            tip = 1;
            idx2 = succ[idx1];
            min = D_(idx1, idx2);
            for (i = succ[idx2]; i < N; i = succ[i])
                if (D_(idx1, i) < min) { min = D_(idx1, i); idx2 = i; }
            nscan++; scan_sum += nactive;
        } else {
            tip -= 3;
            idx1 = chain[tip - 1];
            idx2 = chain[tip];
            min = idx1 < idx2 ? D_(idx1, idx2) : D_(idx2, idx1);
        }
        do {
            chain[tip] = idx2;
            for (i = start; i < idx2; i = succ[i])
                if (D_(i, idx2) < min) { min = D_(i, idx2); idx1 = i; }
            for (i = succ[idx2]; i < N; i = succ[i])
                if (D_(idx2, i) < min) { min = D_(idx2, i); idx1 = i; }
            nscan++; scan_sum += nactive;
            idx2 = idx1;
            idx1 = chain[tip++];
        } while (idx2 != chain[tip - 2]);

        rec[j].a = idx1; rec[j].b = idx2; rec[j].d = min; rec[j].seq = j;

        if (idx1 > idx2) { int64_t t = idx1; idx1 = idx2; idx2 = t; }
        size1 = (double)members[idx1];
        size2 = (double)members[idx2];
        members[idx2] += members[idx1];
        /* remove idx1 from the active list */
        if (idx1 == start) start = succ[idx1];
        else { succ[pred[idx1]] = succ[idx1]; pred[succ[idx1]] = pred[idx1]; }
        nactive--;
        upd_sum += nactive;

#define UPDATE(EXPR_LOW, EXPR_MID, EXPR_HIGH)                                   \
        for (i = start; i < idx1; i = succ[i]) { EXPR_LOW; }                    \
        for (; i < idx2; i = succ[i]) { EXPR_MID; }                             \
        for (i = succ[idx2]; i < N; i = succ[i]) { EXPR_HIGH; }
        switch (method) {
        case M_SINGLE:
#define F(b, a) do { double *_b = &(b); double _a = (a); if (*_b > _a) *_b = _a; } while (0)
            UPDATE(F(D_(i, idx2), D_(i, idx1)), F(D_(i, idx2), D_(idx1, i)), F(D_(idx2, i), D_(idx1, i)))
#undef F
            break;
        case M_COMPLETE:
#define F(b, a) do { double *_b = &(b); double _a = (a); if (*_b < _a) *_b = _a; } while (0)
            UPDATE(F(D_(i, idx2), D_(i, idx1)), F(D_(i, idx2), D_(idx1, i)), F(D_(idx2, i), D_(idx1, i)))
#undef F
            break;
        case M_AVERAGE: {
            double s = size1 / (size1 + size2), t = size2 / (size1 + size2);
#define F(b, a) do { double *_b = &(b); *_b = s * (a) + t * (*_b); } while (0)
            UPDATE(F(D_(i, idx2), D_(i, idx1)), F(D_(i, idx2), D_(idx1, i)), F(D_(idx2, i), D_(idx1, i)))
#undef F
            break;
        }
        case M_WEIGHTED:
#define F(b, a) do { double *_b = &(b); *_b = ((a) + *_b) * .5; } while (0)
            UPDATE(F(D_(i, idx2), D_(i, idx1)), F(D_(i, idx2), D_(idx1, i)), F(D_(idx2, i), D_(idx1, i)))
#undef F
            break;
        case M_WARD:
#define F(b, a) do { double *_b = &(b); double v = (double)members[i];                       \
            *_b = ((v + size1) * (a) - v * min + (v + size2) * (*_b)) / (size1 + size2 + v); } while (0)
            UPDATE(F(D_(i, idx2), D_(i, idx1)), F(D_(i, idx2), D_(idx1, i)), F(D_(idx2, i), D_(idx1, i)))
#undef F
            break;
        default:
            return -3;
        }
    }
    if (method == M_WARD)
        for (int64_t j = 0; j < N - 1; ++j) rec[j].d = sqrt(rec[j].d);

    /* generate_SciPy_dendrogram<sorted=false>: stable sort by height + union-find labels */
    qsort(rec, N - 1, sizeof *rec, cmp_rec);
    int64_t *parent = malloc((2 * N - 1) * sizeof *parent), *size = malloc((2 * N - 1) * sizeof *size);
    for (int64_t k = 0; k < 2 * N - 1; ++k) { parent[k] = k; size[k] = 1; }
    int64_t next = N;
    for (int64_t j = 0; j < N - 1; ++j) {
        int64_t a = uf_find(parent, rec[j].a), b = uf_find(parent, rec[j].b);
        parent[a] = parent[b] = next;
        size[next] = size[a] + size[b];
        Z[4 * j + 0] = (double)(a < b ? a : b);
        Z[4 * j + 1] = (double)(a < b ? b : a);
        Z[4 * j + 2] = rec[j].d;
        Z[4 * j + 3] = (double)size[next];
        next++;
    }
    if (stats) { stats[0] = nscan; stats[1] = scan_sum; stats[2] = N - 1; stats[3] = upd_sum; }
    free(succ); free(pred); free(chain); free(members); free(rec); free(parent); free(size);
    return 0;
}
