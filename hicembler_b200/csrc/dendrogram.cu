// Host-side tail shared by both linkage engines: merge records (a, b, height) -> scipy linkage array.
// Restates fastcluster's generate_SciPy_dendrogram: sqrt for ward, STABLE sort by
// height, union-find relabelling (new ids n, n+1, ...), row = (min root, max root, height, size).
// O(n log n) on n-1 records, so it stays on the host.
#include <math.h>

#include <algorithm>
#include <numeric>
#include <vector>

#include "common.cuh"

namespace hc {

namespace {
struct UF {
    std::vector<int> parent;
    explicit UF(int n) : parent(n) { std::iota(parent.begin(), parent.end(), 0); }
    int find(int x) {
        int root = x;
        while (parent[root] != root) root = parent[root];
        while (parent[x] != root) { int nx = parent[x]; parent[x] = root; x = nx; }
        return root;
    }
};
}  // namespace

// rd is modified in place (sqrt for ward).  If near_ties != nullptr it receives the number of adjacent
// sorted heights closer than rel_tol (relative): the round-parallel engine uses it to certify that the
// sorted merge order cannot depend on last-bit rounding.
void records_to_Z(int n, bool ward, const int *ra, const int *rb, double *rd, double *h_Z, double rel_tol,
                  long long *near_ties) {
    const int m = n - 1;
    if (ward)
        for (int k = 0; k < m; ++k) rd[k] = sqrt(rd[k]);
    // stable sort by height: LSD radix sort of order-preserving 64-bit keys (11-bit digits), ~10x faster
    // than std::stable_sort with an indirect comparison at m = 50 000
    std::vector<int> order(m), tmp_o(m);
    std::vector<unsigned long long> key(m), tmp_k(m);
    for (int k = 0; k < m; ++k) {
        unsigned long long b;
        double v = rd[k] == 0.0 ? 0.0 : rd[k];  // -0.0 sorts with +0.0
        memcpy(&b, &v, 8);
        key[k] = b ^ ((b >> 63) ? ~0ull : 0x8000000000000000ull);
        order[k] = k;
    }
    {
        // LSD radix passes of 11-bit digits, all histograms of a run in ONE pass over the keys (2048-bin tables stay in L1),
        // one scatter pass per digit that actually varies.  First the three TOP digits only (bits 31..63: sign, exponent,
        // 21 mantissa bits): merge heights rarely agree that far, so a stable insertion pass on the full keys finishes the
        // order with next to no moves; if it does meet many inversions (near-equal heights) all six digits are sorted.
        constexpr int BITS = 11, BINS = 1 << BITS;
        auto radix = [&](int first_pass, int n_pass) {
            std::vector<unsigned> cnt((size_t)n_pass * BINS, 0u);
            for (int k = 0; k < m; ++k) {
                const unsigned long long v = key[k];
                for (int ps = 0; ps < n_pass; ++ps) cnt[(size_t)ps * BINS + ((v >> (BITS * (first_pass + ps))) & (BINS - 1))]++;
            }
            for (int ps = 0; ps < n_pass; ++ps) {
                const int sh = BITS * (first_pass + ps);
                unsigned *c = cnt.data() + (size_t)ps * BINS;
                if (m > 0 && c[(key[0] >> sh) & (BINS - 1)] == (unsigned)m) continue;  // all keys share this digit
                unsigned run = 0;
                for (int d = 0; d < BINS; ++d) { const unsigned t = c[d]; c[d] = run; run += t; }
                for (int k = 0; k < m; ++k) {
                    const unsigned pos = c[(key[k] >> sh) & (BINS - 1)]++;
                    tmp_k[pos] = key[k];
                    tmp_o[pos] = order[k];
                }
                key.swap(tmp_k);
                order.swap(tmp_o);
            }
        };
        radix(3, 3);
        long long moves = 0;
        const long long budget = 4ll * m + 64;
        bool done = true;
        for (int k = 1; k < m && done; ++k) {   // stable insertion on the full keys (strict comparison keeps equal keys in order)
            const unsigned long long kk = key[k];
            if (key[k - 1] <= kk) continue;
            const int oo = order[k];
            int j = k;
            while (j > 0 && key[j - 1] > kk) {
                key[j] = key[j - 1];
                order[j] = order[j - 1];
                --j;
                if (++moves > budget) { done = false; break; }
            }
            key[j] = kk;
            order[j] = oo;
        }
        if (!done) radix(0, 3);   // the array is a permutation sorted by the top digits: the low digits first, then stable top
        if (!done) radix(3, 3);
    }
    if (near_ties) {
        long long c = 0;
        for (int k = 1; k < m; ++k) {
            const double lo = rd[order[k - 1]], hi = rd[order[k]];
            if (!(hi - lo > rel_tol * fabs(hi))) c++;
        }
        *near_ties = c;
    }
    UF uf(2 * n - 1);
    std::vector<int> size(2 * n - 1, 1);
    int next = n;
    for (int row = 0; row < m; ++row) {
        const int k = order[row];
        const int x = uf.find(ra[k]), y = uf.find(rb[k]);
        uf.parent[x] = uf.parent[y] = next;
        size[next] = size[x] + size[y];
        h_Z[4 * row + 0] = (double)std::min(x, y);
        h_Z[4 * row + 1] = (double)std::max(x, y);
        h_Z[4 * row + 2] = rd[k];
        h_Z[4 * row + 3] = (double)size[next];
        next++;
    }
}

}  // namespace hc

extern "C" int hc_merges_to_linkage(int32_t n, int ward, const int32_t *h_a, const int32_t *h_b, double *h_d,
                                    double *h_Z, double rel_tol, int64_t *h_near_ties) {
    HC_REQUIRE(n >= 1 && (n < 2 || (h_a && h_b && h_d && h_Z)), "bad arguments");
    if (h_near_ties) *h_near_ties = 0;
    if (n < 2) return HC_OK;
    for (int k = 0; k < n - 1; ++k)
        HC_REQUIRE(h_a[k] >= 0 && h_a[k] < n && h_b[k] >= 0 && h_b[k] < n, "leaf index out of range");
    long long near = 0;
    hc::records_to_Z(n, ward != 0, h_a, h_b, h_d, h_Z, rel_tol, h_near_ties ? &near : nullptr);
    if (h_near_ties) *h_near_ties = near;
    return HC_OK;
}
