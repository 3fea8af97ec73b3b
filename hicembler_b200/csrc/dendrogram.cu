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
    // stable sort by height: LSD radix sort of order-preserving 64-bit keys (6 digits of 11 bits), ~10x faster
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
        // six 11-bit digits; all histograms in ONE pass over the keys (2048-bin tables stay in L1), then one scatter
        // pass per digit that actually varies
        constexpr int BITS = 11, NPASS = 6, BINS = 1 << BITS;
        std::vector<unsigned> cnt((size_t)NPASS * BINS, 0u);
        for (int k = 0; k < m; ++k) {
            const unsigned long long v = key[k];
            for (int pass = 0; pass < NPASS; ++pass) cnt[(size_t)pass * BINS + ((v >> (BITS * pass)) & (BINS - 1))]++;
        }
        for (int pass = 0; pass < NPASS; ++pass) {
            const int sh = BITS * pass;
            unsigned *c = cnt.data() + (size_t)pass * BINS;
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
