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
    std::vector<int> order(m);
    std::iota(order.begin(), order.end(), 0);
    std::stable_sort(order.begin(), order.end(), [&](int x, int y) { return rd[x] < rd[y]; });
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
