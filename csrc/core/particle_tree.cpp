// Bounding-volume tree over the particle table, the second container next to the uniform grid
// (csrc/kernels/particle_grid.cu).  Role of the reference's SolidTree (src/SolidTree.h:24-80, src/SolidTree.hpp:22-85: a
// tree over the ball centres, split along the widest direction, walked on the device without a stack through "back"
// links).  Here: a threaded box tree - every node holds the box of its whole subtree (particle reach included) and the
// index of the node to continue with when the subtree is skipped; children follow their parent in pre-order, so the walk
// is  `i = hit ? (leaf ? skip : i + 1) : skip`  with no stack and no per-level state.  Unlike the grid it needs no cell
// size, so a few large particles among many small ones do not inflate every cell (the grid's cell edge is twice the
// LARGEST reach).  Built on the host from the rows the integrator delivered (n log n, median split by nth_element).
//
// Node layout (8 words of 4 bytes, read by ParticleRange in the generated model code):
//   float lo[3], hi[3];  int skip;  int particle (-1 = inner node)
#include <algorithm>
#include <cmath>
#include <cstring>
#include <vector>
#include "lattice.h"

namespace tclb {

namespace {
struct Builder {
    const double *rows;
    double margin;
    std::vector<int> order;
    std::vector<int> words;                 // 8 per node
    float down(double v) const { float f = (float)v; return (double)f > v ? std::nextafterf(f, -INFINITY) : f; }
    float up(double v) const { float f = (float)v; return (double)f < v ? std::nextafterf(f, INFINITY) : f; }
    static void put_f(std::vector<int> &w, size_t at, float f) { memcpy(&w[at], &f, 4); }
    // builds the subtree over order[b..e) at the next free node, returns its index; `skip` = where a walk continues after it
    int build(int b, int e) {
        const int node = (int)(words.size() / 8);
        words.resize(words.size() + 8);
        double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300}, clo[3] = {1e300, 1e300, 1e300}, chi[3] = {-1e300, -1e300, -1e300};
        for (int k = b; k < e; k++) {
            const double *r = rows + (size_t)order[k] * 16;
            const double reach = r[0] + margin;
            for (int a = 0; a < 3; a++) {
                lo[a] = std::min(lo[a], r[1 + a] - reach); hi[a] = std::max(hi[a], r[1 + a] + reach);
                clo[a] = std::min(clo[a], r[1 + a]); chi[a] = std::max(chi[a], r[1 + a]);
            }
        }
        if (e - b > 1) {
            int axis = 0;
            for (int a = 1; a < 3; a++) if (chi[a] - clo[a] > chi[axis] - clo[axis]) axis = a;
            const int mid = b + (e - b) / 2;
            std::nth_element(order.begin() + b, order.begin() + mid, order.begin() + e,
                             [&](int p, int q) { return rows[(size_t)p * 16 + 1 + axis] < rows[(size_t)q * 16 + 1 + axis]; });
            build(b, mid);
            build(mid, e);
        }
        const size_t at = (size_t)node * 8;
        // a little room: the walk compares single-precision node positions, the exact test is done on the particle itself
        for (int a = 0; a < 3; a++) { put_f(words, at + a, down(lo[a] - 1e-5 * (1 + fabs(lo[a])))); put_f(words, at + 3 + a, up(hi[a] + 1e-5 * (1 + fabs(hi[a])))); }
        words[at + 6] = (int)(words.size() / 8);          // everything after this subtree
        words[at + 7] = e - b == 1 ? order[b] : -1;
        return node;
    }
};
}  // namespace

// returns the node words (8 per node); empty when there are no particles
std::vector<int> build_particle_tree_words(const std::vector<double> &rows16, int n, double margin) {
    Builder B;
    B.rows = rows16.data();
    B.margin = margin;
    B.order.resize(n);
    for (int i = 0; i < n; i++) B.order[i] = i;
    if (n > 0) { B.words.reserve((size_t)(2 * n) * 8); B.build(0, n); }
    return B.words;
}

}  // namespace tclb
