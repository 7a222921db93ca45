/* TEST INFRASTRUCTURE ONLY -- plain-C CPU oracle for SURVEY.md section 8 row A1.
 *
 * Restates gnn_manip/utils/utils.py:64-93 (get_connectivity) together with the scikit-learn
 * KDTree64 arithmetic it relies on (third-party dependency, pinned scikit-learn==0.24.1 in
 * environment.yml:12; 1.9.0 installed here):
 *   - float32 inputs are widened to float64           (sklearn/neighbors/_binary_tree.pxi.tp:1267)
 *   - d2 = ((dx*dx) + dy*dy) + dz*dz in float64        (sklearn/metrics/_dist_metrics.pxd.tp:39-49)
 *   - a point is a neighbour when d2 <= r*r            (_binary_tree.pxi.tp:1952-1967)
 *   - neighbours are ordered by ascending distance     (_binary_tree.pxi.tp:1320-1322)
 *   - the first max_neighbours are kept, query = sender, neighbour = receiver, self included
 *                                                       (utils.py:80-91)
 * Ties in distance are undefined in the reference (unstable sort, sklearn/utils/_sorting.pyx);
 * this oracle breaks them by ascending index.  Parity pinned against sklearn itself in
 * tests/test_oracle.py.  Build: `make -C oracle` (gcc -O2 -ffp-contract=off).
 *
 * Never linked into the product library.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef struct { double d2; int32_t j; } cand_t;

static int cand_cmp(const void *a, const void *b) {
    const cand_t *x = (const cand_t *)a, *y = (const cand_t *)b;
    if (x->d2 < y->d2) return -1;
    if (x->d2 > y->d2) return 1;
    return (x->j > y->j) - (x->j < y->j);
}

/* pos [n,3] f32; nbr [n,K] i32 (filled with receivers, -1 padded); cnt [n] i32. */
int oracle_radius_graph(const float *pos, int n, double r, int K, int32_t *nbr, int32_t *cnt) {
    if (n <= 0) return 0;
    const double r2 = r * r;
    double lo[3], hi[3];
    for (int d = 0; d < 3; ++d) { lo[d] = hi[d] = (double)pos[d]; }
    for (int i = 0; i < n; ++i)
        for (int d = 0; d < 3; ++d) {
            double v = (double)pos[3 * i + d];
            if (!(v == v)) return 2; /* NaN */
            if (v < lo[d]) lo[d] = v;
            if (v > hi[d]) hi[d] = v;
        }
    /* cell edge >= r (slightly larger) and capped grid resolution */
    double cs = r * 1.0001;
    int dim[3];
    for (int d = 0; d < 3; ++d) {
        double ext = hi[d] - lo[d];
        double c = cs;
        if (ext / c > 254.0) c = ext / 254.0;
        if (c > cs) cs = c;
    }
    for (int d = 0; d < 3; ++d) dim[d] = (int)floor((hi[d] - lo[d]) / cs) + 1;
    const long ncell = (long)dim[0] * dim[1] * dim[2];
    int32_t *cell_of = (int32_t *)malloc(sizeof(int32_t) * n);
    int32_t *start = (int32_t *)calloc(ncell + 1, sizeof(int32_t));
    int32_t *order = (int32_t *)malloc(sizeof(int32_t) * n);
    if (!cell_of || !start || !order) return 1;
    for (int i = 0; i < n; ++i) {
        int c[3];
        for (int d = 0; d < 3; ++d) {
            c[d] = (int)floor(((double)pos[3 * i + d] - lo[d]) / cs);
            if (c[d] >= dim[d]) c[d] = dim[d] - 1;
        }
        cell_of[i] = (c[2] * dim[1] + c[1]) * dim[0] + c[0];
        start[cell_of[i] + 1]++;
    }
    for (long c = 0; c < ncell; ++c) start[c + 1] += start[c];
    int32_t *cursor = (int32_t *)malloc(sizeof(int32_t) * ncell);
    memcpy(cursor, start, sizeof(int32_t) * ncell);
    for (int i = 0; i < n; ++i) order[cursor[cell_of[i]]++] = i;
    free(cursor);

    int cap = 256;
    cand_t *buf = (cand_t *)malloc(sizeof(cand_t) * cap);
    for (int i = 0; i < n; ++i) {
        const double px = (double)pos[3 * i], py = (double)pos[3 * i + 1], pz = (double)pos[3 * i + 2];
        int ci = cell_of[i];
        int cx = ci % dim[0], cy = (ci / dim[0]) % dim[1], cz = ci / (dim[0] * dim[1]);
        int m = 0;
        for (int dz = -1; dz <= 1; ++dz) {
            int z = cz + dz; if (z < 0 || z >= dim[2]) continue;
            for (int dy = -1; dy <= 1; ++dy) {
                int y = cy + dy; if (y < 0 || y >= dim[1]) continue;
                int x0 = cx > 0 ? cx - 1 : 0, x1 = cx + 1 < dim[0] ? cx + 1 : dim[0] - 1;
                long c0 = ((long)z * dim[1] + y) * dim[0];
                for (int32_t s = start[c0 + x0]; s < start[c0 + x1 + 1]; ++s) {
                    int j = order[s];
                    double dx = px - (double)pos[3 * j];
                    double dyy = py - (double)pos[3 * j + 1];
                    double dzz = pz - (double)pos[3 * j + 2];
                    double d2 = dx * dx;
                    d2 = d2 + dyy * dyy;
                    d2 = d2 + dzz * dzz;
                    if (d2 <= r2) {
                        if (m == cap) { cap *= 2; buf = (cand_t *)realloc(buf, sizeof(cand_t) * cap); }
                        buf[m].d2 = d2; buf[m].j = j; ++m;
                    }
                }
            }
        }
        qsort(buf, m, sizeof(cand_t), cand_cmp);
        int keep = m < K ? m : K;
        cnt[i] = keep;
        for (int t = 0; t < keep; ++t) nbr[(long)i * K + t] = buf[t].j;
        for (int t = keep; t < K; ++t) nbr[(long)i * K + t] = -1;
    }
    free(buf); free(cell_of); free(start); free(order);
    return 0;
}
