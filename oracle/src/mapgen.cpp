// ORACLE (test infrastructure).  Synthetic Octomap-like maps and the
// fcl::OcTree::toBoxes() view of them (MS:288-306 consumes {x,y,z,size,occupancy,threshold}).
// Octomap facts used: pruned leaves are octree-aligned cubes of
// 2^L finest cells; leaf iteration is depth-first with child bit0=x, bit1=y, bit2=z.
#include "../oracle.h"
#include <algorithm>
#include <cmath>
#include <cstring>
#include <vector>

namespace {

struct Pyr {
    int levels;
    std::vector<std::vector<uint8_t>> full, any;  // per level, dims n>>L (padded cube of side S)
    int S;
};

struct Emit {
    const Pyr& p;
    int nx, ny, nz, max_level;
    double res;
    const double* origin;
    double* boxes;
    int64_t cap, count = 0;
    void go(int L, int bx, int by, int bz) {
        const int side = p.S >> L;
        const size_t id = ((size_t)bz * side + by) * side + bx;
        if (!p.any[L][id]) return;
        if (L <= max_level && p.full[L][id]) {
            if (boxes && count < cap) {
                const double sz = res * (double)(1 << L);
                double* o = boxes + 6 * count;
                o[0] = origin[0] + ((double)(bx << L) + 0.5 * (1 << L)) * res;
                o[1] = origin[1] + ((double)(by << L) + 0.5 * (1 << L)) * res;
                o[2] = origin[2] + ((double)(bz << L) + 0.5 * (1 << L)) * res;
                o[3] = sz;
                o[4] = 0.97;  // occupancy after clamping (octomap default clamp max)
                o[5] = 0.5;   // occupancy threshold (octomap default)
            }
            ++count;
            return;
        }
        if (L == 0) return;
        for (int c = 0; c < 8; ++c) go(L - 1, 2 * bx + (c & 1), 2 * by + ((c >> 1) & 1), 2 * bz + ((c >> 2) & 1));
    }
};

}  // namespace

extern "C" {

int64_t orc_grid_to_boxes(const uint8_t* occ, int nx, int ny, int nz, double res,
                          const double origin[3], int max_level, double* boxes, int64_t cap) {
    int S = 1, levels = 0;
    while (S < std::max(nx, std::max(ny, nz))) { S <<= 1; ++levels; }
    Pyr p;
    p.S = S;
    p.levels = levels;
    p.full.resize(levels + 1);
    p.any.resize(levels + 1);
    p.full[0].assign((size_t)S * S * S, 0);
    for (int z = 0; z < nz; ++z)
        for (int y = 0; y < ny; ++y)
            for (int x = 0; x < nx; ++x)
                p.full[0][((size_t)z * S + y) * S + x] = occ[((size_t)z * ny + y) * nx + x] ? 1 : 0;
    p.any[0] = p.full[0];
    for (int L = 1; L <= levels; ++L) {
        const int side = S >> L, cs = side * 2;
        p.full[L].assign((size_t)side * side * side, 0);
        p.any[L].assign((size_t)side * side * side, 0);
        for (int z = 0; z < side; ++z)
            for (int y = 0; y < side; ++y)
                for (int x = 0; x < side; ++x) {
                    int f = 1, a = 0;
                    for (int c = 0; c < 8; ++c) {
                        const size_t id = ((size_t)(2 * z + ((c >> 2) & 1)) * cs + (2 * y + ((c >> 1) & 1))) * cs + (2 * x + (c & 1));
                        f &= p.full[L - 1][id];
                        a |= p.any[L - 1][id];
                    }
                    p.full[L][((size_t)z * side + y) * side + x] = (uint8_t)f;
                    p.any[L][((size_t)z * side + y) * side + x] = (uint8_t)a;
                }
    }
    Emit e{p, nx, ny, nz, std::min(max_level, levels), res, origin, boxes, cap};
    e.go(levels, 0, 0, 0);
    return e.count;
}

// n^3 map: n_walls slabs normal to x (thickness wall_thick cells) each pierced by one
// door_cells^2 door at a seeded (y,z); then seeded clutter cubes (side 1..4 cells) until
// the occupied fraction reaches `occupancy`.  A 6-cell cube is kept free at both corners
// (start / goal regions).  All draws from MT19937(seed) via MTRand [0,1).
int orc_make_map(uint8_t* occ, int n, double res, uint32_t seed, int n_walls, int wall_thick,
                 int door_cells, double occupancy) {
    (void)res;
    std::memset(occ, 0, (size_t)n * n * n);
    orc_mt* mt = orc_mt_new_seed(seed);
    auto at = [&](int x, int y, int z) -> uint8_t& { return occ[((size_t)z * n + y) * n + x]; };
    auto keep_free = [&](int x, int y, int z) {
        const int m = std::min(6, n / 4);
        return (x < m && y < m && z < m) || (x >= n - m && y >= n - m && z >= n - m);
    };
    for (int w = 0; w < n_walls; ++w) {
        const int x0 = (int)((int64_t)(w + 1) * n / (n_walls + 1)) - wall_thick / 2;
        const int dy = (int)(orc_mt_real(mt) * (n - door_cells));
        const int dz = (int)(orc_mt_real(mt) * (n - door_cells));
        for (int x = std::max(0, x0); x < std::min(n, x0 + wall_thick); ++x)
            for (int y = 0; y < n; ++y)
                for (int z = 0; z < n; ++z) {
                    const bool door = y >= dy && y < dy + door_cells && z >= dz && z < dz + door_cells;
                    if (!door) at(x, y, z) = 1;
                }
    }
    int64_t filled = 0;
    const int64_t total = (int64_t)n * n * n;
    for (int64_t i = 0; i < total; ++i) filled += occ[i];
    const int64_t target = (int64_t)(occupancy * (double)total);
    int64_t guard = 0;
    while (filled < target && guard++ < total * 4) {
        const int s = 1 + (int)(orc_mt_real(mt) * 4);
        const int x0 = (int)(orc_mt_real(mt) * (n - s + 1));
        const int y0 = (int)(orc_mt_real(mt) * (n - s + 1));
        const int z0 = (int)(orc_mt_real(mt) * (n - s + 1));
        for (int z = z0; z < z0 + s; ++z)
            for (int y = y0; y < y0 + s; ++y)
                for (int x = x0; x < x0 + s; ++x)
                    if (!at(x, y, z) && !keep_free(x, y, z)) { at(x, y, z) = 1; ++filled; }
    }
    orc_mt_free(mt);
    return 0;
}

}  // extern "C"
