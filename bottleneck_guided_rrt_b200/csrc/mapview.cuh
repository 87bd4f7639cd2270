// Device view of the occupancy map (bgr_map, map.cu) shared by the translation units that
// query it: level-0 occupancy bricks + per-level leaf markers (layout: see map.cu).
#pragma once
#include "mpr.cuh"

namespace bgr {

constexpr int kMaxLevel = 5;            // largest pruned leaf kept as one box: 32^3 cells
constexpr double kBandTol = 1e-6;       // FCL collision_tolerance (libccd mpr_tolerance)
constexpr double kBandFlag = 2.5e-6;    // inflation used to flag tolerance-determined results
constexpr double kCertMargin = 1e-5;    // closed-form certificates decide only beyond this
constexpr double kSweepMargin = 1e-4;   // broadphase hull inflation

struct LevelGrid {
    uint64_t* bits;
    int nb[3];   // bricks per axis
};

struct MapView {
    double res, inv_res;
    double origin[3];
    int dims[3];
    LevelGrid lv[kMaxLevel + 1];
    unsigned level_mask;   // bit L set: level L has at least one leaf marker
};

__device__ __forceinline__ bool test_bit(const LevelGrid& g, int x, int y, int z) {
    const int64_t w = ((int64_t)(z >> 2) * g.nb[1] + (y >> 2)) * g.nb[0] + (x >> 2);
    const int b = ((z & 3) << 4) | ((y & 3) << 2) | (x & 3);
    return (__ldg(g.bits + w) >> b) & 1ull;
}

// leaf containing occupied finest cell (x,y,z): level and min-corner finest coordinates
__device__ __forceinline__ int find_leaf(const MapView& m, int x, int y, int z, int& lx, int& ly, int& lz) {
    for (int L = kMaxLevel; L >= 1; --L) {
        if (!((m.level_mask >> L) & 1u)) continue;
        if (test_bit(m.lv[L], x >> L, y >> L, z >> L)) {
            lx = (x >> L) << L;
            ly = (y >> L) << L;
            lz = (z >> L) << L;
            return L;
        }
    }
    lx = x; ly = y; lz = z;
    return 0;
}

__device__ __forceinline__ void leaf_box(const MapView& m, int L, int lx, int ly, int lz, D3& c, double& e) {
    const double half_cells = 0.5 * (double)(1 << L);
    c.x = m.origin[0] + ((double)lx + half_cells) * m.res;
    c.y = m.origin[1] + ((double)ly + half_cells) * m.res;
    c.z = m.origin[2] + ((double)lz + half_cells) * m.res;
    e = (m.res * (double)(1 << L)) / 2;
}

}  // namespace bgr

struct bgr_map;
namespace bgr {
MapView map_view(const bgr_map* h);      // map.cu
cudaStream_t map_stream(const bgr_map* h);
int map_device(const bgr_map* h);
// slot for scratch memory another translation unit keeps per map (freed with free_fn by bgr_map_destroy)
void** map_aux(bgr_map* h, void (*free_fn)(void*));
// gate map of a sphere query of this radius (level 0 dilated; built / brought up to date if need be); bits = nullptr
// when the gate does not apply
int map_sphere_gate(bgr_map* h, double radius, LevelGrid* dil);
uint64_t map_signature(const bgr_map* h);   // changes whenever a captured planner step would have to be re-captured
}  // namespace bgr
