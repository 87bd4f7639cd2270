/*
 * oracle/oracle.h — C API of the CPU ORACLE (test infrastructure, NOT product code).
 *
 * This directory is a CPU restatement of the reference planner's two hot paths
 * (indraneelpatil/bottleneck_guided_rrt, abb_manipulator/src/rrt_star_milestones.cpp,
 * "MS" below) and of the third-party arithmetic those paths call (libnabo 1.0.7,
 * FCL 0.5.x over libccd 2.0, Eigen 3, libstdc++ <random>).  None of those libraries is
 * present in /root/reference or in this image, so their algorithms are restated from
 * their published behaviour: PARITY IS UNPINNED against the real libraries (the
 * reference has no tests or golden vectors for this path).  The only piece pinned
 * against reference code is MT19937 (oracle/_ref builds the reference's own mtrand.cpp).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load this library.  The product (bottleneck_guided_rrt_b200/) never does.
 */
#ifndef BGRRT_ORACLE_H
#define BGRRT_ORACLE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---------------- MT19937 (MS:1015-1020 MTRand; MS:1171-1184 std::mt19937) ---------- */
typedef struct orc_mt orc_mt;
orc_mt*  orc_mt_new_seed(uint32_t seed);                       /* mtrand.cpp seed(unsigned long) */
orc_mt*  orc_mt_new_array(const uint32_t* key, int len);       /* mtrand.cpp seed(array,size)    */
void     orc_mt_free(orc_mt*);
uint32_t orc_mt_u32(orc_mt*);                                  /* rand_int32()                   */
double   orc_mt_real(orc_mt*);                                 /* MTRand::operator(): [0,1)      */
double   orc_mt_std_uniform(orc_mt*, double a, double b);      /* std::uniform_real_distribution<double>(a,b)(std::mt19937) */
void     orc_mt_fill_u32(orc_mt*, uint32_t* out, int64_t n);
void     orc_mt_fill_real(orc_mt*, double* out, int64_t n);

/* ---------------- kNN: libnabo semantics (MS:1306-1312) ------------ */
#define ORC_ALLOW_SELF_MATCH 1u
#define ORC_SORT_RESULTS     2u
/* Brute force with libnabo's leaf arithmetic d2=((dx*dx)+dy*dy)+dz*dz in f32 (no FMA),
 * admission rule d2<=maxR2 && (self || d2>FLT_EPSILON), result ascending by (d2, idx).
 * cloud: 3xN column-major (x0,y0,z0,x1,...).  q: 3xQ.  idx/d2: kxQ column-major.
 * max_radii may be NULL (then max_radius applies to every query). */
int orc_knn_brute(const float* cloud, int64_t n, const float* q, int64_t nq, int k,
                  unsigned flags, float max_radius, const float* max_radii,
                  int32_t* idx, float* d2, int threads);

/* libnabo kd-tree restatement (KDTreeUnbalancedPtInLeavesImplicitBoundsStackOpt with
 * IndexHeapBruteForceVector): build + batched knn with OpenMP over query columns. */
typedef struct orc_nabo orc_nabo;
orc_nabo* orc_nabo_create(const float* cloud, int64_t n, int bucket_size);
void      orc_nabo_destroy(orc_nabo*);
/* returns number of touched points (libnabo TOUCH_STATISTICS semantics) or <0 on error */
int64_t   orc_nabo_knn(const orc_nabo*, const float* q, int64_t nq, int k, float eps,
                       unsigned flags, float max_radius, int32_t* idx, float* d2,
                       int threads);

/* Nearest() literal: MS:130-154.  V: Nx3 row-major xyz (parent column dropped).
 * out_idx[i] = last index among float-equal minimal Norm; out_dist[i] = that Norm. */
int orc_nearest(const float* V_xyz, int64_t n, const float* q_xyz, int64_t nq,
                int32_t* out_idx, float* out_dist, int threads);
float orc_norm(const float* a, const float* b);                /* MS:130-134 */
/* calls of orc_norm (process-wide) where float(pow(s,0.5)) != float(sqrt(s)): the reference's pow
 * against the correctly rounded root the CUDA path uses; tests assert it stays 0 */
long long orc_norm_pow_sqrt_mismatches(void);
float orc_pow_root(double s);                                  /* float(pow(s, 0.5)), counted as above */

/* steer: MS:1233-1253.  nearest, rnd: 3 floats; out: 3 floats; returns 1 if rnd kept. */
int orc_steer(const float* nearest, const float* rnd, float eta, float* out);

/* ---------------- collision: FCL-0.5/libccd-MPR restatement (MS:288-351, 385-471) ---- */
/* boxes: Bx6 row-major doubles {x,y,z,size,occupancy,threshold} = fcl::OcTree::toBoxes().
 * mode 0 = reference-literal linear loop over every box with MPR (early break);
 * mode 1 = same loop, boxes farther than a conservative bound are skipped (no MPR).
 * hit[i] = 1 means IN COLLISION (reference polarity).  Empty box list => 0 (free).
 * band (may be NULL): 1 if |exact separation| <= band_tol for the closest contact, by an
 * independent double-precision distance computation (no MPR). */
int orc_collide_spheres(const double* boxes, int64_t nb, const float* centers_xyz,
                        int64_t m, double radius, int mode, double band_tol,
                        uint8_t* hit, uint8_t* band, int threads);
/* shape 0 = fcl::Cylinder(radius, |p2-p1|) with the reference's f32 transform setup,
 * shape 1 = capsule (fcl::Capsule support) on the same frame. */
int orc_collide_edges(const double* boxes, int64_t nb, const float* p1_xyz,
                      const float* p2_xyz, int64_t m, double radius, int shape, int mode,
                      double band_tol, uint8_t* hit, uint8_t* band, int threads);
/* single pair, for known-answer tests: returns 1 hit / 0 miss; iterations via *iters */
int orc_mpr_sphere_box(const double c[3], double r, const double box_c[3],
                       double box_size, int* iters);
int orc_mpr_cyl_box(const float p1[3], const float p2[3], double r, int shape,
                    const double box_c[3], double box_size, int* iters);
/* independent exact separation (>=0; 0 if intersecting) shape vs box, double */
double orc_sep_sphere_box(const double c[3], double r, const double box_c[3], double box_size);
double orc_sep_cyl_box(const float p1[3], const float p2[3], double r, double dr, int shape,
                       const double box_c[3], double box_size);
/* the reference's f32 cylinder frame: out = {cx,cy,cz, qx,qy,qz,qw, lz} as doubles */
void orc_edge_frame(const float p1[3], const float p2[3], double out[8]);

/* ---------------- planner loop (MS:1160-1406 / RS:1140-1300), B candidates per step ---- */
#define ORC_PLAN_EDGE_CHECK 1u   /* beyond reference: isLineValid geometry on candidate parent edges */
#define ORC_PLAN_REWIRE     2u   /* beyond reference: real rewiring (the reference only prints) */
#define ORC_PLAN_QUIRKS     4u   /* MS:1236-1238: `found` rows keep the Nearest index as stored parent */
#define ORC_PLAN_NO_GOAL    8u   /* do not stop / append the goal node (grow to max_nodes) */
#define ORC_PLAN_DEDUP      16u  /* beyond reference: one candidate per distinct goal / milestone sample of a batch */
typedef struct orc_plan_params {   /* same layout as bgr_planner_params (include/bgrrt.h) */
    float start[3], goal[3];
    float eta;            /* MS:100  0.15 */
    float goal_tol;       /* MS:1340 0.4  */
    float goal_chance;    /* RS:67   0.1 (mode 0) */
    int32_t mode;         /* 0 = RS sampler, 1 = MS sampler (0.8/0.5 + 3 milestones) */
    int32_t k;            /* MS:1309 5 */
    int32_t batch;        /* candidates per step; 1 = the reference's sequential loop */
    int32_t edge_shape;   /* 0 cylinder (MS:437), 1 capsule */
    uint32_t flags;
    uint32_t seed;
    uint32_t pad_;
    double state_radius;  /* MS:322 0.1  */
    double edge_radius;   /* MS:437 0.03 */
} orc_plan_params;
typedef struct orc_planner orc_planner;
orc_planner* orc_planner_create(const orc_plan_params* p, const double* boxes, int64_t nb,
                                const float* table_xyz, int64_t nt, const float* milestones_xyz, int nm);
void    orc_planner_destroy(orc_planner*);
/* collision mode of the planner's state / edge checks: 1 = every leaf box (default), 2 = bucket-grid
 * broadphase in front of the same MPR (same booleans; edge band flags not computed) */
void    orc_planner_set_collision_mode(orc_planner*, int mode);
int     orc_planner_step(orc_planner*, int* n_accepted, int* goal_reached);
int     orc_planner_run(orc_planner*, int64_t max_nodes, int64_t max_batches);
int64_t orc_planner_size(const orc_planner*);
void    orc_planner_get_tree(const orc_planner*, float* xyz, int32_t* parent);
/* [0] batches [1] candidates [2] rejected by state check [3] rejected by edge check
 * [4] improved parents [5] rewired nodes [6] in-band collision results [7] goal candidate */
void    orc_planner_stats(const orc_planner*, int64_t out[8]);

/* ---------------- synthetic maps ---------------------------------------- */
/* Occupancy grid (uint8, x fastest: idx = (z*ny + y)*nx + x) -> toBoxes list with
 * octree-aligned pruning up to 2^max_level cells.  Returns number of boxes written
 * (call with boxes==NULL to count). */
int64_t orc_grid_to_boxes(const uint8_t* occ, int nx, int ny, int nz, double res,
                          const double origin[3], int max_level, double* boxes,
                          int64_t cap);
/* seeded synthetic map: wall slabs with doors + clutter to target occupancy */
int orc_make_map(uint8_t* occ, int n, double res, uint32_t seed, int n_walls,
                 int wall_thick, int door_cells, double occupancy);

#ifdef __cplusplus
}
#endif
#endif
