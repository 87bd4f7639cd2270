/*
 * bgrrt.h — C ABI of the B200-native hot paths of the bottleneck-guided RRT* planner.
 *
 * Drop-in boundary for the two data-parallel paths of
 * indraneelpatil/bottleneck_guided_rrt, abb_manipulator/src/rrt_star_milestones.cpp ("MS"):
 *   (1) nearest-neighbour search  — Nearest() MS:138-154 and the libnabo object built at
 *       MS:1306 (Nabo::NNSearchF::createKDTreeLinearHeap) and queried at MS:1312 (knn);
 *   (2) collision queries          — isStateValid() MS:310-351 (FCL Sphere vs every Octomap
 *       leaf box) and isLineValid() MS:385-471 (FCL Cylinder vs every leaf box), fed by
 *       generateBoxesFromOctomap() MS:288-306 / fcl::OcTree::toBoxes().
 * The reference has no FFI; these entry points are what a cgo/ctypes/C++ binding for that
 * call surface binds (INTEGRATION.md shows the planner-side stub).  Plain pointers and
 * sizes only; no C++ or torch types.  Every function returns 0 on success or a negative
 * bgr_status; bgr_last_error() gives the message (thread-local).  There is NO CPU
 * fallback: without a CUDA device every compute call fails with BGR_ERR_CUDA.
 *
 * Layout conventions (libnabo's): a point set is dim x N column-major float, i.e.
 * x0,y0,z0,x1,y1,z1,...; knn results are k x Q column-major (query c owns [c*k, c*k+k)).
 * "_dev" variants take DEVICE pointers and are asynchronous on the handle's stream;
 * the others take HOST pointers, copy with cudaMemcpyAsync straight from / to the caller's
 * buffers and return synchronised.  Large batches are cut into chunks whose copies overlap
 * the kernels of their neighbours: that overlap needs page-locked caller buffers
 * (cudaHostAlloc / cudaHostRegister); pageable buffers give the same results with the
 * copies serialised by the driver.  bgr_nns_knn writes page-locked RESULT buffers from the search
 * kernel itself (no staging copy).
 * Handles are not thread-safe (the reference caller is single-threaded) and belong to the
 * CUDA device that was current when they were created: calls made with another device
 * current return BGR_ERR_STATE.
 */
#ifndef BGRRT_H
#define BGRRT_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define BGR_VERSION 100

typedef enum bgr_status {
    BGR_OK = 0,
    BGR_ERR_ARG = -1,      /* bad argument (libnabo would throw std::runtime_error) */
    BGR_ERR_CUDA = -2,     /* CUDA runtime error / no device */
    BGR_ERR_ALLOC = -3,
    BGR_ERR_UNSUPPORTED = -4,
    BGR_ERR_STATE = -5
} bgr_status;

const char* bgr_last_error(void);
int bgr_version(void);
/* number of CUDA kernels this library has launched in this process (monotonic) */
uint64_t bgr_launch_count(void);
int bgr_device_count(int* count);
/* SM count, SM clock (kHz), L2 bytes of the current device */
int bgr_device_info(int* sms, int* sm_clock_khz, int64_t* l2_bytes);

/* measurement aid (no counterpart in the reference): L2 read bandwidth of the current device in GB/s,
 * best of `reps` launches over a resident buffer of `bytes` (power of two).  mode 0 = coalesced
 * LDG.128 rows, 1 = random 32-byte sectors (8-byte loads, 32 B counted per load), 2 = random 512-byte
 * rows (LDG.128 per lane); L1 bypassed.  bench.py uses it as the peak of roofline.l2. */
int bgr_microbench_l2(int64_t bytes, int mode, int reps, double* gb_per_s);

/* ------------------------------------------------------------------ nearest neighbours */
typedef struct bgr_nns bgr_nns;

/* libnabo search option flags (nabo.h SearchOptionFlags) and creation flags */
#define BGR_ALLOW_SELF_MATCH 1u
#define BGR_SORT_RESULTS 2u      /* accepted; results are always ascending by (d2, index) */
#define BGR_TOUCH_STATISTICS 1u  /* creation flag: knn reports evaluated (query,node) pairs */

/* search methods for bgr_nns_set_method */
#define BGR_METHOD_AUTO 0
#define BGR_METHOD_BRUTE 1   /* exact brute-force scan (FP32-pipe bound) */
#define BGR_METHOD_GRID 2    /* exact grid-bucketed scan: one-pass ring-1 path, ring expansion behind it */
#define BGR_METHOD_GRID_GENERAL 3 /* diagnostic: the ring-expansion search only */

/* Replaces Nabo::NNSearchF::createKDTreeLinearHeap(cloud) (MS:1306).  cloud is copied to
 * an append-only SoA buffer in HBM (libnabo keeps a reference instead).  n may be 0. */
int bgr_nns_create(const float* cloud_3xN, int64_t n, int dim, unsigned create_flags,
                   bgr_nns** out);
int bgr_nns_create_dev(const float* d_cloud_3xN, int64_t n, int dim, unsigned create_flags,
                       void* cuda_stream, bgr_nns** out);
/* Replaces the per-iteration rebuild MS:1277-1306: appends m points (indices continue). */
int bgr_nns_append(bgr_nns* nns, const float* pts_3xM, int64_t m);
int bgr_nns_append_dev(bgr_nns* nns, const float* d_pts_3xM, int64_t m);
int64_t bgr_nns_size(const bgr_nns* nns);
int bgr_nns_set_stream(bgr_nns* nns, void* cuda_stream);
int bgr_nns_set_method(bgr_nns* nns, int method);
/* indices returned by knn/nearest are local index + base (index-range shards, §8e) */
int bgr_nns_set_index_base(bgr_nns* nns, int64_t base);
/* (re)build the grid index now instead of lazily at the next query */
int bgr_nns_build_index(bgr_nns* nns);

/* Replaces nns->knn(q, indices, dists2, k, eps, flags, maxRadius) (MS:1309-1312) and its
 * matrix / per-query-maxRadii overloads.  Exact search only: eps must be 0.
 * d2 = ((dx*dx)+dy*dy)+dz*dz in f32, every operation rounded, no FMA; admission
 * d2<=maxRadius^2 && (ALLOW_SELF_MATCH || d2>FLT_EPSILON); output ascending by (d2,index);
 * unfilled slots idx=-1, d2=+inf.  Errors like libnabo's checkSizesKnn: k > N (or N-1
 * without self match), unknown flag bits.  max_radii (Q floats) may be NULL.
 * touched (may be NULL) receives the number of (query,node) pairs evaluated.
 * Host-pointer form only: d2_kxQ may be NULL when the caller consumes the indices alone (the
 * reference does, MS:1312-1319): the distances are then not copied back (half the D2H bytes). */
int bgr_nns_knn(bgr_nns* nns, const float* q_3xQ, int64_t nq, int k, float eps,
                unsigned flags, float max_radius, const float* max_radii,
                int32_t* idx_kxQ, float* d2_kxQ, uint64_t* touched);
int bgr_nns_knn_dev(bgr_nns* nns, const float* d_q_3xQ, int64_t nq, int k, float eps,
                    unsigned flags, float max_radius, const float* d_max_radii,
                    int32_t* d_idx_kxQ, float* d_d2_kxQ, uint64_t* touched);
/* the same search (eps = 0, no radius limit) with a per-query skip flag on the device: rows of
 * queries with d_skip[q] != 0 are left untouched (grid method: their warps exit at once; the
 * plain scan computes them anyway).  The batched planner passes the state-check result
 * (isStateValid MS:1270 comes before the knn MS:1312: rejected candidates never search). */
int bgr_nns_knn_skip_dev(bgr_nns* nns, const float* d_q_3xQ, int64_t nq, int k, unsigned flags,
                         const uint8_t* d_skip, int32_t* d_idx_kxQ, float* d_d2_kxQ);
/* Allocation-free form of the brute-force kNN scan.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph
 * capture, tight serving loops, harnesses that forbid it).  Same search and same results as
 * bgr_nns_knn_dev with eps = 0 and no per-query radii (exact (d2, index) order, rows padded with
 * idx = -1, d2 = +inf), always by the plain scan; device pointers only; runs on the handle's stream.
 * The per-chunk partial lists of the scan live in ONE fixed 4 MiB `static __device__` array of the
 * library (held by every GPU that loads it until exit) instead of the handle's growable scratch, so
 * the call makes no cudaMalloc / cudaFree and no allocating library call.
 * Limit: nq * k <= BGR_KNN_STATIC_MAX_RESULTS; larger batches return BGR_ERR_ARG (split the batch,
 * or use bgr_nns_knn_dev).  Every call of the process shares the array and it keeps the last
 * call's contents (nothing in it has to start at zero: each call writes every slot it later reads):
 * calls that overlap on the GPU — different streams — race on it and must be serialized.
 * chunks_used (may be NULL) receives the number of node chunks, 1 = the scratch was not needed. */
#define BGR_KNN_STATIC_MAX_RESULTS 32768
int bgr_nns_knn_scan_static_dev(bgr_nns* nns, const float* d_q_3xQ, int64_t nq, int k, unsigned flags,
                                float max_radius, int32_t* d_idx_kxQ, float* d_d2_kxQ, int* chunks_used);

/* Replaces Nearest(V, x) (MS:138-154) for a batch of x: Norm in the reference's mixed
 * precision (f32 differences, f64 squares/sum/sqrt, f32 result), `<=` so the LAST index
 * among float-equal minima wins.  dist may be NULL. */
int bgr_nns_nearest(bgr_nns* nns, const float* q_3xQ, int64_t nq, int32_t* idx, float* dist);
int bgr_nns_nearest_dev(bgr_nns* nns, const float* d_q_3xQ, int64_t nq, int32_t* d_idx,
                        float* d_dist);
/* test / measurement aid — which code paths this handle has run so far: [0] grid index builds,
 * [1] grid_knn_kernel / grid_knn_group_kernel launches (uniform clouds), [2] launches of the crowded-index instantiation,
 * [3] queries the LAST crowded launch handed to grid_knn_heavy_kernel (may exceed its 256 slots: the
 * rest ran the ring search), [4] grid_nearest_kernel launches, [5] nearest_kernel (plain scan)
 * launches, [6] knn_brute_kernel launches, [7] nodes covered by the current index */
int bgr_nns_debug_stats(bgr_nns* nns, uint64_t out[8]);
/* Tile kNN (csrc/tile_knn.cuh): an alternative exact search for large query batches over a fully indexed,
 * uncrowded cloud — one lane per query out of shared-memory tiles.  OFF by default: measured on the
 * configs[1] workload it is slower than the warp-per-query grid search (0.143 ms against 0.100 ms, DESIGN.md
 * section 3.2b), so it is kept as a tested option.  set_tile: tx < 0 = off, all 0 = tile size chosen from the
 * query density, else the tile size in grid cells and the warps per tile CTA.
 * tile_stats: [0] tile searches run so far, [1] queries the LAST one handed back to the warp-per-query
 * search (threshold misses, 26-bit distance ties, overfull tiles), [2] tx | ty<<8 | tz<<16 | warps<<24 of
 * the last one, [3] its tile count.  Results never depend on any of this. */
int bgr_nns_set_tile(bgr_nns* nns, int tx, int ty, int tz, int warps);
int bgr_nns_tile_stats(bgr_nns* nns, uint64_t out[4]);
int bgr_nns_destroy(bgr_nns* nns);

/* Merge G per-shard top-k slabs (each k x Q, ascending, global indices) into one top-k by
 * (d2, index) — the exchange step after the all-gather of an index-range-sharded search. */
int bgr_knn_merge_dev(const int32_t* d_idx_GxkxQ, const float* d_d2_GxkxQ, int g,
                      int64_t nq, int k, int32_t* d_idx_kxQ, float* d_d2_kxQ,
                      void* cuda_stream);
/* same, slab s starting at entry s*slab_stride of both inputs (packed all-gather buffers) */
int bgr_knn_merge_strided_dev(const int32_t* d_idx_in, const float* d_d2_in, int g,
                              int64_t slab_stride, int64_t nq, int k, int32_t* d_idx_kxQ,
                              float* d_d2_kxQ, void* cuda_stream);

/* Exchange step of the index-range-sharded search over NVLink peer memory (one process per
 * GPU; no counterpart in the reference, which is single-process).  Every rank creates a window
 * of 4 regions of max_entries (>= k*Q) 32-bit words, exports it as a CUDA IPC handle
 * (bgr_peer_handle_bytes() bytes; the caller ships the handles between processes, e.g. with
 * torch.distributed.all_gather_object) and opens every peer's handle.  Per search: the local
 * bgr_nns_knn_dev writes into bgr_peer_slab(); bgr_peer_merge_dev then runs ONE kernel that
 * waits for the peers' slabs, merges this rank's query slice [r*Q/G,(r+1)*Q/G) from all G slabs
 * by (d2, index) with peer loads and stores the merged k-lists into every peer's result region,
 * plus a wait kernel; after it (stream order) bgr_peer_result() holds the full k x Q result on
 * every rank.  All ranks must call bgr_peer_merge_dev the same number of times with the same
 * nq and k.  Bounded spins: a peer that never arrives sets *error (bgr_peer_status) instead of
 * hanging the GPU. */
typedef struct bgr_peer bgr_peer;
int bgr_peer_handle_bytes(void);
int bgr_peer_create(int rank, int world, int64_t max_entries, bgr_peer** out);
int bgr_peer_export(bgr_peer* peer, void* handle_out);
int bgr_peer_open(bgr_peer* peer, int peer_rank, const void* handle);
int bgr_peer_slab(bgr_peer* peer, int32_t** d_idx_kxQ, float** d_d2_kxQ);
int bgr_peer_result(bgr_peer* peer, int32_t** d_idx_kxQ, float** d_d2_kxQ);
int bgr_peer_merge_dev(bgr_peer* peer, int64_t nq, int k, void* cuda_stream);
int bgr_peer_status(bgr_peer* peer, void* cuda_stream, int* error);
int bgr_peer_destroy(bgr_peer* peer);

/* ------------------------------------------------------------------ occupancy map */
typedef struct bgr_map bgr_map;

#define BGR_SHAPE_CYLINDER 0 /* fcl::Cylinder(radius, |p2-p1|): the reference, MS:437 */
#define BGR_SHAPE_CAPSULE 1  /* fcl::Capsule on the same frame (conservative superset) */

/* Octomap view: finest resolution res, grid origin (min corner), dims in finest cells. */
int bgr_map_create(double res, const double origin[3], const int dims[3], bgr_map** out);
int bgr_map_set_stream(bgr_map* map, void* cuda_stream);
/* Replaces octomap_Callback + generateBoxesFromOctomap (MS:273-306): boxes is B x 6
 * row-major {x,y,z,size,occupancy,threshold} = fcl::OcTree::toBoxes(); kept if
 * occupancy >= threshold.  Boxes must be octree-aligned cubes of res*2^L.  Clears first. */
int bgr_map_set_boxes(bgr_map* map, const double* boxes_Bx6, int64_t nb);
/* incremental refresh: remove then add leaf boxes (same layout) */
int bgr_map_update_boxes(bgr_map* map, const double* add_Bx6, int64_t n_add,
                         const double* remove_Bx6, int64_t n_remove);
/* map refresh: the new complete leaf list is diffed against the list set by the last
 * set_boxes / refresh_boxes and only the changed leaves are rasterised (counts returned) */
int bgr_map_refresh_boxes(bgr_map* map, const double* boxes_Bx6, int64_t nb, int64_t* n_added,
                          int64_t* n_removed);
int bgr_map_occupied_cells(bgr_map* map, int64_t* cells);
/* copy the occupancy bitmap (one bit per finest cell, 4x4x4 bricks) to host; bytes may be
 * queried with out==NULL */
int bgr_map_download_bitmap(bgr_map* map, uint64_t* out, int64_t* n_words);

/* Replaces isStateValid(point) (MS:310-351) for a batch: hit[i]=1 means IN COLLISION
 * (reference polarity).  in_band[i]=1 when the closest contact lies within the 1e-6 m
 * tolerance band of FCL's MPR solver (result tolerance-determined).  in_band may be NULL.
 * An empty map reports free (the reference returns a stale flag there). */
int bgr_map_collide_spheres(bgr_map* map, const float* c_3xM, int64_t m, double radius,
                            uint8_t* hit, uint8_t* in_band);
int bgr_map_collide_spheres_dev(bgr_map* map, const float* d_c_3xM, int64_t m, double radius,
                                uint8_t* d_hit, uint8_t* d_in_band);
/* Replaces isLineValid(p1,p2) (MS:385-471) for a batch of edges. */
int bgr_map_collide_edges(bgr_map* map, const float* p1_3xM, const float* p2_3xM, int64_t m,
                          double radius, int shape, uint8_t* hit, uint8_t* in_band);
int bgr_map_collide_edges_dev(bgr_map* map, const float* d_p1_3xM, const float* d_p2_3xM,
                              int64_t m, double radius, int shape, uint8_t* d_hit,
                              uint8_t* d_in_band);
/* counters of the last collide_edges call: [0] work items (edge,slab), [1] occupied
 * candidate voxels, [2] exact narrow-phase tests, [3] narrow-phase iteration-cap events */
int bgr_map_last_stats(bgr_map* map, uint64_t out[4]);
/* sweep gate of collide_edges (isLineValid MS:385-471 has no counterpart: the reference loops
 * over every leaf): level 0 dilated by d = ceil((radius + 1e-4) / res) cells decides, from the
 * cells the edge's AXIS crosses alone, the edges that are certainly clear and (with a deep-hit
 * certificate on an axis cell) certainly colliding; only slabs next to a dilated cell run the
 * rectangle phase.  Results are identical with the gate off (set_gate(map, 0): measurement /
 * test aid).  last_gate_stats: [0] (edge,slab) items that ran the rectangle phase in the last
 * call, [1] d of the dilated map currently built (0 = none). */
int bgr_map_set_gate(bgr_map* map, int on);
int bgr_map_last_gate_stats(bgr_map* map, uint64_t out[2]);
/* The gate maps (level 0 dilated by the query radius, one cached per radius class) follow map updates
 * incrementally: update_boxes / refresh_boxes re-dilate only the bricks within d cells of the changed leaves;
 * set_boxes and very large updates rebuild them in full at the next query.  out[0] = full dilations,
 * out[1] = incremental updates so far (test / measurement aid). */
int bgr_map_dilation_stats(bgr_map* map, uint64_t out[2]);
/* the gate map for dilation d (1..3 cells; built or brought up to date if need be), same word layout as
 * bgr_map_download_bitmap (test aid) */
int bgr_map_download_gate(bgr_map* map, int d, uint64_t* out, int64_t* n_words);
/* measurement aid: with timing on, collide_edges records CUDA events between its kernels on the
 * map's stream; last_kernel_ms returns the durations of the last call: [0] edges_sweep_kernel,
 * [1] edges_narrow_kernel, [2] collide_edges_kernel (general / overflow path), [3] finalize */
int bgr_map_set_timing(bgr_map* map, int on);
int bgr_map_last_kernel_ms(bgr_map* map, float out[4]);
int bgr_map_destroy(bgr_map* map);

/* ------------------------------------------------------------------ OcTree ingestion */
/* Replaces octomap_msgs::fullMsgToMap + fcl::OcTree::toBoxes (MS:279,291): a serialised Octomap
 * -> the toBoxes rows {x,y,z,size,occupancy,threshold} of its occupied leaves (occupancy >=
 * occ_threshold, 0.5 in the reference), in leaf-iterator order.  Formats:
 *   BGR_OCTREE_FULL_DATA    octomap_msgs/Octomap.data with binary=false = the data section of an
 *                           .ot file (what the reference receives); res from the message
 *   BGR_OCTREE_BINARY_DATA  Octomap.data with binary=true = the data section of a .bt file
 *   BGR_OCTREE_FILE         a whole .ot / .bt file with its text header (res is read from it;
 *                           ids OcTree, ColorOcTree, OcTreeStamped)
 * boxes may be NULL to query the count (*nb: capacity in, count out).  bbox = metric bounds of
 * the occupied leaves (size the bgr_map from it; the origin must sit on a multiple of 32 cells
 * from the key origin so that pruned leaves stay octree-aligned).  Host code (serial stream). */
#define BGR_OCTREE_FULL_DATA 0
#define BGR_OCTREE_BINARY_DATA 1
#define BGR_OCTREE_FILE 2
int bgr_octree_to_boxes(const void* data, int64_t nbytes, int format, double res, double occ_threshold,
                        double* boxes_Bx6, int64_t* nb, double* res_out, double bbox_min[3],
                        double bbox_max[3]);

/* ------------------------------------------------------------------ path post-processing */
/* Replaces PrunePlan(rrt_path) (MS:562-643): greedy shortcutting from the start — from the
 * current anchor take the FARTHEST waypoint whose isLineValid edge is free.  All P(P-1)/2
 * anchor-waypoint edges are evaluated in one collide_edges launch; the walk then reads the
 * answers.  path is P x 3 host floats; keep receives the kept waypoint indices (capacity
 * *n_keep >= P in, count out); n_band (may be NULL) counts consulted edges whose result is
 * tolerance-determined.  Reproduced quirk: the walk ends when the last kept waypoint has the
 * goal's x coordinate (MS:579).  Where the reference would loop forever (no free successor)
 * this returns BGR_ERR_STATE. */
int bgr_map_prune_path(bgr_map* map, const float* path_3xP, int64_t p, double radius, int shape,
                       int32_t* keep, int64_t* n_keep, int64_t* n_band);
/* Replaces SmoothenPlan(rrt_path) (MS:645-799), arithmetic types as in the reference: P == 3
 * quadratic Bezier (float t), P > 3 cubic Bezier over the FIRST FOUR waypoints (double t),
 * P == 2 straight line resampled; 11 samples.  Host arithmetic (33 numbers).  out may be NULL
 * to query the sample count. */
int bgr_path_smoothen(const double* path_3xP, int64_t p, double* out_3xS, int64_t* n_out);
/* Obstacle clearance of find_labels (MS:956-975) / distance_from_obstacle (RD:890-917) for a
 * batch of points: min over all leaf boxes of fcl::distance(Sphere(radius), Box) in closed
 * form, -1 when the sphere touches or penetrates a leaf, 1000 when the map is empty (MS:962). */
int bgr_map_clearance(bgr_map* map, const double* pts_3xM, int64_t m, double radius, double* dist);
/* Replaces find_labels(final_path) (MS:943-1011) up to the file write: drops the start and the
 * goal waypoint, takes the n_labels interior waypoints of smallest float clearance (ties by
 * index; n_ties reports float-equal neighbours in the selection, where std::sort's order is
 * unspecified) -> idx (into the trimmed path), labels (n_labels x 3 floats), clearance (P-2
 * floats, may be NULL). */
int bgr_map_find_labels(bgr_map* map, const double* path_3xP, int64_t p, double radius, int n_labels,
                        int32_t* idx, float* labels_3xL, float* clearance, int* n_ties);

/* ------------------------------------------------------------------ batched planner step */
/* The extension loop of rrt_star_milestones.cpp main() (MS:1160-1406; rrt_star.cpp
 * RS:1140-1300 for the 10 % goal-bias sampler) for B candidates per launch, evaluated against
 * the tree snapshot at step start; batch = 1 is the reference's sequential loop.  Per
 * candidate: sampler (MS:1171-1219 | RS:1145-1178; ONE seeded MT19937 stream, draw order
 * y, x, (z), two 32-bit words per uniform_real_distribution<double> draw) -> Nearest
 * (MS:138-154) -> steer (MS:1233-1268) -> isStateValid (MS:310-351) -> knn K (MS:1309-1312,
 * when size > K, MS:1273) -> cost_wiring (MS:802-831) / extend_tree (MS:834-853) -> parent
 * rule (MS:1322-1323) -> append (MS:1325-1327) -> goal test (MS:1340) -> goal node
 * (MS:1346-1347).  The tree lives on the device: positions in the planner's own bgr_nns
 * store, parents as int32. */
#define BGR_PLAN_EDGE_CHECK 1u  /* beyond reference: isLineValid geometry on candidate parent edges */
#define BGR_PLAN_REWIRE 2u      /* beyond reference: real rewiring (MS:883/923 only print) */
#define BGR_PLAN_QUIRKS 4u      /* MS:1236-1238: rows of the `found` branch keep the Nearest index as parent */
#define BGR_PLAN_NO_GOAL 8u     /* never stop / append the goal node: grow the tree to max_nodes */
#define BGR_PLAN_DEDUP 16u      /* beyond reference (batch > 1): the goal / milestone samples of one batch are the same
                                 * point and, against one snapshot, the same candidate; keep the first of each and
                                 * reject the copies (counted in rejected_state) instead of appending coincident nodes */
#define BGR_PLAN_NO_PERSIST 32u  /* measurement / test aid: never run steps in the persistent small-batch kernel (batch <= 8,
                                 * no edge checks, tree < 65,536 nodes: whole steps inside one launch, same tree bit for bit) */
#define BGR_PLAN_NO_GRAPH 64u    /* measurement / test aid: never replay a step as a CUDA graph (planner on its own stream, tree below
                                 * 16,384 nodes: the step's ~25 launches are captured once and replayed; same tree bit for bit) */
#define BGR_PLAN_MODE_RS 0      /* rrt_star.cpp sampler: rand_chance > goal_chance ? table : goal */
#define BGR_PLAN_MODE_MS 1      /* rrt_star_milestones.cpp sampler: >0.8 table, (0.5,0.8) goal, else milestone[int z] */

typedef struct bgr_planner_params {
    float start[3], goal[3];
    float eta;            /* MS:100  0.15 */
    float goal_tol;       /* MS:1340 0.4  */
    float goal_chance;    /* RS:67   0.1 (mode RS) */
    int32_t mode;
    int32_t k;            /* MS:1309 5; 1..32 */
    int32_t batch;        /* candidates per step, 1..65536 */
    int32_t edge_shape;   /* BGR_SHAPE_* for BGR_PLAN_EDGE_CHECK */
    uint32_t flags;
    uint32_t seed;        /* MT19937 seed (std::mt19937(seed) / mtrand.cpp seed()) */
    uint32_t pad_;
    double state_radius;  /* MS:322 0.1  */
    double edge_radius;   /* MS:437 0.03 */
} bgr_planner_params;

typedef struct bgr_planner bgr_planner;

/* map is borrowed (must outlive the planner).  table_xyz: the sample domain replacing the
 * reference's reachability table rows (MS:1138-1140,1195-1197), nt x 3 host floats, nt <= 2^24.
 * milestones_xyz: nm x 3 host floats (MS:249-271), nm >= 3 in mode MS. */
int bgr_planner_create(const bgr_planner_params* params, bgr_map* map, const float* table_xyz, int64_t nt,
                       const float* milestones_xyz, int nm, bgr_planner** out);
int bgr_planner_set_stream(bgr_planner* pl, void* cuda_stream);
/* one batch; n_accepted / goal_reached may be NULL */
int bgr_planner_step(bgr_planner* pl, int* n_accepted, int* goal_reached);
/* steps until the goal is reached, the tree has >= max_nodes nodes or max_batches were run */
int bgr_planner_run(bgr_planner* pl, int64_t max_nodes, int64_t max_batches, int* goal_reached);
/* independent planning queries (one planner per start/goal pair; the reference runs one process
 * per query): the queries run concurrently (one host thread and one stream per query in flight,
 * up to 16), so their small per-step kernels overlap on the device.  Each planner's tree is exactly
 * the one bgr_planner_run would grow.  The planners must not share a stream or a map handle.
 * goal_reached: n flags (may be NULL). */
int bgr_planner_run_many(bgr_planner** pls, int n, int64_t max_nodes, int64_t max_batches, int* goal_reached);
/* Re-arm a planner for another query on the same map / sample table / milestones: the tree becomes the single
 * root `start`, the sample stream restarts from `seed`, the statistics are cleared; every buffer is kept.  The
 * tree then grown is the one a freshly created planner with these parameters grows. */
int bgr_planner_reset(bgr_planner* pl, const float start[3], const float goal[3], uint32_t seed);
/* nq independent queries through n_workers planners (one host thread each; the planners must not share a map or a
 * stream): a worker re-arms its planner with the next (start, goal, seed) of the list as soon as it finishes one,
 * so no query waits for the slowest of a group.  starts / goals: 3 x Q column-major.  out_reached[q] / out_nodes[q]:
 * goal flag and tree size of query q (either may be NULL). */
int bgr_planner_run_queue(bgr_planner** pls, int n_workers, const float* starts_3xQ, const float* goals_3xQ,
                          const uint32_t* seeds, int64_t nq, int64_t max_nodes, int64_t max_batches, int* out_reached,
                          int64_t* out_nodes);
int64_t bgr_planner_size(const bgr_planner* pl);
/* tree download: xyz as 3 x N column-major (x0,y0,z0,...), parent[0] = -1 */
int bgr_planner_get_tree(bgr_planner* pl, float* xyz_3xN, int32_t* parent);
/* BuildPlan (MS:156-218): node indices from the root to the last node (the goal node once the
 * goal is reached); returns the length through *n (query with path == NULL) */
int bgr_planner_get_path(bgr_planner* pl, int32_t* path, int64_t* n);
/* [0] batches [1] candidates [2] rejected by the state check [3] rejected by the edge check
 * [4] improved parents [5] rewired nodes [6] in-band collision results [7] goal candidate */
int bgr_planner_stats(bgr_planner* pl, int64_t out[8]);
/* test / measurement aid: [0] launches of the persistent small-batch kernel so far, [1] steps they ran,
 * [2] steps replayed as a CUDA graph */
int bgr_planner_persist_stats(bgr_planner* pl, uint64_t out[3]);
/* the planner's node store (borrowed): knn / nearest queries against the growing tree */
int bgr_planner_nns(bgr_planner* pl, bgr_nns** out);
int bgr_planner_destroy(bgr_planner* pl);

#ifdef __cplusplus
}
#endif
#endif /* BGRRT_H */
