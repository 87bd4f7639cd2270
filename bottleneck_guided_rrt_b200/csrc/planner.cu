// bgr_planner_* — the extension loop of the reference planner, B candidates per launch.
// Reference: abb_manipulator/src/rrt_star_milestones.cpp main() MS:1160-1406 (sampler
// MS:1171-1219, steer MS:1233-1268, cost_wiring MS:802-831, extend_tree MS:834-853, parent rule
// MS:1322-1323, goal test MS:1340, goal node MS:1346-1347) and rrt_star.cpp RS:1145-1178 (10 %
// goal-bias sampler).  Every candidate of a step is evaluated against the tree snapshot at step
// start; batch = 1 is the reference's sequential loop.  The hot primitives are the library's own
// device entry points (Nearest, kNN, sphere / edge collision); this file adds the small
// per-candidate kernels around them and keeps the whole tree on the device:
//   positions  : the planner's bgr_nns store (SoA x/y/z in HBM, append-only)
//   parents    : int32 array (the reference keeps them as float in V[i][3])
//   MT19937    : 624-word state in HBM, advanced by a one-CTA kernel (3-phase parallel twist)
// One 48-byte counter block is read back per step (accepted count, goal flag, statistics).
#include <atomic>
#include <cmath>
#include <cstring>
#include <string>
#include <thread>
#include <vector>
#include "internal.cuh"
#include "spheres.cuh"
#include "warp_topk.cuh"

using namespace bgr;

namespace {

constexpr int kMtN = 624, kMtM = 397;

struct PlanDev {                 // device view shared by the kernels
    bgr_planner_params p;
    const float *x, *y, *z;      // node store
    int32_t* par;
    int64_t n;                   // snapshot size
    const long long* n_dev;      // optional: the snapshot size lives on the device (CUDA-graph replays of a step)
    int B, K, W;                 // candidates, neighbours, stream words per candidate
    int have_knn;
    const uint32_t* words;
    const float* table;
    int64_t nt;
    const float* milestones;
    float *prand, *pnew;
    uint8_t* found;
    int32_t* near;
    uint8_t *hit, *band;
    int32_t* knn_idx;
    float *e1, *e2;
    uint8_t *ehit, *eband;
    float* costs;                // [B][K+1], slot K = via the Nearest node
    float* psis;                 // [B][K+1] full path length (root edge included) via that slot, see chain_cost
    float* psi;                  // [B] full path length of the accepted candidate via its committed parent
    uint8_t* ok;
    int32_t *parent, *new_index;
    float* cost;
    unsigned long long* rkey;    // per node rewire proposal key
    long long* counters;         // [0] accepted [1] goal candidate (-1) [2..] stats deltas [10] distinct table samples
    // Nearest runs on the DISTINCT sample points of the batch: slots 0..n_fixed-1 = goal (+ the 3
    // milestones in MS mode), then the table samples, compacted; qslot[c] = candidate c's slot
    float* uq;
    int32_t* qslot;
    const int32_t* u_near;
    int n_fixed;
};

__device__ __forceinline__ int64_t tree_n(const PlanDev& d) { return d.n_dev ? (int64_t)*d.n_dev : d.n; }

// ------------------------------------------------------------------ MT19937 on the device
__device__ __forceinline__ uint32_t mt_twist(uint32_t u, uint32_t v) {   // mtrand.h:77-82
    return (((u & 0x80000000u) | (v & 0x7FFFFFFFu)) >> 1) ^ ((v & 1u) ? 0x9908B0DFu : 0u);
}
__device__ __forceinline__ uint32_t mt_temper(uint32_t x) {              // mtrand.h:84-93
    x ^= (x >> 11);
    x ^= (x << 7) & 0x9D2C5680u;
    x ^= (x << 15) & 0xEFC60000u;
    return x ^ (x >> 18);
}
// state: [0..623] words, [624] position.  Produces n tempered words in stream order.
// s = the state in shared memory, pos = its position; every thread of the CTA calls this (barriers
// inside); returns the new position.  out may be global or shared memory.
__device__ __forceinline__ int mt_generate(uint32_t* s, int pos, uint32_t* out, int n) {
    const int tid = threadIdx.x;
    int done = 0;
    while (done < n) {
        if (pos == kMtN) {                    // gen_state (mtrand.cpp:12-19), three dependent phases
            uint32_t v = 0;
            // phase 1: i in [0,227): needs old s[i], s[i+1], s[i+397]
            if (tid < kMtN - kMtM) v = s[tid + kMtM] ^ mt_twist(s[tid], s[tid + 1]);
            __syncthreads();
            if (tid < kMtN - kMtM) s[tid] = v;
            __syncthreads();
            // phase 2: i in [227,454): needs old s[i], s[i+1] and NEW s[i-227]
            for (int half = 0; half < 2; ++half) {
                const int lo = (kMtN - kMtM) + half * (kMtN - kMtM);           // 227, 454
                const int hi = min(lo + (kMtN - kMtM), kMtN - 1);             // 454, 623
                const int i = lo + tid;
                if (i < hi) v = s[i - (kMtN - kMtM)] ^ mt_twist(s[i], s[i + 1]);
                __syncthreads();
                if (i < hi) s[i] = v;
                __syncthreads();
            }
            if (tid == 0) s[kMtN - 1] = s[kMtM - 1] ^ mt_twist(s[kMtN - 1], s[0]);
            __syncthreads();
            pos = 0;
        }
        const int take = min(n - done, kMtN - pos);
        for (int i = tid; i < take; i += blockDim.x) out[done + i] = mt_temper(s[pos + i]);
        done += take;
        pos += take;
        __syncthreads();
    }
    return pos;
}
__global__ void __launch_bounds__(256) mt_fill_kernel(uint32_t* __restrict__ state, uint32_t* __restrict__ out, int n) {
    __shared__ uint32_t s[kMtN];
    const int tid = threadIdx.x;
    for (int i = tid; i < kMtN; i += blockDim.x) s[i] = state[i];
    int pos = (int)state[kMtN];
    __syncthreads();
    pos = mt_generate(s, pos, out, n);
    for (int i = tid; i < kMtN; i += blockDim.x) state[i] = s[i];
    if (tid == 0) state[kMtN] = (uint32_t)pos;
}

// libstdc++ uniform_real_distribution<double>(a,b) over mt19937 (MS:1171-1184):
// generate_canonical<double,53> = (w0 + w1*2^32) / 2^64, clamped below 1
__device__ __forceinline__ double std_uniform(uint32_t w0, uint32_t w1, double a, double b) {
    double sum = __dadd_rn(0.0, __dmul_rn((double)w0, 1.0));
    sum = __dadd_rn(sum, __dmul_rn((double)w1, 4294967296.0));
    double r = sum / 18446744073709551616.0;
    if (r >= 1.0) r = 0.99999999999999988897769753748;   // nextafter(1, 0)
    return __dadd_rn(__dmul_rn(r, b - a), a);
}

// Norm(x1,x2) MS:130-134: f32 differences, f64 squares / sum ((x^2+y^2)+z^2) / root, f32 result
__device__ __forceinline__ float norm_ref(float ax, float ay, float az, float bx, float by, float bz) {
    const double dx = (double)__fsub_rn(bx, ax), dy = (double)__fsub_rn(by, ay), dz = (double)__fsub_rn(bz, az);
    const double s = __dadd_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)), __dmul_rn(dz, dz));
    return __double2float_rn(__dsqrt_rn(s));
}

// ------------------------------------------------------------------ per-candidate kernels
// whole warps call this (c >= B: a lane that only takes part in the ballots)
__device__ __forceinline__ void plan_sample_body(const PlanDev& d, const int c, const int lane) {   // MS:1182-1219 | RS:1155-1178
    const bool live = c < d.B;
    const float* src = d.p.goal;
    int slot = 0;                                               // goal
    bool from_table = false;
    if (live) {
        const uint32_t* w = d.words + (size_t)c * d.W;
        const float rand_chance = __double2float_rn(std_uniform(w[0], w[1], 0.0, 1.0));
        long long ri = (long long)(float)(int)std_uniform(w[2], w[3], 0.0, (double)d.nt);   // float random_indice = int(x(gen))
        if (ri > d.nt - 1) ri = d.nt - 1;
        if (d.p.mode == BGR_PLAN_MODE_MS) {
            int mi = (int)std_uniform(w[4], w[5], 0.0, 3.0);
            if (mi > 2) mi = 2;
            if ((double)rand_chance > 0.8) { src = d.table + 3 * ri; from_table = true; }
            else if ((double)rand_chance < 0.8 && (double)rand_chance > 0.5) src = d.p.goal;
            else { src = d.milestones + 3 * mi; slot = 1 + mi; }
        } else {
            if (rand_chance > d.p.goal_chance) { src = d.table + 3 * ri; from_table = true; }
        }
    }
    // table samples take the next free slots (one atomic per warp); goal / milestone samples of the
    // whole batch share their fixed slot: one Nearest query instead of thousands of identical ones
    const unsigned tm = __ballot_sync(kFull, from_table);
    if (tm) {
        int base = 0;
        if (lane == __ffs(tm) - 1) base = (int)atomicAdd((unsigned long long*)&d.counters[10], (unsigned long long)__popc(tm));
        base = __shfl_sync(kFull, base, __ffs(tm) - 1);
        if (from_table) slot = d.n_fixed + base + __popc(tm & ((1u << lane) - 1u));
    }
    if (!live) return;
    d.prand[3 * c] = src[0];
    d.prand[3 * c + 1] = src[1];
    d.prand[3 * c + 2] = src[2];
    d.qslot[c] = slot;
    if (!from_table) atomicMax((unsigned long long*)&d.counters[11 + slot], (unsigned long long)(d.B - c));   // first candidate of the slot
    if (from_table) {
        d.uq[3 * slot] = src[0];
        d.uq[3 * slot + 1] = src[1];
        d.uq[3 * slot + 2] = src[2];
    }
    if (c < d.n_fixed) {                                        // the fixed slots (B >= 1: thread 0 writes all of them)
        for (int f = c; f < d.n_fixed; f += d.B) {
            const float* fs = f == 0 ? d.p.goal : d.milestones + 3 * (f - 1);
            d.uq[3 * f] = fs[0]; d.uq[3 * f + 1] = fs[1]; d.uq[3 * f + 2] = fs[2];
        }
    }
}
__global__ void plan_sample_kernel(PlanDev d) { plan_sample_body(d, blockIdx.x * blockDim.x + threadIdx.x, threadIdx.x & 31); }

// BGR_PLAN_DEDUP: copies of a goal / milestone candidate are rejected like a failed state check
__device__ __forceinline__ void plan_dedup_body(const PlanDev& d, const int c) {
    if (c >= d.B) return;
    const int slot = d.qslot[c];
    if (slot < d.n_fixed && (long long)(d.B - c) != d.counters[11 + slot]) d.hit[c] = 1;
}
__global__ void plan_dedup_kernel(PlanDev d) { plan_dedup_body(d, blockIdx.x * blockDim.x + threadIdx.x); }

// The goal / milestone candidates of a batch that share a sample slot share the nearest node, the
// steered point and therefore the neighbour rows: the search runs for the first of them only
// (kskip marks the others) and plan_knn_spread_kernel copies its rows to the copies.
__global__ void plan_knn_rep_kernel(PlanDev d, uint8_t* __restrict__ kskip) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= d.B) return;
    const int slot = d.qslot[c];
    kskip[c] = (d.hit[c] || (slot < d.n_fixed && (long long)(d.B - c) != d.counters[11 + slot])) ? 1 : 0;
}

__global__ void plan_knn_spread_kernel(PlanDev d, float* __restrict__ knn_d2) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= (int64_t)d.B * d.K) return;
    const int c = (int)(t / d.K), j = (int)(t - (int64_t)c * d.K);
    if (d.hit[c]) return;
    const int slot = d.qslot[c];
    if (slot >= d.n_fixed) return;
    const int first = d.B - (int)d.counters[11 + slot];
    if (first == c) return;
    d.knn_idx[(size_t)c * d.K + j] = d.knn_idx[(size_t)first * d.K + j];
    knn_d2[(size_t)c * d.K + j] = knn_d2[(size_t)first * d.K + j];
}

__device__ __forceinline__ void plan_steer_body(const PlanDev& d, const int c) {   // MS:1233-1268
    if (c >= d.B) return;
    const int32_t ni = d.u_near[d.qslot[c]];
    d.near[c] = ni;
    const float nx = d.x[ni], ny = d.y[ni], nz = d.z[ni];
    const float rx = d.prand[3 * c], ry = d.prand[3 * c + 1], rz = d.prand[3 * c + 2];
    float ox, oy, oz;
    uint8_t found;
    if (norm_ref(nx, ny, nz, rx, ry, rz) <= d.p.eta) {
        ox = rx; oy = ry; oz = rz;
        found = 1;
    } else {
        const float dx = __fsub_rn(rx, nx), dy = __fsub_rn(ry, ny), dz = __fsub_rn(rz, nz);
        const double s = __dadd_rn(__dadd_rn(__dmul_rn((double)dz, (double)dz), __dmul_rn((double)dy, (double)dy)),
                                   __dmul_rn((double)dx, (double)dx));               // MS:1246: z^2 + y^2 + x^2
        const float mag = __double2float_rn(__dsqrt_rn(s));
        ox = __fadd_rn(nx, __fmul_rn(d.p.eta, __fdiv_rn(dx, mag)));
        oy = __fadd_rn(ny, __fmul_rn(d.p.eta, __fdiv_rn(dy, mag)));
        oz = __fadd_rn(nz, __fmul_rn(d.p.eta, __fdiv_rn(dz, mag)));
        found = 0;
    }
    d.pnew[3 * c] = ox; d.pnew[3 * c + 1] = oy; d.pnew[3 * c + 2] = oz;
    d.found[c] = found;
}
__global__ void plan_steer_kernel(PlanDev d) { plan_steer_body(d, blockIdx.x * blockDim.x + threadIdx.x); }

// candidate parent of slot j of candidate c: j < K the j-th neighbour, j == K the Nearest node
__device__ __forceinline__ int32_t slot_node(const PlanDev& d, int c, int j) {
    if (j == d.K) return d.near[c];
    return d.have_knn ? d.knn_idx[(size_t)c * d.K + j] : -1;
}

// The goal / milestone samples of a batch are the same point, hence — against one snapshot — the same candidate:
// same nearest node, steered point, state check, neighbour rows, candidate edges and chain costs.  rep_of(c) = the
// first candidate of c's sample slot; the edge checks and the parent-chain walks run for it only and its copies
// read its rows (in MS mode 80 % of a batch are such copies: 5x fewer walks and edges, identical results).
__device__ __forceinline__ int rep_of(const PlanDev& d, int c) {
    const int slot = d.qslot[c];
    return slot < d.n_fixed ? d.B - (int)d.counters[11 + slot] : c;
}

__global__ void plan_edges_kernel(PlanDev d) {                 // edge list for isLineValid geometry
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= d.B * (d.K + 1)) return;
    const int c = t / (d.K + 1), j = t % (d.K + 1);
    const int32_t ind = (d.hit[c] || rep_of(d, c) != c) ? -1 : slot_node(d, c, j);
    float a[3], b[3];
    if (ind >= 0) {
        a[0] = d.x[ind]; a[1] = d.y[ind]; a[2] = d.z[ind];
        b[0] = d.pnew[3 * c]; b[1] = d.pnew[3 * c + 1]; b[2] = d.pnew[3 * c + 2];
    } else {                                                   // unused slot: an edge far outside every map
        a[0] = a[1] = a[2] = -1.0e9f;
        b[0] = b[1] = -1.0e9f; b[2] = -1.0e9f + 128.f;
    }
    for (int k = 0; k < 3; ++k) { d.e1[3 * (size_t)t + k] = a[k]; d.e2[3 * (size_t)t + k] = b[k]; }
}

// cost_wiring(tree, tree[P], child) MS:802-831: f32 accumulation from the child upwards; the
// walk ends at index <= 0, so the edge into the root is never added.
// *psi (optional) = the same sum PLUS that skipped root edge: the true path length root -> child.
// It never increases along a parent pointer (float addition is monotone), which is what keeps
// batched rewiring acyclic (rewire_proposal).
__device__ float chain_cost(const PlanDev& d, int32_t P, float cx, float cy, float cz, float* psi = nullptr) {
    const float px = d.x[P], py = d.y[P], pz = d.z[P];
    float cost = __fadd_rn(0.f, norm_ref(px, py, pz, cx, cy, cz));
    int cur = d.par[P];
    float lx = px, ly = py, lz = pz;                           // previous node on the walk (P first)
    int64_t guard = tree_n(d) + d.B + 2;                       // a tree walk visits every node at most once
    while (cur > 0 && guard-- > 0) {
        const float qx = d.x[cur], qy = d.y[cur], qz = d.z[cur];
        cost = __fadd_rn(cost, norm_ref(qx, qy, qz, lx, ly, lz));
        lx = qx; ly = qy; lz = qz;
        cur = d.par[cur];
    }
    if (psi) *psi = cur == 0 ? __fadd_rn(cost, norm_ref(d.x[0], d.y[0], d.z[0], lx, ly, lz)) : cost;
    return cost;
}

__device__ __forceinline__ void plan_cost_body(const PlanDev& d, const int t) {
    if (t >= d.B * (d.K + 1)) return;
    const int c = t / (d.K + 1), j = t % (d.K + 1);
    if (d.hit[c] || rep_of(d, c) != c) return;
    const int32_t ind = slot_node(d, c, j);
    if (ind < 0) return;
    d.costs[t] = chain_cost(d, ind, d.pnew[3 * c], d.pnew[3 * c + 1], d.pnew[3 * c + 2], &d.psis[t]);
}
__global__ void plan_cost_kernel(PlanDev d) { plan_cost_body(d, blockIdx.x * blockDim.x + threadIdx.x); }

__device__ __forceinline__ void plan_select_body(const PlanDev& d, const int c) {   // extend_tree MS:838-853 + MS:1322-1323
    if (c >= d.B) return;
    const bool edge_check = d.p.flags & BGR_PLAN_EDGE_CHECK, quirks = d.p.flags & BGR_PLAN_QUIRKS;
    const int K1 = d.K + 1;
    const size_t r1 = (size_t)rep_of(d, c) * K1, rk = (size_t)rep_of(d, c) * d.K;   // rows of the slot's first candidate
    d.ok[c] = 0;
    long long band = d.band[c];
    if (edge_check && !d.hit[c])
        for (int j = 0; j < K1; ++j)
            if ((j == d.K ? d.near[c] : (d.have_knn ? d.knn_idx[rk + j] : -1)) >= 0) band += d.eband[r1 + j];
    if (band) atomicAdd((unsigned long long*)&d.counters[8], (unsigned long long)band);
    if (d.hit[c]) { atomicAdd((unsigned long long*)&d.counters[4], 1ull); return; }
    if (edge_check && d.ehit[r1 + d.K]) { atomicAdd((unsigned long long*)&d.counters[5], 1ull); return; }
    int32_t parent_index = d.near[c];
    float lowest = d.costs[r1 + d.K];                         // current_cost MS:1316
    float mycost = lowest;
    int slot = d.K, pslot = d.K;
    if (d.have_knn) {
        int32_t pc = -5;
        for (int j = 0; j < d.K; ++j) {
            // an unfilled slot (fewer than K admissible nodes: the self-match rule drops the copies that
            // coincide with the candidate) has no cost: plan_cost_kernel left costs[] untouched for it
            const int32_t ind = d.knn_idx[rk + j];
            if (ind < 0) continue;
            if (edge_check && d.ehit[r1 + j]) continue;
            const float temp = d.costs[r1 + j];
            if (temp < lowest) { pc = ind; lowest = temp; pslot = j; }
        }
        if (pc > 0) {
            parent_index = pc;
            mycost = lowest;
            slot = pslot;
            atomicAdd((unsigned long long*)&d.counters[6], 1ull);
        }
    }
    const bool keep_near = quirks && d.found[c];
    d.parent[c] = keep_near ? d.near[c] : parent_index;
    d.cost[c] = mycost;
    d.psi[c] = d.psis[r1 + (keep_near ? d.K : slot)];          // follows the pointer actually committed
    d.ok[c] = 1;
}
__global__ void plan_select_kernel(PlanDev d) { plan_select_body(d, blockIdx.x * blockDim.x + threadIdx.x); }

// Fused forms of the host-driven step (a small step costs ~5 us per DEPENDENT kernel, whatever the kernel does):
// one warp per candidate.  plan_state_kernel = steer, sphere state check, dedup and the kNN skip flag;
// plan_cost_select_kernel = the K+1 chain costs of the candidate on K+1 lanes, then its parent choice.
__global__ void __launch_bounds__(128) plan_state_kernel(PlanDev d, MapView m, LevelGrid dil, uint8_t* kskip) {
    const int c = (int)((blockIdx.x * blockDim.x + threadIdx.x) >> 5), lane = threadIdx.x & 31;
    if (c >= d.B) return;
    if (lane == 0) plan_steer_body(d, c);
    __syncwarp();
    collide_spheres_warp<1>(m, d.pnew, d.B, d.p.state_radius, d.hit, d.band, dil, c, lane);
    __syncwarp();
    if (lane == 0) {
        if (d.p.flags & BGR_PLAN_DEDUP) plan_dedup_body(d, c);
        if (d.have_knn) {
            const int slot = d.qslot[c];
            kskip[c] = (d.hit[c] || (slot < d.n_fixed && (long long)(d.B - c) != d.counters[11 + slot])) ? 1 : 0;
        }
    }
}
__global__ void __launch_bounds__(128) plan_cost_select_kernel(PlanDev d) {
    const int c = (int)((blockIdx.x * blockDim.x + threadIdx.x) >> 5), lane = threadIdx.x & 31;
    if (c >= d.B) return;
    for (int j = lane; j < d.K + 1; j += 32) plan_cost_body(d, c * (d.K + 1) + j);
    __syncwarp();
    if (lane == 0) plan_select_body(d, c);
}

// One CTA: goal cut (first accepted candidate within goal_tol, MS:1340), ordered compaction,
// append to the store (MS:1325-1327) and the goal node (MS:1346-1347).
// every thread of ONE CTA calls this (barriers inside)
__device__ __forceinline__ void plan_commit_body(const PlanDev& d, float* x, float* y, float* z, float4* f4) {
    __shared__ int s_goal;
    __shared__ int warp_tot[32];
    __shared__ int s_base;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const bool no_goal = d.p.flags & BGR_PLAN_NO_GOAL;
    if (tid == 0) { s_goal = 0x7fffffff; s_base = 0; }
    __syncthreads();
    if (!no_goal)
        for (int c = tid; c < d.B; c += blockDim.x)
            if (d.ok[c] && norm_ref(d.p.goal[0], d.p.goal[1], d.p.goal[2], d.pnew[3 * c], d.pnew[3 * c + 1], d.pnew[3 * c + 2]) < d.p.goal_tol)
                atomicMin(&s_goal, c);
    __syncthreads();
    const int goal_c = s_goal;
    for (int c0 = 0; c0 < d.B; c0 += blockDim.x) {
        const int c = c0 + tid;
        const int take = (c < d.B && d.ok[c] && c <= goal_c) ? 1 : 0;
        int inc = take;
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(kFull, inc, o);
            if (lane >= o) inc += t;
        }
        if (lane == 31) warp_tot[warp] = inc;
        __syncthreads();
        if (warp == 0) {
            int w = warp_tot[lane];
            for (int o = 1; o < 32; o <<= 1) {
                const int t = __shfl_up_sync(kFull, w, o);
                if (lane >= o) w += t;
            }
            warp_tot[lane] = w;
        }
        __syncthreads();
        const int base = s_base;
        const int rank = base + (warp ? warp_tot[warp - 1] : 0) + inc - take;
        if (c < d.B) d.new_index[c] = take ? (int32_t)(tree_n(d) + rank) : -1;
        if (take) {
            const int64_t ni = tree_n(d) + rank;
            x[ni] = d.pnew[3 * c]; y[ni] = d.pnew[3 * c + 1]; z[ni] = d.pnew[3 * c + 2];
            f4[ni] = filter_image(d.pnew[3 * c], d.pnew[3 * c + 1], d.pnew[3 * c + 2]);
            d.par[ni] = d.parent[c];
        }
        __syncthreads();
        if (tid == 0) s_base = base + warp_tot[31];
        __syncthreads();
    }
    if (tid == 0) {
        const int acc = s_base;
        d.counters[0] = acc;
        d.counters[1] = goal_c == 0x7fffffff ? -1 : goal_c;
        if (goal_c != 0x7fffffff) {
            const int64_t gi = tree_n(d) + acc;
            x[gi] = d.p.goal[0]; y[gi] = d.p.goal[1]; z[gi] = d.p.goal[2];
            f4[gi] = filter_image(d.p.goal[0], d.p.goal[1], d.p.goal[2]);
            d.par[gi] = (int32_t)(gi - 1);                      // goal.push_back(V.size()-1)
        }
    }
}
__global__ void __launch_bounds__(1024) plan_commit_kernel(PlanDev d, float* x, float* y, float* z, float4* f4) {
    plan_commit_body(d, x, y, z, f4);
}

// rewiring (beyond reference): neighbour `ind` adopts the new node when
// cost(new) + Norm(tree[ind], new) < cost(ind) (and the edge is free); one proposal per node
// survives: the smallest (new cost bits, new index) key.
// Acyclicity: the reference's cost skips the edge into the root (cost_wiring stops at index <= 0),
// so a depth-1 node can look more expensive than its own descendants, and proposals of one batch
// are applied together.  A proposal must therefore also lower the TRUE path length:
// psi(new) < psi(ind), both on the tree as it stands before this batch's rewires.  psi never
// increases along any kept pointer and strictly decreases along every rewired one, so no set of
// simultaneously applied proposals can close a cycle.
__device__ __forceinline__ bool rewire_proposal(const PlanDev& d, int t, int32_t* ind_out, unsigned long long* key) {
    const int c = t / d.K, j = t % d.K;
    const int32_t ni = d.new_index[c];
    if (ni < 0) return false;
    // a copy of a goal / milestone candidate proposes the same new cost with a larger new index than the slot's first
    // candidate (accepted whenever a copy is): its key can never be the minimum
    if (rep_of(d, c) != c) return false;
    const int32_t ind = d.knn_idx[(size_t)c * d.K + j];
    if (ind <= 0 || ind == d.par[ni]) return false;
    if ((d.p.flags & BGR_PLAN_EDGE_CHECK) && d.ehit[(size_t)c * (d.K + 1) + j]) return false;
    const float ix = d.x[ind], iy = d.y[ind], iz = d.z[ind];
    const float nc = __fadd_rn(d.cost[c], norm_ref(ix, iy, iz, d.pnew[3 * c], d.pnew[3 * c + 1], d.pnew[3 * c + 2]));
    float psi_ind;
    const float oc = chain_cost(d, d.par[ind], ix, iy, iz, &psi_ind);
    if (!(nc < oc)) return false;
    if (!(d.psi[c] < psi_ind)) return false;
    *ind_out = ind;
    *key = ((unsigned long long)__float_as_uint(nc) << 32) | (unsigned)ni;
    return true;
}
__device__ __forceinline__ void plan_rewire_propose_body(const PlanDev& d, const int t) {
    if (t >= d.B * d.K) return;
    int32_t ind;
    unsigned long long key;
    if (rewire_proposal(d, t, &ind, &key)) atomicMin(&d.rkey[ind], key);
}
__global__ void plan_rewire_propose_kernel(PlanDev d) { plan_rewire_propose_body(d, blockIdx.x * blockDim.x + threadIdx.x); }
__device__ __forceinline__ void plan_rewire_apply_body(const PlanDev& d, const int t) {
    if (t >= d.B * d.K) return;
    const int c = t / d.K, j = t % d.K;
    const int32_t ni = d.new_index[c];
    if (ni < 0) return;
    const int32_t ind = d.knn_idx[(size_t)c * d.K + j];
    if (ind <= 0 || ind >= tree_n(d)) return;
    const unsigned long long key = d.rkey[ind];
    if (key == ~0ull || (unsigned)(key & 0xffffffffu) != (unsigned)ni) return;
    // several slots of the same candidate can name the same node only once (kNN lists are sets)
    d.par[ind] = ni;
    atomicAdd((unsigned long long*)&d.counters[7], 1ull);
}
__global__ void plan_rewire_apply_kernel(PlanDev d, const int32_t* __restrict__ par_snapshot_unused) {
    plan_rewire_apply_body(d, blockIdx.x * blockDim.x + threadIdx.x);
}
__device__ __forceinline__ void plan_rewire_clear_body(const PlanDev& d, const int t) {
    if (t >= d.B * d.K) return;
    const int32_t ind = d.knn_idx[(size_t)t];
    if (ind > 0 && ind < tree_n(d)) d.rkey[ind] = ~0ull;
}
__global__ void plan_rewire_clear_kernel(PlanDev d) { plan_rewire_clear_body(d, blockIdx.x * blockDim.x + threadIdx.x); }

// ------------------------------------------------------------------ persistent small-batch steps
// A step of a small batch is ~15 tiny dependent launches and a host round trip (88 us at batch 1: the
// reference's sequential loop).  For batches of at most kPersistMaxB candidates without edge checks, on trees
// below kPersistMaxNodes nodes, ONE CTA runs whole steps back to back inside a single launch: the same device
// functions as the per-step kernels above (sample, steer, state check, cost, select, commit, rewire), a
// CTA-wide Nearest and a CTA-wide exact kNN in place of the library launches, barriers in place of kernel
// boundaries, the MT19937 state kept in shared memory.  The tree it grows is the same, bit for bit
// (tests/test_gpu_planner.py); the host reads one control block back per launch instead of per step.
constexpr int kPersistMaxB = 8;
constexpr int kPersistMaxNodes = 65536;
constexpr int kPersistThreads = 1024;

struct PersistCtl {
    long long n;          // tree size (in / out)
    long long steps;      // steps run by this launch (out)
    long long goal;       // goal candidate of the last step, or -1 (out)
    long long stats[5];   // sums over the launch: rejected by state / by edge, parent improved, rewired, in-band events
    long long unique;     // distinct table samples of the last step
};

// Nearest (MS:138-154) over the whole tree for one query: every thread of the CTA calls this
__device__ void nearest_cta(const PlanDev& d, const float* q, int32_t* out_i, float* s_f, int32_t* s_i) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const float qx = q[0], qy = q[1], qz = q[2];
    float bf = __uint_as_float(kInfBits);
    int32_t bi = -1;
    for (int64_t i = tid; i < d.n; i += blockDim.x) {           // ascending: `<=` keeps the last index of equal Norms
        const float f = norm_ref(d.x[i], d.y[i], d.z[i], qx, qy, qz);
        if (f <= bf) { bf = f; bi = (int32_t)i; }
    }
    auto better = [](float of, int32_t oi, float f, int32_t i) { return oi >= 0 && (i < 0 || of < f || (of == f && oi > i)); };
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float of = __shfl_xor_sync(kFull, bf, o);
        const int32_t oi = __shfl_xor_sync(kFull, bi, o);
        if (better(of, oi, bf, bi)) { bf = of; bi = oi; }
    }
    if (lane == 0) { s_f[warp] = bf; s_i[warp] = bi; }
    __syncthreads();
    if (warp == 0) {
        const int nw = blockDim.x >> 5;
        bf = lane < nw ? s_f[lane] : __uint_as_float(kInfBits);
        bi = lane < nw ? s_i[lane] : -1;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float of = __shfl_xor_sync(kFull, bf, o);
            const int32_t oi = __shfl_xor_sync(kFull, bi, o);
            if (better(of, oi, bf, bi)) { bf = of; bi = oi; }
        }
        if (lane == 0) *out_i = bi;
    }
    __syncthreads();
}

// exact kNN of candidate c over the whole tree (libnabo semantics of the planner's call: no self match,
// no radius limit, (d2, index) order): every thread of the CTA calls this
__device__ void knn_cta(const PlanDev& d, const int c, float* knn_d2, uint32_t (*sd)[32], uint32_t (*si)[32]) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nw = blockDim.x >> 5;
    const float qx = d.pnew[3 * c], qy = d.pnew[3 * c + 1], qz = d.pnew[3 * c + 2];
    const uint32_t sent = kInfBits;                               // bits(FLT_MAX) + 1: every finite d2 is below it
    WarpList L;
    L.init(sent);
    uint32_t tau_d = sent, tau_i = 0u;
    float tau_f = __uint_as_float(sent);
    for (int64_t j0 = (int64_t)warp * 32; j0 < d.n; j0 += blockDim.x) {
        const int64_t j = j0 + lane;
        const bool valid = j < d.n;
        const float px = valid ? d.x[j] : 0.f, py = valid ? d.y[j] : 0.f, pz = valid ? d.z[j] : 0.f;
        const float dx = __fsub_rn(qx, px), dy = __fsub_rn(qy, py), dz = __fsub_rn(qz, pz);
        const float dd = __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
        const bool pass = valid && (dd <= tau_f) && (dd > FLT_EPSILON);
        if (__any_sync(kFull, pass)) {
            warp_offer(L, pass, __float_as_uint(dd), (uint32_t)j, tau_d, tau_i, d.K, lane);
            tau_f = __uint_as_float(tau_d);
        }
    }
    sd[warp][lane] = L.d;
    si[warp][lane] = L.i;
    __syncthreads();
    if (warp == 0) {
        for (int w = 1; w < nw; ++w) {
            const uint32_t cd = sd[w][lane], ci = si[w][lane];
            const bool pass = lane < d.K && cd != sent;
            if (__any_sync(kFull, pass)) warp_offer(L, pass, cd, ci, tau_d, tau_i, d.K, lane);
        }
        if (lane < d.K) {
            const bool empty = L.d == sent;
            d.knn_idx[(size_t)c * d.K + lane] = empty ? -1 : (int32_t)L.i;
            knn_d2[(size_t)c * d.K + lane] = empty ? __uint_as_float(kInfBits) : __uint_as_float(L.d);
        }
    }
    __syncthreads();
}

__global__ void __launch_bounds__(kPersistThreads) plan_persist_kernel(PlanDev d0, float* x, float* y, float* z, float4* f4,
                                                                       uint32_t* mt_state, const uint32_t* first_words,
                                                                       float* knn_d2, float* u_dist_unused, int32_t* u_near,
                                                                       MapView m, LevelGrid dil, int max_steps, long long max_nodes,
                                                                       PersistCtl* ctl) {
    __shared__ PlanDev d;
    __shared__ uint32_t mts[kMtN];
    __shared__ uint32_t words[kPersistMaxB * 6];
    __shared__ uint32_t sd[kPersistThreads / 32][32], si[kPersistThreads / 32][32];
    __shared__ float s_f[32];
    __shared__ int32_t s_i[32];
    __shared__ int s_full, s_goal_hit;   // two flags: each is rewritten only many barriers after its last read
    __shared__ long long s_stats[5];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) {
        d = d0;
        d.n = ctl->n;
        d.words = words;
        d.u_near = u_near;
        for (int i = 0; i < 5; ++i) s_stats[i] = 0;
    }
    for (int i = tid; i < kMtN; i += blockDim.x) mts[i] = mt_state[i];
    int mt_pos = (int)mt_state[kMtN];
    __syncthreads();
    const bool rewire = d.p.flags & BGR_PLAN_REWIRE;
    const int B = d.B, K = d.K, K1 = K + 1;
    int steps = 0;
    long long goal = -1, unique = 0;
    for (; steps < max_steps; ++steps) {
        if (tid == 0) {
            s_full = d.n >= max_nodes;
            d.have_knn = d.n > K;                                  // MS:1273
            for (int i = 0; i < 16; ++i) d.counters[i] = 0;
        }
        __syncthreads();
        if (s_full) break;
        // sample-stream words of this step (the step the host had prepared ahead comes first)
        if (steps == 0 && first_words) {
            for (int i = tid; i < B * d.W; i += blockDim.x) words[i] = first_words[i];
            __syncthreads();
        } else {
            mt_pos = mt_generate(mts, mt_pos, words, B * d.W);
        }
        if (warp == 0) plan_sample_body(d, lane, lane);            // B <= 32: one warp
        __syncthreads();
        const int nq_u = d.n_fixed + (int)d.counters[10];
        for (int u = 0; u < nq_u; ++u) nearest_cta(d, d.uq + 3 * u, u_near + u, s_f, s_i);
        if (tid < B) plan_steer_body(d, tid);
        __syncthreads();
        if (warp < B) collide_spheres_warp<1>(m, d.pnew, B, d.p.state_radius, d.hit, d.band, dil, warp, lane);
        __syncthreads();
        if ((d.p.flags & BGR_PLAN_DEDUP) && tid < B) plan_dedup_body(d, tid);
        __syncthreads();
        if (d.have_knn)
            for (int c = 0; c < B; ++c)
                if (!d.hit[c]) knn_cta(d, c, knn_d2, sd, si);      // d.hit[c] is CTA-uniform
        for (int t = tid; t < B * K1; t += blockDim.x) plan_cost_body(d, t);
        __syncthreads();
        if (tid < B) plan_select_body(d, tid);
        __syncthreads();
        plan_commit_body(d, x, y, z, f4);
        __syncthreads();
        if (rewire && d.have_knn) {
            for (int t = tid; t < B * K; t += blockDim.x) plan_rewire_propose_body(d, t);
            __syncthreads();
            for (int t = tid; t < B * K; t += blockDim.x) plan_rewire_apply_body(d, t);
            __syncthreads();
            for (int t = tid; t < B * K; t += blockDim.x) plan_rewire_clear_body(d, t);
            __syncthreads();
        }
        if (tid == 0) {
            goal = d.counters[1];
            unique = d.counters[10];
            d.n += d.counters[0] + (goal >= 0 ? 1 : 0);
            for (int i = 0; i < 5; ++i) s_stats[i] += d.counters[4 + i];
            s_goal_hit = goal >= 0;
        }
        __syncthreads();
        if (s_goal_hit) { ++steps; break; }
    }
    for (int i = tid; i < kMtN; i += blockDim.x) mt_state[i] = mts[i];
    if (tid == 0) {
        mt_state[kMtN] = (uint32_t)mt_pos;
        ctl->n = d.n;
        ctl->steps = steps;
        ctl->goal = goal;
        ctl->unique = unique;
        for (int i = 0; i < 5; ++i) ctl->stats[i] = s_stats[i];
    }
}

void mt_seed_host(uint32_t* s, uint32_t seed) {               // mtrand.cpp:21-31
    s[0] = seed;
    for (int i = 1; i < kMtN; ++i) s[i] = 1812433253u * (s[i - 1] ^ (s[i - 1] >> 30)) + (uint32_t)i;
    s[kMtN] = kMtN;
}

}  // namespace

struct bgr_planner {
    bgr_planner_params p;
    bgr_map* map = nullptr;
    bgr_nns* nns = nullptr;
    cudaStream_t stream = nullptr;
    int64_t nt = 0;
    int nm = 0;
    int goal_reached = 0;
    int64_t stats[8] = {0, 0, 0, 0, 0, 0, 0, -1};
    DevBuf<uint32_t> mt_state, words, words2;   // sample-stream words: this step's and the next one's (filled ahead)
    int words_cur = 0;
    bool words_ahead = false;                   // the current buffer was filled at the end of the previous step
    cudaEvent_t ev_step = nullptr;              // recorded after the counters copy of the step in flight
    DevBuf<float> table, milestones, prand, pnew, near_dist, knn_d2, e1, e2, costs, cost, psis, psi;
    DevBuf<uint8_t> found, hit, kskip, band, ehit, eband, ok;
    DevBuf<int32_t> par, near, knn_idx, parent, new_index, qslot, u_near;
    DevBuf<float> uq, u_dist;
    int64_t pending_n = 0;     // tree size at the launch of the step in flight
    int64_t unique_hint = 0;   // distinct sample points of the previous batch (sizes the next Nearest scan)
    DevBuf<unsigned long long> rkey;
    int64_t rkey_init = 0;       // entries of rkey already set to ~0
    DevBuf<long long> counters;
    PinBuf host_counters;
    DevBuf<PersistCtl> pctl;     // control block of the persistent small-batch kernel
    PinBuf host_pctl;
    PinBuf host_init;            // counters block a replayed step uploads: zeros + the tree size in word 15
    cudaGraphExec_t gexec[2] = {nullptr, nullptr};   // the captured step, one per sample-word buffer
    uint64_t g_sig = 0;          // what the captured launches depend on (graph_signature)
    bool g_warm = false;
    uint64_t n_graph_steps = 0;
    uint64_t n_persist_launches = 0, n_persist_steps = 0;
};

extern "C" {

static void graph_drop(bgr_planner* pl);

int bgr_planner_destroy(bgr_planner* pl) {
    if (!pl) return BGR_OK;
    cudaStreamSynchronize(pl->stream);
    if (pl->nns) bgr_nns_destroy(pl->nns);
    if (pl->ev_step) cudaEventDestroy(pl->ev_step);
    pl->mt_state.release(); pl->words.release(); pl->words2.release(); pl->table.release(); pl->milestones.release();
    pl->prand.release(); pl->pnew.release(); pl->near_dist.release(); pl->knn_d2.release();
    pl->e1.release(); pl->e2.release(); pl->costs.release(); pl->cost.release(); pl->psis.release(); pl->psi.release();
    pl->found.release(); pl->hit.release(); pl->kskip.release(); pl->band.release(); pl->ehit.release(); pl->eband.release();
    pl->ok.release(); pl->par.release(); pl->near.release(); pl->knn_idx.release(); pl->parent.release();
    pl->new_index.release(); pl->rkey.release(); pl->counters.release(); pl->host_counters.release();
    pl->pctl.release(); pl->host_pctl.release(); pl->host_init.release();
    for (int i = 0; i < 2; ++i)
        if (pl->gexec[i]) cudaGraphExecDestroy(pl->gexec[i]);
    delete pl;
    return BGR_OK;
}

int bgr_planner_create(const bgr_planner_params* p, bgr_map* map, const float* table, int64_t nt,
                       const float* milestones, int nm, bgr_planner** out) {
    BGR_ARG(out != nullptr, "out is null");
    *out = nullptr;
    BGR_ARG(p != nullptr && map != nullptr, "null params / map");
    BGR_ARG(table != nullptr && nt >= 1 && nt <= (1 << 24), "sample table must hold 1..2^24 rows");
    BGR_ARG(p->k >= 1 && p->k <= 32, "k must be in 1..32");
    BGR_ARG(p->batch >= 1 && p->batch <= 65536, "batch must be in 1..65536");
    BGR_ARG(p->mode == BGR_PLAN_MODE_RS || p->mode == BGR_PLAN_MODE_MS, "unknown sampler mode %d", p->mode);
    BGR_ARG(p->mode != BGR_PLAN_MODE_MS || (milestones != nullptr && nm >= 3), "mode MS needs >= 3 milestone points");
    BGR_ARG((p->flags & ~127u) == 0, "unknown planner flags %u", p->flags);
    BGR_ARG(p->eta > 0.f && p->state_radius >= 0 && p->edge_radius >= 0, "bad eta / radii");
    bgr_planner* pl = new bgr_planner;
    pl->p = *p;
    pl->map = map;
    pl->nt = nt;
    pl->nm = nm;
    int rc = bgr_nns_create(p->start, 1, 3, 0, &pl->nns);                     // V = [[xs,ys,zs,-1]] MS:1152-1156
    auto fail = [&](int r) { bgr_planner_destroy(pl); return r; };
    if (rc != BGR_OK) return fail(rc);
    cudaStream_t st = pl->stream;
    if ((rc = pl->table.reserve(3 * nt, st, false)) != BGR_OK) return fail(rc);
    if ((rc = pl->milestones.reserve(3 * std::max(nm, 1), st, false)) != BGR_OK) return fail(rc);
    if ((rc = pl->mt_state.reserve(kMtN + 1, st, false)) != BGR_OK) return fail(rc);
    if ((rc = pl->par.reserve(1024, st, false)) != BGR_OK) return fail(rc);
    if ((rc = pl->counters.reserve(16, st, false)) != BGR_OK) return fail(rc);
    if ((rc = pl->host_counters.reserve(16 * sizeof(long long))) != BGR_OK) return fail(rc);
    uint32_t s[kMtN + 1];
    mt_seed_host(s, p->seed);
    const int32_t root_parent = -1;
    if (cudaMemcpyAsync(pl->table.p, table, (size_t)3 * nt * sizeof(float), cudaMemcpyHostToDevice, st) != cudaSuccess ||
        (nm > 0 && cudaMemcpyAsync(pl->milestones.p, milestones, (size_t)3 * nm * sizeof(float), cudaMemcpyHostToDevice, st) != cudaSuccess) ||
        cudaMemcpyAsync(pl->mt_state.p, s, sizeof(s), cudaMemcpyHostToDevice, st) != cudaSuccess ||
        cudaMemcpyAsync(pl->par.p, &root_parent, sizeof(int32_t), cudaMemcpyHostToDevice, st) != cudaSuccess ||
        cudaStreamSynchronize(st) != cudaSuccess) {
        set_error("planner upload failed: %s", cudaGetErrorString(cudaGetLastError()));
        return fail(BGR_ERR_CUDA);
    }
    *out = pl;
    return BGR_OK;
}

int bgr_planner_set_stream(bgr_planner* pl, void* stream) {
    BGR_ARG(pl != nullptr, "null planner");
    BGR_CUDA(cudaStreamSynchronize(pl->stream));
    graph_drop(pl);
    pl->stream = (cudaStream_t)stream;
    BGR_CHECK(bgr_nns_set_stream(pl->nns, stream));
    BGR_CHECK(bgr_map_set_stream(pl->map, stream));
    return BGR_OK;
}

int64_t bgr_planner_size(const bgr_planner* pl) { return pl ? bgr_nns_size(pl->nns) : -1; }

int bgr_planner_nns(bgr_planner* pl, bgr_nns** out) {
    BGR_ARG(pl != nullptr && out != nullptr, "bad arguments");
    *out = pl->nns;
    return BGR_OK;
}

// One step = step_launch (everything enqueued on the planner's stream, counters copied back
// asynchronously) + step_finish (wait for the stream, host bookkeeping).  bgr_planner_run_many
// launches the steps of several independent planners before it waits for any of them.
// buffers of a step on a tree that may grow to n_cap nodes during it (capacity only)
static int step_reserve(bgr_planner* pl, int64_t n_cap) {
    const bgr_planner_params& P = pl->p;
    const int B = P.batch, K = P.k, K1 = K + 1;
    const int W = P.mode == BGR_PLAN_MODE_MS ? 6 : 4;
    const int64_t N = bgr_nns_size(pl->nns);
    const bool edge_check = P.flags & BGR_PLAN_EDGE_CHECK, rewire = P.flags & BGR_PLAN_REWIRE;
    cudaStream_t st = pl->stream;
    BGR_ARG(n_cap < ((int64_t)1 << 31), "tree too large");
    BGR_CHECK(nns_reserve(pl->nns, n_cap));
    BGR_CHECK(pl->par.reserve(n_cap, st, true, N));
    BGR_CHECK(pl->words.reserve((int64_t)B * W, st, false));
    BGR_CHECK(pl->words2.reserve((int64_t)B * W, st, false));
    BGR_CHECK(pl->prand.reserve(3 * (int64_t)B, st, false));
    BGR_CHECK(pl->pnew.reserve(3 * (int64_t)B, st, false));
    BGR_CHECK(pl->near_dist.reserve(B, st, false));
    BGR_CHECK(pl->found.reserve(B, st, false));
    BGR_CHECK(pl->hit.reserve(B, st, false));
    BGR_CHECK(pl->kskip.reserve(B, st, false));
    BGR_CHECK(pl->band.reserve(B, st, false));
    BGR_CHECK(pl->ok.reserve(B, st, false));
    BGR_CHECK(pl->near.reserve(B, st, false));
    const int n_fixed = P.mode == BGR_PLAN_MODE_MS ? 4 : 1;
    BGR_CHECK(pl->qslot.reserve(B, st, false));
    BGR_CHECK(pl->u_near.reserve(B + n_fixed, st, false));
    BGR_CHECK(pl->u_dist.reserve(B + n_fixed, st, false));
    BGR_CHECK(pl->uq.reserve(3 * (int64_t)(B + n_fixed), st, false));
    BGR_CHECK(pl->parent.reserve(B, st, false));
    BGR_CHECK(pl->new_index.reserve(B, st, false));
    BGR_CHECK(pl->cost.reserve(B, st, false));
    BGR_CHECK(pl->knn_idx.reserve((int64_t)B * K, st, false));
    BGR_CHECK(pl->knn_d2.reserve((int64_t)B * K, st, false));
    BGR_CHECK(pl->costs.reserve((int64_t)B * K1, st, false));
    BGR_CHECK(pl->psis.reserve((int64_t)B * K1, st, false));
    BGR_CHECK(pl->psi.reserve(B, st, false));
    if (edge_check) {
        BGR_CHECK(pl->e1.reserve(3 * (int64_t)B * K1, st, false));
        BGR_CHECK(pl->e2.reserve(3 * (int64_t)B * K1, st, false));
        BGR_CHECK(pl->ehit.reserve((int64_t)B * K1, st, false));
        BGR_CHECK(pl->eband.reserve((int64_t)B * K1, st, false));
    }
    if (rewire) {
        const int64_t old_cap = pl->rkey.cap;
        BGR_CHECK(pl->rkey.reserve(n_cap, st, true, pl->rkey_init));
        if (pl->rkey.cap != old_cap || pl->rkey_init < pl->rkey.cap) {
            BGR_CUDA(cudaMemsetAsync(pl->rkey.p + pl->rkey_init, 0xFF, (size_t)(pl->rkey.cap - pl->rkey_init) * sizeof(unsigned long long), st));
            pl->rkey_init = pl->rkey.cap;
        }
    }
    return BGR_OK;
}

// device view of the planner's buffers for a step on the tree of N nodes (call after step_reserve)
static PlanDev step_view(bgr_planner* pl, int64_t N, uint32_t* words_now) {
    const bgr_planner_params& P = pl->p;
    PlanDev d;
    d.p = P;
    float *x, *y, *z;
    float4* f4;
    nns_soa(pl->nns, &x, &y, &z, &f4);
    d.x = x; d.y = y; d.z = z;
    d.par = pl->par.p;
    d.n = N; d.n_dev = nullptr; d.B = P.batch; d.K = P.k; d.W = P.mode == BGR_PLAN_MODE_MS ? 6 : 4; d.have_knn = N > P.k;   // MS:1273
    d.words = words_now; d.table = pl->table.p; d.nt = pl->nt; d.milestones = pl->milestones.p;
    d.prand = pl->prand.p; d.pnew = pl->pnew.p; d.found = pl->found.p; d.near = pl->near.p;
    d.hit = pl->hit.p; d.band = pl->band.p; d.knn_idx = pl->knn_idx.p;
    d.e1 = pl->e1.p; d.e2 = pl->e2.p; d.ehit = pl->ehit.p; d.eband = pl->eband.p;
    d.costs = pl->costs.p; d.psis = pl->psis.p; d.psi = pl->psi.p; d.ok = pl->ok.p; d.parent = pl->parent.p; d.new_index = pl->new_index.p;
    d.cost = pl->cost.p; d.rkey = pl->rkey.p; d.counters = pl->counters.p;
    d.uq = pl->uq.p; d.qslot = pl->qslot.p; d.u_near = pl->u_near.p; d.n_fixed = P.mode == BGR_PLAN_MODE_MS ? 4 : 1;
    return d;
}

// The device work of one step, from the counters block to the counters copy, enqueued on the planner's stream.
// capped = false: host-driven form (launch parameters from the exact tree size N).
// capped = true : the form a CUDA graph replays — the counters block is uploaded from a pinned host block whose last
// word carries N (d.n_dev points at it), and Nearest / kNN are the plain scans laid out for kGraphCap nodes.
constexpr int64_t kGraphCap = 16384;     // below this size the AUTO method is the plain scan anyway
static int step_enqueue(bgr_planner* pl, const PlanDev& d, float* x, float* y, float* z, float4* f4, bool capped) {
    const bgr_planner_params& P = pl->p;
    const int B = P.batch, K = P.k, K1 = K + 1;
    const bool edge_check = P.flags & BGR_PLAN_EDGE_CHECK, rewire = P.flags & BGR_PLAN_REWIRE;
    const bool have_knn = d.have_knn;
    const int n_fixed = d.n_fixed;
    cudaStream_t st = pl->stream;
    const unsigned gb = (unsigned)div_up(B, 256), gbk1 = (unsigned)div_up((int64_t)B * K1, 256), gbk = (unsigned)div_up((int64_t)B * K, 256);
    if (capped) BGR_CUDA(cudaMemcpyAsync(pl->counters.p, pl->host_init.p, 16 * sizeof(long long), cudaMemcpyHostToDevice, st));
    else BGR_CUDA(cudaMemsetAsync(pl->counters.p, 0, 16 * sizeof(long long), st));
    plan_sample_kernel<<<gb, 256, 0, st>>>(d);
    BGR_LAUNCH_CHECK();
    if (capped)
        BGR_CHECK(nns_nearest_scan_cap(pl->nns, pl->uq.p, B + n_fixed, pl->counters.p + 10, n_fixed, pl->u_near.p, pl->u_dist.p,
                                       kGraphCap, pl->counters.p + 15));
    else
        BGR_CHECK(nns_nearest_counted(pl->nns, pl->uq.p, B + n_fixed, pl->unique_hint > 0 ? pl->unique_hint + pl->unique_hint / 4 + 8 : B + n_fixed,
                                      pl->counters.p + 10, n_fixed, pl->u_near.p, pl->u_dist.p));
    {
        // steer + state check + dedup + kNN skip flags, a warp per candidate (candidates the state check rejected
        // never read their neighbour rows, and the copies of a goal / milestone candidate have the rows of the
        // first one: their searches are skipped)
        LevelGrid dil;
        BGR_CHECK(map_sphere_gate(pl->map, P.state_radius, &dil));
        plan_state_kernel<<<(unsigned)div_up(B, 4), 128, 0, st>>>(d, map_view(pl->map), dil, pl->kskip.p);
        BGR_LAUNCH_CHECK();
    }
    if (have_knn) {
        if (capped) BGR_CHECK(nns_knn_brute_cap(pl->nns, pl->pnew.p, B, K, pl->knn_idx.p, pl->knn_d2.p, kGraphCap, pl->counters.p + 15));
        else BGR_CHECK(bgr_nns_knn_skip_dev(pl->nns, pl->pnew.p, B, K, 0u, pl->kskip.p, pl->knn_idx.p, pl->knn_d2.p));
        // (the copies' rows are never read: edges, costs, parent choice and rewire proposals go through rep_of)
    }
    if (edge_check) {
        plan_edges_kernel<<<gbk1, 256, 0, st>>>(d);
        BGR_LAUNCH_CHECK();
        BGR_CHECK(bgr_map_collide_edges_dev(pl->map, pl->e1.p, pl->e2.p, (int64_t)B * K1, P.edge_radius, P.edge_shape,
                                            pl->ehit.p, pl->eband.p));
    }
    if (P.flags & BGR_PLAN_DEDUP) {
        // no live copies (dedup rejected them): costs and parent choice of a candidate in one warp
        plan_cost_select_kernel<<<(unsigned)div_up(B, 4), 128, 0, st>>>(d);
        BGR_LAUNCH_CHECK();
    } else {
        // copies read the rows of their slot's first candidate: all costs first, then the choices
        plan_cost_kernel<<<gbk1, 256, 0, st>>>(d);
        BGR_LAUNCH_CHECK();
        plan_select_kernel<<<gb, 256, 0, st>>>(d);
        BGR_LAUNCH_CHECK();
    }
    plan_commit_kernel<<<1, 1024, 0, st>>>(d, x, y, z, f4);
    BGR_LAUNCH_CHECK();
    if (rewire && have_knn) {
        plan_rewire_propose_kernel<<<gbk, 256, 0, st>>>(d);
        BGR_LAUNCH_CHECK();
        plan_rewire_apply_kernel<<<gbk, 256, 0, st>>>(d, nullptr);
        BGR_LAUNCH_CHECK();
        plan_rewire_clear_kernel<<<gbk, 256, 0, st>>>(d);
        BGR_LAUNCH_CHECK();
    }
    BGR_CUDA(cudaMemcpyAsync(pl->host_counters.p, pl->counters.p, 16 * sizeof(long long), cudaMemcpyDeviceToHost, st));
    return BGR_OK;
}

// A step of a small tree is ~20-30 tiny dependent launches; on a stream of its own the whole sequence is captured
// once (per sample-word buffer) and replayed as ONE graph launch per step while the tree stays below kGraphCap nodes:
// nothing in the captured launches depends on the tree size any more (it is read from the counters block).
static bool graph_ok(const bgr_planner* pl) {
    const bgr_planner_params& P = pl->p;
    const int64_t N = bgr_nns_size(pl->nns);
    static const bool off = getenv("BGR_PLAN_NO_GRAPH") != nullptr;
    return !off && pl->stream != nullptr && !(P.flags & BGR_PLAN_NO_GRAPH) && nns_method(pl->nns) == BGR_METHOD_AUTO &&
           N > P.k && N + P.batch + 1 <= kGraphCap;
}

static void graph_drop(bgr_planner* pl) {
    for (int i = 0; i < 2; ++i)
        if (pl->gexec[i]) { cudaGraphExecDestroy(pl->gexec[i]); pl->gexec[i] = nullptr; }
    pl->g_warm = false;
}

static uint64_t graph_signature(bgr_planner* pl) {
    float *x, *y, *z;
    float4* f4;
    nns_soa(pl->nns, &x, &y, &z, &f4);
    const void *sa, *sb;
    nns_scan_scratch(pl->nns, &sa, &sb);
    uint64_t v = map_signature(pl->map);
    auto mix = [&](const void* p) { v = (v ^ (uint64_t)(uintptr_t)p) * 1099511628211ull; };
    mix(x); mix(y); mix(z); mix(f4); mix(sa); mix(sb); mix(pl->par.p); mix(pl->rkey.p); mix(pl->counters.p);
    mix(pl->words.p); mix(pl->words2.p); mix(pl->knn_idx.p); mix(pl->e1.p); mix(pl->ehit.p); mix(pl->pnew.p);
    mix(pl->uq.p); mix(pl->costs.p); mix(pl->stream); mix(pl->host_counters.p); mix(pl->host_init.p);
    // the kernels take the parameters by value (goal, eta, radii, flags ...): a re-armed planner re-captures
    const unsigned char* pb = reinterpret_cast<const unsigned char*>(&pl->p);
    for (size_t i = 0; i < sizeof(pl->p); ++i) v = (v ^ pb[i]) * 1099511628211ull;
    v = (v ^ (uint64_t)pl->nt) * 1099511628211ull;
    mix(pl->table.p); mix(pl->milestones.p);
    return v;
}

static int step_launch(bgr_planner* pl) {
    const bgr_planner_params& P = pl->p;
    const int B = P.batch;
    const int W = P.mode == BGR_PLAN_MODE_MS ? 6 : 4;
    const int64_t N = bgr_nns_size(pl->nns);
    cudaStream_t st = pl->stream;
    const bool capped = graph_ok(pl);
    // capacity for this step's appends (+ the goal node); the graph form reserves the whole regime at once so that
    // no pointer moves between replays
    BGR_CHECK(step_reserve(pl, capped ? kGraphCap + B + 1 : N + B + 1));
    uint32_t* words_now = pl->words_cur ? pl->words2.p : pl->words.p;
    uint32_t* words_next = pl->words_cur ? pl->words.p : pl->words2.p;
    PlanDev d = step_view(pl, N, words_now);
    float *x, *y, *z;
    float4* f4;
    nns_soa(pl->nns, &x, &y, &z, &f4);
    if (!pl->words_ahead) {
        mt_fill_kernel<<<1, 256, 0, st>>>(pl->mt_state.p, words_now, B * W);
        BGR_LAUNCH_CHECK();
    }
    if (!capped) {
        BGR_CHECK(step_enqueue(pl, d, x, y, z, f4, false));
    } else {
        BGR_CHECK(pl->host_init.reserve(16 * sizeof(long long)));
        long long* hi = (long long*)pl->host_init.p;
        for (int i = 0; i < 15; ++i) hi[i] = 0;
        hi[15] = N;                                   // read by the replayed upload: the host has synchronised the previous step
        d.n_dev = pl->counters.p + 15;
        d.have_knn = 1;
        const int par = pl->words_cur;
        const uint64_t sig = graph_signature(pl);
        if (!pl->g_warm || sig != pl->g_sig) {
            // first step of the regime (or something moved): run the sequence directly — allocations, function
            // attributes and the gate maps settle — and capture from the next step on
            graph_drop(pl);
            BGR_CHECK(step_enqueue(pl, d, x, y, z, f4, true));
            pl->g_sig = graph_signature(pl);
            pl->g_warm = true;
        } else {
            if (!pl->gexec[par]) {
                cudaGraph_t g = nullptr;
                BGR_CUDA(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
                const int rc = step_enqueue(pl, d, x, y, z, f4, true);
                const cudaError_t ce = cudaStreamEndCapture(st, &g);
                if (rc != BGR_OK || ce != cudaSuccess || !g) {
                    if (g) cudaGraphDestroy(g);
                    (void)cudaGetLastError();
                    if (rc != BGR_OK) return rc;
                    set_error("planner step capture failed: %s", cudaGetErrorString(ce));
                    return BGR_ERR_CUDA;
                }
                const cudaError_t ie = cudaGraphInstantiate(&pl->gexec[par], g, 0);
                cudaGraphDestroy(g);
                if (ie != cudaSuccess) {
                    pl->gexec[par] = nullptr;
                    set_error("planner step graph instantiation failed: %s", cudaGetErrorString(ie));
                    return BGR_ERR_CUDA;
                }
            }
            BGR_CUDA(cudaGraphLaunch(pl->gexec[par], st));
            pl->n_graph_steps++;
        }
    }
    if (!pl->ev_step) BGR_CUDA(cudaEventCreateWithFlags(&pl->ev_step, cudaEventDisableTiming));
    BGR_CUDA(cudaEventRecord(pl->ev_step, st));
    // the sample stream does not depend on the tree: the next step's words (a single-CTA, serial MT19937
    // regeneration) are produced behind the counters copy, while the host turns the step around
    mt_fill_kernel<<<1, 256, 0, st>>>(pl->mt_state.p, words_next, B * W);
    BGR_LAUNCH_CHECK();
    pl->words_ahead = true;
    pl->words_cur ^= 1;
    pl->pending_n = N;
    return BGR_OK;
}

static int step_finish(bgr_planner* pl, int* n_accepted, int* goal_reached) {
    const bgr_planner_params& P = pl->p;
    const int B = P.batch;
    const int n_fixed = P.mode == BGR_PLAN_MODE_MS ? 4 : 1;
    const int64_t N = pl->pending_n;
    long long* hc = (long long*)pl->host_counters.p;
    BGR_CUDA(cudaEventSynchronize(pl->ev_step));   // the counters are home; the next step's words may still be in flight
    const int acc = (int)hc[0];
    pl->unique_hint = hc[10] + n_fixed;
    const bool goal = hc[1] >= 0;
    nns_set_size(pl->nns, N + acc + (goal ? 1 : 0));
    pl->stats[0] += 1;
    pl->stats[1] += B;
    pl->stats[2] += hc[4];
    pl->stats[3] += hc[5];
    pl->stats[4] += hc[6];
    pl->stats[5] += hc[7];
    pl->stats[6] += hc[8];
    if (goal) {
        pl->goal_reached = 1;
        pl->stats[7] = hc[1];
    }
    if (n_accepted) *n_accepted = acc;
    if (goal_reached) *goal_reached = pl->goal_reached;
    return BGR_OK;
}

// Small batches on small trees without edge checks: whole steps inside one launch (plan_persist_kernel).
static bool persist_ok(const bgr_planner* pl) {
    const bgr_planner_params& P = pl->p;
    return P.batch <= kPersistMaxB && !(P.flags & (BGR_PLAN_EDGE_CHECK | BGR_PLAN_NO_PERSIST)) &&
           bgr_nns_size(pl->nns) + 2 * (P.batch + 1) <= kPersistMaxNodes;
}

// runs up to max_steps steps (fewer when the tree reaches max_nodes / the persistent regime's node limit, or
// the goal is reached); *steps_done = the number it ran
static int persist_run(bgr_planner* pl, int64_t max_nodes, int64_t max_steps, int64_t* steps_done) {
    const bgr_planner_params& P = pl->p;
    const int B = P.batch;
    const int64_t N = bgr_nns_size(pl->nns);
    cudaStream_t st = pl->stream;
    *steps_done = 0;
    BGR_CHECK(check_device(map_device(pl->map)));
    // a launch runs at most as many steps as the reserved capacity can absorb
    int64_t steps = std::min<int64_t>(max_steps, 4096);
    steps = std::min<int64_t>(steps, (kPersistMaxNodes - N) / (B + 1));
    if (steps <= 0) return BGR_OK;
    BGR_CHECK(step_reserve(pl, N + steps * (B + 1)));
    BGR_CHECK(pl->pctl.reserve(1, st, false));
    BGR_CHECK(pl->host_pctl.reserve(sizeof(PersistCtl)));
    uint32_t* words_now = pl->words_cur ? pl->words2.p : pl->words.p;
    const PlanDev d = step_view(pl, N, words_now);
    float *x, *y, *z;
    float4* f4;
    nns_soa(pl->nns, &x, &y, &z, &f4);
    LevelGrid dil;
    BGR_CHECK(map_sphere_gate(pl->map, P.state_radius, &dil));
    PersistCtl* hc = (PersistCtl*)pl->host_pctl.p;
    std::memset(hc, 0, sizeof(PersistCtl));
    hc->n = N;
    hc->goal = -1;
    BGR_CUDA(cudaMemcpyAsync(pl->pctl.p, hc, sizeof(PersistCtl), cudaMemcpyHostToDevice, st));
    // the map's stream and the planner's are the same stream (bgr_planner_set_stream keeps them so)
    plan_persist_kernel<<<1, kPersistThreads, 0, st>>>(d, x, y, z, f4, pl->mt_state.p, pl->words_ahead ? words_now : nullptr,
                                                      pl->knn_d2.p, pl->u_dist.p, pl->u_near.p, map_view(pl->map), dil,
                                                      (int)steps, (long long)max_nodes, pl->pctl.p);
    BGR_LAUNCH_CHECK();
    BGR_CUDA(cudaMemcpyAsync(hc, pl->pctl.p, sizeof(PersistCtl), cudaMemcpyDeviceToHost, st));
    BGR_CUDA(cudaStreamSynchronize(st));
    pl->n_persist_launches++;
    if (hc->steps <= 0) return BGR_OK;                 // nothing ran: the prepared words (if any) are still unused
    pl->n_persist_steps += (uint64_t)hc->steps;
    pl->words_ahead = false;                           // consumed by the first step; the kernel left the MT state at the stream position
    const int n_fixed = P.mode == BGR_PLAN_MODE_MS ? 4 : 1;
    pl->unique_hint = hc->unique + n_fixed;
    nns_set_size(pl->nns, hc->n);
    pl->stats[0] += hc->steps;
    pl->stats[1] += hc->steps * B;
    for (int i = 0; i < 5; ++i) pl->stats[2 + i] += hc->stats[i];
    if (hc->goal >= 0) {
        pl->goal_reached = 1;
        pl->stats[7] = hc->goal;
    }
    *steps_done = hc->steps;
    return BGR_OK;
}

// advances the planner: one host-driven step, or a launch of persistent steps where they apply
static int advance(bgr_planner* pl, int64_t max_nodes, int64_t batches_left, int64_t* steps_done) {
    if (persist_ok(pl)) {
        BGR_CHECK(persist_run(pl, max_nodes, batches_left, steps_done));
        if (*steps_done > 0) return BGR_OK;
    }
    BGR_CHECK(step_launch(pl));
    BGR_CHECK(step_finish(pl, nullptr, nullptr));
    *steps_done = 1;
    return BGR_OK;
}

int bgr_planner_step(bgr_planner* pl, int* n_accepted, int* goal_reached) {
    BGR_ARG(pl != nullptr, "null planner");
    if (n_accepted) *n_accepted = 0;
    if (goal_reached) *goal_reached = pl->goal_reached;
    if (pl->goal_reached) return BGR_OK;
    BGR_CHECK(step_launch(pl));
    return step_finish(pl, n_accepted, goal_reached);
}

// Independent planning queries (north_star: "independent start/goal planning queries per GPU"):
// the queries run concurrently, each on its own stream, so their small per-step kernels overlap
// on the device.  The planners must not share a stream or a map handle (the map owns per-call
// scratch memory).
int bgr_planner_run_many(bgr_planner** pls, int n, int64_t max_nodes, int64_t max_batches, int* goal_reached) {
    BGR_ARG(pls != nullptr && n >= 0, "bad arguments");
    for (int i = 0; i < n; ++i) {
        BGR_ARG(pls[i] != nullptr, "null planner %d", i);
        for (int j = 0; j < i; ++j)
            BGR_ARG(pls[i] != pls[j] && pls[i]->map != pls[j]->map && pls[i]->stream != pls[j]->stream,
                    "planners %d and %d share a handle, a map or a stream", j, i);
    }
    // one host thread per query in flight (a step is ~30 driver calls: a single thread cannot feed
    // several streams fast enough); each thread runs whole queries, taken from a shared counter
    int dev = 0;
    BGR_CUDA(cudaGetDevice(&dev));
    const int nthreads = std::max(1, std::min(n, std::min(16, (int)std::thread::hardware_concurrency())));
    std::atomic<int> next{0};
    std::vector<int> rcs((size_t)n, BGR_OK);
    std::vector<std::string> errs((size_t)n);
    auto worker = [&]() {
        if (cudaSetDevice(dev) != cudaSuccess) return;
        for (int i = next.fetch_add(1); i < n; i = next.fetch_add(1)) {
            bgr_planner* pl = pls[i];
            for (int64_t b = 0; b < max_batches && !pl->goal_reached && bgr_nns_size(pl->nns) < max_nodes;) {
                int64_t done = 0;
                const int rc = advance(pl, max_nodes, max_batches - b, &done);
                b += done;
                if (rc != BGR_OK) {
                    rcs[(size_t)i] = rc;
                    errs[(size_t)i] = get_error();
                    break;
                }
            }
        }
    };
    if (nthreads == 1) {
        worker();
    } else {
        std::vector<std::thread> th;
        for (int t = 0; t < nthreads; ++t) th.emplace_back(worker);
        for (auto& t : th) t.join();
    }
    for (int i = 0; i < n; ++i)
        if (rcs[(size_t)i] != BGR_OK) {
            set_error("planner %d: %s", i, errs[(size_t)i].c_str());
            return rcs[(size_t)i];
        }
    if (goal_reached)
        for (int i = 0; i < n; ++i) goal_reached[i] = pls[i]->goal_reached;
    return BGR_OK;
}

int bgr_planner_reset(bgr_planner* pl, const float start[3], const float goal[3], uint32_t seed) {
    BGR_ARG(pl != nullptr && start != nullptr && goal != nullptr, "bad arguments");
    cudaStream_t st = pl->stream;
    BGR_CUDA(cudaStreamSynchronize(st));
    for (int a = 0; a < 3; ++a) { pl->p.start[a] = start[a]; pl->p.goal[a] = goal[a]; }
    pl->p.seed = seed;
    BGR_CHECK(nns_reset(pl->nns));
    BGR_CHECK(bgr_nns_append(pl->nns, start, 1));                               // V = [[xs,ys,zs,-1]] MS:1152-1156
    uint32_t s[kMtN + 1];
    mt_seed_host(s, seed);
    const int32_t root_parent = -1;
    BGR_CUDA(cudaMemcpyAsync(pl->mt_state.p, s, sizeof(s), cudaMemcpyHostToDevice, st));
    BGR_CUDA(cudaMemcpyAsync(pl->par.p, &root_parent, sizeof(int32_t), cudaMemcpyHostToDevice, st));
    BGR_CUDA(cudaStreamSynchronize(st));                                        // s / root_parent are stack temporaries
    pl->goal_reached = 0;
    for (int i = 0; i < 7; ++i) pl->stats[i] = 0;
    pl->stats[7] = -1;
    pl->words_cur = 0;
    pl->words_ahead = false;
    pl->pending_n = 0;
    pl->unique_hint = 0;
    return BGR_OK;
}

int bgr_planner_run_queue(bgr_planner** pls, int n_workers, const float* starts, const float* goals, const uint32_t* seeds,
                          int64_t nq, int64_t max_nodes, int64_t max_batches, int* out_reached, int64_t* out_nodes) {
    BGR_ARG(pls != nullptr && n_workers >= 1 && nq >= 0 && (nq == 0 || (starts && goals && seeds)), "bad arguments");
    for (int i = 0; i < n_workers; ++i) {
        BGR_ARG(pls[i] != nullptr, "null planner %d", i);
        for (int j = 0; j < i; ++j)
            BGR_ARG(pls[i] != pls[j] && pls[i]->map != pls[j]->map && pls[i]->stream != pls[j]->stream,
                    "planners %d and %d share a handle, a map or a stream", j, i);
    }
    int dev = 0;
    BGR_CUDA(cudaGetDevice(&dev));
    std::atomic<int64_t> next{0};
    std::vector<int> rcs((size_t)n_workers, BGR_OK);
    std::vector<std::string> errs((size_t)n_workers);
    auto worker = [&](int w) {
        if (cudaSetDevice(dev) != cudaSuccess) return;
        bgr_planner* pl = pls[w];
        for (int64_t q = next.fetch_add(1); q < nq; q = next.fetch_add(1)) {
            int rc = bgr_planner_reset(pl, starts + 3 * q, goals + 3 * q, seeds[q]);
            for (int64_t b = 0; rc == BGR_OK && b < max_batches && !pl->goal_reached && bgr_nns_size(pl->nns) < max_nodes;) {
                int64_t done = 0;
                rc = advance(pl, max_nodes, max_batches - b, &done);
                b += done;
            }
            if (rc != BGR_OK) {
                rcs[(size_t)w] = rc;
                errs[(size_t)w] = get_error();
                return;
            }
            if (out_reached) out_reached[q] = pl->goal_reached;
            if (out_nodes) out_nodes[q] = bgr_nns_size(pl->nns);
        }
    };
    if (n_workers == 1) {
        worker(0);
    } else {
        std::vector<std::thread> th;
        for (int t = 0; t < n_workers; ++t) th.emplace_back(worker, t);
        for (auto& t : th) t.join();
    }
    for (int i = 0; i < n_workers; ++i)
        if (rcs[(size_t)i] != BGR_OK) {
            set_error("worker %d: %s", i, errs[(size_t)i].c_str());
            return rcs[(size_t)i];
        }
    return BGR_OK;
}

int bgr_planner_run(bgr_planner* pl, int64_t max_nodes, int64_t max_batches, int* goal_reached) {
    BGR_ARG(pl != nullptr, "null planner");
    for (int64_t b = 0; b < max_batches && !pl->goal_reached && bgr_nns_size(pl->nns) < max_nodes;) {
        int64_t done = 0;
        BGR_CHECK(advance(pl, max_nodes, max_batches - b, &done));
        b += done;
    }
    if (goal_reached) *goal_reached = pl->goal_reached;
    return BGR_OK;
}

int bgr_planner_get_tree(bgr_planner* pl, float* xyz, int32_t* parent) {
    BGR_ARG(pl != nullptr && xyz != nullptr && parent != nullptr, "bad arguments");
    const int64_t n = bgr_nns_size(pl->nns);
    float *x, *y, *z;
    nns_soa(pl->nns, &x, &y, &z, nullptr);
    std::vector<float> hx(n), hy(n), hz(n);
    BGR_CUDA(cudaMemcpyAsync(hx.data(), x, (size_t)n * 4, cudaMemcpyDeviceToHost, pl->stream));
    BGR_CUDA(cudaMemcpyAsync(hy.data(), y, (size_t)n * 4, cudaMemcpyDeviceToHost, pl->stream));
    BGR_CUDA(cudaMemcpyAsync(hz.data(), z, (size_t)n * 4, cudaMemcpyDeviceToHost, pl->stream));
    BGR_CUDA(cudaMemcpyAsync(parent, pl->par.p, (size_t)n * 4, cudaMemcpyDeviceToHost, pl->stream));
    BGR_CUDA(cudaStreamSynchronize(pl->stream));
    for (int64_t i = 0; i < n; ++i) { xyz[3 * i] = hx[i]; xyz[3 * i + 1] = hy[i]; xyz[3 * i + 2] = hz[i]; }
    return BGR_OK;
}

int bgr_planner_get_path(bgr_planner* pl, int32_t* path, int64_t* n_out) {   // BuildPlan MS:156-171
    BGR_ARG(pl != nullptr && n_out != nullptr, "bad arguments");
    const int64_t n = bgr_nns_size(pl->nns);
    std::vector<int32_t> par(n);
    BGR_CUDA(cudaMemcpyAsync(par.data(), pl->par.p, (size_t)n * 4, cudaMemcpyDeviceToHost, pl->stream));
    BGR_CUDA(cudaStreamSynchronize(pl->stream));
    std::vector<int32_t> rev;
    int64_t cur = n - 1, guard = 0;
    while (cur > 0 && guard++ <= n) {
        rev.push_back((int32_t)cur);
        cur = par[cur];
    }
    BGR_ARG(guard <= n, "parent chain does not reach the root (cycle)");
    rev.push_back(0);
    const int64_t len = (int64_t)rev.size();
    if (path) {
        BGR_ARG(*n_out >= len, "path buffer too small (%lld < %lld)", (long long)*n_out, (long long)len);
        for (int64_t i = 0; i < len; ++i) path[i] = rev[len - 1 - i];
    }
    *n_out = len;
    return BGR_OK;
}

int bgr_planner_persist_stats(bgr_planner* pl, uint64_t out[3]) {
    BGR_ARG(pl != nullptr && out != nullptr, "bad arguments");
    out[0] = pl->n_persist_launches;
    out[1] = pl->n_persist_steps;
    out[2] = pl->n_graph_steps;
    return BGR_OK;
}

int bgr_planner_stats(bgr_planner* pl, int64_t out[8]) {
    BGR_ARG(pl != nullptr && out != nullptr, "bad arguments");
    std::memcpy(out, pl->stats, sizeof(pl->stats));
    return BGR_OK;
}

}  // extern "C"
