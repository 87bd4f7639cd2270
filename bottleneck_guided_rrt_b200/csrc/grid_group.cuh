// Group kNN: the grid search's one-pass ring-1 path with SEVERAL queries per warp, one per group of LQ lanes
// (LQ = 16: two queries per warp, LQ = 8: four).  Reference surface nns->knn, MS:1309-1312; same rows as
// grid_knn_kernel / the brute-force scan.
//
// grid_knn_kernel is bound by instruction issue (80 % of the issue slots busy, ~960 warp instructions per
// query; profiles/r2_ncu_full_q_knn.txt), and a large part of those instructions is per-QUERY bookkeeping that
// all 32 lanes repeat: cell coordinates, the nine row bounds, the nearest unscanned face, the a-priori
// threshold, the retry logic, the kernel prologue.  With a query per lane group every such instruction serves
// 32 / LQ queries.  The per-NODE part: a ring-1 x-row holds ~24 nodes — one 32-lane pass at 75 % of the lanes
// before, 32 / LQ passes of LQ lanes now (LQ = 8: three of the four passes are full).  The selection sorts the
// group's <= 32 (or <= 64) survivors with a bitonic network on packed 32-bit keys (d2 bits, low 6 bits = the
// survivor's slot), all groups of the warp at once; the counting pass of grid_knn_query stays as the exact
// fallback for survivors whose d2 agree in the upper 26 bits.
//
// Exactness: the same argument as the fast path of grid_knn_query (every node of the 3x3x3 block with
// d2 <= tau is in the group's buffer, tau is strictly below the nearest unscanned face, >= k survivors or tau at
// the radius limit => the k smallest keys are the answer).  The packed keys order the survivors exactly unless two
// of the first k + 1 ranks agree in their upper 26 bits; those groups are ranked by counting on the full d2, and equal
// d2 among the first k ranks by the full (d2, index) key.  A group whose query cannot be finished here (empty block, needs ring 2, three
// attempts spent) hands it to the general ring search, run by the whole warp at the end of the kernel.
//
// Used for fully indexed, uncrowded clouds (no unsorted tail, no piles): grid_knn() in tile_knn.cuh.
#pragma once
#include <type_traits>
#include "grid.cuh"

namespace bgr {

// nearest unscanned face after ring 1, squared and shrunk by the rounding slack of the cell computation
// (+inf when the block reaches every border): the arithmetic of grid_knn_query's face2(1)
__device__ __forceinline__ float group_face2(const GridGeom& gm, const float (&qv)[3], const int (&cc)[3]) {
    float b = __uint_as_float(kInfBits);
#pragma unroll
    for (int ax = 0; ax < 3; ++ax) {
        if (cc[ax] - 1 > 0) b = fminf(b, qv[ax] - (gm.mn[ax] + (float)(cc[ax] - 1) * gm.cs));
        if (cc[ax] + 2 < gm.g[ax]) b = fminf(b, (gm.mn[ax] + (float)(cc[ax] + 2) * gm.cs) - qv[ax]);
    }
    if (b == __uint_as_float(kInfBits)) return b;
    const float bs = fmaxf(0.f, b - 2e-3f * gm.cs);
    return bs * bs * (1.f - 2e-6f);
}

// rank += (v < key) as a compare and a predicated add (the plain C form compiles to compare, select, add)
__device__ __forceinline__ void count_below(int& rank, const uint32_t v, const uint32_t key) {
    asm("{\n\t.reg .pred p;\n\tsetp.lt.u32 p, %1, %2;\n\t@p add.s32 %0, %0, 1;\n\t}" : "+r"(rank) : "r"(v), "r"(key));
}

// Ascending bitonic sort of the N (32 or 64) keys a group of LQ lanes holds in N / LQ registers each: position
// P = s LQ + l (register s of lane l) ends up holding rank P.  Strides >= LQ exchange registers of one lane,
// smaller strides are xor shuffles, which stay inside the group.
template <int LQ, int N>
__device__ __forceinline__ void group_sort(uint32_t (&v)[N / LQ], const int l) {
    constexpr int KS = N / LQ;
#pragma unroll
    for (int size = 2; size <= N; size <<= 1) {
#pragma unroll
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
            if (stride >= LQ) {
#pragma unroll
                for (int s = 0; s < KS; ++s) {
                    const int s2 = s ^ (stride / LQ);
                    if (s2 > s) {
                        const bool asc = ((s * LQ) & size) == 0;            // size >= 2 LQ: the bit lies in s
                        const uint32_t lo = min(v[s], v[s2]), hi = max(v[s], v[s2]);
                        v[s] = asc ? lo : hi;
                        v[s2] = asc ? hi : lo;
                    }
                }
            } else {
#pragma unroll
                for (int s = 0; s < KS; ++s) {
                    const uint32_t o = __shfl_xor_sync(kFull, v[s], stride);
                    const bool asc = ((s * LQ + l) & size) == 0;
                    const bool lower = (l & stride) == 0;
                    v[s] = (lower == asc) ? min(v[s], o) : max(v[s], o);
                }
            }
        }
    }
}

#ifndef BGR_GROUP_ROWS
#define BGR_GROUP_ROWS 3
#endif
#ifndef BGR_GROUP_SM_WARPS
#define BGR_GROUP_SM_WARPS 32
#endif
constexpr int kGroupRows = BGR_GROUP_ROWS;     // ring-1 rows whose loads are in flight together (divides 9)
#ifndef BGR_GROUP_WARPS
#define BGR_GROUP_WARPS 4
#endif
constexpr int kGroupLanesDefault = 16;           // lanes per query grid_knn() uses
constexpr int kGroupWarps = BGR_GROUP_WARPS;   // 64 registers, 8 CTAs of 4 warps per SM

// requires: a.n_indexed == a.n > 0, a.fast, a.heavy == nullptr
template <int LQ>
__global__ void __launch_bounds__(kGroupWarps * 32, BGR_GROUP_SM_WARPS / kGroupWarps) grid_knn_group_kernel(GridKnnArgs a) {
    static_assert(LQ == 16 || LQ == 8, "lanes per query");
    constexpr int NG = 32 / LQ;               // queries per warp
    constexpr int RS = (9 + LQ - 1) / LQ;     // ring-1 rows a lane owns
    constexpr int PASSES = 32 / LQ;           // LQ-lane passes over the first 32 nodes of a row
    constexpr int KS = 32 / LQ;               // survivors a lane owns per 32
    __shared__ unsigned long long cand[kGroupWarps][NG][kFastCap];
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const int grp = lane / LQ, l = lane % LQ;
    const unsigned gmask = ((1u << LQ) - 1u) << (grp * LQ);
    const int64_t q0 = ((int64_t)blockIdx.x * kGroupWarps + wib) * NG;
    if (q0 >= a.nq) return;
    const int64_t qi = min(q0 + grp, a.nq - 1);            // missing queries of the last warp repeat the last one, inactive
    bool active = (q0 + grp < a.nq) && !(a.skip && __ldg(a.skip + qi));
    bool failed = false;                                   // this group's query goes to the general search
    unsigned long long touched = 0;

    const GridGeom& gm = a.gm;
    const float qv[3] = {__ldg(a.q + 3 * qi), __ldg(a.q + 3 * qi + 1), __ldg(a.q + 3 * qi + 2)};
    const float qx = qv[0], qy = qv[1], qz = qv[2];
    const bool self = (a.flags & BGR_ALLOW_SELF_MATCH) != 0;
    const uint32_t sent = radius_sentinel_g(a.max_radii ? __ldg(a.max_radii + qi) : a.max_radius);
    const float lim = __uint_as_float(sent - 1u);
    const int cc[3] = {cell_coord(qx, gm.mn[0], gm.inv_cs, gm.g[0]), cell_coord(qy, gm.mn[1], gm.inv_cs, gm.g[1]),
                       cell_coord(qz, gm.mn[2], gm.inv_cs, gm.g[2])};
    const int cx = cc[0], cy = cc[1], cz = cc[2];

    // the 9 x-rows of ring 1: row r = (dy, dz) is owned by lane r % LQ of the group, register slot r / LQ
    int rs[RS], re[RS];
    int n27 = 0;
    unsigned longm[RS];                                    // rows longer than 32 nodes: ballot over their owner lanes (warp-uniform)
#pragma unroll
    for (int t = 0; t < RS; ++t) {
        rs[t] = re[t] = 0;
        const int r = l + LQ * t;
        if (r < 9) {
            const int yy = cy + r % 3 - 1, zz = cz + r / 3 - 1;
            if (yy >= 0 && yy < gm.g[1] && zz >= 0 && zz < gm.g[2]) {
                const int x0 = max(cx - 1, 0), x1 = min(cx + 1, gm.g[0] - 1);
                const int64_t rb = ((int64_t)zz * gm.g[1] + yy) * gm.g[0];
                rs[t] = __ldg(a.cell_start + rb + x0);
                re[t] = __ldg(a.cell_start + rb + x1 + 1);
            }
        }
        n27 += re[t] - rs[t];
        longm[t] = __ballot_sync(kFull, re[t] - rs[t] > 32);
    }
#pragma unroll
    for (int o = LQ / 2; o > 0; o >>= 1) n27 += __shfl_xor_sync(kFull, n27, o);   // xor < LQ stays inside the group
    // bounds of row r (uniform over the warp) for every lane's own group
    auto row_bounds = [&](const int r, int& s, int& e) {
        if (RS == 1) {
            s = __shfl_sync(kFull, rs[0], r, LQ);
            e = __shfl_sync(kFull, re[0], r, LQ);
        } else {
            s = __shfl_sync(kFull, r < LQ ? rs[0] : rs[RS - 1], r % LQ, LQ);
            e = __shfl_sync(kFull, r < LQ ? re[0] : re[RS - 1], r % LQ, LQ);
        }
    };
    const float b2 = group_face2(gm, qv, cc);
    // largest usable threshold: strictly below the unscanned region, within the radius limit
    const float tmax = fminf(lim, b2 == __uint_as_float(kInfBits) ? lim : __uint_as_float(max(__float_as_uint(b2), 1u) - 1u));
    if (active && n27 == 0) { failed = true; active = false; }
    const int nxr = min(cx + 1, gm.g[0] - 1) - max(cx - 1, 0) + 1, nyr = min(cy + 1, gm.g[1] - 1) - max(cy - 1, 0) + 1,
              nzr = min(cz + 1, gm.g[2] - 1) - max(cz - 1, 0) + 1;
    const float vol = (float)(nxr * nyr * nzr) * gm.cs * gm.cs * gm.cs;
    const float r3 = a.fast_target * vol / (4.18879f * (float)max(n27, 1));
    const float r0 = exp2f(__log2f(r3) * 0.33333334f);      // the threshold only steers the work: any value is exact
    float tau = fminf(r0 * r0, tmax);
    unsigned long long* cand_w = cand[wib][grp];
    const uint32_t cand_s = smem_u32(cand_w);
    const unsigned lt_mask = ((1u << l) - 1u) << (grp * LQ);        // the lanes of this group below this one
    const float eps_lo = self ? -1.f : FLT_EPSILON;         // the self-match rule d2 > FLT_EPSILON as one comparison (d2 >= 0 > -1)
    const int last = (int)a.n_indexed - 1;

    for (int attempt = 0; attempt < kFastTries && __any_sync(kFull, active); ++attempt) {
        int c = 0;
        auto consider = [&](const float4 p, bool valid) {
            const float dx = __fsub_rn(qx, p.x), dy = __fsub_rn(qy, p.y), dz = __fsub_rn(qz, p.z);
            const float d = __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
            const bool pass = valid && (d <= tau) && (d > eps_lo);
            const unsigned m = __ballot_sync(kFull, pass);
            const int slot = c + __popc(m & lt_mask);
            // key = (d2 bits << 32) | index: low word first
            if (pass && slot < kFastCap)
                asm volatile("st.shared.v2.u32 [%0], {%1, %2};" ::"r"(cand_s + 8u * (unsigned)slot), "r"(__float_as_uint(p.w)), "r"(__float_as_uint(d)) : "memory");
            c += __popc(m & gmask);
        };
        // rows in groups of three, two LQ-lane passes at a time: six LDG.128 in flight together
#pragma unroll 1
        for (int g = 0; g < 9; g += kGroupRows) {
            int s0[kGroupRows], e0[kGroupRows];
#pragma unroll
            for (int i = 0; i < kGroupRows; ++i) {
                int e;
                row_bounds(g + i, s0[i], e);
                e0[i] = active ? e : s0[i];                                           // a finished group admits nothing
            }
#pragma unroll
            for (int pp = 0; pp < PASSES; pp += 2) {
                float4 p[kGroupRows][2];
#pragma unroll
                for (int i = 0; i < kGroupRows; ++i) {
                    p[i][0] = __ldg(a.pts + min(s0[i] + pp * LQ + l, last));          // lanes past the row's end load a node they ignore
                    p[i][1] = __ldg(a.pts + min(s0[i] + (pp + 1) * LQ + l, last));
                }
#pragma unroll
                for (int i = 0; i < kGroupRows; ++i) {
                    consider(p[i][0], s0[i] + pp * LQ + l < e0[i]);
                    consider(p[i][1], s0[i] + (pp + 1) * LQ + l < e0[i]);
                }
            }
        }
        // rows longer than 32 nodes (a few per cent of the rows): each group walks its own, together
#pragma unroll
        for (int t = 0; t < RS; ++t) {
            unsigned lm = longm[t];
            while (lm) {
                const unsigned mine = lm & gmask;
                const int r = (mine ? (__ffs(mine) - 1) % LQ : 0) + LQ * t;
#pragma unroll
                for (int g2 = 0; g2 < NG; ++g2) {                    // drop the lowest row of every group
                    const unsigned part = lm & (((1u << LQ) - 1u) << (g2 * LQ));
                    lm &= ~(part & (0u - part));
                }
                int s1, e;
                row_bounds(r, s1, e);
                s1 += 32;
                const int e1 = (active && mine) ? e : s1;
                for (int b = s1; __any_sync(kFull, b < e1); b += LQ) {
                    const float4 q4 = __ldg(a.pts + min(b + l, last));
                    consider(q4, b + l < e1);
                }
            }
        }
        if (active) touched += (unsigned long long)n27;
        const bool ok = active && c <= kFastCap && (c >= a.k || (tau >= tmax && tmax >= lim));
        __syncwarp();
        if (__any_sync(kFull, ok)) {
            const uint32_t* cd = reinterpret_cast<const uint32_t*>(cand_w) + 1;
            const bool wide = __any_sync(kFull, ok && c > 32);
            bool by_count = false;                         // this group's survivors have to be ranked exactly, by counting
            // ---- selection: sort 32-bit keys = d2 bits with the low 6 bits replaced by the survivor's slot, bitonic
            // network across the group's lanes x N / LQ registers (min / max and xor shuffles only; N = 32, or 64 when
            // some group of the warp holds more than 32 survivors).  Afterwards register s of lane l holds rank
            // s LQ + l: the row is written in rank order (coalesced), the full key fetched by slot.  The 6 dropped
            // bits can only misorder two survivors whose d2 agree in the upper 26 bits: such a pair among the first
            // k + 1 ranks sends the group to the exact ranking below.
            auto select = [&](auto n_tag) {
                constexpr int N = decltype(n_tag)::value, NS = N / LQ;
                const int cm = ok ? c : 0;
                uint32_t v[NS];
#pragma unroll
                for (int s = 0; s < NS; ++s) {
                    const int pos = l + LQ * s;
                    v[s] = pos < cm ? ((cd[2 * pos] & ~63u) | (uint32_t)pos) : 0xFFFFFFFFu;
                }
                group_sort<LQ, N>(v, l);
                bool tie = false;
#pragma unroll
                for (int s = 0; s < NS; ++s) {
                    if (s * LQ < a.k) {
                        uint32_t nx = __shfl_sync(kFull, v[s], l + 1, LQ);
                        const uint32_t wrap = __shfl_sync(kFull, v[s + 1 < NS ? s + 1 : s], 0, LQ);
                        if (l == LQ - 1) nx = s + 1 < NS ? wrap : 0xFFFFFFFFu;
                        const int r = s * LQ + l;
                        tie |= (r < a.k) && (r + 1 < cm) && ((v[s] >> 6) == (nx >> 6));
                    }
                }
                const bool group_tie = (__ballot_sync(kFull, tie) & gmask) != 0u;
                by_count = ok && group_tie;
                if (ok && !group_tie) {
#pragma unroll
                    for (int s = 0; s < NS; ++s) {
                        const int r = s * LQ + l;
                        if (r < a.k) {
                            if (r < cm) {
                                const unsigned long long key = cand_w[v[s] & 63u];
                                a.idx[qi * a.k + r] = (int32_t)((int64_t)(uint32_t)key + a.index_base);
                                a.d2[qi * a.k + r] = __uint_as_float((uint32_t)(key >> 32));
                            } else {                                   // fewer than k inside the radius limit
                                a.idx[qi * a.k + r] = -1;
                                a.d2[qi * a.k + r] = __uint_as_float(kInfBits);
                            }
                        }
                    }
                    if (a.counter && l == 0) atomicAdd(a.counter, touched);
                    active = false;
                }
            };
            if (wide) select(std::integral_constant<int, 64>());
            else select(std::integral_constant<int, 32>());
            if (__any_sync(kFull, by_count)) {
                // ---- selection, exact ranking by counting (more than 32 survivors in some group, or a possible tie).
                // Lane l owns the survivors l + LQ s of the first 32 (slots 0 .. KS-1) and of the next 32 (slots
                // KS .. 2 KS - 1, only when some group holds more than 32); d2 alone orders them unless two of the first k
                // share a d2.  No warp-wide operation inside the loops: the groups run their own trip counts.
                const int cm = by_count ? c : 0;
                const bool wide2 = __any_sync(kFull, cm > 32);
                uint32_t kd[2 * KS];
                int rk[2 * KS];
#pragma unroll
                for (int s = 0; s < 2 * KS; ++s) { kd[s] = 0xFFFFFFFFu; rk[s] = 0; }
#pragma unroll
                for (int s = 0; s < KS; ++s)
                    if (l + LQ * s < cm) kd[s] = cd[2 * (l + LQ * s)];
#pragma unroll 4
                for (int j = 0; j < cm; ++j) {
                    const uint32_t v = cd[2 * j];
#pragma unroll
                    for (int s = 0; s < KS; ++s) count_below(rk[s], v, kd[s]);
                }
                if (wide2) {
#pragma unroll
                    for (int s = KS; s < 2 * KS; ++s)
                        if (l + LQ * s < cm) kd[s] = cd[2 * (l + LQ * s)];
#pragma unroll 4
                    for (int j = 0; j < cm; ++j) {
                        const uint32_t v = cd[2 * j];
#pragma unroll
                        for (int s = KS; s < 2 * KS; ++s) count_below(rk[s], v, kd[s]);
                    }
                }
                // the first k ranks are a permutation of 0 .. min(c, k) - 1 exactly when no two of them tie
                unsigned bits = 0;
                int below = 0;
#pragma unroll
                for (int s = 0; s < 2 * KS; ++s) {
                    if (s >= KS && !wide2) break;
                    const bool in = (l + LQ * s < cm) && rk[s] < a.k;
                    bits |= in ? 1u << rk[s] : 0u;
                    below += in ? 1 : 0;
                }
                const int want = min(cm, a.k);
                const unsigned all_bits = __reduce_or_sync(gmask, bits);
                const int all_below = (int)__reduce_add_sync(gmask, (unsigned)below);
                const bool tie = by_count && (__popc(all_bits) != want || all_below != want);
                if (tie) {
                    // re-rank by the full (d2, index) key
#pragma unroll
                    for (int s = 0; s < 2 * KS; ++s) {
                        rk[s] = 0;
                        if (l + LQ * s < cm) {
                            const unsigned long long key = cand_w[l + LQ * s];
                            for (int j = 0; j < cm; ++j) rk[s] += cand_w[j] < key ? 1 : 0;
                        }
                    }
                }
                if (by_count) {
#pragma unroll
                    for (int s = 0; s < 2 * KS; ++s) {
                        if (s >= KS && !wide2) break;
                        const int t = l + LQ * s;
                        if (t < cm && rk[s] < a.k) {
                            a.idx[qi * a.k + rk[s]] = (int32_t)((int64_t)(uint32_t)cand_w[t] + a.index_base);
                            a.d2[qi * a.k + rk[s]] = __uint_as_float(kd[s]);
                        }
                    }
                    if (cm < a.k) {                                    // fewer than k inside the radius limit
                        for (int t = cm + l; t < a.k; t += LQ) {
                            a.idx[qi * a.k + t] = -1;
                            a.d2[qi * a.k + t] = __uint_as_float(kInfBits);
                        }
                    }
                    if (a.counter && l == 0) atomicAdd(a.counter, touched);
                    active = false;
                }
            }
        }
        __syncwarp();
        if (active) {
            if (c > kFastCap) {
                const float f = exp2f(__log2f(a.fast_target / (float)c) * 0.33333334f);
                tau = tau * f * f;
            } else if (tau >= tmax) {                              // needs ring 2: general search
                failed = true;
                active = false;
            } else {
                const float f = c > 0 ? exp2f(__log2f((a.fast_target + 6.f) / (float)c) * 0.33333334f) : 2.f;
                tau = fminf(tmax, tau * f * f * 1.15f);
            }
        }
    }
    failed |= active;                                              // attempts spent
    if (failed && a.counter && l == 0 && touched) atomicAdd(a.counter, touched);
    // ---- the general ring search for the queries left over, one after the other with the whole warp
#pragma unroll 1
    for (int g2 = 0; g2 < NG; ++g2) {
        if (__any_sync(kFull, failed && grp == g2)) {
            __syncwarp();
            grid_knn_query<false>(a, q0 + g2, cand[wib][0], lane, false);
            __syncwarp();
        }
    }
}

}  // namespace bgr
