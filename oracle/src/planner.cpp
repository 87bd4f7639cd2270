// ORACLE (test infrastructure).  CPU restatement of the planner loop of the reference,
// abb_manipulator/src/rrt_star_milestones.cpp ("MS") main() MS:1160-1406 and
// rrt_star.cpp ("RS") RS:1140-1300, generalised to B candidates per step evaluated against
// the tree snapshot at step start (B = 1 is the reference's sequential loop):
//   sampler   MS:1171-1219 (0.8 / 0.5 thresholds, 3 milestones)  |  RS:1145-1178 (goal_chance)
//   Nearest   MS:138-154     steer MS:1233-1268     isStateValid MS:1270 -> MS:310-351
//   knn       MS:1309-1312 (flags 0, maxRadius inf; guarded by V.size() > K, MS:1273)
//   cost_wiring MS:802-831   extend_tree MS:834-940 (strict <, rewire only printed)
//   parent rule MS:1322-1323 (parent_candidate > 0), append MS:1325-1327, goal test MS:1340,
//   goal node MS:1346-1347.
// Stated divergences: one seeded MT19937 stream instead of
// mt19937(random_device()) per iteration; the reachability table is a caller-supplied
// sample-domain table; optional beyond-reference steps (flags): edge checks of the candidate
// parent edges with isLineValid geometry (MS:385-471) and real rewiring.
// PARITY IS UNPINNED against a run of the reference itself (it needs ROS/FCL/libnabo).
#include "../oracle.h"
#include <cmath>
#include <cstring>
#include <vector>

struct orc_planner {
    orc_plan_params p;
    std::vector<double> boxes;
    std::vector<float> table, milestones;
    std::vector<float> xyz;        // V[i][0..2]
    std::vector<int32_t> par;      // V[i][3]
    orc_mt* mt;
    int goal_reached;
    int collision_mode = 1;        // oracle collision mode of the state / edge checks (orc_planner_set_collision_mode)
    int64_t stats[8];              // batches, candidates, rejected_state, rejected_edge, improved_parent,
                                   // rewired, band_events, goal_candidate
};

namespace {

// cost_wiring(tree, tree[P], child) MS:802-831: float accumulation from the child upwards;
// the walk stops at index <= 0, so the edge into the root is never added.
// *psi (optional): the same sum plus that skipped root edge = the true path length root -> child
// (used only by the beyond-reference rewiring, to keep it acyclic).
float chain_cost(const orc_planner* pl, int32_t P, const float* child, float* psi = nullptr) {
    const float* V = pl->xyz.data();
    float cost = 0;
    cost = cost + orc_norm(V + 3 * P, child);
    int cur = pl->par[P], l = 0, prev = P;
    while (cur > 0) {
        if (l == 0) {
            cost = cost + orc_norm(V + 3 * cur, V + 3 * P);
            l++;
        } else {
            cost = cost + orc_norm(V + 3 * cur, V + 3 * prev);
        }
        prev = cur;
        cur = pl->par[cur];
    }
    if (psi) {
        float full = cost;
        if (cur == 0) full = full + orc_norm(V, V + 3 * prev);
        *psi = full;
    }
    return cost;
}

}  // namespace

extern "C" {

orc_planner* orc_planner_create(const orc_plan_params* p, const double* boxes, int64_t nb, const float* table,
                                int64_t nt, const float* milestones, int nm) {
    if (!p || nt < 1 || nt > (1 << 24) || p->k < 1 || p->k > 32 || p->batch < 1) return nullptr;
    if (p->mode == 1 && nm < 3) return nullptr;
    orc_planner* pl = new orc_planner;
    pl->p = *p;
    pl->boxes.assign(boxes, boxes + 6 * nb);
    pl->table.assign(table, table + 3 * nt);
    pl->milestones.assign(milestones, milestones + 3 * (size_t)nm);
    pl->xyz.assign(p->start, p->start + 3);                         // MS:1152-1156: V = [[xs,ys,zs,-1]]
    pl->par.assign(1, -1);
    pl->mt = orc_mt_new_seed(p->seed);
    pl->goal_reached = 0;
    std::memset(pl->stats, 0, sizeof(pl->stats));
    pl->stats[7] = -1;
    return pl;
}

void orc_planner_destroy(orc_planner* pl) {
    if (!pl) return;
    orc_mt_free(pl->mt);
    delete pl;
}

// 1 (default) = loop over every leaf box with a bounding cull, tolerance-band flags from the
// independent distance; 2 = bucket-grid broadphase in front of the same MPR (same booleans; sphere
// band flags kept, edge band flags not computed): what makes the 256^3 / batch-4096 shapes of
// BASELINE configs[3] checkable in seconds.
void orc_planner_set_collision_mode(orc_planner* pl, int mode) { pl->collision_mode = mode == 2 ? 2 : 1; }

int64_t orc_planner_size(const orc_planner* pl) { return (int64_t)pl->par.size(); }

void orc_planner_get_tree(const orc_planner* pl, float* xyz, int32_t* parent) {
    std::memcpy(xyz, pl->xyz.data(), pl->xyz.size() * sizeof(float));
    std::memcpy(parent, pl->par.data(), pl->par.size() * sizeof(int32_t));
}

void orc_planner_stats(const orc_planner* pl, int64_t out[8]) { std::memcpy(out, pl->stats, sizeof(pl->stats)); }

int orc_planner_step(orc_planner* pl, int* n_accepted, int* goal_reached) {
    const orc_plan_params& P = pl->p;
    if (n_accepted) *n_accepted = 0;
    if (goal_reached) *goal_reached = pl->goal_reached;
    if (pl->goal_reached) return 0;
    const int B = P.batch, K = P.k;
    const int64_t N = (int64_t)pl->par.size(), T = (int64_t)pl->table.size() / 3, nb = (int64_t)pl->boxes.size() / 6;
    const bool edge_check = P.flags & 1u, rewire = P.flags & 2u, quirks = P.flags & 4u, no_goal = P.flags & 8u, dedup = P.flags & 16u;
    std::vector<float> prand(3 * B), pnew(3 * B);
    std::vector<uint8_t> found(B);
    std::vector<int> slot(B, -1);            // -1 table sample, 0 goal, 1 + i milestone i
    int first_of[4] = {-1, -1, -1, -1};
    // ---- sampler: draw order y, x, (z) — MS:1182-1184 / RS:1155-1156
    for (int c = 0; c < B; ++c) {
        const float rand_chance = (float)orc_mt_std_uniform(pl->mt, 0.0, 1.0);
        int64_t ri = (int64_t)(float)(int)orc_mt_std_uniform(pl->mt, 0.0, (double)T);   // `float random_indice = int(x(gen))`
        if (ri > T - 1) ri = T - 1;
        const float* src;
        if (P.mode == 1) {
            const int mi = (int)orc_mt_std_uniform(pl->mt, 0.0, 3.0);
            if ((double)rand_chance > 0.8) src = pl->table.data() + 3 * ri;
            else if ((double)rand_chance < 0.8 && (double)rand_chance > 0.5) { src = P.goal; slot[c] = 0; }
            else { src = pl->milestones.data() + 3 * (mi > 2 ? 2 : mi); slot[c] = 1 + (mi > 2 ? 2 : mi); }
        } else {
            if (rand_chance > P.goal_chance) src = pl->table.data() + 3 * ri;
            else { src = P.goal; slot[c] = 0; }
        }
        if (slot[c] >= 0 && first_of[slot[c]] < 0) first_of[slot[c]] = c;
        std::memcpy(&prand[3 * c], src, 3 * sizeof(float));
    }
    // ---- Nearest + steer (against the snapshot)
    // (the goal / milestone samples of a batch are the same point: their Nearest is computed once)
    std::vector<int32_t> near(B);
    {
        std::vector<float> uq;
        std::vector<int> uslot(B);
        int fixed_u[4] = {-1, -1, -1, -1};
        for (int c = 0; c < B; ++c) {
            if (slot[c] >= 0 && fixed_u[slot[c]] >= 0) { uslot[c] = fixed_u[slot[c]]; continue; }
            uslot[c] = (int)(uq.size() / 3);
            if (slot[c] >= 0) fixed_u[slot[c]] = uslot[c];
            uq.insert(uq.end(), &prand[3 * c], &prand[3 * c] + 3);
        }
        std::vector<int32_t> unear(uq.size() / 3);
        orc_nearest(pl->xyz.data(), N, uq.data(), (int64_t)unear.size(), unear.data(), nullptr, 0);
        for (int c = 0; c < B; ++c) near[c] = unear[(size_t)uslot[c]];
    }
    for (int c = 0; c < B; ++c)
        found[c] = (uint8_t)orc_steer(&pl->xyz[3 * (size_t)near[c]], &prand[3 * c], P.eta, &pnew[3 * c]);
    // ---- isStateValid (true = collision)
    std::vector<uint8_t> hit(B), band(B);
    orc_collide_spheres(pl->boxes.data(), nb, pnew.data(), B, P.state_radius, pl->collision_mode, 1e-6, hit.data(), band.data(), 0);
    for (int c = 0; c < B; ++c) pl->stats[6] += band[c];
    // ORC_PLAN_DEDUP: the goal / milestone samples of one batch are the same point and, against one
    // snapshot, the same candidate: only the first of each is kept, the copies are rejected here
    if (dedup)
        for (int c = 0; c < B; ++c)
            if (slot[c] >= 0 && first_of[slot[c]] != c) hit[c] = 1;
    // ---- kNN + choose parent
    std::vector<int32_t> knn_idx((size_t)B * K, -1), parent(B);
    std::vector<float> knn_d2((size_t)B * K), cost(B), psi(B);
    const bool have_knn = N > K;                                                    // MS:1273 V.size() > 5 (K = 5)
    if (have_knn) {
        // rejected candidates never read their rows (MS:1270 precedes MS:1312): search the others only
        std::vector<float> kq;
        std::vector<int> kc;
        for (int c = 0; c < B; ++c)
            if (!hit[c]) { kq.insert(kq.end(), &pnew[3 * c], &pnew[3 * c] + 3); kc.push_back(c); }
        const int64_t m = (int64_t)kc.size();
        std::vector<int32_t> ki((size_t)m * K);
        std::vector<float> kd((size_t)m * K);
        if (m) orc_knn_brute(pl->xyz.data(), N, kq.data(), m, K, 0, INFINITY, nullptr, ki.data(), kd.data(), 0);
        for (int64_t i = 0; i < m; ++i) {
            std::memcpy(&knn_idx[(size_t)kc[(size_t)i] * K], &ki[(size_t)i * K], (size_t)K * sizeof(int32_t));
            std::memcpy(&knn_d2[(size_t)kc[(size_t)i] * K], &kd[(size_t)i * K], (size_t)K * sizeof(float));
        }
    }
    // candidate edges for the optional edge check: j < K neighbours, j == K the nearest node
    std::vector<uint8_t> ehit((size_t)B * (K + 1), 0);
    if (edge_check) {
        std::vector<float> e1, e2;
        std::vector<size_t> slot;
        for (int c = 0; c < B; ++c) {
            if (hit[c]) continue;
            for (int j = 0; j <= K; ++j) {
                const int32_t ind = j == K ? near[c] : (have_knn ? knn_idx[(size_t)c * K + j] : -1);
                if (ind < 0) continue;
                e1.insert(e1.end(), &pl->xyz[3 * (size_t)ind], &pl->xyz[3 * (size_t)ind] + 3);
                e2.insert(e2.end(), &pnew[3 * c], &pnew[3 * c] + 3);
                slot.push_back((size_t)c * (K + 1) + j);
            }
        }
        const int64_t m = (int64_t)slot.size();
        std::vector<uint8_t> h(m), bd(m);
        if (m) orc_collide_edges(pl->boxes.data(), nb, e1.data(), e2.data(), m, P.edge_radius, P.edge_shape, pl->collision_mode, 1e-6, h.data(), bd.data(), 0);
        for (int64_t i = 0; i < m; ++i) {
            ehit[slot[i]] = h[i];
            pl->stats[6] += bd[i];
        }
    }
    std::vector<uint8_t> ok(B, 0);
    for (int c = 0; c < B; ++c) {
        pl->stats[1]++;
        if (hit[c]) { pl->stats[2]++; continue; }
        if (edge_check && ehit[(size_t)c * (K + 1) + K]) { pl->stats[3]++; continue; }
        int32_t parent_index = near[c];
        float psi_near = 0, psi_low = 0, psi_tmp = 0;
        float lowest = chain_cost(pl, near[c], &pnew[3 * c], &psi_near);            // current_cost MS:1316
        cost[c] = lowest;
        psi[c] = psi_near;
        if (have_knn) {
            int32_t pc = -5;                                                         // extend_tree MS:838-853
            for (int d = 0; d < K; ++d) {
                const int32_t ind = knn_idx[(size_t)c * K + d];
                if (ind < 0) continue;   // unfilled slot (libnabo's InvalidIndex): the reference would index tree[-1]
                if (edge_check && ehit[(size_t)c * (K + 1) + d]) continue;
                const float temp = chain_cost(pl, ind, &pnew[3 * c], &psi_tmp);
                if (temp < lowest) { pc = ind; lowest = temp; psi_low = psi_tmp; }
            }
            if (pc > 0) {                                                            // MS:1322-1323
                parent_index = pc;
                cost[c] = lowest;
                psi[c] = psi_low;
                pl->stats[4]++;
            }
        }
        // MS:1236-1238: in the `found` branch the row already carries the Nearest index in
        // column 3 before the final push_back, so BuildPlan/cost_wiring keep reading THAT.
        parent[c] = (quirks && found[c]) ? near[c] : parent_index;
        if (quirks && found[c]) psi[c] = psi_near;                                   // psi follows the committed pointer
        ok[c] = 1;
    }
    // ---- append in candidate order; goal test MS:1340 cuts the batch
    std::vector<int32_t> new_index(B, -1);
    int acc = 0, goal_c = -1;
    for (int c = 0; c < B && goal_c < 0; ++c) {
        if (!ok[c]) continue;
        new_index[c] = (int32_t)pl->par.size();
        pl->xyz.insert(pl->xyz.end(), &pnew[3 * c], &pnew[3 * c] + 3);
        pl->par.push_back(parent[c]);
        ++acc;
        if (!no_goal && orc_norm(P.goal, &pnew[3 * c]) < P.goal_tol) goal_c = c;
    }
    // ---- rewiring (beyond reference: MS:883/923 only print): neighbour `ind` adopts the new
    // node when cost(new) + Norm(tree[ind], new) < cost(ind) (and the edge is free); one
    // proposal per node survives: smallest (new cost, new index).
    if (rewire && have_knn) {
        std::vector<float> best(N, INFINITY);
        std::vector<int32_t> who(N, -1);
        std::vector<float> e1, e2;
        struct Prop { int32_t ind, ni; float nc; };
        std::vector<Prop> props;
        for (int c = 0; c < B; ++c) {
            if (new_index[c] < 0) continue;
            for (int d = 0; d < K; ++d) {
                const int32_t ind = knn_idx[(size_t)c * K + d];
                if (ind <= 0 || ind == pl->par[new_index[c]]) continue;
                const float nc = cost[c] + orc_norm(&pl->xyz[3 * (size_t)ind], &pnew[3 * c]);
                float psi_ind = 0;
                const float oc = chain_cost(pl, pl->par[ind], &pl->xyz[3 * (size_t)ind], &psi_ind);
                if (!(nc < oc)) continue;
                // acyclicity (the reference's cost skips the root edge, and a batch applies its
                // proposals together): the true path length must drop too, on the pre-batch tree
                if (!(psi[c] < psi_ind)) continue;
                if (edge_check && ehit[(size_t)c * (K + 1) + d]) continue;
                props.push_back({ind, new_index[c], nc});
            }
        }
        for (const Prop& q : props)
            if (q.nc < best[q.ind] || (q.nc == best[q.ind] && q.ni < who[q.ind])) { best[q.ind] = q.nc; who[q.ind] = q.ni; }
        for (int64_t i = 1; i < N; ++i)
            if (who[i] >= 0) { pl->par[i] = who[i]; pl->stats[5]++; }
    }
    if (goal_c >= 0) {                                                               // MS:1346-1347
        pl->par.push_back(new_index[goal_c]);
        pl->xyz.insert(pl->xyz.end(), P.goal, P.goal + 3);
        pl->goal_reached = 1;
        pl->stats[7] = goal_c;
    }
    pl->stats[0]++;
    if (n_accepted) *n_accepted = acc;
    if (goal_reached) *goal_reached = pl->goal_reached;
    return 0;
}

int orc_planner_run(orc_planner* pl, int64_t max_nodes, int64_t max_batches) {
    for (int64_t b = 0; b < max_batches && !pl->goal_reached && (int64_t)pl->par.size() < max_nodes; ++b)
        orc_planner_step(pl, nullptr, nullptr);
    return pl->goal_reached;
}

}  // extern "C"
