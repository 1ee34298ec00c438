// spif_placement.cu -- host-side shard placement and swap planning behind the C-ABI (no device code).
//
// Reference: the planning half of the cache manager -- sparkinfer_reload_plan (src/llama-sparkinfer.cpp:45-91) turns the
// decayed firing rates into a copy_pair list (ggml-sparkinfer.hpp:16-24).  Re-targeted: every neuron group is
// resident on exactly one GPU shard and the quantity to equalise is the summed score per shard.  Same decisions, tie-breaks
// included, as sparkinfer_b200/placement.py (tests/test_sharding.py compares the two).
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <numeric>
#include <vector>

#include "../../include/spif_b200.h"

extern "C" int spif_initial_placement(const float * group_hotness, int n_g, int n_shards, int32_t * shard_of_group) {
    if (!group_hotness || !shard_of_group) {
        return SPIF_E_NULL;
    }
    if (n_g <= 0 || n_shards <= 0 || n_g % n_shards != 0) {
        return SPIF_E_SHAPE;
    }
    std::vector<int> order(n_g);
    std::iota(order.begin(), order.end(), 0);
    std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return group_hotness[a] > group_hotness[b]; });
    for (int k = 0; k < n_g; ++k) {
        const int rnd = k / n_shards, pos = k % n_shards;
        shard_of_group[order[k]] = (rnd % 2 == 0) ? pos : n_shards - 1 - pos;  // snake deal
    }
    return 0;
}

extern "C" int spif_rebalance_plan(const float * dfr, int32_t * shard_of_group, int n_g, int n_shards, int max_swaps,
                                   float min_gain, int32_t * plan, int * n_swaps) {
    if (!dfr || !shard_of_group || !n_swaps || (max_swaps > 0 && !plan)) {
        return SPIF_E_NULL;
    }
    if (n_g <= 0 || n_shards <= 0) {
        return SPIF_E_SHAPE;
    }
    for (int i = 0; i < n_g; ++i) {
        if (shard_of_group[i] < 0 || shard_of_group[i] >= n_shards) {
            return SPIF_E_SHAPE;
        }
    }
    const double inf = std::numeric_limits<double>::infinity();
    std::vector<double> load(n_shards);
    std::vector<int>    gh, gc, oc;
    std::vector<double> dc;
    int                 n = 0;
    for (; n < max_swaps; ++n) {
        std::fill(load.begin(), load.end(), 0.0);
        for (int i = 0; i < n_g; ++i) {
            load[shard_of_group[i]] += (double) dfr[i];
        }
        const int    hot  = (int) (std::max_element(load.begin(), load.end()) - load.begin());  // first maximum
        const int    cold = (int) (std::min_element(load.begin(), load.end()) - load.begin());  // first minimum
        const double gap  = load[hot] - load[cold];
        if (hot == cold || !(gap > 0.0)) {
            break;
        }
        gh.clear(), gc.clear();
        for (int i = 0; i < n_g; ++i) {
            if (shard_of_group[i] == hot) {
                gh.push_back(i);
            } else if (shard_of_group[i] == cold) {
                gc.push_back(i);
            }
        }
        // moving delta = d[a] - d[b] from hot to cold: the larger load becomes mid + |delta - gap/2| -> the best pair has
        // delta > 0 closest to gap/2; for every a the candidates are the two neighbours of d[a] - gap/2 among the cold scores
        oc.resize(gc.size());
        std::iota(oc.begin(), oc.end(), 0);
        std::stable_sort(oc.begin(), oc.end(), [&](int x, int y) { return dfr[gc[x]] < dfr[gc[y]]; });
        dc.resize(gc.size());
        for (size_t j = 0; j < gc.size(); ++j) {
            dc[j] = (double) dfr[gc[oc[j]]];
        }
        double best = inf;
        int    bi = -1, bj = -1;
        for (int side = 0; side < 2; ++side) {  // side 0: neighbour below the target, side 1: at / above it
            double cbest = inf;
            int    ci = -1, cj = -1;
            for (size_t i = 0; i < gh.size(); ++i) {
                const double da  = (double) dfr[gh[i]];
                long         pos = std::lower_bound(dc.begin(), dc.end(), da - gap / 2) - dc.begin();
                pos              = std::min<long>(std::max<long>(pos - (side == 0 ? 1 : 0), 0), (long) gc.size() - 1);
                const double delta = da - dc[pos];
                const double cost  = (delta > 0.0) ? load[cold] + gap / 2 + std::fabs(delta - gap / 2) : inf;
                if (cost < cbest) {  // first minimum over a
                    cbest = cost, ci = (int) i, cj = oc[pos];
                }
            }
            if (cbest < best) {
                best = cbest, bi = ci, bj = cj;
            }
        }
        if (!(best < inf) || load[hot] - best <= (double) min_gain) {
            break;
        }
        const int a = gh[bi], b = gc[bj];
        shard_of_group[a] = cold;
        shard_of_group[b] = hot;
        plan[4 * n + 0] = a, plan[4 * n + 1] = b, plan[4 * n + 2] = hot, plan[4 * n + 3] = cold;
    }
    *n_swaps = n;
    return 0;
}

// sparkinfer_cache_manager ctor, src/llama-sparkinfer.cpp:183-202: per-layer share of the VRAM budget, then round-robin
extern "C" int spif_cache_group_budget(int64_t vram_budget, int64_t group_bytes, int has_gate, const float * pattern,
                                       int n_layer, int n_group, int32_t * groups_per_layer) {
    if (!pattern || !groups_per_layer) {
        return SPIF_E_NULL;
    }
    if (vram_budget < 0 || group_bytes <= 0 || n_layer <= 0 || n_group <= 0) {
        return SPIF_E_SHAPE;
    }
    const int budget = (int) std::min<int64_t>(vram_budget / (group_bytes * (has_gate ? 3 : 2)), (int64_t) n_layer * n_group);
    int       used   = 0;
    for (int il = 0; il < n_layer; ++il) {
        const int init       = std::min<int>((int) ((float) budget * pattern[il]), n_group);  // int * float, truncated (:185)
        groups_per_layer[il] = init < 0 ? 0 : init;
        used += groups_per_layer[il];
    }
    for (int left = budget - used; left > 0;) {
        const int before = left;
        for (int il = 0; il < n_layer && left > 0; ++il) {
            if (groups_per_layer[il] < n_group) {
                ++groups_per_layer[il];
                --left;
            }
        }
        if (left == before) {
            break;  // every layer is full
        }
    }
    return 0;
}

// reorder_tensor_2d / _1d, src/llama-sparkinfer.cpp:285-311
extern "C" int spif_reorder_rows(const void * src, void * dst, const int32_t * perm, int64_t n_rows, int64_t row_bytes) {
    if (!src || !dst || !perm) {
        return SPIF_E_NULL;
    }
    if (n_rows < 0 || row_bytes <= 0 || src == dst) {
        return SPIF_E_SHAPE;
    }
    for (int64_t i = 0; i < n_rows; ++i) {
        if (perm[i] < 0 || perm[i] >= n_rows) {
            return SPIF_E_SHAPE;
        }
    }
    const unsigned char * s = (const unsigned char *) src;
    unsigned char *       d = (unsigned char *) dst;
    for (int64_t i = 0; i < n_rows; ++i) {
        std::memcpy(d + i * row_bytes, s + (int64_t) perm[i] * row_bytes, (size_t) row_bytes);
    }
    return 0;
}

// ---------------------------------------------------------------------------------------------------------------
// spif_balancer: the host-side state of the online shard balancer -- the counterpart of the reference's per-layer cache_meta /
// group_maps / dfr_decay_pack (src/llama-sparkinfer.h, ggml/include/ggml-sparkinfer.hpp:16-24,143-175) and of the
// plan -> exec -> anchor split of ggml_sparkinfer_reload (ggml-cuda.cu:2499-2606), re-targeted from "which groups sit in the GPU
// cache" to "which GPU shard holds which group":
//   plan    spif_balancer_plan: gathered per-slot scores -> per-layer swap plans (spif_rebalance_plan) -> the moves THIS rank has
//           to execute, in an order both ends of every move agree on (round-major, then layer); the new slot tables are PENDING;
//   exec    the caller moves the rows (NVLink send/recv or peer copies on a side stream) -- device work, not done here;
//   anchor  spif_balancer_commit makes the pending tables current (the reference flips group_maps / neuron_idx after the
//           copies of a window completed); spif_balancer_abort drops them.
// The decay lambda adapts like ggml-sparkinfer.hpp:169-173 (lambda *= 1 +- step, clamped to [0.05, 0.95]); the reference's
// trigger "compute had to wait for the reload" becomes "the balance the last plan promised did not hold until this plan".
struct spif_balancer {
    int   n_layer, n_g, n_shards, rank, n_slot;
    float lambda, lambda_step, drift_tol;
    float promised;  // mean planned imbalance of the last plan, < 0: none yet
    std::vector<int32_t> slots, pending;  // [n_layer][n_shards][n_slot] -> group id
    bool  has_pending;
    long long swaps_done;
};

namespace {
double layer_imbalance(const std::vector<double> & d, const int32_t * sog, int n_g, int n_shards) {
    std::vector<double> load(n_shards, 0.0);
    for (int i = 0; i < n_g; ++i) {
        load[sog[i]] += d[i];
    }
    double mx = 0.0, sum = 0.0;
    for (double v : load) {
        mx = std::max(mx, v), sum += v;
    }
    return mx / std::max(sum / n_shards, 1e-30);
}
}  // namespace

extern "C" int spif_balancer_create(int n_layer, int n_g, int n_shards, int rank, const int32_t * shard_of_group, float lambda0,
                                    float lambda_step, float drift_tol, spif_balancer ** out) {
    if (!shard_of_group || !out) {
        return SPIF_E_NULL;
    }
    if (n_layer <= 0 || n_g <= 0 || n_shards <= 0 || n_g % n_shards != 0 || rank < 0 || rank >= n_shards) {
        return SPIF_E_SHAPE;
    }
    auto * b       = new spif_balancer();
    b->n_layer     = n_layer, b->n_g = n_g, b->n_shards = n_shards, b->rank = rank, b->n_slot = n_g / n_shards;
    b->lambda      = lambda0, b->lambda_step = lambda_step, b->drift_tol = drift_tol;
    b->promised    = -1.f;
    b->has_pending = false;
    b->swaps_done  = 0;
    b->slots.assign((size_t) n_layer * n_g, -1);
    // slot s of shard r initially holds the s-th smallest group id placed on r
    for (int l = 0; l < n_layer; ++l) {
        std::vector<int> fill(n_shards, 0);
        for (int i = 0; i < n_g; ++i) {
            const int r = shard_of_group[(size_t) l * n_g + i];
            if (r < 0 || r >= n_shards || fill[r] >= b->n_slot) {
                delete b;
                return SPIF_E_SHAPE;  // not an equal split
            }
            b->slots[((size_t) l * n_shards + r) * b->n_slot + fill[r]++] = i;
        }
    }
    *out = b;
    return 0;
}

extern "C" void spif_balancer_destroy(spif_balancer * b) {
    delete b;
}

extern "C" int spif_balancer_slots(const spif_balancer * b, int layer, int32_t * slots_out) {
    if (!b || !slots_out) {
        return SPIF_E_NULL;
    }
    if (layer < 0 || layer >= b->n_layer) {
        return SPIF_E_SHAPE;
    }
    std::memcpy(slots_out, b->slots.data() + (size_t) layer * b->n_g, sizeof(int32_t) * (size_t) b->n_g);
    return 0;
}

extern "C" int spif_balancer_state(const spif_balancer * b, float * lambda, int64_t * swaps_done, int * has_pending) {
    if (!b) {
        return SPIF_E_NULL;
    }
    if (lambda) {
        *lambda = b->lambda;
    }
    if (swaps_done) {
        *swaps_done = b->swaps_done;
    }
    if (has_pending) {
        *has_pending = b->has_pending ? 1 : 0;
    }
    return 0;
}

// slot_scores: [n_shards][n_layer][n_slot], the all-gathered decayed firing rates of every shard's slots (host memory).
// moves_out[i] = {layer, round, my slot, peer shard, group leaving, group arriving} for the moves of THIS rank, sorted by
// (round, layer): swap k of a layer may move a group that swap k-1 brought in, so rounds are ordered; layers are independent.
extern "C" int spif_balancer_plan(spif_balancer * b, const float * slot_scores, int max_swaps, float min_gain, spif_move * moves_out,
                                  int moves_cap, int * n_moves, float * imbalance_before, float * imbalance_after,
                                  int * n_swaps_total) {
    if (!b || !slot_scores || !n_moves || (moves_cap > 0 && !moves_out)) {
        return SPIF_E_NULL;
    }
    if (max_swaps < 0 || b->has_pending) {
        return SPIF_E_SHAPE;  // commit or abort the previous plan first
    }
    const int L = b->n_layer, G = b->n_g, S = b->n_shards, NS = b->n_slot;
    std::vector<double>  d(G);
    std::vector<float>   df(G);
    std::vector<int32_t> sog(G), plan((size_t) 4 * std::max(max_swaps, 1));
    std::vector<std::vector<int32_t>> plans(L);
    double before = 0.0, after = 0.0;
    b->pending    = b->slots;
    int total     = 0;
    for (int l = 0; l < L; ++l) {
        int32_t * tab = b->pending.data() + (size_t) l * G;  // [S][NS]
        for (int r = 0; r < S; ++r) {
            for (int s = 0; s < NS; ++s) {
                const int g = tab[r * NS + s];
                d[g]        = (double) slot_scores[((size_t) r * L + l) * NS + s];
                df[g]       = (float) d[g];
                sog[g]      = r;
            }
        }
        before += layer_imbalance(d, sog.data(), G, S);
        int n = 0;
        const int rc = spif_rebalance_plan(df.data(), sog.data(), G, S, max_swaps, min_gain, plan.data(), &n);
        if (rc) {
            b->pending.clear();
            return rc;
        }
        after += layer_imbalance(d, sog.data(), G, S);
        plans[l].assign(plan.begin(), plan.begin() + 4 * n);
        total += n;
    }
    // lambda adaptation against what the previous plan promised
    if (b->lambda_step > 0.f && b->promised >= 0.f) {
        const bool drifted = (float) (before / L) > b->promised * (1.f + b->drift_tol);
        b->lambda = std::min(0.95f, std::max(0.05f, b->lambda * (1.f + (drifted ? b->lambda_step : -b->lambda_step))));
    }
    b->promised = (float) (after / L);
    // moves of this rank, round-major; the pending tables follow every swap so that later swaps find the moved groups
    int nm = 0;
    for (int k = 0; k < max_swaps; ++k) {
        for (int l = 0; l < L; ++l) {
            if ((size_t) 4 * k >= plans[l].size()) {
                continue;
            }
            int32_t *     tab = b->pending.data() + (size_t) l * G;
            const int32_t ga = plans[l][4 * k], gb = plans[l][4 * k + 1], hot = plans[l][4 * k + 2], cold = plans[l][4 * k + 3];
            int sa = -1, sb = -1;
            for (int s = 0; s < NS; ++s) {
                sa = (tab[hot * NS + s] == ga) ? s : sa;
                sb = (tab[cold * NS + s] == gb) ? s : sb;
            }
            if (sa < 0 || sb < 0) {
                b->pending.clear();
                return SPIF_E_SHAPE;
            }
            if (b->rank == hot || b->rank == cold) {
                if (nm < moves_cap) {
                    spif_move & m = moves_out[nm];
                    m.layer = l, m.round = k;
                    m.slot      = (b->rank == hot) ? sa : sb;
                    m.peer      = (b->rank == hot) ? cold : hot;
                    m.group_out = (b->rank == hot) ? ga : gb;
                    m.group_in  = (b->rank == hot) ? gb : ga;
                }
                ++nm;
            }
            tab[hot * NS + sa]  = gb;
            tab[cold * NS + sb] = ga;
        }
    }
    *n_moves = nm;
    if (imbalance_before) {
        *imbalance_before = (float) (before / L);
    }
    if (imbalance_after) {
        *imbalance_after = (float) (after / L);
    }
    if (n_swaps_total) {
        *n_swaps_total = total;
    }
    if (nm > moves_cap) {
        b->pending.clear();
        return SPIF_E_WORKSPACE;  // *n_moves tells the caller how many entries it needs
    }
    b->has_pending = true;
    b->swaps_done += total;
    return 0;
}

extern "C" int spif_balancer_commit(spif_balancer * b) {
    if (!b) {
        return SPIF_E_NULL;
    }
    if (!b->has_pending) {
        return SPIF_E_SHAPE;
    }
    b->slots.swap(b->pending);
    b->pending.clear();
    b->has_pending = false;
    return 0;
}

extern "C" int spif_balancer_abort(spif_balancer * b) {
    if (!b) {
        return SPIF_E_NULL;
    }
    b->pending.clear();
    b->has_pending = false;
    return 0;
}
