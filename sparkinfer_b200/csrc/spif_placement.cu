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
