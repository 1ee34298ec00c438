// spif_placement.cu -- host-side shard placement and swap planning behind the C-ABI (no device code).
//
// Reference: the planning half of the cache manager -- sparkinfer_reload_plan (src/llama-sparkinfer.cpp:45-91) turns the
// decayed firing rates into a copy_pair list (ggml-sparkinfer.hpp:16-24).  Re-targeted: every neuron group is
// resident on exactly one GPU shard and the quantity to equalise is the summed score per shard.  Same decisions, tie-breaks
// included, as sparkinfer_b200/placement.py (tests/test_sharding.py compares the two).
#include <algorithm>
#include <cmath>
#include <cstdint>
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
