/*
 * plan.cpp — msbfs_plan_shares: which rank runs which query groups, and in what order.
 *
 * Replaces the reference's round-robin ownership (`kidx = rank; kidx += world_size`,
 * /root/reference/main.cu:303-307).  Host-only; the one planner behind the C ABI that both
 * the drop-in executable (host/main.cpp) and the benchmark harness call, so the two cannot
 * drift apart.
 *
 * The rule.  Groups are ordered heaviest first by `weights` (msbfs_group_weights: how fast
 * a group's BFS reaches the hubs).  A batch — the groups that advance together as the lanes
 * of one traversal — is a run of consecutive groups of that order: lanes that reach the dense
 * levels at the same depth let the bottom-up kernels stop early for the whole batch.  The
 * cost of a batch is dominated by the lane-count-independent scan of the CSR, so batches are
 * as wide as the hardware word allows (512) unless that would leave a rank with fewer than
 * two batches to balance with AND the batches' weights differ by more than 3 x (then 256), or
 * a rank without any batch (then halved down to 64).  Batches
 * are handed to the ranks
 *   - by measured cost when the caller has one (`batch_cost`, e.g. msbfs_get_batch_ms of a
 *     previous pass over the same plan): longest batch first, each to the rank with the
 *     least work so far;
 *   - else in snake order 0..R-1, R-1..0, ...: the order runs from fast (hub sources) to
 *     slow batches, and the snake gives every rank some of each.
 * Any assignment yields the same sums and the same argmin (groups are independent).
 */
#include "internal.h"

#include <algorithm>
#include <cstdint>
#include <numeric>
#include <vector>

extern "C" int msbfs_plan_lanes(int32_t num_groups, int32_t world) {
    if (world <= 1) return 0; /* one rank: the engine's own choice */
    int lanes = 512;
    auto batches = [&](int l) { return ((int64_t)num_groups + l - 1) / l; };
    if (batches(lanes) < 2 * (int64_t)world) lanes = 256;
    while (lanes > 64 && batches(lanes) < world) lanes /= 2;
    return lanes;
}

extern "C" int msbfs_plan_shares(int32_t num_groups, const int64_t* weights, int32_t world, int32_t lanes,
                                 const double* batch_cost, int32_t num_costs, int32_t* order_out, int32_t* owner_out,
                                 int32_t* lanes_out) {
    if (num_groups < 0 || world < 1 || (num_groups > 0 && (!weights || !order_out || !owner_out)))
        return msbfs_fail(MSBFS_ERR_INVALID, "msbfs_plan_shares: bad argument");
    if (lanes != 0 && lanes != 64 && lanes != 128 && lanes != 256 && lanes != 512)
        return msbfs_fail(MSBFS_ERR_INVALID, "msbfs_plan_shares: lanes must be 0, 64, 128, 256 or 512");
    const int K = num_groups;
    std::vector<int32_t> order((size_t)K);
    std::iota(order.begin(), order.end(), 0);
    std::stable_sort(order.begin(), order.end(), [&](int32_t x, int32_t y) { return weights[x] > weights[y]; });
    int plan_lanes = lanes ? lanes : msbfs_plan_lanes(K, world);
    if (!lanes && plan_lanes == 256 && ((int64_t)K + 511) / 512 >= world) {
        /* One 512-lane batch per rank costs less than two 256-lane ones (the scan of the CSR is paid per
         * batch) — unless the batches differ so much in cost that a rank needs two of them to mix a
         * fast one with a slow one.  The weights tell: on a graph without hubs (Erdos-Renyi: 2 GPUs
         * 40.9 ms with 2 x 256 lanes per rank, 28 ms with 1 x 512) every 512-group batch weighs the
         * same; on a power-law graph the first weighs orders of magnitude more than the last. */
        const int nb512 = (K + 511) / 512;
        long double lo = -1, hi = 0;
        for (int b = 0; b < nb512; ++b) {
            long double w = 0;
            for (int i = b * 512; i < std::min(K, (b + 1) * 512); ++i) w += (long double)weights[order[(size_t)i]];
            if (b == nb512 - 1 && K % 512) w *= 512.0L / (K % 512); /* a partial last batch, scaled up */
            if (lo < 0 || w < lo) lo = w;
            if (w > hi) hi = w;
        }
        if (hi <= 3 * lo) plan_lanes = 512;
    }
    if (lanes_out) *lanes_out = plan_lanes;
    for (int i = 0; i < K; ++i) order_out[i] = order[(size_t)i];
    if (world == 1 || K == 0) {
        for (int i = 0; i < K; ++i) owner_out[i] = 0;
        return MSBFS_OK;
    }
    const int width = plan_lanes ? plan_lanes : 512;
    const int nb = (K + width - 1) / width;
    std::vector<int> batch_owner((size_t)nb, 0);
    if (batch_cost && num_costs == nb) {
        /* longest processing time first onto the least-loaded rank; ties keep the lower index */
        std::vector<int> by_cost((size_t)nb);
        std::iota(by_cost.begin(), by_cost.end(), 0);
        std::stable_sort(by_cost.begin(), by_cost.end(), [&](int x, int y) { return batch_cost[x] > batch_cost[y]; });
        std::vector<double> load((size_t)world, 0.0);
        for (int b : by_cost) {
            int best = 0;
            for (int r = 1; r < world; ++r)
                if (load[(size_t)r] < load[(size_t)best]) best = r;
            batch_owner[(size_t)b] = best;
            load[(size_t)best] += batch_cost[b];
        }
    } else {
        for (int b = 0; b < nb; ++b) {
            const int lap = b / world, pos = b % world;
            batch_owner[(size_t)b] = (lap % 2 == 0) ? pos : world - 1 - pos;
        }
    }
    for (int i = 0; i < K; ++i) owner_out[order[(size_t)i]] = batch_owner[(size_t)(i / width)];
    return MSBFS_OK;
}
