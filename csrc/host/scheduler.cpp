// Client -> GPU placement for the federated workloads (stands in for the TFF simulation executor stack that
// decides where each client's work runs; /root/reference/fed_model.py:207-229 runs them sequentially in one
// process; SURVEY.md 2.C "gRPC / TFF C++ executors", 2.D "Federated"). One process per GPU trains the clients
// assigned to it; a round ends when the slowest rank is done, so the assignment minimises the maximum per-rank
// work: longest-processing-time-first (sort clients by cost, give each to the least loaded rank), which is
// within 4/3 of the optimal makespan and exact for equal-sized clients (then it degenerates to round-robin).
// Deterministic: ties break towards the lower client id / lower rank, so every rank computes the same plan.
#include <algorithm>
#include <cstdint>
#include <numeric>
#include <vector>

#include "host_common.h"
#include "idcb200_host.h"

extern "C" int idcbh_schedule_clients(int n_clients, const long long* cost, int n_ranks, int* rank_of_client,
                                      long long* load_of_rank) {
    if (n_clients < 0 || n_ranks < 1 || (n_clients > 0 && (!cost || !rank_of_client))) {
        idcbh::set_error("schedule_clients: bad arguments");
        return 1;
    }
    std::vector<int> order(static_cast<size_t>(n_clients));
    std::iota(order.begin(), order.end(), 0);
    std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return cost[a] > cost[b]; });
    std::vector<long long> load(static_cast<size_t>(n_ranks), 0);
    std::vector<int> count(static_cast<size_t>(n_ranks), 0);
    for (int c : order) {
        int best = 0;
        for (int r = 1; r < n_ranks; ++r) {
            // least loaded; among equals the one with fewer clients (keeps per-rank client counts even, which the
            // aggregation kernels' [k_local] blocks like), then the lower rank
            if (load[r] < load[best] || (load[r] == load[best] && count[r] < count[best])) best = r;
        }
        rank_of_client[c] = best;
        load[static_cast<size_t>(best)] += cost[c] > 0 ? cost[c] : 0;
        count[static_cast<size_t>(best)] += 1;
    }
    if (load_of_rank)
        for (int r = 0; r < n_ranks; ++r) load_of_rank[r] = load[static_cast<size_t>(r)];
    return 0;
}
