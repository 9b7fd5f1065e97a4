// What knn_graph.cu hands to its callers (api_knn.cu): ONE definition of the structs, so that both sides agree on their
// layout.
#pragma once

#include "common.cuh"

namespace mcb {

// The balanced-rank CSR of knn_graph.py:300-340 plus the preserved edge of every row.
struct BalancedRanks {
    Scratch data, indices, indptr, best_index, best_value, best_raw;  // best_raw: the preserved value before max(., 1)
    Scratch class_rows, class_counts;  // the rows by band-width class (prune_mirror.cuh): int32 [3, n] and int [4] ([3]: a flag)
    int64_t nnz = 0;
    int64_t widest = 0;  // the longest row
};

// knn_graph.py:343-392 (_prune_ranks) after the balanced CSR exists.
struct PrunedRanks {
    Scratch data, indices, indptr;
    int64_t nnz = 0;
    int32_t min_degree = 0;
};

int balance_ranks_launch(const float* ranks, int64_t n, float shift, BalancedRanks* out, cudaStream_t stream);
int prune_ranks_launch(const BalancedRanks& balanced, int64_t n, int64_t incoming_degree, int64_t outgoing_degree,
                       PrunedRanks* out, cudaStream_t stream);

}  // namespace mcb
