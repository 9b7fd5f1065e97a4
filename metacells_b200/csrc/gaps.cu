// compute_cell_gaps (gaps.cpp:6-186; bound as xt.compute_cell_gaps, tools/deviants.py:309-326) on the GPU.
//
// Per candidate metacell and gene the reference sorts the candidate's cells by the gene's log fraction and looks, at the
// low and at the high end of that order, for a gap of more than the gene's minimal fold between a cell and the next
// (`gap_skip_cells` further); the cells beyond the gap get the gap's size as their score.  Two things shape the kernel:
//
//  * The cells are kept in ONE array per candidate that std::sort permutes gene after gene, so where log fractions are
//    equal (cells of equal totals and counts) the order of a gene depends on the genes before it, and it decides which of
//    the equal cells stands next to a gap.  The order is therefore reproduced: one warp per candidate walks the genes in
//    order and its first lane runs libstdc++'s std::sort step for step (std_sort.cuh) on keys staged in shared memory by
//    all 32 lanes.  Candidates are independent of each other (~330 of ~48 cells in a 16 000-cell pile).
//  * Only genes with at least one cell of `umis * 2 + 1e-6 > min_compare_umis` are sorted at all (gaps.cpp:72-84): a first
//    kernel marks them for all candidates in one coalesced pass over the dense UMIs.
#include <algorithm>
#include <vector>

#include "common.cuh"
#include "std_sort.cuh"

namespace mcb {

namespace {

// flags[candidate, gene] = some cell of the candidate has significant UMIs of the gene (gaps.cpp:72-84).
__global__ void significant_genes_kernel(const float* __restrict__ umis, int64_t stride, int64_t genes,
                                         const int32_t* __restrict__ members, const int64_t* __restrict__ member_offsets,
                                         float min_compare_umis, uint8_t* __restrict__ flags) {
    const int candidate = blockIdx.y;
    const int64_t gene = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (gene >= genes) return;
    bool significant = false;
    for (int64_t at = member_offsets[candidate]; at < member_offsets[candidate + 1] && !significant; ++at) {
        const float value = umis[int64_t(members[at]) * stride + gene];
        significant = double(value * 2.0f) + 1e-6 > double(min_compare_umis);
    }
    flags[int64_t(candidate) * genes + gene] = significant;
}

struct GapParams {
    const float* umis;
    const float* log_fraction;
    int64_t stride, genes;
    const int32_t* members;
    const int64_t* member_offsets;
    const int32_t* positions_limit;  // per candidate: min((cells - 1) / 2, max_deviant_cells_count_of_candidate)
    const uint8_t* flags;
    const float* min_gap_per_gene;
    int gap_skip_cells;
    float min_compare_umis;
    float* max_gap_per_cell;
};

__global__ void __launch_bounds__(32) cell_gaps_kernel(GapParams P) {
    extern __shared__ __align__(16) unsigned char shared[];
    const int candidate = blockIdx.x, lane = threadIdx.x;
    const int64_t first_member = P.member_offsets[candidate];
    const int count = int(P.member_offsets[candidate + 1] - first_member);
    const int32_t* members = P.members + first_member;
    // cell_index_of_position, as slots of `members`, each with the bits of its key above it: the sort compares and moves
    // whole entries (the same comparisons and moves as the reference's sort of indices through a comparator that looks the
    // keys up), one shared-memory load per operand instead of two dependent ones
    unsigned long long* order = reinterpret_cast<unsigned long long*>(shared);
    float* key = reinterpret_cast<float*>(order + count);  // the gene's log fraction of every cell of the candidate
    float* count_of = key + count;                         // its UMIs
    float* gap_of = count_of + count;                      // the cell's largest gap so far
    for (int slot = lane; slot < count; slot += 32) {
        order[slot] = (unsigned long long)slot;
        gap_of[slot] = P.max_gap_per_cell[members[slot]];
    }
    const int positions = P.positions_limit[candidate];
    const int skip = P.gap_skip_cells;
    const uint8_t* flags = P.flags + int64_t(candidate) * P.genes;
    for (int64_t base = 0; base < P.genes; base += 32) {
        unsigned pending = __ballot_sync(0xffffffffu, base + lane < P.genes && flags[base + lane] != 0);
        while (pending) {
            const int64_t gene = base + __ffs(pending) - 1;
            pending &= pending - 1;
            __syncwarp();
            for (int slot = lane; slot < count; slot += 32) {
                const int64_t at = int64_t(members[slot]) * P.stride + gene;
                key[slot] = P.log_fraction[at];
                count_of[slot] = P.umis[at];
            }
            __syncwarp();
            for (int place = lane; place < count; place += 32) {
                const unsigned slot = (unsigned)(order[place] & 0xffffu);
                order[place] = ((unsigned long long)__float_as_uint(key[slot]) << 32) | slot;
            }
            __syncwarp();
            if (lane != 0) continue;
            serial_sort(order, count, [](unsigned long long left, unsigned long long right) {
                return __uint_as_float((unsigned)(left >> 32)) < __uint_as_float((unsigned)(right >> 32));
            });
            const float min_gap = P.min_gap_per_gene[gene];
            for (int position = 0; position < positions; ++position) {  // gaps.cpp:106-137: the low end
                const int first = (int)(order[position] & 0xffffu), second = (int)(order[position + skip] & 0xffffu);
                if (double(__fadd_rn(count_of[first], count_of[second])) + 1e-6 < double(P.min_compare_umis)) continue;
                const float full_gap = __fsub_rn(__fsub_rn(key[second], key[first]), min_gap);
                if (full_gap < 0.0f) continue;
                int stop = position + 1;
                if (skip > 1) {
                    const float start_gap =
                        float(2.0 * double(__fsub_rn(key[order[position + 1] & 0xffffu], key[first])) - double(min_gap));
                    if (start_gap < full_gap) stop += 1;
                }
                for (int at = 0; at <= stop; ++at) {
                    const int slot = (int)(order[at] & 0xffffu);
                    gap_of[slot] = fmaxf(gap_of[slot], full_gap);
                }
            }
            for (int offset = 0; offset < positions; ++offset) {  // gaps.cpp:139-178: the high end
                const int position = count - 1 - offset - skip;
                const int first = (int)(order[position] & 0xffffu), second = (int)(order[position + skip] & 0xffffu);
                if (__fadd_rn(count_of[first], count_of[second]) < P.min_compare_umis) continue;
                const float full_gap = __fsub_rn(__fsub_rn(key[second], key[first]), min_gap);
                if (full_gap <= 0.0f) continue;
                int start = position + 1;
                if (skip > 1) {
                    const float start_gap =
                        float(2.0 * double(__fsub_rn(key[order[position + 1] & 0xffffu], key[first])) - double(min_gap));
                    if (start_gap < full_gap) start += 1;
                }
                for (int at = start; at < count; ++at) {
                    const int slot = (int)(order[at] & 0xffffu);
                    gap_of[slot] = fmaxf(gap_of[slot], full_gap);
                }
            }
        }
    }
    __syncwarp();
    for (int slot = lane; slot < count; slot += 32) P.max_gap_per_cell[members[slot]] = gap_of[slot];
}

template <typename T>
int to_device(Scratch& device, const T* host, size_t count, cudaStream_t stream) {
    MCB_TRY(device.alloc(count * sizeof(T), stream));
    if (count) MCB_CUDA(cudaMemcpyAsync(device.ptr, host, count * sizeof(T), cudaMemcpyHostToDevice, stream));
    return MCB_OK;
}

}  // namespace

// umis / log_fraction: device, [cells, stride >= genes]; every other array is the host's.  max_gap_per_cell is read and
// written (the reference takes the maximum with what is there).
int cell_gaps_launch(const float* umis, const float* log_fraction, int64_t cells, int64_t genes, int64_t stride,
                     const int32_t* candidate_of_cells, const uint8_t* deviant_per_cell, const uint8_t* active_per_cell,
                     const float* min_gap_per_gene, int64_t candidates_count, int64_t gap_skip_cells,
                     int64_t max_deviant_cells_count, float max_deviant_cells_fraction, float min_compare_umis,
                     float* max_gap_per_cell, cudaStream_t stream) {
    MCB_REQUIRE(gap_skip_cells >= 1 && gap_skip_cells <= 3, "compute_cell_gaps: gap_skip_cells must be 1, 2 or 3");  // :43-44
    MCB_REQUIRE(cells > 0 && genes > 0 && candidates_count >= 0, "compute_cell_gaps: empty input");
    // gaps.cpp:47-64: the cells of every candidate that are not deviant yet, in cell order; a candidate is looked at when
    // one of its cells is active and at least four are left
    std::vector<std::vector<int32_t>> cells_of{size_t(candidates_count)};
    std::vector<uint8_t> is_active(static_cast<size_t>(candidates_count), uint8_t(0));
    for (int64_t cell = 0; cell < cells; ++cell) {
        const int64_t candidate = candidate_of_cells[cell];
        if (candidate < 0 || candidate >= candidates_count) continue;
        if (active_per_cell[cell]) is_active[size_t(candidate)] = 1;
        if (!deviant_per_cell[cell]) cells_of[size_t(candidate)].push_back(int32_t(cell));
    }
    std::vector<int32_t> members, limits;
    std::vector<int64_t> offsets(1, 0);
    size_t largest = 0;
    for (int64_t candidate = 0; candidate < candidates_count; ++candidate) {
        const auto& list = cells_of[size_t(candidate)];
        const size_t count = list.size();
        if (!is_active[size_t(candidate)] || count < 4) continue;
        const size_t limit = std::min(count - size_t(gap_skip_cells),
                                      std::max(size_t(max_deviant_cells_count), size_t(max_deviant_cells_fraction * count + 0.5)));
        limits.push_back(int32_t(std::min((count - 1) / 2, limit)));
        members.insert(members.end(), list.begin(), list.end());
        offsets.push_back(int64_t(members.size()));
        largest = std::max(largest, count);
    }
    const int64_t listed = int64_t(limits.size());
    if (listed == 0) return MCB_OK;
    MCB_REQUIRE(largest <= 8192, "compute_cell_gaps: a candidate of %lld cells (at most 8192 are supported)", (long long)largest);
    Scratch d_members, d_offsets, d_limits, d_flags, d_min_gap, d_max_gap;
    MCB_TRY(to_device(d_members, members.data(), members.size(), stream));
    MCB_TRY(to_device(d_offsets, offsets.data(), offsets.size(), stream));
    MCB_TRY(to_device(d_limits, limits.data(), limits.size(), stream));
    MCB_TRY(to_device(d_min_gap, min_gap_per_gene, size_t(genes), stream));
    MCB_TRY(to_device(d_max_gap, max_gap_per_cell, size_t(cells), stream));
    MCB_TRY(d_flags.alloc(size_t(listed) * size_t(genes), stream));
    MCB_REQUIRE(listed <= 65535, "compute_cell_gaps: more than 65535 candidates");
    MCB_PROFILED("significant_genes", stream,
                 significant_genes_kernel<<<dim3(unsigned(ceil_div(genes, 256)), unsigned(listed)), 256, 0, stream>>>(
                     umis, stride, genes, d_members.as<int32_t>(), d_offsets.as<int64_t>(), min_compare_umis,
                     d_flags.as<uint8_t>()));
    MCB_LAUNCH_CHECK();
    GapParams params{umis, log_fraction, stride, genes, d_members.as<int32_t>(), d_offsets.as<int64_t>(),
                     d_limits.as<int32_t>(), d_flags.as<uint8_t>(), d_min_gap.as<float>(), int(gap_skip_cells),
                     min_compare_umis, d_max_gap.as<float>()};
    const size_t shared_bytes = (largest * 20 + 15) / 16 * 16;
    MCB_TRY(allow_dynamic_shared(reinterpret_cast<const void*>(cell_gaps_kernel), shared_bytes));
    MCB_PROFILED("cell_gaps", stream, cell_gaps_kernel<<<unsigned(listed), 32, shared_bytes, stream>>>(params));
    MCB_LAUNCH_CHECK();
    MCB_TRY(copy_to_host(max_gap_per_cell, d_max_gap.ptr, size_t(cells) * sizeof(float), stream));
    MCB_CUDA(cudaStreamSynchronize(stream));
    return MCB_OK;
}

}  // namespace mcb
