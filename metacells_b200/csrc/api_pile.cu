// C-ABI of the device-resident pile (see pile.cu and mcb200.h).
#include <algorithm>
#include <cstdlib>
#include <memory>
#include <vector>

#include <cstring>
#include <vector>

#include "common.cuh"
#include "../../include/mcb200.h"

namespace mcb {
int downsample_rows_launch(const float* input, float* output, const int32_t* indptr, const int32_t* samples,
                           int32_t samples_all, int64_t first_row, int64_t last_row, int64_t columns, int64_t max_row_size,
                           uint64_t seed, int use_row_seed, unsigned long long known_max_total, int* deferred_flag,
                           cudaStream_t stream, const int32_t* indices = nullptr, float* column_totals = nullptr,
                           unsigned long long* column_stats = nullptr, float top_threshold = 0.0f, const int32_t* seed_rows = nullptr);
int column_totals_launch(const float* data, const int32_t* indices, int64_t nnz, int64_t columns, float* totals,
                         cudaStream_t stream);
int log_fraction_dense_launch(const float* data, const int32_t* indices, const int32_t* indptr, int64_t rows,
                              const int32_t* slot_of_gene, int64_t features, const float* table, int64_t table_size,
                              float regularization, float* dense, int* outside_table, cudaStream_t stream);
int collect_top_launch(int64_t degree, const float* similarity, int64_t rows, int64_t columns, int64_t row_stride,
                       int32_t* out_indices, float* out_data, int ranks, cudaStream_t stream);
int correlate_dense_with_min(const float* input, int64_t rows, int64_t columns, int64_t row_stride, float* output,
                             int64_t out_stride, int absolute, unsigned int* min_key, cudaStream_t stream);
int knn_graph_consuming(float* similarity, const uint32_t* known_min, int64_t n, int64_t k, double balanced_ranks_factor,
                        double incoming_degree_factor, double outgoing_degree_factor, int64_t min_outgoing_degree,
                        cudaStream_t stream, mcb_graph** out);
}  // namespace mcb

using namespace mcb;

struct mcb_pile {
    int64_t cells = 0, genes = 0, nnz = 0, widest = 0;
    cudaStream_t stream = nullptr;
    // the UMI matrix (CSR float32 / int32 / int32): owned copies, or borrowed device pointers
    Scratch own_data, own_indices, own_indptr;
    const float* data = nullptr;
    const int32_t* indices = nullptr;
    const int32_t* indptr = nullptr;
    Scratch compact_indices, compact_counts, escape_at, escape_values;  // the staging of mcb_pile_create_host_compact
    Scratch downsampled;  // same sparsity structure as `data`
    bool has_downsampled = false;
    Scratch gene_totals;  // float32[genes]: per-gene totals of `downsampled`, added up by the downsample kernels
    // uint64[genes], when mcb_pile_want_gene_statistics asked for them: sum of squares << 16 | cells >= gene_top_threshold
    Scratch gene_stats;
    bool want_gene_stats = false;
    float gene_top_threshold = 0.0f;
    Scratch top_genes;    // int32[top_genes_count] ascending, set by mcb_pile_top_genes
    int64_t top_genes_count = 0;
    // A host pile is uploaded in row chunks on its own stream, so that the downsampling of the first chunks runs
    // under the upload of the later ones: chunk c holds rows [chunk_rows[c], chunk_rows[c + 1]).
    cudaStream_t copy_stream = nullptr;
    std::vector<int64_t> chunk_rows;
    std::vector<cudaEvent_t> chunk_ready;
    unsigned long long max_row_total = 0;  // mcb_pile_set_max_cell_total: spares a kernel + synchronisation
    Scratch status;                        // device int: deferred errors of asynchronous steps
    ~mcb_pile() {
        for (cudaEvent_t event : chunk_ready) cudaEventDestroy(event);
        if (copy_stream) {
            cudaStreamSynchronize(copy_stream);
            cudaStreamDestroy(copy_stream);
        }
    }
};

// The atlas: the cells x genes UMI matrix of a whole divide-and-conquer run (CSR float32 / int32 with 64-bit row
// pointers: a million cells hold more than 2^31 non-zeros), resident in HBM for all phases.  Piles are row subsets
// gathered on the device (mcb_pile_create_from_atlas), so a pile costs an index list on the host link, not its UMIs.
struct mcb_atlas {
    int64_t cells = 0, genes = 0, nnz = 0;
    Scratch own_data, own_indices, own_indptr;  // owned copies (create_host); empty when the arrays are borrowed
    const float* data = nullptr;
    const int32_t* indices = nullptr;
    const int64_t* indptr = nullptr;   // device
    std::vector<int64_t> host_indptr;  // the same row pointers on the host: pile sizes without a device round trip
    int device = 0;
};

namespace {

// Orders the pile's stream after the upload of a host pile (a no-op for a device pile).
int wait_for_upload(mcb_pile* pile) {
    if (!pile->chunk_ready.empty())
        MCB_CUDA(cudaStreamWaitEvent(pile->stream, pile->chunk_ready.back(), 0));
    return MCB_OK;
}

// Deferred errors of the asynchronous steps, read where the caller synchronises anyway.
int check_status(mcb_pile* pile) {
    if (!pile->status.ptr) {
        MCB_CUDA(cudaStreamSynchronize(pile->stream));
        return MCB_OK;
    }
    int status = 0;
    MCB_TRY(copy_to_host(&status, pile->status.ptr, sizeof(int), pile->stream));
    MCB_CUDA(cudaStreamSynchronize(pile->stream));
    MCB_REQUIRE(status != 3, "downsample: a cell totals more than the maximum given to mcb_pile_set_max_cell_total");
    MCB_REQUIRE(status == 0, "downsample: a cell totals 2^32 UMIs or more");
    return MCB_OK;
}

int finish_create(mcb_pile* pile, const int32_t* host_indptr) {
    int64_t widest = 0;
    for (int64_t c = 0; c < pile->cells; ++c) {
        MCB_REQUIRE(host_indptr[c + 1] >= host_indptr[c], "pile: indptr is not monotone at cell %lld", (long long)c);
        widest = std::max<int64_t>(widest, host_indptr[c + 1] - host_indptr[c]);
    }
    pile->widest = widest;
    pile->nnz = host_indptr[pile->cells];
    return MCB_OK;
}

}  // namespace

extern "C" {

int mcb_pile_create_host(const float* data, const int32_t* indices, const int32_t* indptr, int64_t cells, int64_t genes,
                         mcb_pile** out) {
    MCB_REQUIRE(out && cells >= 1 && genes >= 1, "pile: bad arguments");
    std::unique_ptr<mcb_pile> pile(new mcb_pile());
    pile->cells = cells;
    pile->genes = genes;
    MCB_TRY(finish_create(pile.get(), indptr));
    // the calling thread's own stream: piles of different host threads run side by side, and one thread's pile fills
    // the gaps the other's host-side steps leave
    MCB_TRY(thread_stream(&pile->stream));
    cudaStream_t stream = pile->stream;
    MCB_TRY(pile->own_data.alloc((size_t)pile->nnz * sizeof(float), stream));
    MCB_TRY(pile->own_indices.alloc((size_t)pile->nnz * sizeof(int32_t), stream));
    MCB_TRY(pile->own_indptr.alloc((size_t)(cells + 1) * sizeof(int32_t), stream));
    MCB_TRY(pile->status.alloc(sizeof(int), stream));
    MCB_CUDA(cudaMemsetAsync(pile->status.ptr, 0, sizeof(int), stream));
    // Upload on a stream of its own, in row chunks of about equal bytes, each followed by an event.
    MCB_CUDA(cudaStreamCreateWithFlags(&pile->copy_stream, cudaStreamNonBlocking));
    cudaEvent_t allocated;
    MCB_CUDA(cudaEventCreateWithFlags(&allocated, cudaEventDisableTiming));
    MCB_CUDA(cudaEventRecord(allocated, stream));
    MCB_CUDA(cudaStreamWaitEvent(pile->copy_stream, allocated, 0));
    cudaEventDestroy(allocated);
    MCB_CUDA(cudaMemcpyAsync(pile->own_indptr.ptr, indptr, (size_t)(cells + 1) * sizeof(int32_t), cudaMemcpyHostToDevice,
                             pile->copy_stream));
    int chunks = pile->nnz >= (8 << 20) ? 8 : 1;
    if (const char* forced = getenv("MCB_PILE_CHUNKS"))  // developer aid
        chunks = std::max(1, atoi(forced));
    pile->chunk_rows.push_back(0);
    for (int c = 1; c <= chunks; ++c) {
        int64_t row = cells;
        if (c < chunks) {  // the first row at or past this share of the non-zeros
            const int64_t want = pile->nnz / chunks * c;
            row = std::lower_bound(indptr, indptr + cells + 1, (int32_t)want) - indptr;
            row = std::min<int64_t>(std::max<int64_t>(row, pile->chunk_rows.back()), cells);
        }
        const int64_t begin = indptr[pile->chunk_rows.back()], end = indptr[row];
        if (end > begin) {
            MCB_CUDA(cudaMemcpyAsync(pile->own_data.as<float>() + begin, data + begin, (size_t)(end - begin) * sizeof(float),
                                     cudaMemcpyHostToDevice, pile->copy_stream));
            MCB_CUDA(cudaMemcpyAsync(pile->own_indices.as<int32_t>() + begin, indices + begin, (size_t)(end - begin) * sizeof(int32_t),
                                     cudaMemcpyHostToDevice, pile->copy_stream));
        }
        cudaEvent_t ready;
        MCB_CUDA(cudaEventCreateWithFlags(&ready, cudaEventDisableTiming));
        MCB_CUDA(cudaEventRecord(ready, pile->copy_stream));
        pile->chunk_ready.push_back(ready);
        pile->chunk_rows.push_back(row);
    }
    pile->data = pile->own_data.as<float>();
    pile->indices = pile->own_indices.as<int32_t>();
    pile->indptr = pile->own_indptr.as<int32_t>();
    *out = pile.release();
    return MCB_OK;
}

}  // extern "C"

namespace {

// A chunk of a compact pile (16-bit gene indices, 8-bit counts, 255 = the count is in the sorted escape list) widened to
// the float32 / int32 arrays every kernel of the path reads.
__global__ void __launch_bounds__(256) widen_compact_kernel(const uint16_t* __restrict__ indices16, const uint8_t* __restrict__ counts8,
                                                           int64_t begin, int64_t end, const int64_t* __restrict__ escape_at,
                                                           const float* __restrict__ escape_values, int64_t escapes,
                                                           float* __restrict__ data, int32_t* __restrict__ indices) {
    for (int64_t at = begin + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; at < end; at += (int64_t)gridDim.x * blockDim.x) {
        indices[at] = (int32_t)indices16[at];
        const uint8_t count = counts8[at];
        float value = (float)count;
        if (count == 255) {  // rare: a binary search for the entry's place in the escape list
            int64_t low = 0, high = escapes;
            while (low < high) {
                const int64_t middle = (low + high) >> 1;
                if (escape_at[middle] < at) low = middle + 1; else high = middle;
            }
            if (low < escapes && escape_at[low] == at) value = escape_values[low];
        }
        data[at] = value;
    }
}

}  // namespace

extern "C" {

int mcb_pile_create_host_compact(const uint16_t* indices, const uint8_t* counts, const int32_t* indptr, int64_t cells,
                                 int64_t genes, const int64_t* escape_at, const float* escape_values, int64_t escapes,
                                 mcb_pile** out) {
    MCB_REQUIRE(out && cells >= 1 && genes >= 1 && genes <= 65536, "pile: bad arguments (a compact pile has at most 65536 genes)");
    MCB_REQUIRE(escapes >= 0 && (escapes == 0 || (escape_at && escape_values)), "pile: bad escape list");
    for (int64_t e = 1; e < escapes; ++e)
        MCB_REQUIRE(escape_at[e - 1] < escape_at[e], "pile: the escape list must be sorted by position");
    std::unique_ptr<mcb_pile> pile(new mcb_pile());
    pile->cells = cells;
    pile->genes = genes;
    MCB_TRY(finish_create(pile.get(), indptr));
    MCB_TRY(thread_stream(&pile->stream));
    cudaStream_t stream = pile->stream;
    MCB_TRY(pile->own_data.alloc((size_t)pile->nnz * sizeof(float), stream));
    MCB_TRY(pile->own_indices.alloc((size_t)pile->nnz * sizeof(int32_t), stream));
    MCB_TRY(pile->own_indptr.alloc((size_t)(cells + 1) * sizeof(int32_t), stream));
    MCB_TRY(pile->compact_indices.alloc((size_t)pile->nnz * sizeof(uint16_t), stream));
    MCB_TRY(pile->compact_counts.alloc((size_t)pile->nnz, stream));
    MCB_TRY(pile->escape_at.alloc((size_t)std::max<int64_t>(escapes, 1) * sizeof(int64_t), stream));
    MCB_TRY(pile->escape_values.alloc((size_t)std::max<int64_t>(escapes, 1) * sizeof(float), stream));
    MCB_TRY(pile->status.alloc(sizeof(int), stream));
    MCB_CUDA(cudaMemsetAsync(pile->status.ptr, 0, sizeof(int), stream));
    MCB_CUDA(cudaStreamCreateWithFlags(&pile->copy_stream, cudaStreamNonBlocking));
    cudaEvent_t allocated;
    MCB_CUDA(cudaEventCreateWithFlags(&allocated, cudaEventDisableTiming));
    MCB_CUDA(cudaEventRecord(allocated, stream));
    MCB_CUDA(cudaStreamWaitEvent(pile->copy_stream, allocated, 0));
    cudaEventDestroy(allocated);
    cudaStream_t copy = pile->copy_stream;
    MCB_CUDA(cudaMemcpyAsync(pile->own_indptr.ptr, indptr, (size_t)(cells + 1) * sizeof(int32_t), cudaMemcpyHostToDevice, copy));
    if (escapes > 0) {
        MCB_CUDA(cudaMemcpyAsync(pile->escape_at.ptr, escape_at, (size_t)escapes * sizeof(int64_t), cudaMemcpyHostToDevice, copy));
        MCB_CUDA(cudaMemcpyAsync(pile->escape_values.ptr, escape_values, (size_t)escapes * sizeof(float), cudaMemcpyHostToDevice, copy));
    }
    int chunks = pile->nnz >= (8 << 20) ? 8 : 1;
    if (const char* forced = getenv("MCB_PILE_CHUNKS"))  // developer aid
        chunks = std::max(1, atoi(forced));
    pile->chunk_rows.push_back(0);
    for (int c = 1; c <= chunks; ++c) {
        int64_t row = cells;
        if (c < chunks) {
            const int64_t want = pile->nnz / chunks * c;
            row = std::lower_bound(indptr, indptr + cells + 1, (int32_t)want) - indptr;
            row = std::min<int64_t>(std::max<int64_t>(row, pile->chunk_rows.back()), cells);
        }
        const int64_t begin = indptr[pile->chunk_rows.back()], end = indptr[row];
        if (end > begin) {
            MCB_CUDA(cudaMemcpyAsync(pile->compact_indices.as<uint16_t>() + begin, indices + begin,
                                     (size_t)(end - begin) * sizeof(uint16_t), cudaMemcpyHostToDevice, copy));
            MCB_CUDA(cudaMemcpyAsync(pile->compact_counts.as<uint8_t>() + begin, counts + begin, (size_t)(end - begin),
                                     cudaMemcpyHostToDevice, copy));
            const unsigned blocks = (unsigned)std::min<int64_t>(ceil_div(end - begin, 256 * 8), 148 * 8);
            MCB_PROFILED("widen_compact", copy,
                         widen_compact_kernel<<<blocks, 256, 0, copy>>>(
                             pile->compact_indices.as<uint16_t>(), pile->compact_counts.as<uint8_t>(), begin, end,
                             pile->escape_at.as<int64_t>(), pile->escape_values.as<float>(), escapes,
                             pile->own_data.as<float>(), pile->own_indices.as<int32_t>()));
            MCB_LAUNCH_CHECK();
        }
        cudaEvent_t ready;
        MCB_CUDA(cudaEventCreateWithFlags(&ready, cudaEventDisableTiming));
        MCB_CUDA(cudaEventRecord(ready, copy));
        pile->chunk_ready.push_back(ready);
        pile->chunk_rows.push_back(row);
    }
    pile->data = pile->own_data.as<float>();
    pile->indices = pile->own_indices.as<int32_t>();
    pile->indptr = pile->own_indptr.as<int32_t>();
    *out = pile.release();
    return MCB_OK;
}

int mcb_host_pin(const void* buffer, uint64_t bytes) {
    MCB_REQUIRE(buffer != nullptr || bytes == 0, "host_pin: null buffer");
    if (bytes > 0)
        MCB_CUDA(cudaHostRegister(const_cast<void*>(buffer), (size_t)bytes, cudaHostRegisterDefault));
    return MCB_OK;
}

int mcb_host_unpin(const void* buffer) {
    if (buffer != nullptr)
        MCB_CUDA(cudaHostUnregister(const_cast<void*>(buffer)));
    return MCB_OK;
}

int mcb_pile_set_max_cell_total(mcb_pile* pile, uint64_t max_total) {
    MCB_REQUIRE(pile != nullptr, "pile: null handle");
    pile->max_row_total = max_total;
    return MCB_OK;
}

int mcb_pile_create_dev(const float* data, const int32_t* indices, const int32_t* indptr, int64_t cells, int64_t genes,
                        void* stream, mcb_pile** out) {
    MCB_REQUIRE(out && cells >= 1 && genes >= 1, "pile: bad arguments");
    std::unique_ptr<mcb_pile> pile(new mcb_pile());
    pile->cells = cells;
    pile->genes = genes;
    pile->stream = (cudaStream_t)stream;
    std::vector<int32_t> host_indptr((size_t)cells + 1);
    MCB_TRY(copy_to_host(host_indptr.data(), indptr, (size_t)(cells + 1) * sizeof(int32_t), pile->stream));
    MCB_CUDA(cudaStreamSynchronize(pile->stream));
    MCB_TRY(finish_create(pile.get(), host_indptr.data()));
    MCB_TRY(pile->status.alloc(sizeof(int), pile->stream));  // deferred errors of the downsampling (a row above the hint)
    MCB_CUDA(cudaMemsetAsync(pile->status.ptr, 0, sizeof(int), pile->stream));
    pile->data = data;
    pile->indices = indices;
    pile->indptr = indptr;
    *out = pile.release();
    return MCB_OK;
}

int mcb_pile_downsample(mcb_pile* pile, int32_t samples, uint64_t random_seed) {
    MCB_REQUIRE(pile != nullptr, "pile: null handle");
    MCB_REQUIRE(samples >= 0, "pile: negative samples");
    if (!pile->downsampled.ptr)
        MCB_TRY(pile->downsampled.alloc((size_t)pile->nnz * sizeof(float), pile->stream));
    if (!pile->gene_totals.ptr)
        MCB_TRY(pile->gene_totals.alloc((size_t)pile->genes * sizeof(float), pile->stream));
    MCB_CUDA(cudaMemsetAsync(pile->gene_totals.ptr, 0, (size_t)pile->genes * sizeof(float), pile->stream));
    float* totals = pile->gene_totals.as<float>();
    unsigned long long* stats = nullptr;
    if (pile->want_gene_stats) {
        // a downsampled cell holds at most `samples` UMIs of a gene: the packed sum of squares keeps 48 bits
        MCB_REQUIRE((long double)pile->cells * samples * samples < 281474976710656.0L,
                    "pile: gene statistics of %lld cells downsampled to %d UMIs would overflow 48 bits", (long long)pile->cells, samples);
        if (!pile->gene_stats.ptr)
            MCB_TRY(pile->gene_stats.alloc((size_t)pile->genes * sizeof(unsigned long long), pile->stream));
        MCB_CUDA(cudaMemsetAsync(pile->gene_stats.ptr, 0, (size_t)pile->genes * sizeof(unsigned long long), pile->stream));
        stats = pile->gene_stats.as<unsigned long long>();
    }
    const float top_threshold = pile->gene_top_threshold;
    if (pile->chunk_ready.empty()) {
        MCB_TRY(downsample_rows_launch(pile->data, pile->downsampled.as<float>(), pile->indptr, nullptr, samples, 0, pile->cells, 0,
                                       pile->widest, random_seed, 1, pile->max_row_total,
                                       pile->max_row_total ? pile->status.as<int>() : nullptr, pile->stream, pile->indices, totals,
                                       stats, top_threshold));
    } else {
        // Every chunk as soon as it has arrived; with the largest cell total known nothing here waits on the device.
        // Chunks that have arrived already (a pile uploaded under the previous pile's kernels) go in one launch: a
        // launch over an eighth of the cells fills a third of the GPU.
        const size_t chunks = pile->chunk_ready.size();
        for (size_t c = 0; c < chunks;) {
            size_t last = c;
            while (last + 1 < chunks && cudaEventQuery(pile->chunk_ready[last + 1]) == cudaSuccess) ++last;
            MCB_CUDA(cudaStreamWaitEvent(pile->stream, pile->chunk_ready[last], 0));
            MCB_TRY(downsample_rows_launch(pile->data, pile->downsampled.as<float>(), pile->indptr, nullptr, samples,
                                           pile->chunk_rows[c], pile->chunk_rows[last + 1], 0, pile->widest, random_seed, 1,
                                           pile->max_row_total, pile->max_row_total ? pile->status.as<int>() : nullptr,
                                           pile->stream, pile->indices, totals, stats, top_threshold));
            c = last + 1;
        }
    }
    pile->has_downsampled = true;
    return MCB_OK;
}

int mcb_pile_gene_totals(mcb_pile* pile, float* totals) {
    MCB_REQUIRE(pile != nullptr && totals != nullptr, "pile: null argument");
    MCB_TRY(wait_for_upload(pile));
    Scratch device;
    const float* source = pile->gene_totals.as<float>();  // of the downsampled pile: the downsample kernels added them up
    if (!pile->has_downsampled) {
        MCB_TRY(device.alloc((size_t)pile->genes * sizeof(float), pile->stream));
        MCB_TRY(column_totals_launch(pile->data, pile->indices, pile->nnz, pile->genes, device.as<float>(), pile->stream));
        source = device.as<float>();
    }
    MCB_TRY(copy_to_host(totals, source, (size_t)pile->genes * sizeof(float), pile->stream));
    MCB_CUDA(cudaStreamSynchronize(pile->stream));
    return MCB_OK;
}

int mcb_pile_read_downsampled(mcb_pile* pile, float* data) {
    MCB_REQUIRE(pile != nullptr && pile->has_downsampled, "pile: not downsampled yet");
    MCB_TRY(check_status(pile));
    MCB_TRY(copy_to_host(data, pile->downsampled.ptr, (size_t)pile->nnz * sizeof(float), pile->stream));
    MCB_CUDA(cudaStreamSynchronize(pile->stream));
    return MCB_OK;
}

}  // extern "C" (reopened below)

namespace {

__global__ void gene_slots_kernel(const int32_t* __restrict__ feature_genes, int64_t features, int64_t genes,
                                  int32_t* __restrict__ slot_of_gene, int* __restrict__ bad) {
    const int64_t f = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (f >= features) return;
    const int32_t gene = feature_genes[f];
    if (gene < 0 || gene >= genes) {
        atomicOr(bad, 1);
        return;
    }
    if (atomicCAS(&slot_of_gene[gene], -1, (int32_t)f) != -1) atomicOr(bad, 2);
}

// direct.py:242-249: the dense [cells, features] log-fraction operand of the feature genes, left in `d_dense`.
// feature_genes: host pointer, or nullptr to use the list mcb_pile_top_genes left on the device (*features is then set).
// d_flag (2 device ints) receives: [0] a value the table does not cover, [1] a bad feature gene list.
int pile_log_fraction(mcb_pile* pile, const int32_t* feature_genes, int64_t* features_io, const float* log_table,
                      int64_t table_size, float regularization, Scratch& d_dense, Scratch& d_flag) {
    int64_t features = *features_io;
    if (feature_genes == nullptr) {
        MCB_REQUIRE(pile->top_genes_count > 0, "pile: no feature genes given and mcb_pile_top_genes was not called");
        features = pile->top_genes_count;
    }
    MCB_REQUIRE(features >= 1 && features <= pile->genes, "pile: bad feature genes count %lld", (long long)features);
    *features_io = features;
    cudaStream_t stream = pile->stream;
    MCB_TRY(wait_for_upload(pile));
    Scratch d_genes, d_slot;
    const int32_t* genes_device = pile->top_genes.as<int32_t>();
    if (feature_genes != nullptr) {
        MCB_TRY(d_genes.alloc((size_t)features * sizeof(int32_t), stream));
        MCB_TRY(upload_small(d_genes.ptr, feature_genes, (size_t)features * sizeof(int32_t), stream));
        genes_device = d_genes.as<int32_t>();
    }
    MCB_TRY(d_slot.alloc((size_t)pile->genes * sizeof(int32_t), stream));
    MCB_TRY(d_dense.alloc((size_t)pile->cells * features * sizeof(float), stream));
    MCB_TRY(d_flag.alloc(2 * sizeof(int), stream));
    MCB_CUDA(cudaMemsetAsync(d_slot.ptr, 0xff, (size_t)pile->genes * sizeof(int32_t), stream));
    // The log table of the calling thread's previous pile is still on the device when it is the same table (a worker
    // walks piles with one table): no host-to-device copy here, where it would queue behind the next pile's upload.
    const float* table_device = nullptr;
    {
        struct TableCache {
            std::vector<float> host;
            float* device = nullptr;
            int device_index = -1;
        };
        static thread_local TableCache cache;
        int device_index = 0;
        MCB_CUDA(cudaGetDevice(&device_index));
        const bool same = getenv("MCB_NO_TABLE_CACHE") == nullptr && cache.device != nullptr && cache.device_index == device_index &&
                          (int64_t)cache.host.size() == table_size &&
                          memcmp(cache.host.data(), log_table, (size_t)table_size * sizeof(float)) == 0;
        if (!same) {
            MCB_CUDA(cudaStreamSynchronize(stream));  // nothing of this thread still reads the old table
            if (cache.device != nullptr)
                cudaFree(cache.device);
            cache.device = nullptr;
            MCB_CUDA(cudaMalloc(&cache.device, (size_t)table_size * sizeof(float)));
            cache.host.assign(log_table, log_table + table_size);
            cache.device_index = device_index;
            MCB_TRY(upload_small(cache.device, cache.host.data(), (size_t)table_size * sizeof(float), stream));
        }
        table_device = cache.device;
    }
    MCB_CUDA(cudaMemsetAsync(d_flag.ptr, 0, 2 * sizeof(int), stream));
    gene_slots_kernel<<<(unsigned)ceil_div(features, 256), 256, 0, stream>>>(genes_device, features, pile->genes,
                                                                             d_slot.as<int32_t>(), d_flag.as<int>() + 1);
    MCB_LAUNCH_CHECK();
    const float* values = pile->has_downsampled ? pile->downsampled.as<float>() : pile->data;
    MCB_TRY(log_fraction_dense_launch(values, pile->indices, pile->indptr, pile->cells, d_slot.as<int32_t>(), features,
                                      table_device, table_size, regularization, d_dense.as<float>(), d_flag.as<int>(), stream));
    return MCB_OK;
}

// Reads the two flags of pile_log_fraction (synchronises the stream) and turns them into errors.  A value outside
// the table (not a non-negative integer below table_size) is an error: only the table, made by numpy's own log2 on the
// host, is bit-exact with direct.py:245-249; *outside_table (nullable) still reports it.
int check_log_fraction_flags(mcb_pile* pile, Scratch& d_flag, int* outside_table) {
    int flags[2] = {0, 0};
    MCB_TRY(copy_to_host(flags, d_flag.ptr, sizeof(flags), pile->stream));
    MCB_TRY(check_status(pile));  // synchronises the stream
    MCB_REQUIRE((flags[1] & 1) == 0, "pile: a feature gene is out of range");
    MCB_REQUIRE((flags[1] & 2) == 0, "pile: a feature gene is listed twice");
    if (outside_table)
        *outside_table = flags[0];
    MCB_REQUIRE(flags[0] == 0, "pile: a value of a feature gene is not an integer count below the log table's size");
    return MCB_OK;
}

// min_key (optional, device uint32 preset to 0xffffffff) receives the minimum of the similarity.
int pile_similarity(mcb_pile* pile, const int32_t* feature_genes, int64_t features, const float* log_table, int64_t table_size,
                    float regularization, int absolute, float* similarity_dev, int* outside_table, unsigned int* min_key) {
    MCB_REQUIRE(pile != nullptr && log_table && similarity_dev, "pile: null argument");
    cudaStream_t stream = pile->stream;
    trace_lap("similarity: enter", stream);
    Scratch d_dense, d_flag;
    MCB_TRY(pile_log_fraction(pile, feature_genes, &features, log_table, table_size, regularization, d_dense, d_flag));
    trace_lap("similarity: log_fraction", stream);
    MCB_TRY(correlate_dense_with_min(d_dense.as<float>(), pile->cells, features, features, similarity_dev, pile->cells, absolute,
                                     min_key, stream));
    trace_lap("similarity: correlate", stream);
    return check_log_fraction_flags(pile, d_flag, outside_table);
}

}  // namespace

extern "C" {

int mcb_pile_similarity(mcb_pile* pile, const int32_t* feature_genes, int64_t features, const float* log_table,
                        int64_t table_size, float regularization, int absolute, float* similarity_dev, int* outside_table) {
    return pile_similarity(pile, feature_genes, features, log_table, table_size, regularization, absolute, similarity_dev,
                           outside_table, nullptr);
}

int mcb_pile_log_fraction(mcb_pile* pile, const int32_t* feature_genes, int64_t features, const float* log_table,
                          int64_t table_size, float regularization, float* dense_host, int* outside_table) {
    MCB_REQUIRE(pile != nullptr && log_table && dense_host, "pile: null argument");
    Scratch d_dense, d_flag;
    MCB_TRY(pile_log_fraction(pile, feature_genes, &features, log_table, table_size, regularization, d_dense, d_flag));
    MCB_TRY(check_log_fraction_flags(pile, d_flag, outside_table));
    MCB_TRY(copy_to_host(dense_host, d_dense.ptr, (size_t)pile->cells * features * sizeof(float), pile->stream));
    MCB_CUDA(cudaStreamSynchronize(pile->stream));
    return MCB_OK;
}

int mcb_pile_top_genes(mcb_pile* pile, int64_t count, int32_t* genes_host) {
    MCB_REQUIRE(pile != nullptr && count >= 1 && count < pile->genes, "pile: top genes count must be in 1..genes-1");
    cudaStream_t stream = pile->stream;
    MCB_TRY(wait_for_upload(pile));
    Scratch computed, unused;
    const float* totals = pile->gene_totals.as<float>();  // of the downsampled pile: the downsample kernels added them up
    if (!pile->has_downsampled) {
        MCB_TRY(computed.alloc((size_t)pile->genes * sizeof(float), stream));
        MCB_TRY(column_totals_launch(pile->data, pile->indices, pile->nnz, pile->genes, computed.as<float>(), stream));
        totals = computed.as<float>();
    }
    MCB_TRY(unused.alloc((size_t)count * sizeof(float), stream));
    if (pile->top_genes_count != count) {
        Scratch fresh;
        MCB_TRY(fresh.alloc((size_t)count * sizeof(int32_t), stream));
        std::swap(pile->top_genes.ptr, fresh.ptr);
        std::swap(pile->top_genes.stream, fresh.stream);
        pile->top_genes_count = count;
    }
    // the `count` largest totals, equal totals by lower gene index, indices ascending: one row of collect_top
    MCB_TRY(collect_top_launch(count, totals, 1, pile->genes, pile->genes, pile->top_genes.as<int32_t>(),
                               unused.as<float>(), 0, stream));
    if (genes_host != nullptr) {
        MCB_TRY(copy_to_host(genes_host, pile->top_genes.ptr, (size_t)count * sizeof(int32_t), stream));
        MCB_CUDA(cudaStreamSynchronize(stream));
    }
    return MCB_OK;
}

int mcb_pile_knn_graph(mcb_pile* pile, const int32_t* feature_genes, int64_t features, const float* log_table,
                       int64_t table_size, float regularization, int absolute, int64_t k, double balanced_ranks_factor,
                       double incoming_degree_factor, double outgoing_degree_factor, int64_t min_outgoing_degree,
                       int* outside_table, mcb_graph** out) {
    MCB_REQUIRE(pile != nullptr && out != nullptr, "pile: null argument");
    // The similarity never leaves this call: its minimum comes out of the correlate epilogue and the ranks
    // overwrite it in place (the public mcb_dev_knn_graph_f32 must copy and scan its input instead).
    Scratch similarity, min_key;
    MCB_TRY(similarity.alloc((size_t)pile->cells * pile->cells * sizeof(float), pile->stream));
    MCB_TRY(min_key.alloc(sizeof(unsigned int), pile->stream));
    MCB_CUDA(cudaMemsetAsync(min_key.ptr, 0xff, sizeof(unsigned int), pile->stream));
    MCB_TRY(pile_similarity(pile, feature_genes, features, log_table, table_size, regularization, absolute,
                            similarity.as<float>(), outside_table, min_key.as<unsigned int>()));
    return knn_graph_consuming(similarity.as<float>(), min_key.as<uint32_t>(), pile->cells, k, balanced_ranks_factor,
                               incoming_degree_factor, outgoing_degree_factor, min_outgoing_degree, pile->stream, out);
}

}  // extern "C" (reopened below)

namespace {

// One CTA per pile row: the row's values and gene indices are copied from the atlas to the pile's own arrays.
__global__ void __launch_bounds__(256) gather_rows_kernel(const float* __restrict__ data, const int32_t* __restrict__ indices,
                                                         const int64_t* __restrict__ atlas_indptr, const int64_t* __restrict__ rows,
                                                         const int32_t* __restrict__ pile_indptr, int64_t count,
                                                         float* __restrict__ out_data, int32_t* __restrict__ out_indices) {
    for (int64_t r = blockIdx.x; r < count; r += gridDim.x) {
        const int64_t source = atlas_indptr[rows[r]];
        const int32_t target = pile_indptr[r], size = pile_indptr[r + 1] - target;
        for (int e = threadIdx.x; e < size; e += blockDim.x) {
            out_data[target + e] = data[source + e];
            out_indices[target + e] = indices[source + e];
        }
    }
}

// Row sums of a CSR matrix with 64-bit row pointers, one warp per row (UMI counts: exact in float32 in any order).
__global__ void __launch_bounds__(256) atlas_row_totals_kernel(const float* __restrict__ data, const int64_t* __restrict__ indptr,
                                                              int64_t rows, float* __restrict__ totals) {
    const int lane = threadIdx.x & 31;
    const int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (row >= rows) return;
    float sum = 0.0f;
    for (int64_t e = indptr[row] + lane; e < indptr[row + 1]; e += 32) sum += data[e];
    for (int offset = 16; offset > 0; offset >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, offset);
    if (lane == 0) totals[row] = sum;
}

// sums[group_of_row[r], gene] += value for every element of the rows with group_of_row[r] >= 0: the per-metacell UMI
// totals of collect_metacells (pipeline/collect.py:162-201, `ut.sum_groups`), dense float32 [groups, genes] (integers
// far below 2^24: exact in any order).
__global__ void __launch_bounds__(256) atlas_sum_groups_kernel(const float* __restrict__ data, const int32_t* __restrict__ indices,
                                                              const int64_t* __restrict__ indptr, int64_t rows,
                                                              const int32_t* __restrict__ group_of_row, int64_t genes,
                                                              float* __restrict__ sums) {
    const int lane = threadIdx.x & 31;
    const int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (row >= rows) return;
    const int32_t group = group_of_row[row];
    if (group < 0) return;
    float* target = sums + (int64_t)group * genes;
    for (int64_t e = indptr[row] + lane; e < indptr[row + 1]; e += 32) atomicAdd(&target[indices[e]], data[e]);
}

// Non-zeros of every row of a dense matrix, one warp per row.
__global__ void __launch_bounds__(256) dense_row_counts_kernel(const float* __restrict__ dense, int64_t rows, int64_t columns,
                                                              int32_t* __restrict__ counts) {
    const int lane = threadIdx.x & 31;
    const int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (row >= rows) return;
    int count = 0;
    for (int64_t c = lane; c < columns; c += 32) count += dense[row * columns + c] != 0.0f ? 1 : 0;
    count = warp_reduce_sum(count);
    if (lane == 0) counts[row] = count;
}

// The non-zeros of every dense row in column order (scipy's csr_matrix(dense)), one warp per row.
__global__ void __launch_bounds__(256) dense_to_csr_kernel(const float* __restrict__ dense, int64_t rows, int64_t columns,
                                                          const int32_t* __restrict__ indptr, float* __restrict__ data,
                                                          int32_t* __restrict__ indices) {
    const int lane = threadIdx.x & 31;
    const int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (row >= rows) return;
    int at = indptr[row];
    for (int64_t base = 0; base < columns; base += 32) {
        const int64_t c = base + lane;
        const float value = c < columns ? dense[row * columns + c] : 0.0f;
        const unsigned kept = __ballot_sync(0xffffffffu, value != 0.0f);
        if (value != 0.0f) {
            const int slot = at + __popc(kept & ((1u << lane) - 1u));
            data[slot] = value;
            indices[slot] = (int32_t)c;
        }
        at += __popc(kept);
    }
}

// values[g * rows + row] = the (downsampled) count of gene genes[g] in cell `row`, 0 when the cell does not hold it: one
// thread per (gene, cell), a binary search in the cell's sorted gene indices.  The host drops the zeros: what is left is the
// column of the CSC matrix without explicit zeros (computation.py:892-899 eliminates them), in cell order.
__global__ void __launch_bounds__(256) gene_columns_kernel(const float* __restrict__ data, const int32_t* __restrict__ indices,
                                                          const int32_t* __restrict__ indptr, int64_t rows,
                                                          const int32_t* __restrict__ genes, float* __restrict__ values) {
    const int64_t row = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (row >= rows) return;
    const int32_t gene = genes[blockIdx.y];
    const int last = indptr[row + 1];
    int low = indptr[row], high = last;  // the first element with index >= gene
    while (low < high) {
        const int middle = (low + high) >> 1;
        if (indices[middle] < gene)
            low = middle + 1;
        else
            high = middle;
    }
    values[(int64_t)blockIdx.y * rows + row] = (low < last && indices[low] == gene) ? data[low] : 0.0f;
}

// out[i] = the median of values[i .. i + window) for an odd window: the element with at most window / 2 smaller and more
// than window / 2 not larger elements.  One warp per output, the window in registers (window <= 128).
__global__ void __launch_bounds__(256) sliding_median_kernel(const float* __restrict__ values, int64_t outputs, int window,
                                                            float* __restrict__ out) {
    const int lane = threadIdx.x & 31;
    const int64_t i = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (i >= outputs) return;
    float mine[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) mine[u] = lane + 32 * u < window ? values[i + lane + 32 * u] : 0.0f;
    // pandas' rolling functions take infinities for missing values (window._prep_values) and a window with a missing value
    // has no median: NaN, as the reference gets (a gene that is constant over the cells has log2(0) = -inf)
    bool finite = true;
#pragma unroll
    for (int u = 0; u < 4; ++u) finite = finite && (lane + 32 * u >= window || isfinite(mine[u]));
    if (!__all_sync(0xffffffffu, finite)) {
        if (lane == 0) out[i] = __int_as_float(0x7fc00000);
        return;
    }
    const int half = window / 2;
    float median = 0.0f;
    for (int j = 0; j < window; ++j) {
        const float candidate = __shfl_sync(0xffffffffu, mine[j >> 5], j & 31);
        int smaller = 0, not_larger = 0;
#pragma unroll
        for (int u = 0; u < 4; ++u)
            if (lane + 32 * u < window) {
                smaller += mine[u] < candidate ? 1 : 0;
                not_larger += mine[u] <= candidate ? 1 : 0;
            }
        smaller = warp_reduce_sum(smaller);
        not_larger = warp_reduce_sum(not_larger);
        if (smaller <= half && half < not_larger) {  // the same for every lane: the counts are warp sums
            median = candidate;
            break;
        }
    }
    if (lane == 0) out[i] = median;
}

// The same medians for sliding_window_function(values, order_by=...) (utilities/computation.py:1836-1897) without the
// host's gathers: `order` sorts the entries by `order_by`; the window of the entry at sorted place i covers the places
// i - half .. i + half, the windows at both ends padded with the places the reference pads them with (:1877-1883); the
// median goes to the entry's own index, as float64 (what `values - medians` needs).
__global__ void __launch_bounds__(256) sliding_median_ordered_kernel(const float* __restrict__ values, const int64_t* __restrict__ order,
                                                                    int64_t count, int window, double* __restrict__ out) {
    const int lane = threadIdx.x & 31;
    const int64_t i = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (i >= count) return;
    const int half = window / 2;
    float mine[4];
    bool finite = true;
#pragma unroll
    for (int u = 0; u < 4; ++u) {
        const int64_t padded = i + lane + 32 * u;  // place in [before | order | after]
        mine[u] = 0.0f;
        if (lane + 32 * u < window) {
            const int64_t place = padded < half ? window - half + padded
                                  : padded < half + count ? padded - half : count - window + (padded - half - count);
            mine[u] = values[order[place]];
            finite = finite && isfinite(mine[u]);
        }
    }
    const int64_t target = order[i];
    if (!__all_sync(0xffffffffu, finite)) {
        if (lane == 0) out[target] = __longlong_as_double(0x7ff8000000000000ll);
        return;
    }
    float median = 0.0f;
    for (int j = 0; j < window; ++j) {
        const float candidate = __shfl_sync(0xffffffffu, mine[j >> 5], j & 31);
        int smaller = 0, not_larger = 0;
#pragma unroll
        for (int u = 0; u < 4; ++u)
            if (lane + 32 * u < window) {
                smaller += mine[u] < candidate ? 1 : 0;
                not_larger += mine[u] <= candidate ? 1 : 0;
            }
        smaller = warp_reduce_sum(smaller);
        not_larger = warp_reduce_sum(not_larger);
        if (smaller <= half && half < not_larger) {
            median = candidate;
            break;
        }
    }
    if (lane == 0) out[target] = (double)median;
}

int check_atlas_indptr(const int64_t* indptr, int64_t cells) {
    MCB_REQUIRE(indptr[0] == 0, "atlas: indptr[0] must be 0");
    for (int64_t c = 0; c < cells; ++c)
        MCB_REQUIRE(indptr[c + 1] >= indptr[c], "atlas: indptr is not monotone at cell %lld", (long long)c);
    return MCB_OK;
}

}  // namespace

extern "C" {

int mcb_atlas_create_dev(const float* data, const int32_t* indices, const int64_t* indptr_device, const int64_t* indptr_host,
                         int64_t cells, int64_t genes, mcb_atlas** out) {
    MCB_REQUIRE(out && data && indices && indptr_device && indptr_host && cells >= 1 && genes >= 1, "atlas: bad arguments");
    MCB_TRY(check_atlas_indptr(indptr_host, cells));
    std::unique_ptr<mcb_atlas> atlas(new mcb_atlas());
    atlas->cells = cells;
    atlas->genes = genes;
    atlas->nnz = indptr_host[cells];
    atlas->data = data;
    atlas->indices = indices;
    atlas->indptr = indptr_device;
    atlas->host_indptr.assign(indptr_host, indptr_host + cells + 1);
    MCB_CUDA(cudaGetDevice(&atlas->device));
    *out = atlas.release();
    return MCB_OK;
}

int mcb_atlas_create_host(const float* data, const int32_t* indices, const int64_t* indptr, int64_t cells, int64_t genes,
                          mcb_atlas** out) {
    MCB_REQUIRE(out && data && indices && indptr && cells >= 1 && genes >= 1, "atlas: bad arguments");
    MCB_TRY(check_atlas_indptr(indptr, cells));
    std::unique_ptr<mcb_atlas> atlas(new mcb_atlas());
    atlas->cells = cells;
    atlas->genes = genes;
    atlas->nnz = indptr[cells];
    cudaStream_t stream = nullptr;
    MCB_TRY(thread_stream(&stream));
    MCB_TRY(atlas->own_data.alloc((size_t)atlas->nnz * sizeof(float), stream));
    MCB_TRY(atlas->own_indices.alloc((size_t)atlas->nnz * sizeof(int32_t), stream));
    MCB_TRY(atlas->own_indptr.alloc((size_t)(cells + 1) * sizeof(int64_t), stream));
    MCB_CUDA(cudaMemcpyAsync(atlas->own_indptr.ptr, indptr, (size_t)(cells + 1) * sizeof(int64_t), cudaMemcpyHostToDevice, stream));
    // one cudaMemcpyAsync per 256 MiB block: page-locked sources stream at the link's speed, pageable ones are staged
    const int64_t block = 64ll << 20;
    for (int64_t start = 0; start < atlas->nnz; start += block) {
        const size_t count = (size_t)std::min<int64_t>(block, atlas->nnz - start);
        MCB_CUDA(cudaMemcpyAsync(atlas->own_data.as<float>() + start, data + start, count * sizeof(float), cudaMemcpyHostToDevice, stream));
        MCB_CUDA(cudaMemcpyAsync(atlas->own_indices.as<int32_t>() + start, indices + start, count * sizeof(int32_t),
                                 cudaMemcpyHostToDevice, stream));
    }
    MCB_CUDA(cudaStreamSynchronize(stream));
    atlas->data = atlas->own_data.as<float>();
    atlas->indices = atlas->own_indices.as<int32_t>();
    atlas->indptr = atlas->own_indptr.as<int64_t>();
    atlas->host_indptr.assign(indptr, indptr + cells + 1);
    MCB_CUDA(cudaGetDevice(&atlas->device));
    *out = atlas.release();
    return MCB_OK;
}

int64_t mcb_atlas_cells(const mcb_atlas* atlas) { return atlas ? atlas->cells : 0; }
int64_t mcb_atlas_nnz(const mcb_atlas* atlas) { return atlas ? atlas->nnz : 0; }

int mcb_atlas_cell_totals(const mcb_atlas* atlas, float* totals_host) {
    MCB_REQUIRE(atlas != nullptr && totals_host != nullptr, "atlas: null argument");
    cudaStream_t stream = nullptr;
    MCB_TRY(thread_stream(&stream));
    Scratch totals;
    MCB_TRY(totals.alloc((size_t)atlas->cells * sizeof(float), stream));
    MCB_PROFILED("atlas_row_totals", stream,
                 atlas_row_totals_kernel<<<(unsigned)ceil_div(atlas->cells, 8), 256, 0, stream>>>(atlas->data, atlas->indptr, atlas->cells,
                                                                                               totals.as<float>()));
    MCB_LAUNCH_CHECK();
    MCB_TRY(copy_to_host(totals_host, totals.ptr, (size_t)atlas->cells * sizeof(float), stream));
    MCB_CUDA(cudaStreamSynchronize(stream));
    return MCB_OK;
}

int mcb_atlas_sum_groups(const mcb_atlas* atlas, const int32_t* group_of_cells, int64_t groups, float* sums_device, void* stream_) {
    MCB_REQUIRE(atlas != nullptr && group_of_cells != nullptr && sums_device != nullptr && groups >= 1, "atlas: bad arguments");
    cudaStream_t stream = (cudaStream_t)stream_;
    for (int64_t c = 0; c < atlas->cells; ++c)
        MCB_REQUIRE(group_of_cells[c] < groups, "atlas: group %d of cell %lld is out of range", group_of_cells[c], (long long)c);
    Scratch d_groups;
    MCB_TRY(d_groups.alloc((size_t)atlas->cells * sizeof(int32_t), stream));
    MCB_CUDA(cudaMemcpyAsync(d_groups.ptr, group_of_cells, (size_t)atlas->cells * sizeof(int32_t), cudaMemcpyHostToDevice, stream));
    MCB_CUDA(cudaMemsetAsync(sums_device, 0, (size_t)groups * atlas->genes * sizeof(float), stream));
    MCB_PROFILED("atlas_sum_groups", stream,
                 atlas_sum_groups_kernel<<<(unsigned)ceil_div(atlas->cells, 8), 256, 0, stream>>>(
                     atlas->data, atlas->indices, atlas->indptr, atlas->cells, d_groups.as<int32_t>(), atlas->genes, sums_device));
    MCB_LAUNCH_CHECK();
    MCB_CUDA(cudaStreamSynchronize(stream));  // d_groups' host source may be pageable
    return MCB_OK;
}

void mcb_atlas_free(mcb_atlas* atlas) { delete atlas; }

int mcb_pile_create_from_atlas(const mcb_atlas* atlas, const int64_t* rows, int64_t count, mcb_pile** out) {
    MCB_REQUIRE(atlas != nullptr && rows != nullptr && out != nullptr && count >= 1, "pile: bad arguments");
    std::unique_ptr<mcb_pile> pile(new mcb_pile());
    pile->cells = count;
    pile->genes = atlas->genes;
    // the pile's row pointers, from the atlas' row pointers on the host
    std::vector<int32_t> indptr((size_t)count + 1);
    int64_t total = 0, widest = 0;
    indptr[0] = 0;
    for (int64_t r = 0; r < count; ++r) {
        MCB_REQUIRE(rows[r] >= 0 && rows[r] < atlas->cells, "pile: atlas row %lld is out of range", (long long)rows[r]);
        const int64_t size = atlas->host_indptr[rows[r] + 1] - atlas->host_indptr[rows[r]];
        total += size;
        widest = std::max(widest, size);
        MCB_REQUIRE(total < (1ll << 31), "pile: more than 2^31 - 1 non-zeros");
        indptr[r + 1] = (int32_t)total;
    }
    pile->nnz = total;
    pile->widest = widest;
    MCB_TRY(thread_stream(&pile->stream));
    cudaStream_t stream = pile->stream;
    Scratch d_rows;
    MCB_TRY(pile->own_data.alloc((size_t)total * sizeof(float), stream));
    MCB_TRY(pile->own_indices.alloc((size_t)total * sizeof(int32_t), stream));
    MCB_TRY(pile->own_indptr.alloc((size_t)(count + 1) * sizeof(int32_t), stream));
    MCB_TRY(pile->status.alloc(sizeof(int), stream));
    MCB_TRY(d_rows.alloc((size_t)count * sizeof(int64_t), stream));
    MCB_CUDA(cudaMemsetAsync(pile->status.ptr, 0, sizeof(int), stream));
    // the two small lists go through one page-locked staging block of the pile (the sources are the caller's / a local)
    MCB_CUDA(cudaMemcpyAsync(d_rows.ptr, rows, (size_t)count * sizeof(int64_t), cudaMemcpyHostToDevice, stream));
    MCB_CUDA(cudaMemcpyAsync(pile->own_indptr.ptr, indptr.data(), (size_t)(count + 1) * sizeof(int32_t), cudaMemcpyHostToDevice, stream));
    if (total > 0) {
        MCB_PROFILED("gather_rows", stream,
                     gather_rows_kernel<<<(unsigned)std::min<int64_t>(count, 1 << 20), 256, 0, stream>>>(
                         atlas->data, atlas->indices, atlas->indptr, d_rows.as<int64_t>(), pile->own_indptr.as<int32_t>(), count,
                         pile->own_data.as<float>(), pile->own_indices.as<int32_t>()));
        MCB_LAUNCH_CHECK();
    }
    MCB_CUDA(cudaStreamSynchronize(stream));  // `rows` and the local row pointers may be pageable: they have been read now
    pile->data = pile->own_data.as<float>();
    pile->indices = pile->own_indices.as<int32_t>();
    pile->indptr = pile->own_indptr.as<int32_t>();
    *out = pile.release();
    return MCB_OK;
}

int mcb_pile_create_dense_dev(const float* dense, int64_t cells, int64_t genes, mcb_pile** out) {
    MCB_REQUIRE(dense != nullptr && out != nullptr && cells >= 1 && genes >= 1, "pile: bad arguments");
    std::unique_ptr<mcb_pile> pile(new mcb_pile());
    pile->cells = cells;
    pile->genes = genes;
    MCB_TRY(thread_stream(&pile->stream));
    cudaStream_t stream = pile->stream;
    Scratch counts;
    MCB_TRY(counts.alloc((size_t)cells * sizeof(int32_t), stream));
    dense_row_counts_kernel<<<(unsigned)ceil_div(cells, 8), 256, 0, stream>>>(dense, cells, genes, counts.as<int32_t>());
    MCB_LAUNCH_CHECK();
    std::vector<int32_t> indptr((size_t)cells + 1);
    MCB_TRY(copy_to_host(indptr.data() + 1, counts.ptr, (size_t)cells * sizeof(int32_t), stream));
    MCB_CUDA(cudaStreamSynchronize(stream));
    indptr[0] = 0;
    int64_t total = 0, widest = 0;
    for (int64_t c = 0; c < cells; ++c) {
        widest = std::max<int64_t>(widest, indptr[c + 1]);
        total += indptr[c + 1];
        MCB_REQUIRE(total < (1ll << 31), "pile: more than 2^31 - 1 non-zeros");
        indptr[c + 1] = (int32_t)total;
    }
    pile->nnz = total;
    pile->widest = widest;
    MCB_TRY(pile->own_data.alloc((size_t)total * sizeof(float), stream));
    MCB_TRY(pile->own_indices.alloc((size_t)total * sizeof(int32_t), stream));
    MCB_TRY(pile->own_indptr.alloc((size_t)(cells + 1) * sizeof(int32_t), stream));
    MCB_TRY(pile->status.alloc(sizeof(int), stream));
    MCB_CUDA(cudaMemsetAsync(pile->status.ptr, 0, sizeof(int), stream));
    MCB_CUDA(cudaMemcpyAsync(pile->own_indptr.ptr, indptr.data(), (size_t)(cells + 1) * sizeof(int32_t), cudaMemcpyHostToDevice, stream));
    dense_to_csr_kernel<<<(unsigned)ceil_div(cells, 8), 256, 0, stream>>>(dense, cells, genes, pile->own_indptr.as<int32_t>(),
                                                                      pile->own_data.as<float>(), pile->own_indices.as<int32_t>());
    MCB_LAUNCH_CHECK();
    MCB_CUDA(cudaStreamSynchronize(stream));  // the local row pointers have been read; `dense` may be reused by the caller
    pile->data = pile->own_data.as<float>();
    pile->indices = pile->own_indices.as<int32_t>();
    pile->indptr = pile->own_indptr.as<int32_t>();
    *out = pile.release();
    return MCB_OK;
}

int mcb_host_sliding_median_f32(const float* values, int64_t count, int64_t window, float* medians) {
    MCB_REQUIRE(values && medians && window >= 1 && (window & 1) == 1 && window <= 127 && count >= window,
                "sliding_median: needs an odd window of at most 127 and at least one whole window");
    const int64_t outputs = count - window + 1;
    cudaStream_t stream = nullptr;
    MCB_TRY(thread_stream(&stream));
    Scratch d_values, d_out;
    MCB_TRY(d_values.alloc((size_t)count * sizeof(float), stream));
    MCB_TRY(d_out.alloc((size_t)outputs * sizeof(float), stream));
    MCB_TRY(upload_small(d_values.ptr, values, (size_t)count * sizeof(float), stream));
    MCB_PROFILED("sliding_median", stream,
                 sliding_median_kernel<<<(unsigned)ceil_div(outputs, 8), 256, 0, stream>>>(d_values.as<float>(), outputs, (int)window,
                                                                                         d_out.as<float>()));
    MCB_LAUNCH_CHECK();
    MCB_TRY(copy_to_host(medians, d_out.ptr, (size_t)outputs * sizeof(float), stream));
    MCB_CUDA(cudaStreamSynchronize(stream));
    return MCB_OK;
}

int mcb_host_sliding_median_ordered_f32(const float* values, const int64_t* order, int64_t count, int64_t window,
                                        double* medians) {
    MCB_REQUIRE(values && order && medians && window >= 1 && (window & 1) == 1 && window <= 127 && count > window,
                "sliding_median: needs an odd window of at most 127 and more entries than the window");
    cudaStream_t stream = nullptr;
    MCB_TRY(thread_stream(&stream));
    Scratch d_values, d_order, d_out;
    MCB_TRY(d_values.alloc((size_t)count * sizeof(float), stream));
    MCB_TRY(d_order.alloc((size_t)count * sizeof(int64_t), stream));
    MCB_TRY(d_out.alloc((size_t)count * sizeof(double), stream));
    MCB_TRY(upload_small(d_values.ptr, values, (size_t)count * sizeof(float), stream));
    MCB_TRY(upload_small(d_order.ptr, order, (size_t)count * sizeof(int64_t), stream));
    MCB_PROFILED("sliding_median", stream,
                 sliding_median_ordered_kernel<<<(unsigned)ceil_div(count, 8), 256, 0, stream>>>(
                     d_values.as<float>(), d_order.as<int64_t>(), count, (int)window, d_out.as<double>()));
    MCB_LAUNCH_CHECK();
    MCB_TRY(copy_to_host(medians, d_out.ptr, (size_t)count * sizeof(double), stream));
    MCB_CUDA(cudaStreamSynchronize(stream));
    return MCB_OK;
}

int mcb_pile_want_gene_statistics(mcb_pile* pile, float min_top_value) {
    MCB_REQUIRE(pile != nullptr, "pile: null handle");
    MCB_REQUIRE(!pile->has_downsampled, "pile: ask for the gene statistics before mcb_pile_downsample");
    MCB_REQUIRE(pile->cells < 65536, "pile: gene statistics count cells in 16 bits (at most 65535 cells)");
    pile->want_gene_stats = true;
    pile->gene_top_threshold = min_top_value;
    return MCB_OK;
}

int mcb_pile_gene_statistics(mcb_pile* pile, float* totals, uint64_t* squares, int32_t* top_counts) {
    MCB_REQUIRE(pile != nullptr && totals && squares && top_counts, "pile: null argument");
    MCB_REQUIRE(pile->has_downsampled && pile->gene_stats.ptr != nullptr,
                "pile: no gene statistics (mcb_pile_want_gene_statistics, then mcb_pile_downsample)");
    MCB_TRY(check_status(pile));
    MCB_TRY(copy_to_host(totals, pile->gene_totals.ptr, (size_t)pile->genes * sizeof(float), pile->stream));
    MCB_TRY(copy_to_host(squares, pile->gene_stats.ptr, (size_t)pile->genes * sizeof(uint64_t), pile->stream));
    MCB_CUDA(cudaStreamSynchronize(pile->stream));
    for (int64_t gene = 0; gene < pile->genes; ++gene) {
        top_counts[gene] = (int32_t)(squares[gene] & 0xffffu);
        squares[gene] >>= 16;
    }
    return MCB_OK;
}

int mcb_pile_gene_columns(mcb_pile* pile, const int32_t* genes, int64_t count, float* values, int32_t* sizes) {
    MCB_REQUIRE(pile != nullptr && genes && values && sizes && count >= 0, "pile: bad arguments");
    if (count == 0)
        return MCB_OK;
    for (int64_t g = 0; g < count; ++g)
        MCB_REQUIRE(genes[g] >= 0 && genes[g] < pile->genes, "pile: gene %d is out of range", genes[g]);
    cudaStream_t stream = pile->stream;
    MCB_TRY(wait_for_upload(pile));
    Scratch d_genes, d_values;
    MCB_REQUIRE(count <= 65535, "pile: at most 65535 gene columns per call");
    MCB_TRY(d_genes.alloc((size_t)count * sizeof(int32_t), stream));
    MCB_TRY(d_values.alloc((size_t)count * pile->cells * sizeof(float), stream));
    MCB_TRY(upload_small(d_genes.ptr, genes, (size_t)count * sizeof(int32_t), stream));
    const float* source = pile->has_downsampled ? pile->downsampled.as<float>() : pile->data;
    const dim3 grid((unsigned)ceil_div(pile->cells, 256), (unsigned)count);
    MCB_PROFILED("gene_columns", stream,
                 gene_columns_kernel<<<grid, 256, 0, stream>>>(source, pile->indices, pile->indptr, pile->cells,
                                                               d_genes.as<int32_t>(), d_values.as<float>()));
    MCB_LAUNCH_CHECK();
    MCB_TRY(copy_to_host(values, d_values.ptr, (size_t)count * pile->cells * sizeof(float), stream));
    MCB_CUDA(cudaStreamSynchronize(stream));
    for (int64_t g = 0; g < count; ++g) {  // drop the zeros, keep the order
        float* column = values + g * pile->cells;
        int32_t kept = 0;
        for (int64_t row = 0; row < pile->cells; ++row)
            if (column[row] != 0.0f) column[kept++] = column[row];
        std::fill(column + kept, column + pile->cells, 0.0f);
        sizes[g] = kept;
    }
    return MCB_OK;
}

int64_t mcb_pile_nnz(const mcb_pile* pile) { return pile ? pile->nnz : 0; }
int64_t mcb_pile_top_genes_count(const mcb_pile* pile) { return pile ? pile->top_genes_count : 0; }

int mcb_pile_wait_upload(mcb_pile* pile) {
    MCB_REQUIRE(pile != nullptr, "pile: null handle");
    if (pile->copy_stream)
        MCB_CUDA(cudaStreamSynchronize(pile->copy_stream));
    return MCB_OK;
}

void mcb_pile_free(mcb_pile* pile) { delete pile; }

}  // extern "C"

// ---- dissolve: per-gene sums of groups of a pile's cells (tools/dissolve.py:101-102, :187-189) ---------------------------

namespace {

// One warp per cell; float64 atomics: sums of UMI counts are exact (below 2^53) in any order.
__global__ void __launch_bounds__(256) pile_sum_groups_kernel(const float* __restrict__ data, const int32_t* __restrict__ indices,
                                                             const int32_t* __restrict__ indptr, int64_t rows,
                                                             const int32_t* __restrict__ group_of_row, int64_t genes,
                                                             double* __restrict__ sums) {
    const int lane = threadIdx.x & 31;
    const int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (row >= rows) return;
    const int32_t group = group_of_row[row];
    if (group < 0) return;
    double* target = sums + (int64_t)group * genes;
    for (int e = indptr[row] + lane; e < indptr[row + 1]; e += 32) atomicAdd(&target[indices[e]], (double)data[e]);
}

}  // namespace

extern "C" {

int mcb_pile_sum_groups(mcb_pile* pile, const int32_t* group_of_cells, int64_t groups, double* sums) {
    MCB_REQUIRE(pile != nullptr && group_of_cells && sums && groups >= 1, "pile: bad arguments");
    for (int64_t cell = 0; cell < pile->cells; ++cell)
        MCB_REQUIRE(group_of_cells[cell] < groups, "pile: cell %lld is in group %d of %lld", (long long)cell,
                    group_of_cells[cell], (long long)groups);
    cudaStream_t stream = pile->stream;
    MCB_TRY(wait_for_upload(pile));
    Scratch d_groups, d_sums;
    MCB_TRY(d_groups.alloc((size_t)pile->cells * sizeof(int32_t), stream));
    MCB_TRY(d_sums.alloc((size_t)groups * pile->genes * sizeof(double), stream));
    MCB_TRY(upload_small(d_groups.ptr, group_of_cells, (size_t)pile->cells * sizeof(int32_t), stream));
    MCB_CUDA(cudaMemsetAsync(d_sums.ptr, 0, (size_t)groups * pile->genes * sizeof(double), stream));
    MCB_PROFILED("pile_sum_groups", stream,
                 pile_sum_groups_kernel<<<(unsigned)ceil_div(pile->cells, 8), 256, 0, stream>>>(
                     pile->data, pile->indices, pile->indptr, pile->cells, d_groups.as<int32_t>(), pile->genes,
                     d_sums.as<double>()));
    MCB_LAUNCH_CHECK();
    MCB_TRY(copy_to_host(sums, d_sums.ptr, (size_t)groups * pile->genes * sizeof(double), stream));
    MCB_CUDA(cudaStreamSynchronize(stream));
    return MCB_OK;
}

}  // extern "C"

// ---- collect_metacells (pipeline/collect.py:127-232) ------------------------------------------------------------------

namespace {

// One CTA per metacell, its cells one after the other (a barrier between two cells): a cell's stored counts are added to
// the metacell's per-gene UMIs, the genes it still holds after the "zeros" downsampling are counted, and its share of the
// weighted geometric mean of the gene fractions (collect.py:196-225) is added up as
// weight * (log(umis / total + regularization) - log(regularization)):  a gene a cell does not express contributes
// weight * log(regularization), the same for every gene, which the finishing kernel adds for all cells at once.  No
// atomics: the float64 sums are made in cell order, so every run and every rank gets the same bits (the ranks of a walk
// must agree on the grouping they each compute from these).
__global__ void __launch_bounds__(256) collect_cells_kernel(const float* __restrict__ data, const float* __restrict__ downsampled,
                                                           const int32_t* __restrict__ indices, const int32_t* __restrict__ indptr,
                                                           int64_t genes, const int32_t* __restrict__ first_cell_of,
                                                           const double* __restrict__ weight, const double* __restrict__ regularization,
                                                           const double* __restrict__ total, float* __restrict__ sums,
                                                           int32_t* __restrict__ kept, double* __restrict__ logs) {
    const int64_t metacell = blockIdx.x;
    const int64_t base = metacell * genes;
    for (int row = first_cell_of[metacell]; row < first_cell_of[metacell + 1]; ++row) {
        const double w = weight[row], r = regularization[row], t = total[row];
        const double log_r = log(r);
        for (int e = indptr[row] + threadIdx.x; e < indptr[row + 1]; e += blockDim.x) {
            const float umis = data[e];
            if (umis == 0.0f) continue;
            const int64_t at = base + indices[e];
            sums[at] += umis;
            if (downsampled[e] != 0.0f) kept[at] += 1;
            if (logs != nullptr) logs[at] += w * (log((double)umis / t + r) - log_r);
        }
        __syncthreads();
    }
}

// One CTA per metacell: the fractions are finished (collect.py:216-232: the geometric mean less that of the
// regularization, genes without UMIs and negative values zeroed, normalised to 1, float32, times `scale`), the kept
// counts become the zeros.
__global__ void __launch_bounds__(256) collect_finish_kernel(int64_t genes, const float* __restrict__ sums, int32_t* __restrict__ kept_then_zeros,
                                                            const double* __restrict__ logs, const double* __restrict__ mean_log_regularization,
                                                            const double* __restrict__ weight_sum, const double* __restrict__ scale,
                                                            const double* __restrict__ total_umis, const int32_t* __restrict__ cells_of,
                                                            float* __restrict__ fractions) {
    __shared__ double partial[8];
    const int64_t metacell = blockIdx.x;
    const int64_t base = metacell * genes;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const bool geometric = logs != nullptr;
    const double mean_log_r = geometric ? mean_log_regularization[metacell] : 0.0;
    const double of_regularization = geometric ? exp(mean_log_r) : 0.0;
    const double weights = geometric ? weight_sum[metacell] : 1.0;
    double sum = 0.0;
    for (int64_t gene = threadIdx.x; gene < genes; gene += blockDim.x) {
        double fraction;
        if (geometric) {
            fraction = exp(mean_log_r + logs[base + gene] / weights) - of_regularization;
            if (sums[base + gene] == 0.0f || fraction < 0.0) fraction = 0.0;
        } else {
            fraction = (double)sums[base + gene];
        }
        sum += fraction;
        kept_then_zeros[base + gene] = cells_of[metacell] - kept_then_zeros[base + gene];
    }
    for (int offset = 16; offset > 0; offset >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, offset);
    if (lane == 0) partial[warp] = sum;
    __syncthreads();
    sum = 0.0;
    for (int w = 0; w < 8; ++w) sum += partial[w];
    const double factor = scale[metacell];
    for (int64_t gene = threadIdx.x; gene < genes; gene += blockDim.x) {
        double fraction;
        if (geometric) {
            fraction = exp(mean_log_r + logs[base + gene] / weights) - of_regularization;
            if (sums[base + gene] == 0.0f || fraction < 0.0) fraction = 0.0;
            fraction /= sum;
        } else {
            fraction = (double)sums[base + gene] / total_umis[metacell];
        }
        // .astype("float32"), then `*= total_umis` with a numpy int64: a float64 product stored as float32
        fractions[base + gene] = factor == 1.0 ? (float)fraction : (float)((double)(float)fraction * factor);
    }
}

}  // namespace

extern "C" {

int mcb_pile_collect_metacells(mcb_pile* pile, const int32_t* metacell_of_cells, int64_t metacells, const double* weight_of_cells,
                               const double* regularization_of_cells, const double* total_of_cells,
                               const int32_t* samples_of_cells, const int32_t* seed_index_of_cells, uint64_t random_seed,
                               int geometric, const double* mean_log_regularization, const double* weight_sum, const double* scale,
                               const double* total_umis, const int32_t* cells_of_metacells, float* fractions_device,
                               float* sums_device, int32_t* zeros_device) {
    MCB_REQUIRE(pile && metacell_of_cells && metacells >= 1 && weight_of_cells && regularization_of_cells && total_of_cells &&
                    samples_of_cells && seed_index_of_cells && mean_log_regularization && weight_sum && scale && total_umis &&
                    cells_of_metacells && fractions_device && sums_device && zeros_device,
                "collect_metacells: null argument");
    const int64_t cells = pile->cells, genes = pile->genes;
    // the cells of a metacell lie together, the metacells in order: their first cells
    std::vector<int32_t> first_cell_of((size_t)metacells + 1, 0);
    for (int64_t cell = 0; cell < cells; ++cell) {
        const int32_t metacell = metacell_of_cells[cell];
        MCB_REQUIRE(metacell >= 0 && metacell < metacells, "collect_metacells: cell %lld is in metacell %d of %lld", (long long)cell,
                    metacell, (long long)metacells);
        MCB_REQUIRE(cell == 0 || metacell >= metacell_of_cells[cell - 1], "collect_metacells: the cells must be in metacell order");
        first_cell_of[(size_t)metacell + 1] += 1;
    }
    for (int64_t metacell = 0; metacell < metacells; ++metacell) first_cell_of[(size_t)metacell + 1] += first_cell_of[(size_t)metacell];
    cudaStream_t stream = pile->stream;
    MCB_TRY(wait_for_upload(pile));
    Scratch d_metacell, d_weight, d_regularization, d_total, d_samples, d_seed_rows, d_down, d_logs, d_mean, d_weights, d_scale,
        d_totals, d_counts;
    auto upload = [&](Scratch& device, const void* host, size_t bytes) -> int {
        MCB_TRY(device.alloc(bytes, stream));
        MCB_CUDA(cudaMemcpyAsync(device.ptr, host, bytes, cudaMemcpyHostToDevice, stream));
        return MCB_OK;
    };
    MCB_TRY(upload(d_metacell, first_cell_of.data(), ((size_t)metacells + 1) * sizeof(int32_t)));
    MCB_TRY(upload(d_weight, weight_of_cells, (size_t)cells * sizeof(double)));
    MCB_TRY(upload(d_regularization, regularization_of_cells, (size_t)cells * sizeof(double)));
    MCB_TRY(upload(d_total, total_of_cells, (size_t)cells * sizeof(double)));
    MCB_TRY(upload(d_samples, samples_of_cells, (size_t)cells * sizeof(int32_t)));
    MCB_TRY(upload(d_seed_rows, seed_index_of_cells, (size_t)cells * sizeof(int32_t)));
    MCB_TRY(upload(d_mean, mean_log_regularization, (size_t)metacells * sizeof(double)));
    MCB_TRY(upload(d_weights, weight_sum, (size_t)metacells * sizeof(double)));
    MCB_TRY(upload(d_scale, scale, (size_t)metacells * sizeof(double)));
    MCB_TRY(upload(d_totals, total_umis, (size_t)metacells * sizeof(double)));
    MCB_TRY(upload(d_counts, cells_of_metacells, (size_t)metacells * sizeof(int32_t)));
    // the "zeros" downsampling of collect.py:160-171: every metacell's cells as a matrix of their own (the seed of a row is
    // made of its index in that matrix), each to its metacell's quantile of the cells' UMIs
    MCB_TRY(d_down.alloc((size_t)std::max<int64_t>(pile->nnz, 1) * sizeof(float), stream));
    MCB_TRY(downsample_rows_launch(pile->data, d_down.as<float>(), pile->indptr, d_samples.as<int32_t>(), 0, 0, cells, 0,
                                   pile->widest, random_seed, 1, 0, nullptr, stream, nullptr, nullptr, nullptr, 0.0f,
                                   d_seed_rows.as<int32_t>()));
    const size_t entries = (size_t)metacells * genes;
    MCB_CUDA(cudaMemsetAsync(sums_device, 0, entries * sizeof(float), stream));
    MCB_CUDA(cudaMemsetAsync(zeros_device, 0, entries * sizeof(int32_t), stream));
    if (geometric) {
        MCB_TRY(d_logs.alloc(entries * sizeof(double), stream));
        MCB_CUDA(cudaMemsetAsync(d_logs.ptr, 0, entries * sizeof(double), stream));
    }
    MCB_PROFILED("collect_cells", stream,
                 collect_cells_kernel<<<(unsigned)metacells, 256, 0, stream>>>(
                     pile->data, d_down.as<float>(), pile->indices, pile->indptr, genes, d_metacell.as<int32_t>(),
                     d_weight.as<double>(), d_regularization.as<double>(), d_total.as<double>(), sums_device, zeros_device,
                     geometric ? d_logs.as<double>() : nullptr));
    MCB_LAUNCH_CHECK();
    MCB_PROFILED("collect_finish", stream,
                 collect_finish_kernel<<<(unsigned)metacells, 256, 0, stream>>>(
                     genes, sums_device, zeros_device, geometric ? d_logs.as<double>() : nullptr, d_mean.as<double>(),
                     d_weights.as<double>(), d_scale.as<double>(), d_totals.as<double>(), d_counts.as<int32_t>(), fractions_device));
    MCB_LAUNCH_CHECK();
    MCB_CUDA(cudaStreamSynchronize(stream));  // the host arrays above may be pageable
    return MCB_OK;
}

}  // extern "C"

// ---- deviant cells: the dense inputs of compute_cell_gaps, built on the device (tools/deviants.py:270-292) ------------

namespace mcb {
int cell_gaps_launch(const float* umis, const float* log_fraction, int64_t cells, int64_t genes, int64_t stride,
                     const int32_t* candidate_of_cells, const uint8_t* deviant_per_cell, const uint8_t* active_per_cell,
                     const float* min_gap_per_gene, int64_t candidates_count, int64_t gap_skip_cells,
                     int64_t max_deviant_cells_count, float max_deviant_cells_fraction, float min_compare_umis,
                     float* max_gap_per_cell, cudaStream_t stream);
}

struct mcb_gap_inputs {
    int64_t cells = 0, genes = 0;
    Scratch umis, log_fraction;       // float32 [cells, genes]
    Scratch exception_at, exception_x;  // entries whose count is beyond the table: position in the matrix, fraction + regularization
    int64_t exceptions = 0;
};

namespace {

constexpr int64_t GAP_EXCEPTIONS_CAPACITY = 1 << 22;

// One warp per cell: the row of log fractions starts as the cell's value of a zero count, then every stored count is
// written to the dense UMIs and looked up in the cell's row of the table; counts the table does not hold (large, or not
// whole numbers) are listed for the host, which takes their logarithm with numpy's own log2.
__global__ void gap_inputs_kernel(const float* __restrict__ data, const int32_t* __restrict__ indices,
                                  const int32_t* __restrict__ indptr, int64_t cells, int64_t genes,
                                  const float* __restrict__ total_per_cell, const float* __restrict__ regularization_per_cell,
                                  const float* __restrict__ table, int table_counts, float* __restrict__ umis,
                                  float* __restrict__ log_fraction, int64_t* __restrict__ exception_at,
                                  float* __restrict__ exception_x, unsigned long long* __restrict__ exceptions,
                                  int64_t capacity) {
    const int64_t cell = (int64_t(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (cell >= cells) return;
    const float* row_table = table + cell * table_counts;
    const float of_zero = row_table[0];
    float* row_logs = log_fraction + cell * genes;
    float* row_umis = umis + cell * genes;
    for (int64_t gene = lane; gene < genes; gene += 32) {
        row_logs[gene] = of_zero;
        row_umis[gene] = 0.0f;
    }
    __syncwarp();
    const float total = total_per_cell[cell], regularization = regularization_per_cell[cell];
    for (int at = indptr[cell] + lane; at < indptr[cell + 1]; at += 32) {
        const float value = data[at];
        const int gene = indices[at];
        row_umis[gene] = value;
        const int whole = int(value);
        if (value >= 0.0f && value < float(table_counts) && float(whole) == value) {
            row_logs[gene] = row_table[whole];
        } else {
            const unsigned long long slot = atomicAdd(exceptions, 1ull);
            if (slot < (unsigned long long)capacity) {
                exception_at[slot] = cell * genes + gene;
                exception_x[slot] = __fadd_rn(__fdiv_rn(value, total), regularization);
            }
        }
    }
}

__global__ void gap_patch_kernel(const int64_t* __restrict__ exception_at, const float* __restrict__ logs, int64_t count,
                                 float* __restrict__ log_fraction) {
    const int64_t at = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (at < count) log_fraction[exception_at[at]] = logs[at];
}

}  // namespace

extern "C" {

int mcb_pile_gap_inputs_create(mcb_pile* pile, const float* total_umis_per_cell, const float* regularization_per_cell,
                               const float* log_table, int64_t table_counts, mcb_gap_inputs** out) {
    MCB_REQUIRE(pile && total_umis_per_cell && regularization_per_cell && log_table && out, "gap_inputs: null argument");
    MCB_REQUIRE(table_counts >= 1 && table_counts <= (1 << 20), "gap_inputs: table_counts out of range");
    *out = nullptr;
    cudaStream_t stream = pile->stream;
    MCB_TRY(wait_for_upload(pile));
    const int64_t cells = pile->cells, genes = pile->genes;
    std::unique_ptr<mcb_gap_inputs> inputs(new mcb_gap_inputs());
    inputs->cells = cells;
    inputs->genes = genes;
    MCB_TRY(inputs->umis.alloc(size_t(cells) * genes * sizeof(float), stream));
    MCB_TRY(inputs->log_fraction.alloc(size_t(cells) * genes * sizeof(float), stream));
    MCB_TRY(inputs->exception_at.alloc(size_t(GAP_EXCEPTIONS_CAPACITY) * sizeof(int64_t), stream));
    MCB_TRY(inputs->exception_x.alloc(size_t(GAP_EXCEPTIONS_CAPACITY) * sizeof(float), stream));
    Scratch d_totals, d_regularization, d_table, d_count;
    MCB_TRY(d_totals.alloc(size_t(cells) * sizeof(float), stream));
    MCB_TRY(d_regularization.alloc(size_t(cells) * sizeof(float), stream));
    MCB_TRY(d_table.alloc(size_t(cells) * table_counts * sizeof(float), stream));
    MCB_TRY(d_count.alloc(sizeof(unsigned long long), stream));
    MCB_CUDA(cudaMemcpyAsync(d_totals.ptr, total_umis_per_cell, size_t(cells) * sizeof(float), cudaMemcpyHostToDevice, stream));
    MCB_CUDA(cudaMemcpyAsync(d_regularization.ptr, regularization_per_cell, size_t(cells) * sizeof(float),
                             cudaMemcpyHostToDevice, stream));
    MCB_CUDA(cudaMemcpyAsync(d_table.ptr, log_table, size_t(cells) * table_counts * sizeof(float), cudaMemcpyHostToDevice,
                             stream));
    MCB_CUDA(cudaMemsetAsync(d_count.ptr, 0, sizeof(unsigned long long), stream));
    MCB_PROFILED("gap_inputs", stream,
                 gap_inputs_kernel<<<unsigned(ceil_div(cells * 32, 256)), 256, 0, stream>>>(
                     pile->data, pile->indices, pile->indptr, cells, genes, d_totals.as<float>(), d_regularization.as<float>(),
                     d_table.as<float>(), int(table_counts), inputs->umis.as<float>(), inputs->log_fraction.as<float>(),
                     inputs->exception_at.as<int64_t>(), inputs->exception_x.as<float>(), d_count.as<unsigned long long>(),
                     GAP_EXCEPTIONS_CAPACITY));
    MCB_LAUNCH_CHECK();
    unsigned long long count = 0;
    MCB_TRY(copy_to_host(&count, d_count.ptr, sizeof(count), stream));
    MCB_CUDA(cudaStreamSynchronize(stream));
    MCB_REQUIRE(count <= (unsigned long long)GAP_EXCEPTIONS_CAPACITY,
                "gap_inputs: %llu counts are beyond the table of %lld (raise table_counts)", count, (long long)table_counts);
    inputs->exceptions = int64_t(count);
    *out = inputs.release();
    return MCB_OK;
}

int64_t mcb_gap_inputs_exceptions(const mcb_gap_inputs* inputs) { return inputs ? inputs->exceptions : 0; }

int mcb_gap_inputs_read_exceptions(const mcb_gap_inputs* inputs, float* values) {
    MCB_REQUIRE(inputs && (values || inputs->exceptions == 0), "gap_inputs: null argument");
    cudaStream_t stream = nullptr;
    MCB_TRY(thread_stream(&stream));
    MCB_TRY(copy_to_host(values, inputs->exception_x.ptr, size_t(inputs->exceptions) * sizeof(float), stream));
    MCB_CUDA(cudaStreamSynchronize(stream));
    return MCB_OK;
}

int mcb_gap_inputs_patch(mcb_gap_inputs* inputs, const float* logs) {
    MCB_REQUIRE(inputs && (logs || inputs->exceptions == 0), "gap_inputs: null argument");
    if (inputs->exceptions == 0) return MCB_OK;
    cudaStream_t stream = nullptr;
    MCB_TRY(thread_stream(&stream));
    Scratch d_logs;
    MCB_TRY(d_logs.alloc(size_t(inputs->exceptions) * sizeof(float), stream));
    MCB_CUDA(cudaMemcpyAsync(d_logs.ptr, logs, size_t(inputs->exceptions) * sizeof(float), cudaMemcpyHostToDevice, stream));
    MCB_PROFILED("gap_patch", stream,
                 gap_patch_kernel<<<unsigned(ceil_div(inputs->exceptions, 256)), 256, 0, stream>>>(
                     inputs->exception_at.as<int64_t>(), d_logs.as<float>(), inputs->exceptions,
                     inputs->log_fraction.as<float>()));
    MCB_LAUNCH_CHECK();
    MCB_CUDA(cudaStreamSynchronize(stream));
    return MCB_OK;
}

int mcb_gap_inputs_read(const mcb_gap_inputs* inputs, float* umis, float* log_fraction) {
    MCB_REQUIRE(inputs != nullptr, "gap_inputs: null handle");
    cudaStream_t stream = nullptr;
    MCB_TRY(thread_stream(&stream));
    const size_t bytes = size_t(inputs->cells) * inputs->genes * sizeof(float);
    if (umis) MCB_TRY(copy_to_host(umis, inputs->umis.ptr, bytes, stream));
    if (log_fraction) MCB_TRY(copy_to_host(log_fraction, inputs->log_fraction.ptr, bytes, stream));
    MCB_CUDA(cudaStreamSynchronize(stream));
    return MCB_OK;
}

int mcb_gap_inputs_device_arrays(const mcb_gap_inputs* inputs, const float** umis, const float** log_fraction) {
    MCB_REQUIRE(inputs != nullptr, "gap_inputs: null handle");
    if (umis) *umis = inputs->umis.as<float>();
    if (log_fraction) *log_fraction = inputs->log_fraction.as<float>();
    return MCB_OK;
}

int mcb_gap_inputs_cell_gaps(const mcb_gap_inputs* inputs, const int32_t* candidate_of_cells, const uint8_t* deviant_per_cell,
                             const uint8_t* active_per_cell, const float* min_gap_per_gene, int64_t candidates_count,
                             int64_t gap_skip_cells, int64_t max_deviant_cells_count, float max_deviant_cells_fraction,
                             float min_compare_umis, float* max_gap_per_cell) {
    MCB_REQUIRE(inputs && candidate_of_cells && deviant_per_cell && active_per_cell && min_gap_per_gene && max_gap_per_cell,
                "compute_cell_gaps: null argument");
    cudaStream_t stream = nullptr;
    MCB_TRY(thread_stream(&stream));
    return cell_gaps_launch(inputs->umis.as<float>(), inputs->log_fraction.as<float>(), inputs->cells, inputs->genes,
                            inputs->genes, candidate_of_cells, deviant_per_cell, active_per_cell, min_gap_per_gene,
                            candidates_count, gap_skip_cells, max_deviant_cells_count, max_deviant_cells_fraction,
                            min_compare_umis, max_gap_per_cell, stream);
}

void mcb_gap_inputs_free(mcb_gap_inputs* inputs) { delete inputs; }

/* xt.compute_cell_gaps itself (gaps.cpp:6-186) on host arrays: the two dense matrices go up, the gaps come back.  The
 * fraction matrix of the reference's signature is not read by the reference (only its shape is checked) and is left out. */
int mcb_host_compute_cell_gaps(const float* umis_per_gene_per_cell, const float* log_fraction_per_gene_per_cell, int64_t cells,
                               int64_t genes, const int32_t* candidate_of_cells, const uint8_t* deviant_per_cell,
                               const uint8_t* active_per_cell, const float* min_gap_per_gene, int64_t candidates_count,
                               int64_t gap_skip_cells, int64_t max_deviant_cells_count, float max_deviant_cells_fraction,
                               float min_compare_umis, float* max_gap_per_cell) {
    MCB_REQUIRE(umis_per_gene_per_cell && log_fraction_per_gene_per_cell && candidate_of_cells && deviant_per_cell &&
                    active_per_cell && min_gap_per_gene && max_gap_per_cell && cells > 0 && genes > 0,
                "compute_cell_gaps: null argument or empty matrix");
    cudaStream_t stream = nullptr;
    MCB_TRY(thread_stream(&stream));
    Scratch d_umis, d_logs;
    const size_t bytes = size_t(cells) * genes * sizeof(float);
    MCB_TRY(d_umis.alloc(bytes, stream));
    MCB_TRY(d_logs.alloc(bytes, stream));
    MCB_CUDA(cudaMemcpyAsync(d_umis.ptr, umis_per_gene_per_cell, bytes, cudaMemcpyHostToDevice, stream));
    MCB_CUDA(cudaMemcpyAsync(d_logs.ptr, log_fraction_per_gene_per_cell, bytes, cudaMemcpyHostToDevice, stream));
    return cell_gaps_launch(d_umis.as<float>(), d_logs.as<float>(), cells, genes, genes, candidate_of_cells, deviant_per_cell,
                            active_per_cell, min_gap_per_gene, candidates_count, gap_skip_cells, max_deviant_cells_count,
                            max_deviant_cells_fraction, min_compare_umis, max_gap_per_cell, stream);
}

int mcb_dev_compute_cell_gaps(const float* umis_per_gene_per_cell, const float* log_fraction_per_gene_per_cell, int64_t cells,
                              int64_t genes, const int32_t* candidate_of_cells, const uint8_t* deviant_per_cell,
                              const uint8_t* active_per_cell, const float* min_gap_per_gene, int64_t candidates_count,
                              int64_t gap_skip_cells, int64_t max_deviant_cells_count, float max_deviant_cells_fraction,
                              float min_compare_umis, float* max_gap_per_cell, void* stream) {
    MCB_REQUIRE(umis_per_gene_per_cell && log_fraction_per_gene_per_cell && candidate_of_cells && deviant_per_cell &&
                    active_per_cell && min_gap_per_gene && max_gap_per_cell && cells > 0 && genes > 0,
                "compute_cell_gaps: null argument or empty matrix");
    return cell_gaps_launch(umis_per_gene_per_cell, log_fraction_per_gene_per_cell, cells, genes, genes, candidate_of_cells,
                            deviant_per_cell, active_per_cell, min_gap_per_gene, candidates_count, gap_skip_cells,
                            max_deviant_cells_count, max_deviant_cells_fraction, min_compare_umis, max_gap_per_cell,
                            static_cast<cudaStream_t>(stream));
}

}  // extern "C"
