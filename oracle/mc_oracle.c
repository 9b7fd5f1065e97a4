/*
 * mc_oracle.c -- TEST INFRASTRUCTURE ONLY.
 *
 * A plain-C, single-threaded CPU restatement of the reference's per-pile k-NN hot path
 * (tanaylab/metacells v0.9.5).  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may load this; the product (metacells_b200/) never does.
 *
 * Each function cites the reference file:line it restates.  The restatement is pinned against
 * the compiled reference (oracle/_ref, built by oracle/Makefile from the sources in place) and
 * against fixtures in tests/golden/ generated from that build (tests/golden/make_golden.py).
 *
 * Where the reference's result depends on libstdc++ introsort/introselect tie order (unstable
 * std::sort / std::nth_element on value-only comparators: rank.cpp:52-66, top_per.cpp:23,
 * prune_per.cpp:33) the oracle uses the documented rule of this project instead: among exactly
 * equal keys the lower original position comes first.  Tests compare "modulo permutations
 * inside groups of exactly equal keys" against oracle/_ref, and exactly against this file.
 *
 * Build: gcc -O2 -fPIC -shared -o oracle/libmc_oracle.so oracle/mc_oracle.c -lm
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define MINSTD_A 48271ull
#define MINSTD_M 2147483647ull

/* std::minstd_rand seeding and stepping (C++ standard [rand.eng.lcong]). */
static uint32_t minstd_seed(uint64_t seed) {
    uint32_t x = (uint32_t)(seed % MINSTD_M);
    return x == 0 ? 1u : x;
}
static uint32_t minstd_next(uint32_t* state) {
    *state = (uint32_t)((MINSTD_A * (uint64_t)(*state)) % MINSTD_M);
    return *state;
}

/* downsample.cpp:5-8 -- uses floating log2; may over-allocate, harmless. */
static size_t ceil_power_of_two(size_t size) {
    size_t p = 1;
    while (p < size) p <<= 1;
    return p;
}

/*
 * downsample.cpp:81-113 (downsample_slice) with initialize_tree :18-44 and random_sample :46-79.
 * The binary sum tree is laid out level after level (leaves first) exactly as the reference does;
 * the walk decrements every node on the path and goes right when random >= left child.
 * Returns 0 on success.
 */
static int downsample_slice_f32(const float* input, float* output, size_t size, int32_t samples,
                                uint64_t random_seed) {
    if (samples < 0 || size == 0) return 0; /* :86-88: output untouched */
    if (size == 1) {                        /* :90-93 */
        output[0] = (double)samples < (double)input[0] ? (float)samples : input[0];
        return 0;
    }
    size_t leaves = ceil_power_of_two(size);
    size_t tree_size = 2 * leaves - 1;
    uint64_t* tree = (uint64_t*)malloc(tree_size * sizeof(uint64_t));
    if (!tree) return 1;
    for (size_t i = 0; i < size; ++i) tree[i] = (uint64_t)input[i];
    for (size_t i = size; i < leaves; ++i) tree[i] = 0;
    {
        size_t level_base = 0, level_size = leaves;
        while (level_size > 1) {
            size_t next_base = level_base + level_size;
            for (size_t i = 0; i < level_size / 2; ++i)
                tree[next_base + i] = tree[level_base + 2 * i] + tree[level_base + 2 * i + 1];
            level_base = next_base;
            level_size /= 2;
        }
    }
    uint64_t* total = &tree[tree_size - 1];
    if (*total <= (uint64_t)samples) { /* :100-105 */
        if (output != input) memcpy(output, input, size * sizeof(float));
        free(tree);
        return 0;
    }
    for (size_t i = 0; i < size; ++i) output[i] = 0.0f;
    uint32_t state = minstd_seed(random_seed);
    for (int32_t draw = 0; draw < samples; ++draw) {
        int64_t random = (int64_t)(minstd_next(&state) % *total); /* :111 */
        /* random_sample :46-79 */
        size_t size_of_level = 1;
        int64_t base_of_level = (int64_t)tree_size - 1;
        size_t index_in_level = 0;
        for (;;) {
            --tree[base_of_level + (int64_t)index_in_level];
            size_of_level *= 2;
            base_of_level -= (int64_t)size_of_level;
            if (base_of_level < 0) break;
            index_in_level *= 2;
            int64_t right_random = random - (int64_t)tree[base_of_level + (int64_t)index_in_level];
            if (right_random >= 0) {
                ++index_in_level;
                random = right_random;
            }
        }
        output[index_in_level] += 1.0f;
    }
    free(tree);
    return 0;
}

/* downsample.cpp:171-192 (downsample_compressed<float32,int32,float32>), per-band seed :189. */
int mco_downsample_compressed_f32_i32_f32(const float* data, const int32_t* indptr, float* out,
                                          const int32_t* samples, int64_t bands, uint64_t seed) {
    for (int64_t b = 0; b < bands; ++b) {
        uint64_t band_seed = seed == 0 ? 0 : seed + (uint64_t)b * 997ull;
        int32_t lo = indptr[b], hi = indptr[b + 1];
        if (downsample_slice_f32(data + lo, out + lo, (size_t)(hi - lo), samples[b], band_seed)) return 1;
    }
    return 0;
}

/* downsample.cpp:131-151 (downsample_dense<float32,float32>). */
int mco_downsample_dense_f32_f32(const float* input, float* output, const int32_t* samples, int64_t rows,
                                 int64_t cols, uint64_t seed) {
    for (int64_t r = 0; r < rows; ++r) {
        uint64_t row_seed = seed == 0 ? 0 : seed + (uint64_t)r * 997ull;
        if (downsample_slice_f32(input + r * cols, output + r * cols, (size_t)cols, samples[r], row_seed)) return 1;
    }
    return 0;
}

/* downsample.cpp:116-128 (downsample_array<float32,float32>). */
int mco_downsample_array_f32_f32(const float* input, float* output, int64_t size, int32_t samples, uint64_t seed) {
    return downsample_slice_f32(input, output, (size_t)size, samples, seed);
}

/* ------------------------------------------------------------------------------------------ */
/* correlate.cpp                                                                               */
/* ------------------------------------------------------------------------------------------ */

/* correlate.cpp:14-27 (sum_row_values): float64 sums, squares through fma. */
static void sum_row_values_f32(const float* row, int64_t cols, double* sum, double* squared) {
    double s = 0, q = 0;
    for (int64_t c = 0; c < cols; ++c) {
        double v = row[c];
        s += v;
        q = fma(v, v, q);
    }
    *sum = s;
    *squared = q;
}

/*
 * correlate.cpp:29-62 (correlate_two_dense_rows): the reference's float64-accumulating pair
 * kernel.  The 2-row blocked variants (:161-206 generic float32 accumulate, :208-338 AVX2
 * 8-lane float32 accumulate) differ from this by up to 3e-5 / 3e-6 respectively (BASELINE.md
 * section 2); the parity contract is 1e-5 absolute to exact Pearson, so the oracle restates the
 * most accurate of the reference's own kernels for every pair.
 */
static float correlate_two_rows_f32(const float* a, double a_sum, double a_sq, const float* b, double b_sum,
                                    double b_sq, int64_t cols) {
    double both = 0;
    for (int64_t c = 0; c < cols; ++c) both = fma((double)a[c], (double)b[c], both);
    double correlation = (double)cols * both - a_sum * b_sum;
    double a_factor = (double)cols * a_sq - a_sum * a_sum;
    double b_factor = (double)cols * b_sq - b_sum * b_sum;
    double factors = sqrt(a_factor * b_factor);
    if (factors != 0) {
        float r = (float)(correlation / factors);
        if (r > 1.0f) r = 1.0f;
        if (r < -1.0f) r = -1.0f;
        return r;
    }
    return 0.0f;
}

/* correlate.cpp:350-431 (correlate_dense<float32>): diagonal 1, symmetric write. */
int mco_correlate_dense_f32(const float* input, int64_t rows, int64_t cols, int64_t row_stride, float* output) {
    double* sums = (double*)malloc(2 * (size_t)rows * sizeof(double));
    if (!sums && rows) return 1;
    for (int64_t r = 0; r < rows; ++r) sum_row_values_f32(input + r * row_stride, cols, &sums[2 * r], &sums[2 * r + 1]);
    for (int64_t i = 0; i < rows; ++i) {
        output[i * rows + i] = 1.0f;
        for (int64_t j = 0; j < i; ++j) {
            float r = correlate_two_rows_f32(input + i * row_stride, sums[2 * i], sums[2 * i + 1], input + j * row_stride,
                                             sums[2 * j], sums[2 * j + 1], cols);
            output[i * rows + j] = r;
            output[j * rows + i] = r;
        }
    }
    free(sums);
    return 0;
}

/* correlate.cpp:433-494 (cross_correlate_dense<float32>). */
int mco_cross_correlate_dense_f32(const float* first, int64_t first_rows, const float* second, int64_t second_rows,
                                  int64_t cols, float* output) {
    double* sums = (double*)malloc(2 * (size_t)second_rows * sizeof(double));
    if (!sums && second_rows) return 1;
    for (int64_t r = 0; r < second_rows; ++r) sum_row_values_f32(second + r * cols, cols, &sums[2 * r], &sums[2 * r + 1]);
    for (int64_t i = 0; i < first_rows; ++i) {
        double s, q;
        sum_row_values_f32(first + i * cols, cols, &s, &q);
        for (int64_t j = 0; j < second_rows; ++j)
            output[i * second_rows + j] =
                correlate_two_rows_f32(first + i * cols, s, q, second + j * cols, sums[2 * j], sums[2 * j + 1], cols);
    }
    free(sums);
    return 0;
}

/* correlate.cpp:496-527 (pairs_correlate_dense<float32>). */
int mco_pairs_correlate_dense_f32(const float* first, const float* second, int64_t rows, int64_t cols, float* output) {
    for (int64_t i = 0; i < rows; ++i) {
        double s1, q1, s2, q2;
        sum_row_values_f32(first + i * cols, cols, &s1, &q1);
        sum_row_values_f32(second + i * cols, cols, &s2, &q2);
        output[i] = correlate_two_rows_f32(first + i * cols, s1, q1, second + i * cols, s2, q2, cols);
    }
    return 0;
}

/* ------------------------------------------------------------------------------------------ */
/* rank.cpp / top_per.cpp / prune_per.cpp -- ordering helpers                                  */
/* ------------------------------------------------------------------------------------------ */

typedef struct {
    float value;
    int64_t position;
} keyed_t;

/* Descending by value; among exactly equal values the lower position first (project tie rule). */
static int cmp_desc(const void* pa, const void* pb) {
    const keyed_t* a = (const keyed_t*)pa;
    const keyed_t* b = (const keyed_t*)pb;
    if (a->value > b->value) return -1;
    if (a->value < b->value) return 1;
    return (a->position > b->position) - (a->position < b->position);
}
static int cmp_asc(const void* pa, const void* pb) {
    const keyed_t* a = (const keyed_t*)pa;
    const keyed_t* b = (const keyed_t*)pb;
    if (a->value < b->value) return -1;
    if (a->value > b->value) return 1;
    return (a->position > b->position) - (a->position < b->position);
}
static int cmp_i64(const void* pa, const void* pb) {
    int64_t a = *(const int64_t*)pa, b = *(const int64_t*)pb;
    return (a > b) - (a < b);
}

/* rank.cpp:38-87 (rank_matrix<float32>): each row replaced in place by 1-based ranks. */
int mco_rank_matrix_f32(float* matrix, int64_t rows, int64_t cols, int ascending) {
    keyed_t* keys = (keyed_t*)malloc((size_t)cols * sizeof(keyed_t));
    if (!keys && cols) return 1;
    for (int64_t r = 0; r < rows; ++r) {
        float* row = matrix + r * cols;
        for (int64_t c = 0; c < cols; ++c) {
            keys[c].value = row[c];
            keys[c].position = c;
        }
        qsort(keys, (size_t)cols, sizeof(keyed_t), ascending ? cmp_asc : cmp_desc);
        for (int64_t c = 0; c < cols; ++c) row[keys[c].position] = (float)(c + 1);
    }
    free(keys);
    return 0;
}

/* rank.cpp:5-36 (rank_rows<float32>): the rank-th smallest element of each row. */
int mco_rank_rows_f32(const float* matrix, int64_t rows, int64_t cols, float* output, int64_t rank) {
    keyed_t* keys = (keyed_t*)malloc((size_t)cols * sizeof(keyed_t));
    if (!keys && cols) return 1;
    for (int64_t r = 0; r < rows; ++r) {
        for (int64_t c = 0; c < cols; ++c) {
            keys[c].value = matrix[r * cols + c];
            keys[c].position = c;
        }
        qsort(keys, (size_t)cols, sizeof(keyed_t), cmp_asc);
        output[r] = keys[rank].value;
    }
    free(keys);
    return 0;
}

/* top_per.cpp:5-92 (collect_top<float32>): per row the `degree` largest, indices ascending. */
int mco_collect_top_f32(int64_t degree, const float* similarity, int64_t rows, int64_t cols, int32_t* out_indices,
                        float* out_data, int ranks) {
    keyed_t* keys = (keyed_t*)malloc((size_t)cols * sizeof(keyed_t));
    int64_t* kept = (int64_t*)malloc((size_t)degree * sizeof(int64_t));
    if (!keys || !kept) return 1;
    for (int64_t r = 0; r < rows; ++r) {
        const float* row = similarity + r * cols;
        for (int64_t c = 0; c < cols; ++c) {
            keys[c].value = row[c];
            keys[c].position = c;
        }
        qsort(keys, (size_t)cols, sizeof(keyed_t), cmp_desc);
        for (int64_t d = 0; d < degree; ++d) kept[d] = keys[d].position;
        qsort(kept, (size_t)degree, sizeof(int64_t), cmp_i64);
        int32_t* row_indices = out_indices + r * degree;
        float* row_data = out_data + r * degree;
        for (int64_t d = 0; d < degree; ++d) row_indices[d] = (int32_t)kept[d];
        if (!ranks) {
            for (int64_t d = 0; d < degree; ++d) row_data[d] = row[kept[d]];
        } else { /* :44-63: ascending rank 1..degree among the kept values */
            for (int64_t d = 0; d < degree; ++d) {
                keys[d].value = row[kept[d]];
                keys[d].position = d;
            }
            qsort(keys, (size_t)degree, sizeof(keyed_t), cmp_asc);
            for (int64_t d = 0; d < degree; ++d) row_data[keys[d].position] = (float)(d + 1);
        }
    }
    free(keys);
    free(kept);
    return 0;
}

/* prune_per.cpp:5-105 (collect_pruned<float32,int32,int32>). */
int mco_collect_pruned_f32_i32_i32(int64_t degree, const float* data, const int32_t* indices, const int32_t* indptr,
                                   int64_t bands, float* out_data, int32_t* out_indices, int32_t* out_indptr) {
    int64_t start = 0, widest = 0;
    out_indptr[0] = 0;
    for (int64_t b = 0; b < bands; ++b) { /* :85-95 serial scan */
        int64_t size = indptr[b + 1] - indptr[b];
        if (size > widest) widest = size;
        start += size <= degree ? size : degree;
        out_indptr[b + 1] = (int32_t)start;
    }
    keyed_t* keys = (keyed_t*)malloc((size_t)(widest ? widest : 1) * sizeof(keyed_t));
    int64_t* kept = (int64_t*)malloc((size_t)(degree ? degree : 1) * sizeof(int64_t));
    if (!keys || !kept) return 1;
    for (int64_t b = 0; b < bands; ++b) {
        int64_t lo = indptr[b], size = indptr[b + 1] - lo, out = out_indptr[b];
        if (size <= degree) {
            memcpy(out_data + out, data + lo, (size_t)size * sizeof(float));
            memcpy(out_indices + out, indices + lo, (size_t)size * sizeof(int32_t));
            continue;
        }
        for (int64_t p = 0; p < size; ++p) {
            keys[p].value = data[lo + p];
            keys[p].position = p;
        }
        qsort(keys, (size_t)size, sizeof(keyed_t), cmp_desc);
        for (int64_t d = 0; d < degree; ++d) kept[d] = keys[d].position;
        qsort(kept, (size_t)degree, sizeof(int64_t), cmp_i64);
        for (int64_t d = 0; d < degree; ++d) {
            out_indices[out + d] = indices[lo + kept[d]];
            out_data[out + d] = data[lo + kept[d]];
        }
    }
    free(keys);
    free(kept);
    return 0;
}

/*
 * relayout.cpp:73-108 (collect_compressed) followed by :152-167 (sort_compressed_indices), as
 * computation.py:251-303 sequences them.  `out_indptr` has bands_out + 1 entries on return.
 * The serial scatter visits input bands in order, so output bands come out already sorted.
 */
int mco_relayout_f32_i32_i32(const float* data, const int32_t* indices, const int32_t* indptr, int64_t bands_in,
                             int64_t bands_out, float* out_data, int32_t* out_indices, int32_t* out_indptr) {
    int64_t nnz = indptr[bands_in];
    int32_t* cursor = (int32_t*)calloc((size_t)bands_out + 1, sizeof(int32_t));
    if (!cursor) return 1;
    for (int64_t e = 0; e < nnz; ++e) ++cursor[indices[e] + 1];
    for (int64_t b = 0; b < bands_out; ++b) cursor[b + 1] += cursor[b];
    memcpy(out_indptr, cursor, ((size_t)bands_out + 1) * sizeof(int32_t));
    for (int64_t b = 0; b < bands_in; ++b)
        for (int64_t e = indptr[b]; e < indptr[b + 1]; ++e) {
            int32_t at = cursor[indices[e]]++;
            out_indices[at] = (int32_t)b;
            out_data[at] = data[e];
        }
    free(cursor);
    return 0;
}
