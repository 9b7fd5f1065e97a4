/*
 * mc_seeds.c -- TEST INFRASTRUCTURE ONLY.
 *
 * Plain-C restatement of the reference's seed chooser (metacells/choose_seeds.cpp of tanaylab/metacells
 * v0.9.5) for use as the parity oracle of metacells_b200's CUDA seed chooser.
 * Only tests/, __graft_entry__.smoke() and the CPU legs of bench.py may load this; nothing under
 * metacells_b200/ does.
 *
 * Pinned (tests/test_oracle_seeds.py) against the UNMODIFIED reference compiled into oracle/_ref/
 * (`choose_seeds`): `==` on every element of seed_of_nodes and on the returned seeds count, and against
 * the golden fixtures that build produced (tests/golden/seeds.npz).
 *
 * Third-party arithmetic the path depends on, restated here.  The chosen seeds depend on
 * WHERE libstdc++'s selection leaves the elements, not only on which value lands at the nth place
 * (choose_seeds.cpp:68-70 reads an element between two nth_element cuts), so the algorithms are
 * restated step for step from libstdc++ 13 (bits/stl_algo.h, bits/stl_heap.h):
 *   - std::nth_element = __introselect: median of (first+1, mid, last-1) swapped to first, Hoare
 *     partition of [first+1, last) around it, depth limit 2 lg n, then __heap_select; ranges of at most 3
 *     end in an insertion sort.
 *   - std::sort = __introsort_loop (same partition, threshold 16, heap sort at the depth limit) followed by
 *     __final_insertion_sort.
 *   - std::minstd_rand, std::uniform_real_distribution<double> (generate_canonical<double,53>: two draws).
 * Compile with -ffp-contract=off (the reference is built in ISO C++ mode, which disables contraction).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

/* std::minstd_rand */
static uint32_t seeds_rng_seed(uint64_t seed) {
    uint32_t x = (uint32_t)(seed % 2147483647u);
    return x == 0 ? 1 : x;
}

static uint32_t seeds_rng_next(uint32_t* state) {
    *state = (uint32_t)(((uint64_t)*state * 48271u) % 2147483647u);
    return *state;
}

/* random.tcc:3349-3381 with minstd_rand, then uniform_real_distribution::operator() */
static double seeds_uniform_real(uint32_t* rng, double a, double b) {
    const double r = 2147483646.0;
    double sum = 0.0;
    double tmp = 1.0;
    for (int k = 2; k != 0; --k) {
        sum += (double)(seeds_rng_next(rng) - 1u) * tmp;
        tmp *= r;
    }
    double ret = sum / tmp;
    if (ret >= 1.0) ret = nextafter(1.0, 0.0);
    return ret * (b - a) + a;
}

/* ---- libstdc++ selection and sorting over an array of values, value comparator ---------------------- */

typedef int (*less_fn)(const void* context, int64_t left, int64_t right);

typedef struct {
    int64_t* a;
    less_fn less;
    const void* context;
} Range;

#define LESS(r, x, y) ((r)->less((r)->context, (x), (y)))

static void swap_at(Range* r, int64_t i, int64_t j) {
    int64_t t = r->a[i];
    r->a[i] = r->a[j];
    r->a[j] = t;
}

static int64_t floor_lg(int64_t n) { return 63 - __builtin_clzll((unsigned long long)n); }

/* stl_algo.h:84-103 */
static void move_median_to_first(Range* r, int64_t result, int64_t x, int64_t y, int64_t z) {
    int64_t* a = r->a;
    if (LESS(r, a[x], a[y])) {
        if (LESS(r, a[y], a[z])) swap_at(r, result, y);
        else if (LESS(r, a[x], a[z])) swap_at(r, result, z);
        else swap_at(r, result, x);
    } else if (LESS(r, a[x], a[z])) swap_at(r, result, x);
    else if (LESS(r, a[y], a[z])) swap_at(r, result, z);
    else swap_at(r, result, y);
}

/* stl_algo.h:1870-1888 */
static int64_t unguarded_partition(Range* r, int64_t first, int64_t last, int64_t pivot) {
    int64_t* a = r->a;
    for (;;) {
        while (LESS(r, a[first], a[pivot])) ++first;
        --last;
        while (LESS(r, a[pivot], a[last])) --last;
        if (!(first < last)) return first;
        swap_at(r, first, last);
        ++first;
    }
}

/* stl_algo.h:1892-1900 */
static int64_t unguarded_partition_pivot(Range* r, int64_t first, int64_t last) {
    int64_t mid = first + (last - first) / 2;
    move_median_to_first(r, first, first + 1, mid, last - 1);
    return unguarded_partition(r, first + 1, last, first);
}

/* stl_heap.h:135-148 (value compared as the right operand) */
static void push_heap_value(Range* r, int64_t first, int64_t hole, int64_t top, int64_t value) {
    int64_t* a = r->a;
    int64_t parent = (hole - 1) / 2;
    while (hole > top && LESS(r, a[first + parent], value)) {
        a[first + hole] = a[first + parent];
        hole = parent;
        parent = (hole - 1) / 2;
    }
    a[first + hole] = value;
}

/* stl_heap.h:224-249 */
static void adjust_heap(Range* r, int64_t first, int64_t hole, int64_t len, int64_t value) {
    int64_t* a = r->a;
    const int64_t top = hole;
    int64_t child = hole;
    while (child < (len - 1) / 2) {
        child = 2 * (child + 1);
        if (LESS(r, a[first + child], a[first + child - 1])) child--;
        a[first + hole] = a[first + child];
        hole = child;
    }
    if ((len & 1) == 0 && child == (len - 2) / 2) {
        child = 2 * (child + 1);
        a[first + hole] = a[first + child - 1];
        hole = child - 1;
    }
    push_heap_value(r, first, hole, top, value);
}

/* stl_heap.h:254-266 */
static void pop_heap_to(Range* r, int64_t first, int64_t last, int64_t result) {
    int64_t value = r->a[result];
    r->a[result] = r->a[first];
    adjust_heap(r, first, 0, last - first, value);
}

/* stl_heap.h:340-361 */
static void make_heap_range(Range* r, int64_t first, int64_t last) {
    if (last - first < 2) return;
    const int64_t len = last - first;
    int64_t parent = (len - 2) / 2;
    for (;;) {
        adjust_heap(r, first, parent, len, r->a[first + parent]);
        if (parent == 0) return;
        parent--;
    }
}

/* stl_algo.h:1631-1640 */
static void heap_select(Range* r, int64_t first, int64_t middle, int64_t last) {
    make_heap_range(r, first, middle);
    for (int64_t i = middle; i < last; ++i)
        if (LESS(r, r->a[i], r->a[first])) pop_heap_to(r, first, middle, i);
}

/* stl_heap.h:419-428 */
static void sort_heap_range(Range* r, int64_t first, int64_t last) {
    while (last - first > 1) {
        --last;
        pop_heap_to(r, first, last, last);
    }
}

/* stl_algo.h:1792-1807 */
static void unguarded_linear_insert(Range* r, int64_t last) {
    int64_t* a = r->a;
    int64_t value = a[last];
    int64_t next = last - 1;
    while (LESS(r, value, a[next])) {
        a[last] = a[next];
        last = next;
        --next;
    }
    a[last] = value;
}

/* stl_algo.h:1812-1831 */
static void insertion_sort(Range* r, int64_t first, int64_t last) {
    int64_t* a = r->a;
    if (first == last) return;
    for (int64_t i = first + 1; i != last; ++i) {
        if (LESS(r, a[i], a[first])) {
            int64_t value = a[i];
            memmove(a + first + 1, a + first, (size_t)(i - first) * sizeof(int64_t));
            a[first] = value;
        } else {
            unguarded_linear_insert(r, i);
        }
    }
}

/* stl_algo.h:1957-1979 and 4775-4791 */
static void nth_element_range(Range* r, int64_t first, int64_t nth, int64_t last) {
    if (first == last || nth == last) return;
    int64_t depth_limit = floor_lg(last - first) * 2;
    while (last - first > 3) {
        if (depth_limit == 0) {
            heap_select(r, first, nth + 1, last);
            swap_at(r, first, nth);
            return;
        }
        --depth_limit;
        int64_t cut = unguarded_partition_pivot(r, first, last);
        if (cut <= nth) first = cut;
        else last = cut;
    }
    insertion_sort(r, first, last);
}

/* stl_algo.h:1918-1936 */
static void introsort_loop(Range* r, int64_t first, int64_t last, int64_t depth_limit) {
    while (last - first > 16) {
        if (depth_limit == 0) {
            heap_select(r, first, last, last);
            sort_heap_range(r, first, last);
            return;
        }
        --depth_limit;
        int64_t cut = unguarded_partition_pivot(r, first, last);
        introsort_loop(r, cut, last, depth_limit);
        last = cut;
    }
}

/* stl_algo.h:1940-1952, 1854-1865 */
static void sort_range(Range* r, int64_t first, int64_t last) {
    if (first == last) return;
    introsort_loop(r, first, last, floor_lg(last - first) * 2);
    if (last - first > 16) {
        insertion_sort(r, first, first + 16);
        for (int64_t i = first + 16; i != last; ++i) unguarded_linear_insert(r, i);
    } else {
        insertion_sort(r, first, last);
    }
}

/* exported for pinning the two algorithms on their own against std:: (tests build a tiny C++ driver) and
 * for checking the device restatement: keys are compared, positions are permuted. */
static int less_by_key(const void* context, int64_t left, int64_t right) {
    const int64_t* keys = (const int64_t*)context;
    return keys[left] < keys[right];
}

void mco_nth_element_positions(int64_t* positions, const int64_t* keys, int64_t first, int64_t nth, int64_t last) {
    Range r = {positions, less_by_key, keys};
    nth_element_range(&r, first, nth, last);
}

void mco_sort_positions(int64_t* positions, const int64_t* keys, int64_t count) {
    Range r = {positions, less_by_key, keys};
    sort_range(&r, 0, count);
}

/* ---- choose_seeds.cpp ------------------------------------------------------------------------------- */

typedef struct {
    const float* out_w;
    const int32_t* out_idx;
    const int32_t* out_ptr;
    const float* in_w;
    const int32_t* in_idx;
    const int32_t* in_ptr;
    int64_t nodes;
    int32_t* seed_of_nodes;
    int32_t* connected;      /* connected_nodes[node] lives at connected + in_ptr[node] */
    int64_t* connected_size; /* its current size */
    int64_t* candidates;
    int64_t candidates_size;
    int64_t* positions;
} Seeds;

/* choose_seeds.cpp:76-86; returns 0 when the value is missing (the reference exits) */
static int remove_sorted_i32(int32_t* values, int64_t* size, int32_t value) {
    int64_t low = 0, high = *size;
    while (low < high) {
        int64_t mid = low + (high - low) / 2;
        if (values[mid] < value) low = mid + 1;
        else high = mid;
    }
    if (low == *size || values[low] != value) return 0;
    memmove(values + low, values + low + 1, (size_t)(*size - low - 1) * sizeof(int32_t));
    --*size;
    return 1;
}

static int remove_sorted_i64(int64_t* values, int64_t* size, int64_t value) {
    int64_t low = 0, high = *size;
    while (low < high) {
        int64_t mid = low + (high - low) / 2;
        if (values[mid] < value) low = mid + 1;
        else high = mid;
    }
    if (low == *size || values[low] != value) return 0;
    memmove(values + low, values + low + 1, (size_t)(*size - low - 1) * sizeof(int64_t));
    --*size;
    return 1;
}

/* choose_seeds.cpp:46-66: positions ordered by the connected size of their candidate */
static int less_connected_size(const void* context, int64_t left, int64_t right) {
    const Seeds* s = (const Seeds*)context;
    return s->connected_size[s->candidates[left]] < s->connected_size[s->candidates[right]];
}

/* choose_seeds.cpp:28-74 */
static int64_t choose_seed_node(Seeds* s, float min_quantile, float max_quantile, uint32_t* rng, int* failed) {
    const int64_t size = s->candidates_size;
    for (int64_t i = 0; i < size; ++i) s->positions[i] = i;
    /* (size - 1) * quantile: size_t times float32 is a float32 product */
    const int64_t min_rank = (int64_t)floorf((float)(uint64_t)(size - 1) * min_quantile);
    const int64_t max_rank = (int64_t)ceilf((float)(uint64_t)(size - 1) * max_quantile);
    if (!(0 <= min_rank && min_rank <= max_rank && max_rank <= size - 1)) {
        *failed = 1;
        return 0;
    }
    Range r = {s->positions, less_connected_size, s};
    nth_element_range(&r, 0, min_rank, size);
    nth_element_range(&r, min_rank, max_rank, size);
    const int64_t selected = (int64_t)(seeds_rng_next(rng) % (uint64_t)(max_rank + 1 - min_rank));
    return s->candidates[s->positions[min_rank + selected]];
}

/* choose_seeds.cpp:113-117: positions of the seed's incoming band, heavier edges first */
static int heavier_edge(const void* context, int64_t left, int64_t right) {
    const float* weights = (const float*)context;
    return weights[left] > weights[right];
}

/* choose_seeds.cpp:88-151 */
static int store_seed_node(Seeds* s, int64_t seed_index, int64_t seed_node, int64_t seed_size) {
    if (!remove_sorted_i64(s->candidates, &s->candidates_size, seed_node)) return 0;
    const int32_t* incoming_nodes = s->in_idx + s->in_ptr[seed_node];
    const float* incoming_weights = s->in_w + s->in_ptr[seed_node];
    const int64_t incoming_count = s->in_ptr[seed_node + 1] - s->in_ptr[seed_node];
    int64_t kept = 0;
    for (int64_t position = 0; position < incoming_count; ++position)
        if (s->seed_of_nodes[incoming_nodes[position]] < 0) s->positions[kept++] = position;
    if (seed_size < kept) {
        Range r = {s->positions, heavier_edge, incoming_weights};
        nth_element_range(&r, 0, seed_size, kept);
        kept = seed_size;
    }
    if (!(s->seed_of_nodes[seed_node] < 0)) return 0;
    s->seed_of_nodes[seed_node] = (int32_t)seed_index;
    s->connected_size[seed_node] = 0;
    for (int64_t i = 0; i < kept; ++i) {
        const int32_t node = incoming_nodes[s->positions[i]];
        s->connected_size[node] = 0;
        s->seed_of_nodes[node] = (int32_t)seed_index;
    }
    for (int64_t at = s->out_ptr[seed_node]; at < s->out_ptr[seed_node + 1]; ++at) {
        const int32_t other = s->out_idx[at];
        if (s->seed_of_nodes[other] < 0 &&
            !remove_sorted_i32(s->connected + s->in_ptr[other], &s->connected_size[other], (int32_t)seed_node))
            return 0;
    }
    for (int64_t i = 0; i < kept; ++i) {
        const int32_t node = incoming_nodes[s->positions[i]];
        for (int64_t at = s->out_ptr[node]; at < s->out_ptr[node + 1]; ++at) {
            const int32_t other = s->out_idx[at];
            if (s->seed_of_nodes[other] < 0 &&
                !remove_sorted_i32(s->connected + s->in_ptr[other], &s->connected_size[other], node))
                return 0;
        }
    }
    return 1;
}

/* choose_seeds.cpp:154-162 */
static int keep_large_candidates(Seeds* s) {
    int64_t kept = 0;
    for (int64_t i = 0; i < s->candidates_size; ++i)
        if (s->connected_size[s->candidates[i]] != 0) s->candidates[kept++] = s->candidates[i];
    s->candidates_size = kept;
    return kept > 0;
}

/* choose_seeds.cpp:346-356: more incoming x outgoing edges first */
static int better_connected(const void* context, int64_t left, int64_t right) {
    const Seeds* s = (const Seeds*)context;
    const uint64_t left_in = (uint64_t)(s->in_ptr[left + 1] - s->in_ptr[left]);
    const uint64_t right_in = (uint64_t)(s->in_ptr[right + 1] - s->in_ptr[right]);
    const uint64_t left_out = (uint64_t)(s->out_ptr[left + 1] - s->out_ptr[left]);
    const uint64_t right_out = (uint64_t)(s->out_ptr[right + 1] - s->out_ptr[right]);
    return (left_in + 1) * (left_out + 1) > (right_in + 1) * (right_out + 1);
}

/* choose_seeds.cpp:164-228; 1 connected, 0 not yet, -1 the reference's final assert */
static int connect_node(const Seeds* s, int64_t node, int32_t* seed_sizes, float* seed_weights, int64_t seeds_count,
                        uint32_t* rng) {
    if (s->seed_of_nodes[node] >= 0) return 1;
    for (int64_t i = 0; i < seeds_count; ++i) seed_weights[i] = 0.0f;
    double total = 0.0;
    for (int64_t at = s->in_ptr[node]; at < s->in_ptr[node + 1]; ++at) {
        const int32_t seed = s->seed_of_nodes[s->in_idx[at]];
        if (seed >= 0) {
            const float weight = s->in_w[at] / (float)seed_sizes[seed];
            seed_weights[seed] += weight;
            total += weight;
        }
    }
    for (int64_t at = s->out_ptr[node]; at < s->out_ptr[node + 1]; ++at) {
        const int32_t seed = s->seed_of_nodes[s->out_idx[at]];
        if (seed >= 0) {
            const float weight = s->out_w[at] / (float)seed_sizes[seed];
            seed_weights[seed] += weight;
            total += weight;
        }
    }
    if (total == 0.0) return 0;
    double weight = seeds_uniform_real(rng, 0.0, total);
    if (!(0 <= weight && weight <= total)) return -1;
    for (int64_t seed = 0; seed < seeds_count; ++seed) {
        weight -= seed_weights[seed];
        if (weight <= 0.0) {
            s->seed_of_nodes[node] = (int32_t)seed;
            seed_sizes[seed] += 1;
            return 1;
        }
    }
    return -1;
}

/* choose_seeds.cpp:230-262 */
static int complete_seeds(const Seeds* s, int64_t seeds_count, uint32_t* rng) {
    const int64_t nodes = s->nodes;
    float* seed_weights = (float*)calloc((size_t)seeds_count + 1, sizeof(float));
    int32_t* seed_sizes = (int32_t*)calloc((size_t)seeds_count + 1, sizeof(int32_t));
    int64_t* old_list = (int64_t*)malloc((size_t)nodes * sizeof(int64_t));
    int64_t* new_list = (int64_t*)malloc((size_t)nodes * sizeof(int64_t));
    int status = 0;
    if (!seed_weights || !seed_sizes || !old_list || !new_list) {
        status = 1;
        goto done;
    }
    for (int64_t node = 0; node < nodes; ++node)
        if (s->seed_of_nodes[node] >= 0) seed_sizes[s->seed_of_nodes[node]] += 1;
    int64_t new_size = 0;
    for (int64_t node = 0; node < nodes; ++node) {
        const int connected = connect_node(s, node, seed_sizes, seed_weights, seeds_count, rng);
        if (connected < 0) {
            status = 2;
            goto done;
        }
        if (!connected) new_list[new_size++] = node;
    }
    while (new_size > 0) {
        int64_t* swap = old_list;
        old_list = new_list;
        new_list = swap;
        const int64_t old_size = new_size;
        new_size = 0;
        for (int64_t i = 0; i < old_size; ++i) {
            const int connected = connect_node(s, old_list[i], seed_sizes, seed_weights, seeds_count, rng);
            if (connected < 0) {
                status = 2;
                goto done;
            }
            if (!connected) new_list[new_size++] = old_list[i];
        }
        if (!(new_size < old_size)) { /* the reference asserts progress */
            status = 2;
            goto done;
        }
    }
done:
    free(seed_weights);
    free(seed_sizes);
    free(old_list);
    free(new_list);
    return status;
}

/* choose_seeds.cpp:264-372.  0 ok, 1 allocation failure, 2 a FastAssert of the reference fails. */
int mco_choose_seeds(const float* out_w, const int32_t* out_idx, const int32_t* out_ptr, const float* in_w,
                     const int32_t* in_idx, const int32_t* in_ptr, int64_t nodes, uint64_t random_seed,
                     int64_t max_seeds_count, float min_seed_size_quantile, float max_seed_size_quantile,
                     int32_t* seed_of_nodes, int64_t* seeds_count_out) {
    if (nodes <= 0) return 2;
    if (!(0 <= min_seed_size_quantile && min_seed_size_quantile <= max_seed_size_quantile &&
          max_seed_size_quantile <= 1.0))
        return 2;
    Seeds s;
    memset(&s, 0, sizeof(s));
    s.out_w = out_w, s.out_idx = out_idx, s.out_ptr = out_ptr;
    s.in_w = in_w, s.in_idx = in_idx, s.in_ptr = in_ptr;
    s.nodes = nodes;
    s.seed_of_nodes = seed_of_nodes;

    int32_t largest = seed_of_nodes[0];
    for (int64_t node = 1; node < nodes; ++node)
        if (seed_of_nodes[node] > largest) largest = seed_of_nodes[node];
    const int64_t given_seeds_count = (int64_t)largest + 1;
    int64_t seeds_count = given_seeds_count;
    uint32_t rng = seeds_rng_seed(random_seed);
    int status = 0;

    if (given_seeds_count < max_seeds_count) {
        s.connected = (int32_t*)malloc(((size_t)in_ptr[nodes] + 1) * sizeof(int32_t));
        s.connected_size = (int64_t*)calloc((size_t)nodes, sizeof(int64_t));
        s.candidates = (int64_t*)malloc((size_t)nodes * sizeof(int64_t));
        s.positions = (int64_t*)malloc((size_t)nodes * sizeof(int64_t));
        if (!s.connected || !s.connected_size || !s.candidates || !s.positions) {
            status = 1;
            goto done;
        }
        /* collect_connected_nodes, choose_seeds.cpp:5-26 */
        for (int64_t node = 0; node < nodes; ++node) {
            if (seed_of_nodes[node] >= 0) continue;
            int32_t* connected = s.connected + in_ptr[node];
            int64_t count = 0;
            for (int64_t at = in_ptr[node]; at < in_ptr[node + 1]; ++at)
                if (seed_of_nodes[in_idx[at]] < 0) connected[count++] = in_idx[at];
            s.connected_size[node] = count;
        }
        for (int64_t node = 0; node < nodes; ++node)
            if (seed_of_nodes[node] < 0) s.candidates[s.candidates_size++] = node;
        if (!(s.candidates_size >= max_seeds_count - given_seeds_count)) {
            status = 2;
            goto done;
        }
        const int64_t mean_seed_size = s.candidates_size / (max_seeds_count - given_seeds_count);
        if (mean_seed_size < 1) {
            status = 2;
            goto done;
        }
        while (seeds_count < max_seeds_count && keep_large_candidates(&s)) {
            int failed = 0;
            const int64_t seed_node = choose_seed_node(&s, min_seed_size_quantile, max_seed_size_quantile, &rng, &failed);
            if (failed || !store_seed_node(&s, seeds_count, seed_node, mean_seed_size)) {
                status = 2;
                goto done;
            }
            ++seeds_count;
        }
        if (seeds_count < max_seeds_count) {
            s.candidates_size = 0;
            for (int64_t node = 0; node < nodes; ++node)
                if (seed_of_nodes[node] < 0) s.candidates[s.candidates_size++] = node;
            Range r = {s.candidates, better_connected, &s};
            sort_range(&r, 0, s.candidates_size);
            while (s.candidates_size > 0 && seeds_count < max_seeds_count) {
                const int64_t node = s.candidates[--s.candidates_size];
                seed_of_nodes[node] = (int32_t)seeds_count;
                ++seeds_count;
            }
        }
    }
    if (!(seeds_count <= max_seeds_count) && given_seeds_count < max_seeds_count) {
        status = 2;
        goto done;
    }
    status = complete_seeds(&s, seeds_count, &rng);
    *seeds_count_out = seeds_count;
done:
    free(s.connected);
    free(s.connected_size);
    free(s.candidates);
    free(s.positions);
    return status;
}
