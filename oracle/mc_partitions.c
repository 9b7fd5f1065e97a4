/*
 * mc_partitions.c -- TEST INFRASTRUCTURE ONLY.
 *
 * Plain-C restatement of the reference's candidate-metacell partition optimiser
 * (metacells/partitions.cpp of tanaylab/metacells v0.9.5) for use as the
 * parity oracle of metacells_b200's CUDA optimiser.  Only tests/, __graft_entry__.smoke() and the CPU
 * legs of bench.py may load this; nothing under metacells_b200/ does.
 *
 * Pinned (tests/test_oracle_partitions.py) against the UNMODIFIED reference compiled into
 * oracle/_ref/ (`optimize_partitions`, `score_partitions`), `==` on the float64 score and on every
 * element of partition_of_nodes.
 *
 * Third-party arithmetic the path depends on, restated here:
 *   - std::minstd_rand                      : x <- 48271 x mod 2^31-1, seed 0 -> 1.
 *   - std::shuffle (libstdc++ 13 stl_algo.h:3742-3804) and its uniform_int_distribution fallback
 *     branch (uniform_int_dist.h:284-345): pairs of swaps from one draw, rejection sampling.
 *   - std::uniform_real_distribution<double> = generate_canonical<double,53> (random.tcc:3349-3381):
 *     two draws, sum = (x1-1) + (x2-1) r, ret = sum / r^2 with r = 2147483646.
 *   - log2(x) := log(x) * (1.0 / log(2.0)) (extensions.h:131-138) with glibc's double `log`; this
 *     file calls libm's `log` itself (that IS the reference's behaviour on the CPU).
 * Compile with -ffp-contract=off (the reference is built in ISO C++ mode, which disables contraction).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define EPSILON 1e-6 /* extensions.h:131 */

static double g_log2_scale = 0.0;

static double mc_log2(double x) { /* extensions.h:135-138 */
    return log(x) * g_log2_scale;
}

/* partitions.cpp:5-19 */
static double mass_factor(double mass, double low_mass, double target_mass, double high_mass) {
    const double inner_slope = 0.02;
    const double outer_slope = 0.5;
    if (mass < low_mass) {
        return mc_log2((1 - inner_slope) * low_mass / (low_mass + (1 / (1 - outer_slope) - 1) * (low_mass - mass)));
    }
    if (mass < target_mass) {
        return mc_log2(1 - inner_slope * (target_mass - mass) / (target_mass - low_mass));
    }
    if (mass < high_mass) {
        return mc_log2(1 - inner_slope * (mass - target_mass) / (high_mass - target_mass));
    }
    return mc_log2((1 - inner_slope) * low_mass / (low_mass + (1 / (1 - outer_slope) - 1) * (mass - high_mass)));
}

/* partitions.cpp:39-80 */
typedef struct {
    double outgoing;
    double incoming;
    double score;
} NodeScore;

static void update_outgoing(NodeScore* s, int direction, double edge_weight) {
    s->outgoing += direction * edge_weight;
    s->outgoing = s->outgoing < 0.0 ? 0.0 : s->outgoing; /* std::max(x, 0.0) */
}

static void update_incoming(NodeScore* s, int direction, double edge_weight) {
    s->incoming += direction * edge_weight;
    s->incoming = s->incoming < 0.0 ? 0.0 : s->incoming;
}

static int connectivity(const NodeScore* s) { return (int)(s->outgoing > EPSILON) + (int)(s->incoming > 0); }

static double rescore(NodeScore* s) {
    s->score = mc_log2(EPSILON + s->outgoing * s->incoming) / 2.0;
    return s->score;
}

/* std::minstd_rand */
static uint32_t rng_seed(uint64_t seed) {
    uint32_t x = (uint32_t)(seed % 2147483647u);
    return x == 0 ? 1 : x;
}

static uint32_t rng_next(uint32_t* state) {
    *state = (uint32_t)(((uint64_t)*state * 48271u) % 2147483647u);
    return *state;
}

/* uniform_int_dist.h:284-345, fallback branch (generator range 2147483645 is not 2^32-1 or 2^64-1). */
static uint64_t uniform_int(uint32_t* rng, uint64_t low, uint64_t high) {
    const uint64_t urngrange = 2147483646u - 1u;
    const uint64_t urange = high - low;
    uint64_t ret;
    if (urngrange > urange) {
        const uint64_t uerange = urange + 1;
        const uint64_t scaling = urngrange / uerange;
        const uint64_t past = uerange * scaling;
        do {
            ret = (uint64_t)rng_next(rng) - 1u;
        } while (ret >= past);
        ret /= scaling;
    } else if (urngrange < urange) {
        /* upscaling (uniform_int_dist.h:347-371); never reached for fewer than 2^31 nodes. */
        uint64_t tmp;
        do {
            const uint64_t uerngrange = urngrange + 1;
            tmp = uerngrange * uniform_int(rng, 0, urange / uerngrange);
            ret = tmp + ((uint64_t)rng_next(rng) - 1u);
        } while (ret > urange || ret < tmp);
    } else {
        ret = (uint64_t)rng_next(rng) - 1u;
    }
    return ret + low;
}

/* stl_algo.h:3742-3804 */
static void shuffle_indices(int64_t* a, int64_t n, uint32_t* rng) {
    if (n == 0) return;
    const uint64_t urngrange = 2147483646u - 1u;
    const uint64_t urange = (uint64_t)n;
    int64_t tmp;
    if (urngrange / urange >= urange) {
        int64_t i = 1;
        if ((urange % 2) == 0) {
            const uint64_t j = uniform_int(rng, 0, 1);
            tmp = a[i]; a[i] = a[j]; a[j] = tmp;
            ++i;
        }
        while (i != n) {
            const uint64_t swap_range = (uint64_t)i + 1;
            const uint64_t x = uniform_int(rng, 0, swap_range * (swap_range + 1) - 1);
            const uint64_t first = x / (swap_range + 1);
            const uint64_t second = x % (swap_range + 1);
            tmp = a[i]; a[i] = a[first]; a[first] = tmp;
            ++i;
            tmp = a[i]; a[i] = a[second]; a[second] = tmp;
            ++i;
        }
        return;
    }
    for (int64_t i = 1; i != n; ++i) {
        const uint64_t j = uniform_int(rng, 0, (uint64_t)i);
        tmp = a[i]; a[i] = a[j]; a[j] = tmp;
    }
}

/* random.tcc:3349-3381 with minstd_rand, then random.h uniform_real_distribution::operator(). */
static double uniform_real(uint32_t* rng, double a, double b) {
    const double r = 2147483646.0;
    double sum = 0.0;
    double tmp = 1.0;
    for (int k = 2; k != 0; --k) {
        sum += (double)(rng_next(rng) - 1u) * tmp;
        tmp *= r;
    }
    double ret = sum / tmp;
    if (ret >= 1.0) ret = nextafter(1.0, 0.0);
    return ret * (b - a) + a;
}

void mco_log_array(const double* x, double* y, int64_t n) { /* libm `log`, for pinning the device replica */
    for (int64_t i = 0; i < n; ++i) y[i] = log(x[i]);
}

void mco_shuffle(int64_t* a, int64_t n, uint32_t seed, int64_t rounds) { /* for pinning std::shuffle */
    uint32_t rng = rng_seed(seed);
    for (int64_t r = 0; r < rounds; ++r) shuffle_indices(a, n, &rng);
}

/* partitions.cpp:243-320: the optimiser state.  score_of_nodes is [partition][node], as there. */
typedef struct {
    const float* out_w; const int32_t* out_idx; const int32_t* out_ptr;
    const float* in_w; const int32_t* in_idx; const int32_t* in_ptr;
    int64_t nodes;
    const float* mass_of_nodes;
    double low_mass, target_mass, high_mass, low_size, target_size, high_size;
    int32_t* partition_of_nodes;
    const int8_t* hot_mask;
    double* temperature_of_nodes;
    int64_t* size_of_partitions;
    int64_t partitions;
    double incoming_scale;
    NodeScore* scores; /* [partitions][nodes] */
    double* mass_of_partitions;
    double* score_of_partitions;
    /* scratch of optimize() */
    double *total_diffs, *cold_diffs, *hot_diffs, *connectivity_diffs, *disconnected_diffs;
    int64_t* candidates; /* the three candidate lists alias ONE vector (partitions.cpp:26-29 + extensions.h:483-486) */
    int64_t candidates_count;
} Optimizer;

#define SCORE(o, p, n) ((o)->scores[(int64_t)(p) * (o)->nodes + (n)])

static void optimizer_free(Optimizer* o) {
    free(o->temperature_of_nodes); free(o->size_of_partitions); free(o->scores); free(o->mass_of_partitions);
    free(o->score_of_partitions); free(o->total_diffs); free(o->cold_diffs); free(o->hot_diffs);
    free(o->connectivity_diffs); free(o->disconnected_diffs); free(o->candidates);
}

/* Returns 0, or 1 = allocation failure, 2 = a failed FastAssert of the reference (which would exit(1)). */
static int optimizer_init(Optimizer* o) {
    const int64_t n = o->nodes;
    /* initial_size_of_partitions :88-105 */
    int64_t parts = 0;
    for (int64_t i = 0; i < n; ++i) {
        if (o->partition_of_nodes[i] < 0) return 2;
        if (o->partition_of_nodes[i] + 1 > parts) parts = o->partition_of_nodes[i] + 1;
    }
    o->partitions = parts;
    o->size_of_partitions = (int64_t*)calloc((size_t)(parts ? parts : 1), sizeof(int64_t));
    o->temperature_of_nodes = (double*)malloc(sizeof(double) * (size_t)(n ? n : 1));
    o->scores = (NodeScore*)malloc(sizeof(NodeScore) * (size_t)(parts * n > 0 ? parts * n : 1));
    o->mass_of_partitions = (double*)calloc((size_t)(parts ? parts : 1), sizeof(double));
    o->score_of_partitions = (double*)calloc((size_t)(parts ? parts : 1), sizeof(double));
    o->total_diffs = (double*)calloc((size_t)(parts ? parts : 1), sizeof(double));
    o->cold_diffs = (double*)calloc((size_t)(parts ? parts : 1), sizeof(double));
    o->hot_diffs = (double*)calloc((size_t)(parts ? parts : 1), sizeof(double));
    o->connectivity_diffs = (double*)calloc((size_t)(parts ? parts : 1), sizeof(double));
    o->disconnected_diffs = (double*)calloc((size_t)(parts ? parts : 1), sizeof(double));
    o->candidates = (int64_t*)calloc((size_t)(parts ? parts : 1), sizeof(int64_t));
    if (!o->size_of_partitions || !o->temperature_of_nodes || !o->scores || !o->mass_of_partitions ||
        !o->score_of_partitions || !o->total_diffs || !o->cold_diffs || !o->hot_diffs || !o->connectivity_diffs ||
        !o->disconnected_diffs || !o->candidates)
        return 1;
    for (int64_t i = 0; i < n; ++i) {
        o->temperature_of_nodes[i] = 1.0;
        ++o->size_of_partitions[o->partition_of_nodes[i]];
    }
    for (int64_t p = 0; p < parts; ++p)
        if (o->size_of_partitions[p] <= 0) return 2;

    /* initial_incoming_scale :107-128 */
    o->incoming_scale = 0;
    for (int64_t node = 0; node < n; ++node) {
        double total = 0;
        for (int32_t e = o->in_ptr[node]; e < o->in_ptr[node + 1]; ++e) total += o->in_w[e];
        if (!(total > 0)) return 2;
        o->incoming_scale += mc_log2(total);
    }

    /* initial_score_of_nodes_of_partitions :130-206 */
    const double default_score = mc_log2(EPSILON) / 2.0;
    for (int64_t i = 0; i < parts * n; ++i) {
        o->scores[i].outgoing = 0;
        o->scores[i].incoming = 0;
        o->scores[i].score = default_score;
    }
    for (int64_t node = 0; node < n; ++node) {
        const int32_t partition = o->partition_of_nodes[node];
        double total_outgoing = 0;
        for (int32_t e = o->out_ptr[node]; e < o->out_ptr[node + 1]; ++e) {
            const int32_t other = o->out_idx[e];
            const float weight = o->out_w[e];
            total_outgoing += weight;
            const int32_t other_partition = o->partition_of_nodes[other];
            update_outgoing(&SCORE(o, other_partition, node), +1, weight);
            rescore(&SCORE(o, other_partition, node));
            update_incoming(&SCORE(o, partition, other), +1, weight);
            rescore(&SCORE(o, partition, other));
        }
        if (!(total_outgoing >= 1 - EPSILON) || !(total_outgoing < 1 + EPSILON)) return 2;
    }

    /* initial_score_of_partitions :208-221, initial_mass_of_partitions :223-236 */
    for (int64_t node = 0; node < n; ++node) {
        const int32_t partition = o->partition_of_nodes[node];
        o->score_of_partitions[partition] += SCORE(o, partition, node).score;
    }
    for (int64_t node = 0; node < n; ++node) o->mass_of_partitions[o->partition_of_nodes[node]] += o->mass_of_nodes[node];

    if (!(0.0 < o->low_mass) || !(o->low_mass < o->target_mass) || !(o->target_mass < o->high_mass) ||
        !(0.0 < o->low_size) || !(o->low_size < o->target_size) || !(o->target_size < o->high_size))
        return 2;
    return 0;
}

/* partitions.cpp:322-339 */
static double optimizer_score(const Optimizer* o, int with_orphans) {
    const int64_t n = o->nodes;
    double total_score = n * mc_log2((double)n) - o->incoming_scale;
    int64_t orphans = n;
    for (int64_t p = 0; p < o->partitions; ++p) {
        const double score_of_partition = o->score_of_partitions[p];
        const double mass_of_partition = o->mass_of_partitions[p];
        const int64_t size_of_partition = o->size_of_partitions[p];
        total_score += score_of_partition - size_of_partition * mc_log2((double)size_of_partition);
        total_score += size_of_partition * mass_factor(mass_of_partition, o->low_mass, o->target_mass, o->high_mass);
        orphans -= size_of_partition;
    }
    if (with_orphans) {
        total_score += orphans * (mc_log2(EPSILON) / 2.0);
        return total_score / n;
    }
    return total_score / (n - orphans);
}

/* partitions.cpp:647-662 */
static void collect_initial_partition_diffs(Optimizer* o, int64_t node, int32_t current) {
    for (int64_t p = 0; p < o->partitions; ++p) {
        const int direction = 1 - 2 * ((int32_t)p == current);
        const double score = direction * SCORE(o, p, node).score;
        o->cold_diffs[p] = score;
        o->hot_diffs[p] = score;
    }
}

/* partitions.cpp:664-803 */
static void collect_cold_partition_diffs(Optimizer* o, int64_t node, int32_t current) {
    const int32_t* out_idx = o->out_idx + o->out_ptr[node];
    const int32_t* in_idx = o->in_idx + o->in_ptr[node];
    const float* out_w = o->out_w + o->out_ptr[node];
    const float* in_w = o->in_w + o->in_ptr[node];
    const int64_t out_count = o->out_ptr[node + 1] - o->out_ptr[node];
    const int64_t in_count = o->in_ptr[node + 1] - o->in_ptr[node];
    int64_t out_pos = 0, in_pos = 0;
    int32_t out_node = out_idx[0], in_node = in_idx[0];
    float out_weight = out_w[0], in_weight = in_w[0];

    const int current_connectivity = connectivity(&SCORE(o, current, node));
    for (int64_t p = 0; p < o->partitions; ++p) {
        if (p == (int64_t)current) {
            o->connectivity_diffs[p] = 0.0;
            o->disconnected_diffs[p] = 0.0;
            continue;
        }
        const int partition_connectivity = connectivity(&SCORE(o, p, node));
        o->connectivity_diffs[p] = partition_connectivity - current_connectivity;
        if (current_connectivity > 0 && partition_connectivity == 0) {
            o->disconnected_diffs[p] = 1;
        } else if (current_connectivity == 0 && partition_connectivity > 0) {
            o->disconnected_diffs[p] = -1;
        } else {
            o->disconnected_diffs[p] = 0.0;
        }
    }

    while (out_pos < out_count || in_pos < in_count) {
        const int32_t other = out_node < in_node ? out_node : in_node;
        const int32_t other_partition = o->partition_of_nodes[other];
        const int is_outgoing = (int)(out_node == other);
        const int is_incoming = (int)(in_node == other);

        NodeScore other_score = SCORE(o, other_partition, other);
        const double old_score = other_score.score;
        const int old_connectivity = connectivity(&other_score);
        const int direction = 1 - 2 * (other_partition == current);

        update_incoming(&other_score, direction * is_outgoing, out_weight);
        update_outgoing(&other_score, direction * is_incoming, in_weight);
        const double new_score = rescore(&other_score);
        const int new_connectivity = connectivity(&other_score);

        o->cold_diffs[other_partition] += new_score - old_score;
        o->connectivity_diffs[other_partition] += new_connectivity - old_connectivity;
        if (old_connectivity == 0 && new_connectivity > 0) o->disconnected_diffs[other_partition] -= 1;
        if (old_connectivity > 0 && new_connectivity == 0) o->disconnected_diffs[other_partition] += 1;

        out_pos += is_outgoing;
        in_pos += is_incoming;
        if (out_pos < out_count) { out_node = out_idx[out_pos]; out_weight = out_w[out_pos]; }
        else { out_node = (int32_t)o->nodes; out_weight = 0; }
        if (in_pos < in_count) { in_node = in_idx[in_pos]; in_weight = in_w[in_pos]; }
        else { in_node = (int32_t)o->nodes; in_weight = 0; }
    }
}

/* partitions.cpp:805-1005; every push_back there lands in the one aliased vector, in partition order. */
static double collect_candidate_partitions(Optimizer* o, int64_t node, int32_t current, double temperature) {
    const float node_mass = o->mass_of_nodes[node];
    const double current_hot_diff = o->hot_diffs[current];
    const double current_cold_diff = o->cold_diffs[current];
    const int current_connectivity = connectivity(&SCORE(o, current, node));
    const int min_connectivity_diff = current_connectivity == 0 ? 1 : 0;
    double current_adjusted_cold_diff;
    {
        const int64_t old_size = o->size_of_partitions[current];
        const double old_mass = o->mass_of_partitions[current];
        const double old_mass_factor = mass_factor(old_mass, o->low_mass, o->target_mass, o->high_mass);
        const double old_size_factor = mass_factor((double)old_size, o->low_size, o->target_size, o->high_size);
        const double old_score = o->score_of_partitions[current];
        double old_adjusted = old_score - old_size * mc_log2((double)old_size);
        old_adjusted += old_size * (old_size_factor < old_mass_factor ? old_size_factor : old_mass_factor);

        const int64_t new_size = old_size - 1;
        const double new_mass = old_mass - node_mass;
        const double new_mass_factor = mass_factor(new_mass, o->low_mass, o->target_mass, o->high_mass);
        const double new_size_factor = mass_factor((double)new_size, o->low_size, o->target_size, o->high_size);
        const double new_score = old_score + current_cold_diff;
        double new_adjusted = new_score - new_size * mc_log2((double)new_size);
        new_adjusted += new_size * (new_size_factor < new_mass_factor ? new_size_factor : new_mass_factor);
        current_adjusted_cold_diff = new_adjusted - old_adjusted;
    }
    const double current_connectivity_diff = o->connectivity_diffs[current];
    const double current_disconnected_diff = o->disconnected_diffs[current];

    o->candidates_count = 0;
    for (int64_t p = 0; p < o->partitions; ++p) {
        if ((int32_t)p == current) {
            o->total_diffs[p] = 0.0;
            continue;
        }
        const double hot_diff = o->hot_diffs[p];
        const double cold_diff = o->cold_diffs[p];

        const int64_t old_size = o->size_of_partitions[p];
        const double old_mass = o->mass_of_partitions[p];
        const double old_mass_factor = mass_factor(old_mass, o->low_mass, o->target_mass, o->high_mass);
        const double old_size_factor = mass_factor((double)old_size, o->low_size, o->target_size, o->high_size);
        const double old_score = o->score_of_partitions[p];
        double old_adjusted = old_score - old_size * mc_log2((double)old_size);
        old_adjusted += old_size * (old_size_factor < old_mass_factor ? old_size_factor : old_mass_factor);

        const int64_t new_size = old_size + 1;
        const double new_mass = old_mass + node_mass;
        const double new_mass_factor = mass_factor(new_mass, o->low_mass, o->target_mass, o->high_mass);
        const double new_size_factor = mass_factor((double)new_size, o->low_size, o->target_size, o->high_size);
        const double new_score = old_score + cold_diff;
        double new_adjusted = new_score - new_size * mc_log2((double)new_size);
        new_adjusted += new_size * (new_size_factor < new_mass_factor ? new_size_factor : new_mass_factor);

        const double adjusted_connectivity_diff = o->connectivity_diffs[p] + current_connectivity_diff;
        if (o->disconnected_diffs[p] + current_disconnected_diff > 0) continue;
        if (adjusted_connectivity_diff < min_connectivity_diff) continue;

        const double adjusted_cold_diff = new_adjusted - old_adjusted;
        const double total_diff = (current_hot_diff + hot_diff) * temperature +
                                  (current_adjusted_cold_diff + adjusted_cold_diff) * (1.0 - temperature);
        o->total_diffs[p] = total_diff;
        if (total_diff >= EPSILON || min_connectivity_diff > 0) o->candidates[o->candidates_count++] = p;
    }
    return current_cold_diff;
}

/* partitions.cpp:1007-1072 */
static int32_t choose_target_partition(Optimizer* o, uint32_t* rng) {
    if (o->candidates_count == 0) return -1;
    double min_total_diff = 0;
    for (int64_t c = 0; c < o->candidates_count; ++c) {
        const double diff = o->total_diffs[o->candidates[c]];
        if (c == 0 || diff < min_total_diff) min_total_diff = diff;
    }
    double sum_total_diff = 0;
    for (int64_t c = 0; c < o->candidates_count; ++c) sum_total_diff += o->total_diffs[o->candidates[c]] - min_total_diff;
    double chosen = uniform_real(rng, 0.0, sum_total_diff);
    for (int64_t c = 0; c < o->candidates_count; ++c) {
        chosen -= o->total_diffs[o->candidates[c]] - min_total_diff;
        if (chosen <= EPSILON) return (int32_t)o->candidates[c];
    }
    return -2; /* FastAssertCompare(false, ==, true) */
}

/* partitions.cpp:1074-1160 */
static void update_scores_of_nodes(Optimizer* o, int64_t node, int32_t from, int32_t to) {
    const int32_t* out_idx = o->out_idx + o->out_ptr[node];
    const int32_t* in_idx = o->in_idx + o->in_ptr[node];
    const float* out_w = o->out_w + o->out_ptr[node];
    const float* in_w = o->in_w + o->in_ptr[node];
    const int64_t out_count = o->out_ptr[node + 1] - o->out_ptr[node];
    const int64_t in_count = o->in_ptr[node + 1] - o->in_ptr[node];
    int64_t out_pos = 0, in_pos = 0;
    int32_t out_node = out_idx[0], in_node = in_idx[0];
    float out_weight = out_w[0], in_weight = in_w[0];
    while (out_pos < out_count || in_pos < in_count) {
        const int32_t other = out_node < in_node ? out_node : in_node;
        const int is_outgoing = (int)(out_node == other);
        const int is_incoming = (int)(in_node == other);

        NodeScore* from_score = &SCORE(o, from, other);
        update_incoming(from_score, -is_outgoing, out_weight);
        update_outgoing(from_score, -is_incoming, in_weight);
        rescore(from_score);

        NodeScore* to_score = &SCORE(o, to, other);
        update_incoming(to_score, +is_outgoing, out_weight);
        update_outgoing(to_score, +is_incoming, in_weight);
        rescore(to_score);

        out_pos += is_outgoing;
        in_pos += is_incoming;
        if (out_pos < out_count) { out_node = out_idx[out_pos]; out_weight = out_w[out_pos]; }
        else { out_node = (int32_t)o->nodes; out_weight = 0; }
        if (in_pos < in_count) { in_node = in_idx[in_pos]; in_weight = in_w[in_pos]; }
        else { in_node = (int32_t)o->nodes; in_weight = 0; }
    }
}

/* partitions.cpp:558-645.  Returns 1 = moved, 0 = not, <0 = a failed assert of the reference. */
static int improve_node(Optimizer* o, int64_t node, uint32_t* rng, double temperature) {
    const int32_t current = o->partition_of_nodes[node];
    if (o->size_of_partitions[current] < 2) return 0;
    if (o->out_ptr[node + 1] == o->out_ptr[node] || o->in_ptr[node + 1] == o->in_ptr[node]) return -1;
    collect_initial_partition_diffs(o, node, current);
    collect_cold_partition_diffs(o, node, current);
    const double current_cold_diff = collect_candidate_partitions(o, node, current, temperature);
    const int32_t chosen = choose_target_partition(o, rng);
    if (chosen == -2) return -1;
    if (chosen < 0) return connectivity(&SCORE(o, current, node)) > 0 ? 0 : -1;
    const double chosen_cold_diff = o->cold_diffs[chosen];
    update_scores_of_nodes(o, node, current, chosen);
    o->partition_of_nodes[node] = chosen;
    o->size_of_partitions[current] -= 1;
    o->size_of_partitions[chosen] += 1;
    o->score_of_partitions[current] += current_cold_diff;
    o->score_of_partitions[chosen] += chosen_cold_diff;
    const float node_mass = o->mass_of_nodes[node];
    o->mass_of_partitions[current] -= node_mass;
    o->mass_of_partitions[chosen] += node_mass;
    return 1;
}

/* partitions.cpp:388-556.  stats (nullable): [0] passes, [1] improve_node calls, [2] moves, [3] skipped. */
static int optimizer_optimize(Optimizer* o, uint64_t random_seed, double cooldown_pass, double cooldown_node,
                              double cold_temperature, int64_t* stats) {
    const int64_t n = o->nodes;
    uint32_t rng = rng_seed(random_seed);
    int64_t* indices = (int64_t*)malloc(sizeof(int64_t) * (size_t)(n ? n : 1));
    uint8_t* frozen = (uint8_t*)malloc((size_t)(n ? n : 1));
    if (!indices || !frozen) { free(indices); free(frozen); return 1; }
    for (int64_t i = 0; i < n; ++i) indices[i] = i;
    for (int64_t i = 0; i < n; ++i) {
        frozen[i] = 1;
        const int32_t partition = o->partition_of_nodes[i];
        if (0 <= partition && o->hot_mask[partition] == 0) {
            o->temperature_of_nodes[i] = cold_temperature;
            frozen[i] = 1;
        } else if (partition >= -1) {
            o->temperature_of_nodes[i] = 1.0;
            frozen[i] = 0;
        }
    }
    int status = 0;
    if (!(cooldown_node >= 0.0) || !(cooldown_node <= 1.0)) status = 2;
    if (cooldown_pass != 0.0 && (!(cooldown_pass > 0.0) || !(cooldown_pass < 1.0))) status = 2;
    const double cooldown_rate = 1.0 - cooldown_pass / n;
    double temperature = 1.0;
    int did_improve = 1, did_skip = 0;
    while (status == 0 && (temperature > 0 || did_improve)) {
        if (did_improve) {
            did_improve = 0;
            did_skip = 0;
        } else if (did_skip) {
            did_skip = 0;
            for (int64_t i = 0; i < n; ++i)
                if (!frozen[i] && o->temperature_of_nodes[i] < temperature) temperature = o->temperature_of_nodes[i];
        } else {
            temperature = 0;
        }
        shuffle_indices(indices, n, &rng);
        if (stats) ++stats[0];
        for (int64_t position = 0; position < n; ++position) {
            const int64_t node = indices[position];
            temperature *= cooldown_rate;
            if (o->temperature_of_nodes[node] < temperature) {
                did_skip = 1;
                if (stats) ++stats[3];
                continue;
            }
            if (stats) ++stats[1];
            const int moved = improve_node(o, node, &rng, temperature);
            if (moved < 0) { status = 2; break; }
            frozen[node] = 0;
            if (moved) {
                did_improve = 1;
                if (stats) ++stats[2];
            } else {
                o->temperature_of_nodes[node] = temperature * (1 - cooldown_node);
            }
        }
    }
    free(indices);
    free(frozen);
    return status;
}

static void optimizer_bind(Optimizer* o, const float* out_w, const int32_t* out_idx, const int32_t* out_ptr,
                           const float* in_w, const int32_t* in_idx, const int32_t* in_ptr, int64_t nodes,
                           const float* mass_of_nodes, const double* bounds, int32_t* partition_of_nodes,
                           const int8_t* hot_mask) {
    memset(o, 0, sizeof(*o));
    if (g_log2_scale == 0.0) g_log2_scale = 1.0 / log(2.0); /* extensions.h:132 */
    o->out_w = out_w; o->out_idx = out_idx; o->out_ptr = out_ptr;
    o->in_w = in_w; o->in_idx = in_idx; o->in_ptr = in_ptr;
    o->nodes = nodes;
    o->mass_of_nodes = mass_of_nodes;
    o->low_mass = bounds[0]; o->target_mass = bounds[1]; o->high_mass = bounds[2];
    o->low_size = bounds[3]; o->target_size = bounds[4]; o->high_size = bounds[5];
    o->partition_of_nodes = partition_of_nodes;
    o->hot_mask = hot_mask;
}

/* partitions.cpp:1234-1302.  bounds = {low_mass, target_mass, high_mass, low_size, target_size, high_size}. */
int mco_optimize_partitions(const float* out_w, const int32_t* out_idx, const int32_t* out_ptr, const float* in_w,
                            const int32_t* in_idx, const int32_t* in_ptr, int64_t nodes, uint32_t random_seed,
                            const float* mass_of_nodes, const double* bounds, double cooldown_pass, double cooldown_node,
                            int32_t* partition_of_nodes, const int8_t* hot_mask, double cold_temperature, double* score,
                            int64_t* stats) {
    Optimizer o;
    optimizer_bind(&o, out_w, out_idx, out_ptr, in_w, in_idx, in_ptr, nodes, mass_of_nodes, bounds, partition_of_nodes,
                   hot_mask);
    int status = optimizer_init(&o);
    if (status == 0) status = optimizer_optimize(&o, random_seed, cooldown_pass, cooldown_node, cold_temperature, stats);
    if (status == 0) *score = optimizer_score(&o, 1);
    optimizer_free(&o);
    return status;
}

/* partitions.cpp:1304-1338 (the 16-argument C++ form; SURVEY.md 8(c) quirk iii). */
int mco_score_partitions(const float* out_w, const int32_t* out_idx, const int32_t* out_ptr, const float* in_w,
                         const int32_t* in_idx, const int32_t* in_ptr, int64_t nodes, const float* mass_of_nodes,
                         const double* bounds, int32_t* partition_of_nodes, const int8_t* hot_mask, int with_orphans,
                         double* score) {
    Optimizer o;
    optimizer_bind(&o, out_w, out_idx, out_ptr, in_w, in_idx, in_ptr, nodes, mass_of_nodes, bounds, partition_of_nodes,
                   hot_mask);
    int status = optimizer_init(&o);
    if (status == 0) *score = optimizer_score(&o, with_orphans);
    optimizer_free(&o);
    return status;
}
