"""Generates tests/golden/*.npz from the UNMODIFIED reference compiled into oracle/_ref (run in the
build container, where /root/reference exists:  make -C oracle && python tests/golden/make_golden.py).

Every fixture stores the inputs and the reference's outputs of one stage, so the CPU oracle
(oracle/mc_oracle.c, oracle/oracle.py) and the CUDA path can be checked on machines where
/root/reference and oracle/_ref do not exist.  Inputs are small and seeded (123456, the reference's
habitual test seed, tests/test_utilities.py:38).
"""
import os
import sys

import numpy as np
import scipy.sparse as sp

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from oracle import oracle  # noqa: E402
from synth import metacell_bounds, random_partition_graph, seeds_count_for, synth_umis  # noqa: E402

SEED = 123456


def main() -> None:
    ref = oracle.load_ref()
    if ref is None:
        raise SystemExit("oracle/_ref is not built: run `make -C oracle` first")
    ref.set_threads_count(1)

    # ---- downsample ---------------------------------------------------------------------------
    umis = synth_umis(96, 1500, seed=SEED)
    totals = np.asarray(umis.sum(axis=1)).reshape(-1)
    samples = np.full(umis.shape[0], oracle.downsample_samples(totals), dtype="int32")
    samples[3] = -1
    samples[4] = 0
    samples[5] = 10**7
    outputs = {}
    for seed in (0, 1, SEED):
        out = np.full_like(umis.data, -7.0)
        ref.downsample_compressed_float32_t_int32_t_float32_t(umis.data, umis.indptr, out, samples, seed)
        outputs[f"out_seed_{seed}"] = out
    np.savez_compressed(os.path.join(HERE, "downsample.npz"), data=umis.data, indices=umis.indices, indptr=umis.indptr,
                        samples=samples, **outputs)

    # ---- correlate (wheel-equivalent AVX2 build of the reference) -------------------------------
    samples[:] = oracle.downsample_samples(totals)
    down = umis.copy()
    down.data = oracle.downsample_compressed(umis.data, umis.indptr, samples, SEED)
    gene_totals = np.asarray(down.sum(axis=0)).reshape(-1)
    chosen = np.sort(np.argsort(-gene_totals, kind="stable")[:300])
    dense = oracle.log_fraction(down[:, chosen].toarray())
    corr = np.empty((dense.shape[0], dense.shape[0]), dtype="float32")
    ref.correlate_dense_float32_t(dense, corr)
    corr64 = np.empty_like(corr)
    ref.correlate_dense_float64_t(dense.astype("float64"), corr64)
    cross = np.empty((40, 56), dtype="float32")
    ref.cross_correlate_dense_float32_t(np.ascontiguousarray(dense[:40]), np.ascontiguousarray(dense[40:]), cross)
    pairs = np.empty(40, dtype="float32")
    ref.pairs_correlate_dense_float32_t(np.ascontiguousarray(dense[:40]), np.ascontiguousarray(dense[40:80]), pairs)
    np.savez_compressed(os.path.join(HERE, "correlate.npz"), dense=dense, corr=corr, corr64=corr64, cross=cross, pairs=pairs)

    # ---- rank / top / prune / relayout on tie-free inputs (where the reference is unambiguous) ----
    rng = np.random.default_rng(SEED)
    distinct = np.stack([rng.permutation(400) for _ in range(12)]).astype("float32") / 8
    ranks_desc = distinct.copy()
    ref.rank_matrix_float32_t(ranks_desc, False)
    ranks_asc = distinct.copy()
    ref.rank_matrix_float32_t(ranks_asc, True)
    kth = np.empty(12, dtype="float32")
    ref.rank_rows_float32_t(distinct, kth, 37)
    top_indices = np.empty(12 * 25, dtype="int32")
    top_data = np.empty(12 * 25, dtype="float32")
    ref.collect_top_float32_t(25, distinct, top_indices, top_data, False)
    top_rank_data = np.empty(12 * 25, dtype="float32")
    top_rank_indices = np.empty(12 * 25, dtype="int32")
    ref.collect_top_float32_t(25, distinct, top_rank_indices, top_rank_data, True)
    np.savez_compressed(os.path.join(HERE, "rank.npz"), distinct=distinct, ranks_desc=ranks_desc, ranks_asc=ranks_asc, kth=kth,
                        top_indices=top_indices, top_data=top_data, top_rank_data=top_rank_data)

    graph = sp.random(60, 60, density=0.3, format="csr", random_state=np.random.default_rng(SEED), dtype="float64")
    graph.data = rng.permutation(len(graph.data)).astype("float32") + 1  # distinct values
    graph = sp.csr_matrix(graph, dtype="float32")
    graph.sort_indices()
    indices = graph.indices.astype("int32")
    indptr = graph.indptr.astype("int32")
    out_data = np.empty(60 * 7, dtype="float32")
    out_indices = np.empty(60 * 7, dtype="int32")
    out_indptr = np.empty(61, dtype="int32")
    ref.collect_pruned_float32_t_int32_t_int32_t(7, graph.data, indices, indptr, out_data, out_indices, out_indptr)
    nnz = out_indptr[-1]
    counts = np.bincount(indices, minlength=60)
    t_indptr = np.zeros(61, dtype="int32")
    np.cumsum(counts[:-1], out=t_indptr[2:])
    t_data = np.empty_like(graph.data)
    t_indices = np.empty_like(indices)
    ref.collect_compressed_float32_t_int32_t_int32_t(graph.data, indices, indptr, t_data, t_indices, t_indptr[1:])
    ref.sort_compressed_indices_float32_t_int32_t_int32_t(t_data, t_indices, t_indptr, 60)
    np.savez_compressed(os.path.join(HERE, "sparse.npz"), data=graph.data, indices=indices, indptr=indptr,
                        pruned_data=out_data[:nnz], pruned_indices=out_indices[:nnz], pruned_indptr=out_indptr,
                        t_data=t_data, t_indices=t_indices, t_indptr=t_indptr)

    # ---- the k-NN graph of a small pile: reference natives + the numpy glue of knn_graph.py ----------
    # The similarity has no exact ties off the diagonal, and balanced ranks are checked through the
    # oracle's glue with the REFERENCE natives plugged in (see tests/test_oracle.py).
    sim = np.abs(corr64).astype("float32")
    np.fill_diagonal(sim, 1.0)
    work = sim.copy()
    np.fill_diagonal(work, work.min() - 1)
    ref.rank_matrix_float32_t(work, False)
    np.savez_compressed(os.path.join(HERE, "knn.npz"), similarity=sim, outgoing_ranks=work)
    # ---- partitions.cpp: a small pile's graph, seeds from the reference's choose_seeds, then the reference's
    # optimize_partitions / score_partitions for several seeds, hot masks and cold temperatures ------------
    umis = synth_umis(600, 3000, seed=SEED)
    cell_umis = np.asarray(umis.sum(axis=1)).reshape(-1).astype("float32")
    outgoing = oracle.cpu_pile_knn_graph(umis, 400, 20, SEED, ref=ref)
    outgoing.sort_indices()
    incoming = sp.csc_matrix(outgoing)
    incoming.sort_indices()
    bounds = metacell_bounds(48, 48 * float(np.median(cell_umis)))
    seeds = np.full(600, -1, dtype="int32")
    seeds_count = ref.choose_seeds(outgoing.data, outgoing.indices, outgoing.indptr, incoming.data, incoming.indices,
                                   incoming.indptr, SEED, seeds_count_for(float(cell_umis.sum()), 600, bounds), 0.8, 1.0, seeds)
    cases = {}
    hot_all = np.ones(max(seeds_count, 2), dtype="int8")
    hot_some = hot_all.copy()
    hot_some[::3] = 0
    for name, (seed, mask, cold, cooldown_pass, cooldown_node) in {
        "a": (SEED, hot_all, 0.98, 0.02, 0.25), "b": (1, hot_some, 0.5, 0.02, 0.25), "c": (0, hot_all, 1.0, 0.1, 0.0),
    }.items():
        partition = seeds.copy()
        score = ref.optimize_partitions(outgoing.data, outgoing.indices, outgoing.indptr, incoming.data, incoming.indices,
                                        incoming.indptr, seed, cell_umis, *bounds, cooldown_pass, cooldown_node, partition, mask,
                                        cold)
        scored = ref.score_partitions(outgoing.data, outgoing.indices, outgoing.indptr, incoming.data, incoming.indices,
                                      incoming.indptr, cell_umis, *bounds, partition, mask, False)
        cases[f"{name}_params"] = np.array([seed, cold, cooldown_pass, cooldown_node], dtype="float64")
        cases[f"{name}_mask"] = mask
        cases[f"{name}_partition"] = partition
        cases[f"{name}_scores"] = np.array([score, scored], dtype="float64")
    np.savez_compressed(os.path.join(HERE, "partitions.npz"), data=outgoing.data, indices=outgoing.indices, indptr=outgoing.indptr,
                        cell_umis=cell_umis, bounds=np.array(bounds), seeds=seeds, **cases)

    # a random clustered graph with two partitions only (the hot_mask of the reference's caller has >= 2 entries)
    outgoing, incoming, masses, initial = random_partition_graph(300, 12, 2, SEED)
    bounds = (30000.0, 400000.0, 600000.0, 20.0, 150.0, 250.0)
    partition = initial.copy()
    mask = np.ones(2, dtype="int8")
    score = ref.optimize_partitions(outgoing.data, outgoing.indices, outgoing.indptr, incoming.data, incoming.indices,
                                    incoming.indptr, 7, masses, *bounds, 0.05, 0.25, partition, mask, 0.95)
    np.savez_compressed(os.path.join(HERE, "partitions_two.npz"), partition=partition, score=np.array([score]))
    print("golden fixtures written to", HERE)


if __name__ == "__main__":
    main()
