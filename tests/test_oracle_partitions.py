"""CPU tests pinning the partition-optimiser oracle (oracle/mc_partitions.c) and the restated glibc log.

- against the golden fixtures the UNMODIFIED reference produced (tests/golden/partitions*.npz), bit for bit;
- against the compiled reference itself (oracle/_ref) when it is present (build container);
- the device log's operation sequence (metacells_b200/csrc/glibc_log.cuh compiled as plain C++) against libm.
"""
import ctypes
import os
import subprocess

import numpy as np
import pytest
import scipy.sparse as sp

from oracle import oracle
from synth import random_partition_graph

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SEED = 123456


def golden_problem(golden_dir):
    fixture = np.load(os.path.join(golden_dir, "partitions.npz"))
    data, indices, indptr = fixture["data"], fixture["indices"], fixture["indptr"]
    nodes = len(indptr) - 1
    outgoing = sp.csr_matrix((data, indices, indptr), shape=(nodes, nodes))
    incoming = sp.csc_matrix(outgoing)
    incoming.sort_indices()
    return fixture, outgoing, incoming


def test_optimize_partitions_matches_reference_golden(golden_dir):
    fixture, outgoing, incoming = golden_problem(golden_dir)
    bounds = tuple(fixture["bounds"])
    for name in "abc":
        seed, cold, cooldown_pass, cooldown_node = fixture[f"{name}_params"]
        mask = fixture[f"{name}_mask"]
        partition = fixture["seeds"].copy()
        score = oracle.optimize_partitions(outgoing, incoming, int(seed), fixture["cell_umis"], *bounds, cooldown_pass,
                                           cooldown_node, partition, mask, cold)
        assert np.array_equal(partition, fixture[f"{name}_partition"]), name
        assert score == fixture[f"{name}_scores"][0], name
        scored = oracle.score_partitions(outgoing, incoming, fixture["cell_umis"], *bounds, partition, mask, False)
        assert scored == fixture[f"{name}_scores"][1], name


def test_two_partitions_matches_reference_golden(golden_dir):
    expected = np.load(os.path.join(golden_dir, "partitions_two.npz"))
    outgoing, incoming, masses, initial = random_partition_graph(300, 12, 2, SEED)
    partition = initial.copy()
    score = oracle.optimize_partitions(outgoing, incoming, 7, masses, 30000.0, 400000.0, 600000.0, 20.0, 150.0, 250.0, 0.05,
                                       0.25, partition, np.ones(2, dtype="int8"), 0.95)
    assert np.array_equal(partition, expected["partition"])
    assert score == expected["score"][0]


def test_oracle_matches_compiled_reference(ref):
    if ref is None:
        pytest.skip("oracle/_ref is not built here")
    for nodes, degree, partitions, seed in ((200, 8, 5, 1), (400, 16, 9, 2), (257, 6, 3, 3)):
        outgoing, incoming, masses, initial = random_partition_graph(nodes, degree, partitions, seed)
        bounds = (20000.0, 60000.0, 150000.0, 10.0, 40.0, 90.0)
        mask = np.ones(max(partitions, 2), dtype="int8")
        mask[seed % partitions] = 0
        mine, theirs = initial.copy(), initial.copy()
        stats = np.zeros(4, dtype="int64")
        my_score = oracle.optimize_partitions(outgoing, incoming, seed, masses, *bounds, 0.03, 0.25, mine, mask, 0.9, stats)
        their_score = ref.optimize_partitions(outgoing.data, outgoing.indices, outgoing.indptr, incoming.data, incoming.indices,
                                              incoming.indptr, seed, masses, *bounds, 0.03, 0.25, theirs, mask, 0.9)
        assert np.array_equal(mine, theirs)
        assert my_score == their_score
        assert stats[2] > 0  # the case does move nodes
        for with_orphans in (True, False):
            assert oracle.score_partitions(outgoing, incoming, masses, *bounds, mine, mask, with_orphans) == \
                ref.score_partitions(outgoing.data, outgoing.indices, outgoing.indptr, incoming.data, incoming.indices,
                                     incoming.indptr, masses, *bounds, theirs, mask, with_orphans)


def test_reference_asserts_are_errors():
    outgoing, incoming, masses, initial = random_partition_graph(64, 4, 2, 5)
    bad = initial.copy()
    bad[3] = -1
    with pytest.raises(oracle.ReferenceAssert):
        oracle.optimize_partitions(outgoing, incoming, 1, masses, 1.0, 2.0, 3.0, 1.0, 2.0, 3.0, 0.02, 0.25, bad,
                                   np.ones(2, dtype="int8"), 0.9)


def test_std_shuffle_is_a_permutation_and_deterministic():
    for count in (1, 2, 7, 1000, 46341):
        first = oracle.std_shuffle(count, SEED)
        assert np.array_equal(np.sort(first), np.arange(count))
        assert np.array_equal(first, oracle.std_shuffle(count, SEED))
    # libstdc++ 13, std::minstd_rand(1), ten elements (checked against a g++ program in the build container)
    assert oracle.std_shuffle(10, 1).tolist() == [5, 7, 1, 4, 0, 2, 6, 9, 8, 3]


def test_restated_glibc_log_matches_libm(tmp_path):
    """metacells_b200/csrc/glibc_log.cuh, compiled for the host, against the libm `log` the reference calls."""
    library = tmp_path / "host_glibc_log.so"
    compiler = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
    subprocess.check_call([compiler, "-O2", "-mfma", "-ffp-contract=off", "-fPIC", "-shared", "-o", str(library),
                           os.path.join(ROOT, "tests", "host_glibc_log.cpp")])
    harness = ctypes.CDLL(str(library))
    rng = np.random.default_rng(SEED)
    values = np.concatenate([
        rng.random(400_000) * 2, np.exp(rng.normal(0, 20, 400_000)), 1 + rng.normal(0, 0.03, 400_000),
        1e-6 + rng.random(200_000) * rng.random(200_000),  # the EPSILON + outgoing * incoming of rescore()
        np.array([1.0, 0.0, np.inf, 5e-324, 1e-310, 2.0, 1e-6, 0.9375, 1.064697265625, -1.0, np.nan])])
    mine = np.empty_like(values)
    harness.host_glibc_log(ctypes.c_void_p(values.ctypes.data), ctypes.c_void_p(mine.ctypes.data), ctypes.c_longlong(values.size))
    theirs = oracle.glibc_log(values)
    same = (mine.view("uint64") == theirs.view("uint64")) | (np.isnan(mine) & np.isnan(theirs))
    assert same.all(), values[~same][:5]
