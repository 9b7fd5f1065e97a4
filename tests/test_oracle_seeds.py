"""CPU tests pinning the seed-chooser oracle (oracle/mc_seeds.c):

- against the golden vectors the UNMODIFIED reference produced (tests/golden/seeds.npz, partitions.npz), bit for bit;
- against the compiled reference itself (oracle/_ref) when it is present (build container);
- its restatements of libstdc++'s std::nth_element / std::sort against libstdc++ (a tiny C++ driver built here).
"""
import ctypes
import os
import subprocess

import numpy as np
import pytest

from oracle import oracle
from seeds_cases import CASES, case_inputs
from synth import random_partition_graph

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_choose_seeds_matches_reference_golden(golden_dir):
    expected = np.load(os.path.join(golden_dir, "seeds.npz"))
    for name in CASES:
        outgoing, incoming, seed, max_seeds, low, high, seed_of_nodes = case_inputs(name, golden_dir)
        count = oracle.choose_seeds(outgoing, incoming, seed, max_seeds, low, high, seed_of_nodes)
        assert count == expected[f"{name}_count"][0], name
        assert np.array_equal(seed_of_nodes, expected[f"{name}_seeds"]), name
        assert seed_of_nodes.min() >= 0 and seed_of_nodes.max() == count - 1, name


def test_choose_seeds_matches_the_partition_fixture(golden_dir):
    # tests/golden/partitions.npz holds the reference's seeds of the pile graph (quantiles 0.8 .. 1.0)
    fixture = np.load(os.path.join(golden_dir, "partitions.npz"))
    outgoing, incoming = case_inputs("pile_default", golden_dir)[:2]
    seed_of_nodes = np.full(outgoing.shape[0], -1, dtype="int32")
    count = int(fixture["seeds"].max()) + 1
    assert oracle.choose_seeds(outgoing, incoming, 123456, count, 0.8, 1.0, seed_of_nodes) == count
    assert np.array_equal(seed_of_nodes, fixture["seeds"])


def test_choose_seeds_matches_compiled_reference(ref):
    if ref is None:
        pytest.skip("oracle/_ref is not built here")
    for nodes, degree, clusters, seed, max_seeds, low, high in (
        (300, 10, 4, 11, 9, 0.85, 0.95), (1200, 24, 10, 12, 40, 0.1, 0.3), (700, 5, 3, 13, 300, 0.85, 0.95),
        (2500, 50, 25, 14, 52, 0.85, 0.95),
    ):
        outgoing, incoming, _, _ = random_partition_graph(nodes, degree, clusters, seed)
        mine = np.full(nodes, -1, dtype="int32")
        theirs = mine.copy()
        my_count = oracle.choose_seeds(outgoing, incoming, seed, max_seeds, low, high, mine)
        their_count = ref.choose_seeds(outgoing.data, outgoing.indices, outgoing.indptr, incoming.data, incoming.indices,
                                       incoming.indptr, seed, max_seeds, low, high, theirs)
        assert my_count == their_count
        assert np.array_equal(mine, theirs)


def test_reference_asserts_are_errors():
    outgoing, incoming, _, _ = random_partition_graph(64, 4, 2, 5)
    seed_of_nodes = np.full(64, -1, dtype="int32")
    with pytest.raises(oracle.ReferenceAssert):  # more seeds than candidates (choose_seeds.cpp:318)
        oracle.choose_seeds(outgoing, incoming, 1, 65, 0.85, 0.95, seed_of_nodes)
    with pytest.raises(oracle.ReferenceAssert):  # quantiles out of order (choose_seeds.cpp:296-298)
        oracle.choose_seeds(outgoing, incoming, 1, 5, 0.9, 0.5, seed_of_nodes)


@pytest.fixture(scope="module")
def std_select(tmp_path_factory):
    library = str(tmp_path_factory.mktemp("std_select") / "host_std_select.so")
    compiler = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
    subprocess.check_call([compiler, "-O2", "-std=c++14", "-shared", "-fPIC", "-o", library,
                           os.path.join(ROOT, "tests", "host_std_select.cpp")])
    return ctypes.CDLL(library)


def selection_keys(count: int, kind: int, rng) -> np.ndarray:
    if kind == 0:
        keys = rng.integers(0, 10**9, count)
    elif kind == 1:
        keys = rng.integers(0, 5, count)  # wide ties, as the connected sizes of a pile's candidates
    elif kind == 2:
        keys = np.arange(count)
    elif kind == 3:
        keys = np.arange(count)[::-1]
    elif kind == 4:
        keys = np.concatenate([np.arange(count // 2), np.arange(count - count // 2)[::-1]])
    else:
        keys = np.zeros(count)
    return np.ascontiguousarray(keys, dtype="int64")


def test_selection_and_sort_match_libstdcxx(std_select):
    rng = np.random.default_rng(1)
    pointer = lambda array: ctypes.c_void_p(array.ctypes.data)  # noqa: E731
    for count in list(range(1, 40)) + [100, 1000, 4097, 16000]:
        for kind in range(6):
            keys = selection_keys(count, kind, rng)
            for nth in sorted({0, count // 2, count - 1, int(0.85 * (count - 1))}):
                theirs = np.arange(count, dtype="int64")
                std_select.host_nth_element(pointer(theirs), pointer(keys), ctypes.c_int64(0), ctypes.c_int64(nth),
                                            ctypes.c_int64(count))
                assert np.array_equal(oracle.std_nth_element(keys, 0, nth, count), theirs), (count, kind, nth)
            theirs = np.arange(count, dtype="int64")
            std_select.host_sort(pointer(theirs), pointer(keys), ctypes.c_int64(count))
            assert np.array_equal(oracle.std_sort(keys), theirs), (count, kind)
