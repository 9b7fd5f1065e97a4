"""The inputs of the choose_seeds parity cases (shared by the golden generator, the oracle tests and the GPU tests)."""
import os

import numpy as np
import scipy.sparse as sp

from synth import random_partition_graph

# name -> (graph, random_seed, max_seeds_count, min_quantile, max_quantile, given)
#   graph: ("pile",) = the 600-cell pile graph of tests/golden/partitions.npz; ("random", nodes, degree, clusters, seed);
#          ("ties", nodes, degree, clusters, seed) = the same with weights rounded to 1/8 (equal weights at the cut)
#   given: 0 = no seeds given; n = the first nodes // n nodes carry three given seeds
CASES = {
    "pile_default": (("pile",), 123456, 12, 0.85, 0.95, 0),
    "pile_wide": (("pile",), 7, 20, 0.0, 1.0, 0),
    "pile_given": (("pile",), 0, 16, 0.8, 1.0, 10),
    "random_small": (("random", 200, 8, 5, 1), 1, 10, 0.85, 0.95, 0),
    "random_medium": (("random", 3000, 30, 40, 5), 5, 70, 0.85, 0.95, 0),
    "random_given": (("random", 1000, 20, 12, 4), 4, 25, 0.5, 0.5, 10),
    "random_exhausted": (("random", 500, 4, 4, 6), 6, 200, 0.85, 0.95, 0),
    "random_tiny": (("random", 64, 3, 2, 7), 7, 50, 0.85, 0.95, 7),
    "ties": (("ties", 800, 40, 6, 8), 8, 30, 0.85, 0.95, 0),
    "ties_top": (("ties", 800, 40, 6, 9), 9, 11, 1.0, 1.0, 0),
}


def case_inputs(name: str, golden_dir: str):
    graph, seed, max_seeds, low, high, given = CASES[name]
    if graph[0] == "pile":
        fixture = np.load(os.path.join(golden_dir, "partitions.npz"))
        nodes = len(fixture["indptr"]) - 1
        outgoing = sp.csr_matrix((fixture["data"], fixture["indices"], fixture["indptr"]), shape=(nodes, nodes))
    else:
        outgoing = random_partition_graph(*graph[1:])[0]
        if graph[0] == "ties":
            outgoing.data = (np.ceil(outgoing.data * 64 * 8) / 8).astype("float32")
        nodes = outgoing.shape[0]
    incoming = sp.csc_matrix(outgoing)
    incoming.sort_indices()
    seed_of_nodes = np.full(nodes, -1, dtype="int32")
    if given:
        seed_of_nodes[: nodes // given] = np.arange(nodes // given) % 3
    return outgoing, incoming, seed, max_seeds, low, high, seed_of_nodes
