"""Synthetic UMI matrices for tests and bench (SURVEY.md section 8d generator)."""
import numpy as np
import scipy.sparse as sp


def synth_umis(n_cells: int, n_genes: int, seed: int = 123456, depth_scale: float = 1.0) -> sp.csr_matrix:
    """Cell-type mixture, lognormal gene rates, lognormal depth, Poisson counts -> CSR f32/i32/i32."""
    rng = np.random.default_rng(seed)
    n_types = max(4, n_cells // 400)
    base = np.exp(rng.normal(-3.0, 1.6, size=n_genes))
    fold = rng.normal(0.0, 0.7, size=(n_types, n_genes))
    types = rng.integers(0, n_types, size=n_cells)
    depth = rng.lognormal(0.0, 0.4, size=n_cells) * depth_scale
    blocks = []
    step = 2048
    for start in range(0, n_cells, step):
        stop = min(start + step, n_cells)
        rate = base[None, :] * np.exp(fold[types[start:stop]]) * depth[start:stop, None]
        blocks.append(sp.csr_matrix(rng.poisson(rate).astype("float32")))
    matrix = sp.vstack(blocks, format="csr")
    matrix.sort_indices()
    matrix.indices = matrix.indices.astype("int32")
    matrix.indptr = matrix.indptr.astype("int32")
    return matrix


def metacell_bounds(target_size: int, target_umis: float):
    """tools/candidates.py:151-160 with the default merge / split factors (0.5, 2.0) and min size 12:
    (low_umis, target_umis, high_umis, low_size, target_size, high_size)."""
    import math

    low_umis = int(math.floor(target_umis * 0.5))
    high_umis = int(math.ceil(target_umis * 2.0)) - 1
    low_size = max(12, int(math.floor(target_size * 0.5)))
    high_size = max(int(math.ceil(target_size * 2.0)) - 1, low_size + 1)
    return float(low_umis), float(target_umis), float(high_umis), float(low_size), float(target_size), float(high_size)


def seeds_count_for(total_umis: float, total_size: int, bounds) -> int:
    """tools/candidates.py:294-322."""
    import math

    low_umis, _, high_umis, low_size, _, high_size = bounds
    min_by_umis = int(math.ceil(total_umis / high_umis))
    max_by_umis = int(math.ceil(total_umis / low_umis))
    min_by_size = int(math.ceil(total_size / high_size))
    max_by_size = int(math.ceil(total_size / low_size))
    if max_by_size < min_by_umis:
        return int(round((max_by_size + min_by_umis) / 2))
    if max_by_umis < min_by_size:
        return int(round((max_by_umis + min_by_size) / 2))
    return int(round((max(min_by_size, min_by_umis) + min(max_by_size, max_by_umis)) / 2))


def random_partition_graph(nodes: int, degree: int, partitions: int, seed: int):
    """A random clustered graph in the form optimize_partitions takes: outgoing CSR with rows summing to 1
    (float32, sorted indices, no self edges), every node with at least one incoming edge, the incoming CSC,
    float32 node masses and an initial partition with no empty partition."""
    rng = np.random.default_rng(seed)
    truth = rng.integers(0, partitions, size=nodes)
    rows, cols = [], []
    for node in range(nodes):
        same = np.flatnonzero(truth == truth[node])
        same = same[same != node]
        picks = set(rng.choice(same, size=min(len(same), max(1, degree * 3 // 4)), replace=False).tolist()) if len(same) else set()
        while len(picks) < degree:
            other = int(rng.integers(0, nodes))
            if other != node:
                picks.add(other)
        picks.add((node + 1) % nodes)  # a ring: every node has an incoming edge
        picks = sorted(picks)
        rows += [node] * len(picks)
        cols += picks
    data = rng.random(len(rows)).astype("float32") + np.float32(0.05)
    outgoing = sp.csr_matrix((data, (rows, cols)), shape=(nodes, nodes), dtype="float32")
    outgoing.sort_indices()
    sums = np.asarray(outgoing.sum(axis=1)).reshape(-1)
    outgoing.data /= np.repeat(sums, np.diff(outgoing.indptr)).astype("float32")
    outgoing.data = outgoing.data.astype("float32")
    incoming = sp.csc_matrix(outgoing)
    incoming.sort_indices()
    masses = rng.integers(500, 5000, size=nodes).astype("float32")
    initial = rng.permutation(np.arange(nodes) % partitions).astype("int32")
    return outgoing, incoming, masses, initial
