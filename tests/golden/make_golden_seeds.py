"""Golden vectors for choose_seeds (choose_seeds.cpp:264-372), made by the UNMODIFIED reference compiled into
oracle/_ref (run in the build container: python tests/golden/make_golden_seeds.py).  The graphs are rebuilt
from tests/synth.py and tests/golden/partitions.npz by the tests; only the reference's outputs are stored."""
import os
import sys

import numpy as np
import scipy.sparse as sp

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle import oracle  # noqa: E402
from seeds_cases import CASES, case_inputs  # noqa: E402


def main() -> None:
    ref = oracle.load_ref()
    assert ref is not None, "build oracle/_ref first (make -C oracle)"
    stored = {}
    for name in CASES:
        outgoing, incoming, seed, max_seeds, low, high, seed_of_nodes = case_inputs(name, HERE)
        count = ref.choose_seeds(outgoing.data, outgoing.indices, outgoing.indptr, incoming.data, incoming.indices,
                                 incoming.indptr, seed, max_seeds, low, high, seed_of_nodes)
        stored[f"{name}_seeds"] = seed_of_nodes
        stored[f"{name}_count"] = np.array([count], dtype="int64")
        print(name, count)
    np.savez_compressed(os.path.join(HERE, "seeds.npz"), **stored)


if __name__ == "__main__":
    main()
