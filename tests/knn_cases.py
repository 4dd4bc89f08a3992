"""Seeded inputs of the kNN parity cases (dim x N row-per-dimension float32), shared by the CPU and GPU tests and by
tests/golden/make_golden_knn.py."""
import numpy as np

# name -> (dim, ref_nb, query_nb, k, kind)
CASES = {
    "cloud3_k1": (3, 4000, 1500, 1, "cloud"),
    "cloud3_k10": (3, 3000, 700, 10, "cloud"),
    "ties3_k8": (3, 600, 300, 8, "lattice"),     # integer lattice points: many exactly equal distances
    "dim1_k4": (1, 257, 129, 4, "uniform"),
    "dim2_k16": (2, 1025, 130, 16, "uniform"),
    "dim4_k3": (4, 900, 64, 3, "uniform"),
    "dim7_k5": (7, 500, 33, 5, "uniform"),       # generic-dimension kernel
    "k_equals_ref": (3, 20, 50, 20, "uniform"),
    "one_query": (3, 2049, 1, 32, "uniform"),
}


def make(name):
    dim, nr, nq, k, kind = CASES[name]
    rng = np.random.default_rng(abs(hash(name)) % (2 ** 31) if False else sum(map(ord, name)))
    if kind == "cloud":      # metre-scale scene cloud, like the observed / rendered clouds of the cost stage
        ref = rng.normal(0, 0.15, (dim, nr)).astype(np.float32) + np.array([[0.1], [0.0], [0.9]], np.float32)
        query = (ref[:, rng.integers(0, nr, nq)] + rng.normal(0, 0.004, (dim, nq))).astype(np.float32)
    elif kind == "lattice":
        ref = rng.integers(0, 6, (dim, nr)).astype(np.float32)
        query = rng.integers(0, 6, (dim, nq)).astype(np.float32)
    else:
        ref = rng.uniform(-1, 1, (dim, nr)).astype(np.float32)
        query = rng.uniform(-1, 1, (dim, nq)).astype(np.float32)
    return np.ascontiguousarray(ref), np.ascontiguousarray(query), k
