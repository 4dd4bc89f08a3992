"""kNN package API (knncuda.h contract): perch_knn against the reference's own knn_cuda_global / knn_cuda_texture /
knn_cublas compiled unmodified (oracle/_ref/libknn_ref.so), same inputs, host arrays in and out (wall clock of the call),
plus the device time of perch's kernel and its FP32 rate.  Usage (GPU box): python tools/knn_probe.py"""
import contextlib
import json
import os
import sys
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))
from oracle_lib import KNN_REF_SO, ReferenceKnn  # noqa: E402
from perception_b200.renderer import Renderer  # noqa: E402


@contextlib.contextmanager
def quiet():  # the reference printf's on stdout
    sys.stdout.flush()
    old = os.dup(1)
    null = os.open(os.devnull, os.O_WRONLY)
    os.dup2(null, 1)
    try:
        yield
    finally:
        os.dup2(old, 1)
        os.close(null)
        os.close(old)


ref_lib = ReferenceKnn() if KNN_REF_SO.exists() else None
rng = np.random.default_rng(3)
with Renderer(0) as r:
    for ref_nb, query_nb, k in ((4096, 262144, 1), (813, 3000000, 1), (16384, 65536, 4), (16384, 65536, 8), (16384, 65536, 10), (233098, 233098, 10)):
        ref = rng.random((3, ref_nb), np.float32)
        qry = rng.random((3, query_nb), np.float32)
        wall, dev = [], []
        for _ in range(4):
            t0 = time.perf_counter()
            d, i = r.knn(ref, qry, k)
            wall.append(time.perf_counter() - t0)
            dev.append(r.last_timing()[0])
        row = {"ref_nb": ref_nb, "query_nb": query_nb, "k": k, "perch_wall_ms": 1e3 * min(wall[1:]),
               "perch_kernel_ms": float(np.median(dev[1:])),
               "perch_kernel_GFMA_per_s": 3.0 * ref_nb * query_nb / (np.median(dev[1:]) * 1e-3) / 1e9}
        if ref_lib is not None and ref_nb * query_nb < 2 ** 31:  # the reference indexes its distance matrix with int
            for which in ("knn_cuda_global", "knn_cuda_texture", "knn_cublas"):
                ts, ok, same = [], True, None
                for _ in range(3):
                    with quiet():
                        t0 = time.perf_counter()
                        ok, rd, ri = ref_lib.run(which, ref, qry, k)
                        ts.append(time.perf_counter() - t0)
                if ok and which != "knn_cublas":
                    same = bool(np.array_equal(ri, i) and np.array_equal(rd, d))
                row[which + "_wall_ms"] = 1e3 * min(ts[1:]) if ok else None
                if same is not None:
                    row[which + "_bit_equal"] = same
        print(json.dumps(row), flush=True)
