"""Generate tests/golden/knn.npz from the reference's OWN kNN package (oracle/_ref/libknn_ref.so = unmodified
cuda_renderer/src/cuda/knncuda.cu): knn_c (host) and knn_cuda_global (device) outputs for tests/knn_cases.py.
Needs a GPU: gpurun -- 'python tests/golden/make_golden_knn.py gpurun_out/golden'"""
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))

import knn_cases  # noqa: E402
from oracle_lib import Oracle, ReferenceKnn  # noqa: E402


def main(out_dir):
    out = Path(out_dir)
    out.mkdir(parents=True, exist_ok=True)
    ref, orc = ReferenceKnn(), Oracle()
    save = {}
    for name in knn_cases.CASES:
        r, q, k = knn_cases.make(name)
        ok_c, d_c, i_c = ref.run("knn_c", r, q, k)
        ok_g, d_g, i_g = ref.run("knn_cuda_global", r, q, k)
        ok_t, d_t, i_t = ref.run("knn_cuda_texture", r, q, k)
        assert ok_c and ok_g and ok_t, name
        od, oi = orc.knn(r, q, k, fused=True)
        hd, hi = orc.knn(r, q, k, fused=False)
        print(f"{name}: oracle(fused)==knn_cuda_global dist {np.array_equal(od, d_g)} idx {np.array_equal(oi, i_g)} | "
              f"texture==global {np.array_equal(d_t, d_g) and np.array_equal(i_t, i_g)} | "
              f"oracle(host)==knn_c dist {np.array_equal(hd, d_c)} idx {np.array_equal(hi, i_c)}", flush=True)
        save[f"{name}_c_dist"], save[f"{name}_c_index"] = d_c, i_c
        save[f"{name}_global_dist"], save[f"{name}_global_index"] = d_g, i_g
    np.savez_compressed(out / "knn.npz", **save)
    print("wrote", out / "knn.npz")


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else str(Path(__file__).parent))
