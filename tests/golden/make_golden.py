"""Generate tests/golden/*.npz from the reference's OWN CUDA kernels (oracle/_ref/libperch_ref.so, built from
/root/reference by `make -C oracle ref`).  Needs a GPU: run on the B200 box, e.g.
    gpurun -- 'python tests/golden/make_golden.py gpurun_out/golden'
and copy the .npz files into tests/golden/.  Each rasterizer case is run 3 times to expose the reference kernel's
data race (image_renderer.cuh:161-186); the number of pixels that differ between runs is stored as `race_budget`."""
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))

import make_golden_inputs as gi  # noqa: E402
from oracle_lib import Oracle, Reference  # noqa: E402


def main(out_dir):
    out = Path(out_dir)
    out.mkdir(parents=True, exist_ok=True)
    ref, orc = Reference(), Oracle()
    for name in gi.NAMES:
        runs = [gi.run_reference(ref, orc, name) for _ in range(3 if name.startswith("render") else 1)]
        save = {f"ref_{k}": v for k, v in runs[0].items()}
        if name.startswith("render"):
            unstable = sum(int(np.count_nonzero(runs[0]["depth"] != r["depth"])) for r in runs[1:])
            o = gi.run_oracle(orc, name)
            vs_oracle = int(np.count_nonzero(runs[0]["depth"] != o["depth"]))
            save["race_budget"] = np.array(max(4, unstable + vs_oracle))
            print(f"{name}: run-to-run unstable px {unstable}, vs oracle {vs_oracle}")
        np.savez_compressed(out / f"{name}.npz", **save)
        print("wrote", out / f"{name}.npz")


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else str(Path(__file__).parent))
