"""Generate tests/golden/*.npz from the reference's OWN CUDA kernels (oracle/_ref/libperch_ref.so, built from
/root/reference by `make -C oracle ref`).  Needs a GPU: run on the B200 box, e.g.
    gpurun -- 'python tests/golden/make_golden.py gpurun_out/golden [case ...]'
and copy the .npz files into tests/golden/.  EXACT cases (see tests/make_golden_inputs.py) are run twice and must be
reproducible; BUDGET cases (full-mesh renders, where the reference rasterizer races) are run 3 times and the number of
pixels that differ between runs / from the canonical oracle is stored as `race_budget`."""
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))

import make_golden_inputs as gi  # noqa: E402
from oracle_lib import Oracle, Reference  # noqa: E402


def main(out_dir, only=()):
    out = Path(out_dir)
    out.mkdir(parents=True, exist_ok=True)
    ref, orc = Reference(), Oracle()
    for name in (only or gi.NAMES):
        exact = name in gi.EXACT
        runs = [gi.run_reference(ref, orc, name) for _ in range(2 if exact else 3)]
        save = {f"ref_{k}": v for k, v in runs[0].items()}
        o = gi.run_oracle(orc, name)
        if exact:
            stable = all(np.array_equal(runs[0][k], runs[1][k], equal_nan=True) for k in runs[0])
            same = {k: bool(np.array_equal(runs[0][k], o[k], equal_nan=True)) for k in runs[0]}
            print(f"{name}: reproducible={stable} equals_oracle={all(same.values())} {[k for k, v in same.items() if not v]}")
            for k, v in same.items():
                if not v and runs[0][k].shape == o[k].shape:
                    print("   ", k, "differs at", int(np.count_nonzero(runs[0][k] != o[k])), "of", runs[0][k].size)
        else:
            unstable = max(int(np.count_nonzero(runs[0]["depth"] != r["depth"])) for r in runs[1:])
            vs = [int(np.count_nonzero(r["depth"] != o["depth"])) for r in runs]
            lost = int(np.count_nonzero((runs[0]["depth"] == 0) & (o["depth"] > 0)))
            farther = int(np.count_nonzero((runs[0]["depth"] > o["depth"]) & (o["depth"] > 0)))
            nearer = int(np.count_nonzero((runs[0]["depth"] < o["depth"]) & (runs[0]["depth"] > 0)))
            save["race_budget"] = np.array(2 * max(vs) + 16)
            save["race_stats"] = np.array([unstable, max(vs), lost, farther, nearer, int((o["depth"] > 0).sum())])
            print(f"{name}: run-to-run unstable px {unstable}; vs canonical {vs} (cleared {lost}, farther {farther}, "
                  f"nearer {nearer}) of {int((o['depth'] > 0).sum())} object px")
        np.savez_compressed(out / f"{name}.npz", **save)
        print("wrote", out / f"{name}.npz")


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else str(Path(__file__).parent), sys.argv[2:])
