"""Device time of the RENDER stage (full-resolution depth + RGB images, 7 B/px/pose written) against the HBM write
roofline, new rasteriser (raster_full.cu) vs the per-pose kernel (PERCH_IMAGE_LEGACY=1), and a sweep of its two knobs.
Usage (GPU box): python tools/render_probe.py [n_poses]"""
import json
import os
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))
import bench  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 512
r, wl = bench.build_workload("c2", 0, 0, None)
poses, pmm, lab = wl["poses"][:n], wl["pmm"][:n], wl["plabel"][:n]
peak, kind = bench.measured_peak_gbs()


def run(tag, **env):
    old = {k: os.environ.get(k) for k in env}
    os.environ.update({k: str(v) for k, v in env.items()})
    ms = []
    for rep in range(4):
        out = r.render_cuda_multi_unified("RENDER", poses, pmm, lab)
        ms.append(r.last_timing()[0])
    tf = []
    for rep in range(3):
        r.tree_flags(poses, pmm, lab, 1.0)
        tf.append(r.last_timing()[0])
    for k, v in old.items():
        os.environ.pop(k, None) if v is None else os.environ.__setitem__(k, v)
    k = float(np.median(ms[1:]))
    bytes_out = n * bench.W * bench.H * 7
    print(json.dumps({"variant": tag, "stage": "RENDER", "poses": n, "kernel_ms": k, "poses_per_s": n / (k / 1e3),
                      "written_GB_per_s": bytes_out / (k / 1e3) / 1e9, "hbm_peak_GB_per_s": peak,
                      "frac": bytes_out / (k / 1e3) / 1e9 / peak, "tree_flags_ms": float(np.median(tf[1:])),
                      "covered_px_per_pose": float((out["result_depth"] > 0).sum() / n)}), flush=True)
    return out


if len(sys.argv) > 2:  # one variant only (profiling): python tools/render_probe.py 1024 PERCH_FULL_CHUNK_MB=128
    run(sys.argv[2], **dict(kv.split("=") for kv in sys.argv[2:]))
    sys.exit(0)
base = run("legacy", PERCH_IMAGE_LEGACY=1)
new = run("full")
print("bit-equal:", all(np.array_equal(base[k], new[k]) for k in ("result_depth", "result_color")))
for mb in (32, 64, 128, 512):
    run(f"chunk{mb}", PERCH_FULL_CHUNK_MB=mb)
