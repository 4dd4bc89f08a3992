"""Where does the host-API time go?  (run on the GPU box)"""
import sys, time
from pathlib import Path
import numpy as np
ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))
import bench
import torch
from perception_b200.renderer import topk
r, wl = bench.build_workload("c2", 0, 0, None)
P = wl["P"]
kw = dict(stride=wl["stride"], cost_type=wl["cost_type"], calculate_observed_cost=True, sensor_resolution=wl["res"])
pin = lambda a: torch.from_numpy(np.ascontiguousarray(a)).pin_memory().numpy()
hp, hm, hl, ht = pin(wl["poses"]), pin(wl["pmm"]), pin(wl["plabel"]), pin(wl["tot"])
for name, args in (("pinned", (hp, hm, hl, ht)), ("pageable", (wl["poses"], wl["pmm"], wl["plabel"], wl["tot"]))):
    for _ in range(3):
        out = r.render_cuda_multi_unified("COST", *args, **kw)
    t0 = time.perf_counter()
    for _ in range(20):
        out = r.render_cuda_multi_unified("COST", *args, **kw)
    t1 = time.perf_counter()
    for _ in range(20):
        hits = topk(out["rendered_cost"], out["observed_cost"], hm, 1, k=8)
    t2 = time.perf_counter()
    print(f"{name}: render_cuda_multi_unified {1e3*(t1-t0)/20:.3f} ms (kernel {r.last_timing()[0]:.3f} ms), topk {1e3*(t2-t1)/20:.3f} ms")
