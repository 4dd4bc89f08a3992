import sys, time
from pathlib import Path
ROOT = Path("/root/repo"); sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))
import numpy as np
import bench
from perception_b200 import synth
from perception_b200.renderer import Renderer
for name in ("c2", "c5"):
    r, wl = bench.build_workload(name, 0, 0, 2048)
    s = wl["scene"]
    r.close()
    r = Renderer(0)
    def T(f, n=3):
        ts = []
        for _ in range(n):
            t0 = time.perf_counter(); f(); r.synchronize(); ts.append((time.perf_counter() - t0) * 1e3)
        return [round(t, 2) for t in ts]
    print(name, "set_camera", T(lambda: r.set_camera(bench.W, bench.H, wl["proj"], synth.CX, synth.CY, synth.FX, synth.FY, synth.DEPTH_FACTOR)))
    print(name, "set_models", len(s.tris), T(lambda: r.set_models(s.tris, s.tris_model_count)))
    print(name, "set_scene", T(lambda: r.set_scene(s.source_depth, s.source_color, s.source_label if wl["dof6"] else None)))
    print(name, "set_observed", len(wl["obs"]["result_cloud_eigen"]), T(lambda: r.set_observed(wl["obs"]["result_cloud_eigen"], wl["obs"]["result_cloud_color"], wl["obs_label"])))
    kw = dict(stride=wl["stride"], cost_type=wl["cost_type"], calculate_observed_cost=True, sensor_resolution=wl["res"])
    print(name, "first COST call (grid build)", T(lambda: r.render_cuda_multi_unified("COST", wl["poses"], wl["pmm"], wl["plabel"], wl["tot"], **kw)))
    r.close()
