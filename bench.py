#!/usr/bin/env python
"""bench.py -- scored poses/s of the fused render+cost hot path (BASELINE.json metric) on synthetic inputs.

  python bench.py --gpus N --steps K --warmup W            our arm   (torchrun for N > 1, one rank per GPU)
  python bench.py --impl reference ...                     the reference's CPU scoring path (oracle port, host cores)
  python bench.py --impl reference-cuda ...                the reference's own CUDA kernels (oracle/_ref) on the GPU

A step = one pass of the hot path over one batch of candidate poses per GPU (workload C2 of BASELINE.json configs[1]:
16 384 6-DoF poses of a 20 000-triangle mesh against a 640x480 scene, stride 8, label-restricted cost_type 2).
`value` times the kernel with inputs resident in HBM (CUDA events on the library's stream, L2 flushed between steps);
`e2e` times the public host API (pinned host poses in, costs + top-k out, NCCL gather of the per-GPU top-k for N > 1);
its steps are spread over --e2e-contexts contexts (default 2, one host thread each) so one step's copies overlap another's kernel.
Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))

METRIC = "scored poses/sec (render+cost, 640x480)"
UNIT = "poses/s"
W, H = 640, 480

WORKLOADS = {
    # name: (description, n_tris per mesh, poses per GPU per step, stride, dof6, cost_type, sensor_resolution, meshes)
    "c1": ("C1: 3-DoF, 5k-tri mesh, 2k poses, stride 4, cost_type 0", 5000, 2000, 4, False, 0, 0.0075, 1),
    "c2": ("C2: 6-DoF render+cost, 16384 poses x 20k-tri mesh, 640x480, stride 8, cost_type 2", 20000, 16384, 8, True, 2, 0.01, 1),
    # BASELINE configs[2] and [4] are quoted over 8 GPUs: the per-GPU share is the default, --poses runs any other size
    "c3": ("C3: multi-object 3-DoF, 5 meshes x 5k tris, 12500 poses per GPU (100k over 8), stride 4, cost_type 0", 5000, 12500,
           4, False, 0, 0.0075, 5),
    "c5": ("C5: 6-DoF sweep, 21 meshes x 20k tris, 125000 poses per GPU (1M over 8), stride 8, cost_type 2", 20000, 125000, 8,
           True, 2, 0.01, 21),
}


def rank_info():
    return int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))


# ---- clocks --------------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu, self.lines, self.proc = gpu_index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                                          "-i", str(self.gpu)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ---- workload ----------------------------------------------------------------------------------------------------------
def build_workload(name: str, rank: int, device: int, n_poses: int | None, use_product: bool = True):
    """Everything untimed: mesh, scene, observed cloud, poses.  With use_product the scene is rasterised and
    back-projected by libperch_b200 (GPU); the reference arms use the CPU oracle instead so that none of this repo's
    kernels touch their inputs."""
    from perception_b200 import synth
    from perception_b200.renderer import Renderer, compute_proj

    desc, n_tris, P, stride, dof6, cost_type, res, n_models = WORKLOADS[name]
    if n_poses:
        P = n_poses
    proj = compute_proj(synth.camera_K(), W, H)
    # mesh variants from seed 1000+m, r0 in [0.04, 0.12] m
    meshes = [synth.make_mesh(n_tris, seed=1000 + m, r0=0.09 if n_models == 1 else 0.05 + 0.06 * ((m * 7) % n_models) / n_models)
              for m in range(n_models)]
    r = None
    if use_product:
        r = Renderer(device)
        r.set_camera(W, H, proj, synth.CX, synth.CY, synth.FX, synth.FY, synth.DEPTH_FACTOR)

        def render_fn(tris, counts, poses, pmm):
            r.set_models(tris, counts)
            r.set_scene(np.zeros(W * H, np.int32))
            o = r.render_cuda_multi_unified("RENDER", poses, pmm)
            return o["result_depth"], o["result_color"]
    else:
        from oracle_lib import Oracle
        orc = Oracle()

        def render_fn(tris, counts, poses, pmm):
            return orc.render(tris, counts, poses, pmm, W, H, proj)

    scene = synth.make_scene(meshes, render_fn, n_instances=4 if n_models == 1 else min(n_models, 8), seed=42)
    cam = (synth.CX, synth.CY, synth.FX, synth.FY, synth.DEPTH_FACTOR)
    lab_img = scene.source_label if dof6 else None
    bounds = None if dof6 else synth.table_bounds(scene)
    tf = None if dof6 else np.linalg.inv(scene.T_cam_table)
    if use_product:
        r.set_models(scene.tris, scene.tris_model_count)
        r.set_scene(scene.source_depth, scene.source_color, lab_img)
        obs = r.depth2cloud_global(scene.source_depth, scene.source_color, W, H, 1, *cam, stride, lab_img, bounds, tf)
    else:
        o = orc.depth2cloud(scene.source_depth, scene.source_color, W, H, 1, *cam, stride, label_mask=lab_img, bounds=bounds, cam_tf=tf)
        obs = {"result_cloud_eigen": o["aos"], "result_cloud_color": o["rgb"], "rendered_cloud_point_num": o["n"],
               "result_observed_cloud_label": o["label"] if dof6 else None}
    obs_label = obs["result_observed_cloud_label"] if dof6 else None
    if use_product:
        r.set_observed(obs["result_cloud_eigen"], obs["result_cloud_color"], obs_label)
    n_obs = obs["rendered_cloud_point_num"]
    # pose list in per-model blocks, like the reference's per-model batches (search_env.cpp:2504-2525)
    pmm = (np.arange(P, dtype=np.int64) * n_models // P).astype(np.int32)
    n_inst = len(scene.gt_poses)
    if dof6:
        # a model is searched around the instance that shows it (models without an instance borrow one)
        inst_of = [next((g for g in range(n_inst) if scene.gt_models[g] == m), m % n_inst) for m in range(n_models)]
        poses = np.zeros((P, 4, 4), np.float32)
        plabel = np.zeros(P, np.int32)
        for m in range(n_models):
            sel = np.nonzero(pmm == m)[0]
            g = inst_of[m]
            poses[sel] = synth.poses_6dof(scene.gt_poses[g][:3, 3], len(sel), seed=3 + rank + 101 * m)
            plabel[sel] = g
        poses[0] = synth.pose_to_mat4x4(scene.gt_poses[inst_of[0]])
        tot = np.bincount(obs_label, minlength=n_inst).astype(np.float32)[plabel]
    else:
        poses = np.concatenate([synth.poses_3dof(scene, m, int((pmm == m).sum()), seed=1 + rank) for m in range(n_models)])
        plabel = None
        tot = synth.cylinder_counts(scene, obs["result_cloud_eigen"], poses, pmm)
    wl = dict(name=name, desc=desc, n_tris=n_tris, n_models=n_models, P=P, stride=stride, dof6=dof6, cost_type=cost_type, res=res, n_obs=n_obs,
              scene=scene, obs=obs, obs_label=obs_label, poses=poses, pmm=pmm, plabel=plabel, tot=tot, proj=proj)
    return r, wl


def algorithmic_bytes_per_pose(n_tris, n_obs):
    # SURVEY.md 8(d) / BASELINE.md 2c: 40*T_m + 64 + 4*W*H + 15*N_obs + 12
    return 40 * n_tris + 64 + 4 * W * H + 15 * n_obs + 12


def measured_peak_gbs():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        try:
            return float(json.loads(p.read_text())["hbm_gbs"]), "measured"
        except Exception:
            pass
    return 6650.0, "fallback"


# ---- arms ----------------------------------------------------------------------------------------------------------------
def run_ours(args):
    import torch
    import torch.distributed as dist
    from perception_b200 import sharding
    from perception_b200 import synth
    from perception_b200.renderer import Renderer, topk

    rank, local_rank, world = rank_info()
    if world > 1:
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    r, wl = build_workload(args.workload, rank, local_rank, args.poses)
    P = wl["P"]
    kw = dict(stride=wl["stride"], cost_type=wl["cost_type"], calculate_observed_cost=True, sensor_resolution=wl["res"],
              color_distance_threshold=20.0, occlusion_threshold=1.0)

    # device-resident inputs/outputs for `value`
    d_poses = r.device_alloc(P * 64); r.h2d(d_poses, wl["poses"])
    d_pmm = r.device_alloc(P * 4); r.h2d(d_pmm, wl["pmm"])
    d_lab = None
    if wl["plabel"] is not None:
        d_lab = r.device_alloc(P * 4); r.h2d(d_lab, wl["plabel"])
    d_tot = r.device_alloc(P * 4); r.h2d(d_tot, wl["tot"])
    d_costs = r.device_alloc(P * 12)
    d_counts = r.device_alloc(P * 12)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def device_step():
        r.l2_flush()
        r.render_cost_device(d_poses, P, d_pmm, d_lab, d_tot, d_costs, d_counts, **kw)
        r.synchronize()
        return r.last_timing()[0]

    for _ in range(args.warmup):
        device_step()
    sampler = ClockSampler(local_rank)
    barrier()
    sampler.start()
    t0 = time.perf_counter()
    kernel_ms = [device_step() for _ in range(args.steps)]
    barrier()
    wall_device_loop = time.perf_counter() - t0
    total_ms = float(np.sum(kernel_ms))

    # e2e: public host API, pinned host buffers, costs + per-model top-k back on the host, gather across ranks
    pin = lambda a: torch.from_numpy(np.ascontiguousarray(a)).pin_memory().numpy()  # noqa: E731
    h_poses, h_pmm, h_tot = pin(wl["poses"]), pin(wl["pmm"]), pin(wl["tot"])
    h_lab = pin(wl["plabel"]) if wl["plabel"] is not None else None
    K_TOP = 8
    exchange = sharding.HitExchange(1, K_TOP)

    # Every step is one call of the public API on pinned host buffers (H2D of the step's poses, kernel, D2H of its costs)
    # followed by the host top-k.  A caller with many batches to score keeps the GPU busy by driving more than one
    # context (each has its own stream) from its own host thread, so that one step's copies and host work overlap the
    # other's kernel: --e2e-contexts (default 2) such contexts share the steps round-robin; the top-k gather across
    # ranks runs on the main thread, in step order.
    n_ctx = max(1, min(args.e2e_contexts, args.steps))
    ctxs = [r]
    for _ in range(n_ctx - 1):
        r2 = Renderer(local_rank)
        r2.set_camera(W, H, wl["proj"], synth.CX, synth.CY, synth.FX, synth.FY, synth.DEPTH_FACTOR)
        r2.set_models(wl["scene"].tris, wl["scene"].tris_model_count)
        r2.set_scene(wl["scene"].source_depth, wl["scene"].source_color, wl["scene"].source_label if wl["dof6"] else None)
        r2.set_observed(wl["obs"]["result_cloud_eigen"], wl["obs"]["result_cloud_color"], wl["obs_label"])
        ctxs.append(r2)

    def score_step(rr):
        out = rr.render_cuda_multi_unified("COST", h_poses, h_pmm, h_lab, h_tot, **kw)
        return out, topk(out["rendered_cost"], out["observed_cost"], h_pmm, 1, k=K_TOP, pose_id_offset=rank * P)

    def e2e_steps(n_steps):
        """n_steps steps over the contexts; returns the last step's outputs and merged best hit."""
        results = [None] * n_steps
        done = [threading.Event() for _ in range(n_steps)]
        errors = []

        def worker(j):
            try:
                for i in range(j, n_steps, n_ctx):
                    results[i] = score_step(ctxs[j])
                    done[i].set()
            except BaseException as e:  # noqa: BLE001
                errors.append(e)
                for d in done:
                    d.set()

        threads = [threading.Thread(target=worker, args=(j,)) for j in range(n_ctx)]
        for t in threads:
            t.start()
        # the main thread gathers the finished steps in step order (collectives stay ordered across ranks) while the
        # workers are already scoring the next ones
        best = None
        for i in range(n_steps):
            done[i].wait()
            if errors:
                break
            best = sharding.merge_hits(exchange.gather(results[i][1]), K_TOP)[0, 0]  # all_gather of k x 16 B per rank
        for t in threads:
            t.join()
        if errors:
            raise errors[0]
        return results[-1][0], best

    out, best = e2e_steps(max(n_ctx, min(args.warmup, 3)))
    barrier()
    t0 = time.perf_counter()
    out, best = e2e_steps(args.steps)
    barrier()
    e2e_s = time.perf_counter() - t0
    clocks = sampler.stop()

    counts = np.zeros((3, P), np.int32)
    r.d2h(counts, d_counts)
    work = {"rendered_points_per_pose": float(counts[0].mean()), "n_obs": int(wl["n_obs"])}

    if world > 1:
        t = torch.tensor([total_ms, e2e_s], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        total_ms, e2e_s = float(t[0]), float(t[1])

    line = None
    if rank == 0:
        peak, peak_kind = measured_peak_gbs()
        ms_per_step = total_ms / args.steps
        value = world * P * args.steps / (total_ms / 1e3)
        bpp = algorithmic_bytes_per_pose(wl["n_tris"], wl["n_obs"])
        kern_ms = float(np.mean(kernel_ms))
        achieved = P * bpp / (kern_ms / 1e3) / 1e9
        traffic = None
        tfile = ROOT / "profiles" / "dram_traffic.json"
        if tfile.exists():
            try:
                traffic = json.loads(tfile.read_text()).get(args.workload)
            except Exception:
                traffic = None
        # what actually bounds the kernel: warp-instruction issue.  The per-launch instruction count is the
        # smsp__inst_executed.sum of the committed ncu capture of this workload (profiles/kernel_counters.json); the rate
        # is that count over the kernel time measured HERE, against SMs x 4 schedulers x the SM clock sampled in this run.
        issue = None
        cfile = ROOT / "profiles" / "kernel_counters.json"
        if cfile.exists():
            try:
                kc = json.loads(cfile.read_text()).get(args.workload)
                if kc and kc.get("poses") == P and clocks.get("sm_mhz"):
                    sms = int(kc.get("sm_count", 148))
                    rate = kc["warp_inst_per_launch"] / (kern_ms / 1e3)
                    peak_rate = sms * 4 * clocks["sm_mhz"] * 1e6
                    issue = {"bound": "warp-instruction issue", "warp_inst_per_pose": kc["warp_inst_per_launch"] / P,
                             "achieved": rate / 1e9, "peak": peak_rate / 1e9, "unit": "G warp-inst/s", "frac": rate / peak_rate,
                             "source": kc.get("source")}
            except Exception:
                issue = None
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": wl["desc"], "poses_per_gpu_per_step": P, "triangles": wl["n_tris"], "meshes": wl["n_models"], "image": [W, H],
                       "stride": wl["stride"], "n_obs": int(wl["n_obs"]), "l2": "flushed between steps (256 MiB memset)",
                       "timing": "CUDA events on the library stream around the fused kernel, max over ranks"},
            "clocks": clocks,
            "e2e": {"value": world * P * args.steps / e2e_s, "unit": UNIT, "h2d_bytes_per_step": int(P * (64 + 4 + 4 + (4 if h_lab is not None else 0))),
                    "d2h_bytes_per_step": int(P * 24), "ms_per_scene": e2e_s / args.steps * 1e3,
                    "contexts": n_ctx,
                    "how": "every step is one blocking call of the public host API (H2D of its poses, kernel, D2H of its costs) "
                           "+ host top-k; steps alternate between `contexts` contexts, one host thread each"},
            "gpu_launches": args.steps,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic, "peak_kind": f"{peak_kind} copy bandwidth", "bytes_per_pose": bpp,
                         "poses_per_launch": P, "kernel": "pose_kernel<MODE_COST>", "kernel_ms": kern_ms},
            "issue_roofline": issue,
            "work": work,
            "best": {"cost": int(best["cost"]), "pose_id": int(best["pose_id"])},
            "wall_s_device_loop": wall_device_loop,
        }
        single = wl["n_models"] == 1  # the GICP / glue legs are written for the one-mesh workloads
        if world == 1 and not args.no_gicp and single:
            line["gicp"] = gicp_leg(r, wl, out, kw)
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline(wl, budget_s=args.cpu_budget)
        if world == 1 and not args.no_glue and single:
            line["glue"] = glue_leg(r, wl, kw)  # last: re-targets the context at a 3-DoF scene
        if world == 1 and not args.no_ref_cuda:
            line["reference_cuda"] = reference_cuda(wl, batch=args.ref_batch)
        emit(line)
    if world > 1:
        dist.destroy_process_group()
    return line


def gicp_leg(r, wl, out, kw, n_top=4096, iterations=20, reps=3):
    """SURVEY 8(d) GICP measurement (C4-shaped): the n_top best poses of the scored batch by (rendered + observed) cost,
    refined by batched GICP with k = 10 and a FIXED number of iterations (epsilons 0 so every iteration runs).
    Device time of the GICP kernel alone (CUDA events inside the library) and wall time of the whole perch_gicp call
    (lattice render + clouds + GICP + H2D/D2H)."""
    P = wl["P"]
    cost = np.where(out["rendered_cost"] >= 0, out["rendered_cost"] + np.nan_to_num(out["observed_cost"], nan=1e9), 1e9)
    sel = np.sort(np.argsort(cost, kind="stable")[:min(n_top, P)])
    poses, pmm = wl["poses"][sel], wl["pmm"][sel]
    lab = wl["plabel"][sel] if wl["plabel"] is not None else None
    args = dict(stride=kw["stride"], sensor_resolution=kw["sensor_resolution"], occlusion_threshold=kw["occlusion_threshold"],
                icp_max_iterations=iterations, icp_k_correspondences=10, icp_rotation_epsilon=0.0, icp_transformation_epsilon=0.0)
    g = r.gicp(poses, pmm, lab, **args)  # warm-up (target normals are cached after the first call)
    dev, wall = [], []
    for _ in range(reps):
        t0 = time.perf_counter()
        g = r.gicp(poses, pmm, lab, **args)
        wall.append(time.perf_counter() - t0)
        dev.append(r.last_timing()[0])
    k_ms, w_s = float(np.median(dev)), float(np.median(wall))
    n = len(sel)
    res = {"refined_poses_per_s": n / w_s, "kernel_poses_per_s": n / (k_ms / 1e3), "poses": n, "iterations": iterations,
           "kernel_ms": k_ms, "ms_total": w_s * 1e3, "k_correspondences": 10,
           "mean_iterations_run": float(g["iterations"].mean()), "n_obs": int(wl["n_obs"]),
           "note": "top poses of the scored batch; fixed iteration count (epsilons 0)"}
    # BASELINE configs[3]: the same poses against the FULL-resolution scene cloud (stride 1, no label restriction)
    from perception_b200 import synth
    s = wl["scene"]
    cam = (synth.CX, synth.CY, synth.FX, synth.FY, synth.DEPTH_FACTOR)
    dense = r.depth2cloud_global(s.source_depth, s.source_color, W, H, 1, *cam, 1)
    t0 = time.perf_counter()
    r.set_observed(dense["result_cloud_eigen"], dense["result_cloud_color"], None)
    g = r.gicp(poses, pmm, None, **args)  # first call: target normals (10-NN of every scene point) + the search grids
    first_s = time.perf_counter() - t0
    dev, wall = [], []
    for _ in range(reps):
        t0 = time.perf_counter()
        g = r.gicp(poses, pmm, None, **args)
        wall.append(time.perf_counter() - t0)
        dev.append(r.last_timing()[0])
    k_ms, w_s = float(np.median(dev)), float(np.median(wall))
    res["scene_cloud"] = {"n_obs": int(dense["rendered_cloud_point_num"]), "refined_poses_per_s": n / w_s,
                          "kernel_poses_per_s": n / (k_ms / 1e3), "kernel_ms": k_ms, "ms_total": w_s * 1e3,
                          "first_call_ms": first_s * 1e3, "mean_iterations_run": float(g["iterations"].mean()),
                          "note": "target = every valid pixel of the 640x480 scene (C4); first call also builds the "
                                  "target normals and grids"}
    r.set_observed(wl["obs"]["result_cloud_eigen"], wl["obs"]["result_cloud_color"], wl["obs_label"])
    return res


def _quat_from_R(R):
    """(N,3,3) rotations -> (N,4) quaternions (x, y, z, w), Shepperd's method."""
    R = np.asarray(R, np.float64)
    q = np.zeros((len(R), 4))
    tr = R[:, 0, 0] + R[:, 1, 1] + R[:, 2, 2]
    for i in range(len(R)):
        m = R[i]
        if tr[i] > 0:
            s = np.sqrt(tr[i] + 1.0) * 2
            q[i] = [(m[2, 1] - m[1, 2]) / s, (m[0, 2] - m[2, 0]) / s, (m[1, 0] - m[0, 1]) / s, 0.25 * s]
        else:
            k = int(np.argmax([m[0, 0], m[1, 1], m[2, 2]]))
            a, b, c = k, (k + 1) % 3, (k + 2) % 3
            s = np.sqrt(1.0 + m[a, a] - m[b, b] - m[c, c]) * 2
            v = np.zeros(4)
            v[a] = 0.25 * s
            v[b] = (m[b, a] + m[a, b]) / s
            v[c] = (m[c, a] + m[a, c]) / s
            v[3] = (m[c, b] - m[b, c]) / s
            q[i] = v
    return q


def glue_leg(r, wl, kw, reps=5):
    """SURVEY 8(f) rows 1-2, measured next to the hot path (N=1):
    (a) the workload's pose batch as ObjectStates (poses.txt layout: x y z qx qy qz qw, camera frame) through
        perch_score_states -- matrix assembly + label totals on the device, then the same fused kernel;
    (b) a 3-DoF scene from nothing: perch_generate_states (grid + IsValidPose on the device) followed by
        perch_score_states on the resident list, against the CPU restatement of the same two host loops."""
    from oracle_lib import StatesOracle
    from perception_b200 import states as S, synth
    out = {}
    P, scene = wl["P"], wl["scene"]
    cam_to_body = np.array([[0, 0, 1, 0], [-1, 0, 0, 0], [0, -1, 0, 0], [0, 0, 0, 1.0]])
    verts = scene.tris["v"].reshape(-1, 3)
    if wl["dof6"]:
        T = wl["poses"].astype(np.float64)
        q = _quat_from_R(T[:, :3, :3] / 100.0)
        st = np.zeros(P, S.STATE_DTYPE)
        st["x"], st["y"], st["z"] = (T[:, k, 3] / 100.0 for k in range(3))
        st["qx"], st["qy"], st["qz"], st["qw"] = q.T
        st["segmentation_label_id"] = wl["plabel"] + 1
        env = S.Env(cam_to_world=np.linalg.inv(cam_to_body), use_external_pose_list=1, min_neighbor_points_for_valid_pose=50)
        metas = [S.ModelMeta.from_vertices(verts)]
        r.set_env(env, metas)
        lab = wl["obs_label"]
        xyz = wl["obs"]["result_cloud_eigen"]
        parts, ranges = [xyz], [[0, len(xyz)]]
        for l in range(int(lab.max()) + 1):
            pts = xyz[lab == l]
            ranges.append([ranges[-1][1], ranges[-1][1] + len(pts)])
            parts.append(pts)
        r.set_search_cloud(np.concatenate(parts), np.array(ranges, np.int32))
        r.score_states("COST", st, **kw)
        t = []
        for _ in range(reps):
            t0 = time.perf_counter()
            o = r.score_states("COST", st, **kw)
            t.append(time.perf_counter() - t0)
        ref = r.render_cuda_multi_unified("COST", wl["poses"], wl["pmm"], wl["plabel"], wl["tot"], **kw)
        t0 = time.perf_counter()
        valid = r.validate_states(st)
        t_val = time.perf_counter() - t0
        orc = StatesOracle(env, metas, np.concatenate(parts), np.array(ranges, np.int32))
        n_cpu = min(P, 4096)
        t0 = time.perf_counter()
        orc.assemble(st[:n_cpu], seg_counts=np.bincount(lab).astype(np.float32))
        orc.validate(st[:n_cpu])
        t_cpu = time.perf_counter() - t0
        out["score_states_6dof"] = {
            "poses_per_s": P / float(np.median(t)), "ms_per_scene": float(np.median(t)) * 1e3, "states": P,
            "h2d_bytes_per_step": int(P * 88), "d2h_bytes_per_step": int(P * 24),
            "rendered_cost_equal_to_matrix_path": float(np.mean(o["rendered_cost"] == ref["rendered_cost"])),
            "validate_states_per_s": P / t_val, "valid": int(valid.sum()),
            "cpu_restatement_states_per_s": n_cpu / t_cpu, "cpu_sample": f"assemble + IsValidPose of {n_cpu} states, 1 thread"}
    # (b) 3-DoF greedy scene: table bounds +-0.3 m at 5 mm, 32 yaws, one model
    T_table_cam = np.linalg.inv(scene.T_cam_table)
    cam = (synth.CX, synth.CY, synth.FX, synth.FY, synth.DEPTH_FACTOR)
    stride = wl["stride"]
    obs = r.depth2cloud_global(scene.source_depth, scene.source_color, W, H, 1, *cam, stride, None, synth.table_bounds(scene),
                               T_table_cam)
    r.set_scene(scene.source_depth, scene.source_color, None)
    r.set_observed(obs["result_cloud_eigen"], obs["result_cloud_color"], None)
    zrest = float(scene.model_zrest[0])
    pre = np.eye(4, dtype=np.float32)
    pre[2, 3] = zrest
    vp = (verts.astype(np.float64) + [0, 0, zrest]).astype(np.float32)
    env = S.Env(cam_to_world=T_table_cam @ np.linalg.inv(cam_to_body), x_min=-0.3, x_max=0.3, y_min=-0.3, y_max=0.3,
                table_height=0.0, res=0.005, theta_res=np.pi / 16, min_neighbor_points_for_valid_pose=50)
    metas = [S.ModelMeta.from_vertices(vp, pre)]
    world = (obs["result_cloud_eigen"].astype(np.float64) @ T_table_cam[:3, :3].T + T_table_cam[:3, 3]).astype(np.float32)
    world[:, 2] = 0.0
    r.set_env(env, metas)
    r.set_search_cloud(world)
    kw3 = dict(kw, cost_type=0)
    r.generate_states(fetch=False)
    r.score_states("COST", None, **kw3)
    tg, ts = [], []
    for _ in range(reps):
        t0 = time.perf_counter()
        r.generate_states(fetch=False)
        t1 = time.perf_counter()
        o = r.score_states("COST", None, **kw3)
        t2 = time.perf_counter()
        tg.append(t1 - t0)
        ts.append(t2 - t1)
    n_gen = r.n_generated
    n_cand = 121 * 121 * 32
    orc = StatesOracle(env, metas, world)
    t0 = time.perf_counter()
    gen_cpu = orc.generate()
    t_cpu_gen = time.perf_counter() - t0
    assert len(gen_cpu) == n_gen, (len(gen_cpu), n_gen)
    best = int(np.argmin(np.where(o["rendered_cost"] >= 0, o["rendered_cost"] + o["observed_cost"], 1e9)))
    out["greedy_3dof"] = {
        "grid_candidates": n_cand, "valid_states": n_gen, "generate_ms": float(np.median(tg)) * 1e3,
        "score_ms": float(np.median(ts)) * 1e3, "ms_per_scene": float(np.median(tg) + np.median(ts)) * 1e3,
        "candidates_per_s": n_cand / float(np.median(tg)), "scored_poses_per_s": n_gen / float(np.median(ts)),
        "h2d_bytes_per_step": 0, "d2h_bytes_per_step": int(n_gen * 24), "n_obs": int(obs["rendered_cloud_point_num"]),
        "cpu_restatement_generate_ms": t_cpu_gen * 1e3, "cpu_threads": 1, "best_pose_id": best,
        "note": "one model, +-0.3 m at 5 mm x 32 yaws; list generated, validated, assembled and scored on the device"}
    return out


def _baseline_inputs(wl):
    s = wl["scene"]
    if wl["dof6"]:
        return dict(center=None, radius=None, t_tc=None)
    T_tc = np.linalg.inv(s.T_cam_table)
    t = wl["poses"][:, :3, 3].astype(np.float64) / 100.0
    c = (t @ T_tc[:3, :3].T + T_tc[:3, 3])[:, :2].astype(np.float32)
    return dict(center=np.ascontiguousarray(c), radius=np.ascontiguousarray(np.asarray(s.model_radius)[wl["pmm"]], np.float32),
                t_tc=np.ascontiguousarray(T_tc[:3, :4], dtype=np.float32))


def cpu_score(wl, n):
    """Time the CPU port of the reference's non-GPU scoring path (oracle/cpu_baseline.c) on the first n poses."""
    import ctypes as C
    from oracle_lib import ORACLE_SO
    from perception_b200.renderer import TRIANGLE_DTYPE
    lib = C.CDLL(str(ORACLE_SO))
    lib.cpu_baseline_threads.restype = C.c_int
    s = wl["scene"]
    b = _baseline_inputs(wl)
    vp = C.c_void_p
    p = lambda a: None if a is None else a.ctypes.data_as(vp)  # noqa: E731
    tris = np.ascontiguousarray(s.tris, dtype=TRIANGLE_DTYPE)
    cnt = np.ascontiguousarray(s.tris_model_count, np.int32)
    # the sample: the first n poses of a one-mesh workload, n poses spread evenly over the per-model blocks otherwise
    sel = np.arange(n) if wl["n_models"] == 1 else np.linspace(0, wl["P"] - 1, n).astype(np.int64)
    poses = np.ascontiguousarray(wl["poses"][sel], np.float32)
    pmm = np.ascontiguousarray(wl["pmm"][sel], np.int32)
    sd = np.ascontiguousarray(s.source_depth, np.int32)
    sl = np.ascontiguousarray(s.source_label, np.uint8) if wl["dof6"] else None
    pl = np.ascontiguousarray(wl["plabel"][sel], np.int32) if wl["dof6"] else None
    if b["center"] is not None:
        b = dict(b, center=np.ascontiguousarray(b["center"][sel]), radius=np.ascontiguousarray(b["radius"][sel]))
    ox = np.ascontiguousarray(wl["obs"]["result_cloud_eigen"], np.float32)
    ol = np.ascontiguousarray(wl["obs_label"], np.int32) if wl["dof6"] else None
    proj = np.ascontiguousarray(wl["proj"], np.float32)
    from perception_b200 import synth
    tc = np.zeros(n, np.int32); scst = np.zeros(n, np.int32)
    t0 = time.perf_counter()
    lib.cpu_baseline_score(p(tris), C.c_int(len(tris)), p(poses), C.c_int(n), p(pmm), p(cnt), C.c_int(len(cnt)), C.c_int(W),
                           C.c_int(H), p(proj), p(sd), p(sl), p(pl), C.c_float(synth.CX), C.c_float(synth.CY), C.c_float(synth.FX),
                           C.c_float(synth.FY), C.c_float(100.0), C.c_int(wl["stride"]), p(ox), p(ol), C.c_int(len(ox)),
                           p(b["center"]), p(b["radius"]), p(b["t_tc"]), C.c_float(wl["res"]), C.c_float(1.0), C.c_int(10),
                           p(tc), p(scst))
    dt = time.perf_counter() - t0
    return dt, lib.cpu_baseline_threads(), tc, scst


def cpu_baseline(wl, budget_s=15.0):
    dt, cores, _, _ = cpu_score(wl, 32)
    n = int(min(wl["P"], max(32, 32 * budget_s / max(dt, 1e-3))))
    dt, cores, _, _ = cpu_score(wl, n)
    return {"value": n / dt, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": f"{'first ' if wl['n_models'] == 1 else ''}{n} poses of the same workload"
                      f"{'' if wl['n_models'] == 1 else ' (evenly over the meshes)'}, {dt:.1f} s, OpenMP over poses "
                      "(oracle/cpu_baseline.c)"}


def reference_cuda(wl, batch=700):
    """The reference's own CUDA kernels (oracle/_ref/libperch_ref.so) on the same inputs: per-call upload + alloc like
    the reference, shipped gpu_batch_size 700 (pr3_env_config.yaml:10).  Reported beside, same GPU, same job."""
    from oracle_lib import REF_SO, Reference
    if not REF_SO.exists():
        return {"unavailable": "oracle/_ref/libperch_ref.so not built"}
    ref = Reference()
    s = wl["scene"]
    from perception_b200 import synth
    n = min(batch, wl["P"])
    a = dict(tris=s.tris, tris_model_count=s.tris_model_count, poses=wl["poses"][:n], pose_model_map=wl["pmm"][:n], width=W,
             height=H, proj=wl["proj"], source_depth=s.source_depth, source_color=s.source_color,
             source_label=s.source_label if wl["dof6"] else None, pose_label=wl["plabel"][:n] if wl["dof6"] else None,
             stride=wl["stride"], depth_factor=100, cx=synth.CX, cy=synth.CY, fx=synth.FX, fy=synth.FY,
             obs_xyz=wl["obs"]["result_cloud_eigen"], obs_rgb=wl["obs"]["result_cloud_color"],
             obs_label=wl["obs_label"] if wl["dof6"] else None, pose_obs_total=wl["tot"][:n], cost_type=wl["cost_type"],
             calc_observed=True, sensor_resolution=wl["res"], color_distance_threshold=20.0, occlusion_threshold=1.0)
    ref.render_cost(**a)  # warm-up (context, allocator)
    times, phases = [], []
    for _ in range(3):
        t0 = time.perf_counter()
        q = ref.render_cost(**a)
        times.append(time.perf_counter() - t0)
        phases.append(q["ms"])
    dt = float(np.median(times))
    ph = np.median(np.array(phases), 0)
    return {"value": n / dt, "unit": UNIT, "batch": n, "ms_per_batch": dt * 1e3,
            "phase_ms": {"render": float(ph[0]), "cloud": float(ph[1]), "knn": float(ph[2]), "cost": float(ph[3])},
            "note": "reference .cuh kernels compiled unmodified for sm_100a + restated orchestration and brute-force NN"}


def run_reference_cpu(args):
    """--impl reference: the reference's CPU scoring path (port) on the host cores, bounded sample per step."""
    rank, local_rank, world = rank_info()
    if rank != 0:
        return
    r, wl = build_workload(args.workload, 0, local_rank, args.poses, use_product=False)
    dt, cores, _, _ = cpu_score(wl, 16)
    per_step = int(min(wl["P"], max(16, 16 * (60.0 / max(args.steps + args.warmup, 1)) / max(dt, 1e-3))))
    for _ in range(args.warmup):
        cpu_score(wl, per_step)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        cpu_score(wl, per_step)
    el = time.perf_counter() - t0
    v = per_step * args.steps / el
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": el / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": wl["desc"], "poses_per_step_sample": per_step, "triangles": wl["n_tris"], "image": [W, H],
                       "stride": wl["stride"], "n_obs": int(wl["n_obs"])},
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                             "sample": f"{per_step} poses per step of the same workload (oracle/cpu_baseline.c, OpenMP)"},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit(line)


def run_reference_cuda(args):
    rank, local_rank, world = rank_info()
    if rank != 0:
        return
    r, wl = build_workload(args.workload, 0, local_rank, args.poses, use_product=False)
    out = reference_cuda(wl, batch=args.ref_batch)
    emit({"impl": "reference-cuda", "metric": METRIC, "unit": UNIT, **out, "config": {"workload": wl["desc"]}})


_REAL_STDOUT = None


def emit(line: dict):
    """The ONE JSON line goes to the process's original stdout; everything else any library prints while the bench
    runs (e.g. NCCL's version banner) has been routed to stderr by main()."""
    data = (json.dumps(line) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(data.decode()); sys.stdout.flush()
    else:
        os.write(_REAL_STDOUT, data)


def main():
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--e2e-contexts", type=int, default=2,
                    help="contexts (one host thread each) the e2e leg spreads its steps over; 1 = strictly serial calls")
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference", "reference-cuda"])
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS))
    ap.add_argument("--poses", type=int, default=None, help="override poses per GPU per step")
    ap.add_argument("--cpu-budget", type=float, default=15.0)
    ap.add_argument("--ref-batch", type=int, default=700)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-ref-cuda", action="store_true")
    ap.add_argument("--no-gicp", action="store_true")
    ap.add_argument("--no-glue", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 0)
    if args.impl == "reference":
        return run_reference_cpu(args)
    if args.impl == "reference-cuda":
        return run_reference_cuda(args)
    run_ours(args)


if __name__ == "__main__":
    main()
