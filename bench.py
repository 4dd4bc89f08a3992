#!/usr/bin/env python
"""bench.py -- scored poses/s of the fused render+cost hot path (BASELINE.json metric) on synthetic inputs.

  python bench.py --gpus N --steps K --warmup W            our arm   (torchrun for N > 1, one rank per GPU)
  python bench.py --impl reference ...                     the reference's CPU scoring path (oracle port, host cores)
  python bench.py --impl reference-cuda ...                the reference's own CUDA kernels (oracle/_ref) on the GPU

Workload C2 of BASELINE.json configs[1]: 16 384 6-DoF poses of a 20 000-triangle mesh against a 640x480 scene, stride 8,
label-restricted cost_type 2 -- one such batch is a SCENE.  A step = --batches (default 40) scenes with DISTINCT pose
batches per GPU, so that a step is >= 50 ms of device work and the timed region >= 1 s.
`value`  : poses resident in HBM, the step's fused-kernel launches back to back on the library stream, CUDA events around
           them, L2 flushed once per step.
`e2e`    : every scene through the public host call perch_score_scene: pinned host poses in (H2D), fused kernel, per-model
           top-k on the device, NCCL all-gather of the per-rank records (N > 1), costs + merged hits back (D2H).  Reported
           strictly serial (one context, blocking calls) and pipelined (--e2e-contexts contexts, one host thread each).
After the timed loops rank 0 checks what the timed launch of scene 0 produced against the CPU oracle (`parity_check`) and
against the library's unfused kernels on every pose (`unfused_check`).  Prints ONE JSON line on rank 0.
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
def build_workload(name: str, rank: int, device: int, n_poses: int | None, use_product: bool = True, n_batches: int = 1,
                   compute_proj_fn=None):
    """Everything untimed: mesh, scene, observed cloud, poses.  With use_product the scene is rasterised and
    back-projected by libperch_b200 (GPU); the reference arms use the CPU oracle instead so that none of this repo's
    kernels touch their inputs."""
    from perception_b200 import synth

    desc, n_tris, P, stride, dof6, cost_type, res, n_models = WORKLOADS[name]
    if n_poses:
        P = n_poses
    if compute_proj_fn is None:
        from perception_b200.renderer import compute_proj as compute_proj_fn
    if use_product:
        from perception_b200.renderer import Renderer
    proj = compute_proj_fn(synth.camera_K(), W, H)
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
    # further scenes of a step: the same hypotheses displaced (distinct pose batches, same shape of work)
    poses_b, tot_b = [poses], [tot]
    for b in range(1, n_batches):
        rng = np.random.default_rng(7919 * b + rank)
        pb = poses.copy()
        if dof6:
            pb[:, :3, 3] += rng.uniform(-1.0, 1.0, size=(P, 3)).astype(np.float32)          # +-1 cm (pose units are cm)
            tb = tot
        else:
            # rotate every hypothesis about the table normal through the table origin and shift it along the table
            a = 0.002 * b
            Rz = np.array([[np.cos(a), -np.sin(a), 0, 0.0005 * b], [np.sin(a), np.cos(a), 0, 0], [0, 0, 1, 0], [0, 0, 0, 1.0]])
            M = scene.T_cam_table @ Rz @ np.linalg.inv(scene.T_cam_table)
            q = poses.astype(np.float64)
            q[:, :3, :] /= 100.0
            q = M[None] @ q
            q[:, :3, :] *= 100.0
            pb = q.astype(np.float32)
            tb = synth.cylinder_counts(scene, obs["result_cloud_eigen"], pb, pmm)
        poses_b.append(pb)
        tot_b.append(tb)
    wl = dict(name=name, desc=desc, n_tris=n_tris, n_models=n_models, P=P, stride=stride, dof6=dof6, cost_type=cost_type, res=res, n_obs=n_obs,
              scene=scene, obs=obs, obs_label=obs_label, poses=poses, pmm=pmm, plabel=plabel, tot=tot, proj=proj,
              poses_b=poses_b, tot_b=tot_b)
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
KERNEL_SOURCES = ("perception_b200/csrc/pose_kernels.cu", "perception_b200/csrc/exact_math.cuh", "perception_b200/csrc/kernels.cuh",
                  "perception_b200/csrc/context.hpp", "perception_b200/csrc/clusters.hpp", "perception_b200/csrc/clusters.cpp", "Makefile")


def kernel_source_hash() -> str:
    """sha256 over the sources that decide the fused kernel's instruction count: the committed ncu counters
    (profiles/kernel_counters.json) are only used when they were captured from exactly these sources."""
    import hashlib
    h = hashlib.sha256()
    for f in KERNEL_SOURCES:
        h.update((ROOT / f).read_bytes())
    return h.hexdigest()[:16]


def rooflines(args, wl, P, kern_ms_per_launch, clocks):
    """The bound that applies (warp-instruction issue) as `roofline`; SURVEY 8(d)'s HBM contract figure kept
    beside it, labelled for what it is (the working set is L2-resident, so it is not a bound on this design)."""
    peak, peak_kind = measured_peak_gbs()
    bpp = algorithmic_bytes_per_pose(wl["n_tris"], wl["n_obs"])
    achieved = P * bpp / (kern_ms_per_launch / 1e3) / 1e9
    traffic = None
    tfile = ROOT / "profiles" / "dram_traffic.json"
    if tfile.exists():
        try:
            traffic = json.loads(tfile.read_text()).get(args.workload)
        except Exception:
            traffic = None
    hbm = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
           "peak_kind": f"{peak_kind} copy bandwidth", "bytes_per_pose": bpp, "poses_per_launch": P,
           "note": "SURVEY 8(d) algorithmic bytes / kernel time. Inapplicable as a bound: mesh, scene and observed cloud are "
                   "L2-resident across poses" + (f", measured DRAM traffic {traffic} B per launch" if traffic else "")}
    roof = dict(hbm, traffic=traffic, kernel="pose_kernel<MODE_COST>", kernel_ms=kern_ms_per_launch)
    cfile = ROOT / "profiles" / "kernel_counters.json"
    kc = None
    if cfile.exists():
        try:
            kc = json.loads(cfile.read_text()).get(args.workload)
        except Exception:
            kc = None
    src = kernel_source_hash()
    if kc and kc.get("poses") == P and kc.get("source_sha16") == src and clocks.get("sm_mhz"):
        sms = int(kc.get("sm_count", 148))
        rate = kc["warp_inst_per_launch"] / (kern_ms_per_launch / 1e3)
        peak_rate = sms * 4 * clocks["sm_mhz"] * 1e6
        roof = {"bound": "issue", "achieved": rate / 1e9, "peak": peak_rate / 1e9, "unit": "G warp-inst/s",
                "frac": rate / peak_rate, "traffic": traffic, "kernel": "pose_kernel<MODE_COST>", "kernel_ms": kern_ms_per_launch,
                "warp_inst_per_pose": kc["warp_inst_per_launch"] / P, "poses_per_launch": P,
                "peak_kind": f"{sms} SMs x 4 schedulers x {clocks['sm_mhz']:.0f} MHz sampled in this run",
                "counter_source": kc.get("source"), "source_sha16": src, "hbm_contract": hbm}
    else:
        roof["note_issue"] = ("no warp-instruction count for these kernel sources (profiles/kernel_counters.json is for "
                              f"{kc.get('source_sha16') if kc else None}, sources are {src}): issue roofline omitted")
    return roof


def run_ours(args):
    import torch
    import torch.distributed as dist
    from perception_b200 import sharding
    from perception_b200 import synth
    from perception_b200.renderer import HIT_DTYPE, Renderer, comm_unique_id

    rank, local_rank, world = rank_info()
    if world > 1:
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    strong = args.scaling == "strong"
    B = max(1, args.batches)
    n_poses = args.poses
    P_scene = None
    if strong:
        # BASELINE configs[2] / [4]: the scene's pose list is fixed and cut into contiguous blocks, one per GPU
        P_scene = args.poses or {"c3": 100000, "c5": 1000000}.get(args.workload, WORKLOADS[args.workload][2] * 8)
        lo, hi = sharding.shard_range(P_scene, rank, world)
        n_poses = hi - lo
    r, wl = build_workload(args.workload, rank, local_rank, n_poses, n_batches=B)
    P = wl["P"]
    n_models = wl["n_models"]
    kw = dict(stride=wl["stride"], cost_type=wl["cost_type"], calculate_observed_cost=True, sensor_resolution=wl["res"],
              color_distance_threshold=20.0, occlusion_threshold=1.0)
    composed = None
    if args.workload == "c3" and not args.no_compose:
        # "inter-object occlusion" (BASELINE configs[2]): one object is already placed; it is rendered into the source
        # image on the device (perch_compose_source, the reference's tree flow) and occludes every candidate behind it
        sc = wl["scene"]
        composed = dict(poses=synth.pose_to_mat4x4(sc.gt_poses[0])[None], pmm=np.array([sc.gt_models[0]], np.int32))
        r.compose_source(composed["poses"], composed["pmm"], None, mode="merge")
        wl["oracle_source_depth"] = r.get_source_depth().reshape(-1)  # composed image (pinned by tests/test_compose.py)

    # ---- `value`: device-resident inputs / outputs, B launches per step ---------------------------------------------------
    d_poses, d_tot, d_costs, d_counts = [], [], [], []
    for b in range(B):
        dp = r.device_alloc(P * 64); r.h2d(dp, wl["poses_b"][b]); d_poses.append(dp)
        dt = r.device_alloc(P * 4); r.h2d(dt, wl["tot_b"][b]); d_tot.append(dt)
        d_costs.append(r.device_alloc(P * 12))
        d_counts.append(r.device_alloc(P * 12))
    d_pmm = r.device_alloc(P * 4); r.h2d(d_pmm, wl["pmm"])
    d_lab = None
    if wl["plabel"] is not None:
        d_lab = r.device_alloc(P * 4); r.h2d(d_lab, wl["plabel"])

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def device_step():
        r.l2_flush()
        r.timer_start()
        for b in range(B):
            r.render_cost_device(d_poses[b], P, d_pmm, d_lab, d_tot[b], d_costs[b], d_counts[b], **kw)
        r.timer_stop()
        r.synchronize()
        return r.timer_elapsed()

    for _ in range(args.warmup):
        device_step()
    sampler = ClockSampler(local_rank)
    barrier()
    sampler.start()
    t0 = time.perf_counter()
    step_ms = [device_step() for _ in range(args.steps)]
    barrier()
    wall_device_loop = time.perf_counter() - t0
    total_ms = float(np.sum(step_ms))
    counts0 = np.zeros((3, P), np.int32)
    r.d2h(counts0, d_counts[0])          # what the TIMED launch of scene 0 produced
    costs0 = np.zeros((3, P), np.float32)
    r.d2h(costs0, d_costs[0])

    # ---- `e2e`: the public host call, pinned host buffers in and out --------------------------------------------------------
    pin = lambda a: torch.from_numpy(np.ascontiguousarray(a)).pin_memory().numpy()  # noqa: E731
    h_poses = [pin(x) for x in wl["poses_b"]]
    h_tot = [pin(x) for x in wl["tot_b"]]
    h_pmm = pin(wl["pmm"])
    h_lab = pin(wl["plabel"]) if wl["plabel"] is not None else None
    K_TOP = args.topk
    offset = (sharding.shard_range(P_scene, rank, world)[0] if strong else rank * P)

    def pinned_out():
        return {"rendered_cost": pin(np.zeros(P, np.float32)), "observed_cost": pin(np.zeros(P, np.float32)),
                "points_diff_cost": pin(np.zeros(P, np.float32)), "counts": pin(np.zeros((3, P), np.int32))}

    def new_comm(rr):
        if world == 1:
            return
        ids = [comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(ids, src=0)
        rr.comm_init(world, rank, ids[0])

    def clone_context():
        r2 = Renderer(local_rank)
        r2.set_camera(W, H, wl["proj"], synth.CX, synth.CY, synth.FX, synth.FY, synth.DEPTH_FACTOR)
        r2.set_models(wl["scene"].tris, wl["scene"].tris_model_count)
        r2.set_scene(wl["scene"].source_depth, wl["scene"].source_color, wl["scene"].source_label if wl["dof6"] else None)
        r2.set_observed(wl["obs"]["result_cloud_eigen"], wl["obs"]["result_cloud_color"], wl["obs_label"])
        if composed:
            r2.compose_source(composed["poses"], composed["pmm"], None, mode="merge")
        return r2

    n_ctx = max(1, args.e2e_contexts)
    ctxs = [r] + [clone_context() for _ in range(n_ctx - 1)]
    outs = [pinned_out() for _ in ctxs]
    for rr in ctxs:
        new_comm(rr)

    def score_scene(j, b):
        """One scene = one blocking call of the public API on context j: H2D, kernel, device top-k, all-gather, D2H."""
        hits, _ = ctxs[j].score_scene(h_poses[b], h_pmm, h_lab, h_tot[b], n_models=n_models, k=K_TOP, pose_id_offset=offset,
                                      out=outs[j], **kw)
        return hits

    def e2e_steps(n_steps, n_use):
        """n_steps x B scenes dealt round-robin to n_use contexts, one host thread each (n_use = 1: the calling thread)."""
        scenes = [(st, b) for st in range(n_steps) for b in range(B)]
        last = [None] * n_use
        if n_use == 1:
            for _, b in scenes:
                last[0] = score_scene(0, b)
            return last[0]
        errors = []

        def worker(j):
            try:
                for i in range(j, len(scenes), n_use):
                    last[j] = score_scene(j, scenes[i][1])
            except BaseException as e:  # noqa: BLE001
                errors.append(e)

        threads = [threading.Thread(target=worker, args=(j,)) for j in range(n_use)]
        for t in threads:
            t.start()
        for t in threads:
            t.join()
        if errors:
            raise errors[0]
        return last[(len(scenes) - 1) % n_use]

    def timed_e2e(n_use):
        e2e_steps(min(max(args.warmup, 1), 3), n_use)
        barrier()
        t0 = time.perf_counter()
        best = e2e_steps(args.steps, n_use)
        barrier()
        return time.perf_counter() - t0, best

    serial_s, best = timed_e2e(1)
    piped_s = None
    if n_ctx > 1:
        piped_s, best = timed_e2e(n_ctx)
    clocks = sampler.stop()
    e2e_counts_scene_last = outs[0]["counts"].copy()

    work = {"rendered_points_per_pose": float(counts0[0].mean()), "n_obs": int(wl["n_obs"])}

    if world > 1:
        t = torch.tensor([total_ms, serial_s, piped_s or 0.0], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        total_ms, serial_s = float(t[0]), float(t[1])
        piped_s = float(t[2]) if piped_s is not None else None

    line = None
    if rank == 0:
        scenes_per_step = B
        poses_per_step_all = (P_scene if strong else world * P) * B
        ms_per_step = total_ms / args.steps
        value = poses_per_step_all * args.steps / (total_ms / 1e3)
        kern_ms = ms_per_step / B  # per launch of P poses
        e2e_best_s = min(x for x in (serial_s, piped_s) if x is not None)
        h2d = int(P * (64 + 4 + 4 + (4 if h_lab is not None else 0))) * B
        d2h = int(P * 24 + n_models * K_TOP * 16 * world) * B
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong" if strong else "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": wl["desc"], "scenes_per_step": scenes_per_step,
                       "poses_per_gpu_per_scene": P, "poses_per_scene_all_gpus": P_scene if strong else world * P,
                       "poses_per_gpu_per_step": P * B, "triangles": wl["n_tris"], "meshes": wl["n_models"], "image": [W, H],
                       "stride": wl["stride"], "n_obs": int(wl["n_obs"]),
                       "composed_source": bool(composed),
                       "l2": "flushed once per step (256 MiB memset); the step's scenes use distinct pose batches",
                       "timing": "CUDA events on the library stream around the step's fused-kernel launches, max over ranks"},
            "clocks": clocks,
            "timed_region_s": total_ms / 1e3,
            "e2e": {"value": poses_per_step_all * args.steps / e2e_best_s, "unit": UNIT, "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": d2h, "ms_per_scene": e2e_best_s / (args.steps * B) * 1e3,
                    "contexts": n_ctx if (piped_s is not None and piped_s <= serial_s) else 1,
                    "serial": {"value": poses_per_step_all * args.steps / serial_s, "ms_per_scene": serial_s / (args.steps * B) * 1e3,
                               "contexts": 1},
                    "pipelined": None if piped_s is None else {"value": poses_per_step_all * args.steps / piped_s,
                                                               "ms_per_scene": piped_s / (args.steps * B) * 1e3, "contexts": n_ctx},
                    "topk": K_TOP,
                    "how": "every scene is one blocking perch_score_scene call: pinned host poses -> H2D -> fused kernel -> "
                           "per-model top-k on the device -> ncclAllGather of the per-rank records (N > 1) -> D2H of costs, "
                           "counts and merged hits; `serial` = one context, `pipelined` = scenes alternate between "
                           "`contexts` contexts, one host thread each; value = the faster of the two.  Pipelined scenes run on separate "
                           "streams, so kernels of two scenes can share the GPU: where one scene does not fill it (C1: 2 000 CTAs = 1.7 "
                           "waves) the pipelined rate can exceed `value`, which times the launches of ONE stream back to back"},
            "gpu_launches": args.steps * B,
            "roofline": rooflines(args, wl, P, kern_ms, clocks),
            "work": work,
            "best": {"cost": int(best["cost"][0, 0]), "pose_id": int(best["pose_id"][0, 0])},
            "wall_s_device_loop": wall_device_loop,
        }
        if not args.no_parity_check:
            line["parity_check"] = parity_check(wl, counts0, costs0, n=args.parity_poses)
        if world == 1 and not args.no_parity_check and wl["cost_type"] != 1:
            u = r.render_cost_unfused(wl["poses"], wl["pmm"], wl["plabel"], wl["tot"], **kw)
            bad = int((u["counts"] != counts0).any(0).sum() + (~_feq(u["rendered_cost"], costs0[0])).sum())
            line["unfused_check"] = {"poses": int(P), "mismatches": bad, "unfused_ms": r.last_timing()[0],
                                     "what": "scene 0 of the timed step, every pose: fused kernel output vs the library's "
                                             "unfused kernels (full raster, clouds, brute-force NN)"}
        single = wl["n_models"] == 1  # the GICP / glue legs are written for the one-mesh workloads
        if world == 1 and not args.no_gicp and single:
            o0 = {"rendered_cost": costs0[0], "observed_cost": costs0[1]}
            line["gicp"] = gicp_leg(r, wl, o0, kw)
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline(wl, budget_s=args.cpu_budget)
        if world == 1 and not args.no_glue and single:
            line["glue"] = glue_leg(r, wl, kw)  # last: re-targets the context at a 3-DoF scene
        if world == 1 and not args.no_ref_cuda:
            line["reference_cuda"] = reference_cuda(wl, batch=args.ref_batch)
        emit(line)
    if world > 1:
        for rr in ctxs:
            rr.comm_destroy()
        dist.destroy_process_group()
    return line


def _feq(a, b):
    return (a == b) | (np.isnan(a) & np.isnan(b))


def parity_check(wl, counts0, costs0, n=64):
    """What the timed launch of scene 0 wrote (counts + three costs, read back from its device buffers) against the CPU
    oracle (oracle_render + oracle_costs per pose, image_renderer.cuh:59-210, compute_point_clouds.cuh:37-184,
    compute_costs.cuh:161-290) on n poses spread evenly over the batch."""
    from oracle_lib import Oracle
    from perception_b200 import synth
    P = wl["P"]
    sel = np.unique(np.linspace(0, P - 1, min(n, P)).astype(np.int64))
    s = wl["scene"]
    t0 = time.perf_counter()
    o = Oracle().score_poses(s.tris, s.tris_model_count, wl["poses"][sel], wl["pmm"][sel], W, H, wl["proj"],
                             wl.get("oracle_source_depth", s.source_depth), s.source_label if wl["dof6"] else None,
                             wl["plabel"][sel] if wl["plabel"] is not None else None, synth.CX, synth.CY, synth.FX, synth.FY,
                             synth.DEPTH_FACTOR, wl["stride"], wl["obs"]["result_cloud_eigen"], wl["obs"]["result_cloud_color"],
                             wl["obs_label"], wl["tot"][sel], wl["cost_type"], True, wl["res"], 20.0)
    bad = (o["counts"] != counts0[:, sel]).any(0)
    for k, name in enumerate(("rendered_cost", "observed_cost", "points_diff_cost")):
        bad |= ~_feq(o[name], costs0[k, sel])
    return {"poses": int(len(sel)), "mismatches": int(bad.sum()), "oracle_s": time.perf_counter() - t0,
            "what": "scene 0 of the timed step: counts and the three costs of the timed launch, bit for bit, vs oracle/perch_oracle.c"}


def gicp_leg(r, wl, out, kw, n_top=4096, iterations=20, reps=3):
    """SURVEY 8(d) GICP measurement (BASELINE configs[3]): the n_top best poses of the scored batch by (rendered +
    observed) cost, refined by batched GICP with k = 10.
      fixed:      exactly `iterations` linearisations per pose (epsilons 0, stall detection off) -- the figure BASELINE
                  names ("20 iterations"); `roofline` = SURVEY 8(d)'s B_it = 72 B x N_src per pose and iteration over the
                  kernel time, against the measured HBM peak (the working set is L1/L2-resident: latency / issue bound);
      reference_settings: 150-iteration cap, epsilons 2e-3 / 5e-4 (renderer.cu:1696-1699), library defaults otherwise.
    Each against the label's observed cloud (6-DoF restriction) and against a ~300 k-point scene cloud (stride 1, table
    plane extended over the whole image).  Device time of the GICP kernel alone (CUDA events inside the library) and wall
    time of the whole perch_gicp call (lattice render + clouds + GICP + H2D/D2H)."""
    from perception_b200 import synth
    P = wl["P"]
    cost = np.where(out["rendered_cost"] >= 0, out["rendered_cost"] + np.nan_to_num(out["observed_cost"], nan=1e9), 1e9)
    sel = np.sort(np.argsort(cost, kind="stable")[:min(n_top, P)])
    poses, pmm = wl["poses"][sel], wl["pmm"][sel]
    lab = wl["plabel"][sel] if wl["plabel"] is not None else None
    base = dict(stride=kw["stride"], sensor_resolution=kw["sensor_resolution"], occlusion_threshold=kw["occlusion_threshold"],
                icp_k_correspondences=10)
    fixed = dict(base, icp_max_iterations=iterations, icp_rotation_epsilon=0.0, icp_transformation_epsilon=0.0)
    ref_set = dict(base, icp_max_iterations=150, icp_rotation_epsilon=2e-3, icp_transformation_epsilon=5e-4)
    n = len(sel)
    peak, _ = measured_peak_gbs()
    cl = r.render_cuda_multi_unified("CLOUD", poses, pmm, lab, stride=kw["stride"])
    n_src = float(cl["result_cloud_point_num"]) / max(n, 1)

    def timed(labels, args, stall_off):
        if stall_off:
            os.environ["PERCH_GICP_STALL_WINDOW"] = "0"
        try:
            g = r.gicp(poses, pmm, labels, **args)  # warm-up (target normals and grids are cached after the first call)
            dev, wall = [], []
            for _ in range(reps):
                t0 = time.perf_counter()
                g = r.gicp(poses, pmm, labels, **args)
                wall.append(time.perf_counter() - t0)
                dev.append(r.last_timing()[0])
        finally:
            os.environ.pop("PERCH_GICP_STALL_WINDOW", None)
        k_ms, w_s = float(np.median(dev)), float(np.median(wall))
        its = float(g["iterations"].mean())
        bytes_alg = 72.0 * n_src * its * n
        return {"refined_poses_per_s": n / w_s, "kernel_poses_per_s": n / (k_ms / 1e3), "kernel_ms": k_ms, "ms_total": w_s * 1e3,
                "mean_iterations_run": its, "max_iterations_run": int(g["iterations"].max()),
                "stopped_before_cap": float((g["iterations"] < args["icp_max_iterations"]).mean()),
                "roofline": {"bound": "hbm", "achieved": bytes_alg / (k_ms / 1e3) / 1e9, "peak": peak, "unit": "GB/s",
                             "frac": bytes_alg / (k_ms / 1e3) / 1e9 / peak, "bytes_per_pose_iteration": 72.0 * n_src,
                             "note": "SURVEY 8(d) B_it = 72 x N_src; the clouds and the search grid are L1/L2-resident, "
                                     "the kernel is latency / issue bound"}}

    res = {"poses": n, "k_correspondences": 10, "n_src_per_pose": n_src, "n_obs": int(wl["n_obs"]),
           "fixed": dict(timed(lab, fixed, True), iterations=iterations),
           "reference_settings": timed(lab, ref_set, False),
           "note": "top poses of the scored batch; `fixed` = exactly 20 linearisations per pose (BASELINE configs[3]), "
                   "`reference_settings` = 150-iteration cap with the reference's epsilons"}
    # BASELINE configs[3]: the same poses against a ~300k-point scene cloud (stride 1, no label restriction).  The
    # bench scene's table covers 76 % of the image; the table plane is extended over the rest (same noise / dropout model)
    s = wl["scene"]
    cam = (synth.CX, synth.CY, synth.FX, synth.FY, synth.DEPTH_FACTOR)
    T = np.asarray(s.T_cam_table, np.float64)
    nrm, org = T[:3, 2], T[:3, 3]
    uu, vv = np.meshgrid(np.arange(W), np.arange(H))
    den = nrm[0] * (uu - synth.CX) / synth.FX + nrm[1] * (vv - synth.CY) / synth.FY + nrm[2]
    with np.errstate(divide="ignore", invalid="ignore"):
        zc = np.rint(100.0 * float(nrm @ org) / den)
    rng = np.random.default_rng(77)
    fill = np.where((zc > 0) & (zc < 1000), zc + np.rint(rng.normal(0, 0.3, zc.shape)), 0).astype(np.int32)
    fill[rng.uniform(size=fill.shape) < 0.02] = 0
    full_depth = np.where(np.asarray(s.source_depth).reshape(H, W) > 0, np.asarray(s.source_depth).reshape(H, W), fill)
    dense = r.depth2cloud_global(full_depth.reshape(-1), s.source_color, W, H, 1, *cam, 1)
    t0 = time.perf_counter()
    r.set_observed(dense["result_cloud_eigen"], dense["result_cloud_color"], None)
    r.gicp(poses, pmm, None, **fixed)  # first call: target normals (10-NN of every scene point) + the search grids
    first_s = time.perf_counter() - t0
    res["scene_cloud"] = {"n_obs": int(dense["rendered_cloud_point_num"]), "first_call_ms": first_s * 1e3,
                          "fixed": dict(timed(None, fixed, True), iterations=iterations),
                          "reference_settings": timed(None, ref_set, False),
                          "note": "target = every valid pixel of a 640x480 scene whose table plane fills the image (C4); the "
                                  "first call also builds the target normals and grids"}
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
    # all host cores this process may run on, whatever OMP_NUM_THREADS says (torchrun sets it to 1)
    lib.cpu_baseline_set_threads(C.c_int(len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)))
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


def reference_cuda(wl, batch=700, reps=5):
    """The reference's own CUDA kernels (oracle/_ref/libperch_ref.so) on the same inputs, per-call upload + alloc like the
    reference (renderer.cu:1532-1544): at the shipped gpu_batch_size 700 (pr3_env_config.yaml:10) AND at
    min(P, 6 990) -- the largest batch below the reference's int-index overflow at 640x480.  Per batch size:
    one warm-up call, then `reps` calls; CUDA-event time of the four phases (median over reps) and wall time."""
    from oracle_lib import REF_SO, Reference
    if not REF_SO.exists():
        return {"unavailable": "oracle/_ref/libperch_ref.so not built"}
    ref = Reference()
    s = wl["scene"]
    from perception_b200 import synth

    def run(n):
        a = dict(tris=s.tris, tris_model_count=s.tris_model_count, poses=wl["poses"][:n], pose_model_map=wl["pmm"][:n], width=W,
                 height=H, proj=wl["proj"], source_depth=s.source_depth, source_color=s.source_color,
                 source_label=s.source_label if wl["dof6"] else None, pose_label=wl["plabel"][:n] if wl["dof6"] else None,
                 stride=wl["stride"], depth_factor=100, cx=synth.CX, cy=synth.CY, fx=synth.FX, fy=synth.FY,
                 obs_xyz=wl["obs"]["result_cloud_eigen"], obs_rgb=wl["obs"]["result_cloud_color"],
                 obs_label=wl["obs_label"] if wl["dof6"] else None, pose_obs_total=wl["tot"][:n], cost_type=wl["cost_type"],
                 calc_observed=True, sensor_resolution=wl["res"], color_distance_threshold=20.0, occlusion_threshold=1.0)
        ref.render_cost(**a)  # warm-up (context, allocator)
        wall, phases = [], []
        for _ in range(reps):
            t0 = time.perf_counter()
            q = ref.render_cost(**a)
            wall.append(time.perf_counter() - t0)
            phases.append(q["ms"])
        ph = np.median(np.array(phases), 0)
        ev_ms = float(np.median(np.array(phases).sum(1)))
        return {"batch": int(n), "value": n / (ev_ms / 1e3), "unit": UNIT, "event_ms_per_batch": ev_ms,
                "wall_value": n / float(np.median(wall)), "wall_ms_per_batch": float(np.median(wall)) * 1e3,
                "wall_ms_spread": [float(min(wall)) * 1e3, float(max(wall)) * 1e3], "reps": reps,
                "phase_ms": {"render": float(ph[0]), "cloud": float(ph[1]), "knn": float(ph[2]), "cost": float(ph[3])}}

    small = run(min(batch, wl["P"]))
    out = dict(small)
    big_n = min(wl["P"], 6990)
    if big_n > small["batch"]:
        try:
            out["max_batch"] = run(big_n)
        except Exception as e:  # noqa: BLE001
            out["max_batch"] = {"batch": int(big_n), "failed": repr(e)[:200]}
    out["note"] = ("reference .cuh kernels compiled unmodified for sm_100a + restated orchestration and brute-force NN; value = "
                   "poses / CUDA-event time of the call's four phases (uploads and allocations included), median of reps")
    return out


def run_reference_cpu(args):
    """--impl reference: the reference's CPU scoring path (port) on the host cores, bounded sample per step."""
    rank, local_rank, world = rank_info()
    if rank != 0:
        return
    from oracle_lib import Oracle
    r, wl = build_workload(args.workload, 0, local_rank, args.poses, use_product=False, compute_proj_fn=Oracle().compute_proj)
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
    from oracle_lib import Oracle
    r, wl = build_workload(args.workload, 0, local_rank, args.poses, use_product=False, compute_proj_fn=Oracle().compute_proj)
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
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--e2e-contexts", type=int, default=3,
                    help="contexts (one host thread each) the e2e leg spreads its steps over; 1 = strictly serial calls")
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference", "reference-cuda"])
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS))
    ap.add_argument("--poses", type=int, default=None, help="override poses per GPU per step")
    ap.add_argument("--cpu-budget", type=float, default=15.0)
    ap.add_argument("--ref-batch", type=int, default=700)
    ap.add_argument("--batches", type=int, default=40, help="scenes (distinct pose batches, one fused-kernel launch each) per step")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="strong: the scene's pose list is fixed (c3: 100k, c5: 1M, or --poses) and sharded over the GPUs")
    ap.add_argument("--topk", type=int, default=8)
    ap.add_argument("--parity-poses", type=int, default=64)
    ap.add_argument("--no-parity-check", action="store_true")
    ap.add_argument("--no-compose", action="store_true", help="c3: do not place an object into the source image")
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
