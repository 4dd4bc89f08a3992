"""Offline estimate of the cull rates of the hierarchical candidate test (pose_kernels.cu pass A) for different
triangle clusterings.  numpy only; mirrors the device test (bounding sphere -> screen square -> lattice pixel?)."""
import sys, numpy as np
sys.path.insert(0, '.')
from perception_b200 import synth

W, H, S = 640, 480, 8
DIST = 0.8
K = 4 if len(sys.argv) < 2 else int(sys.argv[1])
mesh = synth.make_mesh(20000, seed=1000, r0=0.09)
V = mesh['v'].astype(np.float64)  # T,3,3
T = len(V)
cen = V.mean(1)

def morton_order(c):
    q = ((c - c.min(0)) / (c.max(0) - c.min(0)) * 1023).astype(np.uint64)
    def spread(v):
        v = (v | (v << 16)) & 0x030000FF; v = (v | (v << 8)) & 0x0300F00F
        v = (v | (v << 4)) & 0x030C30C3; v = (v | (v << 2)) & 0x09249249
        return v
    code = spread(q[:, 0]) | (spread(q[:, 1]) << 1) | (spread(q[:, 2]) << 2)
    return np.argsort(code, kind='stable')

def chunks(order, k):
    n = (len(order) + k - 1) // k
    idx = np.minimum(np.arange(n * k), len(order) - 1)
    return order[idx].reshape(n, k)

def greedy(k):
    """seed in Morton order; repeatedly add the unassigned triangle (from nearby cells) that keeps the bounding
    sphere smallest"""
    order = morton_order(cen)
    edge = np.linalg.norm(V[:, 0] - V[:, 1], axis=1).mean()
    cell = 2.5 * edge
    ijk = np.floor((cen - cen.min(0)) / cell).astype(np.int64)
    from collections import defaultdict
    grid = defaultdict(list)
    for t in range(T): grid[tuple(ijk[t])].append(t)
    used = np.zeros(T, bool)
    out = []
    for seed in order:
        if used[seed]: continue
        used[seed] = True
        members = [seed]
        pts = V[seed]
        while len(members) < k:
            c0 = 0.5 * (pts.min(0) + pts.max(0))
            ci = tuple(np.floor((c0 - cen.min(0)) / cell).astype(np.int64))
            best, bt = None, -1
            for dx in (-1, 0, 1):
                for dy in (-1, 0, 1):
                    for dz in (-1, 0, 1):
                        for t in grid.get((ci[0] + dx, ci[1] + dy, ci[2] + dz), ()):
                            if used[t]: continue
                            p2 = np.concatenate([pts, V[t]])
                            c = 0.5 * (p2.min(0) + p2.max(0))
                            r = np.linalg.norm(p2 - c, axis=1).max()
                            if best is None or r < best: best, bt = r, t
            if bt < 0: break
            used[bt] = True; members.append(bt); pts = np.concatenate([pts, V[bt]])
        while len(members) < k: members.append(members[-1])
        out.append(members)
    return np.array(out)

def stats(cl, name):
    pts = V[cl].reshape(len(cl), -1, 3)
    c = 0.5 * (pts.min(1) + pts.max(1))
    r = np.linalg.norm(pts - c[:, None], axis=2).max(1)
    rng = np.random.default_rng(0)
    views = synth.fibonacci_views(512)
    poses = []
    for i in range(0, 512, 37):
        M = np.eye(4); M[:3, :3] = synth.rotation_from_view(views[i], 0.3 * i); M[:3, 3] = (0.05, 0.02, DIST)
        poses.append(M)
    fx, fy, cx, cy = synth.FX, synth.FY, synth.CX, synth.CY
    surv_c, surv_t, hits, ideal_s, ideal_b = [], [], [], [], []
    for P in poses:
        M = P
        R, t = M[:3, :3], M[:3, 3]
        cc = c @ R.T + t
        sx, sy, z = fx * cc[:, 0] / cc[:, 2] + cx, fy * cc[:, 1] / cc[:, 2] + cy, cc[:, 2]
        Rx = r * (fx + abs(sx - cx)) / (z - r) + 0.01
        Ry = r * (fy + abs(sy - cy)) / (z - r) + 0.01
        def lat(lo, hi):  # a multiple of S in (lo-0.52, hi+0.02]
            return np.ceil((lo - 0.52) / S) * S <= hi + 0.02
        keep = lat(sx - Rx, sx + Rx) & lat(sy - Ry, sy + Ry)
        surv_c.append(keep.mean())
        Rx0 = r * fx / (z - r); Ry0 = r * fy / (z - r)
        ideal_s.append((lat(sx - Rx0, sx + Rx0) & lat(sy - Ry0, sy + Ry0)).mean())
        vv = V @ R.T + t
        px, py = fx * vv[..., 0] / vv[..., 2] + cx, fy * vv[..., 1] / vv[..., 2] + cy
        th = lat(px.min(1), px.max(1)) & lat(py.min(1), py.max(1))
        cpx, cpy = px[cl].reshape(len(cl), -1), py[cl].reshape(len(cl), -1)
        ideal_b.append((lat(cpx.min(1), cpx.max(1)) & lat(cpy.min(1), cpy.max(1))).mean())
        hits.append(th.mean())
        surv_t.append(th[np.unique(cl[keep])].mean())
    print(f"{name:10s} k={cl.shape[1]} clusters {len(cl)} mean r {r.mean()*1e3:.2f} mm | cluster survive {np.mean(surv_c):.3f} "
          f"| tri candidates overall {np.mean(hits):.3f}, among surviving clusters {np.mean(surv_t):.3f} "
          f"| sphere w/o tan {np.mean(ideal_s):.3f} | exact cluster bbox {np.mean(ideal_b):.3f} "
          f"| est instr/tri {60/cl.shape[1] + np.mean(surv_c)*100:.1f}")

print("focal", synth.FX, "mean edge mm", np.linalg.norm(V[:, 0] - V[:, 1], axis=1).mean() * 1e3)
stats(chunks(np.arange(T), K), "original")
stats(chunks(morton_order(cen), K), "morton")
stats(greedy(K), "greedy")
