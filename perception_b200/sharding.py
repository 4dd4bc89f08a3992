"""Multi-GPU host logic: candidate poses are independent units, so the pose list is cut into contiguous blocks, one
per rank (one process per GPU), each rank scores its block against replicated models / scene / observed cloud, and the
only exchange is the per-rank top-k hit list (k x 16 bytes per model) followed by a lexicographic (cost, pose id)
merge -- the multi-process form of the caller's per-model argmin (search_env.cpp:2542-2583; the reference itself
walks its gpu_batch_size batches serially on one GPU, search_env.cpp:2504-2525, and has no multi-GPU code).

`torch.distributed` is plumbing only (NCCL on GPU boxes, gloo in the CPU tests)."""
from __future__ import annotations

import numpy as np

HIT_DTYPE = np.dtype([("cost", "<i4"), ("pose_id", "<i4"), ("rendered_cost", "<f4"), ("observed_cost", "<f4")])
EMPTY_COST = np.iinfo(np.int32).max


def shard_range(n_items: int, rank: int, world: int) -> tuple[int, int]:
    """Contiguous block [lo, hi) of rank `rank`: sizes differ by at most one and blocks are in rank order, so global
    pose ids stay monotone in (rank, local index) and first-wins ties survive the merge."""
    if world <= 0 or not (0 <= rank < world) or n_items < 0:
        raise ValueError(f"bad shard request: n={n_items} rank={rank} world={world}")
    base, extra = divmod(n_items, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def merge_hits(all_hits: np.ndarray, k: int) -> np.ndarray:
    """(world, n_models, k_local) hit records -> (n_models, k) best by (cost, pose_id); empty slots (cost INT32_MAX) last."""
    all_hits = np.asarray(all_hits, dtype=HIT_DTYPE)
    if all_hits.ndim != 3:
        raise ValueError("expected (world, n_models, k_local) hit records")
    n_models = all_hits.shape[1]
    out = np.zeros((n_models, k), dtype=HIT_DTYPE)
    out["cost"] = EMPTY_COST
    out["pose_id"] = -1
    for m in range(n_models):
        h = all_hits[:, m, :].reshape(-1)
        h = h[h["cost"] != EMPTY_COST]
        order = np.lexsort((h["pose_id"], h["cost"]))[:k]
        out[m, :len(order)] = h[order]
    return out


def gather_hits(hits: np.ndarray, group=None) -> np.ndarray:
    """all_gather of this rank's (n_models, k) hit records -> (world, n_models, k).  Every rank gets the full list.
    Without an initialised process group this is the single-rank identity."""
    import torch
    import torch.distributed as dist

    hits = np.ascontiguousarray(hits, dtype=HIT_DTYPE)
    if not (dist.is_available() and dist.is_initialized()):
        return hits[None]
    world = dist.get_world_size(group)
    flat = torch.from_numpy(hits.view(np.int32).reshape(-1).copy())
    if dist.get_backend(group) == "nccl":
        flat = flat.cuda()
    parts = [torch.empty_like(flat) for _ in range(world)]
    dist.all_gather(parts, flat, group=group)
    merged = torch.stack(parts).cpu().numpy()
    return merged.view(HIT_DTYPE).reshape((world,) + hits.shape)


class HitExchange:
    """gather_hits with every buffer allocated once: pinned host staging, one device send buffer, one device receive
    buffer for all ranks, `all_gather_into_tensor` (a single NCCL call), one synchronising copy back.  Per step this is
    two small copies and one collective instead of world + 3 allocations and a list-based gather."""

    def __init__(self, n_models: int, k: int, group=None):
        import torch
        import torch.distributed as dist

        self.shape = (n_models, k)
        self.group = group
        self.active = dist.is_available() and dist.is_initialized()
        self.world = dist.get_world_size(group) if self.active else 1
        n = n_models * k * 4  # int32 words per rank
        self.nccl = self.active and dist.get_backend(group) == "nccl"
        self.h_send = torch.empty(n, dtype=torch.int32)
        self.h_recv = torch.empty(self.world * n, dtype=torch.int32)
        if self.nccl:
            self.h_send, self.h_recv = self.h_send.pin_memory(), self.h_recv.pin_memory()
            self.d_send = torch.empty(n, dtype=torch.int32, device="cuda")
            self.d_recv = torch.empty(self.world * n, dtype=torch.int32, device="cuda")

    def gather(self, hits: np.ndarray) -> np.ndarray:
        import torch
        import torch.distributed as dist

        hits = np.ascontiguousarray(hits, dtype=HIT_DTYPE)
        if hits.shape != self.shape:
            raise ValueError(f"expected hit records of shape {self.shape}, got {hits.shape}")
        if not self.active:
            return hits[None]
        self.h_send.numpy()[:] = hits.view(np.int32).reshape(-1)
        if self.nccl:
            self.d_send.copy_(self.h_send, non_blocking=True)
            dist.all_gather_into_tensor(self.d_recv, self.d_send, group=self.group)
            self.h_recv.copy_(self.d_recv, non_blocking=True)
            torch.cuda.current_stream().synchronize()
        else:
            dist.all_gather_into_tensor(self.h_recv, self.h_send, group=self.group)
        return self.h_recv.numpy().copy().view(HIT_DTYPE).reshape((self.world,) + self.shape)


def score_sharded(renderer, poses, pose_model_map, pose_label, pose_observed_points_total, n_models: int, k: int = 1,
                  group=None, **params):
    """Scores this rank's block of the GLOBAL pose list on `renderer` (a perception_b200.Renderer with models, scene and
    observed cloud already set) and returns (local result dict, global (n_models, k) hits, (lo, hi))."""
    import torch.distributed as dist

    from .renderer import topk

    rank, world = (dist.get_rank(group), dist.get_world_size(group)) if (dist.is_available() and dist.is_initialized()) else (0, 1)
    lo, hi = shard_range(len(poses), rank, world)
    sl = slice(lo, hi)
    out = renderer.render_cuda_multi_unified(
        "COST", poses[sl], pose_model_map[sl], None if pose_label is None else pose_label[sl],
        None if pose_observed_points_total is None else pose_observed_points_total[sl], **params)
    hits = topk(out["rendered_cost"], out["observed_cost"], pose_model_map[sl], n_models, k=k, pose_id_offset=lo)
    return out, merge_hits(gather_hits(hits, group), k), (lo, hi)
