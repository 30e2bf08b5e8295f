"""Multi-GPU host side: frames are independent, so ranks own whole frames and the fusion stage itself needs no
collective.  The only exchange is the gather of the (B_local, K, F) detections for reporting, which replaces the
reference's all_gather of the entire raw head-output dict and input batch
(reference src/lib/model/progressBar.py:85-91, :177-183; batch split src/main.py:98-122).

Works with any torch.distributed backend: "nccl" (one process per GPU, NVLink / NVSwitch) in production,
"gloo" in the CPU tests.
"""
from __future__ import annotations

import torch
import torch.distributed as dist


def frame_shard(n_frames: int, rank: int, world: int) -> tuple[int, int]:
    """Contiguous shard [lo, hi) of rank `rank`: the first n_frames % world ranks own one frame more."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError("bad rank/world %d/%d" % (rank, world))
    base, extra = divmod(int(n_frames), world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def shard_batch(tensors: dict, rank: int, world: int) -> dict:
    """Slices every tensor whose leading dimension is the batch (taken from 'heatmap' or the first tensor)."""
    ref = tensors.get("heatmap", next(iter(tensors.values())))
    B = ref.shape[0]
    lo, hi = frame_shard(B, rank, world)
    return {k: (v[lo:hi] if hasattr(v, "shape") and v.ndim >= 1 and v.shape[0] == B else v) for k, v in tensors.items()}


def gather_detections(ret: dict, group=None) -> dict:
    """all_gather of per-rank detection tensors along the batch dimension (uneven shards allowed).
    ret: {name: (B_local, ...)} as returned by fusionDecode; returns {name: (B_total, ...)} on every rank."""
    if not (dist.is_available() and dist.is_initialized()):
        return dict(ret)
    world = dist.get_world_size(group)
    if world == 1:
        return dict(ret)
    any_t = next(iter(ret.values()))
    n_local = torch.tensor([any_t.shape[0]], dtype=torch.int64, device=any_t.device)
    counts = [torch.zeros_like(n_local) for _ in range(world)]
    dist.all_gather(counts, n_local, group=group)
    counts = [int(c.item()) for c in counts]
    n_max = max(counts)
    out = {}
    for name, t in ret.items():
        t = t.contiguous()
        if t.shape[0] < n_max:  # pad to the largest shard: all_gather_into_tensor wants equal shapes
            pad = torch.zeros((n_max - t.shape[0],) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
            t = torch.cat([t, pad], 0)
        buf = torch.empty((world * n_max,) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
        dist.all_gather_into_tensor(buf, t, group=group)
        parts = [buf[r * n_max:r * n_max + counts[r]] for r in range(world)]
        out[name] = torch.cat(parts, 0)
    return out
