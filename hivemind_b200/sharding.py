"""Multi-GPU host logic: one evaluator process per GPU over disjoint game batches.

The leaf-evaluation path has no exchange step: positions are independent, so N GPUs are N
evaluators, each fed its own contiguous shard of the games (the reference also creates one
`Engine(deviceId)` per CUDA device with no device-to-device traffic: engine/src/main.cc:154-171,
engine/src/search/agent.cc:312, 540-549).  The only collective is the weight refresh: rank 0 reads the packed
weight blob and broadcasts it (NCCL over NVLink/NVSwitch on the GPU box, gloo in CPU tests),
outside any timed region.
"""
from __future__ import annotations

from typing import Tuple

import numpy as np


def shard_range(n_items: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous [lo, hi) of rank's items; the first n_items % world ranks hold one more."""
    if not 0 <= rank < world:
        raise ValueError(f"rank {rank} outside world of {world}")
    base, extra = divmod(n_items, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def shard_positions(packed: np.ndarray, moves: np.ndarray, offsets: np.ndarray, rank: int, world: int):
    """Rank's slice of a packed-position batch with its move lists, offsets re-based to 0.

    offsets holds 2 entries per position (board A, board B) plus the end sentinel, exactly as
    hm_engine_enqueue_positions takes them (include/hm_engine.h)."""
    n = packed.shape[0]
    if offsets.shape[0] != 2 * n + 1:
        raise ValueError("offsets must have 2 * n_pos + 1 entries")
    lo, hi = shard_range(n, rank, world)
    o = offsets[2 * lo: 2 * hi + 1].astype(np.int64)
    return packed[lo:hi], moves[o[0]: o[-1]], (o - o[0]).astype(offsets.dtype)


def broadcast_blob(blob: bytes | None, src: int = 0, device=None) -> bytes:
    """Weight refresh: `src`'s packed weight blob to every rank.  Needs an initialised
    torch.distributed process group; `device` is the CUDA device for NCCL groups (None = CPU
    tensors, i.e. gloo)."""
    import torch
    import torch.distributed as dist

    rank = dist.get_rank()
    n = torch.tensor([len(blob) if rank == src else 0], dtype=torch.int64, device=device)
    dist.broadcast(n, src)
    if rank == src:
        t = torch.frombuffer(bytearray(blob), dtype=torch.uint8)
        t = t.to(device) if device is not None else t
    else:
        t = torch.empty(int(n.item()), dtype=torch.uint8, device=device)
    dist.broadcast(t, src)
    return t.cpu().numpy().tobytes()


def gather_ragged(local: np.ndarray, device=None) -> np.ndarray:
    """Concatenate every rank's 1-D result array in rank order (control plane, e.g. collecting
    priors of sharded games on rank 0 for a check; never inside a timed region)."""
    import torch
    import torch.distributed as dist

    world = dist.get_world_size()
    n = torch.tensor([local.shape[0]], dtype=torch.int64, device=device)
    sizes = [torch.zeros_like(n) for _ in range(world)]
    dist.all_gather(sizes, n)
    sizes = [int(s.item()) for s in sizes]
    cap = max(sizes) if sizes else 0
    buf = torch.zeros(cap, dtype=torch.from_numpy(local[:0].copy()).dtype, device=device)
    buf[: local.shape[0]] = torch.from_numpy(np.ascontiguousarray(local)).to(buf.device)
    parts = [torch.empty_like(buf) for _ in range(world)]
    dist.all_gather(parts, buf)
    return np.concatenate([p[:s].cpu().numpy() for p, s in zip(parts, sizes)])
