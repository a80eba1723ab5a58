"""Sample sharding over the GPUs of one box: images are independent through the tower and each
image's tokens are spliced only into its own row, so rows (with all their images) are partitioned across ranks and
NO collective sits on the data path.  torch.distributed is used only to agree on timings (max over ranks)."""
from __future__ import annotations

import os
from typing import List, Tuple


def env_rank_world() -> Tuple[int, int, int]:
    return (int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)))


def shard_rows(n_rows: int, rank: int, world: int) -> range:
    """Contiguous block partition, remainder spread over the first ranks (every row owned exactly once)."""
    base, rem = divmod(n_rows, world)
    start = rank * base + min(rank, rem)
    return range(start, start + base + (1 if rank < rem else 0))


def shard_batch(input_ids, images_per_row: List[int], rank: int, world: int):
    """-> (row range, image range) owned by `rank`: a row travels with all of its images."""
    rows = shard_rows(len(images_per_row), rank, world)
    first_img = sum(images_per_row[: rows.start])
    n_img = sum(images_per_row[rows.start: rows.stop])
    return rows, range(first_img, first_img + n_img)


def max_over_ranks(value: float, device=None) -> float:
    """Max of a host scalar over all ranks (timing plumbing only; gloo on CPU, nccl on GPU)."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return float(value)
    t = torch.tensor([value], dtype=torch.float64, device=device if dist.get_backend() == "nccl" else "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def sum_over_ranks(value: float, device=None) -> float:
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return float(value)
    t = torch.tensor([value], dtype=torch.float64, device=device if dist.get_backend() == "nccl" else "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return float(t.item())
