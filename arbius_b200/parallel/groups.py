"""Process bootstrap: one process per GPU, torch.distributed for rendezvous / barriers / handle exchange
(NCCL on GPU boxes, gloo on CPU) — the product data paths do not go through it."""
from __future__ import annotations

import os
from typing import List, Optional

import torch
import torch.distributed as dist


def local_rank() -> int:
    return int(os.environ.get("LOCAL_RANK", "0"))


def init_distributed(backend: Optional[str] = None) -> int:
    """Initialise from torchrun's environment; returns the world size. Safe to call twice."""
    if dist.is_initialized():
        return dist.get_world_size()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world == 1:
        return 1
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29500")
    use_cuda = torch.cuda.is_available()
    if use_cuda:
        torch.cuda.set_device(local_rank())
    backend = backend or ("nccl" if use_cuda else "gloo")
    kw = {"device_id": torch.device("cuda", local_rank())} if backend == "nccl" else {}
    dist.init_process_group(backend, **kw)
    return dist.get_world_size()


_subset_cache = {}


def subset_group(size: int):
    """The aligned power-of-two subset of ranks this rank belongs to ({1,2,4,8} GPUs).
    Every rank must call this with the same size (new_group is collective)."""
    world = dist.get_world_size() if dist.is_initialized() else 1
    if size >= world or world == 1:
        return None
    key = size
    if key not in _subset_cache:
        groups = []
        for start in range(0, world, size):
            groups.append(dist.new_group(list(range(start, start + size))))
        _subset_cache[key] = groups
    return _subset_cache[key][dist.get_rank() // size]


def subset_ranks(rank: int, size: int) -> List[int]:
    start = rank // size * size
    return list(range(start, start + size))


_strided_cache = {}


def strided_group(stride: int):
    """The ranks congruent to this one modulo ``stride`` (e.g. stride 2: the ranks that run the same guidance branch when
    aligned pairs split the branches — the frame-parallel group of a guidance x frame layout). Collective, like
    ``subset_group``."""
    world = dist.get_world_size() if dist.is_initialized() else 1
    if world == 1 or stride <= 1:
        return None
    if stride not in _strided_cache:
        _strided_cache[stride] = [dist.new_group(list(range(r, world, stride))) for r in range(stride)]
    return _strided_cache[stride][dist.get_rank() % stride]
