"""Env sharding across the GPUs of one box.

Environments are independent (no cross-env reads), so the only multi-GPU logic is (1) which
contiguous env range a rank owns and (2) one small all-reduce of episode statistics / timings.
There is NO collective on the data path.  The reference's analogue is `ParallelEnv`, one OS
process per env over pipes (smarts/env/wrappers/parallel_env.py:96-164); here it is one process
per GPU, `torch.distributed` (nccl on the GPU box, gloo in the CPU tests).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, Sequence, Tuple

import numpy as np


@dataclass(frozen=True)
class EnvShard:
    rank: int
    world: int
    start: int   # first global env index owned by this rank
    count: int   # envs owned
    total: int = 0   # envs of the whole job (0: not recorded)

    @property
    def stop(self) -> int:
        return self.start + self.count

    def local(self, global_env: int) -> int:
        if not (self.start <= global_env < self.stop):
            raise IndexError(f"env {global_env} is not on rank {self.rank} [{self.start}, {self.stop})")
        return global_env - self.start


def shard_envs(n_envs: int, rank: int, world: int) -> EnvShard:
    """Strong-scaling split of `n_envs` global envs: contiguous ranges, sizes differ by <= 1,
    the first `n_envs % world` ranks take the extra env (ParallelEnv keeps env order too)."""
    if world < 1 or not (0 <= rank < world) or n_envs < 0:
        raise ValueError("bad rank/world/n_envs")
    base, extra = divmod(n_envs, world)
    count = base + (1 if rank < extra else 0)
    start = rank * base + min(rank, extra)
    return EnvShard(rank, world, start, count, n_envs)


def weak_shard(envs_per_gpu: int, rank: int, world: int) -> EnvShard:
    """Weak-scaling layout used by bench.py: every rank owns `envs_per_gpu` envs of its own."""
    return EnvShard(rank, world, rank * envs_per_gpu, envs_per_gpu, world * envs_per_gpu)


def shard_from_env(n_envs: int) -> EnvShard:
    """The shard of this process under torchrun (RANK / WORLD_SIZE; a single process owns everything)."""
    import os
    return shard_envs(n_envs, int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")))


def owner_of(global_env: int, n_envs: int, world: int) -> int:
    base, extra = divmod(n_envs, world)
    edge = extra * (base + 1)
    if global_env < edge:
        return global_env // (base + 1)
    return extra + (global_env - edge) // max(base, 1)


STAT_KEYS: Tuple[str, ...] = ("agent_steps", "episodes_done", "reward_sum", "collisions", "off_road")


def reduce_stats(stats: Dict[str, float], max_keys: Sequence[str] = ("elapsed_ms",), device=None) -> Dict[str, float]:
    """All-reduce a small statistics record: SUM for counters, MAX for the keys in `max_keys`
    (device timings are the max over ranks).  Works without an initialised process group
    (single process) and with any backend."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return dict(stats)
    keys = sorted(stats)
    sums = torch.tensor([0.0 if k in max_keys else float(stats[k]) for k in keys], dtype=torch.float64, device=device)
    maxs = torch.tensor([float(stats[k]) if k in max_keys else 0.0 for k in keys], dtype=torch.float64, device=device)
    dist.all_reduce(sums, op=dist.ReduceOp.SUM)
    dist.all_reduce(maxs, op=dist.ReduceOp.MAX)
    return {k: float(maxs[i] if k in max_keys else sums[i]) for i, k in enumerate(keys)}


def gather_rows(x: np.ndarray, device=None) -> np.ndarray:
    """All-gather equally-shaped per-rank rows (e.g. per-GPU throughput for the scaling table)."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return np.asarray(x)[None]
    t = torch.as_tensor(np.ascontiguousarray(x), device=device)
    out = [torch.empty_like(t) for _ in range(dist.get_world_size())]
    dist.all_gather(out, t)
    return np.stack([o.cpu().numpy() for o in out])
