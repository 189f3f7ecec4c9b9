"""Multi-GPU host logic on CPU: env sharding + the statistics all-reduce, gloo, world_size 2.

The data path has no collective (envs are independent); what N>1 adds is the
env-range split, per-shard scenario generation and one small reduce — all covered here without
a GPU.  The reference's analogue is ParallelEnv's process-per-env fan-out
(smarts/env/wrappers/parallel_env.py:96-164)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import load_map
from smarts_b200 import scenario_gen, sharding


def test_shard_envs_partitions_exactly():
    for n in (0, 1, 7, 8, 4096, 65537):
        for world in (1, 2, 3, 8):
            shards = [sharding.shard_envs(n, r, world) for r in range(world)]
            assert shards[0].start == 0 and shards[-1].stop == n
            for a, b in zip(shards, shards[1:]):
                assert a.stop == b.start
            sizes = [s.count for s in shards]
            assert max(sizes) - min(sizes) <= 1 and sum(sizes) == n
            for g in {0, n // 2, n - 1} - {-1}:
                if n:
                    r = sharding.owner_of(g, n, world)
                    assert shards[r].start <= g < shards[r].stop
                    assert shards[r].local(g) == g - shards[r].start
    with pytest.raises(ValueError):
        sharding.shard_envs(4, 2, 2)
    with pytest.raises(IndexError):
        sharding.shard_envs(8, 0, 2).local(5)


def test_weak_shard_layout():
    s = sharding.weak_shard(4096, 3, 8)
    assert (s.start, s.count, s.stop) == (3 * 4096, 4096, 4 * 4096)
    assert s.total == 8 * 4096 and sharding.shard_envs(9, 1, 2).total == 9


def test_shard_from_env_reads_the_torchrun_variables(monkeypatch):
    monkeypatch.setenv("RANK", "1"); monkeypatch.setenv("WORLD_SIZE", "4")
    s = sharding.shard_from_env(10)
    assert (s.rank, s.world, s.start, s.count, s.total) == (1, 4, 3, 3, 10)
    monkeypatch.delenv("RANK"); monkeypatch.delenv("WORLD_SIZE")
    assert sharding.shard_from_env(10).count == 10


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, n_envs, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        sh = sharding.shard_envs(n_envs, rank, world)
        rm = load_map("4lane")
        # per-shard social traffic: the global recording restricted to this rank's env range
        full = scenario_gen.SocialTraffic(rm, n_envs, 5, seed=3).states(2)
        mine = full[sh.start:sh.stop]
        agent_steps = float(mine.shape[0] * 4 * 10)
        red = sharding.reduce_stats({"agent_steps": agent_steps, "episodes_done": float(rank + 1),
                                     "elapsed_ms": 10.0 + rank}, max_keys=("elapsed_ms",))
        rows = sharding.gather_rows(np.array([rank, sh.start, sh.count], np.int64))
        np.savez(os.path.join(out_dir, f"r{rank}.npz"), mine=mine, agent_steps=red["agent_steps"],
                 episodes=red["episodes_done"], elapsed=red["elapsed_ms"], rows=rows)
        dist.barrier()
    finally:
        dist.destroy_process_group()


def test_two_rank_gloo_shards_and_reduces(tmp_path):
    world, n_envs = 2, 9
    mp.spawn(_worker, args=(world, _free_port(), n_envs, str(tmp_path)), nprocs=world, join=True)
    rm = load_map("4lane")
    full = scenario_gen.SocialTraffic(rm, n_envs, 5, seed=3).states(2)
    parts = [np.load(tmp_path / f"r{r}.npz") for r in range(world)]
    assert np.array_equal(np.concatenate([p["mine"] for p in parts]), full)     # every env exactly once, in order
    for p in parts:
        assert float(p["agent_steps"]) == n_envs * 4 * 10                        # SUM over ranks
        assert float(p["episodes"]) == 3.0
        assert float(p["elapsed"]) == 11.0                                       # MAX over ranks
        assert p["rows"].tolist() == [[0, 0, 5], [1, 5, 4]]


def test_reduce_stats_single_process_passthrough():
    assert sharding.reduce_stats({"agent_steps": 5.0, "elapsed_ms": 2.0}) == {"agent_steps": 5.0, "elapsed_ms": 2.0}
    assert sharding.gather_rows(np.arange(3)).tolist() == [[0, 1, 2]]
    assert torch.distributed.is_available()
