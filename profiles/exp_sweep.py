"""BASELINE.json configs[4]: synthetic straight-road sweep, 16 agents + 16 social vehicles per env, the full
sensor suite (nbr r=50, wp 32, LiDAR 300, OGM + DAGM), envs per GPU swept up to 64 k (~1 M agents).
Device-resident timing only (CUDA events over K ticks); prints one JSON line per point.

    python profiles/exp_sweep.py [--grid 64] [--envs 1024,4096,16384,65536] [--steps 10]
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from smarts_b200 import lidar_params, scenario_gen  # noqa: E402
from smarts_b200.road_map import RoadMapData  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--grid", type=int, default=64)
    ap.add_argument("--envs", default="1024,4096,16384,65536")
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--events", action="store_true")
    args = ap.parse_args()
    import torch
    from smarts_b200.engine import Engine
    rm = RoadMapData.load(os.path.join(ROOT, "maps", "straight.npz"))
    A, M, E0 = 16, 16, 1024
    missions = [("west", k % 3, 4.0 + 5.5 * k, "east") for k in range(A)]
    base = scenario_gen.agent_starts(rm, E0, missions, 4, 17, jitter=2.0, speed=8.0)
    soc0 = scenario_gen.SocialTraffic(rm, E0, M, 18).recording(4)
    res = 50.0 / args.grid if args.grid == 64 else 0.1
    for E in [int(x) for x in args.envs.split(",")]:
        reps = (E + E0 - 1) // E0
        init = tuple(np.concatenate([b] * reps)[:E] for b in base)
        cfg = bench.make_cfg(E, A, M, 0, nbr=True, nbr_radius=50.0, wp=True, wp_lookahead=32, lidar=True, wall=1.5,
                             ogm=(args.grid, args.grid, res), dagm=(args.grid, args.grid, res), route_cap=4,
                             events=args.events)
        eng = Engine(cfg, rm, lidar_params.base_rays(lidar_params.BasicLidar))
        eng.reset(*init)
        rec = torch.from_numpy(np.concatenate([soc0] * reps, axis=1)[:, :E]).cuda()
        act = torch.zeros((E, A), dtype=torch.int32, device="cuda")

        def tick(t):
            eng.set_social(rec[t % 4]); eng.step(act); eng.observe()
        for t in range(3):
            tick(t)
        torch.cuda.synchronize()
        eng.set_timing(True)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for t in range(args.steps):
            tick(3 + t)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / args.steps
        kt = {k: round(v[0] / max(v[1], 1), 4) for k, v in eng.collect_timing().items() if v[1]}
        st = eng.state_numpy()["status"]
        b = bench.algorithmic_bytes_per_agent_step(cfg, M, A)
        print(json.dumps({"envs": E, "agents": E * A, "grid": args.grid, "ms_per_step": ms,
                          "agent_steps_per_s": E * A / (ms * 1e-3), "hbm_gbs": b * E * A / (ms * 1e-3) / 1e9,
                          "mem_gb": torch.cuda.memory_allocated() / 1e9 + sum(0 for _ in ()),
                          "status_bits": int(np.bitwise_or.reduce(st.ravel())), "kernel_ms": kt}), flush=True)
        eng.close()
        del rec, act
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
