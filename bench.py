#!/usr/bin/env python
"""bench.py -- agent-steps/s of the batched SMARTS step including LiDAR + OGM (+ DAGM) observations.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload NAME]

One "step" = one tick of every environment of the workload: social-traffic injection, the Lane
controller + chassis step, and every configured sensor written in full.  Default workload is
BASELINE.json configs[2] (the one the metric -- "incl. LiDAR+OGM obs" -- is quoted on):
scenarios/intersections/4lane, 4 agents x 4096 envs, LiDAR 300 rays + OGM and DrivableAreaGridMap
256x256 @ 0.1 m/px, 20 synthetic social vehicles per env.  Under torchrun every rank runs the same
per-GPU workload on its own envs (weak scaling; envs are independent, so NCCL carries only the
episode-statistics all-reduce and the timing max).

Prints ONE JSON line (see the driver contract in the task statement).  `--impl reference` times
the CPU oracle (oracle/, the restated reference path) on all host threads instead.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from smarts_b200 import _abi, lidar_params, scenario_gen  # noqa: E402
from smarts_b200.road_map import RoadMapData  # noqa: E402

METRIC = "agent-steps/sec incl. LiDAR+OGM obs"
UNIT = "agent-steps/s"

WORKLOADS = {
    # BASELINE.json configs[2]
    "4lane_lidar_ogm_dagm_256": dict(
        map="4lane", E=4096, A=4, M=20, seed=11,
        missions=[("edge-south-SN", 1, 10.0, "edge-west-EW"), ("edge-west-WE", 0, 10.0, "edge-east-WE"),
                  ("edge-north-NS", 0, 10.0, "edge-south-NS"), ("edge-east-EW", 1, 10.0, "edge-west-EW")],
        cfg=dict(lidar=True, wall=1.5, ogm=(256, 256, 0.1), dagm=(256, 256, 0.1), route_cap=16),
        desc="scenarios/intersections/4lane, 4 agents x 4096 envs, LiDAR 300 rays + OGM + DAGM 256x256 @0.1 m/px, "
             "20 synthetic social vehicles/env, Lane action keep_lane"),
    # BASELINE.json configs[1]
    "figure_eight_nbr_wp": dict(
        map="figure_eight", E=256, A=4, M=20, seed=7,
        missions=[("gneE20", 0, 5.0, "gneE27"), ("gneE36", 1, 5.0, "gneE25"), ("gneE20", 2, 25.0, "gneE27"),
                  ("gneE36", 0, 30.0, "gneE25")],
        cfg=dict(nbr=True, nbr_radius=50.0, wp=True, wp_lookahead=32, route_cap=24),
        desc="scenarios/figure_eight, 4 agents x 256 envs, NeighborhoodVehicles + Waypoints"),
    # BASELINE.json configs[3] (Lane action here: the Trajectory policy needs the previous observation on the host)
    "minicity_wp_100_social": dict(
        map="minicity", E=1024, A=32, M=100, seed=13, missions="random",
        cfg=dict(wp=True, wp_lookahead=32, route_cap=32),
        desc="scenarios/minicity, 32 agents x 1024 envs, Waypoints obs, 100 synthetic BoxChassis social vehicles/env"),
    # BASELINE.json configs[4], one point of the sweep
    "straight_full_suite_64": dict(
        map="straight", E=4096, A=16, M=16, seed=17,
        missions=[("west", k % 3, 4.0 + 5.5 * k, "east") for k in range(16)],
        cfg=dict(nbr=True, nbr_radius=50.0, wp=True, wp_lookahead=32, lidar=True, wall=1.5, ogm=(64, 64, 50 / 64),
                 dagm=(64, 64, 50 / 64), route_cap=4),
        desc="synthetic straight road, 16 agents x 4096 envs + 16 social/env, nbr + wp + LiDAR + OGM + DAGM 64x64"),
}


def make_cfg(E, A, M, device, **kw):
    c = _abi.SbConfig()
    c.abi_version = _abi.SB_ABI_VERSION
    c.device = device
    c.n_envs, c.n_agents, c.n_social = E, A, M
    c.dt = 0.1
    c.action_space = _abi.SB_ACTION_LANE
    c.traj_len = 10
    c.route_cap = kw.get("route_cap", 16)
    c.nbr_enable, c.nbr_cap, c.nbr_radius = int(kw.get("nbr", False)), 10, kw.get("nbr_radius", 50.0)
    c.wp_enable, c.wp_lookahead, c.wp_out_paths, c.wp_out_len = int(kw.get("wp", False)), kw.get("wp_lookahead", 32), 4, 20
    c.lidar_enable, c.lidar_rays, c.lidar_miss_inf, c.lidar_wall_height = int(kw.get("lidar", False)), 300, 0, kw.get("wall", 0.0)
    if kw.get("ogm"):
        c.ogm_enable, c.ogm_width, c.ogm_height, c.ogm_resolution = (1,) + tuple(kw["ogm"])
    if kw.get("dagm"):
        c.dagm_enable, c.dagm_width, c.dagm_height, c.dagm_resolution = (1,) + tuple(kw["dagm"])
    if kw.get("events"):  # Events + done flags + trip-meter reward, DoneCriteria defaults (agent_interface.py:185-206)
        c.events_enable = 1
        c.done_mask = _abi.SB_EV_COLLISION | _abi.SB_EV_OFF_ROAD | _abi.SB_EV_OFF_ROUTE
        c.max_episode_steps = 0
        c.not_moving_time, c.not_moving_distance = 60.0, 1.0
    return c


def algorithmic_bytes_per_agent_step(cfg, M, A):
    """every input read once, every output written once at the StdObs dtypes."""
    b = 128 + 80 + 1 + 32.0 * M / A
    if cfg.nbr_enable:
        b += 494
    if cfg.wp_enable:
        b += 3280
    if cfg.lidar_enable:
        b += 25 * cfg.lidar_rays
    if cfg.ogm_enable:
        b += cfg.ogm_width * cfg.ogm_height
    if cfg.dagm_enable:
        b += cfg.dagm_width * cfg.dagm_height
    return b


def raster_bytes_per_agent(cfg, V):
    """k_raster: reads the env's vehicle boxes (7 doubles + alive each) and the ego pose once,
    writes both grids once."""
    b = V * 57 + 32
    if cfg.ogm_enable:
        b += cfg.ogm_width * cfg.ogm_height
    if cfg.dagm_enable:
        b += cfg.dagm_width * cfg.dagm_height
    return b


class ClockSampler(threading.Thread):
    """nvidia-smi-equivalent samples (NVML) of SM clock and throttle reasons during the timed region."""

    REASONS = {0x1: "gpu_idle", 0x2: "applications_clocks_setting", 0x4: "sw_power_cap", 0x8: "hw_slowdown",
               0x10: "sync_boost", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
               0x80: "hw_power_brake_slowdown", 0x100: "display_clock_setting"}

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz, self.ok = index, [], set(), None, False
        self._stop_evt = threading.Event()
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception:
            self.ok = False

    def run(self):
        while self.ok and not self._stop_evt.is_set():
            try:
                self.samples.append(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM))
                r = self.nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in self.REASONS.items():
                    if r & bit and name != "gpu_idle":
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.004)

    def stop(self):
        self._stop_evt.set()
        if self.is_alive():
            self.join(timeout=2)
        return {"sm_mhz": float(np.median(self.samples)) if self.samples else None,
                "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(self.samples)}


def build_inputs(w, E, seed_shift=0):
    rm = RoadMapData.load(os.path.join(ROOT, "maps", w["map"] + ".npz"))
    if w["missions"] == "random":
        w = dict(w, missions=scenario_gen.random_missions(rm, w["A"], w["seed"], w["cfg"].get("route_cap", 16)))
    x, y, h, v, route, rlen = scenario_gen.agent_starts(rm, E, w["missions"], w["cfg"].get("route_cap", 16),
                                                         w["seed"] + seed_shift, jitter=25.0, speed=8.0)
    traffic = scenario_gen.SocialTraffic(rm, E, w["M"], w["seed"] + 1 + seed_shift)
    return rm, (x, y, h, v, route, rlen), traffic


def cpu_run(w, E_s, ticks, threads):
    """Oracle (restated reference path) on `threads` host threads: E_s envs x `ticks` ticks."""
    from oracle import oracle_py
    rm, init, traffic = build_inputs(w, E_s)
    cfg = make_cfg(E_s, w["A"], w["M"], 0, **w["cfg"])
    ora = oracle_py.OracleWorld(cfg, rm.arrays(), rm.spacing, lidar_params.base_rays(lidar_params.BasicLidar), threads)
    ora.reset(*init)
    act = np.zeros((E_s, w["A"]), np.int32)
    socs = [traffic.states(t) for t in range(ticks + 1)]
    ora.set_social(socs[0]); ora.step(act); ora.observe()  # warm
    t0 = time.perf_counter()
    for t in range(ticks):
        ora.set_social(socs[t + 1])
        ora.step(act)
        ora.observe()
    dt = time.perf_counter() - t0
    ora.close()
    return E_s * w["A"] * ticks / dt, dt


def run_reference(args, w):
    """--impl reference: the reference's CPU path.  The reference itself (pybullet + Panda3D + SUMO)
    cannot be installed here (wheels absent, no network), so this is the CPU oracle
    port on all host threads, each step a bounded sample of the workload."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    # each step = the whole workload (every env of the per-GPU batch, one tick) unless that would not fit a
    # few minutes on this host, in which case the per-step sample is halved until it does
    E_s = w["E"]
    rate, _ = cpu_run(w, max(threads, 16), 2, threads)
    budget = 150.0 / max(args.steps + args.warmup, 1)
    while E_s * w["A"] / rate > budget and E_s > threads:
        E_s //= 2
    cpu_run(w, E_s, max(args.warmup, 1), threads)
    value, dt = cpu_run(w, E_s, args.steps, threads)
    sample = f"{E_s} envs x {w['A']} agents per step, {args.steps} steps, oracle port on {threads} threads"
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": w["desc"], "sample": sample},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="smarts_b200", choices=["smarts_b200", "reference"])
    ap.add_argument("--workload", default="4lane_lidar_ogm_dagm_256", choices=sorted(WORKLOADS))
    ap.add_argument("--envs", type=int, default=0, help="override envs per GPU")
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--e2e-full-rewrite", action="store_true", help="e2e with the default transport only (every grid byte rewritten per call)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--events", action="store_true", help="also compute Events / dones / trip-meter rewards (k_events)")
    ap.add_argument("--no-kernel-timing", action="store_true",
                    help="do not bracket every kernel with CUDA events inside the timed region (no roofline object)")
    args = ap.parse_args()
    w = dict(WORKLOADS[args.workload])
    if args.envs:
        w["E"] = args.envs
    if args.events:
        w["cfg"] = dict(w["cfg"], events=True)
        w["desc"] += ", events + rewards"
    if args.impl == "reference":
        return run_reference(args, w)

    import torch
    import torch.distributed as dist
    from smarts_b200.engine import Engine

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    E, A, M = w["E"], w["A"], w["M"]
    W_, K = max(args.warmup, 3), args.steps
    rm, init, traffic = build_inputs(w, E, seed_shift=1000 * rank)
    cfg = make_cfg(E, A, M, local, **w["cfg"])
    eng = Engine(cfg, rm, lidar_params.base_rays(lidar_params.BasicLidar))
    eng.reset(*init)
    T = 8
    rec_host = traffic.recording(T)                      # [T, E, M, 8] recorded social traffic
    rec = torch.from_numpy(rec_host).cuda()              # resident in HBM before timing
    actions = torch.zeros((E, A), dtype=torch.int32, device="cuda")  # keep_lane

    def tick(t):
        eng.set_social(rec[t % T])
        eng.step(actions)
        eng.observe()

    for t in range(W_):
        tick(t)
    torch.cuda.synchronize()
    sampler = ClockSampler(local)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    sampler.start()
    eng.set_timing(not args.no_kernel_timing)
    l0 = eng.launch_count
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for t in range(K):
        tick(W_ + t)
    ev1.record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    clocks = sampler.stop()
    ms = ev0.elapsed_time(ev1)
    launches = eng.launch_count - l0
    ktimes = eng.collect_timing()
    eng.set_timing(False)
    stats = torch.tensor([ms, float(E * A * K)], dtype=torch.float64, device="cuda")
    if world > 1:  # max time over ranks; episode statistics (agent-steps) summed -- the only collectives
        tmax = stats[:1].clone()
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        tot = stats[1:].clone()
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
        ms, total_agent_steps = float(tmax.item()), float(tot.item())
    else:
        total_agent_steps = float(E * A * K)
    value = total_agent_steps / (ms * 1e-3)

    # ---- end-to-end through the host-buffer call: H2D actions + social, D2H every observation
    Ke = max(1, args.e2e_steps)
    host_obs = eng.host_obs_buffers(pinned=True)
    act_host = torch.zeros((E, A), dtype=torch.int32).pin_memory()
    soc_host = torch.from_numpy(rec_host).pin_memory()
    h2d = act_host.numel() * 4 + soc_host[0].numel() * 8
    obs_host_bytes = sum(t.numel() * t.element_size() for t in host_obs.values())
    # bytes that cross PCIe: the 0/255 grids travel as 1 bit per pixel (sb_copy_obs_to_host) and are widened
    # to the uint8 StdObs layout by host threads inside the call; every other buffer is copied as is
    d2h = sum(t.numel() * t.element_size() // (8 if k in ("ogm", "dagm") else 1) for k, t in host_obs.items())
    def e2e_run(t_base):
        eng.step_host(act_host.numpy(), soc_host[t_base % T].numpy()); eng.copy_obs_to_host(host_obs)  # warm
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        t0 = time.perf_counter()
        for t in range(Ke):
            eng.step_host(act_host.numpy(), soc_host[(t_base + t + 1) % T].numpy())
            eng.copy_obs_to_host(host_obs)
        torch.cuda.synchronize()
        e2e_t = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(e2e_t, op=dist.ReduceOp.MAX)
        return world * E * A * Ke / float(e2e_t.item())
    # first with every grid byte rewritten per call (the default of sb_copy_obs_to_host), then the way a
    # vectorised env that keeps its pinned staging buffers would run it: sb_set_incremental_host_grids, only the
    # 64-pixel words that changed since the previous call are rewritten (the buffers hold the full images either way)
    e2e_full = e2e_run(0)
    e2e_value = e2e_full
    if not args.e2e_full_rewrite:
        eng.set_incremental_host_grids(True)
        e2e_value = e2e_run(Ke + 1)
        eng.set_incremental_host_grids(False)

    # ---- informational: the same step with Events / dones / trip-meter rewards also computed (k_events)
    ms_events = None
    if not args.events and world == 1:
        cfg_e = make_cfg(E, A, M, local, **dict(w["cfg"], events=True))
        eng_e = Engine(cfg_e, rm, lidar_params.base_rays(lidar_params.BasicLidar))
        eng_e.reset(*init)

        def tick_e(t):
            eng_e.set_social(rec[t % T])
            eng_e.step(actions)
            eng_e.observe()
        for t in range(W_):
            tick_e(t)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for t in range(K):  # the same ticks as the timed region above: the scene gets denser as agents converge
            tick_e(W_ + t)
        e1.record()
        torch.cuda.synchronize()
        ms_events = e0.elapsed_time(e1) / K
        eng_e.close()

    if rank == 0:
        peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
        if os.path.exists(peaks_path):
            peak, peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        else:
            peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
        r_ms, r_n = ktimes["k_raster"]
        dom = "k_raster" if r_n else max(ktimes, key=lambda k: ktimes[k][0])
        d_ms, d_n = ktimes[dom]
        if dom == "k_raster":
            per_launch = raster_bytes_per_agent(cfg, A + M) * E * A
        else:
            per_launch = algorithmic_bytes_per_agent_step(cfg, M, A) * E * A
        achieved = per_launch / (d_ms / d_n * 1e-3) / 1e9 if d_n and d_ms > 0 else float('nan')
        traffic = None  # dram__bytes_read + write per launch of that kernel, from the last committed ncu capture
        cur = os.path.join(ROOT, "profiles", "current.json")
        if os.path.exists(cur) and args.workload == "4lane_lidar_ogm_dagm_256" and not args.envs:
            traffic = json.load(open(cur)).get("kernels", {}).get(dom, {}).get("dram_bytes")
        step_bytes = algorithmic_bytes_per_agent_step(cfg, M, A)
        step_gbs = step_bytes * (E * A * K) / (ms * 1e-3) / 1e9  # this rank's share (weak scaling: per GPU)
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W_,
            "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": w["desc"], "envs_per_gpu": E, "agents_per_env": A, "social_per_env": M,
                       "l2": "outputs (%.2f GB/step/GPU) exceed the 126 MB L2, no flush needed" % (obs_host_bytes / 1e9),
                       "algorithmic_bytes_per_agent_step": step_bytes,
                       "step_hbm_gbs_per_gpu": step_gbs, "step_frac_of_peak": step_gbs / peak,
                       "kernel_ms_per_step": {k: v[0] / max(v[1], 1) for k, v in ktimes.items() if v[1]},
                       "ms_per_step_with_events_and_rewards": ms_events},
            "roofline": {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                         "bytes_per_launch": per_launch, "ms_per_launch": d_ms / max(d_n, 1)},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "host_obs_bytes_per_step": obs_host_bytes, "steps": Ke,
                    "value_full_rewrite": e2e_full,
                    "transport": "grids bit-packed over PCIe, widened to uint8 in host memory by the library; "
                                 + ("every byte rewritten per call" if args.e2e_full_rewrite else
                                    "persistent pinned destination buffers, only the 64-pixel words that changed since "
                                    "the previous call are rewritten (sb_set_incremental_host_grids); "
                                    "value_full_rewrite = the default mode")},
            "gpu_launches": int(launches), "clocks": clocks,
        }
        if world == 1 and not args.no_cpu_baseline:
            threads = os.cpu_count() or 1
            rate, _ = cpu_run(w, max(threads, 16), 2, threads)
            ticks = 20  # ~15 s of CPU work: as many envs of the workload as that allows, 20 ticks each
            E_s = int(min(max(15.0 * rate / (A * ticks), threads), E))
            v, dt = cpu_run(w, E_s, ticks, threads)
            line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": threads, "kind": "port",
                                    "sample": f"{E_s} envs x {A} agents x {ticks} ticks of the same workload "
                                              f"({dt:.1f} s), CPU oracle on {threads} threads"}
        else:
            line["cpu_baseline"] = None
        print(json.dumps(line), flush=True)
    eng.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
