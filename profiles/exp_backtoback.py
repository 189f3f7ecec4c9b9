"""Experiment: does a kernel run slower back-to-back than when ncu serialises it?
Times raster-only / lidar-only / full ticks per iteration with CUDA events and samples NVML
clocks + power at ~1 kHz.  usage (GPU box): python profiles/exp_backtoback.py [iters]"""
import json
import os
import sys
import threading
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from smarts_b200 import lidar_params  # noqa: E402
from smarts_b200.engine import Engine  # noqa: E402


class Nvml(threading.Thread):
    def __init__(self):
        super().__init__(daemon=True)
        import pynvml
        pynvml.nvmlInit()
        self.nv, self.h = pynvml, pynvml.nvmlDeviceGetHandleByIndex(0)
        self.rows, self.stop_evt = [], threading.Event()

    def run(self):
        while not self.stop_evt.is_set():
            try:
                self.rows.append((time.perf_counter(), self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM),
                                  self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_MEM),
                                  self.nv.nvmlDeviceGetPowerUsage(self.h) / 1000.0,
                                  self.nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)))
            except Exception:
                pass
            time.sleep(0.001)

    def window(self, t0, t1):
        r = [x for x in self.rows if t0 <= x[0] <= t1]
        if not r:
            return None
        return {"n": len(r), "sm_mhz_min": min(x[1] for x in r), "sm_mhz_med": float(np.median([x[1] for x in r])),
                "mem_mhz_min": min(x[2] for x in r), "power_w_max": max(x[3] for x in r),
                "reasons_or": hex(int(np.bitwise_or.reduce([x[4] for x in r])))}


def run_case(name, cfg_kw, iters, nv, pause=0.0):
    w = dict(bench.WORKLOADS["4lane_lidar_ogm_dagm_256"])
    w["cfg"] = dict(cfg_kw, route_cap=16)
    E, A, M = w["E"], w["A"], w["M"]
    rm, init, traffic = bench.build_inputs(w, E)
    cfg = bench.make_cfg(E, A, M, 0, **w["cfg"])
    eng = Engine(cfg, rm, lidar_params.base_rays(lidar_params.BasicLidar))
    eng.reset(*init)
    rec = torch.from_numpy(traffic.recording(4)).cuda()
    act = torch.zeros((E, A), dtype=torch.int32, device="cuda")
    eng.set_social(rec[0]); eng.step(act); eng.observe()
    torch.cuda.synchronize()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(iters)]
    t0 = time.perf_counter()
    for i in range(iters):
        evs[i][0].record()
        eng.observe()
        evs[i][1].record()
        if pause:
            torch.cuda.synchronize(); time.sleep(pause)
    torch.cuda.synchronize()
    t1 = time.perf_counter()
    ms = [a.elapsed_time(b) for a, b in evs]
    out = {"case": name, "pause_s": pause, "iters": iters, "first5_ms": [round(x, 4) for x in ms[:5]],
           "last5_ms": [round(x, 4) for x in ms[-5:]], "median_ms": float(np.median(ms)), "min_ms": min(ms),
           "nvml": nv.window(t0, t1)}
    print(json.dumps(out), flush=True)
    eng.close()


if __name__ == "__main__":
    iters = int(sys.argv[1]) if len(sys.argv) > 1 else 200
    nv = Nvml(); nv.start()
    grids = dict(ogm=(256, 256, 0.1), dagm=(256, 256, 0.1))
    run_case("raster_only", grids, iters, nv)
    run_case("raster_only_paused", grids, 20, nv, pause=0.05)
    run_case("lidar_only", dict(lidar=True, wall=1.5), iters, nv)
    run_case("lidar_only_paused", dict(lidar=True, wall=1.5), 20, nv, pause=0.05)
    run_case("observe_all", dict(grids, lidar=True, wall=1.5), iters, nv)
    nv.stop_evt.set()
