"""Experiment: the headline batch as S env shards, each with its own handle and CUDA stream, ticked round-robin --
do the latency-bound kernels of one shard (k_control_law, k_control_paths, k_observe) hide behind the
throughput-bound ones (k_lidar, k_raster) of the others?   python profiles/exp_two_streams.py [S ...]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def run(S, E=4096, W=5, K=20, name=bench.HEADLINE):
    w = dict(bench.WORKLOADS[name])
    parts = [bench.GpuWorkload(w, E // S, 0, 0) for _ in range(S)]
    streams = [torch.cuda.Stream() for _ in range(S)]
    def tick(t):
        for g, st in zip(parts, streams):
            with torch.cuda.stream(st):
                g.tick(t)
    for t in range(W):
        tick(t)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for st in streams:
        st.wait_stream(torch.cuda.current_stream())
    for t in range(W, W + K):
        tick(t)
    for st in streams:
        torch.cuda.current_stream().wait_stream(st)
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / K
    for g in parts:
        g.close()
    return ms


if __name__ == "__main__":
    names = [bench.HEADLINE, "straight_full_suite_64"]
    for name in names:
        for S in [int(x) for x in sys.argv[1:]] or [1, 2, 4]:
            print(name, "shards", S, "ms/step", round(run(S, name=name), 4), flush=True)
