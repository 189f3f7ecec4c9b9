"""Experiment: two env shards, each with a handle; a shard's control kernels (set_social + step: latency-bound, low
occupancy) go to a HIGH-priority stream and its sensor kernels (observe: lidar, raster) to a normal one, and shard 1
starts half a tick after shard 0 -- does one shard's control hide behind the other's raster?
    python profiles/exp_skewed_shards.py"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def run(S, skew, prio, E=4096, W=5, K=20, name=bench.HEADLINE):
    w = dict(bench.WORKLOADS[name])
    parts = [bench.GpuWorkload(w, E // S, 0, 0) for _ in range(S)]
    lo_p, hi_p = torch.cuda.Stream.priority_range() if hasattr(torch.cuda.Stream, "priority_range") else (0, -1)
    hi = [torch.cuda.Stream(priority=-1 if prio else 0) for _ in range(S)]
    lo = [torch.cuda.Stream(priority=0) for _ in range(S)]
    obs_done = [None] * S

    def tick(t):
        for s, g in enumerate(parts):
            if obs_done[s] is not None:
                hi[s].wait_event(obs_done[s])              # the policy reads the observation before it acts
            with torch.cuda.stream(hi[s]):
                if g.traj:
                    g.policy()
                g.eng.set_social(g.rec[t % g.T])
                g.eng.step(g.actions)
                ev = torch.cuda.Event(); ev.record()
            lo[s].wait_event(ev)
            if skew and t == 0 and s > 0:
                lo[s].wait_event(first_ctl)                  # shard s starts its sensors when shard 0 has started its own
            with torch.cuda.stream(lo[s]):
                g.eng.observe()
                obs_done[s] = torch.cuda.Event(); obs_done[s].record()
            if s == 0 and t == 0:
                first_ctl_holder.append(obs_done[0])

    first_ctl_holder = []
    first_ctl = None
    # warm-up (establishes the skew: shard 1's first sensors wait for shard 0's first observation)
    for t in range(W):
        if t == 0 and skew:
            # run shard 0's tick first, then let the others follow
            g = parts[0]
            with torch.cuda.stream(hi[0]):
                g.eng.set_social(g.rec[0]); g.eng.step(g.actions); ev = torch.cuda.Event(); ev.record()
            lo[0].wait_event(ev)
            with torch.cuda.stream(lo[0]):
                g.eng.observe(); obs_done[0] = torch.cuda.Event(); obs_done[0].record()
            first_ctl = ev
            for s in range(1, S):
                g = parts[s]
                hi[s].wait_event(ev)
                with torch.cuda.stream(hi[s]):
                    g.eng.set_social(g.rec[0]); g.eng.step(g.actions); e2 = torch.cuda.Event(); e2.record()
                lo[s].wait_event(e2)
                with torch.cuda.stream(lo[s]):
                    g.eng.observe(); obs_done[s] = torch.cuda.Event(); obs_done[s].record()
            continue
        tick(t)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for st in hi + lo:
        torch.cuda.current_stream().wait_stream(st)
    a.record()
    for st in hi + lo:
        st.wait_stream(torch.cuda.current_stream())
    for t in range(W, W + K):
        tick(t)
    for st in hi + lo:
        torch.cuda.current_stream().wait_stream(st)
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / K
    for g in parts:
        g.close()
    return ms


if __name__ == "__main__":
    for S, skew, prio in [(1, False, False), (2, False, False), (2, True, False), (2, False, True), (2, True, True), (4, True, True)]:
        print("shards", S, "skew", skew, "priority", prio, "ms/step", round(run(S, skew, prio), 4), flush=True)
