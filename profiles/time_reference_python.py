"""the reference's OWN Python for the parts of the path that run here behind the
import shims (tests/golden/ref_shims.py), timed per call on this container's CPU, next to the C++ oracle port doing
the same calls.  Container-only (imports /root/reference); the GPU box cannot run it, so the numbers live in
profiles/reference_python_timing.json and are quoted in BASELINE.md / DESIGN.md.

Timed calls (4lane map, poses near lane points, a generated route):
  * SumoRoadNetwork.waypoint_paths(pose, lookahead=32, route)        -- the Waypoints sensor's query
  * LaneFollowingController.perform_lane_following(...)               -- the Lane action path (its own waypoint query,
                                                                         place_poles gains, control law) on a mock chassis
  * TrajectoryTrackingController.perform_trajectory_tracking_PD(...)  -- the Trajectory action path
  * RoadWaypointsSensor()                                             -- horizon 20
  * RoadMap.nearest_lanes(point, radius=10)                           -- through the restated sumolib rtree stand-in
"""
import json
import math
import os
import sys
import time
import types
from unittest import mock

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, ROOT)
import ref_shims  # noqa: E402

ref_shims.install()
_st = sys.modules["smarts.sstudio.types"]
for _n in ("EntryTactic", "TrapEntryTactic", "Via"):
    if not hasattr(_st, _n):
        setattr(_st, _n, type(_n, (), {}))
import gen_golden as gg  # noqa: E402
from conftest import load_map, make_cfg, oracle_world  # noqa: E402
from smarts.core.coordinates import Heading, Point, Pose  # noqa: E402
from smarts.core.sensors import RoadWaypointsSensor  # noqa: E402


def per_call(fn, n, warm=3):
    for i in range(warm):
        fn(i)
    t0 = time.perf_counter()
    for i in range(n):
        fn(i)
    return (time.perf_counter() - t0) / n


def main():
    rng = np.random.default_rng(1)
    ref_rm, _ = ref_shims.reference_road_map(gg.REF + gg.MAPS["4lane"])
    mine = load_map("4lane")
    route = mine.generate_route(*gg.ROUTES["4lane"][0])
    n = 200
    x, y, h = gg.sample_poses(mine, rng, n)
    poses = [gg.make_pose(x[i], y[i], h[i]) for i in range(n)]
    froute = gg.FakeRoute(ref_rm, route.road_ids)
    out = {"cpu": os.uname().machine, "nproc": os.cpu_count(), "map": "4lane", "calls": n}

    def clear():
        ref_rm._waypoints_cache = type(ref_rm._waypoints_cache)()

    def wp(i):
        clear()
        ref_rm.waypoint_paths(poses[i % n], lookahead=32, route=froute)
    out["waypoint_paths_ms"] = 1e3 * per_call(wp, n)

    def nl(i):
        ref_rm.nearest_lanes.cache_clear()
        ref_rm.nearest_lanes(Point(x[i % n], y[i % n]), radius=10.0)
    out["nearest_lanes_ms"] = 1e3 * per_call(nl, n)

    veh = types.SimpleNamespace()

    def rw(i):
        clear()
        k = i % n
        veh.pose = Pose.from_center((x[k], y[k], 0.0), Heading(h[k]))
        veh.position = np.array([x[k], y[k], 0.0])
        veh.heading = Heading(h[k])
        RoadWaypointsSensor(veh, types.SimpleNamespace(road_map=ref_rm), types.SimpleNamespace(route=froute), horizon=20)()
    out["road_waypoints_ms"] = 1e3 * per_call(rw, 60)

    t0 = time.perf_counter()
    recs = gg.gen_lane_controller(ref_rm, mine, [route], rng, n_steps=20)
    out["lane_following_ms"] = 1e3 * (time.perf_counter() - t0) / len(recs)
    t0 = time.perf_counter()
    recs, _ = gg.gen_traj_controller(rng, n=200)
    out["trajectory_tracking_ms"] = 1e3 * (time.perf_counter() - t0) / len(recs)

    # the C++ oracle port on the same kind of work: one tick of 256 Lane agents with the Waypoints sensor
    cfg = make_cfg(E=64, A=4, wp=True, route_cap=16)
    from smarts_b200 import scenario_gen
    m4 = [("edge-south-SN", 1, 10.0, "edge-west-EW"), ("edge-west-WE", 0, 10.0, "edge-east-WE"),
          ("edge-north-NS", 0, 10.0, "edge-south-NS"), ("edge-east-EW", 1, 10.0, "edge-west-EW")]
    st = scenario_gen.agent_starts(mine, 64, m4, 16, 1, jitter=20.0, speed=6.0)
    ora = oracle_world(mine, cfg, threads=1)
    ora.reset(*st)
    act = np.zeros((64, 4), np.int32)
    ora.step(act); ora.observe()
    t0 = time.perf_counter()
    for _ in range(20):
        ora.step(act)
    t_step = (time.perf_counter() - t0) / (20 * 256)
    t0 = time.perf_counter()
    for _ in range(20):
        ora.observe()
    t_obs = (time.perf_counter() - t0) / (20 * 256)
    out["oracle_port_1_thread"] = {"lane_following_plus_chassis_ms": 1e3 * t_step, "ego_lane_plus_waypoint_paths_ms": 1e3 * t_obs}
    json.dump(out, open(os.path.join(ROOT, "profiles", "reference_python_timing.json"), "w"), indent=1)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
