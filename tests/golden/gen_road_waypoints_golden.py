"""Golden vectors for the RoadWaypoints sensor (k_road_waypoints / the oracle's observe_road_waypoints) from the
reference's OWN code: `RoadWaypointsSensor.__call__` and `paths_for_lane` (smarts/core/sensors.py:991-1043) --
`smarts.core.sensors` imported whole behind tests/golden/ref_shims.py -- running against the reference's own
`SumoRoadNetwork` (`nearest_lane`, `Road.parallel_roads`, `Road.oncoming_roads_at_point`,
`Lane.oncoming_lanes_at_offset`, `Lane.incoming_lanes`, `Lane.offset_along_lane`, `Lane.from_lane_coord`,
`Lane.waypoint_paths_for_pose`, `LanePoints.paths_starting_at_lanepoint`, `_equally_spaced_path`) over the
reference's net.xml files (sumolib's accessors restated in smarts_b200/sumo_net.py, incl. Edge.getFromNode /
Node.getOutgoing for `parallel_roads`).

Poses: near lane points of 4lane / 6lane / figure_eight / loop with lateral and heading jitter (on lanes, between
lanes, inside junctions, beside oncoming roads), each with a generated route or with an empty route, horizons 20
(the AgentInterface default) and 32.  Output: tests/golden/road_waypoints.npz.  Container-only script."""
from __future__ import annotations

import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, HERE)
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import ref_shims  # noqa: E402

ref_shims.install()
_st = sys.modules["smarts.sstudio.types"]
for _n in ("EntryTactic", "TrapEntryTactic", "Via"):
    if not hasattr(_st, _n):
        setattr(_st, _n, type(_n, (), {}))

from smarts.core.coordinates import Heading, Pose  # noqa: E402
from smarts.core.sensors import RoadWaypointsSensor  # noqa: E402

from conftest import load_map  # noqa: E402
from gen_golden import MAPS, REF, ROUTES, FakeRoute, sample_poses  # noqa: E402

N_POSES = {"4lane": 48, "6lane": 24, "figure_eight": 24, "loop": 12}
HORIZONS = (20, 32)


class FakeVehicle:
    def set(self, x, y, h):
        self.pose = Pose.from_center((x, y, 0.0), Heading(h))
        self.position = np.array([x, y, 0.0])
        self.heading = Heading(h)


def main():
    out = {}
    rng = np.random.default_rng(20261105)
    for name, n in N_POSES.items():
        ref_rm, _net = ref_shims.reference_road_map(REF + MAPS[name])
        mine = load_map(name)
        ref_shims.set_tie_ranks(ref_rm, mine)
        routes = [mine.generate_route(a, b).road_ids for a, b in ROUTES.get(name, [])]
        x, y, h = sample_poses(mine, rng, n)
        # a few poses right at lane starts / ends, where the walk back through incoming lanes happens
        ends = rng.integers(0, len(mine.lane_ids), n // 4)
        for k, li in enumerate(ends):
            s = int(mine.lane_lp_start[li])
            if mine.lane_lp_count[li] > 0:
                x[k], y[k], h[k] = mine.lp_x[s] + rng.normal(0, 0.3), mine.lp_y[s] + rng.normal(0, 0.3), mine.lp_heading[s]
        route_of = np.array([(i % (len(routes) + 1)) - 1 for i in range(n)], np.int32)   # -1: empty route
        sim = types.SimpleNamespace(road_map=ref_rm)
        veh = FakeVehicle()
        meta, rows, lanes_meta = [], [], []
        for hz in HORIZONS:
            for i in range(n):
                veh.set(x[i], y[i], h[i])
                ids = routes[route_of[i]] if route_of[i] >= 0 else []
                plan = types.SimpleNamespace(route=FakeRoute(ref_rm, ids))
                ref_rm._waypoints_cache = type(ref_rm._waypoints_cache)()
                sensor = RoadWaypointsSensor(veh, sim, plan, horizon=hz)
                res = sensor()
                for slot, (lane_id, paths) in enumerate(res.lanes.items()):
                    lanes_meta.append((hz, i, slot, mine.lane_index[lane_id], len(paths)))
                    for p, path in enumerate(paths):
                        meta.append((hz, i, mine.lane_index[lane_id], p, len(path), len(rows)))
                        for w in path:
                            rows.append([w.pos[0], w.pos[1], float(w.heading), w.lane_width, w.speed_limit,
                                         mine.lane_index[w.lane_id], w.lane_index])
        out[f"{name}_x"], out[f"{name}_y"], out[f"{name}_h"] = x, y, h
        out[f"{name}_route_of"] = route_of
        out[f"{name}_routes"] = np.array([";".join(r) for r in routes] or [""])
        out[f"{name}_lanes"] = np.array(lanes_meta, np.int64)      # horizon, pose, slot, lane, n paths
        out[f"{name}_meta"] = np.array(meta, np.int64)             # horizon, pose, lane, path, length, first row
        out[f"{name}_rows"] = np.array(rows, np.float64)
        lm = out[f"{name}_lanes"]
        print(f"{name}: {n} poses x {len(HORIZONS)} horizons, {len(lm)} dict entries (max {max(np.bincount(lm[:, 1] + n * (lm[:, 0] == 32)))} lanes "
              f"per call), {len(meta)} paths, {len(rows)} waypoints")
    np.savez_compressed(os.path.join(HERE, "road_waypoints.npz"), **out)


if __name__ == "__main__":
    main()
