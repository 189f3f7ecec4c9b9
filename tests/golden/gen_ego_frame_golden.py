"""Golden vectors for k_ego_frame / the oracle's ego_frame_agent from the reference's OWN
ego_centric_observation_adapter (smarts/core/utils/adapters/ego_centric_adapters.py:62-178), executed in this
container.

The adapter module imports smarts.core.sensors / plan / controllers (which pull in sstudio, shapely, ...), so
the function `ego_centric_observation_adapter` and its helpers `_isnamedtupleinstance` / `_replace` are
compiled from the reference's source text, unmodified, in a namespace holding the names they use: numpy, math,
dataclasses, and the reference's own Heading, position_to_ego_frame and wrap_value (smarts/core/coordinates.py,
smarts/core/utils/math.py:452-487, imported for real).  PositionalGoal / Via / ViaPoint are only used in
isinstance checks and type hints there; stand-in classes are supplied.

Inputs are world-frame observations of the CPU oracle (ego_frame = 0) wrapped in namedtuples with the
Observation field names the adapter touches; outputs are stored in tests/golden/ego_frame.npz in the buffer
layout.  Lidar: with the reference's (inf, inf, inf) misses every point goes through the adapter; with the
zero fill of StdObs only the hit points do (a zero-filled miss is not a point).  Container-only script."""
from __future__ import annotations

import ast
import math
import os
import sys
from collections import namedtuple
from dataclasses import is_dataclass
from dataclasses import replace as dc_replace
from typing import Any, Callable, Dict, Optional, Sequence, Tuple, Union

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, HERE)
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import ref_shims  # noqa: E402

ref_shims.install()

from smarts.core.coordinates import Heading  # noqa: E402
from smarts.core.utils.math import position_to_ego_frame, world_position_from_ego_frame, wrap_value  # noqa: E402

from conftest import load_map, oracle_world  # noqa: E402
import test_ego_frame as tx  # noqa: E402


class _Never:  # isinstance(goal, PositionalGoal) is False for the EndlessGoal stand-in below
    pass


def ref_adapter():
    ns = {"np": np, "math": math, "is_dataclass": is_dataclass, "dc_replace": dc_replace, "Any": Any,
          "Callable": Callable, "Dict": Dict, "Optional": Optional, "Sequence": Sequence, "Tuple": Tuple, "Union": Union,
          "Heading": Heading, "position_to_ego_frame": position_to_ego_frame,
          "world_position_from_ego_frame": world_position_from_ego_frame, "wrap_value": wrap_value,
          "Observation": object, "PositionalGoal": _Never, "Via": _Never, "ViaPoint": _Never, "Vias": _Never, "Goal": _Never}
    path = ref_shims.REF + "/smarts/core/utils/adapters/ego_centric_adapters.py"
    src = open(path).read()
    for node in ast.parse(src).body:
        if isinstance(node, ast.FunctionDef) and node.name in ("_isnamedtupleinstance", "_replace",
                                                               "ego_centric_observation_adapter"):
            exec(compile(ast.get_source_segment(src, node), path, "exec"), ns)
    return ns["ego_centric_observation_adapter"]


Obs = namedtuple("Obs", "ego_vehicle_state neighborhood_vehicle_states waypoint_paths via_data road_waypoints "
                        "lidar_point_cloud drivable_area_grid_map occupancy_grid_map top_down_rgb")
Ego = namedtuple("Ego", "position heading linear_velocity linear_acceleration linear_jerk mission")
MissionT = namedtuple("MissionT", "start via goal")
Start = namedtuple("Start", "position heading")
Veh = namedtuple("Veh", "position heading")
Wp = namedtuple("Wp", "pos heading")
Cam = namedtuple("Cam", "metadata data")
Meta = namedtuple("Meta", "camera_pos camera_heading_in_degrees")


def main():
    adapter = ref_adapter()
    out = {}
    for name in sorted(tx.CASES):
        rm = load_map(name)
        cfg = tx.cfg_for(name, False)
        A, P, L, R, cap = cfg.n_agents, cfg.wp_out_paths, cfg.wp_out_len, cfg.lidar_rays, cfg.nbr_cap
        world = oracle_world(rm, cfg, threads=4)
        mine = oracle_world(rm, tx.cfg_for(name, True), threads=4)
        rec = {k: [] for k in tx.KEYS}
        worst = 0.0
        for t in tx.drive(name, [(world, lambda a: a), (mine, lambda a: a)]):
            o, s, m = world.obs(), world.state(), mine.obs()
            row = {k: o[k].astype(np.float64).copy() for k in tx.KEYS}
            for e in range(tx.E):
                for a in range(A):
                    pos = np.array([s["pos_x"][e, a], s["pos_y"][e, a], s["pos_z"][e, a]])
                    hd = Heading(float(s["heading"][e, a]))
                    n_nb = min(int(o["nbr_count"][e, a]), cap)
                    nvs = [Veh(o["nbr_pos"][e, a, k].copy(), Heading(float(o["nbr_heading"][e, a, k]))) for k in range(n_nb)]
                    n_p = min(int(o["wp_n_paths"][e, a]), P)
                    wps = [[Wp(o["wp_pos"][e, a, p, i, :2].copy(), Heading(float(o["wp_heading"][e, a, p, i])))
                            for i in range(int(o["wp_len"][e, a, p]))] for p in range(n_p)]
                    pts = o["lidar_points"][e, a]
                    idx = np.arange(R) if cfg.lidar_miss_inf else np.flatnonzero(o["lidar_hits"][e, a])
                    lidar = [[pts[i].copy() for i in idx], [bool(o["lidar_hits"][e, a, i]) for i in idx],
                             [(pos.copy(), pos + 1.0) for _ in idx]] if len(idx) else []
                    cam = Cam(Meta((1.0, 2.0, 3.0), 45.0), None)
                    ego = Ego(pos, hd, np.array([1.0, 2.0, 0.0]), np.zeros(3), np.zeros(3),
                              MissionT(Start(pos[:2].copy(), hd), (), object()))
                    with np.errstate(invalid="ignore"):
                        r = adapter(Obs(ego, nvs, wps, None, None, lidar, cam, cam, cam))
                    assert list(r.ego_vehicle_state.position) == [0, 0, 0] and float(r.ego_vehicle_state.heading) == 0.0
                    for k, nv in enumerate(r.neighborhood_vehicle_states):
                        row["nbr_pos"][e, a, k] = nv.position
                        row["nbr_heading"][e, a, k] = float(nv.heading)
                    for p, path in enumerate(r.waypoint_paths):
                        for i, wp in enumerate(path):
                            row["wp_pos"][e, a, p, i, :2] = wp.pos
                            row["wp_heading"][e, a, p, i] = float(wp.heading)
                    if len(idx):
                        for j, i in enumerate(idx):
                            row["lidar_points"][e, a, i] = r.lidar_point_cloud[0][j]
            for k in tx.KEYS:
                rec[k].append(row[k])
                d = np.abs(m[k].astype(np.float64) - row[k])
                assert np.array_equal(np.isnan(m[k]), np.isnan(row[k])), (name, t, k)
                worst = max(worst, float(np.nanmax(d)))
        for k in tx.KEYS:
            out[f"{name}_{k}"] = np.stack(rec[k])
        print(name, "observations:", len(rec["nbr_pos"]), "max |oracle - reference adapter|:", worst,
              "NaN lidar points:", int(np.isnan(out[f"{name}_lidar_points"]).any(axis=-1).sum()))
    np.savez_compressed(os.path.join(HERE, "ego_frame.npz"), **out)


if __name__ == "__main__":
    main()
