"""Golden vectors for k_ttc / the oracle's observe_ttc from the reference's OWN lane_ttc
(smarts/env/custom_observations.py:148-290), executed in this container.

custom_observations.py is loaded on its own (smarts/env/__init__.py needs gym.envs); its import of
smarts.core.sensors.Observation -- a type annotation only -- is satisfied by a stand-in module, because the
real one pulls in smarts.sstudio.  The Waypoint / Heading objects handed to lane_ttc are the reference's own
classes (smarts/core/road_map.py:556-618, coordinates.py:169-252).  Inputs are observations of the CPU
oracle on 4lane / figure_eight (every path and every neighbour kept: wp_out_len = lookahead + 1, 16 paths,
neighbour cap = V), stored next to the outputs in tests/golden/ttc.npz.  Container-only script."""
from __future__ import annotations

import importlib.util
import os
import sys
import types
from collections import namedtuple

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, HERE)
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import ref_shims  # noqa: E402

ref_shims.install()
_sens = types.ModuleType("smarts.core.sensors")
_sens.Observation = object
sys.modules["smarts.core.sensors"] = _sens
sys.modules["gym"].spaces = sys.modules["gym.spaces"]  # the stand-in package: gym.spaces.Dict(...) builds a dummy
_spec = importlib.util.spec_from_file_location("ref_custom_observations",
                                               ref_shims.REF + "/smarts/env/custom_observations.py")
ref_co = importlib.util.module_from_spec(_spec)
_spec.loader.exec_module(ref_co)

from smarts.core.coordinates import Heading  # noqa: E402
from smarts.core.road_map import Waypoint  # noqa: E402

from conftest import load_map, make_cfg, oracle_world  # noqa: E402
from smarts_b200 import scenario_gen  # noqa: E402

Ego = namedtuple("Ego", "position heading speed steering")
Veh = namedtuple("Veh", "position speed lane_id")
Obs = namedtuple("Obs", "ego_vehicle_state waypoint_paths neighborhood_vehicle_states")

CASES = {
    "4lane": [("edge-south-SN", 1, 10.0, "edge-west-EW"), ("edge-west-WE", 0, 10.0, "edge-east-WE"),
              ("edge-north-NS", 0, 10.0, "edge-south-NS"), ("edge-east-EW", 1, 10.0, "edge-west-EW")],
    "figure_eight": [("gneE20", 0, 5.0, "gneE27"), ("gneE36", 1, 5.0, "gneE25"), ("gneE20", 2, 25.0, "gneE27"),
                     ("gneE36", 0, 30.0, "gneE25")],
}


def main():
    out = {}
    for name, missions in CASES.items():
        rm = load_map(name)
        E, A, M = 6, 4, 20
        cfg = make_cfg(E=E, A=A, M=M, nbr=True, nbr_radius=60.0, nbr_cap=A + M, wp=True, wp_lookahead=32, wp_paths=16,
                       wp_len=33, route_cap=24)
        x, y, h, v, route, rlen = scenario_gen.agent_starts(rm, E, missions, 24, 61, jitter=30.0, speed=9.0)
        tr = scenario_gen.SocialTraffic(rm, E, M, 62)
        ora = oracle_world(rm, cfg, threads=4)
        ora.reset(x, y, h, v, route, rlen)
        rows, keep = [], {}
        for t in range(12):
            ora.set_social(tr.states(t))
            ora.step(np.zeros((E, A), np.int32))
            if t % 4 != 3:
                continue
            ora.observe()
            o, s = ora.obs(), ora.state()
            for e in range(E):
                for a in range(A):
                    n = int(o["wp_n_paths"][e, a])
                    cnt = int(o["nbr_count"][e, a])
                    res = np.full(9, np.nan)
                    if n and cnt:
                        paths = [[Waypoint(pos=o["wp_pos"][e, a, p, i, :2].copy(), heading=Heading(float(o["wp_heading"][e, a, p, i])),
                                           lane_id=str(int(o["wp_lane"][e, a, p, i])), lane_width=float(o["wp_lane_width"][e, a, p, i]),
                                           speed_limit=float(o["wp_speed_limit"][e, a, p, i]),
                                           lane_index=int(o["wp_lane_index"][e, a, p, i]))
                                  for i in range(int(o["wp_len"][e, a, p]))] for p in range(n)]
                        vehs = [Veh(tuple(o["nbr_pos"][e, a, k]), float(o["nbr_speed"][e, a, k]),
                                    str(int(o["nbr_lane"][e, a, k])) if o["nbr_lane"][e, a, k] >= 0 else None)
                                for k in range(cnt)]
                        ego = Ego(np.array([s["pos_x"][e, a], s["pos_y"][e, a], 0.0]), Heading(float(s["heading"][e, a])),
                                  float(s["speed"][e, a]), float(s["steering"][e, a]))
                        try:
                            r = ref_co.lane_ttc(Obs(ego, paths, vehs))
                            res = np.concatenate([[1.0], r["angle_error"], r["distance_from_center"], r["ego_lane_dist"], r["ego_ttc"]])
                        except IndexError:  # ego lane index beyond the path list (the reference raises)
                            res = np.full(9, np.inf)
                    else:
                        res = np.zeros(9)
                    rows.append(res)
            for k in ("wp_n_paths", "wp_len", "wp_pos", "wp_heading", "wp_lane_width", "wp_lane", "wp_lane_index",
                      "nbr_count", "nbr_pos", "nbr_speed", "nbr_lane"):
                keep.setdefault(k, []).append(o[k])
            for k in ("pos_x", "pos_y", "heading", "speed"):
                keep.setdefault(k, []).append(s[k][:, :A])
        out[name + "_ref"] = np.array(rows)
        for k, vv in keep.items():
            out[f"{name}_{k}"] = np.stack(vv)
        r = out[name + "_ref"]
        print(name, "agents:", len(r), "valid:", int((r[:, 0] == 1).sum()), "IndexError:", int(np.isinf(r[:, 0]).sum()),
              "with a finite ttc:", int((r[:, 6:9] < 999).any(axis=1).sum()))
    np.savez_compressed(os.path.join(HERE, "ttc.npz"), **out)


if __name__ == "__main__":
    main()
