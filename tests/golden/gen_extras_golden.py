"""Golden vectors for k_extras / the oracle's observe_accelerometer + observe_vias from the reference's OWN
AccelerometerSensor and ViaSensor (smarts/core/sensors.py:1046-1150), executed in this container.

sensors.py cannot be imported whole (it pulls in smarts.sstudio, shapely, ...), so the three class bodies
this needs -- ViaPoint, AccelerometerSensor, ViaSensor -- and plan.py's Via dataclass are compiled from the
reference's source text, unmodified, in a namespace holding the names they use (numpy, deque, lru_cache, the
reference's own squared_dist).  ViaSensor runs against the reference's SumoRoadNetwork
(lane_by_id(...).center_at_point, behind ref_shims).  The vehicle states fed to the sensors are those of the
CPU oracle stepping the scenario of tests/test_extras.py; the sensors' outputs are stored in
tests/golden/extras.npz.  Container-only script."""
from __future__ import annotations

import ast
import os
import sys
import types
from collections import deque
from dataclasses import dataclass
from functools import lru_cache
from typing import Iterable, List, Tuple

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, HERE)
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import ref_shims  # noqa: E402

ref_shims.install()

from smarts.core.utils.math import radians_to_vec, squared_dist  # noqa: E402

from conftest import oracle_world  # noqa: E402
import test_extras as tx  # noqa: E402
from gen_golden import MAPS, REF  # noqa: E402


def ref_classes():
    ns = {"np": np, "deque": deque, "lru_cache": lru_cache, "dataclass": dataclass, "Iterable": Iterable,
          "List": List, "Tuple": Tuple, "squared_dist": squared_dist, "Sensor": object, "Plan": object}

    def pull(path, names):
        src = open(path).read()
        for node in ast.parse(src).body:
            if isinstance(node, ast.ClassDef) and node.name in names:
                seg = ast.get_source_segment(src, node)
                deco = "".join("@" + ast.get_source_segment(src, d) + "\n" for d in node.decorator_list)
                exec(compile(deco + seg, path, "exec"), ns)
    pull(ref_shims.REF + "/smarts/core/plan.py", {"Via"})
    pull(ref_shims.REF + "/smarts/core/sensors.py", {"ViaPoint", "AccelerometerSensor", "ViaSensor"})
    return ns


def main():
    ns = ref_classes()
    Via, Accel, ViaSensor = ns["Via"], ns["AccelerometerSensor"], ns["ViaSensor"]
    out = {}
    for name in sorted(tx.CASES):
        ref_rm, _net = ref_shims.reference_road_map(REF + MAPS[name])
        rm, A, starts, vias, acts = tx.scenario(name)
        E = starts[0].shape[0]
        cfg = tx.extras_cfg(E, A)
        ora = oracle_world(rm, cfg, threads=4)
        lane_ids = rm.lane_ids
        cnt, vlane, vx, vy, vhit, vspd = vias
        sensors = {}

        def new_sensors(e, a):
            mission_vias = tuple(
                Via(lane_id=lane_ids[vlane[e, a, k]], road_id="r", lane_index=k, position=(float(vx[e, a, k]), float(vy[e, a, k])),
                    hit_distance=float(vhit[e, a, k]), required_speed=float(vspd[e, a, k])) for k in range(cnt[e, a]))
            veh = types.SimpleNamespace(position=None, speed=None)
            plan = types.SimpleNamespace(road_map=ref_rm, mission=types.SimpleNamespace(via=mission_vias))
            sensors[e, a] = (Accel(veh), ViaSensor(vehicle=veh, plan=plan, lane_acquisition_range=40, speed_accuracy=1.5),
                             veh, mission_vias)
        for e in range(E):
            for a in range(A):
                new_sensors(e, a)
        rec = {k: [] for k in tx.OUT_KEYS}
        agree = {"n": 0}

        def observe(t):
            s, o = ora.state(), ora.obs()
            row = {k: np.zeros_like(o[k]) for k in tx.OUT_KEYS}
            row["via_near_idx"][:] = -1
            row["via_hit_idx"][:] = -1
            for e in range(E):
                for a in range(A):
                    acc, via, veh, mvias = sensors[e, a]
                    hd, sp, yr = float(s["heading"][e, a]), float(s["speed"][e, a]), float(s["yaw_rate"][e, a])
                    vh = radians_to_vec(hd)
                    lin = np.array((vh[0], vh[1], 0.0)) * sp  # BoxChassis.velocity_vectors, chassis.py:276-280
                    ang = np.array((0.0, 0.0, yr))
                    la, aa, lj, aj = acc(lin, ang, cfg.dt)
                    row["acc_linear"][e, a], row["acc_angular"][e, a] = la, aa
                    row["acc_linear_jerk"][e, a], row["acc_angular_jerk"][e, a] = lj, aj
                    veh.position = np.array([s["pos_x"][e, a], s["pos_y"][e, a], 0.0])
                    veh.speed = sp
                    near, hit = via()
                    # ViaPoint carries no identity: recover the mission index from lane_index (set to k above)
                    row["via_near_count"][e, a] = len(near)
                    row["via_hit_count"][e, a] = len(hit)
                    for j, p in enumerate(near):
                        row["via_near_idx"][e, a, j] = p.lane_index
                    for j, p in enumerate(hit):
                        row["via_hit_idx"][e, a, j] = p.lane_index
            for k in tx.OUT_KEYS:
                rec[k].append(row[k])
                agree["n"] += int(np.allclose(row[k], o[k], rtol=1e-12, atol=1e-12))  # sin / cos: libm vs sb_detmath.h

        # run() resets RESET_ENVS at RESET_AT: hook the oracle's reset to renew those agents' sensors
        class Hooked:
            def reset(self, *a, env_ids=None, **k):
                ora.reset(*a, env_ids=env_ids, **k)
                if env_ids is not None:
                    for e in env_ids:
                        for ag in range(A):
                            new_sensors(int(e), ag)

            def step(self, act):
                ora.step(act)

            def observe(self):
                ora.observe()
        tx.run(Hooked(), starts, vias, acts, observe)
        for k in tx.OUT_KEYS:
            out[f"{name}_{k}"] = np.stack(rec[k])
        total = (tx.TICKS + 1) * len(tx.OUT_KEYS)
        print(name, "observations:", tx.TICKS + 1, "oracle == reference on", agree["n"], "of", total, "arrays;",
              "hits:", int(out[f"{name}_via_hit_count"].sum()), "near (sum):", int(out[f"{name}_via_near_count"].sum()),
              "max |lin acc|:", float(np.abs(out[f"{name}_acc_linear"]).max()))
    np.savez_compressed(os.path.join(HERE, "extras.npz"), **out)


if __name__ == "__main__":
    main()
