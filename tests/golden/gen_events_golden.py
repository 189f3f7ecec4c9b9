"""Golden vectors for the events pass (k_events / the oracle's observe_events) from the reference's OWN code:
`Sensors._is_done_with_events` with everything it calls (`_agent_reached_goal`, `_vehicle_is_off_road`,
`_vehicle_is_on_shoulder`, `_vehicle_is_not_moving`, `_vehicle_is_off_route_and_wrong_way`, `_check_wrong_way_event`;
smarts/core/sensors.py:443-594), `TripMeterSensor`, `DrivenPathSensor`, `SensorState` (sensors.py:609-637, :830-947),
`Vehicle.bounding_box` (vehicle.py:315-332), `Mission.is_complete` / `PositionalGoal.is_reached` (plan.py:95-124,
:201-222), `DoneCriteria` / `EventConfiguration` (agent_interface.py:174-206) -- `smarts.core.sensors` imported whole
behind tests/golden/ref_shims.py -- running against the reference's own `SumoRoadNetwork` (`road_with_point`,
`nearest_lane`, `offset_along_lane`, `oncoming_lanes_at_offset`, `center_pose_at_point`, `waypoint_paths`).

The order of calls per observation is that of `Sensors.observe` (sensors.py:268-372) and `SMARTS._step` / `_reset`
(smarts.py:225-300, :391-430): sensor step -> waypoint query -> trip meter -> driven path -> `_is_done_with_events`.
Vehicle states are the CPU oracle's over the rollout of tests/test_events_golden.py; vehicle contacts (pybullet in the
reference) are injected from the oracle.  Output: tests/golden/events.npz.  Container-only script."""
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
for _n in ("EntryTactic", "TrapEntryTactic", "Via"):  # names plan.py imports from the scenario DSL; never used here
    setattr(_st, _n, type(_n, (), {}))

from smarts.core.agent_interface import DoneCriteria, EventConfiguration  # noqa: E402
from smarts.core.coordinates import Dimensions, Heading, Point, Pose  # noqa: E402
from smarts.core.plan import EndlessGoal, Mission, PositionalGoal, Start  # noqa: E402
from smarts.core.sensors import DrivenPathSensor, Sensors, SensorState, TripMeterSensor  # noqa: E402
from smarts.core.utils.math import rounder_for_dt  # noqa: E402
from smarts.core.vehicle import Vehicle  # noqa: E402

from conftest import oracle_world  # noqa: E402
import test_events_golden as tg  # noqa: E402
from gen_golden import MAPS, REF, FakeRoute  # noqa: E402
from smarts_b200 import _abi  # noqa: E402


class FakeVehicle:
    """The attributes the reference code reads from a Vehicle; bounding_box IS the reference's property."""
    bounding_box = Vehicle.bounding_box

    def __init__(self, vid):
        self.id = vid

    def _assert_initialized(self):
        pass

    def set(self, x, y, z, h, speed, length, width, height):
        self.pose = Pose.from_center((x, y, z), Heading(h))
        self.position = np.array([x, y, z])
        self.heading = Heading(h)
        self.speed = speed
        self.length, self.width, self.height = length, width, height
        self.chassis = types.SimpleNamespace(dimensions=Dimensions(length, width, height))


def events_to_bits(ev):
    b = 0
    b |= _abi.SB_EV_COLLISION if len(ev.collisions) else 0
    b |= _abi.SB_EV_OFF_ROAD if ev.off_road else 0
    b |= _abi.SB_EV_OFF_ROUTE if ev.off_route else 0
    b |= _abi.SB_EV_ON_SHOULDER if ev.on_shoulder else 0
    b |= _abi.SB_EV_WRONG_WAY if ev.wrong_way else 0
    b |= _abi.SB_EV_NOT_MOVING if ev.not_moving else 0
    b |= _abi.SB_EV_REACHED_GOAL if ev.reached_goal else 0
    b |= _abi.SB_EV_REACHED_MAX_STEPS if ev.reached_max_episode_steps else 0
    return b


def main():
    ref_rm, _net = ref_shims.reference_road_map(REF + MAPS["4lane"])
    ref_shims.set_tie_ranks(ref_rm, tg.load_map("4lane"))
    out = {}
    for mode, wp in tg.MODES.items():
        rm, starts, goal, acts, traffic = tg.scenario()
        cfg = tg.events_cfg(wp)
        ora = oracle_world(rm, cfg, threads=4)
        E, A = tg.E, tg.A
        x0, y0, h0, v0, route, rlen = starts
        gx, gy, gr = goal
        iface = types.SimpleNamespace(
            done_criteria=DoneCriteria(collision=True, off_road=True, off_route=True, on_shoulder=True, wrong_way=True,
                                       not_moving=True),
            event_configuration=EventConfiguration(not_moving_time=cfg.not_moving_time,
                                                   not_moving_distance=cfg.not_moving_distance))
        rounder = rounder_for_dt(cfg.dt)
        agents = {}
        clock = np.zeros(E)        # SMARTS._elapsed_sim_time per env (every env is its own SMARTS instance)
        collided = {}

        def new_agent(e, a):
            """Vehicle creation at the mission start (vehicle.py:344-457): fresh sensors and sensor state."""
            veh = FakeVehicle(f"v-{e}-{a}")
            veh.set(x0[e, a], y0[e, a], 0.0, h0[e, a], v0[e, a], 3.68, 1.47, 1.0)
            g = (PositionalGoal(position=Point(gx[e, a], gy[e, a]), radius=gr[e, a]) if gr[e, a] >= 0 else EndlessGoal())
            mission = Mission(start=Start(np.array([x0[e, a], y0[e, a]]), Heading(h0[e, a])), goal=g)
            ids = [rm.road_ids[r] for r in route[e, a, :rlen[e, a]]]
            plan = types.SimpleNamespace(mission=mission, route=FakeRoute(ref_rm, ids), road_map=ref_rm)
            sim = types.SimpleNamespace(
                road_map=ref_rm, scenario=types.SimpleNamespace(road_map=ref_rm), elapsed_sim_time=0.0, resetting=True,
                vehicle_index=types.SimpleNamespace(sensor_state_for_vehicle_id=lambda vid: agents[vid][3]),
                agent_manager=types.SimpleNamespace(agent_interface_for_agent_id=lambda aid: iface,
                                                    ego_agent_ids=set(), agent_ids=set()),
                vehicle_did_collide=lambda vid: collided[vid] > 0,
                vehicle_collisions=lambda vid: [object()] * collided[vid])
            veh.trip_meter_sensor = TripMeterSensor(veh, sim, plan)
            veh.driven_path_sensor = DrivenPathSensor(veh)
            agents[veh.id] = (veh, sim, plan, SensorState(cfg.max_episode_steps or None, plan))

        for e in range(E):
            for a in range(A):
                new_agent(e, a)
        rec = {k: [] for k in tg.OUT_KEYS}
        rec["observation_distance_travelled"] = []
        agree = {"n": 0, "of": 0}

        def observe(t, ids):
            s, o = ora.state(), ora.obs()
            envs = range(E) if ids is None else [int(i) for i in ids]
            row = {k: (rec[k][-1].copy() if rec[k] else np.zeros_like(o[k])) for k in tg.OUT_KEYS}
            obs_dist = rec["observation_distance_travelled"][-1].copy() if rec["observation_distance_travelled"] else np.zeros((E, A))
            for e in envs:
                if ids is not None:      # SMARTS.reset of this env: new vehicles, clock back to 0, no sensor step
                    clock[e] = 0.0
                    for a in range(A):
                        new_agent(e, a)
                elif t >= 0:             # SMARTS._step: clock first (smarts.py:262)
                    clock[e] = rounder(clock[e] + cfg.dt)
                for a in range(A):
                    veh, sim, plan, st = agents[f"v-{e}-{a}"]
                    stepping = ids is None and t >= 0
                    sim.resetting = not stepping
                    sim.elapsed_sim_time = float(clock[e])
                    veh.set(s["pos_x"][e, a], s["pos_y"][e, a], s["pos_z"][e, a], s["heading"][e, a], s["speed"][e, a],
                            s["length"][e, a], s["width"][e, a], s["height"][e, a])
                    collided[veh.id] = int(o["ev_n_collisions"][e, a])
                    if stepping:
                        Sensors.step(sim, st)                                  # agent_manager.step_sensors
                    # Sensors.observe, sensors.py:268-276, :349-353, :367-369
                    if wp:
                        paths = ref_rm.waypoint_paths(veh.pose, lookahead=cfg.wp_lookahead, route=plan.route)
                    else:
                        paths = ref_rm.waypoint_paths(veh.pose, lookahead=1, within_radius=veh.length)
                    if paths:
                        veh.trip_meter_sensor.append_waypoint_if_new(paths[0][0])
                    obs_dist[e, a] = veh.trip_meter_sensor(sim)                 # Observation.distance_travelled
                    veh.driven_path_sensor.track_latest_driven_path(sim)
                    done, events = Sensors._is_done_with_events(sim, "agent", veh, st)
                    row["ev_events"][e, a] = events_to_bits(events)
                    row["ev_done"][e, a] = int(done)
                    row["ev_reward"][e, a] = veh.trip_meter_sensor(increment=True)     # agent_manager.py:233
                    row["ev_distance"][e, a] = veh.trip_meter_sensor()                 # agent_manager.py:234 (score)
                    row["ev_steps"][e, a] = st.steps_completed
            for k in tg.OUT_KEYS:
                rec[k].append(row[k])
                agree["of"] += 1
                same = np.allclose(row[k], o[k], rtol=0, atol=1e-9) if row[k].dtype.kind == "f" else np.array_equal(row[k], o[k])
                agree["n"] += int(same)
                if not same:
                    bad = np.argwhere(~np.isclose(row[k].astype(float), o[k].astype(float), rtol=0, atol=1e-9))
                    print(f"  [{mode}] t={t} ids={ids} {k}: {len(bad)} differ, first {bad[0]}: reference "
                          f"{row[k][tuple(bad[0])]} oracle {o[k][tuple(bad[0])]}")
            rec["observation_distance_travelled"].append(obs_dist)

        tg.run(ora, starts, goal, acts, traffic, observe)
        for k in rec:
            out[f"{mode}_{k}"] = np.stack(rec[k])
        ev = out[f"{mode}_ev_events"]
        print(mode, "observations:", len(rec["ev_events"]), "oracle == reference on", agree["n"], "of", agree["of"], "arrays; bits seen:",
              bin(int(np.bitwise_or.reduce(ev.ravel()))), "done:", int(out[f"{mode}_ev_done"].sum()),
              "max distance:", float(out[f"{mode}_ev_distance"].max()))
        # sensors.py:351 passes `sim` as TripMeterSensor.__call__'s `increment`: the Observation member holds the
        # INCREMENT of this observation, not the total (the total is the score, agent_manager.py:234)
        assert np.allclose(out[f"{mode}_observation_distance_travelled"], out[f"{mode}_ev_reward"])
    np.savez_compressed(os.path.join(HERE, "events.npz"), **out)


if __name__ == "__main__":
    main()
