"""Social-traffic wire format and loaders: everything that turns provider
records into the ``[T, E, M, 8]`` tensor ``sb_set_social`` consumes, one tick at a time
(x, y, heading, speed, length, width, height, alive; centre pose, heading in radians, 0 = +y,
counter-clockwise).  SUMO itself stays out of the kernel path (BASELINE.json north_star): these
functions only convert what a recorded or external provider already produced.

  ProviderState / VehicleState            smarts/core/provider.py:42-67, vehicle.py:53-80
  ExternalProvider.state_update           smarts/core/external_provider.py:44-51
  SUMO state conversion                   smarts/core/sumo_traffic_simulation.py:672-705
      Heading.from_sumo                   smarts/core/coordinates.py:204-210, :187-202
      Pose.from_front_bumper              smarts/core/coordinates.py:302-321
  TrafficHistoryProvider.step             smarts/core/traffic_history_provider.py:96-136
      TrafficHistory.vehicles_active_between / decode_vehicle_type   traffic_history.py:127-146, :208-231
      Dimensions.init_with_defaults       smarts/core/coordinates.py:54-65
"""
from __future__ import annotations

import math
import sqlite3
from dataclasses import dataclass
from typing import Dict, Iterable, List, Optional, Sequence, Tuple

import numpy as np

# VEHICLE_CONFIGS dimensions (length, width, height), smarts/core/vehicle.py:97-140
VEHICLE_DIMS: Dict[str, Tuple[float, float, float]] = {
    "passenger": (3.68, 1.47, 1.4), "bus": (7.0, 2.25, 3.0), "coach": (8.0, 2.4, 3.5),
    "truck": (5.0, 1.91, 1.89), "trailer": (10.0, 2.5, 4.0), "pedestrian": (0.5, 0.5, 1.6),
    "motorcycle": (2.5, 1.0, 1.4),
}
_HISTORY_TYPES = {1: "motorcycle", 2: "passenger", 3: "truck", 4: "pedestrian"}  # traffic_history.py:127-146


@dataclass
class VehicleRecord:
    """What a provider reports per vehicle per tick (VehicleState, vehicle.py:53-80), centre pose."""
    vehicle_id: str
    x: float
    y: float
    heading: float
    speed: float
    length: float
    width: float
    height: float


def heading_from_sumo(angle_deg: float) -> float:
    """Heading.from_sumo: SUMO degrees (0 = north, clockwise) -> radians, counter-clockwise, (-pi, pi]."""
    h = math.radians(angle_deg)
    h = (2 * math.pi - h) % (2 * math.pi)          # _flip_clockwise
    h = h % (2 * math.pi)                          # Heading.__new__
    return h - 2 * math.pi if h > math.pi else h


def centre_from_front_bumper(front_x: float, front_y: float, heading: float, length: float) -> Tuple[float, float]:
    """Pose.from_front_bumper: the box centre sits half a length behind the bumper along the heading."""
    a = (heading + math.pi * 0.5) % (2 * math.pi)  # radians_to_vec, utils/math.py:247-253
    return front_x - math.cos(a) * (0.5 * length), front_y - math.sin(a) * (0.5 * length)


def records_from_sumo(ids: Sequence[str], front_bumper_xy: np.ndarray, angles_deg: Sequence[float],
                      speeds: Sequence[float], vehicle_classes: Sequence[str]) -> List[VehicleRecord]:
    """One TraCI subscription result -> records (sumo_traffic_simulation.py:672-705)."""
    out = []
    for i, vid in enumerate(ids):
        l, w, h = VEHICLE_DIMS[vehicle_classes[i]]
        hd = heading_from_sumo(float(angles_deg[i]))
        cx, cy = centre_from_front_bumper(float(front_bumper_xy[i][0]), float(front_bumper_xy[i][1]), hd, l)
        out.append(VehicleRecord(str(vid), cx, cy, hd, float(speeds[i]), l, w, h))
    return out


def _dims_with_defaults(length, width, height, defaults):
    pick = lambda v, d: d if (not v or v == -1) else float(v)  # noqa: E731  (coordinates.py:59-64)
    return pick(length, defaults[0]), pick(width, defaults[1]), pick(height, defaults[2])


def records_from_history_rows(rows: Iterable[Sequence], id_prefix: str = "history-vehicle-") -> List[VehicleRecord]:
    """rows as TrafficHistory.VehicleRow (id, type, length, width, height, x, y, heading_rad, speed),
    newest first; the first row of a vehicle wins (traffic_history_provider.py:107-111)."""
    seen, out = set(), []
    for vid, vtype, length, width, height, x, y, hd, speed in rows:
        if vid in seen:
            continue
        seen.add(vid)
        cfg = _HISTORY_TYPES.get(int(vtype), "passenger")
        l, w, h = _dims_with_defaults(length, width, height, VEHICLE_DIMS[cfg])
        h_wrapped = float(hd) % (2 * math.pi)
        if h_wrapped > math.pi:
            h_wrapped -= 2 * math.pi
        out.append(VehicleRecord(id_prefix + str(vid), float(x), float(y), h_wrapped, float(speed or 0.0), l, w, h))
    return out


class SlotAssigner:
    """Stable vehicle-id -> slot map for one environment: a vehicle keeps its slot while it is present,
    a departed vehicle's slot is marked dead for one tick (so BoxChassis.yaw_rate restarts, chassis.py:302-308)
    and then reused.  Vehicles beyond M slots are dropped and counted."""

    def __init__(self, M: int):
        self.M = M
        self.slot_of: Dict[str, int] = {}
        self.free = list(range(M - 1, -1, -1))  # kept descending: pop() hands out the lowest free slot
        self.cooling: List[int] = []            # slots vacated last tick: dead for one tick before reuse
        self.dropped = 0

    def tick(self, records: Sequence[VehicleRecord]) -> np.ndarray:
        out = np.zeros((self.M, 8), np.float64)
        present = {r.vehicle_id for r in records}
        self.free = sorted(self.free + self.cooling, reverse=True)
        self.cooling = [self.slot_of.pop(vid) for vid in [v for v in self.slot_of if v not in present]]
        newly = []
        for r in records:
            s = self.slot_of.get(r.vehicle_id)
            if s is None:
                if not self.free:
                    self.dropped += 1
                    continue
                s = self.free.pop()
                newly.append((r.vehicle_id, s))
            out[s] = (r.x, r.y, r.heading, r.speed, r.length, r.width, r.height, 1.0)
        for vid, s in newly:
            self.slot_of[vid] = s
        return out


def recording_from_ticks(ticks: Sequence[Sequence[Sequence[VehicleRecord]]], M: int) -> np.ndarray:
    """ticks[t][e] = records of env e at tick t  ->  [T, E, M, 8]."""
    T, E = len(ticks), len(ticks[0])
    assigners = [SlotAssigner(M) for _ in range(E)]
    out = np.zeros((T, E, M, 8), np.float64)
    for t in range(T):
        for e in range(E):
            out[t, e] = assigners[e].tick(ticks[t][e])
    return out


def _round_param_for_dt(dt: float) -> int:
    """utils/math.py round_param_for_dt: decimals that resolve one time step."""
    strep = np.format_float_positional(dt)
    decimal_digits = len(strep) - strep.index(".") - 1
    return max(0, decimal_digits)


def load_history_sqlite(path: str, dt: float, n_ticks: int, M: int, n_envs: int = 1, start_time: float = 0.0,
                        env_time_offset: float = 0.0) -> np.ndarray:
    """Replay a traffic-history database (schema of smarts/sstudio/genhistories.py:108-135: tables
    Vehicle + Trajectory) as TrafficHistoryProvider.step does, env e starting ``e * env_time_offset``
    seconds later into the recording.  Returns [T, E, M, 8]."""
    rp = _round_param_for_dt(dt)
    con = sqlite3.connect(f"file:{path}?mode=ro", uri=True)
    try:
        q = """SELECT V.id, V.type, V.length, V.width, V.height, T.position_x, T.position_y, T.heading_rad, T.speed
               FROM Vehicle AS V INNER JOIN Trajectory AS T ON V.id = T.vehicle_id
               WHERE T.sim_time > ? AND T.sim_time <= ? ORDER BY T.sim_time DESC"""
        ticks = []
        for t in range(n_ticks):
            per_env = []
            for e in range(n_envs):
                now = round(start_time + e * env_time_offset + t * dt, rp)
                prev = round(now - dt, rp)
                per_env.append(records_from_history_rows(con.execute(q, (prev, now)).fetchall()))
            ticks.append(per_env)
    finally:
        con.close()
    return recording_from_ticks(ticks, M)
