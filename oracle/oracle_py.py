"""ctypes wrapper around ``oracle/_build/liboracle.so`` -- TEST INFRASTRUCTURE ONLY.

Imported by tests/, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline legs; never by
``smarts_b200``.  Mirrors the product's handle API so parity tests read symmetrically.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
import sys
from typing import Dict, Optional

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(_HERE))
from smarts_b200 import _abi  # noqa: E402  (struct layouts only)

_LIB = None


def build(force: bool = False) -> str:
    so = os.path.join(_HERE, "_build", "liboracle.so")
    srcs = [os.path.join(_HERE, f) for f in ("oracle.cpp", "oracle_core.hpp")] + [
        os.path.join(_HERE, "..", "include", f) for f in ("smarts_b200.h", "sb_detmath.h", "sb_lidar_table.h", "sb_vehicle.h")]
    stale = force or not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs)
    if stale:
        subprocess.check_call(["make", "-s", "-C", _HERE], stdout=subprocess.DEVNULL)
    return so


def lib():
    global _LIB
    if _LIB is None:
        L = C.CDLL(build())
        L.or_create.restype = C.c_void_p
        L.or_create.argtypes = [C.POINTER(_abi.SbConfig), C.POINTER(_abi.SbMap), C.POINTER(C.c_double), C.c_int]
        L.or_destroy.argtypes = [C.c_void_p]
        L.or_set_threads.argtypes = [C.c_void_p, C.c_int]
        L.or_reset.argtypes = [C.c_void_p, C.POINTER(C.c_int32), C.c_int, C.POINTER(_abi.SbResetState)]
        L.or_set_social.argtypes = [C.c_void_p, C.POINTER(C.c_double)]
        L.or_step.argtypes = [C.c_void_p, C.c_void_p]
        L.or_observe.argtypes = [C.c_void_p]
        L.or_observe_envs.argtypes = [C.c_void_p, C.POINTER(C.c_int32), C.c_int]
        L.or_get_state.argtypes = [C.c_void_p, C.POINTER(_abi.SbStatePtrs)]
        L.or_get_obs.argtypes = [C.c_void_p, C.POINTER(_abi.SbObsPtrs)]
        L.or_waypoint_paths.restype = C.c_int
        L.or_waypoint_paths.argtypes = [C.c_void_p, C.c_double, C.c_double, C.c_double, C.c_int, C.c_double,
                                        C.POINTER(C.c_int32), C.c_int, C.POINTER(C.c_int32), C.POINTER(C.c_double)]
        L.or_nearest_lane.restype = C.c_int
        L.or_nearest_lane.argtypes = [C.c_void_p, C.c_double, C.c_double, C.c_double, C.POINTER(C.c_double)]
        L.or_lane_probe.argtypes = [C.c_void_p, C.c_int, C.c_double, C.c_double, C.c_double, C.POINTER(C.c_double)]
        L.or_lateral_gains.argtypes = [C.c_double, C.POINTER(C.c_double), C.POINTER(C.c_double)]
        L.or_lateral_gains_v.argtypes = [C.c_int, C.c_double, C.POINTER(C.c_double), C.POINTER(C.c_double)]
        L.or_lane_following.restype = C.c_uint32
        L.or_lane_following.argtypes = [C.c_void_p, C.POINTER(C.c_double), C.POINTER(C.c_int32), C.c_int,
                                        C.c_double, C.c_int, C.c_double, C.POINTER(C.c_double), C.POINTER(C.c_double)]
        L.or_trajectory_tracking.argtypes = [C.POINTER(C.c_double), C.POINTER(C.c_double), C.c_int, C.c_double,
                                             C.POINTER(C.c_double), C.POINTER(C.c_double)]
        L.or_trajectory_tracking_v.argtypes = [C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double), C.c_int, C.c_double,
                                               C.POINTER(C.c_double), C.POINTER(C.c_double)]
        L.or_sincos.argtypes = [C.POINTER(C.c_double), C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double)]
        L.or_atan2.argtypes = [C.POINTER(C.c_double), C.POINTER(C.c_double), C.c_int, C.POINTER(C.c_double)]
        _LIB = L
    return _LIB


def _dp(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def _ip(a):
    return a.ctypes.data_as(C.POINTER(C.c_int32))


def det_sincos(x):
    x = np.ascontiguousarray(x, np.float64)
    s, c = np.empty_like(x), np.empty_like(x)
    lib().or_sincos(_dp(x), x.size, _dp(s), _dp(c))
    return s, c


def det_atan2(y, x):
    y = np.ascontiguousarray(y, np.float64)
    x = np.ascontiguousarray(x, np.float64)
    o = np.empty_like(x)
    lib().or_atan2(_dp(y), _dp(x), x.size, _dp(o))
    return o


def lateral_gains(ts: float, vehicle_type: int = 0):
    a, b = C.c_double(), C.c_double()
    lib().or_lateral_gains_v(vehicle_type, ts, C.byref(a), C.byref(b))
    return a.value, b.value


class OracleWorld:
    """CPU twin of ``smarts_b200.engine.Engine``."""

    def __init__(self, cfg: _abi.SbConfig, map_arrays: Dict[str, np.ndarray], spacing: float,
                 base_rays: Optional[np.ndarray] = None, threads: int = 1):
        self.cfg = cfg
        self._map, self._keep = _abi.make_map_struct(map_arrays, spacing)
        rays = np.ascontiguousarray(base_rays, np.float64) if base_rays is not None else None
        self._h = lib().or_create(C.byref(cfg), C.byref(self._map), _dp(rays) if rays is not None else None,
                                  0 if rays is None else len(rays))
        lib().or_set_threads(self._h, threads)
        self.E, self.A, self.M = cfg.n_envs, cfg.n_agents, cfg.n_social
        self.V = self.A + self.M

    def close(self):
        if self._h:
            lib().or_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def reset(self, x, y, heading, speed, route=None, route_len=None, env_ids=None, goal=None, vias=None,
              keep_social=False):
        x, y, heading, speed = (np.ascontiguousarray(v, np.float64) for v in (x, y, heading, speed))
        rs = _abi.SbResetState()
        rs.keep_social = int(keep_social)
        rs.agent_x, rs.agent_y, rs.agent_heading, rs.agent_speed = _dp(x), _dp(y), _dp(heading), _dp(speed)
        keep = []
        if route is not None:
            route = np.ascontiguousarray(route, np.int32)
            route_len = np.ascontiguousarray(route_len, np.int32)
            keep = [route, route_len]
            rs.route, rs.route_len = _ip(route), _ip(route_len)
        if goal is not None:
            goal = [np.ascontiguousarray(g, np.float64) for g in goal]
            keep += goal
            rs.goal_x, rs.goal_y, rs.goal_radius = (_dp(g) for g in goal)
        if vias is not None:
            vi = [np.ascontiguousarray(v, np.int32) for v in vias[:2]]
            vf = [np.ascontiguousarray(v, np.float64) for v in vias[2:]]
            keep += vi + vf
            rs.via_count, rs.via_lane = _ip(vi[0]), _ip(vi[1])
            rs.via_x, rs.via_y, rs.via_hit_distance, rs.via_required_speed = (_dp(v) for v in vf)
        ids = None
        n = self.E
        if env_ids is not None:
            env_ids = np.ascontiguousarray(env_ids, np.int32)
            ids, n = _ip(env_ids), len(env_ids)
        lib().or_reset(self._h, ids, n, C.byref(rs))
        del keep

    def set_social(self, states):
        s = np.ascontiguousarray(states, np.float64)
        assert s.shape == (self.E, self.M, 8), s.shape
        lib().or_set_social(self._h, _dp(s))

    def step(self, actions):
        a = np.ascontiguousarray(actions)
        lib().or_step(self._h, a.ctypes.data_as(C.c_void_p))

    def observe(self, env_ids=None):
        if env_ids is None:
            lib().or_observe(self._h)
        else:
            ids = np.ascontiguousarray(env_ids, np.int32)
            lib().or_observe_envs(self._h, _ip(ids), len(ids))

    def obs(self) -> Dict[str, np.ndarray]:
        p = _abi.SbObsPtrs()
        lib().or_get_obs(self._h, C.byref(p))
        out = {}
        n = self.E * self.A
        for name, (shape, dt) in _abi.obs_shapes(self.cfg).items():
            ptr = getattr(p, name)
            cnt = n * int(np.prod(shape)) if shape else n
            arr = np.ctypeslib.as_array(C.cast(ptr, C.POINTER(C.c_uint8)), shape=(cnt * np.dtype(dt).itemsize,))
            out[name] = arr.view(dt).reshape((self.E, self.A) + tuple(shape)).copy()
        return out

    def state(self) -> Dict[str, np.ndarray]:
        p = _abi.SbStatePtrs()
        lib().or_get_state(self._h, C.byref(p))
        out = {}
        for name, (per, shape, dt) in _abi.state_shapes(self.cfg).items():
            lead = (self.E, self.V) if per == "v" else (self.E, self.A)
            cnt = int(np.prod(lead)) * (int(np.prod(shape)) if shape else 1)
            ptr = getattr(p, name)
            arr = np.ctypeslib.as_array(C.cast(ptr, C.POINTER(C.c_uint8)), shape=(cnt * np.dtype(dt).itemsize,))
            out[name] = arr.view(dt).reshape(lead + tuple(shape)).copy()
        return out

    # ---- unit probes -------------------------------------------------------------------
    def waypoint_paths(self, x, y, heading, lookahead, route=None, within_radius=5.0):
        """route: None = no route object; [] = empty Route; list of road indices otherwise."""
        lens = np.zeros(_abi.SB_MAX_PATHS, np.int32)
        wp = np.zeros((_abi.SB_MAX_PATHS, _abi.SB_MAX_LOOKAHEAD + 1, 7), np.float64)
        if route is None:
            r, n = None, -1
        else:
            ra = np.ascontiguousarray(route if len(route) else [0], np.int32)
            r, n = _ip(ra), len(route)
        k = lib().or_waypoint_paths(self._h, x, y, heading, lookahead, within_radius, r, n, _ip(lens), _dp(wp))
        return [wp[p, :lens[p]].copy() for p in range(k)]

    def nearest_lane(self, x, y, radius):
        d = C.c_double()
        l = lib().or_nearest_lane(self._h, x, y, radius, C.byref(d))
        return l, d.value

    def lane_probe(self, lane, x, y, off):
        """(offset of (x, y) along `lane`, lane vector at `off`, lane point at `off`, (x, y) on a road?)"""
        out = np.zeros(6)
        lib().or_lane_probe(self._h, int(lane), float(x), float(y), float(off), _dp(out))
        return out[0], out[1:3].copy(), out[3:5].copy(), bool(out[5])

    def lane_following(self, veh7, route, target_speed, lane_change, dt, st):
        veh7 = np.ascontiguousarray(veh7, np.float64)
        st = np.ascontiguousarray(st, np.float64)
        out = np.zeros(3)
        ra = np.ascontiguousarray(route if len(route) else [0], np.int32)
        status = lib().or_lane_following(self._h, _dp(veh7), _ip(ra), len(route), target_speed, lane_change, dt,
                                         _dp(st), _dp(out))
        return out, st, status


def trajectory_tracking(veh7, traj, dt, st, vehicle_type: int = 0):
    veh7 = np.ascontiguousarray(veh7, np.float64)
    traj = np.ascontiguousarray(traj, np.float64)
    st = np.ascontiguousarray(st, np.float64)
    out = np.zeros(3)
    lib().or_trajectory_tracking_v(vehicle_type, _dp(veh7), _dp(traj), traj.shape[1], dt, _dp(st), _dp(out))
    return out, st
