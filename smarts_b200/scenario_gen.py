"""Synthetic scenario inputs: agent start poses / routes and recorded-style social traffic.

SUMO stays out of the kernel path (BASELINE.json north_star): social vehicles are fed per tick
as a ``[E, M, 8]`` tensor, the same record ``ExternalProvider.state_update`` takes
(/root/reference/smarts/core/external_provider.py:44-51 -> VehicleState, vehicle.py:53-80).
This module fabricates such recordings deterministically (fixed seeds):
vehicles ride lane centre lines at constant speed, wrapping at the lane end.
"""
from __future__ import annotations

from typing import List, Optional, Sequence, Tuple

import numpy as np

from .road_map import RoadMapData, vec_to_radians, wrap_heading

# VEHICLE_CONFIGS dimensions (length, width, height), vehicle.py:97-140, drawn with the weights
# (5, 1, 1, 1, 1) of scenarios/minicity/scenario.py:13-27
_DIMS = np.array([[3.68, 1.47, 1.4], [7.0, 2.25, 3.0], [8.0, 2.4, 3.5], [5.0, 1.91, 1.89], [10.0, 2.5, 4.0]])
_WEIGHTS = np.array([5, 1, 1, 1, 1], np.float64) / 9.0


class SocialTraffic:
    """M vehicles per env on lane centre lines; ``states(t)`` gives the [E, M, 8] record of tick t."""

    def __init__(self, rm: RoadMapData, E: int, M: int, seed: int, dt: float = 0.1,
                 lanes: Optional[Sequence[int]] = None, p_alive: float = 1.0):
        rng = np.random.default_rng(seed)
        self.rm, self.E, self.M, self.dt = rm, E, M, dt
        cand = np.asarray(lanes if lanes is not None else
                          [l for l in range(len(rm.lane_width)) if rm.lane_lp_count[l] > 0 and rm.lane_length[l] > 8.0])
        self.lane = cand[rng.integers(0, len(cand), (E, M))]
        # polyline tables per lane
        self._cum, self._xs, self._ys, self._hd = {}, {}, {}, {}
        for l in np.unique(self.lane):
            s, e = int(rm.lane_shape_start[l]), int(rm.lane_shape_start[l + 1])
            xs, ys = rm.shape_x[s:e], rm.shape_y[s:e]
            seg = np.hypot(np.diff(xs), np.diff(ys))
            self._cum[l] = np.concatenate([[0.0], np.cumsum(seg)])
            self._xs[l], self._ys[l] = xs, ys
            self._hd[l] = np.array([wrap_heading(vec_to_radians(dx, dy)) for dx, dy in zip(np.diff(xs), np.diff(ys))])
        self.length = np.vectorize(lambda l: self._cum[l][-1])(self.lane)
        self.s0 = rng.uniform(0, 1, (E, M)) * self.length
        self.u = rng.uniform(5.0, 15.0, (E, M))
        kind = rng.choice(len(_DIMS), size=(E, M), p=_WEIGHTS)
        self.dims = _DIMS[kind]
        self.alive = (rng.uniform(size=(E, M)) < p_alive).astype(np.float64)

    def states(self, t: int) -> np.ndarray:
        E, M = self.E, self.M
        out = np.zeros((E, M, 8), np.float64)
        s = np.mod(self.s0 + self.u * (t * self.dt), self.length)
        for l in np.unique(self.lane):
            msk = self.lane == l
            sv = s[msk]
            cum = self._cum[l]
            out[msk, 0] = np.interp(sv, cum, self._xs[l])
            out[msk, 1] = np.interp(sv, cum, self._ys[l])
            seg = np.clip(np.searchsorted(cum, sv, side="right") - 1, 0, len(cum) - 2)
            out[msk, 2] = self._hd[l][seg]
        out[..., 3] = self.u
        out[..., 4:7] = self.dims
        out[..., 7] = self.alive
        return out

    def recording(self, T: int) -> np.ndarray:
        """[T, E, M, 8] pre-generated recording."""
        return np.stack([self.states(t) for t in range(T)])


def agent_starts(rm: RoadMapData, E: int, missions: Sequence[Tuple[str, int, float, str]], route_cap: int,
                 seed: int, jitter: float = 0.0, speed: float = 0.0):
    """Start poses and routes for A = len(missions) agents per env.

    missions: (begin road, lane index, offset m, end road) like sstudio ``Route(begin=..., end=...)``
    (e.g. scenarios/intersections/4lane/scenario.py:20-23).  ``jitter`` shifts the start offset
    per env (uniform, metres) so envs are not identical."""
    rng = np.random.default_rng(seed)
    A = len(missions)
    x = np.zeros((E, A)); y = np.zeros((E, A)); h = np.zeros((E, A)); v = np.full((E, A), float(speed))
    route = np.full((E, A, route_cap), -1, np.int32)
    rlen = np.zeros((E, A), np.int32)
    for a, (road, lane_idx, off, end) in enumerate(missions):
        r = rm.generate_route(road, end)
        assert 0 < len(r.road_indices) <= route_cap, (road, end, len(r.road_indices))
        route[:, a, :len(r.road_indices)] = r.road_indices
        rlen[:, a] = len(r.road_indices)
        offs = off + (rng.uniform(0, jitter, E) if jitter > 0 else np.zeros(E))
        for e in range(E):
            x[e, a], y[e, a], h[e, a] = rm.lane_start_pose(road, lane_idx, float(offs[e]))
    return x, y, h, v, route, rlen


def random_missions(rm: RoadMapData, A: int, seed: int, route_cap: int, min_len: float = 30.0,
                    tries: int = 4000) -> List[Tuple[str, int, float, str]]:
    """A begin/end pairs joined by a route of at most ``route_cap`` roads, on lanes long enough to start
    10 m in -- stands in for the random missions of large scenarios (scenarios/minicity/scenario.py)."""
    rng = np.random.default_rng(seed)
    normal = [r for r in range(len(rm.road_ids)) if not rm.road_is_junction[r]
              and rm.lane_length[int(rm.road_lane_start[r])] >= min_len and rm.road_lp_count[r] > 0]
    out: List[Tuple[str, int, float, str]] = []
    for _ in range(tries):
        if len(out) == A:
            break
        b, e = (int(t) for t in rng.choice(normal, 2, replace=False))
        r = rm.generate_route(rm.road_ids[b], rm.road_ids[e])
        if not 1 < len(r.road_indices) <= route_cap:
            continue
        lane = int(rng.integers(0, rm.road_lane_count[b]))
        out.append((rm.road_ids[b], lane, 10.0, rm.road_ids[e]))
    if len(out) < A:
        raise RuntimeError(f"only {len(out)} of {A} missions found")
    return out
