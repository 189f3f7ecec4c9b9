"""Host mirror of smarts/core/utils/adapters/ego_centric_adapters.py: the observation adapter that moves an
``Observation`` into the ego frame (origin at the ego, +x along its heading) and the action adapters that move
ego-frame trajectories / target poses back to world space.

Two ways to get ego-centric observations:

* per agent, as in the reference: ``AgentSpec(observation_adapter=obs_adapter, action_adapter=act_adapter)`` with
  ``obs_adapter, act_adapter = get_egocentric_adapters(action_space)`` -- plain numpy on the few hundred values of
  one ``Observation`` object;
* for the whole batch on the device: ``BatchedHiWayEnv(..., ego_centric=True)`` sets ``SbConfig.ego_frame`` and
  ``k_ego_frame`` (csrc/sb_extras.cuh) rewrites the neighbour / waypoint / lidar buffers in place, so
  ``std_obs()`` / ``step_tensor()`` hand out ego-frame tensors without a host round trip.
"""
from __future__ import annotations

import math
from dataclasses import is_dataclass
from dataclasses import replace as dc_replace
from typing import Any, Callable, Optional

import numpy as np

from .agent_interface import ActionSpaceType
from .observation import Observation


def wrap_value(value: float, lo: float, hi: float) -> float:
    """Fold ``value`` into (lo, hi] (utils/math.py:452-461; Python's float ``%``, so lo itself maps to hi)."""
    span = hi - lo
    if value <= lo:
        return hi - (lo - value) % span
    if value > hi:
        return lo + (value - hi) % span
    return value


def _ego_frame_matrix(ego_heading: float) -> np.ndarray:
    """utils/math.py:464-470: rotation by -heading about z."""
    m = np.eye(3)
    m[0, 0] = math.cos(-ego_heading)
    m[0, 1] = -math.sin(-ego_heading)
    m[1, 0] = math.sin(-ego_heading)
    m[1, 1] = math.cos(-ego_heading)
    return m


def position_to_ego_frame(position, ego_position, ego_heading):
    """utils/math.py:473-487."""
    rel = np.asarray(position, np.float64) - np.asarray(ego_position, np.float64)
    return (_ego_frame_matrix(ego_heading) @ rel).tolist()


def world_position_from_ego_frame(position, ego_world_position, ego_world_heading):
    """utils/math.py:490-505."""
    rot = np.linalg.inv(_ego_frame_matrix(ego_world_heading)) @ np.asarray(position, np.float64)
    return (rot + np.asarray(ego_world_position, np.float64)).tolist()


def _replace(obj: Any, **kwargs):
    if is_dataclass(obj):
        return dc_replace(obj, **kwargs)
    if isinstance(obj, tuple) and hasattr(obj, "_replace"):
        return obj._replace(**kwargs)
    raise ValueError("Must be a namedtuple or dataclass.")


def ego_centric_observation_adapter(obs: Observation, *args: Any, **kwargs: Any) -> Observation:
    """ego_centric_adapters.py:62-178.  Members this framework never fills (top_down_rgb) pass
    through as ``None``; the reference would raise on a missing camera, this mirror skips it.  The ego's
    ``mission`` is passed through unchanged: this framework's ``Mission`` record holds road / lane / offset,
    not world poses, so there is nothing to rotate (the reference rotates mission.start / via / goal)."""
    ego = obs.ego_vehicle_state
    position, heading = np.asarray(ego.position, np.float64), float(ego.heading)

    def ego_frame_dynamics(v):
        return None if v is None else np.array([np.linalg.norm(v[:2]), 0, *v[2:]])  # point to X

    def transform(v):
        return position_to_ego_frame(v, position, heading)

    def adjust_heading(h):
        return wrap_value(float(h) - heading, -math.pi, math.pi)

    def replace_via(vp):
        p = transform((*vp.position[:2], 0.0))
        return _replace(vp, position=tuple(p[:2]))

    vd = None
    if obs.via_data:
        vd = _replace(obs.via_data, near_via_points=[replace_via(v) for v in obs.via_data.near_via_points],
                      hit_via_points=[replace_via(v) for v in obs.via_data.hit_via_points])

    def replace_wps(paths):
        return [[_replace(wp, pos=np.asarray(transform(np.append(wp.pos, [0]))[:2]), heading=adjust_heading(wp.heading))
                 for wp in path] for path in paths]

    def replace_metadata(cam):
        if cam is None:
            return None
        return _replace(cam, metadata=_replace(cam.metadata, camera_pos=(0, 0, 0), camera_heading_in_degrees=0))

    def replace_lidar(lidar):
        if lidar is None or len(lidar) == 0:
            return lidar
        with np.errstate(invalid="ignore"):  # misses are (inf, inf, inf)
            return ([np.asarray(transform(p)) for p in lidar[0]], lidar[1],
                    [(np.asarray(transform(a)), np.asarray(transform(b))) for a, b in lidar[2]])

    nvs = obs.neighborhood_vehicle_states
    return _replace(
        obs,
        ego_vehicle_state=_replace(
            ego, position=np.array([0, 0, 0]), heading=0.0,
            linear_velocity=ego_frame_dynamics(ego.linear_velocity),
            linear_acceleration=ego_frame_dynamics(ego.linear_acceleration),
            linear_jerk=ego_frame_dynamics(ego.linear_jerk)),
        neighborhood_vehicle_states=None if nvs is None else [
            _replace(nv, position=np.asarray(transform(nv.position)), heading=adjust_heading(nv.heading)) for nv in nvs],
        lidar_point_cloud=replace_lidar(obs.lidar_point_cloud),
        waypoint_paths=None if obs.waypoint_paths is None else replace_wps(obs.waypoint_paths),
        drivable_area_grid_map=replace_metadata(obs.drivable_area_grid_map),
        occupancy_grid_map=replace_metadata(obs.occupancy_grid_map),
        top_down_rgb=replace_metadata(obs.top_down_rgb),
        road_waypoints=None if not obs.road_waypoints else _replace(
            obs.road_waypoints, lanes={l_id: replace_wps(wps) for l_id, wps in obs.road_waypoints.lanes.items()}),
        via_data=vd,
    )


def _identity(act, _=None):
    return act


def _trajectory_adaption(act, last_obs: Observation):
    """ego_centric_adapters.py:195-213: ego-frame (x, y, heading, ...) sequences -> world frame."""
    ego = last_obs.ego_vehicle_state
    new_pos = np.array([world_position_from_ego_frame([x, y, 0], ego.position, ego.heading)[:2]
                        for x, y in zip(*act[:2])]).T
    new_headings = np.array([wrap_value(h + float(ego.heading), -math.pi, math.pi) for h in act[2]])
    return (*new_pos, new_headings, *act[3:])


def _trajectory_adapter(act, last_obs: Optional[Observation] = None):
    return _trajectory_adaption(act, last_obs) if last_obs else act


def _target_pose_adapter(act, last_obs: Optional[Observation] = None):
    if not last_obs:
        return act
    ego = last_obs.ego_vehicle_state
    out = world_position_from_ego_frame(np.append(act[:2], [0]), ego.position, ego.heading)
    return np.array([*out[:2], wrap_value(float(ego.heading) + act[2], -math.pi, math.pi), act[3]])


class _AdapterPair:
    """ego_centric_adapters.py:280-299: the action adapter converts with the last observation the observation
    adapter saw *before* it was moved into the ego frame."""

    def __init__(self, to_ego: Callable, to_world: Callable):
        self._to_ego, self._to_world, self._seen = to_ego, to_world, None

    def observe(self, obs: Observation):
        self._seen = obs
        return self._to_ego(obs)

    def act(self, action: Any):
        return self._to_world(action, self._seen)


def get_egocentric_adapters(action_space: ActionSpaceType):
    """ego_centric_adapters.py:302-325 -> (observation adapter, action adapter)."""
    m = {ActionSpaceType.Trajectory: _trajectory_adapter, ActionSpaceType.TrajectoryWithTime: _trajectory_adapter,
         ActionSpaceType.MPC: _trajectory_adapter, ActionSpaceType.TargetPose: _target_pose_adapter}
    pair = _AdapterPair(ego_centric_observation_adapter, m.get(action_space, _identity))
    return pair.observe, pair.act
