"""AgentInterface / AgentType / ActionSpaceType / AgentSpec -- the agent-facing configuration
records of the reference, kept name-for-name so a hiway-v0 user changes only the import.

Mirrors (field names, defaults, presets; no code shared):
  /root/reference/smarts/core/agent_interface.py:29-420   sensor configs, DoneCriteria, AgentInterface, AgentType
  /root/reference/smarts/core/controllers/__init__.py:42-54   ActionSpaceType
  /root/reference/smarts/zoo/agent_spec.py:36-119, smarts/core/agent.py:31-58   AgentSpec, Agent

``compile_config`` turns one AgentInterface into the ``SbConfig`` the C-ABI takes
(include/smarts_b200.h); sensors the hot path does not cover (RGB camera, the
MPC / TrajectoryWithTime / MultiTargetPose / Imitation action spaces) raise ``NotImplementedError``
instead of being silently dropped.
"""
from __future__ import annotations

import dataclasses
import inspect
from dataclasses import dataclass, field
from enum import Enum, IntEnum
from typing import Any, Callable, List, Optional, Union

from . import _abi
from .lidar_params import BasicLidar


class ActionSpaceType(Enum):
    Continuous = 0
    Lane = 1
    ActuatorDynamic = 2
    LaneWithContinuousSpeed = 3
    TargetPose = 4
    Trajectory = 5
    MultiTargetPose = 6
    MPC = 7
    TrajectoryWithTime = 8
    Imitation = 9


_SUPPORTED_ACTIONS = {
    ActionSpaceType.Continuous: _abi.SB_ACTION_CONTINUOUS,
    ActionSpaceType.Lane: _abi.SB_ACTION_LANE,
    ActionSpaceType.ActuatorDynamic: _abi.SB_ACTION_ACTUATOR_DYNAMIC,
    ActionSpaceType.LaneWithContinuousSpeed: _abi.SB_ACTION_LANE_WITH_CONTINUOUS_SPEED,
    ActionSpaceType.Trajectory: _abi.SB_ACTION_TRAJECTORY,
}


@dataclass
class DrivableAreaGridMap:
    width: int = 256
    height: int = 256
    resolution: float = 50 / 256


@dataclass
class OGM:
    width: int = 256
    height: int = 256
    resolution: float = 50 / 256


@dataclass
class RGB:
    width: int = 256
    height: int = 256
    resolution: float = 50 / 256


@dataclass
class Lidar:
    sensor_params: Any = None

    def __post_init__(self):
        if self.sensor_params is None:
            self.sensor_params = BasicLidar


@dataclass
class Waypoints:
    lookahead: int = 32


@dataclass
class RoadWaypoints:
    horizon: int = 20


@dataclass
class NeighborhoodVehicles:
    radius: Optional[float] = None


@dataclass
class Accelerometer:
    pass


class AgentType(IntEnum):
    Buddha = 0
    Full = 1
    Standard = 2
    Laner = 3
    Loner = 4
    Tagger = 5
    StandardWithAbsoluteSteering = 6
    LanerWithSpeed = 7
    Tracker = 8
    Boid = 9
    MPCTracker = 10
    TrajectoryInterpolator = 11
    Imitation = 12


@dataclass(frozen=True)
class AgentsListAlive:
    agents_list: List[str]
    minimum_agents_alive_in_list: int


@dataclass(frozen=True)
class AgentsAliveDoneCriteria:
    minimum_ego_agents_alive: Optional[int] = None
    minimum_total_agents_alive: Optional[int] = None
    agent_lists_alive: Optional[List[AgentsListAlive]] = None


@dataclass(frozen=True)
class EventConfiguration:
    not_moving_time: float = 60
    not_moving_distance: float = 1


@dataclass(frozen=True)
class DoneCriteria:
    collision: bool = True
    off_road: bool = True
    off_route: bool = True
    on_shoulder: bool = False
    wrong_way: bool = False
    not_moving: bool = False
    agents_alive: Optional[AgentsAliveDoneCriteria] = None

    def mask(self) -> int:
        """SB_EV_* bits that end an episode (include/smarts_b200.h)."""
        m = 0
        for flag, bit in ((self.collision, _abi.SB_EV_COLLISION), (self.off_road, _abi.SB_EV_OFF_ROAD),
                          (self.off_route, _abi.SB_EV_OFF_ROUTE), (self.on_shoulder, _abi.SB_EV_ON_SHOULDER),
                          (self.wrong_way, _abi.SB_EV_WRONG_WAY), (self.not_moving, _abi.SB_EV_NOT_MOVING)):
            if flag:
                m |= bit
        return m


def _resolve(config, type_):
    if config is True:
        return type_()
    if isinstance(config, type_):
        return config
    return False


@dataclass
class AgentInterface:
    debug: bool = False
    event_configuration: EventConfiguration = field(default_factory=EventConfiguration)
    done_criteria: DoneCriteria = field(default_factory=DoneCriteria)
    max_episode_steps: Optional[int] = None
    neighborhood_vehicles: Union[NeighborhoodVehicles, bool] = False
    waypoints: Union[Waypoints, bool] = False
    road_waypoints: Union[RoadWaypoints, bool] = False
    drivable_area_grid_map: Union[DrivableAreaGridMap, bool] = False
    ogm: Union[OGM, bool] = False
    rgb: Union[RGB, bool] = False
    lidar: Union[Lidar, bool] = False
    action: Optional[ActionSpaceType] = None
    vehicle_type: str = "sedan"
    accelerometer: Union[Accelerometer, bool] = True

    def __post_init__(self):
        self.neighborhood_vehicles = _resolve(self.neighborhood_vehicles, NeighborhoodVehicles)
        self.waypoints = _resolve(self.waypoints, Waypoints)
        self.road_waypoints = _resolve(self.road_waypoints, RoadWaypoints)
        self.drivable_area_grid_map = _resolve(self.drivable_area_grid_map, DrivableAreaGridMap)
        self.ogm = _resolve(self.ogm, OGM)
        self.rgb = _resolve(self.rgb, RGB)
        self.lidar = _resolve(self.lidar, Lidar)
        self.accelerometer = _resolve(self.accelerometer, Accelerometer)
        assert self.vehicle_type in {"sedan", "bus"}

    _PRESETS = {
        AgentType.Buddha: dict(),
        AgentType.Full: dict(neighborhood_vehicles=True, waypoints=True, drivable_area_grid_map=True, ogm=True,
                             rgb=True, lidar=True, action=ActionSpaceType.Continuous),
        AgentType.StandardWithAbsoluteSteering: dict(waypoints=True, neighborhood_vehicles=True,
                                                     action=ActionSpaceType.Continuous),
        AgentType.Standard: dict(waypoints=True, neighborhood_vehicles=True, action=ActionSpaceType.ActuatorDynamic),
        AgentType.Laner: dict(waypoints=True, action=ActionSpaceType.Lane),
        AgentType.LanerWithSpeed: dict(waypoints=True, action=ActionSpaceType.LaneWithContinuousSpeed),
        AgentType.Tracker: dict(waypoints=True, action=ActionSpaceType.Trajectory),
        AgentType.TrajectoryInterpolator: dict(action=ActionSpaceType.TrajectoryWithTime),
        AgentType.MPCTracker: dict(waypoints=True, action=ActionSpaceType.MPC),
        AgentType.Boid: dict(waypoints=True, neighborhood_vehicles=True, action=ActionSpaceType.MultiTargetPose),
        AgentType.Loner: dict(waypoints=True, action=ActionSpaceType.Continuous),
        AgentType.Tagger: dict(waypoints=True, neighborhood_vehicles=True, action=ActionSpaceType.Continuous),
        AgentType.Imitation: dict(neighborhood_vehicles=True, action=ActionSpaceType.Imitation),
    }

    @staticmethod
    def from_type(requested_type: AgentType, **kwargs) -> "AgentInterface":
        try:
            preset = AgentInterface._PRESETS[AgentType(requested_type)]
        except (KeyError, ValueError):
            raise Exception("Unsupported agent type %s" % requested_type)
        return AgentInterface(**preset).replace(**kwargs)

    def replace(self, **kwargs) -> "AgentInterface":
        return dataclasses.replace(self, **kwargs)

    @property
    def action_space(self):
        return self.action


def compile_config(itf: AgentInterface, n_envs: int, n_agents: int, n_social: int, dt: float, *, device: int = 0,
                   route_cap: int = 16, traj_len: int = 10, nbr_cap: int = 10, wp_out_paths: int = 4,
                   wp_out_len: Optional[int] = None, lidar_wall_height: float = 0.0, lidar_miss_inf: int = 1,
                   events: bool = True, via_cap: int = 0, rw_lanes_cap: int = 32,
                   rw_paths_cap: int = 64) -> _abi.SbConfig:
    """AgentInterface (+ HiWayEnv kwargs) -> SbConfig.  ``wp_out_len`` None keeps the sensor's full
    lookahead (what ``Observation.waypoint_paths`` holds); StdObs uses 4 x 20 (format_obs.py:41)."""
    if itf.action is None:
        raise ValueError("AgentInterface.action must be set (the Buddha preset has no controller)")
    if itf.action not in _SUPPORTED_ACTIONS:
        raise NotImplementedError(f"action space {itf.action} is outside the B200 hot path "
                                  f"(supported: {[a.name for a in _SUPPORTED_ACTIONS]})")
    if itf.rgb:
        raise NotImplementedError("RGB camera needs the Panda3D renderer: unavailable on this box (DESIGN.md 9)")
    c = _abi.SbConfig()
    c.abi_version = _abi.SB_ABI_VERSION
    c.device = device
    c.n_envs, c.n_agents, c.n_social = n_envs, n_agents, n_social
    c.dt = dt
    c.vehicle_type = {"sedan": _abi.SB_VEHICLE_SEDAN, "bus": _abi.SB_VEHICLE_BUS}[itf.vehicle_type]  # vehicle.py:393-398
    c.action_space = _SUPPORTED_ACTIONS[itf.action]
    c.traj_len = traj_len
    c.route_cap = route_cap
    if itf.neighborhood_vehicles:
        r = itf.neighborhood_vehicles.radius
        c.nbr_enable, c.nbr_cap, c.nbr_radius = 1, nbr_cap, (-1.0 if r is None else float(r))
    if itf.waypoints:
        la = int(itf.waypoints.lookahead)
        if la > _abi.SB_MAX_LOOKAHEAD:
            raise ValueError(f"Waypoints.lookahead {la} > {_abi.SB_MAX_LOOKAHEAD}")
        c.wp_enable, c.wp_lookahead, c.wp_out_paths = 1, la, wp_out_paths
        c.wp_out_len = la + 1 if wp_out_len is None else wp_out_len  # a path holds lookahead + 1 waypoints
    if itf.lidar:
        p = itf.lidar.sensor_params
        n_az = int((p.end_angle - p.start_angle) / p.angle_resolution)  # lidar.py:93-96
        c.lidar_enable, c.lidar_rays = 1, n_az * len(p.laser_angles)
        c.lidar_miss_inf, c.lidar_wall_height = lidar_miss_inf, lidar_wall_height
    if itf.ogm:
        c.ogm_enable, c.ogm_width, c.ogm_height = 1, itf.ogm.width, itf.ogm.height
        c.ogm_resolution = itf.ogm.resolution
    if itf.drivable_area_grid_map:
        g = itf.drivable_area_grid_map
        c.dagm_enable, c.dagm_width, c.dagm_height, c.dagm_resolution = 1, g.width, g.height, g.resolution
    if itf.waypoints and itf.neighborhood_vehicles:  # StdObs `ttc` (format_obs.py:387-388)
        c.ttc_enable = 1
    if itf.accelerometer:
        c.accel_enable = 1
    if itf.road_waypoints:  # sensors.py:991-1043; a path holds 2 * horizon + 1 waypoints
        hz = int(itf.road_waypoints.horizon)
        if not 1 <= hz <= _abi.SB_MAX_LOOKAHEAD // 2:
            raise ValueError(f"RoadWaypoints.horizon {hz} outside 1 .. {_abi.SB_MAX_LOOKAHEAD // 2}")
        c.rw_enable, c.rw_horizon, c.rw_lanes_cap, c.rw_paths_cap = 1, hz, int(rw_lanes_cap), int(rw_paths_cap)
    if via_cap > 0:  # ViaSensor is always attached (vehicle.py:552-558); it only does work for missions with vias
        c.via_cap, c.via_acquisition_range, c.via_speed_accuracy = int(via_cap), 40.0, 1.5
    if events:
        c.events_enable = 1
        c.done_mask = itf.done_criteria.mask()
        c.max_episode_steps = int(itf.max_episode_steps or 0)
        c.not_moving_time = float(itf.event_configuration.not_moving_time)
        c.not_moving_distance = float(itf.event_configuration.not_moving_distance)
    return c


class Agent:
    """smarts/core/agent.py:31-58: the policy side; ``act(obs)`` returns an action."""

    @classmethod
    def from_function(cls, agent_function: Callable):
        assert callable(agent_function)

        class FunctionAgent(Agent):
            def act(self, obs):
                return agent_function(obs)

        return FunctionAgent()

    def act(self, obs):
        raise NotImplementedError


def _identity(x):
    return x


@dataclass
class AgentSpec:
    """smarts/zoo/agent_spec.py:36-119."""
    interface: Optional[AgentInterface] = None
    agent_builder: Optional[Callable[..., Agent]] = None
    agent_params: Optional[Any] = None
    observation_adapter: Callable = _identity
    action_adapter: Callable = _identity
    reward_adapter: Callable = lambda obs, reward: reward
    info_adapter: Callable = lambda obs, reward, info: info

    def replace(self, **kwargs) -> "AgentSpec":
        return dataclasses.replace(self, **kwargs)

    def build_agent(self) -> Agent:
        if self.agent_builder is None:
            raise ValueError("Can't build agent, no agent builder was supplied")
        if not callable(self.agent_builder):
            raise ValueError(f"agent_builder: {self.agent_builder} is not callable")
        p = self.agent_params
        if p is None:
            return self.agent_builder()
        if isinstance(p, (list, tuple)):
            return self.agent_builder(*p)
        if isinstance(p, dict):
            spec = inspect.getfullargspec(self.agent_builder)
            if spec.varkw is not None:
                return self.agent_builder(**p)
            return self.agent_builder(**{k: p[k] for k in p.keys() & set(spec.args)})
        return self.agent_builder(p)
