"""Observation records handed to agents -- same type and field names as the reference
(/root/reference/smarts/core/sensors.py:44-203, events.py:23-34, road_map.py:556-618,
coordinates.py:47-86, plan.py) so ``obs.ego_vehicle_state.speed`` etc. keep working."""
from __future__ import annotations

from dataclasses import dataclass
from typing import Any, List, NamedTuple, Optional, Sequence, Tuple

import numpy as np


class Dimensions(NamedTuple):
    length: float
    width: float
    height: float

    @property
    def as_lwh(self):
        return (self.length, self.width, self.height)


class Collision(NamedTuple):
    collidee_id: str


class Events(NamedTuple):
    collisions: Sequence
    off_road: bool
    off_route: bool
    on_shoulder: bool
    wrong_way: bool
    not_moving: bool
    reached_goal: bool
    reached_max_episode_steps: bool
    agents_alive_done: bool


class Waypoint(NamedTuple):
    pos: np.ndarray
    heading: float
    lane_id: str
    lane_width: float
    speed_limit: float
    lane_index: int


class VehicleObservation(NamedTuple):
    id: str
    position: Tuple[float, float, float]
    bounding_box: Dimensions
    heading: float
    speed: float
    road_id: Optional[str]
    lane_id: Optional[str]
    lane_index: Optional[int]


class EgoVehicleObservation(NamedTuple):
    id: str
    position: np.ndarray
    bounding_box: Dimensions
    heading: float
    speed: float
    steering: float
    yaw_rate: float
    road_id: Optional[str]
    lane_id: Optional[str]
    lane_index: Optional[int]
    mission: Any
    linear_velocity: np.ndarray
    angular_velocity: np.ndarray
    linear_acceleration: Optional[np.ndarray]
    angular_acceleration: Optional[np.ndarray]
    linear_jerk: Optional[np.ndarray]
    angular_jerk: Optional[np.ndarray]


class GridMapMetadata(NamedTuple):
    created_at: int
    resolution: float
    width: int
    height: int
    camera_pos: Tuple[float, float, float]
    camera_heading_in_degrees: float


class OccupancyGridMap(NamedTuple):
    metadata: GridMapMetadata
    data: np.ndarray


class DrivableAreaGridMap(NamedTuple):
    metadata: GridMapMetadata
    data: np.ndarray


class ViaPoint(NamedTuple):
    """sensors.py:154-165."""
    position: tuple
    lane_index: float
    road_id: str
    required_speed: float


class Vias(NamedTuple):
    near_via_points: List
    hit_via_points: List


class RoadWaypoints(NamedTuple):
    """sensors.py:104-107: lane id -> waypoint paths along that lane."""
    lanes: dict


@dataclass
class Observation:
    dt: float
    step_count: int
    elapsed_sim_time: float
    events: Events
    ego_vehicle_state: EgoVehicleObservation
    neighborhood_vehicle_states: Optional[List[VehicleObservation]]
    waypoint_paths: Optional[List[List[Waypoint]]]
    distance_travelled: float
    lidar_point_cloud: Optional[Tuple[List[np.ndarray], List[bool], List[Tuple[np.ndarray, np.ndarray]]]]
    drivable_area_grid_map: Optional[DrivableAreaGridMap]
    occupancy_grid_map: Optional[OccupancyGridMap]
    top_down_rgb: Optional[Any]
    road_waypoints: Optional[RoadWaypoints]
    via_data: Vias


@dataclass(frozen=True)
class PositionalGoal:
    """plan.py:95-124"""
    position: Tuple[float, float]
    radius: float

    def is_endless(self) -> bool:
        return False


@dataclass(frozen=True)
class EndlessGoal:
    def is_endless(self) -> bool:
        return True


@dataclass(frozen=True)
class Via:
    """smarts/sstudio/types.py:423-435: a point on a lane the agent should pass at ``required_speed``;
    a non-positive ``hit_distance`` means half the lane width (scenario.py:652-676)."""
    road_id: str
    lane_index: int
    lane_offset: float
    required_speed: float
    hit_distance: float = -1


@dataclass(frozen=True)
class Mission:
    """What sstudio's ``Mission(Route(begin=(road, lane, offset), end=(road, lane, offset)))`` pins
    down for one agent (smarts/sstudio/types.py Route/Mission -> smarts/core/plan.py Mission)."""
    begin_road: str
    begin_lane: int
    begin_offset: float
    end_road: str
    end_lane: int = 0
    end_offset: Optional[float] = None     # None: middle of the lane (plan.py:108-111)
    goal_radius: float = 1.0               # PositionalGoal default (plan.py:100)
    endless: bool = False
    start_speed: float = 0.0
    via: tuple = ()                        # of Via (sstudio types.py:423-435)
