"""ctypes mirror of ``include/smarts_b200.h`` (struct layouts and constants only)."""
from __future__ import annotations

import ctypes as C
from typing import Dict

import numpy as np

SB_ABI_VERSION = 7
SB_OK, SB_EINVAL, SB_ECUDA, SB_ESTATE, SB_ENOMEM = 0, -1, -2, -3, -4
SB_NOALLOC_BITS_MAX_BYTES = 16 << 20  # static device staging of sb_copy_grids_to_host_packed_noalloc

SB_ACTION_CONTINUOUS = 0
SB_ACTION_LANE = 1
SB_ACTION_ACTUATOR_DYNAMIC = 2
SB_ACTION_LANE_WITH_CONTINUOUS_SPEED = 3
SB_ACTION_TRAJECTORY = 5

SB_LANE_KEEP, SB_LANE_SLOW_DOWN, SB_LANE_CHANGE_LEFT, SB_LANE_CHANGE_RIGHT = 0, 1, 2, 3

SB_MAX_PATHS = 16
SB_MAX_LOOKAHEAD = 64
SB_MAX_TRAJ = 32
SB_CTRL_STATE = 12
SB_STATUS_PATH_OVERFLOW = 1
SB_STATUS_NO_WAYPOINTS = 2
SB_STATUS_LIDAR_OVERFLOW = 4
SB_STATUS_TTC_TRUNCATED = 8
SB_STATUS_RW_TRUNCATED = 16
SB_VEHICLE_SEDAN, SB_VEHICLE_BUS = 0, 1

SB_EV_COLLISION, SB_EV_OFF_ROAD, SB_EV_OFF_ROUTE, SB_EV_ON_SHOULDER = 1, 2, 4, 8
SB_EV_WRONG_WAY, SB_EV_NOT_MOVING, SB_EV_REACHED_GOAL, SB_EV_REACHED_MAX_STEPS = 16, 32, 64, 128
SB_EV_AGENTS_ALIVE_DONE = 256
SB_DRIVEN_PATH_MAX = 500

_i32, _f64, _u8 = C.c_int32, C.c_double, C.c_uint8
_pi32, _pf64, _pu8 = C.POINTER(C.c_int32), C.POINTER(C.c_double), C.POINTER(C.c_uint8)
_pf32, _pi8, _pu32 = C.POINTER(C.c_float), C.POINTER(C.c_int8), C.POINTER(C.c_uint32)


class SbConfig(C.Structure):
    _fields_ = [
        ("abi_version", _i32), ("device", _i32), ("n_envs", _i32), ("n_agents", _i32), ("n_social", _i32),
        ("dt", _f64), ("substeps", _i32), ("action_space", _i32), ("traj_len", _i32), ("route_cap", _i32),
        ("nbr_enable", _i32), ("nbr_cap", _i32), ("nbr_radius", _f64),
        ("wp_enable", _i32), ("wp_lookahead", _i32), ("wp_out_paths", _i32), ("wp_out_len", _i32),
        ("lidar_enable", _i32), ("lidar_rays", _i32), ("lidar_miss_inf", _i32), ("lidar_wall_height", _f64),
        ("ogm_enable", _i32), ("ogm_width", _i32), ("ogm_height", _i32), ("ogm_resolution", _f64),
        ("dagm_enable", _i32), ("dagm_width", _i32), ("dagm_height", _i32), ("dagm_resolution", _f64),
        ("events_enable", _i32), ("done_mask", C.c_uint32), ("max_episode_steps", _i32),
        ("not_moving_time", _f64), ("not_moving_distance", _f64),
        ("ttc_enable", _i32),
        ("accel_enable", _i32), ("via_cap", _i32), ("via_acquisition_range", _f64), ("via_speed_accuracy", _f64),
        ("ego_frame", _i32),
        ("rw_enable", _i32), ("rw_horizon", _i32), ("rw_lanes_cap", _i32), ("rw_paths_cap", _i32),
        ("vehicle_type", _i32),
    ]


_MAP_PTRS = [
    ("road_lane_start", _pi32), ("road_lane_count", _pi32), ("road_is_junction", _pu8),
    ("road_lp_start", _pi32), ("road_lp_count", _pi32), ("road_first_out", _pi32),
    ("lane_road", _pi32), ("lane_idx_in_road", _pi32),
    ("lane_width", _pf64), ("lane_speed", _pf64), ("lane_length", _pf64),
    ("lane_out_start", _pi32), ("lane_out", _pi32), ("lane_lp_start", _pi32), ("lane_lp_count", _pi32),
    ("lane_shape_start", _pi32),
    ("lp_x", _pf64), ("lp_y", _pf64), ("lp_heading", _pf64), ("lp_inferred", _pu8), ("lp_lane", _pi32),
    ("lp_next_start", _pi32), ("lp_next", _pi32),
    ("shape_x", _pf64), ("shape_y", _pf64),
    ("poly_start", _pi32), ("poly_x", _pf64), ("poly_y", _pf64), ("road_edges", _pf64),
    ("lane_in_start", _pi32), ("lane_in", _pi32), ("road_par_start", _pi32), ("road_par", _pi32),
]


class SbMap(C.Structure):
    _fields_ = [
        ("n_roads", _i32), ("n_lanes", _i32), ("n_lanepoints", _i32), ("n_shape_pts", _i32),
        ("n_polys", _i32), ("n_poly_verts", _i32), ("n_edges", _i32), ("n_lane_out", _i32), ("n_lp_next", _i32),
        ("n_lane_in", _i32), ("n_road_par", _i32),
        ("lanepoint_spacing", _f64),
    ] + _MAP_PTRS


class SbStatePtrs(C.Structure):
    _fields_ = [(n, _pf64) for n in
                ("pos_x", "pos_y", "pos_z", "heading", "speed", "yaw_rate", "length", "width", "height",
                 "cos_h", "sin_h")] + [("alive", _pu8)] + [(n, _pf64) for n in
                ("steering", "v_long", "v_lat", "ctrl", "last_control")] + [
                ("route", _pi32), ("route_len", _pi32), ("status", _pu32)]


class SbObsPtrs(C.Structure):
    _fields_ = [
        ("ego_lane", _pi32),
        ("nbr_count", _pi32), ("nbr_id", _pi32), ("nbr_pos", _pf64), ("nbr_box", _pf32),
        ("nbr_heading", _pf32), ("nbr_speed", _pf32), ("nbr_lane", _pi32), ("nbr_lane_index", _pi8),
        ("wp_n_paths", _pi32), ("wp_len", _pi32), ("wp_pos", _pf64), ("wp_heading", _pf32),
        ("wp_lane_width", _pf32), ("wp_speed_limit", _pf32), ("wp_lane", _pi32), ("wp_lane_index", _pi8),
        ("lidar_points", _pf64), ("lidar_hits", _pi8),
        ("ogm", _pu8), ("dagm", _pu8),
        ("ev_events", _pu32), ("ev_done", _pu8), ("ev_n_collisions", _pi32), ("ev_collidee", _pi32),
        ("ev_distance", _pf64), ("ev_reward", _pf64), ("ev_steps", _pi32),
        ("ttc_angle_error", _pf32), ("ttc_distance_from_center", _pf32), ("ttc_dtc", _pf32), ("ttc_ttc", _pf32),
        ("ttc_valid", _pu8),
        ("acc_linear", _pf64), ("acc_angular", _pf64), ("acc_linear_jerk", _pf64), ("acc_angular_jerk", _pf64),
        ("via_near_count", _pi32), ("via_near_idx", _pi32), ("via_hit_count", _pi32), ("via_hit_idx", _pi32),
        ("rw_n_lanes", _pi32), ("rw_lane", _pi32), ("rw_n_paths", _pi32), ("rw_path_slot", _pi32), ("rw_len", _pi32),
        ("rw_pos", _pf64), ("rw_heading", _pf64), ("rw_lane_width", _pf32), ("rw_speed_limit", _pf32),
        ("rw_wp_lane", _pi32), ("rw_lane_index", _pi8),
    ]


class SbResetState(C.Structure):
    _fields_ = [("agent_x", _pf64), ("agent_y", _pf64), ("agent_heading", _pf64), ("agent_speed", _pf64),
                ("route", _pi32), ("route_len", _pi32),
                ("goal_x", _pf64), ("goal_y", _pf64), ("goal_radius", _pf64),
                ("via_count", _pi32), ("via_lane", _pi32),
                ("via_x", _pf64), ("via_y", _pf64), ("via_hit_distance", _pf64), ("via_required_speed", _pf64),
                ("keep_social", C.c_int32)]


_NP = {_pi32: np.int32, _pf64: np.float64, _pu8: np.uint8, _pf32: np.float32, _pi8: np.int8, _pu32: np.uint32}


def make_map_struct(arrays: Dict[str, np.ndarray], spacing: float):
    """Fill an SbMap from ``RoadMapData.arrays()``.  Returns (struct, keepalive list)."""
    m = SbMap()
    keep = []
    for name, ptype in _MAP_PTRS:
        a = np.ascontiguousarray(arrays[name], dtype=_NP[ptype])
        keep.append(a)
        setattr(m, name, a.ctypes.data_as(ptype))
    m.n_roads = len(arrays["road_lane_start"])
    m.n_lanes = len(arrays["lane_road"])
    m.n_lanepoints = len(arrays["lp_x"])
    m.n_shape_pts = len(arrays["shape_x"])
    m.n_polys = len(arrays["poly_start"]) - 1
    m.n_poly_verts = len(arrays["poly_x"])
    m.n_edges = len(arrays["road_edges"])
    m.n_lane_out = int(arrays["lane_out_start"][-1])
    m.n_lp_next = int(arrays["lp_next_start"][-1])
    m.n_lane_in = int(arrays["lane_in_start"][-1])
    m.n_road_par = int(arrays["road_par_start"][-1])
    m.lanepoint_spacing = float(spacing)
    return m, keep


def obs_shapes(cfg: SbConfig) -> Dict[str, tuple]:
    """Shapes (without the leading E*A) and dtypes of every SbObsPtrs field enabled by cfg."""
    out = {"ego_lane": ((), np.int32)}
    if cfg.nbr_enable:
        c = cfg.nbr_cap
        out.update(nbr_count=((), np.int32), nbr_id=((c,), np.int32), nbr_pos=((c, 3), np.float64),
                   nbr_box=((c, 3), np.float32), nbr_heading=((c,), np.float32), nbr_speed=((c,), np.float32),
                   nbr_lane=((c,), np.int32), nbr_lane_index=((c,), np.int8))
    if cfg.wp_enable:
        p, l = cfg.wp_out_paths, cfg.wp_out_len
        out.update(wp_n_paths=((), np.int32), wp_len=((p,), np.int32), wp_pos=((p, l, 3), np.float64),
                   wp_heading=((p, l), np.float32), wp_lane_width=((p, l), np.float32),
                   wp_speed_limit=((p, l), np.float32), wp_lane=((p, l), np.int32),
                   wp_lane_index=((p, l), np.int8))
    if cfg.lidar_enable:
        r = cfg.lidar_rays
        out.update(lidar_points=((r, 3), np.float64), lidar_hits=((r,), np.int8))
    if cfg.ogm_enable:
        out.update(ogm=((cfg.ogm_height, cfg.ogm_width), np.uint8))
    if cfg.dagm_enable:
        out.update(dagm=((cfg.dagm_height, cfg.dagm_width), np.uint8))
    if cfg.events_enable:
        out.update(ev_events=((), np.uint32), ev_done=((), np.uint8), ev_n_collisions=((), np.int32),
                   ev_collidee=((), np.int32), ev_distance=((), np.float64), ev_reward=((), np.float64),
                   ev_steps=((), np.int32))
    if cfg.ttc_enable:
        out.update(ttc_angle_error=((), np.float32), ttc_distance_from_center=((), np.float32),
                   ttc_dtc=((3,), np.float32), ttc_ttc=((3,), np.float32), ttc_valid=((), np.uint8))
    if cfg.accel_enable:
        out.update(acc_linear=((3,), np.float64), acc_angular=((3,), np.float64),
                   acc_linear_jerk=((3,), np.float64), acc_angular_jerk=((3,), np.float64))
    if cfg.via_cap > 0:
        out.update(via_near_count=((), np.int32), via_near_idx=((cfg.via_cap,), np.int32),
                   via_hit_count=((), np.int32), via_hit_idx=((cfg.via_cap,), np.int32))
    if cfg.rw_enable:
        lc, pc, hz = cfg.rw_lanes_cap, cfg.rw_paths_cap, 2 * cfg.rw_horizon + 1
        out.update(rw_n_lanes=((), np.int32), rw_lane=((lc,), np.int32), rw_n_paths=((), np.int32),
                   rw_path_slot=((pc,), np.int32), rw_len=((pc,), np.int32), rw_pos=((pc, hz, 2), np.float64),
                   rw_heading=((pc, hz), np.float64), rw_lane_width=((pc, hz), np.float32),
                   rw_speed_limit=((pc, hz), np.float32), rw_wp_lane=((pc, hz), np.int32),
                   rw_lane_index=((pc, hz), np.int8))
    return out


def state_shapes(cfg: SbConfig) -> Dict[str, tuple]:
    """(per-what, trailing shape, dtype): 'v' = per vehicle slot [E*V], 'a' = per agent [E*A]."""
    out = {n: ("v", (), np.float64) for n in
           ("pos_x", "pos_y", "pos_z", "heading", "speed", "yaw_rate", "length", "width", "height", "cos_h", "sin_h")}
    out["alive"] = ("v", (), np.uint8)
    for n in ("steering", "v_long", "v_lat"):
        out[n] = ("a", (), np.float64)
    out["ctrl"] = ("a", (SB_CTRL_STATE,), np.float64)
    out["last_control"] = ("a", (3,), np.float64)
    out["route"] = ("a", (max(cfg.route_cap, 1),), np.int32)
    out["route_len"] = ("a", (), np.int32)
    out["status"] = ("a", (), np.uint32)
    return out
