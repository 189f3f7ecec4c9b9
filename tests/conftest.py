import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

GOLDEN = os.path.join(ROOT, "tests", "golden")

from smarts_b200 import _abi, lidar_params  # noqa: E402
from smarts_b200.road_map import RoadMapData  # noqa: E402


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


_MAPS = {}


def load_map(name: str) -> RoadMapData:
    """Flattened maps committed under maps/ (written by tests/golden/gen_golden.py from the
    reference's own net.xml files; /root/reference is never read at test time)."""
    if name not in _MAPS:
        _MAPS[name] = RoadMapData.load(os.path.join(ROOT, "maps", name + ".npz"))
    return _MAPS[name]


def make_cfg(E=1, A=1, M=0, dt=0.1, action_space=_abi.SB_ACTION_LANE, traj_len=10, route_cap=16,
             nbr=False, nbr_radius=50.0, nbr_cap=10, wp=False, wp_lookahead=32, wp_paths=4, wp_len=20,
             lidar=False, lidar_miss_inf=0, wall=0.0, ogm=None, dagm=None, device=0, substeps=0,
             events=False, done_mask=None, max_episode_steps=0, not_moving_time=60.0, not_moving_distance=1.0,
             ttc=False, accel=False, via_cap=0, via_range=40.0, via_speed_accuracy=1.5, ego_frame=False,
             rw=False, rw_horizon=20, rw_lanes=32, rw_paths=64, vehicle_type=0):
    c = _abi.SbConfig()
    c.abi_version = _abi.SB_ABI_VERSION
    c.device = device
    c.n_envs, c.n_agents, c.n_social = E, A, M
    c.dt = dt
    c.substeps = substeps
    c.action_space = action_space
    c.traj_len = traj_len
    c.route_cap = route_cap
    c.nbr_enable, c.nbr_cap, c.nbr_radius = int(nbr), nbr_cap, nbr_radius
    c.wp_enable, c.wp_lookahead, c.wp_out_paths, c.wp_out_len = int(wp), wp_lookahead, wp_paths, wp_len
    c.lidar_enable, c.lidar_rays, c.lidar_miss_inf, c.lidar_wall_height = int(lidar), 300, lidar_miss_inf, wall
    if ogm:
        c.ogm_enable, c.ogm_width, c.ogm_height, c.ogm_resolution = 1, ogm[0], ogm[1], ogm[2]
    if dagm:
        c.dagm_enable, c.dagm_width, c.dagm_height, c.dagm_resolution = 1, dagm[0], dagm[1], dagm[2]
    if events:
        c.events_enable = 1
        # DoneCriteria defaults (agent_interface.py:185-206): collision, off_road, off_route
        c.done_mask = (_abi.SB_EV_COLLISION | _abi.SB_EV_OFF_ROAD | _abi.SB_EV_OFF_ROUTE) if done_mask is None else done_mask
        c.max_episode_steps = max_episode_steps
        c.not_moving_time, c.not_moving_distance = not_moving_time, not_moving_distance
    c.ttc_enable = int(ttc)
    c.accel_enable = int(accel)
    c.via_cap, c.via_acquisition_range, c.via_speed_accuracy = via_cap, via_range, via_speed_accuracy
    c.ego_frame = int(ego_frame)
    c.rw_enable, c.rw_horizon, c.rw_lanes_cap, c.rw_paths_cap = int(rw), rw_horizon, rw_lanes, rw_paths
    c.vehicle_type = vehicle_type
    return c


def oracle_world(rm: RoadMapData, cfg, threads=1, rays=None):
    from oracle import oracle_py
    if rays is None:
        rays = lidar_params.base_rays(lidar_params.BasicLidar)
    return oracle_py.OracleWorld(cfg, rm.arrays(), rm.spacing, rays, threads=threads)
