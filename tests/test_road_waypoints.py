"""RoadWaypoints sensor (/root/reference/smarts/core/sensors.py:991-1043): the oracle against
tests/golden/road_waypoints.npz -- outputs of the reference's own RoadWaypointsSensor on its own SumoRoadNetwork
(generator tests/golden/gen_road_waypoints_golden.py) -- and the CUDA path against the oracle and the fixture."""
import os

import numpy as np
import pytest

from conftest import GOLDEN, load_map, make_cfg, oracle_world
from smarts_b200 import _abi

MAPS = ("4lane", "6lane", "figure_eight", "loop")
HORIZONS = (20, 32)
ROUTE_CAP = 16


def fixture():
    return np.load(os.path.join(GOLDEN, "road_waypoints.npz"))


def poses(z, name, rm):
    x, y, h = z[f"{name}_x"], z[f"{name}_y"], z[f"{name}_h"]
    n = len(x)
    routes = [[rm.road_index[r] for r in s.split(";")] for s in z[f"{name}_routes"] if s]
    route = np.full((1, n, ROUTE_CAP), -1, np.int32)
    rlen = np.zeros((1, n), np.int32)
    for i, k in enumerate(z[f"{name}_route_of"]):
        if k >= 0:
            rlen[0, i] = len(routes[k])
            route[0, i, :len(routes[k])] = routes[k]
    return x[None], y[None], h[None], np.zeros((1, n)), route, rlen


def as_dicts(obs, n, hz):
    """[{lane: [path rows]}] per agent from the flat arrays: the reference's RoadWaypoints.lanes."""
    out = []
    for i in range(n):
        d = {}
        lanes = obs["rw_lane"][0, i]
        assert obs["rw_n_lanes"][0, i] <= len(lanes) and obs["rw_n_paths"][0, i] <= obs["rw_len"].shape[2]
        for s in range(obs["rw_n_lanes"][0, i]):
            d[int(lanes[s])] = []
        for p in range(obs["rw_n_paths"][0, i]):
            ln = int(obs["rw_len"][0, i, p])
            assert ln <= 2 * hz + 1
            rows = np.column_stack([obs["rw_pos"][0, i, p, :ln, 0], obs["rw_pos"][0, i, p, :ln, 1],
                                    obs["rw_heading"][0, i, p, :ln], obs["rw_lane_width"][0, i, p, :ln],
                                    obs["rw_speed_limit"][0, i, p, :ln], obs["rw_wp_lane"][0, i, p, :ln],
                                    obs["rw_lane_index"][0, i, p, :ln]])
            d[int(lanes[obs["rw_path_slot"][0, i, p]])].append(rows)
        out.append(d)
    return out


def check_against_fixture(z, name, hz, obs, n):
    got = as_dicts(obs, n, hz)
    lanes, meta, rows = z[f"{name}_lanes"], z[f"{name}_meta"], z[f"{name}_rows"]
    for i in range(n):
        want_lanes = lanes[(lanes[:, 0] == hz) & (lanes[:, 1] == i)]
        # dict keys in the reference's insertion order, and the number of paths under each
        assert list(got[i].keys()) == [int(l) for l in want_lanes[:, 3]], (name, hz, i)
        for _, _, _, lane, npaths in want_lanes:
            assert len(got[i][int(lane)]) == npaths, (name, hz, i, lane)
        for _, _, lane, p, ln, first in meta[(meta[:, 0] == hz) & (meta[:, 1] == i)]:
            g, w = got[i][int(lane)][p], rows[first:first + ln]
            assert g.shape == w.shape, (name, hz, i, lane, p)
            assert np.array_equal(g[:, 5:], w[:, 5:]), (name, hz, i, lane, p)           # lane ids, lane indices
            np.testing.assert_allclose(g[:, :3], w[:, :3], rtol=0, atol=1e-9)           # x, y, heading
            np.testing.assert_allclose(g[:, 3:5], w[:, 3:5], rtol=1e-6)                 # float32 width / speed limit


@pytest.mark.parametrize("name", MAPS)
@pytest.mark.parametrize("hz", HORIZONS)
def test_oracle_matches_reference_road_waypoints(name, hz):
    z, rm = fixture(), load_map(name)
    x, y, h, v, route, rlen = poses(z, name, rm)
    n = x.shape[1]
    cfg = make_cfg(E=1, A=n, rw=True, rw_horizon=hz, rw_lanes=32, rw_paths=64, route_cap=ROUTE_CAP)
    ora = oracle_world(rm, cfg)
    ora.reset(x, y, h, v, route, rlen)
    ora.observe()
    assert not (ora.state()["status"] & _abi.SB_STATUS_RW_TRUNCATED).any()
    check_against_fixture(z, name, hz, ora.obs(), n)


def test_oracle_flags_truncation():
    z, rm = fixture(), load_map("4lane")
    x, y, h, v, route, rlen = poses(z, "4lane", rm)
    n = x.shape[1]
    cfg = make_cfg(E=1, A=n, rw=True, rw_horizon=20, rw_lanes=4, rw_paths=6, route_cap=ROUTE_CAP)
    ora = oracle_world(rm, cfg)
    ora.reset(x, y, h, v, route, rlen)
    ora.observe()
    o, st = ora.obs(), ora.state()["status"][0]
    over = (o["rw_n_lanes"][0] > 4) | (o["rw_n_paths"][0] > 6)
    assert over.any() and np.array_equal((st & _abi.SB_STATUS_RW_TRUNCATED) != 0, over)


# ------------------------------------------------------------------------------------------ CUDA path
def gpu_engine(rm, cfg):
    from smarts_b200 import lidar_params
    from smarts_b200.engine import Engine
    return Engine(cfg, rm, lidar_params.base_rays(lidar_params.BasicLidar))


@pytest.mark.gpu
@pytest.mark.parametrize("name", MAPS)
@pytest.mark.parametrize("hz", HORIZONS)
def test_gpu_matches_reference_road_waypoints(name, hz):
    """k_road_waypoints against the reference's own sensor outputs, and bit for bit against the oracle."""
    from test_gpu_parity import assert_same
    z, rm = fixture(), load_map(name)
    x, y, h, v, route, rlen = poses(z, name, rm)
    n = x.shape[1]
    cfg = make_cfg(E=1, A=n, rw=True, rw_horizon=hz, rw_lanes=32, rw_paths=64, route_cap=ROUTE_CAP)
    eng, ora = gpu_engine(rm, cfg), oracle_world(rm, cfg)
    for w in (eng, ora):
        w.reset(x, y, h, v, route, rlen)
        w.observe()
    got = eng.obs_numpy()
    check_against_fixture(z, name, hz, got, n)
    assert_same(got, ora.obs())
    assert_same(eng.state_numpy(), ora.state())
    eng.close()


@pytest.mark.gpu
@pytest.mark.parametrize("name,ego_frame", [("minicity", False), ("4lane", True), ("figure_eight", False)])
def test_gpu_road_waypoints_rollout_matches_oracle(name, ego_frame):
    """Random poses on and off the lanes (minicity: 2 695 lanes, segment grid path), a Lane-action rollout, small caps on
    one of the runs so that the truncation flag and the capped writes are compared too."""
    import torch
    from test_gpu_parity import assert_same
    rm = load_map(name)
    rng = np.random.default_rng(7)
    E, A = 6, 8
    idx = rng.integers(0, len(rm.lp_x), (E, A))
    off = rng.normal(0, 1.5, (E, A))
    hd = rm.lp_heading[idx]
    x = rm.lp_x[idx] + off * np.cos(hd)
    y = rm.lp_y[idx] + off * np.sin(hd)
    h = hd + rng.normal(0, 0.2, (E, A))
    v = np.full((E, A), 6.0)
    small = name == "4lane"
    cfg = make_cfg(E=E, A=A, rw=True, rw_horizon=20, rw_lanes=8 if small else 32, rw_paths=12 if small else 64,
                   route_cap=ROUTE_CAP, wp=True, ego_frame=ego_frame)
    eng, ora = gpu_engine(rm, cfg), oracle_world(rm, cfg, threads=4)
    for w in (eng, ora):
        w.reset(x, y, h, v)
    act = np.zeros((E, A), np.int32)
    for t in range(4):
        eng.step(torch.from_numpy(act).cuda())
        ora.step(act)
        eng.observe()
        ora.observe()
        assert_same(eng.obs_numpy(), ora.obs())
        assert_same(eng.state_numpy(), ora.state())
    st = ora.state()["status"]
    assert bool((st & _abi.SB_STATUS_RW_TRUNCATED).any()) == small
    assert (ora.obs()["rw_n_paths"] > 0).any()
    eng.close()


@pytest.mark.gpu
def test_hiway_env_road_waypoints_observation():
    """AgentInterface(road_waypoints=...) through the hiway-v0 surface: Observation.road_waypoints is the reference's
    RoadWaypoints(lanes={lane_id: [[Waypoint]]}) (sensors.py:104-107, :1000-1011)."""
    from smarts_b200.agent_interface import ActionSpaceType, AgentInterface, AgentSpec, RoadWaypoints
    from smarts_b200.hiway_env import HiWayEnv, Scenario
    from smarts_b200.observation import Mission
    itf = AgentInterface(road_waypoints=RoadWaypoints(horizon=20), waypoints=True, action=ActionSpaceType.Lane,
                         max_episode_steps=5)
    sc = Scenario(map="4lane", missions={"a0": Mission("edge-south-SN", 1, 30.0, "edge-west-EW"),
                                         "a1": Mission("edge-west-WE", 0, 30.0, "edge-east-WE")})
    env = HiWayEnv([sc], {k: AgentSpec(interface=itf) for k in ("a0", "a1")}, seed=3)
    obs = env.reset()
    obs, _, _, _ = env.step({k: "keep_lane" for k in obs})
    for k, ob in obs.items():
        rw = ob.road_waypoints
        assert rw is not None and len(rw.lanes) >= 2
        ego_lane = ob.ego_vehicle_state.lane_id
        assert ego_lane in rw.lanes and len(rw.lanes[ego_lane]) >= 1
        path = rw.lanes[ego_lane][0]
        assert 1 <= len(path) <= 41 and path[0].lane_id.split("_")[0] == ego_lane.split("_")[0]
        # the lane's paths start `horizon` metres behind the ego (or at the lane start)
        d0 = np.linalg.norm(path[0].pos - ob.ego_vehicle_state.position[:2])
        assert d0 <= 20.0 + 2.0
    env.close()


def test_old_format_map_is_flagged_and_refused_for_road_waypoints(tmp_path):
    """An .npz written before the lane_in / road_par tables existed loads (with a warning) for every other sensor,
    but is refused for RoadWaypoints: its parallel-road table is empty and its incoming lanes are not in
    getIncoming() order, so the sensor's lane set would differ from the reference's (sensors.py:1007-1011)."""
    from smarts_b200.road_map import RoadMapData
    src = np.load(os.path.join(os.path.dirname(GOLDEN), "..", "maps", "4lane.npz"), allow_pickle=False)
    old = {k: src[k] for k in src.files if k not in ("lane_in_start", "lane_in", "road_par_start", "road_par")}
    path = str(tmp_path / "old.npz")
    np.savez_compressed(path, **old)
    with pytest.warns(UserWarning, match="RoadWaypoints tables"):
        rm = RoadMapData.load(path)
    assert not rm.rw_tables_complete
    assert len(rm.lane_in_start) == len(rm.lane_road) + 1  # still a usable map for the other sensors
    with pytest.raises(ValueError, match="RoadWaypoints"):
        rm.require_road_waypoint_tables()
    new = load_map("4lane")
    assert new.rw_tables_complete
    new.require_road_waypoint_tables()
