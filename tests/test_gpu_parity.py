"""GPU parity: the CUDA path (through the C-ABI, via smarts_b200.engine's ctypes loader) against
the CPU oracle on identical seeded inputs.  Everything is compared BIT-EXACT -- integers and
bytes by contract, floating point because both sides evaluate the same binary64 operation
sequences without FMA contraction (include/sb_detmath.h); the 1e-4 relative tolerance BASELINE
allows for poses / speeds / LiDAR distances is therefore met with zero error."""
import numpy as np
import pytest

from conftest import load_map, make_cfg, oracle_world
from smarts_b200 import _abi, lidar_params, scenario_gen

pytestmark = pytest.mark.gpu

MISSIONS_4LANE = [("edge-south-SN", 1, 10.0, "edge-west-EW"), ("edge-west-WE", 0, 10.0, "edge-east-WE"),
                  ("edge-north-NS", 0, 10.0, "edge-south-NS"), ("edge-east-EW", 1, 10.0, "edge-west-EW")]
MISSIONS_LOOP = [("445633931", 0, 10.0, "445633932"), ("445633931", 2, 40.0, "445633932")]
MISSIONS_F8 = [("gneE20", 0, 5.0, "gneE27"), ("gneE36", 1, 5.0, "gneE25"), ("gneE20", 2, 25.0, "gneE27"),
               ("gneE36", 0, 30.0, "gneE25")]


def engine(rm, cfg, rays=None):
    from smarts_b200.engine import Engine
    return Engine(cfg, rm, lidar_params.base_rays(lidar_params.BasicLidar) if rays is None else rays)


def assert_same(got: dict, want: dict, keys=None):
    for k in (keys or want.keys()):
        g, w = got[k], want[k]
        assert g.shape == w.shape, k
        if g.dtype.kind == "f":
            same = (g.view(np.uint8) == w.view(np.uint8)).all() or np.array_equal(g, w)  # -0.0 == 0.0 tolerated
        else:
            same = np.array_equal(g.view(w.dtype) if g.dtype != w.dtype else g, w)
        if not same:
            bad = np.argwhere(g != w)
            raise AssertionError(f"{k}: {len(bad)} mismatches, first at {bad[0]}: got {g[tuple(bad[0])]} want {w[tuple(bad[0])]}")


def run_pair(rm, cfg, missions, ticks, seed, action_fn, jitter=20.0, check_every=1, social_alive=0.9, rays=None):
    import torch
    E, M = cfg.n_envs, cfg.n_social
    x, y, h, v, route, rlen = scenario_gen.agent_starts(rm, E, missions, cfg.route_cap, seed, jitter=jitter, speed=5.0)
    traffic = scenario_gen.SocialTraffic(rm, E, M, seed + 1, p_alive=social_alive) if M else None
    eng = engine(rm, cfg, rays)
    ora = oracle_world(rm, cfg, threads=8, rays=rays)
    eng.reset(x, y, h, v, route, rlen)
    ora.reset(x, y, h, v, route, rlen)
    for t in range(ticks):
        act = action_fn(t, ora)
        if traffic is not None:
            soc = traffic.states(t)
            eng.set_social(torch.from_numpy(soc).cuda())
            ora.set_social(soc)
        eng.step(torch.from_numpy(act).cuda())
        ora.step(act)
        if t % check_every == 0 or t == ticks - 1:
            eng.observe()
            ora.observe()
            assert_same(eng.state_numpy(), ora.state())
            assert_same(eng.obs_numpy(), ora.obs())
    return eng, ora


def lane_actions(E, A, seed):
    rng = np.random.default_rng(seed)

    def fn(t, _):
        return rng.integers(0, 4, (E, A)).astype(np.int32)
    return fn


def test_config2_figure_eight_neighbors_waypoints():
    """BASELINE configs[1] shape: 4 agents x envs, NeighborhoodVehicles + Waypoints, recorded traffic."""
    rm = load_map("figure_eight")
    E, A, M = 24, 4, 20
    cfg = make_cfg(E=E, A=A, M=M, nbr=True, nbr_radius=50.0, wp=True, wp_lookahead=32, route_cap=24)
    run_pair(rm, cfg, MISSIONS_F8, ticks=40, seed=7, action_fn=lane_actions(E, A, 3))


def test_config3_4lane_lidar_ogm_dagm():
    """BASELINE configs[2] shape: LiDAR 300 rays + OGM + DAGM 256x256 @ 0.1 m/px (few envs)."""
    rm = load_map("4lane")
    E, A, M = 6, 4, 20
    cfg = make_cfg(E=E, A=A, M=M, lidar=True, wall=1.5, ogm=(256, 256, 0.1), dagm=(256, 256, 0.1), route_cap=16)
    eng, ora = run_pair(rm, cfg, MISSIONS_4LANE, ticks=30, seed=11, action_fn=lane_actions(E, A, 5), check_every=5)
    o = ora.obs()
    assert o["ogm"].max() == 255 and o["dagm"].max() == 255 and o["lidar_hits"].sum() > 0
    assert 0.05 < (o["dagm"] == 255).mean() < 0.95


@pytest.mark.parametrize("table", ["hdl32e", "ring"])
def test_lidar_other_ray_tables(table):
    """Ray tables other than BasicLidar: VelodyneHDL32E (36 azimuth groups x 24 laser angles, 100 m reach: the
    interval lists are rebuilt per chunk of groups) and a flat ring of 250 distinct azimuths x 2 laser angles
    plus straight-up / straight-down rays (every group holds two rays; vertical rays take the ground-only path)."""
    rm = load_map("4lane")
    E, A, M = 4, 4, 20
    if table == "hdl32e":
        rays = lidar_params.base_rays(lidar_params.VelodyneHDL32E)
    else:
        az = np.linspace(0.0, 2 * np.pi, 250, endpoint=False)
        rays = np.concatenate([
            np.stack([30 * np.cos(az), 30 * np.sin(az), np.zeros_like(az)], 1),
            np.stack([25 * np.cos(az), 25 * np.sin(az), np.full_like(az, -2.0)], 1),
            np.array([[0.0, 0.0, -5.0], [0.0, 0.0, 5.0], [0.0, 0.0, -0.5]])])
    cfg = make_cfg(E=E, A=A, M=M, lidar=True, lidar_miss_inf=1, wall=1.5, route_cap=16)
    cfg.lidar_rays = len(rays)
    eng, ora = run_pair(rm, cfg, MISSIONS_4LANE, ticks=12, seed=23, action_fn=lane_actions(E, A, 9), check_every=3,
                        rays=rays)
    o = ora.obs()
    assert 0 < o["lidar_hits"].mean() < 1


def test_config1_loop_laner_full_suite():
    """BASELINE configs[0] shape: Laner agent, keep_lane, waypoints + LiDAR + OGM 64x64 @ 50/64."""
    rm = load_map("loop")
    E, A, M = 4, 2, 20
    cfg = make_cfg(E=E, A=A, M=M, nbr=True, wp=True, lidar=True, lidar_miss_inf=1, ogm=(64, 64, 50 / 64),
                   dagm=(64, 64, 50 / 64), route_cap=8)

    def keep(t, _):
        return np.zeros((E, A), np.int32)
    run_pair(rm, cfg, MISSIONS_LOOP, ticks=120, seed=42, action_fn=keep, check_every=10)


def test_lane_with_continuous_speed_and_rollout_stays_on_road():
    rm = load_map("4lane")
    E, A = 8, 4
    cfg = make_cfg(E=E, A=A, M=0, action_space=_abi.SB_ACTION_LANE_WITH_CONTINUOUS_SPEED, wp=True, route_cap=16)
    rng = np.random.default_rng(1)

    def fn(t, _):
        a = np.zeros((E, A, 2))
        a[..., 0] = rng.uniform(0, 14, (E, A))
        a[..., 1] = rng.integers(-1, 2, (E, A))
        return a
    run_pair(rm, cfg, MISSIONS_4LANE, ticks=60, seed=2, action_fn=fn, check_every=6)


def test_trajectory_tracking_action_space():
    """BASELINE configs[3] shape: Trajectory action = first 10 waypoints of path 0 at 50/3.6 m/s
    (examples/trajectory_tracking_agent.py:16-37), fed back from the oracle's own observation."""
    rm = load_map("figure_eight")
    E, A, N = 6, 4, 10
    cfg = make_cfg(E=E, A=A, M=10, action_space=_abi.SB_ACTION_TRAJECTORY, traj_len=N, wp=True, wp_len=20,
                   route_cap=24)
    ora0 = oracle_world(rm, cfg)
    x, y, h, v, route, rlen = scenario_gen.agent_starts(rm, E, MISSIONS_F8, cfg.route_cap, 13, jitter=20.0, speed=5.0)
    ora0.reset(x, y, h, v, route, rlen)
    ora0.observe()

    def fn(t, ora):
        ora.observe()
        o = ora.obs()
        act = np.zeros((E, A, 4, N))
        act[:, :, 0] = o["wp_pos"][:, :, 0, :N, 0]
        act[:, :, 1] = o["wp_pos"][:, :, 0, :N, 1]
        act[:, :, 2] = o["wp_heading"][:, :, 0, :N].astype(np.float64)
        act[:, :, 3] = 50 / 3.6
        return act
    run_pair(rm, cfg, MISSIONS_F8, ticks=50, seed=13, action_fn=fn, check_every=5)


@pytest.mark.parametrize("space", [_abi.SB_ACTION_CONTINUOUS, _abi.SB_ACTION_ACTUATOR_DYNAMIC])
def test_continuous_action_spaces(space):
    rm = load_map("straight")
    E, A = 16, 3
    cfg = make_cfg(E=E, A=A, M=6, action_space=space, nbr=True, nbr_radius=-1.0, wp=True, ogm=(64, 64, 0.5),
                   route_cap=4)
    rng = np.random.default_rng(9)
    missions = [("west", 0, 5.0, "east"), ("west", 1, 15.0, "east"), ("west", 2, 25.0, "east")]

    def fn(t, _):
        a = rng.uniform(-0.2, 1.2, (E, A, 3))
        a[..., 2] = rng.uniform(-1.3, 1.3, (E, A))
        return a
    run_pair(rm, cfg, missions, ticks=40, seed=17, action_fn=fn, check_every=4)


def test_edge_cases_empty_social_dead_slots_and_partial_reset():
    import torch
    rm = load_map("4lane")
    E, A, M = 5, 4, 8
    cfg = make_cfg(E=E, A=A, M=M, nbr=True, nbr_radius=30.0, nbr_cap=3, wp=True, lidar=True,
                   ogm=(128, 64, 0.25), dagm=(64, 128, 0.25), route_cap=16)
    x, y, h, v, route, rlen = scenario_gen.agent_starts(rm, E, MISSIONS_4LANE, cfg.route_cap, 3, jitter=30.0, speed=8.0)
    eng, ora = engine(rm, cfg), oracle_world(rm, cfg)
    for w in (eng, ora):
        w.reset(x, y, h, v, route, rlen)
    # no social set at all: every social slot dead
    eng.observe(); ora.observe()
    assert_same(eng.obs_numpy(), ora.obs())
    assert (ora.obs()["nbr_count"] <= A - 1).all()
    # ragged: some slots alive, vehicles stacked on the ego (lidar origin inside a box, OGM overlap)
    soc = scenario_gen.SocialTraffic(rm, E, M, 5, p_alive=0.5).states(0)
    soc[0, 0, :3] = [x[0, 0], y[0, 0], h[0, 0]]
    soc[0, 0, 7] = 1.0
    soc[1, :, 7] = 0.0
    act = np.zeros((E, A), np.int32)
    for t in range(3):
        eng.set_social(torch.from_numpy(soc).cuda()); ora.set_social(soc)
        eng.step(torch.from_numpy(act).cuda()); ora.step(act)
    eng.observe(); ora.observe()
    assert_same(eng.state_numpy(), ora.state())
    assert_same(eng.obs_numpy(), ora.obs())
    # partial reset of two envs (auto-reset path of ParallelEnv, parallel_env.py:267-326)
    ids = np.array([3, 1], np.int32)
    for w in (eng, ora):
        w.reset(x[ids], y[ids], h[ids], v[ids], route[ids], rlen[ids], env_ids=ids)
    eng.observe(); ora.observe()
    assert_same(eng.state_numpy(), ora.state())
    assert_same(eng.obs_numpy(), ora.obs())


def test_off_map_and_no_route_agents():
    """Agents far from any lane and agents with an empty Route (-> _resolve_in_junction /
    closest_lanepoints fallback, sumo_road_network.py:822-840)."""
    import torch
    rm = load_map("4lane")
    E, A = 6, 4
    cfg = make_cfg(E=E, A=A, M=0, nbr=True, wp=True, dagm=(64, 64, 0.5), route_cap=16)
    rng = np.random.default_rng(4)
    x = rng.uniform(100, 200, (E, A)); y = rng.uniform(-110, -10, (E, A))
    x[0, :] = [500.0, -300.0, 150.0, 150.0]; y[0, :] = [500.0, 40.0, -60.0, -59.0]
    h = rng.uniform(-3.1, 3.1, (E, A)); v = np.zeros((E, A))
    eng, ora = engine(rm, cfg), oracle_world(rm, cfg)
    for w in (eng, ora):
        w.reset(x, y, h, v, None, None)
    act = np.zeros((E, A), np.int32)
    for t in range(5):
        eng.step(torch.from_numpy(act).cuda()); ora.step(act)
        eng.observe(); ora.observe()
        assert_same(eng.state_numpy(), ora.state())
        assert_same(eng.obs_numpy(), ora.obs())


@pytest.mark.parametrize("name,wp", [("figure_eight", True), ("loop", True), ("minicity", True), ("minicity", False)])
def test_unrouted_queries_at_random_positions(name, wp):
    """Closest-lane-point queries without a route at positions anywhere on and around the map: on a lane, a few
    metres off, and beyond the lane-point grid's three rings (the table scan takes over there); with and without
    the waypoint sensor (without it the events pass makes the lookahead-1 trip-meter query)."""
    import torch
    rm = load_map(name)
    E, A = 64, 4
    cfg = make_cfg(E=E, A=A, M=0, nbr=False, wp=wp, wp_lookahead=16, wp_paths=8, wp_len=17, route_cap=16, events=True)
    rng = np.random.default_rng(21)
    on = rng.integers(0, len(rm.lp_x), (E, A))
    spread = np.where(rng.random((E, A)) < 0.6, rng.normal(0, 2.0, (E, A)), rng.uniform(-60, 60, (E, A)))
    ang = rng.uniform(-np.pi, np.pi, (E, A))
    x = rm.lp_x[on] + spread * np.cos(ang); y = rm.lp_y[on] + spread * np.sin(ang)
    x[0, 0], y[0, 0] = rm.lp_x.min() - 400.0, rm.lp_y.max() + 400.0      # far outside the grid
    h = rng.uniform(-3.1, 3.1, (E, A)); v = np.zeros((E, A))
    eng, ora = engine(rm, cfg), oracle_world(rm, cfg, threads=8)
    for w in (eng, ora):
        w.reset(x, y, h, v, None, None)
    act = np.zeros((E, A), np.int32)
    for t in range(2):
        eng.step(torch.from_numpy(act).cuda()); ora.step(act)
        eng.observe(); ora.observe()
        assert_same(eng.state_numpy(), ora.state())
        assert_same(eng.obs_numpy(), ora.obs())
    if wp:
        n = ora.obs()["wp_n_paths"]
        assert (n > 0).mean() > 0.5


def test_full_size_grid_properties():
    """At BASELINE's full grid size the oracle is too slow for many envs, so check size-independent
    properties on a larger batch: determinism (two runs bit-identical), OGM == 255 at the ego's
    own projected centre (test_observations.py:179-187), DAGM covers the ego when on road, and
    translation invariance (shifting nothing but env index leaves results unchanged)."""
    import torch
    rm = load_map("4lane")
    E, A, M = 256, 4, 20
    cfg = make_cfg(E=E, A=A, M=M, lidar=True, ogm=(256, 256, 0.1), dagm=(256, 256, 0.1), route_cap=16)
    x, y, h, v, route, rlen = scenario_gen.agent_starts(rm, E, MISSIONS_4LANE, cfg.route_cap, 21, jitter=25.0, speed=6.0)
    # env e and env e + E/2 are identical copies
    half = E // 2
    for arr in (x, y, h, v, route, rlen):
        arr[half:] = arr[:half]
    soc = scenario_gen.SocialTraffic(rm, E, M, 22).states(0)
    soc[half:] = soc[:half]
    outs = []
    for rep in range(2):
        eng = engine(rm, cfg)
        eng.reset(x, y, h, v, route, rlen)
        eng.set_social(torch.from_numpy(soc).cuda())
        eng.step(torch.zeros((E, A), dtype=torch.int32, device="cuda"))
        eng.observe()
        outs.append(eng.obs_numpy())
        eng.close()
    for k in outs[0]:
        assert np.array_equal(outs[0][k], outs[1][k]), k
        assert np.array_equal(outs[0][k][:half], outs[0][k][half:]), k
    ogm, dagm = outs[0]["ogm"], outs[0]["dagm"]
    assert (ogm[:, :, 127:129, 127:129] == 255).all()
    assert (dagm[:, :, 127:129, 127:129] == 255).all()
    assert set(np.unique(ogm)) <= {0, 255} and set(np.unique(dagm)) <= {0, 255}


def _goals(rm, E, missions, radius=3.0):
    """PositionalGoal at the middle of lane 0 of each mission's end road (plan.py:95-115)."""
    A = len(missions)
    gx = np.zeros((E, A)); gy = np.zeros((E, A)); gr = np.full((E, A), radius)
    for a, (_, _, _, end) in enumerate(missions):
        lane = rm.lane_of_road(end, 0)
        px, py, _ = rm.lane_start_pose(end, 0, 0.5 * float(rm.lane_length[lane]))
        gx[:, a], gy[:, a] = px, py
    return gx, gy, gr


@pytest.mark.parametrize("wp", [True, False])
def test_events_rewards_dones_match_oracle(wp):
    """Events bits, done flags, trip-meter reward / distance, contacts --
    bit-exact against the oracle over a rollout with random lane commands (agents drift off route,
    collide with traffic, some reach their goal, all hit max_episode_steps)."""
    import torch
    rm = load_map("4lane")
    E, A, M = 24, 4, 12
    cfg = make_cfg(E=E, A=A, M=M, wp=wp, nbr=True, route_cap=16, events=True, max_episode_steps=55,
                   not_moving_time=2.0, not_moving_distance=1.0,
                   done_mask=_abi.SB_EV_COLLISION | _abi.SB_EV_OFF_ROAD | _abi.SB_EV_OFF_ROUTE | _abi.SB_EV_WRONG_WAY)
    x, y, h, v, route, rlen = scenario_gen.agent_starts(rm, E, MISSIONS_4LANE, cfg.route_cap, 31, jitter=40.0, speed=7.0)
    goal = _goals(rm, E, MISSIONS_4LANE)
    goal[2][E // 2:] = -1.0  # half the envs: endless missions
    traffic = scenario_gen.SocialTraffic(rm, E, M, 32, p_alive=0.9)
    eng, ora = engine(rm, cfg), oracle_world(rm, cfg, threads=8)
    for w in (eng, ora):
        w.reset(x, y, h, v, route, rlen, goal=goal)
    eng.observe(); ora.observe()  # the reset observation: no tick, nobody done
    assert_same(eng.obs_numpy(), ora.obs())
    assert not ora.obs()["ev_done"].any()
    rng = np.random.default_rng(33)
    seen = 0
    for t in range(60):
        act = rng.integers(0, 4, (E, A)).astype(np.int32)
        act[:, 3] = _abi.SB_LANE_SLOW_DOWN if t > 20 else _abi.SB_LANE_KEEP  # agent 3 stops: not_moving
        soc = traffic.states(t)
        eng.set_social(torch.from_numpy(soc).cuda()); ora.set_social(soc)
        eng.step(torch.from_numpy(act).cuda()); ora.step(act)
        eng.observe(); ora.observe()
        got, want = eng.obs_numpy(), ora.obs()
        assert_same(got, want)
        seen |= int(np.bitwise_or.reduce(want["ev_events"].ravel()))
        if t == 20:  # partial reset restarts the trip meter / step counter of those envs
            ids = np.array([2, 5], np.int32)
            for w in (eng, ora):
                w.reset(x[ids], y[ids], h[ids], v[ids], route[ids], rlen[ids], env_ids=ids,
                        goal=tuple(g[ids] for g in goal))
    want = ora.obs()
    assert (want["ev_steps"].ravel() >= 39).all()
    assert want["ev_done"][want["ev_steps"] >= 55].all()  # max_episode_steps reached (envs 2, 5 restarted at t = 20)
    assert want["ev_distance"].max() > 20.0
    for bit in (_abi.SB_EV_COLLISION, _abi.SB_EV_NOT_MOVING, _abi.SB_EV_REACHED_MAX_STEPS):
        assert seen & bit, f"rollout never raised event bit {bit}"


def test_partial_reset_observes_only_the_reset_envs():
    """sb_observe_envs: after a partial reset only the reset envs are observed (ParallelEnv envs are independent,
    parallel_env.py:96-127).  CUDA == oracle bit for bit through the reset, and every env that was NOT reset holds
    exactly what a run without any reset holds -- stateful sensors included (driven path / not_moving window, trip
    meter, accelerometer history, vias' consumed set, step counters) and the ev_* outputs of the last tick."""
    import torch
    rm = load_map("4lane")
    E, A, M = 8, 4, 10
    cfg = make_cfg(E=E, A=A, M=M, nbr=True, wp=True, lidar=True, wall=1.5, ogm=(64, 64, 0.4), dagm=(64, 64, 0.4),
                   route_cap=16, events=True, max_episode_steps=0, not_moving_time=1.0, not_moving_distance=0.5,
                   ttc=True, accel=True, ego_frame=True)
    x, y, h, v, route, rlen = scenario_gen.agent_starts(rm, E, MISSIONS_4LANE, cfg.route_cap, 61, jitter=30.0, speed=6.0)
    traffic = scenario_gen.SocialTraffic(rm, E, M, 62)
    ids = np.array([1, 4, 6], np.int32)
    keep = np.setdiff1d(np.arange(E), ids)
    runs = {}
    for name in ("reset", "plain"):
        eng, ora = engine(rm, cfg), oracle_world(rm, cfg, threads=8)
        for w in (eng, ora):
            w.reset(x, y, h, v, route, rlen)
            w.observe()
        rng = np.random.default_rng(63)
        hist = []
        for t in range(24):
            act = rng.integers(0, 4, (E, A)).astype(np.int32)
            act[:, 2] = _abi.SB_LANE_SLOW_DOWN  # agent 2 stops: not_moving depends on the driven-path window
            soc = traffic.states(t)
            eng.set_social(torch.from_numpy(soc).cuda()); ora.set_social(soc)
            eng.step(torch.from_numpy(act).cuda()); ora.step(act)
            eng.observe(); ora.observe()
            if name == "reset" and t in (7, 15):
                for w in (eng, ora):
                    w.reset(x[ids], y[ids], h[ids], v[ids], route[ids], rlen[ids], env_ids=ids, keep_social=True)
                    w.observe(ids)
            got, want = eng.obs_numpy(), ora.obs()
            assert_same(got, want)
            assert_same(eng.state_numpy(), ora.state())
            hist.append(got)
        runs[name] = hist
        eng.close()
    for t, (a, b) in enumerate(zip(runs["reset"], runs["plain"])):
        for k in a:
            assert np.array_equal(a[k][keep], b[k][keep], equal_nan=a[k].dtype.kind == "f"), (t, k)
    last = runs["reset"][-1]
    assert last["ev_steps"][ids].max() == 24 - 16 and last["ev_steps"][keep].min() == 24
    assert (last["ev_events"][keep][:, 2] & _abi.SB_EV_NOT_MOVING).all()


def test_events_off_road_and_wrong_way_poses():
    """Hand-placed poses: on the lane, on the shoulder, off the road, facing against the lane."""
    rm = load_map("4lane")
    E, A = 2, 4
    cfg = make_cfg(E=E, A=A, M=0, route_cap=16, events=True)
    x0, y0, h0 = rm.lane_start_pose("edge-south-SN", 0, 30.0)
    fx, fy = -np.sin(h0), np.cos(h0)      # forward
    rx, ry = np.cos(h0), np.sin(h0)        # "side" axis
    x = np.array([[x0, x0 - 1.2 * rx, x0 - 30 * rx, x0]] * E)
    y = np.array([[y0, y0 - 1.2 * ry, y0 - 30 * ry, y0]] * E)
    h = np.array([[h0, h0, h0, h0 + np.pi]] * E)
    v = np.zeros((E, A))
    eng, ora = engine(rm, cfg), oracle_world(rm, cfg)
    for w in (eng, ora):
        w.reset(x, y, h, v, None, None)
        w.observe()
    got, want = eng.obs_numpy(), ora.obs()
    assert_same(got, want)
    ev = want["ev_events"][0]
    assert not ev[0] & (_abi.SB_EV_OFF_ROAD | _abi.SB_EV_WRONG_WAY)
    assert ev[2] & _abi.SB_EV_OFF_ROAD and ev[2] & _abi.SB_EV_ON_SHOULDER
    assert ev[3] & _abi.SB_EV_WRONG_WAY
    _ = fx, fy


@pytest.mark.parametrize("pinned,share", [(True, None), (False, None), (True, "0.4"), (True, "0")])
def test_host_buffer_call_matches_device_views(pinned, share, monkeypatch):
    """sb_step_host + sb_copy_obs_to_host (the reference-facing call with HOST buffers) against the
    zero-copy device views and the oracle: the grids cross PCIe bit-packed and are widened on the
    host, every other buffer is copied as is -- results must be identical, for pinned and for
    unaligned pageable destinations, on grids whose size is not a multiple of the chunking."""
    import torch
    if share is not None:  # part of the agents' grids as plain byte copies, the rest packed
        monkeypatch.setenv("SB_XFER_PACKED_SHARE", share)
    rm = load_map("4lane")
    E, A, M = 37, 4, 9
    cfg = make_cfg(E=E, A=A, M=M, nbr=True, wp=True, lidar=True, wall=1.5, ogm=(96, 40, 0.3), dagm=(160, 72, 0.2),
                   route_cap=16, events=True)
    x, y, h, v, route, rlen = scenario_gen.agent_starts(rm, E, MISSIONS_4LANE, cfg.route_cap, 41, jitter=30.0, speed=6.0)
    traffic = scenario_gen.SocialTraffic(rm, E, M, 42)
    eng, ora = engine(rm, cfg), oracle_world(rm, cfg, threads=8)
    for w in (eng, ora):
        w.reset(x, y, h, v, route, rlen)
    bufs = eng.host_obs_buffers(pinned=pinned)
    if not pinned:  # misaligned destinations: views one byte into a larger buffer
        for k in ("ogm", "dagm"):
            raw = torch.empty(bufs[k].numel() + 1, dtype=torch.uint8)
            bufs[k] = raw[1:].view(bufs[k].shape)
    act = np.random.default_rng(43).integers(0, 4, (E, A)).astype(np.int32)
    for t in range(4):
        soc = traffic.states(t)
        eng.step_host(act, soc)
        eng.copy_obs_to_host(bufs)
        ora.set_social(soc); ora.step(act); ora.observe()
    got = {k: t.numpy() for k, t in bufs.items()}
    assert_same(got, ora.obs())
    assert_same(got, eng.obs_numpy())
    assert got["ogm"].any() and got["dagm"].any()


@pytest.mark.parametrize("pinned", [True, False])
def test_packed_host_grids_unpack_to_the_uint8_images(pinned):
    """sb_copy_obs_to_host_packed: the grids arrive as 1-bit images in the layout smarts_b200.h documents
    (numpy.unpackbits(..., bitorder="little") * 255 is the uint8 image); every other member as in the uint8 call."""
    rm = load_map("4lane")
    E, A, M = 19, 4, 9
    cfg = make_cfg(E=E, A=A, M=M, nbr=True, lidar=True, wall=1.5, ogm=(96, 40, 0.3), dagm=(160, 72, 0.2), route_cap=16)
    x, y, h, v, route, rlen = scenario_gen.agent_starts(rm, E, MISSIONS_4LANE, cfg.route_cap, 61, jitter=30.0, speed=6.0)
    traffic = scenario_gen.SocialTraffic(rm, E, M, 62)
    eng = engine(rm, cfg)
    eng.reset(x, y, h, v, route, rlen)
    bufs = eng.host_obs_buffers(pinned=pinned, packed_grids=True)
    assert tuple(bufs["ogm"].shape) == (E, A, 40, 96 // 8) and tuple(bufs["dagm"].shape) == (E, A, 72, 160 // 8)
    full = eng.host_obs_buffers(pinned=pinned)
    act = np.random.default_rng(63).integers(0, 4, (E, A)).astype(np.int32)
    for t in range(3):
        eng.step_host(act, traffic.states(t))
        eng.copy_obs_to_host(bufs, packed=True)
        eng.copy_obs_to_host(full)
        for k, t_full in full.items():
            got = bufs[k].numpy()
            if k in ("ogm", "dagm"):
                got = eng.unpack_grid(got)
            assert np.array_equal(got, t_full.numpy(), equal_nan=got.dtype.kind == "f"), k
    assert full["ogm"].numpy().any() and full["dagm"].numpy().any()


@pytest.mark.parametrize("pinned", [True, False])
def test_incremental_host_grids_hold_the_full_images(pinned):
    """sb_set_incremental_host_grids: with the same destination buffers call after call only the changed 64-pixel
    words are rewritten -- the buffers must still equal the device images after every call, including after a
    reset (large change), a switch to other buffers and back, and turning the mode off and on."""
    import torch
    rm = load_map("4lane")
    E, A, M = 21, 4, 9
    cfg = make_cfg(E=E, A=A, M=M, ogm=(96, 40, 0.3), dagm=(160, 72, 0.2), route_cap=16)
    x, y, h, v, route, rlen = scenario_gen.agent_starts(rm, E, MISSIONS_4LANE, cfg.route_cap, 51, jitter=30.0, speed=6.0)
    traffic = scenario_gen.SocialTraffic(rm, E, M, 52)
    eng = engine(rm, cfg)
    eng.reset(x, y, h, v, route, rlen)
    eng.set_incremental_host_grids(True)

    def buffers():
        b = eng.host_obs_buffers(pinned=pinned)
        if not pinned:
            for k in ("ogm", "dagm"):
                raw = torch.empty(b[k].numel() + 1, dtype=torch.uint8)
                b[k] = raw[1:].view(b[k].shape)
        for k in ("ogm", "dagm"):
            b[k].fill_(0x5a)  # garbage: the first call into a buffer must write all of it
        return b
    bufs, other = buffers(), buffers()
    act = np.random.default_rng(53).integers(0, 4, (E, A)).astype(np.int32)
    changed = 0
    for t in range(14):
        if t == 5:
            eng.reset(x[:7], y[:7], h[:7], v[:7], route[:7], rlen[:7], env_ids=np.arange(7, dtype=np.int32))
        if t == 11:
            eng.set_incremental_host_grids(False); eng.set_incremental_host_grids(True)
        dst = other if t in (7, 8) else bufs      # other buffers for two calls, then back to the stale first pair
        before = dst["dagm"].numpy().copy()
        eng.step_host(act, traffic.states(t))
        eng.copy_obs_to_host(dst)
        dev = eng.obs_numpy()
        for k in ("ogm", "dagm"):
            assert np.array_equal(dst[k].numpy(), dev[k]), (t, k)
        changed += int((before != dst["dagm"].numpy()).any())
    assert changed >= 10 and dev["ogm"].any() and dev["dagm"].any()


def test_config4_minicity_trajectory_tracking_100_social():
    """BASELINE configs[3] shape: minicity (52 k lane points), 32 TrajectoryTracking agents per env with
    100 BoxChassis social vehicles; action = first 10 waypoints of path 0 at 50/3.6 m/s
    (examples/trajectory_tracking_agent.py:16-37)."""
    rm = load_map("minicity")
    E, A, M, N = 2, 32, 100, 10
    cfg = make_cfg(E=E, A=A, M=M, action_space=_abi.SB_ACTION_TRAJECTORY, traj_len=N, wp=True, wp_len=20,
                   nbr=True, route_cap=32, events=True)
    missions = scenario_gen.random_missions(rm, A, 13, cfg.route_cap)

    import torch
    x, y, h, v, route, rlen = scenario_gen.agent_starts(rm, E, missions, cfg.route_cap, 13, jitter=5.0, speed=5.0)
    traffic = scenario_gen.SocialTraffic(rm, E, M, 14, p_alive=0.9)
    eng, ora = engine(rm, cfg), oracle_world(rm, cfg, threads=8)
    for w in (eng, ora):
        w.reset(x, y, h, v, route, rlen)
        w.observe()  # both sides observe exactly once per tick: the trip meter / step counter are stateful
    for t in range(24):
        o = ora.obs()
        act = np.zeros((E, A, 4, N))
        act[:, :, 0] = o["wp_pos"][:, :, 0, :N, 0]
        act[:, :, 1] = o["wp_pos"][:, :, 0, :N, 1]
        act[:, :, 2] = o["wp_heading"][:, :, 0, :N].astype(np.float64)
        act[:, :, 3] = 50 / 3.6
        soc = traffic.states(t)
        eng.set_social(torch.from_numpy(soc).cuda()); ora.set_social(soc)
        eng.step(torch.from_numpy(act).cuda()); ora.step(act)
        eng.observe(); ora.observe()
        assert_same(eng.state_numpy(), ora.state())
        assert_same(eng.obs_numpy(), ora.obs())
    assert (ora.obs()["wp_len"][:, :, 0] == 20).mean() > 0.8 and ora.obs()["ev_distance"].mean() > 5.0


def test_config5_straight_full_sensor_suite():
    """BASELINE configs[4] shape: straight road, 16 agents + 16 social per env, every sensor on
    (nbr r=50, wp 32, LiDAR 300, OGM + DAGM at 64x64 and at 256x256)."""
    rm = load_map("straight")
    E, A, M = 6, 16, 16
    missions = [("west", k % 3, 4.0 + 5.5 * k, "east") for k in range(A)]
    for grid in ((64, 64, 50 / 64), (256, 256, 0.1)):
        cfg = make_cfg(E=E, A=A, M=M, nbr=True, nbr_radius=50.0, wp=True, wp_lookahead=32, lidar=True, wall=1.5,
                       ogm=grid, dagm=grid, route_cap=4, events=True)
        eng, ora = run_pair(rm, cfg, missions, ticks=12, seed=17, action_fn=lane_actions(E, A, 6), jitter=2.0,
                            check_every=4)
        o = ora.obs()
        assert (o["nbr_count"] > 0).all() and o["lidar_hits"].sum() > 0 and (o["dagm"] == 255).any()


def test_maximum_grid_sizes_and_multi_pass_rows():
    """Grids at the limits of the rasteriser: 1024 columns (four 256-column passes), more than 256 rows (several
    row passes), the narrowest legal width, very coarse and very fine resolutions -- bit-exact against the oracle."""
    import torch
    rm = load_map("4lane")
    E, A, M = 2, 4, 20
    x, y, h, v, route, rlen = scenario_gen.agent_starts(rm, E, MISSIONS_4LANE, 16, 51, jitter=60.0, speed=9.0)
    soc = scenario_gen.SocialTraffic(rm, E, M, 52).states(3)
    act = np.zeros((E, A), np.int32)
    for ogm, dagm in (((1024, 300, 0.05), (32, 1024, 0.4)), ((512, 513, 0.11), (1024, 1024, 0.03)),
                      ((32, 1, 2.0), (64, 700, 1.7))):
        cfg = make_cfg(E=E, A=A, M=M, ogm=ogm, dagm=dagm, route_cap=16)
        eng, ora = engine(rm, cfg), oracle_world(rm, cfg, threads=8)
        for w in (eng, ora):
            w.reset(x, y, h, v, route, rlen)
        eng.set_social(torch.from_numpy(soc).cuda()); ora.set_social(soc)
        for t in range(2):
            eng.step(torch.from_numpy(act).cuda()); ora.step(act)
        eng.observe(); ora.observe()
        got, want = eng.obs_numpy(), ora.obs()
        assert_same(got, want)
        assert want["dagm"].any()
        eng.close(); ora.close()


def test_grid_widths_that_are_not_multiples_of_32():
    """Any width from 1 to 1024 (AgentInterface takes what the user asks for, agent_interface.py:62-79): the bit image
    is padded to whole words, the byte image is cropped -- bit-exact against the oracle, on the device views and
    through the host-buffer call (such a grid travels as plain bytes unless W * H is a multiple of 32)."""
    import torch
    rm = load_map("4lane")
    E, A, M = 3, 4, 20
    x, y, h, v, route, rlen = scenario_gen.agent_starts(rm, E, MISSIONS_4LANE, 16, 81, jitter=40.0, speed=7.0)
    soc = scenario_gen.SocialTraffic(rm, E, M, 82).states(2)
    act = np.zeros((E, A), np.int32)
    for ogm, dagm in (((100, 75, 0.3), (50, 33, 0.6)), ((1, 5, 3.0), (333, 64, 0.15)), ((257, 300, 0.1), (96, 31, 0.4))):
        cfg = make_cfg(E=E, A=A, M=M, ogm=ogm, dagm=dagm, route_cap=16)
        eng, ora = engine(rm, cfg), oracle_world(rm, cfg, threads=8)
        for w in (eng, ora):
            w.reset(x, y, h, v, route, rlen)
        ora.set_social(soc); ora.step(act); ora.observe()
        bufs = eng.host_obs_buffers()
        eng.step_host(act, soc)
        eng.copy_obs_to_host(bufs)
        got, want = eng.obs_numpy(), ora.obs()
        assert_same(got, want)
        assert_same({k: t.numpy() for k, t in bufs.items()}, want)
        assert want["dagm"].any() and tuple(got["ogm"].shape[-2:]) == (ogm[1], ogm[0])
        eng.close(); ora.close()


def test_captured_tick_replays_bit_identically():
    """Engine.capture_tick: the tick as a CUDA graph over fixed action / social tensors gives the states and
    observations of the eager calls, tick for tick."""
    import torch
    rm = load_map("4lane")
    E, A, M = 33, 4, 9
    cfg = make_cfg(E=E, A=A, M=M, nbr=True, wp=True, lidar=True, wall=1.5, ogm=(64, 64, 0.5), dagm=(96, 72, 0.3),
                   route_cap=16, events=True)
    x, y, h, v, route, rlen = scenario_gen.agent_starts(rm, E, MISSIONS_4LANE, cfg.route_cap, 71, jitter=30.0, speed=6.0)
    traffic = scenario_gen.SocialTraffic(rm, E, M, 72)
    acts = np.random.default_rng(73).integers(0, 4, (8, E, A)).astype(np.int32)
    eager, graphed = engine(rm, cfg), engine(rm, cfg)
    for w in (eager, graphed):
        w.reset(x, y, h, v, route, rlen)
    act_buf = torch.zeros((E, A), dtype=torch.int32, device="cuda")
    soc_buf = torch.zeros((E, M, 8), dtype=torch.float64, device="cuda")
    tick = None
    for t in range(8):
        a, s = torch.from_numpy(acts[t]).cuda(), torch.from_numpy(traffic.states(t)).cuda()
        eager.set_social(s); eager.step(a); eager.observe()
        act_buf.copy_(a); soc_buf.copy_(s)
        if t == 0:      # modules loaded by an eager tick, then the capture
            graphed.set_social(soc_buf); graphed.step(act_buf); graphed.observe()
            tick = graphed.capture_tick(act_buf, soc_buf)
            assert tick.launches >= 5
        else:
            n0 = graphed.launch_count
            tick.replay()
            assert graphed.launch_count - n0 == tick.launches
        torch.cuda.synchronize()
        assert_same(graphed.obs_numpy(), eager.obs_numpy())
        assert_same(graphed.state_numpy(), eager.state_numpy())


def test_lane_action_strings_use_the_reference_dispatch_table():
    """Controllers.perform_action for ActionSpaceType.Lane (/root/reference/smarts/core/controllers/__init__.py:125-144):
    keep_lane -> (target_speed 15, lane_change 0), slow_down -> (0, 0), change_lane_left -> (12.5, +1),
    change_lane_right -> (12.5, -1).  The Lane action space must move the vehicles exactly as
    LaneWithContinuousSpeed does with those tuples (same controller behind both, :100-124)."""
    import torch
    from smarts_b200.agent_interface import ActionSpaceType
    from smarts_b200.hiway_env import encode_actions
    table = {"keep_lane": (15, 0), "slow_down": (0, 0), "change_lane_left": (12.5, 1), "change_lane_right": (12.5, -1)}
    rm = load_map("4lane")
    E, A = 6, 4
    x, y, h, v, route, rlen = scenario_gen.agent_starts(rm, E, MISSIONS_4LANE, 16, 3, jitter=20.0, speed=5.0)
    names = list(table)
    rng = np.random.default_rng(5)
    engs = [engine(rm, make_cfg(E=E, A=A, action_space=sp, route_cap=16))
            for sp in (_abi.SB_ACTION_LANE, _abi.SB_ACTION_LANE_WITH_CONTINUOUS_SPEED)]
    for g in engs:
        g.reset(x, y, h, v, route, rlen)
    for t in range(25):
        pick = [names[i] for i in rng.integers(0, 4, E * A)]
        a_lane = encode_actions(ActionSpaceType.Lane, pick, 10).reshape(E, A)
        a_speed = encode_actions(ActionSpaceType.LaneWithContinuousSpeed, [table[p] for p in pick], 10).reshape(E, A, 2)
        engs[0].step(torch.from_numpy(a_lane).cuda())
        engs[1].step(torch.from_numpy(a_speed).cuda())
    s0, s1 = engs[0].state_numpy(), engs[1].state_numpy()
    assert_same(s0, s1, keys=("pos_x", "pos_y", "heading", "speed", "steering", "last_control"))
    assert np.ptp(s0["speed"][:, :A]) > 1.0          # the four commands did drive the agents apart
    for g in engs:
        g.close()


def test_headline_batch_at_full_size_matches_oracle():
    """BASELINE configs[2] at its full size -- 4096 envs x 4 agents, 20 social vehicles, LiDAR 300 rays (road-edge
    walls on) + OGM + DAGM 256 x 256 @ 0.1 m/px -- three ticks, every state and observation tensor bit for bit against
    the oracle (16 384 agents, 2.27 GB of observations per tick; the oracle takes about half a second per tick on the
    box's host cores).  Tick 2 is compared after the agents have moved off their reset poses."""
    import os
    import torch
    rm = load_map("4lane")
    E, A, M = 4096, 4, 20
    cfg = make_cfg(E=E, A=A, M=M, lidar=True, wall=1.5, ogm=(256, 256, 0.1), dagm=(256, 256, 0.1), route_cap=16)
    x, y, h, v, route, rlen = scenario_gen.agent_starts(rm, E, MISSIONS_4LANE, cfg.route_cap, 11, jitter=25.0, speed=8.0)
    traffic = scenario_gen.SocialTraffic(rm, E, M, 12)
    eng = engine(rm, cfg)
    ora = oracle_world(rm, cfg, threads=os.cpu_count() or 8)
    eng.reset(x, y, h, v, route, rlen)
    ora.reset(x, y, h, v, route, rlen)
    act = np.zeros((E, A), np.int32)
    for t in range(3):
        soc = traffic.states(t)
        eng.set_social(torch.from_numpy(soc).cuda())
        ora.set_social(soc)
        eng.step(torch.from_numpy(act).cuda())
        ora.step(act)
    eng.observe()
    ora.observe()
    assert_same(eng.state_numpy(), ora.state())
    want = ora.obs()
    obs = eng.obs()                                     # device tensors: compared key by key to bound host memory
    for k, wv in want.items():
        got = obs[k].cpu().numpy()
        assert_same({k: got}, {k: wv})
        del got
    assert want["ogm"].any() and want["dagm"].any() and want["lidar_hits"].any()
    eng.close()


def test_host_buffer_copy_orders_itself_after_a_side_stream():
    """sb_step / sb_observe queued on a non-blocking side stream, then sb_copy_obs_to_host (which copies on the legacy
    stream): the copy must wait for the observation kernels of the side stream.  A long chain of ticks is queued first
    so that the kernels are still running when the copy call is made."""
    import torch
    rm = load_map("4lane")
    E, A, M = 64, 4, 20
    cfg = make_cfg(E=E, A=A, M=M, lidar=True, wall=1.5, ogm=(256, 256, 0.1), dagm=(256, 256, 0.1), route_cap=16)
    x, y, h, v, route, rlen = scenario_gen.agent_starts(rm, E, MISSIONS_4LANE, cfg.route_cap, 5, jitter=20.0, speed=6.0)
    traffic = scenario_gen.SocialTraffic(rm, E, M, 6)
    eng, ora = engine(rm, cfg), oracle_world(rm, cfg, threads=8)
    eng.reset(x, y, h, v, route, rlen)
    ora.reset(x, y, h, v, route, rlen)
    act = np.zeros((E, A), np.int32)
    side = torch.cuda.Stream()
    socs = [torch.from_numpy(traffic.states(t)).cuda() for t in range(12)]
    acts = torch.from_numpy(act).cuda()
    torch.cuda.synchronize()
    with torch.cuda.stream(side):
        for t in range(12):
            eng.set_social(socs[t])
            eng.step(acts)
            eng.observe()
    bufs = eng.host_obs_buffers(pinned=True)
    eng.copy_obs_to_host(bufs)                      # no torch-side synchronisation in between
    for t in range(12):
        ora.set_social(traffic.states(t))
        ora.step(act)
        ora.observe()
    want = ora.obs()
    assert_same({k: bufs[k].numpy() for k in want}, want)
    eng.close()


def test_host_buffer_copy_after_the_callers_stream_is_gone():
    """The library keeps no handle of the caller's stream: ticks are queued on a raw non-blocking stream through the
    C-ABI, the stream is destroyed at once (CUDA lets its queued work finish), and sb_copy_obs_to_host still
    delivers the last tick's observations -- it waits on an event the handle owns, recorded behind those kernels."""
    import ctypes as C
    import torch
    try:
        from cuda.bindings import runtime as cudart
    except ImportError:
        cudart = pytest.importorskip("cuda.cudart")
    rm = load_map("4lane")
    E, A, M = 64, 4, 20
    cfg = make_cfg(E=E, A=A, M=M, lidar=True, wall=1.5, ogm=(256, 256, 0.1), dagm=(256, 256, 0.1), route_cap=16)
    x, y, h, v, route, rlen = scenario_gen.agent_starts(rm, E, MISSIONS_4LANE, cfg.route_cap, 7, jitter=20.0, speed=6.0)
    traffic = scenario_gen.SocialTraffic(rm, E, M, 8)
    eng, ora = engine(rm, cfg), oracle_world(rm, cfg, threads=8)
    eng.reset(x, y, h, v, route, rlen)
    ora.reset(x, y, h, v, route, rlen)
    act = np.zeros((E, A), np.int32)
    socs = [torch.from_numpy(traffic.states(t)).cuda() for t in range(12)]
    acts = torch.from_numpy(act).cuda()
    torch.cuda.synchronize()
    err, raw = cudart.cudaStreamCreateWithFlags(cudart.cudaStreamNonBlocking)
    assert int(err) == 0
    st = C.c_void_p(int(raw))
    lib, hdl = eng._lib, eng._h
    for t in range(12):
        assert lib.sb_set_social(hdl, C.c_void_p(socs[t].data_ptr()), st) == 0
        assert lib.sb_step(hdl, C.c_void_p(acts.data_ptr()), st) == 0
        assert lib.sb_observe(hdl, st) == 0
    assert int(cudart.cudaStreamDestroy(raw)[0]) == 0
    bufs = eng.host_obs_buffers(pinned=True)
    eng.copy_obs_to_host(bufs)
    for t in range(12):
        ora.set_social(traffic.states(t))
        ora.step(act)
        ora.observe()
    want = ora.obs()
    assert_same({k: bufs[k].numpy() for k in want}, want)
    eng.close()


@pytest.mark.parametrize("A,cap", [(6, None), (5, "8"), (1, None)])
def test_lidar_tile_remainders_and_list_overflow_fallback(A, cap, monkeypatch):
    """k_lidar_tiles casts for four agents of an env per block: A = 6 leaves a tile of two, A = 5 a tile of one; with the
    block's interval list capped at 8 entries (SB_LIDAR_LIST_CAP, read at sb_create) every tile overflows and falls back
    to the one-agent chunked cast; A = 1 takes the one-agent kernel outright.  All bit-identical to the oracle."""
    if cap is not None:
        monkeypatch.setenv("SB_DEBUG_HOOKS", "1")
        monkeypatch.setenv("SB_LIDAR_LIST_CAP", cap)
    rm = load_map("4lane")
    E, M = 5, 18
    missions = (MISSIONS_4LANE + MISSIONS_4LANE)[:A]
    cfg = make_cfg(E=E, A=A, M=M, lidar=True, lidar_miss_inf=1, wall=1.5, route_cap=16)
    eng, ora = run_pair(rm, cfg, missions, ticks=16, seed=41 + A, action_fn=lane_actions(E, A, 3), check_every=4, jitter=30.0)
    assert 0 < ora.obs()["lidar_hits"].mean() < 1
    eng.close()


@pytest.mark.gpu
def test_observing_the_envs_in_two_halves_ticks_every_env_once():
    """sb_step, then sb_observe_envs on one half of the envs and on the other half: the tick belongs to each agent,
    not to the handle, so the second half's step counters advance and its dones are raised exactly as in one
    sb_observe over all envs (the first partial observe must not use up the others' tick; SensorState.step,
    sensors.py:609-637).  A third observe without a step in between is the `sim.resetting` kind for everyone."""
    import torch
    rm = load_map("4lane")
    E, A, M = 6, 4, 6
    cfg = make_cfg(E=E, A=A, M=M, nbr=True, wp=True, route_cap=16, events=True, max_episode_steps=3)
    x, y, h, v, route, rlen = scenario_gen.agent_starts(rm, E, MISSIONS_4LANE, cfg.route_cap, 91, jitter=30.0, speed=6.0)
    traffic = scenario_gen.SocialTraffic(rm, E, M, 92)
    first, second = np.array([0, 2, 5], np.int32), np.array([1, 3, 4], np.int32)
    split, whole, ora = engine(rm, cfg), engine(rm, cfg), oracle_world(rm, cfg, threads=4)
    for w in (split, whole, ora):
        w.reset(x, y, h, v, route, rlen)
        w.observe()
    act = np.zeros((E, A), np.int32)
    for t in range(3):
        soc = traffic.states(t)
        for w in (split, whole):
            w.set_social(torch.from_numpy(soc).cuda())
            w.step(torch.from_numpy(act).cuda())
        ora.set_social(soc); ora.step(act)
        whole.observe()
        for w in (split, ora):
            w.observe(first)
            w.observe(second)
        got = split.obs_numpy()
        assert_same(got, whole.obs_numpy())
        assert_same(got, ora.obs())
        assert (got["ev_steps"] == t + 1).all()
    assert got["ev_done"].all() and (got["ev_events"] & _abi.SB_EV_REACHED_MAX_STEPS).all()
    for w in (split, ora):
        w.observe(second)  # no step since: nothing advances, no done is raised
    got = split.obs_numpy()
    assert_same(got, ora.obs())
    assert (got["ev_steps"] == 3).all() and not got["ev_done"][second].any() and got["ev_done"][first].all()
    for w in (split, whole):
        w.close()


@pytest.mark.gpu
def test_lidar_debug_hooks_are_gated_and_validated(monkeypatch):
    """SB_LIDAR_* only count together with SB_DEBUG_HOOKS=1, and then garbage fails sb_create with the variable named
    in the message instead of quietly selecting the per-agent LiDAR fallback."""
    from smarts_b200.engine import SmartsB200Error
    rm = load_map("4lane")
    cfg = make_cfg(E=2, A=4, M=4, lidar=True, wall=1.5, route_cap=16)
    monkeypatch.setenv("SB_LIDAR_LIST_CAP", "zero")
    engine(rm, cfg).close()  # ignored: the hooks are off
    monkeypatch.setenv("SB_DEBUG_HOOKS", "1")
    with pytest.raises(SmartsB200Error, match="SB_LIDAR_LIST_CAP"):
        engine(rm, cfg)
    monkeypatch.setenv("SB_LIDAR_LIST_CAP", "8")
    monkeypatch.setenv("SB_LIDAR_TILE_MIN_A", "-3")
    with pytest.raises(SmartsB200Error, match="SB_LIDAR_TILE_MIN_A"):
        engine(rm, cfg)
