"""Configuration-space fuzz: random handle shapes and sensor subsets, CUDA == oracle bit for bit.

The hand-written parity tests pin the BASELINE shapes; this one walks the space between them -- env / agent / social
counts from 1 upwards (one agent per env, agent counts that leave partial lidar tiles, no social vehicles at all),
every map, every action space, grids of odd sizes on both sides of the 128-row switch between k_raster_slices and
k_raster, each sensor on or off, partial observes (sb_observe_envs) and partial resets in the middle of the run.  Every
case is seeded; a failure names its case number.  (The one-agent-per-env index bug of round 1 is the kind of thing this
catches: no BASELINE-shaped test had A = 1 on the GPU.)"""
import numpy as np
import pytest

from conftest import load_map, make_cfg, oracle_world
from smarts_b200 import _abi, lidar_params, scenario_gen

pytestmark = pytest.mark.gpu

MAPS = ["4lane", "loop", "figure_eight", "straight", "6lane"]
SPACES = [_abi.SB_ACTION_LANE, _abi.SB_ACTION_LANE_WITH_CONTINUOUS_SPEED, _abi.SB_ACTION_TRAJECTORY,
          _abi.SB_ACTION_CONTINUOUS, _abi.SB_ACTION_ACTUATOR_DYNAMIC]


def draw_case(i):
    rng = np.random.default_rng(9000 + i)
    name = MAPS[i % len(MAPS)]
    E = int(rng.integers(1, 8))
    A = int(rng.choice([1, 1, 2, 3, 4, 5, 6, 7, 9]))
    M = int(rng.choice([0, 0, 1, 5, 12, 30]))
    space = SPACES[int(rng.integers(0, len(SPACES)))]
    flag = lambda p=0.5: bool(rng.uniform() < p)
    nbr, wp = flag(), flag()
    if space == _abi.SB_ACTION_TRAJECTORY:
        wp = True                                     # the policy below follows the waypoint observation

    def grid():
        if not flag(0.6):
            return None
        return (int(rng.choice([8, 31, 33, 64, 100, 160, 257])), int(rng.choice([5, 32, 64, 96, 128, 129, 200])),
                float(rng.choice([0.1, 0.25, 0.5, 1.0])))
    kw = dict(E=E, A=A, M=M, action_space=space, traj_len=int(rng.integers(2, 13)), route_cap=24,
              nbr=nbr, nbr_radius=float(rng.choice([-1.0, 10.0, 50.0])), nbr_cap=int(rng.choice([1, 4, 10])),
              wp=wp, wp_lookahead=int(rng.choice([1, 8, 32, 33, 64])), wp_paths=int(rng.choice([1, 4, 6])),
              wp_len=int(rng.choice([1, 20, 65])),
              lidar=flag(), lidar_miss_inf=int(flag()), wall=float(rng.choice([0.0, 1.5])), ogm=grid(), dagm=grid(),
              events=flag(), max_episode_steps=int(rng.choice([0, 3])), ttc=nbr and wp and flag(), accel=flag(),
              ego_frame=flag(0.3), rw=flag(0.3), rw_horizon=int(rng.choice([5, 20, 32])),
              rw_lanes=int(rng.choice([2, 32])), rw_paths=int(rng.choice([3, 64])),
              vehicle_type=int(flag(0.25)))
    return name, kw, rng


def actions_for(space, E, A, N, rng, ora):
    if space == _abi.SB_ACTION_LANE:
        return rng.integers(0, 4, (E, A)).astype(np.int32)
    if space == _abi.SB_ACTION_LANE_WITH_CONTINUOUS_SPEED:
        a = np.zeros((E, A, 2))
        a[..., 0] = rng.uniform(0, 16, (E, A))
        a[..., 1] = rng.integers(-1, 2, (E, A))
        return a
    if space == _abi.SB_ACTION_TRAJECTORY:
        o = ora.obs()
        a = np.zeros((E, A, 4, N))
        L = o["wp_pos"].shape[3]
        n = np.clip(o["wp_len"][:, :, 0], 1, min(N, L))
        idx = np.minimum(np.arange(N)[None, None, :], (n - 1)[:, :, None])
        a[:, :, 0] = np.take_along_axis(o["wp_pos"][:, :, 0, :, 0], idx, 2)
        a[:, :, 1] = np.take_along_axis(o["wp_pos"][:, :, 0, :, 1], idx, 2)
        a[:, :, 2] = np.take_along_axis(o["wp_heading"][:, :, 0, :], idx, 2)
        a[:, :, 3] = rng.uniform(2, 15)
        return a
    a = np.zeros((E, A, 3))
    a[..., 0] = rng.uniform(0, 1, (E, A))
    a[..., 1] = rng.uniform(0, 1, (E, A)) * (rng.uniform(size=(E, A)) < 0.2)
    a[..., 2] = rng.uniform(-1, 1, (E, A))
    return a


@pytest.mark.parametrize("case", range(80))
def test_random_configuration_matches_oracle(case):
    import torch
    from smarts_b200.engine import Engine
    from test_gpu_parity import assert_same
    name, kw, rng = draw_case(case)
    rm = load_map(name)
    E, A, M = kw["E"], kw["A"], kw["M"]
    cfg = make_cfg(**kw)
    rays = lidar_params.base_rays(lidar_params.VelodyneHDL32E if case % 4 == 3 else lidar_params.BasicLidar)
    cfg.lidar_rays = len(rays)
    missions = scenario_gen.random_missions(rm, A, 77 + case, cfg.route_cap, min_len=15.0)
    x, y, h, v, route, rlen = scenario_gen.agent_starts(rm, E, missions, cfg.route_cap, case, jitter=20.0, speed=5.0)
    off = rng.uniform(size=(E, A)) < 0.2                # some agents start beside their lane, a few far off the map
    x = x + off * rng.normal(0, 3.0, (E, A))
    x[rng.uniform(size=(E, A)) < 0.03] += 500.0
    traffic = scenario_gen.SocialTraffic(rm, E, M, case + 1, p_alive=0.8) if M else None
    eng, ora = Engine(cfg, rm, rays), oracle_world(rm, cfg, threads=4, rays=rays)
    for w in (eng, ora):
        w.reset(x, y, h, v, route, rlen)
        w.observe()

    def compare(tag):
        try:
            assert_same(eng.state_numpy(), ora.state())
            assert_same(eng.obs_numpy(), ora.obs())
        except AssertionError as err:
            raise AssertionError(f"case {case} ({name}, {tag}): {err}\n{kw}") from None
    compare("reset")
    for t in range(6):
        act = actions_for(cfg.action_space, E, A, cfg.traj_len, rng, ora)
        if traffic is not None:
            soc = traffic.states(t)
            eng.set_social(torch.from_numpy(soc).cuda())
            ora.set_social(soc)
        eng.step(torch.from_numpy(act).cuda())
        ora.step(act)
        if t == 3 and E > 1:                             # a partial reset followed by a partial observe
            ids = np.sort(rng.choice(E, max(1, E // 2), replace=False)).astype(np.int32)
            for w in (eng, ora):
                w.reset(x[ids], y[ids], h[ids], v[ids], route[ids], rlen[ids], env_ids=ids)
                w.observe(env_ids=ids)
            compare(f"tick {t}, partial reset of {ids.tolist()}")
            continue
        eng.observe()
        ora.observe()
        compare(f"tick {t}")
    eng.close()
