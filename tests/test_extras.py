"""AccelerometerSensor and ViaSensor (smarts/core/sensors.py:1046-1150): the oracle against
golden outputs of the reference's own classes (tests/golden/gen_extras_golden.py -> extras.npz), and (gpu) the
CUDA kernel k_extras against the oracle, bit for bit."""
import os

import numpy as np
import pytest

from conftest import GOLDEN, load_map, make_cfg, oracle_world
from smarts_b200 import _abi, scenario_gen

CASES = {
    "4lane": [("edge-south-SN", 1, 10.0, "edge-west-EW"), ("edge-west-WE", 0, 10.0, "edge-east-WE"),
              ("edge-north-NS", 0, 10.0, "edge-south-NS"), ("edge-east-EW", 1, 10.0, "edge-west-EW")],
    "figure_eight": [("gneE20", 0, 5.0, "gneE27"), ("gneE36", 1, 5.0, "gneE25"), ("gneE20", 2, 25.0, "gneE27"),
                     ("gneE36", 0, 30.0, "gneE25")],
}
VIA_CAP = 8
TICKS = 40
RESET_AT, RESET_ENVS = 17, [1, 4]
OUT_KEYS = ("acc_linear", "acc_angular", "acc_linear_jerk", "acc_angular_jerk",
            "via_near_count", "via_near_idx", "via_hit_count", "via_hit_idx")


def extras_cfg(E, A, **kw):
    return make_cfg(E=E, A=A, M=0, route_cap=24, accel=True, via_cap=VIA_CAP, via_range=40.0, via_speed_accuracy=1.5,
                    action_space=_abi.SB_ACTION_LANE, **kw)


def make_vias(rm, x, y, seed):
    """Mission.via lists: per agent 3..VIA_CAP lane points of the map -- most ahead of the start within 45 m, a
    few anywhere on the map (beyond the 40 m acquisition range) -- with hit distances of 1-4 m and required
    speeds around the cruising speed, some far from it."""
    rng = np.random.default_rng(seed)
    E, A = x.shape
    cnt = rng.integers(3, VIA_CAP + 1, (E, A)).astype(np.int32)
    lane = np.full((E, A, VIA_CAP), -1, np.int32)
    vx = np.zeros((E, A, VIA_CAP)); vy = np.zeros_like(vx); hit = np.zeros_like(vx); spd = np.zeros_like(vx)
    for e in range(E):
        for a in range(A):
            d = np.hypot(rm.lp_x - x[e, a], rm.lp_y - y[e, a])
            close = np.flatnonzero(d < 45.0)
            for k in range(cnt[e, a]):
                i = int(rng.choice(close)) if (k % 4 != 3 and len(close)) else int(rng.integers(0, len(rm.lp_x)))
                lane[e, a, k] = rm.lp_lane[i]
                vx[e, a, k] = rm.lp_x[i] + rng.normal(0, 0.2)
                vy[e, a, k] = rm.lp_y[i] + rng.normal(0, 0.2)
                hit[e, a, k] = rng.uniform(1.0, 4.0)
                spd[e, a, k] = rng.choice([6.0, 9.0, 12.0, 30.0])
    return cnt, lane, vx, vy, hit, spd


def scenario(name, E=6, seed=81):
    rm = load_map(name)
    A = len(CASES[name])
    x, y, h, v, route, rlen = scenario_gen.agent_starts(rm, E, CASES[name], 24, seed, jitter=30.0, speed=9.0)
    vias = make_vias(rm, x, y, seed + 1)
    acts = np.random.default_rng(seed + 2).integers(0, 4, (TICKS, E, A)).astype(np.int32)
    return rm, A, (x, y, h, v, route, rlen), vias, acts


def run(world, starts, vias, acts, observe, to_dev=lambda a: a):
    """reset -> observe, then TICKS x (step, observe); RESET_ENVS restart at RESET_AT.  observe(t) is called
    after every observation (t = -1 for the reset observation)."""
    x, y, h, v, route, rlen = starts
    world.reset(x, y, h, v, route, rlen, vias=vias)
    world.observe()
    observe(-1)
    for t in range(len(acts)):
        if t == RESET_AT:
            ids = np.asarray(RESET_ENVS, np.int32)
            world.reset(x[ids], y[ids], h[ids], v[ids], route[ids], rlen[ids], env_ids=ids,
                        vias=tuple(q[ids] for q in vias))
        world.step(to_dev(acts[t]))
        world.observe()
        observe(t)


@pytest.mark.parametrize("name", sorted(CASES))
def test_oracle_extras_match_reference_sensors(name):
    z = np.load(os.path.join(GOLDEN, "extras.npz"))
    rm, A, starts, vias, acts = scenario(name)
    E = starts[0].shape[0]
    ora = oracle_world(rm, extras_cfg(E, A), threads=4)
    got = {k: [] for k in OUT_KEYS}

    def grab(t):
        o = ora.obs()
        for k in OUT_KEYS:
            got[k].append(o[k])
    run(ora, starts, vias, acts, grab)
    for k in OUT_KEYS:
        g, w = np.stack(got[k]), z[f"{name}_{k}"]
        assert g.shape == w.shape, k
        if g.dtype.kind == "f":
            # same binary64 operations, but the velocity vector's sin / cos come from sb_detmath.h here and
            # from libm there: one ulp of a 9 m/s component, divided by dt = 0.1, is 2e-14
            assert np.allclose(g, w, rtol=1e-12, atol=1e-12), (k, np.abs(g - w).max())
        else:
            assert np.array_equal(g, w), k
    assert np.abs(z[f"{name}_acc_linear"]).max() > 0.1 and np.abs(z[f"{name}_acc_angular_jerk"]).max() > 0
    assert z[f"{name}_via_hit_count"].sum() >= 3 and (z[f"{name}_via_near_count"] >= 2).any()
    assert (z[f"{name}_via_near_count"] < np.asarray(vias[0])[None]).any(), "no via was ever out of range"


def test_via_arguments_are_validated():
    rm, A, starts, vias, acts = scenario("4lane", E=2)
    cfg = extras_cfg(2, A)
    assert cfg.via_cap == VIA_CAP and _abi.obs_shapes(cfg)["via_near_idx"] == ((VIA_CAP,), np.int32)
    assert _abi.obs_shapes(cfg)["acc_linear_jerk"] == ((3,), np.float64)
    assert "via_near_idx" not in _abi.obs_shapes(make_cfg(E=2, A=A))


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(CASES))
def test_gpu_extras_match_oracle(name):
    import torch
    from smarts_b200 import lidar_params
    from smarts_b200.engine import Engine
    rm, A, starts, vias, acts = scenario(name, E=32, seed=91)
    E = starts[0].shape[0]
    cfg = extras_cfg(E, A, events=True)
    eng = Engine(cfg, rm, lidar_params.base_rays(lidar_params.BasicLidar))
    ora = oracle_world(rm, cfg, threads=8)
    seen = {"hits": 0, "near": 0}

    def compare(t):
        got, want = eng.obs_numpy(), ora.obs()
        for k in want:
            assert np.array_equal(got[k].view(np.uint8), want[k].view(np.uint8)), (t, k)
        seen["hits"] += int(want["via_hit_count"].sum())
        seen["near"] += int(want["via_near_count"].sum())
    # the two worlds advance in lock step: drive them from one loop
    class Both:
        def reset(self, *a, **k):
            eng.reset(*a, **k); ora.reset(*a, **k)

        def step(self, act):
            eng.step(torch.from_numpy(act).cuda()); ora.step(act)

        def observe(self):
            eng.observe(); ora.observe()
    run(Both(), starts, vias, acts, compare)
    assert seen["hits"] > 0 and seen["near"] > 0
    # host-buffer call carries the new members too
    bufs = eng.host_obs_buffers(pinned=False)
    eng.copy_obs_to_host(bufs)
    want = ora.obs()
    for k in OUT_KEYS:
        assert np.array_equal(bufs[k].numpy(), want[k]), k


@pytest.mark.gpu
def test_gpu_rejects_bad_vias():
    from smarts_b200 import lidar_params
    from smarts_b200.engine import Engine, SmartsB200Error as SbError
    rm, A, starts, vias, acts = scenario("4lane", E=2)
    eng = Engine(extras_cfg(2, A), rm, lidar_params.base_rays(lidar_params.BasicLidar))
    x, y, h, v, route, rlen = starts
    bad = list(vias)
    bad[1] = np.full_like(vias[1], 10 ** 6)
    with pytest.raises(SbError):
        eng.reset(x, y, h, v, route, rlen, vias=tuple(bad))
    bad = list(vias)
    bad[0] = np.full_like(vias[0], VIA_CAP + 1)
    with pytest.raises(SbError):
        eng.reset(x, y, h, v, route, rlen, vias=tuple(bad))
    eng.reset(x, y, h, v, route, rlen, vias=vias)
