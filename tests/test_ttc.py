"""lane_ttc (StdObs `ttc`, smarts/env/custom_observations.py:148-290): the oracle against golden outputs of the
reference's own function (tests/golden/gen_ttc_golden.py -> ttc.npz), and (gpu) the CUDA kernel against the oracle."""
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


def _cfg(E, A, M, **kw):
    return make_cfg(E=E, A=A, M=M, nbr=True, nbr_radius=60.0, nbr_cap=A + M, wp=True, wp_lookahead=32, wp_paths=16,
                    wp_len=33, route_cap=24, ttc=True, **kw)


@pytest.mark.parametrize("name", sorted(CASES))
def test_oracle_ttc_matches_reference_lane_ttc(name):
    z = np.load(os.path.join(GOLDEN, "ttc.npz"))
    rm = load_map(name)
    E, A, M = 6, 4, 20
    cfg = _cfg(E, A, M)
    x, y, h, v, route, rlen = scenario_gen.agent_starts(rm, E, CASES[name], 24, 61, jitter=30.0, speed=9.0)
    tr = scenario_gen.SocialTraffic(rm, E, M, 62)
    ora = oracle_world(rm, cfg, threads=4)
    ora.reset(x, y, h, v, route, rlen)
    got, k = [], 0
    for t in range(12):
        ora.set_social(tr.states(t))
        ora.step(np.zeros((E, A), np.int32))
        if t % 4 != 3:
            continue
        ora.observe()
        o = ora.obs()
        for key in ("wp_pos", "wp_lane", "nbr_pos", "nbr_lane", "nbr_count"):  # same inputs as the golden run
            assert np.array_equal(o[key], z[f"{name}_{key}"][k]), key
        k += 1
        got.append(np.concatenate([o["ttc_valid"][..., None].astype(np.float64), o["ttc_angle_error"][..., None],
                                   o["ttc_distance_from_center"][..., None], o["ttc_dtc"], o["ttc_ttc"]], axis=-1).reshape(-1, 9))
        assert not (ora.state()["status"] & _abi.SB_STATUS_TTC_TRUNCATED).any()  # nothing was capped in this set-up
    got, want = np.concatenate(got), z[name + "_ref"]
    assert got.shape == want.shape and np.isfinite(want).all()
    assert np.array_equal(got[:, 0], want[:, 0])
    assert np.allclose(got[:, 1:], want[:, 1:], rtol=1e-5, atol=1e-6)  # float32 outputs of binary64 arithmetic
    hit = (want[:, 0] == 1) & ((want[:, 6:9] > 0) & (want[:, 6:9] < 1000)).any(axis=1)
    assert hit.sum() >= 10, "golden set holds too few real time-to-collision values"


@pytest.mark.gpu
@pytest.mark.parametrize("capped", [False, True])
def test_gpu_ttc_matches_oracle(capped):
    import torch
    from smarts_b200.engine import Engine
    from smarts_b200 import lidar_params
    rm = load_map("figure_eight")
    E, A, M = 16, 4, 20
    cfg = _cfg(E, A, M) if not capped else make_cfg(E=E, A=A, M=M, nbr=True, nbr_radius=60.0, nbr_cap=3, wp=True,
                                                     wp_lookahead=32, wp_paths=2, wp_len=20, route_cap=24, ttc=True)
    x, y, h, v, route, rlen = scenario_gen.agent_starts(rm, E, CASES["figure_eight"], 24, 71, jitter=30.0, speed=9.0)
    tr = scenario_gen.SocialTraffic(rm, E, M, 72)
    eng = Engine(cfg, rm, lidar_params.base_rays(lidar_params.BasicLidar))
    ora = oracle_world(rm, cfg, threads=8)
    for w in (eng, ora):
        w.reset(x, y, h, v, route, rlen)
    rng = np.random.default_rng(73)
    for t in range(20):
        soc = tr.states(t)
        act = rng.integers(0, 4, (E, A)).astype(np.int32)
        eng.set_social(torch.from_numpy(soc).cuda()); ora.set_social(soc)
        eng.step(torch.from_numpy(act).cuda()); ora.step(act)
        eng.observe(); ora.observe()
        got, want = eng.obs_numpy(), ora.obs()
        for k in want:
            assert np.array_equal(got[k].view(np.uint8), want[k].view(np.uint8)), (t, k)
        assert np.array_equal(eng.state_numpy()["status"], ora.state()["status"])
    st = ora.state()["status"]
    assert bool((st & _abi.SB_STATUS_TTC_TRUNCATED).any()) == capped
    assert want["ttc_valid"].any() and ((want["ttc_ttc"] > 0) & (want["ttc_ttc"] < 1000)).any()
