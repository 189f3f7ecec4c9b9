"""BoxChassis: the oracle's social-vehicle update against the reference's own class
(tests/golden/gen_boxchassis_golden.py -> boxchassis.npz): heading wrap, yaw rate from the signed heading difference,
0 where the reference reports None (a chassis that was just created), velocity vectors."""
import os

import numpy as np

from conftest import GOLDEN, load_map, make_cfg, oracle_world

DT = 0.1
E, A, M, T = 3, 1, 8, 14


def social_recording():
    """[T, E, M, 8] injected states: headings that drift, jump and cross +-pi (some given outside (-pi, pi], as a
    recording in SUMO's convention converted without wrapping would be), speeds that change, and vehicles that leave
    and re-enter."""
    rng = np.random.default_rng(5)
    s = np.zeros((T, E, M, 8))
    h0 = rng.uniform(-np.pi, np.pi, (E, M))
    rate = rng.normal(0, 0.8, (E, M))
    h0[0, 0], rate[0, 0] = np.pi - 0.05, 0.4          # crosses +pi
    h0[0, 1], rate[0, 1] = -np.pi + 0.05, -0.4        # crosses -pi
    h0[1, 0], rate[1, 0] = 5.0, 0.3                   # outside the wrapped range from the start
    for t in range(T):
        s[t, ..., 0] = 150.0 + 3.0 * t + rng.normal(0, 0.1, (E, M))
        s[t, ..., 1] = -60.0 + rng.normal(0, 0.1, (E, M))
        s[t, ..., 2] = h0 + rate * DT * t + (rng.random((E, M)) < 0.1) * rng.normal(0, 1.5, (E, M))
        s[t, ..., 3] = np.abs(8.0 + rng.normal(0, 2.0, (E, M)))
        s[t, ..., 4:7] = (3.68, 1.47, 1.4)
        s[t, ..., 7] = 1.0
    s[4:7, 0, 2, 7] = 0.0      # leaves for three ticks, then re-enters
    s[0:3, 1, 3, 7] = 0.0      # enters late
    s[9, 2, 4, 7] = 0.0        # one tick away
    return s


def test_oracle_social_update_matches_reference_boxchassis():
    z = np.load(os.path.join(GOLDEN, "boxchassis.npz"))
    rec = social_recording()
    rm = load_map("4lane")
    ora = oracle_world(rm, make_cfg(E=E, A=A, M=M, dt=DT))
    x, y, h = (np.array([[v]] * E) for v in rm.lane_start_pose("edge-south-SN", 0, 10.0))
    ora.reset(x, y, h, np.zeros((E, A)))
    for t in range(T):
        ora.set_social(rec[t])
        st = ora.state()
        alive = rec[t, ..., 7] != 0
        assert np.array_equal(st["alive"][:, A:] != 0, alive)
        for key, tol in (("heading", 0.0), ("speed", 0.0), ("yaw_rate", 1e-12)):
            got, want = st[key][:, A:][alive], z[key][t][alive]
            assert np.allclose(got, want, rtol=0, atol=tol), (t, key, np.abs(got - want).max())
        # yaw_rate is None in the reference right after creation (chassis.py:302-308): 0 here
        assert (st["yaw_rate"][:, A:][alive & (z["has_rate"][t] == 0)] == 0).all()
        # linear velocity = heading vector * speed (chassis.py:275-288); the heading vector is (-sin h, cos h)
        lin = np.stack([-st["sin_h"][:, A:] * st["speed"][:, A:], st["cos_h"][:, A:] * st["speed"][:, A:]], -1)
        assert np.allclose(lin[alive], z["linear_velocity"][t][alive][:, :2], rtol=0, atol=1e-12)
    assert np.abs(z["yaw_rate"]).max() > 1.0 and (z["has_rate"] == 0).sum() > E * M
