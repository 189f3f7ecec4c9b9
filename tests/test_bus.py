"""vehicle_type="bus" (AgentInterface.vehicle_type, /root/reference/smarts/core/agent_interface.py:246-250;
vehicle.py:393-406: models/bus.urdf + the `bus` block of models/controller_parameters.yaml).  The oracle's controllers
with the bus record of include/sb_vehicle.h against tests/golden/controllers_bus.npz -- outputs of the reference's own
LaneFollowingController / TrajectoryTrackingController / place_poles run with the bus's length, masses and gains
(tests/golden/gen_bus_golden.py) -- and the CUDA path bit for bit against the oracle.  The bus's single-track chassis
 has no reference envelope test: its dynamics are the sedan's equations with bus.urdf's constants."""
import os

import numpy as np
import pytest

from conftest import GOLDEN, load_map, make_cfg, oracle_world
from oracle import oracle_py
from smarts_b200 import _abi

BUS = _abi.SB_VEHICLE_BUS


def test_bus_lateral_gains_match_place_poles():
    g = np.load(os.path.join(GOLDEN, "controllers_bus.npz"))["gains"]
    for v, k0, k1, hg, lg in g:
        h, l = oracle_py.lateral_gains(float(v), BUS)
        assert abs(h - hg) < 1e-9 and abs(l - lg) < 1e-9
    sedan = [oracle_py.lateral_gains(float(v)) for v in g[:, 0]]
    assert any(abs(a[1] - b) > 1e-3 for a, b in zip(sedan, g[:, 4]))     # not the sedan's gains under another name


def test_bus_lane_following_matches_reference():
    rec = np.load(os.path.join(GOLDEN, "controllers_bus.npz"))["lane_4lane"]
    rm = load_map("4lane")
    q = np.load(os.path.join(GOLDEN, "queries_4lane.npz"))
    routes = [[rm.road_index[r] for r in s.split(";")] for s in q["routes"] if s]
    w = oracle_world(rm, make_cfg(E=1, A=1, M=0, vehicle_type=BUS))
    for r in rec:
        ridx, ts, lc = int(r[0]), float(r[1]), int(r[2])
        st = np.zeros(12)
        st[:8] = r[10:18]
        out, st2, status = w.lane_following(r[3:10], routes[ridx], ts, lc, 0.1, st)
        assert status == 0
        np.testing.assert_allclose(st2[:5], r[18:23], rtol=1e-9, atol=1e-10)
        assert st2[7] == r[25]
        exp_brake = 0.0 if (r[27] > 0 and r[8] < 1 / 36) else r[27]
        np.testing.assert_allclose(out, [r[26], exp_brake, r[28]], rtol=1e-9, atol=1e-10)


def test_bus_trajectory_tracking_matches_reference():
    g = np.load(os.path.join(GOLDEN, "controllers_bus.npz"))
    differs = 0
    for r, traj in zip(g["traj_rec"], g["trajs"]):
        st = np.zeros(12)
        st[:7] = r[7:14]
        out, st2 = oracle_py.trajectory_tracking(r[0:7], traj, 0.1, st.copy(), BUS)
        np.testing.assert_allclose(st2[:7], r[14:21], rtol=1e-9, atol=1e-10)
        exp_brake = 0.0 if (r[22] > 0 and r[5] < 1 / 36) else r[22]
        np.testing.assert_allclose(out, [r[21], exp_brake, r[23]], rtol=1e-9, atol=1e-10)
        out_sedan, _ = oracle_py.trajectory_tracking(r[0:7], traj, 0.1, st.copy())
        differs += not np.allclose(out, out_sedan)
    assert differs > 50                                                   # the bus block really is in use


def test_bus_box_and_config():
    from smarts_b200.agent_interface import ActionSpaceType, AgentInterface, compile_config
    c = compile_config(AgentInterface(action=ActionSpaceType.Lane, vehicle_type="bus"), 1, 1, 0, 0.1)
    assert c.vehicle_type == BUS
    rm = load_map("4lane")
    w = oracle_world(rm, make_cfg(E=1, A=1, vehicle_type=BUS))
    w.reset(np.array([[150.0]]), np.array([[-80.0]]), np.zeros((1, 1)), np.zeros((1, 1)))
    s = w.state()
    assert (s["length"][0, 0], s["width"][0, 0], s["height"][0, 0]) == (7.0, 2.2026, 2.4)    # bus.urdf:13-15


@pytest.mark.gpu
@pytest.mark.parametrize("space", ["lane", "trajectory"])
def test_gpu_bus_rollout_matches_oracle(space):
    """Bus agents among social traffic with every sensor that looks at vehicle boxes (neighbours, LiDAR, OGM, events'
    contacts): CUDA == oracle bit for bit, Lane and Trajectory action spaces."""
    from test_gpu_parity import MISSIONS_4LANE, lane_actions, run_pair
    rm = load_map("4lane")
    E, A, M = 6, 4, 12
    if space == "lane":
        cfg = make_cfg(E=E, A=A, M=M, vehicle_type=BUS, nbr=True, wp=True, lidar=True, wall=1.5, ogm=(64, 64, 0.5),
                       events=True, route_cap=16)
        eng, ora = run_pair(rm, cfg, MISSIONS_4LANE, ticks=40, seed=31, action_fn=lane_actions(E, A, 4), check_every=5)
    else:
        N = 10
        cfg = make_cfg(E=E, A=A, M=M, vehicle_type=BUS, action_space=_abi.SB_ACTION_TRAJECTORY, traj_len=N, wp=True,
                       wp_len=20, route_cap=16)

        def fn(t, ora):
            o = ora.obs()
            a = np.zeros((E, A, 4, N))
            n = np.clip(o["wp_len"][:, :, 0], 1, N)
            idx = np.minimum(np.arange(N)[None, None, :], (n - 1)[:, :, None])
            a[:, :, 0] = np.take_along_axis(o["wp_pos"][:, :, 0, :, 0], idx, 2)
            a[:, :, 1] = np.take_along_axis(o["wp_pos"][:, :, 0, :, 1], idx, 2)
            a[:, :, 2] = np.take_along_axis(o["wp_heading"][:, :, 0, :], idx, 2)
            a[:, :, 3] = 30 / 3.6
            return a
        eng, ora = run_pair(rm, cfg, MISSIONS_4LANE, ticks=40, seed=32, action_fn=fn, check_every=5)
    s = ora.state()
    assert (s["length"][:, :A] == 7.0).all() and np.ptp(s["pos_x"][:, :A]) > 1.0
    eng.close()
