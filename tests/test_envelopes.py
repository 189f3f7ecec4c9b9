"""The reference's envelope tests for the controller + chassis pair, re-run on the kinematic stand-in (SURVEY 8(c),
A.2): the pybullet multibody cannot be stepped here, so step-level parity of row a5 is unpinnable -- what
the reference itself asserts about its dynamics are these envelopes, with these thresholds:

  * smarts/core/tests/test_trajectory_controller.py:95-167 -- the PD trajectory tracker on half circles of radius
    10 / 20 m at yaw rates 0.1 / 0.15 / 0.2 rad/s, 15-point trajectories rebuilt every step: final error <= 10 m;
  * smarts/core/tests/test_controller_lane.py:39-147 -- the lane-following controller through the 90-degree turn of
    maps/turning_radius/map{55,78,90,107,128}.net.xml, mission edge-west-WE (lane 0, 10 m) -> edge-south-NS (lane 0,
    40 m), actions keep_lane (Laner) and (10, 0), (16, 0) (LanerWithSpeed): goal reached within 500 steps, speed never
    below 5 km/h once above it, mean speed > 5 m/s, lateral error to the current lane's first waypoint < 2.2 m at
    every step and < 1 m on average.

Run on the CPU oracle here and on the CUDA path under -m gpu (the same code runs both worlds)."""
import math

import numpy as np
import pytest

from conftest import load_map, make_cfg, oracle_world
from smarts_b200 import _abi

DT = 0.1


# ---------------------------------------------------------------------------------------- trajectory tracking
def build_trajectory(radius, omega, step_num, n=15):
    """test_trajectory_controller.py:107-137 (omega_1 == omega_2 there, so alph is 0)."""
    k = np.arange(n) + step_num
    a = k * omega * DT
    return np.stack([-(radius - radius * np.cos(a)), radius * np.sin(a), a, np.full(n, radius * omega)])


CIRCLES = [(r, w) for r in (10, 20) for w in (0.1, 0.15, 0.2)]      # np.arange(10, 30, 10) x np.arange(0.1, 0.21, 0.05)


def circle_errors(world, to_dev=lambda a: a, state=lambda w: w.state()):
    E = len(CIRCLES)
    z = np.zeros((E, 1))
    world.reset(z, z, z, z)       # Pose.from_center((0, 0, 0), Heading(0)), at rest
    n_steps = [int(0.5 * 3.14 / (w * DT)) for _, w in CIRCLES]
    err = np.zeros(E)
    for step in range(max(n_steps)):
        traj = np.stack([build_trajectory(r, w, step) for r, w in CIRCLES])[:, None]    # [E, 1, 4, 15]
        world.step(to_dev(np.ascontiguousarray(traj)))
        s = state(world)
        for e in range(E):
            if step == n_steps[e] - 1:
                err[e] = math.hypot(s["pos_x"][e, 0] - traj[e, 0, 0, 0], s["pos_y"][e, 0] - traj[e, 0, 1, 0])
    return err


def circle_cfg():
    return make_cfg(E=len(CIRCLES), A=1, M=0, dt=DT, action_space=_abi.SB_ACTION_TRAJECTORY, traj_len=15)


def test_trajectory_tracking_envelope_oracle():
    err = circle_errors(oracle_world(load_map("straight"), circle_cfg()))
    assert (err <= 10.0).all(), err


@pytest.mark.gpu
def test_trajectory_tracking_envelope_gpu():
    import torch
    from smarts_b200.engine import Engine
    eng = Engine(circle_cfg(), load_map("straight"))
    err = circle_errors(eng, to_dev=lambda a: torch.from_numpy(a).cuda(), state=lambda w: w.state_numpy())
    assert (err <= 10.0).all(), err
    want = circle_errors(oracle_world(load_map("straight"), circle_cfg()))
    assert np.array_equal(err, want)
    eng.close()


# ---------------------------------------------------------------------------------------- lane following
TURNS = [55, 78, 90, 107, 128]
# keep_lane is (15, 0) (controllers/__init__.py:136-137): the Lane space is run by itself, the two
# LanerWithSpeed actions share a world
LANE_RUNS = [("lane", _abi.SB_ACTION_LANE, [_abi.SB_LANE_KEEP]),
             ("speed", _abi.SB_ACTION_LANE_WITH_CONTINUOUS_SPEED, [(10.0, 0.0), (16.0, 0.0)])]


def lane_world_inputs(rm, space, actions):
    E = len(actions)
    cfg = make_cfg(E=E, A=1, M=0, dt=DT, action_space=space, wp=True, wp_lookahead=32, wp_paths=4, wp_len=33,
                   route_cap=8, events=True, done_mask=0, max_episode_steps=5000)
    route = rm.generate_route("edge-west-WE", "edge-south-NS")
    x, y, h = rm.lane_start_pose("edge-west-WE", 0, 10.0)
    gx, gy, _ = rm.lane_start_pose("edge-south-NS", 0, 40.0)
    r = np.full((E, 1, 8), -1, np.int32)
    r[:, 0, :len(route.road_indices)] = route.road_indices
    starts = (np.full((E, 1), x), np.full((E, 1), y), np.full((E, 1), h), np.zeros((E, 1)), r,
              np.full((E, 1), len(route.road_indices), np.int32))
    goal = (np.full((E, 1), gx), np.full((E, 1), gy), np.full((E, 1), 2.0))     # scenario.py:691 radius = 2
    if space == _abi.SB_ACTION_LANE:
        act = np.asarray(actions, np.int32).reshape(E, 1)
    else:
        act = np.asarray(actions, np.float64).reshape(E, 1, 2)
    return cfg, starts, goal, act


def lane_envelope(world, starts, goal, act, to_dev=lambda a: a, obs=lambda w: w.obs(), state=lambda w: w.state()):
    """test_controller_lane.py:100-147, per env; returns the per-env record of the asserted quantities."""
    E = act.shape[0]
    world.reset(*starts, goal=goal)
    world.observe()
    detected = np.zeros(E, bool); reached = np.zeros(E, bool)
    speed = [[] for _ in range(E)]; lateral = [[] for _ in range(E)]
    for _ in range(500):
        o, s = obs(world), state(world)
        for e in range(E):
            if reached[e]:
                continue
            px, py, sp = s["pos_x"][e, 0], s["pos_y"][e, 0], s["speed"][e, 0]
            if sp > 5 / 3.6:
                detected[e] = True
            if detected[e]:
                speed[e].append(sp)
            n = min(int(o["wp_n_paths"][e, 0]), 4)
            first = o["wp_pos"][e, 0, :n, 0, :2]
            cur = int(np.argmin(np.hypot(first[:, 0] - px, first[:, 1] - py)))      # find_current_lane
            # Waypoint.signed_lateral_error (road_map.py:590-606): signed distance to the line through the waypoint
            # along its heading
            wh = float(o["wp_heading"][e, 0, cur, 0])
            dx, dy = -math.sin(wh), math.cos(wh)
            lateral[e].append(abs((px - first[cur, 0]) * dy - (py - first[cur, 1]) * dx))
            if o["ev_events"][e, 0] & _abi.SB_EV_REACHED_GOAL:
                reached[e] = True
        if reached.all():
            break
        world.step(to_dev(act))
        world.observe()
    return reached, speed, lateral


def check_lane_envelope(reached, speed, lateral, tag):
    for e in range(len(speed)):
        assert reached[e], f"{tag}[{e}]: didn't reach goal"
        assert min(speed[e]) > 5 / 3.6, f"{tag}[{e}]: speed dropped below minimum"
        assert sum(speed[e]) / len(speed[e]) > 5, f"{tag}[{e}]: average speed {sum(speed[e]) / len(speed[e])}"
        assert max(lateral[e]) < 2.2, f"{tag}[{e}]: lateral error {max(lateral[e])}"
        assert sum(lateral[e]) / len(lateral[e]) < 1, f"{tag}[{e}]: average lateral error"


@pytest.mark.parametrize("turn", TURNS)
@pytest.mark.parametrize("run", LANE_RUNS, ids=[r[0] for r in LANE_RUNS])
def test_lane_following_envelope_oracle(turn, run):
    rm = load_map(f"turning_radius_{turn}")
    cfg, starts, goal, act = lane_world_inputs(rm, run[1], run[2])
    check_lane_envelope(*lane_envelope(oracle_world(rm, cfg), starts, goal, act), f"map{turn}/{run[0]}")


@pytest.mark.gpu
@pytest.mark.parametrize("turn", TURNS)
@pytest.mark.parametrize("run", LANE_RUNS, ids=[r[0] for r in LANE_RUNS])
def test_lane_following_envelope_gpu(turn, run):
    import torch
    from smarts_b200.engine import Engine
    rm = load_map(f"turning_radius_{turn}")
    cfg, starts, goal, act = lane_world_inputs(rm, run[1], run[2])
    eng = Engine(cfg, rm)
    out = lane_envelope(eng, starts, goal, act, to_dev=lambda a: torch.from_numpy(a).cuda(),
                        obs=lambda w: w.obs_numpy(), state=lambda w: w.state_numpy())
    check_lane_envelope(*out, f"map{turn}/{run[0]}")
    eng.close()


# ---------------------------------------------------------------------------------------- steering direction
def steering_direction(world, to_dev=lambda a: a, state=lambda w: w.state()):
    """smarts/core/tests/test_dynamics_backend.py:53-66: a chassis at rest, 100 Bullet steps (240 Hz) per phase -- here
    5 ticks of 0.1 s = 120 sub-steps -- with steering 0, then +1, then -1 (throttle and brake 0): the steering value
    read back is 0, then positive (a right turn), then negative."""
    z = np.zeros((1, 1))
    world.reset(z, z, z + math.pi * 0.5, z)
    out = []
    for steering in (0.0, 1.0, -1.0):
        act = np.zeros((1, 1, 3))
        act[..., 2] = steering
        for _ in range(5):
            world.step(to_dev(act))
        out.append(float(state(world)["steering"][0, 0]))
    return out


def check_steering(s):
    assert math.isclose(s[0], 0, rel_tol=1e-2, abs_tol=1e-12)
    assert s[1] > 0 and s[2] < 0
    # the joint settles on the commanded angle: steering * max_steering / steering_gear_ratio (chassis.py:790-793)
    assert math.isclose(s[1], 12.56 / 17.4, rel_tol=1e-3) and math.isclose(s[2], -12.56 / 17.4, rel_tol=1e-2)


def test_steering_direction_oracle():
    cfg = make_cfg(E=1, A=1, action_space=_abi.SB_ACTION_CONTINUOUS)
    check_steering(steering_direction(oracle_world(load_map("straight"), cfg)))


@pytest.mark.gpu
def test_steering_direction_gpu():
    import torch
    from smarts_b200 import lidar_params
    from smarts_b200.engine import Engine
    cfg = make_cfg(E=1, A=1, action_space=_abi.SB_ACTION_CONTINUOUS)
    eng = Engine(cfg, load_map("straight"), lidar_params.base_rays(lidar_params.BasicLidar))
    check_steering(steering_direction(eng, to_dev=lambda a: torch.from_numpy(a).cuda(), state=lambda w: w.state_numpy()))
    eng.close()
