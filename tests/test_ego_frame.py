"""ego_centric_observation_adapter (smarts/core/utils/adapters/ego_centric_adapters.py:62-178) on the observation
buffers (`SbConfig.ego_frame`): the oracle against golden outputs of the reference's own adapter function
(tests/golden/gen_ego_frame_golden.py -> ego_frame.npz), and (gpu) k_ego_frame against the oracle, bit for bit."""
import os

import numpy as np
import pytest

from conftest import GOLDEN, load_map, make_cfg, oracle_world
from smarts_b200 import scenario_gen

CASES = {
    "4lane": [("edge-south-SN", 1, 10.0, "edge-west-EW"), ("edge-west-WE", 0, 10.0, "edge-east-WE"),
              ("edge-north-NS", 0, 10.0, "edge-south-NS"), ("edge-east-EW", 1, 10.0, "edge-west-EW")],
    "figure_eight": [("gneE20", 0, 5.0, "gneE27"), ("gneE36", 1, 5.0, "gneE25"), ("gneE20", 2, 25.0, "gneE27"),
                     ("gneE36", 0, 30.0, "gneE25")],
}
KEYS = ("nbr_pos", "nbr_heading", "wp_pos", "wp_heading", "lidar_points")
E, M, TICKS, OBSERVE_AT = 4, 12, 12, (0, 5, 11)


def cfg_for(name, ego_frame, E_=E, **kw):
    # 4lane keeps the reference's (inf, inf, inf) lidar misses, figure_eight the zero fill of StdObs
    return make_cfg(E=E_, A=len(CASES[name]), M=M, nbr=True, nbr_radius=60.0, nbr_cap=8, wp=True, wp_lookahead=24,
                    wp_paths=4, wp_len=25, route_cap=24, lidar=True, lidar_miss_inf=int(name == "4lane"),
                    ego_frame=ego_frame, **kw)


def drive(name, world_pairs, E_=E, seed=101):
    """Steps every world in ``world_pairs`` [(world, to_device)] through the same scenario; yields after each
    observation listed in OBSERVE_AT."""
    rm = load_map(name)
    A = len(CASES[name])
    x, y, h, v, route, rlen = scenario_gen.agent_starts(rm, E_, CASES[name], 24, seed, jitter=30.0, speed=9.0)
    tr = scenario_gen.SocialTraffic(rm, E_, M, seed + 1)
    acts = np.random.default_rng(seed + 2).integers(0, 4, (TICKS, E_, A)).astype(np.int32)
    for w, _ in world_pairs:
        w.reset(x, y, h, v, route, rlen)
    for t in range(TICKS):
        soc = tr.states(t)
        for w, dev in world_pairs:
            w.set_social(dev(soc))
            w.step(dev(acts[t]))
        if t in OBSERVE_AT:
            for w, _ in world_pairs:
                w.observe()
            yield t


@pytest.mark.parametrize("name", sorted(CASES))
def test_oracle_ego_frame_matches_reference_adapter(name):
    z = np.load(os.path.join(GOLDEN, "ego_frame.npz"))
    rm = load_map(name)
    ora = oracle_world(rm, cfg_for(name, True), threads=4)
    k = 0
    for t in drive(name, [(ora, lambda a: a)]):
        o = ora.obs()
        for key in KEYS:
            want = z[f"{name}_{key}"][k]
            got = o[key].astype(np.float64)
            assert got.shape == want.shape, key
            assert np.array_equal(np.isnan(got), np.isnan(want)), (t, key)
            # float32 heading members hold the rounded value of the same binary64 arithmetic; the rotation's
            # sin / cos are sb_detmath.h here and libm there, and numpy's 3x3 product may fuse multiply-adds
            tol = 1e-6 if key.endswith("heading") else 1e-9
            assert np.allclose(got, want, rtol=0, atol=tol, equal_nan=True), (t, key, np.nanmax(np.abs(got - want)))
        k += 1
    assert k == len(OBSERVE_AT)
    lid = z[f"{name}_lidar_points"]
    assert np.isnan(lid).any() == (name == "4lane") and np.isfinite(lid).any()
    assert (np.abs(z[f"{name}_nbr_pos"]) > 1.0).any() and (np.abs(z[f"{name}_wp_pos"][..., 0]) > 5.0).any()


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(CASES))
def test_gpu_ego_frame_matches_oracle(name):
    import torch
    from smarts_b200 import lidar_params
    from smarts_b200.engine import Engine
    rm = load_map(name)
    E_ = 24
    cfg = cfg_for(name, True, E_=E_, events=True, ttc=True)   # ttc / events still see the world-frame buffers
    eng = Engine(cfg, rm, lidar_params.base_rays(lidar_params.BasicLidar))
    ora = oracle_world(rm, cfg, threads=8)
    plain = oracle_world(rm, cfg_for(name, False, E_=E_, events=True, ttc=True), threads=8)
    n = 0
    for t in drive(name, [(eng, lambda a: torch.from_numpy(a).cuda()), (ora, lambda a: a), (plain, lambda a: a)], E_=E_,
                   seed=111):
        got, want, world = eng.obs_numpy(), ora.obs(), plain.obs()
        for k in want:
            assert np.array_equal(got[k].view(np.uint8), want[k].view(np.uint8)), (t, k)
        for k in want:  # everything but the five rewritten members is what the world-frame run gives
            if k not in KEYS:
                assert np.array_equal(want[k].view(np.uint8), world[k].view(np.uint8)), (t, k)
        assert not np.array_equal(want["nbr_pos"], world["nbr_pos"])
        n += 1
    assert n == len(OBSERVE_AT)


def test_host_adapters_round_trip():
    """ego_centric.py mirrors utils/math.py:452-505 and the adapter pairing of ego_centric_adapters.py:280-325 (`_AdapterPair`)."""
    from smarts_b200 import ego_centric as ec
    from smarts_b200.agent_interface import ActionSpaceType
    from smarts_b200.observation import (EgoVehicleObservation, Dimensions, Events, Observation, VehicleObservation, Vias,
                                         ViaPoint, Waypoint)
    rng = np.random.default_rng(5)
    pos, hd = np.array([12.0, -7.5, 0.0]), 2.3
    for _ in range(20):
        p = rng.normal(0, 50, 3)
        q = ec.position_to_ego_frame(p, pos, hd)
        assert np.allclose(ec.world_position_from_ego_frame(q, pos, hd), p, atol=1e-9)
    # +x of the ego frame is the heading direction radians_to_vec(h) = (-sin h, cos h)
    ahead = pos + 4.0 * np.array([-np.sin(hd), np.cos(hd), 0.0])
    assert np.allclose(ec.position_to_ego_frame(ahead, pos, hd + np.pi / 2), [4.0, 0.0, 0.0], atol=1e-12)
    assert ec.wrap_value(3.5, -np.pi, np.pi) == pytest.approx(3.5 - 2 * np.pi) and ec.wrap_value(-np.pi, -np.pi, np.pi) == np.pi
    ego = EgoVehicleObservation(id="A", position=pos, bounding_box=Dimensions(3.68, 1.47, 1.0), heading=hd, speed=5.0,
                                steering=0.0, yaw_rate=0.0, road_id="r", lane_id="l", lane_index=0, mission=None,
                                linear_velocity=np.array([3.0, 4.0, 0.0]), angular_velocity=np.zeros(3),
                                linear_acceleration=np.array([0.0, 2.0, 0.0]), angular_acceleration=np.zeros(3),
                                linear_jerk=None, angular_jerk=None)
    nv = VehicleObservation(id="s", position=np.array([20.0, 1.0, 0.0]), bounding_box=Dimensions(4, 2, 1.5), heading=-2.9,
                            speed=3.0, road_id="r", lane_id="l", lane_index=1)
    wp = Waypoint(pos=np.array([13.0, -6.0]), heading=2.4, lane_id="l", lane_width=3.2, speed_limit=13.9, lane_index=0)
    obs = Observation(dt=0.1, step_count=1, elapsed_sim_time=0.1, events=None, ego_vehicle_state=ego,
                      neighborhood_vehicle_states=[nv], waypoint_paths=[[wp]], distance_travelled=0.0,
                      lidar_point_cloud=([np.array([np.inf] * 3), ahead], [False, True], [(pos, ahead)] * 2),
                      drivable_area_grid_map=None, occupancy_grid_map=None, top_down_rgb=None, road_waypoints=None,
                      via_data=Vias([ViaPoint((ahead[0], ahead[1]), 0, "r", 6.0)], []))
    oa, aa = ec.get_egocentric_adapters(ActionSpaceType.Trajectory)
    out = oa(obs)
    assert list(out.ego_vehicle_state.position) == [0, 0, 0] and out.ego_vehicle_state.heading == 0.0
    assert np.allclose(out.ego_vehicle_state.linear_velocity, [5.0, 0.0, 0.0]) and out.ego_vehicle_state.linear_jerk is None
    assert np.allclose(out.neighborhood_vehicle_states[0].position, ec.position_to_ego_frame(nv.position, pos, hd))
    assert out.neighborhood_vehicle_states[0].heading == pytest.approx(ec.wrap_value(-2.9 - hd, -np.pi, np.pi))
    assert out.waypoint_paths[0][0].heading == pytest.approx(0.1) and out.waypoint_paths[0][0].lane_width == 3.2
    assert np.isnan(out.lidar_point_cloud[0][0]).all() and np.allclose(out.lidar_point_cloud[2][0][0], 0.0)
    assert np.allclose(out.via_data.near_via_points[0].position, ec.position_to_ego_frame(ahead, pos, hd)[:2])
    assert obs.ego_vehicle_state.heading == hd                      # the input is not modified
    # the action adapter maps an ego-frame trajectory back with the last unmodified observation
    ex, ey = np.array([1.0, 2.0, 3.0]), np.array([0.0, 0.5, 1.0])
    x, y, h, sp = aa((ex, ey, np.array([0.0, 0.1, 0.2]), np.array([5.0, 5.0, 5.0])))
    for i in range(3):
        assert np.allclose(ec.position_to_ego_frame([x[i], y[i], 0.0], pos, hd)[:2], [ex[i], ey[i]], atol=1e-9)
    assert h[1] == pytest.approx(ec.wrap_value(hd + 0.1, -np.pi, np.pi))
    assert ec.get_egocentric_adapters(ActionSpaceType.Lane)[1]("keep_lane") == "keep_lane"


@pytest.mark.gpu
def test_env_ego_centric_tensor_path_matches_host_adapter():
    """BatchedHiWayEnv(ego_centric=True).std_obs() == the host adapter applied to the world-frame Observations of
    an identical env."""
    import torch
    from smarts_b200 import ego_centric as ec
    from smarts_b200.agent_interface import Agent, AgentInterface, AgentSpec, AgentType, NeighborhoodVehicles
    from smarts_b200.hiway_env import BatchedHiWayEnv, Scenario
    from smarts_b200.observation import Mission
    rm = load_map("4lane")
    missions = {"A0": Mission("edge-south-SN", 1, 10.0, "edge-west-EW", start_speed=5.0),
                "A1": Mission("edge-west-WE", 0, 10.0, "edge-east-WE", start_speed=5.0)}
    itf = AgentInterface.from_type(AgentType.Laner, neighborhood_vehicles=NeighborhoodVehicles(radius=50), lidar=True)
    specs = {i: AgentSpec(interface=itf, agent_builder=Agent.from_function, agent_params=lambda _: "keep_lane")
             for i in missions}
    E_ = 3
    mk = lambda ego: BatchedHiWayEnv(Scenario(rm, missions, scenario_gen.SocialTraffic(rm, E_, 8, 8), n_social=8), specs,  # noqa: E731
                                     n_envs=E_, wp_out_len=20, start_jitter=10.0, ego_centric=ego)
    world, egoenv = mk(False), mk(True)
    obs = world.reset(); egoenv.reset()
    for _ in range(4):
        acts = [{i: "keep_lane" for i in ob} for ob in obs]
        obs, _, _, _ = world.step(acts); egoenv.step(acts)
    so = egoenv.std_obs()
    torch.cuda.synchronize()
    assert not so["ego"]["pos"].any() and not so["ego"]["heading"].any()
    compared = seen_nb = 0
    for e in range(E_):
        for a, aid in enumerate(missions):
            if aid not in obs[e]:   # the agent's episode ended (contact with the synthetic traffic)
                continue
            compared += 1
            want = ec.ego_centric_observation_adapter(obs[e][aid])
            nb = want.neighborhood_vehicle_states
            got = so["neighbors"]["pos"][e, a, :len(nb)].cpu().numpy()
            seen_nb += len(nb)
            if nb:
                assert np.allclose(got, [v.position for v in nb], atol=1e-9)
                assert np.allclose(so["neighbors"]["heading"][e, a, :len(nb)].cpu().numpy(), [v.heading for v in nb], atol=1e-6)
            p0 = want.waypoint_paths[0]
            n = min(len(p0), 20)
            assert np.allclose(so["waypoints"]["pos"][e, a, 0, :n, :2].cpu().numpy(), [w.pos for w in p0[:n]], atol=1e-9)
            hits = np.asarray(want.lidar_point_cloud[1], bool)
            pc = so["lidar"]["point_cloud"][e, a].cpu().numpy()
            assert np.allclose(pc[hits], np.asarray(want.lidar_point_cloud[0])[hits], atol=1e-9)
            assert not pc[~hits].any()                                   # misses stay zero in StdObs
            rv = np.asarray([r[1] for r in want.lidar_point_cloud[2]])
            assert np.allclose(so["lidar"]["ray_vector"][e, a].cpu().numpy(), rv, atol=1e-9)
            assert np.allclose(so["ego"]["linear_velocity"][e, a].cpu().numpy(), want.ego_vehicle_state.linear_velocity, atol=1e-5)
    assert compared >= 3 and seen_nb > 0
    world.close(); egoenv.close()
