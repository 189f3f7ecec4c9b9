"""Host-side mirror of the reference's agent / environment API: AgentInterface presets,
AgentSpec, action encoding, AgentInterface -> SbConfig; plus (gpu) the hiway-v0 step() contract."""
import numpy as np
import pytest

from conftest import load_map
from smarts_b200 import _abi, scenario_gen
from smarts_b200.agent_interface import (OGM, Agent, AgentInterface, AgentSpec, AgentType, ActionSpaceType,
                                         DoneCriteria, DrivableAreaGridMap, Lidar, NeighborhoodVehicles, Waypoints,
                                         compile_config)
from smarts_b200.hiway_env import (BatchedHiWayEnv, HiWayEnv, Scenario, SMARTSDestroyedError, SMARTSNotSetupError,
                                   encode_actions, make)
from smarts_b200.observation import Mission, Observation, Via

MISSIONS = {"A0": Mission("edge-south-SN", 1, 10.0, "edge-west-EW", start_speed=5.0),
            "A1": Mission("edge-west-WE", 0, 10.0, "edge-east-WE", start_speed=5.0),
            "A2": Mission("edge-north-NS", 0, 10.0, "edge-south-NS", start_speed=5.0),
            "A3": Mission("edge-east-EW", 1, 10.0, "edge-west-EW", endless=True, start_speed=5.0)}


def test_agent_type_presets_match_reference_table():
    # agent_interface.py:299-392
    laner = AgentInterface.from_type(AgentType.Laner)
    assert laner.waypoints == Waypoints(32) and laner.action == ActionSpaceType.Lane and not laner.lidar
    full = AgentInterface.from_type(AgentType.Full, max_episode_steps=7)
    assert full.ogm == OGM() and full.lidar and full.rgb and full.max_episode_steps == 7
    assert AgentInterface.from_type(AgentType.Tracker).action == ActionSpaceType.Trajectory
    assert AgentInterface.from_type(AgentType.Standard).action == ActionSpaceType.ActuatorDynamic
    assert AgentInterface.from_type(AgentType.LanerWithSpeed).action == ActionSpaceType.LaneWithContinuousSpeed
    assert AgentInterface(waypoints=True).waypoints.lookahead == 32
    assert AgentInterface(waypoints=17).waypoints is False  # _resolve_config: neither True nor the type
    assert laner.replace(ogm=True).ogm.resolution == 50 / 256 and laner.ogm is False
    assert [t.value for t in ActionSpaceType][:6] == [0, 1, 2, 3, 4, 5]  # controllers/__init__.py:42-54
    with pytest.raises(Exception):
        AgentInterface.from_type(99)


def test_compile_config_from_interface():
    itf = AgentInterface.from_type(AgentType.Laner, lidar=True, ogm=OGM(64, 64, 50 / 64),
                                   neighborhood_vehicles=NeighborhoodVehicles(radius=50), max_episode_steps=100,
                                   done_criteria=DoneCriteria(off_route=False, wrong_way=True))
    c = compile_config(itf, 8, 2, 20, 0.1)
    assert (c.n_envs, c.n_agents, c.n_social, c.action_space) == (8, 2, 20, _abi.SB_ACTION_LANE)
    assert (c.wp_enable, c.wp_lookahead, c.wp_out_paths, c.wp_out_len) == (1, 32, 4, 33)
    assert (c.lidar_enable, c.lidar_rays) == (1, 300)      # BasicLidar: 6 azimuths x 50 lasers
    assert (c.ogm_enable, c.ogm_width, c.ogm_height) == (1, 64, 64) and c.ogm_resolution == 50 / 64
    assert c.dagm_enable == 0 and c.nbr_radius == 50.0 and c.nbr_cap == 10
    assert c.done_mask == _abi.SB_EV_COLLISION | _abi.SB_EV_OFF_ROAD | _abi.SB_EV_WRONG_WAY
    assert c.max_episode_steps == 100 and c.not_moving_time == 60.0
    assert c.accel_enable == 1 and c.via_cap == 0          # accelerometer defaults to True (agent_interface.py:260)
    cv = compile_config(itf, 8, 2, 20, 0.1, via_cap=3)
    assert (cv.via_cap, cv.via_acquisition_range, cv.via_speed_accuracy) == (3, 40.0, 1.5)   # vehicle.py:552-558
    assert compile_config(AgentInterface.from_type(AgentType.Tagger), 1, 1, 0, 0.1).nbr_radius == -1.0  # radius=None
    for bad in (AgentType.Full, AgentType.MPCTracker, AgentType.Boid, AgentType.Buddha):
        with pytest.raises((NotImplementedError, ValueError)):
            compile_config(AgentInterface.from_type(bad), 1, 1, 0, 0.1)


def test_encode_actions_layouts():
    a = encode_actions(ActionSpaceType.Lane, ["keep_lane", "slow_down", "change_lane_left", "change_lane_right"], 10)
    assert a.dtype == np.int32 and a.tolist() == [0, 1, 2, 3]
    with pytest.raises(ValueError):
        encode_actions(ActionSpaceType.Lane, ["fly"], 10)
    a = encode_actions(ActionSpaceType.LaneWithContinuousSpeed, [(12.5, -1), (3, 0)], 10)
    assert a.shape == (2, 2) and a.dtype == np.float64 and a[0].tolist() == [12.5, -1.0]
    a = encode_actions(ActionSpaceType.Continuous, [np.array([0.5, 0.0, -0.2])], 10)
    assert a.shape == (1, 3)
    traj = [np.arange(4), np.arange(4) * 2.0, np.zeros(4), np.full(4, 13.9)]
    a = encode_actions(ActionSpaceType.Trajectory, [traj], 6)
    assert a.shape == (1, 4, 6) and a[0, 0].tolist() == [0, 1, 2, 3, 3, 3] and a[0, 3, 5] == 13.9


def test_agent_spec_build_agent_variants():
    class P(Agent):
        def __init__(self, k=1, j=2):
            self.k, self.j = k, j

        def act(self, obs):
            return self.k
    assert AgentSpec(agent_builder=P).build_agent().k == 1
    assert AgentSpec(agent_builder=P, agent_params=(5, 6)).build_agent().j == 6
    assert AgentSpec(agent_builder=P, agent_params={"k": 9, "zzz": 0}).build_agent().k == 9  # unknown keys dropped
    assert AgentSpec(agent_builder=Agent.from_function, agent_params=lambda _: "keep_lane").build_agent().act(None) == "keep_lane"
    with pytest.raises(ValueError):
        AgentSpec().build_agent()
    spec = AgentSpec(interface=AgentInterface.from_type(AgentType.Laner))
    assert spec.observation_adapter(3) == 3 and spec.reward_adapter(None, 2.0) == 2.0
    assert spec.info_adapter(None, 0, {"a": 1}) == {"a": 1}


def test_env_requires_gpu_and_valid_specs():
    import torch
    spec = AgentSpec(interface=AgentInterface.from_type(AgentType.Laner))
    with pytest.raises(ValueError):
        BatchedHiWayEnv(Scenario("4lane", missions={}), {"A0": spec})
    if not torch.cuda.is_available():
        from smarts_b200.engine import SmartsB200Error
        with pytest.raises(SmartsB200Error):  # no CPU fallback
            HiWayEnv([Scenario("4lane", missions=MISSIONS)], {"A0": spec})
    with pytest.raises(ValueError):
        make("smarts.env:other-v0")


def _specs(itf, ids=("A0", "A1", "A2", "A3")):
    return {i: AgentSpec(interface=itf, agent_builder=Agent.from_function, agent_params=lambda _: "keep_lane",
                         reward_adapter=lambda obs, r: 2.0 * r) for i in ids}


@pytest.mark.gpu
def test_hiway_env_step_contract():
    """reset()/step() return shapes, adapters, dones["__all__"], close() idempotence, errors."""
    rm = load_map("4lane")
    itf = AgentInterface.from_type(AgentType.Laner, neighborhood_vehicles=True, lidar=True, ogm=OGM(64, 64, 0.5),
                                   drivable_area_grid_map=DrivableAreaGridMap(64, 64, 0.5), max_episode_steps=12)
    specs = _specs(itf)
    traffic = scenario_gen.SocialTraffic(rm, 1, 10, 3)
    env = make("smarts.env:hiway-v0", scenarios=[Scenario(rm, MISSIONS, traffic, n_social=10)], agent_specs=specs)
    with pytest.raises(SMARTSNotSetupError):
        env.step({})
    obs = env.reset()
    assert set(obs) == set(specs) and all(isinstance(o, Observation) for o in obs.values())
    o = obs["A0"]
    assert o.step_count == 0 and o.ego_vehicle_state.lane_id is not None and o.ego_vehicle_state.road_id == "edge-south-SN"
    assert len(o.waypoint_paths) >= 1 and max(len(p) for p in o.waypoint_paths) == 33   # lookahead 32 + 1
    assert o.occupancy_grid_map.data.shape == (64, 64, 1) and o.occupancy_grid_map.data[32, 32, 0] == 255
    assert o.drivable_area_grid_map.metadata.resolution == 0.5
    pts, hits, rays = o.lidar_point_cloud
    assert len(pts) == len(hits) == len(rays) == 300
    agents = {i: s.build_agent() for i, s in specs.items()}
    total = {i: 0.0 for i in specs}
    for t in range(12):
        acts = {i: agents[i].act(obs[i]) for i in obs}
        obs, rewards, dones, infos = env.step(acts)
        for i, r in rewards.items():
            total[i] += r
            assert infos[i]["env_obs"].step_count == t + 1
    assert dones["__all__"] and all(dones[i] for i in specs)                        # max_episode_steps
    assert obs["A0"].events.reached_max_episode_steps
    for i in specs:  # reward adapter doubled the trip-meter increments; score is the trip meter
        assert total[i] == pytest.approx(2.0 * infos[i]["score"], rel=1e-12) and infos[i]["score"] > 3.0
    env.close(); env.close()
    with pytest.raises(SMARTSDestroyedError):
        env.step({})


@pytest.mark.gpu
def test_hiway_env_vias_and_accelerometer():
    """Mission vias reach Observation.via_data (near sorted by distance, hit once at the required speed) and the
    accelerometer members of the ego state are finite differences of its velocity vectors."""
    rm = load_map("4lane")
    itf = AgentInterface.from_type(AgentType.Laner, max_episode_steps=60)
    specs = _specs(itf)
    # a dry run without vias tells how fast A0 is at each point of its (straight) approach
    env = make("smarts.env:hiway-v0", scenarios=[Scenario(rm, MISSIONS)], agent_specs=specs)
    obs, track = env.reset(), []
    for t in range(25):
        obs, _, _, infos = env.step({i: "keep_lane" for i in obs})
        track.append((infos["A0"]["score"], obs["A0"].ego_vehicle_state.speed))   # score = the trip meter's total
    env.close()
    (d1, v1), (d2, v2) = track[8], track[20]
    assert d2 - d1 > 6.0
    missions = dict(MISSIONS)
    missions["A0"] = Mission("edge-south-SN", 1, 10.0, "edge-west-EW", start_speed=5.0,
                             via=(Via("edge-south-SN", 1, 10.0 + d2, v2 + 1.0, hit_distance=2.0),
                                  Via("edge-south-SN", 1, 10.0 + d1, v1 - 1.0),
                                  Via("edge-south-SN", 0, 10.0 + d1, 30.0, hit_distance=5.0)))
    env = make("smarts.env:hiway-v0", scenarios=[Scenario(rm, missions)], agent_specs=specs)
    obs = env.reset()
    near = obs["A0"].via_data.near_via_points
    assert [p.required_speed for p in near] == [v1 - 1.0, 30.0, v2 + 1.0] and obs["A0"].via_data.hit_via_points == []
    assert obs["A1"].via_data.near_via_points == []
    assert np.array_equal(obs["A0"].ego_vehicle_state.linear_acceleration, np.zeros(3))
    hits, prev_v, prev_a = [], obs["A0"].ego_vehicle_state.linear_velocity, None
    for t in range(40):
        obs, _, dones, _ = env.step({i: "keep_lane" for i in obs})
        if "A0" not in obs:
            break
        ego = obs["A0"].ego_vehicle_state
        acc = (ego.linear_velocity - prev_v) / 0.1
        assert np.allclose(ego.linear_acceleration, acc, rtol=1e-9, atol=1e-9)
        if prev_a is not None:
            assert np.allclose(ego.linear_jerk, acc - prev_a, rtol=1e-9, atol=1e-9)
        prev_v, prev_a = ego.linear_velocity, acc
        hits += [p.required_speed for p in obs["A0"].via_data.hit_via_points]
    assert sorted(hits) == sorted([v1 - 1.0, v2 + 1.0])   # each collected exactly once; the 30 m/s via never matches the speed
    so = env.batched.std_obs()
    assert so["ego"]["linear_acceleration"].dtype.is_floating_point and tuple(so["ego"]["linear_jerk"].shape) == (1, 4, 3)
    env.close()


@pytest.mark.gpu
def test_batched_env_auto_reset_and_tensor_path_agree():
    """ParallelEnv surface: per-env dicts, auto-reset on dones['__all__']; the device-tensor path gives the
    same rewards as the dict path."""
    import torch
    rm = load_map("4lane")
    itf = AgentInterface.from_type(AgentType.LanerWithSpeed, max_episode_steps=5)
    E = 6
    sc = Scenario(rm, MISSIONS, None, 0)
    env = BatchedHiWayEnv(sc, _specs(itf), n_envs=E, start_jitter=20.0, seed=5)
    assert env.batch_size == E and env.seed(5) == [5 + i for i in range(E)]
    obs = env.reset()
    assert len(obs) == E and set(obs[0]) == set(MISSIONS)
    rew_hist = []
    for t in range(5):
        obs, rew, dones, infos = env.step([{i: (8.0, 0) for i in ob} for ob in obs])
        rew_hist.append(np.array([[rew[e][i] for i in MISSIONS] for e in range(E)]))
    assert all(d["__all__"] for d in dones)
    assert all(ob["A0"].step_count == 0 for ob in obs)            # auto-reset: fresh observations
    obs, rew, dones, infos = env.step([{i: (8.0, 0) for i in ob} for ob in obs])
    assert not any(d["__all__"] for d in dones) and obs[0]["A1"].step_count == 1
    env.close()
    env2 = BatchedHiWayEnv(sc, _specs(itf), n_envs=E, start_jitter=20.0, seed=5)
    env2.reset_tensor()
    act = torch.tensor([[8.0, 0.0]], dtype=torch.float64, device="cuda").repeat(E * 4, 1)
    for t in range(5):
        out = env2.step_tensor(act)
        torch.cuda.synchronize()
        assert np.array_equal(2.0 * out["ev_reward"].cpu().numpy(), rew_hist[t])
    assert len(env2.auto_reset_done_envs()) == E
    env2.close()


@pytest.mark.gpu
def test_sharded_envs_reproduce_the_unsharded_job():
    """BatchedHiWayEnv(shard=...): the two halves of a 7-env job (as two ranks would own them) start and run
    exactly like envs [0, 4) and [4, 7) of the unsharded job, and count their agent-steps / episodes."""
    import torch
    from smarts_b200 import sharding
    rm = load_map("4lane")
    itf = AgentInterface.from_type(AgentType.LanerWithSpeed, max_episode_steps=4)
    sc = Scenario(rm, MISSIONS, None, 0)
    whole = BatchedHiWayEnv(sc, _specs(itf), n_envs=7, start_jitter=20.0, seed=5, shard=sharding.shard_envs(7, 0, 1))
    parts = [BatchedHiWayEnv(sc, _specs(itf), n_envs=sharding.shard_envs(7, r, 2).count, start_jitter=20.0, seed=5,
                             shard=sharding.shard_envs(7, r, 2)) for r in range(2)]
    assert [p.E for p in parts] == [4, 3] and parts[1].seed(5) == [9, 10, 11]
    for env in [whole] + parts:
        env.seed(5)
        env.reset_tensor()
    for t in range(4):
        outs = []
        for env in [whole] + parts:
            act = torch.tensor([[8.0, 0.0]], dtype=torch.float64, device="cuda").repeat(env.E * 4, 1)
            o = env.step_tensor(act)
            torch.cuda.synchronize()
            outs.append({k: o[k].cpu().numpy() for k in ("ev_reward", "ev_distance", "wp_pos")})
        for k in outs[0]:
            assert np.array_equal(outs[0][k], np.concatenate([outs[1][k], outs[2][k]]), equal_nan=True), (t, k)
    done = [len(env.auto_reset_done_envs()) for env in [whole] + parts]
    assert done == [7, 4, 3]
    st = [env.episode_stats() for env in [whole] + parts]
    assert st[0] == {"agent_steps": 7 * 4 * 4.0, "episodes_done": 7.0}
    assert st[1]["agent_steps"] + st[2]["agent_steps"] == st[0]["agent_steps"]
    with pytest.raises(ValueError):
        BatchedHiWayEnv(sc, _specs(itf), n_envs=5, shard=sharding.shard_envs(7, 0, 2))
    for env in [whole] + parts:
        env.close()


@pytest.mark.gpu
def test_scenario_cycle_replaces_the_map_on_reset():
    """hiway_env.py:265-285: reset() takes the next scenario of the cycle; here that means another road map on the
    same handle (sb_upload_map again), and the library refuses to step until every env was reset onto it."""
    from smarts_b200.engine import SmartsB200Error
    fe = {"A0": Mission("gneE20", 0, 5.0, "gneE27", start_speed=4.0), "A1": Mission("gneE36", 1, 5.0, "gneE25", start_speed=4.0)}
    fl = {"A0": MISSIONS["A0"], "A1": MISSIONS["A1"]}
    itf = AgentInterface.from_type(AgentType.Laner, neighborhood_vehicles=True)
    specs = _specs(itf, ("A0", "A1"))
    scs = [Scenario("4lane", fl, None, 0), Scenario("figure_eight", fe, None, 0)]
    env = HiWayEnv(scs, specs, shuffle_scenarios=False)
    seen = []
    for k in range(4):                      # 4lane, figure_eight, 4lane, figure_eight
        obs = env.reset()
        for t in range(3):
            obs, rew, dones, infos = env.step({i: "keep_lane" for i in obs})
        seen.append({i: (o.ego_vehicle_state.road_id, o.ego_vehicle_state.position.copy(),
                         o.waypoint_paths[0][-1].pos.copy()) for i, o in obs.items()})
    assert seen[0]["A0"][0] == "edge-south-SN" and seen[1]["A0"][0] == "gneE20"
    for i in ("A0", "A1"):                  # a cycle later the same scenario gives the same episode
        for a, b in ((0, 2), (1, 3)):
            assert seen[a][i][0] == seen[b][i][0] and np.array_equal(seen[a][i][1], seen[b][i][1])
            assert np.array_equal(seen[a][i][2], seen[b][i][2])
    fresh = HiWayEnv([scs[1]], specs)       # and the same as a handle that only ever saw that map
    obs = fresh.reset()
    for t in range(3):
        obs, _, _, _ = fresh.step({i: "keep_lane" for i in obs})
    for i in ("A0", "A1"):
        assert np.array_equal(obs[i].ego_vehicle_state.position, seen[1][i][1])
        assert np.array_equal(obs[i].waypoint_paths[0][-1].pos, seen[1][i][2])
    fresh.close()
    eng = env.batched.engine                # the C-ABI's guard
    eng.upload_map(load_map("4lane"))
    with pytest.raises(SmartsB200Error):
        eng.observe()
    env.close()
    with pytest.raises(ValueError):         # social-slot count is a size of the handle
        e2 = HiWayEnv([scs[0]], specs)
        try:
            e2.batched.set_scenario(Scenario("4lane", fl, None, 3))
        finally:
            e2.close()


@pytest.mark.gpu
def test_agents_with_different_sensor_sets_share_a_handle():
    """vehicle.py:470-560 attaches sensors per agent: A0 asks for lidar + OGM, A1 for neighbours only.  The handle
    computes the union; each Observation carries what its own interface subscribes to, with the values a
    single-interface env gives."""
    rm = load_map("4lane")
    base = AgentInterface.from_type(AgentType.Laner, accelerometer=False)
    i0 = base.replace(lidar=True, ogm=OGM(64, 64, 0.5))
    i1 = base.replace(neighborhood_vehicles=NeighborhoodVehicles(radius=400), accelerometer=True)
    fl = {"A0": MISSIONS["A0"], "A1": MISSIONS["A1"]}
    specs = {"A0": AgentSpec(interface=i0), "A1": AgentSpec(interface=i1)}
    env = HiWayEnv([Scenario(rm, fl, None, 0)], specs)
    both = i0.replace(neighborhood_vehicles=NeighborhoodVehicles(radius=400), accelerometer=True)
    ref = HiWayEnv([Scenario(rm, fl, None, 0)], {"A0": AgentSpec(interface=both), "A1": AgentSpec(interface=both)})
    obs, want = env.reset(), ref.reset()
    for t in range(3):
        obs, _, _, _ = env.step({i: "keep_lane" for i in obs})
        want, _, _, _ = ref.step({i: "keep_lane" for i in want})
    a0, a1 = obs["A0"], obs["A1"]
    assert a0.lidar_point_cloud is not None and a0.occupancy_grid_map is not None
    assert a0.neighborhood_vehicle_states is None and a0.ego_vehicle_state.linear_acceleration is None
    assert a1.lidar_point_cloud is None and a1.occupancy_grid_map is None
    assert len(a1.neighborhood_vehicle_states) == len(want["A1"].neighborhood_vehicle_states) == 1
    assert np.array_equal(a1.ego_vehicle_state.linear_acceleration, want["A1"].ego_vehicle_state.linear_acceleration)
    assert np.array_equal(a0.occupancy_grid_map.data, want["A0"].occupancy_grid_map.data)
    assert np.array_equal(np.asarray(a0.lidar_point_cloud[0]), np.asarray(want["A0"].lidar_point_cloud[0]))
    assert np.array_equal(a0.waypoint_paths[0][0].pos, want["A0"].waypoint_paths[0][0].pos)
    env.close(); ref.close()
    with pytest.raises(NotImplementedError):    # same sensor, other parameters
        HiWayEnv([Scenario(rm, fl, None, 0)], {"A0": AgentSpec(interface=i0),
                                               "A1": AgentSpec(interface=base.replace(ogm=OGM(32, 32, 1.0)))})
    with pytest.raises(NotImplementedError):    # not a sensor choice
        HiWayEnv([Scenario(rm, fl, None, 0)], {"A0": AgentSpec(interface=i0),
                                               "A1": AgentSpec(interface=base.replace(max_episode_steps=7))})


@pytest.mark.gpu
def test_agents_alive_done_criteria():
    """DoneCriteria.agents_alive (sensors.py:404-441): once A0 has reached its goal and left, the tick after it the
    watch list [A0] has no agent alive -> every remaining agent is done with events.agents_alive_done."""
    from smarts_b200.agent_interface import AgentsAliveDoneCriteria, AgentsListAlive
    rm = load_map("4lane")
    crit = DoneCriteria(agents_alive=AgentsAliveDoneCriteria(agent_lists_alive=[AgentsListAlive(["A0"], 1)]))
    itf = AgentInterface.from_type(AgentType.LanerWithSpeed, done_criteria=crit)
    missions = dict(MISSIONS, A0=Mission("edge-south-SN", 1, 10.0, "edge-south-SN", end_lane=1, end_offset=22.0,
                                         goal_radius=2.0, start_speed=5.0))
    env = BatchedHiWayEnv(Scenario(rm, missions, None, 0), _specs(itf), n_envs=3, auto_reset=False)
    obs = env.reset()
    left_at = None
    for t in range(60):
        obs, rew, dones, infos = env.step([{i: (8.0, 0) for i in ob} for ob in obs])
        if left_at is None:
            assert not any(o.events.agents_alive_done for ob in obs for o in ob.values())
            if dones[0].get("A0"):
                assert obs[0]["A0"].events.reached_goal and not dones[0]["__all__"]
                left_at = t
        else:
            assert t == left_at + 1
            for e in range(3):
                assert set(obs[e]) == {"A1", "A2", "A3"} and dones[e]["__all__"]
                assert all(o.events.agents_alive_done for o in obs[e].values()) and all(dones[e][i] for i in obs[e])
            so = env.std_obs()
            assert so["events"]["agents_alive_done"].cpu().numpy().all()
            break
    assert left_at is not None
    env.close()
    # minimum_total_agents_alive: four agents enter the first tick alive, the minimum is five
    itf5 = AgentInterface.from_type(AgentType.LanerWithSpeed, done_criteria=DoneCriteria(
        agents_alive=AgentsAliveDoneCriteria(minimum_total_agents_alive=5)))
    env = BatchedHiWayEnv(Scenario(rm, MISSIONS, None, 0), _specs(itf5), n_envs=2, auto_reset=False)
    obs = env.reset()
    assert not obs[0]["A0"].events.agents_alive_done
    obs, rew, dones, infos = env.step([{i: (8.0, 0) for i in ob} for ob in obs])
    assert all(d["__all__"] for d in dones) and obs[1]["A3"].events.agents_alive_done
    env.close()


@pytest.mark.gpu
def test_staggered_auto_reset_leaves_running_envs_alone():
    """Envs that finish on different ticks: an env that is auto-reset must not disturb the others (in the reference
    every ParallelEnv env is its own process).  Env 1 stands still and is ended by `not_moving` every few ticks
    while envs 0 and 2 drive on; their observations equal those of a run without auto-reset."""
    rm = load_map("4lane")
    from smarts_b200.agent_interface import EventConfiguration
    itf = AgentInterface.from_type(AgentType.LanerWithSpeed, neighborhood_vehicles=True,
                                   done_criteria=DoneCriteria(not_moving=True, off_route=False),
                                   event_configuration=EventConfiguration(not_moving_time=0.5, not_moving_distance=0.5))
    E, M = 3, 6
    sc = Scenario(rm, MISSIONS, scenario_gen.SocialTraffic(rm, E, M, 17), n_social=M)
    hist = {}
    for auto in (True, False):
        env = BatchedHiWayEnv(sc, _specs(itf), n_envs=E, start_jitter=15.0, seed=9, auto_reset=auto)
        obs = env.reset()
        rows, resets = [], 0
        for t in range(20):
            acts = [{i: ((0.0, 0) if e == 1 else (9.0, 0)) for i in MISSIONS} for e in range(E)]
            obs, rew, dones, infos = env.step([{i: a for i, a in acts[e].items() if i in obs[e]} for e in range(E)])
            resets += int(dones[1]["__all__"])
            row = {}
            for e in (0, 2):
                for i in infos[e]:
                    o = infos[e][i]["env_obs"]
                    eg = o.ego_vehicle_state
                    row[e, i] = (eg.position.copy(), o.distance_travelled, o.events, eg.linear_acceleration.copy(),
                                 eg.linear_jerk.copy(), [n.id for n in o.neighborhood_vehicle_states], rew[e][i])
            rows.append(row)
        hist[auto] = rows
        if auto:
            assert resets >= 2          # env 1 was ended by not_moving and restarted more than once
            assert any(n.id.startswith("social") for o in obs[1].values() for n in o.neighborhood_vehicle_states), \
                "the reset observation still sees the injected traffic"
        env.close()
    for t, (a, b) in enumerate(zip(hist[True], hist[False])):
        assert a.keys() == b.keys() and len(a) == 8
        for key in a:
            for u, w in zip(a[key], b[key]):
                assert np.array_equal(u, w) if isinstance(u, np.ndarray) else u == w, (t, key)


@pytest.mark.gpu
def test_ego_centric_adapters_are_per_env():
    """get_egocentric_adapters keeps the last observation between observe and act (ego_centric_adapters.py:280-299);
    with E > 1 every env must convert its ego-frame trajectory with ITS OWN ego pose."""
    from smarts_b200 import ego_centric
    rm = load_map("4lane")
    itf = AgentInterface.from_type(AgentType.Tracker)
    oa, aa = ego_centric.get_egocentric_adapters(ActionSpaceType.Trajectory)
    spec = AgentSpec(interface=itf, observation_adapter=oa, action_adapter=aa)
    E = 3
    env = BatchedHiWayEnv(Scenario(rm, {"A0": MISSIONS["A0"]}, None, 0), {"A0": spec}, n_envs=E, start_jitter=40.0, seed=3)
    obs = env.reset()
    assert all(np.allclose(o["A0"].ego_vehicle_state.position, 0) for o in obs)      # ego frame
    world = env._host_state
    ego_traj = [np.array([2.0, 4.0, 6.0]), np.zeros(3), np.zeros(3), np.full(3, 8.0)]  # straight ahead, ego frame
    xs = []
    for e in range(E):
        tx, ty, th, tv = env._spec(e, "A0").action_adapter(ego_traj)
        hd = world["heading"][e, 0]
        pos = np.array([world["pos_x"][e, 0], world["pos_y"][e, 0], world["pos_z"][e, 0]])
        ex, ey = ego_centric.world_position_from_ego_frame([2.0, 0.0, 0.0], pos, hd)[:2]   # with env e's own pose
        assert np.allclose([tx[0], ty[0]], [ex, ey], atol=1e-9) and np.allclose(th, hd)
        xs.append((tx[0], ty[0]))
    assert len({(round(a, 3), round(b, 3)) for a, b in xs}) == E                          # the envs differ
    obs, _, _, _ = env.step([{"A0": ego_traj}] * E)
    env.close()


@pytest.mark.gpu
def test_std_obs_layout_matches_format_obs():
    """StdObs keys / dtypes / trailing shapes (format_obs.py:45-197, :313-376) on the device, and agreement
    with the per-agent Observation objects of the dict path."""
    import torch
    rm = load_map("4lane")
    itf = AgentInterface.from_type(AgentType.Laner, neighborhood_vehicles=NeighborhoodVehicles(radius=50), lidar=True,
                                   ogm=OGM(64, 64, 0.5), drivable_area_grid_map=DrivableAreaGridMap(64, 64, 0.5))
    E = 3
    env = BatchedHiWayEnv(Scenario(rm, MISSIONS, scenario_gen.SocialTraffic(rm, E, 12, 8), n_social=12), _specs(itf),
                          n_envs=E, wp_out_len=20, start_jitter=10.0)
    obs = env.reset()
    for _ in range(3):
        obs, _, _, _ = env.step([{i: "keep_lane" for i in ob} for ob in obs])
    so = env.std_obs()
    torch.cuda.synchronize()
    A = 4
    want = {"dist": ((), torch.float32), "dagm": ((64, 64, 1), torch.uint8), "ogm": ((64, 64, 1), torch.uint8)}
    for k, (shp, dt) in want.items():
        assert tuple(so[k].shape) == (E, A) + shp and so[k].dtype == dt, k
    ego = {"angular_acceleration": (3,), "angular_jerk": (3,), "angular_velocity": (3,), "box": (3,), "heading": (),
           "lane_index": (), "linear_acceleration": (3,), "linear_jerk": (3,), "linear_velocity": (3,), "pos": (3,),
           "speed": (), "steering": (), "yaw_rate": ()}
    assert set(so["ego"]) == set(ego)
    for k, shp in ego.items():
        dt = torch.float64 if k == "pos" else (torch.int8 if k == "lane_index" else torch.float32)
        assert tuple(so["ego"][k].shape) == (E, A) + shp and so["ego"][k].dtype == dt, k
    assert set(so["events"]) == {"agents_alive_done", "collisions", "not_moving", "off_road", "off_route", "on_shoulder",
                                 "reached_goal", "reached_max_episode_steps", "wrong_way"}
    assert all(v.dtype == torch.int8 and tuple(v.shape) == (E, A) for v in so["events"].values())
    lid = so["lidar"]
    assert tuple(lid["hit"].shape) == (E, A, 300) and lid["hit"].dtype == torch.int8
    for k in ("point_cloud", "ray_origin", "ray_vector"):
        assert tuple(lid[k].shape) == (E, A, 300, 3) and lid[k].dtype == torch.float64
    assert torch.isfinite(lid["point_cloud"]).all()
    nb = so["neighbors"]
    assert tuple(nb["pos"].shape) == (E, A, 10, 3) and nb["pos"].dtype == torch.float64 and nb["lane_index"].dtype == torch.int8
    wp = so["waypoints"]
    assert tuple(wp["pos"].shape) == (E, A, 4, 20, 3) and tuple(wp["heading"].shape) == (E, A, 4, 20)
    ttc = so["ttc"]  # enabled because both the waypoint and the neighbour sensor are on (format_obs.py:387-388)
    assert set(ttc) == {"angle_error", "distance_from_center", "dtc", "ttc"}
    assert tuple(ttc["dtc"].shape) == (E, A, 3) and tuple(ttc["angle_error"].shape) == (E, A)
    assert all(v.dtype == torch.float32 for v in ttc.values())
    # agreement with the Observation objects
    o = obs[1]["A2"]
    assert np.allclose(so["ego"]["pos"][1, 2].cpu().numpy(), o.ego_vehicle_state.position)
    assert so["ego"]["lane_index"][1, 2].item() == o.ego_vehicle_state.lane_index
    assert np.float32(o.distance_travelled) == so["dist"][1, 2].item()
    assert np.array_equal(so["ogm"][1, 2].cpu().numpy(), o.occupancy_grid_map.data)
    assert np.allclose(lid["ray_vector"][1, 2, 7].cpu().numpy(), o.lidar_point_cloud[2][7][1])
    env.close()
