"""``hiway-v0`` on the B200 step: the gym-style environment the reference registers as
``smarts.env:hiway-v0`` (/root/reference/smarts/env/hiway_env.py:43-291, smarts/env/__init__.py:20-23)
and its batched form ``ParallelEnv`` (/root/reference/smarts/env/wrappers/parallel_env.py:129-262),
both driving ONE ``Engine`` handle (E environments on one GPU) instead of E simulator processes.

    env = HiWayEnv(scenarios=[Scenario("4lane", missions={...})], agent_specs={"A0": spec, ...})
    obs = env.reset();  obs, rewards, dones, infos = env.step({"A0": "keep_lane", ...})

    envs = BatchedHiWayEnv(scenario, agent_specs, n_envs=4096)            # ParallelEnv surface
    obs_tensors = envs.step_tensor(action_tensor)                          # zero-copy device tensors

What the reference builds from a scenario directory (sstudio pickles, SUMO route files) is given
here explicitly as a ``Scenario`` record: the flattened map, one ``Mission`` per agent and the
social-traffic source (SUMO stays out of the kernel path, BASELINE.json north_star).  There is no
CPU fallback: constructing either class without the CUDA library or a GPU raises.
"""
from __future__ import annotations

import copy
import dataclasses
import itertools
import math
import os
from dataclasses import dataclass, field
from typing import Any, Callable, Dict, List, Optional, Sequence, Tuple, Union

import numpy as np

from . import _abi, lidar_params
from .agent_interface import ActionSpaceType, AgentInterface, AgentSpec, compile_config
from .observation import (Collision, Dimensions, DrivableAreaGridMap, EgoVehicleObservation, Events,
                          GridMapMetadata, Mission, Observation, OccupancyGridMap, RoadWaypoints, VehicleObservation, Via,
                          ViaPoint, Vias, Waypoint)
from .road_map import RoadMapData

_MAPS_DIR = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "maps")
_LANE_COMMANDS = {"keep_lane": _abi.SB_LANE_KEEP, "slow_down": _abi.SB_LANE_SLOW_DOWN,
                  "change_lane_left": _abi.SB_LANE_CHANGE_LEFT, "change_lane_right": _abi.SB_LANE_CHANGE_RIGHT}


class SMARTSNotSetupError(Exception):
    """step() before reset() (smarts/core/smarts.py:207-208)."""


class SMARTSDestroyedError(Exception):
    """use after close() (smarts/core/smarts.py:705-709)."""


@dataclass
class Scenario:
    """One scenario: map + agent missions + social traffic source.

    ``map`` is a ``RoadMapData``, the name of a flattened map under ``maps/`` ("4lane", "loop" ...)
    or the path of a SUMO ``map.net.xml``.  ``traffic`` is ``None`` (no social vehicles), a
    ``[T, E, M, 8]`` recording, or a callable ``t -> [E, M, 8]`` (the ``ExternalProvider`` record,
    external_provider.py:44-51)."""
    map: Union[str, RoadMapData]
    missions: Dict[str, Mission] = field(default_factory=dict)
    traffic: Any = None
    n_social: int = 0

    def road_map(self) -> RoadMapData:
        if isinstance(self.map, RoadMapData):
            return self.map
        p = self.map
        if p.endswith(".net.xml"):
            return RoadMapData.from_file(p)
        if not p.endswith(".npz"):
            p = os.path.join(_MAPS_DIR, os.path.basename(p.rstrip("/")) + ".npz")
        return RoadMapData.load(p)


def encode_actions(space: ActionSpaceType, actions: Sequence[Any], traj_len: int) -> np.ndarray:
    """Reference action values -> the flat array ``sb_step`` takes (include/smarts_b200.h).
    Lane strings as in controllers/__init__.py:125-144; Trajectory = 4 sequences x, y, heading, speed."""
    n = len(actions)
    if space == ActionSpaceType.Lane:
        out = np.zeros(n, np.int32)
        for i, a in enumerate(actions):
            if isinstance(a, str):
                if a not in _LANE_COMMANDS:
                    raise ValueError(f"unknown lane action {a!r}")
                out[i] = _LANE_COMMANDS[a]
            else:
                out[i] = int(a)
        return out
    if space == ActionSpaceType.LaneWithContinuousSpeed:
        return np.asarray([[float(a[0]), float(a[1])] for a in actions], np.float64).reshape(n, 2)
    if space in (ActionSpaceType.Continuous, ActionSpaceType.ActuatorDynamic):
        return np.asarray([[float(a[0]), float(a[1]), float(a[2])] for a in actions], np.float64).reshape(n, 3)
    if space == ActionSpaceType.Trajectory:
        out = np.zeros((n, 4, traj_len), np.float64)
        for i, a in enumerate(actions):
            t = np.asarray(a, np.float64)
            if t.shape[0] != 4 or t.shape[1] < 2:
                raise ValueError("Trajectory action must be 4 sequences (x, y, heading, speed) of >= 2 points")
            k = min(traj_len, t.shape[1])
            out[i, :, :k] = t[:, :k]
            out[i, :, k:] = t[:, k - 1:k]  # pad by repeating the last point
        return out
    raise NotImplementedError(space)


class BatchedHiWayEnv:
    """E copies of one scenario on one GPU behind the ``ParallelEnv`` surface
    (parallel_env.py:129-262): ``step(Sequence[Dict])``, ``reset()``, ``seed()``, ``batch_size``,
    auto-reset of an env when its ``dones["__all__"]`` is raised.  ``step_tensor`` is the same step
    without leaving the device."""

    def __init__(self, scenario: Scenario, agent_specs: Dict[str, AgentSpec], n_envs: int = 1, *,
                 fixed_timestep_sec: float = 0.1, seed: int = 42, headless: bool = True, device: int = 0,
                 auto_reset: bool = True, start_jitter: float = 0.0, route_cap: int = 16, traj_len: int = 10,
                 lidar_wall_height: float = 0.0, wp_out_len: Optional[int] = None, sim_name: Optional[str] = None,
                 ego_centric: bool = False, shard=None, min_via_cap: int = 0, **_ignored_reference_kwargs):
        from .engine import Engine  # raises without the CUDA library / a GPU: no CPU fallback
        if not agent_specs:
            raise ValueError("agent_specs is empty")
        self._agent_specs = dict(agent_specs)
        self.agent_ids = list(agent_specs.keys())
        # ParallelEnv gives every env process its own copy of the AgentSpecs (parallel_env.py:96-127), so adapters
        # that remember something between observe and act (the ego-centric pair keeps the last ego pose) never mix
        # envs up.  Same here: one deep copy of the four adapters per env (the agent builder / params are shared).
        self._env_specs: List[Dict[str, AgentSpec]] = [dict(agent_specs)] if int(n_envs) == 1 else [
            {aid: self._own_adapters(s) for aid, s in agent_specs.items()} for _ in range(int(n_envs))]
        itfs = [s.interface for s in agent_specs.values()]
        if any(i is None for i in itfs):
            raise ValueError("every AgentSpec needs an interface")
        # One SbConfig per handle: the handle computes the UNION of the agents' sensors for every agent and each
        # agent's Observation shows what its own interface subscribes to (vehicle.py:470-560 attaches sensors per
        # agent).  Everything that is not an on/off sensor choice must agree.
        self.interface: AgentInterface = self._union_interface(itfs)
        self._own_itf: List[AgentInterface] = itfs
        self.scenario = scenario
        self.road_map = scenario.road_map()
        missing = [a for a in self.agent_ids if a not in scenario.missions]
        if missing:
            raise ValueError(f"scenario has no mission for agents {missing}")
        # shard (sharding.EnvShard): this process owns envs [shard.start, shard.stop) of a job of shard.total envs --
        # one process per GPU, no data-path collective; start jitter and seeds are functions of the GLOBAL env index,
        # so a job gives the same episodes whatever the number of GPUs it is split over
        self.shard = shard
        if shard is not None:
            if int(n_envs) != shard.count:
                raise ValueError(f"n_envs {n_envs} != shard.count {shard.count}")
            if not shard.total:
                raise ValueError("shard.total is not set (use sharding.shard_envs / weak_shard)")
        self.E, self.A, self.M = int(n_envs), len(self.agent_ids), int(scenario.n_social)
        self._stats = {"agent_steps": 0.0, "episodes_done": 0.0}
        self.dt = float(fixed_timestep_sec)
        self._seed = int(seed)
        self._jitter = float(start_jitter)
        self._auto_reset = auto_reset
        self.cfg = compile_config(self.interface, self.E, self.A, self.M, self.dt, device=device,
                                  route_cap=route_cap, traj_len=traj_len, lidar_wall_height=lidar_wall_height,
                                  wp_out_len=wp_out_len,
                                  via_cap=max([len(scenario.missions[a].via) for a in self.agent_ids] + [int(min_via_cap)]))
        # ego_centric: the tensor path (std_obs / step_tensor) hands out ego-frame buffers (k_ego_frame); the dict
        # path keeps world-frame Observations -- use ego_centric.get_egocentric_adapters in the AgentSpec there
        self.ego_centric = bool(ego_centric)
        self.cfg.ego_frame = int(self.ego_centric)
        rays = lidar_params.base_rays(self.interface.lidar.sensor_params) if self.interface.lidar else None
        self._base_rays = rays
        self._engine: Optional[Engine] = Engine(self.cfg, self.road_map, rays)
        self._torch = self._engine._torch
        self._start = self._build_starts(self._seed)
        self._setup = False
        self._tick = np.zeros(self.E, np.int64)           # per-env step_count
        self._global_tick = 0
        self._done_agents = np.zeros((self.E, self.A), bool)
        self._aad = None                                  # [E] bool on the device: agents_alive criterion of the last tick
        self._host: Dict[str, np.ndarray] = {}

    _MASKABLE = ("neighborhood_vehicles", "lidar", "ogm", "drivable_area_grid_map", "accelerometer", "road_waypoints")

    @classmethod
    def _union_interface(cls, itfs: Sequence[AgentInterface]) -> AgentInterface:
        first = itfs[0]
        if all(i is first or repr(i) == repr(first) for i in itfs[1:]):
            return first
        union = {}
        for name in cls._MASKABLE:
            on = [getattr(i, name) for i in itfs if getattr(i, name)]
            if any(repr(c) != repr(on[0]) for c in on[1:]):
                raise NotImplementedError(f"agents enable {name} with different parameters: one configuration per "
                                          f"sensor and handle (use one env per interface)")
            union[name] = on[0] if on else getattr(first, name)
        rest = [dataclasses.replace(i, **union) for i in itfs]
        if any(repr(r) != repr(rest[0]) for r in rest[1:]):
            # (the waypoint sensor is not maskable: without it the trip meter reads another query, sensors.py:337-345)
            raise NotImplementedError("agent interfaces may differ in which of " + ", ".join(cls._MASKABLE) +
                                      " they enable, nothing else (use one env per interface)")
        return rest[0]

    def set_scenario(self, scenario: Scenario):
        """The next scenario of a scenario cycle (hiway_env.py:265-285 -> SMARTS.reset(scenario) -> SMARTS.setup,
        smarts.py:411-460): same agents and social-slot count (both are sizes of the handle), possibly another map,
        other missions and traffic.  Takes effect with the next ``reset`` -- which the library insists on when the
        map changed."""
        missing = [a for a in self.agent_ids if a not in scenario.missions]
        if missing:
            raise ValueError(f"scenario has no mission for agents {missing}")
        if int(scenario.n_social) != self.M:
            raise ValueError(f"scenario has {scenario.n_social} social slots, the handle was built for {self.M}")
        if max([len(scenario.missions[a].via) for a in self.agent_ids] + [0]) > self.cfg.via_cap:
            raise ValueError("scenario has more vias per mission than the handle was built for")
        rm = scenario.road_map()
        if rm is not self.road_map:
            self.engine.upload_map(rm)
            self.road_map = rm
            self._setup = False
        self.scenario = scenario
        self._start = self._build_starts(self._seed)

    @staticmethod
    def _own_adapters(spec: AgentSpec) -> AgentSpec:
        memo: Dict[int, Any] = {}  # shared: bound methods of one object stay bound to one (new) object
        return dataclasses.replace(spec, **{k: copy.deepcopy(getattr(spec, k), memo) for k in
                                            ("observation_adapter", "action_adapter", "reward_adapter", "info_adapter")})

    # -- gym-ish surface -------------------------------------------------------------------
    @property
    def batch_size(self) -> int:
        return self.E

    def seed(self, seed: int) -> List[int]:
        """parallel_env.py:246-256: env i gets seed + i."""
        self._seed = int(seed)
        self._start = self._build_starts(self._seed)
        first = self.shard.start if self.shard is not None else 0
        return [seed + first + i for i in range(self.E)]

    def close(self, terminate: bool = False):
        if self._engine is not None:
            self._engine.close()
            self._engine = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def engine(self):
        if self._engine is None:
            raise SMARTSDestroyedError("environment was closed")
        return self._engine

    # -- scenario -> reset arrays ------------------------------------------------------------
    def _build_starts(self, seed: int):
        rm, E, A = self.road_map, self.E, self.A
        rng = np.random.default_rng(seed)
        rc = self.cfg.route_cap
        x = np.zeros((E, A)); y = np.zeros((E, A)); h = np.zeros((E, A)); v = np.zeros((E, A))
        gx = np.zeros((E, A)); gy = np.zeros((E, A)); gr = np.full((E, A), -1.0)
        route = np.full((E, A, rc), -1, np.int32)
        rlen = np.zeros((E, A), np.int32)
        for a, aid in enumerate(self.agent_ids):
            m = self.scenario.missions[aid]
            r = rm.generate_route(m.begin_road, m.end_road)
            if not 0 < len(r.road_indices) <= rc:
                raise ValueError(f"route of {aid}: {len(r.road_indices)} roads (route_cap {rc})")
            route[:, a, :len(r.road_indices)] = r.road_indices
            rlen[:, a] = len(r.road_indices)
            if self._jitter > 0 and self.shard is not None:  # drawn for the whole job, this shard's slice kept
                jit = rng.uniform(0, self._jitter, self.shard.total)[self.shard.start:self.shard.stop]
            else:
                jit = rng.uniform(0, self._jitter, E) if self._jitter > 0 else np.zeros(E)
            offs = m.begin_offset + jit
            for e in range(E):
                x[e, a], y[e, a], h[e, a] = rm.lane_start_pose(m.begin_road, m.begin_lane, float(offs[e]))
            v[:, a] = m.start_speed
            if not m.endless:
                lane = rm.lane_of_road(m.end_road, m.end_lane)
                off = 0.5 * float(rm.lane_length[lane]) if m.end_offset is None else m.end_offset
                px, py, _ = rm.lane_start_pose(m.end_road, m.end_lane, off)
                gx[:, a], gy[:, a], gr[:, a] = px, py, m.goal_radius
        return dict(x=x, y=y, h=h, v=v, route=route, rlen=rlen, goal=(gx, gy, gr), vias=self._build_vias())

    def _build_vias(self):
        """Scenario.to_scenario_via (scenario.py:652-676): sstudio vias -> lane, position, hit distance."""
        C_ = self.cfg.via_cap
        if C_ <= 0:
            return None
        rm, E, A = self.road_map, self.E, self.A
        cnt = np.zeros((E, A), np.int32)
        lane = np.full((E, A, C_), -1, np.int32)
        vx = np.zeros((E, A, C_)); vy = np.zeros_like(vx); hit = np.zeros_like(vx); spd = np.zeros_like(vx)
        for a, aid in enumerate(self.agent_ids):
            for k, via in enumerate(self.scenario.missions[aid].via):
                l = rm.lane_of_road(via.road_id, via.lane_index)
                px, py, _ = rm.lane_start_pose(via.road_id, via.lane_index, float(via.lane_offset))
                lane[:, a, k], vx[:, a, k], vy[:, a, k] = l, px, py
                hit[:, a, k] = via.hit_distance if via.hit_distance > 0 else 0.5 * float(rm.lane_width[l])
                spd[:, a, k] = via.required_speed
            cnt[:, a] = len(self.scenario.missions[aid].via)
        return cnt, lane, vx, vy, hit, spd

    def _social(self, t: int) -> Optional[np.ndarray]:
        tr = self.scenario.traffic
        if self.M == 0 or tr is None:
            return None
        if callable(tr):
            s = tr(t)
        elif hasattr(tr, "states"):
            s = tr.states(t)
        else:
            s = tr[t % len(tr)]
        return np.ascontiguousarray(s, np.float64)

    def _reset_envs(self, env_ids: Optional[np.ndarray]):
        """SMARTS.reset of the listed envs (None: all).  An auto-reset (env_ids given) leaves the envs' social
        slots as the last injection wrote them: the recorded traffic is driven by the global tick and keeps
        running, so the reset observation still sees it."""
        s = self._start
        sel = slice(None) if env_ids is None else env_ids
        self.engine.reset(s["x"][sel], s["y"][sel], s["h"][sel], s["v"][sel], s["route"][sel], s["rlen"][sel],
                          env_ids=env_ids, goal=tuple(g[sel] for g in s["goal"]),
                          vias=None if s["vias"] is None else tuple(q[sel] for q in s["vias"]),
                          keep_social=env_ids is not None)
        self._tick[sel] = 0
        self._done_agents[sel] = False
        if self._aad is not None:
            if env_ids is None:
                self._aad = None
            else:
                self._aad[self._torch.from_numpy(np.asarray(env_ids, np.int64)).to(self._aad.device)] = False

    # -- device-resident step (the scalable API) ----------------------------------------------
    def reset_tensor(self) -> Dict[str, Any]:
        self._reset_envs(None)
        self._global_tick = 0
        soc = self._social(0)
        if soc is not None:
            self.engine.set_social(self._torch.from_numpy(soc).to(self.engine.device))
        self.engine.observe()
        self._setup = True
        return self.engine.obs()

    def step_tensor(self, actions, social=None) -> Dict[str, Any]:
        """actions: CUDA tensor in the ``sb_step`` layout; social: CUDA ``[E, M, 8]`` f64 or None (then the
        scenario's traffic source is uploaded).  Returns the zero-copy observation tensors
        (``ev_reward`` / ``ev_done`` included); envs whose agents are all done are reset in place when
        ``auto_reset`` is on, after the returned views were filled for this tick."""
        if not self._setup:
            raise SMARTSNotSetupError("reset() must be called before step()")
        eng = self.engine
        if social is None:
            soc = self._social(self._global_tick)
            if soc is not None:
                social = self._torch.from_numpy(soc).to(eng.device)
        if social is not None:
            eng.set_social(social)
        eng.step(actions)
        eng.observe()
        self._global_tick += 1
        self._tick += 1
        self._stats["agent_steps"] += float(self.E * self.A - int(self._done_agents.sum()))
        obs = eng.obs()
        if "ev_done" in obs:  # done agents leave the simulation after this observation (agent_manager teardown):
            alive = eng.state()["alive"]  # a device-side op on the zero-copy views, no synchronisation
            crit = self.interface.done_criteria.agents_alive
            if crit:  # DoneCriteria.agents_alive: a host-layer criterion over the agents that entered this tick alive
                before = alive[:, :self.A] != 0
                self._aad = self._agents_alive_done(before, crit)
                obs["ev_done"] |= (self._aad[:, None] & before).to(obs["ev_done"].dtype)
            alive[:, :self.A] &= (obs["ev_done"] == 0).to(alive.dtype)
        return obs

    def _agents_alive_done(self, alive, crit):
        """Sensors._agents_alive_done_check (sensors.py:404-441) per env, on the device: ``alive`` [E, A] bool =
        agent_manager.agent_ids at observation time.  Every agent here is an ego agent, so the ego and total
        minima read the same count.  Returns [E] bool (the criterion is the shared interface's: the same for every
        agent of an env)."""
        n = alive.sum(1)
        res = self._torch.zeros_like(n, dtype=self._torch.bool)
        if crit.minimum_ego_agents_alive:
            res |= n < int(crit.minimum_ego_agents_alive)
        if crit.minimum_total_agents_alive:
            res |= n < int(crit.minimum_total_agents_alive)
        for lst in crit.agent_lists_alive or ():
            assert isinstance(lst.agents_list, (list, set, tuple)), "Please specify a list of agent ids to watch"
            assert isinstance(lst.minimum_agents_alive_in_list, int) and lst.minimum_agents_alive_in_list >= 0
            idx = [self.agent_ids.index(i) for i in lst.agents_list if i in self.agent_ids]
            cnt = alive[:, idx].sum(1) if idx else self._torch.zeros_like(n)
            res |= cnt < lst.minimum_agents_alive_in_list
        return res

    def auto_reset_done_envs(self) -> np.ndarray:
        """Reset every env whose agents are all done (parallel_env.py:300-326); returns their ids."""
        done = self.engine.obs()["ev_done"]
        self._torch.cuda.synchronize(self.engine.device)
        d = done.cpu().numpy().astype(bool) | self._done_agents
        self._done_agents = d
        ids = np.nonzero(d.all(axis=1))[0].astype(np.int32)
        self._stats["episodes_done"] += float(len(ids))
        if len(ids) and self._auto_reset:
            self._reset_envs(ids)
            self.engine.observe(ids)  # only the reset envs: the others' sensor state and buffers stay as they are
        return ids

    @classmethod
    def sharded(cls, scenario: Scenario, agent_specs: Dict[str, AgentSpec], n_envs_total: int, **kw):
        """One process per GPU under torchrun: this rank's contiguous share of ``n_envs_total`` envs on
        ``cuda:LOCAL_RANK`` (sharding.shard_from_env; ParallelEnv's analogue is one process per env)."""
        import os
        from . import sharding
        sh = sharding.shard_from_env(int(n_envs_total))
        return cls(scenario, agent_specs, sh.count, device=int(os.environ.get("LOCAL_RANK", "0")), shard=sh, **kw)

    def episode_stats(self, reduce: bool = True) -> Dict[str, float]:
        """Agent-steps taken and episodes finished so far, summed over the ranks of the job (the one collective:
        a small all-reduce, sharding.reduce_stats)."""
        from . import sharding
        if not reduce:
            return dict(self._stats)
        return sharding.reduce_stats(self._stats, max_keys=(), device=self._torch.device("cuda", self.cfg.device))

    # -- StdObs: the dense, batched observation of format_obs.py, on the device -------------------
    def std_obs(self) -> Dict[str, Any]:
        """The current observation of every agent as ``StdObs`` (smarts/env/wrappers/format_obs.py:45-197):
        same keys, dtypes and trailing shapes, with a leading ``[E, A]`` batch and every tensor resident on the
        GPU.  The big members (grids, lidar point cloud, waypoints, neighbours) are the library's own buffers,
        not copies; the small ``ego`` / ``events`` members are assembled by a few torch ops.  Use
        ``wp_out_len=20`` at construction for the 4 x 20 waypoint block of StdObs."""
        t = self._torch
        o, s, itf, A = self.engine.obs(), self.engine.state(), self.interface, self.A
        ego = {k: s[k][:, :A] for k in ("pos_x", "pos_y", "pos_z", "heading", "speed", "yaw_rate", "length", "width", "height")}
        f32 = t.float32
        zeros3 = t.zeros((self.E, A, 3), dtype=f32, device=self.engine.device)
        lane = o["ego_lane"].long()
        lane_idx = t.as_tensor(self.road_map.lane_idx_in_road, device=self.engine.device)
        ev = o["ev_events"]
        bit = lambda b: ((ev & b) != 0).to(t.int8)  # noqa: E731
        acc = (lambda k: o[k].to(f32)) if self.cfg.accel_enable else (lambda k: zeros3)  # noqa: E731
        out: Dict[str, Any] = {
            # Observation.distance_travelled as the reference fills it: sensors.py:351 calls the trip meter with `sim`
            # in the `increment` position, so the member is this observation's INCREMENT (= the reward), not the total
            "dist": o["ev_reward"].to(f32),
            "ego": {
                "angular_acceleration": acc("acc_angular"), "angular_jerk": acc("acc_angular_jerk"),
                "linear_acceleration": acc("acc_linear"), "linear_jerk": acc("acc_linear_jerk"),
                "angular_velocity": t.stack([t.zeros_like(ego["yaw_rate"]), t.zeros_like(ego["yaw_rate"]), ego["yaw_rate"]], -1).to(f32),
                "box": t.stack([ego["length"], ego["width"], ego["height"]], -1).to(f32),
                "heading": ego["heading"].to(f32),
                "lane_index": t.where(lane >= 0, lane_idx[lane.clamp(min=0)], t.zeros_like(lane)).to(t.int8),
                "linear_velocity": t.stack([-t.sin(ego["heading"]) * ego["speed"], t.cos(ego["heading"]) * ego["speed"],
                                            t.zeros_like(ego["speed"])], -1).to(f32),
                "pos": t.stack([ego["pos_x"], ego["pos_y"], ego["pos_z"]], -1),
                "speed": ego["speed"].to(f32), "steering": s["steering"].to(f32), "yaw_rate": ego["yaw_rate"].to(f32),
            },
            "events": {"agents_alive_done": (t.zeros_like(ev, dtype=t.int8) if self._aad is None
                                             else self._aad[:, None].expand_as(ev).to(t.int8)), "collisions": bit(_abi.SB_EV_COLLISION),
                       "not_moving": bit(_abi.SB_EV_NOT_MOVING), "off_road": bit(_abi.SB_EV_OFF_ROAD),
                       "off_route": bit(_abi.SB_EV_OFF_ROUTE), "on_shoulder": bit(_abi.SB_EV_ON_SHOULDER),
                       "reached_goal": bit(_abi.SB_EV_REACHED_GOAL),
                       "reached_max_episode_steps": bit(_abi.SB_EV_REACHED_MAX_STEPS), "wrong_way": bit(_abi.SB_EV_WRONG_WAY)},
        }
        if itf.drivable_area_grid_map:
            out["dagm"] = o["dagm"].unsqueeze(-1)
        if itf.ogm:
            out["ogm"] = o["ogm"].unsqueeze(-1)
        if itf.lidar:  # misses are zeros in StdObs (format_obs.py:437-443): construct with lidar_miss_inf=0 or clean here
            pc = o["lidar_points"]
            if self.cfg.lidar_miss_inf:
                pc = t.nan_to_num(pc, nan=0.0, posinf=0.0, neginf=0.0)
            origin = out["ego"]["pos"].clone()
            origin[..., 2] += 1.0
            rays = t.as_tensor(self._base_rays, device=self.engine.device)
            n = rays.shape[0]
            org = origin.unsqueeze(2).expand(self.E, A, n, 3)
            end = org + rays
            if self.ego_centric:  # transform(ray_start), transform(ray_end), ego_centric_adapters.py:118-122
                org, end = self._to_ego_frame(org, ego), self._to_ego_frame(end, ego)
            out["lidar"] = {"hit": o["lidar_hits"], "point_cloud": pc, "ray_origin": org, "ray_vector": end}
        if itf.neighborhood_vehicles:
            out["neighbors"] = {"box": o["nbr_box"], "heading": o["nbr_heading"], "lane_index": o["nbr_lane_index"],
                                "pos": o["nbr_pos"], "speed": o["nbr_speed"]}
        if self.cfg.ttc_enable:  # zeros where the reference's _std_ttc returns None (o["ttc_valid"] == 0)
            out["ttc"] = {"angle_error": o["ttc_angle_error"], "distance_from_center": o["ttc_distance_from_center"],
                          "dtc": o["ttc_dtc"], "ttc": o["ttc_ttc"]}
        if itf.waypoints:
            out["waypoints"] = {"heading": o["wp_heading"], "lane_index": o["wp_lane_index"], "lane_width": o["wp_lane_width"],
                                "pos": o["wp_pos"], "speed_limit": o["wp_speed_limit"]}
        if self.ego_centric:  # ego_centric_adapters.py:140-147: the ego sits at the origin, facing +x
            g = out["ego"]
            for k in ("linear_velocity", "linear_acceleration", "linear_jerk"):  # ego_frame_dynamics
                v = g[k]
                g[k] = t.stack([t.linalg.vector_norm(v[..., :2], dim=-1), t.zeros_like(v[..., 0]), v[..., 2]], -1)
            g["pos"] = t.zeros_like(g["pos"])
            g["heading"] = t.zeros_like(g["heading"])
        return out

    def _to_ego_frame(self, pts, ego):
        """position_to_ego_frame (utils/math.py:473-487) for [E, A, n, 3] points, on the device."""
        t = self._torch
        h = -ego["heading"]
        c, s_ = t.cos(h)[..., None], t.sin(h)[..., None]
        rel = pts - t.stack([ego["pos_x"], ego["pos_y"], ego["pos_z"]], -1)[:, :, None, :]
        return t.stack([c * rel[..., 0] - s_ * rel[..., 1], s_ * rel[..., 0] + c * rel[..., 1], rel[..., 2]], -1)

    # -- host-side Observation objects ---------------------------------------------------------
    def _fetch(self):
        self._host = self.engine.obs_numpy()
        self._host_state = self.engine.state_numpy()
        self._host_aad = self._aad.cpu().numpy() if self._aad is not None else np.zeros(self.E, bool)

    def _vehicle_id(self, e: int, slot: int) -> str:
        return self.agent_ids[slot] if slot < self.A else f"social-{slot - self.A}"

    def _observation(self, e: int, a: int) -> Observation:
        o, s, itf, rm = self._host, self._host_state, self._own_itf[a], self.road_map
        px, py, pz = s["pos_x"][e, a], s["pos_y"][e, a], s["pos_z"][e, a]
        hd, sp = float(s["heading"][e, a]), float(s["speed"][e, a])
        el = int(o["ego_lane"][e, a])
        lane_id = rm.lane_ids[el] if el >= 0 else None
        road_id = rm.road_ids[int(rm.lane_road[el])] if el >= 0 else None
        lane_index = int(rm.lane_idx_in_road[el]) if el >= 0 else None
        ego = EgoVehicleObservation(
            id=self.agent_ids[a], position=np.array([px, py, pz]),
            bounding_box=Dimensions(float(s["length"][e, a]), float(s["width"][e, a]), float(s["height"][e, a])),
            heading=hd, speed=sp, steering=float(s["steering"][e, a]), yaw_rate=float(s["yaw_rate"][e, a]),
            road_id=road_id, lane_id=lane_id, lane_index=lane_index, mission=self.scenario.missions[self.agent_ids[a]],
            linear_velocity=np.array([-math.sin(hd) * sp, math.cos(hd) * sp, 0.0]),
            angular_velocity=np.array([0.0, 0.0, float(s["yaw_rate"][e, a])]),
            linear_acceleration=o["acc_linear"][e, a].copy() if itf.accelerometer else None,
            angular_acceleration=o["acc_angular"][e, a].copy() if itf.accelerometer else None,
            linear_jerk=o["acc_linear_jerk"][e, a].copy() if itf.accelerometer else None,
            angular_jerk=o["acc_angular_jerk"][e, a].copy() if itf.accelerometer else None)
        nbrs = None
        if itf.neighborhood_vehicles:
            nbrs = []
            for k in range(min(int(o["nbr_count"][e, a]), self.cfg.nbr_cap)):
                nl = int(o["nbr_lane"][e, a, k])
                nbrs.append(VehicleObservation(
                    id=self._vehicle_id(e, int(o["nbr_id"][e, a, k])), position=tuple(o["nbr_pos"][e, a, k]),
                    bounding_box=Dimensions(*[float(t) for t in o["nbr_box"][e, a, k]]),
                    heading=float(o["nbr_heading"][e, a, k]), speed=float(o["nbr_speed"][e, a, k]),
                    road_id=rm.road_ids[int(rm.lane_road[nl])] if nl >= 0 else None,
                    lane_id=rm.lane_ids[nl] if nl >= 0 else None,
                    lane_index=int(o["nbr_lane_index"][e, a, k]) if nl >= 0 else None))
        wps = None
        if itf.waypoints:
            wps = []
            for p in range(self.cfg.wp_out_paths):
                n = int(o["wp_len"][e, a, p])
                if n == 0:
                    continue
                wps.append([Waypoint(pos=o["wp_pos"][e, a, p, i, :2].copy(), heading=float(o["wp_heading"][e, a, p, i]),
                                     lane_id=rm.lane_ids[int(o["wp_lane"][e, a, p, i])],
                                     lane_width=float(o["wp_lane_width"][e, a, p, i]),
                                     speed_limit=float(o["wp_speed_limit"][e, a, p, i]),
                                     lane_index=int(o["wp_lane_index"][e, a, p, i])) for i in range(n)])
        rwps = None
        if itf.road_waypoints:  # sensors.py:1000-1011: {lane_id: [[Waypoint]]} in the sensor's insertion order
            lanes = {rm.lane_ids[int(l)]: [] for l in o["rw_lane"][e, a, :min(int(o["rw_n_lanes"][e, a]), self.cfg.rw_lanes_cap)]}
            keys = list(lanes)
            for p in range(min(int(o["rw_n_paths"][e, a]), self.cfg.rw_paths_cap)):
                n = int(o["rw_len"][e, a, p])
                lanes[keys[int(o["rw_path_slot"][e, a, p])]].append(
                    [Waypoint(pos=o["rw_pos"][e, a, p, i].copy(), heading=float(o["rw_heading"][e, a, p, i]),
                              lane_id=rm.lane_ids[int(o["rw_wp_lane"][e, a, p, i])],
                              lane_width=float(o["rw_lane_width"][e, a, p, i]),
                              speed_limit=float(o["rw_speed_limit"][e, a, p, i]),
                              lane_index=int(o["rw_lane_index"][e, a, p, i])) for i in range(n)])
            rwps = RoadWaypoints(lanes=lanes)
        lidar = None
        if itf.lidar:  # (points, hits, [(origin, end)]) as lidar.py:115-134
            origin = np.array([px, py, pz + 1.0])
            pts = o["lidar_points"][e, a]
            hits = o["lidar_hits"][e, a].astype(bool)
            lidar = ([p for p in pts], [bool(t) for t in hits], [(origin, origin + r) for r in self._base_rays])

        def grid(key, cfg_w, cfg_h, res, cls):
            meta = GridMapMetadata(created_at=int(self._tick[e]), resolution=res, width=cfg_w, height=cfg_h,
                                   camera_pos=(px, py, 20.0 * max(s["length"][e, a], s["width"][e, a], s["height"][e, a])),
                                   camera_heading_in_degrees=math.degrees(hd))
            return cls(metadata=meta, data=o[key][e, a][..., None])
        ogm = grid("ogm", itf.ogm.width, itf.ogm.height, itf.ogm.resolution, OccupancyGridMap) if itf.ogm else None
        dagm = (grid("dagm", itf.drivable_area_grid_map.width, itf.drivable_area_grid_map.height,
                     itf.drivable_area_grid_map.resolution, DrivableAreaGridMap)
                if itf.drivable_area_grid_map else None)
        bits = int(o["ev_events"][e, a])
        cols = []
        if bits & _abi.SB_EV_COLLISION:
            cols = [Collision(collidee_id=self._vehicle_id(e, int(o["ev_collidee"][e, a])))]
        events = Events(collisions=cols, off_road=bool(bits & _abi.SB_EV_OFF_ROAD),
                        off_route=bool(bits & _abi.SB_EV_OFF_ROUTE), on_shoulder=bool(bits & _abi.SB_EV_ON_SHOULDER),
                        wrong_way=bool(bits & _abi.SB_EV_WRONG_WAY), not_moving=bool(bits & _abi.SB_EV_NOT_MOVING),
                        reached_goal=bool(bits & _abi.SB_EV_REACHED_GOAL),
                        reached_max_episode_steps=bool(bits & _abi.SB_EV_REACHED_MAX_STEPS),
                        agents_alive_done=bool(self._host_aad[e]))
        return Observation(dt=self.dt, step_count=int(self._tick[e]), elapsed_sim_time=float(self._tick[e]) * self.dt,
                           events=events, ego_vehicle_state=ego, neighborhood_vehicle_states=nbrs,
                           waypoint_paths=wps,
                           # sensors.py:351 `vehicle.trip_meter_sensor(sim)`: `sim` lands in `increment`, so the
                           # reference's Observation holds the increment; the total is infos["score"] (agent_manager.py:234)
                           distance_travelled=float(o["ev_reward"][e, a]),
                           lidar_point_cloud=lidar, drivable_area_grid_map=dagm, occupancy_grid_map=ogm,
                           top_down_rgb=None, road_waypoints=rwps, via_data=self._vias(e, a))

    def _vias(self, e: int, a: int) -> Vias:
        """Observation.via_data (sensors.py:337-347) from the kernel's index lists."""
        o = self._host
        if "via_near_idx" not in o:
            return Vias([], [])
        mv = self.scenario.missions[self.agent_ids[a]].via
        pos = self._start["vias"]

        def point(k):
            return ViaPoint((float(pos[2][e, a, k]), float(pos[3][e, a, k])), mv[k].lane_index, mv[k].road_id,
                            mv[k].required_speed)
        near = [point(int(k)) for k in o["via_near_idx"][e, a, :int(o["via_near_count"][e, a])]]
        hit = [point(int(k)) for k in o["via_hit_idx"][e, a, :int(o["via_hit_count"][e, a])]]
        return Vias(near, hit)

    def _env_observations(self, e: int) -> Dict[str, Observation]:
        return {aid: self._observation(e, a) for a, aid in enumerate(self.agent_ids) if not self._done_agents[e, a]}

    # -- ParallelEnv surface -------------------------------------------------------------------
    def reset(self) -> List[Dict[str, Any]]:
        self.reset_tensor()
        self._fetch()
        return [{aid: self._spec(e, aid).observation_adapter(ob) for aid, ob in self._env_observations(e).items()}
                for e in range(self.E)]

    def _spec(self, e: int, aid: str) -> AgentSpec:
        return self._env_specs[e if self.E > 1 else 0][aid]

    def step(self, actions: Sequence[Dict[str, Any]]):
        """actions: one ``{agent_id: action}`` dict per env.  Returns (observations, rewards, dones, infos),
        each a list over envs (parallel_env.py:188-216).  Agents that are done drop out of the dicts
        (agent_manager.py:161-240); an env whose agents are all done raises ``dones["__all__"]`` and,
        with ``auto_reset``, restarts -- its observation entry is then the reset observation."""
        if self._engine is None:
            raise SMARTSDestroyedError("environment was closed")
        if not self._setup:
            raise SMARTSNotSetupError("reset() must be called before step()")
        if len(actions) != self.E:
            raise ValueError(f"expected {self.E} action dicts, got {len(actions)}")
        space = self.interface.action
        flat: List[Any] = []
        default = {ActionSpaceType.Lane: "keep_lane", ActionSpaceType.LaneWithContinuousSpeed: (0.0, 0)}.get(space)
        for e, env_act in enumerate(actions):
            assert isinstance(env_act, dict) and all(isinstance(k, str) for k in env_act), "Expected Dict[str, any]"
            for a, aid in enumerate(self.agent_ids):
                if aid in env_act:
                    flat.append(self._spec(e, aid).action_adapter(env_act[aid]))
                elif self._done_agents[e, a] or default is not None:
                    flat.append(default if default is not None else self._neutral_action())
                else:
                    raise KeyError(f"env {e}: no action for live agent {aid}")
        arr = encode_actions(space, flat, self.cfg.traj_len)
        try:
            self.step_tensor(self._torch.from_numpy(arr).to(self.engine.device))
            self._fetch()
        except Exception:
            self.close()  # smarts.py:214-227: tear down, re-raise
            raise
        obs_l, rew_l, done_l, info_l = [], [], [], []
        newly = self._host["ev_done"].astype(bool) & ~self._done_agents
        for e in range(self.E):
            obs_e, rew_e, done_e, info_e = {}, {}, {}, {}
            for a, aid in enumerate(self.agent_ids):
                if self._done_agents[e, a]:
                    continue
                spec = self._spec(e, aid)
                ob = self._observation(e, a)
                r = float(self._host["ev_reward"][e, a])
                info = {"score": float(self._host["ev_distance"][e, a]), "env_obs": ob}
                rew_e[aid] = spec.reward_adapter(ob, r)
                obs_e[aid] = spec.observation_adapter(ob)
                info_e[aid] = spec.info_adapter(ob, r, info)
                done_e[aid] = bool(newly[e, a])
            self._done_agents[e] |= newly[e]
            done_e["__all__"] = bool(self._done_agents[e].all())
            obs_l.append(obs_e); rew_l.append(rew_e); done_l.append(done_e); info_l.append(info_e)
        # (done agents left the simulation inside step_tensor: vehicle removed, agent_manager teardown)
        finished = np.nonzero(self._done_agents.all(axis=1))[0].astype(np.int32)
        self._stats["episodes_done"] += float(len(finished))
        if len(finished) and self._auto_reset:
            self._reset_envs(finished)
            self.engine.observe(finished)  # the reset envs only (ParallelEnv envs are independent)
            self._fetch()
            for e in finished:
                obs_l[e] = {aid: self._spec(int(e), aid).observation_adapter(ob)
                            for aid, ob in self._env_observations(int(e)).items()}
        return obs_l, rew_l, done_l, info_l

    def _neutral_action(self):
        if self.interface.action == ActionSpaceType.Trajectory:
            return np.zeros((4, self.cfg.traj_len))
        return (0.0, 0.0, 0.0)


class HiWayEnv:
    """The single-environment ``hiway-v0`` surface (hiway_env.py:43-291): dict in, dicts out."""

    metadata = {"render.modes": ["human"]}

    def __init__(self, scenarios: Sequence[Scenario], agent_specs: Dict[str, AgentSpec], sim_name: Optional[str] = None,
                 shuffle_scenarios: bool = True, headless: bool = True, fixed_timestep_sec: Optional[float] = None,
                 seed: int = 42, timestep_sec: Optional[float] = None, **kwargs):
        if not scenarios:
            raise ValueError("scenarios is empty")
        dt = fixed_timestep_sec or timestep_sec or 0.1  # hiway_env.py:112-113
        self._agent_specs = agent_specs
        # Scenario.scenario_variations (scenario.py:137-166): the roots shuffled once, then cycled; reset() takes the next
        order = list(scenarios)
        if shuffle_scenarios:
            np.random.default_rng(seed).shuffle(order)
        self._scenarios_iterator = itertools.cycle(order)
        via_cap = max(len(m.via) for sc in order for m in sc.missions.values()) if any(sc.missions for sc in order) else 0
        self._batched: Optional[BatchedHiWayEnv] = BatchedHiWayEnv(
            order[0], agent_specs, n_envs=1, fixed_timestep_sec=dt, seed=seed, auto_reset=False,
            min_via_cap=via_cap, **kwargs)
        self._dones_registered = 0

    @property
    def agent_specs(self):
        return self._agent_specs

    @property
    def batched(self) -> BatchedHiWayEnv:
        if self._batched is None:
            raise SMARTSDestroyedError("environment was closed")
        return self._batched

    def seed(self, seed: int) -> int:
        self.batched.seed(seed)
        return seed

    def reset(self) -> Dict[str, Any]:
        """hiway_env.py:265-285: the next scenario of the cycle, then SMARTS.reset."""
        scenario = next(self._scenarios_iterator)
        if scenario is not self.batched.scenario:
            self.batched.set_scenario(scenario)
        self._dones_registered = 0
        return self.batched.reset()[0]

    def step(self, agent_actions: Dict[str, Any]):
        obs, rew, done, info = self.batched.step([agent_actions])
        observations, rewards, dones, infos = obs[0], rew[0], done[0], info[0]
        dones.pop("__all__", None)
        for d in dones.values():
            self._dones_registered += 1 if d else 0
        dones["__all__"] = self._dones_registered >= len(self._agent_specs)  # hiway_env.py:258-261
        return observations, rewards, dones, infos

    def render(self, mode="human"):
        pass

    def close(self):
        if self._batched is not None:
            self._batched.close()
            self._batched = None


def make(id: str = "smarts.env:hiway-v0", **kwargs) -> HiWayEnv:
    """``gym.make("smarts.env:hiway-v0", scenarios=..., agent_specs=...)`` (smarts/env/__init__.py:20-23);
    gym itself is not needed for the call."""
    if id.split(":")[-1] != "hiway-v0":
        raise ValueError(f"unknown environment id {id!r}")
    return HiWayEnv(**kwargs)
