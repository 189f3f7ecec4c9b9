"""Thin ctypes loader for ``libsmartsb200.so`` (the C-ABI of include/smarts_b200.h).

PyTorch is used only to own device memory for caller-side tensors (actions, social states) and
to wrap the library's device buffers zero-copy (``__cuda_array_interface__``).  There is no CPU
fallback: if the shared library is missing or CUDA is unavailable, construction raises.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Dict, Optional

import numpy as np

from . import _abi
from .road_map import RoadMapData

_LIB = None
# SMARTS_B200_LIB: a variant build of the same sources (kernel A/B experiments, profiles/README.md)
_LIB_PATH = os.environ.get("SMARTS_B200_LIB") or os.path.join(os.path.dirname(os.path.abspath(__file__)), "libsmartsb200.so")

EXPORTS = [
    "sb_abi_version", "sb_create", "sb_destroy", "sb_upload_map", "sb_upload_lidar_rays", "sb_reset",
    "sb_set_social", "sb_step", "sb_observe", "sb_observe_envs", "sb_step_host", "sb_copy_obs_to_host", "sb_get_obs",
    "sb_get_state", "sb_launch_count", "sb_last_error", "sb_set_timing", "sb_collect_timing",
    "sb_set_incremental_host_grids", "sb_copy_obs_to_host_packed", "sb_copy_grids_to_host_packed_noalloc",
]
KERNEL_KINDS = ["k_set_social", "k_control", "k_observe", "k_lidar", "k_raster", "k_events", "k_pack_bits", "k_ttc", "k_extras", "k_ego_frame", "k_road_waypoints"]


class SmartsB200Error(RuntimeError):
    pass


def load_library(path: Optional[str] = None):
    """dlopen the CUDA library and declare every prototype.  Raises if it is not built."""
    global _LIB
    if _LIB is not None and path is None:
        return _LIB
    p = path or _LIB_PATH
    if not os.path.exists(p):
        raise SmartsB200Error(
            f"{p} not found: build it with `python -m smarts_b200.build` (nvcc, sm_100a). "
            "smarts_b200 has no CPU fallback.")
    L = C.CDLL(p)
    H = C.c_void_p
    L.sb_abi_version.restype = C.c_int
    L.sb_create.argtypes = [C.POINTER(_abi.SbConfig), C.POINTER(H)]
    L.sb_destroy.argtypes = [H]
    L.sb_destroy.restype = None
    L.sb_upload_map.argtypes = [H, C.POINTER(_abi.SbMap)]
    L.sb_upload_lidar_rays.argtypes = [H, C.POINTER(C.c_double), C.c_int32]
    L.sb_reset.argtypes = [H, C.POINTER(C.c_int32), C.c_int32, C.POINTER(_abi.SbResetState)]
    L.sb_set_social.argtypes = [H, C.c_void_p, C.c_void_p]
    L.sb_step.argtypes = [H, C.c_void_p, C.c_void_p]
    L.sb_observe.argtypes = [H, C.c_void_p]
    L.sb_observe_envs.argtypes = [H, C.POINTER(C.c_int32), C.c_int32, C.c_void_p]
    L.sb_step_host.argtypes = [H, C.c_void_p, C.c_int64, C.c_void_p]
    L.sb_copy_obs_to_host.argtypes = [H, C.POINTER(_abi.SbObsPtrs)]
    L.sb_copy_obs_to_host_packed.argtypes = [H, C.POINTER(_abi.SbObsPtrs)]
    L.sb_copy_obs_to_host_packed.restype = C.c_int
    L.sb_copy_grids_to_host_packed_noalloc.argtypes = [H, C.c_void_p, C.c_void_p]
    L.sb_copy_grids_to_host_packed_noalloc.restype = C.c_int
    L.sb_get_obs.argtypes = [H, C.POINTER(_abi.SbObsPtrs)]
    L.sb_get_state.argtypes = [H, C.POINTER(_abi.SbStatePtrs)]
    L.sb_launch_count.argtypes = [H]
    L.sb_launch_count.restype = C.c_int64
    L.sb_last_error.argtypes = [H]
    L.sb_last_error.restype = C.c_char_p
    L.sb_set_timing.argtypes = [H, C.c_int]
    L.sb_set_timing.restype = C.c_int
    L.sb_collect_timing.argtypes = [H, C.POINTER(C.c_double), C.POINTER(C.c_int64)]
    L.sb_collect_timing.restype = C.c_int
    L.sb_set_incremental_host_grids.argtypes = [H, C.c_int]
    L.sb_set_incremental_host_grids.restype = C.c_int
    for f in ("sb_create", "sb_upload_map", "sb_upload_lidar_rays", "sb_reset", "sb_set_social", "sb_step",
              "sb_observe", "sb_observe_envs", "sb_step_host", "sb_copy_obs_to_host", "sb_get_obs", "sb_get_state"):
        getattr(L, f).restype = C.c_int
    if L.sb_abi_version() != _abi.SB_ABI_VERSION:
        raise SmartsB200Error("libsmartsb200.so ABI version mismatch; rebuild")
    if path is None:
        _LIB = L
    return L


class _DevArray:
    """Minimal ``__cuda_array_interface__`` carrier for a library-owned device buffer."""

    def __init__(self, ptr: int, shape, dtype, owner):
        self._owner = owner
        self.__cuda_array_interface__ = {
            "shape": tuple(int(s) for s in shape), "typestr": np.dtype(dtype).str,
            "data": (int(ptr), False), "version": 2, "strides": None,
        }


class TickGraph:
    """See ``Engine.capture_tick``."""

    def __init__(self, eng: "Engine", actions, social):
        torch = eng._torch
        if eng.M and social is None:
            raise ValueError("this engine has social vehicles: pass the tensor their states are written to")
        self.eng, self.actions, self.social = eng, actions, social
        before = int(eng._lib.sb_launch_count(eng._h))
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            if social is not None:
                eng.set_social(social)
            eng.step(actions)
            eng.observe()
        self.launches = int(eng._lib.sb_launch_count(eng._h)) - before
        eng._graph_launches -= self.launches  # the capture itself launched nothing

    def replay(self):
        self.graph.replay()
        self.eng._graph_launches += self.launches


class Engine:
    """One handle = E batched environments on one GPU."""

    def __init__(self, cfg: _abi.SbConfig, road_map: RoadMapData, base_rays: Optional[np.ndarray] = None):
        import torch
        if not torch.cuda.is_available():
            raise SmartsB200Error("CUDA device required: smarts_b200 has no CPU fallback")
        self._torch = torch
        self._lib = load_library()
        self.cfg = cfg
        self.road_map = road_map
        self.E, self.A, self.M = cfg.n_envs, cfg.n_agents, cfg.n_social
        self.V = self.A + self.M
        self.device = torch.device("cuda", cfg.device)
        self._h = C.c_void_p()
        rc = self._lib.sb_create(C.byref(cfg), C.byref(self._h))
        self._check(rc, "sb_create")
        self.upload_map(road_map)
        if cfg.lidar_enable:
            if base_rays is None:
                raise SmartsB200Error("lidar enabled but no base ray table given")
            r = np.ascontiguousarray(base_rays, np.float64)
            self._check(self._lib.sb_upload_lidar_rays(self._h, r.ctypes.data_as(C.POINTER(C.c_double)), len(r)),
                        "sb_upload_lidar_rays")
        self._obs = None
        self._state = None
        self._graph_launches = 0

    # -- plumbing ------------------------------------------------------------------------
    def upload_map(self, road_map: RoadMapData):
        """Load (or, on a live handle, replace) the road map.  After a replacement every env must be reset
        (``reset`` without ``env_ids``) before the next step / observe -- the library refuses anything else."""
        if self.cfg.rw_enable:
            road_map.require_road_waypoint_tables()
        m, keep = _abi.make_map_struct(road_map.arrays(), road_map.spacing)
        self._check(self._lib.sb_upload_map(self._h, C.byref(m)), "sb_upload_map")
        del keep
        self.road_map = road_map

    def _check(self, rc: int, what: str):
        if rc != 0:
            msg = self._lib.sb_last_error(self._h)
            raise SmartsB200Error(f"{what} failed ({rc}): {msg.decode() if msg else ''}")

    def close(self):
        if getattr(self, "_h", None) is not None and self._h:
            self._obs = self._state = None
            self._lib.sb_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _wrap(self, ptr, shape, dtype):
        addr = C.cast(ptr, C.c_void_p).value
        if np.dtype(dtype) == np.uint32:  # torch has no first-class uint32; same bits
            dtype = np.int32
        return self._torch.as_tensor(_DevArray(addr, shape, dtype, self), device=self.device)

    @property
    def launch_count(self) -> int:
        return int(self._lib.sb_launch_count(self._h)) + self._graph_launches

    def set_timing(self, on: bool):
        self._check(self._lib.sb_set_timing(self._h, int(on)), "sb_set_timing")

    def collect_timing(self) -> Dict[str, tuple]:
        """{kernel: (total ms, launches)} since the last collect (synchronises)."""
        ms = (C.c_double * len(KERNEL_KINDS))()
        cnt = (C.c_int64 * len(KERNEL_KINDS))()
        self._check(self._lib.sb_collect_timing(self._h, ms, cnt), "sb_collect_timing")
        return {k: (ms[i], int(cnt[i])) for i, k in enumerate(KERNEL_KINDS)}

    def _stream_ptr(self):
        return C.c_void_p(self._torch.cuda.current_stream(self.device).cuda_stream)

    # -- the step ------------------------------------------------------------------------
    def reset(self, x, y, heading, speed, route=None, route_len=None, env_ids=None, goal=None, vias=None,
              keep_social=False):
        """Place agents (arrays [n, A]); ``route`` int32 [n, A, route_cap] road indices (-1 pad);
        ``goal`` = (x, y, radius) arrays [n, A] of PositionalGoals (radius < 0: EndlessGoal);
        ``vias`` = (count [n, A], lane [n, A, via_cap], x, y, hit_distance, required_speed) of Mission.via."""
        f = lambda a: np.ascontiguousarray(a, np.float64)  # noqa: E731
        x, y, heading, speed = f(x), f(y), f(heading), f(speed)
        dp = lambda a: a.ctypes.data_as(C.POINTER(C.c_double))  # noqa: E731
        ip = lambda a: a.ctypes.data_as(C.POINTER(C.c_int32))  # noqa: E731
        rs = _abi.SbResetState()
        rs.keep_social = int(keep_social)
        rs.agent_x, rs.agent_y, rs.agent_heading, rs.agent_speed = dp(x), dp(y), dp(heading), dp(speed)
        if route is not None:
            route = np.ascontiguousarray(route, np.int32)
            route_len = np.ascontiguousarray(route_len, np.int32)
            rs.route, rs.route_len = ip(route), ip(route_len)
        if goal is not None:
            gx, gy, gr = (f(g) for g in goal)
            assert gx.size == gy.size == gr.size == x.size
            rs.goal_x, rs.goal_y, rs.goal_radius = dp(gx), dp(gy), dp(gr)
        if vias is not None:
            vc, vl = (np.ascontiguousarray(v, np.int32) for v in vias[:2])
            vx, vy, vh, vs = (f(v) for v in vias[2:])
            assert vl.size == vx.size == vy.size == vh.size == vs.size == vc.size * self.cfg.via_cap
            rs.via_count, rs.via_lane = ip(vc), ip(vl)
            rs.via_x, rs.via_y, rs.via_hit_distance, rs.via_required_speed = dp(vx), dp(vy), dp(vh), dp(vs)
        ids, n = None, self.E
        if env_ids is not None:
            env_ids = np.ascontiguousarray(env_ids, np.int32)
            ids, n = ip(env_ids), len(env_ids)
        assert x.size == n * self.A, (x.shape, n, self.A)
        self._check(self._lib.sb_reset(self._h, ids, n, C.byref(rs)), "sb_reset")

    def set_social(self, states):
        """states: CUDA float64 tensor [E, M, 8] (x, y, heading, speed, length, width, height, alive)."""
        if self.M == 0:
            return
        assert states.is_cuda and states.dtype == self._torch.float64 and states.is_contiguous()
        assert tuple(states.shape) == (self.E, self.M, 8), states.shape
        self._check(self._lib.sb_set_social(self._h, C.c_void_p(states.data_ptr()), self._stream_ptr()), "sb_set_social")

    def step(self, actions):
        """actions: CUDA tensor laid out as the action space demands (see smarts_b200.h sb_step)."""
        assert actions.is_cuda and actions.is_contiguous()
        self._check(self._lib.sb_step(self._h, C.c_void_p(actions.data_ptr()), self._stream_ptr()), "sb_step")

    def observe(self, env_ids=None):
        """Sensors.observe for every agent, or for the agents of ``env_ids`` only (the reset observation of an
        auto-reset: the other envs' sensor state and observation buffers stay untouched)."""
        if env_ids is None:
            self._check(self._lib.sb_observe(self._h, self._stream_ptr()), "sb_observe")
            return
        ids = np.ascontiguousarray(env_ids, np.int32)
        self._check(self._lib.sb_observe_envs(self._h, ids.ctypes.data_as(C.POINTER(C.c_int32)), len(ids),
                                              self._stream_ptr()), "sb_observe_envs")

    def capture_tick(self, actions, social=None) -> "TickGraph":
        """One tick (set_social + step + observe) captured into a CUDA graph over the given device tensors: the
        graph reads ``actions`` / ``social`` at replay time, so write the next tick's values into the same
        tensors and call ``.replay()``.  For small batches the tick is launch-bound (figure_eight x 256 envs:
        three kernels of ~0.04 ms); the graph removes the per-launch host cost.  Run one eager tick first so
        that every kernel's module is loaded outside the capture."""
        return TickGraph(self, actions, social)

    def step_host(self, actions: np.ndarray, social: Optional[np.ndarray]):
        """The reference-facing call with HOST buffers: H2D copies + set_social + step + observe + sync."""
        a = np.ascontiguousarray(actions)
        s = None
        if self.M:
            s = np.ascontiguousarray(social, np.float64)
            assert s.shape == (self.E, self.M, 8)
        self._check(self._lib.sb_step_host(self._h, a.ctypes.data_as(C.c_void_p), a.nbytes,
                                           s.ctypes.data_as(C.c_void_p) if s is not None else None), "sb_step_host")

    def host_obs_buffers(self, pinned: bool = True, packed_grids: bool = False) -> Dict[str, "np.ndarray"]:
        """Allocate host destination buffers (pinned by default) for ``copy_obs_to_host``; with ``packed_grids``
        the grids are [E, A, H, W // 8] uint8 bit images for ``copy_obs_to_host(..., packed=True)``."""
        out = {}
        for name, (shape, dt) in _abi.obs_shapes(self.cfg).items():
            if packed_grids and name in ("ogm", "dagm"):
                if (shape[0] * shape[1]) % 32 or shape[1] % 8:
                    raise SmartsB200Error("packed grids need a width that is a multiple of 8 and W * H a multiple of 32")
                shape = (shape[0], shape[1] // 8)
            full = (self.E, self.A) + tuple(shape)
            t = self._torch.empty(full, dtype=getattr(self._torch, np.dtype(dt).name), pin_memory=pinned)
            out[name] = t
        return out

    def set_incremental_host_grids(self, on: bool = True):
        """Promise that the same host grid buffers are passed to ``copy_obs_to_host`` every time and are not
        written to in between: only the 64-pixel words that changed are rewritten (smarts_b200.h)."""
        self._check(self._lib.sb_set_incremental_host_grids(self._h, int(on)), "sb_set_incremental_host_grids")

    @staticmethod
    def unpack_grid(bits: np.ndarray) -> np.ndarray:
        """[..., H, W // 8] bit image of ``copy_obs_to_host(packed=True)`` -> the [..., H, W] 0 / 255 uint8 image."""
        return np.unpackbits(np.asarray(bits), axis=-1, bitorder="little") * np.uint8(255)

    def copy_obs_to_host(self, bufs: Dict[str, "object"], packed: bool = False):
        """``packed``: the grids are delivered as 1-bit images (sb_copy_obs_to_host_packed; ``unpack_grid`` widens)."""
        p = _abi.SbObsPtrs()
        for name, t in bufs.items():
            field_t = dict(_abi.SbObsPtrs._fields_)[name]
            setattr(p, name, C.cast(C.c_void_p(t.data_ptr()), field_t))
        if packed:
            self._check(self._lib.sb_copy_obs_to_host_packed(self._h, C.byref(p)), "sb_copy_obs_to_host_packed")
        else:
            self._check(self._lib.sb_copy_obs_to_host(self._h, C.byref(p)), "sb_copy_obs_to_host")

    def copy_grids_to_host_packed_noalloc(self, ogm_bits=None, dagm_bits=None):
        """The 1-bit grid images of ``copy_obs_to_host(packed=True)`` without a single allocation inside the library
        (sb_copy_grids_to_host_packed_noalloc): for processes whose allocator already owns all of HBM, loops that
        cannot afford cudaMalloc's device-wide synchronisation, or code running beside another thread's stream
        capture.  ``ogm_bits`` / ``dagm_bits``: host uint8 tensors [E, A, H, W // 8] (``host_obs_buffers(packed_grids=
        True)``); ``None`` skips that grid.  The staging array is 16 MiB per GPU and shared by every engine on it:
        larger batches raise, and overlapping calls from two threads must be serialised by the caller."""
        ptr = lambda t: None if t is None else C.c_void_p(t.data_ptr())
        self._check(self._lib.sb_copy_grids_to_host_packed_noalloc(self._h, ptr(ogm_bits), ptr(dagm_bits)),
                    "sb_copy_grids_to_host_packed_noalloc")

    # -- zero-copy views -----------------------------------------------------------------
    def obs(self) -> Dict[str, "object"]:
        if self._obs is None:
            p = _abi.SbObsPtrs()
            self._check(self._lib.sb_get_obs(self._h, C.byref(p)), "sb_get_obs")
            self._obs = {name: self._wrap(getattr(p, name), (self.E, self.A) + tuple(shape), dt)
                         for name, (shape, dt) in _abi.obs_shapes(self.cfg).items()}
        return self._obs

    def state(self) -> Dict[str, "object"]:
        if self._state is None:
            p = _abi.SbStatePtrs()
            self._check(self._lib.sb_get_state(self._h, C.byref(p)), "sb_get_state")
            out = {}
            for name, (per, shape, dt) in _abi.state_shapes(self.cfg).items():
                lead = (self.E, self.V) if per == "v" else (self.E, self.A)
                out[name] = self._wrap(getattr(p, name), lead + tuple(shape), dt)
            self._state = out
        return self._state

    def obs_numpy(self) -> Dict[str, np.ndarray]:
        self._torch.cuda.synchronize(self.device)
        return {k: v.cpu().numpy() for k, v in self.obs().items()}

    def state_numpy(self) -> Dict[str, np.ndarray]:
        self._torch.cuda.synchronize(self.device)
        return {k: v.cpu().numpy() for k, v in self.state().items()}
