"""sb_copy_grids_to_host_packed_noalloc -- packed OGM / drivable-area delivery for callers that cannot allocate.

Who that is: an RL trainer whose caching allocator was given the whole of HBM before this library was loaded (a
cudaMalloc inside sb_copy_obs_to_host_packed then fails), a real-time loop that cannot take the device-wide
synchronisation every cudaMalloc / cudaFree implies, and any thread that runs while another thread of the process
holds a stream capture open in the global mode (cudaMalloc is an error there).  For them the library keeps a 16 MiB
`static __device__` array as the bit image's staging area and never allocates in this call.

The file sorts last on purpose: the array is per-GPU state that lives as long as the process, so the test has to hold
both in a fresh process (first call on the device resolves the array's address) and after every other GPU test has
used the device (the array then holds whatever earlier calls left there).  A failure here must not hide the others
under `pytest -x`.
"""
import ctypes as C

import numpy as np
import pytest

from conftest import load_map, make_cfg
from smarts_b200 import _abi, lidar_params, scenario_gen

MISSIONS_4LANE = [("edge-south-SN", 1, 10.0, "edge-west-EW"), ("edge-west-WE", 0, 10.0, "edge-east-WE"),
                  ("edge-north-NS", 0, 10.0, "edge-south-NS"), ("edge-east-EW", 1, 10.0, "edge-west-EW")]


def _engine(rm, cfg):
    from smarts_b200.engine import Engine
    return Engine(cfg, rm, lidar_params.base_rays(lidar_params.BasicLidar))


@pytest.mark.gpu
def test_noalloc_packed_grids_equal_the_allocating_call_twice_in_a_row():
    """Two ticks, two calls in a row through the C-ABI: each time the bit images must be identical, byte for byte
    (integer work -- the tolerance of the other grid tests is zero), to what sb_copy_obs_to_host_packed delivers
    for the same device images, and must unpack to the uint8 grids of the device views.  The second call runs on
    an array that still holds the first call's bits, and on a second, smaller engine that shares the array."""
    rm = load_map("4lane")
    E, A, M = 19, 4, 9
    cfg = make_cfg(E=E, A=A, M=M, nbr=True, lidar=True, wall=1.5, ogm=(96, 40, 0.3), dagm=(160, 72, 0.2), route_cap=16)
    x, y, h, v, route, rlen = scenario_gen.agent_starts(rm, E, MISSIONS_4LANE, cfg.route_cap, 71, jitter=30.0, speed=6.0)
    traffic = scenario_gen.SocialTraffic(rm, E, M, 72)
    eng = _engine(rm, cfg)
    eng.reset(x, y, h, v, route, rlen)
    want = eng.host_obs_buffers(pinned=True, packed_grids=True)
    want = {k: want[k] for k in ("ogm", "dagm")}
    got = {k: t.clone().pin_memory() for k, t in want.items()}
    act = np.random.default_rng(73).integers(0, 4, (E, A)).astype(np.int32)
    seen = []
    for t in range(2):
        eng.step_host(act, traffic.states(t))
        for b in got.values():
            b.fill_(0xA5)  # a word the call failed to write would show
        eng.copy_grids_to_host_packed_noalloc(got["ogm"], got["dagm"])
        eng.copy_obs_to_host(want, packed=True)
        dev = eng.obs_numpy()
        for k in ("ogm", "dagm"):
            assert np.array_equal(got[k].numpy(), want[k].numpy()), (t, k)
            assert np.array_equal(eng.unpack_grid(got[k].numpy()), dev[k]), (t, k)
        seen.append(got["ogm"].numpy().copy())
    assert seen[0].any() and not np.array_equal(seen[0], seen[1])  # the two calls really delivered different images
    # one grid only, into pageable memory, from another engine on the same GPU (same array, other sizes)
    cfg2 = make_cfg(E=3, A=2, M=0, dagm=(64, 64, 0.25), route_cap=16)
    x2, y2, h2, v2, r2, l2 = scenario_gen.agent_starts(rm, 3, MISSIONS_4LANE[:2], cfg2.route_cap, 74, jitter=10.0, speed=5.0)
    eng2 = _engine(rm, cfg2)
    eng2.reset(x2, y2, h2, v2, r2, l2)
    eng2.step_host(np.zeros((3, 2), np.int32), None)
    bits = eng2.host_obs_buffers(pinned=False, packed_grids=True)["dagm"]
    eng2.copy_grids_to_host_packed_noalloc(None, bits)
    assert np.array_equal(eng2.unpack_grid(bits.numpy()), eng2.obs_numpy()["dagm"])
    eng2.close()
    eng.close()


@pytest.mark.gpu
def test_noalloc_rejects_what_does_not_fit_the_static_array():
    """Beyond SB_NOALLOC_BITS_MAX_BYTES the call answers SB_EINVAL with both numbers in the message and launches
    nothing; so does a destination for a grid that is off, and a call without any destination."""
    rm = load_map("4lane")
    E, A = 600, 4  # 2400 agents x 2 x 256 x 256 / 8 = 39.3 MB of bits > 16 MiB
    cfg = make_cfg(E=E, A=A, ogm=(256, 256, 0.1), dagm=(256, 256, 0.1), route_cap=16)
    x, y, h, v, route, rlen = scenario_gen.agent_starts(rm, E, MISSIONS_4LANE, cfg.route_cap, 81, jitter=30.0, speed=6.0)
    eng = _engine(rm, cfg)
    eng.reset(x, y, h, v, route, rlen)
    eng.step_host(np.zeros((E, A), np.int32), None)
    lib, hdl = eng._lib, eng._h
    before = lib.sb_launch_count(hdl)
    dummy = np.zeros(64, np.uint8)
    p = dummy.ctypes.data_as(C.c_void_p)  # never written: every call below is refused before any GPU work
    assert lib.sb_copy_grids_to_host_packed_noalloc(hdl, p, p) == _abi.SB_EINVAL
    msg = lib.sb_last_error(hdl).decode()
    assert str(E * A * 2 * 256 * 256 // 8) in msg and str(_abi.SB_NOALLOC_BITS_MAX_BYTES) in msg, msg
    assert lib.sb_copy_grids_to_host_packed_noalloc(hdl, None, None) == _abi.SB_EINVAL
    assert lib.sb_launch_count(hdl) == before and not dummy.any()
    eng.close()
    cfg = make_cfg(E=2, A=1, ogm=(64, 64, 0.5), route_cap=16)
    eng = _engine(rm, cfg)
    assert eng._lib.sb_copy_grids_to_host_packed_noalloc(eng._h, None, p) == _abi.SB_EINVAL  # dagm is not enabled
    eng.close()
