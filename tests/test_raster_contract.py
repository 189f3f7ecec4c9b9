"""The grid-map contract of DESIGN.md section 5 checked from the other side: the oracle rasterises in snapped pixel
space with exact integer edge functions; here every pixel centre is taken BACK to the world by the camera of
/root/reference/smarts/core/renderer.py:325-336 + sensors.py:706, :748 (orthographic film W*res x H*res centred on the ego,
+x of the image to the ego's right, row 0 ahead of it) and tested against the world-space shapes in floating point --
vehicle boxes for the OGM (ego included, test_observations.py:179-187), the map's drivable-area polygons for the
DrivableAreaGridMap.  Pixels whose centre lies within 2 mm of a shape's edge are excused (snapping to 1/256 px moves an
edge by at most res/512).  CPU only; the CUDA path equals the oracle bit for bit (tests/test_gpu_parity.py)."""
import numpy as np
import pytest

from conftest import load_map, make_cfg, oracle_world
from smarts_b200 import scenario_gen


def inside_convex(px, py, vx, vy, margin):
    """(inside, near_edge) of points (px, py) for the convex polygon (vx, vy), either orientation."""
    n = len(vx)
    area2 = sum(vx[i] * vy[(i + 1) % n] - vx[(i + 1) % n] * vy[i] for i in range(n))
    sg = 1.0 if area2 > 0 else -1.0
    inside = np.ones(px.shape, bool)
    near = np.zeros(px.shape, bool)
    for i in range(n):
        j = (i + 1) % n
        ex, ey = vx[j] - vx[i], vy[j] - vy[i]
        ln = np.hypot(ex, ey)
        if ln == 0:
            continue
        d = sg * (ex * (py - vy[i]) - ey * (px - vx[i])) / ln     # signed distance to the edge's line, + = inside
        inside &= d >= 0
        near |= np.abs(d) < margin
    return inside, near & (np.abs(px - np.mean(vx)) < 1e3)          # (near a line of the polygon)


@pytest.mark.parametrize("W,H,res", [(64, 64, 0.5), (96, 48, 0.25)])
def test_grid_maps_agree_with_world_space_geometry(W, H, res):
    rm = load_map("4lane")
    E, A, M = 3, 3, 8
    cfg = make_cfg(E=E, A=A, M=M, ogm=(W, H, res), dagm=(W, H, res), route_cap=16)
    missions = [("edge-south-SN", 1, 60.0, "edge-west-EW"), ("edge-west-WE", 0, 70.0, "edge-east-WE"),
                ("edge-north-NS", 0, 65.0, "edge-south-NS")]
    x, y, h, v, route, rlen = scenario_gen.agent_starts(rm, E, missions, 16, 4, jitter=15.0, speed=5.0)
    h = h + np.random.default_rng(3).normal(0, 0.4, h.shape)       # images rotated against the lanes
    ora = oracle_world(rm, cfg)
    ora.reset(x, y, h, v, route, rlen)
    ora.set_social(scenario_gen.SocialTraffic(rm, E, M, 9).states(0))
    ora.observe()
    o, s = ora.obs(), ora.state()
    cols, rows = np.meshgrid(np.arange(W), np.arange(H))
    fx = (cols + 0.5 - 0.5 * W) * res          # film coordinates of the pixel centres: right, ahead
    fy = (0.5 * H - (rows + 0.5)) * res
    polys = [(rm.poly_x[a:b], rm.poly_y[a:b]) for a, b in zip(rm.poly_start[:-1], rm.poly_start[1:])]
    checked = covered = 0
    for e in range(E):
        for a in range(A):
            ch, sh = s["cos_h"][e, a], s["sin_h"][e, a]     # heading 0 looks along +y; the image's right is (cos h, sin h)
            wx = s["pos_x"][e, a] + fx * ch - fy * sh
            wy = s["pos_y"][e, a] + fx * sh + fy * ch
            # OGM: every alive vehicle's box
            want = np.zeros((H, W), bool)
            excuse = np.zeros((H, W), bool)
            for j in range(A + M):
                if not s["alive"][e, j]:
                    continue
                c, sn = s["cos_h"][e, j], s["sin_h"][e, j]
                hw, hl = 0.5 * s["width"][e, j], 0.5 * s["length"][e, j]
                bx = [s["pos_x"][e, j] + (p * c - q * sn) for p, q in ((-hw, -hl), (hw, -hl), (hw, hl), (-hw, hl))]
                by = [s["pos_y"][e, j] + (p * sn + q * c) for p, q in ((-hw, -hl), (hw, -hl), (hw, hl), (-hw, hl))]
                ins, near = inside_convex(wx, wy, bx, by, 0.002)
                want |= ins
                excuse |= near
            got = o["ogm"][e, a] == 255
            assert np.array_equal(got[~excuse], want[~excuse]), ("ogm", e, a)
            assert got[H // 2 - 1:H // 2 + 1, W // 2 - 1:W // 2 + 1].all()          # the ego's own box at the centre
            checked += (~excuse).sum(); covered += want.sum()
            # drivable area: the map's polygons
            want = np.zeros((H, W), bool)
            excuse = np.zeros((H, W), bool)
            for vx, vy in polys:
                if not (vx.min() < wx.max() and vx.max() > wx.min() and vy.min() < wy.max() and vy.max() > wy.min()):
                    continue
                ins, near = inside_convex(wx, wy, vx, vy, 0.002)
                want |= ins
                excuse |= near
            got = o["dagm"][e, a] == 255
            assert np.array_equal(got[~excuse], want[~excuse]), ("dagm", e, a)
            checked += (~excuse).sum(); covered += want.sum()
    assert checked > 0.9 * 2 * E * A * W * H and covered > 0.05 * E * A * W * H
