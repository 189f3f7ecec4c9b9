"""Row a12 (DrivableAreaGridMap) pinned by a reference-held artefact: tests/golden/ultra_2lane_t/map.glb is the
reference's OWN road mesh (`SumoRoadNetwork.to_glb`, smarts/core/sumo_road_network.py:760-800, shipped next to the
map.net.xml it was built from under ultra/tests/task/train_mock_2lane_t_70kmh_p-test-flow-0/; both files are copied
verbatim).  That mesh is what the reference's renderer draws into the drivable-area camera (renderer.py:228-249,
sensors.py:728-758).  Here the same net.xml goes through THIS repo's polygon builder (smarts_b200/road_map.py), the
oracle / the CUDA kernel rasterise those polygons into ego-centred 1024 x 1024 grids at 0.1 m/px that tile the whole
map, and every pixel is compared with the mesh's triangles rasterised at the same pixel centres.  Pixels within one
pixel of the mesh outline are excused (the GL fill rule of the reference is not reproducible here); everywhere else
the two must agree, and the number of differing outline pixels is bounded."""
import json
import os
import struct

import numpy as np
import pytest

from conftest import GOLDEN, make_cfg, oracle_world
from smarts_b200 import lidar_params
from smarts_b200.road_map import RoadMapData
from test_raster_contract import inside_convex

DIR = os.path.join(GOLDEN, "ultra_2lane_t")
W = H = 1024
RES = 0.1


def glb_triangles(path):
    """Binary glTF 2.0 (one mesh, uint32 indices, float32 positions, no node transforms -- trimesh's export):
    triangles as world (x, y).  The mesh lies in the glTF x-z plane with z = -y (sumo_road_network.py:775-783)."""
    b = open(path, "rb").read()
    magic, ver, _ = struct.unpack("<III", b[:12])
    assert magic == 0x46546C67 and ver == 2
    n = struct.unpack("<I", b[12:16])[0]
    js = json.loads(b[20:20 + n])
    binc = b[20 + n + 8:]
    assert all("matrix" not in nd and "rotation" not in nd and "translation" not in nd for nd in js["nodes"])
    prim = js["meshes"][0]["primitives"][0]

    def view(acc, dtype, width):
        a = js["accessors"][acc]
        v = js["bufferViews"][a["bufferView"]]
        off = v.get("byteOffset", 0) + a.get("byteOffset", 0)
        return np.frombuffer(binc, dtype, a["count"] * width, off).reshape(a["count"], width)

    assert js["accessors"][prim["indices"]]["componentType"] == 5125 and prim.get("mode", 4) == 4
    idx = view(prim["indices"], np.uint32, 1).reshape(-1, 3)
    pos = view(prim["attributes"]["POSITION"], np.float32, 3).astype(np.float64)
    assert np.abs(pos[:, 1]).max() < 1e-6  # flat
    return np.stack([pos[idx, 0], -pos[idx, 2]], axis=-1)  # [T, 3, 2]


def tile_centres(tris):
    lo, hi = tris.reshape(-1, 2).min(0) - 2.0, tris.reshape(-1, 2).max(0) + 2.0
    span = W * RES
    nx, ny = int(np.ceil((hi[0] - lo[0]) / span)), int(np.ceil((hi[1] - lo[1]) / span))
    return [(lo[0] + (i + 0.5) * span, lo[1] + (j + 0.5) * span) for j in range(ny) for i in range(nx)]


def mesh_raster(tris, cx, cy):
    """The mesh at the pixel centres of an ego-centred grid with heading 0 (image right = +x, row 0 = +y side)."""
    cols, rows = np.meshgrid(np.arange(W), np.arange(H))
    wx = cx + (cols + 0.5 - 0.5 * W) * RES
    wy = cy + (0.5 * H - (rows + 0.5)) * RES
    out = np.zeros((H, W), bool)
    for t in tris:
        c0 = int(np.floor((t[:, 0].min() - wx[0, 0]) / RES)) - 1
        c1 = int(np.ceil((t[:, 0].max() - wx[0, 0]) / RES)) + 1
        r0 = int(np.floor((wy[0, 0] - t[:, 1].max()) / RES)) - 1
        r1 = int(np.ceil((wy[0, 0] - t[:, 1].min()) / RES)) + 1
        c0, c1, r0, r1 = max(c0, 0), min(c1, W), max(r0, 0), min(r1, H)
        if c0 >= c1 or r0 >= r1:
            continue
        ins, _ = inside_convex(wx[r0:r1, c0:c1], wy[r0:r1, c0:c1], list(t[:, 0]), list(t[:, 1]), 0.0)
        out[r0:r1, c0:c1] |= ins
    return out


def outline_band(img):
    """Pixels within one pixel (8-neighbourhood) of a 0 / 1 transition of img."""
    p = np.pad(img, 1, mode="edge")
    nb = [p[1 + dr:1 + dr + H, 1 + dc:1 + dc + W] for dr in (-1, 0, 1) for dc in (-1, 0, 1)]
    return np.logical_or.reduce(nb) != np.logical_and.reduce(nb)


def check_against_mesh(dagm, tris, centres):
    covered = differ = 0
    for k, (cx, cy) in enumerate(centres):
        want = mesh_raster(tris, cx, cy)
        got = dagm[k] == 255
        band = outline_band(want)
        bad = (got != want) & ~band
        assert not bad.any(), f"tile {k}: {int(bad.sum())} pixels away from the mesh outline differ, first at {np.argwhere(bad)[0]}"
        covered += int(want.sum())
        differ += int((got != want).sum())
    return covered, differ


def world_inputs(rm, centres):
    E = len(centres)
    x = np.array([[c[0]] for c in centres]); y = np.array([[c[1]] for c in centres])
    z = np.zeros((E, 1))
    return x, y, z, z.copy(), np.full((E, 1, 4), -1, np.int32), np.zeros((E, 1), np.int32)


def test_repo_polygons_cover_what_the_reference_mesh_covers():
    tris = glb_triangles(os.path.join(DIR, "map.glb"))
    assert len(tris) == 232
    rm = RoadMapData.from_file(os.path.join(DIR, "map.net.xml"))
    centres = tile_centres(tris)
    cfg = make_cfg(E=len(centres), A=1, dagm=(W, H, RES), route_cap=4)
    ora = oracle_world(rm, cfg, threads=8)
    ora.reset(*world_inputs(rm, centres))
    ora.observe()
    covered, differ = check_against_mesh(ora.obs()["dagm"][:, 0], tris, centres)
    # 0.1 m pixels over the whole T-junction: ~4e5 road pixels (4 177 m^2), a handful of outline pixels differ
    assert 380_000 < covered < 460_000, covered
    assert differ <= 64, (differ, covered)


@pytest.mark.gpu
def test_gpu_dagm_against_the_reference_mesh():
    from smarts_b200.engine import Engine
    tris = glb_triangles(os.path.join(DIR, "map.glb"))
    rm = RoadMapData.from_file(os.path.join(DIR, "map.net.xml"))
    centres = tile_centres(tris)
    cfg = make_cfg(E=len(centres), A=1, dagm=(W, H, RES), route_cap=4)
    eng, ora = Engine(cfg, rm, lidar_params.base_rays(lidar_params.BasicLidar)), oracle_world(rm, cfg, threads=8)
    for w in (eng, ora):
        w.reset(*world_inputs(rm, centres))
        w.observe()
    got = eng.obs_numpy()["dagm"]
    assert np.array_equal(got, ora.obs()["dagm"])
    covered, differ = check_against_mesh(got[:, 0], tris, centres)
    assert covered > 380_000 and differ <= 64, (differ, covered)
    eng.close()
