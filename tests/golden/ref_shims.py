"""Import the reference's own Python for the hot path behind import shims.

TEST INFRASTRUCTURE, container-only: this file is used by ``gen_golden.py`` to *generate* the
fixtures committed next to it.  Nothing under ``tests/`` reads /root/reference at run time on
the GPU box; the product never imports this.

What is shimmed (absent wheels):
  cached_property -> functools.cached_property;  shapely / trimesh / pybullet / sh / gym ->
  empty stand-ins (never called on the paths exercised);  sumolib -> the accessor surface
  restated in ``smarts_b200/sumo_net.py`` with the geometry helpers mapped onto the reference's
  own in-repo equivalents (smarts/core/utils/math.py:300-433).
What is NOT shimmed: smarts.core.{lanepoints, sumo_road_network, road_map, coordinates,
controllers.*, lidar, utils.math} run as the reference wrote them (scipy KD-trees are real).
"""
from __future__ import annotations

import functools
import importlib
import sys
import types
from dataclasses import dataclass
from typing import Optional

REF = "/root/reference"


class _Anything:
    def __init__(self, *a, **k):
        pass

    def __call__(self, *a, **k):
        return _Anything()

    def __getattr__(self, name):
        if name.startswith("__"):
            raise AttributeError(name)
        return _Anything()


class _StubModule(types.ModuleType):
    def __getattr__(self, name):
        if name.startswith("__"):
            raise AttributeError(name)
        cls = type(name, (_Anything,), {})
        setattr(self, name, cls)
        return cls


def _stub(name: str) -> types.ModuleType:
    m = _StubModule(name)
    m.__path__ = []  # behave as a package
    sys.modules[name] = m
    return m


def install():
    if "smarts" in sys.modules and getattr(sys.modules["smarts"], "_sb_shim", False):
        return
    smarts = types.ModuleType("smarts")
    smarts.__path__ = [REF + "/smarts"]
    smarts.VERSION = "0.6.1rc1"
    smarts._sb_shim = True
    sys.modules["smarts"] = smarts

    cp = types.ModuleType("cached_property")
    cp.cached_property = functools.cached_property
    sys.modules["cached_property"] = cp

    for name in [
        "shapely", "shapely.geometry", "shapely.geometry.base", "shapely.ops", "shapely.affinity",
        "trimesh", "trimesh.scene", "trimesh.exchange", "trimesh.exchange.gltf", "trimesh.visual",
        "pybullet", "pybullet_utils", "pybullet_utils.bullet_client", "sh", "gym", "gym.spaces",
        "sumolib", "sumolib.net", "sumolib.net.edge", "sumolib.net.lane", "sumolib.geomhelper",
        "yaml",
    ]:
        if name == "yaml":
            continue
        _stub(name)
    sys.modules["trimesh.exchange"].gltf = sys.modules["trimesh.exchange.gltf"]

    # pybullet: the only function the LiDAR ray table needs (lidar.py:101).  Bullet's
    # b3Quaternion::setEulerZYX with (yaw=z, pitch=y, roll=x); formula recalled, not executed
    # from pybullet (module absent).
    import math

    def getQuaternionFromEuler(e):
        roll, pitch, yaw = e
        hy, hp, hr = yaw * 0.5, pitch * 0.5, roll * 0.5
        cy, sy, cp_, sp, cr, sr = math.cos(hy), math.sin(hy), math.cos(hp), math.sin(hp), math.cos(hr), math.sin(hr)
        return (
            sr * cp_ * cy - cr * sp * sy,
            cr * sp * cy + sr * cp_ * sy,
            cr * cp_ * sy - sr * sp * cy,
            cr * cp_ * cy + sr * sp * sy,
        )

    sys.modules["pybullet"].getQuaternionFromEuler = getQuaternionFromEuler
    sys.modules["pybullet"].MAX_RAY_INTERSECTION_BATCH_SIZE = 16384

    # smarts.core.utils.sumo normally locates SUMO_HOME; give it the stub directly.
    importlib.import_module("smarts.core")
    importlib.import_module("smarts.core.utils")
    # sumolib.geomhelper -> the reference's in-repo equivalents (utils/math.py:300-390)
    rmath = importlib.import_module("smarts.core.utils.math")
    gh = sys.modules["sumolib.geomhelper"]
    gh.polygonOffsetWithMinimumDistanceToPoint = (
        lambda point, shape, perpendicular=False: rmath.polygon_offset_with_minimum_distance_to_point(point, shape)
    )
    gh.positionAtShapeOffset = rmath.position_at_shape_offset
    gh.distance = rmath.euclidean_distance
    gh.distancePointToPolygon = rmath.distance_point_to_polygon
    sys.modules["sumolib"].geomhelper = gh
    su = types.ModuleType("smarts.core.utils.sumo")
    su.sumolib = sys.modules["sumolib"]
    sys.modules["smarts.core.utils.sumo"] = su
    # same treatment for the pybullet redirect module (smarts/core/utils/pybullet.py)
    pb = sys.modules["pybullet"]
    pb.pybullet = pb
    pb.bullet_client = sys.modules["pybullet_utils.bullet_client"]
    sys.modules["smarts.core.utils.pybullet"] = pb
    sys.modules["smarts.core.utils"].pybullet = pb

    # sstudio.types drags in the whole scenario DSL; only MapSpec is needed.
    sst = types.ModuleType("smarts.sstudio")
    sst.__path__ = []
    sys.modules["smarts.sstudio"] = sst
    sstt = types.ModuleType("smarts.sstudio.types")

    @dataclass(frozen=True)
    class MapSpec:
        source: str
        lanepoint_spacing: Optional[float] = None
        default_lane_width: Optional[float] = None
        shift_to_origin: bool = False

    sstt.MapSpec = MapSpec
    sys.modules["smarts.sstudio.types"] = sstt


def canonical_kd_ties():
    """scipy's cKDTree returns equidistant neighbours in an order that depends on the tree's internal partition
    (not on anything the reference specifies); lane maps are full of exactly coincident lane points -- the last
    point of a lane and the first of its successor -- so which of the two comes first decides which lane a vehicle
    standing there is "on".  The fixtures canonicalise that one third-party degree of freedom: among neighbours at
    exactly the same distance the lower index comes first (the rule of the oracle and the kernels).
    Everything else about the query, and all of the reference's own code around it, runs unchanged."""
    import numpy as np
    lp_mod = importlib.import_module("smarts.core.lanepoints")
    base = lp_mod.KDTree
    if getattr(base, "_sb_canonical", False):
        return

    class CanonicalKDTree(base):
        _sb_canonical = True

        def query(self, x, k=1, *a, **kw):
            x = np.asarray(x, float)
            kk = int(min(self.n, k + 8))
            d, i = super().query(x.reshape(-1, self.m), kk, *a, **kw)
            d, i = np.asarray(d).reshape(len(d), -1), np.asarray(i).reshape(len(i), -1)
            rank = getattr(self, "_sb_rank", None)        # set_tie_ranks: index of the point in the flattened map
            order = np.lexsort((i if rank is None else rank[np.minimum(i, self.n - 1)], d), axis=1)  # distance, then rank
            d, i = np.take_along_axis(d, order, 1)[:, :k], np.take_along_axis(i, order, 1)[:, :k]
            if k == 1:
                d, i = d[:, 0], i[:, 0]
            if x.ndim == 1:
                d, i = d[0], i[0]
            return d, i

    lp_mod.KDTree = CanonicalKDTree


def set_tie_ranks(ref_rm, mine):
    """Give every KD-tree of the reference map the rank that breaks distance ties: the index of the lane point in
    the flattened map `mine` (lanes in net.xml order, points along the lane), i.e. "lowest index wins" in the
    numbering the oracle and the kernels use.  The reference's own global order is the traversal order of its
    lane-point builder and plays no role in its results except through scipy's arbitrary tie order."""
    import numpy as np
    lps = ref_rm._lanepoints
    lane_index = {lid: i for i, lid in enumerate(mine.lane_ids)}
    rank_of = {}
    for lane_id, llps in lps._lanepoints_by_lane_id.items():
        start = int(mine.lane_lp_start[lane_index[lane_id]])
        assert len(llps) == int(mine.lane_lp_count[lane_index[lane_id]]), lane_id
        for k, llp in enumerate(llps):
            p = llp.lp.pose.as_position2d()
            assert abs(p[0] - mine.lp_x[start + k]) < 1e-9 and abs(p[1] - mine.lp_y[start + k]) < 1e-9, (lane_id, k)
            rank_of[id(llp)] = start + k
    def attach(tree, llps):
        tree._sb_rank = np.array([rank_of[id(l)] for l in llps], np.int64)
    attach(lps._lanepoints_kd_tree, lps._linked_lanepoints)
    for lane_id, tree in lps._lanepoints_kd_tree_by_lane_id.items():
        attach(tree, lps._lanepoints_by_lane_id[lane_id])
    for edge_id, tree in lps._lanepoints_kd_tree_by_edge_id.items():
        attach(tree, lps._lanepoints_by_edge_id[edge_id])


class RefGraph:
    """``sumolib.net.Net`` stand-in over ``smarts_b200.sumo_net.SumoNet`` for the reference's
    ``SumoRoadNetwork``; ``getNeighboringLanes`` restates sumolib's rtree + distance filter with
    the reference's own ``distance_point_to_polygon`` (utils/math.py:414-433)."""

    def __init__(self, net):
        self._net = net
        self._shifted_by_smarts = False

    def __getattr__(self, name):
        return getattr(self._net, name)

    def getNeighboringLanes(self, x, y, r=0.1, includeJunctions=True, allowFallback=True):
        from smarts.core.utils.math import distance_point_to_polygon

        lanes = []
        for e in self._net.getEdges(True):
            for l in e.getLanes():
                d = distance_point_to_polygon((x, y), l.getShape(includeJunctions))
                if d < r:
                    lanes.append((l, d))
        return lanes


def reference_road_map(net_file: str, spacing: float = 1.0):
    """The reference ``SumoRoadNetwork`` (its own lane points, KD-trees, waypoint code) over a
    map parsed by ``smarts_b200.sumo_net``."""
    install()
    canonical_kd_ties()
    sys.path.insert(0, "/root/repo")
    from smarts_b200.sumo_net import SumoNet
    from smarts.core.sumo_road_network import SumoRoadNetwork
    from smarts.sstudio.types import MapSpec

    net = SumoNet(net_file)
    spec = MapSpec(source=net_file, lanepoint_spacing=spacing)
    return SumoRoadNetwork(RefGraph(net), net_file, spec), net
