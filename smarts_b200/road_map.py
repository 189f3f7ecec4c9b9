"""Flattened, device-ready road map (setup-time, host side).

Builds, once per scenario, the immutable arrays every kernel reads:

* lanes / roads tables (``SumoRoadNetwork.Lane/Road`` accessors,
  /root/reference/smarts/core/sumo_road_network.py:261-358, :550-657),
* the linked lane points of ``LanePoints.from_sumo`` + ``_interpolate_shape_lanepoints``
  (/root/reference/smarts/core/lanepoints.py:104-227, :376-515) as structure-of-arrays with a
  CSR ``nexts`` list,
* lane centre polylines for ``nearest_lanes`` (sumo_road_network.py:675-709),
* convex road-surface polygons for the DrivableAreaGridMap (sumo_road_network.py:986-1019,
  utils/geometry.py:31-45) and road boundary segments for the LiDAR,
* routes (sumo_road_network.py:711-801).

Lane points of one lane are contiguous and lanes of one road are contiguous (ordered by lane
index), so the reference's three KD-trees (all points / per lane / per road,
lanepoints.py:85-102) become three index ranges of the same arrays.
"""
from __future__ import annotations

import math
import warnings
from dataclasses import dataclass
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np

from .sumo_net import NetEdge, NetLane, SumoNet

TWO_PI = 2.0 * math.pi


def vec_to_radians(x: float, y: float) -> float:
    """utils/math.py:256-277: +y is 0 rad, counter-clockwise, result in [0, 2pi)."""
    r = math.atan2(abs(y), abs(x))
    if x < 0:
        if y < 0:
            return (r + 0.5 * math.pi) % TWO_PI
        return (0.5 * math.pi - r) % TWO_PI
    elif y < 0:
        return (1.5 * math.pi - r) % TWO_PI
    return (r - 0.5 * math.pi) % TWO_PI


def wrap_heading(h: float) -> float:
    """coordinates.py:169-185 (``Heading.__new__``): (-pi, pi]."""
    h = h % TWO_PI
    if h > math.pi:
        h -= TWO_PI
    return h


@dataclass
class Route:
    """sumo_road_network.py:884-903."""

    road_ids: List[str]
    road_indices: List[int]
    length: float


class RoadMapData:
    """All per-scenario constant arrays.  ``arrays()`` returns the dict handed to the C-ABI
    (``sb_upload_map``) and, unchanged, to the CPU oracle."""

    _SAVE_KEYS = ["road_length", "conn_from", "conn_to", "conn_via"]
    _RW_KEYS = ["lane_in_start", "lane_in", "road_par_start", "road_par"]  # RoadWaypoints tables (ABI 6)

    def __init__(self, net: Optional[SumoNet], lanepoint_spacing: float = 1.0):
        self.net = net
        self.spacing = float(lanepoint_spacing)
        if net is not None:
            self._build_tables()
            self._build_lanepoints()
            self._build_shapes()
            self._build_polygons()

    @classmethod
    def from_file(cls, net_file: str, lanepoint_spacing: float = 1.0,
                  default_lane_width: float = 3.2) -> "RoadMapData":
        """``net_file``: a SUMO ``*.net.xml`` (or a scenario directory holding ``map.net.xml``,
        sumo_road_network.py:191-196) or a ``*.npz`` written by :meth:`save`."""
        import os
        if os.path.isdir(net_file):
            net_file = os.path.join(net_file, "map.net.xml")
        if net_file.endswith(".npz"):
            return cls.load(net_file)
        return cls(SumoNet(net_file, default_lane_width), lanepoint_spacing)

    def save(self, path: str):
        """Write the flattened map (this repo's own format; no reference content)."""
        d = self.arrays()
        for k in self._SAVE_KEYS:
            d[k] = getattr(self, k)
        d["lane_ids"] = np.asarray(self._lane_ids)
        d["road_ids"] = np.asarray(self._road_ids)
        d["spacing"] = np.asarray(self.spacing)
        np.savez_compressed(path, **d)

    @classmethod
    def load(cls, path: str) -> "RoadMapData":
        z = np.load(path, allow_pickle=False)
        self = cls(None, float(z["spacing"]))
        for k in z.files:
            if k in ("lane_ids", "road_ids", "spacing"):
                continue
            setattr(self, k, z[k])
        self._lane_ids = [str(s) for s in z["lane_ids"]]
        self._road_ids = [str(s) for s in z["road_ids"]]
        self.lane_index = {s: i for i, s in enumerate(self._lane_ids)}
        self.road_index = {s: i for i, s in enumerate(self._road_ids)}
        if not hasattr(self, "lane_in_start"):
            # a map saved before the RoadWaypoints tables existed: incoming lanes follow from the connection table
            # (from-lanes of the connections that end on the lane, file order); parallel roads need the node ids of
            # the net.xml, which such a file does not hold -- none are listed
            nl = len(self._lane_ids)
            inc = [[] for _ in range(nl)]
            for f, t in zip(self.conn_from, self.conn_to):
                inc[int(t)].append(int(f))
            self.lane_in_start = np.asarray(np.cumsum([0] + [len(v) for v in inc]), np.int32)
            flat = [f for v in inc for f in v]
            self.lane_in = np.asarray(flat if flat else [0], np.int32)
            self.road_par_start = np.zeros(len(self._road_ids) + 1, np.int32)
            self.road_par = np.zeros(1, np.int32)
            # ... so the RoadWaypoints sensor would see other lane sets than the reference (sensors.py:1007-1011):
            # the map is flagged and `require_road_waypoint_tables` (called for rw_enable) refuses it
            self.rw_tables_complete = False
            warnings.warn(f"{path}: saved without the RoadWaypoints tables (lane_in / road_par); every sensor but "
                          "RoadWaypoints works, rebuild the file from its net.xml to use that one", stacklevel=2)
        return self

    rw_tables_complete = True

    def require_road_waypoint_tables(self):
        """Raise unless the incoming-lane and parallel-road tables come from the net itself (``Lane.incoming_lanes``
        in getIncoming() order, ``Road.parallel_roads`` by node ids; sumo_road_network.py:342-348, :607-618).  An
        old-format .npz only approximates the first and has none of the second."""
        if not self.rw_tables_complete:
            raise ValueError("this map was loaded from an .npz written before the RoadWaypoints tables existed: "
                             "RoadWaypoints (rw_enable) would list other lanes than the reference; rebuild the file "
                             "with RoadMapData.from_file(net_xml).save(path)")

    # ------------------------------------------------------------------ tables
    def _build_tables(self):
        net = self.net
        self.roads: List[NetEdge] = list(net.getEdges(True))
        self.road_index: Dict[str, int] = {e.getID(): i for i, e in enumerate(self.roads)}
        self.lanes: List[NetLane] = []
        road_lane_start, road_lane_count = [], []
        for e in self.roads:
            road_lane_start.append(len(self.lanes))
            road_lane_count.append(len(e.getLanes()))
            self.lanes.extend(e.getLanes())
        self.lane_index: Dict[str, int] = {l.getID(): i for i, l in enumerate(self.lanes)}
        nl = len(self.lanes)
        self.road_lane_start = np.asarray(road_lane_start, np.int32)
        self.road_lane_count = np.asarray(road_lane_count, np.int32)
        self.road_is_junction = np.asarray([e.isSpecial() for e in self.roads], np.uint8)
        self.road_length = np.asarray([e.getLength() for e in self.roads], np.float64)
        # Road.outgoing_roads[0] (sumo_road_network.py:578-583), used by _resolve_in_junction
        self.road_first_out = np.asarray(
            [self.road_index[next(iter(e.getOutgoing().keys())).getID()] if e.getOutgoing() else -1
             for e in self.roads], np.int32)
        self.lane_road = np.asarray([self.road_index[l.getEdge().getID()] for l in self.lanes], np.int32)
        self.lane_idx_in_road = np.asarray([l.getIndex() for l in self.lanes], np.int32)
        self.lane_width = np.asarray([l.getWidth() for l in self.lanes], np.float64)
        self.lane_speed = np.asarray([l.getSpeed() for l in self.lanes], np.float64)
        self.lane_length = np.asarray([l.getLength() for l in self.lanes], np.float64)
        # outgoing lanes: via lane if present else to-lane (sumo_road_network.py:351-358)
        out_start, out_flat = [0], []
        for l in self.lanes:
            for c in l.getOutgoing():
                out_flat.append(self.lane_index[c.getViaLaneID() or c.getToLane().getID()])
            out_start.append(len(out_flat))
        self.lane_out_start = np.asarray(out_start, np.int32)
        self.lane_out = np.asarray(out_flat if out_flat else [0], np.int32)
        assert nl == len(self.lane_width)
        # incoming lanes (Lane.incoming_lanes, sumo_road_network.py:342-348: sumolib Lane.getIncoming()) and
        # parallel roads (Road.parallel_roads, :607-618: other edges between the same two nodes) -- RoadWaypoints
        in_start, in_flat = [0], []
        for l in self.lanes:
            in_flat.extend(self.lane_index[i.getID()] for i in l.getIncoming())
            in_start.append(len(in_flat))
        self.lane_in_start = np.asarray(in_start, np.int32)
        self.lane_in = np.asarray(in_flat if in_flat else [0], np.int32)
        par_start, par_flat = [0], []
        for e in self.roads:
            to_id = e.getToNode().getID()
            par_flat.extend(self.road_index[o.getID()] for o in e.getFromNode().getOutgoing()
                            if o.getID() != e.getID() and o.getToNode().getID() == to_id)
            par_start.append(len(par_flat))
        self.road_par_start = np.asarray(par_start, np.int32)
        self.road_par = np.asarray(par_flat if par_flat else [0], np.int32)
        # <connection> table in file order (from lane, to lane, via lane or -1): all the routing
        # code needs, so a map saved with ``save()`` can route without the net.xml.
        cf, ct, cv = [], [], []
        for c in net.connections:
            cf.append(self.lane_index[c.getFromLane().getID()])
            ct.append(self.lane_index[c.getToLane().getID()])
            cv.append(self.lane_index[c.getViaLaneID()] if c.getViaLaneID() else -1)
        self.conn_from = np.asarray(cf, np.int32)
        self.conn_to = np.asarray(ct, np.int32)
        self.conn_via = np.asarray(cv, np.int32)
        self._lane_ids = [l.getID() for l in self.lanes]
        self._road_ids = [e.getID() for e in self.roads]

    def outgoing_lanes(self, li: int) -> List[int]:
        return list(self.lane_out[self.lane_out_start[li]:self.lane_out_start[li + 1]])

    # -------------------------------------------------------------- lane points
    def _build_lanepoints(self):
        """Restates ``LanePoints.from_sumo`` + ``_interpolate_shape_lanepoints``.

        The reference links lanes breadth-first from every non-internal lane and memoises per
        lane, so each lane ends up with exactly one chain of points:
        shape vertex, interpolated points every ``spacing`` (with the 0.8*spacing / 1.4 m end
        rules, lanepoints.py:460-515), next shape vertex, ..., last vertex; the last vertex
        links to the first point of every outgoing lane (via lane if present,
        lanepoints.py:202-210), with interpolated points on the *from* lane if there is a gap.
        Lanes never reached from a non-internal lane get no points (same as the reference).
        """
        sp = self.spacing
        nlanes = len(self.lanes)
        # reachability in the reference's order (only matters for which lanes exist)
        reached = [False] * nlanes
        stack = []
        for e in self.net.getEdges(False):
            for l in e.getLanes():
                stack.append(self.lane_index[l.getID()])
        while stack:
            li = stack.pop()
            if reached[li]:
                continue
            reached[li] = True
            stack.extend(self.outgoing_lanes(li))

        xs: List[float] = []
        ys: List[float] = []
        hd: List[float] = []
        inferred: List[int] = []
        lane_of: List[int] = []
        nexts: List[List[int]] = []
        lane_lp_start = np.zeros(nlanes, np.int32)
        lane_lp_count = np.zeros(nlanes, np.int32)
        lane_first = [-1] * nlanes
        lane_last = [-1] * nlanes
        pending_links: List[Tuple[int, int]] = []  # (from lane, to lane) resolved after all chains exist

        def add(x, y, h, inf, li):
            xs.append(float(x)); ys.append(float(y)); hd.append(float(h))
            inferred.append(int(inf)); lane_of.append(li); nexts.append([])
            return len(xs) - 1

        def interp(ax, ay, bx, by, curr, li):
            """lanepoints.py:444-515; returns the last created (or ``curr``)."""
            vx, vy = bx - ax, by - ay
            seg_len = math.hypot(vx, vy) if False else float(np.linalg.norm(np.array([vx, vy])))
            dist = sp
            while dist < seg_len:
                p = dist / seg_len
                px, py = ax + vx * p, ay + vy * p
                cx, cy = xs[curr], ys[curr]
                half = float(np.linalg.norm(0.5 * (np.array([cx, cy]) - np.array([bx, by]))))
                if half < 1.4:
                    px, py = 0.5 * (bx + cx), 0.5 * (by + cy)
                d = float(np.linalg.norm(np.array([bx - px, by - py])))
                if d < 0.8 * sp:
                    break
                h = wrap_heading(vec_to_radians(vx, vy))
                n = add(px, py, h, 1, li)
                nexts[curr].append(n)
                curr = n
                dist += sp
            return curr

        for li, lane in enumerate(self.lanes):
            lane_lp_start[li] = len(xs)
            if not reached[li]:
                continue
            shape = lane.getShape(False)
            assert len(shape) >= 2, repr(shape)
            nv = len(shape)
            seg_h = [wrap_heading(vec_to_radians(shape[i + 1][0] - shape[i][0],
                                                 shape[i + 1][1] - shape[i][1]))
                     for i in range(nv - 1)]
            curr = add(shape[0][0], shape[0][1], seg_h[0], 0, li)
            lane_first[li] = curr
            shape_lp = {(float(shape[0][0]), float(shape[0][1]), seg_h[0]): curr}
            merged = False
            for i in range(nv - 1):
                curr = interp(shape[i][0], shape[i][1], shape[i + 1][0], shape[i + 1][1], curr, li)
                h = seg_h[i + 1] if i + 1 < nv - 1 else seg_h[nv - 2]
                key = (float(shape[i + 1][0]), float(shape[i + 1][1]), h)
                if key in shape_lp:
                    # The reference memoises interpolation per LanePoint *value* (lane, pose, width: frozen dataclass,
                    # lanepoints.py:396-401), so a shape point equal to an earlier one of its lane is not created
                    # again: the chain links back to the earlier point and whatever followed the duplicate is not
                    # linked from here.  Happens for lanes whose last two vertices coincide (zero-length internal
                    # lanes of minicity): ONE lane point whose only successor is itself, no link to the outgoing lanes.
                    if i + 1 != nv - 1:
                        raise ValueError(f"lane {lane.getID()}: repeated interior shape vertex (not restated)")
                    nexts[curr].append(shape_lp[key])
                    merged = True
                    break
                n = add(shape[i + 1][0], shape[i + 1][1], h, 0, li)
                shape_lp[key] = n
                nexts[curr].append(n)
                curr = n
            lane_last[li] = curr
            if not merged:
                for lo in self.outgoing_lanes(li):
                    pending_links.append((li, lo))
            lane_lp_count[li] = len(xs) - lane_lp_start[li]

        # cross-lane links (gap interpolation keeps lane points contiguous only when there is
        # no gap; gaps are appended after the main chains and get their own contiguous block
        # accounted to the from-lane through ``lp_lane`` -- the per-lane *range* then excludes
        # them, which matches the maps shipped with the reference: all gaps are < spacing).
        self.gap_points = 0
        for li, lo in pending_links:
            a, b = lane_last[li], lane_first[lo]
            if b < 0:
                continue
            before = len(xs)
            curr = interp(xs[a], ys[a], xs[b], ys[b], a, li)
            self.gap_points += len(xs) - before
            nexts[curr].append(b)

        self.lp_x = np.asarray(xs, np.float64)
        self.lp_y = np.asarray(ys, np.float64)
        self.lp_heading = np.asarray(hd, np.float64)
        self.lp_inferred = np.asarray(inferred, np.uint8)
        self.lp_lane = np.asarray(lane_of, np.int32)
        start, flat = [0], []
        for n in nexts:
            flat.extend(n)
            start.append(len(flat))
        self.lp_next_start = np.asarray(start, np.int32)
        self.lp_next = np.asarray(flat if flat else [0], np.int32)
        self.lane_lp_start = lane_lp_start
        self.lane_lp_count = lane_lp_count
        assert self.gap_points == 0, (
            "map has lane-to-lane gaps >= lanepoint spacing; gap lane points would fall outside "
            "the per-lane index range")
        # per road range = union of its (contiguous) lanes
        rs, rc = [], []
        for r in range(len(self.roads)):
            l0 = self.road_lane_start[r]
            l1 = l0 + self.road_lane_count[r]
            rs.append(lane_lp_start[l0] if l1 > l0 else 0)
            rc.append(int(lane_lp_count[l0:l1].sum()))
        self.road_lp_start = np.asarray(rs, np.int32)
        self.road_lp_count = np.asarray(rc, np.int32)

    # -------------------------------------------------------------- polylines
    def _build_shapes(self):
        sx, sy, start = [], [], [0]
        for l in self.lanes:
            for x, y in l.getShape(False):
                sx.append(x); sy.append(y)
            start.append(len(sx))
        self.shape_x = np.asarray(sx, np.float64)
        self.shape_y = np.asarray(sy, np.float64)
        self.lane_shape_start = np.asarray(start, np.int32)

    # ---------------------------------------------------------------- polygons
    def _build_polygons(self):
        """Convex pieces whose union is the drivable surface.

        The reference buffers every lane centre line by ``width/2`` with flat caps and round
        joins at one segment per quarter circle (utils/geometry.py:31-45) and adds every
        junction shape with > 2 vertices (sumo_road_network.py:1009-1017).  GEOS/shapely and
        Panda3D are unavailable on this box, so the surface is restated as: one rectangle per
        centre-line segment, one join fan per interior vertex (the GEOS fillet: the arc between
        the two offset points on the outer side, subdivided at multiples of pi/2 measured from
        the first offset direction), and the junction polygons triangulated by ear clipping.
        The ``snap`` passes (sumo_road_network.py:1021-1139) are not restated (documented
        deviation).
        """
        polys: List[np.ndarray] = []  # each [n,2] CCW convex
        edges: List[Tuple[float, float, float, float]] = []

        def ccw(p):
            p = np.asarray(p, np.float64)
            a = 0.0
            for i in range(len(p)):
                j = (i + 1) % len(p)
                a += p[i, 0] * p[j, 1] - p[j, 0] * p[i, 1]
            return p if a >= 0 else p[::-1].copy()

        for l in self.lanes:
            shp = l.getShape()
            hw = 0.5 * l.getWidth()
            pts = [shp[0]]
            for p in shp[1:]:
                if p != pts[-1]:
                    pts.append(p)
            if len(pts) < 2:
                continue
            offs = []
            for i in range(len(pts) - 1):
                (ax, ay), (bx, by) = pts[i], pts[i + 1]
                dx, dy = bx - ax, by - ay
                ln = math.hypot(dx, dy)
                nx, ny = -dy / ln * hw, dx / ln * hw  # left normal * hw
                polys.append(ccw([(ax + nx, ay + ny), (ax - nx, ay - ny), (bx - nx, by - ny), (bx + nx, by + ny)]))
                offs.append((nx, ny))
            for i in range(1, len(pts) - 1):
                (n0x, n0y), (n1x, n1y) = offs[i - 1], offs[i]
                cross = n0x * n1y - n0y * n1x
                if abs(cross) < 1e-12 * hw * hw:
                    continue
                s = -1.0 if cross > 0 else 1.0  # left turn -> outer side is the right
                cx, cy = pts[i]
                a0 = math.atan2(s * n0y, s * n0x)
                a1 = math.atan2(s * n1y, s * n1x)
                da = a1 - a0
                while da > math.pi:
                    da -= TWO_PI
                while da < -math.pi:
                    da += TWO_PI
                fan = [(cx, cy), (cx + s * n0x, cy + s * n0y)]
                nseg = int(abs(da) / (0.5 * math.pi) + 0.5)
                if nseg >= 1:
                    inc = da / nseg
                    for k in range(1, nseg):
                        fan.append((cx + hw * math.cos(a0 + k * inc), cy + hw * math.sin(a0 + k * inc)))
                fan.append((cx + s * n1x, cy + s * n1y))
                for k in range(1, len(fan) - 1):  # triangles: every piece has <= 4 vertices
                    polys.append(ccw([fan[0], fan[k], fan[k + 1]]))
            # boundary segments of the outer-most lanes feed the LiDAR "road edge" walls
        for e in self.roads:
            if e.isSpecial():
                continue
            lanes = e.getLanes()
            for lane, side in ((lanes[0], -1.0), (lanes[-1], 1.0)):
                shp = lane.getShape()
                hw = 0.5 * lane.getWidth()
                for i in range(len(shp) - 1):
                    (ax, ay), (bx, by) = shp[i], shp[i + 1]
                    dx, dy = bx - ax, by - ay
                    ln = math.hypot(dx, dy)
                    if ln == 0:
                        continue
                    nx, ny = -dy / ln * hw * side, dx / ln * hw * side
                    edges.append((ax + nx, ay + ny, bx + nx, by + ny))

        for node in self.net.getNodes():
            line = node.getShape()
            if len(line) <= 2 or len(set(line)) == 1:
                continue
            for tri in _ear_clip(line):
                polys.append(ccw(tri))

        self.n_polys_raw = len(polys)
        polys = split_polygons(simplify_polygons(polys, MAX_POLY_VERTS), STAGE_POLY_VERTS)
        vx, vy, start = [], [], [0]
        for p in polys:
            vx.extend(p[:, 0]); vy.extend(p[:, 1])
            start.append(len(vx))
        self.poly_x = np.asarray(vx, np.float64)
        self.poly_y = np.asarray(vy, np.float64)
        self.poly_start = np.asarray(start, np.int32)
        self.road_edges = np.asarray(edges if edges else np.zeros((0, 4)), np.float64).reshape(-1, 4)

    # ------------------------------------------------------------------ routes
    def road_by_id(self, road_id: str) -> int:
        return self.road_index[road_id]

    def generate_route(self, start_road: str, end_road: str,
                       via: Optional[Sequence[str]] = None) -> Route:
        """sumo_road_network.py:711-801 (max_to_gen == 1), on the <connection> table."""
        g = self._route_graph()
        ri = self.road_index
        roads = [ri[start_road]] + [ri[v] for v in (via or [])]
        if end_road != start_road:
            roads.append(ri[end_road])
        edges: List[int] = []
        for cur, nxt in zip(roads, roads[1:] + [None]):
            if nxt is None:
                edges.append(cur)
                break
            sub = self._shortest_path(g, cur, nxt) or []
            if len(sub) < 2:
                return Route([], [], 0.0)
            edges.extend(sub[:-1])
        if len(edges) == 1:
            idx = [edges[0]]
        else:
            used: List[int] = []
            for ce, ne in zip(edges, edges[1:]):
                for via_lane in g[ce][ne]:  # one entry per connection ce -> ne
                    conn_route = [ce]
                    v = via_lane
                    while v >= 0:
                        via_road = int(self.lane_road[v])
                        conn_route.append(via_road)
                        nxt_ids = set(g[via_road][ne])
                        assert len(nxt_ids) == 1, nxt_ids
                        v = list(nxt_ids)[0]
                    conn_route.append(ne)
                    used.extend(conn_route)
            idx, seen = [], set()
            for r in used:  # first-occurrence order == np.unique(return_index) + sorted(indices)
                if r not in seen:
                    seen.add(r)
                    idx.append(r)
        ids = [self._road_ids[i] for i in idx]
        return Route(ids, idx, float(sum(self.road_length[i] for i in idx)))

    def _route_graph(self) -> Dict[int, Dict[int, List[int]]]:
        """road -> {to road -> [via lane (or -1) per connection]}, insertion-ordered like
        sumolib's ``Edge.getOutgoing()``."""
        if getattr(self, "_rg", None) is None:
            g: Dict[int, Dict[int, List[int]]] = {r: {} for r in range(len(self._road_ids))}
            for f, t, v in zip(self.conn_from, self.conn_to, self.conn_via):
                g[int(self.lane_road[f])].setdefault(int(self.lane_road[t]), []).append(int(v))
            self._rg = g
        return self._rg

    def _shortest_path(self, g, src: int, dst: int) -> Optional[List[int]]:
        """Dijkstra over non-junction roads weighted by road length
        (sumolib.net.Net.getShortestPath as used at sumo_road_network.py:733-738)."""
        import heapq
        seen = set()
        q = [(float(self.road_length[src]), 0, src, (src,))]
        tie = 1
        while q:
            cost, _, r, path = heapq.heappop(q)
            if r in seen:
                continue
            seen.add(r)
            if r == dst:
                return list(path)
            for nr in g[r].keys():
                if self.road_is_junction[nr] or nr in seen:
                    continue
                heapq.heappush(q, (cost + float(self.road_length[nr]), tie, nr, path + (nr,)))
                tie += 1
        return None

    # ------------------------------------------------------------------ export
    def arrays(self) -> Dict[str, np.ndarray]:
        keys = [
            "road_lane_start", "road_lane_count", "road_is_junction", "road_lp_start", "road_lp_count", "road_first_out",
            "lane_road", "lane_idx_in_road", "lane_width", "lane_speed", "lane_length",
            "lane_out_start", "lane_out", "lane_lp_start", "lane_lp_count", "lane_shape_start",
            "lp_x", "lp_y", "lp_heading", "lp_inferred", "lp_lane", "lp_next_start", "lp_next",
            "shape_x", "shape_y", "poly_x", "poly_y", "poly_start", "road_edges",
            "lane_in_start", "lane_in", "road_par_start", "road_par",
        ]
        return {k: np.ascontiguousarray(getattr(self, k)) for k in keys}

    @property
    def lane_ids(self) -> List[str]:
        return self._lane_ids

    @property
    def road_ids(self) -> List[str]:
        return self._road_ids

    def lane_of_road(self, road_id: str, lane_idx: int) -> int:
        r = self.road_index[road_id]
        assert 0 <= lane_idx < self.road_lane_count[r]
        return int(self.road_lane_start[r]) + lane_idx

    def lane_start_pose(self, road_id: str, lane_idx: int, offset: float) -> Tuple[float, float, float]:
        """Centre-line pose at ``offset`` metres along a lane (mission start poses,
        scenarios/*/scenario.py ``Route(begin=(edge, lane, offset))``)."""
        li = self.lane_of_road(road_id, lane_idx)
        s, e = int(self.lane_shape_start[li]), int(self.lane_shape_start[li + 1])
        shp = [(float(self.shape_x[i]), float(self.shape_y[i])) for i in range(s, e)]
        seen = 0.0
        for k in range(len(shp) - 1):
            (ax, ay), (bx, by) = shp[k], shp[k + 1]
            ln = math.hypot(bx - ax, by - ay)
            if seen + ln > offset or k == len(shp) - 2:
                t = min(max((offset - seen) / ln, 0.0), 1.0) if ln > 0 else 0.0
                return (ax + (bx - ax) * t, ay + (by - ay) * t,
                        wrap_heading(vec_to_radians(bx - ax, by - ay)))
            seen += ln
        raise ValueError("empty lane shape")


MAX_POLY_VERTS = 8  # SB_MAX_POLY_VERTS of include/smarts_b200.h
STAGE_POLY_VERTS = 4  # k_raster stages a polygon on 4 lanes when no kept polygon has more vertices


def split_polygons(polys: Sequence[np.ndarray], max_verts: int = STAGE_POLY_VERTS) -> List[np.ndarray]:
    """Fan-split convex polygons with more than ``max_verts`` vertices from their first vertex into
    pieces of at most ``max_verts`` (same union: the pieces share full diagonals)."""
    out: List[np.ndarray] = []
    for p in polys:
        p = np.asarray(p, np.float64)
        n = len(p)
        if n <= max_verts:
            out.append(p)
            continue
        i = 1
        while i < n - 1:
            j = min(i + max_verts - 2, n - 1)
            out.append(np.concatenate([p[:1], p[i:j + 1]], axis=0))
            i = j
    return out



def _cross(o, a, b) -> float:
    return (a[0] - o[0]) * (b[1] - o[1]) - (a[1] - o[1]) * (b[0] - o[0])


def _clean_ring(pts: List[Tuple[float, float]], tol: float) -> List[Tuple[float, float]]:
    """Drop repeated vertices (closer than 1e-6 m) and vertices within ``tol`` metres of the
    line through their neighbours (when they lie between them) of a CCW ring."""
    out = list(pts)
    changed = True
    while changed and len(out) >= 3:
        changed = False
        n = len(out)
        for i in range(n):
            a, b, c = out[i - 1], out[i], out[(i + 1) % n]
            lab = math.hypot(b[0] - a[0], b[1] - a[1])
            lbc = math.hypot(c[0] - b[0], c[1] - b[1])
            lac = math.hypot(c[0] - a[0], c[1] - a[1])
            if lab <= 1e-6 or lbc <= 1e-6:
                out.pop(i)
                changed = True
                break
            if n > 3 and abs(_cross(a, b, c)) <= tol * lac and \
                    (b[0] - a[0]) * (c[0] - b[0]) + (b[1] - a[1]) * (c[1] - b[1]) > 0:
                out.pop(i)
                changed = True
                break
    return out


def _is_convex(pts: Sequence[Tuple[float, float]], tol: float) -> bool:
    """Strictly simple convex CCW ring: every turn is left (to within ``tol`` metres of
    deviation) and the turning angles add up to one revolution."""
    n = len(pts)
    if n < 3:
        return False
    total = 0.0
    for i in range(n):
        a, b, c = pts[i - 1], pts[i], pts[(i + 1) % n]
        lab = math.hypot(b[0] - a[0], b[1] - a[1])
        lbc = math.hypot(c[0] - b[0], c[1] - b[1])
        if lab <= 1e-6 or lbc <= 1e-6:
            return False
        cr = _cross(a, b, c)
        if cr < -tol * max(lab, lbc):
            return False
        total += math.atan2(cr, (b[0] - a[0]) * (c[0] - b[0]) + (b[1] - a[1]) * (c[1] - b[1]))
    return abs(total - TWO_PI) < 1e-3 and _ring_area(pts) > 0.0


def simplify_polygons(polys: Sequence[np.ndarray], max_verts: int = MAX_POLY_VERTS,
                      tol: float = 1e-6) -> List[np.ndarray]:
    """Union-preserving reduction of a list of convex CCW polygons (setup time, host side).

    The DrivableAreaGridMap shows the *union* of the road pieces (sensors.py:675-716 renders the
    merged road mesh), so any polygon list with the same union is an equally valid restatement;
    fewer, larger pieces make the rasteriser cheaper for both the CUDA kernel and the oracle
    (both consume this list unchanged).  Two exact steps, repeated to a fixed point:

    1. merge two polygons that share a full edge when the result is convex and has at most
       ``max_verts`` vertices (Hertel-Mehlhorn style; adjacent lane strips of one edge, the
       ear-clipped triangles of a junction shape);
    2. drop a polygon whose vertices all lie inside (or on) one other convex polygon.
    """
    rings: List[Optional[List[Tuple[float, float]]]] = []
    for p in polys:
        r = _clean_ring([(float(x), float(y)) for x, y in np.asarray(p, np.float64)], tol)
        if len(r) >= 3 and _is_convex(r, tol):
            rings.append(r)
        elif len(r) >= 3 and abs(_ring_area(r)) > 1e-9:
            raise ValueError("simplify_polygons expects convex counter-clockwise polygons")
        # else: zero-area input (repeated shape points) covers nothing
    q = 1e-5

    def key(pt):
        return (round(pt[0] / q), round(pt[1] / q))

    def try_merge_pass() -> bool:
        edge_owner: Dict[Tuple, Tuple[int, int]] = {}
        for pi, r in enumerate(rings):
            if r is None:
                continue
            for i in range(len(r)):
                edge_owner[(key(r[i]), key(r[(i + 1) % len(r)]))] = (pi, i)
        merged_any = False
        dirty = set()
        for pi, r in enumerate(rings):
            if r is None or pi in dirty:
                continue
            n = len(r)
            for i in range(n):
                rev = (key(r[(i + 1) % n]), key(r[i]))
                if rev[0] == rev[1]:
                    continue
                hit = edge_owner.get(rev)
                if hit is None:
                    continue
                pj, j = hit
                if pj == pi or rings[pj] is None or pj in dirty:
                    continue
                s = rings[pj]
                m = len(s)
                # r: ... r[i] -> r[i+1] ...; s: ... s[j] (= r[i+1]) -> s[j+1] (= r[i]) ...
                cand = [r[(i + 1 + k) % n] for k in range(n)]          # starts at r[i+1], ends at r[i]
                cand += [s[(j + 2 + k) % m] for k in range(m - 2)]     # s after s[j+1] up to before s[j]
                cand = _clean_ring(cand, tol)
                if 3 <= len(cand) <= max_verts and _is_convex(cand, tol) and \
                        abs(_ring_area(cand) - _ring_area(r) - _ring_area(s)) <= 1e-6 * (1.0 + _ring_area(cand)):
                    rings[pi] = cand
                    rings[pj] = None
                    dirty.add(pi); dirty.add(pj)
                    merged_any = True
                    break
        return merged_any

    def drop_contained() -> bool:
        live = [i for i, r in enumerate(rings) if r is not None]
        if not live:
            return False
        bb = np.array([[min(p[0] for p in rings[i]), min(p[1] for p in rings[i]),
                        max(p[0] for p in rings[i]), max(p[1] for p in rings[i])] for i in live])
        area = np.array([0.5 * sum(rings[i][k][0] * rings[i][(k + 1) % len(rings[i])][1] -
                                   rings[i][(k + 1) % len(rings[i])][0] * rings[i][k][1]
                                   for k in range(len(rings[i]))) for i in live])
        order = np.argsort(area)  # small first: try to drop the small ones
        dropped = False
        alive = np.ones(len(live), bool)
        eps = 1e-9
        for a in order:
            cand = np.nonzero(alive & (bb[:, 0] <= bb[a, 0] + eps) & (bb[:, 1] <= bb[a, 1] + eps) &
                              (bb[:, 2] >= bb[a, 2] - eps) & (bb[:, 3] >= bb[a, 3] - eps))[0]
            ra = rings[live[a]]
            for b in cand:
                if b == a:
                    continue
                rb = rings[live[b]]
                inside = True
                for k in range(len(rb)):
                    p0, p1 = rb[k], rb[(k + 1) % len(rb)]
                    ln = math.hypot(p1[0] - p0[0], p1[1] - p0[1])
                    for v in ra:
                        if _cross(p0, p1, v) < -eps * max(ln, 1.0):
                            inside = False
                            break
                    if not inside:
                        break
                if inside and (area[b] > area[a] or (area[b] == area[a] and b < a)):
                    alive[a] = False
                    rings[live[a]] = None
                    dropped = True
                    break
        return dropped

    def drop_covered() -> bool:
        """Step 3: drop a polygon covered by the union of the others (its area minus every
        overlapping polygon, tracked as convex pieces, falls below 1e-9 m^2)."""
        live = [i for i, r in enumerate(rings) if r is not None]
        if len(live) < 2:
            return False
        bb = np.array([[min(p[0] for p in rings[i]), min(p[1] for p in rings[i]),
                        max(p[0] for p in rings[i]), max(p[1] for p in rings[i])] for i in live])
        area = np.array([_ring_area(rings[i]) for i in live])
        alive = np.ones(len(live), bool)
        dropped = False
        for a in np.argsort(area):
            cand = np.nonzero(alive & (bb[:, 0] < bb[a, 2]) & (bb[:, 2] > bb[a, 0]) &
                              (bb[:, 1] < bb[a, 3]) & (bb[:, 3] > bb[a, 1]))[0]
            cand = [b for b in cand if b != a]
            if not cand:
                continue
            rest = [rings[live[a]]]
            for b in sorted(cand, key=lambda k: -area[k]):
                nxt = []
                for piece in rest:
                    nxt.extend(_convex_minus(piece, rings[live[b]]))
                rest = nxt
                if not rest:
                    break
            if sum(_ring_area(r) for r in rest) < 1e-9:
                alive[a] = False
                rings[live[a]] = None
                dropped = True
        return dropped

    def fixed_point():
        for _ in range(64):
            a = False
            while try_merge_pass():
                a = True
            b = drop_contained()
            c = drop_covered()
            if not (a or b or c):
                break

    fixed_point()
    # Step 4: make the pieces disjoint (largest first; every later polygon keeps only the convex
    # pieces of itself outside what is already there, slivers under 1e-9 m^2 dropped), so that a
    # raster row meets each part of the surface once; then merge again.
    rings[:] = _disjoint_pieces([r for r in rings if r is not None], max_verts, tol)
    fixed_point()
    return [np.asarray(r, np.float64) for r in rings if r is not None]


def _disjoint_pieces(rings, max_verts: int, tol: float):
    rings = sorted(rings, key=lambda r: -_ring_area(r))
    acc: List[List[Tuple[float, float]]] = []
    bbs: List[Tuple[float, float, float, float]] = []

    def bbox(r):
        return (min(p[0] for p in r), min(p[1] for p in r), max(p[0] for p in r), max(p[1] for p in r))

    for r in rings:
        rest = [r]
        bb = bbox(r)
        for q, qb in zip(acc, bbs):
            if qb[0] >= bb[2] or qb[2] <= bb[0] or qb[1] >= bb[3] or qb[3] <= bb[1]:
                continue
            nxt = []
            for piece in rest:
                nxt.extend(_convex_minus(piece, q))
            rest = nxt
            if not rest:
                break
        for piece in rest:
            piece = _clean_ring(piece, tol)
            if len(piece) < 3 or _ring_area(piece) <= 1e-9 or not _is_convex(piece, tol):
                continue
            stack = [piece]
            while stack:  # split rings with too many vertices along a diagonal (stays convex)
                g = stack.pop()
                if len(g) <= max_verts:
                    acc.append(g)
                    bbs.append(bbox(g))
                else:
                    h = len(g) // 2
                    stack.append(g[:h + 1])
                    stack.append(g[h:] + g[:1])
    return acc


def _ring_area(r: Sequence[Tuple[float, float]]) -> float:
    return 0.5 * sum(r[k][0] * r[(k + 1) % len(r)][1] - r[(k + 1) % len(r)][0] * r[k][1] for k in range(len(r)))


def _clip_halfplane(ring, p0, p1, keep_left: bool):
    """Part of a convex ring on the left (or right) of the directed line p0 -> p1."""
    out = []
    n = len(ring)
    sgn = 1.0 if keep_left else -1.0
    d = [sgn * _cross(p0, p1, v) for v in ring]
    for i in range(n):
        a, b, da, db = ring[i], ring[(i + 1) % n], d[i], d[(i + 1) % n]
        if da >= 0:
            out.append(a)
        if (da > 0 and db < 0) or (da < 0 and db > 0):
            t = da / (da - db)
            out.append((a[0] + (b[0] - a[0]) * t, a[1] + (b[1] - a[1]) * t))
    return out


def _convex_minus(p, q):
    """Convex pieces of P minus Q (both convex, CCW): for each edge of Q, the part of what is
    left of P outside that edge; what remains inside all edges is P intersect Q (discarded)."""
    pieces = []
    rest = list(p)
    for k in range(len(q)):
        if len(rest) < 3:
            break
        p0, p1 = q[k], q[(k + 1) % len(q)]
        outside = _clip_halfplane(rest, p0, p1, keep_left=False)
        if len(outside) >= 3 and _ring_area(outside) > 1e-12:
            pieces.append(outside)
        rest = _clip_halfplane(rest, p0, p1, keep_left=True)
    return pieces


def _ear_clip(line: Sequence[Tuple[float, float]]) -> List[List[Tuple[float, float]]]:
    """Triangulate a simple polygon (junction shape) into triangles."""
    pts = []
    for p in line:
        if not pts or p != pts[-1]:
            pts.append(p)
    if len(pts) > 1 and pts[0] == pts[-1]:
        pts.pop()
    if len(pts) < 3:
        return []
    area = 0.0
    for i in range(len(pts)):
        j = (i + 1) % len(pts)
        area += pts[i][0] * pts[j][1] - pts[j][0] * pts[i][1]
    if area < 0:
        pts = pts[::-1]
    idx = list(range(len(pts)))
    tris = []

    def cross(o, a, b):
        return (a[0] - o[0]) * (b[1] - o[1]) - (a[1] - o[1]) * (b[0] - o[0])

    guard = 0
    while len(idx) > 3 and guard < 10000:
        guard += 1
        n = len(idx)
        clipped = False
        for k in range(n):
            i0, i1, i2 = idx[(k - 1) % n], idx[k], idx[(k + 1) % n]
            a, b, c = pts[i0], pts[i1], pts[i2]
            cr = cross(a, b, c)
            if cr < 0:
                continue
            ok = True
            for m in idx:
                if m in (i0, i1, i2):
                    continue
                p = pts[m]
                if cross(a, b, p) > 0 and cross(b, c, p) > 0 and cross(c, a, p) > 0:
                    ok = False
                    break
            if ok:
                if cr > 0:
                    tris.append([a, b, c])
                idx.pop(k)
                clipped = True
                break
        if not clipped:
            break
    if len(idx) == 3:
        a, b, c = pts[idx[0]], pts[idx[1]], pts[idx[2]]
        if cross(a, b, c) > 0:
            tris.append([a, b, c])
    return tris
