"""SUMO ``net.xml`` reader (setup-time, host side).

The reference reads maps through ``sumolib.net.readNet(net_file, withInternal=True)``
(/root/reference/smarts/core/sumo_road_network.py:154) and then only uses a handful of
accessors (``Lane.getShape/getWidth/getSpeed/getIndex/getLength/getOutgoing/getIncoming``,
``Edge.getLanes/isSpecial/getOutgoing``, ``Connection.getViaLaneID/getToLane``,
``Node.getShape``; call sites lanepoints.py:136-210, sumo_road_network.py:281-358,
:561, :689-695, :986-1019).  sumolib (eclipse-sumo 1.10.0, setup.py:28) is not vendored in
the reference, so this module restates exactly that accessor surface on top of
``xml.etree``; nothing here runs per tick.
"""
from __future__ import annotations

import heapq
import xml.etree.ElementTree as ET
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence, Tuple

DEFAULT_LANE_WIDTH = 3.2  # sumo_road_network.py:80


def _parse_shape(text: str) -> List[Tuple[float, float]]:
    pts = []
    for tok in text.split():
        c = tok.split(",")
        pts.append((float(c[0]), float(c[1])))
    return pts


@dataclass
class Connection:
    """One ``<connection>`` element (sumolib.net.connection.Connection)."""

    from_lane: "NetLane"
    to_lane: "NetLane"
    via_lane_id: str  # "" when absent

    # sumolib-style accessors so the golden generator can hand these objects to the
    # reference code unchanged.
    def getToLane(self):
        return self.to_lane

    def getFromLane(self):
        return self.from_lane

    def getViaLaneID(self):
        return self.via_lane_id

    def getFrom(self):
        return self.from_lane.edge

    def getTo(self):
        return self.to_lane.edge


@dataclass(eq=False)
class NetLane:
    lane_id: str
    index: int
    speed: float
    length: float
    width: float
    shape: List[Tuple[float, float]]
    edge: "NetEdge" = None
    outgoing: List[Connection] = field(default_factory=list)
    incoming_lanes: List["NetLane"] = field(default_factory=list)

    def getID(self):
        return self.lane_id

    def getIndex(self):
        return self.index

    def getSpeed(self):
        return self.speed

    def getLength(self):
        return self.length

    def getWidth(self):
        return self.width

    def getEdge(self):
        return self.edge

    def getShape(self, includeJunctions: bool = False):
        # See DESIGN.md "third-party behaviour": plain shape; junction positions are not
        # appended (sumolib 1.10 default ``includeJunctions=False``).
        return self.shape

    def getOutgoing(self):
        return self.outgoing

    def getIncoming(self, onlyDirect: bool = False):
        return self.incoming_lanes

    def getConnection(self, to_lane):
        for c in self.outgoing:
            if c.to_lane is to_lane:
                return c
        return None


@dataclass(eq=False)
class NetEdge:
    edge_id: str
    function: str
    from_node: str
    to_node: str
    lanes: List[NetLane] = field(default_factory=list)
    # insertion-ordered like sumolib's dict: to-edge -> [Connection]
    outgoing: Dict["NetEdge", List[Connection]] = field(default_factory=dict)
    incoming: Dict["NetEdge", List[Connection]] = field(default_factory=dict)
    from_node_obj: "NetNode" = None
    to_node_obj: "NetNode" = None

    def getID(self):
        return self.edge_id

    def getLanes(self):
        return self.lanes

    def getLane(self, idx):
        return self.lanes[idx]

    def getFunction(self):
        return self.function

    def isSpecial(self):
        # sumolib: function in ("internal", "crossing", "walkingarea", "connector")
        return self.function != ""

    def getLength(self):
        return self.lanes[0].length

    def getOutgoing(self):
        return self.outgoing

    def getIncoming(self):
        return self.incoming

    def getConnections(self, to_edge):
        return self.outgoing.get(to_edge, [])

    def getFromNode(self):
        return self.from_node_obj

    def getToNode(self):
        return self.to_node_obj


@dataclass(eq=False)
class NetNode:
    node_id: str
    node_type: str
    coord: Tuple[float, float]
    shape: List[Tuple[float, float]]
    outgoing: List[NetEdge] = field(default_factory=list)  # edges leaving / entering, in edge file order
    incoming: List[NetEdge] = field(default_factory=list)

    def getID(self):
        return self.node_id

    def getOutgoing(self):
        return self.outgoing

    def getIncoming(self):
        return self.incoming

    def getType(self):
        return self.node_type

    def getShape(self):
        return self.shape

    def getCoord(self):
        return self.coord


class SumoNet:
    """Parsed road network; duck-types the part of ``sumolib.net.Net`` the reference uses."""

    def __init__(self, net_file: str, default_lane_width: float = DEFAULT_LANE_WIDTH):
        self.net_file = net_file
        self.edges: List[NetEdge] = []
        self.nodes: List[NetNode] = []
        self.connections: List[Connection] = []  # <connection> elements in file order
        self._edge_by_id: Dict[str, NetEdge] = {}
        self._lane_by_id: Dict[str, NetLane] = {}
        self._parse(net_file, default_lane_width)

    # -- parsing -------------------------------------------------------------------------
    def _parse(self, net_file: str, default_lane_width: float):
        root = ET.parse(net_file).getroot()
        conns = []
        for el in root:
            if el.tag == "edge":
                fn = el.get("function", "")
                edge = NetEdge(el.get("id"), fn, el.get("from", ""), el.get("to", ""))
                for lel in el.findall("lane"):
                    w = lel.get("width")
                    lane = NetLane(
                        lane_id=lel.get("id"),
                        index=int(lel.get("index")),
                        speed=float(lel.get("speed")),
                        length=float(lel.get("length")),
                        width=float(w) if w is not None else default_lane_width,
                        shape=_parse_shape(lel.get("shape")),
                        edge=edge,
                    )
                    edge.lanes.append(lane)
                    self._lane_by_id[lane.lane_id] = lane
                edge.lanes.sort(key=lambda l: l.index)
                self.edges.append(edge)
                self._edge_by_id[edge.edge_id] = edge
            elif el.tag == "junction":
                shape = _parse_shape(el.get("shape", "")) if el.get("shape") else []
                self.nodes.append(
                    NetNode(
                        el.get("id"),
                        el.get("type", ""),
                        (float(el.get("x")), float(el.get("y"))),
                        shape,
                    )
                )
            elif el.tag == "connection":
                conns.append(el)
        for el in conns:
            fe = self._edge_by_id.get(el.get("from"))
            te = self._edge_by_id.get(el.get("to"))
            if fe is None or te is None:
                continue
            fl = fe.lanes[int(el.get("fromLane"))]
            tl = te.lanes[int(el.get("toLane"))]
            c = Connection(fl, tl, el.get("via", "") or "")
            self.connections.append(c)
            fl.outgoing.append(c)
            tl.incoming_lanes.append(fl)
            fe.outgoing.setdefault(te, []).append(c)
            te.incoming.setdefault(fe, []).append(c)

        # Edge.getFromNode / getToNode and Node.getOutgoing / getIncoming (Road.parallel_roads,
        # sumo_road_network.py:607-618).  sumolib's reader gives an internal edge ":<junction>_<k>" the junction
        # itself as both end nodes, and links every edge to its nodes as it is added (file order).
        node_by_id = {n.node_id: n for n in self.nodes}
        for e in self.edges:
            f_id, t_id = e.from_node, e.to_node
            if e.function in ("internal", "crossing", "walkingarea"):
                f_id = t_id = e.edge_id[1:e.edge_id.rfind("_")]
            for attr, nid, lst in (("from_node_obj", f_id, "outgoing"), ("to_node_obj", t_id, "incoming")):
                n = node_by_id.get(nid)
                if n is None:  # sumolib's addNode creates a bare node for an id the file never declares
                    n = node_by_id[nid] = NetNode(nid, "", (0.0, 0.0), [])
                setattr(e, attr, n)
                getattr(n, lst).append(e)

    # -- sumolib.net.Net surface ---------------------------------------------------------
    def getEdges(self, withInternal: bool = True) -> List[NetEdge]:
        if withInternal:
            return self.edges
        return [e for e in self.edges if e.function == ""]

    def getEdge(self, edge_id: str) -> Optional[NetEdge]:
        return self._edge_by_id.get(edge_id)

    def getLane(self, lane_id: str) -> Optional[NetLane]:
        return self._lane_by_id.get(lane_id)

    def hasEdge(self, edge_id: str) -> bool:
        return edge_id in self._edge_by_id

    def getNodes(self) -> List[NetNode]:
        return self.nodes

    def getBoundary(self):
        xs, ys = [], []
        for e in self.edges:
            for l in e.lanes:
                for x, y in l.shape:
                    xs.append(x)
                    ys.append(y)
        for n in self.nodes:
            xs.append(n.coord[0])
            ys.append(n.coord[1])
        return (min(xs), min(ys), max(xs), max(ys))

    def getShortestPath(self, from_edge: NetEdge, to_edge: NetEdge):
        """Dijkstra over normal edges weighted by edge length (sumolib.net.Net.getShortestPath;
        used by sumo_road_network.py:733-738).  Returns (edge list | None, cost)."""
        seen = set()
        q = [(from_edge.getLength(), 0, from_edge, (from_edge,))]
        tie = 1
        while q:
            cost, _, e, path = heapq.heappop(q)
            if e in seen:
                continue
            seen.add(e)
            if e is to_edge:
                return list(path), cost
            for ne in e.outgoing.keys():
                if ne.function != "" or ne in seen:
                    continue
                heapq.heappush(q, (cost + ne.getLength(), tie, ne, path + (ne,)))
                tie += 1
        return None, 1e400
