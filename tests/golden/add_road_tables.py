"""Add the RoadWaypoints tables (incoming lanes per lane, parallel roads per road: smarts_b200.road_map._RW_KEYS)
to the committed maps/*.npz from the reference's net.xml files, leaving every other array as it is.
Container-only (reads /root/reference).

    python tests/golden/add_road_tables.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from smarts_b200.road_map import RoadMapData  # noqa: E402
from smarts_b200.sumo_net import SumoNet  # noqa: E402

REF = "/root/reference/"
MAPS = {
    "loop": "scenarios/loop/map.net.xml",
    "figure_eight": "scenarios/figure_eight/map.net.xml",
    "4lane": "scenarios/intersections/4lane/map.net.xml",
    "straight": "scenarios/straight/map.net.xml",
    "6lane": "smarts/core/tests/maps/6lane.net.xml",
    "minicity": "scenarios/minicity/map.net.xml",
}
MAPS.update({f"turning_radius_{r}": f"smarts/core/tests/maps/turning_radius/map{r}.net.xml" for r in (55, 78, 90, 107, 128)})


def main():
    for name, rel in MAPS.items():
        path = os.path.join(ROOT, "maps", name + ".npz")
        rm = RoadMapData.load(path)
        fresh = RoadMapData.__new__(RoadMapData)
        fresh.net = SumoNet(REF + rel)
        fresh._build_tables()
        assert fresh._lane_ids == rm._lane_ids and fresh._road_ids == rm._road_ids, name
        assert np.array_equal(fresh.lane_out, rm.lane_out) and np.array_equal(fresh.conn_from, rm.conn_from), name
        for k in RoadMapData._RW_KEYS:
            setattr(rm, k, getattr(fresh, k))
        rm.save(path)
        chk = RoadMapData.load(path)
        print(f"{name}: {len(rm.lane_ids)} lanes, {int(rm.lane_in_start[-1])} incoming links, "
              f"{int(rm.road_par_start[-1])} parallel-road entries;",
              "reload ok" if all(np.array_equal(getattr(chk, k), getattr(rm, k)) for k in RoadMapData._RW_KEYS) else "RELOAD MISMATCH")


if __name__ == "__main__":
    main()
