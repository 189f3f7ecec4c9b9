"""Flatten the five maps of the reference's lane-following envelope test
(smarts/core/tests/test_controller_lane.py:60-68: smarts/core/tests/maps/turning_radius/map{55,78,90,107,128}.net.xml)
into maps/turning_radius_<r>.npz with smarts_b200.road_map.  Container-only (reads /root/reference)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from smarts_b200.road_map import RoadMapData  # noqa: E402

for r in (55, 78, 90, 107, 128):
    rm = RoadMapData.from_file(f"/root/reference/smarts/core/tests/maps/turning_radius/map{r}.net.xml")
    rm.save(os.path.join(ROOT, "maps", f"turning_radius_{r}.npz"))
    print(r, "lanes", len(rm.lane_ids), "lane points", len(rm.lp_x), "polygons", len(rm.poly_start) - 1, "roads", rm.road_ids)
