"""Re-run smarts_b200.road_map.simplify_polygons over the drivable-area polygons of the committed
maps/*.npz (idempotent maintenance script; the maps themselves come from gen_golden.py, which
reads the reference's net.xml files and already simplifies when it builds a RoadMapData).

    python tests/golden/resimplify_maps.py [map ...]
"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from smarts_b200.road_map import RoadMapData, simplify_polygons, split_polygons  # noqa: E402


def main(names):
    for name in names:
        path = os.path.join(ROOT, "maps", name + ".npz")
        rm = RoadMapData.load(path)
        ps, px, py = rm.poly_start, rm.poly_x, rm.poly_y
        polys = [np.stack([px[ps[p]:ps[p + 1]], py[ps[p]:ps[p + 1]]], 1) for p in range(len(ps) - 1)]
        t = time.time()
        out = split_polygons(simplify_polygons(polys))
        vx, vy, st = [], [], [0]
        for q in out:
            vx.extend(q[:, 0]); vy.extend(q[:, 1]); st.append(len(vx))
        rm.poly_x, rm.poly_y, rm.poly_start = np.asarray(vx), np.asarray(vy), np.asarray(st, np.int32)
        rm.save(path)
        print(f"{name}: {len(polys)} -> {len(out)} polygons ({time.time() - t:.1f} s)")


if __name__ == "__main__":
    main(sys.argv[1:] or ["4lane", "loop", "figure_eight", "straight", "6lane", "minicity"])
