"""Golden waypoint / nearest-lane queries on scenarios/minicity (BASELINE configs[3]'s map: 2 695 lanes, 52 k lane
points, internal lanes, 2 113 groups of exactly coincident lane points) from the reference's own
`SumoRoadNetwork.waypoint_paths` / `nearest_lanes`, as tests/golden/gen_golden.py does for the small maps.  The one
difference: the order among lane points at exactly the same distance -- scipy's cKDTree leaves it to its internal
partition -- is canonicalised to "lowest index in the flattened map first" (ref_shims.canonical_kd_ties /
set_tie_ranks), the rule of the oracle and the kernels.  Output: tests/golden/queries_minicity.npz.  Container-only."""
from __future__ import annotations

import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)

import gen_golden as gg  # noqa: E402

gg.ROUTES["minicity"] = [("4212249#1", "-34081173#1"), ("27616622#1", "-354354332#2"), ("33023072#2", "695649266")]


def main():
    rng = np.random.default_rng(20261104)
    rec, _ = gg.gen_map_queries("minicity", gg.MAPS["minicity"], rng, n=24, tie_ranks=True, save_map="--save-map" in sys.argv)
    np.savez_compressed(os.path.join(HERE, "queries_minicity.npz"), **rec)
    print("minicity paths:", len(rec["wp_meta"]), "rows:", len(rec["wp_rows"]), "nearest lanes found:",
          int((rec["nl_lane"] >= 0).sum()), "of", rec["nl_lane"].size)


if __name__ == "__main__":
    main()
