"""Rebuild the waypoint / lane query fixtures from /root/reference in memory and compare them with the committed
queries_*.npz (container-only: needs the reference tree).  Nothing is written.  Exit code 1 on any difference.

    python tests/golden/regen_check.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, HERE)

import gen_golden as gg  # noqa: E402


def main() -> int:
    rng = np.random.default_rng(20241104)  # gen_golden.main()'s seed; the maps draw from one stream, in MAPS order
    bad = 0
    for name, rel in gg.MAPS.items():
        if name == "minicity":  # its fixture comes from gen_minicity_golden.py
            continue
        rec, _ = gg.gen_map_queries(name, rel, rng, tie_ranks=(name == "4lane"), save_map=False)
        z = np.load(os.path.join(HERE, f"queries_{name}.npz"))
        diff = [k for k in z.files
                if not np.array_equal(np.asarray(rec[k]), z[k], equal_nan=np.asarray(rec[k]).dtype.kind == "f")]
        same = not diff and set(rec) == set(z.files)
        print(f"queries_{name}.npz:", "identical" if same else f"DIFFERS in {diff}")
        bad += not same
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
