"""Golden vectors for smarts_b200/traffic.py from the reference's OWN code, run in this container behind
the import shims of ref_shims.py: Heading.from_sumo + Pose.from_front_bumper + VEHICLE_CONFIGS exactly as
sumo_traffic_simulation.py:672-705 uses them.  Writes tests/golden/traffic.npz.  Container-only: reads
/root/reference; the tests read only the committed .npz."""
from __future__ import annotations

import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

import ref_shims  # noqa: E402

ref_shims.install()

from smarts.core.coordinates import Heading, Pose  # noqa: E402


def reference_vehicle_dims():
    """VEHICLE_CONFIGS of smarts/core/vehicle.py:97-140, read from the reference's source text (importing the
    module pulls in smarts.sstudio, which needs packages absent here)."""
    import re
    src = open(os.path.join(ref_shims.REF, "smarts/core/vehicle.py")).read()
    pat = re.compile(r'"(\w+)":\s*VehicleConfig\(.*?Dimensions\(length=([\d.]+),\s*width=([\d.]+),\s*height=([\d.]+)\)', re.S)
    return {m.group(1): tuple(float(m.group(k)) for k in (2, 3, 4)) for m in pat.finditer(src)}


def main():
    rng = np.random.default_rng(20241105)
    n = 64
    VEHICLE_DIMS = reference_vehicle_dims()
    assert len(VEHICLE_DIMS) == 7, VEHICLE_DIMS
    classes = sorted(VEHICLE_DIMS.keys())
    vclass = rng.choice(classes, n)
    front = rng.uniform(-500, 500, (n, 2))
    angle = np.concatenate([[0.0, 90.0, 180.0, 270.0, 360.0, -45.0, 725.5], rng.uniform(-360, 720, n - 7)])
    speed = rng.uniform(0, 30, n)
    centre = np.zeros((n, 2)); heading = np.zeros(n); dims = np.zeros((n, 3))
    for i in range(n):
        h = Heading.from_sumo(float(angle[i]))
        d = VEHICLE_DIMS[str(vclass[i])]
        pose = Pose.from_front_bumper(front[i], h, d[0])
        centre[i] = pose.position[:2]
        heading[i] = float(h)
        dims[i] = d
    np.savez_compressed(os.path.join(HERE, "traffic.npz"), front=front, angle=angle, speed=speed, vclass=vclass,
                        centre=centre, heading=heading, dims=dims)
    print("traffic.npz:", n, "vehicles")


if __name__ == "__main__":
    main()
