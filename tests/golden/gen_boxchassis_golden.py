"""Golden vectors for k_set_social / the oracle's or_set_social from the reference's OWN BoxChassis
(smarts/core/chassis.py:187-320): `control(pose, speed, dt)`, `yaw_rate` (min_angles_difference_signed of the last two
headings over dt; None right after creation), `velocity_vectors` (heading vector * speed; difference of the heading
vectors over dt), `speed`, `pose`.  The class runs unmodified behind tests/golden/ref_shims.py with a do-nothing bullet
client (its pybullet calls only move a collision body).  A vehicle that disappears and comes back gets a new chassis, as
a vehicle re-entering the simulation does (vehicle_index.py builds a new Vehicle).  Output: tests/golden/boxchassis.npz.
Container-only script."""
from __future__ import annotations

import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, HERE)
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import ref_shims  # noqa: E402

ref_shims.install()

from smarts.core.chassis import BoxChassis  # noqa: E402
from smarts.core.coordinates import Dimensions, Heading, Pose  # noqa: E402

import test_boxchassis as tb  # noqa: E402


def main():
    states = tb.social_recording()
    T, E, M, _ = states.shape
    dt = tb.DT
    yaw = np.zeros((T, E, M)); heading = np.zeros((T, E, M)); speed = np.zeros((T, E, M))
    lin = np.zeros((T, E, M, 3)); ang = np.zeros((T, E, M, 3)); has_rate = np.zeros((T, E, M), np.uint8)
    chassis = {}
    client = ref_shims._Anything()
    for t in range(T):
        for e in range(E):
            for m in range(M):
                x, y, h, v, l, w, hh, alive = states[t, e, m]
                if not alive:
                    chassis.pop((e, m), None)
                    continue
                pose = Pose.from_center((x, y, 0.0), Heading(h))
                if (e, m) not in chassis:
                    chassis[e, m] = BoxChassis(pose, v, Dimensions(l, w, hh), client)   # control(pose, speed), dt = 0
                else:
                    chassis[e, m].control(pose, v, dt)
                c = chassis[e, m]
                yr = c.yaw_rate
                has_rate[t, e, m] = yr is not None
                yaw[t, e, m] = 0.0 if yr is None else yr
                heading[t, e, m] = float(c.pose.heading)
                speed[t, e, m] = c.speed
                lv, av = c.velocity_vectors
                lin[t, e, m], ang[t, e, m] = lv, av
    np.savez_compressed(os.path.join(HERE, "boxchassis.npz"), yaw_rate=yaw, heading=heading, speed=speed,
                        linear_velocity=lin, angular_velocity=ang, has_rate=has_rate)
    print("boxchassis.npz:", T, "ticks,", int(has_rate.sum()), "yaw rates, max |yaw rate|", float(np.abs(yaw).max()))


if __name__ == "__main__":
    main()
