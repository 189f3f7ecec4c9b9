"""Golden vectors for vehicle_type="bus" (AgentInterface.vehicle_type, smarts/core/vehicle.py:393-406) from the
reference's OWN controllers, as tests/golden/gen_golden.py produces them for the sedan:
  * LaneFollowingController.perform_lane_following closed loop on a mock chassis that reports the bus's
    Vehicle.length (7 m, models/bus.urdf:13-15) and chassis.mass_and_inertia (6000 kg, 26 920 kg m^2, bus.urdf:19-21);
  * TrajectoryTrackingController.perform_trajectory_tracking_PD with the `bus` control block of
    models/controller_parameters.yaml (read from the reference's file);
  * scipy.signal.place_poles gains of calculate_lateral_gains for the bus's linearisation.
Output: tests/golden/controllers_bus.npz.  Container-only script."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import gen_golden as gg  # noqa: E402  (installs the import shims)

BUS = dict(length=7.0, mass_and_inertia=(6000.0, 26920.0))


def main():
    rng = np.random.default_rng(20261105)
    rm, mine, routes = gg.gen_map_queries("4lane", gg.MAPS["4lane"], rng, n=2, save_map=False)[1]
    lane = gg.gen_lane_controller(rm, mine, routes, rng, **BUS)
    traj_rec, trajs = gg.gen_traj_controller(rng, vehicle_type="bus")
    gains = gg.gen_gains(lh=BUS["length"] / 2, m=BUS["mass_and_inertia"][0], iz=BUS["mass_and_inertia"][1])
    np.savez_compressed(os.path.join(HERE, "controllers_bus.npz"), lane_4lane=lane, traj_rec=traj_rec, trajs=trajs, gains=gains)
    print("bus controllers:", lane.shape, traj_rec.shape, gains.shape)


if __name__ == "__main__":
    main()
