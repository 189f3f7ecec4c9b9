"""LiDAR sensor parameters and the base-ray table (host side, computed once).

Mirrors ``SensorParams`` / ``BasicLidar`` / ``VelodyneHDL32E``
(/root/reference/smarts/core/lidar_sensor_params.py:25-56) and the ray construction of
``Lidar._compute_rays`` (/root/reference/smarts/core/lidar.py:89-113).
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np


@dataclass(frozen=True)
class SensorParams:
    """Lidar sensor parameters (same fields as the reference dataclass)."""

    start_angle: float
    end_angle: float
    laser_angles: np.ndarray
    angle_resolution: float
    max_distance: float
    noise_mu: float
    noise_sigma: float


VelodyneHDL32E = SensorParams(
    start_angle=0,
    end_angle=2 * np.pi,
    laser_angles=np.linspace(-np.radians(30.67), np.radians(10.67), 24),
    angle_resolution=0.1728,
    max_distance=100,
    noise_mu=0,
    noise_sigma=0.078,
)

BasicLidar = SensorParams(
    start_angle=0,
    end_angle=2 * np.pi,
    laser_angles=np.linspace(-np.radians(4), np.radians(10), 50),
    angle_resolution=1,
    max_distance=20,
    noise_mu=0,
    noise_sigma=0.078,
)


def base_rays(params: SensorParams) -> np.ndarray:
    """[R, 3] end points of the rays relative to the sensor origin, in the order of
    ``itertools.product(yaws, rolls)`` (lidar.py:98-100).

    The reference rotates (0, max_distance, 0) by ``getQuaternionFromEuler((roll, 0, yaw))`` but
    feeds Bullet's [x, y, z, w] quaternion to ``rotate_quat`` which treats index 0 as the scalar
    (utils/math.py:109-133).  Carried through literally that gives the closed form below (roll
    acts as azimuth, the negated laser angle as elevation; rays live in the world frame).
    Checked against the reference's own ``_compute_rays`` in tests/golden/controllers.npz.
    """
    n_rays = int((params.end_angle - params.start_angle) / params.angle_resolution)
    yaws = -1 * np.asarray(params.laser_angles, np.float64)
    rolls = np.arange(n_rays) * params.angle_resolution
    eps = np.repeat(yaws, n_rays)
    az = np.tile(rolls, len(yaws))
    d = float(params.max_distance)
    return np.stack([-d * np.cos(eps) * np.sin(az), -d * np.cos(eps) * np.cos(az), d * np.sin(eps)], axis=1)
