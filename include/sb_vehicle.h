/*
 * sb_vehicle.h -- the per-vehicle-type constants of the controlled (agent) vehicles, in ONE place for the CUDA
 * library and the CPU oracle.  AgentInterface.vehicle_type (smarts/core/agent_interface.py:246-250) selects the URDF
 * (smarts/core/vehicle.py:393-398: "sedan" -> models/vehicle.urdf, "bus" -> models/bus.urdf) and the block of
 * models/controller_parameters.yaml handed to the chassis and to the trajectory-tracking controller
 * (vehicle.py:404-406, chassis.py:343-362, trajectory_tracking_controller.py:187-205).
 *
 * Plain C, evaluated on the host (no FMA contraction: x86-64 baseline / -ffp-contract=off), then passed by value
 * to the kernels, so both sides hold bit-identical doubles.
 */
#ifndef SB_VEHICLE_H
#define SB_VEHICLE_H

#define SB_VEHICLE_SEDAN 0
#define SB_VEHICLE_BUS 1

typedef struct SbVehicle {
  /* chassis box: the URDF collision box (vehicle.urdf:12-16 "1.47 3.68 1.0" at z 0.6, bus.urdf:13-15 "2.2026 7 2.4"
   * at z 1.4) -- Vehicle.length / width / height, and the box other vehicles' lidars and contacts see */
  double length, width, height, box_z0;
  /* LaneFollowingController.calculate_lateral_gains (lane_following_controller.py:387-389): vehicle.length / 2 and
   * chassis.mass_and_inertia = getDynamicsInfo of the chassis link (vehicle.urdf:19-21, bus.urdf:19-21) */
  double half_len, link_mass, link_izz, road_stiffness;
  /* chassis block of controller_parameters.yaml */
  double max_steering_wheel; /* max_steering / steering_gear_ratio (chassis.py:611-613) */
  double steer_limit;        /* steer joint limit (vehicle.urdf:101, bus.urdf:101) */
  double wheel_radius, max_torque, max_btorque;
  /* single-track stand-in for the Bullet multibody (DESIGN.md "Vehicle model") */
  double eff_mass;       /* all links + the wheels' rotational inertia 4 Izz / r^2 */
  double axle_to_cg;     /* wheel joints at y = +-1.5 (vehicle.urdf:96-135) / +-2.5 (bus.urdf:96-135) */
  double mass, izz;      /* chassis + wheels; yaw inertia with the wheels as point masses */
  double corner_stiffness, tyre_force_max;
  /* control block of controller_parameters.yaml as perform_trajectory_tracking_PD reads it (:187-205) */
  double final_steering_filter_constant, throttle_filter_constant, velocity_gain, velocity_integral_gain,
      traction_gain, initial_look_ahead_distant, derivative_activation, velocity_damping_gain, windup_gain;
  int speed_reduction_activation;
  int type;
} SbVehicle;

static inline void sb_vehicle(int type, SbVehicle *v) {
  const double r = 0.31265; /* wheel_radius, both blocks */
  v->type = type;
  v->road_stiffness = 100000.0; /* sim.road_stiffness: plane.urdf:26 contact stiffness */
  v->corner_stiffness = 100000.0;
  v->steer_limit = 0.8727;
  v->wheel_radius = r;
  v->max_steering_wheel = 12.56 / 17.4;
  v->velocity_integral_gain = 0;
  v->derivative_activation = 1;
  v->velocity_damping_gain = 0.001;
  v->windup_gain = 0.01;
  if (type == SB_VEHICLE_BUS) {
    v->length = 7.0;
    v->width = 2.2026;
    v->height = 2.4;
    v->box_z0 = 0.2; /* box centre at z 1.4, 2.4 tall */
    v->half_len = 7.0 / 2;
    v->link_mass = 6000.0;
    v->link_izz = 26920.0;
    v->max_torque = 2100.0;
    v->max_btorque = 1200.0;
    /* 1 (base) + 6000 (chassis) + 2 * 5 (steer links) + 4 * 25 (wheels), wheel Izz 4 (bus.urdf:7-81) */
    v->eff_mass = 6111.0 + 4.0 * 4.0 / (r * r);
    v->axle_to_cg = 2.5;
    v->mass = 6000.0 + 4.0 * 25.0;
    v->izz = 26920.0 + 4.0 * 25.0 * (2.5 * 2.5 + 0.5 * 0.5);
    v->final_steering_filter_constant = 27.5;
    v->throttle_filter_constant = 2;
    v->velocity_gain = 6.1;
    v->traction_gain = 8;
    v->initial_look_ahead_distant = 3;
    v->speed_reduction_activation = 0;
  } else {
    v->length = 3.68;
    v->width = 1.47;
    v->height = 1.0;
    v->box_z0 = 0.1; /* box centre at z 0.6, 1.0 tall */
    v->half_len = 1.84;
    v->link_mass = 2356.0;
    v->link_izz = 2681.95008628;
    v->max_torque = 1600.0;
    v->max_btorque = 1400.0;
    /* 1 + 2356 + 2 * 1 + 4 * 15 kg, wheel Izz 2 (vehicle.urdf) */
    v->eff_mass = 2419.0 + 4.0 * 2.0 / (r * r);
    v->axle_to_cg = 1.5;
    v->mass = 2356.0 + 4.0 * 15.0;
    v->izz = 2681.95008628 + 4.0 * 15.0 * (1.5 * 1.5 + 0.5 * 0.5);
    v->final_steering_filter_constant = 23.5;
    v->throttle_filter_constant = 22.5;
    v->velocity_gain = 5.1;
    v->traction_gain = 6;
    v->initial_look_ahead_distant = 6;
    v->speed_reduction_activation = 1;
  }
  v->tyre_force_max = 3.0 * v->mass * 9.8 / 2.0; /* per axle: plane.urdf:24 against a wheel friction of 1, g = 9.8 */
}

#endif /* SB_VEHICLE_H */
