// sb_control.cuh -- action-space controllers and the kinematic chassis step (device, scalar).
//   Controllers.perform_action            smarts/core/controllers/__init__.py:61-152
//   LaneFollowingController               smarts/core/controllers/lane_following_controller.py:63-437
//   TrajectoryTrackingController (PD)     smarts/core/controllers/trajectory_tracking_controller.py:176-473
//   ActuatorDynamicController             smarts/core/controllers/actuator_dynamic_controller.py:51-80
//   AckermannChassis.control / read-backs smarts/core/chassis.py:493-566, :678-794
#pragma once

#include "sb_device.cuh"

#include "../../include/sb_vehicle.h"

// Vehicle-type constants (sedan / bus) arrive in DCfg::veh, filled on the host by sb_vehicle() of include/sb_vehicle.h
// -- the oracle holds the same record.
#define SB_STEER_JOINT_GAIN 0.1  // POSITION_CONTROL default positionGain
#define SB_DYN_SPEED 1.0         // m/s: below it the tyres roll without slip

struct DEgo {
  double x, y, h, speed, yaw_rate, v_long, v_lat;
};

// chassis.py:678-718
__device__ __forceinline__ void d_apply_control(const DEgo& v, double thr, double brake, double steer, double* out3) {
  if (brake > 0 && v.v_long < 1.0 / 36.0) brake = 0;
  out3[0] = thr;
  out3[1] = brake;
  out3[2] = steer;
}

// trajectory_tracking_controller.py:443-473
__device__ double d_curvature(const double* tx, const double* ty, const double* th, int n, int offset,
                              int num_points) {
  if (n <= num_points + offset) return 1e20;
  double hs = 0.0, ds = 0.0;
  for (int i = 0; i < num_points; ++i) {
    hs += d_angdiff(th[i + 1 + offset], th[i + offset]);
    double dx = tx[i + offset] - tx[i + offset + 1], dy = ty[i + offset] - ty[i + offset + 1];
    ds += fabs(sqrt(dx * dx + dy * dy));
  }
  if (hs == 0.0) return 1e20;
  return ds / hs;
}

// lane_following_controller.py:376-437: single-input pole placement == Ackermann's formula.
__device__ void d_lateral_gains(double ts, const SbVehicle& vp, double* heading_gain, double* lateral_gain) {
  if (!(ts > 0)) {
    *heading_gain = 0.01;
    *lateral_gain = 0.36;
    return;
  }
  const double lh = vp.half_len, mass = vp.link_mass, iz = vp.link_izz, c = vp.road_stiffness;
  double Am[4][4] = {{0, ts, 0, ts},
                     {0, 0, 1, 0},
                     {0, 0, -(2 * c * (lh * lh)) / (ts * iz), 0},
                     {0, 0, -1, -2 * c / (mass * ts)}};
  double B[4] = {0, 0, lh * c / iz, c / (mass * ts)};
  const double poles[4] = {-35, -15, -2, -3};
  double Cm[4][4];
  double v[4] = {B[0], B[1], B[2], B[3]};
  for (int col = 0; col < 4; ++col) {
    for (int r = 0; r < 4; ++r) Cm[r][col] = v[r];
    double w[4];
    for (int r = 0; r < 4; ++r) {
      double s = 0;
      for (int k = 0; k < 4; ++k) s += Am[r][k] * v[k];
      w[r] = s;
    }
    for (int r = 0; r < 4; ++r) v[r] = w[r];
  }
  double Ph[4][4] = {{1, 0, 0, 0}, {0, 1, 0, 0}, {0, 0, 1, 0}, {0, 0, 0, 1}};
  for (int p = 0; p < 4; ++p) {
    double R[4][4];
    for (int r = 0; r < 4; ++r)
      for (int k = 0; k < 4; ++k) {
        double s = 0;
        for (int j = 0; j < 4; ++j) s += Ph[r][j] * (Am[j][k] - (j == k ? poles[p] : 0.0));
        R[r][k] = s;
      }
    for (int r = 0; r < 4; ++r)
      for (int k = 0; k < 4; ++k) Ph[r][k] = R[r][k];
  }
  double Mx[4][5];
  for (int r = 0; r < 4; ++r) {
    for (int k = 0; k < 4; ++k) Mx[r][k] = Cm[k][r];
    Mx[r][4] = (r == 3) ? 1.0 : 0.0;
  }
  for (int col = 0; col < 4; ++col) {
    int piv = col;
    for (int r = col + 1; r < 4; ++r)
      if (fabs(Mx[r][col]) > fabs(Mx[piv][col])) piv = r;
    if (piv != col)
      for (int k = 0; k < 5; ++k) {
        double t = Mx[piv][k];
        Mx[piv][k] = Mx[col][k];
        Mx[col][k] = t;
      }
    for (int r = col + 1; r < 4; ++r) {
      double f = Mx[r][col] / Mx[col][col];
      for (int k = col; k < 5; ++k) Mx[r][k] -= f * Mx[col][k];
    }
  }
  double yv[4];
  for (int r = 3; r >= 0; --r) {
    double s = Mx[r][4];
    for (int k = r + 1; k < 4; ++k) s -= Mx[r][k] * yv[k];
    yv[r] = s / Mx[r][r];
  }
  double K0 = 0, K1 = 0;
  for (int j = 0; j < 4; ++j) {
    K0 += yv[j] * Ph[j][0];
    K1 += yv[j] * Ph[j][1];
  }
  *heading_gain = d_clip(K1, 0.02, 0.04);
  *lateral_gain = d_clip(K0, 3.4, 4.1);
}

// lane_following_controller.py:105-361 once the waypoint path (x, y, heading; L points) is known.
__device__ void d_lane_following_law(const DEgo& v, const SbVehicle& vp, const double* wx, const double* wy,
                                     const double* wh, int L, double target_speed, int lane_change, double dt,
                                     double* st, double* out3) {
  double ewma = 0.0;
  for (int i = L - 2; i >= 0; --i) ewma = d_lerp(ewma, d_degrees(fabs(d_wrap_heading(wh[i] - wh[i + 1]))), 0.03);
  double curv = d_clip(ewma / 2.5, 0, 1);
  int npts = L < 10 ? L : 10;
  double lac = fabs(d_curvature(wx, wy, wh, npts, 4, 5));
  if (lac <= 2) {
    st[5] = wx[4];
    st[6] = wy[4];
    st[7] = 1.0;
  }
  int la = curv > 0.5 ? 3 : 4;
  if (la > L - 1) la = L - 1;
  double ref_h = wh[0];
  double ddx = wx[la] - v.x, ddy = wy[la] - v.y;
  double lad = sqrt(ddx * ddx + ddy * ddy);
  double sh, ch;
  sb_sincos(v.h, &sh, &ch);
  double lx = v.x - lad * sh, ly = v.y + lad * ch;
  double raw;
  if (curv < 0.3) {
    raw = -3.6 * 1.8 * (v.speed - target_speed);
  } else if (curv > 0.3 && curv < 0.8) {
    raw = -0.6 * 3.6 * (v.speed - d_clip(target_speed, 0, 6.94));
  } else {
    raw = -0.6 * 3.6 * (v.speed - d_clip(target_speed, 0, 5.56));
  }
  double speed_error = v.speed - target_speed;
  st[1] += speed_error * dt;
  double damping = (speed_error - st[4]) / dt;
  double lfc = 1.5;
  if (v.speed < 8 || target_speed < 6) lfc = 0;
  raw += (-0.2 * damping - 0.1 * st[1]) + fabs(lfc * sb_sin(st[2] * vp.max_steering_wheel));
  st[4] = speed_error;
  if (st[7] != 0.0) {
    double mx = v.x - st[5], my = v.y - st[6];
    if (sqrt(mx * mx + my * my) < 2) ref_h = wh[la];
  }
  double Kh, Kl;
  d_lateral_gains(target_speed, vp, &Kh, &Kl);
  double dvx, dvy;
  d_radians_to_vec(wh[la], &dvx, &dvy);
  double lat_err = d_signed_dist_to_line(lx, ly, wx[la], wy[la], dvx, dvy);
  double Rc = d_curvature(wx, wy, wh, npts, 0, 5);
  double brake = 0, thr;
  if (raw < 0) {
    brake = d_clip(-raw, 0, 1);
    thr = 0;
  } else {
    double tg;
    if (v.speed > 70 / 3.6 && fabs(Rc) <= 1e3)
      tg = 4.5;
    else if (40 / 3.6 <= v.speed && v.speed <= 70 / 3.6 && fabs(Rc) <= 3)
      tg = 2.5;
    else
      tg = 0.5;
    thr = d_clip(raw - tg * 3.6 * fabs(v.v_lat), 0, 1);
  }
  st[0] += dt * lat_err;
  double ffg = fabs(Rc) < 7 ? 0.45 : 0.15;
  double ff = 1 * ffg * (1 / Rc) * (v.speed * v.speed);
  double ns = d_clip(v.speed * 3.6 / 100, 0, 1);
  double hsg = -d_lerp(0.5, 14, ns);
  double yrg = d_lerp(5.75, 11.75, ns);
  double lsg = d_clip(d_lerp(-1, 14, ns), 1, 2);
  double max_s = 1;
  if (fabs(Rc) > 1e7 && lane_change != 0) {
    hsg = -4.95;
    yrg = 1;
    lsg = 0.22;
    max_s = 0.12;
  }
  double herr = d_angdiff(sb_pymod(v.h, SB_TWO_PI), ref_h);
  double steer =
      d_clip(-hsg * d_degrees(Kh) * herr + lsg * Kl * lat_err + yrg * v.yaw_rate + 0.3 * st[0] - ff, -max_s, max_s);
  st[2] = d_low_pass(steer, st[2], 5.5, dt, -1.0, 0.0);
  st[3] = d_low_pass(thr, st[3], 2, dt, 0.0, 0.0);
  d_apply_control(v, st[3], brake, st[2], out3);
}

// trajectory_tracking_controller.py:176-441 (the vehicle type's control block of controller_parameters.yaml)
__device__ void d_trajectory_tracking(const DEgo& v, const SbVehicle& vp, const double* traj, int N, double dt,
                                      double* st, double* out3) {
  const double* tx = traj;
  const double* ty = traj + N;
  const double* th = traj + 2 * N;
  const double* tv = traj + 3 * N;
  double throttle_filter_constant = vp.throttle_filter_constant;
  const bool speed_reduction = vp.speed_reduction_activation != 0;
  double lateral_gain = 0.61, heading_gain = 0.01, led_gain = 0.15, hed_gain = 0.5;
  double ns = d_clip((3.6 * tv[0] - 20) / (80 - 20), 0, 1);
  if (v.speed > 70 / 3.6) {
    lateral_gain = 1.51;
    hed_gain = 0.1;
  }
  double steering_filter_constant = d_lerp(12, vp.final_steering_filter_constant, ns);
  if (fabs(d_angdiff(th[N - 1], th[0])) > 2) throttle_filter_constant = 2.5;
  if (fabs(d_curvature(tx, ty, th, N, 0, 3)) < 150) {
    heading_gain = 0.05;
    led_gain = 0.015;
    hed_gain = 0.05;
  }
  double heading_error = d_angdiff(sb_pymod(v.h, SB_TWO_PI), th[0]);
  double look_dist = vp.initial_look_ahead_distant;
  int look_pts = (int)vp.initial_look_ahead_distant;
  double ahead_curv = fabs(d_curvature(tx, ty, th, N, 4, 5));
  if (ahead_curv < 30 && speed_reduction) {
    look_dist = 1;
    look_pts = 1;
  }
  int j = look_pts < N - 1 ? look_pts : N - 1;
  double pvx, pvy;
  d_radians_to_vec(th[j], &pvx, &pvy);
  double sh, ch;
  sb_sincos(v.h, &sh, &ch);
  double lx = v.x - look_dist * sh, ly = v.y + look_dist * ch;
  double lateral_error = d_signed_dist_to_line(lx, ly, tx[j], ty[j], pvx, pvy);
  double Rc = d_curvature(tx, ty, th, N, 0, 5);
  double derivative_term = hed_gain * v.yaw_rate + led_gain * (lateral_error - st[1]) / dt;
  double ff = 0.1 * (1 / Rc) * (v.speed * v.speed);
  double steering_raw =
      d_clip(vp.derivative_activation * derivative_term + d_degrees(heading_gain * heading_error) +
                 1 * lateral_gain * lateral_error - ff,
             -1, 1);
  st[5] = d_low_pass(steering_raw, st[5], steering_filter_constant, dt, -1.0, 0.0);
  double desired_speed = tv[N - 1];
  if (ahead_curv < 30 && speed_reduction)
    desired_speed = d_clip(0.8 * desired_speed, 0, 8.3);
  else if (ahead_curv < 100 && speed_reduction)
    desired_speed *= 0.8;
  double velocity_error = v.speed - desired_speed;
  double ved = (velocity_error - st[2]) / dt;
  double raw_throttle = 3.6 * (-0.5 * vp.velocity_gain * velocity_error -
                               vp.velocity_integral_gain * (st[3] + vp.windup_gain * st[4]) - vp.velocity_damping_gain * ved);
  st[2] = velocity_error;
  st[4] = d_clip(raw_throttle, -1, 1) - raw_throttle;
  st[6] = d_low_pass(raw_throttle, st[6], throttle_filter_constant, dt, -1.0, -vp.traction_gain * fabs(v.v_lat));
  double thr, brake;
  if (st[6] > 0) {
    brake = 0;
    thr = d_clip(st[6], 0, 1);
  } else {
    brake = d_clip(-st[6], 0, 1);
    thr = 0;
  }
  st[0] = heading_error;
  st[1] = lateral_error;
  st[3] += (v.speed - desired_speed) * dt;
  d_apply_control(v, thr, brake, st[5], out3);
}

// Kinematic stand-in for the Bullet AckermannChassis step (DESIGN.md "Vehicle model").
__device__ void d_chassis_step(const DState& s, const SbVehicle& vp, int ai, int vi, double dt, int substeps) {
  const double* c = &s.last_control[(size_t)ai * 3];
  double target = -c[2] * vp.max_steering_wheel;  // joint angle, left positive (chassis.py:790-793)
  target = d_clip(target, -vp.steer_limit, vp.steer_limit);
  const double ax = 4.0 * (c[0] * vp.max_torque - c[1] * vp.max_btorque) / vp.wheel_radius / vp.eff_mass;
  const double hdt = dt / (double)substeps;
  double x = s.pos_x[vi], y = s.pos_y[vi], h = s.heading[vi];
  double vx = s.v_long[ai], vy = -s.v_lat[ai], om = s.yaw_rate[vi], delta = -s.steering[ai];
  for (int k = 0; k < substeps; ++k) {
    delta = delta + SB_STEER_JOINT_GAIN * (target - delta);
    double sd, cd;
    sb_sincos(delta, &sd, &cd);
    if (vx > SB_DYN_SPEED) {
      const double af = (vy + vp.axle_to_cg * om) / vx - delta, ar = (vy - vp.axle_to_cg * om) / vx;
      const double fyf = d_clip(-vp.corner_stiffness * af, -vp.tyre_force_max, vp.tyre_force_max) * cd;
      const double fyr = d_clip(-vp.corner_stiffness * ar, -vp.tyre_force_max, vp.tyre_force_max);
      const double ay = (fyf + fyr) / vp.mass - vx * om;
      const double dom = (vp.axle_to_cg * fyf - vp.axle_to_cg * fyr) / vp.izz;
      vy = vy + ay * hdt;
      om = om + dom * hdt;
    } else {  // rolling without slip
      om = vx * (sd / cd) / (2.0 * vp.axle_to_cg);
      vy = om * vp.axle_to_cg;
    }
    vx = vx + (ax + vy * om) * hdt;
    if (vx < 0) vx = 0;
    double sn, cs;
    sb_sincos(h, &sn, &cs);
    x = x + (-(vx * sn) - vy * cs) * hdt;
    y = y + (vx * cs - vy * sn) * hdt;
    h = h + om * hdt;
  }
  h = d_wrap_heading(h);
  s.pos_x[vi] = x;
  s.pos_y[vi] = y;
  s.heading[vi] = h;
  s.speed[vi] = sqrt(vx * vx + vy * vy);
  s.yaw_rate[vi] = om;
  sb_sincos(h, &s.sin_h[vi], &s.cos_h[vi]);
  s.steering[ai] = -delta;  // chassis.py:513-528
  s.v_long[ai] = vx;        // chassis.py:558-566
  s.v_lat[ai] = -vy;
}
