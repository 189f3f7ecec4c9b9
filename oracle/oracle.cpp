/*
 * oracle.cpp -- second half of the CPU oracle: controllers, chassis, sensors and the C entry
 * points used by tests/ and bench.py (ctypes).  TEST INFRASTRUCTURE ONLY -- see the header of
 * oracle_core.hpp for what may call this and for the parity status of each part.
 */
#include "oracle_core.hpp"

#include <atomic>
#include <thread>

namespace {

// envs are independent, so the CPU baseline runs them on plain std::threads (dynamic chunks)
template <class F>
void parallel_for(int n, int threads, int chunk, F fn) {
  if (threads <= 1 || n <= chunk) {
    for (int i = 0; i < n; ++i) fn(i);
    return;
  }
  std::atomic<int> next(0);
  std::vector<std::thread> pool;
  for (int t = 0; t < threads; ++t)
    pool.emplace_back([&]() {
      for (;;) {
        int b = next.fetch_add(chunk);
        if (b >= n) break;
        int e = b + chunk < n ? b + chunk : n;
        for (int i = b; i < e; ++i) fn(i);
      }
    });
  for (auto& th : pool) th.join();
}

// vehicle-type constants: World::veh, filled by sb_vehicle() of include/sb_vehicle.h (the CUDA library holds the same record)
constexpr double kSteerJointGain = 0.1;  // POSITION_CONTROL default positionGain: 10 % of the error per sub-step
constexpr double kDynSpeed = 1.0;        // m/s: below it the tyres roll without slip (the slip angles divide by speed)

struct Ego {
  double x, y, h, speed, yaw_rate, v_long, v_lat;
};

// AckermannChassis.control, chassis.py:678-718: brake is dropped below 1/36 m/s.
inline void apply_control(const Ego& v, double thr, double brake, double steer, double* out3) {
  if (brake > 0 && v.v_long < 1.0 / 36.0) brake = 0;
  out3[0] = thr;
  out3[1] = brake;
  out3[2] = steer;
}

// lane_following_controller.py:63-365.  st: [0] lateral_integral_error [1] integral_speed_error
// [2] steering_state [3] throttle_state [4] speed_error [5,6] min_curvature_location [7] valid.
// _update_target_lane_if_reached_end_of_lane (:439-473) only rewrites state.target_lane_id, which
// feeds nothing but itself, so it is not restated.
uint32_t lane_following(const Map& mp, const SbVehicle& vp, const Ego& v, const int32_t* route, int route_len,
                        double target_speed, int lane_change, double dt, double* st, double* out3) {
  static thread_local PathSet ps;
  waypoint_paths(mp, v.x, v.y, v.h, 16, 5.0, route, route_len, &ps);
  uint32_t status = ps.overflow ? SB_STATUS_PATH_OVERFLOW : 0u;
  if (ps.n == 0) {
    apply_control(v, 0, 0, 0, out3);
    return status | SB_STATUS_NO_WAYPOINTS;
  }
  // find_current_lane, :367-374
  int current = 0;
  double bd = 0;
  for (int i = 0; i < ps.n; ++i) {
    double dx = ps.wp[i][0].x - v.x, dy = ps.wp[i][0].y - v.y;
    double d = std::sqrt(dx * dx + dy * dy);
    if (i == 0 || d < bd) {
      current = i;
      bd = d;
    }
  }
  int sel = current + lane_change;
  sel = sel < 0 ? 0 : (sel > ps.n - 1 ? ps.n - 1 : sel);
  const Waypoint* wp = ps.wp[sel];
  const int L = ps.len[sel];

  double ewma = 0.0;
  for (int i = L - 2; i >= 0; --i)
    ewma = lerp(ewma, degrees(std::fabs(wrap_heading(wp[i].heading - wp[i + 1].heading))), 0.03);
  double curv = clipd(ewma / 2.5, 0, 1);
  int npts = L < 10 ? L : 10;
  double tx[10], ty[10], th[10];
  for (int i = 0; i < npts; ++i) {
    tx[i] = wp[i].x;
    ty[i] = wp[i].y;
    th[i] = wp[i].heading;
  }
  double lac = std::fabs(curvature_calculation(tx, ty, th, npts, 4));
  if (lac <= 2) {
    st[5] = wp[4].x;
    st[6] = wp[4].y;
    st[7] = 1.0;
  }
  int la = curv > 0.5 ? 3 : 4;
  if (la > L - 1) la = L - 1;
  double ref_h = wp[0].heading;
  double ddx = wp[la].x - v.x, ddy = wp[la].y - v.y;
  double lad = std::sqrt(ddx * ddx + ddy * ddy);
  double sh, ch;
  sb_sincos(v.h, &sh, &ch);
  double lx = v.x - lad * sh, ly = v.y + lad * ch;

  double raw;
  if (curv < 0.3) {
    raw = -3.6 * 1.8 * (v.speed - target_speed);
  } else if (curv > 0.3 && curv < 0.8) {
    raw = -0.6 * 3.6 * (v.speed - clipd(target_speed, 0, 6.94));
  } else {
    raw = -0.6 * 3.6 * (v.speed - clipd(target_speed, 0, 5.56));
  }
  double speed_error = v.speed - target_speed;
  st[1] += speed_error * dt;
  double damping = (speed_error - st[4]) / dt;
  double lfc = 1.5;
  if (v.speed < 8 || target_speed < 6) lfc = 0;
  raw += (-0.2 * damping - 0.1 * st[1]) + std::fabs(lfc * sb_sin(st[2] * vp.max_steering_wheel));
  st[4] = speed_error;
  if (st[7] != 0.0) {
    double mx = v.x - st[5], my = v.y - st[6];
    if (std::sqrt(mx * mx + my * my) < 2) ref_h = wp[la].heading;
  }
  double Kh, Kl;
  lateral_gains(target_speed, vp, &Kh, &Kl);
  double dvx, dvy;
  radians_to_vec(wp[la].heading, &dvx, &dvy);
  double lat_err = signed_dist_to_line(lx, ly, wp[la].x, wp[la].y, dvx, dvy);
  double Rc = curvature_calculation(tx, ty, th, npts);
  double brake = 0, thr;
  if (raw < 0) {
    brake = clipd(-raw, 0, 1);
    thr = 0;
  } else {
    double tg;
    if (v.speed > 70 / 3.6 && std::fabs(Rc) <= 1e3)
      tg = 4.5;
    else if (40 / 3.6 <= v.speed && v.speed <= 70 / 3.6 && std::fabs(Rc) <= 3)
      tg = 2.5;
    else
      tg = 0.5;
    thr = clipd(raw - tg * 3.6 * std::fabs(v.v_lat), 0, 1);
  }
  st[0] += dt * lat_err;
  double ffg = std::fabs(Rc) < 7 ? 0.45 : 0.15;
  double ff = 1 * ffg * (1 / Rc) * (v.speed * v.speed);
  double ns = clipd(v.speed * 3.6 / 100, 0, 1);
  double hsg = -lerp(0.5, 14, ns);
  double yrg = lerp(5.75, 11.75, ns);
  double lsg = clipd(lerp(-1, 14, ns), 1, 2);
  double max_s = 1;
  if (std::fabs(Rc) > 1e7 && lane_change != 0) {
    hsg = -4.95;
    yrg = 1;
    lsg = 0.22;
    max_s = 0.12;
  }
  double herr = angdiff(sb_pymod(v.h, kTwoPi), ref_h);
  double steer = clipd(-hsg * degrees(Kh) * herr + lsg * Kl * lat_err + yrg * v.yaw_rate + 0.3 * st[0] - ff, -max_s,
                       max_s);
  st[2] = low_pass(steer, st[2], 5.5, dt);
  st[3] = low_pass(thr, st[3], 2, dt, 0);
  apply_control(v, st[3], brake, st[2], out3);
  return status;
}

// trajectory_tracking_controller.py:176-331 with the sedan parameters of
// models/controller_parameters.yaml:3-16.  st: [0] heading_error [1] lateral_error
// [2] velocity_error [3] integral_velocity_error [4] integral_windup_error [5] steering_state
// [6] throttle_state.  traj = x[N], y[N], heading[N], speed[N].
void trajectory_tracking(const SbVehicle& vp, const Ego& v, const double* traj, int N, double dt, double* st, double* out3) {
  const double* tx = traj;
  const double* ty = traj + N;
  const double* th = traj + 2 * N;
  const double* tv = traj + 3 * N;
  const double final_steering_filter_constant = vp.final_steering_filter_constant;
  double throttle_filter_constant = vp.throttle_filter_constant;
  const double velocity_gain = vp.velocity_gain, velocity_integral_gain = vp.velocity_integral_gain,
               traction_gain = vp.traction_gain;
  const double initial_look_ahead_distant = vp.initial_look_ahead_distant, derivative_activation = vp.derivative_activation;
  const bool speed_reduction = vp.speed_reduction_activation != 0;
  const double velocity_damping_gain = vp.velocity_damping_gain, windup_gain = vp.windup_gain;

  double lateral_gain = 0.61, heading_gain = 0.01, led_gain = 0.15, hed_gain = 0.5;
  double ns = clipd((3.6 * tv[0] - 20) / (80 - 20), 0, 1);
  if (v.speed > 70 / 3.6) {
    lateral_gain = 1.51;
    hed_gain = 0.1;
  }
  double steering_filter_constant = lerp(12, final_steering_filter_constant, ns);
  if (std::fabs(angdiff(th[N - 1], th[0])) > 2) throttle_filter_constant = 2.5;
  if (std::fabs(curvature_calculation(tx, ty, th, N, 0, 3)) < 150) {
    heading_gain = 0.05;
    led_gain = 0.015;
    hed_gain = 0.05;
  }
  // calculate_heading_lateral_error, :397-441
  double heading_error = angdiff(sb_pymod(v.h, kTwoPi), th[0]);
  double look_dist = initial_look_ahead_distant;
  int look_pts = (int)initial_look_ahead_distant;
  double ahead_curv = std::fabs(curvature_calculation(tx, ty, th, N, 4));
  if (ahead_curv < 30 && speed_reduction) {
    look_dist = 1;
    look_pts = 1;
  }
  int j = look_pts < N - 1 ? look_pts : N - 1;
  double pvx, pvy;
  radians_to_vec(th[j], &pvx, &pvy);
  double sh, ch;
  sb_sincos(v.h, &sh, &ch);
  double lx = v.x - look_dist * sh, ly = v.y + look_dist * ch;
  double lateral_error = signed_dist_to_line(lx, ly, tx[j], ty[j], pvx, pvy);

  double Rc = curvature_calculation(tx, ty, th, N);
  double derivative_term = hed_gain * v.yaw_rate + led_gain * (lateral_error - st[1]) / dt;
  double ff = 0.1 * (1 / Rc) * (v.speed * v.speed);
  double steering_raw = clipd(
      derivative_activation * derivative_term + degrees(heading_gain * heading_error) + 1 * lateral_gain * lateral_error - ff,
      -1, 1);
  st[5] = low_pass(steering_raw, st[5], steering_filter_constant, dt);

  // calculate_raw_throttle_feedback, :333-395
  double desired_speed = tv[N - 1];
  if (ahead_curv < 30 && speed_reduction)
    desired_speed = clipd(0.8 * desired_speed, 0, 8.3);
  else if (ahead_curv < 100 && speed_reduction)
    desired_speed *= 0.8;
  double velocity_error = v.speed - desired_speed;
  double ved = (velocity_error - st[2]) / dt;
  double raw_throttle = 3.6 * (-0.5 * velocity_gain * velocity_error -
                               velocity_integral_gain * (st[3] + windup_gain * st[4]) - velocity_damping_gain * ved);
  st[2] = velocity_error;
  st[4] = clipd(raw_throttle, -1, 1) - raw_throttle;
  st[6] = low_pass(raw_throttle, st[6], throttle_filter_constant, dt, -1, -traction_gain * std::fabs(v.v_lat));
  double thr, brake;
  if (st[6] > 0) {
    brake = 0;
    thr = clipd(st[6], 0, 1);
  } else {
    brake = clipd(-st[6], 0, 1);
    thr = 0;
  }
  st[0] = heading_error;
  st[1] = lateral_error;
  st[3] += (v.speed - desired_speed) * dt;
  apply_control(v, thr, brake, st[5], out3);
}

// Stand-in for the pybullet AckermannChassis step (DESIGN.md "Vehicle model"): a single-track
// ("bicycle") model with yaw inertia and linear, friction-limited tyres above kDynSpeed, the kinematic bicycle
// below it, and a first-order steering joint; `substeps` explicit Euler sub-steps (240 Hz, smarts.py:607-615).
// Body frame: x forward, y left, yaw counter-clockwise; delta = steering joint angle, left positive.
void chassis_step(World& w, int ai, int vi) {
  const SbVehicle& vp = w.veh;
  const double kMaxSteeringWheel = vp.max_steering_wheel, kSteerLimit = vp.steer_limit, kMaxTorque = vp.max_torque,
               kMaxBrakeTorque = vp.max_btorque, kWheelRadius = vp.wheel_radius, kEffMass = vp.eff_mass,
               kRearAxleToCg = vp.axle_to_cg, kCornerStiffness = vp.corner_stiffness, kTyreForceMax = vp.tyre_force_max,
               kMass = vp.mass, kIzz = vp.izz;
  const double* c = &w.last_control[(size_t)ai * 3];
  double target = -c[2] * kMaxSteeringWheel;  // chassis.py:790-793 (joint angle, left positive)
  target = clipd(target, -kSteerLimit, kSteerLimit);
  const double ax = 4.0 * (c[0] * kMaxTorque - c[1] * kMaxBrakeTorque) / kWheelRadius / kEffMass;
  const int n = w.substeps;
  const double hdt = w.cfg.dt / (double)n;
  double x = w.pos_x[vi], y = w.pos_y[vi], h = w.heading[vi];
  double vx = w.v_long[ai], vy = -w.v_lat[ai], om = w.yaw_rate[vi], delta = -w.steering[ai];
  for (int k = 0; k < n; ++k) {
    delta = delta + kSteerJointGain * (target - delta);
    double sd, cd;
    sb_sincos(delta, &sd, &cd);
    if (vx > kDynSpeed) {
      const double af = (vy + kRearAxleToCg * om) / vx - delta, ar = (vy - kRearAxleToCg * om) / vx;
      const double fyf = clipd(-kCornerStiffness * af, -kTyreForceMax, kTyreForceMax) * cd;
      const double fyr = clipd(-kCornerStiffness * ar, -kTyreForceMax, kTyreForceMax);
      const double ay = (fyf + fyr) / kMass - vx * om;
      const double dom = (kRearAxleToCg * fyf - kRearAxleToCg * fyr) / kIzz;
      vy = vy + ay * hdt;
      om = om + dom * hdt;
    } else {  // rolling without slip
      om = vx * (sd / cd) / (2.0 * kRearAxleToCg);
      vy = om * kRearAxleToCg;
    }
    vx = vx + (ax + vy * om) * hdt;
    if (vx < 0) vx = 0;
    double sn, cs;
    sb_sincos(h, &sn, &cs);
    x = x + (-(vx * sn) - vy * cs) * hdt;
    y = y + (vx * cs - vy * sn) * hdt;
    h = h + om * hdt;
  }
  h = wrap_heading(h);
  w.pos_x[vi] = x;
  w.pos_y[vi] = y;
  w.heading[vi] = h;
  w.speed[vi] = std::sqrt(vx * vx + vy * vy);
  w.yaw_rate[vi] = om;
  sb_sincos(h, &w.sin_h[vi], &w.cos_h[vi]);
  w.steering[ai] = -delta;  // chassis.py:513-528
  w.v_long[ai] = vx;        // chassis.py:558-566
  w.v_lat[ai] = -vy;
}

// ---------------------------------------------------------------- sensors
// smarts.py:1191-1208 + sensors.py:241-266
void observe_neighbors(World& w, int e, int a) {
  const int V = w.V, cap = w.cfg.nbr_cap;
  const int ai = e * w.A + a, vi = e * V + a;
  int cnt = 0;
  for (int k = 0; k < cap; ++k) {
    size_t o = (size_t)ai * cap + k;
    w.nbr_id[o] = -1;
    w.nbr_lane[o] = -1;
    w.nbr_lane_index[o] = 0;
    w.nbr_heading[o] = 0;
    w.nbr_speed[o] = 0;
    for (int c = 0; c < 3; ++c) {
      w.nbr_pos[o * 3 + c] = 0;
      w.nbr_box[o * 3 + c] = 0;
    }
  }
  for (int j = 0; j < V; ++j) {
    if (j == a) continue;
    int vj = e * V + j;
    if (!w.alive[vj]) continue;
    if (w.cfg.nbr_radius >= 0) {
      double dx = w.pos_x[vj] - w.pos_x[vi], dy = w.pos_y[vj] - w.pos_y[vi], dz = w.pos_z[vj] - w.pos_z[vi];
      double d = std::sqrt((dx * dx + dy * dy) + dz * dz);
      if (!(d <= w.cfg.nbr_radius)) continue;
    }
    if (cnt < cap) {
      size_t o = (size_t)ai * cap + cnt;
      w.nbr_id[o] = j;
      w.nbr_pos[o * 3 + 0] = w.pos_x[vj];
      w.nbr_pos[o * 3 + 1] = w.pos_y[vj];
      w.nbr_pos[o * 3 + 2] = w.pos_z[vj];
      w.nbr_box[o * 3 + 0] = (float)w.length[vj];
      w.nbr_box[o * 3 + 1] = (float)w.width[vj];
      w.nbr_box[o * 3 + 2] = (float)w.height[vj];
      w.nbr_heading[o] = (float)w.heading[vj];
      w.nbr_speed[o] = (float)w.speed[vj];
      int lane = nearest_lane(w.map, w.pos_x[vj], w.pos_y[vj], w.length[vi]);
      w.nbr_lane[o] = lane;
      w.nbr_lane_index[o] = lane >= 0 ? (int8_t)w.map.lane_idx_in_road[lane] : 0;
    }
    ++cnt;
  }
  w.nbr_count[ai] = cnt;
}

// waypoint_paths[0][0] feeds TripMeterSensor.append_waypoint_if_new (sensors.py:340-341)
void stash_trip_waypoint(World& w, int ai, const PathSet& ps, int slot = 0) {
  double* tw = &w.trip_wp[(size_t)ai * 8 + slot * 4];
  if (ps.n == 0 || ps.len[0] == 0) {
    tw[3] = -1.0;
    return;
  }
  const Waypoint& q = ps.wp[0][0];
  tw[0] = q.x;
  tw[1] = q.y;
  tw[2] = q.heading;
  tw[3] = (double)q.lane;
}

// sensors.py:972-988 -> waypoint_paths; StdObs layout format_obs.py:546-604
void observe_waypoints(World& w, int e, int a) {
  static thread_local PathSet ps;
  const int ai = e * w.A + a, vi = e * w.V + a;
  const int P = w.cfg.wp_out_paths, L = w.cfg.wp_out_len;
  waypoint_paths(w.map, w.pos_x[vi], w.pos_y[vi], w.heading[vi], w.cfg.wp_lookahead, 5.0,
                 &w.route[(size_t)ai * w.cfg.route_cap], w.route_len[ai], &ps);
  if (ps.overflow) w.status[ai] |= SB_STATUS_PATH_OVERFLOW;
  w.wp_n_paths[ai] = ps.total;
  if (w.cfg.events_enable) stash_trip_waypoint(w, ai, ps);
  for (int p = 0; p < P; ++p) {
    int len = p < ps.n ? std::min(ps.len[p], L) : 0;
    w.wp_len[(size_t)ai * P + p] = len;
    for (int i = 0; i < L; ++i) {
      size_t o = ((size_t)ai * P + p) * L + i;
      if (i < len) {
        const Waypoint& q = ps.wp[p][i];
        w.wp_pos[o * 3 + 0] = q.x;
        w.wp_pos[o * 3 + 1] = q.y;
        w.wp_pos[o * 3 + 2] = 0.0;
        w.wp_heading[o] = (float)q.heading;
        w.wp_lane_width[o] = (float)q.lane_width;
        w.wp_speed_limit[o] = (float)q.speed_limit;
        w.wp_lane[o] = q.lane;
        w.wp_lane_index[o] = (int8_t)q.lane_index;
      } else {
        w.wp_pos[o * 3 + 0] = w.wp_pos[o * 3 + 1] = w.wp_pos[o * 3 + 2] = 0.0;
        w.wp_heading[o] = w.wp_lane_width[o] = w.wp_speed_limit[o] = 0.f;
        w.wp_lane[o] = -1;
        w.wp_lane_index[o] = 0;
      }
    }
  }
}

// lidar.py:74-134 + sensors.py:797-827.  Hittable set: ground plane z = 0
// (plane.urdf), every other alive vehicle's box (agents z in [0.1, 1.1] above their base,
// vehicle.urdf:12-16; social boxes z in [0.2, 0.2 + height], utils/bullet.py:169-177) and,
// when lidar_wall_height > 0, vertical walls on the road boundary segments.
//
// Hit contract (DESIGN.md "LiDAR"): the ray table is factored into azimuth groups by
// include/sb_lidar_table.h; a ray is  origin + s (ux, uy, slope),  0 <= s <= hlen.  Against a box the
// horizontal slab interval [s0, s1] is found in the box frame with one reciprocal per axis, the
// height slab with the ray's inv_slope; the hit is the entry of the intersection of the two.  A wall
// is the degenerate interval s0 = s1 at the 2-D ray / segment intersection with heights [0, wall].
namespace {
inline void lidar_entry(double s0, double s1, double zlo, double zhi, double oz, const SbRayRec& q, double* sbest) {
  double e0 = s0, e1 = s1;
  if (q.dz == 0.0) {
    if (oz < zlo || oz > zhi) return;
  } else {
    double za = (zlo - oz) * q.inv_slope, zb = (zhi - oz) * q.inv_slope;
    if (za > zb) std::swap(za, zb);
    if (za > e0) e0 = za;
    if (zb < e1) e1 = zb;
  }
  if (e0 > e1 || e0 < 0.0 || e0 > q.hlen) return;  // origin inside the box: no hit
  if (e0 < *sbest) *sbest = e0;
}
}  // namespace

void observe_lidar(World& w, int e, int a) {
  const int R = w.cfg.lidar_rays, V = w.V;
  const int ai = e * w.A + a, vi = e * V + a;
  const double ox = w.pos_x[vi], oy = w.pos_y[vi], oz = w.pos_z[vi] + 1.0;
  const double inf = std::numeric_limits<double>::infinity();
  const double miss = w.cfg.lidar_miss_inf ? inf : 0.0;
  for (int r = 0; r < R; ++r) {
    const SbRayRec& q = w.ray_rec[r];
    const size_t o = (size_t)ai * R + r;
    if (q.group < 0) {  // straight up / down: the ground plane only
      double t = 2.0;
      if (q.dz < 0.0) {
        t = (0.0 - oz) / q.dz;
        if (!(t >= 0.0 && t <= 1.0)) t = 2.0;
      }
      if (t <= 1.0) {
        w.lidar_hits[o] = 1;
        w.lidar_points[o * 3 + 0] = ox;
        w.lidar_points[o * 3 + 1] = oy;
        w.lidar_points[o * 3 + 2] = oz + t * q.dz;
      } else {
        w.lidar_hits[o] = 0;
        w.lidar_points[o * 3 + 0] = w.lidar_points[o * 3 + 1] = w.lidar_points[o * 3 + 2] = miss;
      }
      continue;
    }
    const double ux = w.ray_group_u[2 * q.group], uy = w.ray_group_u[2 * q.group + 1];
    double sbest = inf;
    if (q.dz < 0.0) {
      const double sg = (0.0 - oz) * q.inv_slope;
      if (sg >= 0.0 && sg <= q.hlen) sbest = sg;
    }
    for (int j = 0; j < V; ++j) {
      if (j == a) continue;
      int vj = e * V + j;
      if (!w.alive[vj]) continue;
      double c = w.cos_h[vj], s = w.sin_h[vj];
      double rx = ox - w.pos_x[vj], ry = oy - w.pos_y[vj];
      // box frame: axis 0 = right (width), axis 1 = forward (length)
      double lo_[2] = {rx * c + ry * s, ry * c - rx * s};
      double ld_[2] = {ux * c + uy * s, uy * c - ux * s};
      double hw = 0.5 * w.width[vj], hl = 0.5 * w.length[vj];
      double mn[2] = {-hw, -hl}, mx[2] = {hw, hl};
      double s0 = -inf, s1 = inf;
      bool ok = true;
      for (int k = 0; k < 2 && ok; ++k) {
        if (ld_[k] == 0.0) {
          if (lo_[k] < mn[k] || lo_[k] > mx[k]) ok = false;
        } else {
          double inv = 1.0 / ld_[k];
          double sa = (mn[k] - lo_[k]) * inv, sb = (mx[k] - lo_[k]) * inv;
          if (sa > sb) std::swap(sa, sb);
          if (sa > s0) s0 = sa;
          if (sb < s1) s1 = sb;
        }
      }
      if (!ok || s0 > s1) continue;
      double z0 = w.pos_z[vj] + (j < w.A ? w.veh.box_z0 : 0.2);
      lidar_entry(s0, s1, z0, z0 + w.height[vj], oz, q, &sbest);
    }
    if (w.cfg.lidar_wall_height > 0.0) {
      const int NE = w.map.m.n_edges;
      for (int k = 0; k < NE; ++k) {
        const double* ed = &w.map.road_edges[(size_t)k * 4];
        double ex = ed[2] - ed[0], ey = ed[3] - ed[1];
        double den = ux * ey - uy * ex;
        if (den == 0.0) continue;
        double qx = ed[0] - ox, qy = ed[1] - oy;
        double inv = 1.0 / den;
        double sw = (qx * ey - qy * ex) * inv;
        double u = (qx * uy - qy * ux) * inv;
        if (sw < 0.0 || u < 0.0 || u > 1.0) continue;
        lidar_entry(sw, sw, 0.0, w.cfg.lidar_wall_height, oz, q, &sbest);
      }
    }
    if (sbest <= q.hlen) {
      w.lidar_hits[o] = 1;
      w.lidar_points[o * 3 + 0] = ox + sbest * ux;
      w.lidar_points[o * 3 + 1] = oy + sbest * uy;
      w.lidar_points[o * 3 + 2] = oz + sbest * q.slope;
    } else {
      w.lidar_hits[o] = 0;
      w.lidar_points[o * 3 + 0] = w.lidar_points[o * 3 + 1] = w.lidar_points[o * 3 + 2] = miss;
    }
  }
}

// Raster contract (DESIGN.md "Grid maps"): camera = renderer.py:325-336, :383-386 (orthographic,
// film W*res x H*res centred on the ego, rotated by the ego heading, rows flipped so row 0 is
// ahead: sensors.py:706, :748).  World vertices go to pixel space in binary64, are snapped to
// 1/256 px (round-half-even), and a pixel is covered iff its centre satisfies every edge function
// of the convex polygon (inclusive), evaluated in exact integer arithmetic.
struct Cam {
  double px, py, c, s, res, halfw, halfh;
};
inline void to_pixel(const Cam& cam, double x, double y, int64_t* U, int64_t* Vv) {
  double dx = x - cam.px, dy = y - cam.py;
  double xr = dx * cam.c + dy * cam.s;
  double yr = dy * cam.c - dx * cam.s;
  double u = xr / cam.res + cam.halfw;
  double v = cam.halfh - yr / cam.res;
  double su = rint(u * 256.0), sv = rint(v * 256.0);
  const double lim = 16777216.0;  // 2^24 units = 65536 px
  su = su > lim ? lim : (su < -lim ? -lim : su);
  sv = sv > lim ? lim : (sv < -lim ? -lim : sv);
  *U = (int64_t)su;
  *Vv = (int64_t)sv;
}

void fill_convex(uint8_t* img, int W, int H, const int64_t* U, const int64_t* Vv, int n) {
  int64_t area2 = 0;
  for (int i = 0; i < n; ++i) {
    int j = (i + 1) % n;
    area2 += U[i] * Vv[j] - U[j] * Vv[i];
  }
  if (area2 == 0) return;
  int64_t sg = area2 > 0 ? 1 : -1;
  int64_t umin = U[0], umax = U[0], vmin = Vv[0], vmax = Vv[0];
  for (int i = 1; i < n; ++i) {
    umin = std::min(umin, U[i]); umax = std::max(umax, U[i]);
    vmin = std::min(vmin, Vv[i]); vmax = std::max(vmax, Vv[i]);
  }
  // pixel centre c*256+128 within [min,max]
  auto lo = [](int64_t m) { int64_t t = m - 128; return t <= 0 ? (int64_t)0 : (t + 255) / 256; };
  auto hi = [](int64_t m) { int64_t t = m - 128; return t < 0 ? (int64_t)-1 : t / 256; };
  int64_t c0 = lo(umin), c1 = std::min<int64_t>(hi(umax), W - 1);
  int64_t r0 = lo(vmin), r1 = std::min<int64_t>(hi(vmax), H - 1);
  for (int64_t r = r0; r <= r1; ++r) {
    int64_t py = r * 256 + 128;
    for (int64_t c = c0; c <= c1; ++c) {
      int64_t px = c * 256 + 128;
      bool in = true;
      for (int i = 0; i < n && in; ++i) {
        int j = (i + 1) % n;
        int64_t ev = (U[j] - U[i]) * (py - Vv[i]) - (Vv[j] - Vv[i]) * (px - U[i]);
        if (sg * ev < 0) in = false;
      }
      if (in) img[r * W + c] = 255;
    }
  }
}

// sensors.py:719-758 (vehicles only, ego included: test_observations.py:179-187)
void observe_ogm(World& w, int e, int a) {
  const int W = w.cfg.ogm_width, H = w.cfg.ogm_height, V = w.V;
  const int ai = e * w.A + a, vi = e * V + a;
  uint8_t* img = &w.ogm[(size_t)ai * W * H];
  std::memset(img, 0, (size_t)W * H);
  Cam cam{w.pos_x[vi], w.pos_y[vi], w.cos_h[vi], w.sin_h[vi], w.cfg.ogm_resolution, 0.5 * W, 0.5 * H};
  for (int j = 0; j < V; ++j) {
    int vj = e * V + j;
    if (!w.alive[vj]) continue;
    double c = w.cos_h[vj], s = w.sin_h[vj];
    double hw = 0.5 * w.width[vj], hl = 0.5 * w.length[vj];
    const double sa[4] = {-hw, hw, hw, -hw}, sb_[4] = {-hl, -hl, hl, hl};
    int64_t U[4], Vv[4];
    for (int k = 0; k < 4; ++k) {
      double x = w.pos_x[vj] + (sa[k] * c - sb_[k] * s);
      double y = w.pos_y[vj] + (sa[k] * s + sb_[k] * c);
      to_pixel(cam, x, y, &U[k], &Vv[k]);
    }
    fill_convex(img, W, H, U, Vv, 4);
  }
}

// sensors.py:675-716 (road surface only)
void observe_dagm(World& w, int e, int a) {
  const int W = w.cfg.dagm_width, H = w.cfg.dagm_height;
  const int ai = e * w.A + a, vi = e * w.V + a;
  uint8_t* img = &w.dagm[(size_t)ai * W * H];
  std::memset(img, 0, (size_t)W * H);
  Cam cam{w.pos_x[vi], w.pos_y[vi], w.cos_h[vi], w.sin_h[vi], w.cfg.dagm_resolution, 0.5 * W, 0.5 * H};
  const Map& mp = w.map;
  for (int p = 0; p < mp.m.n_polys; ++p) {
    int s = mp.poly_start[p], n = mp.poly_start[p + 1] - s;
    if (n < 3 || n > 8) continue;
    int64_t U[8], Vv[8];
    for (int k = 0; k < n; ++k) to_pixel(cam, mp.poly_x[s + k], mp.poly_y[s + k], &U[k], &Vv[k]);
    fill_convex(img, W, H, U, Vv, n);
  }
}

// ---------------------------------------------------------------- events
// sumolib 1.10.0 geomhelper restatements (package not vendored in the reference, setup.py:28)
inline double line_offset(double px, double py, double ax, double ay, double bx, double by) {
  double dx = ax - bx, dy = ay - by;
  double d = std::sqrt(dx * dx + dy * dy);
  double u = ((px - ax) * (bx - ax)) + ((py - ay) * (by - ay));
  if (d == 0.0 || u < 0.0 || u > d * d) return u < 0.0 ? 0.0 : d;
  return u / d;
}
inline bool geo_isclose(double a, double b) { return std::fabs(a - b) <= 1e-9 * std::max(std::fabs(a), std::fabs(b)); }
inline void position_at_offset(double ax, double ay, double bx, double by, double off, double* x, double* y) {
  if (geo_isclose(off, 0.0)) { *x = ax; *y = ay; return; }
  double dx = ax - bx, dy = ay - by;
  double dist = std::sqrt(dx * dx + dy * dy);
  if (geo_isclose(dist, off)) { *x = bx; *y = by; return; }
  *x = ax + (bx - ax) * (off / dist);
  *y = ay + (by - ay) * (off / dist);
}
// Lane.offset_along_lane, sumo_road_network.py:476-491
double offset_along_lane(const Map& mp, int lane, double px, double py) {
  int s = mp.lane_shape_start[lane], e = mp.lane_shape_start[lane + 1];
  double seen = 0.0, min_dist = 1e300, min_off = -1.0;
  for (int i = s; i + 1 < e; ++i) {
    double ax = mp.shape_x[i], ay = mp.shape_y[i], bx = mp.shape_x[i + 1], by = mp.shape_y[i + 1];
    double pos = line_offset(px, py, ax, ay, bx, by);
    double qx, qy;
    position_at_offset(ax, ay, bx, by, pos, &qx, &qy);
    double ex = px - qx, ey = py - qy;
    double dist = std::sqrt(ex * ex + ey * ey);
    if (dist < min_dist) {
      min_dist = dist;
      min_off = pos + seen;
    }
    double dx = ax - bx, dy = ay - by;
    seen += std::sqrt(dx * dx + dy * dy);
  }
  return min_off;
}
// Lane.from_lane_coord, sumo_road_network.py:502-506
void from_lane_coord(const Map& mp, int lane, double off, double* x, double* y) {
  int s = mp.lane_shape_start[lane], e = mp.lane_shape_start[lane + 1];
  double seen = 0.0;
  for (int i = s; i + 1 < e; ++i) {
    double ax = mp.shape_x[i], ay = mp.shape_y[i], bx = mp.shape_x[i + 1], by = mp.shape_y[i + 1];
    double dx = ax - bx, dy = ay - by;
    double len = std::sqrt(dx * dx + dy * dy);
    if (seen + len > off) {
      position_at_offset(ax, ay, bx, by, off - seen, x, y);
      return;
    }
    seen += len;
  }
  *x = mp.shape_x[e - 1];
  *y = mp.shape_y[e - 1];
}
// RoadMap.Lane.vector_at_offset, road_map.py:377-388
void vector_at_offset(const Map& mp, int lane, double off, double* vx, double* vy) {
  double len = mp.lane_length[lane], s0, s1;
  if (off >= len) {
    s0 = len - 1;
    s1 = len;
  } else {
    s0 = off;
    s1 = off + 1;
  }
  s0 = s0 > 0 ? s0 : 0;
  double x0, y0, x1, y1;
  from_lane_coord(mp, lane, s0, &x0, &y0);
  from_lane_coord(mp, lane, s1, &x1, &y1);
  *vx = x1 - x0;
  *vy = y1 - y0;
}
// utils/math.py:256-277
double vec_to_radians(double x, double y) {
  double r = sb_atan2(std::fabs(y), std::fabs(x));
  if (x < 0) {
    if (y < 0) return sb_pymod(r + 0.5 * kPi, kTwoPi);
    return sb_pymod(0.5 * kPi - r, kTwoPi);
  } else if (y < 0) {
    return sb_pymod(1.5 * kPi - r, kTwoPi);
  }
  return sb_pymod(r - 0.5 * kPi, kTwoPi);
}
// sumo_road_network.py:703-709
bool road_with_point(const Map& mp, double px, double py) {
  const double radius = std::max(5.0, 2 * 3.2);
  for (int l = 0; l < mp.m.n_lanes; ++l) {
    double d = dist_point_to_polyline(mp, l, px, py);
    if (d < 0.0 || !(d < radius)) continue;
    if (d < 0.5 * mp.lane_width[l] + 1e-1) return true;
  }
  return false;
}
// vehicle-vehicle contact: smarts.py:1270-1291 with the 0.05 m leeway of chassis.py:76-79,
// as a 2-D separating-axis test on the heading-aligned boxes
bool boxes_touch(double ax, double ay, double ac, double as, double al, double aw, double bx, double by, double bc,
                 double bs, double bl, double bw) {
  const double afx = -as, afy = ac, asx = ac, asy = as;
  const double bfx = -bs, bfy = bc, bsx = bc, bsy = bs;
  const double dx = bx - ax, dy = by - ay;
  const double hal = 0.5 * al, haw = 0.5 * aw, hbl = 0.5 * bl, hbw = 0.5 * bw;
  const double leeway = 0.05;
  const double ff = std::fabs(afx * bfx + afy * bfy), fs = std::fabs(afx * bsx + afy * bsy);
  const double sf = std::fabs(asx * bfx + asy * bfy), ss = std::fabs(asx * bsx + asy * bsy);
  if (std::fabs(dx * afx + dy * afy) > (hal + (hbl * ff + hbw * fs)) + leeway) return false;
  if (std::fabs(dx * asx + dy * asy) > (haw + (hbl * sf + hbw * ss)) + leeway) return false;
  if (std::fabs(dx * bfx + dy * bfy) > (hbl + (hal * ff + haw * sf)) + leeway) return false;
  if (std::fabs(dx * bsx + dy * bsy) > (hbw + (hal * fs + haw * ss)) + leeway) return false;
  return true;
}

// Sensors.observe tail: trip meter, driven path, _is_done_with_events (sensors.py:340-396, :443-594)
void observe_events(World& w, int e, int a) {
  const Map& mp = w.map;
  const int ai = e * w.A + a, vi = e * w.V + a;
  const bool ticked = w.ticked[ai] != 0;  // per agent: observing a subset of the envs leaves the others' tick pending
  w.ticked[ai] = 0;
  const double px = w.pos_x[vi], py = w.pos_y[vi], ph = w.heading[vi];
  const double len = w.length[vi], wid = w.width[vi], ch = w.cos_h[vi], sh = w.sin_h[vi];
  const int32_t* route = &w.route[(size_t)ai * w.cfg.route_cap];
  const int route_len = w.route_len[ai];
  uint32_t bits = 0;
  const int steps = w.steps[ai] + (ticked ? 1 : 0);
  w.steps[ai] = steps;
  w.ev_steps[ai] = steps;
  {  // TripMeterSensor, sensors.py:880-947
    const double* tw = &w.trip_wp[(size_t)ai * 8];
    const int tlane = (int)tw[3];
    double dist = w.trip_dist[ai], last = w.trip_last[ai];
    if (!w.trip_init[ai]) {  // TripMeterSensor.__init__ (sensors.py:885-898): its own unrouted query
      w.trip_init[ai] = 1;
      const double* cw = w.cfg.wp_enable ? tw + 4 : tw;
      if ((int)cw[3] >= 0) {
        w.trip_has[ai] = 1;
        w.trip_x[ai] = cw[0];
        w.trip_y[ai] = cw[1];
        w.trip_h[ai] = cw[2];
      }
    }
    if (tlane >= 0) {
      last = dist;
      const bool fixed_route = w.goal_r[ai] >= 0.0;
      const bool count = !fixed_route || in_route(route, route_len, mp.lane_road[tlane]);
      if (!w.trip_has[ai]) {
        if (count) {
          w.trip_has[ai] = 1;
          w.trip_x[ai] = tw[0];
          w.trip_y[ai] = tw[1];
          w.trip_h[ai] = tw[2];
        }
      } else {
        double ddx = tw[0] - w.trip_x[ai], ddy = tw[1] - w.trip_y[ai];
        double d = std::sqrt(ddx * ddx + ddy * ddy);
        if (d > 0.5 && count) {
          double hvx, hvy;
          radians_to_vec(w.trip_h[ai], &hvx, &hvy);
          double dot = hvx * ddx + hvy * ddy;
          double sg = dot > 0 ? 1.0 : (dot < 0 ? -1.0 : 0.0);
          dist += sg * d;
          w.trip_x[ai] = tw[0];
          w.trip_y[ai] = tw[1];
          w.trip_h[ai] = tw[2];
        }
      }
    }
    w.trip_dist[ai] = dist;
    w.trip_last[ai] = last;
    w.ev_distance[ai] = dist;
    w.ev_reward[ai] = dist - last;
  }
  {  // DrivenPathSensor + _vehicle_is_not_moving, sensors.py:830-877, :509-524
    const int n = w.path_n[ai];
    double* ring = &w.path_cum[(size_t)ai * w.ring];
    double cum = 0.0;
    if (n > 0) {
      double qx = px - w.last_px[ai], qy = py - w.last_py[ai];
      cum = ring[(n - 1) % w.ring] + std::sqrt(qx * qx + qy * qy);
    }
    ring[n % w.ring] = cum;
    w.path_n[ai] = n + 1;
    w.last_px[ai] = px;
    w.last_py[ai] = py;
    // _vehicle_is_not_moving (sensors.py:509-524) over the entries whose timestamp >= now - not_moving_time
    // (sensors.py:866-871); entry j was recorded at clock value sb_clock(j): the comparison is made on those
    // binary64 values exactly as the reference makes it (0.1-second clocks do not subtract exactly)
    const double now = sb_clock(steps, w.dt_num, w.dt_den);
    if (!(now < w.cfg.not_moving_time)) {
      const double threshold = now - w.cfg.not_moving_time;
      int back = std::min(std::min(w.ring - 1, n), steps);
      while (back > 0 && !(sb_clock(steps - back, w.dt_num, w.dt_den) >= threshold)) --back;
      double moved = cum - ring[(n - back) % w.ring];
      if (moved < w.cfg.not_moving_distance) bits |= SB_EV_NOT_MOVING;
    }
  }
  if (w.goal_r[ai] >= 0.0) {  // plan.py:117-121
    double gx = px - w.goal_x[ai], gy = py - w.goal_y[ai];
    if (gx * gx + gy * gy <= w.goal_r[ai] * w.goal_r[ai]) bits |= SB_EV_REACHED_GOAL;
  }
  if (w.cfg.max_episode_steps > 0 && steps >= w.cfg.max_episode_steps) bits |= SB_EV_REACHED_MAX_STEPS;
  {  // contacts
    int cnt = 0, first = -1;
    for (int j = 0; j < w.V; ++j) {
      int vj = e * w.V + j;
      if (j == a || !w.alive[vj]) continue;
      if (boxes_touch(px, py, ch, sh, len, wid, w.pos_x[vj], w.pos_y[vj], w.cos_h[vj], w.sin_h[vj], w.length[vj],
                      w.width[vj])) {
        if (first < 0) first = j;
        ++cnt;
      }
    }
    w.ev_n_collisions[ai] = cnt;
    w.ev_collidee[ai] = first;
    if (cnt) bits |= SB_EV_COLLISION;
  }
  {  // off road / on shoulder: sensors.py:496-507; corners vehicle.py:316-332, utils/math.py:436-444
    if (!road_with_point(mp, px, py)) bits |= SB_EV_OFF_ROAD;
    const double cw[4] = {-0.5, 0.5, 0.5, -0.5}, cl[4] = {0.5, 0.5, -0.5, -0.5};
    bool shoulder = false;
    for (int k = 0; k < 4; ++k) {
      double ox = (px + cw[k] * wid) - px, oy = (py + cl[k] * len) - py;
      double qx = (px + ch * ox) + sh * oy;
      double qy = (py + (-sh) * ox) + ch * oy;
      if (!road_with_point(mp, qx, qy)) shoulder = true;
    }
    if (shoulder) bits |= SB_EV_ON_SHOULDER;
  }
  {  // sensors.py:526-594
    double radius = std::sqrt(len * len + wid * wid) * 0.5 + 5;
    int nl = nearest_lane(mp, px, py, radius);
    if (nl < 0) {
      bits |= SB_EV_OFF_ROUTE;
    } else {
      int road = mp.lane_road[nl];
      bool in_j = mp.road_is_junction[road] != 0;
      double off = offset_along_lane(mp, nl, px, py), mvx, mvy;
      vector_at_offset(mp, nl, off, &mvx, &mvy);
      if (!in_j) {
        double th = wrap_heading(vec_to_radians(mvx, mvy));
        double rel = wrap_heading(ph - th);
        if (std::fabs(rel) > 0.5 * kPi) bits |= SB_EV_WRONG_WAY;
      }
      if (!(route_len <= 0 || in_route(route, route_len, road) || in_j)) {
        // oncoming_lanes_at_offset, sumo_road_network.py:371-396
        double cx, cy;
        from_lane_coord(mp, nl, off, &cx, &cy);
        double near_r = 1.1 * mp.lane_width[nl];
        double my_norm = std::sqrt(mvx * mvx + mvy * mvy);
        bool found = false;
        if (my_norm != 0.0) {
          for (int l = 0; l < mp.m.n_lanes && !found; ++l) {
            if (l == nl) continue;
            double d = dist_point_to_polyline(mp, l, cx, cy);
            if (d < 0.0 || !(d < near_r)) continue;
            double sl = offset_along_lane(mp, l, cx, cy), lvx, lvy;
            vector_at_offset(mp, l, sl, &lvx, &lvy);
            double lv_norm = std::sqrt(lvx * lvx + lvy * lvy);
            if (lv_norm == 0.0) continue;
            double ang = (mvx * lvx + mvy * lvy) / (my_norm * lv_norm);
            if (ang < -0.995562 && in_route(route, route_len, mp.lane_road[l])) found = true;
          }
        }
        if (!found) bits |= SB_EV_OFF_ROUTE;
      }
    }
  }
  w.ev_events[ai] = bits;
  const uint32_t mask = w.cfg.done_mask | SB_EV_REACHED_GOAL | SB_EV_REACHED_MAX_STEPS;
  w.ev_done[ai] = (ticked && (bits & mask)) ? 1 : 0;
}

// lane_ttc, smarts/env/custom_observations.py:148-290 (the `ttc` member of StdObs, format_obs.py:546-556),
// on the agent's own observation buffers.
void observe_ttc(World& w, int e, int a) {
  const int ai = e * w.A + a, vi = e * w.V + a;
  const int P = w.cfg.wp_out_paths, L = w.cfg.wp_out_len, cap = w.cfg.nbr_cap;
  const double ex = w.pos_x[vi], ey = w.pos_y[vi], eh = w.heading[vi], espeed = w.speed[vi];
  const int total_paths = w.wp_n_paths[ai];
  const int n = std::min(std::min(total_paths, P), 32);
  const int cnt = w.nbr_count[ai], nv = std::min(cnt, cap);
  bool truncated = total_paths > P || cnt > cap;
  auto row = [&](int p) { return ((size_t)ai * P + p) * L; };
  auto plen = [&](int p) { return w.wp_len[(size_t)ai * P + p]; };
  bool any_path = false;
  for (int p = 0; p < n; ++p) {
    if (plen(p) > 0) any_path = true;
    if (plen(p) == L && L < w.cfg.wp_lookahead + 1) truncated = true;
  }
  w.ttc_valid[ai] = 0;
  w.ttc_angle_error[ai] = w.ttc_dist_center[ai] = 0.f;
  for (int k = 0; k < 3; ++k) w.ttc_dtc[(size_t)ai * 3 + k] = w.ttc_ttc[(size_t)ai * 3 + k] = 0.f;
  if (cnt == 0 || n == 0 || !any_path) return;
  // closest first waypoint (first minimum wins)
  int p0 = -1;
  double d0 = 0.0;
  for (int p = 0; p < n; ++p) {
    if (plen(p) <= 0) continue;
    double dx = w.wp_pos[row(p) * 3 + 0] - ex, dy = w.wp_pos[row(p) * 3 + 1] - ey;
    double d = std::sqrt(dx * dx + dy * dy);
    if (p0 < 0 || d < d0) {
      p0 = p;
      d0 = d;
    }
  }
  const size_t crow = row(p0);
  const double cwx = w.wp_pos[crow * 3 + 0], cwy = w.wp_pos[crow * 3 + 1], cwh = (double)w.wp_heading[crow];
  const int li = (int)w.wp_lane_index[crow];
  double hvx, hvy;
  radians_to_vec(cwh, &hvx, &hvy);
  const double signed_dist = signed_dist_to_line(ex, ey, cwx, cwy, hvx, hvy);
  const double norm_dist = signed_dist / ((double)w.wp_lane_width[crow] * 0.5);
  const double angle_error = wrap_heading(cwh - eh);
  std::vector<double> ttc_by(n, 1000.0), dist_by(n, 1.0);
  for (int k = 0; k < nv; ++k) {
    const size_t q = (size_t)ai * cap + k;
    const int vlane = w.nbr_lane[q];
    if (vlane < 0) continue;
    const double vx = w.nbr_pos[q * 3 + 0], vy = w.nbr_pos[q * 3 + 1];
    int bp = -1;
    double bd = 0.0, bdist = 0.0;
    for (int p = 0; p < n; ++p) {
      double cum = 0.0, px = 0.0, py = 0.0;
      for (int i = 0; i < plen(p); ++i) {
        const double wx = w.wp_pos[(row(p) + i) * 3 + 0], wy = w.wp_pos[(row(p) + i) * 3 + 1];
        if (i > 0) {
          const double sx = wx - px, sy = wy - py;
          cum += std::sqrt(sx * sx + sy * sy);
        }
        px = wx;
        py = wy;
        if (w.wp_lane[row(p) + i] != vlane) continue;
        const double dx = wx - vx, dy = wy - vy;
        const double d = std::sqrt(dx * dx + dy * dy);
        if (bp < 0 || d < bd) {
          bp = p;
          bd = d;
          bdist = cum;
        }
      }
    }
    if (bp < 0 || bd > 2.0) continue;
    double rel = (espeed - (double)w.nbr_speed[q]) * 1000.0 / 3600.0;
    if (std::fabs(rel) < 1e-5) rel = 1e-5;
    double ttc = bdist / rel;
    ttc /= 10.0;
    if (ttc <= 0.0) continue;
    const double ld = bdist / 100.0;
    dist_by[bp] = std::min(dist_by[bp], ld);
    ttc_by[bp] = std::min(ttc_by[bp], ttc);
  }
  for (int k = 0; k < 3; ++k) {
    const int src = li - 1 + k;
    if (src >= 0 && src < n) {
      w.ttc_ttc[(size_t)ai * 3 + k] = (float)ttc_by[src];
      w.ttc_dtc[(size_t)ai * 3 + k] = (float)dist_by[src];
    }
  }
  if (li < 0 || li >= n) truncated = true;
  w.ttc_valid[ai] = 1;
  w.ttc_angle_error[ai] = (float)angle_error;
  w.ttc_dist_center[ai] = (float)norm_dist;
  if (truncated) w.status[ai] |= SB_STATUS_TTC_TRUNCATED;
}

// AccelerometerSensor.__call__, sensors.py:1053-1084, fed as sensors.py:294-299 with the ego's
// linear velocity (radians_to_vec(heading) * speed, chassis.py:276-280) and angular velocity
// (0, 0, yaw_rate); dt = sim.last_dt = the fixed time step (smarts.py:122, :261).
void observe_accelerometer(World& w, int e, int a) {
  const int ai = e * w.A + a, vi = e * w.V + a;
  double hvx, hvy;
  radians_to_vec(w.heading[vi], &hvx, &hvy);
  auto& lv = w.acc_lin_hist[ai];
  auto& av = w.acc_ang_hist[ai];
  lv.push_back({hvx * w.speed[vi], hvy * w.speed[vi], 0.0});
  if (lv.size() > 3) lv.pop_front();  // deque(maxlen=3)
  av.push_back({0.0, 0.0, w.yaw_rate[vi]});
  if (av.size() > 3) av.pop_front();
  const double dt = w.cfg.dt;
  std::array<double, 3> lin_acc{0, 0, 0}, ang_acc{0, 0, 0}, lin_jerk{0, 0, 0}, ang_jerk{0, 0, 0};
  if (dt != 0.0) {
    auto accel = [dt](const std::deque<std::array<double, 3>>& d, std::array<double, 3>* acc,
                      std::array<double, 3>* jerk) {
      const size_t n = d.size();
      if (n < 2) return;
      for (int k = 0; k < 3; ++k) (*acc)[k] = (d[n - 1][k] - d[n - 2][k]) / dt;
      if (n < 3) return;
      for (int k = 0; k < 3; ++k) (*jerk)[k] = (*acc)[k] - (d[n - 2][k] - d[n - 3][k]) / dt;  // sic: no / dt
    };
    accel(lv, &lin_acc, &lin_jerk);
    accel(av, &ang_acc, &ang_jerk);
  }
  for (int k = 0; k < 3; ++k) {
    w.acc_linear[(size_t)ai * 3 + k] = lin_acc[k];
    w.acc_angular[(size_t)ai * 3 + k] = ang_acc[k];
    w.acc_linear_jerk[(size_t)ai * 3 + k] = lin_jerk[k];
    w.acc_angular_jerk[(size_t)ai * 3 + k] = ang_jerk[k];
  }
}

// ViaSensor.__call__, sensors.py:1104-1147
void observe_vias(World& w, int e, int a) {
  const int ai = e * w.A + a, vi = e * w.V + a, C = w.cfg.via_cap;
  const double px = w.pos_x[vi], py = w.pos_y[vi];
  const double range = w.cfg.via_acquisition_range;
  struct Near { int idx; double d2; };
  std::vector<Near> near_points;
  std::vector<int> hit_points;
  const auto& vias = w.vias[ai];
  for (int k = 0; k < (int)vias.size(); ++k) {
    const World::ViaRec& via = vias[k];
    // lane.center_at_point(position), road_map.py:357-360
    double cx, cy;
    from_lane_coord(w.map, via.lane, offset_along_lane(w.map, via.lane, px, py), &cx, &cy);
    const double lx = cx - px, ly = cy - py;  // squared_dist, utils/math.py:152-160
    if (lx * lx + ly * ly > range * range) continue;
    const double dx = via.x - px, dy = via.y - py;
    const double d2 = dx * dx + dy * dy;
    near_points.push_back({k, d2});
    const bool speed_ok = std::fabs(w.speed[vi] - via.required_speed) <=
                          w.cfg.via_speed_accuracy + 1e-5 * std::fabs(via.required_speed);  // np.isclose
    if (d2 <= via.hit_distance * via.hit_distance && !w.via_consumed[ai].count(k) && speed_ok) {
      w.via_consumed[ai].insert(k);
      hit_points.push_back(k);
    }
  }
  std::stable_sort(near_points.begin(), near_points.end(), [](const Near& p, const Near& q) { return p.d2 < q.d2; });
  for (int k = 0; k < C; ++k) {
    w.via_near_idx[(size_t)ai * C + k] = k < (int)near_points.size() ? near_points[k].idx : -1;
    w.via_hit_idx[(size_t)ai * C + k] = k < (int)hit_points.size() ? hit_points[k] : -1;
  }
  w.via_near_count[ai] = (int)near_points.size();
  w.via_hit_count[ai] = (int)hit_points.size();
}

// ---------------------------------------------------------------- RoadWaypointsSensor
// nearest_lanes(point, radius) as the sorted list (sumo_road_network.py:675-701); equal distances by lane index
std::vector<int> nearest_lanes_sorted(const Map& mp, double px, double py, double radius) {
  std::vector<std::pair<double, int>> c;
  for (int l = 0; l < mp.m.n_lanes; ++l) {
    double d = dist_point_to_polyline(mp, l, px, py);
    if (d < 0.0 || !(d < radius)) continue;
    c.emplace_back(d, l);
  }
  std::sort(c.begin(), c.end());
  std::vector<int> out;
  for (auto& q : c) out.push_back(q.second);
  return out;
}

// Lane.oncoming_lanes_at_offset, sumo_road_network.py:370-396
std::vector<int> oncoming_lanes_at_offset(const Map& mp, int lane, double off) {
  std::vector<int> result;
  const double radius = 1.1 * mp.lane_width[lane];
  double cx, cy, mvx, mvy;
  from_lane_coord(mp, lane, off, &cx, &cy);
  std::vector<int> nearby = nearest_lanes_sorted(mp, cx, cy, radius);
  if (nearby.empty()) return result;
  vector_at_offset(mp, lane, off, &mvx, &mvy);
  const double my_norm = std::sqrt(mvx * mvx + mvy * mvy);
  if (my_norm == 0.0) return result;
  for (int l : nearby) {
    if (l == lane) continue;
    double sl = offset_along_lane(mp, l, cx, cy), lvx, lvy;
    vector_at_offset(mp, l, sl, &lvx, &lvy);
    const double lv_norm = std::sqrt(lvx * lvx + lvy * lvy);
    if (lv_norm == 0.0) continue;
    const double ang = (mvx * lvx + mvy * lvy) / (my_norm * lv_norm);
    if (ang < -0.995562) result.push_back(l);
  }
  return result;
}

// `cache_*`: SumoRoadNetwork._WaypointsCache (sumo_road_network.py:1229-1276) as it acts inside one sensor call.  The
// reference keeps the paths of the last query point keyed by the start lane's INDEX IN ITS ROAD, so two lanes of
// different roads that have the same index and start the query from exactly the same point (internal lanes fanning
// out of one approach lane) share an entry: the second one is answered with the first one's paths -- which are
// lane_paths_at(first lane, point), so only the lane that filled an entry needs remembering.  The cache is taken
// as empty when the call begins (what other vehicles left in it is not modelled).
struct RwOut {
  World& w;
  int ai, n_paths = 0;
  bool truncated = false;
  bool cache_valid = false;
  double cache_x = 0.0, cache_y = 0.0;
  std::vector<std::array<int, 2>> cache_starts;  // lane index in road, lane that filled the entry
};

// RoadWaypointsSensor.paths_for_lane, sensors.py:1013-1040.  Every visit moves the cache; only a lane's last visit
// is what the dict keeps (`emit`), written as the paths of lane slot `slot`.
void rw_paths_for_lane(RwOut& o, int slot, bool emit, int lane, double px, double py, bool have_overflow,
                       double overflow, int depth) {
  World& w = o.w;
  const Map& mp = w.map;
  const int horizon = w.cfg.rw_horizon;
  double start;
  if (!have_overflow) {
    start = offset_along_lane(mp, lane, px, py) - (double)horizon;
  } else {
    start = mp.lane_length[lane] + overflow;
  }
  const int i0 = mp.lane_in_start[lane], i1 = mp.lane_in_start[lane + 1];
  if (start < 0 && i1 > i0) {
    if (depth + 1 >= SB_RW_MAX_DEPTH) {
      o.truncated = true;
      return;
    }
    for (int i = i0; i < i1; ++i) rw_paths_for_lane(o, slot, emit, mp.lane_in[i], px, py, true, start, depth + 1);
    return;
  }
  start = std::max(0.0, start);
  double sx, sy;
  from_lane_coord(mp, lane, start, &sx, &sy);
  if (mp.lane_lp_count[lane] <= 0) return;
  const int key = mp.lane_idx_in_road[lane];  // llp.lp.lane.index of the closest lane point ON this lane
  int source = lane;
  bool hit = false;
  if (o.cache_valid && sx == o.cache_x && sy == o.cache_y) {
    for (auto& c : o.cache_starts)
      if (c[0] == key) {
        source = c[1];
        hit = true;
      }
  } else {
    o.cache_valid = true;
    o.cache_x = sx;
    o.cache_y = sy;
    o.cache_starts.clear();
  }
  if (!hit) o.cache_starts.push_back({key, lane});
  if (!emit) return;
  static thread_local PathSet ps;
  ps.n = ps.total = 0;
  ps.overflow = false;
  const int rl = w.route_len[o.ai];
  lane_paths_at(mp, source, sx, sy, 2 * horizon, rl > 0 ? &w.route[(size_t)o.ai * w.cfg.route_cap] : nullptr,
                rl > 0 ? rl : 0, &ps);
  if (ps.overflow) o.truncated = true;
  const int Pc = w.cfg.rw_paths_cap, Hz = 2 * horizon + 1;
  for (int k = 0; k < ps.n; ++k) {
    const int p = o.n_paths++;
    if (p >= Pc) {
      o.truncated = true;
      continue;
    }
    const size_t q = (size_t)o.ai * Pc + p;
    w.rw_path_slot[q] = slot;
    w.rw_len[q] = ps.len[k];
    for (int i = 0; i < ps.len[k] && i < Hz; ++i) {
      const Waypoint& wp = ps.wp[k][i];
      const size_t t = q * Hz + i;
      w.rw_pos[t * 2] = wp.x;
      w.rw_pos[t * 2 + 1] = wp.y;
      w.rw_heading[t] = wp.heading;
      w.rw_lane_width[t] = (float)wp.lane_width;
      w.rw_speed_limit[t] = (float)wp.speed_limit;
      w.rw_wp_lane[t] = wp.lane;
      w.rw_lane_index[t] = (int8_t)wp.lane_index;
    }
  }
}

// RoadWaypointsSensor.__call__, sensors.py:1000-1011
void observe_road_waypoints(World& w, int e, int a) {
  const int ai = e * w.A + a, vi = e * w.V + a;
  const Map& mp = w.map;
  const int Lc = w.cfg.rw_lanes_cap, Pc = w.cfg.rw_paths_cap, Hz = 2 * w.cfg.rw_horizon + 1;
  for (int i = 0; i < Lc; ++i) w.rw_lane[(size_t)ai * Lc + i] = -1;
  for (int p = 0; p < Pc; ++p) {
    const size_t q = (size_t)ai * Pc + p;
    w.rw_path_slot[q] = -1;
    w.rw_len[q] = 0;
    for (int i = 0; i < Hz; ++i) {
      const size_t t = q * Hz + i;
      w.rw_pos[t * 2] = w.rw_pos[t * 2 + 1] = w.rw_heading[t] = 0.0;
      w.rw_lane_width[t] = w.rw_speed_limit[t] = 0.f;
      w.rw_wp_lane[t] = -1;
      w.rw_lane_index[t] = 0;
    }
  }
  w.rw_n_lanes[ai] = w.rw_n_paths[ai] = 0;
  const double px = w.pos_x[vi], py = w.pos_y[vi];
  const int nl = nearest_lane(mp, px, py, std::max(10.0, 2 * 3.2));
  if (nl < 0) return;
  const int road = mp.lane_road[nl];
  std::vector<int> croads(1, road);
  for (int i = mp.road_par_start[road]; i < mp.road_par_start[road + 1]; ++i) croads.push_back(mp.road_par[i]);
  // Road.oncoming_roads_at_point, sumo_road_network.py:595-605
  for (int l = mp.road_lane_start[road]; l < mp.road_lane_start[road] + mp.road_lane_count[road]; ++l) {
    const double off = offset_along_lane(mp, l, px, py);  // lane.to_lane_coord(point).s
    for (int ol : oncoming_lanes_at_offset(mp, l, off))
      if (mp.lane_road[ol] != road) croads.push_back(mp.lane_road[ol]);
  }
  RwOut out{w, ai};
  if ((int)croads.size() > SB_RW_MAX_ROADS) {
    croads.resize(SB_RW_MAX_ROADS);
    out.truncated = true;
  }
  // lane_paths[lane.lane_id] = paths_for_lane(lane) over every lane of every road of the list: a road listed twice
  // has its lanes computed twice; the dict keeps the first position and the last value of a key
  std::vector<int> visits, lanes;
  for (int r : croads)
    for (int l = mp.road_lane_start[r]; l < mp.road_lane_start[r] + mp.road_lane_count[r]; ++l) {
      visits.push_back(l);
      if (std::find(lanes.begin(), lanes.end(), l) == lanes.end()) lanes.push_back(l);
    }
  w.rw_n_lanes[ai] = (int)lanes.size();
  if ((int)lanes.size() > Lc) out.truncated = true;
  for (int s = 0; s < std::min((int)lanes.size(), Lc); ++s) w.rw_lane[(size_t)ai * Lc + s] = lanes[s];
  for (size_t v = 0; v < visits.size(); ++v) {
    const int l = visits[v];
    const bool last = std::find(visits.begin() + v + 1, visits.end(), l) == visits.end();
    const int slot = (int)(std::find(lanes.begin(), lanes.end(), l) - lanes.begin());
    rw_paths_for_lane(out, slot, last && slot < Lc, l, px, py, false, 0.0, 0);
  }
  w.rw_n_paths[ai] = out.n_paths;
  if (out.truncated) w.status[ai] |= SB_STATUS_RW_TRUNCATED;
}

// ego_centric_observation_adapter, smarts/core/utils/adapters/ego_centric_adapters.py:62-178, on the
// observation arrays of one agent: neighbours (:160-167), lidar points (:113-123), waypoints (:85-96).
// wrap_value: utils/math.py:452-461; position_to_ego_frame / _gen_ego_frame_matrix: utils/math.py:464-487.
double wrap_value(double value, double lo, double hi) {
  double v = value;
  const double diff = hi - lo;
  if (value <= lo) v = hi - sb_pymod(lo - value, diff);
  if (value > hi) v = lo + sb_pymod(value - hi, diff);
  return v;
}

void position_to_ego_frame(double* pos, const double ego_position[3], double ego_heading) {
  double m[3][3] = {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}};  // np.eye(3)
  double sn, cs;
  sb_sincos(-ego_heading, &sn, &cs);
  m[0][0] = cs;
  m[0][1] = -sn;
  m[1][0] = sn;
  m[1][1] = cs;
  const double rel[3] = {pos[0] - ego_position[0], pos[1] - ego_position[1], pos[2] - ego_position[2]};
  if (!(std::isfinite(rel[0]) && std::isfinite(rel[1]) && std::isfinite(rel[2]))) {
    // a lidar miss (inf, inf, inf): every row of the product holds a 0 * inf
    pos[0] = pos[1] = pos[2] = std::numeric_limits<double>::quiet_NaN();
    return;
  }
  // the third column of rows 0-1 and the first two of row 2 are zero: their products add +-0
  pos[0] = m[0][0] * rel[0] + m[0][1] * rel[1];
  pos[1] = m[1][0] * rel[0] + m[1][1] * rel[1];
  pos[2] = rel[2];
}

void ego_frame_agent(World& w, int e, int a) {
  const int ai = e * w.A + a, vi = e * w.V + a;
  const SbConfig& c = w.cfg;
  const double position[3] = {w.pos_x[vi], w.pos_y[vi], w.pos_z[vi]};
  const double heading = w.heading[vi];
  if (c.lidar_enable)
    for (int r = 0; r < c.lidar_rays; ++r) {
      const size_t q = (size_t)ai * c.lidar_rays + r;
      if (!w.lidar_hits[q] && !c.lidar_miss_inf) continue;  // zero-filled misses of StdObs are not points
      position_to_ego_frame(&w.lidar_points[q * 3], position, heading);
    }
  if (c.nbr_enable)
    for (int k = 0; k < std::min(w.nbr_count[ai], c.nbr_cap); ++k) {
      const size_t q = (size_t)ai * c.nbr_cap + k;
      position_to_ego_frame(&w.nbr_pos[q * 3], position, heading);
      w.nbr_heading[q] = (float)wrap_value((double)w.nbr_heading[q] - heading, -SB_PI, SB_PI);
    }
  if (c.wp_enable) {
    const int P = c.wp_out_paths, L = c.wp_out_len;
    for (int p = 0; p < std::min(w.wp_n_paths[ai], P); ++p)
      for (int i = 0; i < w.wp_len[(size_t)ai * P + p]; ++i) {
        const size_t q = ((size_t)ai * P + p) * L + i;
        double v[3] = {w.wp_pos[q * 3], w.wp_pos[q * 3 + 1], 0.0};  // np.append(wp.pos, [0])
        position_to_ego_frame(v, position, heading);
        w.wp_pos[q * 3] = v[0];
        w.wp_pos[q * 3 + 1] = v[1];
        w.wp_heading[q] = (float)wrap_value((double)w.wp_heading[q] - heading, -SB_PI, SB_PI);
      }
  }
  if (c.rw_enable) {  // :98-108
    const int Pc = c.rw_paths_cap, Hz = 2 * c.rw_horizon + 1;
    for (int p = 0; p < std::min(w.rw_n_paths[ai], Pc); ++p)
      for (int i = 0; i < std::min(w.rw_len[(size_t)ai * Pc + p], Hz); ++i) {
        const size_t q = ((size_t)ai * Pc + p) * Hz + i;
        double v[3] = {w.rw_pos[q * 2], w.rw_pos[q * 2 + 1], 0.0};
        position_to_ego_frame(v, position, heading);
        w.rw_pos[q * 2] = v[0];
        w.rw_pos[q * 2 + 1] = v[1];
        w.rw_heading[q] = wrap_value(w.rw_heading[q] - heading, -SB_PI, SB_PI);
      }
  }
}

void observe_agent(World& w, int e, int a) {
  const int ai = e * w.A + a, vi = e * w.V + a;
  // sensors.py:277-285, radius None -> max(10, 2 * default lane width) (sumo_road_network.py:679-680)
  w.ego_lane[ai] = nearest_lane(w.map, w.pos_x[vi], w.pos_y[vi], std::max(10.0, 2 * 3.2));
  if (w.cfg.nbr_enable) observe_neighbors(w, e, a);
  if (w.cfg.wp_enable) observe_waypoints(w, e, a);
  if (w.cfg.events_enable && (!w.cfg.wp_enable || !w.trip_init[ai])) {
    // the lookahead-1 query within the vehicle's length, no route: every observation of an agent without a
    // waypoint sensor (sensors.py:289-295), and TripMeterSensor.__init__ of every agent (sensors.py:885-898;
    // slot 1: the first observation after a reset stands at the pose the vehicle was created with)
    static thread_local PathSet ps;
    waypoint_paths(w.map, w.pos_x[vi], w.pos_y[vi], w.heading[vi], 1, w.length[vi], nullptr, -1, &ps);
    stash_trip_waypoint(w, ai, ps, w.cfg.wp_enable ? 1 : 0);
  }
  if (w.cfg.ttc_enable) observe_ttc(w, e, a);
  if (w.cfg.accel_enable) observe_accelerometer(w, e, a);
  if (w.cfg.via_cap > 0) observe_vias(w, e, a);
  if (w.cfg.rw_enable) observe_road_waypoints(w, e, a);
  if (w.cfg.lidar_enable) observe_lidar(w, e, a);
  if (w.cfg.ogm_enable) observe_ogm(w, e, a);
  if (w.cfg.dagm_enable) observe_dagm(w, e, a);
}

}  // namespace

// ==================================================================== C entry points (ctypes)
extern "C" {

void* or_create(const SbConfig* cfg, const SbMap* m, const double* base_rays, int n_rays) {
  World* w = new World();
  w->cfg = *cfg;
  sb_vehicle(cfg->vehicle_type, &w->veh);
  Map& mp = w->map;
  mp.m = *m;
  cp(mp.road_lane_start, m->road_lane_start, m->n_roads);
  cp(mp.road_lane_count, m->road_lane_count, m->n_roads);
  cp(mp.road_is_junction, m->road_is_junction, m->n_roads);
  cp(mp.road_lp_start, m->road_lp_start, m->n_roads);
  cp(mp.road_lp_count, m->road_lp_count, m->n_roads);
  cp(mp.road_first_out, m->road_first_out, m->n_roads);
  cp(mp.lane_road, m->lane_road, m->n_lanes);
  cp(mp.lane_idx_in_road, m->lane_idx_in_road, m->n_lanes);
  cp(mp.lane_width, m->lane_width, m->n_lanes);
  cp(mp.lane_speed, m->lane_speed, m->n_lanes);
  cp(mp.lane_length, m->lane_length, m->n_lanes);
  cp(mp.lane_out_start, m->lane_out_start, m->n_lanes + 1);
  cp(mp.lane_out, m->lane_out, std::max(m->n_lane_out, 1));
  cp(mp.lane_lp_start, m->lane_lp_start, m->n_lanes);
  cp(mp.lane_lp_count, m->lane_lp_count, m->n_lanes);
  cp(mp.lane_shape_start, m->lane_shape_start, m->n_lanes + 1);
  cp(mp.lp_x, m->lp_x, m->n_lanepoints);
  cp(mp.lp_y, m->lp_y, m->n_lanepoints);
  cp(mp.lp_heading, m->lp_heading, m->n_lanepoints);
  cp(mp.lp_inferred, m->lp_inferred, m->n_lanepoints);
  cp(mp.lp_lane, m->lp_lane, m->n_lanepoints);
  cp(mp.lp_next_start, m->lp_next_start, m->n_lanepoints + 1);
  cp(mp.lp_next, m->lp_next, std::max(m->n_lp_next, 1));
  cp(mp.shape_x, m->shape_x, m->n_shape_pts);
  cp(mp.shape_y, m->shape_y, m->n_shape_pts);
  cp(mp.poly_start, m->poly_start, m->n_polys + 1);
  cp(mp.poly_x, m->poly_x, m->n_poly_verts);
  cp(mp.poly_y, m->poly_y, m->n_poly_verts);
  cp(mp.road_edges, m->road_edges, (size_t)m->n_edges * 4);
  cp(mp.lane_in_start, m->lane_in_start, m->n_lanes + 1);
  cp(mp.lane_in, m->lane_in, std::max(m->n_lane_in, 1));
  cp(mp.road_par_start, m->road_par_start, m->n_roads + 1);
  cp(mp.road_par, m->road_par, std::max(m->n_road_par, 1));
  if (base_rays && n_rays > 0) {
    w->rays.assign(base_rays, base_rays + (size_t)n_rays * 3);
    w->ray_rec.resize(n_rays);
    w->ray_group_u.assign((size_t)n_rays * 2, 0.0);
    sb_lidar_factor(base_rays, n_rays, w->ray_rec.data(), w->ray_group_u.data());
  }
  w->E = cfg->n_envs;
  w->A = cfg->n_agents;
  w->M = cfg->n_social;
  w->V = w->A + w->M;
  w->substeps = cfg->substeps > 0 ? cfg->substeps : std::max(1, (int)(cfg->dt * 240));
  size_t nv = (size_t)w->E * w->V, na = (size_t)w->E * w->A;
  for (auto* v : {&w->pos_x, &w->pos_y, &w->pos_z, &w->heading, &w->speed, &w->yaw_rate, &w->length, &w->width,
                  &w->height, &w->cos_h, &w->sin_h})
    v->assign(nv, 0.0);
  for (size_t i = 0; i < nv; ++i) w->cos_h[i] = 1.0;
  w->alive.assign(nv, 0);
  w->steering.assign(na, 0.0);
  w->v_long.assign(na, 0.0);
  w->v_lat.assign(na, 0.0);
  w->ctrl.assign(na * SB_CTRL_STATE, 0.0);
  w->last_control.assign(na * 3, 0.0);
  w->route.assign(na * std::max(cfg->route_cap, 1), -1);
  w->route_len.assign(na, 0);
  w->status.assign(na, 0);
  w->ticked.assign(na, 0);
  w->ego_lane.assign(na, -1);
  if (cfg->nbr_enable) {
    size_t n = na * cfg->nbr_cap;
    w->nbr_count.assign(na, 0);
    w->nbr_id.assign(n, -1);
    w->nbr_lane.assign(n, -1);
    w->nbr_pos.assign(n * 3, 0);
    w->nbr_box.assign(n * 3, 0);
    w->nbr_heading.assign(n, 0);
    w->nbr_speed.assign(n, 0);
    w->nbr_lane_index.assign(n, 0);
  }
  if (cfg->wp_enable) {
    size_t np = na * cfg->wp_out_paths, n = np * cfg->wp_out_len;
    w->wp_n_paths.assign(na, 0);
    w->wp_len.assign(np, 0);
    w->wp_pos.assign(n * 3, 0);
    w->wp_heading.assign(n, 0);
    w->wp_lane_width.assign(n, 0);
    w->wp_speed_limit.assign(n, 0);
    w->wp_lane.assign(n, -1);
    w->wp_lane_index.assign(n, 0);
  }
  if (cfg->lidar_enable) {
    w->lidar_points.assign(na * cfg->lidar_rays * 3, 0);
    w->lidar_hits.assign(na * cfg->lidar_rays, 0);
  }
  if (cfg->events_enable) {
    double win = std::floor(cfg->not_moving_time / cfg->dt + 1e-9);
    w->ring = (int)(win + 1 < SB_DRIVEN_PATH_MAX - 1 ? win + 1 : SB_DRIVEN_PATH_MAX - 1) + 1;  // one spare interval
    sb_clock_decimals(cfg->dt, &w->dt_num, &w->dt_den);
    w->trip_init.assign(na, 0);
    w->trip_has.assign(na, 0);
    w->ev_done.assign(na, 0);
    for (auto* v : {&w->trip_x, &w->trip_y, &w->trip_h, &w->trip_dist, &w->trip_last, &w->last_px, &w->last_py,
                    &w->goal_x, &w->goal_y, &w->ev_distance, &w->ev_reward})
      v->assign(na, 0.0);
    w->goal_r.assign(na, -1.0);
    w->trip_wp.assign(na * 8, 0.0);
    w->path_cum.assign(na * w->ring, 0.0);
    for (auto* v : {&w->steps, &w->path_n, &w->ev_n_collisions, &w->ev_collidee, &w->ev_steps}) v->assign(na, 0);
    w->ev_events.assign(na, 0);
  }
  if (cfg->ttc_enable) {
    w->ttc_angle_error.assign(na, 0.f);
    w->ttc_dist_center.assign(na, 0.f);
    w->ttc_dtc.assign(na * 3, 0.f);
    w->ttc_ttc.assign(na * 3, 0.f);
    w->ttc_valid.assign(na, 0);
  }
  if (cfg->accel_enable) {
    w->acc_lin_hist.assign(na, {});
    w->acc_ang_hist.assign(na, {});
    for (auto* v : {&w->acc_linear, &w->acc_angular, &w->acc_linear_jerk, &w->acc_angular_jerk}) v->assign(na * 3, 0.0);
  }
  if (cfg->via_cap > 0) {
    w->vias.assign(na, {});
    w->via_consumed.assign(na, {});
    w->via_near_count.assign(na, 0);
    w->via_hit_count.assign(na, 0);
    w->via_near_idx.assign(na * cfg->via_cap, -1);
    w->via_hit_idx.assign(na * cfg->via_cap, -1);
  }
  if (cfg->rw_enable) {
    const size_t np = na * cfg->rw_paths_cap, n = np * (2 * cfg->rw_horizon + 1);
    w->rw_n_lanes.assign(na, 0);
    w->rw_n_paths.assign(na, 0);
    w->rw_lane.assign(na * cfg->rw_lanes_cap, -1);
    w->rw_path_slot.assign(np, -1);
    w->rw_len.assign(np, 0);
    w->rw_pos.assign(n * 2, 0.0);
    w->rw_heading.assign(n, 0.0);
    w->rw_lane_width.assign(n, 0.f);
    w->rw_speed_limit.assign(n, 0.f);
    w->rw_wp_lane.assign(n, -1);
    w->rw_lane_index.assign(n, 0);
  }
  if (cfg->ogm_enable) w->ogm.assign(na * cfg->ogm_width * cfg->ogm_height, 0);
  if (cfg->dagm_enable) w->dagm.assign(na * cfg->dagm_width * cfg->dagm_height, 0);
  return w;
}

void or_destroy(void* h) { delete static_cast<World*>(h); }
void or_set_threads(void* h, int n) { static_cast<World*>(h)->threads = n < 1 ? 1 : n; }

// SMARTS.reset for a subset of envs: agents placed at their mission start, sedan chassis
// dimensions (vehicle.urdf:12-16), social slots cleared.
void or_reset(void* h, const int32_t* env_ids, int n, const SbResetState* init) {
  World& w = *static_cast<World*>(h);
  if (!env_ids) n = w.E;
  for (int i = 0; i < n; ++i) {
    int e = env_ids ? env_ids[i] : i;
    for (int a = 0; a < w.A; ++a) {
      int ai = e * w.A + a, vi = e * w.V + a;
      size_t src = (size_t)i * w.A + a;
      w.pos_x[vi] = init->agent_x[src];
      w.pos_y[vi] = init->agent_y[src];
      w.pos_z[vi] = 0.0;
      w.heading[vi] = wrap_heading(init->agent_heading[src]);
      w.speed[vi] = init->agent_speed[src];
      w.yaw_rate[vi] = 0.0;
      w.length[vi] = w.veh.length;
      w.width[vi] = w.veh.width;
      w.height[vi] = w.veh.height;
      sb_sincos(w.heading[vi], &w.sin_h[vi], &w.cos_h[vi]);
      w.alive[vi] = 1;
      w.steering[ai] = 0.0;
      w.v_long[ai] = w.speed[vi];
      w.v_lat[ai] = 0.0;
      for (int k = 0; k < SB_CTRL_STATE; ++k) w.ctrl[(size_t)ai * SB_CTRL_STATE + k] = 0.0;
      for (int k = 0; k < 3; ++k) w.last_control[(size_t)ai * 3 + k] = 0.0;
      w.status[ai] = 0;
      w.ticked[ai] = 0;
      int rc = w.cfg.route_cap;
      for (int k = 0; k < rc; ++k) w.route[(size_t)ai * rc + k] = init->route ? init->route[src * rc + k] : -1;
      w.route_len[ai] = init->route_len ? init->route_len[src] : 0;
      if (w.cfg.events_enable) {
        w.trip_init[ai] = w.trip_has[ai] = 0;
        w.trip_x[ai] = w.trip_y[ai] = w.trip_h[ai] = w.trip_dist[ai] = w.trip_last[ai] = 0.0;
        w.trip_wp[(size_t)ai * 8 + 3] = -1.0;
        w.trip_wp[(size_t)ai * 8 + 7] = -1.0;
        w.steps[ai] = w.path_n[ai] = 0;
        w.last_px[ai] = w.last_py[ai] = 0.0;
        const bool goals = init->goal_x && init->goal_y && init->goal_radius;
        w.goal_x[ai] = goals ? init->goal_x[src] : 0.0;
        w.goal_y[ai] = goals ? init->goal_y[src] : 0.0;
        w.goal_r[ai] = goals ? init->goal_radius[src] : -1.0;
      }
      if (w.cfg.accel_enable) {  // a new vehicle carries a new AccelerometerSensor
        w.acc_lin_hist[ai].clear();
        w.acc_ang_hist[ai].clear();
      }
      if (w.cfg.via_cap > 0) {
        w.vias[ai].clear();
        w.via_consumed[ai].clear();
        const int C = w.cfg.via_cap;
        const int cnt = init->via_count ? init->via_count[src] : 0;
        for (int k = 0; k < cnt && k < C; ++k) {
          const size_t q = src * C + k;
          w.vias[ai].push_back({init->via_lane[q], init->via_x[q], init->via_y[q], init->via_hit_distance[q],
                                init->via_required_speed[q]});
        }
      }
    }
    for (int j = w.A; j < w.V && !init->keep_social; ++j) {
      int vj = e * w.V + j;
      w.alive[vj] = 0;
      w.pos_x[vj] = w.pos_y[vj] = w.pos_z[vj] = w.heading[vj] = w.speed[vj] = w.yaw_rate[vj] = 0.0;
      w.length[vj] = w.width[vj] = w.height[vj] = 0.0;
      w.cos_h[vj] = 1.0;
      w.sin_h[vj] = 0.0;
    }
  }
}

// BoxChassis.control + yaw_rate, chassis.py:211-217, :302-308.  states: host [E, M, 8].
void or_set_social(void* h, const double* states) {
  World& w = *static_cast<World*>(h);
  for (int e = 0; e < w.E; ++e)
    for (int m = 0; m < w.M; ++m) {
      const double* s = &states[((size_t)e * w.M + m) * 8];
      int vj = e * w.V + w.A + m;
      bool was = w.alive[vj] != 0;
      bool now = s[7] != 0.0;
      double hn = wrap_heading(s[2]);
      w.yaw_rate[vj] = (was && now) ? angdiff(hn, w.heading[vj]) / w.cfg.dt : 0.0;
      w.pos_x[vj] = s[0];
      w.pos_y[vj] = s[1];
      w.pos_z[vj] = 0.0;
      w.heading[vj] = hn;
      w.speed[vj] = s[3];
      w.length[vj] = s[4];
      w.width[vj] = s[5];
      w.height[vj] = s[6];
      w.alive[vj] = now ? 1 : 0;
      sb_sincos(hn, &w.sin_h[vj], &w.cos_h[vj]);
    }
}

// SMARTS._step_providers: controllers then the chassis step (smarts.py:1048-1110)
void or_step(void* h, const void* actions) {
  World& w = *static_cast<World*>(h);
  const int as = w.cfg.action_space;
  parallel_for(w.E, w.threads, 4, [&](int e) {
    for (int a = 0; a < w.A; ++a) {
      int ai = e * w.A + a, vi = e * w.V + a;
      Ego v{w.pos_x[vi], w.pos_y[vi], w.heading[vi], w.speed[vi], w.yaw_rate[vi], w.v_long[ai], w.v_lat[ai]};
      double* st = &w.ctrl[(size_t)ai * SB_CTRL_STATE];
      double* out3 = &w.last_control[(size_t)ai * 3];
      const int32_t* route = &w.route[(size_t)ai * w.cfg.route_cap];
      if (as == SB_ACTION_LANE) {
        int cmd = static_cast<const int32_t*>(actions)[ai];
        // controllers/__init__.py:125-144
        double ts = 15;
        int lc = 0;
        if (cmd == SB_LANE_SLOW_DOWN) ts = 0;
        else if (cmd == SB_LANE_CHANGE_LEFT) { ts = 12.5; lc = 1; }
        else if (cmd == SB_LANE_CHANGE_RIGHT) { ts = 12.5; lc = -1; }
        w.status[ai] |= lane_following(w.map, w.veh, v, route, w.route_len[ai], ts, lc, w.cfg.dt, st, out3);
      } else if (as == SB_ACTION_LANE_WITH_CONTINUOUS_SPEED) {
        const double* ac = static_cast<const double*>(actions) + (size_t)ai * 2;
        w.status[ai] |= lane_following(w.map, w.veh, v, route, w.route_len[ai], ac[0], (int)ac[1], w.cfg.dt, st, out3);
      } else if (as == SB_ACTION_TRAJECTORY) {
        const int N = w.cfg.traj_len;
        trajectory_tracking(w.veh, v, static_cast<const double*>(actions) + (size_t)ai * 4 * N, N, w.cfg.dt, st, out3);
      } else if (as == SB_ACTION_CONTINUOUS) {
        const double* ac = static_cast<const double*>(actions) + (size_t)ai * 3;
        apply_control(v, clipd(ac[0], 0, 1), clipd(ac[1], 0, 1), clipd(ac[2], -1, 1), out3);
      } else if (as == SB_ACTION_ACTUATOR_DYNAMIC) {
        // actuator_dynamic_controller.py:51-80
        const double* ac = static_cast<const double*>(actions) + (size_t)ai * 3;
        double steering = clipd((1 - 0.001) * st[0] + clipd(ac[2], -1, 1) * w.cfg.dt, -1, 1);
        apply_control(v, clipd(ac[0], 0, 1), clipd(ac[1], 0, 1), steering, out3);
        st[0] = steering;
      }
      chassis_step(w, ai, vi);
    }
  });
  std::fill(w.ticked.begin(), w.ticked.end(), (uint8_t)1);
}

// Sensors.observe for the agents of the listed envs (env_ids == nullptr: every env).  Envs are independent
// simulations in the reference (one SMARTS instance each, parallel_env.py:96-127): observing one never touches
// the sensor state of another.
void or_observe_envs(void* h, const int32_t* env_ids, int n) {
  World& w = *static_cast<World*>(h);
  const int ne = env_ids ? n : w.E;
  auto env_of = [&](int i) { return env_ids ? env_ids[i / w.A] : i / w.A; };
  parallel_for(ne * w.A, w.threads, 1, [&](int i) { observe_agent(w, env_of(i), i % w.A); });
  if (w.cfg.events_enable) {
    parallel_for(ne * w.A, w.threads, 1, [&](int i) { observe_events(w, env_of(i), i % w.A); });
  }
  if (w.cfg.ego_frame) parallel_for(ne * w.A, w.threads, 1, [&](int i) { ego_frame_agent(w, env_of(i), i % w.A); });
}

void or_observe(void* h) { or_observe_envs(h, nullptr, 0); }

void or_get_state(void* h, SbStatePtrs* o) {
  World& w = *static_cast<World*>(h);
  o->pos_x = w.pos_x.data(); o->pos_y = w.pos_y.data(); o->pos_z = w.pos_z.data();
  o->heading = w.heading.data(); o->speed = w.speed.data(); o->yaw_rate = w.yaw_rate.data();
  o->length = w.length.data(); o->width = w.width.data(); o->height = w.height.data();
  o->cos_h = w.cos_h.data(); o->sin_h = w.sin_h.data(); o->alive = w.alive.data();
  o->steering = w.steering.data(); o->v_long = w.v_long.data(); o->v_lat = w.v_lat.data();
  o->ctrl = w.ctrl.data(); o->last_control = w.last_control.data();
  o->route = w.route.data(); o->route_len = w.route_len.data(); o->status = w.status.data();
}

void or_get_obs(void* h, SbObsPtrs* o) {
  World& w = *static_cast<World*>(h);
  std::memset(o, 0, sizeof(*o));
  o->ego_lane = w.ego_lane.data();
  o->nbr_count = w.nbr_count.data(); o->nbr_id = w.nbr_id.data(); o->nbr_pos = w.nbr_pos.data();
  o->nbr_box = w.nbr_box.data(); o->nbr_heading = w.nbr_heading.data(); o->nbr_speed = w.nbr_speed.data();
  o->nbr_lane = w.nbr_lane.data(); o->nbr_lane_index = w.nbr_lane_index.data();
  o->wp_n_paths = w.wp_n_paths.data(); o->wp_len = w.wp_len.data(); o->wp_pos = w.wp_pos.data();
  o->wp_heading = w.wp_heading.data(); o->wp_lane_width = w.wp_lane_width.data();
  o->wp_speed_limit = w.wp_speed_limit.data(); o->wp_lane = w.wp_lane.data();
  o->wp_lane_index = w.wp_lane_index.data();
  o->lidar_points = w.lidar_points.data(); o->lidar_hits = w.lidar_hits.data();
  o->ogm = w.ogm.data(); o->dagm = w.dagm.data();
  o->ev_events = w.ev_events.data(); o->ev_done = w.ev_done.data(); o->ev_n_collisions = w.ev_n_collisions.data();
  o->ev_collidee = w.ev_collidee.data(); o->ev_distance = w.ev_distance.data(); o->ev_reward = w.ev_reward.data();
  o->ev_steps = w.ev_steps.data();
  o->ttc_angle_error = w.ttc_angle_error.data(); o->ttc_distance_from_center = w.ttc_dist_center.data();
  o->ttc_dtc = w.ttc_dtc.data(); o->ttc_ttc = w.ttc_ttc.data(); o->ttc_valid = w.ttc_valid.data();
  o->acc_linear = w.acc_linear.data(); o->acc_angular = w.acc_angular.data();
  o->acc_linear_jerk = w.acc_linear_jerk.data(); o->acc_angular_jerk = w.acc_angular_jerk.data();
  o->via_near_count = w.via_near_count.data(); o->via_near_idx = w.via_near_idx.data();
  o->via_hit_count = w.via_hit_count.data(); o->via_hit_idx = w.via_hit_idx.data();
  o->rw_n_lanes = w.rw_n_lanes.data(); o->rw_lane = w.rw_lane.data(); o->rw_n_paths = w.rw_n_paths.data();
  o->rw_path_slot = w.rw_path_slot.data(); o->rw_len = w.rw_len.data(); o->rw_pos = w.rw_pos.data();
  o->rw_heading = w.rw_heading.data(); o->rw_lane_width = w.rw_lane_width.data();
  o->rw_speed_limit = w.rw_speed_limit.data(); o->rw_wp_lane = w.rw_wp_lane.data();
  o->rw_lane_index = w.rw_lane_index.data();
}

// ---- unit-level probes used to pin the oracle against the reference's golden vectors ----
// RoadMap.waypoint_paths.  Outputs: n paths (return), lens[SB_MAX_PATHS], wp[SB_MAX_PATHS][65][7]
// (x, y, heading, lane_width, speed_limit, lane, lane_index).
int or_waypoint_paths(void* h, double x, double y, double heading, int lookahead, double within_radius,
                      const int32_t* route, int route_len, int32_t* lens, double* wp_out) {
  World& w = *static_cast<World*>(h);
  static thread_local PathSet ps;
  waypoint_paths(w.map, x, y, heading, lookahead, within_radius, route, route_len, &ps);
  for (int p = 0; p < ps.n; ++p) {
    lens[p] = ps.len[p];
    for (int i = 0; i < ps.len[p]; ++i) {
      double* o = &wp_out[((size_t)p * (SB_MAX_LOOKAHEAD + 1) + i) * 7];
      const Waypoint& q = ps.wp[p][i];
      o[0] = q.x; o[1] = q.y; o[2] = q.heading; o[3] = q.lane_width; o[4] = q.speed_limit;
      o[5] = q.lane; o[6] = q.lane_index;
    }
  }
  return ps.n;
}

int or_nearest_lane(void* h, double x, double y, double radius, double* dist) {
  return nearest_lane(static_cast<World*>(h)->map, x, y, radius, dist);
}

// lane-coordinate probes (RoadMap.Lane.offset_along_lane / vector_at_offset / from_lane_coord, road_map.py:346-417;
// SumoRoadNetwork.road_with_point, sumo_road_network.py:703-709): out = {s of (x, y), vector at `off` x, y,
// point at `off` x, y, on-road flag of (x, y)}
void or_lane_probe(void* h, int lane, double x, double y, double off, double* out) {
  const Map& mp = static_cast<World*>(h)->map;
  out[0] = offset_along_lane(mp, lane, x, y);
  vector_at_offset(mp, lane, off, &out[1], &out[2]);
  from_lane_coord(mp, lane, off, &out[3], &out[4]);
  out[5] = road_with_point(mp, x, y) ? 1.0 : 0.0;
}

void or_lateral_gains_v(int vehicle_type, double target_speed, double* heading_gain, double* lateral_gain) {
  SbVehicle vp;
  sb_vehicle(vehicle_type, &vp);
  lateral_gains(target_speed, vp, heading_gain, lateral_gain);
}
void or_lateral_gains(double target_speed, double* heading_gain, double* lateral_gain) {
  or_lateral_gains_v(SB_VEHICLE_SEDAN, target_speed, heading_gain, lateral_gain);
}

// one controller call on an explicit vehicle state (golden checks vs the reference controllers)
uint32_t or_lane_following(void* h, const double* veh7, const int32_t* route, int route_len, double target_speed,
                           int lane_change, double dt, double* st, double* out3) {
  Ego v{veh7[0], veh7[1], veh7[2], veh7[3], veh7[4], veh7[5], veh7[6]};
  World& w = *static_cast<World*>(h);
  return lane_following(w.map, w.veh, v, route, route_len, target_speed, lane_change, dt, st, out3);
}

void or_trajectory_tracking_v(int vehicle_type, const double* veh7, const double* traj, int N, double dt, double* st,
                               double* out3) {
  Ego v{veh7[0], veh7[1], veh7[2], veh7[3], veh7[4], veh7[5], veh7[6]};
  SbVehicle vp;
  sb_vehicle(vehicle_type, &vp);
  trajectory_tracking(vp, v, traj, N, dt, st, out3);
}
void or_trajectory_tracking(const double* veh7, const double* traj, int N, double dt, double* st, double* out3) {
  or_trajectory_tracking_v(SB_VEHICLE_SEDAN, veh7, traj, N, dt, st, out3);
}

void or_sincos(const double* x, int n, double* s, double* c) {
  for (int i = 0; i < n; ++i) sb_sincos(x[i], &s[i], &c[i]);
}
void or_atan2(const double* y, const double* x, int n, double* out) {
  for (int i = 0; i < n; ++i) out[i] = sb_atan2(y[i], x[i]);
}

}  // extern "C"
