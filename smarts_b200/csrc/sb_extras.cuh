// sb_extras.cuh -- k_extras: the two small stateful ego sensors of SURVEY §8(f) row 4.
//
//  * AccelerometerSensor (smarts/core/sensors.py:1046-1084): the last three linear / angular velocity
//    vectors of the ego; acceleration = (v[-1] - v[-2]) / dt, jerk = acc - previous acc (no second
//    division by dt in the reference, :1073 and :1082).  The velocity vectors are the ones the host
//    mirror reports in EgoVehicleObservation: linear = radians_to_vec(heading) * speed
//    (chassis.py:276-280), angular = (0, 0, yaw_rate).
//  * ViaSensor (sensors.py:1090-1150): for every via of the agent's mission (plan.py:180-188) the
//    closest centre-line point of the via's lane to the ego (Lane.center_at_point, road_map.py:357-360);
//    vias within the acquisition range are "near" (sorted by distance to the ego, stable); a near via
//    within its hit distance, not collected before and with ego speed np.isclose to the required speed
//    (atol = speed accuracy, rtol = 1e-5) is "hit" and recorded as consumed.
//
// One thread per agent: both are a few dozen flops on state that is already in L2.
#pragma once

struct DExtras {
  // accelerometer state: lin_hist[ai][k] = (vx, vy) k + 1 observations ago, ang_hist likewise (z)
  double *lin_hist, *ang_hist;
  int32_t* hist_n;  // observations seen since the reset, capped at 2
  double *acc_lin, *acc_ang, *jerk_lin, *jerk_ang;  // [E*A, 3]
  // vias
  int32_t *via_count, *via_lane;
  double *via_x, *via_y, *via_hit, *via_speed;
  uint32_t* via_consumed;  // bit k: via k collected
  int32_t *near_count, *near_idx, *hit_count, *hit_idx;
};

__global__ void __launch_bounds__(128) k_extras(DCfg c, DMap m, DState s, DExtras x) {
  int e, a;
  if (!d_work_agent(c, blockIdx.x * blockDim.x + threadIdx.x, &e, &a)) return;
  const int ai = e * c.A + a, vi = e * c.V + a;
  if (c.accel_enable) {
    double hvx, hvy;
    d_radians_to_vec(s.heading[vi], &hvx, &hvy);
    const double sp = s.speed[vi];
    const double lin[3] = {hvx * sp, hvy * sp, 0.0};
    const double ang[3] = {0.0, 0.0, s.yaw_rate[vi]};
    const int n = x.hist_n[ai];
    double* lh = x.lin_hist + (size_t)ai * 4;
    double* ah = x.ang_hist + (size_t)ai * 2;
    double la[3] = {0.0, 0.0, 0.0}, aa[3] = {0.0, 0.0, 0.0}, lj[3] = {0.0, 0.0, 0.0}, aj[3] = {0.0, 0.0, 0.0};
    if (n >= 1 && c.dt != 0.0) {
      const double l1[3] = {lh[0], lh[1], 0.0}, a1[3] = {0.0, 0.0, ah[0]};
      for (int k = 0; k < 3; ++k) {
        la[k] = (lin[k] - l1[k]) / c.dt;
        aa[k] = (ang[k] - a1[k]) / c.dt;
      }
      if (n >= 2) {
        const double l2[3] = {lh[2], lh[3], 0.0}, a2[3] = {0.0, 0.0, ah[1]};
        for (int k = 0; k < 3; ++k) {
          lj[k] = la[k] - (l1[k] - l2[k]) / c.dt;
          aj[k] = aa[k] - (a1[k] - a2[k]) / c.dt;
        }
      }
    }
    for (int k = 0; k < 3; ++k) {
      x.acc_lin[(size_t)ai * 3 + k] = la[k];
      x.acc_ang[(size_t)ai * 3 + k] = aa[k];
      x.jerk_lin[(size_t)ai * 3 + k] = lj[k];
      x.jerk_ang[(size_t)ai * 3 + k] = aj[k];
    }
    lh[2] = lh[0];
    lh[3] = lh[1];
    lh[0] = lin[0];
    lh[1] = lin[1];
    ah[1] = ah[0];
    ah[0] = ang[2];
    x.hist_n[ai] = n < 2 ? n + 1 : 2;
  }
  if (c.via_cap > 0) {
    const int C = c.via_cap;
    const double px = s.pos_x[vi], py = s.pos_y[vi], sp = s.speed[vi];
    const int nv = x.via_count[ai];
    const double range_sq = c.via_range * c.via_range;
    uint32_t consumed = x.via_consumed[ai];
    int32_t* near = x.near_idx + (size_t)ai * C;
    int32_t* hit = x.hit_idx + (size_t)ai * C;
    double key[SB_MAX_VIAS];
    int nn = 0, nh = 0;
    for (int k = 0; k < nv; ++k) {
      const size_t q = (size_t)ai * C + k;
      const int lane = x.via_lane[q];
      double cx, cy;
      ev_from_lane_coord(m, lane, ev_offset_along_lane(m, lane, px, py), &cx, &cy);
      const double lx = cx - px, ly = cy - py;
      if (lx * lx + ly * ly > range_sq) continue;
      const double dx = x.via_x[q] - px, dy = x.via_y[q] - py;
      const double d2 = dx * dx + dy * dy;
      // sorted(near_points, key=squared_dist(point.position, vehicle_position)): stable insertion
      int pos = nn;
      while (pos > 0 && key[pos - 1] > d2) {
        key[pos] = key[pos - 1];
        near[pos] = near[pos - 1];
        --pos;
      }
      key[pos] = d2;
      near[pos] = k;
      ++nn;
      const double hd = x.via_hit[q], rs = x.via_speed[q];
      if (d2 <= hd * hd && !((consumed >> k) & 1u) && fabs(sp - rs) <= c.via_speed_acc + 1e-5 * fabs(rs)) {
        consumed |= 1u << k;
        hit[nh++] = k;
      }
    }
    for (int k = nn; k < C; ++k) near[k] = -1;
    for (int k = nh; k < C; ++k) hit[k] = -1;
    x.near_count[ai] = nn;
    x.hit_count[ai] = nh;
    x.via_consumed[ai] = consumed;
  }
}

// new AccelerometerSensor / ViaSensor for the agents of the reset envs (vehicle.py:344-457, :552-558)
__global__ void k_reset_extras(DCfg c, DExtras x, const int32_t* env_ids, int n, const int32_t* vcount,
                               const int32_t* vlane, const double* vx, const double* vy, const double* vhit,
                               const double* vspeed) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n * c.A) return;
  const int i = t / c.A, a = t - i * c.A;
  const int e = env_ids ? env_ids[i] : i;
  const int ai = e * c.A + a;
  if (c.accel_enable) {
    x.hist_n[ai] = 0;
    for (int k = 0; k < 4; ++k) x.lin_hist[(size_t)ai * 4 + k] = 0.0;
    x.ang_hist[(size_t)ai * 2] = x.ang_hist[(size_t)ai * 2 + 1] = 0.0;
  }
  if (c.via_cap > 0) {
    const int C = c.via_cap;
    int cnt = vcount ? vcount[t] : 0;
    cnt = cnt < 0 ? 0 : (cnt > C ? C : cnt);
    x.via_count[ai] = cnt;
    x.via_consumed[ai] = 0u;
    for (int k = 0; k < C; ++k) {
      const size_t q = (size_t)ai * C + k, src = (size_t)t * C + k;
      const bool on = k < cnt;
      x.via_lane[q] = on ? vlane[src] : -1;
      x.via_x[q] = on ? vx[src] : 0.0;
      x.via_y[q] = on ? vy[src] : 0.0;
      x.via_hit[q] = on ? vhit[src] : 0.0;
      x.via_speed[q] = on ? vspeed[src] : 0.0;
    }
  }
}

// ---------------------------------------------------------------------------------------------
// k_ego_frame: ego_centric_observation_adapter (smarts/core/utils/adapters/ego_centric_adapters.py:62-178)
// on the observation buffers, in place, after every consumer of the world-frame values has run.
//   transform(v)      = position_to_ego_frame(v, ego.position, ego.heading), utils/math.py:464-487:
//                       rotation by -heading of (v - ego position), z kept
//   adjust_heading(h) = wrap_value(h - ego.heading, -pi, pi), utils/math.py:452-461
// One CTA per agent, threads stride over the agent's lidar points, neighbours and waypoints.
#define SB_EGO_THREADS 128

__device__ __forceinline__ double d_wrap_value_pi(double value) {
  double v = value;
  if (value <= -SB_PI) v = SB_PI - sb_pymod(-SB_PI - value, SB_TWO_PI);
  if (value > SB_PI) v = -SB_PI + sb_pymod(value - SB_PI, SB_TWO_PI);
  return v;
}

__device__ __forceinline__ void d_to_ego_frame(double* p, double ex, double ey, double ez, double cs, double sn) {
  const double rx = p[0] - ex, ry = p[1] - ey, rz = p[2] - ez;
  if (!(isfinite(rx) && isfinite(ry) && isfinite(rz))) {  // inf through the matrix product: 0 * inf
    p[0] = p[1] = p[2] = __longlong_as_double(0x7ff8000000000000LL);
    return;
  }
  // [[cos(-h), -sin(-h), 0], [sin(-h), cos(-h), 0], [0, 0, 1]] @ rel
  p[0] = cs * rx + (-sn) * ry;
  p[1] = sn * rx + cs * ry;
  p[2] = rz;
}

__global__ void __launch_bounds__(SB_EGO_THREADS) k_ego_frame(DCfg c, DState s, DObs o, DRoadWp rw) {
  int e, a;
  if (!d_work_agent(c, blockIdx.x, &e, &a)) return;
  const int ai = e * c.A + a, vi = e * c.V + a;
  const double ex = s.pos_x[vi], ey = s.pos_y[vi], ez = s.pos_z[vi], eh = s.heading[vi];
  double sn, cs;
  sb_sincos(-eh, &sn, &cs);
  if (c.lidar_enable) {
    const int R = c.lidar_rays;
    for (int r = threadIdx.x; r < R; r += SB_EGO_THREADS) {
      const size_t q = (size_t)ai * R + r;
      if (!o.lidar_hits[q] && !c.lidar_miss_inf) continue;  // StdObs zero fill (format_obs.py:437-443) stays
      d_to_ego_frame(o.lidar_points + q * 3, ex, ey, ez, cs, sn);
    }
  }
  if (c.nbr_enable) {
    const int cnt = o.nbr_count[ai], n = cnt < c.nbr_cap ? cnt : c.nbr_cap;
    for (int k = threadIdx.x; k < n; k += SB_EGO_THREADS) {
      const size_t q = (size_t)ai * c.nbr_cap + k;
      d_to_ego_frame(o.nbr_pos + q * 3, ex, ey, ez, cs, sn);
      o.nbr_heading[q] = (float)d_wrap_value_pi((double)o.nbr_heading[q] - eh);
    }
  }
  if (c.wp_enable) {
    const int P = c.wp_out_paths, L = c.wp_out_len;
    const int tp = o.wp_n_paths[ai], np_ = tp < P ? tp : P;
    for (int t = threadIdx.x; t < np_ * L; t += SB_EGO_THREADS) {
      const int p = t / L, i = t - p * L;
      if (i >= o.wp_len[(size_t)ai * P + p]) continue;
      const size_t q = ((size_t)ai * P + p) * L + i;
      // transform(np.append(wp.pos, [0]))[:2]: the waypoint sits at z = 0, the stored z stays 0
      double v[3] = {o.wp_pos[q * 3 + 0], o.wp_pos[q * 3 + 1], 0.0};
      d_to_ego_frame(v, ex, ey, ez, cs, sn);
      o.wp_pos[q * 3 + 0] = v[0];
      o.wp_pos[q * 3 + 1] = v[1];
      o.wp_heading[q] = (float)d_wrap_value_pi((double)o.wp_heading[q] - eh);
    }
  }
  if (c.rw_enable) {  // ego_centric_adapters.py:98-108
    const int Pc = c.rw_paths_cap, Hz = 2 * c.rw_horizon + 1;
    const int tp = rw.n_paths[ai], np_ = tp < Pc ? tp : Pc;
    for (int t = threadIdx.x; t < np_ * Hz; t += SB_EGO_THREADS) {
      const int p = t / Hz, i = t - p * Hz;
      if (i >= rw.len[(size_t)ai * Pc + p]) continue;
      const size_t q = ((size_t)ai * Pc + p) * Hz + i;
      double v[3] = {rw.pos[q * 2 + 0], rw.pos[q * 2 + 1], 0.0};
      d_to_ego_frame(v, ex, ey, ez, cs, sn);
      rw.pos[q * 2 + 0] = v[0];
      rw.pos[q * 2 + 1] = v[1];
      rw.heading[q] = d_wrap_value_pi(rw.heading[q] - eh);
    }
  }
}
