// sb_ttc.cuh -- k_ttc: lane_ttc of smarts/env/custom_observations.py:148-290, the `ttc` member of
// StdObs (smarts/env/wrappers/format_obs.py:546-556).  One warp per agent, lane p owns waypoint path p.
//
// Reads the agent's own observation buffers (waypoint paths, neighbours with their lanes) and the
// ego state; writes angle_error, distance_from_center, dtc[3], ttc[3] (right / current / left lane)
// as float32, computed in binary64 in the reference's operation order.  The reference works on the
// uncapped Observation; when the neighbour block, the path list or the path length was capped for
// this agent (or the ego lane index points past the path list, an IndexError there) the result is
// still the formula on what is in the buffers and SB_STATUS_TTC_TRUNCATED is raised.
#pragma once

struct DTtcOut {
  float *angle_error, *dist_center, *dtc, *ttc;
  uint8_t* valid;
};

#define SB_TTC_WARPS 4

__global__ void __launch_bounds__(SB_TTC_WARPS * 32) k_ttc(DCfg c, DState s, DObs o, DTtcOut out) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  int e, a;
  if (!d_work_agent(c, blockIdx.x * SB_TTC_WARPS + warp, &e, &a)) return;
  const int ai = e * c.A + a, vi = e * c.V + a;
  const int P = c.wp_out_paths, L = c.wp_out_len, cap = c.nbr_cap;
  const double ex = s.pos_x[vi], ey = s.pos_y[vi], eh = s.heading[vi], espeed = s.speed[vi];
  const int total_paths = o.wp_n_paths[ai];
  int n = total_paths < P ? total_paths : P;
  n = n < 32 ? n : 32;
  const int cnt = o.nbr_count[ai];
  const int nv = cnt < cap ? cnt : cap;
  bool truncated = total_paths > P || cnt > cap;
  const size_t prow = ((size_t)ai * P + lane) * L;
  const int len = lane < n ? o.wp_len[(size_t)ai * P + lane] : 0;
  if (lane < n && len == L && L < c.wp_lookahead + 1) truncated = true;
  // `if not obs.neighborhood_vehicle_states or not obs.waypoint_paths: return None`
  const bool any_path = __any_sync(SB_FULL_MASK, len > 0);
  if (cnt == 0 || n == 0 || !any_path) {
    if (lane == 0) {
      out.valid[ai] = 0;
      out.angle_error[ai] = 0.f;
      out.dist_center[ai] = 0.f;
      for (int k = 0; k < 3; ++k) out.dtc[(size_t)ai * 3 + k] = out.ttc[(size_t)ai * 3 + k] = 0.f;
    }
    return;
  }
  // closest first waypoint: min(wps, key=dist_to(ego.position)), first minimum wins
  double d0 = 1e300;
  int p0 = 0x7fffffff;
  if (len > 0) {
    const double dx = o.wp_pos[prow * 3 + 0] - ex, dy = o.wp_pos[prow * 3 + 1] - ey;
    d0 = sqrt(dx * dx + dy * dy);
    p0 = lane;
  }
  warp_argmin(d0, p0);
  const size_t crow = ((size_t)ai * P + p0) * L;
  const double cwx = o.wp_pos[crow * 3 + 0], cwy = o.wp_pos[crow * 3 + 1], cwh = (double)o.wp_heading[crow];
  const int ego_lane_index = (int)o.wp_lane_index[crow];
  double hvx, hvy;
  d_radians_to_vec(cwh, &hvx, &hvy);
  const double signed_dist = d_signed_dist_to_line(ex, ey, cwx, cwy, hvx, hvy);
  const double norm_dist = signed_dist / ((double)o.wp_lane_width[crow] * 0.5);
  const double angle_error = d_wrap_heading(cwh - eh);  // closest_wp.heading.relative_to(ego.heading)

  // _ttc_by_path: lane p keeps the minima of path p
  double ttc_p = 1000.0, dist_p = 1.0;
  for (int k = 0; k < nv; ++k) {
    const size_t q = (size_t)ai * cap + k;
    const int vlane = o.nbr_lane[q];
    if (vlane < 0) continue;  // lane_id None matches no waypoint (warp-uniform)
    const double vx = o.nbr_pos[q * 3 + 0], vy = o.nbr_pos[q * 3 + 1];
    // nearest waypoint on the vehicle's lane over all paths: ties go to the first (path, index)
    double bd = 1e300, bdist = 0.0;
    double cum = 0.0, px = 0.0, py = 0.0;
    for (int i = 0; i < len; ++i) {
      const double wx = o.wp_pos[(prow + i) * 3 + 0], wy = o.wp_pos[(prow + i) * 3 + 1];
      if (i > 0) {
        const double sx = wx - px, sy = wy - py;
        cum += sqrt(sx * sx + sy * sy);
      }
      px = wx;
      py = wy;
      if (o.wp_lane[prow + i] == vlane) {
        const double dx = wx - vx, dy = wy - vy;
        const double d = sqrt(dx * dx + dy * dy);
        if (d < bd) {
          bd = d;
          bdist = cum;
        }
      }
    }
    double wd = bd;
    int wp_ = bd < 1e299 ? lane : 0x7fffffff;
    warp_argmin(wd, wp_);
    if (wp_ == 0x7fffffff) continue;  // not on a nearby lane
    if (wd > 2.0) continue;           // too far from the path
    if (lane == wp_) {
      double rel = (espeed - (double)o.nbr_speed[q]) * 1000.0 / 3600.0;
      if (fabs(rel) < 1e-5) rel = 1e-5;
      double ttc = bdist / rel;
      ttc /= 10.0;
      if (ttc > 0.0) {
        const double ld = bdist / 100.0;
        dist_p = ld < dist_p ? ld : dist_p;
        ttc_p = ttc < ttc_p ? ttc : ttc_p;
      }
    }
  }
  // _ego_ttc_calc: index the per-path lists by the ego lane index
  const int li = ego_lane_index;
  float ttc3[3] = {0.f, 0.f, 0.f}, dtc3[3] = {0.f, 0.f, 0.f};
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    const int src = li - 1 + k;
    const bool ok = src >= 0 && src < n;
    const double tv = __shfl_sync(SB_FULL_MASK, ttc_p, ok ? src : 0);
    const double dv = __shfl_sync(SB_FULL_MASK, dist_p, ok ? src : 0);
    if (ok) {
      ttc3[k] = (float)tv;
      dtc3[k] = (float)dv;
    }
  }
  if (li < 0 || li >= n) truncated = true;  // IndexError in the reference
  if (lane == 0) {
    out.valid[ai] = 1;
    out.angle_error[ai] = (float)angle_error;
    out.dist_center[ai] = (float)norm_dist;
    for (int k = 0; k < 3; ++k) {
      out.dtc[(size_t)ai * 3 + k] = dtc3[k];
      out.ttc[(size_t)ai * 3 + k] = ttc3[k];
    }
  }
  if (__any_sync(SB_FULL_MASK, truncated) && lane == 0) s.status[ai] |= SB_STATUS_TTC_TRUNCATED;
}
