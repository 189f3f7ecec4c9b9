// sb_kernels.cuh -- the per-tick kernels.
//
//   k_set_social      BoxChassis.control for injected traffic             (chassis.py:211-217, :302-308)
//   k_control_paths   Lane / LaneWithContinuousSpeed: waypoint path search    (warp per agent)
//   k_control_law     ... re-spacing, steering/throttle law, chassis           (thread per agent)
//   k_control_simple  Trajectory / Continuous / ActuatorDynamic + chassis  (thread per agent)
//   k_observe         ego lane, NeighborhoodVehicles, Waypoints            (warp per agent)
//   k_lidar           LidarSensor ray casting                              (block per agent)
//   k_raster          OGM + DrivableAreaGridMap                            (block per agent)
//   k_events          Events / done / trip-meter reward / contacts         (warp per agent, sb_events.cuh)
//   k_ttc             lane_ttc (StdObs ttc)                                (warp per agent, sb_ttc.cuh)
//   k_extras          accelerometer + via sensor                            (thread per agent, sb_extras.cuh)
//   k_road_waypoints  RoadWaypoints sensor                                 (warp per agent, sb_roadwp.cuh)
#pragma once

#include "sb_control.cuh"
#include "sb_waypoints.cuh"

struct DObs {
  int32_t* ego_lane;
  int32_t *nbr_count, *nbr_id;
  double* nbr_pos;
  float *nbr_box, *nbr_heading, *nbr_speed;
  int32_t* nbr_lane;
  int8_t* nbr_lane_index;
  int32_t *wp_n_paths, *wp_len;
  double* wp_pos;
  float *wp_heading, *wp_lane_width, *wp_speed_limit;
  int32_t* wp_lane;
  int8_t* wp_lane_index;
  double* lidar_points;
  int8_t* lidar_hits;
  uint8_t *ogm, *dagm;
};

struct DCfg {
  int32_t E, A, M, V;
  double dt;
  int32_t substeps, action_space, traj_len, route_cap;
  int32_t nbr_enable, nbr_cap;
  double nbr_radius;
  int32_t wp_enable, wp_lookahead, wp_out_paths, wp_out_len;
  int32_t lidar_enable, lidar_rays, lidar_miss_inf;
  double lidar_wall_height;
  int32_t ogm_enable, ogm_w, ogm_h;
  double ogm_res;
  int32_t dagm_enable, dagm_w, dagm_h;
  double dagm_res;
  int32_t events_enable, max_steps, ttc_enable, accel_enable, via_cap, ego_frame;
  int32_t rw_enable, rw_horizon, rw_lanes_cap, rw_paths_cap;
  SbVehicle veh;  // the agents' vehicle type (include/sb_vehicle.h)
  uint32_t done_mask;
  double nm_time, nm_dist, via_range, via_speed_acc;
  // sb_observe_envs: the envs the observation kernels work on (device array; nullptr = every env, in order)
  const int32_t* env_list;
  int32_t n_list;
  uint32_t a_magic;  // floor(2^32 / A) + 1 when idx / A == umulhi(idx, a_magic) for every work index, else 0
};

// work item idx -> (env, agent) of an observation kernel; false past the last one
__device__ __forceinline__ bool d_work_agent(const DCfg& c, int idx, int* e, int* a) {
  const int ne = c.env_list ? c.n_list : c.E;
  if (idx >= ne * c.A) return false;
  const int q = c.a_magic ? (int)__umulhi((uint32_t)idx, c.a_magic) : idx / c.A;
  *a = idx - q * c.A;
  *e = c.env_list ? c.env_list[q] : q;
  return true;
}

// --------------------------------------------------------------------------------- social
__global__ void k_set_social(DCfg c, DState s, const double* __restrict__ in) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= c.E * c.M) return;
  int e = i / c.M, m = i - e * c.M;
  const double* p = in + (size_t)i * 8;
  int vj = e * c.V + c.A + m;
  bool was = s.alive[vj] != 0;
  bool now = p[7] != 0.0;
  double hn = d_wrap_heading(p[2]);
  s.yaw_rate[vj] = (was && now) ? d_angdiff(hn, s.heading[vj]) / c.dt : 0.0;
  s.pos_x[vj] = p[0];
  s.pos_y[vj] = p[1];
  s.pos_z[vj] = 0.0;
  s.heading[vj] = hn;
  s.speed[vj] = p[3];
  s.length[vj] = p[4];
  s.width[vj] = p[5];
  s.height[vj] = p[6];
  s.alive[vj] = now ? 1 : 0;
  sb_sincos(hn, &s.sin_h[vj], &s.cos_h[vj]);
}

// --------------------------------------------------------------------------------- control
// Lane / LaneWithContinuousSpeed in two launches.  k_control_paths (warp per agent): the
// waypoint path search -- strided argmin scans, lane-level branch enumeration -- and the choice
// of the lane to follow; the chosen path's lane points go to a small global record.
// k_control_law (thread per agent): the inherently serial part -- re-spacing of that one path,
// the steering / throttle law, the substepped chassis -- so its dependent instruction chain is
// issued once per agent instead of once per warp with 31 lanes idle.
#define SB_CTRL_WARPS 4
#define SB_CTRL_LAW_THREADS 32
#define SB_CTRL_LOOKAHEAD 16

// k_control_paths -> k_control_law: the waypoints of the path the controller follows
struct CtrlPick {
  double wx[SB_CTRL_LOOKAHEAD + 1], wy[SB_CTRL_LOOKAHEAD + 1], wh[SB_CTRL_LOOKAHEAD + 1];
  int32_t len;  // waypoints on the path; 0: no waypoints found
};

// controllers/__init__.py:125-152: Lane strings / LaneWithContinuousSpeed tuple -> (target speed, lane change)
__device__ __forceinline__ void d_lane_action(const DCfg& c, const void* actions, int ai, double* ts, int* lc) {
  *ts = 15;
  *lc = 0;
  if (c.action_space == SB_ACTION_LANE) {
    int cmd = static_cast<const int32_t*>(actions)[ai];
    if (cmd == SB_LANE_SLOW_DOWN) *ts = 0;
    else if (cmd == SB_LANE_CHANGE_LEFT) { *ts = 12.5; *lc = 1; }
    else if (cmd == SB_LANE_CHANGE_RIGHT) { *ts = 12.5; *lc = -1; }
  } else {
    const double* ac = static_cast<const double*>(actions) + (size_t)ai * 2;
    *ts = ac[0];
    *lc = (int)ac[1];
  }
}

#ifndef SB_CTRL_MIN_BLOCKS
#define SB_CTRL_MIN_BLOCKS 12  // 40 registers (92 B of spills): 48 resident warps hide the L1 / L2 latency of the path search; 1 -> 12: 0.107 -> 0.095 ms on the headline, 0.319 -> 0.272 ms on the full suite at 65 k agents (profiles/README.md r2y)
#endif
using CtrlWarpPaths = WarpPathsT<SB_CTRL_LOOKAHEAD>;  // 2.6 KB per warp instead of 7.4 KB

__global__ void __launch_bounds__(SB_CTRL_WARPS * 32, SB_CTRL_MIN_BLOCKS)
k_control_paths(DCfg c, DMap m, DState s, const void* __restrict__ actions, CtrlPick* __restrict__ picks) {
  __shared__ CtrlWarpPaths wps[SB_CTRL_WARPS];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int ai = blockIdx.x * SB_CTRL_WARPS + warp;
  if (ai >= c.E * c.A) return;
  CtrlWarpPaths* wp = &wps[warp];
  const int e = ai / c.A, a = ai - e * c.A, vi = e * c.V + a;
  const double vx = s.pos_x[vi], vy = s.pos_y[vi], vh = s.heading[vi];
  double ts;
  int lc;
  d_lane_action(c, actions, ai, &ts, &lc);
  wp_waypoint_paths(m, vx, vy, vh, SB_CTRL_LOOKAHEAD, 5.0, s.route + (size_t)ai * c.route_cap, s.route_len[ai], wp,
                    lane);
  __syncwarp();
  const int n = wp->n;
  if (lane == 0) {
    uint32_t bits = (wp->overflow ? SB_STATUS_PATH_OVERFLOW : 0u) | (n == 0 ? SB_STATUS_NO_WAYPOINTS : 0u);
    if (bits) s.status[ai] |= bits;
    if (n == 0) picks[ai].len = 0;
  }
  if (n == 0) return;
  int rank = 0x7fffffff;
  double d0 = 1e300;
  if (lane < n) {  // find_current_lane, lane_following_controller.py:367-374
    double fx, fy;
    wp_first_point(m, wp->lp[lane], wp->len[lane], vx, vy, &fx, &fy);
    rank = wp_rank(m, wp, lane);
    double dx = fx - vx, dy = fy - vy;
    d0 = sqrt(dx * dx + dy * dy);
  }
  double dmin = d0;
  int cur = rank;
  warp_argmin(dmin, cur);
  int sel = cur + lc;
  sel = sel < 0 ? 0 : (sel > n - 1 ? n - 1 : sel);
  // the selected path is interpolated by the whole warp, one waypoint per lane
  const unsigned who = __ballot_sync(SB_FULL_MASK, rank == sel);
  const int src = __ffs(who) - 1;
  CtrlPick* pk = &picks[ai];
  const int L = wp_equally_spaced_warp(
      m, wp->lp[src], wp->len[src], vx, vy, SB_CTRL_LOOKAHEAD + 1,
      [&](int i, const DWaypoint& w) {
        pk->wx[i] = w.x;
        pk->wy[i] = w.y;
        pk->wh[i] = w.heading;
      },
      wp, lane);
  if (lane == 0) pk->len = L;
}

__global__ void __launch_bounds__(SB_CTRL_LAW_THREADS)
k_control_law(DCfg c, DMap m, DState s, const void* __restrict__ actions, const CtrlPick* __restrict__ picks) {
  const int ai = blockIdx.x * SB_CTRL_LAW_THREADS + threadIdx.x;
  if (ai >= c.E * c.A) return;
  s.ticked[ai] = 1;
  const int e = ai / c.A, a = ai - e * c.A, vi = e * c.V + a;
  DEgo v{s.pos_x[vi], s.pos_y[vi], s.heading[vi], s.speed[vi], s.yaw_rate[vi], s.v_long[ai], s.v_lat[ai]};
  double* st = s.ctrl + (size_t)ai * SB_CTRL_STATE;
  double* out3 = s.last_control + (size_t)ai * 3;
  const CtrlPick& pk = picks[ai];
  if (pk.len == 0) {
    d_apply_control(v, 0, 0, 0, out3);
  } else {
    double ts;
    int lc;
    d_lane_action(c, actions, ai, &ts, &lc);
    d_lane_following_law(v, c.veh, pk.wx, pk.wy, pk.wh, pk.len, ts, lc, c.dt, st, out3);
  }
  d_chassis_step(s, c.veh, ai, vi, c.dt, c.substeps);
}

__global__ void k_control_simple(DCfg c, DState s, const double* __restrict__ actions) {
  int ai = blockIdx.x * blockDim.x + threadIdx.x;
  if (ai >= c.E * c.A) return;
  s.ticked[ai] = 1;
  const int e = ai / c.A, a = ai - e * c.A, vi = e * c.V + a;
  DEgo v{s.pos_x[vi], s.pos_y[vi], s.heading[vi], s.speed[vi], s.yaw_rate[vi], s.v_long[ai], s.v_lat[ai]};
  double* st = s.ctrl + (size_t)ai * SB_CTRL_STATE;
  double* out3 = s.last_control + (size_t)ai * 3;
  if (c.action_space == SB_ACTION_TRAJECTORY) {
    const int N = c.traj_len;
    double traj[4 * SB_MAX_TRAJ];
    const double* src = actions + (size_t)ai * 4 * N;
    for (int i = 0; i < 4 * N; ++i) traj[i] = src[i];
    d_trajectory_tracking(v, c.veh, traj, N, c.dt, st, out3);
  } else if (c.action_space == SB_ACTION_CONTINUOUS) {  // controllers/__init__.py:94-99
    const double* ac = actions + (size_t)ai * 3;
    d_apply_control(v, d_clip(ac[0], 0, 1), d_clip(ac[1], 0, 1), d_clip(ac[2], -1, 1), out3);
  } else {  // actuator_dynamic_controller.py:51-80
    const double* ac = actions + (size_t)ai * 3;
    double steering = d_clip((1 - 0.001) * st[0] + d_clip(ac[2], -1, 1) * c.dt, -1, 1);
    d_apply_control(v, d_clip(ac[0], 0, 1), d_clip(ac[1], 0, 1), steering, out3);
    st[0] = steering;
  }
  d_chassis_step(s, c.veh, ai, vi, c.dt, c.substeps);
}

// --------------------------------------------------------------------------------- observe
// utils/math.py:348-367 + :393-411
__device__ __forceinline__ double d_dist_point_to_segment(double px, double py, double ax, double ay, double bx,
                                                          double by) {
  double dx = ax - bx, dy = ay - by;
  double d = sqrt(dx * dx + dy * dy);
  double u = ((px - ax) * (bx - ax)) + ((py - ay) * (by - ay));
  double offset;
  if (d == 0.0 || u < 0.0 || u > d * d) {
    offset = (u < 0.0) ? 0.0 : d;
  } else {
    offset = u / d;
  }
  if (offset == 0.0) {
    double ex = px - ax, ey = py - ay;
    return sqrt(ex * ex + ey * ey);
  }
  double uu = offset / d;
  double ix = ax + uu * (bx - ax), iy = ay + uu * (by - ay);
  double ex = px - ix, ey = py - iy;
  return sqrt(ex * ex + ey * ey);
}

// Segment range a query at (px, py) with `radius` has to look at: the whole table, the list of
// the point's grid cell, or nothing (the point is further than the radius from every segment).
// Conservative by construction, so results equal the full scan's.
__device__ __forceinline__ void d_seg_candidates(const DMap& m, double px, double py, double radius,
                                                 const int32_t** items, int* n) {
  *items = nullptr;
  *n = m.n_shape_pts - 1;
  if (m.seg_nx == 0 || !(radius <= m.seg_rmax)) return;
  const double gx = (px - m.seg_x0) * m.seg_inv_cell, gy = (py - m.seg_y0) * m.seg_inv_cell;
  if (!(gx >= 0.0 && gy >= 0.0 && gx < (double)m.seg_nx && gy < (double)m.seg_ny)) {
    *n = 0;  // the grid covers every segment's box grown by seg_rmax
    return;
  }
  const int cell = (int)gy * m.seg_nx + (int)gx;
  const int s0 = m.seg_cell_start[cell];
  *items = m.seg_items + s0;
  *n = m.seg_cell_start[cell + 1] - s0;
}

// nearest_lanes(point, radius)[0] (sumo_road_network.py:675-701): the minimum over lanes of the
// minimum over segments is the minimum over all segments; ties go to the lowest lane index.
__device__ int d_nearest_lane_warp(const DMap& m, double px, double py, double radius, int lane) {
  double bd = 1e300;
  int bl = 0x7fffffff;
  const int32_t* items;
  int n;
  d_seg_candidates(m, px, py, radius, &items, &n);
  for (int t = lane; t < n; t += 32) {
    const int i = items ? items[t] : t;
    int l = m.shape_lane[i];
    if (l != m.shape_lane[i + 1]) continue;
    double d = d_dist_point_to_segment(px, py, m.shape_x[i], m.shape_y[i], m.shape_x[i + 1], m.shape_y[i + 1]);
    if (d < bd || (d == bd && l < bl)) {
      bd = d;
      bl = l;
    }
  }
  warp_argmin(bd, bl);
  if (bl == 0x7fffffff || !(bd < radius)) return -1;
  return bl;
}

#define SB_OBS_WARPS 4

// nearest_lane(vehicle position, radius = sedan length) of every alive vehicle slot, once per env instead of once
// per (agent, neighbour) pair (sensors.py:243-266); warp per vehicle.
__global__ void __launch_bounds__(SB_OBS_WARPS * 32) k_vehicle_lanes(DCfg c, DMap m, DState s, int32_t* veh_lane) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int idx = blockIdx.x * SB_OBS_WARPS + warp;
  if (idx >= (c.env_list ? c.n_list : c.E) * c.V) return;
  const int q = idx / c.V;
  const int vj = (c.env_list ? c.env_list[q] : q) * c.V + (idx - q * c.V);
  int nl = -1;
  if (s.alive[vj]) nl = d_nearest_lane_warp(m, s.pos_x[vj], s.pos_y[vj], c.veh.length, lane);
  if (lane == 0) veh_lane[vj] = nl;
}


#ifndef SB_OBS_MIN_BLOCKS
#define SB_OBS_MIN_BLOCKS 8  // 64 registers: 0.628 -> 0.567 ms on the full suite at 65 k agents (10: 0.569, 12: 0.654 -- spills), profiles/README.md r2y
#endif
// LA: the per-warp path record is sized for lookaheads up to LA (32 covers the Waypoints sensor's default; 64 the cap)
template <int LA>
__global__ void __launch_bounds__(SB_OBS_WARPS * 32, SB_OBS_MIN_BLOCKS)
k_observe(DCfg c, DMap m, DState s, DObs o, double* __restrict__ trip_wp, const uint8_t* __restrict__ trip_init,
          const int32_t* __restrict__ veh_lane) {
  using WarpPaths = WarpPathsT<LA>;
  __shared__ WarpPaths wps[SB_OBS_WARPS];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  int e, a;
  if (!d_work_agent(c, blockIdx.x * SB_OBS_WARPS + warp, &e, &a)) return;
  const int ai = e * c.A + a, vi = e * c.V + a;
  const double x = s.pos_x[vi], y = s.pos_y[vi], z = s.pos_z[vi], h = s.heading[vi];

  // ego lane, sensors.py:277-285 (radius None -> max(10, 2 * 3.2))
  {
    int el = d_nearest_lane_warp(m, x, y, 10.0, lane);
    if (lane == 0) o.ego_lane[ai] = el;
  }

  if (c.nbr_enable) {  // smarts.py:1191-1208, sensors.py:241-266
    const int cap = c.nbr_cap;
    const double ego_len = s.length[vi];
    int cnt = 0;
    for (int base = 0; base < c.V; base += 32) {
      int j = base + lane;
      bool ok = false;
      int vj = e * c.V + j;
      if (j < c.V && j != a && s.alive[vj]) {
        ok = true;
        if (c.nbr_radius >= 0) {
          double dx = s.pos_x[vj] - x, dy = s.pos_y[vj] - y, dz = s.pos_z[vj] - z;
          double d = sqrt((dx * dx + dy * dy) + dz * dz);
          ok = d <= c.nbr_radius;
        }
      }
      unsigned bal = __ballot_sync(SB_FULL_MASK, ok);
      int pos = cnt + __popc(bal & ((1u << lane) - 1u));
      if (ok && pos < cap) {
        size_t q = (size_t)ai * cap + pos;
        o.nbr_id[q] = j;
        o.nbr_pos[q * 3 + 0] = s.pos_x[vj];
        o.nbr_pos[q * 3 + 1] = s.pos_y[vj];
        o.nbr_pos[q * 3 + 2] = s.pos_z[vj];
        o.nbr_box[q * 3 + 0] = (float)s.length[vj];
        o.nbr_box[q * 3 + 1] = (float)s.width[vj];
        o.nbr_box[q * 3 + 2] = (float)s.height[vj];
        o.nbr_heading[q] = (float)s.heading[vj];
        o.nbr_speed[q] = (float)s.speed[vj];
      }
      cnt += __popc(bal);
    }
    if (lane == 0) o.nbr_count[ai] = cnt;
    __syncwarp();
    int nw = cnt < cap ? cnt : cap;
    for (int k = 0; k < nw; ++k) {
      size_t q = (size_t)ai * cap + k;
      int vj = e * c.V + o.nbr_id[q];
      // the lane of a vehicle does not depend on who looks at it (the radius is the sedan's length for every
      // agent): k_vehicle_lanes has it when the env has fewer vehicles than neighbour slots to fill
      int nl = (veh_lane && ego_len == c.veh.length) ? veh_lane[vj]
                                                        : d_nearest_lane_warp(m, s.pos_x[vj], s.pos_y[vj], ego_len, lane);
      if (lane == 0) {
        o.nbr_lane[q] = nl;
        o.nbr_lane_index[q] = nl >= 0 ? (int8_t)m.lane_idx_in_road[nl] : 0;
      }
    }
    for (int k = nw + lane; k < cap; k += 32) {
      size_t q = (size_t)ai * cap + k;
      o.nbr_id[q] = -1;
      o.nbr_lane[q] = -1;
      o.nbr_lane_index[q] = 0;
      o.nbr_heading[q] = 0.f;
      o.nbr_speed[q] = 0.f;
      for (int t = 0; t < 3; ++t) {
        o.nbr_pos[q * 3 + t] = 0.0;
        o.nbr_box[q * 3 + t] = 0.f;
      }
    }
  }

  if (c.events_enable && (!c.wp_enable || !trip_init[ai])) {
    // The lookahead-1 query within the vehicle's length, no route; only path 0's first waypoint is needed.
    // Every observation of an agent without a waypoint sensor (sensors.py:289-295), and TripMeterSensor.__init__
    // of every agent (sensors.py:885-898; slot 1: the first observation after a reset stands at the pose the
    // vehicle was created with).
    WarpPaths* wp = &wps[warp];
    wp_waypoint_paths(m, x, y, h, 1, s.length[vi], nullptr, -1, wp, lane);
    __syncwarp();
    const int n = wp->n;
    double* tw = trip_wp + (size_t)ai * 8 + (c.wp_enable ? 4 : 0);
    if (n == 0 && lane == 0) tw[3] = -1.0;
    const unsigned first = __ballot_sync(SB_FULL_MASK, lane < n && wp_rank(m, wp, lane) == 0);
    if (first) {
      const int p = __ffs(first) - 1;
      wp_equally_spaced_warp(
          m, wp->lp[p], wp->len[p], x, y, 1,
          [&](int, const DWaypoint& w) {
            tw[0] = w.x;
            tw[1] = w.y;
            tw[2] = w.heading;
            tw[3] = (double)w.lane;
          },
          wp, lane);
    }
  }

  if (c.wp_enable) {  // sensors.py:972-988
    WarpPaths* wp = &wps[warp];
    const int P = c.wp_out_paths, L = c.wp_out_len;
    wp_waypoint_paths(m, x, y, h, c.wp_lookahead, 5.0, s.route + (size_t)ai * c.route_cap, s.route_len[ai], wp,
                      lane);
    __syncwarp();
    const int n = wp->n;
    if (lane == 0) {
      o.wp_n_paths[ai] = wp->total;
      if (wp->overflow) s.status[ai] |= SB_STATUS_PATH_OVERFLOW;
      if (c.events_enable && n == 0) trip_wp[(size_t)ai * 8 + 3] = -1.0;
    }
    // each path is interpolated by the whole warp (wp_equally_spaced_warp); lane k ranks path k
    const int my_rank = lane < n ? wp_rank(m, wp, lane) : 0x7fffffff;
    for (int p = 0; p < n; ++p) {
      const int rank = __shfl_sync(SB_FULL_MASK, my_rank, p);
      if (rank >= P) continue;
      const size_t row = ((size_t)ai * P + rank) * L;
      const bool stash = c.events_enable && rank == 0;  // waypoint_paths[0][0] feeds the trip meter (sensors.py:340-341)
      const int got = wp_equally_spaced_warp(
          m, wp->lp[p], wp->len[p], x, y, L,
          [&](int i, const DWaypoint& w) {
            if (stash && i == 0) {
              double* tw = trip_wp + (size_t)ai * 8;
              tw[0] = w.x;
              tw[1] = w.y;
              tw[2] = w.heading;
              tw[3] = (double)w.lane;
            }
            size_t q = row + i;
            o.wp_pos[q * 3 + 0] = w.x;
            o.wp_pos[q * 3 + 1] = w.y;
            o.wp_pos[q * 3 + 2] = 0.0;
            o.wp_heading[q] = (float)w.heading;
            o.wp_lane_width[q] = (float)w.lane_width;
            o.wp_speed_limit[q] = (float)w.speed_limit;
            o.wp_lane[q] = w.lane;
            o.wp_lane_index[q] = (int8_t)w.lane_index;
          },
          wp, lane);
      const int len = got < L ? got : L;
      if (lane == 0) o.wp_len[(size_t)ai * P + rank] = len;
      for (int i = len + lane; i < L; i += 32) {
        size_t q = row + i;
        o.wp_pos[q * 3 + 0] = o.wp_pos[q * 3 + 1] = o.wp_pos[q * 3 + 2] = 0.0;
        o.wp_heading[q] = o.wp_lane_width[q] = o.wp_speed_limit[q] = 0.f;
        o.wp_lane[q] = -1;
        o.wp_lane_index[q] = 0;
      }
    }
    // paths that do not exist: zero rows
    int kept = n < P ? n : P;
    for (int p = kept; p < P; ++p) {
      if (lane == 0) o.wp_len[(size_t)ai * P + p] = 0;
      size_t row = ((size_t)ai * P + p) * L;
      for (int i = lane; i < L; i += 32) {
        size_t q = row + i;
        o.wp_pos[q * 3 + 0] = o.wp_pos[q * 3 + 1] = o.wp_pos[q * 3 + 2] = 0.0;
        o.wp_heading[q] = o.wp_lane_width[q] = o.wp_speed_limit[q] = 0.f;
        o.wp_lane[q] = -1;
        o.wp_lane_index[q] = 0;
      }
    }
  }
}

// --------------------------------------------------------------------------------- events
#include "sb_events.cuh"
#include "sb_roadwp.cuh"
#include "sb_ttc.cuh"
#include "sb_extras.cuh"

// --------------------------------------------------------------------------------- lidar
// k_lidar lives in sb_lidar.cuh (azimuth-sorted warps, sector / per-ray culls, exact binary64 tests).
#include "sb_lidar.cuh"

// --------------------------------------------------------------------------------- raster
// k_raster lives in sb_raster.cuh (row-per-thread exact integer spans).
#include "sb_raster.cuh"

