// sb_kernels.cuh -- the per-tick kernels.
//
//   k_set_social      BoxChassis.control for injected traffic             (chassis.py:211-217, :302-308)
//   k_control_lane    Lane / LaneWithContinuousSpeed action path + chassis (warp per agent)
//   k_control_simple  Trajectory / Continuous / ActuatorDynamic + chassis  (thread per agent)
//   k_observe         ego lane, NeighborhoodVehicles, Waypoints            (warp per agent)
//   k_lidar           LidarSensor ray casting                              (block per agent)
//   k_raster          OGM + DrivableAreaGridMap                            (block per agent)
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
};

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
#define SB_CTRL_WARPS 4

__global__ void __launch_bounds__(SB_CTRL_WARPS * 32)
k_control_lane(DCfg c, DMap m, DState s, const void* __restrict__ actions) {
  __shared__ WarpPaths wps[SB_CTRL_WARPS];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int ai = blockIdx.x * SB_CTRL_WARPS + warp;
  if (ai >= c.E * c.A) return;
  const int e = ai / c.A, a = ai - e * c.A, vi = e * c.V + a;
  DEgo v{s.pos_x[vi], s.pos_y[vi], s.heading[vi], s.speed[vi], s.yaw_rate[vi], s.v_long[ai], s.v_lat[ai]};
  double ts = 15;
  int lc = 0;
  if (c.action_space == SB_ACTION_LANE) {  // controllers/__init__.py:125-144
    int cmd = static_cast<const int32_t*>(actions)[ai];
    if (cmd == SB_LANE_SLOW_DOWN) ts = 0;
    else if (cmd == SB_LANE_CHANGE_LEFT) { ts = 12.5; lc = 1; }
    else if (cmd == SB_LANE_CHANGE_RIGHT) { ts = 12.5; lc = -1; }
  } else {
    const double* ac = static_cast<const double*>(actions) + (size_t)ai * 2;
    ts = ac[0];
    lc = (int)ac[1];
  }
  WarpPaths* wp = &wps[warp];
  wp_waypoint_paths(m, v.x, v.y, v.h, 16, 5.0, s.route + (size_t)ai * c.route_cap, s.route_len[ai], wp, lane);
  __syncwarp();
  const int n = wp->n;
  double* st = s.ctrl + (size_t)ai * SB_CTRL_STATE;
  double* out3 = s.last_control + (size_t)ai * 3;
  if (lane == 0 && wp->overflow) s.status[ai] |= SB_STATUS_PATH_OVERFLOW;
  if (n == 0) {
    if (lane == 0) {
      s.status[ai] |= SB_STATUS_NO_WAYPOINTS;
      d_apply_control(v, 0, 0, 0, out3);
      d_chassis_step(s, ai, vi, c.dt, c.substeps);
    }
    return;
  }
  double wx[17], wy[17], wh[17];
  int L = 0, rank = 0x7fffffff;
  double d0 = 1e300;
  if (lane < n) {
    L = wp_equally_spaced<17>(m, wp->lp[lane], wp->len[lane], v.x, v.y, 17, [&](int i, const DWaypoint& w) {
      wx[i] = w.x;
      wy[i] = w.y;
      wh[i] = w.heading;
    });
    rank = wp_rank(m, wp, lane);
    double dx = wx[0] - v.x, dy = wy[0] - v.y;
    d0 = sqrt(dx * dx + dy * dy);  // find_current_lane, :367-374
  }
  double dmin = d0;
  int cur = rank;
  warp_argmin(dmin, cur);
  int sel = cur + lc;
  sel = sel < 0 ? 0 : (sel > n - 1 ? n - 1 : sel);
  if (rank == sel) {
    d_lane_following_law(v, wx, wy, wh, L, ts, lc, c.dt, st, out3);
    d_chassis_step(s, ai, vi, c.dt, c.substeps);
  }
}

__global__ void k_control_simple(DCfg c, DState s, const double* __restrict__ actions) {
  int ai = blockIdx.x * blockDim.x + threadIdx.x;
  if (ai >= c.E * c.A) return;
  const int e = ai / c.A, a = ai - e * c.A, vi = e * c.V + a;
  DEgo v{s.pos_x[vi], s.pos_y[vi], s.heading[vi], s.speed[vi], s.yaw_rate[vi], s.v_long[ai], s.v_lat[ai]};
  double* st = s.ctrl + (size_t)ai * SB_CTRL_STATE;
  double* out3 = s.last_control + (size_t)ai * 3;
  if (c.action_space == SB_ACTION_TRAJECTORY) {
    const int N = c.traj_len;
    double traj[4 * SB_MAX_TRAJ];
    const double* src = actions + (size_t)ai * 4 * N;
    for (int i = 0; i < 4 * N; ++i) traj[i] = src[i];
    d_trajectory_tracking(v, traj, N, c.dt, st, out3);
  } else if (c.action_space == SB_ACTION_CONTINUOUS) {  // controllers/__init__.py:94-99
    const double* ac = actions + (size_t)ai * 3;
    d_apply_control(v, d_clip(ac[0], 0, 1), d_clip(ac[1], 0, 1), d_clip(ac[2], -1, 1), out3);
  } else {  // actuator_dynamic_controller.py:51-80
    const double* ac = actions + (size_t)ai * 3;
    double steering = d_clip((1 - 0.001) * st[0] + d_clip(ac[2], -1, 1) * c.dt, -1, 1);
    d_apply_control(v, d_clip(ac[0], 0, 1), d_clip(ac[1], 0, 1), steering, out3);
    st[0] = steering;
  }
  d_chassis_step(s, ai, vi, c.dt, c.substeps);
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

// nearest_lanes(point, radius)[0] (sumo_road_network.py:675-701): the minimum over lanes of the
// minimum over segments is the minimum over all segments; ties go to the lowest lane index.
__device__ int d_nearest_lane_warp(const DMap& m, double px, double py, double radius, int lane) {
  double bd = 1e300;
  int bl = 0x7fffffff;
  for (int i = lane; i + 1 < m.n_shape_pts; i += 32) {
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

__global__ void __launch_bounds__(SB_OBS_WARPS * 32) k_observe(DCfg c, DMap m, DState s, DObs o) {
  __shared__ WarpPaths wps[SB_OBS_WARPS];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int ai = blockIdx.x * SB_OBS_WARPS + warp;
  if (ai >= c.E * c.A) return;
  const int e = ai / c.A, a = ai - e * c.A, vi = e * c.V + a;
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
      int nl = d_nearest_lane_warp(m, s.pos_x[vj], s.pos_y[vj], ego_len, lane);
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
    }
    int rank = lane < n ? wp_rank(m, wp, lane) : 0x7fffffff;
    if (rank < P) {
      size_t row = ((size_t)ai * P + rank) * L;
      int got = wp_equally_spaced<SB_MAX_LOOKAHEAD + 1>(
          m, wp->lp[lane], wp->len[lane], x, y, L, [&](int i, const DWaypoint& w) {
            size_t q = row + i;
            o.wp_pos[q * 3 + 0] = w.x;
            o.wp_pos[q * 3 + 1] = w.y;
            o.wp_pos[q * 3 + 2] = 0.0;
            o.wp_heading[q] = (float)w.heading;
            o.wp_lane_width[q] = (float)w.lane_width;
            o.wp_speed_limit[q] = (float)w.speed_limit;
            o.wp_lane[q] = w.lane;
            o.wp_lane_index[q] = (int8_t)w.lane_index;
          });
      int len = got < L ? got : L;
      o.wp_len[(size_t)ai * P + rank] = len;
      for (int i = len; i < L; ++i) {
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

// --------------------------------------------------------------------------------- lidar
#define SB_LIDAR_THREADS 320
#define SB_LIDAR_MAX_BOXES 160
#define SB_LIDAR_MAX_EDGES 256

struct LBox {
  double x, y, c, s, hw, hl, z0, z1;
};

// lidar.py:74-134: closest hit of every ray against ground plane, vehicle boxes, edge walls.
__global__ void __launch_bounds__(SB_LIDAR_THREADS)
k_lidar(DCfg c, DMap m, DState s, DObs o, const double* __restrict__ rays, float reach_xy) {
  // reach_xy: longest horizontal extent of any table ray (+ margin), for conservative culling
  __shared__ LBox boxes[SB_LIDAR_MAX_BOXES];
  __shared__ double edges[SB_LIDAR_MAX_EDGES][4];
  __shared__ int n_boxes, n_edges;
  const int ai = blockIdx.x;
  const int e = ai / c.A, a = ai - e * c.A, vi = e * c.V + a;
  const double ox = s.pos_x[vi], oy = s.pos_y[vi], oz = s.pos_z[vi] + 1.0;
  if (threadIdx.x == 0) {
    n_boxes = 0;
    n_edges = 0;
  }
  __syncthreads();
  // stage the env's other vehicles (order is irrelevant: the result is a minimum)
  for (int j = threadIdx.x; j < c.V; j += blockDim.x) {
    int vj = e * c.V + j;
    if (j == a || !s.alive[vj]) continue;
    {  // a hit lies within reach_xy of the origin and within the half diagonal of the box centre
      float bx = (float)(s.pos_x[vj] - ox), by = (float)(s.pos_y[vj] - oy);
      float fw = (float)s.width[vj], fl = (float)s.length[vj];
      float lim = reach_xy + 0.5f * sqrtf(fw * fw + fl * fl) + 0.25f;
      if (bx * bx + by * by > lim * lim) continue;
    }
    int k = atomicAdd(&n_boxes, 1);
    if (k < SB_LIDAR_MAX_BOXES) {
      double z0 = s.pos_z[vj] + (j < c.A ? 0.1 : 0.2);
      boxes[k] = LBox{s.pos_x[vj], s.pos_y[vj], s.cos_h[vj], s.sin_h[vj], 0.5 * s.width[vj], 0.5 * s.length[vj],
                      z0,          z0 + s.height[vj]};
    }
  }
  if (c.lidar_wall_height > 0.0) {
    // conservative cull: segments whose bounding circle reaches the 20 m + margin disc
    for (int k = threadIdx.x; k < m.n_edges; k += blockDim.x) {
      const double* ed = m.road_edges + (size_t)k * 4;
      float mx = 0.5f * (float)(ed[0] + ed[2]) - (float)ox, my = 0.5f * (float)(ed[1] + ed[3]) - (float)oy;
      float hx = 0.5f * (float)(ed[2] - ed[0]), hy = 0.5f * (float)(ed[3] - ed[1]);
      float reach = sqrtf(hx * hx + hy * hy) + reach_xy + 0.25f;
      if (mx * mx + my * my > reach * reach) continue;
      int q = atomicAdd(&n_edges, 1);
      if (q < SB_LIDAR_MAX_EDGES) {
        edges[q][0] = ed[0];
        edges[q][1] = ed[1];
        edges[q][2] = ed[2];
        edges[q][3] = ed[3];
      }
    }
  }
  __syncthreads();
  const int nb = n_boxes < SB_LIDAR_MAX_BOXES ? n_boxes : SB_LIDAR_MAX_BOXES;
  const int ne = n_edges < SB_LIDAR_MAX_EDGES ? n_edges : SB_LIDAR_MAX_EDGES;
  if (threadIdx.x == 0 && (n_boxes > SB_LIDAR_MAX_BOXES || n_edges > SB_LIDAR_MAX_EDGES))
    s.status[ai] |= SB_STATUS_LIDAR_OVERFLOW;
  const double inf = __longlong_as_double(0x7ff0000000000000LL);
  const double miss = c.lidar_miss_inf ? inf : 0.0;
  for (int r = threadIdx.x; r < c.lidar_rays; r += blockDim.x) {
    const double dx = rays[r * 3 + 0], dy = rays[r * 3 + 1], dz = rays[r * 3 + 2];
    double tbest = 2.0;
    if (dz < 0.0) {
      double t = (0.0 - oz) / dz;
      if (t >= 0.0 && t <= 1.0) tbest = t;
    }
    for (int k = 0; k < nb; ++k) {
      const LBox b = boxes[k];
      double rx = ox - b.x, ry = oy - b.y;
      double lo[3] = {rx * b.c + ry * b.s, ry * b.c - rx * b.s, oz};
      double ld[3] = {dx * b.c + dy * b.s, dy * b.c - dx * b.s, dz};
      double mn[3] = {-b.hw, -b.hl, b.z0}, mxv[3] = {b.hw, b.hl, b.z1};
      double t0 = -inf, t1 = inf;
      bool ok = true;
#pragma unroll
      for (int q = 0; q < 3; ++q) {
        if (ld[q] == 0.0) {
          if (lo[q] < mn[q] || lo[q] > mxv[q]) ok = false;
        } else {
          double ta = (mn[q] - lo[q]) / ld[q], tb = (mxv[q] - lo[q]) / ld[q];
          if (ta > tb) {
            double t = ta;
            ta = tb;
            tb = t;
          }
          if (ta > t0) t0 = ta;
          if (tb < t1) t1 = tb;
        }
      }
      if (!ok || t0 > t1 || t0 < 0.0 || t0 > 1.0) continue;
      if (t0 < tbest) tbest = t0;
    }
    for (int k = 0; k < ne; ++k) {
      double ex = edges[k][2] - edges[k][0], ey = edges[k][3] - edges[k][1];
      double den = dx * ey - dy * ex;
      if (den == 0.0) continue;
      double qx = edges[k][0] - ox, qy = edges[k][1] - oy;
      double t = (qx * ey - qy * ex) / den;
      double u = (qx * dy - qy * dx) / den;
      if (t < 0.0 || t > 1.0 || u < 0.0 || u > 1.0) continue;
      double zz = oz + t * dz;
      if (zz < 0.0 || zz > c.lidar_wall_height) continue;
      if (t < tbest) tbest = t;
    }
    size_t q = (size_t)ai * c.lidar_rays + r;
    if (tbest <= 1.0) {
      o.lidar_hits[q] = 1;
      o.lidar_points[q * 3 + 0] = ox + tbest * dx;
      o.lidar_points[q * 3 + 1] = oy + tbest * dy;
      o.lidar_points[q * 3 + 2] = oz + tbest * dz;
    } else {
      o.lidar_hits[q] = 0;
      o.lidar_points[q * 3 + 0] = miss;
      o.lidar_points[q * 3 + 1] = miss;
      o.lidar_points[q * 3 + 2] = miss;
    }
  }
}

// --------------------------------------------------------------------------------- raster
// k_raster lives in sb_raster.cuh (row-per-thread spans).  The first-generation version below
// (warp per polygon, int64 divisions, atomicOr) is kept for reference only; profiles/README.md
// has its numbers (1.80 ms/step at 16384 agents, 18 % of HBM peak, barrier-stall bound).
#include "sb_raster.cuh"
#if 0
// Contract (DESIGN.md "Grid maps"): vertices -> pixel space in binary64, snapped to 1/256 px; a
// pixel is set iff its centre satisfies all edge functions of the convex polygon (inclusive),
// in exact integer arithmetic.  Here each (polygon, row) task solves the edge inequalities for
// the covered column span with exact integer division and ORs the span into a 1-bit image in
// shared memory; the image is then expanded to 0/255 bytes with 16-byte streaming stores.
#define SB_RASTER_THREADS 256
#define SB_RASTER_MAX_POLYS 256

struct PixPoly {
  long long U[4], V[4];
  int n;
  long long sg;  // orientation sign, 0 = degenerate
};

struct Cam {
  double px, py, c, s, res, halfw, halfh;
};

__device__ __forceinline__ void d_to_pixel(const Cam& cam, double x, double y, long long* U, long long* V) {
  double dx = x - cam.px, dy = y - cam.py;
  double xr = dx * cam.c + dy * cam.s;
  double yr = dy * cam.c - dx * cam.s;
  double u = xr / cam.res + cam.halfw;
  double v = cam.halfh - yr / cam.res;
  double su = rint(u * 256.0), sv = rint(v * 256.0);
  const double lim = 1073741824.0;
  su = su > lim ? lim : (su < -lim ? -lim : su);
  sv = sv > lim ? lim : (sv < -lim ? -lim : sv);
  *U = (long long)su;
  *V = (long long)sv;
}

__device__ __forceinline__ long long d_floor_div(long long a, long long b) {  // b > 0
  long long q = a / b;
  if ((a % b) != 0 && a < 0) --q;
  return q;
}
__device__ __forceinline__ long long d_ceil_div(long long a, long long b) {  // b > 0
  long long q = a / b;
  if ((a % b) != 0 && a > 0) ++q;
  return q;
}

__device__ __forceinline__ void d_finish_poly(PixPoly& p) {
  long long area2 = 0;
  for (int i = 0; i < p.n; ++i) {
    int j = (i + 1 == p.n) ? 0 : i + 1;
    area2 += p.U[i] * p.V[j] - p.U[j] * p.V[i];
  }
  p.sg = area2 > 0 ? 1 : (area2 < 0 ? -1 : 0);
}

// rasterise the polygons staged in shared memory into the bit image (W/32 words per row)
__device__ void d_fill_polys(const PixPoly* polys, int np, uint32_t* bits, int W, int H) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
  const int wpr = W >> 5;
  for (int p = warp; p < np; p += nwarps) {
    const PixPoly& P = polys[p];
    if (P.sg == 0) continue;
    long long vmin = P.V[0], vmax = P.V[0], umin = P.U[0], umax = P.U[0];
    for (int i = 1; i < P.n; ++i) {
      vmin = P.V[i] < vmin ? P.V[i] : vmin;
      vmax = P.V[i] > vmax ? P.V[i] : vmax;
      umin = P.U[i] < umin ? P.U[i] : umin;
      umax = P.U[i] > umax ? P.U[i] : umax;
    }
    long long r0 = d_ceil_div(vmin - 128, 256), r1 = d_floor_div(vmax - 128, 256);
    if (r0 < 0) r0 = 0;
    if (r1 > H - 1) r1 = H - 1;
    if (d_floor_div(umax - 128, 256) < 0 || d_ceil_div(umin - 128, 256) > W - 1) continue;
    for (long long r = r0 + lane; r <= r1; r += 32) {
      long long py = r * 256 + 128;
      long long c0 = 0, c1 = W - 1;
      bool empty = false;
      for (int i = 0; i < P.n; ++i) {
        int j = (i + 1 == P.n) ? 0 : i + 1;
        long long dxe = P.U[j] - P.U[i], dye = P.V[j] - P.V[i];
        // sg * (dxe*(py-Vi) - dye*(px-Ui)) >= 0   <=>   a*px + b >= 0
        long long aa = -P.sg * dye;
        long long bb = P.sg * (dxe * (py - P.V[i]) + dye * P.U[i]);
        if (aa > 0) {
          long long lo = d_ceil_div(-bb - 128 * aa, 256 * aa);
          if (lo > c0) c0 = lo;
        } else if (aa < 0) {
          long long hi = d_floor_div(bb + 128 * aa, -256 * aa);
          if (hi < c1) c1 = hi;
        } else if (bb < 0) {
          empty = true;
        }
      }
      if (empty || c0 > c1) continue;
      int w0 = (int)(c0 >> 5), w1 = (int)(c1 >> 5);
      uint32_t* row = bits + (int)r * wpr;
      for (int w = w0; w <= w1; ++w) {
        uint32_t msk = 0xffffffffu;
        if (w == w0) msk &= 0xffffffffu << (c0 & 31);
        if (w == w1) msk &= 0xffffffffu >> (31 - (c1 & 31));
        atomicOr(&row[w], msk);
      }
    }
  }
}

__device__ __forceinline__ uint32_t d_expand4(uint32_t b) {  // 4 bits -> 4 bytes of 0/255
  uint32_t x = (b & 1u) | ((b & 2u) << 7) | ((b & 4u) << 14) | ((b & 8u) << 21);
  return x * 255u;
}

// write the H x W byte image from the bit image, 16 pixels (one uint4) per thread per step
__device__ void d_store_image(const uint32_t* bits, uint8_t* dst, int W, int H) {
  const int n16 = (W * H) >> 4;
  uint4* out = reinterpret_cast<uint4*>(dst);
  for (int i = threadIdx.x; i < n16; i += blockDim.x) {
    uint32_t word = bits[i >> 1];
    uint32_t hw = (i & 1) ? (word >> 16) : (word & 0xffffu);
    uint4 v;
    v.x = d_expand4(hw & 15u);
    v.y = d_expand4((hw >> 4) & 15u);
    v.z = d_expand4((hw >> 8) & 15u);
    v.w = d_expand4((hw >> 12) & 15u);
    __stcs(out + i, v);
  }
}

// dynamic shared memory: bit image (max(ogm, dagm) W*H/8 bytes)
__global__ void __launch_bounds__(SB_RASTER_THREADS) k_raster(DCfg c, DMap m, DState s, DObs o) {
  extern __shared__ uint32_t bits[];
  __shared__ PixPoly polys[SB_RASTER_MAX_POLYS];
  __shared__ int n_polys;
  const int ai = blockIdx.x;
  const int e = ai / c.A, a = ai - e * c.A, vi = e * c.V + a;
  const double px = s.pos_x[vi], py = s.pos_y[vi], ch = s.cos_h[vi], sh = s.sin_h[vi];

  if (c.ogm_enable) {  // sensors.py:719-758: vehicles only, ego included
    const int W = c.ogm_w, H = c.ogm_h;
    Cam cam{px, py, ch, sh, c.ogm_res, 0.5 * W, 0.5 * H};
    for (int i = threadIdx.x; i < (W * H) >> 5; i += blockDim.x) bits[i] = 0u;
    for (int base = 0; base < c.V; base += SB_RASTER_MAX_POLYS) {
      int cntp = c.V - base < SB_RASTER_MAX_POLYS ? c.V - base : SB_RASTER_MAX_POLYS;
      __syncthreads();
      for (int t = threadIdx.x; t < cntp; t += blockDim.x) {
        int vj = e * c.V + base + t;
        PixPoly P;
        P.n = 4;
        if (s.alive[vj]) {
          double cc = s.cos_h[vj], ss = s.sin_h[vj];
          double hw = 0.5 * s.width[vj], hl = 0.5 * s.length[vj];
          const double sa[4] = {-hw, hw, hw, -hw}, sbv[4] = {-hl, -hl, hl, hl};
          for (int k = 0; k < 4; ++k) {
            double x = s.pos_x[vj] + (sa[k] * cc - sbv[k] * ss);
            double y = s.pos_y[vj] + (sa[k] * ss + sbv[k] * cc);
            d_to_pixel(cam, x, y, &P.U[k], &P.V[k]);
          }
          d_finish_poly(P);
        } else {
          P.sg = 0;
        }
        polys[t] = P;
      }
      __syncthreads();
      d_fill_polys(polys, cntp, bits, W, H);
    }
    __syncthreads();
    d_store_image(bits, o.ogm + (size_t)ai * W * H, W, H);
    __syncthreads();
  }

  if (c.dagm_enable) {  // sensors.py:675-716: road surface only
    const int W = c.dagm_w, H = c.dagm_h;
    Cam cam{px, py, ch, sh, c.dagm_res, 0.5 * W, 0.5 * H};
    for (int i = threadIdx.x; i < (W * H) >> 5; i += blockDim.x) bits[i] = 0u;
    // conservative world-space cull radius: half diagonal of the film + 1 m
    const float rad = 0.5f * (float)c.dagm_res * sqrtf((float)(W * W + H * H)) + 1.0f;
    const float fx = (float)px, fy = (float)py;
    for (int base = 0; base < m.n_polys; base += blockDim.x) {
      if (threadIdx.x == 0) n_polys = 0;
      __syncthreads();
      int p = base + threadIdx.x;
      if (p < m.n_polys) {
        const float* bb = m.poly_bbox + (size_t)p * 4;
        bool hit = !(bb[0] > fx + rad || bb[2] < fx - rad || bb[1] > fy + rad || bb[3] < fy - rad);
        if (hit) {
          int st = m.poly_start[p], n = m.poly_start[p + 1] - st;
          if (n >= 3 && n <= 4) {
            PixPoly P;
            P.n = n;
            for (int k = 0; k < n; ++k) d_to_pixel(cam, m.poly_x[st + k], m.poly_y[st + k], &P.U[k], &P.V[k]);
            d_finish_poly(P);
            int q = atomicAdd(&n_polys, 1);
            polys[q] = P;  // at most blockDim.x == SB_RASTER_MAX_POLYS per round
          }
        }
      }
      __syncthreads();
      d_fill_polys(polys, n_polys, bits, W, H);
      __syncthreads();
    }
    d_store_image(bits, o.dagm + (size_t)ai * W * H, W, H);
  }
}
#endif  // first-generation rasteriser
