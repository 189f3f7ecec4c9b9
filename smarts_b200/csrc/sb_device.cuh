// sb_device.cuh -- device-side data layout and scalar helpers shared by all kernels.
//
// Numeric contract: binary64, no FMA contraction (this translation unit is compiled with
// -fmad=false), transcendental functions from include/sb_detmath.h only.  See DESIGN.md.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/sb_detmath.h"
#include "../../include/smarts_b200.h"

#define SB_FULL_MASK 0xffffffffu

// Device copy of SbMap (same field meaning; plus shape_lane = lane of every shape vertex and
// world-space AABBs of the drivable-area polygons for culling).
struct DMap {
  int32_t n_roads, n_lanes, n_lanepoints, n_shape_pts, n_polys, n_poly_verts, n_edges;
  int32_t lp_chained;  // every lane point but the last of its lane has the single successor index + 1
  int32_t road_lp_tiled;  // a road's lane-point range is its lanes' ranges back to back, in lane order
  // uniform grid over the lane points (wp_closest_with_pose): row-major cells of lpg_cell metres, items = lane
  // point indices sorted by cell (ascending inside a cell); lpg_nx == 0: no grid
  double lpg_x0, lpg_y0, lpg_cell, lpg_inv_cell;
  int32_t lpg_nx, lpg_ny;
  const int32_t *lpg_cell_start, *lpg_items;
  const double *lpg_px, *lpg_py;  // lp_x / lp_y of the items, in item order (no dependent gather)
  double lanepoint_spacing;
  const int32_t *road_lane_start, *road_lane_count, *road_lp_start, *road_lp_count, *road_first_out;
  const uint8_t* road_is_junction;
  const int32_t *lane_road, *lane_idx_in_road, *lane_out_start, *lane_out, *lane_lp_start, *lane_lp_count,
      *lane_shape_start;
  const double *lane_width, *lane_speed, *lane_length;
  const double *lp_x, *lp_y, *lp_heading;
  const uint8_t* lp_inferred;
  const int32_t *lp_lane, *lp_next_start, *lp_next;
  const double *shape_x, *shape_y;
  const int32_t* shape_lane;
  const int32_t* poly_start;
  const double *poly_x, *poly_y;
  const float* poly_bbox;  // [n_polys,4] xmin,ymin,xmax,ymax rounded outwards
  const double* road_edges;
  // Uniform grid over the lane-centre segments (built in sb_upload_map): cell (cx, cy) lists every
  // segment whose bounding box comes within seg_rmax of the cell, so a nearest-lane query of
  // radius <= seg_rmax reads one list.  seg_nx == 0: no grid (small maps scan everything).
  double seg_x0, seg_y0, seg_inv_cell, seg_rmax;
  int32_t seg_nx, seg_ny;
  const int32_t *seg_cell_start, *seg_items;
  const double* road_lp_bbox;  // [n_roads,4] xmin, ymin, xmax, ymax of the road's lane points
  const int32_t *lane_in_start, *lane_in, *road_par_start, *road_par;  // RoadWaypoints tables (SbMap)
};

struct DState {
  double *pos_x, *pos_y, *pos_z, *heading, *speed, *yaw_rate, *length, *width, *height, *cos_h, *sin_h;
  uint8_t* alive;
  double *steering, *v_long, *v_lat, *ctrl, *last_control;
  int32_t *route, *route_len;
  uint32_t* status;
  // [E*A] 1: a control kernel has moved this agent since its last events pass (SensorState.step is due); k_events
  // consumes it.  Per agent, not per handle: observing some envs must not use up the tick of the others.
  uint8_t* ticked;
};

// ---- scalar restatements (reference file:line in the comments) -------------------------------
__device__ __forceinline__ double d_clip(double v, double lo, double hi) { return v < lo ? lo : (v > hi ? hi : v); }
// utils/math.py:206-216
__device__ __forceinline__ double d_lerp(double a, double b, double p) { return a * (1.0 - p) + b * p; }
// coordinates.py:175-181
__device__ __forceinline__ double d_wrap_heading(double h) {
  double v = sb_pymod(h, SB_TWO_PI);
  if (v > SB_PI) v -= SB_TWO_PI;
  return v;
}
// utils/math.py:447-449
__device__ __forceinline__ double d_angdiff(double a, double b) { return sb_pymod((a - b) + SB_PI, SB_TWO_PI) - SB_PI; }
// utils/math.py:247-253
__device__ __forceinline__ void d_radians_to_vec(double r, double* vx, double* vy) {
  double a = sb_pymod(r + SB_PI * 0.5, SB_TWO_PI);
  sb_sincos(a, vy, vx);
}
__device__ __forceinline__ double d_degrees(double r) { return r * (180.0 / SB_PI); }
// utils/math.py:219-244
__device__ __forceinline__ double d_low_pass(double in, double st, double k, double dt, double lo, double raw) {
  st += dt * k * (in - st);
  return d_clip(st + raw, lo, 1.0);
}
// utils/math.py:163-185
__device__ __forceinline__ double d_signed_dist_to_line(double px, double py, double lx, double ly, double dx,
                                                        double dy) {
  double p2x = lx + dx, p2y = ly + dy;
  double u = fabs(((dy * px - dx * py) + p2x * ly) - p2y * lx);
  double d = u / sqrt(dx * dx + dy * dy);
  double dot = (px - lx) * (-dy) + (py - ly) * dx;
  double sg = dot > 0 ? 1.0 : (dot < 0 ? -1.0 : 0.0);
  return d * sg;
}

// lexicographic (value, index) minimum across the warp; every lane gets the result.  Values are distances
// (>= 0, or the 1e300 "nothing" sentinel) and indices are >= 0, so both order like their bit patterns as
// unsigned integers: three redux.sync minima (high word, low word among the high-word winners, index among
// the value winners) replace fifteen shuffles with their compares.
__device__ __forceinline__ void warp_argmin(double& d, int& i) {
  const unsigned long long b = d == 0.0 ? 0ull : (unsigned long long)__double_as_longlong(d);
  const unsigned hi = (unsigned)(b >> 32), lo = (unsigned)b;
  const unsigned mhi = __reduce_min_sync(SB_FULL_MASK, hi);
  const unsigned mlo = __reduce_min_sync(SB_FULL_MASK, hi == mhi ? lo : 0xffffffffu);
  const bool win = hi == mhi && lo == mlo;
  const unsigned mi = __reduce_min_sync(SB_FULL_MASK, win ? (unsigned)i : 0xffffffffu);
  d = __longlong_as_double((long long)(((unsigned long long)mhi << 32) | mlo));
  i = (int)mi;
}
