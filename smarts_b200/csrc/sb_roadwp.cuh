// sb_roadwp.cuh -- k_road_waypoints: RoadWaypointsSensor (smarts/core/sensors.py:991-1043), warp per agent.
//
//   reference                                               here
//   nearest_lane(vehicle point) -> road                     d_nearest_lane_warp
//   [road] + road.parallel_roads                            road_par CSR (sumo_road_network.py:607-618)
//        + road.oncoming_roads_at_point(point)              rw_oncoming_roads: per lane of the road the lanes within
//                                                           1.1 lane widths of the ego's projection, nearest first
//                                                           (selection rounds over the segment grid cell), each
//                                                           tested for an opposite direction (:370-396, :595-605)
//   for road: for lane: lanes[lane_id] = paths_for_lane     the visit loop: a road listed twice is walked twice; the
//                                                           dict keeps a key's first position and last value, so
//                                                           only a road's last listing writes paths
//   paths_for_lane: back `horizon` metres along the lane,   explicit stack (SB_RW_MAX_DEPTH) over lane_in CSR; the
//        through incoming lanes when it starts before 0     serial lane geometry runs on lane 0 and is broadcast
//   lane.waypoint_paths_for_pose(start, 2 * horizon, route) wp_closest_in_range + wp_enumerate_chained +
//                                                           wp_equally_spaced_warp of sb_waypoints.cuh
//   SumoRoadNetwork._WaypointsCache (:1229-1276)            rw cache record: query point + (lane index -> lane that
//                                                           filled the entry); see oracle.cpp RwOut
// Bit-identical to oracle.cpp observe_road_waypoints (same binary64 operations in the same order).
// included by sb_kernels.cuh after sb_events.cuh (lane-coordinate helpers) and sb_waypoints.cuh
#pragma once

#define SB_RW_WARPS 2
#define SB_RW_CACHE 16   // entries of the emulated waypoint cache (lanes of one index starting at one point)
#define SB_RW_NEAR 32    // lanes looked at per oncoming_lanes_at_offset query

struct DRoadWp {
  int32_t *n_lanes, *lane, *n_paths, *path_slot, *len, *wp_lane;
  double *pos, *heading;
  float *lane_width, *speed_limit;
  int8_t* lane_index;
};

struct RwScratch {
  int32_t roads[SB_RW_MAX_ROADS];      // [road] + parallel + oncoming, duplicates kept
  int32_t slot_base[SB_RW_MAX_ROADS];  // lane slot of the road's first lane (first listing), -1 for a later listing
  int32_t n_roads;
  int32_t stk_lane[SB_RW_MAX_DEPTH], stk_cur[SB_RW_MAX_DEPTH];
  double stk_start[SB_RW_MAX_DEPTH];
  int32_t cache_key[SB_RW_CACHE], cache_lane[SB_RW_CACHE];
  int32_t near_taken[SB_RW_NEAR];
};

// Road.oncoming_roads_at_point for one lane of the ego's road: appends to sc->roads.  Returns false when a cap was hit.
__device__ bool rw_oncoming_roads(const DMap& m, int road, int l, double px, double py, RwScratch* sc, int* n_roads,
                                  int lane) {
  bool ok = true;
  double off = 0.0, cx = 0.0, cy = 0.0, mvx = 0.0, mvy = 0.0;
  if (lane == 0) {
    off = ev_offset_along_lane(m, l, px, py);  // lane.to_lane_coord(point).s
    ev_from_lane_coord(m, l, off, &cx, &cy);
    ev_vector_at_offset(m, l, off, &mvx, &mvy);
  }
  cx = __shfl_sync(SB_FULL_MASK, cx, 0);
  cy = __shfl_sync(SB_FULL_MASK, cy, 0);
  mvx = __shfl_sync(SB_FULL_MASK, mvx, 0);
  mvy = __shfl_sync(SB_FULL_MASK, mvy, 0);
  const double radius = 1.1 * m.lane_width[l];
  const double my_norm = sqrt(mvx * mvx + mvy * mvy);
  if (my_norm == 0.0) return true;
  const int32_t* items;
  int n;
  d_seg_candidates(m, cx, cy, radius, &items, &n);
  double last_d = -1.0;
  int last_l = -1, n_taken = 0;
  for (;;) {
    // next (distance, lane) pair in ascending order over the segments within the radius
    double bd = 1e300;
    int bl = 0x7fffffff;
    for (int t = lane; t < n; t += 32) {
      const int i = items ? items[t] : t;
      const int sl = m.shape_lane[i];
      if (sl != m.shape_lane[i + 1]) continue;
      const double d = d_dist_point_to_segment(cx, cy, m.shape_x[i], m.shape_y[i], m.shape_x[i + 1], m.shape_y[i + 1]);
      if (!(d < radius)) continue;
      if (!(d > last_d || (d == last_d && sl > last_l))) continue;
      if (d < bd || (d == bd && sl < bl)) {
        bd = d;
        bl = sl;
      }
    }
    warp_argmin(bd, bl);
    if (bl == 0x7fffffff) break;
    last_d = bd;
    last_l = bl;
    bool seen = bl == l;  // `if lane == self: continue`
    for (int t = 0; t < n_taken && !seen; ++t) seen = sc->near_taken[t] == bl;  // a farther segment of a lane already met
    if (seen) continue;
    if (n_taken >= SB_RW_NEAR) {
      ok = false;
      break;
    }
    __syncwarp();
    if (lane == 0) sc->near_taken[n_taken] = bl;
    __syncwarp();
    ++n_taken;
    int opposite = 0;
    if (lane == 0) {
      const double s2 = ev_offset_along_lane(m, bl, cx, cy);
      double lvx, lvy;
      ev_vector_at_offset(m, bl, s2, &lvx, &lvy);
      const double lv_norm = sqrt(lvx * lvx + lvy * lvy);
      if (lv_norm != 0.0) {
        const double ang = (mvx * lvx + mvy * lvy) / (my_norm * lv_norm);
        opposite = ang < -0.995562;  // cos(175 deg), sumo_road_network.py:385
      }
    }
    opposite = __shfl_sync(SB_FULL_MASK, opposite, 0);
    const int oroad = m.lane_road[bl];
    if (opposite && oroad != road) {
      if (*n_roads >= SB_RW_MAX_ROADS) {
        ok = false;
        break;
      }
      if (lane == 0) sc->roads[*n_roads] = oroad;
      ++*n_roads;
    }
  }
  __syncwarp();
  return ok;
}

__global__ void __launch_bounds__(SB_RW_WARPS * 32) k_road_waypoints(DCfg c, DMap m, DState s, DRoadWp o) {
  __shared__ WarpPaths wps[SB_RW_WARPS];
  __shared__ RwScratch scs[SB_RW_WARPS];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  int e, a;
  if (!d_work_agent(c, blockIdx.x * SB_RW_WARPS + warp, &e, &a)) return;
  WarpPaths* wp = &wps[warp];
  RwScratch* sc = &scs[warp];
  const int ai = e * c.A + a, vi = e * c.V + a;
  const int Lc = c.rw_lanes_cap, Pc = c.rw_paths_cap, horizon = c.rw_horizon, Hz = 2 * horizon + 1;
  const double px = s.pos_x[vi], py = s.pos_y[vi];

  // every entry of the agent's buffers is rewritten each call (the oracle's zero fill)
  for (int i = lane; i < Lc; i += 32) o.lane[(size_t)ai * Lc + i] = -1;
  for (int p = lane; p < Pc; p += 32) {
    o.path_slot[(size_t)ai * Pc + p] = -1;
    o.len[(size_t)ai * Pc + p] = 0;
  }
  {
    const size_t base = (size_t)ai * Pc * Hz;
    for (int t = lane; t < Pc * Hz; t += 32) {
      o.pos[(base + t) * 2] = 0.0;
      o.pos[(base + t) * 2 + 1] = 0.0;
      o.heading[base + t] = 0.0;
      o.lane_width[base + t] = 0.f;
      o.speed_limit[base + t] = 0.f;
      o.wp_lane[base + t] = -1;
      o.lane_index[base + t] = 0;
    }
  }
  __syncwarp();

  const int nl = d_nearest_lane_warp(m, px, py, 10.0, lane);  // radius None -> max(10, 2 * 3.2)
  if (nl < 0) {
    if (lane == 0) o.n_lanes[ai] = o.n_paths[ai] = 0;
    return;
  }
  bool truncated = false;
  const int road = m.lane_road[nl];
  int n_roads = 1;
  if (lane == 0) sc->roads[0] = road;
  for (int i = m.road_par_start[road]; i < m.road_par_start[road + 1]; ++i) {
    if (n_roads >= SB_RW_MAX_ROADS) {
      truncated = true;
      break;
    }
    if (lane == 0) sc->roads[n_roads] = m.road_par[i];
    ++n_roads;
  }
  __syncwarp();
  if (!truncated)
    for (int l = m.road_lane_start[road]; l < m.road_lane_start[road] + m.road_lane_count[road]; ++l)
      if (!rw_oncoming_roads(m, road, l, px, py, sc, &n_roads, lane)) {
        truncated = true;  // the oracle cuts the list at SB_RW_MAX_ROADS the same way
        break;
      }
  __syncwarp();
  // lane slots: a road's lanes take consecutive slots at its first listing
  int n_lanes = 0;
  for (int r = 0; r < n_roads; ++r) {
    const int rd = sc->roads[r];
    bool first = true;
    for (int q = 0; q < r && first; ++q) first = sc->roads[q] != rd;
    if (lane == 0) sc->slot_base[r] = first ? n_lanes : -1;
    if (first) {
      const int l0 = m.road_lane_start[rd], cnt = m.road_lane_count[rd];
      for (int k = lane; k < cnt; k += 32)
        if (n_lanes + k < Lc) o.lane[(size_t)ai * Lc + n_lanes + k] = l0 + k;
      n_lanes += cnt;
    }
  }
  __syncwarp();
  if (n_lanes > Lc) truncated = true;

  const int rl = s.route_len[ai];
  const int32_t* filt = rl > 0 ? s.route + (size_t)ai * c.route_cap : nullptr;
  const int nf = rl > 0 ? rl : 0;
  int n_paths = 0;
  bool cache_valid = false;
  double cache_x = 0.0, cache_y = 0.0;
  int n_cache = 0;
  for (int r = 0; r < n_roads; ++r) {
    const int rd = sc->roads[r];
    bool last = true;
    int base = -1;
    for (int q = r + 1; q < n_roads && last; ++q) last = sc->roads[q] != rd;
    for (int q = 0; q <= r && base < 0; ++q)
      if (sc->roads[q] == rd) base = sc->slot_base[q];
    const int l0 = m.road_lane_start[rd], l1 = l0 + m.road_lane_count[rd];
    for (int l = l0; l < l1; ++l) {
      const int slot = base + (l - l0);
      const bool emit = last && slot < Lc;
      // ---- paths_for_lane(l), sensors.py:1013-1040
      int depth = 0;
      {
        double st = 0.0;
        if (lane == 0) st = ev_offset_along_lane(m, l, px, py) - (double)horizon;
        st = __shfl_sync(SB_FULL_MASK, st, 0);
        __syncwarp();
        if (lane == 0) {
          sc->stk_lane[0] = l;
          sc->stk_start[0] = st;
          sc->stk_cur[0] = -1;
        }
        __syncwarp();
      }
      while (depth >= 0) {
        const int cl = sc->stk_lane[depth];
        const double start = sc->stk_start[depth];
        const int cur = sc->stk_cur[depth];
        const int i0 = m.lane_in_start[cl], i1 = m.lane_in_start[cl + 1];
        bool leaf = false;
        if (cur < 0) {  // entering
          if (start < 0 && i1 > i0) {
            if (depth + 1 >= SB_RW_MAX_DEPTH) {
              truncated = true;
              --depth;
              continue;
            }
            __syncwarp();
            if (lane == 0) sc->stk_cur[depth] = i0;
            __syncwarp();
            continue;
          }
          leaf = true;
        }
        if (!leaf) {
          __syncwarp();
          if (cur < i1) {
            const int child = m.lane_in[cur];
            if (lane == 0) {
              sc->stk_cur[depth] = cur + 1;
              sc->stk_lane[depth + 1] = child;
              sc->stk_start[depth + 1] = m.lane_length[child] + start;
              sc->stk_cur[depth + 1] = -1;
            }
            ++depth;
          } else {
            --depth;
          }
          __syncwarp();
          continue;
        }
        // ---- leaf: lane.waypoint_paths_for_pose(from_lane_coord(max(0, start)), 2 * horizon, route)
        --depth;
        double sx = 0.0, sy = 0.0;
        if (lane == 0) ev_from_lane_coord(m, cl, start > 0.0 ? start : 0.0, &sx, &sy);
        sx = __shfl_sync(SB_FULL_MASK, sx, 0);
        sy = __shfl_sync(SB_FULL_MASK, sy, 0);
        if (m.lane_lp_count[cl] <= 0) continue;
        const int key = m.lane_idx_in_road[cl];
        int source = cl;
        bool hit = false;
        if (cache_valid && sx == cache_x && sy == cache_y) {
          for (int t = 0; t < n_cache; ++t)
            if (sc->cache_key[t] == key) {
              source = sc->cache_lane[t];
              hit = true;
            }
        } else {
          cache_valid = true;
          cache_x = sx;
          cache_y = sy;
          n_cache = 0;
        }
        if (!hit) {
          if (n_cache < SB_RW_CACHE) {
            __syncwarp();
            if (lane == 0) {
              sc->cache_key[n_cache] = key;
              sc->cache_lane[n_cache] = cl;
            }
            __syncwarp();
            ++n_cache;
          } else {
            truncated = true;
          }
        }
        if (!emit) continue;
        __syncwarp();
        if (lane == 0) {
          wp->n = 0;
          wp->total = 0;
          wp->overflow = 0;
        }
        __syncwarp();
        const int lp = wp_closest_in_range(m, m.lane_lp_start[source], m.lane_lp_count[source], sx, sy, lane);
        if (m.lp_chained) {
          wp_enumerate_chained(m, lp, 2 * horizon, filt, nf, wp, lane);
        } else if (lane == 0) {
          wp_enumerate_paths(m, lp, 2 * horizon, filt, nf, wp);
        }
        __syncwarp();
        if (wp->overflow) truncated = true;
        const int np_ = wp->n;
        for (int k = 0; k < np_; ++k) {
          const int p = n_paths++;
          if (p >= Pc) {
            truncated = true;
            continue;
          }
          const size_t q = (size_t)ai * Pc + p;
          const size_t base_t = q * Hz;
          const int len = wp_equally_spaced_warp(
              m, wp->lp[k], wp->len[k], sx, sy, Hz,
              [&](int i, const DWaypoint& w) {
                const size_t t = base_t + i;
                o.pos[t * 2] = w.x;
                o.pos[t * 2 + 1] = w.y;
                o.heading[t] = w.heading;
                o.lane_width[t] = (float)w.lane_width;
                o.speed_limit[t] = (float)w.speed_limit;
                o.wp_lane[t] = w.lane;
                o.lane_index[t] = (int8_t)w.lane_index;
              },
              wp, lane);
          if (lane == 0) {
            o.path_slot[q] = slot;
            o.len[q] = len;
          }
        }
        __syncwarp();
      }
    }
  }
  if (lane == 0) {
    o.n_lanes[ai] = n_lanes;
    o.n_paths[ai] = n_paths;
    if (truncated) s.status[ai] |= SB_STATUS_RW_TRUNCATED;
  }
}
