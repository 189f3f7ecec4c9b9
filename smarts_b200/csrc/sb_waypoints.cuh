// sb_waypoints.cuh -- warp-cooperative RoadMap.waypoint_paths
// (smarts/core/sumo_road_network.py:815-882, :1278-1437; smarts/core/lanepoints.py:517-692).
//
// One warp serves one agent.  The three KD-tree queries of the reference become strided argmin
// scans over index ranges of the lane-point arrays (all 32 lanes), the branch enumeration is a
// depth-first walk on lane 0 (same leaf order as the reference's level-by-level expansion), and
// the equal-arc-length re-spacing runs one path per lane.
#pragma once

#include "sb_device.cuh"

#define SB_WP_CAND 64
#define SB_LPG_RINGS 3  // lane-point grid: squares of up to 3 cells around the query before falling back to the table

// LA = the longest lookahead the kernel asks for: the per-warp record is sized by it (the controller's 16-point query
// needs 2.6 KB where the sensors' 64-point one needs 7.4 KB, and the record's size bounds the resident blocks)
template <int LA>
struct WarpPathsT {
  int32_t lp[SB_MAX_PATHS][LA + 1];  // lane-point indices of each path
  int32_t len[SB_MAX_PATHS];
  int32_t n;         // paths stored
  int32_t total;     // paths found
  int32_t overflow;  // more than SB_MAX_PATHS
  int32_t jr[2];     // junction route of _resolve_in_junction
  // lane-level walk stack of wp_enumerate_chained (depth counts lanes, not lane points)
  int32_t seg_first[LA + 1], seg_cnt[LA + 1], seg_before[LA + 1], seg_cursor[LA + 1];
  uint8_t seg_had[LA + 1];
  // candidate set of wp_closest_with_pose (lane points within a small disc around the nearest one)
  double cand_d[SB_WP_CAND];
  int32_t cand_i[SB_WP_CAND];
  // wp_equally_spaced_warp: arc length and heading-unwrap offset of the kept lane points, and their indices
  double es_cum[LA + 1], es_acc[LA + 1];
  int32_t es_kept[LA + 1];
};
using WarpPaths = WarpPathsT<SB_MAX_LOOKAHEAD>;

struct DWaypoint {
  double x, y, heading, lane_width, speed_limit;
  int32_t lane, lane_index;
};

__device__ __forceinline__ double wp_sqdist(const DMap& m, int i, double x, double y) {
  double dx = m.lp_x[i] - x, dy = m.lp_y[i] - y;
  return dx * dx + dy * dy;
}

// KD-tree k=1 over [start, start+count): lowest index wins ties.  All lanes return the result.
__device__ int wp_closest_in_range(const DMap& m, int start, int count, double x, double y, int lane) {
  double bd = 1e300;
  int bi = 0x7fffffff;
  for (int i = start + lane; i < start + count; i += 32) {
    double d = wp_sqdist(m, i, x, y);
    if (d < bd) {
      bd = d;
      bi = i;
    }
  }
  warp_argmin(bd, bi);
  return bi == 0x7fffffff ? -1 : bi;
}

// lanepoints.py:526-623: closest_lanepoints([pose], within_radius, maximum_count=10)[0].
// The KD-tree's k = 10 query is the 10 smallest (distance, index) pairs.  One scan finds the nearest
// point, a second one gathers every point within 4 m more than that into shared memory (ballot
// compaction); when that set holds at least k points the k smallest are among them, and the ten
// selection rounds run on <= 64 candidates instead of the whole table.  Otherwise (sparse or very
// dense neighbourhood) the rounds scan the table, as before.  Same distances, same ties: same result.
template <class WP>
__device__ int wp_closest_with_pose(const DMap& m, double x, double y, double h, double within_radius, int lane,
                                    WP* wp) {
  const int P = m.n_lanepoints;
  const int k = P < 10 ? P : 10;
  int my_idx = -1;
  double my_d2 = 0.0;
  int n = 0;
  int n_cand = -1;  // < 0: scan the table
  bool gathered = false;
  if (m.lpg_nx > 0) {
    // Lane-point grid: the nearest point is final once it is closer than the scanned square's margin (every
    // unscanned point is at least r cells away); the gather then visits the cells the reach disc can touch.
    // Same candidate set as the two table scans below -- which remain the path for far-away queries.
    const double gx = (x - m.lpg_x0) * m.lpg_inv_cell, gy = (y - m.lpg_y0) * m.lpg_inv_cell;
    if (fabs(gx) < 1e6 && fabs(gy) < 1e6) {
      const int qx = (int)floor(gx), qy = (int)floor(gy);
      bool have = false;
      double bd = 0.0;
      for (int r = 1; r <= SB_LPG_RINGS && !have; ++r) {
        double d = 1e300;
        int bi = 0x7fffffff;
        const int cx0 = max(qx - r, 0), cx1 = min(qx + r, m.lpg_nx - 1);
        if (cx0 <= cx1)
          for (int cy = max(qy - r, 0); cy <= min(qy + r, m.lpg_ny - 1); ++cy) {
            const int a = m.lpg_cell_start[cy * m.lpg_nx + cx0], b = m.lpg_cell_start[cy * m.lpg_nx + cx1 + 1];
            for (int t = a + lane; t < b; t += 32) {
              const int i = m.lpg_items[t];
              const double ddx = m.lpg_px[t] - x, ddy = m.lpg_py[t] - y;  // lp_x / lp_y in item order
              const double dd = ddx * ddx + ddy * ddy;
              if (dd < d || (dd == d && i < bi)) {
                d = dd;
                bi = i;
              }
            }
          }
        warp_argmin(d, bi);
        const double lim = (double)r * m.lpg_cell;
        if (bi != 0x7fffffff && d < lim * lim * (1.0 - 1e-9)) {
          have = true;
          bd = d;
        }
      }
      const double reach = sqrt(bd) + 4.0, T = reach * reach;
      const int R2 = (int)floor((reach * (1.0 + 1e-9) + 1e-9) * m.lpg_inv_cell) + 1;
      if (have && R2 <= 2 * SB_LPG_RINGS) {
        int cnt = 0;
        const int cx0 = max(qx - R2, 0), cx1 = min(qx + R2, m.lpg_nx - 1);
        if (cx0 <= cx1)
          for (int cy = max(qy - R2, 0); cy <= min(qy + R2, m.lpg_ny - 1); ++cy) {
            const int a = m.lpg_cell_start[cy * m.lpg_nx + cx0], b = m.lpg_cell_start[cy * m.lpg_nx + cx1 + 1];
            for (int t0 = a; t0 < b; t0 += 32) {
              const int t = t0 + lane;
              double d = 0.0;
              int i = 0;
              bool in = false;
              if (t < b) {
                i = m.lpg_items[t];
                const double ddx = m.lpg_px[t] - x, ddy = m.lpg_py[t] - y;
                d = ddx * ddx + ddy * ddy;
                in = d <= T;
              }
              const unsigned bal = __ballot_sync(SB_FULL_MASK, in);
              const int pos = cnt + __popc(bal & ((1u << lane) - 1u));
              if (in && pos < SB_WP_CAND) {
                wp->cand_d[pos] = d;
                wp->cand_i[pos] = i;
              }
              cnt += __popc(bal);
            }
          }
        __syncwarp();
        if (cnt >= k && cnt <= SB_WP_CAND) n_cand = cnt;
        gathered = true;
      }
    }
  }
  if (!gathered) {
    double bd = 1e300;
    int bi = 0x7fffffff;
    for (int i = lane; i < P; i += 32) {
      double d = wp_sqdist(m, i, x, y);
      if (d < bd) {
        bd = d;
        bi = i;
      }
    }
    warp_argmin(bd, bi);
    if (bi == 0x7fffffff) return -1;
    const double reach = sqrt(bd) + 4.0, T = reach * reach;
    int cnt = 0;
    for (int i0 = 0; i0 < P; i0 += 32) {
      const int i = i0 + lane;
      double d = 0.0;
      bool in = false;
      if (i < P) {
        d = wp_sqdist(m, i, x, y);
        in = d <= T;
      }
      const unsigned bal = __ballot_sync(SB_FULL_MASK, in);
      const int pos = cnt + __popc(bal & ((1u << lane) - 1u));
      if (in && pos < SB_WP_CAND) {
        wp->cand_d[pos] = d;
        wp->cand_i[pos] = i;
      }
      cnt += __popc(bal);
    }
    __syncwarp();
    if (cnt >= k && cnt <= SB_WP_CAND) n_cand = cnt;
  }
  double last_d = -1.0;
  int last_i = -1;
  for (int r = 0; r < k; ++r) {
    double bd = 1e300;
    int bi = 0x7fffffff;
    if (n_cand >= 0) {
      for (int t = lane; t < n_cand; t += 32) {
        const double d = wp->cand_d[t];
        const int i = wp->cand_i[t];
        const bool after = (d > last_d) || (d == last_d && i > last_i);
        if (r > 0 && !after) continue;
        if (d < bd || (d == bd && i < bi)) {
          bd = d;
          bi = i;
        }
      }
    } else {
      for (int i = lane; i < P; i += 32) {
        double d = wp_sqdist(m, i, x, y);
        bool after = (d > last_d) || (d == last_d && i > last_i);
        if (r > 0 && !after) continue;
        if (d < bd) {
          bd = d;
          bi = i;
        }
      }
    }
    warp_argmin(bd, bi);
    if (bi == 0x7fffffff) break;
    if (lane == n) {  // lane r keeps the r-th nearest point
      my_idx = bi;
      my_d2 = bd;
    }
    ++n;
    last_d = bd;
    last_i = bi;
  }
  __syncwarp();
  // min over the kept points of (squared distance + |heading difference|), the first one on ties
  // (the sort key of lanepoints.py:582-589): lane r scores the r-th nearest, one (key, rank) arg-min picks
  const double r2 = within_radius * within_radius;
  double key = 1e300;
  int rank = 0x7fffffff;
  if (lane < n && !(within_radius >= 0.0 && !(my_d2 <= r2 || lane == 0))) {
    const double rel = d_wrap_heading(h - m.lp_heading[my_idx]);
    key = my_d2 + fabs(rel);
    rank = lane;
  }
  warp_argmin(key, rank);
  if (rank == 0x7fffffff) return -1;
  return __shfl_sync(SB_FULL_MASK, my_idx, rank);
}

__device__ __forceinline__ bool wp_in_route(const int32_t* route, int n, int road) {
  for (int i = 0; i < n; ++i)
    if (route[i] == road) return true;
  return false;
}

// lanepoints.py:646-692 as a depth-first walk (lane 0 only).  Appends to wpaths.
template <class WP>
__device__ void wp_enumerate_paths(const DMap& m, int start, int lookahead, const int32_t* filt, int nf,
                                   WP* wp) {
  int32_t node[SB_MAX_LOOKAHEAD + 1];
  int32_t cursor[SB_MAX_LOOKAHEAD + 1];
  uint8_t had_child[SB_MAX_LOOKAHEAD + 1];
  int d = 0, memo_lane = -1;
  bool memo_ok = false;
  node[0] = start;
  cursor[0] = m.lp_next_start[start];
  had_child[0] = 0;
  while (d >= 0) {
    bool emit = false;
    if (d == lookahead) {
      emit = true;
    } else {
      bool found = false;
      int end = m.lp_next_start[node[d] + 1];
      while (cursor[d] < end) {
        int nlp = m.lp_next[cursor[d]++];
        int nlane = m.lp_lane[nlp];
        if (nf > 0) {
          // the filter is a function of the lane only: successive lane points of one lane reuse it
          if (nlane != memo_lane) {
            int road = m.lane_road[nlane];
            bool ok = wp_in_route(filt, nf, road);
            if (ok && road != filt[nf - 1]) {
              bool any = false;
              for (int o = m.lane_out_start[nlane]; o < m.lane_out_start[nlane + 1]; ++o)
                if (wp_in_route(filt, nf, m.lane_road[m.lane_out[o]])) any = true;
              ok = any;
            }
            memo_lane = nlane;
            memo_ok = ok;
          }
          if (!memo_ok) continue;
        }
        had_child[d] = 1;
        ++d;
        node[d] = nlp;
        cursor[d] = m.lp_next_start[nlp];
        had_child[d] = 0;
        found = true;
        break;
      }
      if (found) continue;
      if (!had_child[d]) emit = true;
    }
    if (emit) {
      wp->total++;
      if (wp->n < SB_MAX_PATHS) {
        int k = wp->n++;
        wp->len[k] = d + 1;
        for (int i = 0; i <= d; ++i) wp->lp[k][i] = node[i];
      } else {
        wp->overflow = 1;
      }
    }
    --d;
  }
}

// route filter of paths_starting_at_lanepoint (lanepoints.py:665-679) for a lane
__device__ __forceinline__ bool wp_lane_ok(const DMap& m, const int32_t* filt, int nf, int nlane) {
  const int road = m.lane_road[nlane];
  if (!wp_in_route(filt, nf, road)) return false;
  if (road == filt[nf - 1]) return true;
  for (int o = m.lane_out_start[nlane]; o < m.lane_out_start[nlane + 1]; ++o)
    if (wp_in_route(filt, nf, m.lane_road[m.lane_out[o]])) return true;
  return false;
}

// Same enumeration on a "chained" map (every lane point but the last of its lane has exactly one
// successor, the next index -- checked on the host at upload; true of every map built by
// road_map.py): a path is a run of consecutive lane points per lane, so the walk branches per
// LANE (2-4 steps for a 16 m lookahead) instead of per lane point.  Leaf order and leaf rule are
// those of the point-level walk above.  Executed by the whole warp uniformly; the lanes share
// the work of writing a path out.
template <class WP>
__device__ void wp_enumerate_chained(const DMap& m, int start, int lookahead, const int32_t* filt, int nf,
                                     WP* wp, int lane) {
  int n = wp->n, total = wp->total, overflow = wp->overflow;
  const int need = lookahead + 1;
  int d = 0;
  bool entering = true;
  wp->seg_first[0] = start;
  wp->seg_before[0] = 0;
  while (d >= 0) {
    bool emit = false;
    const int first = wp->seg_first[d], before = wp->seg_before[d];
    if (entering) {
      const int l = m.lp_lane[first];
      const int avail = m.lane_lp_start[l] + m.lane_lp_count[l] - first;
      int take = min(avail, need - before);
      // the start lane itself is never filtered, but its next lane point (same lane) is
      const bool blocked = d == 0 && nf > 0 && avail > 1 && need > 1 && !wp_lane_ok(m, filt, nf, l);
      if (blocked) take = 1;
      wp->seg_cnt[d] = take;
      entering = false;
      if (blocked || before + take == need) {
        emit = true;
      } else {
        wp->seg_cursor[d] = m.lp_next_start[first + take - 1];
        wp->seg_had[d] = 0;
      }
    }
    if (!emit) {
      const int cnt = wp->seg_cnt[d];
      const int endc = m.lp_next_start[first + cnt];
      int cur = wp->seg_cursor[d];
      bool found = false;
      while (cur < endc) {
        const int child = m.lp_next[cur++];
        if (nf > 0 && !wp_lane_ok(m, filt, nf, m.lp_lane[child])) continue;
        found = true;
        wp->seg_cursor[d] = cur;
        wp->seg_had[d] = 1;
        ++d;
        wp->seg_first[d] = child;
        wp->seg_before[d] = before + cnt;
        entering = true;
        break;
      }
      if (found) continue;
      if (!wp->seg_had[d]) emit = true;
    }
    if (emit) {
      ++total;
      if (n < SB_MAX_PATHS) {
        const int len = wp->seg_before[d] + wp->seg_cnt[d];
        for (int j = lane; j < len; j += 32) {
          int i = 0;
          while (i < d && j >= wp->seg_before[i] + wp->seg_cnt[i]) ++i;
          wp->lp[n][j] = wp->seg_first[i] + (j - wp->seg_before[i]);
        }
        if (lane == 0) wp->len[n] = len;
        ++n;
      } else {
        overflow = 1;
      }
    }
    --d;
  }
  __syncwarp();
  if (lane == 0) {
    wp->n = n;
    wp->total = total;
    wp->overflow = overflow;
  }
}

// Closest lane point of every lane of `road` -> per_lane[0 .. lanes) (-1 for a lane without points); returns
// the closest of the whole road.  On a tiled map (DMap::road_lp_tiled) that is exactly the scan of the road's
// own range: the lane ranges follow each other in index order, and ties keep the earlier (lower-index) lane.
__device__ int wp_closest_per_lane(const DMap& m, int road, double x, double y, int lane, int32_t* per_lane) {
  const int l0 = m.road_lane_start[road], l1 = l0 + m.road_lane_count[road];
  int best = -1;
  double bd = 0.0;
  for (int l = l0; l < l1; ++l) {
    const int cnt = m.lane_lp_count[l];
    int lp = -1;
    if (cnt > 0) {
      lp = wp_closest_in_range(m, m.lane_lp_start[l], cnt, x, y, lane);
      const double d = wp_sqdist(m, lp, x, y);
      if (best < 0 || d < bd) {
        best = lp;
        bd = d;
      }
    }
    if (lane == 0) per_lane[l - l0] = lp;
  }
  __syncwarp();
  return best;
}

// Lane._waypoint_paths_at for every lane of `road` (sumo_road_network.py:429-446, :838-839, :880-881).
// `per_lane`: closest lane point of each lane when the caller has already scanned this road, else null.
template <class WP>
__device__ void wp_paths_on_road(const DMap& m, int road, double x, double y, int lookahead, const int32_t* filt,
                                 int nf, WP* wp, int lane, const int32_t* per_lane = nullptr) {
  int l0 = m.road_lane_start[road], l1 = l0 + m.road_lane_count[road];
  for (int l = l0; l < l1; ++l) {
    int cnt = m.lane_lp_count[l];
    if (cnt <= 0) continue;
    int lp = per_lane ? per_lane[l - l0] : wp_closest_in_range(m, m.lane_lp_start[l], cnt, x, y, lane);
    if (m.lp_chained) {
      wp_enumerate_chained(m, lp, lookahead, filt, nf, wp, lane);
    } else if (lane == 0) {
      wp_enumerate_paths(m, lp, lookahead, filt, nf, wp);
    }
    __syncwarp();
  }
}

// RoadMap.waypoint_paths: fills wp (shared memory, one per warp).  route_len < 0: no route
// object; == 0: empty Route (-> _resolve_in_junction).
template <class WP>
__device__ void wp_waypoint_paths(const DMap& m, double x, double y, double h, int lookahead, double within_radius,
                                  const int32_t* route, int route_len, WP* wp, int lane) {
  if (lane == 0) {
    wp->n = 0;
    wp->total = 0;
    wp->overflow = 0;
  }
  __syncwarp();
  const int32_t* roads = route;
  int nr = route_len;
  if (route_len == 0) {
    int lp = wp_closest_with_pose(m, x, y, h, -1.0, lane, wp);
    nr = 0;
    if (lp >= 0) {
      int road = m.lane_road[m.lp_lane[lp]];
      if (m.road_is_junction[road]) {
        int fo = m.road_first_out[road];
        if (lane == 0) {
          wp->jr[0] = road;
          wp->jr[1] = fo;
        }
        nr = fo >= 0 ? 2 : 1;
      }
    }
    __syncwarp();
    roads = wp->jr;
  }
  if (route_len >= 0 && nr > 0) {
    // The reference queries one KD-tree per route road and keeps the first road with the smallest
    // distance.  Roads whose lane-point bounding box is further away than a distance already achieved
    // cannot win (nor tie), so they are not scanned: squared box distances go to shared memory, the road
    // with the smallest one is scanned first for an upper bound, then the roads are visited in order.
    double ub = 1e300;
    const bool prune = nr > 1 && nr <= SB_WP_CAND;
    // the road scanned for the bound is scanned lane by lane and remembered: the loop below reuses its result
    // and, when it wins, so does wp_paths_on_road -- one pass over its lane points instead of three
    int cached_road = -1, cached_lp = -1;
    if (!prune && nr == 1 && m.road_lp_tiled && m.road_lp_count[roads[0]] > 0) {
      cached_road = roads[0];
      cached_lp = wp_closest_per_lane(m, cached_road, x, y, lane, wp->cand_i);
    }
    if (prune) {
      double lb_best = 1e300;
      int r_best = 0x7fffffff;
      for (int r = lane; r < nr; r += 32) {
        const int road = roads[r];
        double lb = 1e300;
        if (m.road_lp_count[road] > 0) {
          const double* bb = m.road_lp_bbox + (size_t)road * 4;
          const double ex = fmax(fmax(bb[0] - x, x - bb[2]), 0.0), ey = fmax(fmax(bb[1] - y, y - bb[3]), 0.0);
          lb = ex * ex + ey * ey;
        }
        wp->cand_d[r] = lb;
        if (lb < lb_best) {
          lb_best = lb;
          r_best = r;
        }
      }
      warp_argmin(lb_best, r_best);
      __syncwarp();
      if (r_best != 0x7fffffff && lb_best < 1e299) {
        const int road = roads[r_best];
        int lp;
        if (m.road_lp_tiled) {
          lp = wp_closest_per_lane(m, road, x, y, lane, wp->cand_i);
          cached_road = road;
          cached_lp = lp;
        } else {
          lp = wp_closest_in_range(m, m.road_lp_start[road], m.road_lp_count[road], x, y, lane);
        }
        ub = wp_sqdist(m, lp, x, y) * (1.0 + 1e-9);  // guard: bounds and distances round independently
      }
    }
    int best = -1;
    double bd = 0.0;
    for (int r = 0; r < nr; ++r) {
      int road = roads[r];
      int cnt = m.road_lp_count[road];
      if (cnt <= 0) continue;
      if (prune && wp->cand_d[r] > ub) continue;
      int lp = road == cached_road ? cached_lp : wp_closest_in_range(m, m.road_lp_start[road], cnt, x, y, lane);
      double d = sqrt(wp_sqdist(m, lp, x, y));
      if (best < 0 || d < bd) {
        best = lp;
        bd = d;
      }
    }
    if (best >= 0) {
      const int on = m.lane_road[m.lp_lane[best]];
      wp_paths_on_road(m, on, x, y, lookahead, roads, nr, wp, lane, on == cached_road ? wp->cand_i : nullptr);
      return;
    }
  }
  int lp = wp_closest_with_pose(m, x, y, h, within_radius, lane, wp);
  if (lp < 0) return;
  wp_paths_on_road(m, m.lane_road[m.lp_lane[lp]], x, y, lookahead, nullptr, 0, wp, lane);
}

// rank of path k after sorted(paths, key=first waypoint's lane_index) (stable)
template <class WP>
__device__ __forceinline__ int wp_rank(const DMap& m, const WP* wp, int k) {
  int key = m.lane_idx_in_road[m.lp_lane[wp->lp[k][0]]];
  int rank = 0;
  for (int j = 0; j < wp->n; ++j) {
    int kj = m.lane_idx_in_road[m.lp_lane[wp->lp[j][0]]];
    if (kj < key || (kj == key && j < k)) ++rank;
  }
  return rank;
}

// rise / run of one interpolation interval (run > 0: cum is strictly increasing across it).  A zero
// rise -- y, heading, width and speed limit along any straight constant-width lane -- gives a zero
// slope of the same sign, exactly what the division returns, without going through the division's
// out-of-range slow path (ptxas sends a zero numerator there: ~75 instructions per quotient).
__device__ __forceinline__ double wp_slope(double rise, double run) {
  const bool flat = rise == 0.0;
  double num = flat ? 1.0 : rise;
  asm volatile("" : "+d"(num));  // opaque: otherwise nvcc folds the two selects back into rise / run
  const double q = num / run;
  return flat ? rise : q;
}

// First waypoint of wp_equally_spaced_warp() for `path` without building the rest (find_current_lane,
// lane_following_controller.py:367-374, only needs this one): the vehicle's projection on the
// first lane point's heading line, or the lane point itself for a path of <= spacing points.
__device__ void wp_first_point(const DMap& m, const int32_t* path, int n, double vx, double vy, double* x, double* y) {
  int mm = 0;  // kept lane points, counted only until the test below is decided
  for (int i = 0; i < n && (double)mm <= m.lanepoint_spacing; ++i)
    if (!(m.lp_inferred[path[i]] && 0 < i && i < n - 1)) ++mm;
  const int lp0 = path[0];
  const double lx = m.lp_x[lp0], ly = m.lp_y[lp0];
  if ((double)mm <= m.lanepoint_spacing) {
    *x = lx;
    *y = ly;
    return;
  }
  double hvx, hvy;
  d_radians_to_vec(m.lp_heading[lp0], &hvx, &hvy);
  const double proj = (vx - lx) * hvx + (vy - ly) * hvy;
  *x = lx + proj * hvx;
  *y = ly + proj * hvy;
}

// _equally_spaced_path (sumo_road_network.py:1312-1437) for ONE path, executed by the whole warp (all
// arguments warp-uniform).  Calls emit(i, waypoint) for the first min(n_out, n) waypoints, each from the
// lane that computed it, and returns the path's waypoint count.
//   reference (serial)                                  here
//   drop inferred interior lane points                  ballot compaction of the kept indices
//   inplace_unwrap of the headings (utils/math.py:537)  per-lane correction; lane 0 sums the offsets in order
//   cumulative arc length                               per-lane segment length; lane 0 sums in order
//   np.interp at i * step for x, y, heading, width,     lane i interpolates waypoint i (i + 32, ...): the
//   speed; lane of the last point with cum < s          interval is found by bisection on the arc lengths
// The two running sums are the only order-dependent roundings, so they stay serial; everything else is the
// reference's operation on the same operands, hence bit-identical to the oracle's serial restatement:
//   * j  (last kept point with cum[j] <= s) and jd (last with cum[jd] < s, else 0) do not depend on
//     the previous waypoint because cum and s are both non-decreasing;
//   * rh[i] = lp_heading + acc[i] is formed where it is used (same single addition).
template <class Emit, class WP>
__device__ int wp_equally_spaced_warp(const DMap& m, const int32_t* path, int n, double vx, double vy, int n_out,
                                      Emit emit, WP* wp, int lane) {
  double* cum = wp->es_cum;
  double* acc = wp->es_acc;
  int32_t* kept = wp->es_kept;
  __syncwarp();  // scratch may still be read by the previous path's lanes
  int mm = 0;
  for (int base = 0; base < n; base += 32) {
    const int i = base + lane;
    bool keep = false;
    int lp = 0;
    if (i < n) {
      lp = path[i];
      keep = !(m.lp_inferred[lp] && 0 < i && i < n - 1);
    }
    const unsigned b = __ballot_sync(SB_FULL_MASK, keep);
    if (keep) kept[mm + __popc(b & ((1u << lane) - 1u))] = lp;
    mm += __popc(b);
  }
  __syncwarp();
  const int lp0 = path[0];
  const double lx = m.lp_x[lp0], ly = m.lp_y[lp0], rh0 = m.lp_heading[lp0];
  double hvx, hvy;
  d_radians_to_vec(rh0, &hvx, &hvy);
  const double proj = (vx - lx) * hvx + (vy - ly) * hvy;
  const double x0 = lx + proj * hvx, y0 = ly + proj * hvy;  // rx[0], ry[0]
  // per kept point i >= 1: segment length to its predecessor and the inplace_unwrap correction
  bool any_corr = false;
  for (int i = lane; i < mm; i += 32) {
    double d = 0.0, corr = 0.0;
    if (i > 0) {
      const int a = kept[i - 1], b = kept[i];
      const double px = i == 1 ? x0 : m.lp_x[a], py = i == 1 ? y0 : m.lp_y[a];
      const double dx = m.lp_x[b] - px, dy = m.lp_y[b] - py;
      d = sqrt(dx * dx + dy * dy);
      const double dd = m.lp_heading[b] - m.lp_heading[a];
      double ddmod = sb_pymod(dd + SB_PI, SB_TWO_PI) - SB_PI;
      if (ddmod == -SB_PI && dd > 0) ddmod = SB_PI;
      corr = ddmod - dd;
      if (fabs(dd) < SB_PI) corr = 0.0;
    }
    cum[i] = d;
    acc[i] = corr;
    any_corr |= corr != 0.0;
  }
  any_corr = __any_sync(SB_FULL_MASK, any_corr);
  __syncwarp();
  if (lane == 0) {
    double c = 0.0;
    for (int i = 1; i < mm; ++i) {
      c = c + cum[i];
      cum[i] = c;
    }
    if (any_corr) {
      double a = 0.0;
      for (int i = 1; i < mm; ++i) {
        a += acc[i];
        acc[i] = a;
      }
    }
  }
  __syncwarp();
  if ((double)mm <= m.lanepoint_spacing) {  // sic (:1379)
    if (lane == 0 && n_out > 0) {
      const int l = m.lp_lane[lp0];
      DWaypoint w{lx, ly, rh0, m.lane_width[l], m.lane_speed[l], l, m.lane_idx_in_road[l]};
      emit(0, w);
    }
    return 1;
  }
  const double total = cum[mm - 1];
  const double step = total / (double)(n - 1);
  const int lim = n < n_out ? n : n_out;
  // kept point k as the serial version holds it in rx / ry / rh
  auto px = [&](int k) { return k == 0 ? x0 : m.lp_x[kept[k]]; };
  auto py = [&](int k) { return k == 0 ? y0 : m.lp_y[kept[k]]; };
  auto ph = [&](int k) { return k == 0 ? rh0 : m.lp_heading[kept[k]] + acc[k]; };
  for (int i = lane; i < lim; i += 32) {
    const double s = (i == n - 1) ? total : (double)i * step;
    int lo = 0, hi = mm - 1;  // largest j with cum[j] <= s (cum[0] = 0 <= s)
    while (lo < hi) {
      const int mid = (lo + hi + 1) >> 1;
      if (cum[mid] <= s) lo = mid; else hi = mid - 1;
    }
    const int j = lo;
    DWaypoint w;
    if (j >= mm - 1 || cum[j] == s) {
      const int l = m.lp_lane[kept[j]];
      w.x = px(j);
      w.y = py(j);
      w.heading = ph(j);
      w.lane_width = m.lane_width[l];
      w.speed_limit = m.lane_speed[l];
    } else {
      const double den = cum[j + 1] - cum[j], ds = s - cum[j];
      const int la = m.lp_lane[kept[j]], lb = m.lp_lane[kept[j + 1]];
      const double xa = px(j), ya = py(j), ha = ph(j);
      w.x = wp_slope(px(j + 1) - xa, den) * ds + xa;
      w.y = wp_slope(py(j + 1) - ya, den) * ds + ya;
      w.heading = wp_slope(ph(j + 1) - ha, den) * ds + ha;
      const double wa = m.lane_width[la], wb = m.lane_width[lb], sa = m.lane_speed[la], sb_ = m.lane_speed[lb];
      w.lane_width = wp_slope(wb - wa, den) * ds + wa;
      w.speed_limit = wp_slope(sb_ - sa, den) * ds + sa;
    }
    w.heading = d_wrap_heading(w.heading);
    int jd = j;  // largest jd with cum[jd] < s, else 0
    while (jd > 0 && !(cum[jd] < s)) --jd;
    w.lane = m.lp_lane[kept[jd]];
    w.lane_index = m.lane_idx_in_road[w.lane];
    emit(i, w);
  }
  return n;
}
