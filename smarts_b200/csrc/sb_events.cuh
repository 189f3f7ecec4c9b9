// sb_events.cuh -- k_events: Events / done criteria / trip-meter reward / vehicle-vehicle contact
// (SURVEY 8(f) rows 1-2), one warp per agent, launched at the end of sb_observe.
//
//   Sensors._is_done_with_events            smarts/core/sensors.py:443-487
//   _vehicle_is_off_road / _on_shoulder     sensors.py:496-507, sumo_road_network.py:703-709
//   _vehicle_is_not_moving                  sensors.py:509-524, DrivenPathSensor :830-877
//   _vehicle_is_off_route_and_wrong_way     sensors.py:526-594
//   TripMeterSensor                         sensors.py:880-947; reward agent_manager.py:233-234
//   PositionalGoal.is_reached               plan.py:117-121
//   SMARTS._process_collisions              smarts.py:1270-1291, chassis.py:60-80 (0.05 m leeway)
//   sumolib 1.10.0 geomhelper (lineOffsetWithMinimumDistanceToPoint,
//   polygonOffsetWithMinimumDistanceToPoint, positionAtOffset, positionAtShapeOffset): restated,
//   the package is not vendored in the reference (setup.py:28).
#pragma once

// included by sb_kernels.cuh after DCfg / DObs / d_nearest_lane_warp

struct DEvents {
  // per agent [E*A]
  uint8_t *trip_init, *trip_has;
  double *trip_x, *trip_y, *trip_h, *trip_dist, *trip_last;
  // [E*A,2,4] x, y, heading, lane (-1: no waypoint), written by k_observe: [0] waypoint_paths[0][0] of this
  // observation, [1] the waypoint of TripMeterSensor.__init__'s own unrouted query (first observation after a reset)
  double* trip_wp;
  int32_t *steps, *path_n;
  double *path_cum;  // [E*A, ring] cumulative driven distance, ring buffer
  double *last_px, *last_py;
  double *goal_x, *goal_y, *goal_r;
  int32_t ring;      // entries in the ring = not-moving window steps + 2 (<= SB_DRIVEN_PATH_MAX)
  long long dt_num;  // the simulation clock after k ticks is sb_clock(k, dt_num, dt_den) (include/sb_detmath.h)
  double dt_den;
};

struct DEvOut {
  uint32_t* events;
  uint8_t* done;
  int32_t *n_collisions, *collidee;
  double *distance, *reward;
  int32_t* steps;
};

// sumolib.geomhelper.lineOffsetWithMinimumDistanceToPoint(perpendicular=False)
__device__ __forceinline__ double ev_line_offset(double px, double py, double ax, double ay, double bx, double by) {
  double dx = ax - bx, dy = ay - by;
  double d = sqrt(dx * dx + dy * dy);
  double u = ((px - ax) * (bx - ax)) + ((py - ay) * (by - ay));
  if (d == 0.0 || u < 0.0 || u > d * d) return u < 0.0 ? 0.0 : d;
  return u / d;
}

__device__ __forceinline__ bool ev_isclose(double a, double b) {  // sumolib geomhelper.isclose, rel_tol 1e-9
  double m = fabs(a) > fabs(b) ? fabs(a) : fabs(b);
  return fabs(a - b) <= 1e-9 * m;
}

// sumolib.geomhelper.positionAtOffset
__device__ __forceinline__ void ev_position_at_offset(double ax, double ay, double bx, double by, double off,
                                                      double* x, double* y) {
  if (ev_isclose(off, 0.0)) { *x = ax; *y = ay; return; }
  double dx = ax - bx, dy = ay - by;
  double dist = sqrt(dx * dx + dy * dy);
  if (ev_isclose(dist, off)) { *x = bx; *y = by; return; }
  *x = ax + (bx - ax) * (off / dist);
  *y = ay + (by - ay) * (off / dist);
}

// Lane.offset_along_lane (sumo_road_network.py:476-491): polygonOffsetWithMinimumDistanceToPoint
__device__ double ev_offset_along_lane(const DMap& m, int lane, double px, double py) {
  const int s = m.lane_shape_start[lane], e = m.lane_shape_start[lane + 1];
  double seen = 0.0, min_dist = 1e300, min_off = -1.0;
  for (int i = s; i + 1 < e; ++i) {
    double ax = m.shape_x[i], ay = m.shape_y[i], bx = m.shape_x[i + 1], by = m.shape_y[i + 1];
    double pos = ev_line_offset(px, py, ax, ay, bx, by);
    double qx, qy;
    ev_position_at_offset(ax, ay, bx, by, pos, &qx, &qy);
    double ex = px - qx, ey = py - qy;
    double dist = sqrt(ex * ex + ey * ey);
    if (dist < min_dist) {
      min_dist = dist;
      min_off = pos + seen;
    }
    double dx = ax - bx, dy = ay - by;
    seen += sqrt(dx * dx + dy * dy);
  }
  return min_off;
}

// Lane.from_lane_coord (sumo_road_network.py:502-506): positionAtShapeOffset
__device__ void ev_from_lane_coord(const DMap& m, int lane, double off, double* x, double* y) {
  const int s = m.lane_shape_start[lane], e = m.lane_shape_start[lane + 1];
  double seen = 0.0;
  for (int i = s; i + 1 < e; ++i) {
    double ax = m.shape_x[i], ay = m.shape_y[i], bx = m.shape_x[i + 1], by = m.shape_y[i + 1];
    double dx = ax - bx, dy = ay - by;
    double len = sqrt(dx * dx + dy * dy);
    if (seen + len > off) {
      ev_position_at_offset(ax, ay, bx, by, off - seen, x, y);
      return;
    }
    seen += len;
  }
  *x = m.shape_x[e - 1];
  *y = m.shape_y[e - 1];
}

// RoadMap.Lane.vector_at_offset (road_map.py:377-388)
__device__ void ev_vector_at_offset(const DMap& m, int lane, double off, double* vx, double* vy) {
  const double len = m.lane_length[lane];
  double s0, s1;
  if (off >= len) {
    s0 = len - 1;
    s1 = len;
  } else {
    s0 = off;
    s1 = off + 1;
  }
  s0 = s0 > 0 ? s0 : 0;
  double x0, y0, x1, y1;
  ev_from_lane_coord(m, lane, s0, &x0, &y0);
  ev_from_lane_coord(m, lane, s1, &x1, &y1);
  *vx = x1 - x0;
  *vy = y1 - y0;
}

// utils/math.py:256-277
__device__ __forceinline__ double ev_vec_to_radians(double x, double y) {
  double r = sb_atan2(fabs(y), fabs(x));
  if (x < 0) {
    if (y < 0) return sb_pymod(r + 0.5 * SB_PI, SB_TWO_PI);
    return sb_pymod(0.5 * SB_PI - r, SB_TWO_PI);
  } else if (y < 0) {
    return sb_pymod(1.5 * SB_PI - r, SB_TWO_PI);
  }
  return sb_pymod(r - 0.5 * SB_PI, SB_TWO_PI);
}

// utils/math.py:414-433 for one lane (serial)
__device__ double ev_dist_to_lane(const DMap& m, int lane, double px, double py) {
  const int s = m.lane_shape_start[lane], e = m.lane_shape_start[lane + 1];
  double best = 1e300;
  for (int i = s; i + 1 < e; ++i) {
    double d = d_dist_point_to_segment(px, py, m.shape_x[i], m.shape_y[i], m.shape_x[i + 1], m.shape_y[i + 1]);
    if (d < best) best = d;
  }
  return best;
}

// 2-D oriented-box contact with the 0.05 m leeway of chassis.py:76-79, as a separating-axis
// test: the boxes are in contact unless some box axis separates them by more than the leeway.
__device__ __forceinline__ bool ev_boxes_touch(double ax, double ay, double ac, double as, double al, double aw,
                                               double bx, double by, double bc, double bs, double bl, double bw) {
  // heading h: forward = (-sin h, cos h), left = (-cos h, -sin h)  (coordinates.py:169-185)
  const double afx = -as, afy = ac, asx = ac, asy = as;
  const double bfx = -bs, bfy = bc, bsx = bc, bsy = bs;
  const double dx = bx - ax, dy = by - ay;
  const double hal = 0.5 * al, haw = 0.5 * aw, hbl = 0.5 * bl, hbw = 0.5 * bw;
  const double leeway = 0.05;
  const double ff = fabs(afx * bfx + afy * bfy), fs = fabs(afx * bsx + afy * bsy);
  const double sf = fabs(asx * bfx + asy * bfy), ss = fabs(asx * bsx + asy * bsy);
  if (fabs(dx * afx + dy * afy) > (hal + (hbl * ff + hbw * fs)) + leeway) return false;
  if (fabs(dx * asx + dy * asy) > (haw + (hbl * sf + hbw * ss)) + leeway) return false;
  if (fabs(dx * bfx + dy * bfy) > (hbl + (hal * ff + haw * sf)) + leeway) return false;
  if (fabs(dx * bsx + dy * bsy) > (hbw + (hal * fs + haw * ss)) + leeway) return false;
  return true;
}

#define SB_EV_WARPS 4

#ifndef SB_EV_MIN_BLOCKS
#define SB_EV_MIN_BLOCKS 10  // 48 registers: 0.110 -> 0.087 ms on the headline with events (8: 0.091, 12: 0.087), profiles/README.md r2y
#endif
__global__ void __launch_bounds__(SB_EV_WARPS * 32, SB_EV_MIN_BLOCKS)
k_events(DCfg c, DMap m, DState s, DEvents ev, DEvOut o) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  int e, a;
  if (!d_work_agent(c, blockIdx.x * SB_EV_WARPS + warp, &e, &a)) return;
  const int ai = e * c.A + a, vi = e * c.V + a;
  // this agent's tick flag: read by lane 0, handed to the warp, used up (one warp per agent here and one thread
  // per agent in the control kernels, so nobody else touches the byte)
  int ticked = 0;
  if (lane == 0) {
    ticked = s.ticked[ai];
    s.ticked[ai] = 0;
  }
  ticked = __shfl_sync(0xffffffffu, ticked, 0);
  const double px = s.pos_x[vi], py = s.pos_y[vi], ph = s.heading[vi];
  const double len = s.length[vi], wid = s.width[vi];
  const double ch = s.cos_h[vi], sh = s.sin_h[vi];
  const int32_t* route = s.route + (size_t)ai * c.route_cap;
  const int route_len = s.route_len[ai];
  uint32_t bits = 0;

  // ---- SensorState.step + trip meter + driven path (serial, lane 0) -----------------------
  int steps = 0;
  double nm_dist = 0.0;
  bool nm_valid = false;
  if (lane == 0) {
    steps = ev.steps[ai] + (ticked ? 1 : 0);
    ev.steps[ai] = steps;
    o.steps[ai] = steps;
    const double* tw = ev.trip_wp + (size_t)ai * 8;
    const int tlane = (int)tw[3];
    double dist = ev.trip_dist[ai], last = ev.trip_last[ai];
    if (!ev.trip_init[ai]) {  // TripMeterSensor.__init__ (sensors.py:885-898): its own unrouted lookahead-1 query
      ev.trip_init[ai] = 1;
      const double* cw = c.wp_enable ? tw + 4 : tw;
      if ((int)cw[3] >= 0) {
        ev.trip_has[ai] = 1;
        ev.trip_x[ai] = cw[0];
        ev.trip_y[ai] = cw[1];
        ev.trip_h[ai] = cw[2];
      }
    }
    if (tlane >= 0) {
      const double wx = tw[0], wy = tw[1], wh = tw[2];
      // append_waypoint_if_new, sensors.py:900-930
      last = dist;
      const bool fixed_route = ev.goal_r[ai] >= 0.0;
      const bool count = !fixed_route || wp_in_route(route, route_len, m.lane_road[tlane]);
      if (!ev.trip_has[ai]) {
        if (count) {
          ev.trip_has[ai] = 1;
          ev.trip_x[ai] = wx;
          ev.trip_y[ai] = wy;
          ev.trip_h[ai] = wh;
        }
      } else {
        const double rx = ev.trip_x[ai], ry = ev.trip_y[ai];
        const double ddx = wx - rx, ddy = wy - ry;
        const double d = sqrt(ddx * ddx + ddy * ddy);
        if (d > 0.5 && count) {
          double hvx, hvy;
          d_radians_to_vec(ev.trip_h[ai], &hvx, &hvy);
          const double dot = hvx * ddx + hvy * ddy;
          const double sg = dot > 0 ? 1.0 : (dot < 0 ? -1.0 : 0.0);
          dist += sg * d;
          ev.trip_x[ai] = wx;
          ev.trip_y[ai] = wy;
          ev.trip_h[ai] = wh;
        }
      }
    }
    ev.trip_dist[ai] = dist;
    ev.trip_last[ai] = last;
    o.distance[ai] = dist;
    o.reward[ai] = dist - last;

    // DrivenPathSensor.track_latest_driven_path (sensors.py:842-848) as a ring of cumulative length
    const int n = ev.path_n[ai];
    double* ring = ev.path_cum + (size_t)ai * ev.ring;
    double cum = 0.0;
    if (n > 0) {
      const double qx = px - ev.last_px[ai], qy = py - ev.last_py[ai];
      cum = ring[(n - 1) % ev.ring] + sqrt(qx * qx + qy * qy);
    }
    ring[n % ev.ring] = cum;
    ev.path_n[ai] = n + 1;
    ev.last_px[ai] = px;
    ev.last_py[ai] = py;
    // _vehicle_is_not_moving (sensors.py:509-524) over the entries whose timestamp >= now - not_moving_time
    // (sensors.py:866-871).  Entry j carries the clock value sb_clock(j); the comparison is made on those binary64
    // values as the reference makes it (a 0.1-second clock does not subtract exactly, so the window is one entry
    // shorter on some ticks than not_moving_time / dt)
    const double now = sb_clock(steps, ev.dt_num, ev.dt_den);
    if (!(now < c.nm_time)) {
      const double threshold = now - c.nm_time;
      int back = min(min(ev.ring - 1, n), steps);
      while (back > 0 && !(sb_clock(steps - back, ev.dt_num, ev.dt_den) >= threshold)) --back;
      nm_dist = cum - ring[(n - back) % ev.ring];
      nm_valid = true;
    }
    if (nm_valid && nm_dist < c.nm_dist) bits |= SB_EV_NOT_MOVING;
    // PositionalGoal.is_reached, plan.py:117-121
    const double gr = ev.goal_r[ai];
    if (gr >= 0.0) {
      const double gx = px - ev.goal_x[ai], gy = py - ev.goal_y[ai];
      if (gx * gx + gy * gy <= gr * gr) bits |= SB_EV_REACHED_GOAL;
    }
    if (c.max_steps > 0 && steps >= c.max_steps) bits |= SB_EV_REACHED_MAX_STEPS;
  }

  // ---- vehicle-vehicle contact ------------------------------------------------------------
  {
    int cnt = 0, first = 0x7fffffff;
    for (int base = 0; base < c.V; base += 32) {
      const int j = base + lane, vj = e * c.V + j;
      bool hit = false;
      if (j < c.V && j != a && s.alive[vj])
        hit = ev_boxes_touch(px, py, ch, sh, len, wid, s.pos_x[vj], s.pos_y[vj], s.cos_h[vj], s.sin_h[vj],
                             s.length[vj], s.width[vj]);
      const unsigned bal = __ballot_sync(SB_FULL_MASK, hit);
      if (bal && first == 0x7fffffff) first = base + __ffs(bal) - 1;
      cnt += __popc(bal);
    }
    if (lane == 0) {
      o.n_collisions[ai] = cnt;
      o.collidee[ai] = cnt ? first : -1;
      if (cnt) bits |= SB_EV_COLLISION;
    }
  }

  // ---- off road (centre) / on shoulder (the four Vehicle.bounding_box corners) --------------
  {
    // vehicle.py:316-332 with rotate_around_point (utils/math.py:436-444) exactly as written there
    double qx[5], qy[5];
    qx[0] = px;
    qy[0] = py;
    const double cw[4] = {-0.5, 0.5, 0.5, -0.5}, cl[4] = {0.5, 0.5, -0.5, -0.5};
    for (int k = 0; k < 4; ++k) {
      const double ox = (px + cw[k] * wid) - px, oy = (py + cl[k] * len) - py;
      qx[k + 1] = (px + ch * ox) + sh * oy;
      qy[k + 1] = (py + (-sh) * ox) + ch * oy;
    }
    unsigned on = 0;  // bit k: point k lies on some road
    const double rad = 6.4;  // max(5, 2 * default lane width 3.2), sumo_road_network.py:705
    if (m.seg_nx != 0 && rad <= m.seg_rmax) {
      for (int k = 0; k < 5; ++k) {  // each point reads the segment list of its own grid cell
        const int32_t* items;
        int n;
        d_seg_candidates(m, qx[k], qy[k], rad, &items, &n);
        for (int t = lane; t < n; t += 32) {
          const int i = items ? items[t] : t;
          const int l = m.shape_lane[i];
          if (l != m.shape_lane[i + 1]) continue;
          const double d = d_dist_point_to_segment(qx[k], qy[k], m.shape_x[i], m.shape_y[i], m.shape_x[i + 1],
                                                   m.shape_y[i + 1]);
          if (d < rad && d < 0.5 * m.lane_width[l] + 1e-1) on |= 1u << k;
        }
      }
    } else {
      for (int i = lane; i + 1 < m.n_shape_pts; i += 32) {
        const int l = m.shape_lane[i];
        if (l != m.shape_lane[i + 1]) continue;
        const double thr = 0.5 * m.lane_width[l] + 1e-1;
        const double ax = m.shape_x[i], ay = m.shape_y[i], bx = m.shape_x[i + 1], by = m.shape_y[i + 1];
#pragma unroll
        for (int k = 0; k < 5; ++k) {
          const double d = d_dist_point_to_segment(qx[k], qy[k], ax, ay, bx, by);
          if (d < rad && d < thr) on |= 1u << k;
        }
      }
    }
    on = __reduce_or_sync(SB_FULL_MASK, on);
    if (!(on & 1u)) bits |= SB_EV_OFF_ROAD;
    if ((on & 0x1eu) != 0x1eu) bits |= SB_EV_ON_SHOULDER;
  }

  // ---- off route / wrong way -----------------------------------------------------------------
  {
    const double radius = sqrt(len * len + wid * wid) * 0.5 + 5;
    const int nl = d_nearest_lane_warp(m, px, py, radius, lane);
    if (nl < 0) {
      bits |= SB_EV_OFF_ROUTE;
    } else {
      const int road = m.lane_road[nl];
      const bool in_j = m.road_is_junction[road] != 0;
      double off = 0.0, mvx = 0.0, mvy = 0.0;
      if (lane == 0) {
        off = ev_offset_along_lane(m, nl, px, py);
        ev_vector_at_offset(m, nl, off, &mvx, &mvy);
      }
      off = __shfl_sync(SB_FULL_MASK, off, 0);
      mvx = __shfl_sync(SB_FULL_MASK, mvx, 0);
      mvy = __shfl_sync(SB_FULL_MASK, mvy, 0);
      if (!in_j) {  // _check_wrong_way_event, sensors.py:576-594
        const double th = d_wrap_heading(ev_vec_to_radians(mvx, mvy));
        const double rel = d_wrap_heading(ph - th);
        if (fabs(rel) > 0.5 * SB_PI) bits |= SB_EV_WRONG_WAY;
      }
      if (!(route_len <= 0 || wp_in_route(route, route_len, road) || in_j)) {
        // oncoming_lanes_at_offset (sumo_road_network.py:371-396): a lane of the route running the
        // other way next to this one keeps the vehicle on route
        double cx = 0.0, cy = 0.0;
        if (lane == 0) ev_from_lane_coord(m, nl, off, &cx, &cy);
        cx = __shfl_sync(SB_FULL_MASK, cx, 0);
        cy = __shfl_sync(SB_FULL_MASK, cy, 0);
        const double near_r = 1.1 * m.lane_width[nl];
        const double my_norm = sqrt(mvx * mvx + mvy * mvy);
        bool found = false;
        if (my_norm != 0.0) {
          for (int l = lane; l < m.n_lanes; l += 32) {
            if (l == nl) continue;
            if (!wp_in_route(route, route_len, m.lane_road[l])) continue;
            if (!(ev_dist_to_lane(m, l, cx, cy) < near_r)) continue;
            const double sl = ev_offset_along_lane(m, l, cx, cy);
            double lvx, lvy;
            ev_vector_at_offset(m, l, sl, &lvx, &lvy);
            const double lv_norm = sqrt(lvx * lvx + lvy * lvy);
            if (lv_norm == 0.0) continue;
            const double ang = (mvx * lvx + mvy * lvy) / (my_norm * lv_norm);
            if (ang < -0.995562) found = true;
          }
        }
        if (!__any_sync(SB_FULL_MASK, found)) bits |= SB_EV_OFF_ROUTE;
      }
    }
  }

  // the warp-uniform bits are on every lane; lane 0 also holds its serial ones
  if (lane == 0) {
    o.events[ai] = bits;
    const uint32_t mask = c.done_mask | SB_EV_REACHED_GOAL | SB_EV_REACHED_MAX_STEPS;
    o.done[ai] = (ticked && (bits & mask)) ? 1 : 0;  // `not sim.resetting`, sensors.py:464
  }
}
