/*
 * oracle_core.hpp (+ oracle.cpp) -- CPU restatement of the SMARTS per-tick step, used as the
 * parity CHECKER.  This half: map queries, waypoints, controller helpers.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in the product (smarts_b200/) imports, links or executes
 * this file; only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs do.  Plain scalar code, one agent at a time, following the reference function by
 * function (citations are relative to /root/reference).  It shares with the CUDA kernels only
 * include/smarts_b200.h (struct layouts) and include/sb_detmath.h (libm-level primitives).
 *
 * Parity status (see DESIGN.md "Oracle"):
 *   pinned by the reference's own Python run here behind import shims (tests/golden/):
 *     lane points, waypoint_paths/_equally_spaced_path, nearest_lanes math, both controllers,
 *     the LiDAR ray table, neighbour radius filter;
 *   UNPINNED (third-party engine absent: pybullet / Panda3D): chassis integration, ray hits,
 *     OGM / DAGM rasters -- these restate the documented geometry.
 *
 * Build: g++ -O2 -ffp-contract=off -fopenmp -shared -fPIC (oracle/Makefile).
 */
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <array>
#include <cstring>
#include <deque>
#include <set>
#include <limits>
#include <vector>

#include "../include/sb_detmath.h"
#include "../include/sb_lidar_table.h"
#include "../include/sb_vehicle.h"
#include "../include/smarts_b200.h"

namespace {

constexpr double kPi = SB_PI;
constexpr double kTwoPi = SB_TWO_PI;

// ---------------------------------------------------------------- small math restatements
inline double clipd(double v, double lo, double hi) { return v < lo ? lo : (v > hi ? hi : v); }
// utils/math.py:206-216
inline double lerp(double a, double b, double p) { return a * (1.0 - p) + b * p; }
// coordinates.py:175-181 Heading.__new__
inline double wrap_heading(double h) {
  double v = sb_pymod(h, kTwoPi);
  if (v > kPi) v -= kTwoPi;
  return v;
}
// utils/math.py:447-449
inline double angdiff(double a, double b) { return sb_pymod((a - b) + kPi, kTwoPi) - kPi; }
// utils/math.py:247-253
inline void radians_to_vec(double r, double* vx, double* vy) {
  double a = sb_pymod(r + kPi * 0.5, kTwoPi);
  sb_sincos(a, vy, vx);
}
inline double degrees(double r) { return r * (180.0 / kPi); }
// utils/math.py:219-244
inline double low_pass(double in, double st, double k, double dt, double lo = -1.0, double raw = 0.0) {
  st += dt * k * (in - st);
  return clipd(st + raw, lo, 1.0);
}
// utils/math.py:163-185
inline double signed_dist_to_line(double px, double py, double lx, double ly, double dx, double dy) {
  double p2x = lx + dx, p2y = ly + dy;
  double u = std::fabs(((dy * px - dx * py) + p2x * ly) - p2y * lx);
  double d = u / std::sqrt(dx * dx + dy * dy);
  double nx = -dy, ny = dx;
  double dot = (px - lx) * nx + (py - ly) * ny;
  double sg = dot > 0 ? 1.0 : (dot < 0 ? -1.0 : 0.0);
  return d * sg;
}

struct Map {
  SbMap m;  // pointers redirected into the vectors below
  std::vector<int32_t> road_lane_start, road_lane_count, road_lp_start, road_lp_count, road_first_out;
  std::vector<uint8_t> road_is_junction;
  std::vector<int32_t> lane_road, lane_idx_in_road, lane_out_start, lane_out, lane_lp_start,
      lane_lp_count, lane_shape_start;
  std::vector<double> lane_width, lane_speed, lane_length;
  std::vector<double> lp_x, lp_y, lp_heading;
  std::vector<uint8_t> lp_inferred;
  std::vector<int32_t> lp_lane, lp_next_start, lp_next;
  std::vector<double> shape_x, shape_y, poly_x, poly_y, road_edges;
  std::vector<int32_t> poly_start;
  std::vector<int32_t> lane_in_start, lane_in, road_par_start, road_par;  // RoadWaypoints tables
};

template <class T>
void cp(std::vector<T>& dst, const T* src, size_t n) {
  dst.assign(src, src + n);
}

struct Waypoint {
  double x, y, heading, lane_width, speed_limit;
  int32_t lane;
  int32_t lane_index;
};

struct World {
  SbConfig cfg;
  SbVehicle veh;  // the agents' vehicle type (include/sb_vehicle.h)
  Map map;
  long long dt_num = 1;  // the simulation clock after k ticks is sb_clock(k, dt_num, dt_den) (include/sb_detmath.h)
  double dt_den = 10.0;
  std::vector<double> rays;  // [R,3]
  std::vector<SbRayRec> ray_rec;      // include/sb_lidar_table.h
  std::vector<double> ray_group_u;    // [groups,2]
  int E, A, M, V;
  int substeps;
  // state
  std::vector<double> pos_x, pos_y, pos_z, heading, speed, yaw_rate, length, width, height, cos_h, sin_h;
  std::vector<uint8_t> alive;
  std::vector<double> steering, v_long, v_lat, ctrl, last_control;
  std::vector<int32_t> route, route_len;
  std::vector<uint32_t> status;
  // obs
  std::vector<int32_t> ego_lane, nbr_count, nbr_id, nbr_lane, wp_n_paths, wp_len, wp_lane;
  std::vector<double> nbr_pos, wp_pos, lidar_points;
  std::vector<float> nbr_box, nbr_heading, nbr_speed, wp_heading, wp_lane_width, wp_speed_limit;
  std::vector<int8_t> nbr_lane_index, wp_lane_index, lidar_hits;
  std::vector<uint8_t> ogm, dagm;
  // events / trip meter / driven path (sensors.py:443-594, :830-947)
  std::vector<uint8_t> trip_init, trip_has, ev_done;
  std::vector<double> trip_x, trip_y, trip_h, trip_dist, trip_last, trip_wp, path_cum, last_px, last_py, goal_x,
      goal_y, goal_r, ev_distance, ev_reward;
  std::vector<int32_t> steps, path_n, ev_n_collisions, ev_collidee, ev_steps;
  std::vector<uint32_t> ev_events;
  std::vector<float> ttc_angle_error, ttc_dist_center, ttc_dtc, ttc_ttc;
  std::vector<uint8_t> ttc_valid;
  // AccelerometerSensor (sensors.py:1046-1084): the two deques of each agent, and its outputs
  std::vector<std::deque<std::array<double, 3>>> acc_lin_hist, acc_ang_hist;
  std::vector<double> acc_linear, acc_angular, acc_linear_jerk, acc_angular_jerk;  // [E*A,3]
  // ViaSensor (sensors.py:1090-1150): mission vias, the consumed set, and the outputs
  struct ViaRec { int lane; double x, y, hit_distance, required_speed; };
  std::vector<std::vector<ViaRec>> vias;
  std::vector<std::set<int>> via_consumed;
  std::vector<int32_t> via_near_count, via_near_idx, via_hit_count, via_hit_idx;
  // RoadWaypointsSensor (sensors.py:991-1043)
  std::vector<int32_t> rw_n_lanes, rw_lane, rw_n_paths, rw_path_slot, rw_len, rw_wp_lane;
  std::vector<double> rw_pos, rw_heading;
  std::vector<float> rw_lane_width, rw_speed_limit;
  std::vector<int8_t> rw_lane_index;
  int ring = 1;
  std::vector<uint8_t> ticked;  // [E*A] stepped since the agent's last events pass (SensorState.step is due)
  int threads = 1;
};

// ---------------------------------------------------------------- nearest lane (a9)
// utils/math.py:348-367 + :393-411 (in-repo equivalents of sumolib.geomhelper)
double dist_point_to_segment(double px, double py, double ax, double ay, double bx, double by) {
  double dx = ax - bx, dy = ay - by;
  double d = std::sqrt(dx * dx + dy * dy);
  double u = ((px - ax) * (bx - ax)) + ((py - ay) * (by - ay));
  double offset;
  if (d == 0.0 || u < 0.0 || u > d * d) {
    offset = (u < 0.0) ? 0.0 : d;
  } else {
    offset = u / d;
  }
  if (offset == 0.0) {
    double ex = px - ax, ey = py - ay;
    return std::sqrt(ex * ex + ey * ey);
  }
  double uu = offset / d;
  double ix = ax + uu * (bx - ax), iy = ay + uu * (by - ay);
  double ex = px - ix, ey = py - iy;
  return std::sqrt(ex * ex + ey * ey);
}

// utils/math.py:414-433 with perpendicular=False
double dist_point_to_polyline(const Map& mp, int lane, double px, double py) {
  int s = mp.lane_shape_start[lane], e = mp.lane_shape_start[lane + 1];
  double best = -1.0;
  for (int i = s; i + 1 < e; ++i) {
    double d = dist_point_to_segment(px, py, mp.shape_x[i], mp.shape_y[i], mp.shape_x[i + 1], mp.shape_y[i + 1]);
    if (best < 0.0 || d < best) best = d;
  }
  return best;
}

// sumo_road_network.py:675-701 + road_map.py:91-96: first of the candidates sorted by distance.
// Ties (unspecified in the reference: rtree order) resolve to the lowest lane index.
int nearest_lane(const Map& mp, double px, double py, double radius, double* dist_out = nullptr) {
  int best = -1;
  double bd = 0.0;
  for (int l = 0; l < mp.m.n_lanes; ++l) {
    double d = dist_point_to_polyline(mp, l, px, py);
    if (d < 0.0) continue;
    if (!(d < radius)) continue;
    if (best < 0 || d < bd) {
      best = l;
      bd = d;
    }
  }
  if (dist_out) *dist_out = bd;
  return best;
}

// ---------------------------------------------------------------- lane points / waypoints (a8)
inline double sqdist(const Map& mp, int i, double x, double y) {
  double dx = mp.lp_x[i] - x, dy = mp.lp_y[i] - y;
  return dx * dx + dy * dy;
}

// KD-tree k=1 query over an index range (lanepoints.py:517-524); lowest index wins ties.
int closest_lp_in_range(const Map& mp, int start, int count, double x, double y) {
  int best = -1;
  double bd = 0.0;
  for (int i = start; i < start + count; ++i) {
    double d = sqdist(mp, i, x, y);
    if (best < 0 || d < bd) {
      best = i;
      bd = d;
    }
  }
  return best;
}

// lanepoints.py:526-623 closest_lanepoints([pose], within_radius, maximum_count=10) -> first
int closest_lanepoint_with_pose(const Map& mp, double x, double y, double h, double within_radius) {
  const int P = mp.m.n_lanepoints;
  int k = std::min(10, P);
  int idx[10];
  double d2[10];
  int n = 0;
  // k nearest by (d2, index): selection, ascending
  double last_d = -1.0;
  int last_i = -1;
  for (int r = 0; r < k; ++r) {
    int best = -1;
    double bd = 0.0;
    for (int i = 0; i < P; ++i) {
      double d = sqdist(mp, i, x, y);
      bool after = (d > last_d) || (d == last_d && i > last_i);
      if (r > 0 && !after) continue;
      if (best < 0 || d < bd) {
        best = i;
        bd = d;
      }
    }
    if (best < 0) break;
    idx[n] = best;
    d2[n] = bd;
    ++n;
    last_d = bd;
    last_i = best;
  }
  // radius filter keeps index 0 always (lanepoints.py:548-561)
  int keep[10];
  int nk = 0;
  if (within_radius >= 0.0) {
    double r2 = within_radius * within_radius;
    for (int i = 0; i < n; ++i)
      if (d2[i] <= r2 || i == 0) keep[nk++] = i;
  } else {
    for (int i = 0; i < n; ++i) keep[nk++] = i;
  }
  // stable sort by d2 + |heading.relative_to(lp.heading)| (lanepoints.py:581-590) -> first
  int best = -1;
  double bk = 0.0;
  for (int j = 0; j < nk; ++j) {
    int i = keep[j];
    double rel = wrap_heading(h - mp.lp_heading[idx[i]]);
    double key = d2[i] + std::fabs(rel);
    if (best < 0 || key < bk) {
      best = idx[i];
      bk = key;
    }
  }
  return best;
}

bool in_route(const int32_t* route, int n, int road) {
  for (int i = 0; i < n; ++i)
    if (route[i] == road) return true;
  return false;
}

// lanepoints.py:646-692.  Returns number of paths; paths[k][0..len[k]) are lane-point indices.
int paths_starting_at(const Map& mp, int start, int lookahead, const int32_t* filt, int nf,
                      int32_t paths[][SB_MAX_LOOKAHEAD + 1], int* len, int max_paths, bool* overflow) {
  std::vector<std::vector<int32_t>> cur(1, std::vector<int32_t>(1, start));
  for (int it = 0; it < lookahead; ++it) {
    std::vector<std::vector<int32_t>> nxt;
    for (auto& path : cur) {
      int tail = path.back();
      int added = 0;
      for (int j = mp.lp_next_start[tail]; j < mp.lp_next_start[tail + 1]; ++j) {
        int nlp = mp.lp_next[j];
        int nlane = mp.lp_lane[nlp];
        int road = mp.lane_road[nlane];
        if (nf > 0 && !in_route(filt, nf, road)) continue;
        if (nf > 0 && road != filt[nf - 1]) {
          bool any = false;
          for (int o = mp.lane_out_start[nlane]; o < mp.lane_out_start[nlane + 1]; ++o)
            if (in_route(filt, nf, mp.lane_road[mp.lane_out[o]])) any = true;
          if (!any) continue;
        }
        std::vector<int32_t> np = path;
        np.push_back(nlp);
        nxt.push_back(std::move(np));
        ++added;
      }
      if (!added) nxt.push_back(path);
    }
    cur.swap(nxt);
  }
  int n = 0;
  for (auto& p : cur) {
    if (n >= max_paths) {
      *overflow = true;
      break;
    }
    len[n] = (int)p.size();
    for (size_t i = 0; i < p.size(); ++i) paths[n][i] = p[i];
    ++n;
  }
  return n;
}

// sumo_road_network.py:1312-1437
int equally_spaced_path(const Map& mp, const int32_t* path, int n, double vx, double vy, Waypoint* out) {
  double rx[SB_MAX_LOOKAHEAD + 1], ry[SB_MAX_LOOKAHEAD + 1], rh[SB_MAX_LOOKAHEAD + 1], rw[SB_MAX_LOOKAHEAD + 1],
      rs[SB_MAX_LOOKAHEAD + 1];
  int32_t rlane[SB_MAX_LOOKAHEAD + 1];
  if (n <= 0) return 0;
  rh[0] = 0.0;  // (n >= 1 keeps path[0]; this only tells the compiler so)
  int m = 0;
  for (int i = 0; i < n; ++i) {
    int lp = path[i];
    if (mp.lp_inferred[lp] && 0 < i && i < n - 1) continue;
    int lane = mp.lp_lane[lp];
    rx[m] = mp.lp_x[lp];
    ry[m] = mp.lp_y[lp];
    rh[m] = mp.lp_heading[lp];
    rw[m] = mp.lane_width[lane];
    rs[m] = mp.lane_speed[lane];
    rlane[m] = lane;
    ++m;
  }
  // inplace_unwrap, utils/math.py:537-550
  {
    double acc = 0.0;
    double prev = rh[0];
    for (int i = 1; i < m; ++i) {
      double cur = rh[i];
      double dd = cur - prev;
      double ddmod = sb_pymod(dd + kPi, kTwoPi) - kPi;
      if (ddmod == -kPi && dd > 0) ddmod = kPi;
      double corr = ddmod - dd;
      if (std::fabs(dd) < kPi) corr = 0.0;
      acc += corr;
      prev = cur;
      rh[i] = cur + acc;
    }
  }
  double hvx, hvy;
  radians_to_vec(rh[0], &hvx, &hvy);
  int lp0 = path[0];
  double lx = mp.lp_x[lp0], ly = mp.lp_y[lp0];
  double proj = (vx - lx) * hvx + (vy - ly) * hvy;
  rx[0] = lx + proj * hvx;
  ry[0] = ly + proj * hvy;
  double cum[SB_MAX_LOOKAHEAD + 1];
  cum[0] = 0.0;
  for (int i = 1; i < m; ++i) {
    double dx = rx[i] - rx[i - 1], dy = ry[i] - ry[i - 1];
    cum[i] = cum[i - 1] + std::sqrt(dx * dx + dy * dy);
  }
  if ((double)m <= mp.m.lanepoint_spacing) {  // sic: a count compared with the spacing (:1379)
    int lane = mp.lp_lane[lp0];
    out[0] = Waypoint{lx, ly, mp.lp_heading[lp0], mp.lane_width[lane], mp.lane_speed[lane], lane,
                      mp.lane_idx_in_road[lane]};
    return 1;
  }
  double total = cum[m - 1];
  double step = total / (double)(n - 1);
  int jd = 0;
  for (int i = 0; i < n; ++i) {
    double s = (i == n - 1) ? total : (double)i * step;  // np.linspace(0, total, n)
    // np.interp
    double vals[5];
    const double* fp[5] = {rx, ry, rh, rw, rs};
    int j;
    if (s > cum[m - 1]) {
      j = m - 1;
    } else {
      j = 0;
      while (j + 1 < m && cum[j + 1] <= s) ++j;
    }
    for (int v = 0; v < 5; ++v) {
      if (j >= m - 1 || cum[j] == s) {
        vals[v] = fp[v][j];
      } else {
        double slope = (fp[v][j + 1] - fp[v][j]) / (cum[j + 1] - cum[j]);
        vals[v] = slope * (s - cum[j]) + fp[v][j];
      }
    }
    // discrete variables (:1404-1417)
    while (jd + 1 < m && s > cum[jd + 1]) ++jd;
    int lane = rlane[jd];
    out[i] = Waypoint{vals[0], vals[1], wrap_heading(vals[2]), vals[3], vals[4], lane, mp.lane_idx_in_road[lane]};
  }
  return n;
}

struct PathSet {
  int n = 0;         // paths kept (<= SB_MAX_PATHS)
  int total = 0;     // paths found
  int len[SB_MAX_PATHS];
  Waypoint wp[SB_MAX_PATHS][SB_MAX_LOOKAHEAD + 1];
  bool overflow = false;
};

// Lane._waypoint_paths_at, sumo_road_network.py:429-446 + :1278-1310
void lane_paths_at(const Map& mp, int lane, double x, double y, int lookahead, const int32_t* filt, int nf,
                   PathSet* ps) {
  if (mp.lane_lp_count[lane] <= 0) return;
  int lp = closest_lp_in_range(mp, mp.lane_lp_start[lane], mp.lane_lp_count[lane], x, y);
  int32_t paths[SB_MAX_PATHS][SB_MAX_LOOKAHEAD + 1];
  int len[SB_MAX_PATHS];
  bool of = false;
  int np = paths_starting_at(mp, lp, lookahead, filt, nf, paths, len, SB_MAX_PATHS, &of);
  if (of) ps->overflow = true;
  for (int k = 0; k < np; ++k) {
    ps->total++;
    if (ps->n >= SB_MAX_PATHS) {
      ps->overflow = true;
      continue;
    }
    ps->len[ps->n] = equally_spaced_path(mp, paths[k], len[k], x, y, ps->wp[ps->n]);
    ps->n++;
  }
}

void sort_paths(PathSet* ps) {
  // sorted(..., key=lambda p: p[0].lane_index): stable insertion sort
  for (int i = 1; i < ps->n; ++i) {
    int j = i;
    while (j > 0 && ps->wp[j - 1][0].lane_index > ps->wp[j][0].lane_index) {
      std::swap(ps->len[j - 1], ps->len[j]);
      for (int t = 0; t <= SB_MAX_LOOKAHEAD; ++t) std::swap(ps->wp[j - 1][t], ps->wp[j][t]);
      --j;
    }
  }
}

// sumo_road_network.py:815-882.  route_len < 0 means "no route object" (the lookahead=1
// fallback of sensors.py:271-275); route_len == 0 is an empty Route.
void waypoint_paths(const Map& mp, double x, double y, double h, int lookahead,
                    double within_radius, const int32_t* route, int route_len, PathSet* ps) {
  ps->n = ps->total = 0;
  ps->overflow = false;
  int32_t jr[2];
  const int32_t* roads = route;
  int nr = route_len;
  if (route_len == 0) {
    // _resolve_in_junction, :842-860
    int lp = closest_lanepoint_with_pose(mp, x, y, h, -1.0);
    nr = 0;
    if (lp >= 0) {
      int road = mp.lane_road[mp.lp_lane[lp]];
      if (mp.road_is_junction[road]) {
        jr[nr++] = road;
        if (mp.road_first_out[road] >= 0) jr[nr++] = mp.road_first_out[road];
      }
    }
    roads = jr;
  }
  if (route_len >= 0 && nr > 0) {
    // _waypoint_paths_along_route, :862-882
    int best = -1;
    double bd = 0.0;
    for (int r = 0; r < nr; ++r) {
      int road = roads[r];
      if (mp.road_lp_count[road] <= 0) continue;
      int lp = closest_lp_in_range(mp, mp.road_lp_start[road], mp.road_lp_count[road], x, y);
      double d = std::sqrt(sqdist(mp, lp, x, y));
      if (best < 0 || d < bd) {
        best = lp;
        bd = d;
      }
    }
    if (best >= 0) {
      int road = mp.lane_road[mp.lp_lane[best]];
      for (int l = mp.road_lane_start[road]; l < mp.road_lane_start[road] + mp.road_lane_count[road]; ++l)
        lane_paths_at(mp, l, x, y, lookahead, roads, nr, ps);
      sort_paths(ps);
      return;
    }
  }
  int lp = closest_lanepoint_with_pose(mp, x, y, h, within_radius);
  if (lp < 0) return;
  int road = mp.lane_road[mp.lp_lane[lp]];
  for (int l = mp.road_lane_start[road]; l < mp.road_lane_start[road] + mp.road_lane_count[road]; ++l)
    lane_paths_at(mp, l, x, y, lookahead, nullptr, 0, ps);
  sort_paths(ps);
}

// ---------------------------------------------------------------- controllers (a2-a4)
// trajectory_tracking_controller.py:443-473
double curvature_calculation(const double* tx, const double* ty, const double* th, int n, int offset = 0,
                             int num_points = 5) {
  if (n <= num_points + offset) return 1e20;
  double hs = 0.0, ds = 0.0;
  for (int i = 0; i < num_points; ++i) {
    hs += angdiff(th[i + 1 + offset], th[i + offset]);
    double dx = tx[i + offset] - tx[i + offset + 1], dy = ty[i + offset] - ty[i + offset + 1];
    ds += std::fabs(std::sqrt(dx * dx + dy * dy));
  }
  if (hs == 0.0) return 1e20;
  return ds / hs;
}

// lane_following_controller.py:376-437.  place_poles for a single-input system is unique, so
// it is evaluated with Ackermann's formula K = e_n^T C^-1 phi(A) (tests pin it to
// scipy.signal.place_poles).  Constants: vehicle.length/2 = 1.84, chassis link mass 2356,
// Izz 2681.95008628 (models/vehicle.urdf:18-22), ground contact stiffness 1e5 (models/plane.urdf:26).
void lateral_gains(double ts, const SbVehicle& vp, double* heading_gain, double* lateral_gain) {
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
  // controllability matrix columns B, AB, A^2B, A^3B
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
  // phi(A) = prod (A - p_i I)
  double Ph[4][4] = {{1, 0, 0, 0}, {0, 1, 0, 0}, {0, 0, 1, 0}, {0, 0, 0, 1}};
  for (int p = 0; p < 4; ++p) {
    double T[4][4], R[4][4];
    for (int r = 0; r < 4; ++r)
      for (int k = 0; k < 4; ++k) T[r][k] = Am[r][k] - (r == k ? poles[p] : 0.0);
    for (int r = 0; r < 4; ++r)
      for (int k = 0; k < 4; ++k) {
        double s = 0;
        for (int j = 0; j < 4; ++j) s += Ph[r][j] * T[j][k];
        R[r][k] = s;
      }
    std::memcpy(Ph, R, sizeof(R));
  }
  // solve y^T C = e4^T  <=>  C^T y = e4, Gaussian elimination with partial pivoting
  double Mx[4][5];
  for (int r = 0; r < 4; ++r) {
    for (int k = 0; k < 4; ++k) Mx[r][k] = Cm[k][r];
    Mx[r][4] = (r == 3) ? 1.0 : 0.0;
  }
  for (int col = 0; col < 4; ++col) {
    int piv = col;
    for (int r = col + 1; r < 4; ++r)
      if (std::fabs(Mx[r][col]) > std::fabs(Mx[piv][col])) piv = r;
    if (piv != col)
      for (int k = 0; k < 5; ++k) std::swap(Mx[piv][k], Mx[col][k]);
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
  double K[4];
  for (int k = 0; k < 4; ++k) {
    double s = 0;
    for (int j = 0; j < 4; ++j) s += yv[j] * Ph[j][k];
    K[k] = s;
  }
  *heading_gain = clipd(K[1], 0.02, 0.04);
  *lateral_gain = clipd(K[0], 3.4, 4.1);
}

}  // namespace

