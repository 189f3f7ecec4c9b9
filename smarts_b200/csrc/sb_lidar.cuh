// sb_lidar.cuh -- LidarSensor ray casting (block per agent; azimuth groups x primitives, then thread per ray).
//   Lidar.compute_point_cloud / _compute_rays   smarts/core/lidar.py:58-134
//   LidarSensor                                 smarts/core/sensors.py:797-827
//
// Result contract (DESIGN.md "LiDAR"; the oracle evaluates it ray by ray): closest hit of every table ray
// against the ground plane, every other alive vehicle's box and (lidar_wall_height > 0) the road-edge
// walls.  The ray table is a product of azimuths and laser angles (lidar.py:98-100: BasicLidar 6 x 50,
// VelodyneHDL32E 36 x 24); include/sb_lidar_table.h factors it once, on upload, into azimuth groups, and a ray
// is  origin + s (ux, uy, slope),  0 <= s <= hlen.  All rays of a group share their vertical plane, so
//
//   stage   (thread = vehicle / edge)  primitives that can reach the agent's disc go to shared memory
//           (float circle test; the result is a minimum, so a conservative cull cannot change it);
//   plane   (thread = (group, primitive))  the horizontal slab interval [s0, s1] of the group's plane with the
//           box, in the box frame, one binary64 reciprocal per axis -- or the single crossing s of a wall
//           segment -- appended to the group's list in shared memory: 6 x ~40 pairs for BasicLidar instead
//           of 300 x ~40 ray / primitive tests with their six divisions each;
//   ray     (thread = ray)  walks its group's list: the height slab is two multiplications by the ray's
//           inv_slope, the hit is the entry of the intersection of the two intervals.  No division.
//
// A table with many azimuths and few laser angles degenerates gracefully: groups are processed in
// chunks sized to the interval list, and the work per ray approaches the unfactored test.
#pragma once

#include "sb_device.cuh"
#include "../../include/sb_lidar_table.h"

#define SB_LIDAR_THREADS 320
#define SB_LIDAR_MAX_BOXES 160
#define SB_LIDAR_MAX_EDGES 256
#define SB_LIDAR_MAX_RAYS 4096
#define SB_LIDAR_ENTRIES 768       // (group, primitive) intervals held at a time
#define SB_LIDAR_CHUNK_GROUPS 64   // groups per pass over the primitives (<= SB_LIDAR_ENTRIES / primitives)

struct __align__(16) LBox {
  double x, y, c, s, hw, hl, z0, z1;
};
struct __align__(16) LEntry {  // horizontal interval along the group's direction and the primitive's heights
  double s0, s1, zlo, zhi;
};

// Device-side ray table: records sorted by group (vertical rays last), the groups' unit directions and
// the first sorted ray of every group.
struct LTable {
  const SbRayRec* rec;         // [R] sorted
  const double2* group_u;      // [G]
  const int32_t* group_start;  // [G + 1]; rays [group_start[G], R) are vertical
  int32_t n_groups;
  double max_hlen;             // longest horizontal reach of any ray
  const double* edge_inv;      // [G, n_edges] 1 / (u x edge) or NaN (k_lidar_edge_inv); null: divide in the kernel
};

// shared-memory working set of one agent's cast (k_sense carves it out of the raster kernel's dynamic buffer)
struct __align__(16) LidarSmem {
  LBox boxes[SB_LIDAR_MAX_BOXES];
  double edges[SB_LIDAR_MAX_EDGES][4];
  LEntry entries[SB_LIDAR_ENTRIES];
  int cnt[SB_LIDAR_CHUNK_GROUPS];
  int n_boxes, n_edges;
};

// One agent's cast by the whole block (any blockDim.x >= SB_LIDAR_CHUNK_GROUPS).
__device__ __forceinline__ void d_lidar_agent(const DCfg& c, const DMap& m, const DState& s, const DObs& o,
                                              const LTable& tab, float reach_xy, int e, int a, LidarSmem* sm) {
  // reach_xy: max_hlen + margin as a float, for the block-level cull
  LBox* boxes = sm->boxes;
  double(*edges)[4] = sm->edges;
  LEntry* entries = sm->entries;
  int* cnt = sm->cnt;
  int& n_boxes = sm->n_boxes;
  int& n_edges = sm->n_edges;
  const int ai = e * c.A + a, vi = e * c.V + a;
  const double ox = s.pos_x[vi], oy = s.pos_y[vi], oz = s.pos_z[vi] + 1.0;
  if (threadIdx.x == 0) {
    n_boxes = 0;
    n_edges = 0;
  }
  if (threadIdx.x < SB_LIDAR_CHUNK_GROUPS) cnt[threadIdx.x] = 0;
  __syncthreads();
  // stage the env's other vehicles (order is irrelevant: the result is a minimum); every field is
  // loaded before the cull decision so that the block pays one round of global-load latency
  for (int j = threadIdx.x; j < c.V; j += blockDim.x) {
    const int vj = e * c.V + j;
    const bool alive = s.alive[vj] != 0;
    const double vx = s.pos_x[vj], vy = s.pos_y[vj], vz = s.pos_z[vj], vc = s.cos_h[vj], vs = s.sin_h[vj];
    const double vw = s.width[vj], vl = s.length[vj], vhh = s.height[vj];
    if (j == a || !alive) continue;
    // a hit lies within reach_xy of the origin and within the half diagonal of the box centre
    const float bx = (float)(vx - ox), by = (float)(vy - oy);
    const float fw = (float)vw, fl = (float)vl;
    const float hd = 0.5f * sqrtf(fw * fw + fl * fl);
    const float lim = reach_xy + hd + 0.25f;
    if (bx * bx + by * by > lim * lim) continue;
    const int k = atomicAdd(&n_boxes, 1);
    if (k < SB_LIDAR_MAX_BOXES) {
      const double z0 = vz + (j < c.A ? c.veh.box_z0 : 0.2);  // agents: URDF collision box; social: utils/bullet.py:169-177
      boxes[k] = LBox{vx, vy, vc, vs, 0.5 * vw, 0.5 * vl, z0, z0 + vhh};
    }
  }
  if (c.lidar_wall_height > 0.0) {
    // segments whose bounding circle reaches the disc; taken from the far end of the block so
    // that they are staged by other warps than the vehicles, at the same time
    for (int k = blockDim.x - 1 - threadIdx.x; k < m.n_edges; k += blockDim.x) {
      const double4 ed = reinterpret_cast<const double4*>(m.road_edges)[k];
      const float ax = (float)(ed.x - ox), ay = (float)(ed.y - oy), bx = (float)(ed.z - ox), by = (float)(ed.w - oy);
      const float mx = 0.5f * (ax + bx), my = 0.5f * (ay + by), hx = 0.5f * (bx - ax), hy = 0.5f * (by - ay);
      const float hl = sqrtf(hx * hx + hy * hy);
      const float reach = hl + reach_xy + 0.25f;
      if (mx * mx + my * my > reach * reach) continue;
      const int q = atomicAdd(&n_edges, 1);
      if (q < SB_LIDAR_MAX_EDGES) {
        edges[q][0] = ed.x;
        edges[q][1] = ed.y;
        edges[q][2] = ed.z;
        edges[q][3] = ed.w;
      }
    }
  }
  __syncthreads();
  const int nb = n_boxes < SB_LIDAR_MAX_BOXES ? n_boxes : SB_LIDAR_MAX_BOXES;
  const int ne = n_edges < SB_LIDAR_MAX_EDGES ? n_edges : SB_LIDAR_MAX_EDGES;
  if (threadIdx.x == 0 && (n_boxes > SB_LIDAR_MAX_BOXES || n_edges > SB_LIDAR_MAX_EDGES))
    s.status[ai] |= SB_STATUS_LIDAR_OVERFLOW;
  const int nprim = nb + ne;
  const double inf = __longlong_as_double(0x7ff0000000000000LL);
  const double miss = c.lidar_miss_inf ? inf : 0.0;
  const int R = c.lidar_rays, G = tab.n_groups;
  int gpc = SB_LIDAR_ENTRIES / (nprim > 0 ? nprim : 1);  // nprim <= 416 < SB_LIDAR_ENTRIES: at least one group
  gpc = gpc < SB_LIDAR_CHUNK_GROUPS ? gpc : SB_LIDAR_CHUNK_GROUPS;

  for (int g0 = 0; g0 < G; g0 += gpc) {
    const int ng = G - g0 < gpc ? G - g0 : gpc;
    // plane pass: the interval of every (group, primitive) pair of this chunk
    for (int p = threadIdx.x; p < ng * nprim; p += blockDim.x) {
      const int gl = p / nprim, k = p - gl * nprim;
      const double2 u = tab.group_u[g0 + gl];
      LEntry en;
      if (k < nb) {
        const LBox b = boxes[k];
        const double rx = ox - b.x, ry = oy - b.y;
        const double lo0 = rx * b.c + ry * b.s, lo1 = ry * b.c - rx * b.s;  // box frame: right, forward
        const double ld0 = u.x * b.c + u.y * b.s, ld1 = u.y * b.c - u.x * b.s;
        double s0 = -inf, s1 = inf;
        bool ok = true;
        if (ld0 == 0.0) {
          if (lo0 < -b.hw || lo0 > b.hw) ok = false;
        } else {
          const double inv = 1.0 / ld0;
          double sa = (-b.hw - lo0) * inv, sb = (b.hw - lo0) * inv;
          if (sa > sb) {
            const double t = sa;
            sa = sb;
            sb = t;
          }
          if (sa > s0) s0 = sa;
          if (sb < s1) s1 = sb;
        }
        if (ld1 == 0.0) {
          if (lo1 < -b.hl || lo1 > b.hl) ok = false;
        } else {
          const double inv = 1.0 / ld1;
          double sa = (-b.hl - lo1) * inv, sb = (b.hl - lo1) * inv;
          if (sa > sb) {
            const double t = sa;
            sa = sb;
            sb = t;
          }
          if (sa > s0) s0 = sa;
          if (sb < s1) s1 = sb;
        }
        // (s1 < 0, s0 > max_hlen: no ray of the group can enter the box within its reach)
        if (!ok || s0 > s1 || s1 < 0.0 || s0 > tab.max_hlen) continue;
        en = LEntry{s0, s1, b.z0, b.z1};
      } else {
        const double* ed = edges[k - nb];
        const double ex = ed[2] - ed[0], ey = ed[3] - ed[1];
        const double den = u.x * ey - u.y * ex;
        if (den == 0.0) continue;
        const double qx = ed[0] - ox, qy = ed[1] - oy;
        const double inv = 1.0 / den;
        const double sw = (qx * ey - qy * ex) * inv;
        const double uu = (qx * u.y - qy * u.x) * inv;
        if (sw < 0.0 || uu < 0.0 || uu > 1.0 || sw > tab.max_hlen) continue;
        en = LEntry{sw, sw, 0.0, c.lidar_wall_height};
      }
      entries[gl * nprim + atomicAdd(&cnt[gl], 1)] = en;
    }
    __syncthreads();
    // ray pass: every thread writes its own ray's point and hit flag
    const int s_end = tab.group_start[g0 + ng];
    for (int si = tab.group_start[g0] + threadIdx.x; si < s_end; si += blockDim.x) {
      const SbRayRec q = tab.rec[si];
      const int gl = q.group - g0;
      double sbest = inf;
      if (q.dz < 0.0) {
        const double sg = (0.0 - oz) * q.inv_slope;
        if (sg >= 0.0 && sg <= q.hlen) sbest = sg;
      }
      const LEntry* ent = entries + gl * nprim;
      const int n = cnt[gl];
      for (int i = 0; i < n; ++i) {
        const LEntry en = ent[i];
        double e0 = en.s0, e1 = en.s1;
        if (q.dz == 0.0) {
          if (oz < en.zlo || oz > en.zhi) continue;
        } else {
          double za = (en.zlo - oz) * q.inv_slope, zb = (en.zhi - oz) * q.inv_slope;
          if (za > zb) {
            const double t = za;
            za = zb;
            zb = t;
          }
          if (za > e0) e0 = za;
          if (zb < e1) e1 = zb;
        }
        if (e0 > e1 || e0 < 0.0 || e0 > q.hlen) continue;  // e0 < 0: origin inside the box, no hit
        if (e0 < sbest) sbest = e0;
      }
      const size_t w = (size_t)ai * R + q.index;
      if (sbest <= q.hlen) {
        const double2 u = tab.group_u[q.group];
        o.lidar_hits[w] = 1;
        o.lidar_points[w * 3 + 0] = ox + sbest * u.x;
        o.lidar_points[w * 3 + 1] = oy + sbest * u.y;
        o.lidar_points[w * 3 + 2] = oz + sbest * q.slope;
      } else {
        o.lidar_hits[w] = 0;
        o.lidar_points[w * 3 + 0] = miss;
        o.lidar_points[w * 3 + 1] = miss;
        o.lidar_points[w * 3 + 2] = miss;
      }
    }
    if (g0 + gpc < G) {  // another chunk: the lists are rebuilt
      __syncthreads();
      if (threadIdx.x < SB_LIDAR_CHUNK_GROUPS) cnt[threadIdx.x] = 0;
      __syncthreads();
    }
  }
  // vertical rays (no horizontal extent): the ground plane only
  for (int si = tab.group_start[G] + threadIdx.x; si < R; si += blockDim.x) {
    const SbRayRec q = tab.rec[si];
    double t = 2.0;
    if (q.dz < 0.0) {
      t = (0.0 - oz) / q.dz;
      if (!(t >= 0.0 && t <= 1.0)) t = 2.0;
    }
    const size_t w = (size_t)ai * R + q.index;
    if (t <= 1.0) {
      o.lidar_hits[w] = 1;
      o.lidar_points[w * 3 + 0] = ox;
      o.lidar_points[w * 3 + 1] = oy;
      o.lidar_points[w * 3 + 2] = oz + t * q.dz;
    } else {
      o.lidar_hits[w] = 0;
      o.lidar_points[w * 3 + 0] = miss;
      o.lidar_points[w * 3 + 1] = miss;
      o.lidar_points[w * 3 + 2] = miss;
    }
  }
}

__global__ void __launch_bounds__(SB_LIDAR_THREADS, 4)
k_lidar(DCfg c, DMap m, DState s, DObs o, LTable tab, float reach_xy) {
  __shared__ LidarSmem sm;
  int e, a;
  if (!d_work_agent(c, blockIdx.x, &e, &a)) return;
  d_lidar_agent(c, m, s, o, tab, reach_xy, e, a, &sm);
}

// ---------------------------------------------------------------------------------------------
// k_lidar_tiles: one block casts for up to SB_LIDAR_TILE agents of ONE env.
//
// The rays are fixed in the world frame (lidar.py:108-113 only moves the origin), so what a (group, vehicle)
// pair needs a division for -- the reciprocals of the group direction's components in the box frame -- is the
// same for every agent of the env; and the one of a (group, road edge) pair never changes at all
// (LTable::edge_inv, built when both the map and the ray table are on the device).  The block therefore
//   P0  stages the env's alive vehicles once,
//   P1  computes the (group, vehicle) reciprocals once, and culls vehicles / edges per agent into index lists,
//   P2  turns every (agent, group, near primitive) into its horizontal interval with multiplications only and
//       links the survivors (about one pair in sixteen) into a list per (agent, group),
//   P3  casts tile x R rays, each walking the list of its (agent, group),
// with three barriers for the whole tile instead of two per agent plus a staging round.  The arithmetic of an
// interval and of a hit is, operation for operation, the one of d_lidar_agent (a hit distance is a minimum:
// the order in which intervals are met does not matter), so results stay bit-identical to the oracle.
// A tile whose intervals do not fit SB_LIDAR_LIST falls back to d_lidar_agent, agent after agent.
#define SB_LIDAR_TILE 4
#ifndef SB_LIDAR_LIST
#define SB_LIDAR_LIST 640        // intervals held per block
#endif
#ifndef SB_LIDAR_RECIP
#define SB_LIDAR_RECIP 768       // (group, vehicle) reciprocal pairs held per block
#endif
#ifndef SB_LIDAR_TILE_BLOCKS
#define SB_LIDAR_TILE_BLOCKS 3  // 62 registers, no spills: 0.1206 -> 0.1167 ms against 4 blocks at 48 (5 at 32: 0.1282)
#endif
#define SB_LIDAR_MAX_GROUPS 64

struct __align__(16) LidarTile {
  LBox boxes[SB_LIDAR_MAX_BOXES];
  double2 recip[SB_LIDAR_RECIP];           // 1 / ld0, 1 / ld1 (NaN: the component is zero)
  LEntry entries[SB_LIDAR_LIST];
  int32_t next[SB_LIDAR_LIST];
  int32_t head[SB_LIDAR_TILE][SB_LIDAR_MAX_GROUPS];
  uint16_t near_box[SB_LIDAR_TILE][SB_LIDAR_MAX_BOXES];
  uint16_t near_edge[SB_LIDAR_TILE][SB_LIDAR_MAX_EDGES];
  int16_t box_slot[SB_LIDAR_MAX_BOXES];    // vehicle slot of a staged box
  double org[SB_LIDAR_TILE][3];
  int32_t nb[SB_LIDAR_TILE], ne[SB_LIDAR_TILE];
  int32_t total, n_boxes;
};

union LidarShared {
  LidarTile tile;
  LidarSmem one;
};

static_assert(SB_LIDAR_TILE == 4, "the flattened item index of k_lidar_tiles is written out for four agents");

__global__ void __launch_bounds__(SB_LIDAR_THREADS, SB_LIDAR_TILE_BLOCKS)
k_lidar_tiles(DCfg c, DMap m, DState s, DObs o, LTable tab, float reach_xy, int tiles_per_env, int list_cap) {
  extern __shared__ __align__(16) unsigned char lidar_raw[];
  LidarShared* sh = reinterpret_cast<LidarShared*>(lidar_raw);
  LidarTile& t = sh->tile;
  const int ne_envs = c.env_list ? c.n_list : c.E;
  const int qe = blockIdx.x / tiles_per_env;
  if (qe >= ne_envs) return;
  const int e = c.env_list ? c.env_list[qe] : qe;
  const int a0 = (blockIdx.x - qe * tiles_per_env) * SB_LIDAR_TILE;
  const int na = min(SB_LIDAR_TILE, c.A - a0);
  const int tid = threadIdx.x, T = blockDim.x;
  const int R = c.lidar_rays, G = tab.n_groups;
  const double inf = __longlong_as_double(0x7ff0000000000000LL);
  const double nan = __longlong_as_double(0x7ff8000000000000LL);

  // ---- P0: the env's alive vehicles, the tile's origins
  if (tid == 0) {
    t.total = 0;
    t.n_boxes = 0;
  }
  if (tid < SB_LIDAR_TILE) {
    t.nb[tid] = 0;
    t.ne[tid] = 0;
    if (tid < na) {
      const int vi = e * c.V + a0 + tid;
      t.org[tid][0] = s.pos_x[vi];
      t.org[tid][1] = s.pos_y[vi];
      t.org[tid][2] = s.pos_z[vi] + 1.0;
    }
  }
  for (int i = tid; i < SB_LIDAR_TILE * SB_LIDAR_MAX_GROUPS; i += T) (&t.head[0][0])[i] = -1;
  __syncthreads();
  for (int j = tid; j < c.V; j += T) {
    const int vj = e * c.V + j;
    const bool alive = s.alive[vj] != 0;
    const double vx = s.pos_x[vj], vy = s.pos_y[vj], vz = s.pos_z[vj], vc = s.cos_h[vj], vs = s.sin_h[vj];
    const double vw = s.width[vj], vl = s.length[vj], vhh = s.height[vj];
    if (!alive) continue;
    const int k = atomicAdd(&t.n_boxes, 1);  // V <= SB_LIDAR_MAX_BOXES on this path
    const double z0 = vz + (j < c.A ? c.veh.box_z0 : 0.2);  // agents: URDF collision box; social: utils/bullet.py:169-177
    t.boxes[k] = LBox{vx, vy, vc, vs, 0.5 * vw, 0.5 * vl, z0, z0 + vhh};
    t.box_slot[k] = (int16_t)j;
  }
  __syncthreads();

  // ---- P1: reciprocals per (group, vehicle); near lists per agent
  const int nbx = t.n_boxes;
  for (int p = tid; p < G * nbx; p += T) {
    const int g = p / nbx, k = p - g * nbx;
    const double2 u = tab.group_u[g];
    const LBox& b = t.boxes[k];
    const double ld0 = u.x * b.c + u.y * b.s, ld1 = u.y * b.c - u.x * b.s;
    t.recip[p] = make_double2(ld0 == 0.0 ? nan : 1.0 / ld0, ld1 == 0.0 ? nan : 1.0 / ld1);
  }
  for (int p = tid; p < na * nbx; p += T) {
    const int al = p / nbx, k = p - al * nbx;
    if (t.box_slot[k] == a0 + al) continue;  // the ego's own box
    const LBox& b = t.boxes[k];
    // a hit lies within reach_xy of the origin and within the half diagonal of the box centre
    const float bx = (float)(b.x - t.org[al][0]), by = (float)(b.y - t.org[al][1]);
    const float fw = (float)(2.0 * b.hw), fl = (float)(2.0 * b.hl);
    const float hd = 0.5f * sqrtf(fw * fw + fl * fl);
    const float lim = reach_xy + hd + 0.25f;
    if (bx * bx + by * by > lim * lim) continue;
    t.near_box[al][atomicAdd(&t.nb[al], 1)] = (uint16_t)k;
  }
  if (c.lidar_wall_height > 0.0) {
    // segments whose bounding circle reaches an agent's disc; from the far end of the block, so that other
    // warps than the ones above take them
    for (int p = T - 1 - tid; p < na * m.n_edges; p += T) {
      const int al = p / m.n_edges, k = p - al * m.n_edges;
      const double4 ed = reinterpret_cast<const double4*>(m.road_edges)[k];
      const double ox = t.org[al][0], oy = t.org[al][1];
      const float ax = (float)(ed.x - ox), ay = (float)(ed.y - oy), bx = (float)(ed.z - ox), by = (float)(ed.w - oy);
      const float mx = 0.5f * (ax + bx), my = 0.5f * (ay + by), hx = 0.5f * (bx - ax), hy = 0.5f * (by - ay);
      const float hl = sqrtf(hx * hx + hy * hy);
      const float reach = hl + reach_xy + 0.25f;
      if (mx * mx + my * my > reach * reach) continue;
      const int q = atomicAdd(&t.ne[al], 1);
      if (q < SB_LIDAR_MAX_EDGES) t.near_edge[al][q] = (uint16_t)k;
    }
  }
  __syncthreads();

  // ---- P2: intervals of every (agent, group, near primitive), the tile's items flattened
  int cnt_a[SB_LIDAR_TILE];  // near primitives of each agent (static indexing only: stays in registers)
#pragma unroll
  for (int al = 0; al < SB_LIDAR_TILE; ++al) {
    cnt_a[al] = 0;
    if (al < na) {
      const int ne = t.ne[al] < SB_LIDAR_MAX_EDGES ? t.ne[al] : SB_LIDAR_MAX_EDGES;
      cnt_a[al] = t.nb[al] + ne;
      if (tid == 0 && t.ne[al] > SB_LIDAR_MAX_EDGES) s.status[e * c.A + a0 + al] |= SB_STATUS_LIDAR_OVERFLOW;
    }
  }
  const int c1 = G * cnt_a[0], c2 = c1 + G * cnt_a[1], c3 = c2 + G * cnt_a[2], c4 = c3 + G * cnt_a[3];
  for (int p = tid; p < c4; p += T) {
    const int al = (p >= c1) + (p >= c2) + (p >= c3);
    const int base = al == 0 ? 0 : al == 1 ? c1 : al == 2 ? c2 : c3;
    const int nprim = al == 0 ? cnt_a[0] : al == 1 ? cnt_a[1] : al == 2 ? cnt_a[2] : cnt_a[3];
    const int r = p - base;
    const int nbl = t.nb[al];
    const int g = r / nprim, k = r - g * nprim;
    const double ox = t.org[al][0], oy = t.org[al][1];
    LEntry en;
    if (k < nbl) {
      const int kb = t.near_box[al][k];
      const LBox b = t.boxes[kb];
      const double2 inv = t.recip[g * nbx + kb];
      const double rx = ox - b.x, ry = oy - b.y;
      const double lo0 = rx * b.c + ry * b.s, lo1 = ry * b.c - rx * b.s;  // box frame: right, forward
      double s0 = -inf, s1 = inf;
      bool ok = true;
      if (inv.x != inv.x) {  // ld0 == 0
        if (lo0 < -b.hw || lo0 > b.hw) ok = false;
      } else {
        double sa = (-b.hw - lo0) * inv.x, sb = (b.hw - lo0) * inv.x;
        if (sa > sb) {
          const double tmp = sa;
          sa = sb;
          sb = tmp;
        }
        if (sa > s0) s0 = sa;
        if (sb < s1) s1 = sb;
      }
      if (inv.y != inv.y) {  // ld1 == 0
        if (lo1 < -b.hl || lo1 > b.hl) ok = false;
      } else {
        double sa = (-b.hl - lo1) * inv.y, sb = (b.hl - lo1) * inv.y;
        if (sa > sb) {
          const double tmp = sa;
          sa = sb;
          sb = tmp;
        }
        if (sa > s0) s0 = sa;
        if (sb < s1) s1 = sb;
      }
      if (!ok || s0 > s1 || s1 < 0.0 || s0 > tab.max_hlen) continue;
      en = LEntry{s0, s1, b.z0, b.z1};
    } else {
      const int ke = t.near_edge[al][k - nbl];
      const double4 ed = reinterpret_cast<const double4*>(m.road_edges)[ke];
      const double2 u = tab.group_u[g];
      const double ex = ed.z - ed.x, ey = ed.w - ed.y;
      double inv;
      if (tab.edge_inv) {
        inv = tab.edge_inv[(size_t)g * m.n_edges + ke];
        if (inv != inv) continue;  // den == 0
      } else {
        const double den = u.x * ey - u.y * ex;
        if (den == 0.0) continue;
        inv = 1.0 / den;
      }
      const double qx = ed.x - ox, qy = ed.y - oy;
      const double sw = (qx * ey - qy * ex) * inv;
      const double uu = (qx * u.y - qy * u.x) * inv;
      if (sw < 0.0 || uu < 0.0 || uu > 1.0 || sw > tab.max_hlen) continue;
      en = LEntry{sw, sw, 0.0, c.lidar_wall_height};
    }
    const int idx = atomicAdd(&t.total, 1);
    if (idx < list_cap) {
      t.entries[idx] = en;
      t.next[idx] = atomicExch(&t.head[al][g], idx);
    }
  }
  __syncthreads();
  if (t.total > list_cap) {  // (uniform) too many intervals for the list: one agent at a time, chunked
    __syncthreads();
    for (int al = 0; al < na; ++al) {
      d_lidar_agent(c, m, s, o, tab, reach_xy, e, a0 + al, &sh->one);
      __syncthreads();
    }
    return;
  }

  // ---- P3: rays; a thread keeps its ray record for every agent of the tile
  const double miss = c.lidar_miss_inf ? inf : 0.0;
  const int v0 = tab.group_start[G];  // first vertical ray of the sorted table
  for (int si = tid; si < R; si += T) {
    const SbRayRec q = tab.rec[si];
    const bool vertical = si >= v0;
    double2 u = make_double2(0.0, 0.0);
    if (!vertical) u = tab.group_u[q.group];
    for (int al = 0; al < na; ++al) {
      const double ox = t.org[al][0], oy = t.org[al][1], oz = t.org[al][2];
      const size_t w = (size_t)(e * c.A + a0 + al) * R + q.index;
      double px = miss, py = miss, pz = miss;
      bool hit = false;
      if (vertical) {  // no horizontal extent: the ground plane only
        double tt = 2.0;
        if (q.dz < 0.0) {
          tt = (0.0 - oz) / q.dz;
          if (!(tt >= 0.0 && tt <= 1.0)) tt = 2.0;
        }
        if (tt <= 1.0) {
          hit = true;
          px = ox;
          py = oy;
          pz = oz + tt * q.dz;
        }
      } else {
        double sbest = inf;
        if (q.dz < 0.0) {
          const double sg = (0.0 - oz) * q.inv_slope;
          if (sg >= 0.0 && sg <= q.hlen) sbest = sg;
        }
        for (int i = t.head[al][q.group]; i >= 0; i = t.next[i]) {
          const LEntry en = t.entries[i];
          double e0 = en.s0, e1 = en.s1;
          if (q.dz == 0.0) {
            if (oz < en.zlo || oz > en.zhi) continue;
          } else {
            double za = (en.zlo - oz) * q.inv_slope, zb = (en.zhi - oz) * q.inv_slope;
            if (za > zb) {
              const double tmp = za;
              za = zb;
              zb = tmp;
            }
            if (za > e0) e0 = za;
            if (zb < e1) e1 = zb;
          }
          if (e0 > e1 || e0 < 0.0 || e0 > q.hlen) continue;  // e0 < 0: origin inside the box, no hit
          if (e0 < sbest) sbest = e0;
        }
        if (sbest <= q.hlen) {
          hit = true;
          px = ox + sbest * u.x;
          py = oy + sbest * u.y;
          pz = oz + sbest * q.slope;
        }
      }
      o.lidar_hits[w] = hit ? 1 : 0;
      o.lidar_points[w * 3 + 0] = px;
      o.lidar_points[w * 3 + 1] = py;
      o.lidar_points[w * 3 + 2] = pz;
    }
  }
}

// 1 / (u x edge) of every (group, road edge) pair, NaN where the cross product is zero
__global__ void k_lidar_edge_inv(const double* __restrict__ road_edges, int n_edges, const double2* __restrict__ group_u,
                                 int n_groups, double* __restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_groups * n_edges) return;
  const int g = i / n_edges, k = i - g * n_edges;
  const double4 ed = reinterpret_cast<const double4*>(road_edges)[k];
  const double2 u = group_u[g];
  const double ex = ed.z - ed.x, ey = ed.w - ed.y;
  const double den = u.x * ey - u.y * ex;
  out[i] = den == 0.0 ? __longlong_as_double(0x7ff8000000000000LL) : 1.0 / den;
}
