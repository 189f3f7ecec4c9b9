// sb_lidar.cuh -- LidarSensor ray casting (block per agent, thread per ray, azimuth-sorted warps).
//   Lidar.compute_point_cloud / _compute_rays   smarts/core/lidar.py:58-134
//   LidarSensor                                 smarts/core/sensors.py:797-827
//
// Result contract (the oracle evaluates it ray by ray): closest hit of every table ray against
// the ground plane, every other alive vehicle's box and (lidar_wall_height > 0) the road-edge
// walls, by the binary64 slab / segment formulas below.  The result is a minimum over
// primitives, so WHICH primitives are tested exactly is free as long as no true hit is dropped.
// Three conservative culls, each cheaper than the one after it, leave ~2 exact tests per ray:
//
//   block   vehicles / edge segments that cannot reach the 20 m disc of the agent are never
//           staged (float, circle test);
//   warp    rays are handed to threads in azimuth order (permutation + one bounding sector per
//           32 rays, built on the host when the ray table is uploaded), so a warp's rays fan a
//           narrow sector: each lane tests one staged primitive against the sector, a ballot
//           gives the warp's candidate mask -- no list, no shared-memory traffic;
//   ray     distance from the box centre to the ray's line / both edge end points on one side
//           of it (float, a handful of FMAs) before the binary64 test with its divisions.
//
// Every thread writes its own ray's point and hit flag as soon as it is known (no block barrier
// after the ray loop: a warp that is done retires instead of waiting for the slowest one).
#pragma once

#include "sb_device.cuh"

#define SB_LIDAR_THREADS 320
#define SB_LIDAR_MAX_BOXES 160
#define SB_LIDAR_MAX_EDGES 256
#define SB_LIDAR_MAX_RAYS 4096

struct LBox {
  double x, y, c, s, hw, hl, z0, z1;
};
struct __align__(16) LBoxF {  // centre relative to the ray origin, bounding radius + margin
  float bx, by, rad, pad;
};
struct __align__(16) LEdgeF {  // end points relative to the ray origin
  float ax, ay, bx, by;
};

// Bounding sector of 32 azimuth-sorted rays: unit centre direction (ucx, ucy) and the half
// angle as (cos, sin); cos < -1 marks a chunk too wide to cull by direction.
struct LSector {
  float ucx, ucy, ca, sa;
};

#define SB_LIDAR_MARGIN 0.01f

// disc (bx, by, rad) vs sector: distance to the nearer boundary ray <= rad (conservative)
__device__ __forceinline__ bool d_sector_disc(const LSector& q, float bx, float by, float rad) {
  if (q.ca < -1.5f) return true;
  const float a = bx * q.ucx + by * q.ucy, p = fabsf(bx * q.ucy - by * q.ucx);
  const float dist_out = p * q.ca - a * q.sa, t = a * q.ca + p * q.sa;
  return dist_out <= rad && t >= -rad;
}

// segment vs sector (an intersection of two half planes for half angles < 90 deg): rejected when
// both end points violate the same half plane, or both lie behind the origin
__device__ __forceinline__ bool d_sector_segment(const LSector& q, const LEdgeF& e) {
  if (q.ca < -1.5f) return true;
  const float mu = SB_LIDAR_MARGIN;
  const float wpx = q.ucx * q.ca - q.ucy * q.sa, wpy = q.ucx * q.sa + q.ucy * q.ca;  // uc rotated by +a
  const float wmx = q.ucx * q.ca + q.ucy * q.sa, wmy = q.ucy * q.ca - q.ucx * q.sa;  // uc rotated by -a
  const float m0 = wmx * e.ay - wmy * e.ax, m1 = wmx * e.by - wmy * e.bx;            // >= 0 inside
  const float p0 = wpx * e.ay - wpy * e.ax, p1 = wpx * e.by - wpy * e.bx;            // <= 0 inside
  const float f0 = q.ucx * e.ax + q.ucy * e.ay, f1 = q.ucx * e.bx + q.ucy * e.by;
  if (m0 < -mu && m1 < -mu) return false;
  if (p0 > mu && p1 > mu) return false;
  if (f0 < -mu && f1 < -mu) return false;
  return true;
}

__global__ void __launch_bounds__(SB_LIDAR_THREADS, 4)
k_lidar(DCfg c, DMap m, DState s, DObs o, const double* __restrict__ rays, const int32_t* __restrict__ order,
        const LSector* __restrict__ sectors, float reach_xy) {
  // reach_xy: longest horizontal extent of any table ray (+ margin), for the block-level cull
  __shared__ LBox boxes[SB_LIDAR_MAX_BOXES];
  __shared__ LBoxF boxf[SB_LIDAR_MAX_BOXES];
  __shared__ double edges[SB_LIDAR_MAX_EDGES][4];
  __shared__ LEdgeF edgef[SB_LIDAR_MAX_EDGES];
  __shared__ float2 edged[SB_LIDAR_MAX_EDGES];
  __shared__ int n_boxes, n_edges;
  const int ai = blockIdx.x;
  const int e = ai / c.A, a = ai - e * c.A, vi = e * c.V + a;
  const int lane = threadIdx.x & 31;
  const double ox = s.pos_x[vi], oy = s.pos_y[vi], oz = s.pos_z[vi] + 1.0;
  if (threadIdx.x == 0) {
    n_boxes = 0;
    n_edges = 0;
  }
  // this thread's first ray: its table entries are requested before the staging below, so that their
  // latency is paid together with the vehicle / edge loads instead of after the barrier
  const int R = c.lidar_rays;
  const int si_first = threadIdx.x < R ? (int)threadIdx.x : R - 1;
  int r = order[si_first];
  LSector sec = sectors[si_first >> 5];
  double dx = rays[r * 3 + 0], dy = rays[r * 3 + 1], dz = rays[r * 3 + 2];
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
      const double z0 = vz + (j < c.A ? 0.1 : 0.2);
      boxes[k] = LBox{vx, vy, vc, vs, 0.5 * vw, 0.5 * vl, z0, z0 + vhh};
      boxf[k] = LBoxF{bx, by, hd * 1.0001f + SB_LIDAR_MARGIN, 0.f};
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
        edgef[q] = LEdgeF{ax, ay, bx, by};
        const float il = hl > 1e-12f ? 1.0f / hl : 0.f;  // unit direction (0 for a degenerate segment: no cull)
        edged[q] = make_float2(hx * il, hy * il);
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

  for (int base = 0; base < R; base += blockDim.x) {  // blockDim is a multiple of 32: chunks stay warp-aligned
    const int si = base + threadIdx.x;
    const bool act = si < R;
    if (base > 0) {  // (the first chunk's entries were loaded above)
      r = order[act ? si : R - 1];
      sec = sectors[(act ? si : R - 1) >> 5];
      dx = rays[r * 3 + 0];
      dy = rays[r * 3 + 1];
      dz = rays[r * 3 + 2];
    }
    // unit horizontal direction of this ray (float); a vertical ray takes no per-ray cull
    const float fdx = (float)dx, fdy = (float)dy;
    const float len = sqrtf(fdx * fdx + fdy * fdy);
    const bool flat = len > 1e-6f;
    const float inv = flat ? 1.0f / len : 0.f;
    const float ux = fdx * inv, uy = fdy * inv;
    double tbest = 2.0;
    if (dz < 0.0) {
      double t = (0.0 - oz) / dz;
      if (t >= 0.0 && t <= 1.0) tbest = t;
    }
    for (int kb = 0; kb < nb; kb += 32) {
      bool cand = false;
      if (kb + lane < nb) {
        const LBoxF f = boxf[kb + lane];
        cand = d_sector_disc(sec, f.bx, f.by, f.rad);
      }
      unsigned mask = __ballot_sync(SB_FULL_MASK, cand);
      while (mask) {
        const int k = kb + __ffs(mask) - 1;
        mask &= mask - 1;
        const LBoxF f = boxf[k];
        if (!act || (flat && fabsf(f.bx * uy - f.by * ux) > f.rad)) continue;
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
        if (!ok || t0 > t1 || t0 < 0.0 || t0 > 1.0) continue;  // origin inside a box: no hit
        if (t0 < tbest) tbest = t0;
      }
    }
    for (int kb = 0; kb < ne; kb += 32) {
      bool cand = false;
      if (kb + lane < ne) cand = d_sector_segment(sec, edgef[kb + lane]);
      unsigned mask = __ballot_sync(SB_FULL_MASK, cand);
      while (mask) {
        const int k = kb + __ffs(mask) - 1;
        mask &= mask - 1;
        const LEdgeF f = edgef[k];
        if (!act) continue;
        if (flat) {
          // both end points strictly on one side of the ray's line
          const float s0 = ux * f.ay - uy * f.ax, s1 = ux * f.by - uy * f.bx;
          if ((s0 > SB_LIDAR_MARGIN && s1 > SB_LIDAR_MARGIN) || (s0 < -SB_LIDAR_MARGIN && s1 < -SB_LIDAR_MARGIN))
            continue;
          // the ray meets the segment's line behind the origin or beyond its reach:
          // distance along the ray = tn / dn (tn: origin-to-line distance, dn: sine of the angle)
          const float2 ed = edged[k];
          float tn = f.ax * ed.y - f.ay * ed.x, dn = ux * ed.y - uy * ed.x;
          if (dn < 0.f) {
            tn = -tn;
            dn = -dn;
          }
          if ((ed.x != 0.f || ed.y != 0.f) && (tn < -SB_LIDAR_MARGIN || tn > len * dn + SB_LIDAR_MARGIN)) continue;
        }
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
    }
    // each warp writes its own rays as soon as they are done: no block barrier, no staging of t;
    // the agent's 7.5 KB of points fill whole L2 lines within the block's lifetime, so the stride-6
    // ray order of an azimuth-sorted warp costs L2 transactions, not DRAM traffic
    if (act) {
      const size_t q = (size_t)ai * R + r;
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
}
