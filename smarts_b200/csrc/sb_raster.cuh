// sb_raster.cuh -- OGM + DrivableAreaGridMap rasterisation (block per agent).
//   OGMSensor / DrivableAreaGridMapSensor      smarts/core/sensors.py:675-758
//   camera (orthographic, ego-centred, ego-heading-aligned, rows flipped)
//                                               smarts/core/renderer.py:325-336, :383-386
//
// Contract (DESIGN.md "Grid maps"; identical in oracle/oracle.cpp): polygon vertices go to pixel
// space in binary64, are snapped to 1/256 px (round-half-even) and clamped to +-2^24; a pixel is
// set iff its centre (256c+128, 256r+128) satisfies every edge function of the convex polygon
// (inclusive), in exact integer arithmetic.
//
// Implementation: one thread owns one image row.  Polygons are staged in shared memory as
// per-edge linear coefficients; for each polygon whose row range contains its row the thread
// solves the <= 4 edge inequalities for the covered column span and ORs the span into the row's
// bit words held in registers -- no atomics, no cross-thread dependence.  All quantities are
// integers below 2^52 carried in binary64, where +, -, * are exact and a correctly-rounded
// division followed by floor/ceil gives the exact integer quotient (|num| < 2^52: the rounding
// error of num/den is below 1/(2 den), less than the distance of a non-integer quotient to the
// nearest integer).  The bit image then goes through shared memory so that the 0/255 byte image
// is written with fully coalesced 16-byte streaming stores.
#pragma once

#include "sb_device.cuh"

#define SB_RASTER_THREADS 256
#define SB_RASTER_MAX_POLYS 256
#define SB_RASTER_CLAMP 16777216.0  // 2^24 units = 65536 px

struct REdge {
  double A, B0, B1;  // A*px + (B0 + B1*row) >= 0
};
struct RPoly {
  int32_t r0, r1;  // first / last row whose centre can be covered (r1 < r0: nothing)
  int32_t pad0, pad1;
  REdge e[4];
};

struct Cam {
  double px, py, c, s, res, halfw, halfh;
};

__device__ __forceinline__ void d_to_pixel(const Cam& cam, double x, double y, double* U, double* V) {
  double dx = x - cam.px, dy = y - cam.py;
  double xr = dx * cam.c + dy * cam.s;
  double yr = dy * cam.c - dx * cam.s;
  double u = xr / cam.res + cam.halfw;
  double v = cam.halfh - yr / cam.res;
  double su = rint(u * 256.0), sv = rint(v * 256.0);
  const double lim = SB_RASTER_CLAMP;
  *U = su > lim ? lim : (su < -lim ? -lim : su);
  *V = sv > lim ? lim : (sv < -lim ? -lim : sv);
}

// snapped vertices -> row range + edge coefficients (all values exact integers in binary64)
__device__ __forceinline__ void d_make_rpoly(const double* U, const double* V, int n, int W, int H, RPoly* out) {
  double area2 = 0.0;
  double umin = U[0], umax = U[0], vmin = V[0], vmax = V[0];
  for (int i = 0; i < n; ++i) {
    int j = (i + 1 == n) ? 0 : i + 1;
    area2 += U[i] * V[j] - U[j] * V[i];
    umin = fmin(umin, U[i]);
    umax = fmax(umax, U[i]);
    vmin = fmin(vmin, V[i]);
    vmax = fmax(vmax, V[i]);
  }
  RPoly P;
  P.pad0 = P.pad1 = 0;
  double r0 = ceil((vmin - 128.0) / 256.0), r1 = floor((vmax - 128.0) / 256.0);
  double c0 = ceil((umin - 128.0) / 256.0), c1 = floor((umax - 128.0) / 256.0);
  if (r0 < 0.0) r0 = 0.0;
  if (r1 > (double)(H - 1)) r1 = (double)(H - 1);
  if (area2 == 0.0 || r0 > r1 || c1 < 0.0 || c0 > (double)(W - 1)) {
    P.r0 = 1;
    P.r1 = 0;
  } else {
    P.r0 = (int)r0;
    P.r1 = (int)r1;
  }
  const double sg = area2 > 0.0 ? 1.0 : -1.0;
  for (int i = 0; i < 4; ++i) {
    if (i < n) {
      int j = (i + 1 == n) ? 0 : i + 1;
      double dxe = U[j] - U[i], dye = V[j] - V[i];
      P.e[i].A = -(sg * dye);
      P.e[i].B1 = sg * (dxe * 256.0);
      P.e[i].B0 = sg * (dxe * (128.0 - V[i]) + dye * U[i]);
    } else {
      P.e[i].A = 0.0;
      P.e[i].B0 = 0.0;
      P.e[i].B1 = 0.0;
    }
  }
  *out = P;
}

// OR the staged polygons' spans on row `r` (columns [cb, cb+256)) into the thread's row bits
__device__ __forceinline__ void d_row_spans(const RPoly* polys, int np, int r, int cb, int W, unsigned long long* w4) {
  const double rr = (double)r;
  const double wlast = (double)(W - 1);
  for (int p = 0; p < np; ++p) {
    const RPoly& P = polys[p];
    if (r < P.r0 || r > P.r1) continue;
    double c0 = 0.0, c1 = wlast;
    bool empty = false;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const double A = P.e[i].A;
      const double bb = P.e[i].B0 + P.e[i].B1 * rr;
      if (A > 0.0) {
        double lo = ceil((-bb - 128.0 * A) / (256.0 * A));
        c0 = fmax(c0, lo);
      } else if (A < 0.0) {
        double hi = floor((bb + 128.0 * A) / (-256.0 * A));
        c1 = fmin(c1, hi);
      } else if (bb < 0.0) {
        empty = true;
      }
    }
    if (empty || c0 > c1) continue;
    int lo = (int)c0 - cb, hi = (int)c1 - cb;
    if (lo < 0) lo = 0;
    if (hi > 255) hi = 255;
    if (lo > hi) continue;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      int wl = lo - 64 * k, wh = hi - 64 * k;
      if (wl < 0) wl = 0;
      if (wh > 63) wh = 63;
      if (wl <= wh) w4[k] |= (~0ull << wl) & (~0ull >> (63 - wh));
    }
  }
}

__device__ __forceinline__ uint32_t d_expand4(uint32_t b) {  // low 4 bits -> 4 bytes of 0/255
  return ((b * 0x00204081u) & 0x01010101u) * 255u;
}

// rows x W bit image (wpr words per row) -> byte image, 16 pixels per thread per step
__device__ __forceinline__ void d_store_rows(const uint32_t* bits, uint8_t* dst, int n_pixels) {
  const int n16 = n_pixels >> 4;
  uint4* out = reinterpret_cast<uint4*>(dst);
  for (int i = threadIdx.x; i < n16; i += SB_RASTER_THREADS) {
    uint32_t word = bits[i >> 1];
    uint32_t hw = (i & 1) ? (word >> 16) : word;
    uint4 v;
    v.x = d_expand4(hw & 15u);
    v.y = d_expand4((hw >> 4) & 15u);
    v.z = d_expand4((hw >> 8) & 15u);
    v.w = d_expand4((hw >> 12) & 15u);
    __stcs(out + i, v);
  }
}

// One grid (OGM when `vehicles`, DAGM otherwise) of agent `ai`.
template <bool VEHICLES>
__device__ void d_raster_grid(const DCfg& c, const DMap& m, const DState& s, int e, const Cam& cam, int W, int H,
                              uint8_t* dst, uint32_t* bits, RPoly* polys, int* n_staged) {
  const int wpr = W >> 5;
  const int n_src = VEHICLES ? c.V : m.n_polys;
  const float rad = 0.5f * (float)cam.res * sqrtf((float)W * (float)W + (float)H * (float)H) + 1.0f;
  const float fx = (float)cam.px, fy = (float)cam.py;
  for (int rg = 0; rg < H; rg += SB_RASTER_THREADS) {
    const int r = rg + threadIdx.x;
    const int rows = H - rg < SB_RASTER_THREADS ? H - rg : SB_RASTER_THREADS;
    for (int cb = 0; cb < W; cb += 256) {
      unsigned long long w4[4] = {0ull, 0ull, 0ull, 0ull};
      for (int base = 0; base < n_src; base += SB_RASTER_MAX_POLYS) {
        __syncthreads();  // previous round's readers are done
        if (threadIdx.x == 0) *n_staged = 0;
        __syncthreads();
        const int p = base + threadIdx.x;
        if (p < n_src) {
          double U[4], V[4];
          int n = 0;
          if (VEHICLES) {
            const int vj = e * c.V + p;
            if (s.alive[vj]) {
              const double cc = s.cos_h[vj], ss = s.sin_h[vj];
              const double hw = 0.5 * s.width[vj], hl = 0.5 * s.length[vj];
              const double sa[4] = {-hw, hw, hw, -hw}, sbv[4] = {-hl, -hl, hl, hl};
              n = 4;
#pragma unroll
              for (int k = 0; k < 4; ++k) {
                double x = s.pos_x[vj] + (sa[k] * cc - sbv[k] * ss);
                double y = s.pos_y[vj] + (sa[k] * ss + sbv[k] * cc);
                d_to_pixel(cam, x, y, &U[k], &V[k]);
              }
            }
          } else {
            const float* bb = m.poly_bbox + (size_t)p * 4;
            if (!(bb[0] > fx + rad || bb[2] < fx - rad || bb[1] > fy + rad || bb[3] < fy - rad)) {
              const int st = m.poly_start[p];
              n = m.poly_start[p + 1] - st;
              if (n < 3 || n > 4) n = 0;
              for (int k = 0; k < n; ++k) d_to_pixel(cam, m.poly_x[st + k], m.poly_y[st + k], &U[k], &V[k]);
            }
          }
          if (n) {
            RPoly P;
            d_make_rpoly(U, V, n, W, H, &P);
            if (P.r0 <= P.r1) polys[atomicAdd(n_staged, 1)] = P;  // order is irrelevant (OR)
          }
        }
        __syncthreads();
        if (r < H) d_row_spans(polys, *n_staged, r, cb, W, w4);
      }
      if (r < H) {
        uint32_t* row = bits + threadIdx.x * wpr + (cb >> 5);
        const int nw = (W - cb) >= 256 ? 8 : ((W - cb) >> 5);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          if (2 * k < nw) row[2 * k] = (uint32_t)w4[k];
          if (2 * k + 1 < nw) row[2 * k + 1] = (uint32_t)(w4[k] >> 32);
        }
      }
    }
    __syncthreads();
    d_store_rows(bits, dst + (size_t)rg * W, rows * W);
    __syncthreads();
  }
}

// dynamic shared memory: bit image of one 256-row group, min(H,256) * W / 8 bytes
__global__ void __launch_bounds__(SB_RASTER_THREADS) k_raster(DCfg c, DMap m, DState s, DObs o) {
  extern __shared__ uint32_t bits[];
  __shared__ RPoly polys[SB_RASTER_MAX_POLYS];
  __shared__ int n_staged;
  const int ai = blockIdx.x;
  const int e = ai / c.A, a = ai - e * c.A, vi = e * c.V + a;
  const double px = s.pos_x[vi], py = s.pos_y[vi], ch = s.cos_h[vi], sh = s.sin_h[vi];
  if (c.ogm_enable) {  // sensors.py:719-758: vehicles only, ego included
    Cam cam{px, py, ch, sh, c.ogm_res, 0.5 * c.ogm_w, 0.5 * c.ogm_h};
    d_raster_grid<true>(c, m, s, e, cam, c.ogm_w, c.ogm_h, o.ogm + (size_t)ai * c.ogm_w * c.ogm_h, bits, polys,
                        &n_staged);
  }
  if (c.dagm_enable) {  // sensors.py:675-716: road surface only
    Cam cam{px, py, ch, sh, c.dagm_res, 0.5 * c.dagm_w, 0.5 * c.dagm_h};
    d_raster_grid<false>(c, m, s, e, cam, c.dagm_w, c.dagm_h, o.dagm + (size_t)ai * c.dagm_w * c.dagm_h, bits,
                         polys, &n_staged);
  }
}
