// sb_raster.cuh -- OGM + DrivableAreaGridMap rasterisation (block per agent, thread per row).
//   OGMSensor / DrivableAreaGridMapSensor      smarts/core/sensors.py:675-758
//   camera (orthographic, ego-centred, ego-heading-aligned, rows flipped)
//                                               smarts/core/renderer.py:325-336, :383-386
//
// Contract (DESIGN.md "Grid maps"; the oracle evaluates it pixel by pixel): polygon vertices go
// to pixel space in binary64, are snapped to 1/256 px (round-half-even) and clamped to +-2^24; a
// pixel (c, r) is set iff its centre (256c+128, 256r+128) satisfies every edge function
//     F_e(c, r) = A_e (256c+128) + D_e (256r) + B0_e >= 0          (all integers, |F| < 2^52)
// of some polygon.  The result is a set of integers, so HOW a row's covered span is found is
// free; this file finds it without divisions in the per-row loop.  Both grids of an agent are
// produced by one pass:
//
//   cull    (thread = source)  vehicles (OGM) and map polygons (DAGM) that can touch the film;
//   vertex  (thread = polygon vertex)  world -> snapped pixel coordinates, int32 in shared memory;
//   edge    (thread = polygon edge)  A, D, B0 as integers, plus a 32.32 fixed-point line
//           x*(r) = e0 + ek r for the real column where F_e changes sign along a row;
//   rows    (thread = image row)  k = round(x*(r)) by integer multiply-adds, then ONE exact
//           evaluation of F_e(k, r) in 64-bit integers decides between k and k+-1 (the estimate
//           is within 2^-20 of the true crossing, so the answer is always one of the two); the
//           span [max lower bounds, min upper bounds] is OR-ed into the row's bit words held in
//           registers -- no atomics, no cross-thread dependence;
//   store   bit image -> 0/255 bytes through a 256-entry byte-expansion table in shared memory,
//           32 pixels per thread per 256-bit store (STG.256), rows fully coalesced.
//
// An edge that is horizontal in pixel space (A == 0) is a per-row yes/no.  A polygon whose
// crossing line leaves the +-2^29 px range inside the image (nearly horizontal edge) takes an
// exact binary64-division path instead (flag SLOW; the branch is warp-uniform).
#pragma once

#include "sb_device.cuh"

#define SB_RASTER_THREADS 256
#ifndef SB_RASTER_ROUND
#define SB_RASTER_ROUND 48  // polygons staged per round (with 5 blocks / SM: 32 / 40 / 48 / 56 / 64 -> 0.383 / 0.380 / 0.380 / 0.379 / 0.385 ms)
#endif
#ifndef SB_RASTER_BLOCKS
#define SB_RASTER_BLOCKS 5  // resident blocks per SM the register budget is set for: 48 registers, 32 B of spills; a block's life
                           // is a latency chain (cull -> stage -> spans -> stores), so the SM wants many of them: 3 / 4 / 5 / 6
                           // blocks -> 0.442 / 0.394 / 0.380 / 0.395 ms (6 needs 40 registers: 72 B of spills)
#endif
#define SB_RASTER_SLICE_ROUND 32  // ... per round of a slice of k_raster_slices (small grids: four staging areas per block)
#ifndef SB_RASTER_SMALL_ROWS
#define SB_RASTER_SMALL_ROWS 64  // polygons up to this many rows go warp-per-polygon, taller ones row-per-thread
#endif
#ifndef SB_RASTER_STORE_UNROLL
#define SB_RASTER_STORE_UNROLL 4  // steps of the bit -> byte store loop in flight
#endif
#define SB_RASTER_CLAMP 16777216.0        // 2^24 units = 65536 px
#define SB_RASTER_FAST_LIMIT 536870912.0  // 2^29

struct __align__(16) REdgeI {
  int32_t A, D;      // F = A*(256c+128) + D*(256r) + B0
  long long B0;
  long long e0, ek;  // 32.32 fixed point: (x* + 0.5) at r = 0, and d x*/d r
};
struct __align__(8) RPolyI {
  int16_t r0, r1;  // rows whose centre can be covered (r1 < r0: nothing)
  uint8_t n;       // edges
  uint8_t img;     // 0 = OGM, 1 = DAGM
  uint8_t slow;    // exact-division path
  uint8_t pad;
};

struct Cam {
  double px, py, c, s, res, halfw, halfh;
};

// Grid geometry, derived once on the host (sb_create) and passed by value: plain constant-bank
// operands in the kernel instead of selects on the config re-evaluated inside its loops.
struct RasterDims {
  int32_t W0, H0, W1, H1;        // 0 when the grid is disabled
  int32_t wpr0, wpr1;            // 32-pixel words per row
  int32_t stride0, stride1;      // words per bit-image row in shared memory: wpr rounded up to 8, + 1 (bank skew)
  int32_t rows0, rows1;          // rows per pass: min(H, slice_threads)
  int32_t n_veh, n_src;          // OGM sources (vehicle slots), all sources (+ map polygons)
  int32_t bits_off, n_bit_words; // byte offset of bit image 0 in a slice's buffer; words of both images
  int32_t slice_threads, slice_bytes;  // threads per agent (256: block per agent) and bytes of one slice's buffer
  float rad0, rad1;              // world-space cull radii: the film's half diagonal + 1 m
};

// staging area of one round of polygons
template <int T, int R>  // T threads cull into the list, R polygons are staged per round
struct RasterStageT {
  static constexpr int ROUND = R;
  REdgeI edges[R][SB_MAX_POLY_VERTS];
  RPolyI hdr[R];
  int32_t list[T + R];
  int32_t big[R];  // staged polygons tall enough for the row-per-thread pass
  int32_t n_list;
  int32_t n_big[2];  // by round parity: the other one is zeroed while a round is being rasterised
  int32_t wide;      // a kept polygon has more than 4 vertices: stage with 8 lanes per polygon instead of 4
};

using RasterStage = RasterStageT<SB_RASTER_THREADS, SB_RASTER_ROUND>;

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

struct RasterJob {
  Cam cam[2];          // [0] OGM, [1] DAGM
  int W[2], H[2];
  bool on[2];
  int n_veh;           // sources [0, n_veh) are the env's vehicles, [n_veh, n_veh + n_polys) map polygons
  int n_src;
};

// Stage phase: slot = (list entry, vertex k).  The G slots of a polygon (4 when every kept polygon
// has at most 4 vertices -- vehicles, and the maps built by road_map.py -- else 8) are G
// consecutive lanes of one warp, so the polygon never leaves the register file: world -> snapped pixel coordinates
// per lane, neighbour vertex / signed area / bounding box by shuffles inside the group of 8, then
// each lane writes the record of edge k -> k+1 (integer edge function + 32.32 crossing line).
template <int G, int T, class ST>
__device__ __forceinline__ void d_stage_phase(const DCfg& c, const DMap& m, const DState& s, int e,
                                              const RasterJob& job, ST* st, int first, int count, int par,
                                              int tid) {
  const int lane = tid & 31, gbase = lane & ~(G - 1);
  for (int it = 0; it * T < count * G; ++it) {
    const int slot = tid + it * T;
    const int li = slot / G, k = slot % G;
    int n = 0, img = 0;
    double x = 0.0, y = 0.0;
    if (li < count) {
      const int src = st->list[first + li];
      if (src < job.n_veh) {
        n = 4;
        if (k < 4) {
          const int vj = e * c.V + src;
          const double cc = s.cos_h[vj], ss = s.sin_h[vj];
          const double hw = 0.5 * s.width[vj], hl = 0.5 * s.length[vj];
          const double sa = (k == 1 || k == 2) ? hw : -hw, sb = (k >= 2) ? hl : -hl;
          x = s.pos_x[vj] + (sa * cc - sb * ss);
          y = s.pos_y[vj] + (sa * ss + sb * cc);
        }
      } else {
        img = 1;
        const int p = src - job.n_veh;
        const int p0 = m.poly_start[p];
        n = m.poly_start[p + 1] - p0;
        if (k < n) {
          x = m.poly_x[p0 + k];
          y = m.poly_y[p0 + k];
        }
      }
    }
    int U = 0, V = 0;
    if (k < n) {
      double Ud, Vd;
      d_to_pixel(job.cam[img], x, y, &Ud, &Vd);
      U = (int32_t)Ud;
      V = (int32_t)Vd;
    }
    // lanes past the last vertex mirror vertex 0: neutral for min / max, zero area term
    const int U0 = __shfl_sync(SB_FULL_MASK, U, gbase), V0 = __shfl_sync(SB_FULL_MASK, V, gbase);
    if (k >= n) {
      U = U0;
      V = V0;
    }
    const int nxt = (k + 1 >= n) ? gbase : lane + 1;
    const int Uj = __shfl_sync(SB_FULL_MASK, U, nxt), Vj = __shfl_sync(SB_FULL_MASK, V, nxt);
    long long area2 = (k < n) ? (long long)U * Vj - (long long)Uj * V : 0ll;
    int umin = U, umax = U, vmin = V, vmax = V;
#pragma unroll
    for (int off = 1; off < G; off <<= 1) {
      area2 += __shfl_xor_sync(SB_FULL_MASK, area2, off);
      umin = min(umin, __shfl_xor_sync(SB_FULL_MASK, umin, off));
      umax = max(umax, __shfl_xor_sync(SB_FULL_MASK, umax, off));
      vmin = min(vmin, __shfl_xor_sync(SB_FULL_MASK, vmin, off));
      vmax = max(vmax, __shfl_xor_sync(SB_FULL_MASK, vmax, off));
    }
    const int W = job.W[img], H = job.H[img];
    int r0 = (vmin - 128 + 255) >> 8, r1 = (vmax - 128) >> 8;  // rows with vmin <= 256r+128 <= vmax
    const int c0 = (umin - 128 + 255) >> 8, c1 = (umax - 128) >> 8;
    r0 = max(r0, 0);
    r1 = min(r1, H - 1);
    const bool valid = li < count && n >= 3 && area2 != 0 && r0 <= r1 && c1 >= 0 && c0 <= W - 1;
    bool slow = false;
    if (valid && k < n) {
      const int sg = area2 > 0 ? 1 : -1;
      const int dxe = Uj - U, dye = Vj - V;
      REdgeI q;
      q.A = -sg * dye;
      q.D = sg * dxe;
      q.B0 = (long long)sg * ((long long)dxe * (128 - V) + (long long)dye * U);
      q.e0 = 0;
      q.ek = 0;
      if (q.A != 0) {
        const double A = (double)q.A;
        const double e0 = -((double)(q.B0 + 128ll * q.A)) / (256.0 * A), ek = -(double)q.D / A;
        if (fabs(e0) + fabs(ek) * (double)H < SB_RASTER_FAST_LIMIT) {
          q.e0 = __double2ll_rn((e0 + 0.5) * 4294967296.0);
          q.ek = __double2ll_rn(ek * 4294967296.0);
        } else {
          slow = true;
        }
      }
      st->edges[li][k] = q;
    }
    // any edge of the polygon on the exact-division path puts the whole polygon there
    const unsigned slow_mask = __ballot_sync(SB_FULL_MASK, slow);
    if (k == 0 && li < count) {
      RPolyI P;
      P.r0 = (int16_t)(valid ? r0 : 1);
      P.r1 = (int16_t)(valid ? r1 : 0);
      P.n = (uint8_t)n;
      P.img = (uint8_t)img;
      P.slow = (uint8_t)(((slow_mask >> gbase) & ((1u << G) - 1u)) != 0u);
      P.pad = 0;
      st->hdr[li] = P;
      if (valid && r1 - r0 + 1 > SB_RASTER_SMALL_ROWS) st->big[atomicAdd(&st->n_big[par], 1)] = li;
    }
  }
}

// Column span [c0, c1] of staged polygon `li` on row r (c0 > c1: empty).
template <class ST>
__device__ __forceinline__ void d_poly_span(const ST* st, int li, const RPolyI P, int r, int W, int* c0o,
                                            int* c1o) {
  const int r256 = r << 8;
  const REdgeI* e = st->edges[li];
  int c0 = 0, c1 = W - 1;
  if (!P.slow) {
    for (int i = 0; i < P.n; ++i) {
      const REdgeI q = e[i];
      const long long base = (long long)q.D * (long long)r256 + q.B0;
      if (q.A == 0) {  // horizontal edge: the row is in or out (same branch for every lane)
        if (base < 0) c1 = -1;
        continue;
      }
      int k = (int)((q.e0 + q.ek * (long long)r) >> 32);
      k = min(max(k, -2), 1025);
      const long long F = (long long)q.A * (long long)(k * 256 + 128) + base;
      if (q.A > 0) c0 = max(c0, k + (F < 0 ? 1 : 0));  // smallest c with F >= 0
      else c1 = min(c1, k - (F < 0 ? 1 : 0));          // largest c with F >= 0
    }
  } else {  // exact binary64 division (|numerator| < 2^52: floor/ceil of the rounded quotient is exact)
    double d0 = 0.0, d1 = (double)(W - 1);
    for (int i = 0; i < P.n; ++i) {
      const REdgeI q = e[i];
      const double A = (double)q.A, bb = (double)q.B0 + (double)q.D * (double)r256;
      if (A > 0.0) d0 = fmax(d0, ceil((-bb - 128.0 * A) / (256.0 * A)));
      else if (A < 0.0) d1 = fmin(d1, floor((bb + 128.0 * A) / (-256.0 * A)));
      else if (bb < 0.0) d1 = -1.0;
    }
    c0 = (int)d0;
    c1 = (int)d1;
  }
  *c0o = c0;
  *c1o = c1;
}

// OR columns [lo, hi1) (lo < hi1) into a bit-image row in shared memory.  Only the words the span touches are
// visited (a lane strip is one or two of a row's eight); atomicOr because the short-polygon pass of another warp
// may be writing the same word.
__device__ __forceinline__ void d_or_span(uint32_t* row, int lo, int hi1) {
  const int j0 = lo >> 5, j1 = (hi1 - 1) >> 5;
  const uint32_t m0 = 0xffffffffu << (lo & 31), m1 = 0xffffffffu >> (31 - ((hi1 - 1) & 31));
  atomicOr(&row[j0], j0 == j1 ? (m0 & m1) : m0);
  for (int j = j0 + 1; j <= j1; ++j) atomicOr(&row[j], j == j1 ? m1 : 0xffffffffu);
}

// Tall polygons: OR their spans on row `r` (columns [cb, cb+256)) into the row of the shared-memory bit image
// (every thread owns a row).
template <class ST>
__device__ __forceinline__ void d_big_rows(const RasterJob& job, const ST* st, int par, int r, int cb,
                                           uint32_t* row0, uint32_t* row1) {
  const int nbig = st->n_big[par];
  for (int i = 0; i < nbig; ++i) {
    const int p = st->big[i];
    const RPolyI P = st->hdr[p];
    if (r < P.r0 || r > P.r1) continue;
    int c0, c1;
    d_poly_span(st, p, P, r, job.W[P.img], &c0, &c1);
    const int lo = max(c0, cb), hi1 = min(c1, cb + 255) + 1;
    if (lo >= hi1) continue;
    d_or_span(P.img == 0 ? row0 : row1, lo, hi1);
  }
}

// Short polygons (vehicles, junction slivers): GS lanes per polygon, a lane per row, GS rows at a time (a car is
// 19 to 47 rows at 0.1 m per pixel: a warp per polygon; 3 to 7 rows on a 64 x 64 grid over 50 m: eight lanes).  The span goes into the
// shared-memory bit image with atomicOr (one or two words for a vehicle).  Rows [rg, rg+T), columns [cb, cb+256).
template <int T, class ST>
__device__ __forceinline__ void d_small_polys(const RasterJob& job, const ST* st, int count, int rg, int cb,
                                              uint32_t* bits0, uint32_t* bits1, int stride0, int stride1, int tid) {
  constexpr int GS = T == SB_RASTER_THREADS ? 32 : 8;
  for (int li = tid / GS; li < count; li += T / GS) {
    const RPolyI P = st->hdr[li];
    if (P.r1 < P.r0 || P.r1 - P.r0 + 1 > SB_RASTER_SMALL_ROWS) continue;
    const int ra = max((int)P.r0, rg), rb = min((int)P.r1, rg + T - 1);
    const int W = job.W[P.img];
    uint32_t* img = P.img == 0 ? bits0 - rg * stride0 : bits1 - rg * stride1;
    const int stride = P.img == 0 ? stride0 : stride1;
    for (int r = ra + (tid & (GS - 1)); r <= rb; r += GS) {
      int c0, c1;
      d_poly_span(st, li, P, r, W, &c0, &c1);
      c0 = max(c0, cb);
      c1 = min(c1, cb + 255);
      if (c0 <= c1) d_or_span(img + r * stride, c0, c1 + 1);
    }
  }
}

__device__ __forceinline__ uint32_t d_expand4(uint32_t b) {  // low 4 bits -> 4 bytes of 0/255
  return ((b * 0x00204081u) & 0x01010101u) * 255u;
}

// rows x W bit image (stride words per row in shared memory) -> byte image; one 32-pixel word
// per thread per step, written with a single 256-bit store.  A warp whose 32 words are all
// 0 or all ~0 (most of an occupancy grid) skips the table.
template <int T>
__device__ __forceinline__ void d_store_rows(const uint32_t* bits, const uint2* lut, uint8_t* dst, int rows, int W, int wpr,
                                             int stride, int tid) {
  if (W & 31) {  // a width that is not a multiple of 32 pixels: rows are not 32-byte aligned -- byte stores, cropped
    for (int i = tid; i < rows * W; i += T) {
      const int row = i / W, col = i - row * W;
      dst[i] = ((bits[row * stride + (col >> 5)] >> (col & 31)) & 1u) ? 255 : 0;
    }
    return;
  }
  const int n_words = rows * wpr;
  if ((wpr & (wpr - 1)) == 0 && wpr <= T && (n_words & (T - 1)) == 0) {
    // Common shapes (W a power of two up to 8192 px, rows * W/32 a multiple of 256): this thread's word keeps
    // its column and moves 256 / wpr rows per step, every thread is active in every step -- no index
    // carry, no tail predicate, and four steps' shared-memory loads are issued before their stores.
    const int sh = __ffs(wpr) - 1;
    const uint32_t* src = bits + (tid >> sh) * stride + (tid & (wpr - 1));
    const int dsrc = (T >> sh) * stride;
    uint8_t* p = dst + (size_t)tid * 32;
    const int iters = n_words / T;
    constexpr int kStoreUnroll = SB_RASTER_STORE_UNROLL;
#pragma unroll kStoreUnroll
    for (int k = 0; k < iters; ++k) {
      const uint32_t word = src[k * dsrc];
      uint8_t* q = p + (size_t)k * (T * 32);
      if (__all_sync(SB_FULL_MASK, word + 1u <= 1u)) {
        asm volatile("st.global.v8.b32 [%0], {%1,%1,%1,%1,%1,%1,%1,%1};" ::"l"(q), "r"(word) : "memory");
      } else {
        const uint2 a = lut[word & 0xffu], b = lut[(word >> 8) & 0xffu], c = lut[(word >> 16) & 0xffu], d = lut[word >> 24];
        asm volatile("st.global.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(q), "r"(a.x), "r"(a.y), "r"(b.x),
                     "r"(b.y), "r"(c.x), "r"(c.y), "r"(d.x), "r"(d.y)
                     : "memory");
      }
    }
    return;
  }
  // running (row, word-in-row) of this thread's word: 256 words further every step
  int row = tid / wpr, j = tid - row * wpr;
  const int drow = T / wpr, dj = T - drow * wpr;
  const uint32_t* src = bits + row * stride + j;
  const int dsrc = drow * stride + dj;
  uint8_t* p = dst + (size_t)tid * 32;
  for (int i = tid; i - tid < n_words; i += T) {
    const bool act = i < n_words;
    const uint32_t word = act ? *src : 0u;
    if (__all_sync(SB_FULL_MASK, word + 1u <= 1u)) {
      if (act)
        asm volatile("st.global.v8.b32 [%0], {%1,%1,%1,%1,%1,%1,%1,%1};" ::"l"(p), "r"(word) : "memory");
    } else if (act) {
      const uint2 a = lut[word & 0xffu], b = lut[(word >> 8) & 0xffu], c = lut[(word >> 16) & 0xffu], d = lut[word >> 24];
      asm volatile("st.global.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "r"(a.x), "r"(a.y), "r"(b.x),
                   "r"(b.y), "r"(c.x), "r"(c.y), "r"(d.x), "r"(d.y)
                   : "memory");
    }
    p += T * 32;
    j += dj;
    src += dsrc;
    if (j >= wpr) {  // carry into the next row (never taken when wpr divides 256)
      j -= wpr;
      src += stride - wpr;
    }
  }
}

#define SB_RASTER_LUT_BYTES 2048

// Dynamic shared memory: [byte-expansion table | staging area | bit image 0 | bit image 1]
// (bit images: min(H, 256) rows of RasterDims::stride words).  Sized by sb_create in sb_api.cu.
// T threads (a whole block, or one of its 256 / T slices for grids of at most 128 rows: barrier `bar` is the
// slice's own) rasterise both grids of agent (e, a).  `lut` is the block's byte-expansion table, `buf` the
// slice's [staging area | bit image 0 | bit image 1].
template <int T>
__device__ __forceinline__ void d_slice_sync(int bar) {
  if (T == SB_RASTER_THREADS) __syncthreads();
  else asm volatile("bar.sync %0, %1;" ::"r"(bar), "n"(T) : "memory");
}

template <int T, class ST>
__device__ __forceinline__ void d_raster_agent(const DCfg& c, const DMap& m, const DState& s, const DObs& o,
                                               const RasterDims& g, int e, int a, const uint2* lut, unsigned char* buf,
                                               RasterJob& sjob, int tid, int bar) {
  ST* st = reinterpret_cast<ST*>(buf);
  const int ai = e * c.A + a, vi = e * c.V + a;
  const double px = s.pos_x[vi], py = s.pos_y[vi], ch = s.cos_h[vi], sh = s.sin_h[vi];
  // the job record is indexed by image at run time: it lives in shared memory (a per-thread copy
  // would go to local memory -- 144 B x 256 threads x 16 384 CTAs of extra DRAM write traffic)
  if (tid == 0) {
    RasterJob& j = sjob;
    j.on[0] = g.W0 != 0;  // sensors.py:719-758: vehicles only, ego included
    j.on[1] = g.W1 != 0;  // sensors.py:675-716: road surface only
    j.W[0] = g.W0;
    j.H[0] = g.H0;
    j.W[1] = g.W1;
    j.H[1] = g.H1;
    j.cam[0] = Cam{px, py, ch, sh, c.ogm_res, 0.5 * c.ogm_w, 0.5 * c.ogm_h};
    j.cam[1] = Cam{px, py, ch, sh, c.dagm_res, 0.5 * c.dagm_w, 0.5 * c.dagm_h};
    j.n_veh = g.n_veh;
    j.n_src = g.n_src;
  }
  const RasterJob& job = sjob;
  const int W0 = g.W0, H0 = g.H0, W1 = g.W1, H1 = g.H1, n_veh = g.n_veh, n_src = g.n_src;
  const int Hmax = max(H0, H1), Wmax = max(W0, W1);
  const int wpr0 = g.wpr0, wpr1 = g.wpr1;
  uint32_t* bits0 = reinterpret_cast<uint32_t*>(buf + g.bits_off);
  const int stride0 = g.stride0, stride1 = g.stride1;
  uint32_t* bits1 = bits0 + g.rows0 * stride0;
  const int n_bit_words = g.n_bit_words;
  const float rad0 = g.rad0, rad1 = g.rad1;
  const float fx = (float)px, fy = (float)py;

  for (int rg = 0; rg < Hmax; rg += T) {
    const int r = rg + tid;
    for (int i = tid; i * 4 < n_bit_words; i += T)  // (the buffer is padded to 16 bytes)
      reinterpret_cast<uint4*>(bits0)[i] = make_uint4(0u, 0u, 0u, 0u);
    for (int cb = 0; cb < Wmax; cb += 256) {
      if (tid == 0) {
        st->n_list = 0;
        st->n_big[0] = 0;
        st->wide = 0;
      }
      d_slice_sync<T>(bar);
      int par = 0;
      for (int base = 0; base < n_src; base += T) {
        const int p = base + tid;
        bool keep = false;
        if (p < n_veh) {
          const int vj = e * c.V + p;
          const bool alive = s.alive[vj] != 0;  // every field loaded up front: one round of load latency
          const float dx = (float)s.pos_x[vj] - fx, dy = (float)s.pos_y[vj] - fy;
          const float fl = (float)s.length[vj], fw = (float)s.width[vj];
          const float lim = rad0 + 0.5f * sqrtf(fl * fl + fw * fw) + 0.5f;
          keep = alive && dx * dx + dy * dy <= lim * lim;
        } else if (p < n_src) {
          const int q = p - n_veh;
          const float4 bb = reinterpret_cast<const float4*>(m.poly_bbox)[q];
          const int n = m.poly_start[q + 1] - m.poly_start[q];
          keep = !(bb.x > fx + rad1 || bb.z < fx - rad1 || bb.y > fy + rad1 || bb.w < fy - rad1) && n >= 3 &&
                 n <= SB_MAX_POLY_VERTS;
        }
        if (keep) st->list[atomicAdd(&st->n_list, 1)] = p;
        if (keep && p >= n_veh && m.poly_start[p - n_veh + 1] - m.poly_start[p - n_veh] > 4) st->wide = 1;
        d_slice_sync<T>(bar);
        // stage + rasterise in rounds; a remainder < ROUND waits for the next cull pass
        const bool last = base + T >= n_src;
        int n_list = st->n_list;
        while (n_list >= ST::ROUND || (last && n_list > 0)) {
          const int count = min(n_list, ST::ROUND);
          if (st->wide) d_stage_phase<8, T, ST>(c, m, s, e, job, st, n_list - count, count, par, tid);
          else d_stage_phase<4, T, ST>(c, m, s, e, job, st, n_list - count, count, par, tid);
          d_slice_sync<T>(bar);
          if (tid == 0) st->n_big[par ^ 1] = 0;
          d_small_polys<T, ST>(job, st, count, rg, cb, bits0, bits1, stride0, stride1, tid);
          d_big_rows(job, st, par, r, cb, bits0 + tid * stride0, bits1 + tid * stride1);
          n_list -= count;
          par ^= 1;
          d_slice_sync<T>(bar);
        }
        if (!last) {  // carry the remainder over to the next cull pass
          if (tid == 0) st->n_list = n_list;
          d_slice_sync<T>(bar);
        }
      }
    }
    d_slice_sync<T>(bar);
    if (rg < H0)
      d_store_rows<T>(bits0, lut, o.ogm + (size_t)ai * W0 * H0 + (size_t)rg * W0, min(H0 - rg, T), W0, wpr0,
                      stride0, tid);
    if (rg < H1)
      d_store_rows<T>(bits1, lut, o.dagm + (size_t)ai * W1 * H1 + (size_t)rg * W1, min(H1 - rg, T), W1, wpr1,
                      stride1, tid);
    if (rg + T < Hmax) d_slice_sync<T>(bar);
  }
}

// One agent per block (grids taller than 128 rows) ...
__global__ void __launch_bounds__(SB_RASTER_THREADS, SB_RASTER_BLOCKS)
k_raster(DCfg c, DMap m, DState s, DObs o, RasterDims g) {
  extern __shared__ __align__(16) unsigned char smem[];
  __shared__ RasterJob sjob;
  uint2* lut = reinterpret_cast<uint2*>(smem);
  lut[threadIdx.x] = make_uint2(d_expand4(threadIdx.x & 15u), d_expand4(threadIdx.x >> 4));
  int e, a;
  if (!d_work_agent(c, blockIdx.x, &e, &a)) return;
  d_raster_agent<SB_RASTER_THREADS, RasterStage>(c, m, s, o, g, e, a, lut, smem + SB_RASTER_LUT_BYTES, sjob, threadIdx.x, 0);
}

// ... or 256 / T agents per block, each rasterised by its own slice of T = 32 / 64 / 128 threads (T >= rows) with
// its own staging area, bit images and named barrier: a 64 x 64 grid is 4 KB of output, and a block per agent
// would leave three quarters of its threads without a row and the SM without enough stores in flight.
#ifndef SB_RASTER_SLICE_BLOCKS
#define SB_RASTER_SLICE_BLOCKS 4  // 64 registers (5 blocks at 48: 0.437 -> 0.452 ms on the full suite, 6 at 40: 0.508)
#endif
template <int T>
__global__ void __launch_bounds__(SB_RASTER_THREADS, SB_RASTER_SLICE_BLOCKS)
k_raster_slices(DCfg c, DMap m, DState s, DObs o, RasterDims g, int n_work) {
  extern __shared__ __align__(16) unsigned char smem[];
  __shared__ RasterJob sjob[SB_RASTER_THREADS / T];
  uint2* lut = reinterpret_cast<uint2*>(smem);
  lut[threadIdx.x] = make_uint2(d_expand4(threadIdx.x & 15u), d_expand4(threadIdx.x >> 4));
  __syncthreads();
  const int slice = threadIdx.x / T, tid = threadIdx.x - slice * T;
  const int work = blockIdx.x * (SB_RASTER_THREADS / T) + slice;
  int e, a;
  if (work >= n_work || !d_work_agent(c, work, &e, &a)) return;
  d_raster_agent<T, RasterStageT<T, SB_RASTER_SLICE_ROUND>>(c, m, s, o, g, e, a, lut, smem + SB_RASTER_LUT_BYTES + (size_t)slice * g.slice_bytes, sjob[slice],
                    tid, 1 + slice);
}
