/*
 * sb_lidar_table.h -- factorisation of a lidar base-ray table (Lidar._compute_rays,
 * smarts/core/lidar.py:89-113) into azimuth groups.
 *
 * The reference's tables are products: BasicLidar = 6 azimuths x 50 laser angles, VelodyneHDL32E =
 * 36 x 24 (itertools.product(yaws, rolls), lidar.py:98-100; closed form in SURVEY.md A.6).  Every ray of
 * one azimuth lies in the same vertical plane, so its horizontal intersection with a vehicle box or a
 * road-edge wall is shared by all of that azimuth's rays; only the height test is per ray.  This header is
 * the single definition of that factorisation, included by the CUDA library (sb_api.cu, host side) and by
 * the CPU oracle (oracle/oracle.cpp): both derive the same doubles from the same table.
 *
 * A ray (dx, dy, dz) becomes
 *     group g   : unit horizontal direction (ux, uy) -- the first ray seen with that direction,
 *                 matched to 1e-12 per component;
 *     hlen      : sqrt(dx^2 + dy^2), the ray's horizontal reach;
 *     slope     : dz / hlen   (height gained per metre travelled horizontally);
 *     inv_slope : hlen / dz   (0 when dz == 0).
 * The ray is then the set  origin + s * (ux, uy, slope),  0 <= s <= hlen,  which differs from the table
 * entry by at most 1e-12 relative in direction (exactly nothing for the two reference tables up to the
 * rounding of their own sin / cos products).  A ray with hlen == 0 (straight up or down) has group -1 and
 * is tested against the ground plane only.
 *
 * Plain C, no FMA contraction assumed on the host (x86-64 baseline has none; the oracle is built with
 * -ffp-contract=off).
 */
#ifndef SB_LIDAR_TABLE_H
#define SB_LIDAR_TABLE_H

#include <math.h>
#include <stdint.h>

typedef struct SbRayRec {
  double hlen, slope, inv_slope, dz;
  int32_t group; /* -1: vertical ray */
  int32_t index; /* row of the ray in the caller's table (output position) */
} SbRayRec;

/* rec[n]: records in table order; group_u[2 * n]: unit directions of the groups found (first n_groups
 * pairs are valid).  Returns n_groups. */
static inline int sb_lidar_factor(const double *rays, int n, SbRayRec *rec, double *group_u) {
  int n_groups = 0;
  for (int i = 0; i < n; ++i) {
    const double dx = rays[i * 3 + 0], dy = rays[i * 3 + 1], dz = rays[i * 3 + 2];
    const double hlen = sqrt(dx * dx + dy * dy);
    SbRayRec q;
    q.hlen = hlen;
    q.dz = dz;
    q.index = i;
    if (!(hlen > 0.0)) {
      q.group = -1;
      q.slope = 0.0;
      q.inv_slope = 0.0;
      rec[i] = q;
      continue;
    }
    const double ux = dx / hlen, uy = dy / hlen;
    int g = -1;
    for (int k = 0; k < n_groups; ++k)
      if (fabs(group_u[2 * k] - ux) <= 1e-12 && fabs(group_u[2 * k + 1] - uy) <= 1e-12) {
        g = k;
        break;
      }
    if (g < 0) {
      g = n_groups++;
      group_u[2 * g] = ux;
      group_u[2 * g + 1] = uy;
    }
    q.group = g;
    q.slope = dz / hlen;
    q.inv_slope = dz != 0.0 ? hlen / dz : 0.0;
    rec[i] = q;
  }
  return n_groups;
}

#endif /* SB_LIDAR_TABLE_H */
