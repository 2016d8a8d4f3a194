// cull_rows.cuh -- which tiles of a Gaussian's tile rect can it reach alpha >= 1/255 in?
// Answered per TILE ROW from the CullRec conic, in O(rows) instead of O(tiles).
//
// Region (see cull_record(), surfel_preprocess.cu):  {f <= 0} u disk, f the point conic of
// {rho3d <= c2}: f = u^2 + v^2 - c2 w^2 with (u, v, w) linear in the pixel, i.e. the inside of a
// double cone cut by the image plane -- an ellipse, or the two convex insides of a hyperbola.
// For a fixed y, f is a quadratic in x:  A x^2 + 2 b x + c,  b = B y + D,  c = C y^2 + 2 E y + F,
// with discriminant disc(y) = b^2 - A c = p y^2 + 2 q y + r:
//   A > 0:  the row's set is [x-(y), x+(y)] where disc >= 0, else empty;
//   A < 0:  it is everything outside the gap (x_lo(y), x_hi(y)) where disc > 0, else everything.
// Over a band of rows [ya, yb] the extreme of each root curve lies at ya, at yb, or at a point
// where the conic has a vertical tangent (dx/dy = 0), so sampling the curves at those y (and at
// the tips, where disc = 0) gives the exact x-extent of region /\ band (A > 0) or the exact common
// gap (A < 0).  All in fp64 from the normalised fp64 coefficients; the band is the tile row
// dilated by half a pixel plus a small margin, like the per-block test of tile_hits_kernel.
#pragma once
#include "common.cuh"

namespace b200gs {

struct RowCull {
  int mode;              // 0 nothing, 1 everything, 2 conic with A > 0, 3 conic with A < 0
  double A, B, D;        // b(y) = B y + D
  double p, q, r;        // disc(y) = p y^2 + 2 q y + r
  double ys[4];          // special sample rows: vertical-tangent points, then tips / disc vertex
  int ns;
  double cx, cy, r2;     // low-pass disk
};

constexpr double ROW_MARGIN = 0.5 + 1e-3;   // half a pixel (pixel centres vs. the dilated rectangle) + slack

__device__ __forceinline__ void rowcull_init(const CullRec& c, RowCull& rc) {
  rc.cx = c.cx; rc.cy = c.cy; rc.r2 = c.r2; rc.ns = 0;
  rc.A = c.A; rc.B = c.B; rc.D = c.D;
  rc.p = rc.q = rc.r = 0.0;
  if (c.flag != 0.0f) { rc.mode = (c.flag == CULL_NOTHING) ? 0 : 1; return; }
  const double A = c.A, B = c.B, C = c.C, D = c.D, E = c.E, F = c.F;
  rc.mode = 1;
  if (!(fabs(A) > 1e-9) || !(fabs(C) > 1e-9)) return;          // (near-)degenerate in x or y: keep everything
  const double p = B * B - A * C;
  if (!(fabs(p) > 1e-12)) return;                                // parabola
  rc.p = p; rc.q = B * D - A * E; rc.r = D * D - A * F;
  // vertical tangents: disc_x(x) = p x^2 + 2 qx x + rx = 0, y = -(B x + E) / C
  const double qx = B * E - C * D, rx = E * E - C * F;
  const double dx = qx * qx - p * rx;
  if (dx >= 0.0) {
    const double s = sqrt(dx);
    const double x1 = (-qx - s) / p, x2 = (-qx + s) / p;
    rc.ys[rc.ns++] = -(B * x1 + E) / C;
    rc.ys[rc.ns++] = -(B * x2 + E) / C;
  }
  if (A > 0.0) {
    // tips (y extremes of the conic): disc(y) = 0
    const double dy = rc.q * rc.q - p * rc.r;
    if (dy >= 0.0) {
      const double s = sqrt(dy);
      rc.ys[rc.ns++] = (-rc.q - s) / p;
      rc.ys[rc.ns++] = (-rc.q + s) / p;
    }
    rc.mode = 2;
  } else {
    rc.ys[rc.ns++] = -rc.q / p;                                  // vertex of disc(y): where the gap is most likely to close
    rc.mode = 3;
  }
  for (int i = 0; i < rc.ns; ++i)
    if (!(fabs(rc.ys[i]) < 1e300)) { rc.mode = 1; return; }      // NaN / inf: keep everything
}

// Tile columns of tile row `ty` (clipped to [x0, x1)) the Gaussian can reach, as up to two disjoint
// ascending inclusive ranges [a1, b1], [a2, b2] from the conic (empty when a > b) plus the disk's
// range [d0, d1] (its tiles may overlap the conic ranges; the caller skips those).
__device__ __forceinline__ void rowcull_ranges(const RowCull& rc, int ty, int x0, int x1, int& a1, int& b1, int& a2, int& b2,
                                               int& d0, int& d1) {
  a1 = 0; b1 = -1; a2 = 0; b2 = -1; d0 = 0; d1 = -1;
  if (rc.mode == 0) return;
  if (rc.mode == 1) { a1 = x0; b1 = x1 - 1; return; }
  const double ya = 16.0 * ty - ROW_MARGIN, yb = 16.0 * ty + 15.0 + ROW_MARGIN;
  // disk
  {
    const double dy = fmax(fmax(ya - rc.cy, rc.cy - yb), 0.0);
    const double h2 = rc.r2 - dy * dy;
    if (h2 >= 0.0) {
      const double h = sqrt(h2);
      d0 = max(x0, (int)ceil((rc.cx - h - 15.0 - ROW_MARGIN) * 0.0625));
      d1 = min(x1 - 1, (int)floor((rc.cx + h + ROW_MARGIN) * 0.0625));
    }
  }
  double lo = 1e300, hi = -1e300;     // mode 2: hull of the row sets; mode 3: gap (lo = max x_lo, hi = min x_hi)
  bool everything = false;
  if (rc.mode == 3) { lo = -1e300; hi = 1e300; }
  const double invA = 1.0 / rc.A;
#pragma unroll
  for (int k = 0; k < 6; ++k) {
    double y;
    bool special = false;
    if (k == 0) y = ya;
    else if (k == 1) y = yb;
    else {
      if (k - 2 >= rc.ns) continue;
      y = rc.ys[k - 2];
      if (!(y > ya && y < yb)) continue;
      special = true;
    }
    double dsc = (rc.p * y + 2.0 * rc.q) * y + rc.r;
    const double b = rc.B * y + rc.D;
    if (rc.mode == 2) {
      if (dsc < 0.0) { if (!special) continue; dsc = 0.0; }   // special points lie on the conic by construction
      const double s = sqrt(dsc);
      lo = fmin(lo, (-b - s) * invA);
      hi = fmax(hi, (-b + s) * invA);
    } else {
      if (!(dsc > 0.0)) { everything = true; continue; }
      const double s = sqrt(dsc);
      lo = fmax(lo, (-b + s) * invA);      // A < 0: (-b + s)/A is the smaller root
      hi = fmin(hi, (-b - s) * invA);
    }
  }
  if (rc.mode == 2) {
    if (lo > hi) return;                                   // the conic misses this row
    a1 = max(x0, (int)ceil(fmax((lo - 15.0 - ROW_MARGIN) * 0.0625, -1e9)));
    b1 = min(x1 - 1, (int)floor(fmin((hi + ROW_MARGIN) * 0.0625, 1e9)));
  } else {
    if (everything || !(lo < hi)) { a1 = x0; b1 = x1 - 1; return; }
    // kept: x <= lo and x >= hi
    a1 = x0;
    b1 = min(x1 - 1, (int)floor(fmin(fmax((lo + ROW_MARGIN) * 0.0625, -1e9), 1e9)));
    a2 = max(x0, (int)ceil(fmax(fmin((hi - 15.0 - ROW_MARGIN) * 0.0625, 1e9), -1e9)));
    b2 = x1 - 1;
    if (a2 <= b1) { b1 = x1 - 1; a2 = 0; b2 = -1; }       // the gap is narrower than a tile: everything
  }
}

// number of tiles the ranges cover, the disk's tiles counted once
__device__ __forceinline__ int rowcull_count(int a1, int b1, int a2, int b2, int d0, int d1) {
  int n = max(0, b1 - a1 + 1) + max(0, b2 - a2 + 1);
  for (int x = d0; x <= d1; ++x)
    if (!((x >= a1 && x <= b1) || (x >= a2 && x <= b2))) ++n;
  return n;
}

}  // namespace b200gs
