// footprint.cuh — tile footprint of one projected splat row ("Spec T" in DESIGN.md).
//
// Every operation is an explicit round-to-nearest IEEE intrinsic (no FMA contraction, IEEE
// division and square root), so the tile lists are BIT-EXACT against the CPU oracle, which
// evaluates the same sequence with gcc -ffp-contract=off.  The support of a splat is the set of
// pixel centres where opacity·exp(-q/2) >= alpha_threshold, i.e. q <= k2 = 2·ln(opacity/thr):
// an opacity-aware bound that is exact for the rasteriser's own alpha cut, so tile culling never
// changes the image.  ln is evaluated by a fixed polynomial (det_logf) for the same reason.
#pragma once
#include "common.cuh"

__device__ __forceinline__ float det_logf(float x) {
  uint32_t b = __float_as_uint(x);
  int e = (int)(b >> 23) - 127;
  float m = __uint_as_float((b & 0x007fffffu) | 0x3f800000u);
  if (m > 1.41421356f) { m = __fmul_rn(m, 0.5f); e += 1; }
  float f = __fsub_rn(m, 1.0f);
  float s = __fdiv_rn(f, __fadd_rn(2.0f, f));
  float s2 = __fmul_rn(s, s);
  float p = 0.11111111f;
  p = __fadd_rn(__fmul_rn(p, s2), 0.14285715f);
  p = __fadd_rn(__fmul_rn(p, s2), 0.2f);
  p = __fadd_rn(__fmul_rn(p, s2), 0.33333334f);
  p = __fadd_rn(__fmul_rn(p, s2), 1.0f);
  float t = __fmul_rn(__fmul_rn(2.0f, s), p);
  float el = __fmul_rn((float)e, 0.69314718f);
  return __fadd_rn(el, t);
}

struct Footprint {
  int tx0, tx1, ty0, ty1;
  float mx, my, k2, syx, g, inv_ca, dy_right;  // per-row span constants, see row_span
  bool valid;
};

__device__ __forceinline__ Footprint row_footprint(float mx, float my, float ca, float cb, float cc, float op,
                                                   float depth, float thr, int W, int H) {
  Footprint fp;
  fp.valid = false;
  fp.tx0 = fp.ty0 = 0; fp.tx1 = fp.ty1 = -1;
  fp.mx = mx; fp.my = my; fp.k2 = 0.f; fp.syx = 0.f; fp.g = 0.f; fp.inv_ca = 0.f; fp.dy_right = 0.f;
  if (!(op > thr)) return fp;
  if (!(depth > 0.0f) || !(depth < 3.0e38f)) return fp;
  if (!(fabsf(mx) < 1.0e6f) || !(fabsf(my) < 1.0e6f)) return fp;
  float detc = __fsub_rn(__fmul_rn(ca, cc), __fmul_rn(cb, cb));
  if (!(detc > 0.0f) || !(ca > 0.0f) || !(cc > 0.0f)) return fp;
  float k2 = __fmul_rn(2.0f, det_logf(__fdiv_rn(op, thr)));
  k2 = __fadd_rn(__fmul_rn(k2, 1.0001f), 1.0e-5f);
  float ex = __fsqrt_rn(__fdiv_rn(__fmul_rn(k2, cc), detc));
  float ey = __fsqrt_rn(__fdiv_rn(__fmul_rn(k2, ca), detc));
  if (!(ex < 1.0e6f) || !(ey < 1.0e6f)) return fp;
  float x_lo = __fsub_rn(mx, ex), x_hi = __fadd_rn(mx, ex);
  float y_lo = __fsub_rn(my, ey), y_hi = __fadd_rn(my, ey);
  int TX = (W + GS_TILE - 1) / GS_TILE, TY = (H + GS_TILE - 1) / GS_TILE;
  int tx0 = (int)ceilf(__fdiv_rn(__fsub_rn(x_lo, 15.5f), 16.0f));
  int tx1 = (int)floorf(__fdiv_rn(__fsub_rn(x_hi, 0.5f), 16.0f));
  int ty0 = (int)ceilf(__fdiv_rn(__fsub_rn(y_lo, 15.5f), 16.0f));
  int ty1 = (int)floorf(__fdiv_rn(__fsub_rn(y_hi, 0.5f), 16.0f));
  tx0 = max(tx0, 0); ty0 = max(ty0, 0);
  tx1 = min(tx1, TX - 1); ty1 = min(ty1, TY - 1);
  if (tx0 > tx1 || ty0 > ty1) return fp;
  fp.tx0 = tx0; fp.tx1 = tx1; fp.ty0 = ty0; fp.ty1 = ty1;
  fp.k2 = k2;
  // for a fixed dy the ellipse spans dx = syx·dy ± sqrt((k2 − g·dy²)·inv_ca); its rightmost point (dx = +ex) lies at
  // dy_right = −(cb/cc)·ex, its leftmost at −dy_right
  fp.syx = -__fdiv_rn(cb, ca);
  fp.g = __fdiv_rn(detc, ca);
  fp.inv_ca = __fdiv_rn(1.0f, ca);
  fp.dy_right = __fmul_rn(-__fdiv_rn(cb, cc), ex);
  fp.valid = true;
  return fp;
}

__device__ __forceinline__ float clampf_(float v, float lo, float hi) { return v < lo ? lo : (v > hi ? hi : v); }

// Tiles [txa, txb] of tile row ty whose pixel-centre rectangle the k2-ellipse reaches (empty if txa > txb).
// The ellipse cut by the band dy ∈ [dyl, dyr] is convex, so its x-range is [min x_left, max x_right] with
// x_right(dy) = syx·dy + hw(dy) concave (maximal at the band point closest to dy_right) and x_left convex:
// two square roots per tile ROW instead of a closest-point test per TILE.
__device__ __forceinline__ void row_span(const Footprint &fp, int ty, int H, int &txa, int &txb) {
  txa = 0; txb = -1;
  int py1 = min(ty * GS_TILE + GS_TILE - 1, H - 1);
  float dyl = __fsub_rn(__fadd_rn((float)(ty * GS_TILE), 0.5f), fp.my);
  float dyr = __fsub_rn(__fadd_rn((float)py1, 0.5f), fp.my);
  float dy0 = clampf_(0.0f, dyl, dyr);
  float t0 = __fmul_rn(__fmul_rn(fp.g, dy0), dy0);
  if (!(__fsub_rn(fp.k2, t0) >= 0.0f)) return;  // the band misses the ellipse
  float dyR = clampf_(fp.dy_right, dyl, dyr), dyL = clampf_(-fp.dy_right, dyl, dyr);
  float rR = __fsub_rn(fp.k2, __fmul_rn(__fmul_rn(fp.g, dyR), dyR));
  float rL = __fsub_rn(fp.k2, __fmul_rn(__fmul_rn(fp.g, dyL), dyL));
  if (rR < 0.0f) rR = 0.0f;
  if (rL < 0.0f) rL = 0.0f;
  float xr = __fadd_rn(__fmul_rn(fp.syx, dyR), __fsqrt_rn(__fmul_rn(rR, fp.inv_ca)));
  float xl = __fsub_rn(__fmul_rn(fp.syx, dyL), __fsqrt_rn(__fmul_rn(rL, fp.inv_ca)));
  float x_lo = __fadd_rn(fp.mx, xl), x_hi = __fadd_rn(fp.mx, xr);
  int a = (int)ceilf(__fmul_rn(__fsub_rn(x_lo, 15.5f), 0.0625f));
  int b = (int)floorf(__fmul_rn(__fsub_rn(x_hi, 0.5f), 0.0625f));
  txa = max(a, fp.tx0);
  txb = min(b, fp.tx1);
}

__device__ __forceinline__ int footprint_count(const Footprint &fp, int W, int H) {
  (void)W;
  if (!fp.valid) return 0;
  int c = 0;
  for (int ty = fp.ty0; ty <= fp.ty1; ++ty) {
    int a, b;
    row_span(fp, ty, H, a, b);
    c += max(0, b - a + 1);
  }
  return c;
}

// ---- per-row tile record (16 B) -------------------------------------------------------------------------
// The cull already visits every tile of a row's rect; it keeps the outcome so that the mapper never repeats the
// float tests: rect origin/shape plus, for rects of <= 64 tiles, the exact hit bitmask (bit r*w+c = tile
// (tx0+c, ty0+r)).  Rows with larger rects ("big") are re-tested per 8x8-tile super-tile by the mapper.
//   x = tx0 | ty0 << 16      y = w | h << 8 | nsuper << 16 (bit 31: big)      z,w = mask lo/hi
constexpr int GS_SUPER = 8;  // super-tile edge in tiles (8x8 tiles = 64 sub-tiles = one 64-bit mask)
constexpr uint32_t GS_REC_BIG = 0x80000000u;

struct RecView {
  int tx0, ty0, w, h, nsuper;
  bool big;
  unsigned long long mask;
};

__device__ __forceinline__ RecView rec_view(const uint4 &r) {
  RecView v;
  v.tx0 = (int)(r.x & 0xffffu); v.ty0 = (int)(r.x >> 16);
  v.w = (int)(r.y & 0xffu); v.h = (int)((r.y >> 8) & 0xffu);
  v.nsuper = (int)((r.y >> 16) & 0x7fffu);
  v.big = (r.y & GS_REC_BIG) != 0;
  v.mask = ((unsigned long long)r.w << 32) | r.z;
  return v;
}

// hit mask of a SMALL row restricted to super-tile (sx, sy): bit (ty&7)*8 + (tx&7)
__device__ __forceinline__ unsigned long long rec_submask(const RecView &v, int sx, int sy) {
  unsigned long long sub = 0ull;
  int rel = v.tx0 - GS_SUPER * sx;
  unsigned long long rowsel = (v.w >= 64) ? ~0ull : ((1ull << v.w) - 1ull);
  int r0 = max(0, GS_SUPER * sy - v.ty0), r1 = min(v.h - 1, GS_SUPER * sy + GS_SUPER - 1 - v.ty0);
  for (int r = r0; r <= r1; ++r) {
    unsigned long long bits = (v.mask >> (r * v.w)) & rowsel;
    unsigned long long seg = rel >= 0 ? (rel < 64 ? bits << rel : 0ull) : (-rel < 64 ? bits >> (-rel) : 0ull);
    sub |= (seg & 0xffull) << (GS_SUPER * ((v.ty0 + r) & (GS_SUPER - 1)));
  }
  return sub;
}

// exact hit mask of ANY valid footprint restricted to super-tile (sx, sy) (used for big rows)
__device__ __forceinline__ unsigned long long footprint_submask(const Footprint &fp, int sx, int sy, int W, int H) {
  (void)W;
  unsigned long long sub = 0ull;
  int ty_lo = max(fp.ty0, GS_SUPER * sy), ty_hi = min(fp.ty1, GS_SUPER * sy + GS_SUPER - 1);
  for (int ty = ty_lo; ty <= ty_hi; ++ty) {
    int a, b;
    row_span(fp, ty, H, a, b);
    a = max(a, GS_SUPER * sx); b = min(b, GS_SUPER * sx + GS_SUPER - 1);
    if (b >= a) {
      unsigned long long bits = ((1ull << (b - a + 1)) - 1ull) << (a & (GS_SUPER - 1));
      sub |= bits << (GS_SUPER * (ty & (GS_SUPER - 1)));
    }
  }
  return sub;
}

// Builds the record and returns the tile count.
__device__ __forceinline__ int footprint_record(const Footprint &fp, int W, int H, uint4 &rec) {
  rec = make_uint4(0u, 0u, 0u, 0u);
  if (!fp.valid) return 0;
  int w = fp.tx1 - fp.tx0 + 1, h = fp.ty1 - fp.ty0 + 1;
  int count = 0;
  int ns = (fp.tx1 / GS_SUPER - fp.tx0 / GS_SUPER + 1) * (fp.ty1 / GS_SUPER - fp.ty0 / GS_SUPER + 1);  // conservative
  if (w * h <= 64) {
    unsigned long long mask = 0ull;
    for (int ty = fp.ty0; ty <= fp.ty1; ++ty) {
      int a, b;
      row_span(fp, ty, H, a, b);
      if (b >= a) mask |= ((b - a + 1 >= 64) ? ~0ull : ((1ull << (b - a + 1)) - 1ull)) << ((ty - fp.ty0) * w + (a - fp.tx0));
    }
    count = __popcll(mask);
    if (count == 0) return 0;
    rec = make_uint4((uint32_t)fp.tx0 | ((uint32_t)fp.ty0 << 16), (uint32_t)w | ((uint32_t)h << 8) | ((uint32_t)ns << 16),
                     (uint32_t)mask, (uint32_t)(mask >> 32));
  } else {
    for (int ty = fp.ty0; ty <= fp.ty1; ++ty) {
      int a, b;
      row_span(fp, ty, H, a, b);
      count += max(0, b - a + 1);
    }
    if (count == 0) return 0;
    rec = make_uint4((uint32_t)fp.tx0 | ((uint32_t)fp.ty0 << 16), ((uint32_t)ns << 16) | GS_REC_BIG, 0u, 0u);
  }
  return count;
}
