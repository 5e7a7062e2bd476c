// common.cuh — shared helpers for libgsb200 (sm_100a).  No torch types anywhere in csrc/.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include <atomic>

#include "../../include/gsb200.h"

#define GS_FULL 0xffffffffu

void gs_set_error(const char *fmt, ...);
extern std::atomic<unsigned long long> g_gs_launches;  // gs_launch_count(): diagnostics, bumped by GS_LAUNCH_CHECK

#define GS_REQUIRE(cond, msg)                 \
  do {                                        \
    if (!(cond)) {                            \
      gs_set_error("%s: %s", __func__, msg);  \
      return GS_ERR_INVALID;                  \
    }                                         \
  } while (0)

#define GS_LAUNCH_CHECK()                                                             \
  do {                                                                                \
    g_gs_launches.fetch_add(1, std::memory_order_relaxed);                            \
    cudaError_t e__ = cudaGetLastError();                                             \
    if (e__ != cudaSuccess) {                                                         \
      gs_set_error("%s: CUDA error %s (%s:%d)", __func__, cudaGetErrorString(e__), __FILE__, __LINE__); \
      return GS_ERR_CUDA;                                                             \
    }                                                                                 \
  } while (0)

#define GS_CUDA(call)                                                                 \
  do {                                                                                \
    cudaError_t e__ = (call);                                                         \
    if (e__ != cudaSuccess) {                                                         \
      gs_set_error("%s: %s failed: %s", __func__, #call, cudaGetErrorString(e__));    \
      return GS_ERR_CUDA;                                                             \
    }                                                                                 \
  } while (0)

static inline int gs_div_up(int64_t a, int64_t b) { return (int)((a + b - 1) / b); }
static inline size_t gs_align256(size_t x) { return (x + 255) & ~(size_t)255; }

// camera in kernel-parameter form
struct CamDev {
  float T[12];  // rows 0..2 of T_camera_world
  float fx, fy, cx, cy;
  int W, H;
  float near_plane, far_plane;
  float limx, limy;  // frustum clamp on x/z, y/z
  float cam_pos[3];
};

struct CfgDev {
  float thr, amax, tmin, blur;
  int antialias;
};

static inline CamDev gs_make_cam(const gs_camera *c, const gs_raster_config *cfg) {
  CamDev d;
  for (int i = 0; i < 12; ++i) d.T[i] = c->T_camera_world[i];
  d.fx = c->fx; d.fy = c->fy; d.cx = c->cx; d.cy = c->cy;
  d.W = c->width; d.H = c->height;
  d.near_plane = c->near_plane; d.far_plane = c->far_plane;
  d.limx = cfg->frustum_clamp * 0.5f * (float)c->width / c->fx;
  d.limy = cfg->frustum_clamp * 0.5f * (float)c->height / c->fy;
  const float *T = c->T_camera_world;
  for (int k = 0; k < 3; ++k) d.cam_pos[k] = -(T[k] * T[3] + T[4 + k] * T[7] + T[8 + k] * T[11]);
  return d;
}

static inline CfgDev gs_make_cfg(const gs_raster_config *c) {
  CfgDev d;
  d.thr = c->alpha_threshold; d.amax = c->clamp_max_alpha;
  d.tmin = 1.0f - c->saturate_threshold; d.blur = c->blur_cov; d.antialias = c->antialias;
  return d;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(GS_FULL, v, o);
  return v;
}

// exp_lerp(t, a, b) = m + log(lerp(e^(a-m), e^(b-m), t)), m = max(a, b)  (reference util/misc.py:56-59, with torch.lerp's
// two branches).  One of the two exponentials is always e^0 = 1 and exp_lerp(t, a, a) = a + log(1) = a: both are taken
// as shortcuts, which is bit-identical to evaluating them.  Shared by gs_accumulate_stats and gs_exchange_fold so the
// sequential and the data-parallel statistics are the same arithmetic.
__device__ __forceinline__ float gs_exp_lerp(float t, float a, float b) {
  if (a == b) return a;
  float m = fmaxf(a, b);
  float ea = (a >= b) ? 1.0f : expf(a - m), eb = (b >= a) ? 1.0f : expf(b - m);
  float l = t < 0.5f ? ea + t * (eb - ea) : eb - (eb - ea) * (1.0f - t);
  return m + logf(l);
}

// ---- exclusive scan of int32 (scan.cu) -------------------------------------------------------
// out[i] = sum_{j<i} in[idx ? idx[j] : j]; total (nullable) receives the full sum.
// n_dev (nullable) bounds the element count on the device (<= n).  partials: >= gs_scan_partials(n) ints.
// as_flag != 0 scans the predicate (value > 0) instead of the value (stream-compaction ranks).
int64_t gs_scan_partials(int64_t n);
int gs_exclusive_scan_i32(const int32_t *in, const int32_t *idx, int32_t *out, int64_t n, const int32_t *n_dev,
                          int32_t *partials, int32_t *total, cudaStream_t st, int as_flag = 0);

// ---- decoupled look-back (single-pass chained scan) ------------------------------------------------------
// A status word carries a 30-bit count and a 2-bit flag in ONE 32-bit store, so no fence is needed:
// 0 = not yet published, LOCAL = this tile's own aggregate, INCL = inclusive prefix up to and including the tile.
// Tiles take their index from an atomic ticket, so every predecessor of a running tile has already started.
constexpr uint32_t GS_LB_LOCAL = 1u << 30, GS_LB_INCL = 2u << 30, GS_LB_VAL = (1u << 30) - 1u;

__device__ __forceinline__ uint32_t gs_ld_status(const uint32_t *p) { return *(const volatile uint32_t *)p; }
__device__ __forceinline__ void gs_st_status(uint32_t *p, uint32_t v) { *(volatile uint32_t *)p = v; }

// Whole-warp look-back over a scalar status array: returns the exclusive prefix of `tile` (all lanes).
__device__ __forceinline__ uint32_t gs_lookback_warp(const uint32_t *status, int tile) {
  int lane = threadIdx.x & 31;
  uint32_t prefix = 0;
  for (int base = tile - 1; base >= 0; base -= 32) {
    int idx = base - lane;
    uint32_t s = GS_LB_INCL;  // lanes before tile 0 act as an inclusive zero
    if (idx >= 0) {
      do { s = gs_ld_status(status + idx); } while ((s >> 30) == 0);
    }
    unsigned incl = __ballot_sync(GS_FULL, (s >> 30) == 2u);
    int first = incl ? (__ffs(incl) - 1) : 32;
    uint32_t c = (lane <= first) ? (s & GS_LB_VAL) : 0u;
    prefix += __reduce_add_sync(GS_FULL, c);
    if (incl) break;
  }
  return prefix;
}

// ---- tile info (tilemap.cu): per-row tile record (16 B) and per-row tile count ---------------------
struct TileInfo {
  uint4 *recs;       // n
  int32_t *counts;   // n
  size_t bytes;
};
TileInfo gs_tile_info_carve(void *base, int64_t n, int width, int height);
