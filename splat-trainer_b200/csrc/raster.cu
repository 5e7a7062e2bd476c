// raster.cu — tile rasteriser forward / backward (render_projected at reference
// scene/mlp_scene.py:418 and its autograd backward via trainer/trainer.py:512).
//
// Design (B200): one CTA of 64 threads per 16x16 tile, every thread owns a 2x2 pixel quad, so a
// warp covers 16x8 pixels.  The tile's sorted splat records are staged through shared memory in
// batches of 64 (one gathered 36-B record per thread, conic pre-scaled by -log2(e)/2 so the inner
// loop is 3 FMA + one ex2.approx per pixel, and a per-splat log2(thr/opacity) bound that rejects a
// pixel before the SFU is touched).  Forward: front-to-back, block vote (__syncthreads_and) ends
// the tile when all 256 pixels are saturated.  Backward: back-to-front, the 4 pixels of a thread
// are summed in registers, then a 13-shuffle transpose-reduction leaves the 12 per-splat sums on
// 12 lanes, which issue one coalesced vector of fp32 reductions (RED) per (warp, splat) — instead
// of 12 x 5 butterfly shuffles and 12 separate atomics.  No tensor cores: FMA/SFU/LSU bound.
#include "raster_common.cuh"

// raster_rec.cu: the record-staged asynchronous kernels (taken when aux->records is given)
int gs_raster_fwd_rec(const gs_raster_config *h_cfg, const gs_raster_aux *aux, const float *depth, const int32_t *o2p,
                      const int32_t *ranges, int W, int H, float *image, float *final_T, int32_t *n_contrib, float *visibility,
                      float *median, cudaStream_t st);
int gs_raster_bwd_rec(const gs_raster_config *h_cfg, const gs_raster_aux *aux, const int32_t *o2p, const int32_t *ranges,
                      int W, int H, const float *final_T, const int32_t *n_contrib, const float *d_image, float *acc,
                      int acc_stride, float *vis_plane, cudaStream_t st);

struct TileCtx {
  int begin, n;
  float fx0, fy0;
  bool in[4];
};

__device__ __forceinline__ TileCtx tile_ctx(const int32_t *__restrict__ ranges, int W, int H, int TX) {
  TileCtx c;
  int tile = blockIdx.x;
  int tx = tile % TX, ty = tile / TX;
  int lx = threadIdx.x & 7, ly = threadIdx.x >> 3;
  int px0 = tx * GS_TILE + 2 * lx, py0 = ty * GS_TILE + 2 * ly;
  c.fx0 = (float)px0 + 0.5f;
  c.fy0 = (float)py0 + 0.5f;
  c.in[0] = px0 < W && py0 < H;
  c.in[1] = px0 + 1 < W && py0 < H;
  c.in[2] = px0 < W && py0 + 1 < H;
  c.in[3] = px0 + 1 < W && py0 + 1 < H;
  c.begin = ranges[2 * tile];
  c.n = ranges[2 * tile + 1] - c.begin;
  return c;
}

__device__ __forceinline__ int64_t pixel_index(int p, int W, int TX) {
  int tile = blockIdx.x;
  int tx = tile % TX, ty = tile / TX;
  int lx = threadIdx.x & 7, ly = threadIdx.x >> 3;
  int px = tx * GS_TILE + 2 * lx + (p & 1), py = ty * GS_TILE + 2 * ly + (p >> 1);
  return (int64_t)py * W + px;
}

// stage one splat record into shared memory (conic pre-scaled for exp2)
// Returns the (slightly loosened) half-height of the splat's support {alpha >= thr} around its mean row: used only to
// skip whole warps (16x8 pixel halves of the tile) the splat cannot reach; the exact per-pixel test still decides.
template <int C>
__device__ __forceinline__ float stage_splat(int slot, int id, const float *__restrict__ g2d, const float *__restrict__ feat,
                                             float thr, float4 *sA, float4 *sB, float (*sF)[RT]) {
  const float2 *g = reinterpret_cast<const float2 *>(g2d + (int64_t)GS_G2D_STRIDE * id);
  float2 a = __ldg(g), b = __ldg(g + 1), c = __ldg(g + 2);
  float op = c.y;
  // Spec R blend test, decided in the exponent domain with IEEE round-to-nearest ops only (bit-identical in the
  // oracle): a pixel blends iff pcut <= p' <= 0, p' = log2(e)·power, pcut = log2(thr/op)  (<=> op·exp(power) >= thr)
  float pcut = (op > 0.f) ? __fmul_rn(det_logf(__fdiv_rn(thr, op)), LOG2E) : 1.0f;
  sA[slot] = make_float4(a.x, a.y, __fmul_rn(-0.5f * LOG2E, b.x), __fmul_rn(-LOG2E, b.y));
  sB[slot] = make_float4(__fmul_rn(-0.5f * LOG2E, c.x), op, pcut, __int_as_float(id));
  const float *f = feat + (int64_t)C * id;
#pragma unroll
  for (int ch = 0; ch < C; ++ch) sF[ch][slot] = __ldg(f + ch);
  // y half-extent of {q <= k2}: sqrt(k2 * Sigma_yy), Sigma_yy = ca / det(conic), k2 = 2 ln(op/thr) = -2 ln2 pcut
  float det = b.x * c.x - b.y * b.y;
  float k2 = fmaxf(-2.0f * LN2 * pcut, 0.f);
  return (det > 0.f) ? sqrtf(k2 * b.x / det) * 1.001f + 0.01f : 3.0e38f;
}

// per-warp candidate mask of a staged batch: bit k set iff staged splat k can reach this warp's 8 pixel rows
__device__ __forceinline__ void publish_masks(bool staged, float my, float ey, float row0, unsigned (*s_mask)[2]) {
  // warp r covers pixel-centre rows row0 + 8r + 0.5 .. row0 + 8r + 7.5
  bool t0 = staged && (my + ey >= row0 + 0.5f) && (my - ey <= row0 + 7.5f);
  bool t1 = staged && (my + ey >= row0 + 8.5f) && (my - ey <= row0 + 15.5f);
  unsigned m0 = __ballot_sync(GS_FULL, t0), m1 = __ballot_sync(GS_FULL, t1);
  if ((threadIdx.x & 31) == 0) {
    s_mask[0][threadIdx.x >> 5] = m0;
    s_mask[1][threadIdx.x >> 5] = m1;
  }
}

template <int C, bool VIS, bool MEDIAN>
__global__ void __launch_bounds__(RT) k_raster_fwd(CfgDev cfg, const float *__restrict__ g2d, const float *__restrict__ feat,
                                                   const float *__restrict__ depth, const int32_t *__restrict__ o2p,
                                                   const int32_t *__restrict__ ranges, int W, int H, int TX,
                                                   float *__restrict__ image, float *__restrict__ final_T,
                                                   int32_t *__restrict__ n_contrib, float *__restrict__ vis,
                                                   float *__restrict__ median) {
  __shared__ float4 sA[RT];
  __shared__ float4 sB[RT];
  __shared__ float sF[C][RT];
  __shared__ float sD[MEDIAN ? RT : 1];
  __shared__ unsigned s_mask[2][2];
  TileCtx tc = tile_ctx(ranges, W, H, TX);
  const int wid = threadIdx.x >> 5;
  const float row0 = (float)((blockIdx.x / TX) * GS_TILE);
  float T[4] = {1.f, 1.f, 1.f, 1.f};
  float col[4][C];
  float med[4] = {0.f, 0.f, 0.f, 0.f};
  int last[4] = {0, 0, 0, 0};
  bool done[4];
#pragma unroll
  for (int p = 0; p < 4; ++p) {
    done[p] = !tc.in[p];
#pragma unroll
    for (int ch = 0; ch < C; ++ch) col[p][ch] = 0.f;
  }
  const float fx[2] = {tc.fx0, tc.fx0 + 1.0f}, fy[2] = {tc.fy0, tc.fy0 + 1.0f};

  for (int base = 0; base < tc.n; base += RT) {
    bool all_done = done[0] && done[1] && done[2] && done[3];
    if (__syncthreads_and(all_done)) break;  // also orders the previous batch's reads before the new writes
    int j = base + threadIdx.x;
    float ey = 0.f, my = 0.f;
    if (j < tc.n) {
      int id = o2p[tc.begin + j];
      ey = stage_splat<C>(threadIdx.x, id, g2d, feat, cfg.thr, sA, sB, sF);
      my = sA[threadIdx.x].y;
      if (MEDIAN) sD[threadIdx.x] = depth[id];
    }
    publish_masks(j < tc.n, my, ey, row0, s_mask);
    __syncthreads();
    unsigned long long cand = (unsigned long long)s_mask[wid][0] | ((unsigned long long)s_mask[wid][1] << 32);
    while (cand) {
      if (!VIS && all_done) break;
      int k = __ffsll((long long)cand) - 1;
      cand &= cand - 1;
      float4 A = sA[k], B = sB[k];
      float dx[2] = {fx[0] - A.x, fx[1] - A.x}, dy[2] = {fy[0] - A.y, fy[1] - A.y};
      float ax[2] = {__fmul_rn(__fmul_rn(A.z, dx[0]), dx[0]), __fmul_rn(__fmul_rn(A.z, dx[1]), dx[1])};
      float cy[2] = {__fmul_rn(__fmul_rn(B.x, dy[0]), dy[0]), __fmul_rn(__fmul_rn(B.x, dy[1]), dy[1])};
      float bx[2] = {__fmul_rn(A.w, dx[0]), __fmul_rn(A.w, dx[1])};
      float vsum = 0.f;
#pragma unroll
      for (int p = 0; p < 4; ++p) {
        float pw = __fmaf_rn(bx[p & 1], dy[p >> 1], __fadd_rn(ax[p & 1], cy[p >> 1]));
        if (!done[p] && pw <= 0.f && pw >= B.z) {
          float a = fminf(cfg.amax, B.y * ex2_approx(pw));
          float w = a * T[p];
#pragma unroll
          for (int ch = 0; ch < C; ++ch) col[p][ch] += w * sF[ch][k];
          float Tn = T[p] * (1.0f - a);
          if (MEDIAN && T[p] >= 0.5f && Tn < 0.5f) med[p] = sD[k];
          T[p] = Tn;
          last[p] = base + k + 1;
          if (Tn < cfg.tmin) done[p] = true;
          vsum += w;
        }
      }
      if (VIS) {
        if (__any_sync(GS_FULL, vsum > 0.f)) {
          float s = warp_sum(vsum);
          if ((threadIdx.x & 31) == 0) atomicAdd(vis + __float_as_int(B.w), s);
        }
      } else {
        all_done = done[0] && done[1] && done[2] && done[3];
      }
    }
  }
#pragma unroll
  for (int p = 0; p < 4; ++p) {
    if (!tc.in[p]) continue;
    int64_t pix = pixel_index(p, W, TX);
#pragma unroll
    for (int ch = 0; ch < C; ++ch) image[pix * C + ch] = col[p][ch];
    final_T[pix] = T[p];
    n_contrib[pix] = last[p];
    if (MEDIAN) median[pix] = med[p];
  }
}

template <int C>
__global__ void __launch_bounds__(RT) k_raster_bwd(CfgDev cfg, const float *__restrict__ g2d, const float *__restrict__ feat,
                                                   const int32_t *__restrict__ o2p, const int32_t *__restrict__ ranges,
                                                   int W, int H, int TX, const float *__restrict__ final_T,
                                                   const int32_t *__restrict__ n_contrib, const float *__restrict__ d_image,
                                                   float *__restrict__ acc, int acc_stride, float *__restrict__ vis_plane) {
  constexpr int NV = (9 + C <= 12) ? 12 : 16;
  __shared__ float4 sA[RT];
  __shared__ float4 sB[RT];
  __shared__ float sF[C][RT];
  __shared__ int s_max[RT / 32];
  __shared__ unsigned s_mask[2][2];
  TileCtx tc = tile_ctx(ranges, W, H, TX);
  const int wid = threadIdx.x >> 5;
  const float row0 = (float)((blockIdx.x / TX) * GS_TILE);
  float T[4], last_a[4] = {0.f, 0.f, 0.f, 0.f};
  // per pixel: dL/dC and the SCALAR projections onto it of the colour composited behind the current splat (R) and of
  // the last blended splat's colour (lastq): dL/dα needs only (f − rec)·dL/dC, so the per-channel recurrence of the
  // usual formulation collapses to one scalar recurrence (fewer registers, fewer FMAs)
  float dC[4][C], R[4], lastq[4];
  int nc[4];
  int mynmax = 0;
#pragma unroll
  for (int p = 0; p < 4; ++p) {
    T[p] = 1.f; nc[p] = 0; R[p] = 0.f; lastq[p] = 0.f;
#pragma unroll
    for (int ch = 0; ch < C; ++ch) dC[p][ch] = 0.f;
    if (tc.in[p]) {
      int64_t pix = pixel_index(p, W, TX);
      T[p] = final_T[pix];
      nc[p] = n_contrib[pix];
#pragma unroll
      for (int ch = 0; ch < C; ++ch) dC[p][ch] = d_image[pix * C + ch];
    }
    mynmax = max(mynmax, nc[p]);
  }
  int wmax = __reduce_max_sync(GS_FULL, mynmax);
  if ((threadIdx.x & 31) == 0) s_max[threadIdx.x >> 5] = wmax;
  __syncthreads();
  int top = 0;
#pragma unroll
  for (int i = 0; i < RT / 32; ++i) top = max(top, s_max[i]);
  top = min(top, tc.n);
  const float fx[2] = {tc.fx0, tc.fx0 + 1.0f}, fy[2] = {tc.fy0, tc.fy0 + 1.0f};
  const int lane = threadIdx.x & 31;

  for (; top > 0; top -= RT) {
    int lo = max(0, top - RT), cnt = top - lo;
    __syncthreads();
    float ey = 0.f, my = 0.f;
    if ((int)threadIdx.x < cnt) {
      int id = o2p[tc.begin + lo + threadIdx.x];
      ey = stage_splat<C>(threadIdx.x, id, g2d, feat, cfg.thr, sA, sB, sF);
      my = sA[threadIdx.x].y;
    }
    publish_masks((int)threadIdx.x < cnt, my, ey, row0, s_mask);
    __syncthreads();
    unsigned long long cand = (unsigned long long)s_mask[wid][0] | ((unsigned long long)s_mask[wid][1] << 32);
    {  // this warp's own pixels stop at wmax: later list positions cannot contribute to it
      int wn = wmax - lo;
      if (wn < RT) cand &= (wn > 0) ? ((1ull << wn) - 1ull) : 0ull;
    }
    while (cand) {
      int k = 63 - __clzll((long long)cand);
      cand ^= 1ull << k;
      int pos = lo + k;
      float4 A = sA[k], B = sB[k];
      float dx[2] = {fx[0] - A.x, fx[1] - A.x}, dy[2] = {fy[0] - A.y, fy[1] - A.y};
      float ax[2] = {__fmul_rn(__fmul_rn(A.z, dx[0]), dx[0]), __fmul_rn(__fmul_rn(A.z, dx[1]), dx[1])};
      float cy[2] = {__fmul_rn(__fmul_rn(B.x, dy[0]), dy[0]), __fmul_rn(__fmul_rn(B.x, dy[1]), dy[1])};
      float bx[2] = {__fmul_rn(A.w, dx[0]), __fmul_rn(A.w, dx[1])};
      float v[NV];
#pragma unroll
      for (int i = 0; i < NV; ++i) v[i] = 0.f;
      bool touched = false;
#pragma unroll
      for (int p = 0; p < 4; ++p) {
        float pw = __fmaf_rn(bx[p & 1], dy[p >> 1], __fadd_rn(ax[p & 1], cy[p >> 1]));
        if (pos < nc[p] && pw <= 0.f && pw >= B.z) {
          float G = ex2_approx(pw);
          float a = B.y * G;
          bool clamped = a > cfg.amax;
          a = fminf(a, cfg.amax);
          touched = true;
          float Tp = __fdividef(T[p], 1.0f - a);
          T[p] = Tp;
          float w = a * Tp;
          float q = 0.f;
#pragma unroll
          for (int ch = 0; ch < C; ++ch) {
            q += sF[ch][k] * dC[p][ch];
            v[6 + ch] += w * dC[p][ch];
          }
          R[p] = last_a[p] * lastq[p] + (1.0f - last_a[p]) * R[p];
          lastq[p] = q;
          float dLda = (q - R[p]) * Tp;
          last_a[p] = a;
          v[8 + C] += w;
          v[6 + C] += fabsf(a * dLda);
          if (!clamped) {
            float ddx = dx[p & 1], ddy = dy[p >> 1];
            float dp = B.y * dLda * G * LN2;  // dL/dp', p' = log2(e)·power
            float gx = -dp * (2.0f * A.z * ddx + A.w * ddy);
            float gy = -dp * (2.0f * B.x * ddy + A.w * ddx);
            v[0] += gx; v[1] += gy;
            v[2] += dp * ddx * ddx; v[3] += dp * ddx * ddy; v[4] += dp * ddy * ddy;
            v[5] += G * dLda;
            v[7 + C] += fabsf(gx) + fabsf(gy);
          }
        }
      }
      if (__any_sync(GS_FULL, touched)) {
        v[2] *= -0.5f * LOG2E; v[3] *= -LOG2E; v[4] *= -0.5f * LOG2E;
        int index; bool valid;
        float tot = (NV == 12) ? multi_reduce12(v, lane, index, valid) : multi_reduce16(v, lane, index, valid);
        if (valid && index < 9 + C && tot != 0.f) {
          int id = __float_as_int(B.w);
          float *dst = (vis_plane && index == 8 + C) ? vis_plane + id : acc + (int64_t)acc_stride * id + index;
          atomicAdd(dst, tot);
        }
      }
    }
  }
}

extern "C" int gs_raster_acc_stride(int channels) { return (9 + channels + 3) / 4 * 4; }

template <int C>
static int launch_fwd(const gs_raster_config *h_cfg, const float *g2d, const float *features, const float *depth,
                      const int32_t *o2p, const int32_t *ranges, int W, int H, float *image, float *final_T,
                      int32_t *n_contrib, float *visibility, float *median, cudaStream_t st) {
  CfgDev cfg = gs_make_cfg(h_cfg);
  int TX = (W + GS_TILE - 1) / GS_TILE, TY = (H + GS_TILE - 1) / GS_TILE;
  int tiles = TX * TY;
  bool vis = visibility != nullptr, med = median != nullptr;
#define GS_FWD(V, M) k_raster_fwd<C, V, M><<<tiles, RT, 0, st>>>(cfg, g2d, features, depth, o2p, ranges, W, H, TX, image, final_T, n_contrib, visibility, median)
  if (vis && med) GS_FWD(true, true);
  else if (vis) GS_FWD(true, false);
  else if (med) GS_FWD(false, true);
  else GS_FWD(false, false);
#undef GS_FWD
  GS_LAUNCH_CHECK();
  return GS_OK;
}

extern "C" int gs_raster_fwd(const gs_raster_config *h_cfg, int channels, int64_t m, const float *g2d, const float *features,
                             const float *depth, const int32_t *overlap_to_point, const int32_t *tile_ranges, int width,
                             int height, float *image, float *final_T, int32_t *n_contrib, float *visibility,
                             float *median_depth, const gs_raster_aux *aux, void *stream) {
  (void)m;
  GS_REQUIRE(h_cfg && h_cfg->tile_size == GS_TILE, "tile_size must be 16");
  GS_REQUIRE(width > 0 && height > 0, "bad image size");
  GS_REQUIRE(median_depth == nullptr || depth != nullptr, "median depth needs depth");
  cudaStream_t st = (cudaStream_t)stream;
  if (aux && aux->records) {
    GS_REQUIRE(channels == 3, "staged records carry 3 channels");
    return gs_raster_fwd_rec(h_cfg, aux, depth, overlap_to_point, tile_ranges, width, height, image, final_T, n_contrib,
                             visibility, median_depth, st);
  }
  GS_REQUIRE(!aux || (!aux->tile_order && !aux->counters), "tile_order / counters need staged records");
  switch (channels) {
    case 1: return launch_fwd<1>(h_cfg, g2d, features, depth, overlap_to_point, tile_ranges, width, height, image, final_T, n_contrib, visibility, median_depth, st);
    case 3: return launch_fwd<3>(h_cfg, g2d, features, depth, overlap_to_point, tile_ranges, width, height, image, final_T, n_contrib, visibility, median_depth, st);
    case 4: return launch_fwd<4>(h_cfg, g2d, features, depth, overlap_to_point, tile_ranges, width, height, image, final_T, n_contrib, visibility, median_depth, st);
    default: gs_set_error("gs_raster_fwd: channels must be 1, 3 or 4"); return GS_ERR_UNSUPPORTED;
  }
}

extern "C" int gs_raster_bwd(const gs_raster_config *h_cfg, int channels, int64_t m, const float *g2d, const float *features,
                             const int32_t *overlap_to_point, const int32_t *tile_ranges, int width, int height,
                             const float *final_T, const int32_t *n_contrib, const float *d_image, float *acc,
                             int acc_stride, float *vis_plane, const gs_raster_aux *aux, void *stream) {
  (void)m;
  GS_REQUIRE(h_cfg && h_cfg->tile_size == GS_TILE, "tile_size must be 16");
  GS_REQUIRE(acc_stride >= 9 + channels, "acc_stride too small");
  cudaStream_t st = (cudaStream_t)stream;
  if (aux && aux->records) {
    GS_REQUIRE(channels == 3, "staged records carry 3 channels");
    return gs_raster_bwd_rec(h_cfg, aux, overlap_to_point, tile_ranges, width, height, final_T, n_contrib, d_image, acc,
                             acc_stride, vis_plane, st);
  }
  GS_REQUIRE(!aux || !aux->tile_order, "tile_order needs staged records");
  CfgDev cfg = gs_make_cfg(h_cfg);
  int TX = (width + GS_TILE - 1) / GS_TILE, TY = (height + GS_TILE - 1) / GS_TILE;
  int tiles = TX * TY;
  switch (channels) {
    case 1: k_raster_bwd<1><<<tiles, RT, 0, st>>>(cfg, g2d, features, overlap_to_point, tile_ranges, width, height, TX, final_T, n_contrib, d_image, acc, acc_stride, vis_plane); break;
    case 3: k_raster_bwd<3><<<tiles, RT, 0, st>>>(cfg, g2d, features, overlap_to_point, tile_ranges, width, height, TX, final_T, n_contrib, d_image, acc, acc_stride, vis_plane); break;
    case 4: k_raster_bwd<4><<<tiles, RT, 0, st>>>(cfg, g2d, features, overlap_to_point, tile_ranges, width, height, TX, final_T, n_contrib, d_image, acc, acc_stride, vis_plane); break;
    default: gs_set_error("gs_raster_bwd: channels must be 1, 3 or 4"); return GS_ERR_UNSUPPORTED;
  }
  GS_LAUNCH_CHECK();
  return GS_OK;
}
