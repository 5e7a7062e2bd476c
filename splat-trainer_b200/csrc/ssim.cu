// ssim.cu — multi-scale SSIM loss, forward + backward (reference trainer/trainer.py:450-462 with
// self.ssim = fused_ssim(padding="valid"), :112).  fused-ssim is an absent third-party dependency: its published
// algorithm is restated (11x11 Gaussian window sigma 1.5, zero padded, C1 = 0.01^2, C2 = 0.03^2, map cropped by 5).
//
// Images are H x W x C channels-last fp32 (level 0: the raw render, clamped to [0,1] on load like the reference's
// post_activation; target fp32 or uint8).  One CTA = one 16x16 pixel tile: the 26x26xC halo tile is staged in shared
// memory once (coalesced rows), then a horizontal and a vertical 11-tap pass run out of shared memory for all channels
// at once, register-tiled: a thread filters 4 consecutive outputs from 14 loaded inputs.
// forward: per-pixel partial derivatives of the map (w.r.t. mu1, E[pp], E[pt]; zero outside the valid crop) + the
//          sum of the map over the valid crop;
// backward: the three partial maps blurred with the same window, combined into d pred.
// Coarser levels: 2x2 average pooling (F.avg_pool2d(2), floor sizes) done by the finer level's forward kernel from the tile
// it has staged; its transpose is folded into the next finer level's backward kernel (the backward runs coarse to fine).  HBM traffic ~ 40 B per pixel-channel per level.
#include <math.h>

#include "common.cuh"

namespace {

constexpr int ST = 16;            // tile side
constexpr int SR = 5;             // window radius
constexpr int SH = ST + 2 * SR;   // halo tile side (26)

struct Win { float g[11]; };

__device__ __forceinline__ float clamp01(float v) { return fminf(fmaxf(v, 0.f), 1.f); }

// x / 255 for the integer values a uint8 target holds, correctly rounded like the IEEE division the oracle does, without
// the division's slow-path check: one multiply by RN(1/255) and one exact-remainder correction (Markstein)
__device__ __forceinline__ float u8_unit(float x) {
  const float r = 1.0f / 255.0f;
  float q = x * r;
  return fmaf(fmaf(-q, 255.0f, x), r, q);
}

// 1 / x for x well inside the normal range (here x >= C1 * C2): MUFU.RCP refined by one Newton step
__device__ __forceinline__ float rcp_newton(float x) {
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return fmaf(r, fmaf(-x, r, 1.0f), r);
}

constexpr int SG = 4;             // outputs per thread along the filtered axis (register tiling: 14 loads feed 4 outputs)

// 4 outputs of an 11-tap filter from 14 consecutive inputs
__device__ __forceinline__ void taps4(const Win &win, const float *v, float *out) {
#pragma unroll
  for (int o = 0; o < SG; ++o) {
    float a = 0.f;
#pragma unroll
    for (int k = 0; k < 11; ++k) a += win.g[k] * v[o + k];
    out[o] = a;
  }
}

// dynamic shared memory: sP | sT (SH*SH*C each) | sHz (5 * SH*ST*C)
template <int C>
__global__ void __launch_bounds__(ST * ST) k_ssim_fwd(Win win, int H, int W, const float *__restrict__ pred, int clamp_pred,
                                                      const float *__restrict__ target, const uint8_t *__restrict__ target_u8,
                                                      float *__restrict__ dm1, float *__restrict__ dm2, float *__restrict__ dm3,
                                                      float *__restrict__ sum_out, float *__restrict__ lm_sums,
                                                      float *__restrict__ pool_p, float *__restrict__ pool_t, int hc, int wc) {
  extern __shared__ float ssim_smem[];
  float *sP = ssim_smem, *sT = sP + SH * SH * C, *sHz = sT + SH * SH * C;
  constexpr int HZ = SH * ST * C;  // one horizontally filtered plane
  __shared__ float s_red[ST * ST / 32];
  float lm1 = 0.f, lm2 = 0.f;  // fused L1 / MSE photometric loss of the tile's own pixels (lm_sums != NULL, level 0)
  const int tid = threadIdx.x;
  const int tx0 = blockIdx.x * ST, ty0 = blockIdx.y * ST;
  {  // halo tile load, fully unrolled so that every global load is in flight before the first use (the tile is a
     // handful of loads per thread: a rolled loop pays one DRAM round trip per iteration)
    constexpr int NL = (SH * SH * C + ST * ST - 1) / (ST * ST);
    float pl[NL], tl[NL];
#pragma unroll
    for (int it = 0; it < NL; ++it) {
      int i = tid + it * ST * ST;
      int c = i % C, rx = (i / C) % SH, ry = i / (C * SH);
      int gx = tx0 - SR + rx, gy = ty0 - SR + ry;
      pl[it] = 0.f; tl[it] = 0.f;
      if (i < SH * SH * C && gx >= 0 && gx < W && gy >= 0 && gy < H) {
        int64_t gi = ((int64_t)gy * W + gx) * C + c;
        pl[it] = pred[gi];
        tl[it] = target_u8 ? (float)target_u8[gi] : target[gi];
      }
    }
#pragma unroll
    for (int it = 0; it < NL; ++it) {
      int i = tid + it * ST * ST;
      if (i < SH * SH * C) {
        sP[i] = clamp_pred ? clamp01(pl[it]) : pl[it];
        sT[i] = target_u8 ? u8_unit(tl[it]) : tl[it];
        // gs_loss_l1_mse's sums over the 16x16 pixels this tile owns: the values are in registers anyway, so the separate
        // loss kernel and its sweep of the image are gone (its gradient is formed by the level-0 backward, which holds
        // the same pixel again when it writes d pred)
        int rx = (i / C) % SH, ry = i / (C * SH);
        int gx = tx0 - SR + rx, gy = ty0 - SR + ry;
        if (lm_sums && rx >= SR && rx < SR + ST && ry >= SR && ry < SR + ST && gx < W && gy < H) {
          float dd = sP[i] - sT[i];  // (level 0: sP holds the clamped render)
          lm1 += fabsf(dd);
          lm2 += dd * dd;
        }
      }
    }
  }
  __syncthreads();
  // next coarser level's input (F.avg_pool2d(2), floor sizes) of the tile's own 16x16 pixels: the staged tile holds them, so
  // the pooling kernel and its sweep of both images are gone.  Same order of additions as the oracle's pooling loop.
  if (pool_p && tid < (ST / 2) * (ST / 2) * C) {
    const int c = tid % C, x = (tid / C) % (ST / 2), y = tid / (C * (ST / 2));
    const int ox = (tx0 >> 1) + x, oy = (ty0 >> 1) + y;
    if (ox < wc && oy < hc) {
      float ps = 0.f, ts = 0.f;
#pragma unroll
      for (int dy = 0; dy < 2; ++dy)
#pragma unroll
        for (int dx = 0; dx < 2; ++dx) {
          const int j = ((SR + 2 * y + dy) * SH + SR + 2 * x + dx) * C + c;
          ps += sP[j];
          ts += sT[j];
        }
      const int64_t o = ((int64_t)oy * wc + ox) * C + c;
      pool_p[o] = 0.25f * ps;
      pool_t[o] = 0.25f * ts;
    }
  }
  // horizontal pass: a thread filters 4 consecutive x of one (halo row, channel) for the 5 quantities p, t, pp, tt, pt
  for (int i = tid; i < SH * (ST / SG) * C; i += ST * ST) {
    int c = i % C, xg = (i / C) % (ST / SG), ry = i / (C * (ST / SG));
    float pv[SG + 10], tv[SG + 10], tmp[SG + 10], out[SG];
#pragma unroll
    for (int k = 0; k < SG + 10; ++k) {
      int j = (ry * SH + xg * SG + k) * C + c;
      pv[k] = sP[j]; tv[k] = sT[j];
    }
    const int o0 = (ry * ST + xg * SG) * C + c;
    taps4(win, pv, out);
#pragma unroll
    for (int o = 0; o < SG; ++o) sHz[0 * HZ + o0 + o * C] = out[o];
    taps4(win, tv, out);
#pragma unroll
    for (int o = 0; o < SG; ++o) sHz[1 * HZ + o0 + o * C] = out[o];
#pragma unroll
    for (int k = 0; k < SG + 10; ++k) tmp[k] = pv[k] * pv[k];
    taps4(win, tmp, out);
#pragma unroll
    for (int o = 0; o < SG; ++o) sHz[2 * HZ + o0 + o * C] = out[o];
#pragma unroll
    for (int k = 0; k < SG + 10; ++k) tmp[k] = tv[k] * tv[k];
    taps4(win, tmp, out);
#pragma unroll
    for (int o = 0; o < SG; ++o) sHz[3 * HZ + o0 + o * C] = out[o];
#pragma unroll
    for (int k = 0; k < SG + 10; ++k) tmp[k] = pv[k] * tv[k];
    taps4(win, tmp, out);
#pragma unroll
    for (int o = 0; o < SG; ++o) sHz[4 * HZ + o0 + o * C] = out[o];
  }
  __syncthreads();
  // vertical pass: a thread owns 4 consecutive rows of one (column, channel)
  const float C1 = 0.01f * 0.01f, C2 = 0.03f * 0.03f;
  float local = 0.f;
  for (int i = tid; i < ST * (ST / SG) * C; i += ST * ST) {
    int c = i % C, lx = (i / C) % ST, yg = i / (C * ST);
    float q[5][SG];
#pragma unroll
    for (int a = 0; a < 5; ++a) {
      float v[SG + 10];
#pragma unroll
      for (int k = 0; k < SG + 10; ++k) v[k] = sHz[a * HZ + ((yg * SG + k) * ST + lx) * C + c];
      taps4(win, v, q[a]);
    }
    const int gx = tx0 + lx;
#pragma unroll
    for (int o = 0; o < SG; ++o) {
      const int gy = ty0 + yg * SG + o;
      const bool inside = gx < W && gy < H;
      const bool valid = inside && gx >= SR && gx < W - SR && gy >= SR && gy < H - SR;
      float m1 = q[0][o], m2 = q[1][o];
      float s1 = q[2][o] - m1 * m1, s2 = q[3][o] - m2 * m2, s12 = q[4][o] - m1 * m2;
      float A = 2.f * m1 * m2 + C1, B = 2.f * s12 + C2, Cc = m1 * m1 + m2 * m2 + C1, D = s1 + s2 + C2;
      // one reciprocal for the three quotients of the map and its partial derivatives
      float icd = rcp_newton(Cc * D);
      float rC = D * icd, rD = Cc * icd;  // 1 / Cc, 1 / D
      float map = A * B * icd;
      if (valid) local += map;
      if (inside) {
        int64_t gi = ((int64_t)gy * W + gx) * C + c;
        float w = valid ? 1.f : 0.f;
        // d map / d mu1 with E[pp], E[pt] fixed (s1, s12 depend on mu1), d map / d E[pp], d map / d E[pt]
        dm1[gi] = w * (2.f * m2 * (B - A) * icd + 2.f * m1 * map * (rD - rC));
        dm2[gi] = w * (-map * rD);
        dm3[gi] = w * (2.f * A * icd);
      }
    }
  }
  local = warp_sum(local);
  if ((tid & 31) == 0) s_red[tid >> 5] = local;
  __syncthreads();
  if (tid == 0) {
    float t = 0.f;
    for (int k = 0; k < ST * ST / 32; ++k) t += s_red[k];
    if (t != 0.f) atomicAdd(sum_out, t);
  }
  if (lm_sums) {
    __shared__ float s_lm[2][ST * ST / 32];
    lm1 = warp_sum(lm1);
    lm2 = warp_sum(lm2);
    if ((tid & 31) == 0) { s_lm[0][tid >> 5] = lm1; s_lm[1][tid >> 5] = lm2; }
    __syncthreads();
    if (tid < 2) {
      float t = 0.f;
      for (int k = 0; k < ST * ST / 32; ++k) t += s_lm[tid][k];
      atomicAdd(lm_sums + tid, t);
    }
  }
}

// d[i] (+)= scale * (G*dm1 + 2 p G*dm2 + t G*dm3); level 0 honours the clamp (no gradient where the raw render left [0,1])
// dynamic shared memory: sM (3 * SH*SH*C) | sHz (3 * SH*ST*C)
template <int C>
__global__ void __launch_bounds__(ST * ST) k_ssim_bwd(Win win, int H, int W, const float *__restrict__ pred, int clamp_pred,
                                                      const float *__restrict__ target, const uint8_t *__restrict__ target_u8,
                                                      const float *__restrict__ dm1, const float *__restrict__ dm2,
                                                      const float *__restrict__ dm3, float scale, float *__restrict__ d_out,
                                                      int accumulate, const float *__restrict__ d_coarse, int hc, int wc,
                                                      int lm, float w_l1, float w_mse, float inv_n) {
  extern __shared__ float ssim_smem[];
  constexpr int HM = SH * SH * C, HZ = SH * ST * C;
  float *sM = ssim_smem, *sHz = sM + 3 * HM;
  const int tid = threadIdx.x;
  const int tx0 = blockIdx.x * ST, ty0 = blockIdx.y * ST;
  {  // halo tile load of the three partial maps, unrolled (all loads in flight together)
    constexpr int NL = (HM + ST * ST - 1) / (ST * ST);
    float la[NL], lb[NL], ld[NL];
#pragma unroll
    for (int it = 0; it < NL; ++it) {
      int i = tid + it * ST * ST;
      int c = i % C, rx = (i / C) % SH, ry = i / (C * SH);
      int gx = tx0 - SR + rx, gy = ty0 - SR + ry;
      la[it] = 0.f; lb[it] = 0.f; ld[it] = 0.f;
      if (i < HM && gx >= 0 && gx < W && gy >= 0 && gy < H) {
        int64_t gi = ((int64_t)gy * W + gx) * C + c;
        la[it] = dm1[gi]; lb[it] = dm2[gi]; ld[it] = dm3[gi];
      }
    }
#pragma unroll
    for (int it = 0; it < NL; ++it) {
      int i = tid + it * ST * ST;
      if (i < HM) { sM[i] = la[it]; sM[HM + i] = lb[it]; sM[2 * HM + i] = ld[it]; }
    }
  }
  __syncthreads();
  for (int i = tid; i < SH * (ST / SG) * C; i += ST * ST) {
    int c = i % C, xg = (i / C) % (ST / SG), ry = i / (C * (ST / SG));
    const int o0 = (ry * ST + xg * SG) * C + c;
#pragma unroll
    for (int a = 0; a < 3; ++a) {
      float v[SG + 10], out[SG];
#pragma unroll
      for (int k = 0; k < SG + 10; ++k) v[k] = sM[a * HM + (ry * SH + xg * SG + k) * C + c];
      taps4(win, v, out);
#pragma unroll
      for (int o = 0; o < SG; ++o) sHz[a * HZ + o0 + o * C] = out[o];
    }
  }
  __syncthreads();
  for (int i = tid; i < ST * (ST / SG) * C; i += ST * ST) {
    int c = i % C, lx = (i / C) % ST, yg = i / (C * ST);
    float q[3][SG];
#pragma unroll
    for (int a = 0; a < 3; ++a) {
      float v[SG + 10];
#pragma unroll
      for (int k = 0; k < SG + 10; ++k) v[k] = sHz[a * HZ + ((yg * SG + k) * ST + lx) * C + c];
      taps4(win, v, q[a]);
    }
    const int gx = tx0 + lx;
#pragma unroll
    for (int o = 0; o < SG; ++o) {
      const int gy = ty0 + yg * SG + o;
      if (!(gx < W && gy < H)) continue;
      int64_t gi = ((int64_t)gy * W + gx) * C + c;
      float raw = pred[gi];
      float p = clamp_pred ? clamp01(raw) : raw;
      float t = target_u8 ? u8_unit((float)target_u8[gi]) : target[gi];
      float d = scale * (q[0][o] + 2.f * p * q[1][o] + t * q[2][o]);
      // transpose of the 2x2 average pooling: a quarter of the next coarser level's gradient (already complete: the
      // backward runs coarse to fine), for pixels inside the pooled area (floor sizes)
      if (d_coarse && (gy >> 1) < hc && (gx >> 1) < wc) d += 0.25f * d_coarse[((int64_t)(gy >> 1) * wc + (gx >> 1)) * C + c];
      if (lm) {  // + the L1 / MSE gradient of gs_loss_l1_mse on the same pixel (gs_photometric_loss: d pred is written once)
        float dd = p - t;
        float sg = dd > 0.f ? 1.f : (dd < 0.f ? -1.f : 0.f);
        d += (w_l1 * sg + w_mse * 2.0f * dd) * inv_n;
      }
      if (clamp_pred && !(raw >= 0.f && raw <= 1.f)) d = 0.f;
      if (accumulate) d_out[gi] += d; else d_out[gi] = d;
    }
  }
}

Win make_window() {
  Win w;
  double t[11], s = 0;
  for (int i = 0; i < 11; ++i) { t[i] = exp(-(double)((i - 5) * (i - 5)) / (2.0 * 1.5 * 1.5)); s += t[i]; }
  for (int i = 0; i < 11; ++i) w.g[i] = (float)(t[i] / s);
  return w;
}

size_t level_floats(int H, int W, int C, int l) {
  for (int k = 0; k < l; ++k) { H /= 2; W /= 2; }
  return (size_t)(H > 0 ? H : 1) * (W > 0 ? W : 1) * C;
}

}  // namespace

// workspace: per level 3 partial maps; per level l >= 1 also pred_l, target_l, d_l
extern "C" size_t gs_ssim_workspace_bytes(int height, int width, int channels, int levels) {
  if (height <= 0 || width <= 0 || channels <= 0 || levels < 1) return 0;
  size_t b = 0;
  for (int l = 0; l < levels; ++l) b += (l == 0 ? 3 : 6) * gs_align256(level_floats(height, width, channels, l) * 4);
  return b + 256;
}

template <int C>
static int ssim_run(int H, int W, int levels, const float *pred, const float *target, const uint8_t *target_u8, float weight,
                    float *d_pred, float *ssim_sums, void *workspace, cudaStream_t st, float w_l1, float w_mse, float *lm_sums) {
  Win win = make_window();
  const size_t fwd_smem = (size_t)(2 * SH * SH + 5 * SH * ST) * C * sizeof(float);
  const size_t bwd_smem = (size_t)(3 * SH * SH + 3 * SH * ST) * C * sizeof(float);
  GS_CUDA(cudaFuncSetAttribute(k_ssim_fwd<C>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fwd_smem));
  GS_CUDA(cudaFuncSetAttribute(k_ssim_bwd<C>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bwd_smem));
  uintptr_t a = ((uintptr_t)workspace + 255) & ~(uintptr_t)255;
  const float *p[8], *t[8];
  float *d[8], *dm[8][3];
  int hs[8], ws[8];
  p[0] = pred; t[0] = target; d[0] = d_pred; hs[0] = H; ws[0] = W;
  for (int l = 0; l < levels; ++l) {
    if (l > 0) { hs[l] = hs[l - 1] / 2; ws[l] = ws[l - 1] / 2; }
    size_t nl = gs_align256(level_floats(H, W, C, l) * 4);
    for (int q = 0; q < 3; ++q) { dm[l][q] = (float *)a; a += nl; }
    if (l > 0) {
      p[l] = (float *)a; t[l] = (float *)(a + nl);
      d[l] = (float *)(a + 2 * nl);
      a += 3 * nl;
    }
  }
  GS_CUDA(cudaMemsetAsync(ssim_sums, 0, (size_t)levels * sizeof(float), st));
  // (running the coarse levels on a helper stream beside level 0 was measured: no gain — every kernel here is issue-bound
  // and fills the machine on its own, and the fork / join events cost what the overlap of the tails saves)
  cudaStream_t sc = st;
  auto fwd = [&](int l, cudaStream_t s_) -> int {
    if (hs[l] <= 0 || ws[l] <= 0) return GS_OK;
    dim3 grid((ws[l] + ST - 1) / ST, (hs[l] + ST - 1) / ST);
    const bool lm = l == 0 && lm_sums != nullptr;
    const bool pool = l + 1 < levels && hs[l + 1] > 0 && ws[l + 1] > 0;  // this level's forward also pools for the next
    k_ssim_fwd<C><<<grid, ST * ST, fwd_smem, s_>>>(win, hs[l], ws[l], p[l], l == 0, t[l], l == 0 ? target_u8 : nullptr, dm[l][0],
                                            dm[l][1], dm[l][2], ssim_sums + l, lm ? lm_sums : nullptr,
                                            pool ? (float *)p[l + 1] : nullptr, pool ? (float *)t[l + 1] : nullptr,
                                            pool ? hs[l + 1] : 0, pool ? ws[l + 1] : 0);
    GS_LAUNCH_CHECK();
    return GS_OK;
  };
  const bool lm0 = lm_sums != nullptr;
  auto bwd = [&](int l, cudaStream_t s_) -> int {
    if (hs[l] <= 0 || ws[l] <= 0) return GS_OK;
    dim3 grid((ws[l] + ST - 1) / ST, (hs[l] + ST - 1) / ST);
    double count = (hs[l] > 10 && ws[l] > 10) ? (double)(hs[l] - 10) * (ws[l] - 10) * C : 0.0;
    // loss = weight * mean_l (1 - ssim_l), ssim_l = sum_l / count_l
    float scale = count > 0 ? (float)(-(double)weight / ((double)levels * count)) : 0.f;
    bool has_coarse = l + 1 < levels && hs[l + 1] > 0 && ws[l + 1] > 0;
    k_ssim_bwd<C><<<grid, ST * ST, bwd_smem, s_>>>(win, hs[l], ws[l], p[l], l == 0, t[l], l == 0 ? target_u8 : nullptr, dm[l][0],
                                            dm[l][1], dm[l][2], scale, d[l], l == 0 && !lm0, has_coarse ? d[l + 1] : nullptr,
                                            has_coarse ? hs[l + 1] : 0, has_coarse ? ws[l + 1] : 0, l == 0 && lm0, w_l1, w_mse,
                                            1.0f / (float)((int64_t)H * W * C));
    GS_LAUNCH_CHECK();
    return GS_OK;
  };
  // forward fine to coarse (level l writes level l + 1's pooled inputs), backward coarse to fine (level l adds the
  // upsampled gradient of l + 1)
  int rc;
  for (int l = 0; l < levels; ++l)
    if ((rc = fwd(l, st)) != GS_OK) return rc;
  if (d_pred)
    for (int l = levels - 1; l >= 1; --l)
      if ((rc = bwd(l, sc)) != GS_OK) return rc;
  if (!d_pred) return GS_OK;
  return bwd(0, st);
}

static int loss_entry(const char *who, int height, int width, int channels, int levels, const float *pred, const float *target,
                      const uint8_t *target_u8, float weight, float *d_pred, float *ssim_sums, void *workspace,
                      size_t workspace_bytes, float w_l1, float w_mse, float *l1_mse_sums, cudaStream_t st) {
  GS_REQUIRE(pred && (target || target_u8) && ssim_sums && workspace, "null argument");
  GS_REQUIRE(height > 0 && width > 0, "bad image size");
  GS_REQUIRE(levels >= 1 && levels <= 8, "levels must be 1..8");
  if (workspace_bytes < gs_ssim_workspace_bytes(height, width, channels, levels)) {
    gs_set_error("%s: workspace too small", who);
    return GS_ERR_WORKSPACE;
  }
  if (l1_mse_sums) GS_CUDA(cudaMemsetAsync(l1_mse_sums, 0, 2 * sizeof(float), st));
  switch (channels) {
    case 1: return ssim_run<1>(height, width, levels, pred, target, target_u8, weight, d_pred, ssim_sums, workspace, st, w_l1, w_mse, l1_mse_sums);
    case 3: return ssim_run<3>(height, width, levels, pred, target, target_u8, weight, d_pred, ssim_sums, workspace, st, w_l1, w_mse, l1_mse_sums);
    case 4: return ssim_run<4>(height, width, levels, pred, target, target_u8, weight, d_pred, ssim_sums, workspace, st, w_l1, w_mse, l1_mse_sums);
    default: gs_set_error("%s: channels must be 1, 3 or 4", who); return GS_ERR_UNSUPPORTED;
  }
}

extern "C" int gs_ssim_loss(int height, int width, int channels, int levels, const float *pred, const float *target,
                            const uint8_t *target_u8, float weight, float *d_pred, float *ssim_sums, void *workspace,
                            size_t workspace_bytes, void *stream) {
  return loss_entry("gs_ssim_loss", height, width, channels, levels, pred, target, target_u8, weight, d_pred, ssim_sums, workspace,
                    workspace_bytes, 0.f, 0.f, nullptr, (cudaStream_t)stream);
}

extern "C" int gs_photometric_loss(int height, int width, int channels, int levels, const float *pred, const float *target,
                                   const uint8_t *target_u8, float w_l1, float w_mse, float w_ssim, float *d_pred,
                                   float *l1_mse_sums, float *ssim_sums, void *workspace, size_t workspace_bytes, void *stream) {
  GS_REQUIRE(d_pred && l1_mse_sums, "null argument");
  return loss_entry("gs_photometric_loss", height, width, channels, levels, pred, target, target_u8, w_ssim, d_pred, ssim_sums,
                    workspace, workspace_bytes, w_l1, w_mse, l1_mse_sums, (cudaStream_t)stream);
}
