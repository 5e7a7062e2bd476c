// ssim.cu — multi-scale SSIM loss, forward + backward (reference trainer/trainer.py:450-462 with
// self.ssim = fused_ssim(padding="valid"), :112).  fused-ssim is an absent third-party dependency: its published
// algorithm is restated (11x11 Gaussian window sigma 1.5, zero padded, C1 = 0.01^2, C2 = 0.03^2, map cropped by 5).
//
// Images are H x W x C channels-last fp32 (level 0: the raw render, clamped to [0,1] on load like the reference's
// post_activation; target fp32 or uint8).  One CTA = one 16x16 pixel tile: the 26x26xC halo tile is staged in shared
// memory once (coalesced rows), then per channel a horizontal and a vertical 11-tap pass run out of shared memory.
// forward: per-pixel partial derivatives of the map (w.r.t. mu1, E[pp], E[pt]; zero outside the valid crop) + the
//          sum of the map over the valid crop;
// backward: the three partial maps blurred with the same window, combined into d pred.
// Coarser levels: 2x2 average pooling forward; its transpose is folded into the next finer level's backward kernel
// (the backward runs coarse to fine).  HBM traffic ~ 40 B per pixel-channel per level.
#include <math.h>

#include "common.cuh"

namespace {

constexpr int ST = 16;            // tile side
constexpr int SR = 5;             // window radius
constexpr int SH = ST + 2 * SR;   // halo tile side (26)

struct Win { float g[11]; };

__device__ __forceinline__ float clamp01(float v) { return fminf(fmaxf(v, 0.f), 1.f); }

template <int C>
__global__ void __launch_bounds__(ST * ST) k_ssim_fwd(Win win, int H, int W, const float *__restrict__ pred, int clamp_pred,
                                                      const float *__restrict__ target, const uint8_t *__restrict__ target_u8,
                                                      float *__restrict__ dm1, float *__restrict__ dm2, float *__restrict__ dm3,
                                                      float *__restrict__ sum_out) {
  __shared__ float sP[SH * SH * C];
  __shared__ float sT[SH * SH * C];
  __shared__ float sHz[5][SH * ST];
  __shared__ float s_red[ST * ST / 32];
  const int tid = threadIdx.x;
  const int tx0 = blockIdx.x * ST, ty0 = blockIdx.y * ST;
  for (int i = tid; i < SH * SH * C; i += ST * ST) {
    int c = i % C, rx = (i / C) % SH, ry = i / (C * SH);
    int gx = tx0 - SR + rx, gy = ty0 - SR + ry;
    float p = 0.f, t = 0.f;
    if (gx >= 0 && gx < W && gy >= 0 && gy < H) {
      int64_t gi = ((int64_t)gy * W + gx) * C + c;
      p = pred[gi];
      if (clamp_pred) p = clamp01(p);
      t = target_u8 ? (float)target_u8[gi] / 255.0f : target[gi];
    }
    sP[i] = p; sT[i] = t;
  }
  __syncthreads();
  const int lx = tid % ST, ly = tid / ST;
  const int gx = tx0 + lx, gy = ty0 + ly;
  const bool inside = gx < W && gy < H;
  const bool valid = inside && gx >= SR && gx < W - SR && gy >= SR && gy < H - SR;
  const float C1 = 0.01f * 0.01f, C2 = 0.03f * 0.03f;
  float local = 0.f;
  for (int c = 0; c < C; ++c) {
    for (int i = tid; i < SH * ST; i += ST * ST) {
      int x = i % ST, ry = i / ST;
      float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f, a4 = 0.f;
#pragma unroll
      for (int k = 0; k < 11; ++k) {
        float p = sP[(ry * SH + x + k) * C + c], t = sT[(ry * SH + x + k) * C + c], g = win.g[k];
        a0 += g * p; a1 += g * t; a2 += g * p * p; a3 += g * t * t; a4 += g * p * t;
      }
      sHz[0][i] = a0; sHz[1][i] = a1; sHz[2][i] = a2; sHz[3][i] = a3; sHz[4][i] = a4;
    }
    __syncthreads();
    float m1 = 0.f, m2 = 0.f, e11 = 0.f, e22 = 0.f, e12 = 0.f;
#pragma unroll
    for (int k = 0; k < 11; ++k) {
      int i = (ly + k) * ST + lx;
      float g = win.g[k];
      m1 += g * sHz[0][i]; m2 += g * sHz[1][i]; e11 += g * sHz[2][i]; e22 += g * sHz[3][i]; e12 += g * sHz[4][i];
    }
    float s1 = e11 - m1 * m1, s2 = e22 - m2 * m2, s12 = e12 - m1 * m2;
    float A = 2.f * m1 * m2 + C1, B = 2.f * s12 + C2, Cc = m1 * m1 + m2 * m2 + C1, D = s1 + s2 + C2;
    float icd = 1.0f / (Cc * D);
    float map = A * B * icd;
    if (valid) local += map;
    if (inside) {
      int64_t gi = ((int64_t)gy * W + gx) * C + c;
      float w = valid ? 1.f : 0.f;
      // d map / d mu1 with E[pp], E[pt] fixed (s1, s12 depend on mu1), d map / d E[pp], d map / d E[pt]
      dm1[gi] = w * (2.f * m2 * (B - A) * icd - 2.f * m1 * map / Cc + 2.f * m1 * map / D);
      dm2[gi] = w * (-map / D);
      dm3[gi] = w * (2.f * A * icd);
    }
    __syncthreads();
  }
  local = warp_sum(local);
  if ((tid & 31) == 0) s_red[tid >> 5] = local;
  __syncthreads();
  if (tid == 0) {
    float t = 0.f;
    for (int k = 0; k < ST * ST / 32; ++k) t += s_red[k];
    if (t != 0.f) atomicAdd(sum_out, t);
  }
}

// d[i] (+)= scale * (G*dm1 + 2 p G*dm2 + t G*dm3); level 0 honours the clamp (no gradient where the raw render left [0,1])
template <int C>
__global__ void __launch_bounds__(ST * ST) k_ssim_bwd(Win win, int H, int W, const float *__restrict__ pred, int clamp_pred,
                                                      const float *__restrict__ target, const uint8_t *__restrict__ target_u8,
                                                      const float *__restrict__ dm1, const float *__restrict__ dm2,
                                                      const float *__restrict__ dm3, float scale, float *__restrict__ d_out,
                                                      int accumulate, const float *__restrict__ d_coarse, int hc, int wc) {
  __shared__ float sM[3][SH * SH * C];
  __shared__ float sHz[3][SH * ST];
  const int tid = threadIdx.x;
  const int tx0 = blockIdx.x * ST, ty0 = blockIdx.y * ST;
  for (int i = tid; i < SH * SH * C; i += ST * ST) {
    int c = i % C, rx = (i / C) % SH, ry = i / (C * SH);
    int gx = tx0 - SR + rx, gy = ty0 - SR + ry;
    float a = 0.f, b = 0.f, d = 0.f;
    if (gx >= 0 && gx < W && gy >= 0 && gy < H) {
      int64_t gi = ((int64_t)gy * W + gx) * C + c;
      a = dm1[gi]; b = dm2[gi]; d = dm3[gi];
    }
    sM[0][i] = a; sM[1][i] = b; sM[2][i] = d;
  }
  __syncthreads();
  const int lx = tid % ST, ly = tid / ST;
  const int gx = tx0 + lx, gy = ty0 + ly;
  const bool inside = gx < W && gy < H;
  for (int c = 0; c < C; ++c) {
    for (int i = tid; i < SH * ST; i += ST * ST) {
      int x = i % ST, ry = i / ST;
      float a0 = 0.f, a1 = 0.f, a2 = 0.f;
#pragma unroll
      for (int k = 0; k < 11; ++k) {
        int j = (ry * SH + x + k) * C + c;
        float g = win.g[k];
        a0 += g * sM[0][j]; a1 += g * sM[1][j]; a2 += g * sM[2][j];
      }
      sHz[0][i] = a0; sHz[1][i] = a1; sHz[2][i] = a2;
    }
    __syncthreads();
    float c1 = 0.f, c2 = 0.f, c3 = 0.f;
#pragma unroll
    for (int k = 0; k < 11; ++k) {
      int i = (ly + k) * ST + lx;
      float g = win.g[k];
      c1 += g * sHz[0][i]; c2 += g * sHz[1][i]; c3 += g * sHz[2][i];
    }
    if (inside) {
      int64_t gi = ((int64_t)gy * W + gx) * C + c;
      float raw = pred[gi];
      float p = clamp_pred ? clamp01(raw) : raw;
      float t = target_u8 ? (float)target_u8[gi] / 255.0f : target[gi];
      float d = scale * (c1 + 2.f * p * c2 + t * c3);
      // transpose of the 2x2 average pooling: a quarter of the next coarser level's gradient (already complete: the
      // backward runs coarse to fine), for pixels inside the pooled area (floor sizes)
      if (d_coarse && (gy >> 1) < hc && (gx >> 1) < wc) d += 0.25f * d_coarse[((int64_t)(gy >> 1) * wc + (gx >> 1)) * C + c];
      if (clamp_pred && !(raw >= 0.f && raw <= 1.f)) d = 0.f;
      if (accumulate) d_out[gi] += d; else d_out[gi] = d;
    }
    __syncthreads();
  }
}

// F.avg_pool2d(kernel 2, stride 2) of prediction and target together (floor sizes)
__global__ void __launch_bounds__(256) k_pool2(int h, int w, int C, int Wf, const float *__restrict__ p_in, int clamp_pred,
                                               const float *__restrict__ t_in, const uint8_t *__restrict__ t_u8,
                                               float *__restrict__ p_out, float *__restrict__ t_out) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (int64_t)h * w * C) return;
  int c = (int)(i % C);
  int x = (int)((i / C) % w), y = (int)(i / ((int64_t)C * w));
  float ps = 0.f, ts = 0.f;
#pragma unroll
  for (int dy = 0; dy < 2; ++dy)
#pragma unroll
    for (int dx = 0; dx < 2; ++dx) {
      int64_t j = ((int64_t)(2 * y + dy) * Wf + 2 * x + dx) * C + c;
      float p = p_in[j];
      ps += clamp_pred ? clamp01(p) : p;
      ts += t_u8 ? (float)t_u8[j] / 255.0f : t_in[j];
    }
  p_out[i] = 0.25f * ps;
  t_out[i] = 0.25f * ts;
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
                    float *d_pred, float *ssim_sums, void *workspace, cudaStream_t st) {
  Win win = make_window();
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
      float *pl = (float *)a, *tl = (float *)(a + nl);
      d[l] = (float *)(a + 2 * nl);
      a += 3 * nl;
      p[l] = pl; t[l] = tl;
      int64_t cnt = (int64_t)hs[l] * ws[l] * C;
      if (cnt > 0) {
        k_pool2<<<gs_div_up(cnt, 256), 256, 0, st>>>(hs[l], ws[l], C, ws[l - 1], p[l - 1], l == 1, t[l - 1],
                                                     l == 1 ? target_u8 : nullptr, pl, tl);
        GS_LAUNCH_CHECK();
      }
    }
  }
  GS_CUDA(cudaMemsetAsync(ssim_sums, 0, (size_t)levels * sizeof(float), st));
  for (int l = 0; l < levels; ++l) {
    if (hs[l] <= 0 || ws[l] <= 0) continue;
    dim3 grid((ws[l] + ST - 1) / ST, (hs[l] + ST - 1) / ST);
    k_ssim_fwd<C><<<grid, ST * ST, 0, st>>>(win, hs[l], ws[l], p[l], l == 0, t[l], l == 0 ? target_u8 : nullptr, dm[l][0],
                                            dm[l][1], dm[l][2], ssim_sums + l);
    GS_LAUNCH_CHECK();
  }
  if (!d_pred) return GS_OK;
  for (int l = levels - 1; l >= 0; --l) {  // coarse to fine: level l adds the upsampled gradient of level l + 1
    if (hs[l] <= 0 || ws[l] <= 0) continue;
    dim3 grid((ws[l] + ST - 1) / ST, (hs[l] + ST - 1) / ST);
    double count = (hs[l] > 10 && ws[l] > 10) ? (double)(hs[l] - 10) * (ws[l] - 10) * C : 0.0;
    // loss = weight * mean_l (1 - ssim_l), ssim_l = sum_l / count_l
    float scale = count > 0 ? (float)(-(double)weight / ((double)levels * count)) : 0.f;
    bool has_coarse = l + 1 < levels && hs[l + 1] > 0 && ws[l + 1] > 0;
    k_ssim_bwd<C><<<grid, ST * ST, 0, st>>>(win, hs[l], ws[l], p[l], l == 0, t[l], l == 0 ? target_u8 : nullptr, dm[l][0],
                                            dm[l][1], dm[l][2], scale, d[l], l == 0, has_coarse ? d[l + 1] : nullptr,
                                            has_coarse ? hs[l + 1] : 0, has_coarse ? ws[l + 1] : 0);
    GS_LAUNCH_CHECK();
  }
  return GS_OK;
}

extern "C" int gs_ssim_loss(int height, int width, int channels, int levels, const float *pred, const float *target,
                            const uint8_t *target_u8, float weight, float *d_pred, float *ssim_sums, void *workspace,
                            size_t workspace_bytes, void *stream) {
  GS_REQUIRE(pred && (target || target_u8) && ssim_sums && workspace, "null argument");
  GS_REQUIRE(height > 0 && width > 0, "bad image size");
  GS_REQUIRE(levels >= 1 && levels <= 8, "levels must be 1..8");
  if (workspace_bytes < gs_ssim_workspace_bytes(height, width, channels, levels)) { gs_set_error("gs_ssim_loss: workspace too small"); return GS_ERR_WORKSPACE; }
  cudaStream_t st = (cudaStream_t)stream;
  switch (channels) {
    case 1: return ssim_run<1>(height, width, levels, pred, target, target_u8, weight, d_pred, ssim_sums, workspace, st);
    case 3: return ssim_run<3>(height, width, levels, pred, target, target_u8, weight, d_pred, ssim_sums, workspace, st);
    case 4: return ssim_run<4>(height, width, levels, pred, target, target_u8, weight, d_pred, ssim_sums, workspace, st);
    default: gs_set_error("gs_ssim_loss: channels must be 1, 3 or 4"); return GS_ERR_UNSUPPORTED;
  }
}
