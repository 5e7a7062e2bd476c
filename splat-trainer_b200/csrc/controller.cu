// controller.cu — the two per-step controller / scene side passes that touch every splat:
//   gs_mcmc_noise   MCMCController.step, noise branch (reference controller/mcmc_controller.py:91-100 with
//                   util/misc.py:20-37 soft_lt and gaussians/split.py:31-36 sample_gaussians)
//   gs_reg_loss     MLPScene.compute_reg / reg_loss (reference scene/mlp_scene.py:246-288, util/misc.py:68-69 saturate)
// Both are single streaming sweeps (HBM bound); rows the step does not touch are skipped after one 2- or 4-byte read.
#include "common.cuh"
#include "reg.cuh"

// point_basis (reference gaussians/split.py:15-19): B[3*j+k] = R[j][k] * max(exp(s_k), eps)
__device__ __forceinline__ void ctl_point_basis(const float *__restrict__ ls, const float *__restrict__ q, float eps, float *B) {
  float x = q[0], y = q[1], z = q[2], w = q[3];
  float ql = sqrtf(x * x + y * y + z * z + w * w);
  float d = fmaxf(ql, 1e-12f);
  x /= d; y /= d; z /= d; w /= d;
  float R[9];
  R[0] = 1.f - 2.f * (y * y + z * z); R[1] = 2.f * (x * y - w * z); R[2] = 2.f * (x * z + w * y);
  R[3] = 2.f * (x * y + w * z); R[4] = 1.f - 2.f * (x * x + z * z); R[5] = 2.f * (y * z - w * x);
  R[6] = 2.f * (x * z - w * y); R[7] = 2.f * (y * z + w * x); R[8] = 1.f - 2.f * (x * x + y * y);
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    float s = fmaxf(expf(ls[k]), eps);
#pragma unroll
    for (int j = 0; j < 3; ++j) B[3 * j + k] = R[3 * j + k] * s;
  }
}

__global__ void __launch_bounds__(256) k_mcmc_noise(int64_t n, const float *__restrict__ alpha_logit,
                                                    const float *__restrict__ log_scaling, const float *__restrict__ rotation,
                                                    const int16_t *__restrict__ points_in_view, int min_views, float half_thr,
                                                    float noise_level, const float *__restrict__ randn,
                                                    float *__restrict__ position) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  if (!((int)points_in_view[i] > min_views)) return;
  float op = 1.0f / (1.0f + expf(-alpha_logit[i]));
  float gt = 1.0f / (1.0f + expf(-((op - half_thr) * 16.0f / half_thr)));  // soft_gt(op, thr/2, margin 16)
  float level = (1.0f - gt) * noise_level;
  float B[9];
  ctl_point_basis(log_scaling + 3 * i, rotation + 4 * i, 1e-4f, B);
  float s0 = randn[3 * i] * level, s1 = randn[3 * i + 1] * level, s2 = randn[3 * i + 2] * level;
#pragma unroll
  for (int j = 0; j < 3; ++j) position[3 * i + j] += B[3 * j] * s0 + B[3 * j + 1] * s1 + B[3 * j + 2] * s2;
}

extern "C" int gs_mcmc_noise(int64_t n, const float *alpha_logit, const float *log_scaling, const float *rotation,
                             const int16_t *points_in_view, int min_views, float opacity_threshold, float noise_level,
                             const float *randn, float *position, void *stream) {
  GS_REQUIRE(alpha_logit && log_scaling && rotation && points_in_view && randn && position, "null argument");
  GS_REQUIRE(opacity_threshold > 0.f, "opacity_threshold must be positive");
  if (n <= 0) return GS_OK;
  k_mcmc_noise<<<gs_div_up(n, 256), 256, 0, (cudaStream_t)stream>>>(n, alpha_logit, log_scaling, rotation, points_in_view,
                                                                    min_views, 0.5f * opacity_threshold, noise_level, randn,
                                                                    position);
  GS_LAUNCH_CHECK();
  return GS_OK;
}

// ---- regularisation --------------------------------------------------------------------------------------------
// pass 1: Σ over blended rows of w·(scale, opacity, aspect) terms and their count
__global__ void __launch_bounds__(256) k_reg_sums(int64_t n, const float *__restrict__ g2d, const float *__restrict__ depth,
                                                  const float *__restrict__ log_scaling, const float *__restrict__ vis,
                                                  int vis_stride, float *__restrict__ sums) {
  float a0 = 0.f, a1 = 0.f, a2 = 0.f, cnt = 0.f;
  // 4 consecutive rows per thread: the sweep is one read of the visibility plane (blended rows are ~1-50 % of it)
  for (int64_t i0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4; i0 < n; i0 += (int64_t)gridDim.x * blockDim.x * 4) {
    float w4[4];
    if (vis_stride == 1 && i0 + 3 < n) {
      float4 q = *reinterpret_cast<const float4 *>(vis + i0);
      w4[0] = q.x; w4[1] = q.y; w4[2] = q.z; w4[3] = q.w;
    } else {
#pragma unroll
      for (int k = 0; k < 4; ++k) w4[k] = (i0 + k < n) ? vis[(int64_t)vis_stride * (i0 + k)] : 0.f;
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      float w = w4[k];
      if (!(w > 0.f)) continue;
      int64_t i = i0 + k;
      float op = g2d[GS_G2D_STRIDE * i + 5];
      if (!(op > 0.f)) continue;
      RegTerms t = reg_terms(op, depth[i], log_scaling + 3 * i);
      a0 += w * t.ns; a1 += w * t.sat * t.ns; a2 += w * t.aspect; cnt += 1.f;
    }
  }
  a0 = warp_sum(a0); a1 = warp_sum(a1); a2 = warp_sum(a2); cnt = warp_sum(cnt);
  __shared__ float sh[4][8];
  int wp = threadIdx.x >> 5;
  if ((threadIdx.x & 31) == 0) { sh[0][wp] = a0; sh[1][wp] = a1; sh[2][wp] = a2; sh[3][wp] = cnt; }
  __syncthreads();
  if (threadIdx.x < 4) {
    float t = 0.f;
    for (int k = 0; k < 8; ++k) t += sh[threadIdx.x][k];
    if (t != 0.f) atomicAdd(sums + threadIdx.x, t);
  }
}

// pass 2: gradients of Σ_k weight_k · mean_k over the blended rows (w constant)
__global__ void __launch_bounds__(256) k_reg_grads(int64_t n, float w_scale, float w_opacity, float w_aspect,
                                                   const float *__restrict__ g2d, const float *__restrict__ depth,
                                                   const float *__restrict__ log_scaling, const float *__restrict__ vis,
                                                   int vis_stride, const float *__restrict__ sums, float *__restrict__ d_opacity,
                                                   int d_opacity_stride, float *__restrict__ d_depth,
                                                   float *__restrict__ d_log_scaling) {
  int64_t i0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (i0 >= n) return;
  float w4[4];
  if (vis_stride == 1 && i0 + 3 < n) {
    float4 q = *reinterpret_cast<const float4 *>(vis + i0);
    w4[0] = q.x; w4[1] = q.y; w4[2] = q.z; w4[3] = q.w;
  } else {
#pragma unroll
    for (int k = 0; k < 4; ++k) w4[k] = (i0 + k < n) ? vis[(int64_t)vis_stride * (i0 + k)] : 0.f;
  }
  if (!(w4[0] > 0.f || w4[1] > 0.f || w4[2] > 0.f || w4[3] > 0.f)) return;
  float m = sums[3];
  if (!(m > 0.f)) return;
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    float w = w4[q];
    if (!(w > 0.f)) continue;
    int64_t i = i0 + q;
    float op = g2d[GS_G2D_STRIDE * i + 5];
    if (!(op > 0.f)) continue;
    float c = w / m;
    float d = depth[i];
    RegTerms t = reg_terms(op, d, log_scaling + 3 * i);
    RegGrad g = reg_grad(t, c, w_scale, w_opacity, w_aspect);
    if (d_opacity) d_opacity[(int64_t)d_opacity_stride * i] += g.d_op;
    if (d_depth) d_depth[i] = g.d_depth;
    if (d_log_scaling) {
#pragma unroll
      for (int k = 0; k < 3; ++k) d_log_scaling[3 * i + k] += g.d_ls[k];
    }
  }
}

extern "C" int gs_reg_loss(int64_t n, float w_scale, float w_opacity, float w_aspect, const float *g2d, const float *depth,
                           const float *log_scaling, const float *vis, int vis_stride, float *sums, int compute_sums,
                           float *d_opacity, int d_opacity_stride, float *d_depth, float *d_log_scaling, void *stream) {
  GS_REQUIRE(g2d && depth && log_scaling && vis && sums, "null argument");
  if (n <= 0) return GS_OK;
  cudaStream_t st = (cudaStream_t)stream;
  if (compute_sums) {
    GS_CUDA(cudaMemsetAsync(sums, 0, 4 * sizeof(float), st));
    int64_t want = (n + 1023) / 1024;  // few CTAs: every CTA ends with 4 same-sector atomics, which serialise in L2
    k_reg_sums<<<(int)(want < 148 * 4 ? want : 148 * 4), 256, 0, st>>>(n, g2d, depth, log_scaling, vis, vis_stride, sums);
    GS_LAUNCH_CHECK();
  }
  if (d_opacity || d_depth || d_log_scaling) {
    k_reg_grads<<<gs_div_up((n + 3) / 4, 256), 256, 0, st>>>(n, w_scale, w_opacity, w_aspect, g2d, depth, log_scaling, vis, vis_stride,
                                                   sums, d_opacity, d_opacity_stride, d_depth, d_log_scaling);
    GS_LAUNCH_CHECK();
  }
  return GS_OK;
}
