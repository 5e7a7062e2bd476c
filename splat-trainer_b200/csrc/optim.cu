// optim.cu — photometric loss, visibility-aware per-point Adam/LaProp, and the controller's running
// statistics (reference trainer/trainer.py:465-476, scene/mlp_scene.py:214-239,
// controller/point_state.py:34-50).
//
// All three are streaming, HBM-bound kernels.  The optimiser handles one parameter group per launch;
// a row is owned by LANES consecutive lanes (LANES = 1 for dim <= 4, 16 for the SH feature rows) so
// that a warp's loads of a wide row are contiguous 64-B segments, rows the step does not touch
// (visible <= 0) are never read, and the consumed gradient row is cleared in the same pass (the
// reference's dense zero_grad, scene/mlp_scene.py:155-161, costs 4·D·N bytes of writes per step).
#include "common.cuh"

// ---- loss ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_loss_l1_mse(int64_t n, const float *__restrict__ pred, const float *__restrict__ target,
                                                     const uint8_t *__restrict__ target_u8, float w_l1, float w_mse,
                                                     float inv_n, float *__restrict__ d_pred, float *__restrict__ sums) {
  float s1 = 0.f, s2 = 0.f;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    float p = pred[i];
    bool inside = (p >= 0.f) && (p <= 1.f);
    float c = fminf(fmaxf(p, 0.f), 1.f);
    float t = target_u8 ? (float)target_u8[i] / 255.0f : target[i];
    float d = c - t;
    s1 += fabsf(d);
    s2 += d * d;
    float sg = d > 0.f ? 1.f : (d < 0.f ? -1.f : 0.f);
    if (d_pred) d_pred[i] = inside ? (w_l1 * sg + w_mse * 2.0f * d) * inv_n : 0.f;
  }
  s1 = warp_sum(s1);
  s2 = warp_sum(s2);
  __shared__ float sh[2][8];
  int w = threadIdx.x >> 5;
  if ((threadIdx.x & 31) == 0) { sh[0][w] = s1; sh[1][w] = s2; }
  __syncthreads();
  if (threadIdx.x < 2) {
    float t = 0.f;
    for (int k = 0; k < 8; ++k) t += sh[threadIdx.x][k];
    atomicAdd(sums + threadIdx.x, t);
  }
}

extern "C" int gs_loss_l1_mse(int64_t n, const float *pred, const float *target, const uint8_t *target_u8, float w_l1,
                              float w_mse, float *d_pred, float *sums, void *stream) {
  GS_REQUIRE(pred && sums && (target || target_u8), "null argument");
  if (n <= 0) return GS_OK;
  int blocks = (int)((n + 255) / 256);
  if (blocks > 148 * 16) blocks = 148 * 16;
  k_loss_l1_mse<<<blocks, 256, 0, (cudaStream_t)stream>>>(n, pred, target, target_u8, w_l1, w_mse, 1.0f / (float)n, d_pred, sums);
  GS_LAUNCH_CHECK();
  return GS_OK;
}

// ---- optimiser -------------------------------------------------------------------------------------
struct OptDev {
  int mode;
  float beta1, beta2, eps, vis_beta, vis_smooth, grad_clip;
  int bias_correction;
};

__global__ void __launch_bounds__(256) k_vis_update(int64_t n, float vis_beta, const float *__restrict__ visible,
                                                    float *__restrict__ running_vis, float *__restrict__ step) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  float v = visible[i];
  if (!(v > 0.f)) return;
  step[i] += 1.0f;
  running_vis[i] = vis_beta * running_vis[i] + (1.0f - vis_beta) * v;
}

// indexes of rows with visible > 0 (ascending) + their count, all on the device
__global__ void __launch_bounds__(256) k_vis_flags(int64_t n, const float *__restrict__ visible, int32_t *__restrict__ flags) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) flags[i] = visible[i] > 0.f ? 1 : 0;
}
__global__ void __launch_bounds__(256) k_vis_indexes(int64_t n, const float *__restrict__ visible,
                                                     const int32_t *__restrict__ rank, int32_t *__restrict__ indexes) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n && visible[i] > 0.f) indexes[rank[i]] = (int32_t)i;
}

extern "C" size_t gs_visible_indexes_workspace_bytes(int64_t n) {
  size_t nn = (size_t)(n > 0 ? n : 1);
  return 2 * gs_align256(nn * 4) + gs_align256((size_t)gs_scan_partials(nn) * 4) + 256;
}

extern "C" int gs_visible_indexes(int64_t n, const float *visible, int32_t *indexes, int32_t *count_dev, void *workspace,
                                  size_t workspace_bytes, void *stream) {
  GS_REQUIRE(n >= 0 && n < ((int64_t)1 << 31), "n must fit int32");
  GS_REQUIRE(count_dev, "null count");
  cudaStream_t st = (cudaStream_t)stream;
  if (n == 0) { GS_CUDA(cudaMemsetAsync(count_dev, 0, 4, st)); return GS_OK; }
  GS_REQUIRE(visible && indexes && workspace, "null argument");
  if (workspace_bytes < gs_visible_indexes_workspace_bytes(n)) { gs_set_error("gs_visible_indexes: workspace too small"); return GS_ERR_WORKSPACE; }
  uintptr_t a = ((uintptr_t)workspace + 255) & ~(uintptr_t)255;
  int32_t *flags = (int32_t *)a;
  int32_t *rank = (int32_t *)(a + gs_align256((size_t)n * 4));
  int32_t *partials = (int32_t *)(a + 2 * gs_align256((size_t)n * 4));
  int blocks = gs_div_up(n, 256);
  k_vis_flags<<<blocks, 256, 0, st>>>(n, visible, flags);
  int rc = gs_exclusive_scan_i32(flags, nullptr, rank, n, nullptr, partials, count_dev, st);
  if (rc != GS_OK) return rc;
  k_vis_indexes<<<blocks, 256, 0, st>>>(n, visible, rank, indexes);
  GS_LAUNCH_CHECK();
  return GS_OK;
}

// Fused start of the optimiser step: gs_optim_vis_update + gs_visible_indexes in ONE pass over `visible`
// (single-pass chained-scan compaction): for rows with visible > 0: step += 1, running_vis = lerp(vis_beta), and the
// ascending row list / its length are written.
constexpr int OB_THREADS = 256;
constexpr int OB_ITEMS = 8;
constexpr int OB_TILE = OB_THREADS * OB_ITEMS;

__global__ void __launch_bounds__(OB_THREADS) k_optim_begin(int64_t n, float vis_beta, const float *__restrict__ visible,
                                                            float *__restrict__ running_vis, float *__restrict__ step,
                                                            uint32_t *__restrict__ status, uint32_t *__restrict__ ticket,
                                                            int32_t *__restrict__ indexes, int32_t *__restrict__ count_dev) {
  __shared__ int s_warp[8];
  __shared__ int s_tile;
  __shared__ uint32_t s_prefix;
  if (threadIdx.x == 0) s_tile = (int)atomicAdd(ticket, 1u);
  __syncthreads();
  const int tile = s_tile;
  int64_t base = (int64_t)tile * OB_TILE + (int64_t)threadIdx.x * OB_ITEMS;
  float v[OB_ITEMS];
  if (base + OB_ITEMS <= n) {
    const float4 *p = reinterpret_cast<const float4 *>(visible + base);
    float4 a = p[0], b = p[1];
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
  } else {
#pragma unroll
    for (int i = 0; i < OB_ITEMS; ++i) v[i] = (base + i < n) ? visible[base + i] : 0.f;
  }
  int sum = 0;
#pragma unroll
  for (int i = 0; i < OB_ITEMS; ++i) sum += v[i] > 0.f;
  // block exclusive scan (256 threads)
  int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  int inc = sum;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    int t = __shfl_up_sync(GS_FULL, inc, o);
    if (lane >= o) inc += t;
  }
  if (lane == 31) s_warp[w] = inc;
  __syncthreads();
  int wo = 0, tot = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    int c = s_warp[i];
    if (i < w) wo += c;
    tot += c;
  }
  int ex = wo + inc - sum;
  if (threadIdx.x == 0) gs_st_status(status + tile, (uint32_t)tot | (tile == 0 ? GS_LB_INCL : GS_LB_LOCAL));
  if (threadIdx.x < 32) {
    uint32_t prefix = gs_lookback_warp(status, tile);
    if (threadIdx.x == 0) {
      if (tile > 0) gs_st_status(status + tile, (prefix + (uint32_t)tot) | GS_LB_INCL);
      s_prefix = prefix;
      if (tile == (int)gridDim.x - 1) *count_dev = (int32_t)(prefix + (uint32_t)tot);
    }
  }
  __syncthreads();
  uint32_t pos = s_prefix + (uint32_t)ex;
#pragma unroll
  for (int i = 0; i < OB_ITEMS; ++i) {
    if (v[i] > 0.f) {
      int64_t r = base + i;
      indexes[pos++] = (int32_t)r;
      step[r] += 1.0f;
      running_vis[r] = vis_beta * running_vis[r] + (1.0f - vis_beta) * v[i];
    }
  }
}

extern "C" size_t gs_optim_begin_workspace_bytes(int64_t n) {
  size_t tiles = (size_t)((n > 0 ? n : 1) + OB_TILE - 1) / OB_TILE;
  return gs_align256((64 + tiles) * 4) + 256;
}

extern "C" int gs_optim_begin(const gs_optim_config *c, int64_t n, const float *visible, float *running_vis, float *step,
                              int32_t *indexes, int32_t *count_dev, void *workspace, size_t workspace_bytes, void *stream) {
  GS_REQUIRE(c && count_dev, "null argument");
  GS_REQUIRE(n >= 0 && n < ((int64_t)1 << 30), "n must be < 2^30");
  cudaStream_t st = (cudaStream_t)stream;
  if (n == 0) { GS_CUDA(cudaMemsetAsync(count_dev, 0, 4, st)); return GS_OK; }
  GS_REQUIRE(visible && running_vis && step && indexes && workspace, "null argument");
  if (workspace_bytes < gs_optim_begin_workspace_bytes(n)) { gs_set_error("gs_optim_begin: workspace too small"); return GS_ERR_WORKSPACE; }
  uintptr_t a = ((uintptr_t)workspace + 255) & ~(uintptr_t)255;
  int tiles = gs_div_up(n, OB_TILE);
  GS_CUDA(cudaMemsetAsync((void *)a, 0, (64 + (size_t)tiles) * 4, st));
  uint32_t *ticket = (uint32_t *)a, *status = ticket + 64;
  k_optim_begin<<<tiles, OB_THREADS, 0, st>>>(n, c->vis_beta, visible, running_vis, step, status, ticket, indexes, count_dev);
  GS_LAUNCH_CHECK();
  return GS_OK;
}

extern "C" int gs_optim_vis_update(const gs_optim_config *c, int64_t n, const float *visible, float *running_vis,
                                   float *step, void *stream) {
  GS_REQUIRE(c && visible && running_vis && step, "null argument");
  if (n <= 0) return GS_OK;
  k_vis_update<<<gs_div_up(n, 256), 256, 0, (cudaStream_t)stream>>>(n, c->vis_beta, visible, running_vis, step);
  GS_LAUNCH_CHECK();
  return GS_OK;
}

// point_basis (reference gaussians/split.py:15-19): B[3*j+k] = R[j][k] * max(exp(s_k), eps)
__device__ __forceinline__ void point_basis(const float *__restrict__ ls, const float *__restrict__ q, float eps, float *B) {
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

// KIND: 0 scalar, 1 vector, 2 local_vector.  LANES lanes share a row; lane l owns elements l, l+LANES, ...
// EPL = elements per lane (ceil(dim / LANES)).
template <int KIND, int LANES, int EPL>
__device__ __forceinline__ void optim_row(const OptDev &c, int64_t i, int l, bool row_ok, int dim, float lr,
                                          float *__restrict__ param, float *__restrict__ grad, float *__restrict__ exp_avg,
                                          float *__restrict__ exp_avg_sq, const float *__restrict__ visible,
                                          const float *__restrict__ running_vis, const float *__restrict__ step,
                                          const float *__restrict__ basis_ls, const float *__restrict__ basis_q,
                                          int post, float post_lo, float post_hi, int zero_grad) {
  float w = row_ok ? visible[i] : 0.f;
  bool active = row_ok && (w > 0.f);
  // all lanes of a row agree on `active`; sub-warp shuffles below use the full mask, inactive rows just idle
  if (LANES == 1 && !active) return;  // (returns from optim_row only)
  float bc1 = 1.f, bc2 = 1.f, lam = 0.f, inv_w = 0.f;
  if (active) {
    float ts = step[i];
    float rv = running_vis[i];
    if (c.bias_correction) {
      rv = rv / (1.0f - powf(c.vis_beta, ts));
      bc1 = 1.0f - powf(c.beta1, ts);
      bc2 = 1.0f - powf(c.beta2, ts);
    }
    lam = fminf(w / (rv + c.vis_smooth), 1.0f);  // Spec O: a row seen less than usual takes a smaller step, never a larger one
    inv_w = 1.0f / (w + c.vis_smooth);
  }
  float g[EPL], gt[EPL], p[EPL], m1[EPL];
  float B[9];
  int64_t row = i * dim;
#pragma unroll
  for (int j = 0; j < EPL; ++j) {
    int k = l + j * LANES;
    bool ok = active && k < dim;
    g[j] = ok ? grad[row + k] : 0.f;
    p[j] = ok ? param[row + k] : 0.f;
    m1[j] = ok ? exp_avg[row + k] : 0.f;
  }
  if (KIND == 2) {  // LANES == 1, dim == 3
    point_basis(basis_ls + 3 * i, basis_q + 4 * i, 1e-4f, B);
#pragma unroll
    for (int k = 0; k < 3; ++k) gt[k] = (B[k] * g[0] + B[3 + k] * g[1] + B[6 + k] * g[2]) * inv_w;
  } else {
#pragma unroll
    for (int j = 0; j < EPL; ++j) gt[j] = g[j] * inv_w;
  }
  if (c.grad_clip > 0.f) {
#pragma unroll
    for (int j = 0; j < EPL; ++j) gt[j] = fminf(fmaxf(gt[j], -c.grad_clip), c.grad_clip);
  }
  float vshared = 0.f;
  if (KIND != 0) {
    float ms = 0.f;
#pragma unroll
    for (int j = 0; j < EPL; ++j) ms += gt[j] * gt[j];  // lanes past dim hold 0
#pragma unroll
    for (int o = LANES / 2; o > 0; o >>= 1) ms += __shfl_xor_sync(GS_FULL, ms, o);
    ms /= (float)dim;
    if (active) {
      vshared = c.beta2 * exp_avg_sq[i] + (1.0f - c.beta2) * ms;
      if (l == 0) exp_avg_sq[i] = vshared;
    }
  }
  float upd[EPL];
#pragma unroll
  for (int j = 0; j < EPL; ++j) {
    int k = l + j * LANES;
    bool ok = active && k < dim;
    float v2 = vshared;
    if (KIND == 0 && ok) {
      v2 = c.beta2 * exp_avg_sq[row + k] + (1.0f - c.beta2) * gt[j] * gt[j];
      exp_avg_sq[row + k] = v2;
    }
    float denom = sqrtf(v2 / bc2) + c.eps;
    if (c.mode == 0) {
      m1[j] = c.beta1 * m1[j] + (1.0f - c.beta1) * gt[j];
      upd[j] = (m1[j] / bc1) / denom;
    } else {
      m1[j] = c.beta1 * m1[j] + (1.0f - c.beta1) * gt[j] / denom;
      upd[j] = m1[j] / bc1;
    }
    if (ok) exp_avg[row + k] = m1[j];
  }
  float sc = lr * lam;
  if (KIND == 2) {
#pragma unroll
    for (int j = 0; j < 3; ++j) p[j] -= sc * (B[3 * j] * upd[0] + B[3 * j + 1] * upd[1] + B[3 * j + 2] * upd[2]);
  } else {
#pragma unroll
    for (int j = 0; j < EPL; ++j) p[j] -= sc * upd[j];
  }
  if (post == GS_POST_NORMALIZE) {
    float ss = 0.f;
#pragma unroll
    for (int j = 0; j < EPL; ++j) ss += (l + j * LANES < dim) ? p[j] * p[j] : 0.f;
#pragma unroll
    for (int o = LANES / 2; o > 0; o >>= 1) ss += __shfl_xor_sync(GS_FULL, ss, o);
    float len = fmaxf(sqrtf(ss), 1e-12f);
#pragma unroll
    for (int j = 0; j < EPL; ++j) p[j] /= len;
  } else if (post == GS_POST_CLAMP) {
#pragma unroll
    for (int j = 0; j < EPL; ++j) p[j] = fminf(fmaxf(p[j], post_lo), post_hi);
  }
#pragma unroll
  for (int j = 0; j < EPL; ++j) {
    int k = l + j * LANES;
    if (active && k < dim) {
      param[row + k] = p[j];
      if (zero_grad) grad[row + k] = 0.f;
    }
  }
}

// VECTOR groups with wide rows (dim a multiple of 4, >= 16: the 48 SH coefficients, the 16 point features): four lanes
// share a row and move it as 16-byte words (lane l owns words l, l+4, ...), so a warp has 8 rows and up to
// 9 x 512 B of loads in flight instead of 2 rows / 9 x 128 B — the sweep is bound by memory-level parallelism.
// Same arithmetic per element as optim_row<1, ...>; only the partition of the mean-square sum over lanes differs.
template <int EPL4>
__device__ __forceinline__ void optim_row_v4(const OptDev &c, int64_t i, int l, bool row_ok, int dim, float lr,
                                             float *__restrict__ param, float *__restrict__ grad, float *__restrict__ exp_avg,
                                             float *__restrict__ exp_avg_sq, const float *__restrict__ visible,
                                             const float *__restrict__ running_vis, const float *__restrict__ step,
                                             int post, float post_lo, float post_hi, int zero_grad) {
  float w = row_ok ? visible[i] : 0.f;
  bool active = row_ok && (w > 0.f);
  float bc1 = 1.f, bc2 = 1.f, lam = 0.f, inv_w = 0.f;
  if (active) {
    float ts = step[i];
    float rv = running_vis[i];
    if (c.bias_correction) {
      rv = rv / (1.0f - powf(c.vis_beta, ts));
      bc1 = 1.0f - powf(c.beta1, ts);
      bc2 = 1.0f - powf(c.beta2, ts);
    }
    lam = fminf(w / (rv + c.vis_smooth), 1.0f);
    inv_w = 1.0f / (w + c.vis_smooth);
  }
  const int words = dim >> 2;
  float4 *gp = reinterpret_cast<float4 *>(grad + i * dim), *pp = reinterpret_cast<float4 *>(param + i * dim);
  float4 *mp = reinterpret_cast<float4 *>(exp_avg + i * dim);
  float4 g[EPL4], p[EPL4], m1[EPL4];
#pragma unroll
  for (int j = 0; j < EPL4; ++j) {
    int k = l + 4 * j;
    bool ok = active && k < words;
    const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
    g[j] = ok ? gp[k] : z;
    p[j] = ok ? pp[k] : z;
    m1[j] = ok ? mp[k] : z;
  }
  auto clipf = [&](float x) { x *= inv_w; return c.grad_clip > 0.f ? fminf(fmaxf(x, -c.grad_clip), c.grad_clip) : x; };
  float ms = 0.f;
#pragma unroll
  for (int j = 0; j < EPL4; ++j) {
    g[j].x = clipf(g[j].x); g[j].y = clipf(g[j].y); g[j].z = clipf(g[j].z); g[j].w = clipf(g[j].w);
    ms += g[j].x * g[j].x + g[j].y * g[j].y + g[j].z * g[j].z + g[j].w * g[j].w;  // words past dim hold 0
  }
  ms += __shfl_xor_sync(GS_FULL, ms, 2);
  ms += __shfl_xor_sync(GS_FULL, ms, 1);
  ms /= (float)dim;
  float v2 = 0.f;
  if (active) {
    v2 = c.beta2 * exp_avg_sq[i] + (1.0f - c.beta2) * ms;
    if (l == 0) exp_avg_sq[i] = v2;
  }
  const float denom = sqrtf(v2 / bc2) + c.eps;
  const float sc = lr * lam;
  auto upd = [&](float &m, float gt, float &pv) {
    float u;
    if (c.mode == 0) {
      m = c.beta1 * m + (1.0f - c.beta1) * gt;
      u = (m / bc1) / denom;
    } else {
      m = c.beta1 * m + (1.0f - c.beta1) * gt / denom;
      u = m / bc1;
    }
    pv -= sc * u;
  };
#pragma unroll
  for (int j = 0; j < EPL4; ++j) {
    upd(m1[j].x, g[j].x, p[j].x); upd(m1[j].y, g[j].y, p[j].y); upd(m1[j].z, g[j].z, p[j].z); upd(m1[j].w, g[j].w, p[j].w);
  }
  if (post == GS_POST_NORMALIZE) {
    float ss = 0.f;
#pragma unroll
    for (int j = 0; j < EPL4; ++j) ss += p[j].x * p[j].x + p[j].y * p[j].y + p[j].z * p[j].z + p[j].w * p[j].w;
    ss += __shfl_xor_sync(GS_FULL, ss, 2);
    ss += __shfl_xor_sync(GS_FULL, ss, 1);
    float inv = 1.0f / fmaxf(sqrtf(ss), 1e-12f);
#pragma unroll
    for (int j = 0; j < EPL4; ++j) { p[j].x *= inv; p[j].y *= inv; p[j].z *= inv; p[j].w *= inv; }
  } else if (post == GS_POST_CLAMP) {
    auto cl = [&](float x) { return fminf(fmaxf(x, post_lo), post_hi); };
#pragma unroll
    for (int j = 0; j < EPL4; ++j) { p[j].x = cl(p[j].x); p[j].y = cl(p[j].y); p[j].z = cl(p[j].z); p[j].w = cl(p[j].w); }
  }
#pragma unroll
  for (int j = 0; j < EPL4; ++j) {
    int k = l + 4 * j;
    if (active && k < words) {
      mp[k] = m1[j];
      pp[k] = p[j];
      if (zero_grad) gp[k] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
  }
}

// Grid-stride over rows.  Dense mode (indexes == NULL): row r is Gaussian r, rows with visible <= 0 idle.
// Indexed mode: row r is Gaussian indexes[r], r < *count_dev (the device-side `visible.nonzero()` of the
// reference, scene/mlp_scene.py:217, with no host read-back) — only the rows the step touches are ever read.
struct OptGroupDev {
  int variant, dim, post;
  float lr, post_lo, post_hi;
  float *param, *grad, *exp_avg, *exp_avg_sq;
  const float *basis_ls, *basis_q;
};
struct OptGroupsDev { OptGroupDev g[GS_OPTIM_MAX_GROUPS]; };

template <int KIND, int LANES, int EPL>
__device__ __forceinline__ void optim_sweep(const OptDev &c, int64_t n, const int32_t *__restrict__ indexes,
                                            const int32_t *__restrict__ count_dev, const OptGroupDev &g,
                                            const float *__restrict__ visible, const float *__restrict__ running_vis,
                                            const float *__restrict__ step, int zero_grad) {
  int64_t rows = indexes ? min((int64_t)*count_dev, n) : n;
  int64_t total = rows * LANES;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  int lane = threadIdx.x & 31;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t - lane < total; t += stride) {  // warp-uniform trip count
    int64_t r = t / LANES;
    bool row_ok = r < rows;
    int64_t i = row_ok ? (indexes ? (int64_t)indexes[r] : r) : 0;
    optim_row<KIND, LANES, EPL>(c, i, (int)(t % LANES), row_ok, g.dim, g.lr, g.param, g.grad, g.exp_avg, g.exp_avg_sq, visible,
                                running_vis, step, g.basis_ls, g.basis_q, g.post, g.post_lo, g.post_hi, zero_grad);
  }
}

template <int EPL4>
__device__ __forceinline__ void optim_sweep_v4(const OptDev &c, int64_t n, const int32_t *__restrict__ indexes,
                                               const int32_t *__restrict__ count_dev, const OptGroupDev &g,
                                               const float *__restrict__ visible, const float *__restrict__ running_vis,
                                               const float *__restrict__ step, int zero_grad) {
  int64_t rows = indexes ? min((int64_t)*count_dev, n) : n;
  int64_t total = rows * 4;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  int lane = threadIdx.x & 31;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t - lane < total; t += stride) {  // warp-uniform trip count
    int64_t r = t >> 2;
    bool row_ok = r < rows;
    int64_t i = row_ok ? (indexes ? (int64_t)indexes[r] : r) : 0;
    optim_row_v4<EPL4>(c, i, (int)(t & 3), row_ok, g.dim, g.lr, g.param, g.grad, g.exp_avg, g.exp_avg_sq, visible, running_vis,
                       step, g.post, g.post_lo, g.post_hi, zero_grad);
  }
}

// variant = the (KIND, LANES, EPL) instantiation a group runs with
__host__ __device__ inline int optim_variant(int kind, int dim) {
  if (kind == GS_GROUP_LOCAL_VECTOR) return 0;
  if (kind == GS_GROUP_VECTOR && dim >= 16 && dim % 4 == 0) return 7 + (dim <= 16 ? 0 : (dim <= 32 ? 1 : (dim <= 48 ? 2 : 3)));
  int w = dim <= 4 ? 0 : (dim <= 16 ? 1 : 2);
  return 1 + (kind == GS_GROUP_SCALAR ? 0 : 3) + w;
}
__host__ inline int optim_variant_lanes(int variant) {
  const int lanes[11] = {1, 1, 4, 16, 1, 4, 16, 4, 4, 4, 4};
  return lanes[variant];
}

// One launch for several parameter groups: blockIdx.y selects the group (a CTA-uniform read of the parameter block).
__global__ void __launch_bounds__(256) k_optim_groups(OptDev c, int64_t n, const int32_t *__restrict__ indexes,
                                                      const int32_t *__restrict__ count_dev, OptGroupsDev gs,
                                                      const float *__restrict__ visible, const float *__restrict__ running_vis,
                                                      const float *__restrict__ step, int zero_grad) {
  const OptGroupDev &g = gs.g[blockIdx.y];
  switch (g.variant) {
    case 0: optim_sweep<2, 1, 3>(c, n, indexes, count_dev, g, visible, running_vis, step, zero_grad); break;
    case 1: optim_sweep<0, 1, 4>(c, n, indexes, count_dev, g, visible, running_vis, step, zero_grad); break;
    case 2: optim_sweep<0, 4, 4>(c, n, indexes, count_dev, g, visible, running_vis, step, zero_grad); break;
    case 3: optim_sweep<0, 16, 4>(c, n, indexes, count_dev, g, visible, running_vis, step, zero_grad); break;
    case 4: optim_sweep<1, 1, 4>(c, n, indexes, count_dev, g, visible, running_vis, step, zero_grad); break;
    case 5: optim_sweep<1, 4, 4>(c, n, indexes, count_dev, g, visible, running_vis, step, zero_grad); break;
    default: optim_sweep<1, 16, 4>(c, n, indexes, count_dev, g, visible, running_vis, step, zero_grad); break;
  }
}

// the wide-row VECTOR groups (variants 7..10) in their own kernel: they hold a row in twelve 16-byte registers per array
// and would cost the narrow sweeps above their occupancy
__global__ void __launch_bounds__(256) k_optim_groups_v4(OptDev c, int64_t n, const int32_t *__restrict__ indexes,
                                                         const int32_t *__restrict__ count_dev, OptGroupsDev gs,
                                                         const float *__restrict__ visible, const float *__restrict__ running_vis,
                                                         const float *__restrict__ step, int zero_grad) {
  const OptGroupDev &g = gs.g[blockIdx.y];
  switch (g.variant) {
    case 7: optim_sweep_v4<1>(c, n, indexes, count_dev, g, visible, running_vis, step, zero_grad); break;
    case 8: optim_sweep_v4<2>(c, n, indexes, count_dev, g, visible, running_vis, step, zero_grad); break;
    case 9: optim_sweep_v4<3>(c, n, indexes, count_dev, g, visible, running_vis, step, zero_grad); break;
    default: optim_sweep_v4<4>(c, n, indexes, count_dev, g, visible, running_vis, step, zero_grad); break;
  }
}

static int optim_check_group(const gs_optim_group &q) {
  GS_REQUIRE(q.param && q.grad && q.exp_avg && q.exp_avg_sq, "null argument");
  GS_REQUIRE(q.dim >= 1 && q.dim <= 64, "dim must be 1..64");
  GS_REQUIRE(q.kind >= 0 && q.kind <= 2, "bad group kind");
  GS_REQUIRE(q.kind != GS_GROUP_LOCAL_VECTOR || (q.dim == 3 && q.basis_ls && q.basis_q), "local_vector needs dim 3 and a basis");
  return GS_OK;
}

extern "C" int gs_optim_step_groups(const gs_optim_config *h, int64_t n, const int32_t *indexes, const int32_t *count_dev,
                                    int n_groups, const gs_optim_group *groups, const float *visible, const float *running_vis,
                                    const float *step, int zero_grad, void *stream) {
  GS_REQUIRE(h && groups && visible && running_vis && step, "null argument");
  GS_REQUIRE(n_groups >= 1 && n_groups <= GS_OPTIM_MAX_GROUPS, "1..GS_OPTIM_MAX_GROUPS groups per call");
  GS_REQUIRE((indexes == nullptr) == (count_dev == nullptr), "indexes and count_dev go together");
  OptGroupsDev gs[2];  // [0] narrow sweeps, [1] wide-row vector sweeps
  int cnt[2] = {0, 0};
  int64_t want[2] = {1, 1};
  for (int k = 0; k < n_groups; ++k) {
    const gs_optim_group &q = groups[k];
    int rc = optim_check_group(q);
    if (rc != GS_OK) return rc;
    // the groups of one call run concurrently: none may update a plane another one reads as its basis
    for (int j = 0; j < n_groups; ++j)
      GS_REQUIRE(!(j != k && groups[j].kind == GS_GROUP_LOCAL_VECTOR && (groups[j].basis_ls == q.param || groups[j].basis_q == q.param)),
                 "a group of the call updates the basis of another one: put it in a later call");
    int variant = optim_variant(q.kind, q.dim);
    if (variant >= 7 && ((((uintptr_t)q.param | (uintptr_t)q.grad | (uintptr_t)q.exp_avg) & 15u) != 0))
      variant = 1 + 3 + (q.dim <= 4 ? 0 : (q.dim <= 16 ? 1 : 2));  // rows not 16-byte aligned: the scalar-lane sweep
    const int wsel = variant >= 7 ? 1 : 0;
    OptGroupDev &d = gs[wsel].g[cnt[wsel]++];
    d.variant = variant;
    d.dim = q.dim; d.post = q.post; d.lr = q.lr; d.post_lo = q.post_lo; d.post_hi = q.post_hi;
    d.param = q.param; d.grad = q.grad; d.exp_avg = q.exp_avg; d.exp_avg_sq = q.exp_avg_sq;
    d.basis_ls = q.basis_ls; d.basis_q = q.basis_q;
    int64_t w = (n * optim_variant_lanes(d.variant) + 255) / 256;
    want[wsel] = w > want[wsel] ? w : want[wsel];
  }
  if (n <= 0) return GS_OK;
  OptDev c = {h->mode, h->beta1, h->beta2, h->eps, h->vis_beta, h->vis_smooth, h->grad_clip, h->bias_correction};
  // grid: per group enough CTAs to fill the machine (148 SMs x 8 resident CTAs of 256 threads), never more than the
  // work; the sweeps are latency bound (scattered rows), so the widest group sets the pace and the CTAs of narrow
  // groups that find no rows leave at once
  if (cnt[0] > 0) {
    dim3 grid((unsigned)(want[0] < 148 * 8 ? want[0] : 148 * 8), (unsigned)cnt[0]);
    k_optim_groups<<<grid, 256, 0, (cudaStream_t)stream>>>(c, n, indexes, count_dev, gs[0], visible, running_vis, step, zero_grad);
    GS_LAUNCH_CHECK();
  }
  if (cnt[1] > 0) {
    dim3 grid((unsigned)(want[1] < 148 * 8 ? want[1] : 148 * 8), (unsigned)cnt[1]);
    k_optim_groups_v4<<<grid, 256, 0, (cudaStream_t)stream>>>(c, n, indexes, count_dev, gs[1], visible, running_vis, step, zero_grad);
    GS_LAUNCH_CHECK();
  }
  return GS_OK;
}

extern "C" int gs_optim_step(const gs_optim_config *h, int kind, int64_t n, const int32_t *indexes, const int32_t *count_dev,
                             int dim, float lr, float *param, float *grad,
                             float *exp_avg, float *exp_avg_sq, const float *visible, const float *running_vis,
                             const float *step, const float *basis_ls, const float *basis_q, int post, float post_lo,
                             float post_hi, int zero_grad, void *stream) {
  gs_optim_group q = {kind, dim, lr, param, grad, exp_avg, exp_avg_sq, basis_ls, basis_q, post, post_lo, post_hi};
  return gs_optim_step_groups(h, n, indexes, count_dev, 1, &q, visible, running_vis, step, zero_grad, stream);
}

// ---- statistics --------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_accumulate_stats(int64_t m, const int64_t *__restrict__ idx, const float *__restrict__ g2d,
                                                          const float *__restrict__ vis, int vis_stride,
                                                          const float *__restrict__ heur, int heur_stride, float split_alpha,
                                                          float prune_alpha, float *__restrict__ prune_cost,
                                                          float *__restrict__ split_score, float *__restrict__ max_scale_px,
                                                          int16_t *__restrict__ points_in_view, float *__restrict__ visibility,
                                                          float *__restrict__ scene_visible, float *__restrict__ clear_acc,
                                                          int clear_stride, float *__restrict__ clear_vis) {
  int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= m) return;
  const float2 *g = reinterpret_cast<const float2 *>(g2d + GS_G2D_STRIDE * r);
  float2 cc_op = g[2];
  if (!(cc_op.y > 0.f)) return;  // not in view
  float2 ca_cb = g[1];
  int64_t i = idx ? idx[r] : r;
  float ca = ca_cb.x, cb = ca_cb.y, cc = cc_op.x;
  float detc = ca * cc - cb * cb;
  float cxx = cc / detc, cyy = ca / detc, cxy = -cb / detc;
  float mid = 0.5f * (cxx + cyy);
  float disc = mid * mid - (cxx * cyy - cxy * cxy);
  float lam = mid + sqrtf(fmaxf(disc, 0.f));
  float scale_px = sqrtf(fmaxf(lam, 0.f));
  if (scale_px > max_scale_px[i]) max_scale_px[i] = scale_px;
  float v = vis[(int64_t)vis_stride * r];
  // fused path (clear_acc given: heur IS the raster accumulator): the heuristics of a row no pixel blended are exactly
  // zero (the raster backward is their only writer and the projection zeroed the row), so that sector is not fetched
  // for the ~97 % of in-view rows that were not blended
  float h_prune = 0.f, h_split = 0.f;
  if (v > 0.f || !clear_acc) { h_prune = heur[(int64_t)heur_stride * r]; h_split = heur[(int64_t)heur_stride * r + 1]; }
  if (v > 0.f) {
    points_in_view[i] += 1;
    visibility[i] += v;
    if (scene_visible) scene_visible[i] += v;
  }
  float s_old = split_score[i], p_old = prune_cost[i];
  float s_new = gs_exp_lerp(split_alpha, s_old, h_split), p_new = gs_exp_lerp(prune_alpha, p_old, h_prune);
  if (s_new != s_old) split_score[i] = s_new;  // a never-blended row stays 0: do not dirty its sector
  if (p_new != p_old) prune_cost[i] = p_new;
  if (v > 0.f) {
    if (clear_acc) {
      float *a = clear_acc + (int64_t)clear_stride * r;
      for (int k = 0; k < clear_stride; ++k) a[k] = 0.f;
    }
    if (clear_vis) clear_vis[r] = 0.f;
  }
}

extern "C" int gs_accumulate_stats(int64_t m, const int64_t *idx, const float *g2d, const float *vis, int vis_stride,
                                   const float *heur, int heur_stride, float split_alpha, float prune_alpha,
                                   float *prune_cost, float *split_score, float *max_scale_px, int16_t *points_in_view,
                                   float *visibility, float *scene_visible, float *clear_acc, int clear_stride,
                                   float *clear_vis, void *stream) {
  GS_REQUIRE(g2d && vis && heur && prune_cost && split_score && max_scale_px && points_in_view && visibility, "null argument");
  if (m <= 0) return GS_OK;
  k_accumulate_stats<<<gs_div_up(m, 256), 256, 0, (cudaStream_t)stream>>>(m, idx, g2d, vis, vis_stride, heur, heur_stride,
                                                                          split_alpha, prune_alpha, prune_cost, split_score,
                                                                          max_scale_px, points_in_view, visibility, scene_visible, clear_acc, clear_stride, clear_vis);
  GS_LAUNCH_CHECK();
  return GS_OK;
}
