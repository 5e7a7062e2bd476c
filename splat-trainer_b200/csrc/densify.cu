// densify.cu — split / clone / prune as stream-compaction kernels.
//   gs_select_topk_mask  take_n                      reference controller/target_controller.py:150-160
//   gs_compact_*         points[keep_mask]           reference scene/mlp_scene.py:308
//   gs_split_positions   split_gaussians_uniform     reference gaussians/split.py:87-113
//   gs_gather_rows       repeat_interleave gather    reference gaussians/split.py:42-44
// The reference does two full argsorts of N values per densify step; here the k-th value is found by
// a 4-pass MSB radix select (4 reads of N keys, 256-bin histograms) and ties at the threshold are
// resolved by index with one scan, so exactly min(k, n) rows are selected and the result is
// deterministic on every rank.  Integer / byte work, bit-exact against the oracle.
#include "common.cuh"

// float -> uint key whose unsigned order equals the oracle's comparison (ascending, NaN last, -0 == +0)
__device__ __forceinline__ uint32_t order_key(float v, int descending) {
  if (v != v) return 0xffffffffu;
  if (descending) v = -v;
  if (v == 0.f) v = 0.f;
  uint32_t u = __float_as_uint(v);
  u = (u & 0x80000000u) ? ~u : (u | 0x80000000u);
  return u == 0xffffffffu ? 0xfffffffeu : u;  // unreachable for non-NaN input; keeps NaN strictly last
}

// state[0] = prefix (high bits decided so far), state[1] = remaining k, state[2] = number of ties to take
__global__ void __launch_bounds__(256) k_select_hist(const float *__restrict__ values, int64_t n, int descending, int pass,
                                                     const uint32_t *__restrict__ state, uint32_t *__restrict__ hist) {
  __shared__ uint32_t h[256];
  h[threadIdx.x] = 0;
  __syncthreads();
  int shift = 24 - 8 * pass;
  uint32_t prefix = state[0];
  uint32_t pmask = pass == 0 ? 0u : (0xffffffffu << (shift + 8));
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    uint32_t key = order_key(values[i], descending);
    if ((key & pmask) == prefix) atomicAdd(&h[(key >> shift) & 255u], 1u);
  }
  __syncthreads();
  if (h[threadIdx.x]) atomicAdd(&hist[pass * 256 + threadIdx.x], h[threadIdx.x]);
}

__global__ void k_select_digit(int pass, uint32_t *__restrict__ state, const uint32_t *__restrict__ hist) {
  if (threadIdx.x != 0) return;
  uint32_t k = state[1];
  int shift = 24 - 8 * pass;
  uint32_t run = 0;
  int d = 0;
  for (; d < 256; ++d) {
    uint32_t c = hist[pass * 256 + d];
    if (run + c >= k) break;
    run += c;
  }
  if (d == 256) d = 255;
  state[0] |= (uint32_t)d << shift;
  state[1] = k - run;  // k-th element is the (k-run)-th inside this bucket
  if (pass == 3) state[2] = k - run;
}

__global__ void __launch_bounds__(256) k_select_flags(const float *__restrict__ values, int64_t n, int descending,
                                                      const uint32_t *__restrict__ state, int32_t *__restrict__ tie_flags,
                                                      uint8_t *__restrict__ mask) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint32_t key = order_key(values[i], descending), thr = state[0];
  tie_flags[i] = key == thr;
  mask[i] = key < thr;
}

__global__ void __launch_bounds__(256) k_select_ties(int64_t n, const uint32_t *__restrict__ state,
                                                     const int32_t *__restrict__ tie_flags, const int32_t *__restrict__ tie_rank,
                                                     uint8_t *__restrict__ mask) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  if (tie_flags[i] && (uint32_t)tie_rank[i] < state[2]) mask[i] = 1;
}

struct SelWs {
  uint32_t *state, *hist;
  int32_t *flags, *rank, *partials;
  size_t total;
};
static SelWs sel_carve(void *base, int64_t n) {
  SelWs w;
  char *p = (char *)base;
  auto take = [&](size_t b) { char *r = p; p += gs_align256(b); return r; };
  size_t nn = (size_t)(n > 0 ? n : 1);
  w.state = (uint32_t *)take(16);
  w.hist = (uint32_t *)take(4 * 256 * 4);
  w.flags = (int32_t *)take(nn * 4);
  w.rank = (int32_t *)take(nn * 4);
  w.partials = (int32_t *)take((size_t)gs_scan_partials(nn) * 4);
  w.total = (size_t)(p - (char *)base);
  return w;
}

extern "C" size_t gs_select_topk_workspace_bytes(int64_t n) { return sel_carve(nullptr, n).total + 256; }

extern "C" int gs_select_topk_mask(const float *values, int64_t n, int64_t k, int descending, uint8_t *mask, void *workspace,
                                   size_t workspace_bytes, void *stream) {
  GS_REQUIRE(n >= 0 && n < ((int64_t)1 << 31), "n must fit int32");
  cudaStream_t st = (cudaStream_t)stream;
  if (n == 0) return GS_OK;
  GS_REQUIRE(values && mask && workspace, "null argument");
  if (k <= 0) { GS_CUDA(cudaMemsetAsync(mask, 0, (size_t)n, st)); return GS_OK; }
  if (k >= n) { GS_CUDA(cudaMemsetAsync(mask, 1, (size_t)n, st)); return GS_OK; }
  uintptr_t aligned = ((uintptr_t)workspace + 255) & ~(uintptr_t)255;
  SelWs w = sel_carve((void *)aligned, n);
  if (workspace_bytes < w.total + (aligned - (uintptr_t)workspace)) {
    gs_set_error("gs_select_topk_mask: workspace too small");
    return GS_ERR_WORKSPACE;
  }
  uint32_t init[4] = {0u, (uint32_t)k, 0u, 0u};
  GS_CUDA(cudaMemcpyAsync(w.state, init, 16, cudaMemcpyHostToDevice, st));
  GS_CUDA(cudaMemsetAsync(w.hist, 0, 4 * 256 * 4, st));
  int blocks = gs_div_up(n, 256);
  int hb = blocks < 148 * 8 ? blocks : 148 * 8;
  for (int pass = 0; pass < 4; ++pass) {
    k_select_hist<<<hb, 256, 0, st>>>(values, n, descending, pass, w.state, w.hist);
    k_select_digit<<<1, 32, 0, st>>>(pass, w.state, w.hist);
  }
  k_select_flags<<<blocks, 256, 0, st>>>(values, n, descending, w.state, w.flags, mask);
  int rc = gs_exclusive_scan_i32(w.flags, nullptr, w.rank, n, nullptr, w.partials, nullptr, st);
  if (rc != GS_OK) return rc;
  k_select_ties<<<blocks, 256, 0, st>>>(n, w.state, w.flags, w.rank, mask);
  GS_LAUNCH_CHECK();
  return GS_OK;
}

// ---- compaction ------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_mask_to_flags(const uint8_t *__restrict__ mask, int64_t n, int32_t *__restrict__ flags) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) flags[i] = mask[i] != 0;
}
__global__ void __launch_bounds__(256) k_finish_slots(const uint8_t *__restrict__ mask, int64_t n, int32_t *__restrict__ slots) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n && !mask[i]) slots[i] = -1;
}

extern "C" size_t gs_compact_workspace_bytes(int64_t n) {
  size_t nn = (size_t)(n > 0 ? n : 1);
  return gs_align256(nn * 4) + gs_align256((size_t)gs_scan_partials(nn) * 4) + 256;
}

extern "C" int gs_compact_slots(const uint8_t *keep_mask, int64_t n, int32_t *slots, int32_t *count_dev, void *workspace,
                                size_t workspace_bytes, void *stream) {
  GS_REQUIRE(n >= 0 && n < ((int64_t)1 << 31), "n must fit int32");
  cudaStream_t st = (cudaStream_t)stream;
  if (n == 0) { if (count_dev) GS_CUDA(cudaMemsetAsync(count_dev, 0, 4, st)); return GS_OK; }
  GS_REQUIRE(keep_mask && slots && workspace, "null argument");
  if (workspace_bytes < gs_compact_workspace_bytes(n)) { gs_set_error("gs_compact_slots: workspace too small"); return GS_ERR_WORKSPACE; }
  uintptr_t aligned = ((uintptr_t)workspace + 255) & ~(uintptr_t)255;
  int32_t *flags = (int32_t *)aligned;
  int32_t *partials = (int32_t *)(aligned + gs_align256((size_t)n * 4));
  int blocks = gs_div_up(n, 256);
  k_mask_to_flags<<<blocks, 256, 0, st>>>(keep_mask, n, flags);
  int rc = gs_exclusive_scan_i32(flags, nullptr, slots, n, nullptr, partials, count_dev, st);
  if (rc != GS_OK) return rc;
  k_finish_slots<<<blocks, 256, 0, st>>>(keep_mask, n, slots);
  GS_LAUNCH_CHECK();
  return GS_OK;
}

template <typename W>
__global__ void __launch_bounds__(256) k_compact_rows(const W *__restrict__ src, W *__restrict__ dst, int64_t n, int words,
                                                      const int32_t *__restrict__ slots) {
  int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n * words) return;
  int64_t row = t / words;
  int col = (int)(t - row * words);
  int s = slots[row];
  if (s >= 0) dst[(int64_t)s * words + col] = src[t];
}

extern "C" int gs_compact_rows(const void *src, void *dst, int64_t n, int row_bytes, const int32_t *slots, void *stream) {
  GS_REQUIRE(row_bytes > 0, "row_bytes must be positive");
  if (n <= 0) return GS_OK;
  GS_REQUIRE(src && dst && slots, "null argument");
  cudaStream_t st = (cudaStream_t)stream;
  bool w4 = row_bytes % 4 == 0 && ((uintptr_t)src % 4 == 0) && ((uintptr_t)dst % 4 == 0);
  if (w4) {
    int words = row_bytes / 4;
    k_compact_rows<uint32_t><<<gs_div_up(n * words, 256), 256, 0, st>>>((const uint32_t *)src, (uint32_t *)dst, n, words, slots);
  } else {
    k_compact_rows<uint8_t><<<gs_div_up(n * row_bytes, 256), 256, 0, st>>>((const uint8_t *)src, (uint8_t *)dst, n, row_bytes, slots);
  }
  GS_LAUNCH_CHECK();
  return GS_OK;
}

// ---- split -----------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128) k_split_positions(int64_t s, const int64_t *__restrict__ split_idx,
                                                         const float *__restrict__ position, const float *__restrict__ log_scaling,
                                                         const float *__restrict__ rotation, const float *__restrict__ uniform,
                                                         float sep, float *__restrict__ out_position,
                                                         float *__restrict__ out_log_scaling, int64_t out_offset) {
  int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= s) return;
  int64_t i = split_idx[j];
  const float eps = 1e-4f;
  const float lg = -0.34657359027997264f;  // log(1/sqrt(2))
  float ls[3] = {log_scaling[3 * i], log_scaling[3 * i + 1], log_scaling[3 * i + 2]};
  int axis = 0;
  if (uniform) {
    float sc[3], tot = 0.f;
#pragma unroll
    for (int k = 0; k < 3; ++k) { sc[k] = fmaxf(expf(ls[k]), eps); tot += sc[k]; }
    float u = uniform[j] * tot;
    axis = u < sc[0] ? 0 : (u < sc[0] + sc[1] ? 1 : 2);
  } else {
    if (ls[1] > ls[axis]) axis = 1;
    if (ls[2] > ls[axis]) axis = 2;
  }
  float x = rotation[4 * i], y = rotation[4 * i + 1], z = rotation[4 * i + 2], w = rotation[4 * i + 3];
  float d = fmaxf(sqrtf(x * x + y * y + z * z + w * w), 1e-12f);
  x /= d; y /= d; z /= d; w /= d;
  float R[9];
  R[0] = 1.f - 2.f * (y * y + z * z); R[1] = 2.f * (x * y - w * z); R[2] = 2.f * (x * z + w * y);
  R[3] = 2.f * (x * y + w * z); R[4] = 1.f - 2.f * (x * x + z * z); R[5] = 2.f * (y * z - w * x);
  R[6] = 2.f * (x * z - w * y); R[7] = 2.f * (y * z + w * x); R[8] = 1.f - 2.f * (x * x + y * y);
  float sa = fmaxf(expf(ls[axis]), eps);
#pragma unroll
  for (int c = 0; c < 2; ++c) {
    float v = c == 0 ? -sep : sep;
    int64_t o = out_offset + 2 * j + c;
#pragma unroll
    for (int r = 0; r < 3; ++r) out_position[3 * o + r] = position[3 * i + r] + (R[3 * r + axis] * sa) * v;
#pragma unroll
    for (int k = 0; k < 3; ++k) out_log_scaling[3 * o + k] = ls[k] + (k == axis ? lg : 0.f);
  }
}

extern "C" int gs_split_positions(int64_t s, const int64_t *split_idx, const float *position, const float *log_scaling,
                                  const float *rotation, const float *uniform, float sep, float *out_position,
                                  float *out_log_scaling, int64_t out_offset, void *stream) {
  if (s <= 0) return GS_OK;
  GS_REQUIRE(split_idx && position && log_scaling && rotation && out_position && out_log_scaling, "null argument");
  k_split_positions<<<gs_div_up(s, 128), 128, 0, (cudaStream_t)stream>>>(s, split_idx, position, log_scaling, rotation, uniform,
                                                                         sep, out_position, out_log_scaling, out_offset);
  GS_LAUNCH_CHECK();
  return GS_OK;
}

template <typename W>
__global__ void __launch_bounds__(256) k_gather_rows(const W *__restrict__ src, W *__restrict__ dst, int64_t s,
                                                     const int64_t *__restrict__ idx, int words, int repeat, int64_t out_offset) {
  int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t per = (int64_t)words * repeat;
  if (t >= s * per) return;
  int64_t j = t / per;
  int rem = (int)(t - j * per);
  int col = rem % words;
  dst[out_offset * words + t] = src[idx[j] * words + col];
}

extern "C" int gs_gather_rows(const void *src, void *dst, int64_t s, const int64_t *idx, int row_bytes, int repeat,
                              int64_t out_offset, void *stream) {
  GS_REQUIRE(row_bytes > 0 && repeat >= 1, "bad row_bytes/repeat");
  if (s <= 0) return GS_OK;
  GS_REQUIRE(src && dst && idx, "null argument");
  cudaStream_t st = (cudaStream_t)stream;
  bool w4 = row_bytes % 4 == 0 && ((uintptr_t)src % 4 == 0) && ((uintptr_t)dst % 4 == 0);
  if (w4) {
    int words = row_bytes / 4;
    k_gather_rows<uint32_t><<<gs_div_up(s * words * repeat, 256), 256, 0, st>>>((const uint32_t *)src, (uint32_t *)dst, s, idx, words, repeat, out_offset);
  } else {
    k_gather_rows<uint8_t><<<gs_div_up(s * (int64_t)row_bytes * repeat, 256), 256, 0, st>>>((const uint8_t *)src, (uint8_t *)dst, s, idx, row_bytes, repeat, out_offset);
  }
  GS_LAUNCH_CHECK();
  return GS_OK;
}
