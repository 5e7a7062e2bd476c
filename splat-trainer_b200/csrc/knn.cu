// knn.cu — estimate_scale: mean distance to the num_neighbors nearest neighbours of every point,
// reference gaussians/loading.py:45-61 (pykeops Kmin over the full N x N distance matrix), used to initialise
// log_scaling from the point cloud (gaussians/loading.py:23-34).
//
// Instead of N^2 distances: a uniform grid over the bounding box (~2 points per cell), points sorted by cell with the
// library's radix sort, and per point a search over growing cubic shells of cells that stops once the current
// (k+1)-th smallest squared distance is inside the explored radius.  The squared distance is summed in the reference's
// order with round-to-nearest mul / add, so the k+1 smallest values — and the result — equal the brute-force oracle's
// bit for bit; self (distance 0) is among the candidates and the smallest is dropped, exactly like the reference.
#include <math.h>

#include "common.cuh"

extern "C" size_t gs_radix_sort_workspace_bytes(int64_t n);
extern "C" int gs_radix_sort_pairs(uint32_t *keys_in, uint32_t *vals_in, uint32_t *keys_out, uint32_t *vals_out, int64_t n,
                                   const int32_t *n_dev, int begin_bit, int end_bit, void *workspace, size_t workspace_bytes,
                                   void *stream);

namespace {

constexpr int KNN_MAX = 8;  // num_neighbors <= 8

struct Grid {
  float lo[3], cell, inv_cell;
  int dim[3];
};

__global__ void __launch_bounds__(256) k_bbox(int64_t n, const float *__restrict__ p, float *__restrict__ partial) {
  float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
#pragma unroll
    for (int q = 0; q < 3; ++q) { float v = p[3 * i + q]; lo[q] = fminf(lo[q], v); hi[q] = fmaxf(hi[q], v); }
  __shared__ float s[6][8];
#pragma unroll
  for (int q = 0; q < 3; ++q) {
    for (int o = 16; o > 0; o >>= 1) {
      lo[q] = fminf(lo[q], __shfl_xor_sync(GS_FULL, lo[q], o));
      hi[q] = fmaxf(hi[q], __shfl_xor_sync(GS_FULL, hi[q], o));
    }
    if ((threadIdx.x & 31) == 0) { s[q][threadIdx.x >> 5] = lo[q]; s[3 + q][threadIdx.x >> 5] = hi[q]; }
  }
  __syncthreads();
  if (threadIdx.x < 6) {
    float v = s[threadIdx.x][0];
    for (int w = 1; w < 8; ++w) v = threadIdx.x < 3 ? fminf(v, s[threadIdx.x][w]) : fmaxf(v, s[threadIdx.x][w]);
    partial[6 * blockIdx.x + threadIdx.x] = v;
  }
}

__device__ __forceinline__ void cell_of(const Grid &g, const float *x, int *c) {
#pragma unroll
  for (int q = 0; q < 3; ++q) {
    int v = (int)floorf((x[q] - g.lo[q]) * g.inv_cell);
    c[q] = min(max(v, 0), g.dim[q] - 1);
  }
}

__global__ void __launch_bounds__(256) k_cells(int64_t n, Grid g, const float *__restrict__ p, uint32_t *__restrict__ keys,
                                               uint32_t *__restrict__ ids) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  float x[3] = {p[3 * i], p[3 * i + 1], p[3 * i + 2]};
  int c[3];
  cell_of(g, x, c);
  keys[i] = (uint32_t)((c[2] * g.dim[1] + c[1]) * g.dim[0] + c[0]);
  ids[i] = (uint32_t)i;
}

__global__ void __launch_bounds__(256) k_bounds(int64_t n, const uint32_t *__restrict__ keys, int32_t *__restrict__ start,
                                                int32_t *__restrict__ end) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint32_t k = keys[i];
  if (i == 0 || keys[i - 1] != k) start[k] = (int32_t)i;
  if (i == n - 1 || keys[i + 1] != k) end[k] = (int32_t)i + 1;
}

// sorted copy of the points: neighbours of a cell are contiguous
__global__ void __launch_bounds__(256) k_gather(int64_t n, const float *__restrict__ p, const uint32_t *__restrict__ ids,
                                                float *__restrict__ sorted) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  int64_t j = ids[i];
  sorted[3 * i] = p[3 * j]; sorted[3 * i + 1] = p[3 * j + 1]; sorted[3 * i + 2] = p[3 * j + 2];
}

__global__ void __launch_bounds__(128) k_knn(int64_t n, Grid g, int K, const float *__restrict__ sorted,
                                             const uint32_t *__restrict__ ids, const int32_t *__restrict__ start,
                                             const int32_t *__restrict__ end, float *__restrict__ scales) {
  int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= n) return;
  float x[3] = {sorted[3 * s], sorted[3 * s + 1], sorted[3 * s + 2]};
  int c[3];
  cell_of(g, x, c);
  float best[KNN_MAX + 1];
#pragma unroll
  for (int q = 0; q <= KNN_MAX; ++q) best[q] = INFINITY;
  int rmax = max(max(g.dim[0], g.dim[1]), g.dim[2]);
  for (int r = 0; r <= rmax; ++r) {
    int z0 = max(c[2] - r, 0), z1 = min(c[2] + r, g.dim[2] - 1);
    int y0 = max(c[1] - r, 0), y1 = min(c[1] + r, g.dim[1] - 1);
    int x0 = max(c[0] - r, 0), x1 = min(c[0] + r, g.dim[0] - 1);
    for (int cz = z0; cz <= z1; ++cz)
      for (int cy = y0; cy <= y1; ++cy) {
        bool face = (abs(cz - c[2]) == r) || (abs(cy - c[1]) == r);
        // on a z / y face of the shell every x cell belongs to it; otherwise only the two x ends
        for (int cx = x0; cx <= x1; cx += (face ? 1 : max(x1 - x0, 1))) {
          if (!face && abs(cx - c[0]) != r) continue;
          int64_t cell = ((int64_t)cz * g.dim[1] + cy) * g.dim[0] + cx;
          int b = start[cell];
          if (b < 0) continue;
          int e = end[cell];
          for (int j = b; j < e; ++j) {
            float d = 0.f;
#pragma unroll
            for (int q = 0; q < 3; ++q) {
              float t = __fsub_rn(x[q], sorted[3 * (int64_t)j + q]);
              d = __fadd_rn(d, __fmul_rn(t, t));
            }
            if (d < best[K - 1]) {
#pragma unroll
              for (int q = KNN_MAX; q > 0; --q) {  // insertion into the ascending list (fixed trip count: stays in registers)
                if (q < K) {
                  float below = best[q - 1];  // still the old value: slots are rewritten from the top down
                  best[q] = below > d ? below : fminf(best[q], d);
                }
              }
              best[0] = fminf(best[0], d);
            }
          }
        }
      }
    // every unexplored cell is farther than r * cell in at least one axis
    float reach = (float)r * g.cell;
    if (best[K - 1] <= reach * reach * 0.9999f) break;
  }
  float m = 0.f;
  for (int q = 1; q < K; ++q) m += sqrtf(best[q]);
  scales[ids[s]] = m / (float)(K - 1);
}

}  // namespace

static size_t knn_max_cells(int64_t n) {  // the grid aims at n/2 cells; flat or clumpy clouds may need a few times more
  size_t want = (size_t)(n < 1024 ? 1024 : n) * 4;
  return want < ((size_t)1 << 25) ? want : ((size_t)1 << 25);
}

extern "C" size_t gs_knn_scale_workspace_bytes(int64_t n) {
  size_t cells = knn_max_cells(n);
  return gs_align256(1024 * 6 * sizeof(float)) + 4 * gs_align256((size_t)n * 4) + gs_align256((size_t)n * 12) +
         2 * gs_align256(cells * 4) + gs_align256(gs_radix_sort_workspace_bytes(n));
}

extern "C" int gs_knn_scale(int64_t n, const float *points, int num_neighbors, float *scales, void *workspace,
                            size_t workspace_bytes, void *stream) {
  GS_REQUIRE(points && scales && workspace, "null argument");
  GS_REQUIRE(num_neighbors >= 1 && num_neighbors <= KNN_MAX, "1 <= num_neighbors <= 8");
  GS_REQUIRE(n < ((int64_t)1 << 31), "n must fit 31 bits");
  if (workspace_bytes < gs_knn_scale_workspace_bytes(n)) { gs_set_error("gs_knn_scale: workspace too small"); return GS_ERR_WORKSPACE; }
  if (n <= 0) return GS_OK;
  cudaStream_t st = (cudaStream_t)stream;
  unsigned char *w = (unsigned char *)workspace;
  float *partial = (float *)w; w += gs_align256(1024 * 6 * sizeof(float));
  uint32_t *keys_a = (uint32_t *)w; w += gs_align256((size_t)n * 4);
  uint32_t *ids_a = (uint32_t *)w; w += gs_align256((size_t)n * 4);
  uint32_t *keys_b = (uint32_t *)w; w += gs_align256((size_t)n * 4);
  uint32_t *ids_b = (uint32_t *)w; w += gs_align256((size_t)n * 4);
  float *sorted = (float *)w; w += gs_align256((size_t)n * 12);
  int32_t *start = (int32_t *)w; w += gs_align256(knn_max_cells(n) * 4);
  int32_t *end = (int32_t *)w; w += gs_align256(knn_max_cells(n) * 4);
  void *sort_ws = w;
  // bounding box: one host read (initialisation-time call; the reference reads the cloud on the host anyway)
  int blocks = (int)((n + 255) / 256 < 1024 ? (n + 255) / 256 : 1024);
  k_bbox<<<blocks, 256, 0, st>>>(n, points, partial);
  GS_LAUNCH_CHECK();
  static float h_partial[1024 * 6];
  GS_CUDA(cudaMemcpyAsync(h_partial, partial, (size_t)blocks * 6 * sizeof(float), cudaMemcpyDeviceToHost, st));
  GS_CUDA(cudaStreamSynchronize(st));
  float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
  for (int b = 0; b < blocks; ++b)
    for (int q = 0; q < 3; ++q) { lo[q] = fminf(lo[q], h_partial[6 * b + q]); hi[q] = fmaxf(hi[q], h_partial[6 * b + 3 + q]); }
  GS_REQUIRE(isfinite(lo[0]) && isfinite(hi[0]) && isfinite(lo[1]) && isfinite(hi[1]) && isfinite(lo[2]) && isfinite(hi[2]),
             "points must be finite");
  // ~2 points per cell on average, at most 2^25 cells
  double ext[3], vol = 1.0;
  for (int q = 0; q < 3; ++q) { ext[q] = fmax((double)hi[q] - lo[q], 1e-6); vol *= ext[q]; }
  double cell = cbrt(vol / fmax((double)n / 2.0, 1.0));
  Grid g;
  for (;;) {  // the cells must cover the whole box: grow the cell until the grid fits the table
    double cells = 1.0, widest = 0.0;
    for (int q = 0; q < 3; ++q) {
      double dq = fmax(ceil(ext[q] / cell), 1.0);
      widest = fmax(widest, dq);
      cells *= dq;
      g.dim[q] = (int)fmin(dq, 1 << 20);
    }
    if (cells <= (double)knn_max_cells(n) && widest <= 1024.0) break;
    cell *= 1.26;
  }
  for (int q = 0; q < 3; ++q) g.lo[q] = lo[q];
  g.cell = (float)cell; g.inv_cell = (float)(1.0 / cell);
  size_t cells = (size_t)g.dim[0] * g.dim[1] * g.dim[2];
  int bits = 1;
  while (((size_t)1 << bits) < cells) ++bits;
  GS_CUDA(cudaMemsetAsync(start, 0xff, cells * 4, st));
  k_cells<<<gs_div_up(n, 256), 256, 0, st>>>(n, g, points, keys_a, ids_a);
  GS_LAUNCH_CHECK();
  int rc = gs_radix_sort_pairs(keys_a, ids_a, keys_b, ids_b, n, nullptr, 0, bits, sort_ws, gs_radix_sort_workspace_bytes(n), st);
  if (rc != GS_OK) return rc;
  k_bounds<<<gs_div_up(n, 256), 256, 0, st>>>(n, keys_b, start, end);
  k_gather<<<gs_div_up(n, 256), 256, 0, st>>>(n, points, ids_b, sorted);
  k_knn<<<gs_div_up(n, 128), 128, 0, st>>>(n, g, num_neighbors + 1, sorted, ids_b, start, end, scales);
  GS_LAUNCH_CHECK();
  return GS_OK;
}
