// cluster.cu — point clustering for the view-similarity features: assign_clusters / kmeans_iter,
// reference visibility/cluster.py:139-171 (pykeops LazyTensor argmin of ((x_i - c_j)^2).sum(-1)), called by
// PointClusters.cluster (:22) on the splat positions (1024 clusters, 100 iterations) at every evaluation.
//
// assign: one thread per 4 points, all centroids in shared memory (12 KB at 1024 x 3), every centroid read is a
// broadcast; the squared distance is summed in the reference's order with explicit round-to-nearest mul / add (no FMA
// contraction), ties go to the lowest index, a NaN centroid is never nearest: labels equal the oracle's bit for bit.
// update: per-cluster sums by fire-and-forget global fp32 reductions (RED.ADD; fp32 atomics on SHARED memory would be
// CAS loops) into a k x (d+1) table, then centroid = sum / count (0/0 = NaN for an emptied cluster, like the reference).
#include "common.cuh"

namespace {

constexpr int KM_THREADS = 256, KM_PTS = 4, KM_MAX_D = 8;

template <int D>
__global__ void __launch_bounds__(KM_THREADS) k_assign(int64_t n, int k, const float *__restrict__ x,
                                                       const float *__restrict__ centroids, int32_t *__restrict__ labels,
                                                       float *__restrict__ sums) {
  extern __shared__ float sc[];
  for (int i = threadIdx.x; i < k * D; i += KM_THREADS) sc[i] = centroids[i];
  __syncthreads();
  for (int64_t i0 = ((int64_t)blockIdx.x * KM_THREADS + threadIdx.x) * KM_PTS; i0 < n;
       i0 += (int64_t)gridDim.x * KM_THREADS * KM_PTS) {
    float p[KM_PTS][D], best[KM_PTS];
    int bj[KM_PTS];
#pragma unroll
    for (int a = 0; a < KM_PTS; ++a) {
      best[a] = INFINITY; bj[a] = 0;
#pragma unroll
      for (int q = 0; q < D; ++q) p[a][q] = (i0 + a < n) ? x[(i0 + a) * D + q] : 0.f;
    }
    for (int j = 0; j < k; ++j) {
      float c[D];
#pragma unroll
      for (int q = 0; q < D; ++q) c[q] = sc[j * D + q];
#pragma unroll
      for (int a = 0; a < KM_PTS; ++a) {
        float s = 0.f;
#pragma unroll
        for (int q = 0; q < D; ++q) {
          float t = __fsub_rn(p[a][q], c[q]);
          s = __fadd_rn(s, __fmul_rn(t, t));
        }
        if (s < best[a]) { best[a] = s; bj[a] = j; }
      }
    }
#pragma unroll
    for (int a = 0; a < KM_PTS; ++a) {
      if (i0 + a >= n) continue;
      labels[i0 + a] = bj[a];
      if (sums) {
        float *dst = sums + (int64_t)bj[a] * (D + 1);
#pragma unroll
        for (int q = 0; q < D; ++q) atomicAdd(dst + q, p[a][q]);
        atomicAdd(dst + D, 1.0f);
      }
    }
  }
}

__global__ void k_update(int k, int d, float *__restrict__ centroids, float *__restrict__ sums) {
  int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= k) return;
  float cnt = sums[(int64_t)j * (d + 1) + d];
  for (int q = 0; q < d; ++q) {
    centroids[(int64_t)j * d + q] = __fdiv_rn(sums[(int64_t)j * (d + 1) + q], cnt);  // 0/0 = NaN: emptied cluster
    sums[(int64_t)j * (d + 1) + q] = 0.f;
  }
  sums[(int64_t)j * (d + 1) + d] = 0.f;
}

template <int D>
int launch_assign(int64_t n, int k, const float *x, const float *centroids, int32_t *labels, float *sums, cudaStream_t st) {
  int64_t want = (n + KM_THREADS * KM_PTS - 1) / (KM_THREADS * KM_PTS);
  int grid = (int)(want < 148 * 8 ? want : 148 * 8);
  k_assign<D><<<grid, KM_THREADS, (size_t)k * D * sizeof(float), st>>>(n, k, x, centroids, labels, sums);
  return 0;
}

int dispatch_assign(int64_t n, int d, int k, const float *x, const float *centroids, int32_t *labels, float *sums,
                    cudaStream_t st) {
  switch (d) {
    case 1: return launch_assign<1>(n, k, x, centroids, labels, sums, st);
    case 2: return launch_assign<2>(n, k, x, centroids, labels, sums, st);
    case 3: return launch_assign<3>(n, k, x, centroids, labels, sums, st);
    case 4: return launch_assign<4>(n, k, x, centroids, labels, sums, st);
    case 5: return launch_assign<5>(n, k, x, centroids, labels, sums, st);
    case 6: return launch_assign<6>(n, k, x, centroids, labels, sums, st);
    case 7: return launch_assign<7>(n, k, x, centroids, labels, sums, st);
    default: return launch_assign<8>(n, k, x, centroids, labels, sums, st);
  }
}

}  // namespace

extern "C" size_t gs_kmeans_workspace_bytes(int k, int d) { return gs_align256((size_t)k * (d + 1) * sizeof(float)); }

extern "C" int gs_assign_clusters(int64_t n, int d, int k, const float *x, const float *centroids, int32_t *labels,
                                  void *stream) {
  GS_REQUIRE(x && centroids && labels, "null argument");
  GS_REQUIRE(d >= 1 && d <= KM_MAX_D && k >= 1 && (size_t)k * d * sizeof(float) <= 48 * 1024, "1 <= d <= 8, k*d*4 <= 48 KB");
  if (n <= 0) return GS_OK;
  dispatch_assign(n, d, k, x, centroids, labels, nullptr, (cudaStream_t)stream);
  GS_LAUNCH_CHECK();
  return GS_OK;
}

extern "C" int gs_kmeans(int64_t n, int d, int k, int iters, const float *x, float *centroids, int32_t *labels,
                         void *workspace, size_t workspace_bytes, void *stream) {
  GS_REQUIRE(x && centroids && labels && workspace, "null argument");
  GS_REQUIRE(d >= 1 && d <= KM_MAX_D && k >= 1 && (size_t)k * d * sizeof(float) <= 48 * 1024, "1 <= d <= 8, k*d*4 <= 48 KB");
  if (workspace_bytes < gs_kmeans_workspace_bytes(k, d)) {
    gs_set_error("gs_kmeans: workspace too small");
    return GS_ERR_WORKSPACE;
  }
  if (n <= 0 || iters <= 0) return GS_OK;
  cudaStream_t st = (cudaStream_t)stream;
  float *sums = (float *)workspace;
  GS_CUDA(cudaMemsetAsync(sums, 0, (size_t)k * (d + 1) * sizeof(float), st));
  for (int it = 0; it < iters; ++it) {  // stream-ordered, no host synchronisation between iterations
    dispatch_assign(n, d, k, x, centroids, labels, sums, st);
    k_update<<<gs_div_up(k, 128), 128, 0, st>>>(k, d, centroids, sums);
  }
  GS_LAUNCH_CHECK();
  return GS_OK;
}

// ---- per-view cluster features + frustum visibility counts (SURVEY §8f-3 remainder) ---------------------------------
// PointClusters.view_features (reference visibility/cluster.py:40-55): features[label[i]] += vis[i] over the rows with
// vis[i] > threshold.  A segmented reduction over the visibility plane of one render: block-local partial sums in shared
// memory (one fp32 slot per cluster), one global RED per non-empty (block, cluster).
constexpr int VF_THREADS = 256;
constexpr int VF_ITEMS = 16;

__global__ void __launch_bounds__(VF_THREADS) k_view_features(int64_t n, const float *__restrict__ vis, const int64_t *__restrict__ idx,
                                                              int64_t v, const int32_t *__restrict__ labels, float threshold,
                                                              int k, float *__restrict__ out) {
  extern __shared__ float s_feat[];
  for (int c = threadIdx.x; c < k; c += VF_THREADS) s_feat[c] = 0.f;
  __syncthreads();
  const int64_t count = idx ? v : n;
  int64_t base = (int64_t)blockIdx.x * VF_THREADS * VF_ITEMS;
  for (int it = 0; it < VF_ITEMS; ++it) {
    int64_t j = base + (int64_t)it * VF_THREADS + threadIdx.x;
    if (j >= count) break;
    float w = vis[j];
    if (w > threshold) {
      int64_t row = idx ? idx[j] : j;
      atomicAdd(&s_feat[labels[row]], w);
    }
  }
  __syncthreads();
  for (int c = threadIdx.x; c < k; c += VF_THREADS) {
    float s = s_feat[c];
    if (s != 0.f) atomicAdd(out + c, s);
  }
}

extern "C" int gs_view_features(int64_t n, const float *vis, const int64_t *indexes, int64_t v, const int32_t *labels,
                                float threshold, int num_clusters, float *features, void *stream) {
  GS_REQUIRE(vis && labels && features, "null argument");
  GS_REQUIRE(num_clusters >= 1 && num_clusters <= 12288, "1 <= num_clusters <= 12288");
  cudaStream_t st = (cudaStream_t)stream;
  GS_CUDA(cudaMemsetAsync(features, 0, (size_t)num_clusters * sizeof(float), st));
  int64_t count = indexes ? v : n;
  if (count <= 0) return GS_OK;
  k_view_features<<<gs_div_up(count, VF_THREADS * VF_ITEMS), VF_THREADS, (size_t)num_clusters * sizeof(float), st>>>(
      n, vis, indexes, v, labels, threshold, num_clusters, features);
  GS_LAUNCH_CHECK();
  return GS_OK;
}

// point_visibility (reference visibility/query_points.py:62-92): in how many cameras does a point project inside the
// image with near < depth < far.  cameras: (C, 20) floats per camera = image_t_world (row-major 4x4 = K·T_camera_world)
// | width, height, near, far.  A thread per point, the camera table in shared memory.
constexpr int PV_THREADS = 256;
constexpr int PV_CAM_FLOATS = 20;

__global__ void __launch_bounds__(PV_THREADS) k_point_visibility(int64_t n, const float *__restrict__ points, int num_cameras,
                                                                 const float *__restrict__ cameras, int32_t *__restrict__ counts) {
  extern __shared__ float s_cam[];
  int64_t i = (int64_t)blockIdx.x * PV_THREADS + threadIdx.x;
  float x = 0.f, y = 0.f, z = 0.f;
  if (i < n) { x = points[3 * i]; y = points[3 * i + 1]; z = points[3 * i + 2]; }
  int total = 0;
  const int chunk = 512;  // cameras per shared-memory refill (40 KB)
  for (int c0 = 0; c0 < num_cameras; c0 += chunk) {
    int cn = min(chunk, num_cameras - c0);
    __syncthreads();
    for (int t = threadIdx.x; t < cn * PV_CAM_FLOATS; t += PV_THREADS) s_cam[t] = cameras[(size_t)c0 * PV_CAM_FLOATS + t];
    __syncthreads();
    for (int c = 0; c < cn; ++c) {
      const float *M = s_cam + c * PV_CAM_FLOATS;
      // the reference's matrix-vector product, term by term in row order (no contraction)
      float px = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(M[0], x), __fmul_rn(M[1], y)), __fmul_rn(M[2], z)), M[3]);
      float py = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(M[4], x), __fmul_rn(M[5], y)), __fmul_rn(M[6], z)), M[7]);
      float d = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(M[8], x), __fmul_rn(M[9], y)), __fmul_rn(M[10], z)), M[11]);
      float u = __fdiv_rn(px, d), w = __fdiv_rn(py, d);
      total += (u >= 0.f) && (u < M[16]) && (w >= 0.f) && (w < M[17]) && (d > M[18]) && (d < M[19]);
    }
  }
  if (i < n) counts[i] = total;
}

extern "C" int gs_point_visibility(int64_t n, const float *points, int num_cameras, const float *cameras, int32_t *counts,
                                   void *stream) {
  GS_REQUIRE(points && cameras && counts && num_cameras >= 0, "null argument");
  if (n <= 0) return GS_OK;
  int cn = num_cameras < 512 ? (num_cameras > 0 ? num_cameras : 1) : 512;
  k_point_visibility<<<gs_div_up(n, PV_THREADS), PV_THREADS, (size_t)cn * PV_CAM_FLOATS * sizeof(float), (cudaStream_t)stream>>>(
      n, points, num_cameras, cameras, counts);
  GS_LAUNCH_CHECK();
  return GS_OK;
}
