// exchange.cu — camera-batch data parallelism: the per-image exchange record.
//
// The reference sums `.grad` over the images of a batch sequentially and folds the controller statistics image by
// image (trainer/trainer.py:500-514, controller/point_state.py:34-50; exp_lerp is order dependent).  Under camera-
// batch DP every rank renders its own image, PACKS what that image contributed into a compact record, and every rank
// then FOLDS all records of the step in image order — so the replicas stay bit-identical to each other and equal to
// the sequential path.  Only rows a pixel actually blended travel (C2: ~47k of 3M rows, 12 MB instead of the 720 MB
// dense gradient buffer), plus three bit/rank planes of n/32 words each.
//
// The fold kernels take one pointer per image: records may live in a PEER GPU's memory (NVLink loads through
// symmetric-memory mappings, no staging copy) or in a local all-gathered buffer.  Every remote load of a fold kernel
// is independent of the others (rank planes instead of searches), so one NVLink round trip is exposed per stage.
//
// Record layout (every section 256-B aligned; `cap` rows, D = Σ group dims floats per row):
//   header   int32 count (rows with visibility > 0, unclamped), int32 overflow (count > cap)
//   inview   uint32[(n+31)/32]   bit i: row i was in view of this image (opacity > 0)
//   visbits  uint32[(n+31)/32]   bit i: row i was blended by a pixel (visibility > 0)
//   base     uint32[(n+31)/32]   number of visible rows before word w (rank of row i = base + popc(bits below i))
//   idx      int32[cap]          ascending visible rows
//   aux      float4[cap]         (visibility, prune heuristic, split heuristic, 0)
//   rows     float[cap·D]        gradient rows, one plane of cap·dim per parameter group
#include "common.cuh"

namespace {

struct RecView {
  int32_t *header;
  uint32_t *inview, *visbits, *base;
  int32_t *idx;
  float4 *aux;
  float *rows;
};

__host__ __device__ inline size_t al256(size_t x) { return (x + 255) & ~(size_t)255; }

__host__ __device__ inline RecView rec_view(void *rec, int64_t n, int64_t cap) {
  size_t words = (size_t)((n + 31) / 32);
  char *p = (char *)rec;
  RecView v;
  v.header = (int32_t *)p; p += 256;
  v.inview = (uint32_t *)p; p += al256(words * 4);
  v.visbits = (uint32_t *)p; p += al256(words * 4);
  v.base = (uint32_t *)p; p += al256(words * 4);
  v.idx = (int32_t *)p; p += al256((size_t)cap * 4);
  v.aux = (float4 *)p; p += al256((size_t)cap * 16);
  v.rows = (float *)p;
  return v;
}

struct GroupsDev {
  int count, total;
  int dim[GS_EXCHANGE_MAX_GROUPS];
  int start[GS_EXCHANGE_MAX_GROUPS];  // first flattened element of the group
  float *grad[GS_EXCHANGE_MAX_GROUPS];
};

struct RecordsDev {
  int count;
  const void *rec[GS_EXCHANGE_MAX_IMAGES];
};

__device__ __forceinline__ float exp_lerp_x(float t, float a, float b) {  // reference util/misc.py:56-59 (= optim.cu)
  float m = fmaxf(a, b);
  float ea = expf(a - m), eb = expf(b - m);
  float l = t < 0.5f ? ea + t * (eb - ea) : eb - (eb - ea) * (1.0f - t);
  return m + logf(l);
}

constexpr int PK_THREADS = 256;
constexpr int PK_ITERS = 8;
constexpr int PK_TILE = PK_THREADS * PK_ITERS;  // rows per CTA; a warp covers one 32-row word per iteration

// Pass A of the pack: one sweep over the image's rows.  Ordered single-pass compaction (chained scan over CTAs) of the
// visible rows, the three bit/rank planes, the rank's own max_scale_px, and the clearing of the raster accumulator.
__global__ void __launch_bounds__(PK_THREADS) k_pack_stats(int64_t n, int64_t cap, const float *__restrict__ g2d,
                                                           const float *__restrict__ vis, int vis_stride,
                                                           const float *__restrict__ heur, int heur_stride,
                                                           float *__restrict__ max_scale_px, float *__restrict__ clear_acc,
                                                           int clear_stride, float *__restrict__ clear_vis, RecView rv,
                                                           uint32_t *__restrict__ status, uint32_t *__restrict__ ticket) {
  __shared__ int s_cnt[PK_ITERS * 8];
  __shared__ int s_tile;
  __shared__ uint32_t s_prefix;
  if (threadIdx.x == 0) s_tile = (int)atomicAdd(ticket, 1u);
  __syncthreads();
  const int tile = s_tile;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  float v[PK_ITERS], hp[PK_ITERS], hs[PK_ITERS];
  unsigned vb[PK_ITERS];
#pragma unroll
  for (int it = 0; it < PK_ITERS; ++it) {
    int64_t r = (int64_t)tile * PK_TILE + it * PK_THREADS + threadIdx.x;
    bool inview = false;
    v[it] = 0.f; hp[it] = 0.f; hs[it] = 0.f;
    if (r < n) {
      const float2 *g = reinterpret_cast<const float2 *>(g2d + GS_G2D_STRIDE * r);
      float2 cc_op = g[2];
      if (cc_op.y > 0.f) {
        inview = true;
        float2 ca_cb = g[1];
        float ca = ca_cb.x, cb = ca_cb.y, cc = cc_op.x;
        float detc = ca * cc - cb * cb;
        float cxx = cc / detc, cyy = ca / detc, cxy = -cb / detc;
        float mid = 0.5f * (cxx + cyy);
        float disc = mid * mid - (cxx * cyy - cxy * cxy);
        float lam = mid + sqrtf(fmaxf(disc, 0.f));
        float scale_px = sqrtf(fmaxf(lam, 0.f));
        if (scale_px > max_scale_px[r]) max_scale_px[r] = scale_px;
        float vv = vis[(int64_t)vis_stride * r];
        if (vv > 0.f) {
          v[it] = vv;
          hp[it] = heur[(int64_t)heur_stride * r];
          hs[it] = heur[(int64_t)heur_stride * r + 1];
          if (clear_acc) {
            float *a = clear_acc + (int64_t)clear_stride * r;
            for (int k = 0; k < clear_stride; ++k) a[k] = 0.f;
          }
          if (clear_vis) clear_vis[r] = 0.f;
        }
      }
    }
    unsigned ib = __ballot_sync(GS_FULL, inview);
    vb[it] = __ballot_sync(GS_FULL, v[it] > 0.f);
    int64_t r0 = (int64_t)tile * PK_TILE + it * PK_THREADS + w * 32;  // first row of this warp's word
    if (lane == 0) {
      s_cnt[it * 8 + w] = __popc(vb[it]);
      if (r0 < n) { rv.inview[r0 >> 5] = ib; rv.visbits[r0 >> 5] = vb[it]; }
    }
  }
  __syncthreads();
  // exclusive scan of the 64 word counts (order: iteration-major = ascending rows)
  int tot = 0;
  if (threadIdx.x < 32) {
    int a = s_cnt[2 * lane], b = s_cnt[2 * lane + 1];
    int inc = a + b;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      int t = __shfl_up_sync(GS_FULL, inc, o);
      if (lane >= o) inc += t;
    }
    tot = __shfl_sync(GS_FULL, inc, 31);
    int ex = inc - (a + b);
    s_cnt[2 * lane] = ex;
    s_cnt[2 * lane + 1] = ex + a;
    if (lane == 0) gs_st_status(status + tile, (uint32_t)tot | (tile == 0 ? GS_LB_INCL : GS_LB_LOCAL));
    uint32_t prefix = gs_lookback_warp(status, tile);
    if (lane == 0) {
      if (tile > 0) gs_st_status(status + tile, (prefix + (uint32_t)tot) | GS_LB_INCL);
      s_prefix = prefix;
      if (tile == (int)gridDim.x - 1) {
        uint32_t total = prefix + (uint32_t)tot;
        rv.header[0] = (int32_t)total;
        rv.header[1] = total > (uint32_t)cap ? 1 : 0;
      }
    }
  }
  __syncthreads();
#pragma unroll
  for (int it = 0; it < PK_ITERS; ++it) {
    int64_t r = (int64_t)tile * PK_TILE + it * PK_THREADS + threadIdx.x;
    uint32_t wbase = s_prefix + (uint32_t)s_cnt[it * 8 + w];
    int64_t r0 = (int64_t)tile * PK_TILE + it * PK_THREADS + w * 32;
    if (lane == 0 && r0 < n) rv.base[r0 >> 5] = wbase;
    if (v[it] > 0.f) {
      uint32_t pos = wbase + __popc(vb[it] & ((1u << lane) - 1u));
      if (pos < (uint32_t)cap) {
        rv.idx[pos] = (int32_t)r;
        rv.aux[pos] = make_float4(v[it], hp[it], hs[it], 0.f);
      }
    }
  }
}

// Pass B of the pack: move the gradient rows of the visible list into the record's planes and clear them at the
// source, so the dense gradient buffer is all-zero again when the fold adds every image's rows in image order.
__global__ void __launch_bounds__(256) k_pack_rows(int64_t cap, RecView rv, GroupsDev gd) {
  int count = min(rv.header[0], (int)cap);
  int warps = (gridDim.x * blockDim.x) >> 5;
  int lane = threadIdx.x & 31;
  for (int j = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; j < count; j += warps) {
    int64_t id = rv.idx[j];
    for (int e = lane; e < gd.total; e += 32) {
      int g = 0;
#pragma unroll
      for (int k = 1; k < GS_EXCHANGE_MAX_GROUPS; ++k) g += (k < gd.count && e >= gd.start[k]);
      int k = e - gd.start[g], dim = gd.dim[g];
      float *src = gd.grad[g] + id * dim + k;
      rv.rows[(int64_t)cap * gd.start[g] + (int64_t)j * dim + k] = *src;
      *src = 0.f;
    }
  }
}

// Fold, statistics: one thread per row, images strictly in order.  All loads from the (possibly remote) records of a
// chunk of images are issued before the first dependent use.
constexpr int FS_CHUNK = 8;
__global__ void __launch_bounds__(256) k_fold_stats(int64_t n, int64_t cap, RecordsDev recs, float split_alpha, float prune_alpha,
                                                    float *__restrict__ prune_cost, float *__restrict__ split_score,
                                                    int16_t *__restrict__ points_in_view, float *__restrict__ visibility,
                                                    float *__restrict__ scene_visible) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int64_t w = i >> 5;
  const unsigned bit = 1u << (i & 31), below = bit - 1u;
  bool loaded = false;
  float split = 0.f, prune = 0.f;
  for (int j0 = 0; j0 < recs.count; j0 += FS_CHUNK) {
    unsigned iv[FS_CHUNK], vv[FS_CHUNK], bb[FS_CHUNK];
    float4 ax[FS_CHUNK];
#pragma unroll
    for (int c = 0; c < FS_CHUNK; ++c) {
      iv[c] = vv[c] = bb[c] = 0u;
      if (j0 + c < recs.count) {
        RecView rv = rec_view(const_cast<void *>(recs.rec[j0 + c]), n, cap);
        iv[c] = rv.inview[w]; vv[c] = rv.visbits[w]; bb[c] = rv.base[w];
      }
    }
#pragma unroll
    for (int c = 0; c < FS_CHUNK; ++c) {
      ax[c] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (vv[c] & bit) {
        uint32_t pos = bb[c] + __popc(vv[c] & below);
        if (pos < (uint32_t)cap) ax[c] = rec_view(const_cast<void *>(recs.rec[j0 + c]), n, cap).aux[pos];
      }
    }
#pragma unroll
    for (int c = 0; c < FS_CHUNK; ++c) {
      if (!(iv[c] & bit)) continue;
      if (!loaded) { split = split_score[i]; prune = prune_cost[i]; loaded = true; }
      if (ax[c].x > 0.f) {
        points_in_view[i] += 1;
        visibility[i] += ax[c].x;
        if (scene_visible) scene_visible[i] += ax[c].x;
      }
      split = exp_lerp_x(split_alpha, split, ax[c].z);
      prune = exp_lerp_x(prune_alpha, prune, ax[c].y);
    }
  }
  if (loaded) { split_score[i] = split; prune_cost[i] = prune; }
}

// Fold, gradient rows: a CTA owns a contiguous range of Gaussian rows (a multiple of 32, so the record positions of
// the range come straight from the rank plane) and adds image after image into the dense gradient planes.
__global__ void __launch_bounds__(256) k_fold_rows(int64_t n, int64_t cap, int64_t rows_per_cta, RecordsDev recs, GroupsDev gd) {
  __shared__ int s_lo[GS_EXCHANGE_MAX_IMAGES], s_hi[GS_EXCHANGE_MAX_IMAGES];
  int64_t a = (int64_t)blockIdx.x * rows_per_cta, b = min(a + rows_per_cta, n);
  if (a >= n) return;
  for (int j = threadIdx.x; j < recs.count; j += blockDim.x) {
    RecView rv = rec_view(const_cast<void *>(recs.rec[j]), n, cap);
    int count = min(rv.header[0], (int)cap);
    int lo = (int)rv.base[a >> 5];
    int hi = (b >= n) ? count : (int)rv.base[b >> 5];
    s_lo[j] = min(lo, count);
    s_hi[j] = min(hi, count);
  }
  __syncthreads();
  for (int j = 0; j < recs.count; ++j) {
    RecView rv = rec_view(const_cast<void *>(recs.rec[j]), n, cap);
    int lo = s_lo[j], m = s_hi[j] - lo;
    for (int t = threadIdx.x; t < m * gd.total; t += blockDim.x) {
      int r = lo + t / gd.total, e = t % gd.total;
      int g = 0;
#pragma unroll
      for (int k = 1; k < GS_EXCHANGE_MAX_GROUPS; ++k) g += (k < gd.count && e >= gd.start[k]);
      int k = e - gd.start[g], dim = gd.dim[g];
      float val = rv.rows[(int64_t)cap * gd.start[g] + (int64_t)r * dim + k];
      int64_t id = rv.idx[r];
      gd.grad[g][id * dim + k] += val;
    }
    __syncthreads();  // image j's adds are complete (and visible block-wide) before image j+1 touches the same rows
  }
}

int make_groups(const gs_exchange_groups *h, GroupsDev *gd) {
  if (!h || h->count < 1 || h->count > GS_EXCHANGE_MAX_GROUPS) return -1;
  gd->count = h->count;
  int s = 0;
  for (int g = 0; g < GS_EXCHANGE_MAX_GROUPS; ++g) {
    if (g < h->count) {
      if (h->dim[g] < 1 || !h->grad[g]) return -1;
      gd->dim[g] = h->dim[g]; gd->start[g] = s; gd->grad[g] = h->grad[g];
      s += h->dim[g];
    } else {
      gd->dim[g] = 1; gd->start[g] = 1 << 30; gd->grad[g] = nullptr;
    }
  }
  gd->total = s;
  return 0;
}

}  // namespace

extern "C" size_t gs_exchange_record_bytes(int64_t n, int64_t cap, int row_floats) {
  if (n < 0 || cap < 0 || row_floats < 0) return 0;
  size_t words = (size_t)((n + 31) / 32);
  return 256 + 3 * al256(words * 4) + al256((size_t)cap * 4) + al256((size_t)cap * 16) + al256((size_t)cap * row_floats * 4);
}

extern "C" size_t gs_exchange_pack_workspace_bytes(int64_t n) {
  size_t tiles = (size_t)((n > 0 ? n : 1) + PK_TILE - 1) / PK_TILE;
  return gs_align256((64 + tiles) * 4) + 256;
}

extern "C" int gs_exchange_pack(int64_t n, const float *g2d, const float *vis, int vis_stride, const float *heur,
                                int heur_stride, float *max_scale_px, float *clear_acc, int clear_stride, float *clear_vis,
                                const gs_exchange_groups *groups, void *record, int64_t cap, void *workspace,
                                size_t workspace_bytes, void *stream) {
  GS_REQUIRE(n > 0 && n < ((int64_t)1 << 30), "n must be in [1, 2^30)");
  GS_REQUIRE(cap > 0 && cap < ((int64_t)1 << 30), "bad capacity");
  GS_REQUIRE(g2d && vis && heur && max_scale_px && record && workspace, "null argument");
  GroupsDev gd;
  GS_REQUIRE(make_groups(groups, &gd) == 0, "bad parameter groups");
  if (workspace_bytes < gs_exchange_pack_workspace_bytes(n)) { gs_set_error("gs_exchange_pack: workspace too small"); return GS_ERR_WORKSPACE; }
  cudaStream_t st = (cudaStream_t)stream;
  uintptr_t a = ((uintptr_t)workspace + 255) & ~(uintptr_t)255;
  int tiles = gs_div_up(n, PK_TILE);
  GS_CUDA(cudaMemsetAsync((void *)a, 0, (64 + (size_t)tiles) * 4, st));
  uint32_t *ticket = (uint32_t *)a, *status = ticket + 64;
  RecView rv = rec_view(record, n, cap);
  k_pack_stats<<<tiles, PK_THREADS, 0, st>>>(n, cap, g2d, vis, vis_stride, heur, heur_stride, max_scale_px, clear_acc,
                                             clear_stride, clear_vis, rv, status, ticket);
  GS_LAUNCH_CHECK();
  k_pack_rows<<<148 * 4, 256, 0, st>>>(cap, rv, gd);
  GS_LAUNCH_CHECK();
  return GS_OK;
}

extern "C" int gs_exchange_fold(int64_t n, int n_images, const void *const *records, int64_t cap,
                                const gs_exchange_groups *groups, float split_alpha, float prune_alpha, float *prune_cost,
                                float *split_score, int16_t *points_in_view, float *visibility, float *scene_visible,
                                void *stream) {
  GS_REQUIRE(n > 0 && n < ((int64_t)1 << 30), "n must be in [1, 2^30)");
  GS_REQUIRE(n_images >= 1 && n_images <= GS_EXCHANGE_MAX_IMAGES, "n_images must be 1..GS_EXCHANGE_MAX_IMAGES");
  GS_REQUIRE(records && prune_cost && split_score && points_in_view && visibility, "null argument");
  GroupsDev gd;
  GS_REQUIRE(make_groups(groups, &gd) == 0, "bad parameter groups");
  RecordsDev rd;
  rd.count = n_images;
  for (int j = 0; j < GS_EXCHANGE_MAX_IMAGES; ++j) {
    rd.rec[j] = j < n_images ? records[j] : nullptr;
    GS_REQUIRE(j >= n_images || records[j], "null record");
  }
  cudaStream_t st = (cudaStream_t)stream;
  k_fold_stats<<<gs_div_up(n, 256), 256, 0, st>>>(n, cap, rd, split_alpha, prune_alpha, prune_cost, split_score, points_in_view,
                                                  visibility, scene_visible);
  GS_LAUNCH_CHECK();
  int ctas = 148 * 8;
  int64_t rows_per_cta = ((n + ctas - 1) / ctas + 31) / 32 * 32;
  int grid = gs_div_up(n, rows_per_cta);
  k_fold_rows<<<grid, 256, 0, st>>>(n, cap, rows_per_cta, rd, gd);
  GS_LAUNCH_CHECK();
  return GS_OK;
}
