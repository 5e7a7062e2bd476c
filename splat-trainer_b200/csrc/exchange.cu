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
//   rows     float[cap·P]        gradient rows, row-major: D gradient floats (group g's dim floats start at Σ dims before
//                                g), then the row's visibility, padded to P = 16-float multiple: a row is one or two
//                                coalesced (NVLink-friendly) warp loads
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

// floats per record row: the D gradient floats + the row's visibility (element D), rounded up to 16 floats
__host__ __device__ inline int row_pad(int d) { return (d + 1 + 15) / 16 * 16; }

struct GroupsDev {
  int count, total;
  int dim[GS_EXCHANGE_MAX_GROUPS];
  int start[GS_EXCHANGE_MAX_GROUPS];  // first flattened element of the group
  float *grad[GS_EXCHANGE_MAX_GROUPS];
  float *vis;  // fold only: the scene's `visible` plane, fed by element `total` of every record row (nullable)
};

// group of flattened row element e -> (offset inside the group, dim, dense plane).  Static indices only: a dynamically
// indexed kernel-parameter array becomes a divergent constant-bank load per lane (measured: 430 stall cycles/instr).
__device__ __forceinline__ void group_of(const GroupsDev &gd, int e, int &k_in, int &dim, float *&plane) {
  int start = 0;
  dim = gd.dim[0]; plane = gd.grad[0];
#pragma unroll
  for (int k = 1; k < GS_EXCHANGE_MAX_GROUPS; ++k) {
    if (k < gd.count && e >= gd.start[k]) { dim = gd.dim[k]; plane = gd.grad[k]; start = gd.start[k]; }
  }
  k_in = e - start;
}

struct PeersDev {
  int count;
  char *base[GS_EXCHANGE_MAX_PEERS];
  size_t plane_offset[GS_EXCHANGE_MAX_GROUPS];
};

struct RecordsDev {
  int count;
  const void *rec[GS_EXCHANGE_MAX_IMAGES];
};

constexpr int PK_THREADS = 256;
constexpr int PK_ITERS = 4;
constexpr int PK_TILE = PK_THREADS * PK_ITERS;  // rows per CTA; a warp covers one 32-row word per iteration

// Pack pass 1: one sweep over the image's rows, no inter-CTA dependency: in-view / visible bit planes, the per-word
// visible count (left in the `base` plane for pass 2), and the rank's own max_scale_px.  Every row's loads are issued
// before the first dependent use (the sweep is latency bound otherwise).
__global__ void __launch_bounds__(PK_THREADS) k_pack_sweep(int64_t n, const float *__restrict__ g2d, const float *__restrict__ vis,
                                                           int vis_stride, float *__restrict__ max_scale_px, RecView rv) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  float2 ca_cb[PK_ITERS], cc_op[PK_ITERS];
  float v[PK_ITERS], ms[PK_ITERS];
#pragma unroll
  for (int it = 0; it < PK_ITERS; ++it) {
    int64_t r = (int64_t)blockIdx.x * PK_TILE + it * PK_THREADS + threadIdx.x;
    ca_cb[it] = make_float2(1.f, 0.f); cc_op[it] = make_float2(1.f, 0.f); v[it] = 0.f; ms[it] = 0.f;
    if (r < n) {
      const float2 *g = reinterpret_cast<const float2 *>(g2d + GS_G2D_STRIDE * r);
      ca_cb[it] = g[1]; cc_op[it] = g[2];
      v[it] = vis[(int64_t)vis_stride * r];
      ms[it] = max_scale_px[r];
    }
  }
#pragma unroll
  for (int it = 0; it < PK_ITERS; ++it) {
    int64_t r = (int64_t)blockIdx.x * PK_TILE + it * PK_THREADS + threadIdx.x;
    bool inview = cc_op[it].y > 0.f;  // rows past n carry opacity 0
    if (inview) {
      float ca = ca_cb[it].x, cb = ca_cb[it].y, cc = cc_op[it].x;
      float detc = ca * cc - cb * cb;
      float cxx = cc / detc, cyy = ca / detc, cxy = -cb / detc;
      float mid = 0.5f * (cxx + cyy);
      float disc = mid * mid - (cxx * cyy - cxy * cxy);
      float lam = mid + sqrtf(fmaxf(disc, 0.f));
      float scale_px = sqrtf(fmaxf(lam, 0.f));
      if (scale_px > ms[it]) max_scale_px[r] = scale_px;
    }
    unsigned ib = __ballot_sync(GS_FULL, inview);
    unsigned vb = __ballot_sync(GS_FULL, inview && v[it] > 0.f);
    int64_t r0 = (int64_t)blockIdx.x * PK_TILE + it * PK_THREADS + w * 32;  // first row of this warp's word
    if (lane == 0 && r0 < n) {
      rv.inview[r0 >> 5] = ib; rv.visbits[r0 >> 5] = vb; rv.base[r0 >> 5] = (uint32_t)__popc(vb);
    }
  }
}

// Pack pass 2: exclusive scan of the word counts in place (chained scan over a few dozen CTAs of 4096 words each — the
// look-back chain is short because a CTA covers 131072 rows) + the record header.
constexpr int PB_THREADS = 256;
constexpr int PB_ITEMS = 16;
constexpr int PB_TILE = PB_THREADS * PB_ITEMS;
__global__ void __launch_bounds__(PB_THREADS) k_pack_base(int64_t words, int64_t cap, RecView rv, uint32_t *__restrict__ status,
                                                          uint32_t *__restrict__ ticket) {
  __shared__ uint32_t s_warp[PB_THREADS / 32];
  __shared__ uint32_t s_prefix;
  __shared__ int s_tile;
  if (threadIdx.x == 0) s_tile = (int)atomicAdd(ticket, 1u);
  __syncthreads();
  const int tile = s_tile;  // ticket order: every predecessor of a running tile has already started
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  int64_t first = (int64_t)tile * PB_TILE + (int64_t)threadIdx.x * PB_ITEMS;
  uint32_t c[PB_ITEMS];
  uint32_t sum = 0;
#pragma unroll
  for (int k = 0; k < PB_ITEMS; k += 4) {
    uint4 q = make_uint4(0u, 0u, 0u, 0u);
    if (first + k + 3 < words) q = *reinterpret_cast<const uint4 *>(rv.base + first + k);
    else {
      if (first + k < words) q.x = rv.base[first + k];
      if (first + k + 1 < words) q.y = rv.base[first + k + 1];
      if (first + k + 2 < words) q.z = rv.base[first + k + 2];
    }
    c[k] = q.x; c[k + 1] = q.y; c[k + 2] = q.z; c[k + 3] = q.w;
    sum += q.x + q.y + q.z + q.w;
  }
  uint32_t inc = sum;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    uint32_t t = __shfl_up_sync(GS_FULL, inc, o);
    if (lane >= o) inc += t;
  }
  if (lane == 31) s_warp[w] = inc;
  __syncthreads();
  uint32_t wo = 0, tot = 0;
#pragma unroll
  for (int k = 0; k < PB_THREADS / 32; ++k) {
    uint32_t x = s_warp[k];
    if (k < w) wo += x;
    tot += x;
  }
  if (threadIdx.x == 0) gs_st_status(status + tile, tot | (tile == 0 ? GS_LB_INCL : GS_LB_LOCAL));
  if (threadIdx.x < 32) {
    uint32_t prefix = gs_lookback_warp(status, tile);
    if (lane == 0) {
      if (tile > 0) gs_st_status(status + tile, (prefix + tot) | GS_LB_INCL);
      s_prefix = prefix;
      if (tile == (int)gridDim.x - 1) {
        uint32_t total = prefix + tot;
        rv.header[0] = (int32_t)total;
        rv.header[1] = total > (uint32_t)cap ? 1 : 0;
      }
    }
  }
  __syncthreads();
  uint32_t run = s_prefix + wo + inc - sum;
#pragma unroll
  for (int k = 0; k < PB_ITEMS; ++k) {
    if (first + k < words) rv.base[first + k] = run;
    run += c[k];
  }
}

// Pack pass 3: the ascending list of visible rows with their (visibility, prune, split) triple, straight from the bit
// and rank planes; clears the consumed raster accumulator rows like gs_accumulate_stats does.  A thread scans 4 rows.
__global__ void __launch_bounds__(256) k_pack_list(int64_t n, int64_t cap, const float *__restrict__ vis, int vis_stride,
                                                   const float *__restrict__ heur, int heur_stride, float *__restrict__ clear_acc,
                                                   int clear_stride, float *__restrict__ clear_vis, RecView rv, GroupsDev gd) {
  int64_t r0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (r0 >= n) return;
  unsigned vb = rv.visbits[r0 >> 5];
  unsigned mine = (vb >> (r0 & 31)) & 0xfu;
  if (!mine) return;
  uint32_t base = rv.base[r0 >> 5];
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    if (!(mine & (1u << q))) continue;
    int64_t r = r0 + q;
    unsigned bit = 1u << (r & 31);
    uint32_t pos = base + __popc(vb & (bit - 1u));
    float v = vis[(int64_t)vis_stride * r];
    float hp = heur[(int64_t)heur_stride * r], hs = heur[(int64_t)heur_stride * r + 1];
    if (pos < (uint32_t)cap) {
      rv.idx[pos] = (int32_t)r;
      rv.aux[pos] = make_float4(v, hp, hs, 0.f);
    } else {
      // the record is full (header[1] reports it; the Trainer regrows the record): this row is dropped from the step.
      // Clear its gradient at the source so it cannot leak into a later step's fold.
      for (int e = 0; e < gd.total; ++e) {
        int k, dim; float *plane;
        group_of(gd, e, k, dim, plane);
        plane[r * dim + k] = 0.f;
      }
    }
    if (clear_acc) {
      float *a = clear_acc + (int64_t)clear_stride * r;
      for (int k = 0; k < clear_stride; ++k) a[k] = 0.f;
    }
    if (clear_vis) clear_vis[r] = 0.f;
  }
}

// Pack pass 4: move the gradient rows of the visible list into the record (row-major) and clear them at the source, so
// the dense gradient buffer is all-zero again when the fold adds every image's rows in image order.  A warp per row.
__global__ void __launch_bounds__(256) k_pack_rows(int64_t cap, RecView rv, GroupsDev gd) {
  int count = min(rv.header[0], (int)cap);
  int warps = (gridDim.x * blockDim.x) >> 5;
  int lane = threadIdx.x & 31;
  const int pad = row_pad(gd.total);
  for (int j = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; j < count; j += warps) {
    int64_t id = rv.idx[j];
    for (int e = lane; e < pad; e += 32) {
      float val = 0.f;
      float *src = nullptr;
      if (e < gd.total) {
        int k, dim; float *plane;
        group_of(gd, e, k, dim, plane);
        src = plane + id * dim + k;
        val = __ldcg(src);
      } else if (e == gd.total) {
        val = rv.aux[j].x;  // the row's visibility travels with its gradients (the fold sums it into `visible`)
      }
      rv.rows[(int64_t)j * pad + e] = val;
      // clear AFTER the dependent record store: a store issued right behind the load of the same address stalled the
      // L1 miss path (measured 220 us instead of 15 us for this kernel)
      if (src) __stcg(src, 0.f);
    }
  }
}

// Fold, statistics: a thread owns 4 consecutive rows (one float4 of each statistic, all in one 32-row word), images
// strictly in order.  The rows' own statistics and all loads from the (possibly remote) records of a chunk of images
// are issued before the first dependent use.
constexpr int FS_CHUNK = 4;
__global__ void __launch_bounds__(256) k_fold_stats(int64_t n, int64_t cap, RecordsDev recs, float split_alpha, float prune_alpha,
                                                    float *__restrict__ prune_cost, float *__restrict__ split_score,
                                                    int16_t *__restrict__ points_in_view, float *__restrict__ visibility,
                                                    int64_t row_begin, int64_t row_end) {
  int64_t i0 = row_begin + ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;  // row_begin is a multiple of 4
  if (i0 >= row_end) return;
  const int64_t w = i0 >> 5;
  const int sh = (int)(i0 & 31);
  const size_t words = (size_t)((n + 31) / 32);
  const size_t plane = al256(words * 4);
  const size_t aux_off = 256 + 3 * plane + al256((size_t)cap * 4);
  const bool full = i0 + 3 < row_end;
  float split[4], prune[4];
  if (full) {
    float4 a = *reinterpret_cast<const float4 *>(split_score + i0), b = *reinterpret_cast<const float4 *>(prune_cost + i0);
    split[0] = a.x; split[1] = a.y; split[2] = a.z; split[3] = a.w;
    prune[0] = b.x; prune[1] = b.y; prune[2] = b.z; prune[3] = b.w;
  } else {
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      split[q] = (i0 + q < row_end) ? split_score[i0 + q] : 0.f;
      prune[q] = (i0 + q < row_end) ? prune_cost[i0 + q] : 0.f;
    }
  }
  unsigned any = 0u;
  for (int j0 = 0; j0 < recs.count; j0 += FS_CHUNK) {
    unsigned iv[FS_CHUNK], vv[FS_CHUNK], bb[FS_CHUNK];
#pragma unroll
    for (int c = 0; c < FS_CHUNK; ++c) {
      iv[c] = vv[c] = bb[c] = 0u;
      if (j0 + c < recs.count) {
        const char *base = (const char *)recs.rec[j0 + c] + 256;
        iv[c] = *(const uint32_t *)(base + 4 * w);
        vv[c] = *(const uint32_t *)(base + plane + 4 * w);
        bb[c] = *(const uint32_t *)(base + 2 * plane + 4 * w);
      }
    }
#pragma unroll
    for (int c = 0; c < FS_CHUNK; ++c) {
      unsigned in4 = (iv[c] >> sh) & 0xfu;
      if (!in4) continue;
      any |= in4;
      unsigned vis4 = (vv[c] >> sh) & 0xfu;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        if (!(in4 & (1u << q))) continue;
        float hp = 0.f, hs = 0.f;
        if (vis4 & (1u << q)) {
          uint32_t pos = bb[c] + __popc(vv[c] & ((1u << (sh + q)) - 1u));
          if (pos < (uint32_t)cap) {
            float4 ax = *(const float4 *)((const char *)recs.rec[j0 + c] + aux_off + 16 * (size_t)pos);
            if (ax.x > 0.f) {
              points_in_view[i0 + q] += 1;
              visibility[i0 + q] += ax.x;
            }
            hp = ax.y; hs = ax.z;
          }
        }
        split[q] = gs_exp_lerp(split_alpha, split[q], hs);
        prune[q] = gs_exp_lerp(prune_alpha, prune[q], hp);
      }
    }
  }
  if (!any) return;
  if (full) {
    *reinterpret_cast<float4 *>(split_score + i0) = make_float4(split[0], split[1], split[2], split[3]);
    *reinterpret_cast<float4 *>(prune_cost + i0) = make_float4(prune[0], prune[1], prune[2], prune[3]);
  } else {
#pragma unroll
    for (int q = 0; q < 4; ++q)
      if (i0 + q < row_end) { split_score[i0 + q] = split[q]; prune_cost[i0 + q] = prune[q]; }
  }
}

// Fold, gradient rows: a warp owns one 32-row word.  Lane j first fetches image j's visible word and rank base (ONE
// round trip for all images), then for every row some image blended each lane keeps two elements of the dense
// gradient row in registers and adds the images' record rows in image order — no barrier, no atomics.  A record row
// is one coalesced 8-B-per-lane warp load (256 B for SH3), the unit NVLink moves efficiently.
__global__ void __launch_bounds__(256, 8) k_fold_rows(int64_t n, int64_t cap, RecordsDev recs, GroupsDev gd, int64_t word_begin,
                                                      int64_t word_end) {
  const int lane = threadIdx.x & 31;
  const int64_t words = (n + 31) / 32;
  const size_t plane = al256((size_t)words * 4);
  const size_t rows_off = 256 + 3 * plane + al256((size_t)cap * 4) + al256((size_t)cap * 16);
  const int pad = row_pad(gd.total);
  const int64_t warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  // this lane's two elements of the first 64 floats of a row (all of it for SH degree <= 3): group plane, dim, offset
  int hk[2], hdim[2];
  float *hplane[2];
#pragma unroll
  for (int q = 0; q < 2; ++q) {
    const int e = 2 * lane + q;
    hk[q] = 0; hdim[q] = 0; hplane[q] = nullptr;
    if (e < gd.total) group_of(gd, e, hk[q], hdim[q], hplane[q]);
    else if (e == gd.total) { hdim[q] = 1; hplane[q] = gd.vis; }
  }
  for (int64_t w = word_begin + (((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5); w < word_end; w += warps) {
    unsigned vb0 = 0u, bs0 = 0u, vb1 = 0u, bs1 = 0u;  // images lane and lane + 32
    if (lane < recs.count) {
      const char *rec = (const char *)recs.rec[lane] + 256;
      vb0 = *(const uint32_t *)(rec + plane + 4 * w); bs0 = *(const uint32_t *)(rec + 2 * plane + 4 * w);
    }
    if (lane + 32 < recs.count) {
      const char *rec = (const char *)recs.rec[lane + 32] + 256;
      vb1 = *(const uint32_t *)(rec + plane + 4 * w); bs1 = *(const uint32_t *)(rec + 2 * plane + 4 * w);
    }
    unsigned any = __reduce_or_sync(GS_FULL, vb0 | vb1);
    while (any) {
      const int b = __ffs(any) - 1;
      any &= any - 1;
      const unsigned bit = 1u << b;
      const int64_t id = w * 32 + b;
      const unsigned m0 = __ballot_sync(GS_FULL, (vb0 & bit) != 0u), m1 = __ballot_sync(GS_FULL, (vb1 & bit) != 0u);
      const unsigned p0 = bs0 + __popc(vb0 & (bit - 1u)), p1 = bs1 + __popc(vb1 & (bit - 1u));  // row position in my images
      for (int e0 = 0; e0 < pad; e0 += 64) {  // warp-uniform trip count: the shuffles below need every lane
        const int e = e0 + 2 * lane;
        const bool act = e < pad;
        // the two elements e, e+1 of the row: their group planes in the dense gradient buffer
        float *dst[2];
#pragma unroll
        for (int q = 0; q < 2; ++q) {
          if (e0 == 0) {
            dst[q] = hplane[q] ? hplane[q] + id * hdim[q] + hk[q] : nullptr;
          } else {
            dst[q] = nullptr;
            if (e + q < gd.total) {
              int k, dim; float *plane;
              group_of(gd, e + q, k, dim, plane);
              dst[q] = plane + id * dim + k;
            } else if (e + q == gd.total && gd.vis) {
              dst[q] = gd.vis + id;
            }
          }
        }
        float a0 = dst[0] ? *dst[0] : 0.f, a1 = dst[1] ? *dst[1] : 0.f;
        // the row's images in batches of 4: all record loads of a batch are in flight together (one NVLink round
        // trip), the adds then run in image order
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          unsigned m = half ? m1 : m0;
          const unsigned pmine = half ? p1 : p0;
          while (m) {
            float2 v[4];
#pragma unroll
            for (int c = 0; c < 4; ++c) {
              v[c] = make_float2(0.f, 0.f);
              const bool have = m != 0u;
              const int j = have ? __ffs(m) - 1 : 0;
              m &= m - 1;  // (0 & anything stays 0)
              const unsigned pos = __shfl_sync(GS_FULL, pmine, j);
              if (have && act && pos < (unsigned)cap)
                v[c] = *(const float2 *)((const char *)recs.rec[j + 32 * half] + rows_off + 4 * ((size_t)pos * pad + e));
            }
#pragma unroll
            for (int c = 0; c < 4; ++c) { a0 += v[c].x; a1 += v[c].y; }
          }
        }
        if (dst[0]) *dst[0] = a0;
        if (dst[1]) *dst[1] = a1;
      }
    }
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
  gd->vis = nullptr;
  return 0;
}

}  // namespace

extern "C" size_t gs_exchange_record_bytes(int64_t n, int64_t cap, int row_floats) {
  if (n < 0 || cap < 0 || row_floats < 0) return 0;
  size_t words = (size_t)((n + 31) / 32);
  return 256 + 3 * al256(words * 4) + al256((size_t)cap * 4) + al256((size_t)cap * 16) + al256((size_t)cap * row_pad(row_floats) * 4);
}

extern "C" size_t gs_exchange_pack_workspace_bytes(int64_t n) {
  size_t words = (size_t)(((n > 0 ? n : 1) + 31) / 32);
  size_t tiles = (words + PB_TILE - 1) / PB_TILE;
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
  int64_t words = (n + 31) / 32;
  int tiles = gs_div_up(words, PB_TILE);
  GS_CUDA(cudaMemsetAsync((void *)a, 0, (64 + (size_t)tiles) * 4, st));
  uint32_t *ticket = (uint32_t *)a, *status = ticket + 64;
  RecView rv = rec_view(record, n, cap);
  k_pack_sweep<<<gs_div_up(n, PK_TILE), PK_THREADS, 0, st>>>(n, g2d, vis, vis_stride, max_scale_px, rv);
  GS_LAUNCH_CHECK();
  k_pack_base<<<tiles, PB_THREADS, 0, st>>>(words, cap, rv, status, ticket);
  GS_LAUNCH_CHECK();
  k_pack_list<<<gs_div_up((n + 3) / 4, 256), 256, 0, st>>>(n, cap, vis, vis_stride, heur, heur_stride, clear_acc, clear_stride, clear_vis, rv, gd);
  GS_LAUNCH_CHECK();
  {  // one warp per record row where the machine has room for them (the count is only known on the device)
    int64_t want = (cap + 7) / 8;
    k_pack_rows<<<(int)(want < 148 * 32 ? want : 148 * 32), 256, 0, st>>>(cap, rv, gd);
  }
  GS_LAUNCH_CHECK();
  return GS_OK;
}

extern "C" int gs_exchange_fold(int64_t n, int n_images, const void *const *records, int64_t cap,
                                const gs_exchange_groups *groups, float split_alpha, float prune_alpha, float *prune_cost,
                                float *split_score, int16_t *points_in_view, float *visibility, int64_t stats_begin,
                                int64_t stats_end, int64_t grad_begin, int64_t grad_end, float *scene_visible, void *stream) {
  GS_REQUIRE(n > 0 && n < ((int64_t)1 << 30), "n must be in [1, 2^30)");
  GS_REQUIRE(n_images >= 1 && n_images <= GS_EXCHANGE_MAX_IMAGES, "n_images must be 1..GS_EXCHANGE_MAX_IMAGES");
  GS_REQUIRE(records && prune_cost && split_score && points_in_view && visibility, "null argument");
  GS_REQUIRE(stats_begin >= 0 && stats_begin <= stats_end && stats_end <= n, "bad statistics row range");
  GS_REQUIRE(stats_begin % 4 == 0 && (stats_end % 4 == 0 || stats_end == n), "statistics row range must be 4-row aligned");
  GS_REQUIRE(grad_begin >= 0 && grad_begin <= grad_end && grad_end <= n, "bad gradient row range");
  GS_REQUIRE(grad_begin % 32 == 0 && (grad_end % 32 == 0 || grad_end == n), "gradient row range must be 32-row aligned");
  GroupsDev gd;
  GS_REQUIRE(make_groups(groups, &gd) == 0, "bad parameter groups");
  gd.vis = scene_visible;
  RecordsDev rd;
  rd.count = n_images;
  for (int j = 0; j < GS_EXCHANGE_MAX_IMAGES; ++j) {
    rd.rec[j] = j < n_images ? records[j] : nullptr;
    GS_REQUIRE(j >= n_images || records[j], "null record");
  }
  cudaStream_t st = (cudaStream_t)stream;
  if (stats_end > stats_begin) {
    int64_t threads = (stats_end - stats_begin + 3) / 4;
    k_fold_stats<<<gs_div_up(threads, 256), 256, 0, st>>>(n, cap, rd, split_alpha, prune_alpha, prune_cost, split_score,
                                                          points_in_view, visibility, stats_begin, stats_end);
    GS_LAUNCH_CHECK();
  }
  if (grad_end > grad_begin) {
    int64_t wb = grad_begin / 32, we = (grad_end + 31) / 32, want = (we - wb + 7) / 8;
    k_fold_rows<<<(int)(want < 148 * 64 ? want : 148 * 64), 256, 0, st>>>(n, cap, rd, gd, wb, we);
    GS_LAUNCH_CHECK();
  }
  return GS_OK;
}

// Row-sharded optimiser: the rank that owns a row slice has stepped it; the rows it touched are
// written into every peer's copy of the parameter planes over NVLink.  One grid row per parameter group: thread t of a
// group handles element t % dim of the (t / dim)-th touched row, so runs of neighbouring rows leave as full sectors.
__global__ void __launch_bounds__(256) k_publish_rows(const int32_t *__restrict__ indexes, const int32_t *__restrict__ count,
                                                      int64_t row_begin, GroupsDev gd, PeersDev peers) {
  const int g = blockIdx.y;
  if (g >= gd.count) return;
  const int dim = gd.dim[g];
  const float *src = gd.grad[g];  // (here: the PARAMETER plane of the group)
  const int64_t total = (int64_t)count[0] * dim;
  const size_t plane_off = peers.plane_offset[g];
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
    const int64_t i = t / dim;
    const int k = (int)(t - i * dim);
    const int64_t at = (row_begin + indexes[i]) * dim + k;
    const float v = src[at];
#pragma unroll 1
    for (int q = 0; q < peers.count; ++q) reinterpret_cast<float *>(peers.base[q] + plane_off)[at] = v;
  }
}

extern "C" int gs_exchange_publish(const int32_t *indexes, const int32_t *count, int64_t max_rows, int64_t row_begin,
                                   const gs_exchange_groups *param_groups, const void *local_base, int n_peers,
                                   void *const *peer_bases, void *stream) {
  GS_REQUIRE(indexes && count && param_groups && local_base && peer_bases, "null argument");
  GS_REQUIRE(n_peers >= 0 && n_peers <= GS_EXCHANGE_MAX_PEERS, "n_peers must be 0..GS_EXCHANGE_MAX_PEERS");
  GS_REQUIRE(max_rows >= 0 && row_begin >= 0, "bad row range");
  if (n_peers == 0 || max_rows == 0) return GS_OK;
  GroupsDev gd;
  GS_REQUIRE(make_groups(param_groups, &gd) == 0, "bad parameter groups");
  PeersDev pd;
  pd.count = n_peers;
  for (int q = 0; q < GS_EXCHANGE_MAX_PEERS; ++q) {
    pd.base[q] = q < n_peers ? (char *)peer_bases[q] : nullptr;
    GS_REQUIRE(q >= n_peers || peer_bases[q], "null peer base");
  }
  for (int g = 0; g < GS_EXCHANGE_MAX_GROUPS; ++g) {
    pd.plane_offset[g] = 0;
    if (g < gd.count) {
      ptrdiff_t off = (const char *)gd.grad[g] - (const char *)local_base;
      GS_REQUIRE(off >= 0, "a parameter plane lies before local_base");
      pd.plane_offset[g] = (size_t)off;
    }
  }
  int64_t want = (max_rows * 48 + 255) / 256;
  dim3 grid((unsigned)(want < 148 * 4 ? (want > 0 ? want : 1) : 148 * 4), (unsigned)gd.count);
  k_publish_rows<<<grid, 256, 0, (cudaStream_t)stream>>>(indexes, count, row_begin, gd, pd);
  GS_LAUNCH_CHECK();
  return GS_OK;
}
