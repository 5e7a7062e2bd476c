// raster_rec.cu — the training path's rasteriser: record-staged, asynchronous, persistent (SURVEY §8 a9, a10;
// render_projected at reference scene/mlp_scene.py:418 and its backward via trainer/trainer.py:512).
//
// Same per-pixel arithmetic and decisions as raster.cu (Spec R); what differs is how a tile's splats reach the SM:
//   * every in-view splat has a 48-byte staged record (make_record) written ONCE per image by the projection (or by
//     gs_raster_records), so staging a batch is a pure gather: three 16-byte cp.async (LDGSTS) per splat, no arithmetic;
//   * two shared-memory stages, completion tracked by one mbarrier per stage (cp.async.mbarrier.arrive.noinc): batch
//     k+1 is in flight while batch k is composited; the list index of batch k+2 is prefetched into a register;
//   * persistent CTAs pull tiles from a heaviest-first queue (gs_tile_order), so the longest tile starts first
//     instead of forming the tail of the launch;
//   * every warp builds its own candidate mask from the staged records (no cross-warp publish, one barrier per batch).
#include "raster_common.cuh"

// Pixel ownership inside a tile: warp w covers the 16x8 pixel rows [8w, 8w+8) of the tile; lane (lx = lane & 7, ly = lane >> 3)
// owns the FOUR pixels (lx + 8a, 8w + ly + 4b), a, b in {0,1} — one pixel in each 8x4 quarter of the warp's region — instead
// of a compact 2x2 quad.  Pixel index p = a + 2b is therefore "quarter p" for every lane of the warp: a splat that misses a
// quarter leaves NO lane active in that p-block and the warp skips the block, where compact quads would run all four
// blocks with idle lanes.  The lane still shares two dx and two dy values over its pixels, and a lane row is 8
// consecutive pixels (coalesced image / gradient accesses).
#ifndef GS_QUARTERS
#define GS_QUARTERS 1  // 0 = compact 2x2 quads per lane (the layout of raster.cu), kept for A/B measurements
#endif
constexpr int QX = GS_QUARTERS ? 8 : 1, QY = GS_QUARTERS ? 4 : 1;  // distance between a lane's pixels
constexpr int LSX = GS_QUARTERS ? 1 : 2, LSY = GS_QUARTERS ? 1 : 2;  // distance between neighbouring lanes' first pixels

struct TileGeo {
  int begin, n;
  float fx0, fy0, row0;
  bool in[4];
  int px0, py0;
};

__device__ __forceinline__ TileGeo tile_geo(int tile, const int32_t *__restrict__ ranges, int W, int H, int TX) {
  TileGeo c;
  int tx = tile % TX, ty = tile / TX;
  int lx = threadIdx.x & 7, ly = (threadIdx.x >> 3) & 3, w = threadIdx.x >> 5;
  c.px0 = tx * GS_TILE + LSX * lx;
  c.py0 = ty * GS_TILE + 8 * w + LSY * ly;
  c.fx0 = (float)c.px0 + 0.5f;
  c.fy0 = (float)c.py0 + 0.5f;
  c.row0 = (float)(ty * GS_TILE);
  c.in[0] = c.px0 < W && c.py0 < H;
  c.in[1] = c.px0 + QX < W && c.py0 < H;
  c.in[2] = c.px0 < W && c.py0 + QY < H;
  c.in[3] = c.px0 + QX < W && c.py0 + QY < H;
  c.begin = ranges[2 * tile];
  c.n = ranges[2 * tile + 1] - c.begin;
  return c;
}

__device__ __forceinline__ int64_t geo_pixel(const TileGeo &c, int p, int W) {
  return (int64_t)(c.py0 + QY * (p >> 1)) * W + c.px0 + QX * (p & 1);
}

// two stages of 64 records, as three planes of 16-byte words
struct __align__(16) RecStages {
  float4 w0[2][RT];
  float4 w1[2][RT];
  float4 w2[2][RT];
};

// gather the record of row `id` into slot threadIdx.x of `stage`; every thread arrives (with or without a copy)
__device__ __forceinline__ void issue_batch(const float *__restrict__ records, int id, RecStages *S, int stage, uint32_t bar) {
  if (id >= 0) {
    const float *src = records + (int64_t)GS_REC_FLOATS * id;
    cp_async16(smem_u32(&S->w0[stage][threadIdx.x]), src);
    cp_async16(smem_u32(&S->w1[stage][threadIdx.x]), src + 4);
    cp_async16(smem_u32(&S->w2[stage][threadIdx.x]), src + 8);
    cp_async_arrive(bar);  // arrives once the three copies have landed
  } else {
    mbar_arrive(bar);      // nothing to copy: a plain arrival
  }
}

// this warp's candidates among the `cnt` staged records: bit k set iff record k can reach the warp's 8 pixel rows
__device__ __forceinline__ unsigned long long warp_candidates(const RecStages *S, int stage, int cnt, float row_lo) {
  const int lane = threadIdx.x & 31;
  float my0 = S->w0[stage][lane].y, ey0 = S->w2[stage][lane].w;
  float my1 = S->w0[stage][lane + 32].y, ey1 = S->w2[stage][lane + 32].w;
  // the warp covers pixel-centre rows row_lo + 0.5 .. row_lo + 7.5
  bool t0 = lane < cnt && (my0 + ey0 >= row_lo + 0.5f) && (my0 - ey0 <= row_lo + 7.5f);
  bool t1 = lane + 32 < cnt && (my1 + ey1 >= row_lo + 0.5f) && (my1 - ey1 <= row_lo + 7.5f);
  return (unsigned long long)__ballot_sync(GS_FULL, t0) | ((unsigned long long)__ballot_sync(GS_FULL, t1) << 32);
}

// ---- staged records of a whole image (operator path / colour-MLP scenes; the SH scene gets them from gs_project_fwd) --
__global__ void __launch_bounds__(256) k_raster_records(float thr, int64_t m, const float *__restrict__ g2d,
                                                        const float *__restrict__ feat, float *__restrict__ records) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= m) return;
  const float2 *g = reinterpret_cast<const float2 *>(g2d + (int64_t)GS_G2D_STRIDE * i);
  float2 a = g[0], b = g[1], c = g[2];
  if (!(c.y > 0.f)) return;
  const float *f = feat + 3 * i;
  SplatRec R = make_record(a.x, a.y, b.x, b.y, c.x, c.y, thr, (int)i, f[0], f[1], f[2]);
  float4 *dst = reinterpret_cast<float4 *>(records + (int64_t)GS_REC_FLOATS * i);
  dst[0] = R.w0; dst[1] = R.w1; dst[2] = R.w2;
}

extern "C" int gs_raster_records(const gs_raster_config *h_cfg, int64_t m, const float *g2d, const float *features,
                                 float *records, void *stream) {
  GS_REQUIRE(h_cfg && g2d && features && records, "null argument");
  if (m <= 0) return GS_OK;
  k_raster_records<<<gs_div_up(m, 256), 256, 0, (cudaStream_t)stream>>>(h_cfg->alpha_threshold, m, g2d, features, records);
  GS_LAUNCH_CHECK();
  return GS_OK;
}

// ---- heaviest-first tile queue: single-CTA counting sort of the tiles by their work estimate ----------------------
constexpr int TO_THREADS = 1024;
constexpr int TO_BINS = 1024;

__global__ void __launch_bounds__(TO_THREADS) k_tile_order(int T, const int32_t *__restrict__ work, int stride,
                                                           int32_t *__restrict__ order, uint32_t *__restrict__ ticket) {
  __shared__ int s_hist[TO_BINS];
  __shared__ int s_warp[32];
  __shared__ int s_shift;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  auto work_of = [&](int t) { return stride == 2 ? work[2 * t + 1] - work[2 * t] : work[t]; };
  int mx = 0;
  for (int t = threadIdx.x; t < T; t += TO_THREADS) mx = max(mx, work_of(t));
  mx = __reduce_max_sync(GS_FULL, mx);
  if (lane == 0) s_warp[w] = mx;
  s_hist[threadIdx.x] = 0;
  __syncthreads();
  if (threadIdx.x == 0) {
    int m = 0;
    for (int i = 0; i < 32; ++i) m = max(m, s_warp[i]);
    int shift = 0;
    while ((m >> shift) >= TO_BINS) ++shift;
    s_shift = shift;
    *ticket = 0u;
  }
  __syncthreads();
  const int shift = s_shift;
  for (int t = threadIdx.x; t < T; t += TO_THREADS) atomicAdd(&s_hist[TO_BINS - 1 - (work_of(t) >> shift)], 1);  // bin 0 = heaviest
  __syncthreads();
  // exclusive scan of the 1024 bins (one per thread)
  int v = s_hist[threadIdx.x], inc = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    int t = __shfl_up_sync(GS_FULL, inc, o);
    if (lane >= o) inc += t;
  }
  if (lane == 31) s_warp[w] = inc;
  __syncthreads();
  if (w == 0) {
    int x = s_warp[lane], xi = x;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      int t = __shfl_up_sync(GS_FULL, xi, o);
      if (lane >= o) xi += t;
    }
    s_warp[lane] = xi - x;
  }
  __syncthreads();
  s_hist[threadIdx.x] = s_warp[w] + inc - v;
  __syncthreads();
  for (int t = threadIdx.x; t < T; t += TO_THREADS) {
    int pos = atomicAdd(&s_hist[TO_BINS - 1 - (work_of(t) >> shift)], 1);
    order[pos] = t;
  }
}

extern "C" int gs_tile_order(int tiles, const int32_t *work, int work_stride, int32_t *order, uint32_t *ticket, void *stream) {
  GS_REQUIRE(tiles > 0 && work && order && ticket, "null argument");
  GS_REQUIRE(work_stride == 1 || work_stride == 2, "work_stride must be 1 (tile_work) or 2 (tile_ranges)");
  k_tile_order<<<1, TO_THREADS, 0, (cudaStream_t)stream>>>(tiles, work, work_stride, order, ticket);
  GS_LAUNCH_CHECK();
  return GS_OK;
}

// ---- persistent tile queue ------------------------------------------------------------------------------------------
// order == NULL: one tile per CTA (tile = blockIdx.x).  Otherwise CTAs pull queue positions from *ticket; the position
// of the NEXT tile is fetched by thread 0 at the start of a tile and handed over at the tile's closing barrier.
struct TileQueue {
  const int32_t *order;
  uint32_t *ticket;
  int T;
  int q, nq, flip;
};

__device__ __forceinline__ void queue_begin(TileQueue &Q, const int32_t *order, uint32_t *ticket, int T, int *s_next) {
  Q.order = order; Q.ticket = ticket; Q.T = T; Q.flip = 0; Q.nq = T;
  if (order) {
    if (threadIdx.x == 0) s_next[0] = (int)atomicAdd(ticket, 1u);
    __syncthreads();
    Q.q = s_next[0];
  } else {
    Q.q = blockIdx.x;
  }
}
// call at the start of a tile (thread 0 starts fetching the next position; the atomic's latency hides under the tile)
__device__ __forceinline__ void queue_prefetch(TileQueue &Q) {
  if (Q.order && threadIdx.x == 0) Q.nq = (int)atomicAdd(Q.ticket, 1u);
}
// call right before the tile's closing __syncthreads ...
__device__ __forceinline__ void queue_publish(TileQueue &Q, int *s_next) {
  Q.flip ^= 1;
  if (Q.order && threadIdx.x == 0) s_next[Q.flip] = Q.nq;
}
// ... and right after it
__device__ __forceinline__ void queue_advance(TileQueue &Q, const int *s_next) { Q.q = Q.order ? s_next[Q.flip] : Q.T; }

// ---- forward -----------------------------------------------------------------------------------------------------------
template <bool VIS, bool MEDIAN, bool COUNT>
__global__ void __launch_bounds__(RT) k_raster_fwd_rec(CfgDev cfg, const float *__restrict__ records, const float *__restrict__ depth,
                                                       const int32_t *__restrict__ o2p, const int32_t *__restrict__ ranges,
                                                       int W, int H, int TX, int T, float *__restrict__ image,
                                                       float *__restrict__ final_T, int32_t *__restrict__ n_contrib,
                                                       float *__restrict__ vis, float *__restrict__ median,
                                                       const int32_t *__restrict__ order, uint32_t *__restrict__ ticket,
                                                       int32_t *__restrict__ tile_work, unsigned long long *__restrict__ counters) {
  __shared__ RecStages S;
  __shared__ __align__(8) unsigned long long s_bar[2];
  __shared__ int s_next[2];
  __shared__ int s_wmax[2][RT / 32];
  const int wid = threadIdx.x >> 5;
  const uint32_t bar0 = smem_u32(&s_bar[0]);
  if (threadIdx.x == 0) { mbar_init(bar0, RT); mbar_init(bar0 + 8u, RT); }
  __syncthreads();
  uint32_t phase = 0;  // bit s: parity of the next completion of stage s
  unsigned long long c_iter = 0, c_hit = 0, c_pix = 0, c_rec = 0;
  TileQueue Q;
  queue_begin(Q, order, ticket, T, s_next);
  while (Q.q < T) {
    const int tile = order ? order[Q.q] : Q.q;
    queue_prefetch(Q);
    TileGeo tc = tile_geo(tile, ranges, W, H, TX);
    const float row_lo = tc.row0 + 8.0f * (float)wid;
    float Tr[4] = {1.f, 1.f, 1.f, 1.f};
    float col[4][3];
    float med[4] = {0.f, 0.f, 0.f, 0.f};
    int last[4] = {0, 0, 0, 0};
    bool done[4];
#pragma unroll
    for (int p = 0; p < 4; ++p) {
      done[p] = !tc.in[p];
      col[p][0] = col[p][1] = col[p][2] = 0.f;
    }
    const float fx[2] = {tc.fx0, tc.fx0 + (float)QX}, fy[2] = {tc.fy0, tc.fy0 + (float)QY};
    const int nb = (tc.n + RT - 1) / RT;
    if (nb > 0) {
      int id = ((int)threadIdx.x < tc.n) ? o2p[tc.begin + threadIdx.x] : -1;
      issue_batch(records, id, &S, 0, bar0);
      id = (RT + (int)threadIdx.x < tc.n) ? o2p[tc.begin + RT + threadIdx.x] : -1;
      for (int b = 0; b < nb; ++b) {
        const int s = b & 1;
        if (b + 1 < nb) {  // batch b+1 flies while batch b is composited; the index of batch b+2 is prefetched
          issue_batch(records, id, &S, s ^ 1, bar0 + 8u * (uint32_t)(s ^ 1));
          int j = (b + 2) * RT + threadIdx.x;
          id = (j < tc.n) ? o2p[tc.begin + j] : -1;
        }
        mbar_wait(bar0 + 8u * (uint32_t)s, (phase >> s) & 1u);
        phase ^= 1u << s;
        const int cnt = min(RT, tc.n - b * RT);
        if (COUNT && threadIdx.x == 0) c_rec += cnt;
        unsigned long long cand = warp_candidates(&S, s, cnt, row_lo);
        bool all_done = done[0] && done[1] && done[2] && done[3];
        while (cand) {
          if (COUNT) {  // the counters use warp-wide votes: leave the loop together (finished lanes blend nothing anyway)
            if (!VIS && __all_sync(GS_FULL, all_done)) break;
          } else if (!VIS && all_done) {
            break;
          }
          int k = __ffsll((long long)cand) - 1;
          cand &= cand - 1;
          float4 A = S.w0[s][k], B = S.w1[s][k], F = S.w2[s][k];
          float dx[2] = {fx[0] - A.x, fx[1] - A.x}, dy[2] = {fy[0] - A.y, fy[1] - A.y};
          float ax[2] = {__fmul_rn(__fmul_rn(A.z, dx[0]), dx[0]), __fmul_rn(__fmul_rn(A.z, dx[1]), dx[1])};
          float cy[2] = {__fmul_rn(__fmul_rn(B.x, dy[0]), dy[0]), __fmul_rn(__fmul_rn(B.x, dy[1]), dy[1])};
          float bx[2] = {__fmul_rn(A.w, dx[0]), __fmul_rn(A.w, dx[1])};
          float vsum = 0.f;
          int npix = 0;
#pragma unroll
          for (int p = 0; p < 4; ++p) {
            float pw = __fmaf_rn(bx[p & 1], dy[p >> 1], __fadd_rn(ax[p & 1], cy[p >> 1]));
            if (!done[p] && pw <= 0.f && pw >= B.z) {
              float a = fminf(cfg.amax, B.y * ex2_approx(pw));
              float w = a * Tr[p];
              col[p][0] += w * F.x; col[p][1] += w * F.y; col[p][2] += w * F.z;
              float Tn = Tr[p] * (1.0f - a);
              if (MEDIAN && Tr[p] >= 0.5f && Tn < 0.5f) med[p] = depth[__float_as_int(B.w)];
              Tr[p] = Tn;
              last[p] = b * RT + k + 1;
              if (Tn < cfg.tmin) done[p] = true;
              vsum += w;
              if (COUNT) ++npix;
            }
          }
          if (COUNT) {
            unsigned any = __ballot_sync(GS_FULL, npix > 0);
            int tot = __reduce_add_sync(GS_FULL, npix);
            if ((threadIdx.x & 31) == 0) { c_iter += 1; c_hit += any != 0u; c_pix += tot; }
          }
          if (VIS) {
            if (__any_sync(GS_FULL, vsum > 0.f)) {
              float sum = warp_sum(vsum);
              if ((threadIdx.x & 31) == 0) atomicAdd(vis + __float_as_int(B.w), sum);
            }
          } else {
            all_done = done[0] && done[1] && done[2] && done[3];
          }
        }
        all_done = done[0] && done[1] && done[2] && done[3];
        // the vote also orders this batch's reads of stage s before the copies of batch b+2 into it
        if (__syncthreads_and(all_done)) {
          if (b + 1 < nb) {  // drain the batch already in flight before the stages are reused
            mbar_wait(bar0 + 8u * (uint32_t)(s ^ 1), (phase >> (s ^ 1)) & 1u);
            phase ^= 1u << (s ^ 1);
          }
          break;
        }
      }
    }
    int tmax = 0;
#pragma unroll
    for (int p = 0; p < 4; ++p) {
      tmax = max(tmax, last[p]);
      if (!tc.in[p]) continue;
      int64_t pix = geo_pixel(tc, p, W);
      image[pix * 3] = col[p][0]; image[pix * 3 + 1] = col[p][1]; image[pix * 3 + 2] = col[p][2];
      final_T[pix] = Tr[p];
      n_contrib[pix] = last[p];
      if (MEDIAN) median[pix] = med[p];
    }
    tmax = __reduce_max_sync(GS_FULL, tmax);
    queue_publish(Q, s_next);
    if ((threadIdx.x & 31) == 0) s_wmax[Q.flip][wid] = tmax;
    __syncthreads();
    if (tile_work && threadIdx.x == 0) tile_work[tile] = max(s_wmax[Q.flip][0], s_wmax[Q.flip][1]);
    queue_advance(Q, s_next);
  }
  if (COUNT && (threadIdx.x & 31) == 0) {
    atomicAdd(counters + 0, c_iter); atomicAdd(counters + 1, c_hit); atomicAdd(counters + 2, c_pix);
    if (threadIdx.x == 0) atomicAdd(counters + 3, c_rec);
  }
}

// ---- backward ----------------------------------------------------------------------------------------------------------
// Back-to-front from the tile's largest n_contrib.  Per pixel: transmittance T, dL/dC and ONE scalar R = (colour composited
// behind the current splat)·dL/dC, updated eagerly after each blend (R <- a·q + (1-a)·R with q = f·dL/dC), which is the
// recurrence of raster.cu evaluated one step earlier: same values, two registers per pixel fewer.
__global__ void __launch_bounds__(RT) k_raster_bwd_rec(CfgDev cfg, const float *__restrict__ records, const int32_t *__restrict__ o2p,
                                                       const int32_t *__restrict__ ranges, int W, int H, int TX, int T,
                                                       const float *__restrict__ final_T, const int32_t *__restrict__ n_contrib,
                                                       const float *__restrict__ d_image, float *__restrict__ acc, int acc_stride,
                                                       float *__restrict__ vis_plane, const int32_t *__restrict__ order,
                                                       uint32_t *__restrict__ ticket) {
  __shared__ RecStages S;
  __shared__ __align__(8) unsigned long long s_bar[2];
  __shared__ int s_next[2];
  __shared__ int s_max[2][RT / 32];
  const int wid = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t bar0 = smem_u32(&s_bar[0]);
  if (threadIdx.x == 0) { mbar_init(bar0, RT); mbar_init(bar0 + 8u, RT); }
  __syncthreads();
  uint32_t phase = 0;
  TileQueue Q;
  queue_begin(Q, order, ticket, T, s_next);
  while (Q.q < T) {
    const int tile = order ? order[Q.q] : Q.q;
    queue_prefetch(Q);
    TileGeo tc = tile_geo(tile, ranges, W, H, TX);
    const float row_lo = tc.row0 + 8.0f * (float)wid;
    float Tr[4], R[4], dC[4][3];
    int nc[4];
    int mynmax = 0;
#pragma unroll
    for (int p = 0; p < 4; ++p) {
      Tr[p] = 1.f; nc[p] = 0; R[p] = 0.f;
      dC[p][0] = dC[p][1] = dC[p][2] = 0.f;
      if (tc.in[p]) {
        int64_t pix = geo_pixel(tc, p, W);
        Tr[p] = final_T[pix];
        nc[p] = n_contrib[pix];
        dC[p][0] = d_image[pix * 3]; dC[p][1] = d_image[pix * 3 + 1]; dC[p][2] = d_image[pix * 3 + 2];
      }
      mynmax = max(mynmax, nc[p]);
    }
    const int wmax = __reduce_max_sync(GS_FULL, mynmax);
    if (lane == 0) s_max[Q.flip][wid] = wmax;
    __syncthreads();
    const int top = min(max(s_max[Q.flip][0], s_max[Q.flip][1]), tc.n);
    const float fx[2] = {tc.fx0, tc.fx0 + (float)QX}, fy[2] = {tc.fy0, tc.fy0 + (float)QY};
    const int nb = (top + RT - 1) / RT;
    if (nb > 0) {
      // batch b covers list positions [lo_b, hi_b), hi_b = top - b·RT, lo_b = max(0, hi_b - RT); slot j <-> lo_b + j
      int lo = max(0, top - RT);
      int id = (lo + (int)threadIdx.x < top) ? o2p[tc.begin + lo + threadIdx.x] : -1;
      issue_batch(records, id, &S, 0, bar0);
      {
        int hi1 = top - RT, lo1 = max(0, hi1 - RT);
        id = (nb > 1 && lo1 + (int)threadIdx.x < hi1) ? o2p[tc.begin + lo1 + threadIdx.x] : -1;
      }
      for (int b = 0; b < nb; ++b) {
        const int s = b & 1;
        if (b + 1 < nb) {
          issue_batch(records, id, &S, s ^ 1, bar0 + 8u * (uint32_t)(s ^ 1));
          int hi2 = top - (b + 2) * RT, lo2 = max(0, hi2 - RT);
          id = (b + 2 < nb && lo2 + (int)threadIdx.x < hi2) ? o2p[tc.begin + lo2 + threadIdx.x] : -1;
        }
        mbar_wait(bar0 + 8u * (uint32_t)s, (phase >> s) & 1u);
        phase ^= 1u << s;
        const int hi = top - b * RT;
        lo = max(0, hi - RT);
        const int cnt = hi - lo;
        unsigned long long cand = warp_candidates(&S, s, cnt, row_lo);
        {  // this warp's own pixels stop at wmax: later list positions cannot contribute to it
          int wn = wmax - lo;
          if (wn < RT) cand &= (wn > 0) ? ((1ull << wn) - 1ull) : 0ull;
        }
        while (cand) {
          int k = 63 - __clzll((long long)cand);
          cand ^= 1ull << k;
          const int pos = lo + k;
          float4 A = S.w0[s][k], B = S.w1[s][k], F = S.w2[s][k];
          float dx[2] = {fx[0] - A.x, fx[1] - A.x}, dy[2] = {fy[0] - A.y, fy[1] - A.y};
          float ax[2] = {__fmul_rn(__fmul_rn(A.z, dx[0]), dx[0]), __fmul_rn(__fmul_rn(A.z, dx[1]), dx[1])};
          float cy[2] = {__fmul_rn(__fmul_rn(B.x, dy[0]), dy[0]), __fmul_rn(__fmul_rn(B.x, dy[1]), dy[1])};
          float bx[2] = {__fmul_rn(A.w, dx[0]), __fmul_rn(A.w, dx[1])};
          float v[12];
#pragma unroll
          for (int i = 0; i < 12; ++i) v[i] = 0.f;
          bool touched = false;
          // per-splat and per-row factors of the position gradient, hoisted out of the pixels:
          //   dL/dmean = -dp·(2A'·dx + B'·dy, 2C'·dy + B'·dx) = ndp·(tx, ty),  split_score += |dp|·(|tx| + |ty|)
          const float nopl = -(B.y * LN2);  // dp = dL/dp' = opacity·ln2·(G·dL/dalpha); ndp = -dp
          const float A2 = 2.0f * A.z;
          const float bdy[2] = {A.w * dy[0], A.w * dy[1]}, c2dy[2] = {2.0f * B.x * dy[0], 2.0f * B.x * dy[1]};
#pragma unroll
          for (int p = 0; p < 4; ++p) {
            float pw = __fmaf_rn(bx[p & 1], dy[p >> 1], __fadd_rn(ax[p & 1], cy[p >> 1]));
            if (pos < nc[p] && pw <= 0.f && pw >= B.z) {
              float G = ex2_approx(pw);
              float a = B.y * G;
              bool clamped = a > cfg.amax;
              a = fminf(a, cfg.amax);
              touched = true;
              float Tp = __fdividef(Tr[p], 1.0f - a);
              Tr[p] = Tp;
              float w = a * Tp;
              float q = F.x * dC[p][0] + F.y * dC[p][1] + F.z * dC[p][2];
              v[6] = fmaf(w, dC[p][0], v[6]); v[7] = fmaf(w, dC[p][1], v[7]); v[8] = fmaf(w, dC[p][2], v[8]);
              float dLda = (q - R[p]) * Tp;
              R[p] = fmaf(a, q - R[p], R[p]);  // a·q + (1-a)·R
              v[11] += w;
              v[9] += fabsf(a * dLda);
              if (!clamped) {
                float ddx = dx[p & 1], ddy = dy[p >> 1];
                float GdL = G * dLda;
                float ndp = GdL * nopl;
                float tx = fmaf(A2, ddx, bdy[p >> 1]), ty = fmaf(A.w, ddx, c2dy[p >> 1]);
                v[0] = fmaf(ndp, tx, v[0]); v[1] = fmaf(ndp, ty, v[1]);
                float ndx = ndp * ddx, ndy = ndp * ddy;  // v[2..4] accumulate -dp·(dx², dx·dy, dy²): sign folded in below
                v[2] = fmaf(ndx, ddx, v[2]); v[3] = fmaf(ndx, ddy, v[3]); v[4] = fmaf(ndy, ddy, v[4]);
                v[5] += GdL;
                v[10] = fmaf(fabsf(ndp), fabsf(tx) + fabsf(ty), v[10]);
              }
            }
          }
          if (__any_sync(GS_FULL, touched)) {
            v[2] *= 0.5f * LOG2E; v[3] *= LOG2E; v[4] *= 0.5f * LOG2E;  // d/dconic from d/d(conic pre-scaled by -log2e/2, -log2e)
            int index; bool valid;
            float tot = multi_reduce12(v, lane, index, valid);
            if (valid && tot != 0.f) {
              int rid = __float_as_int(B.w);
              float *dst = (vis_plane && index == 11) ? vis_plane + rid : acc + (int64_t)acc_stride * rid + index;
              atomicAdd(dst, tot);
            }
          }
        }
        __syncthreads();  // reads of stage s done before batch b+2 is copied into it
      }
    }
    queue_publish(Q, s_next);
    __syncthreads();
    queue_advance(Q, s_next);
  }
}

// ---- launches ----------------------------------------------------------------------------------------------------------
static int sm_count() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (n <= 0) n = 148;
  }
  return n;
}

template <typename K>
static int persistent_grid(K kernel, int tiles) {
  int per_sm = 0;
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, RT, 0);
  if (per_sm <= 0) per_sm = 8;
  int g = sm_count() * per_sm;
  return g < tiles ? g : tiles;
}

int gs_raster_fwd_rec(const gs_raster_config *h_cfg, const gs_raster_aux *aux, const float *depth, const int32_t *o2p,
                      const int32_t *ranges, int W, int H, float *image, float *final_T, int32_t *n_contrib, float *visibility,
                      float *median, cudaStream_t st) {
  CfgDev cfg = gs_make_cfg(h_cfg);
  int TX = (W + GS_TILE - 1) / GS_TILE, TY = (H + GS_TILE - 1) / GS_TILE;
  int tiles = TX * TY;
  bool vis = visibility != nullptr, med = median != nullptr, cnt = aux->counters != nullptr;
  if (aux->tile_order && !aux->ticket) { gs_set_error("gs_raster_fwd: tile_order needs a ticket"); return GS_ERR_INVALID; }
#define GS_FWD(V, M, CN)                                                                                                       \
  do {                                                                                                                         \
    auto kern = k_raster_fwd_rec<V, M, CN>;                                                                                    \
    int grid = aux->tile_order ? persistent_grid(kern, tiles) : tiles;                                                         \
    kern<<<grid, RT, 0, st>>>(cfg, aux->records, depth, o2p, ranges, W, H, TX, tiles, image, final_T, n_contrib, visibility,   \
                              median, aux->tile_order, aux->ticket, aux->tile_work, aux->counters);                            \
  } while (0)
  if (cnt) {
    if (vis && med) GS_FWD(true, true, true);
    else if (vis) GS_FWD(true, false, true);
    else if (med) GS_FWD(false, true, true);
    else GS_FWD(false, false, true);
  } else {
    if (vis && med) GS_FWD(true, true, false);
    else if (vis) GS_FWD(true, false, false);
    else if (med) GS_FWD(false, true, false);
    else GS_FWD(false, false, false);
  }
#undef GS_FWD
  GS_LAUNCH_CHECK();
  return GS_OK;
}

int gs_raster_bwd_rec(const gs_raster_config *h_cfg, const gs_raster_aux *aux, const int32_t *o2p, const int32_t *ranges,
                      int W, int H, const float *final_T, const int32_t *n_contrib, const float *d_image, float *acc,
                      int acc_stride, float *vis_plane, cudaStream_t st) {
  CfgDev cfg = gs_make_cfg(h_cfg);
  int TX = (W + GS_TILE - 1) / GS_TILE, TY = (H + GS_TILE - 1) / GS_TILE;
  int tiles = TX * TY;
  if (aux->tile_order && !aux->ticket) { gs_set_error("gs_raster_bwd: tile_order needs a ticket"); return GS_ERR_INVALID; }
  int grid = aux->tile_order ? persistent_grid(k_raster_bwd_rec, tiles) : tiles;
  k_raster_bwd_rec<<<grid, RT, 0, st>>>(cfg, aux->records, o2p, ranges, W, H, TX, tiles, final_T, n_contrib, d_image, acc,
                                        acc_stride, vis_plane, aux->tile_order, aux->ticket);
  GS_LAUNCH_CHECK();
  return GS_OK;
}
