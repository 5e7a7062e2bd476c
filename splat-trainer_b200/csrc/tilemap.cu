// tilemap.cu — tile-overlap mapping (mapper inside render_projected, reference
// scene/mlp_scene.py:418,377).  The result equals a stable sort of all (row, tile) overlaps by the key
// (tile << 32 | float_bits(depth)), ties by row — but no O-sized pair list is ever sorted:
//   1. per row: a 16-B tile record (rect + exact hit mask), the tile count and per-tile totals (REDs) — produced by
//      the cull in gs_project_fwd, or by k_tile_info here;
//   2. live rows are compacted (one chained-scan kernel) and sorted by depth (onesweep, 32-bit keys);
//   3. in depth order every row emits one pair per 8x8-tile SUPER-tile it hits (O_s ~ O/6) — sorted by super-tile id;
//   4. per-tile totals are scanned into tile_ranges (so the ranges need no pass over the list);
//   5. k_expand: 256·I super-items per CTA, each thread turns its record into the 64-bit mask of sub-tiles hit, a
//      32x64 bit-matrix transpose across the warp (10 shuffles) yields for every sub-tile the lanes hitting it —
//      counts and stable ranks by popc — chunks of one super-tile chain by decoupled look-back, and row ids
//      go straight to their final slot (4 B per overlap written once).
// Everything stays on the device: totals are device counters, buffers are sized by the caller's max_overlaps.
#include "common.cuh"
#include "footprint.cuh"

int gs_radix_sort_prepare(void *workspace, int64_t n, int begin_bit, int end_bit, cudaStream_t st);
int gs_radix_sort_pingpong(uint32_t *ka, uint32_t *va, uint32_t *kb, uint32_t *vb, int64_t n, const int32_t *n_dev,
                           int begin_bit, int end_bit, void *workspace, int *result_in_b, cudaStream_t st, bool hist_ready);

// ---- tile info: what the cull knows about every row ------------------------------------------------------
TileInfo gs_tile_info_carve(void *base, int64_t n, int width, int height) {
  TileInfo t;
  char *p = (char *)base;
  size_t nn = (size_t)(n > 0 ? n : 1);
  (void)width; (void)height;
  t.recs = (uint4 *)p; p += gs_align256(nn * 16);
  t.counts = (int32_t *)p; p += gs_align256(nn * 4);
  t.bytes = (size_t)(p - (char *)base);
  return t;
}

extern "C" size_t gs_tile_info_bytes(int64_t n, int width, int height) {
  return gs_tile_info_carve(nullptr, n, width, height).bytes;
}

__device__ __forceinline__ Footprint load_footprint(const float *__restrict__ g2d, const float *__restrict__ depth,
                                                    int64_t i, float thr, int W, int H) {
  const float2 *g = reinterpret_cast<const float2 *>(g2d + GS_G2D_STRIDE * i);
  float2 a = g[0], b = g[1], c = g[2];
  return row_footprint(a.x, a.y, b.x, b.y, c.x, c.y, depth[i], thr, W, H);
}

// per row: tiles touched
__global__ void __launch_bounds__(256) k_tile_count(const float *__restrict__ g2d, const float *__restrict__ depth,
                                                    int64_t m, float thr, int W, int H, int32_t *__restrict__ counts) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= m) return;
  Footprint fp = load_footprint(g2d, depth, i, thr, W, H);
  counts[i] = footprint_count(fp, W, H);
}

// per row: record + count (operator path; the fused path gets them from gs_project_fwd)
__global__ void __launch_bounds__(256) k_tile_info(const float *__restrict__ g2d, const float *__restrict__ depth,
                                                   int64_t m, float thr, int W, int H, uint4 *__restrict__ recs,
                                                   int32_t *__restrict__ counts) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= m) return;
  Footprint fp = load_footprint(g2d, depth, i, thr, W, H);
  uint4 rec;
  int c = footprint_record(fp, W, H, rec);
  counts[i] = c;
  if (c > 0) recs[i] = rec;
}

__global__ void __launch_bounds__(256) k_sum_i32(const int32_t *__restrict__ v, int64_t n, int32_t *__restrict__ out) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int x = i < n ? v[i] : 0;
  x = __reduce_add_sync(GS_FULL, x);
  __shared__ int s[8];
  if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = x;
  __syncthreads();
  if (threadIdx.x == 0) {
    int t = 0;
    for (int k = 0; k < 8; ++k) t += s[k];
    if (t) atomicAdd(out, t);
  }
}

__device__ __forceinline__ int block_scan_256(int v, int *s_warp, int &block_total) {  // exclusive, 256 threads
  int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  int inc = v;
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
  block_total = tot;
  return wo + inc - v;
}

// ---- live-row compaction (single pass, decoupled look-back) -------------------------------------------
// rows that touch at least one tile, compacted in row order: (depth bits, row) at their rank; live count -> *live_out
constexpr int LC_THREADS = 256;
constexpr int LC_ITEMS = 8;
constexpr int LC_TILE = LC_THREADS * LC_ITEMS;

__global__ void __launch_bounds__(LC_THREADS) k_live_compact(const int32_t *__restrict__ counts, const float *__restrict__ depth,
                                                             int64_t m, uint32_t *__restrict__ status, uint32_t *__restrict__ ticket,
                                                             uint32_t *__restrict__ depth_keys, uint32_t *__restrict__ rows,
                                                             int32_t *__restrict__ live_out, uint32_t *__restrict__ ghist) {
  __shared__ int s_warp[8];
  __shared__ int s_tile;
  __shared__ uint32_t s_prefix;
  __shared__ uint32_t s_hist[4][256];  // digit histograms of this tile's live depth keys (the depth sort's 4 passes)
  if (threadIdx.x == 0) s_tile = (int)atomicAdd(ticket, 1u);
#pragma unroll
  for (int p = 0; p < 4; ++p) s_hist[p][threadIdx.x] = 0u;
  __syncthreads();
  const int tile = s_tile;
  int64_t base = (int64_t)tile * LC_TILE + (int64_t)threadIdx.x * LC_ITEMS;
  int c[LC_ITEMS];
  if (base + LC_ITEMS <= m) {
    const int4 *p = reinterpret_cast<const int4 *>(counts + base);
    int4 a = p[0], b = p[1];
    c[0] = a.x; c[1] = a.y; c[2] = a.z; c[3] = a.w; c[4] = b.x; c[5] = b.y; c[6] = b.z; c[7] = b.w;
  } else {
#pragma unroll
    for (int i = 0; i < LC_ITEMS; ++i) c[i] = (base + i < m) ? counts[base + i] : 0;
  }
  // the depth keys of the thread's rows are fetched now, live or not (two 16-byte loads that run under the scan and the
  // look-back), instead of one dependent 4-byte load per live row after the look-back
  uint32_t dk[LC_ITEMS];
  if (base + LC_ITEMS <= m) {
    const uint4 *p = reinterpret_cast<const uint4 *>(depth + base);
    uint4 a = p[0], b = p[1];
    dk[0] = a.x; dk[1] = a.y; dk[2] = a.z; dk[3] = a.w; dk[4] = b.x; dk[5] = b.y; dk[6] = b.z; dk[7] = b.w;
  } else {
#pragma unroll
    for (int i = 0; i < LC_ITEMS; ++i) dk[i] = (base + i < m) ? __float_as_uint(depth[base + i]) : 0u;
  }
  int sum = 0;
#pragma unroll
  for (int i = 0; i < LC_ITEMS; ++i) sum += c[i] > 0;
  int tot;
  int ex = block_scan_256(sum, s_warp, tot);
  if (threadIdx.x == 0) gs_st_status(status + tile, (uint32_t)tot | (tile == 0 ? GS_LB_INCL : GS_LB_LOCAL));
  if (threadIdx.x < 32) {
    uint32_t prefix = gs_lookback_warp(status, tile);
    if (threadIdx.x == 0) {
      if (tile > 0) gs_st_status(status + tile, (prefix + (uint32_t)tot) | GS_LB_INCL);
      s_prefix = prefix;
      if (tile == (int)gridDim.x - 1) *live_out = (int32_t)(prefix + (uint32_t)tot);
    }
  }
  __syncthreads();
  uint32_t pos = s_prefix + (uint32_t)ex;
#pragma unroll
  for (int i = 0; i < LC_ITEMS; ++i) {
    if (c[i] > 0) {
      uint32_t key = dk[i];
      depth_keys[pos] = key;
      rows[pos] = (uint32_t)(base + i);
      ++pos;
      // the key is in a register anyway: count its four digits here instead of re-reading every key in a histogram
      // kernel of the sort (one launch and one sweep of the keys less)
      atomicAdd(&s_hist[0][key & 255u], 1u);
      atomicAdd(&s_hist[1][(key >> 8) & 255u], 1u);
      atomicAdd(&s_hist[2][(key >> 16) & 255u], 1u);
      atomicAdd(&s_hist[3][key >> 24], 1u);
    }
  }
  __syncthreads();
#pragma unroll
  for (int p = 0; p < 4; ++p) {
    uint32_t h = s_hist[p][threadIdx.x];
    if (h) atomicAdd(ghist + p * 256 + threadIdx.x, h);
  }
}

// ---- super-tile pairs (single pass: chained scan of the per-row super-tile counts in depth order + the pairs) ------
// Thread j = j-th live row of the depth order; it writes one (super-tile id, row) pair per 8x8-tile super-tile the row
// hits, at its scanned offset.
// The last live tile publishes scounters: [0] total pairs, [1] overflow flag, [2] min(total, capacity).
constexpr int EM_THREADS = 256;
constexpr int EM_ITEMS = 4;                       // live rows per thread: a quarter of the CTAs in the look-back chain, four
constexpr int EM_TILE = EM_THREADS * EM_ITEMS;    // independent record gathers in flight per thread

__global__ void __launch_bounds__(EM_THREADS) k_super_emit(const float *__restrict__ g2d, const float *__restrict__ depth,
                                                           const uint4 *__restrict__ recs, const int32_t *__restrict__ live_dev,
                                                           int64_t m, float thr, int W, int H, int SXN,
                                                           const uint32_t *__restrict__ sorted_rows, int64_t cap,
                                                           uint32_t *__restrict__ status, uint32_t *__restrict__ ticket,
                                                           uint32_t *__restrict__ skeys, uint32_t *__restrict__ svals,
                                                           int32_t *__restrict__ scounters, int NS,
                                                           uint32_t *__restrict__ ghist, int32_t *__restrict__ scount) {
  extern __shared__ uint32_t s_cnt[];  // NS: pairs of this tile per super-tile
  __shared__ int s_warp[8];
  __shared__ int s_tile;
  __shared__ uint32_t s_prefix;
  if (threadIdx.x == 0) s_tile = (int)atomicAdd(ticket, 1u);
  for (int i = threadIdx.x; i < NS; i += EM_THREADS) s_cnt[i] = 0u;
  __syncthreads();
  const int tile = s_tile;
  int64_t live = min((int64_t)*live_dev, m);
  int64_t tbase = (int64_t)tile * EM_TILE;
  if (tbase >= live) return;
  const int64_t j0 = tbase + (int64_t)threadIdx.x * EM_ITEMS;
  int64_t row[EM_ITEMS];
  uint4 rec[EM_ITEMS];
  int ns[EM_ITEMS], sum = 0;
#pragma unroll
  for (int i = 0; i < EM_ITEMS; ++i) row[i] = (j0 + i < live) ? (int64_t)sorted_rows[j0 + i] : -1;
#pragma unroll
  for (int i = 0; i < EM_ITEMS; ++i) rec[i] = row[i] >= 0 ? recs[row[i]] : make_uint4(0u, 0u, 0u, 0u);
#pragma unroll
  for (int i = 0; i < EM_ITEMS; ++i) {
    ns[i] = row[i] >= 0 ? rec_view(rec[i]).nsuper : 0;
    sum += ns[i];
  }
  int tot;
  int ex = block_scan_256(sum, s_warp, tot);
  if (threadIdx.x == 0) gs_st_status(status + tile, (uint32_t)tot | (tile == 0 ? GS_LB_INCL : GS_LB_LOCAL));
  if (threadIdx.x < 32) {
    uint32_t prefix = gs_lookback_warp(status, tile);
    if (threadIdx.x == 0) {
      if (tile > 0) gs_st_status(status + tile, (prefix + (uint32_t)tot) | GS_LB_INCL);
      s_prefix = prefix;
      if (tbase + EM_TILE >= live) {  // last live tile
        int64_t o = (int64_t)prefix + tot;
        scounters[0] = (int32_t)o;
        scounters[1] = (o > cap) ? 1 : 0;
        scounters[2] = (int32_t)min(o, cap);
      }
    }
  }
  __syncthreads();
  int64_t o = (int64_t)s_prefix + ex;
  const int lane = threadIdx.x & 31;
#pragma unroll
  for (int i = 0; i < EM_ITEMS; ++i) {
    int sx0 = 0, sx1 = -1, sy0 = 0, sy1 = -1;
    if (ns[i] > 0) {
      RecView v = rec_view(rec[i]);
      if (v.big) {
        Footprint fp = load_footprint(g2d, depth, row[i], thr, W, H);
        sx0 = fp.tx0 / GS_SUPER; sx1 = fp.tx1 / GS_SUPER; sy0 = fp.ty0 / GS_SUPER; sy1 = fp.ty1 / GS_SUPER;
      } else {
        sx0 = v.tx0 / GS_SUPER; sx1 = (v.tx0 + v.w - 1) / GS_SUPER; sy0 = v.ty0 / GS_SUPER; sy1 = (v.ty0 + v.h - 1) / GS_SUPER;
      }
    }
    // a row that spans many super-tiles (a splat filling the screen emits one pair per super-tile) is written by the whole
    // warp together, 32 pairs per round, instead of by its own lane alone while the other 31 wait
    const bool wide = ns[i] >= 16;
    if (ns[i] > 0 && !wide) {
      int64_t oo = o;
      for (int sy = sy0; sy <= sy1; ++sy)
        for (int sx = sx0; sx <= sx1; ++sx) {
          int sid = sy * SXN + sx;
          if (oo < cap) {
            skeys[oo] = (uint32_t)sid;
            svals[oo] = (uint32_t)row[i];
            atomicAdd(&s_cnt[sid], 1u);
          }
          ++oo;
        }
    }
    unsigned todo = __ballot_sync(GS_FULL, wide);
    while (todo) {
      const int src = __ffs(todo) - 1;
      todo &= todo - 1;
      const uint32_t r = __shfl_sync(GS_FULL, (uint32_t)row[i], src);
      const int x0 = __shfl_sync(GS_FULL, sx0, src), y0 = __shfl_sync(GS_FULL, sy0, src);
      const int nx = __shfl_sync(GS_FULL, sx1 - sx0 + 1, src), count = __shfl_sync(GS_FULL, ns[i], src);
      const long long base = __shfl_sync(GS_FULL, (long long)o, src);
      for (int t = lane; t < count; t += 32) {
        int sid = (y0 + t / nx) * SXN + x0 + t % nx;
        if (base + t < cap) {
          skeys[base + t] = (uint32_t)sid;
          svals[base + t] = r;
          atomicAdd(&s_cnt[sid], 1u);
        }
      }
    }
    o += ns[i];
  }
  // per-super-tile totals (the segment table of the sorted list is their scan: no search over the sorted keys) and the
  // digit histograms of the super-tile sort (its histogram kernel and sweep of the keys are gone)
  __syncthreads();
  for (int i = threadIdx.x; i < NS; i += EM_THREADS) {
    uint32_t c = s_cnt[i];
    if (c) {
      atomicAdd((unsigned *)scount + i, c);
      atomicAdd(ghist + (i & 255), c);
      if (NS > 256) atomicAdd(ghist + 256 + (i >> 8), c);
    }
  }
}


// ---- expansion: super-tile items -> final per-tile lists ---------------------------------------------------
constexpr int EX_THREADS = 256;
constexpr int EX_WARPS = EX_THREADS / 32;
constexpr int EX_ITEMS = 2;                       // super-items per thread
constexpr int EX_CHUNK = EX_THREADS * EX_ITEMS;   // super-items per CTA
constexpr int EX_MAX_SUPER = 4096;                // super-tiles whose tables fit shared memory (image up to ~8K x 8K)

// Segment tables of the sorted super-tile list (one block): stab[0..NS] = seg_begin = exclusive scan of the per-super-tile
// pair counts k_super_emit accumulated (no search over the sorted keys), stab[NS+1 .. 2NS+1] = chunk_begin (exclusive scan
// of ceil(segment / chunk)); both end with their grand total.
__global__ void __launch_bounds__(1024) k_seg_tables(const int32_t *__restrict__ scount, int NS, int32_t *__restrict__ stab) {
  __shared__ int s_warp[33];
  __shared__ int s_carry[2];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  if (threadIdx.x < 2) s_carry[threadIdx.x] = 0;
  __syncthreads();
  for (int base = 0; base < NS; base += 1024) {
    int s = base + threadIdx.x;
    int cnt = s < NS ? scount[s] : 0;
#pragma unroll
    for (int which = 0; which < 2; ++which) {  // 0: pairs per super-tile, 1: chunks per super-tile
      int v = which == 0 ? cnt : (cnt + EX_CHUNK - 1) / EX_CHUNK;
      int inc = v;
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
        if (lane == 31) s_warp[32] = xi;
      }
      __syncthreads();
      if (s < NS) stab[which * (NS + 1) + s] = s_carry[which] + s_warp[w] + inc - v;
      __syncthreads();
      if (threadIdx.x == 0) s_carry[which] += s_warp[32];
      __syncthreads();
    }
  }
  if (threadIdx.x == 0) {
    stab[NS] = s_carry[0];
    stab[2 * NS + 1] = s_carry[1];
  }
}

// 32x32 bit-matrix transpose across the warp: on return bit i of lane L = bit L of lane i on entry (5 shuffles)
__device__ __forceinline__ uint32_t warp_transpose32(uint32_t x, int lane) {
#pragma unroll
  for (int j = 16; j >= 1; j >>= 1) {
    uint32_t lowm = (j == 16) ? 0x0000ffffu : (j == 8) ? 0x00ff00ffu : (j == 4) ? 0x0f0f0f0fu : (j == 2) ? 0x33333333u : 0x55555555u;
    uint32_t y = __shfl_xor_sync(GS_FULL, x, j);
    x = (lane & j) ? ((x & ~lowm) | ((y >> j) & lowm)) : ((x & lowm) | ((y << j) & ~lowm));
  }
  return x;
}

struct ChunkCtx {
  int chunk, sid, sx, sy, item0, seg_end;
};

// which super-tile segment a chunk belongs to: last s with chunk_begin[s] <= chunk (table staged in shared memory)
__device__ __forceinline__ void chunk_table(const int32_t *__restrict__ stab, int NS, int32_t *cb_smem) {
  for (int i = threadIdx.x; i <= NS; i += EX_THREADS) cb_smem[i] = stab[NS + 1 + i];
  __syncthreads();
}

__device__ __forceinline__ void chunk_ctx(const int32_t *__restrict__ stab, int NS, int SXN, const int32_t *cb_smem, int chunk,
                                          ChunkCtx &c) {
  c.chunk = chunk;
  int lo = 0, hi = NS;
  while (hi - lo > 1) {
    int mid = (lo + hi) >> 1;
    if (cb_smem[mid] <= c.chunk) lo = mid; else hi = mid;
  }
  c.sid = lo;
  c.sx = lo % SXN; c.sy = lo / SXN;
  c.seg_end = stab[lo + 1];
  c.item0 = stab[lo] + (c.chunk - cb_smem[lo]) * EX_CHUNK + (threadIdx.x >> 5) * (32 * EX_ITEMS);
}

// COUNT pass (WRITE = false): each thread turns its super-items into 64-bit masks of sub-tiles hit (stored for the
// write pass) and the CTA counts the hits per sub-tile -> chunk_counts[chunk][64].
// WRITE pass: chunk_counts now holds each chunk's exclusive prefix within its super-tile; every hit goes to
// ranges[tile].begin + prefix + (hits of the sub-tile by earlier warps / rounds / lanes of the chunk).
template <bool WRITE>
__global__ void __launch_bounds__(EX_THREADS) k_expand(const float *__restrict__ g2d, const float *__restrict__ depth,
                                                       const uint4 *__restrict__ recs, float thr, int W, int H, int SXN, int NS,
                                                       const uint32_t *__restrict__ svals, const int32_t *__restrict__ stab,
                                                       unsigned long long *__restrict__ masks, int32_t *__restrict__ chunk_counts,
                                                       const int32_t *__restrict__ tile_begin, int64_t max_overlaps,
                                                       int32_t *__restrict__ out) {
  extern __shared__ int32_t ex_smem[];               // chunk_begin table: NS + 1 ints
  __shared__ uint32_t sT[EX_WARPS][EX_ITEMS][64];    // transposed hit words: [s] = lanes of the warp hitting sub-tile s
  __shared__ uint32_t sPos[EX_WARPS][EX_ITEMS][64];  // hits of (warp, sub-tile) in earlier rounds
  __shared__ uint32_t sCnt[EX_WARPS][64];
  // persistent CTAs: the grid is sized for the machine, not for the list capacity (a launch with one CTA per POSSIBLE chunk
  // spent most of its time starting CTAs that found nothing to do); the chunk table is staged once per CTA
  // (the write pass is launched with one CTA per POSSIBLE chunk of the list's capacity: the ones past the real count leave
  // before staging the table — on C2 that is 48 k of 58 k CTAs)
  if (WRITE && (int)blockIdx.x >= stab[2 * NS + 1]) return;
  chunk_table(stab, NS, ex_smem);
  const int total_chunks = ex_smem[NS];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  for (int chunk = blockIdx.x; chunk < total_chunks; chunk += gridDim.x) {
  ChunkCtx cc;
  chunk_ctx(stab, NS, SXN, ex_smem, chunk, cc);
  unsigned long long m[EX_ITEMS];
  uint32_t row[EX_ITEMS];
  uint32_t c0 = 0, c1 = 0;
#pragma unroll
  for (int r = 0; r < EX_ITEMS; ++r) {
    int idx = cc.item0 + r * 32 + lane;
    m[r] = 0ull;
    row[r] = 0u;
    if (idx < cc.seg_end) {
      row[r] = svals[idx];
      if (WRITE) {
        m[r] = masks[idx];
      } else {
        RecView v = rec_view(recs[row[r]]);
        if (v.big) {
          const float2 *g = reinterpret_cast<const float2 *>(g2d + GS_G2D_STRIDE * (int64_t)row[r]);
          float2 a = g[0], b = g[1], c = g[2];
          Footprint fp = row_footprint(a.x, a.y, b.x, b.y, c.x, c.y, depth[row[r]], thr, W, H);
          m[r] = fp.valid ? footprint_submask(fp, cc.sx, cc.sy, W, H) : 0ull;
        } else {
          m[r] = rec_submask(v, cc.sx, cc.sy);
        }
        masks[idx] = m[r];
      }
    }
    uint32_t tl = warp_transpose32((uint32_t)m[r], lane), th = warp_transpose32((uint32_t)(m[r] >> 32), lane);
    if (WRITE) {
      sT[w][r][lane] = tl;
      sT[w][r][32 + lane] = th;
      sPos[w][r][lane] = c0;
      sPos[w][r][32 + lane] = c1;
    }
    c0 += __popc(tl);
    c1 += __popc(th);
  }
  sCnt[w][lane] = c0;
  sCnt[w][32 + lane] = c1;
  __syncthreads();
  if (threadIdx.x < 64) {
    const int s = threadIdx.x;
    uint32_t run = 0, wo[EX_WARPS];
#pragma unroll
    for (int i = 0; i < EX_WARPS; ++i) { wo[i] = run; run += sCnt[i][s]; }
    if (!WRITE) {
      chunk_counts[(size_t)cc.chunk * 64 + s] = (int32_t)run;
    } else {
      int TX = (W + GS_TILE - 1) / GS_TILE, TY = (H + GS_TILE - 1) / GS_TILE;
      int tx = GS_SUPER * cc.sx + (s & (GS_SUPER - 1)), ty = GS_SUPER * cc.sy + (s >> 3);
      uint32_t tbegin = (tx < TX && ty < TY) ? (uint32_t)tile_begin[ty * TX + tx] : 0u;
      uint32_t base = tbegin + (uint32_t)chunk_counts[(size_t)cc.chunk * 64 + s];
#pragma unroll
      for (int i = 0; i < EX_WARPS; ++i) sCnt[i][s] = base + wo[i];
    }
  }
  if (WRITE) {
  __syncthreads();
  const unsigned lt = (1u << lane) - 1u;
#pragma unroll
  for (int r = 0; r < EX_ITEMS; ++r) {
    unsigned long long mm = m[r];
    while (mm) {
      int s = __ffsll((long long)mm) - 1;
      mm &= mm - 1ull;
      uint32_t pos = sCnt[w][s] + sPos[w][r][s] + __popc(sT[w][r][s] & lt);
      if ((int64_t)pos < max_overlaps) out[pos] = (int32_t)row[r];
    }
  }
  }
  __syncthreads();  // the shared arrays are reused by the next chunk
  }
}

// One block per super-tile, 16 chunk groups x 64 sub-tiles: turns the chunk counts of the super-tile into exclusive
// prefixes (in place) and the column totals into per-tile totals.  The last block to finish scans the totals into
// tile_ranges (+ per-tile first slots) and the counters: [0] total overlaps O, [1] overflow flag, [2] min(O, capacity).
constexpr int CP_GROUPS = 16;
constexpr int CP_THREADS = CP_GROUPS * 64;

__global__ void __launch_bounds__(CP_THREADS) k_chunk_prefix(const int32_t *__restrict__ stab, int NS, int SXN, int TX, int TY,
                                                             int32_t *__restrict__ chunk_counts, int32_t *__restrict__ totals,
                                                             uint32_t *__restrict__ done, int64_t max_overlaps,
                                                             const int32_t *__restrict__ scounters, int32_t *__restrict__ ranges,
                                                             int32_t *__restrict__ counters) {
  __shared__ int s_g[CP_GROUPS][64];
  __shared__ int s_warp[33];
  __shared__ long long s_carry;
  __shared__ bool s_last;
  const int sid = blockIdx.x, s = threadIdx.x & 63, g = threadIdx.x >> 6;
  const int c0 = stab[NS + 1 + sid], c1 = stab[NS + 2 + sid];
  const int len = (c1 - c0 + CP_GROUPS - 1) / CP_GROUPS;
  const int b0 = min(c0 + g * len, c1), b1 = min(b0 + len, c1);
  int sum = 0;
  for (int c = b0; c < b1; ++c) sum += chunk_counts[(size_t)c * 64 + s];
  s_g[g][s] = sum;
  __syncthreads();
  int run = 0, tot = 0;
#pragma unroll
  for (int i = 0; i < CP_GROUPS; ++i) {
    int v = s_g[i][s];
    if (i < g) run += v;
    tot += v;
  }
  for (int c = b0; c < b1; ++c) {
    int v = chunk_counts[(size_t)c * 64 + s];
    chunk_counts[(size_t)c * 64 + s] = run;
    run += v;
  }
  if (g == 0) {
    int tx = GS_SUPER * (sid % SXN) + (s & (GS_SUPER - 1)), ty = GS_SUPER * (sid / SXN) + (s >> 3);
    if (tx < TX && ty < TY) totals[ty * TX + tx] = tot;
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    s_last = atomicAdd(done, 1u) == (uint32_t)(NS - 1);
    s_carry = 0;
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  // last block: exclusive scan of the T tile totals, 4 tiles per thread per round
  const int T = TX * TY;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  for (int base = 0; base < T; base += CP_THREADS * 4) {
    int t0 = base + threadIdx.x * 4;
    int v[4], ssum = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) { v[i] = (t0 + i < T) ? __ldcg(totals + t0 + i) : 0; ssum += v[i]; }
    int inc = ssum;
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
      if (lane == 31) s_warp[32] = xi;
    }
    __syncthreads();
    long long b = s_carry + s_warp[w] + inc - ssum;
    int round_total = s_warp[32];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      if (t0 + i < T) {
        long long lo = min(b, (long long)max_overlaps), hi = min(b + v[i], (long long)max_overlaps);
        ranges[2 * (t0 + i)] = hi > lo ? (int32_t)lo : 0;  // empty tiles read [0,0)
        ranges[2 * (t0 + i) + 1] = hi > lo ? (int32_t)hi : 0;
        totals[t0 + i] = (int32_t)lo;  // from here on: the tile's first slot (clamped to the capacity)
      }
      b += v[i];
    }
    __syncthreads();
    if (threadIdx.x == 0) s_carry += round_total;
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    long long o = s_carry;
    counters[0] = (int32_t)min(o, (long long)0x7fffffff);
    counters[1] = (o > max_overlaps || scounters[1] != 0) ? 1 : 0;
    counters[2] = (int32_t)min(o, (long long)max_overlaps);
  }
}

// ---- host side ---------------------------------------------------------------------------------------------
// workspace layout (all 256-B aligned)
struct TileWs {
  TileInfo info;        // used when the caller passes no tile info
  uint32_t *dk[2];      // m each: depth keys ping/pong
  uint32_t *dv[2];      // m each: rows ping/pong
  uint32_t *sk[2];      // cap each: super-tile ids ping/pong
  uint32_t *sv[2];      // cap each: rows ping/pong
  unsigned long long *masks;  // cap: sub-tile hit mask of every super-item
  int32_t *chunk_counts;      // ex_chunks * 64
  int32_t *totals;            // T: per-tile totals, then per-tile first slots
  int32_t *stab;              // 2 NS + 2
  uint32_t *zero;       // zeroed every call: tickets[64] | scounters[4] | compaction / emission status words
  size_t zero_bytes;
  int ns, lc_tiles, em_tiles, ex_chunks;
  void *sort_ws;
  size_t total;
};

static TileWs carve(void *base, int64_t m, int64_t maxo, int width, int height) {
  TileWs w;
  char *p = (char *)base;
  auto take = [&](size_t bytes) { char *r = p; p += gs_align256(bytes); return r; };
  size_t mm = (size_t)(m > 0 ? m : 1), oo = (size_t)(maxo > 0 ? maxo : 1);
  w.info = gs_tile_info_carve(p, m, width, height);
  p += w.info.bytes;
  w.dk[0] = (uint32_t *)take(mm * 4); w.dk[1] = (uint32_t *)take(mm * 4);
  w.dv[0] = (uint32_t *)take(mm * 4); w.dv[1] = (uint32_t *)take(mm * 4);
  w.sk[0] = (uint32_t *)take(oo * 4); w.sk[1] = (uint32_t *)take(oo * 4);
  w.sv[0] = (uint32_t *)take(oo * 4); w.sv[1] = (uint32_t *)take(oo * 4);
  w.masks = (unsigned long long *)take(oo * 8);
  int TX = (width + GS_TILE - 1) / GS_TILE, TY = (height + GS_TILE - 1) / GS_TILE;
  w.ns = ((TX + GS_SUPER - 1) / GS_SUPER) * ((TY + GS_SUPER - 1) / GS_SUPER);
  w.lc_tiles = (int)((mm + LC_TILE - 1) / LC_TILE);
  w.em_tiles = (int)((mm + EM_TILE - 1) / EM_TILE);
  w.ex_chunks = (int)((oo + EX_CHUNK - 1) / EX_CHUNK) + w.ns;
  w.chunk_counts = (int32_t *)take((size_t)w.ex_chunks * 64 * 4);
  w.totals = (int32_t *)take((size_t)TX * TY * 4);
  w.stab = (int32_t *)take((size_t)(2 * w.ns + 2) * 4);
  w.zero_bytes = ((size_t)64 + 4 + w.lc_tiles + w.em_tiles + w.ns) * 4;
  w.zero = (uint32_t *)take(w.zero_bytes);
  int64_t big = m > maxo ? m : maxo;
  w.sort_ws = take(gs_radix_sort_workspace_bytes(big));
  w.total = (size_t)(p - (char *)base);
  return w;
}

extern "C" size_t gs_tile_map_workspace_bytes(int64_t m, int64_t max_overlaps, int width, int height) {
  return carve(nullptr, m, max_overlaps, width, height).total + 256;
}

extern "C" int gs_tile_count(const gs_raster_config *h_cfg, int64_t m, const float *g2d, const float *depth, int width,
                             int height, int32_t *counts, int32_t *counters, void *stream) {
  GS_REQUIRE(h_cfg && h_cfg->tile_size == GS_TILE, "tile_size must be 16");
  cudaStream_t st = (cudaStream_t)stream;
  GS_CUDA(cudaMemsetAsync(counters, 0, 4 * sizeof(int32_t), st));
  if (m <= 0) return GS_OK;
  k_tile_count<<<gs_div_up(m, 256), 256, 0, st>>>(g2d, depth, m, h_cfg->alpha_threshold, width, height, counts);
  GS_LAUNCH_CHECK();
  k_sum_i32<<<gs_div_up(m, 256), 256, 0, st>>>(counts, m, counters);
  GS_LAUNCH_CHECK();
  return GS_OK;
}

extern "C" int gs_tile_map(const gs_raster_config *h_cfg, int64_t m, const float *g2d, const float *depth,
                           const void *tile_info_in, int width, int height, int64_t max_overlaps, void *workspace,
                           size_t workspace_bytes, int32_t *overlap_to_point, int32_t *tile_ranges, int32_t *counters,
                           void *stream) {
  GS_REQUIRE(h_cfg && h_cfg->tile_size == GS_TILE, "tile_size must be 16");
  GS_REQUIRE(m >= 0 && m < ((int64_t)1 << 30) && max_overlaps >= 0 && max_overlaps < ((int64_t)1 << 30), "m and max_overlaps must be < 2^30");
  GS_REQUIRE(width > 0 && height > 0, "bad image size");
  cudaStream_t st = (cudaStream_t)stream;
  int TX = (width + GS_TILE - 1) / GS_TILE, TY = (height + GS_TILE - 1) / GS_TILE;
  int64_t T = (int64_t)TX * TY;
  GS_REQUIRE(TX < 65536 && TY < 65536, "image too large");
  uintptr_t aligned = ((uintptr_t)workspace + 255) & ~(uintptr_t)255;
  TileWs w = carve((void *)aligned, m, max_overlaps, width, height);
  if (workspace_bytes < w.total + (aligned - (uintptr_t)workspace)) {
    gs_set_error("gs_tile_map: workspace too small (%zu < %zu)", workspace_bytes, w.total + 256);
    return GS_ERR_WORKSPACE;
  }
  if (w.ns > EX_MAX_SUPER) { gs_set_error("gs_tile_map: more than %d super-tiles", EX_MAX_SUPER); return GS_ERR_UNSUPPORTED; }
  GS_CUDA(cudaMemsetAsync(counters, 0, 4 * sizeof(int32_t), st));
  if (m == 0) {
    GS_CUDA(cudaMemsetAsync(tile_ranges, 0, (size_t)T * 2 * sizeof(int32_t), st));
    return GS_OK;
  }
  float thr = h_cfg->alpha_threshold;
  TileInfo info = w.info;
  if (tile_info_in) {
    info = gs_tile_info_carve((void *)tile_info_in, m, width, height);  // the fused path: filled by gs_project_fwd
  } else {
    k_tile_info<<<gs_div_up(m, 256), 256, 0, st>>>(g2d, depth, m, thr, width, height, info.recs, info.counts);
    GS_LAUNCH_CHECK();
  }
  GS_CUDA(cudaMemsetAsync(w.zero, 0, w.zero_bytes, st));
  uint32_t *tickets = w.zero;
  int32_t *scounters = (int32_t *)(w.zero + 64);
  uint32_t *lc_status = w.zero + 68;
  uint32_t *em_status = lc_status + w.lc_tiles;
  int32_t *scount = (int32_t *)(em_status + w.em_tiles);  // pairs per super-tile
  // (a) compact the live rows (>= 1 tile) in row order, then sort them by depth: stable 32-bit LSD sort, ties keep
  //     row order.  Culled / off-screen rows never enter the sort.
  int32_t *live = counters + 3;
  int rc = gs_radix_sort_prepare(w.sort_ws, m, 0, 32, st);
  if (rc != GS_OK) return rc;
  k_live_compact<<<w.lc_tiles, LC_THREADS, 0, st>>>(info.counts, depth, m, lc_status, tickets + 0, w.dk[0], w.dv[0], live,
                                                    (uint32_t *)w.sort_ws);
  GS_LAUNCH_CHECK();
  int in_b = 0;
  rc = gs_radix_sort_pingpong(w.dk[0], w.dv[0], w.dk[1], w.dv[1], m, live, 0, 32, w.sort_ws, &in_b, st, true);
  if (rc != GS_OK) return rc;
  uint32_t *srows = w.dv[in_b];
  // (b) one (super-tile, row) pair per super-tile a row hits, in depth order; stable sort by super-tile id
  int SXN = (TX + GS_SUPER - 1) / GS_SUPER;
  int bits = 1;
  while ((1 << bits) < w.ns) ++bits;
  rc = gs_radix_sort_prepare(w.sort_ws, max_overlaps, 0, bits, st);  // after the depth sort: the two sorts share the workspace
  if (rc != GS_OK) return rc;
  k_super_emit<<<w.em_tiles, EM_THREADS, (size_t)w.ns * sizeof(uint32_t), st>>>(
      g2d, depth, info.recs, live, m, thr, width, height, SXN, srows, max_overlaps, em_status, tickets + 1, w.sk[0], w.sv[0],
      scounters, w.ns, (uint32_t *)w.sort_ws, scount);
  GS_LAUNCH_CHECK();
  rc = gs_radix_sort_pingpong(w.sk[0], w.sv[0], w.sk[1], w.sv[1], max_overlaps, scounters + 2, 0, bits, w.sort_ws, &in_b, st,
                              max_overlaps > 0);
  if (rc != GS_OK) return rc;
  // (c) segment / chunk tables, per-chunk sub-tile counts, their prefixes + tile ranges, then the final lists
  k_seg_tables<<<1, 1024, 0, st>>>(scount, w.ns, w.stab);
  GS_LAUNCH_CHECK();
  size_t ex_smem = (size_t)(w.ns + 1) * sizeof(int32_t);
  int ex_grid = w.ex_chunks < 148 * 6 ? w.ex_chunks : 148 * 6;
  // the write pass measured faster with one CTA per possible chunk (77 vs 89 us at C2-sparse: its scattered 4-byte stores
  // want every resident CTA slot refilled at once), the count pass with persistent CTAs (56 -> 39 us)
  int ex_grid_w = w.ex_chunks;
  k_expand<false><<<ex_grid, EX_THREADS, ex_smem, st>>>(g2d, depth, info.recs, thr, width, height, SXN, w.ns, w.sv[in_b],
                                                            w.stab, w.masks, w.chunk_counts, nullptr, max_overlaps, nullptr);
  GS_LAUNCH_CHECK();
  k_chunk_prefix<<<w.ns, CP_THREADS, 0, st>>>(w.stab, w.ns, SXN, TX, TY, w.chunk_counts, w.totals, tickets + 2, max_overlaps,
                                      scounters, tile_ranges, counters);
  GS_LAUNCH_CHECK();
  if (max_overlaps == 0) return GS_OK;
  k_expand<true><<<ex_grid_w, EX_THREADS, ex_smem, st>>>(g2d, depth, info.recs, thr, width, height, SXN, w.ns, w.sv[in_b],
                                                           w.stab, w.masks, w.chunk_counts, w.totals, max_overlaps,
                                                           overlap_to_point);
  GS_LAUNCH_CHECK();
  return GS_OK;
}
