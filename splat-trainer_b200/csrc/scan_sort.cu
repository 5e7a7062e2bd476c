// scan_sort.cu — exclusive scan and stable LSD radix sort of (u32 key, u32 value) pairs.
//
// The scan is a deterministic three-kernel scheme (tile reduce -> scan of tile totals -> tile apply); the sort is
// an onesweep LSD radix sort (one histogram read + one chained-scan kernel per 8-bit pass).  HBM-bound integer work.  The sort
// serves SURVEY §8 a8 (the (tile, depth) key sort inside render_projected, reference
// scene/mlp_scene.py:418): rows are sorted by the 32 depth bits (4 passes over M rows), overlaps
// are emitted in that order and stably sorted by tile id (1–2 passes over O pairs), which equals
// one stable sort by the 64-bit key tile<<32|depth but moves ~4.5x fewer bytes.
#include <stdarg.h>

#include "common.cuh"

// ---- error plumbing ---------------------------------------------------------------------------
static thread_local char g_err[512] = "";
void gs_set_error(const char *fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
extern "C" const char *gs_last_error(void) { return g_err; }
// diagnostics only: kernels this library launched in this process (every launch site passes through GS_LAUNCH_CHECK)
std::atomic<unsigned long long> g_gs_launches{0};
extern "C" unsigned long long gs_launch_count(void) { return g_gs_launches.load(std::memory_order_relaxed); }
extern "C" int gs_version(void) { return GSB200_VERSION; }
extern "C" void gs_default_raster_config(gs_raster_config *c) {
  c->tile_size = GS_TILE;
  c->alpha_threshold = 1.0f / 255.0f;
  c->clamp_max_alpha = 0.99f;
  c->saturate_threshold = 0.9999f;
  c->blur_cov = 0.3f;
  c->antialias = 0;
  c->compute_visibility = 0;
  c->compute_point_heuristic = 0;
  c->render_median_depth = 0;
  c->frustum_clamp = 1.3f;
}

// ---- exclusive scan ---------------------------------------------------------------------------
constexpr int SC_THREADS = 256;
constexpr int SC_ITEMS = 16;
constexpr int SC_TILE = SC_THREADS * SC_ITEMS;

int64_t gs_scan_partials(int64_t n) { return (n + SC_TILE - 1) / SC_TILE + 8; }

__device__ __forceinline__ int block_exclusive_scan(int v, int *s_warp, int &block_total) {
  // inclusive warp scan
  int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  int inc = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    int t = __shfl_up_sync(GS_FULL, inc, o);
    if (lane >= o) inc += t;
  }
  if (lane == 31) s_warp[w] = inc;
  __syncthreads();
  if (w == 0) {
    int nw = blockDim.x >> 5;
    int x = lane < nw ? s_warp[lane] : 0;
    int xi = x;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      int t = __shfl_up_sync(GS_FULL, xi, o);
      if (lane >= o) xi += t;
    }
    if (lane < nw) s_warp[lane] = xi - x;  // exclusive warp offsets
    if (lane == nw - 1) s_warp[32] = xi;
  }
  __syncthreads();
  block_total = s_warp[32];
  return s_warp[w] + inc - v;
}

__global__ void __launch_bounds__(SC_THREADS) k_scan_reduce(const int32_t *__restrict__ in, const int32_t *__restrict__ idx,
                                                            int64_t n_cap, const int32_t *__restrict__ n_dev,
                                                            int32_t *__restrict__ partials, int as_flag) {
  __shared__ int s_warp[33];
  int64_t n = n_dev ? min((int64_t)*n_dev, n_cap) : n_cap;
  int64_t base = (int64_t)blockIdx.x * SC_TILE;
  int sum = 0;
#pragma unroll
  for (int i = 0; i < SC_ITEMS; ++i) {
    int64_t j = base + i * SC_THREADS + threadIdx.x;
    if (j < n) {
      int v = idx ? in[idx[j]] : in[j];
      sum += as_flag ? (v > 0) : v;
    }
  }
  int tot;
  block_exclusive_scan(sum, s_warp, tot);
  if (threadIdx.x == 0) partials[blockIdx.x] = tot;
}

// single block: exclusive scan of the tile totals in place; grand total -> total
__global__ void __launch_bounds__(1024) k_scan_partials(int32_t *__restrict__ partials, int nblk, int32_t *__restrict__ total) {
  __shared__ int s_warp[33];
  __shared__ int s_carry;
  if (threadIdx.x == 0) s_carry = 0;
  __syncthreads();
  for (int base = 0; base < nblk; base += 1024) {
    int j = base + threadIdx.x;
    int v = j < nblk ? partials[j] : 0;
    int tot;
    int ex = block_exclusive_scan(v, s_warp, tot);
    int carry = s_carry;
    if (j < nblk) partials[j] = carry + ex;
    __syncthreads();
    if (threadIdx.x == 0) s_carry = carry + tot;
    __syncthreads();
  }
  if (threadIdx.x == 0 && total) *total = s_carry;
}

__global__ void __launch_bounds__(SC_THREADS) k_scan_apply(const int32_t *__restrict__ in, const int32_t *__restrict__ idx,
                                                           int32_t *__restrict__ out, int64_t n_cap,
                                                           const int32_t *__restrict__ n_dev,
                                                           const int32_t *__restrict__ partials, int as_flag) {
  __shared__ int s_warp[33];
  int64_t n = n_dev ? min((int64_t)*n_dev, n_cap) : n_cap;
  int64_t base = (int64_t)blockIdx.x * SC_TILE + (int64_t)threadIdx.x * SC_ITEMS;
  int v[SC_ITEMS];
  int sum = 0;
#pragma unroll
  for (int i = 0; i < SC_ITEMS; ++i) {
    int64_t j = base + i;
    v[i] = (j < n) ? (idx ? in[idx[j]] : in[j]) : 0;
    if (as_flag) v[i] = v[i] > 0;
    sum += v[i];
  }
  int tot;
  int ex = block_exclusive_scan(sum, s_warp, tot) + partials[blockIdx.x];
#pragma unroll
  for (int i = 0; i < SC_ITEMS; ++i) {
    int64_t j = base + i;
    if (j < n) out[j] = ex;
    ex += v[i];
  }
}

int gs_exclusive_scan_i32(const int32_t *in, const int32_t *idx, int32_t *out, int64_t n, const int32_t *n_dev,
                          int32_t *partials, int32_t *total, cudaStream_t st, int as_flag) {
  if (n <= 0) {
    if (total) cudaMemsetAsync(total, 0, sizeof(int32_t), st);
    return GS_OK;
  }
  int nblk = gs_div_up(n, SC_TILE);
  k_scan_reduce<<<nblk, SC_THREADS, 0, st>>>(in, idx, n, n_dev, partials, as_flag);
  k_scan_partials<<<1, 1024, 0, st>>>(partials, nblk, total);
  k_scan_apply<<<nblk, SC_THREADS, 0, st>>>(in, idx, out, n, n_dev, partials, as_flag);
  GS_LAUNCH_CHECK();
  return GS_OK;
}

// ---- radix sort (onesweep) ------------------------------------------------------------------------
// One read of the keys builds the digit histograms of ALL passes (k_os_hist); each pass is then ONE
// kernel: a tile of 4096 pairs is ranked in registers (ballot-based digit matching, stable), its 256
// digit counts are chained to the previous tiles by decoupled look-back, the pairs are re-ordered by
// digit in shared memory and written out as contiguous runs.  Per pass each pair is read once and
// written once (16 B) — the HBM floor of an LSD pass — with 1 launch instead of 5.
constexpr int OS_THREADS = 256;
constexpr int OS_ITEMS = 16;
constexpr int OS_TILE = OS_THREADS * OS_ITEMS;  // 4096 pairs per CTA
constexpr int OS_WARPS = OS_THREADS / 32;
constexpr int OS_MAX_PASSES = 4;

// lanes of the warp holding the same digit as this lane: NB ballots, no match.any (FULL = every lane valid)
template <int NB, bool FULL>
__device__ __forceinline__ unsigned os_match(uint32_t d, bool valid) {
  unsigned peers = 0xffffffffu;
  if (!FULL) {
    unsigned vm = __ballot_sync(GS_FULL, valid);
    peers = valid ? vm : ~vm;
  }
#pragma unroll
  for (int b = 0; b < NB; ++b) {
    bool bit = (d >> b) & 1u;
    unsigned m = __ballot_sync(GS_FULL, bit);
    peers &= bit ? m : ~m;
  }
  return peers;
}

// Digit histograms of all NP passes in one read of the keys.  No atomics and no warp matching on the data path
// (smem atomics cost 2 cycles per lane): every LANE keeps private 8-bit counters, laid out so that the 32 lanes of
// a warp hit 32 different bytes ((warp, pass, digit) row = 32 bytes), and a lane counts at most 255 keys per block.
constexpr int OH_THREADS = 128;
constexpr int OH_WARPS = OH_THREADS / 32;
constexpr int OH_PER_LANE = 192;                       // <= 255 keys per lane per block
constexpr int OH_BLOCK_ITEMS = OH_THREADS * OH_PER_LANE;

template <int NP>
__global__ void __launch_bounds__(OH_THREADS) k_os_hist(const uint32_t *__restrict__ keys, int64_t n_cap,
                                                        const int32_t *__restrict__ n_dev, int begin_bit, int end_bit,
                                                        uint32_t *__restrict__ ghist) {
  extern __shared__ uint32_t oh_smem[];                // [OH_WARPS][NP][256][8 words]
  uint8_t *cnt = reinterpret_cast<uint8_t *>(oh_smem);
  int64_t n = n_dev ? min((int64_t)*n_dev, n_cap) : n_cap;
  int64_t base = (int64_t)blockIdx.x * OH_BLOCK_ITEMS;
  if (base >= n) return;
  for (int i = threadIdx.x; i < OH_WARPS * NP * 256 * 8; i += OH_THREADS) oh_smem[i] = 0;
  __syncthreads();
  int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  uint8_t *mine = cnt + (size_t)w * NP * 256 * 32 + lane;
  int64_t wbase = base + (int64_t)w * (32 * OH_PER_LANE);
  constexpr int U = 8;
#pragma unroll 1
  for (int r0 = 0; r0 < OH_PER_LANE; r0 += U) {
    uint32_t k[U];
    bool ok[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      int64_t j = wbase + (int64_t)(r0 + u) * 32 + lane;
      ok[u] = j < n;
      k[u] = ok[u] ? keys[j] : 0u;
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      if (!ok[u]) continue;
#pragma unroll
      for (int p = 0; p < NP; ++p) {
        int shift = begin_bit + 8 * p;
        int bits = min(8, end_bit - shift);
        uint32_t d = (k[u] >> shift) & ((1u << bits) - 1u);
        uint8_t *c = mine + (p * 256 + d) * 32;
        *c = (uint8_t)(*c + 1);
      }
    }
  }
  __syncthreads();
  // reduce: thread t sums the 32 byte counters of (warp, pass, digit) rows t, t+128, ...
  for (int row = threadIdx.x; row < NP * 256; row += OH_THREADS) {
    uint32_t tot = 0;
#pragma unroll
    for (int ww = 0; ww < OH_WARPS; ++ww) {
      const uint32_t *r = oh_smem + ((size_t)ww * NP * 256 + row) * 8;
#pragma unroll
      for (int q = 0; q < 8; ++q) tot = __dp4a(r[q], 0x01010101u, tot);
    }
    if (tot) atomicAdd(ghist + row, tot);
  }
}

// stable in-warp ranks of the tile's pairs (warp-striped: warp w owns items [w*512, w*512+512), round r = 32 items)
template <int NB, bool FULL>
__device__ __forceinline__ void os_rank(const uint32_t *k, uint16_t *rank, uint32_t (*wcnt)[256], int w, int lane,
                                        int shift, uint32_t mask, int valid_in_warp) {
  unsigned lt = (1u << lane) - 1u;
#pragma unroll
  for (int r = 0; r < OS_ITEMS; ++r) {
    bool valid = FULL || (r * 32 + lane < valid_in_warp);
    uint32_t d = (k[r] >> shift) & mask;
    unsigned peers = os_match<NB, FULL>(d, valid);
    int leader = __ffs(peers) - 1;
    uint32_t pre = 0;
    if (valid && lane == leader) {
      pre = wcnt[w][d];
      wcnt[w][d] = pre + __popc(peers);
    }
    pre = __shfl_sync(GS_FULL, pre, leader);
    rank[r] = (uint16_t)(pre + __popc(peers & lt));
    __syncwarp();
  }
}

template <int NB>
__global__ void __launch_bounds__(OS_THREADS, 3) k_os_pass(const uint32_t *__restrict__ keys_in, const uint32_t *__restrict__ vals_in,
                                                           uint32_t *__restrict__ keys_out, uint32_t *__restrict__ vals_out,
                                                           int n_cap, const int32_t *__restrict__ n_dev, int shift,
                                                           const uint32_t *__restrict__ ghist, uint32_t *__restrict__ status,
                                                           uint32_t *__restrict__ ticket) {
  constexpr uint32_t mask = (1u << NB) - 1u;
  __shared__ uint32_t s_keys[OS_TILE];
  __shared__ uint32_t s_vals[OS_TILE];
  __shared__ uint32_t wcnt[OS_WARPS][256];
  __shared__ uint32_t s_binoff[256];
  __shared__ uint32_t s_gbase[256];  // global position of local slot 0 of the digit's run (mod 2^32)
  __shared__ int s_warp[33];
  __shared__ int s_tile;
  if (threadIdx.x == 0) s_tile = (int)atomicAdd(ticket, 1u);
#pragma unroll
  for (int i = 0; i < OS_WARPS; ++i) wcnt[i][threadIdx.x] = 0;
  __syncthreads();
  const int tile = s_tile;
  const int n = n_dev ? min(*n_dev, n_cap) : n_cap;
  const int base = tile * OS_TILE;
  if (base >= n) return;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int cnt = min(OS_TILE, n - base);
  uint32_t k[OS_ITEMS], v[OS_ITEMS];
  uint16_t rank[OS_ITEMS];
  const int wbase = base + w * (OS_ITEMS * 32);
  const int valid_in_warp = cnt - w * (OS_ITEMS * 32);
  if (cnt == OS_TILE) {
#pragma unroll
    for (int r = 0; r < OS_ITEMS; ++r) { k[r] = keys_in[wbase + r * 32 + lane]; v[r] = vals_in[wbase + r * 32 + lane]; }
    os_rank<NB, true>(k, rank, wcnt, w, lane, shift, mask, valid_in_warp);
  } else {
#pragma unroll
    for (int r = 0; r < OS_ITEMS; ++r) {
      bool valid = r * 32 + lane < valid_in_warp;
      k[r] = valid ? keys_in[wbase + r * 32 + lane] : 0xffffffffu;
      v[r] = valid ? vals_in[wbase + r * 32 + lane] : 0u;
    }
    os_rank<NB, false>(k, rank, wcnt, w, lane, shift, mask, valid_in_warp);
  }
  __syncthreads();
  // thread d owns digit d: per-warp exclusive offsets, tile count, chained prefix over the earlier tiles
  const int d = threadIdx.x;
  uint32_t run = 0;
#pragma unroll
  for (int i = 0; i < OS_WARPS; ++i) {
    uint32_t c = wcnt[i][d];
    wcnt[i][d] = run;
    run += c;
  }
  uint32_t *my_status = status + (size_t)tile * 256 + d;
  gs_st_status(my_status, run | (tile == 0 ? GS_LB_INCL : GS_LB_LOCAL));
  int tot;
  uint32_t binoff = (uint32_t)block_exclusive_scan((int)run, s_warp, tot);
  __syncthreads();  // s_warp reused below
  int gtot;
  uint32_t gex = (uint32_t)block_exclusive_scan((int)ghist[d], s_warp, gtot);
  // look-back, 8 predecessors per round trip (all tiles of a small sort run in one wave, so a one-at-a-time walk
  // would cost O(sqrt(tiles)) dependent L2 latencies)
  uint32_t prefix = 0;
  constexpr int LBW = 8;
  for (int p = tile - 1; p >= 0; p -= LBW) {
    uint32_t sv[LBW];
#pragma unroll
    for (int i = 0; i < LBW; ++i) sv[i] = (p - i >= 0) ? gs_ld_status(status + (size_t)(p - i) * 256 + d) : GS_LB_INCL;
    bool done = false;
#pragma unroll
    for (int i = 0; i < LBW; ++i) {
      if (done) break;
      while ((sv[i] >> 30) == 0) sv[i] = gs_ld_status(status + (size_t)(p - i) * 256 + d);
      prefix += sv[i] & GS_LB_VAL;
      done = (sv[i] >> 30) == 2u;
    }
    if (done) break;
  }
  if (tile > 0) gs_st_status(my_status, (prefix + run) | GS_LB_INCL);
  s_binoff[d] = binoff;
  s_gbase[d] = gex + prefix - binoff;
  __syncthreads();
#pragma unroll
  for (int r = 0; r < OS_ITEMS; ++r) {
    if (r * 32 + lane < valid_in_warp) {
      uint32_t dd = (k[r] >> shift) & mask;
      uint32_t pos = s_binoff[dd] + wcnt[w][dd] + rank[r];
      s_keys[pos] = k[r];
      s_vals[pos] = v[r];
    }
  }
  __syncthreads();
#pragma unroll
  for (int r = 0; r < OS_ITEMS; ++r) {
    int i = r * OS_THREADS + threadIdx.x;
    if (i < cnt) {
      uint32_t key = s_keys[i];
      uint32_t pos = s_gbase[(key >> shift) & mask] + (uint32_t)i;
      keys_out[pos] = key;
      vals_out[pos] = s_vals[i];
    }
  }
}

static size_t os_header_bytes() { return gs_align256((size_t)(OS_MAX_PASSES * 256 + 64) * sizeof(uint32_t)); }

extern "C" size_t gs_radix_sort_workspace_bytes(int64_t n) {
  int64_t nblk = (n + OS_TILE - 1) / OS_TILE;
  if (nblk < 1) nblk = 1;
  return os_header_bytes() + gs_align256((size_t)OS_MAX_PASSES * nblk * 256 * sizeof(uint32_t));
}

// internal: ping-pong sort; returns (via *result_in_b) which buffer pair holds the result
// zero the sort's header (digit histograms, tickets) and look-back status words.  A caller that fills the digit
// histograms itself (ghist[p*256 + digit], p = pass over bits [begin_bit + 8p, ...)) calls this first, then its own
// kernel, then gs_radix_sort_pingpong(hist_ready = true).
int gs_radix_sort_prepare(void *workspace, int64_t n, int begin_bit, int end_bit, cudaStream_t st) {
  if (n <= 0 || end_bit <= begin_bit) return GS_OK;
  int nblk = gs_div_up(n, OS_TILE);
  int passes = (end_bit - begin_bit + 7) / 8;
  GS_CUDA(cudaMemsetAsync(workspace, 0, os_header_bytes() + (size_t)passes * nblk * 256 * sizeof(uint32_t), st));
  return GS_OK;
}

int gs_radix_sort_pingpong(uint32_t *ka, uint32_t *va, uint32_t *kb, uint32_t *vb, int64_t n, const int32_t *n_dev,
                           int begin_bit, int end_bit, void *workspace, int *result_in_b, cudaStream_t st, bool hist_ready) {
  *result_in_b = 0;
  if (n <= 0 || end_bit <= begin_bit) return GS_OK;
  if (n >= ((int64_t)1 << 30)) { gs_set_error("radix sort: n must be < 2^30"); return GS_ERR_INVALID; }
  int nblk = gs_div_up(n, OS_TILE);
  int passes = (end_bit - begin_bit + 7) / 8;
  uint32_t *ghist = (uint32_t *)workspace;
  uint32_t *tickets = ghist + OS_MAX_PASSES * 256;
  uint32_t *status = (uint32_t *)((char *)workspace + os_header_bytes());
  if (!hist_ready) {
    int rc0 = gs_radix_sort_prepare(workspace, n, begin_bit, end_bit, st);
    if (rc0 != GS_OK) return rc0;
  }
  int hgrid = gs_div_up(n, OH_BLOCK_ITEMS);
  size_t hsm = (size_t)OH_WARPS * passes * 256 * 32;
#define GS_HIST(NP)                                                                                            \
  do {                                                                                                         \
    GS_CUDA(cudaFuncSetAttribute(k_os_hist<NP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)hsm));       \
    k_os_hist<NP><<<hgrid, OH_THREADS, hsm, st>>>(ka, n, n_dev, begin_bit, end_bit, ghist);                     \
  } while (0)
  if (!hist_ready) {
    switch (passes) {
      case 1: GS_HIST(1); break;
      case 2: GS_HIST(2); break;
      case 3: GS_HIST(3); break;
      default: GS_HIST(4); break;
    }
  }
#undef GS_HIST
  uint32_t *ks = ka, *vs = va, *kd = kb, *vd = vb;
  for (int p = 0; p < passes; ++p) {
    int bit = begin_bit + 8 * p;
    int bits = min(8, end_bit - bit);
    uint32_t *stp = status + (size_t)p * nblk * 256;
#define GS_PASS(NB) k_os_pass<NB><<<nblk, OS_THREADS, 0, st>>>(ks, vs, kd, vd, (int)n, n_dev, bit, ghist + p * 256, stp, tickets + p)
    switch (bits) {
      case 1: GS_PASS(1); break;
      case 2: GS_PASS(2); break;
      case 3: GS_PASS(3); break;
      case 4: GS_PASS(4); break;
      case 5: GS_PASS(5); break;
      case 6: GS_PASS(6); break;
      case 7: GS_PASS(7); break;
      default: GS_PASS(8); break;
    }
#undef GS_PASS
    uint32_t *t;
    t = ks; ks = kd; kd = t;
    t = vs; vs = vd; vd = t;
    *result_in_b ^= 1;
  }
  GS_LAUNCH_CHECK();
  return GS_OK;
}

extern "C" int gs_radix_sort_pairs(uint32_t *keys_in, uint32_t *vals_in, uint32_t *keys_out, uint32_t *vals_out, int64_t n,
                                   const int32_t *n_dev, int begin_bit, int end_bit, void *workspace,
                                   size_t workspace_bytes, void *stream) {
  GS_REQUIRE(n >= 0 && begin_bit >= 0 && end_bit <= 32 && begin_bit <= end_bit, "bad range");
  GS_REQUIRE(n < ((int64_t)1 << 31), "n must fit int32");
  if (workspace_bytes < gs_radix_sort_workspace_bytes(n)) {
    gs_set_error("gs_radix_sort_pairs: workspace too small");
    return GS_ERR_WORKSPACE;
  }
  cudaStream_t st = (cudaStream_t)stream;
  int in_b = 0;
  int rc = gs_radix_sort_pingpong(keys_in, vals_in, keys_out, vals_out, n, n_dev, begin_bit, end_bit, workspace, &in_b, st, false);
  if (rc != GS_OK) return rc;
  if (!in_b && n > 0) {  // even number of passes (or none): result sits in the input pair
    GS_CUDA(cudaMemcpyAsync(keys_out, keys_in, (size_t)n * 4, cudaMemcpyDeviceToDevice, st));
    GS_CUDA(cudaMemcpyAsync(vals_out, vals_in, (size_t)n * 4, cudaMemcpyDeviceToDevice, st));
  }
  return GS_OK;
}
