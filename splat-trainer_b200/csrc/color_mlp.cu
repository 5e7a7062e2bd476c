// color_mlp.cu — the colour MLP path of MLPScene: eval_colors -> ColorModel.forward, reference
// scene/mlp_scene.py:352-368, scene/color_model.py:124-149, scene/mlp/torch_mlp.py:52-66 (AffineMLP), :89-97 (GLU layer).
//
// Per in-view splat: LayerNorm([point feature 16, GLO feature 32]) -> base MLP (48 -> GLU 32 -> 4) and directional MLP
// (x = f * a + b with (a, b) = Linear(real SH degree 5 of the view direction), then 48 -> GLU 32 -> 4);
// colour = sigmoid(rgb) * exp(intensity) (+ the same with intensity - 2 for the specular term).
// The reference runs this under torch.autocast(float16): every Linear has fp16 inputs / weights / outputs with fp32
// accumulation.  That is exactly the tensor-core contract, so the five Linear layers run as m16n8k16 fp16 MMAs with
// fp32 accumulators, one warp per 32 rows, activations chained in registers (the accumulator fragment of one layer is
// the A fragment of the next), weights resident in shared memory as fp16.  Nothing but the final colour leaves the SM.
#include <cuda_fp16.h>

#include "common.cuh"

namespace {

constexpr int PF = 16, GF = 32, F = PF + GF, H = 32, DEG = 5, S = (DEG + 1) * (DEG + 1);  // config/scene/mlp.yaml:33-43
constexpr int KP = 56;  // row pitch (halfs) of 48-wide fp16 tiles: conflict-free fragment loads (28 words: 4-bank steps)
constexpr int HP = 40;  // row pitch (halfs) of 32-wide fp16 tiles
constexpr int CM_THREADS = 128, CM_WARPS = CM_THREADS / 32;  // backward CTA

struct ColorDev {
  const float *w[10];  // base_w1 base_b1 base_w2 base_b2 dir_wsh dir_bsh dir_w1 dir_b1 dir_w2 dir_b2
};

// fp16 weights + fp16-rounded biases of the whole model, shared by the CTA
struct Weights {
  __half w1b[2 * H * KP], w1d[2 * H * KP], wsh[2 * F * KP], w2b[8 * HP], w2d[8 * HP];
  float b1b[2 * H], b1d[2 * H], bsh[2 * F], b2b[8], b2d[8];
};

// per-warp staging of one 32-row group
struct Stage {
  __half xh[32 * KP];  // LayerNorm output, fp16 (A operand of the base layer)
  __half yh[32 * KP];  // SH coefficients, fp16, zero padded to 48 (A operand of the direction encoder)
  float xf[32 * KP];   // LayerNorm output, fp32 (x = f * a + b multiplies the fp32 feature, color_model.py / torch_mlp.py:65)
  // the outputs of the two MLPs (2 x 32 x 8 floats) reuse the bytes of xh once its fragments are in registers
  uint16_t list[512];  // the warp's compacted row offsets of its current 512-row slice
};
constexpr int FW_SLICE = 512;
constexpr int FW_THREADS = 384, FW_WARPS = FW_THREADS / 32;  // forward: 12 warps share one copy of the weights per SM

__device__ __forceinline__ float hr(float v) { return __half2float(__float2half_rn(v)); }
__device__ __forceinline__ uint32_t pack2(float a, float b) {
  __half2 h = __floats2half2_rn(a, b);
  return *reinterpret_cast<uint32_t *>(&h);
}
// two values rounded to fp16 and back with ONE packed conversion each way (the scalar conversions issue on the quarter-rate
// conversion pipe; the packed forms do not)
__device__ __forceinline__ float2 hr2(float a, float b) { return __half22float2(__floats2half2_rn(a, b)); }
__device__ __forceinline__ float sigmoidf(float v) { return __fdividef(1.0f, 1.0f + __expf(-v)); }

__device__ __forceinline__ void mma16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

// A fragment (16 rows x 16 k) of a row-major fp16 tile: rows r0 + g / g + 8, columns k0 + 2t (+8)
__device__ __forceinline__ void load_a(uint32_t (&a)[4], const __half *tile, int pitch, int r0, int k0, int g, int t) {
  const __half *p = tile + (r0 + g) * pitch + k0 + 2 * t;
  a[0] = *reinterpret_cast<const uint32_t *>(p);
  a[1] = *reinterpret_cast<const uint32_t *>(p + 8 * pitch);
  a[2] = *reinterpret_cast<const uint32_t *>(p + 8);
  a[3] = *reinterpret_cast<const uint32_t *>(p + 8 * pitch + 8);
}
// B fragment (16 k x 8 n) of y = x W^T from W stored row-major (out, in): B[k][n] = W[n0 + n][k0 + k]
__device__ __forceinline__ void load_b(uint32_t &b0, uint32_t &b1, const __half *w, int pitch, int n0, int k0, int g, int t) {
  const __half *p = w + (n0 + g) * pitch + k0 + 2 * t;
  b0 = *reinterpret_cast<const uint32_t *>(p);
  b1 = *reinterpret_cast<const uint32_t *>(p + 8);
}

// Real spherical harmonics up to degree DEG in the reference's convention (scene/mlp/rsh.py: index l(l+1)+m,
// Condon-Shortley phase), by the standard recurrences — fully unrolled, all constants fold at compile time.
__host__ __device__ constexpr double cm_K(int l, int m) {  // sqrt((2l+1)/(4 pi) (l-m)!/(l+m)!)
  double r = (2 * l + 1) / (4.0 * 3.14159265358979323846);
  for (int k = l - m + 1; k <= l + m; ++k) r /= k;
  // constexpr sqrt by Newton iterations
  double x = r > 1 ? r : 1;
  for (int i = 0; i < 60; ++i) x = 0.5 * (x + r / x);
  return x;
}
__device__ __forceinline__ void rsh5(float x, float y, float z, float *Y) {
  float c[DEG + 1], s[DEG + 1];
  c[0] = 1.f; s[0] = 0.f;
#pragma unroll
  for (int m = 1; m <= DEG; ++m) { c[m] = x * c[m - 1] - y * s[m - 1]; s[m] = x * s[m - 1] + y * c[m - 1]; }
#pragma unroll
  for (int m = 0; m <= DEG; ++m) {
    float tmm = 1.f;
#pragma unroll
    for (int k = 1; k <= m; ++k) tmm *= (float)(2 * k - 1);
    float tp2 = 0.f, tp = tmm;
#pragma unroll
    for (int l = m; l <= DEG; ++l) {
      float T = l == m ? tmm : (l == m + 1 ? (2 * m + 1) * z * tmm : ((2 * l - 1) * z * tp - (l + m - 1) * tp2) / (float)(l - m));
      if (l > m) { tp2 = tp; tp = T; }
      if (m == 0) {
        Y[l * (l + 1)] = (float)cm_K(l, 0) * T;
      } else {
        float k = (float)(1.41421356237309504880 * cm_K(l, m)) * ((m & 1) ? -1.f : 1.f);
        Y[l * (l + 1) + m] = k * c[m] * T;
        Y[l * (l + 1) - m] = k * s[m] * T;
      }
    }
  }
}

__device__ __forceinline__ void load_weights(Weights &W, const ColorDev &m) {
  uint32_t *z = reinterpret_cast<uint32_t *>(&W);
  for (int i = threadIdx.x; i < (int)(sizeof(Weights) / 4); i += blockDim.x) z[i] = 0u;  // the pads must be zero
  __syncthreads();
  for (int i = threadIdx.x; i < 2 * H * F; i += blockDim.x) {
    int o = i / F, k = i - o * F;
    W.w1b[o * KP + k] = __float2half_rn(m.w[0][i]);
    W.w1d[o * KP + k] = __float2half_rn(m.w[6][i]);
  }
  for (int i = threadIdx.x; i < 2 * F * S; i += blockDim.x) {
    int o = i / S, k = i - o * S;
    W.wsh[o * KP + k] = __float2half_rn(m.w[4][i]);
  }
  for (int i = threadIdx.x; i < 4 * H; i += blockDim.x) {
    int o = i / H, k = i - o * H;
    W.w2b[o * HP + k] = __float2half_rn(m.w[2][i]);
    W.w2d[o * HP + k] = __float2half_rn(m.w[8][i]);
  }
  for (int i = threadIdx.x; i < 2 * H; i += blockDim.x) { W.b1b[i] = hr(m.w[1][i]); W.b1d[i] = hr(m.w[7][i]); }
  for (int i = threadIdx.x; i < 2 * F; i += blockDim.x) W.bsh[i] = hr(m.w[5][i]);
  if (threadIdx.x < 4) { W.b2b[threadIdx.x] = hr(m.w[3][threadIdx.x]); W.b2d[threadIdx.x] = hr(m.w[9][threadIdx.x]); }
  __syncthreads();
}

// lane = row: LayerNorm of [feature, glo] and the SH coefficients of the view direction into the warp's staging tiles.
// Returns 1 / sigma of the row (the backward needs it).
template <typename StageT>
__device__ __forceinline__ float stage_row(StageT &st, int lane, bool valid, int64_t i, const float *__restrict__ feature,
                                           const float *__restrict__ position, const float (&cam)[3],
                                           const float *__restrict__ s_glo) {
  float v[PF];
  float px = 0.f, py = 0.f, pz = 1.f;
  if (valid) {
    const float4 *f4 = reinterpret_cast<const float4 *>(feature + (int64_t)PF * i);
#pragma unroll
    for (int q = 0; q < PF / 4; ++q) {
      float4 a = f4[q];
      v[4 * q] = a.x; v[4 * q + 1] = a.y; v[4 * q + 2] = a.z; v[4 * q + 3] = a.w;
    }
    px = position[3 * i] - cam[0]; py = position[3 * i + 1] - cam[1]; pz = position[3 * i + 2] - cam[2];
  } else {
#pragma unroll
    for (int q = 0; q < PF; ++q) v[q] = 0.f;
  }
  float mean = 0.f;
#pragma unroll
  for (int q = 0; q < PF; ++q) mean += v[q];
  for (int q = 0; q < GF; ++q) mean += s_glo[q];
  mean *= (1.0f / F);
  float var = 0.f;
#pragma unroll
  for (int q = 0; q < PF; ++q) var += (v[q] - mean) * (v[q] - mean);
  for (int q = 0; q < GF; ++q) var += (s_glo[q] - mean) * (s_glo[q] - mean);
  float inv = 1.0f / sqrtf(var * (1.0f / F) + 1e-5f);
  float *xf = st.xf + lane * KP;
  __half *xh = st.xh + lane * KP;
#pragma unroll
  for (int q = 0; q < PF; q += 2) {
    float a = (v[q] - mean) * inv, b = (v[q + 1] - mean) * inv;
    *reinterpret_cast<float2 *>(xf + q) = make_float2(a, b);
    *reinterpret_cast<__half2 *>(xh + q) = __floats2half2_rn(a, b);
  }
  for (int q = 0; q < GF; q += 2) {
    float a = (s_glo[q] - mean) * inv, b = (s_glo[q + 1] - mean) * inv;
    *reinterpret_cast<float2 *>(xf + PF + q) = make_float2(a, b);
    *reinterpret_cast<__half2 *>(xh + PF + q) = __floats2half2_rn(a, b);
  }
  float len = fmaxf(sqrtf(px * px + py * py + pz * pz), 1e-12f);  // F.normalize: x / max(|x|, eps)
  float Y[S];
  rsh5(px / len, py / len, pz / len, Y);
  __half *yh = st.yh + lane * KP;
#pragma unroll
  for (int q = 0; q < S; q += 2) *reinterpret_cast<__half2 *>(yh + q) = __floats2half2_rn(Y[q], Y[q + 1]);
#pragma unroll
  for (int q = S; q < F; q += 2) *reinterpret_cast<__half2 *>(yh + q) = __floats2half2_rn(0.f, 0.f);
  return inv;
}

// direction encoder + affine modulation for one 16-row tile: AX = fragments of x = f * a + b, (a, b) = Wsh Y + bsh
template <bool SAVE, typename StageT>
__device__ __forceinline__ void affine_tile(const Weights &W, StageT &st, int r0, int g, int t, uint32_t (&AX)[3][4],
                                            uint32_t (*av)[2] = nullptr) {
  uint32_t AY[3][4];
#pragma unroll
  for (int ks = 0; ks < 3; ++ks) load_a(AY[ks], st.yh, KP, r0, 16 * ks, g, t);
#pragma unroll
  for (int j = 0; j < 3; ++j) {
#pragma unroll
    for (int q = 0; q < 2; ++q) {
      int col = 8 * (2 * j + q) + 2 * t;
      float ca[4] = {W.bsh[col], W.bsh[col + 1], W.bsh[col], W.bsh[col + 1]};
      float cb[4] = {W.bsh[F + col], W.bsh[F + col + 1], W.bsh[F + col], W.bsh[F + col + 1]};
#pragma unroll
      for (int ks = 0; ks < 3; ++ks) {
        uint32_t b0, b1r;
        load_b(b0, b1r, W.wsh, KP, 8 * (2 * j + q), 16 * ks, g, t);
        mma16816(ca, AY[ks], b0, b1r);
        load_b(b0, b1r, W.wsh, KP, F + 8 * (2 * j + q), 16 * ks, g, t);
        mma16816(cb, AY[ks], b0, b1r);
      }
      float2 lo = *reinterpret_cast<const float2 *>(st.xf + (r0 + g) * KP + col);
      float2 hi = *reinterpret_cast<const float2 *>(st.xf + (r0 + g + 8) * KP + col);
      AX[j][2 * q] = pack2(lo.x * hr(ca[0]) + hr(cb[0]), lo.y * hr(ca[1]) + hr(cb[1]));
      AX[j][2 * q + 1] = pack2(hi.x * hr(ca[2]) + hr(cb[2]), hi.y * hr(ca[3]) + hr(cb[3]));
      if (SAVE) {  // backward: keep a (fp16) and leave x as a tile for the dW pass (over the rows of f it replaces)
        av[2 * j + q][0] = pack2(ca[0], ca[1]);
        av[2 * j + q][1] = pack2(ca[2], ca[3]);
        *reinterpret_cast<uint32_t *>(st.xh + (r0 + g) * KP + col) = AX[j][2 * q];
        *reinterpret_cast<uint32_t *>(st.xh + (r0 + g + 8) * KP + col) = AX[j][2 * q + 1];
      }
    }
  }
}

// GLU layer + output layer of one MLP (and, below, the affine step) for BOTH 16-row tiles of a warp's 32 rows at once:
// every weight fragment is loaded once and feeds two independent accumulator chains (half the shared-memory traffic,
// twice the MMAs in flight).  A = input fragments (3 k-steps of 16 per tile); the fp16-rounded outputs (4 valid columns)
// go to `out` (32 x 8 floats).
__device__ __forceinline__ void mlp_tile2(const __half *__restrict__ w1, const float *__restrict__ b1,
                                          const __half *__restrict__ w2, const float *__restrict__ b2, float *__restrict__ out,
                                          int g, int t, const uint32_t (&A)[2][3][4]) {
  uint32_t A2[2][2][4];
#pragma unroll
  for (int j = 0; j < 2; ++j) {
#pragma unroll
    for (int q = 0; q < 2; ++q) {
      int col = 8 * (2 * j + q) + 2 * t;
      float ca[2][4], cb[2][4];
#pragma unroll
      for (int mt = 0; mt < 2; ++mt) {
        ca[mt][0] = ca[mt][2] = b1[col]; ca[mt][1] = ca[mt][3] = b1[col + 1];
        cb[mt][0] = cb[mt][2] = b1[H + col]; cb[mt][1] = cb[mt][3] = b1[H + col + 1];
      }
#pragma unroll
      for (int ks = 0; ks < 3; ++ks) {
        uint32_t b0, b1r, c0, c1r;
        load_b(b0, b1r, w1, KP, 8 * (2 * j + q), 16 * ks, g, t);
        load_b(c0, c1r, w1, KP, H + 8 * (2 * j + q), 16 * ks, g, t);
        mma16816(ca[0], A[0][ks], b0, b1r);
        mma16816(ca[1], A[1][ks], b0, b1r);
        mma16816(cb[0], A[0][ks], c0, c1r);
        mma16816(cb[1], A[1][ks], c0, c1r);
      }
#pragma unroll
      for (int mt = 0; mt < 2; ++mt) {
        float2 a01 = hr2(ca[mt][0], ca[mt][1]), a23 = hr2(ca[mt][2], ca[mt][3]);
        float2 b01 = hr2(cb[mt][0], cb[mt][1]), b23 = hr2(cb[mt][2], cb[mt][3]);
        A2[mt][j][2 * q] = pack2(a01.x * sigmoidf(b01.x), a01.y * sigmoidf(b01.y));
        A2[mt][j][2 * q + 1] = pack2(a23.x * sigmoidf(b23.x), a23.y * sigmoidf(b23.y));
      }
    }
  }
  float co[2][4];
#pragma unroll
  for (int mt = 0; mt < 2; ++mt) { co[mt][0] = co[mt][2] = b2[2 * t]; co[mt][1] = co[mt][3] = b2[2 * t + 1]; }
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    uint32_t b0, b1r;
    load_b(b0, b1r, w2, HP, 0, 16 * j, g, t);
    mma16816(co[0], A2[0][j], b0, b1r);
    mma16816(co[1], A2[1][j], b0, b1r);
  }
  if (t < 2) {
#pragma unroll
    for (int mt = 0; mt < 2; ++mt) {
      *reinterpret_cast<float2 *>(out + (16 * mt + g) * 8 + 2 * t) = hr2(co[mt][0], co[mt][1]);
      *reinterpret_cast<float2 *>(out + (16 * mt + g + 8) * 8 + 2 * t) = hr2(co[mt][2], co[mt][3]);
    }
  }
}

template <typename StageT>
__device__ __forceinline__ void affine_tile2(const Weights &W, StageT &st, int g, int t, uint32_t (&AX)[2][3][4]) {
  uint32_t AY[2][3][4];
#pragma unroll
  for (int mt = 0; mt < 2; ++mt)
#pragma unroll
    for (int ks = 0; ks < 3; ++ks) load_a(AY[mt][ks], st.yh, KP, 16 * mt, 16 * ks, g, t);
#pragma unroll
  for (int j = 0; j < 3; ++j) {
#pragma unroll
    for (int q = 0; q < 2; ++q) {
      int col = 8 * (2 * j + q) + 2 * t;
      float ca[2][4], cb[2][4];
#pragma unroll
      for (int mt = 0; mt < 2; ++mt) {
        ca[mt][0] = ca[mt][2] = W.bsh[col]; ca[mt][1] = ca[mt][3] = W.bsh[col + 1];
        cb[mt][0] = cb[mt][2] = W.bsh[F + col]; cb[mt][1] = cb[mt][3] = W.bsh[F + col + 1];
      }
#pragma unroll
      for (int ks = 0; ks < 3; ++ks) {
        uint32_t b0, b1r, c0, c1r;
        load_b(b0, b1r, W.wsh, KP, 8 * (2 * j + q), 16 * ks, g, t);
        load_b(c0, c1r, W.wsh, KP, F + 8 * (2 * j + q), 16 * ks, g, t);
        mma16816(ca[0], AY[0][ks], b0, b1r);
        mma16816(ca[1], AY[1][ks], b0, b1r);
        mma16816(cb[0], AY[0][ks], c0, c1r);
        mma16816(cb[1], AY[1][ks], c0, c1r);
      }
#pragma unroll
      for (int mt = 0; mt < 2; ++mt) {
        float2 lo = *reinterpret_cast<const float2 *>(st.xf + (16 * mt + g) * KP + col);
        float2 hi = *reinterpret_cast<const float2 *>(st.xf + (16 * mt + g + 8) * KP + col);
        float2 a01 = hr2(ca[mt][0], ca[mt][1]), a23 = hr2(ca[mt][2], ca[mt][3]);
        float2 b01 = hr2(cb[mt][0], cb[mt][1]), b23 = hr2(cb[mt][2], cb[mt][3]);
        AX[mt][j][2 * q] = pack2(lo.x * a01.x + b01.x, lo.y * a01.y + b01.y);
        AX[mt][j][2 * q + 1] = pack2(hi.x * a23.x + b23.x, hi.y * a23.y + b23.y);
      }
    }
  }
}

// block compaction of the rows of one chunk that have work (in view / blended), like k_project_bwd
template <int CHUNK, typename T>
__device__ __forceinline__ int compact_chunk(int64_t base, int64_t n, const float *__restrict__ mask, int mask_stride,
                                             T *s_list, int *s_count) {
  if (threadIdx.x == 0) *s_count = 0;
  __syncthreads();
  for (int k = threadIdx.x; k < CHUNK; k += blockDim.x) {
    int64_t i = base + k;
    if (i < n && (!mask || mask[(int64_t)mask_stride * i] > 0.f)) s_list[atomicAdd(s_count, 1)] = (T)k;
  }
  __syncthreads();
  return *s_count;
}

__global__ void __launch_bounds__(FW_THREADS, 1) k_color_fwd(ColorDev m, int64_t n, const float *__restrict__ in_view,
                                                          int in_view_stride, const float *__restrict__ feature,
                                                          const float *__restrict__ position, float cx, float cy, float cz,
                                                          const float *__restrict__ glo, float specular_weight,
                                                          float *__restrict__ rgb, float *__restrict__ specular) {
  extern __shared__ __align__(16) unsigned char smem[];
  Weights &W = *reinterpret_cast<Weights *>(smem);
  Stage *stages = reinterpret_cast<Stage *>(smem + ((sizeof(Weights) + 15) & ~(size_t)15));
  float *s_glo = reinterpret_cast<float *>(stages + FW_WARPS);
  load_weights(W, m);
  if (threadIdx.x < GF) s_glo[threadIdx.x] = glo[threadIdx.x];
  __syncthreads();
  const float cam[3] = {cx, cy, cz};
  int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  Stage &st = stages[warp];
  float *so = reinterpret_cast<float *>(st.xh);
  // Every warp works alone from here on (no CTA barrier): it compacts the in-view rows of its own 512-row slice by
  // ballot and walks them 32 at a time, so warps never wait for each other's uneven slices.
  int64_t stride = (int64_t)gridDim.x * FW_WARPS * FW_SLICE;
  for (int64_t base = ((int64_t)blockIdx.x * FW_WARPS + warp) * FW_SLICE; base < n; base += stride) {
    int cnt = 0;
#pragma unroll 4
    for (int k = 0; k < FW_SLICE; k += 32) {
      int64_t r = base + k + lane;
      bool on = r < n && (!in_view || in_view[(int64_t)in_view_stride * r] > 0.f);
      unsigned bal = __ballot_sync(GS_FULL, on);
      if (on) st.list[cnt + __popc(bal & ((1u << lane) - 1u))] = (uint16_t)(k + lane);
      cnt += __popc(bal);
    }
    __syncwarp();
    for (int k0 = 0; k0 < cnt; k0 += 32) {
      bool valid = k0 + lane < cnt;
      int64_t i = valid ? base + st.list[k0 + lane] : 0;
      stage_row(st, lane, valid, i, feature, position, cam, s_glo);
      __syncwarp();
      {
        uint32_t A[2][3][4];
#pragma unroll
        for (int mt = 0; mt < 2; ++mt)
#pragma unroll
          for (int ks = 0; ks < 3; ++ks) load_a(A[mt][ks], st.xh, KP, 16 * mt, 16 * ks, g, t);
        __syncwarp();  // xh is in registers: its bytes now take the MLP outputs
        mlp_tile2(W.w1b, W.b1b, W.w2b, W.b2b, so, g, t, A);
        affine_tile2(W, st, g, t, A);
        mlp_tile2(W.w1d, W.b1d, W.w2d, W.b2d, so + 32 * 8, g, t, A);
      }
      __syncwarp();
      if (valid) {
        const float *o = so + lane * 8, *o2 = so + 32 * 8 + lane * 8;
        float lum = expf(o[0]), lum2 = expf(o2[0] - 2.0f);  // luminance_activation, intensity_bias -2 for the specular term
#pragma unroll
        for (int c = 0; c < 3; ++c) {
          float d = lum / (1.0f + expf(-o[1 + c])), sp = lum2 / (1.0f + expf(-o2[1 + c]));
          rgb[3 * i + c] = d + specular_weight * sp;
          if (specular) specular[3 * i + c] = sp;
        }
      }
      __syncwarp();
    }
  }
}

size_t color_smem_bytes() {
  return ((sizeof(Weights) + 15) & ~(size_t)15) + FW_WARPS * sizeof(Stage) + GF * sizeof(float) + 16;
}

int color_check(const gs_color_mlp *m) {
  if (!m) { gs_set_error("gs_color_mlp: null model"); return GS_ERR_INVALID; }
  if (m->point_features != PF || m->glo_features != GF || m->hidden != H || m->sh_degree != DEG) {
    gs_set_error("gs_color_mlp: only point_features 16, glo_features 32, hidden 32, sh_degree 5 (config/scene/mlp.yaml) is instantiated");
    return GS_ERR_UNSUPPORTED;
  }
  for (int k = 0; k < 10; ++k)
    if (!m->weights[k]) { gs_set_error("gs_color_mlp: null weight pointer %d", k); return GS_ERR_INVALID; }
  return GS_OK;
}


// ---- backward ---------------------------------------------------------------------------------------------------
// Rows a pixel blended only.  Per 32-row group a warp recomputes the forward, back-propagates through both MLPs with
// the same fragment chaining (dH is the A operand of dIn = dH W1), leaves dH / dE / dO and the layer inputs as fp16 tiles
// in shared memory and forms the weight gradients dW = dH^T In as MMAs over the 32 rows (operands transposed on the fly by
// ldmatrix.trans; a constant-1 column appended to every input tile turns the bias gradient into one more column of dW).
// The dW pass is cooperative: after a round of 128 rows (32 per warp) each warp owns one 16-output slab of every weight
// matrix and reduces it over all 128 staged rows into the CTA's fp32 accumulators in shared memory (plain
// read-modify-write: the slab has one owner), which are added to global memory once per CTA.
constexpr int TP = 72;    // pitch (halfs) of W1^T (in, out) and of the dH tile (64 wide)
constexpr int OP = 24;    // pitch of W2^T (hidden, out) and of the dO tile (4 valid of 16)
constexpr int EP = 104;   // pitch of the dE tile (96 wide); 208 B rows, shared with the fp32 df tile (52 floats)
constexpr int DFP = 52;
constexpr int CMB_CHUNK = 4096;  // backward: few rows of a chunk were blended, so a round of 128 needs a longer chunk
// Loss scale of the fp16 backward operands (dO, dH, dE).  Per-splat colour gradients are 1e-8..1e-2: unscaled they sit in
// (or below) fp16's subnormal range — the reference's own autocast backward has no GradScaler and loses them.  Everything
// downstream of dO is linear in it, so the fp32 results are simply multiplied by 1/scale at the end.
constexpr float GRAD_SCALE = 4096.0f, GRAD_UNSCALE = 1.0f / 4096.0f, HALF_SAT = 60000.0f;

struct WeightsT {
  __half w1bT[F * TP], w1dT[F * TP], w2bT[H * OP], w2dT[H * OP];
};

struct GradAcc {  // per-CTA fp32 accumulators, shapes of the reference's parameters
  float w1b[2 * H * F], b1b[2 * H], w2b[4 * H], b2b[4], wsh[2 * F * S], bsh[2 * F], w1d[2 * H * F], b1d[2 * H], w2d[4 * H], b2d[4];
  float glo[GF];
};

struct StageB {
  __half xh[32 * KP];     // f (fp16) with a ones column at 48; rows are overwritten by x for the directional phase
  __half yh[32 * KP];     // SH coefficients with a ones column at 36
  float xf[32 * KP];
  float o[2][32 * 8];
  __half dos[2][32 * OP]; // dO of the two MLPs
  __half gs[32 * HP];     // GLU output with a ones column at 32
  __half dhs[32 * TP];    // dH
  __half des[32 * EP];    // dE = [da, db]; the same bytes first hold the base MLP's df (fp32, pitch DFP)
  float inv[32];
  int64_t row[32];
};

__device__ __forceinline__ void ldsm_x4_trans(uint32_t (&r)[4], const __half *p) {
  uint32_t a = (uint32_t)__cvta_generic_to_shared(p);
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];\n"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(a));
}
__device__ __forceinline__ void ldsm_x2_trans(uint32_t &r0, uint32_t &r1, const __half *p) {
  uint32_t a = (uint32_t)__cvta_generic_to_shared(p);
  asm volatile("ldmatrix.sync.aligned.m8n8.x2.trans.shared.b16 {%0,%1}, [%2];\n" : "=r"(r0), "=r"(r1) : "r"(a));
}

// dw[m][n] += sum over the CTA's 128 staged rows of At[row][m] * Bt[row][n] for the 16 outputs m of tile `mo`
// (m < m_valid, n < n_valid); column n == n_valid of Bt is the constant 1: that column of the product is the bias
// gradient.  A0 / B0 point into the first warp's stage, the other stages follow at sizeof(StageB).  The calling warp OWNS
// the tile's rows of dw: plain read-modify-write, no atomics (float atomics on shared memory are CAS loops).
__device__ __forceinline__ void dw_coop(float *dw, float *dbias, int m_valid, int n_valid, const __half *A0, int pa, int mo,
                                        const __half *B0, int pb, int n_tiles, int lane) {
  int g = lane >> 2, t = lane & 3, id = lane >> 3, r = lane & 7;
  uint32_t a[2 * CM_WARPS][4];
#pragma unroll
  for (int sg = 0; sg < CM_WARPS; ++sg) {
    const __half *At = reinterpret_cast<const __half *>(reinterpret_cast<const char *>(A0) + sg * sizeof(StageB));
#pragma unroll
    for (int kt = 0; kt < 2; ++kt) ldsm_x4_trans(a[2 * sg + kt], At + (16 * kt + (id >> 1) * 8 + r) * pa + 16 * mo + (id & 1) * 8);
  }
  for (int ni = 0; ni < n_tiles; ++ni) {
    float c[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int sg = 0; sg < CM_WARPS; ++sg) {
      const __half *Bt = reinterpret_cast<const __half *>(reinterpret_cast<const char *>(B0) + sg * sizeof(StageB));
#pragma unroll
      for (int kt = 0; kt < 2; ++kt) {
        uint32_t b0, b1;
        ldsm_x2_trans(b0, b1, Bt + (16 * kt + (id & 1) * 8 + r) * pb + 8 * ni);
        mma16816(c, a[2 * sg + kt], b0, b1);
      }
    }
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      int m = 16 * mo + g + ((e >> 1) << 3), nn = 8 * ni + 2 * t + (e & 1);
      if (m < m_valid) {
        if (nn < n_valid) dw[m * n_valid + nn] += c[e];
        else if (nn == n_valid) dbias[m] += c[e];
      }
    }
  }
}

// One MLP, one 16-row tile: recompute the GLU layer from the input fragments A, back-propagate dO (tile `dos`) to
// din = d loss / d input (C layout, 6 column tiles), leave g (GLU output) and dH as fp16 tiles for the dW pass.
__device__ __forceinline__ void bwd_model(const __half *__restrict__ w1, const float *__restrict__ b1,
                                          const __half *__restrict__ w1T, const __half *__restrict__ w2T,
                                          const __half *__restrict__ dos, __half *__restrict__ gs, __half *__restrict__ dhs,
                                          int r0, int g, int t, const uint32_t (&A)[3][4], float (&din)[6][4]) {
  uint32_t DO[4];
  load_a(DO, dos, OP, r0, 0, g, t);
  uint32_t DH[4][4];
#pragma unroll
  for (int j = 0; j < 2; ++j) {
#pragma unroll
    for (int q = 0; q < 2; ++q) {
      int nt = 2 * j + q, col = 8 * nt + 2 * t;
      float ca[4] = {b1[col], b1[col + 1], b1[col], b1[col + 1]};
      float cb[4] = {b1[H + col], b1[H + col + 1], b1[H + col], b1[H + col + 1]};
#pragma unroll
      for (int ks = 0; ks < 3; ++ks) {
        uint32_t b0, b1r;
        load_b(b0, b1r, w1, KP, 8 * nt, 16 * ks, g, t);
        mma16816(ca, A[ks], b0, b1r);
        load_b(b0, b1r, w1, KP, H + 8 * nt, 16 * ks, g, t);
        mma16816(cb, A[ks], b0, b1r);
      }
      float dg[4] = {0.f, 0.f, 0.f, 0.f};
      {
        uint32_t b0, b1r;
        load_b(b0, b1r, w2T, OP, 8 * nt, 0, g, t);
        mma16816(dg, DO, b0, b1r);
      }
      float gv[4], da[4], db[4];
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        float a = hr(ca[e]), b = hr(cb[e]);
        float sb = 1.0f / (1.0f + __expf(-b));
        gv[e] = a * sb;
        da[e] = dg[e] * sb;
        db[e] = dg[e] * a * sb * (1.0f - sb);
      }
      *reinterpret_cast<uint32_t *>(gs + (r0 + g) * HP + col) = pack2(gv[0], gv[1]);
      *reinterpret_cast<uint32_t *>(gs + (r0 + g + 8) * HP + col) = pack2(gv[2], gv[3]);
      uint32_t a01 = pack2(da[0], da[1]), a23 = pack2(da[2], da[3]), b01 = pack2(db[0], db[1]), b23 = pack2(db[2], db[3]);
      *reinterpret_cast<uint32_t *>(dhs + (r0 + g) * TP + col) = a01;
      *reinterpret_cast<uint32_t *>(dhs + (r0 + g + 8) * TP + col) = a23;
      *reinterpret_cast<uint32_t *>(dhs + (r0 + g) * TP + H + col) = b01;
      *reinterpret_cast<uint32_t *>(dhs + (r0 + g + 8) * TP + H + col) = b23;
      DH[j][2 * q] = a01; DH[j][2 * q + 1] = a23;
      DH[2 + j][2 * q] = b01; DH[2 + j][2 * q + 1] = b23;
    }
  }
#pragma unroll
  for (int ni = 0; ni < 6; ++ni) {
    din[ni][0] = din[ni][1] = din[ni][2] = din[ni][3] = 0.f;
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) {
      uint32_t b0, b1r;
      load_b(b0, b1r, w1T, TP, 8 * ni, 16 * kk, g, t);
      mma16816(din[ni], DH[kk], b0, b1r);
    }
  }
}


struct ColorGradDev {
  float *w[10];
};

__device__ __forceinline__ float2 unpack2(uint32_t v) { return __half22float2(*reinterpret_cast<__half2 *>(&v)); }

__global__ void __launch_bounds__(CM_THREADS, 1) k_color_bwd(ColorDev m, ColorGradDev gd, int64_t n, const float *__restrict__ row_w,
                                                             int row_w_stride, const float *__restrict__ feature,
                                                             const float *__restrict__ position, float cx, float cy, float cz,
                                                             const float *__restrict__ glo, float specular_weight,
                                                             const float *__restrict__ d_rgb, int d_rgb_stride, float w_spec_reg,
                                                             const float *__restrict__ reg_sums, float *__restrict__ d_feature,
                                                             float *__restrict__ d_glo, int accumulate) {
  extern __shared__ __align__(16) unsigned char smem[];
  unsigned char *sp = smem;
  Weights &W = *reinterpret_cast<Weights *>(sp); sp += (sizeof(Weights) + 15) & ~(size_t)15;
  WeightsT &WT = *reinterpret_cast<WeightsT *>(sp); sp += (sizeof(WeightsT) + 15) & ~(size_t)15;
  GradAcc &acc = *reinterpret_cast<GradAcc *>(sp); sp += (sizeof(GradAcc) + 15) & ~(size_t)15;
  StageB *stages = reinterpret_cast<StageB *>(sp); sp += CM_WARPS * sizeof(StageB);
  uint16_t *s_list = reinterpret_cast<uint16_t *>(sp); sp += CMB_CHUNK * sizeof(uint16_t);
  float *s_glo = reinterpret_cast<float *>(sp); sp += GF * sizeof(float);
  int *s_count = reinterpret_cast<int *>(sp);
  load_weights(W, m);
  {
    uint32_t *z = reinterpret_cast<uint32_t *>(&WT);
    for (int i = threadIdx.x; i < (int)(sizeof(WeightsT) / 4); i += CM_THREADS) z[i] = 0u;
    float *za = reinterpret_cast<float *>(&acc);
    for (int i = threadIdx.x; i < (int)(sizeof(GradAcc) / 4); i += CM_THREADS) za[i] = 0.f;
    __syncthreads();
    for (int i = threadIdx.x; i < 2 * H * F; i += CM_THREADS) {
      int o = i / F, k = i - o * F;
      WT.w1bT[k * TP + o] = __float2half_rn(m.w[0][i]);
      WT.w1dT[k * TP + o] = __float2half_rn(m.w[6][i]);
    }
    for (int i = threadIdx.x; i < 4 * H; i += CM_THREADS) {
      int o = i / H, k = i - o * H;
      WT.w2bT[k * OP + o] = __float2half_rn(m.w[2][i]);
      WT.w2dT[k * OP + o] = __float2half_rn(m.w[8][i]);
    }
  }
  if (threadIdx.x < GF) s_glo[threadIdx.x] = glo[threadIdx.x];
  __syncthreads();
  // specular regulariser (reference scene/mlp_scene.py:262,267): weight * mean_i (w_i * sum_c |specular_ic|), specular > 0
  float spec_scale = 0.f;
  if (reg_sums && w_spec_reg != 0.f) {
    float cnt = reg_sums[3];
    spec_scale = cnt > 0.f ? w_spec_reg / cnt : 0.f;
  }
  const float cam[3] = {cx, cy, cz};
  int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  StageB &st = stages[warp];
  for (int64_t base = (int64_t)blockIdx.x * CMB_CHUNK; base < n; base += (int64_t)gridDim.x * CMB_CHUNK) {
    int cnt = compact_chunk<CMB_CHUNK>(base, n, row_w, row_w_stride, s_list, s_count);
    // rounds of 128 rows (32 per warp); every warp takes part in every round: the weight-gradient pass is cooperative
    for (int kr = 0; kr < cnt; kr += CM_WARPS * 32) {
      int k0 = kr + warp * 32;
      bool valid = k0 + lane < cnt;
      int64_t i = valid ? base + s_list[k0 + lane] : 0;
      float inv = stage_row(st, lane, valid, i, feature, position, cam, s_glo);
      {  // constant-1 columns: bias gradients come out of the dW products
        __half2 one = __floats2half2_rn(1.f, 0.f), zero = __floats2half2_rn(0.f, 0.f);
        __half2 *px = reinterpret_cast<__half2 *>(st.xh + lane * KP + F);
        px[0] = one; px[1] = zero; px[2] = zero; px[3] = zero;
        *reinterpret_cast<__half2 *>(st.yh + lane * KP + S) = one;
        __half2 *pg = reinterpret_cast<__half2 *>(st.gs + lane * HP + H);
        pg[0] = one; pg[1] = zero; pg[2] = zero; pg[3] = zero;
      }
      st.inv[lane] = inv;
      st.row[lane] = valid ? i : (int64_t)-1;
      float dcol[2][3] = {{0.f, 0.f, 0.f}, {0.f, 0.f, 0.f}};
      if (valid) {
        const float *dr = d_rgb + (int64_t)d_rgb_stride * i;
        float wr = row_w[(int64_t)row_w_stride * i];
#pragma unroll
        for (int c = 0; c < 3; ++c) {
          dcol[0][c] = GRAD_SCALE * dr[c];
          dcol[1][c] = GRAD_SCALE * (specular_weight * dr[c] + spec_scale * wr);
        }
      }
      __syncwarp();
      // forward (outputs only), both 16-row tiles at once like k_color_fwd
      {
        uint32_t A[2][3][4];
#pragma unroll
        for (int mt = 0; mt < 2; ++mt)
#pragma unroll
          for (int ks = 0; ks < 3; ++ks) load_a(A[mt][ks], st.xh, KP, 16 * mt, 16 * ks, g, t);
        mlp_tile2(W.w1b, W.b1b, W.w2b, W.b2b, st.o[0], g, t, A);
        affine_tile2(W, st, g, t, A);
        mlp_tile2(W.w1d, W.b1d, W.w2d, W.b2d, st.o[1], g, t, A);
      }
      __syncwarp();
      // lane = row: gradient at the MLP outputs through luminance_activation
#pragma unroll
      for (int md = 0; md < 2; ++md) {
        const float *o = st.o[md] + lane * 8;
        float lum = expf(o[0] - (md ? 2.0f : 0.0f));
        float d0 = 0.f, dc[3];
#pragma unroll
        for (int c = 0; c < 3; ++c) {
          float sg = 1.0f / (1.0f + expf(-o[1 + c]));
          d0 += dcol[md][c] * sg * lum;
          dc[c] = dcol[md][c] * sg * (1.0f - sg) * lum;
        }
        d0 = fminf(fmaxf(d0, -HALF_SAT), HALF_SAT);  // saturate instead of overflowing to inf (pathological gradients only)
#pragma unroll
        for (int c = 0; c < 3; ++c) dc[c] = fminf(fmaxf(dc[c], -HALF_SAT), HALF_SAT);
        __half2 *pd = reinterpret_cast<__half2 *>(st.dos[md] + lane * OP);
        pd[0] = __floats2half2_rn(d0, dc[0]);
        pd[1] = __floats2half2_rn(dc[1], dc[2]);
#pragma unroll
        for (int q = 2; q < 8; ++q) pd[q] = __floats2half2_rn(0.f, 0.f);
      }
      __syncwarp();
      float *dfs = reinterpret_cast<float *>(st.des);
      // base MLP
#pragma unroll 1
      for (int mt = 0; mt < 2; ++mt) {
        int r0 = 16 * mt;
        uint32_t A[3][4];
#pragma unroll
        for (int ks = 0; ks < 3; ++ks) load_a(A[ks], st.xh, KP, r0, 16 * ks, g, t);
        float din[6][4];
        bwd_model(W.w1b, W.b1b, WT.w1bT, WT.w2bT, st.dos[0], st.gs, st.dhs, r0, g, t, A, din);
#pragma unroll
        for (int ni = 0; ni < 6; ++ni) {
          *reinterpret_cast<float2 *>(dfs + (r0 + g) * DFP + 8 * ni + 2 * t) = make_float2(din[ni][0], din[ni][1]);
          *reinterpret_cast<float2 *>(dfs + (r0 + g + 8) * DFP + 8 * ni + 2 * t) = make_float2(din[ni][2], din[ni][3]);
        }
      }
      __syncthreads();
      dw_coop(acc.w1b, acc.b1b, 2 * H, F, stages[0].dhs, TP, warp, stages[0].xh, KP, 7, lane);
      if (warp == 0) dw_coop(acc.w2b, acc.b2b, 4, H, stages[0].dos[0], OP, 0, stages[0].gs, HP, 5, lane);
      __syncthreads();
      // directional MLP
#pragma unroll 1
      for (int mt = 0; mt < 2; ++mt) {
        int r0 = 16 * mt;
        uint32_t AX[3][4], av[6][2];
        affine_tile<true>(W, st, r0, g, t, AX, av);
        float dx[6][4];
        bwd_model(W.w1d, W.b1d, WT.w1dT, WT.w2dT, st.dos[1], st.gs, st.dhs, r0, g, t, AX, dx);
        float df[6][4];
#pragma unroll
        for (int ni = 0; ni < 6; ++ni) {
          float2 lo = *reinterpret_cast<const float2 *>(dfs + (r0 + g) * DFP + 8 * ni + 2 * t);
          float2 hi = *reinterpret_cast<const float2 *>(dfs + (r0 + g + 8) * DFP + 8 * ni + 2 * t);
          df[ni][0] = lo.x; df[ni][1] = lo.y; df[ni][2] = hi.x; df[ni][3] = hi.y;
        }
        __syncwarp();  // every lane has read the base MLP's df rows before dE overwrites the same bytes
        float s1[2] = {0.f, 0.f}, s2[2] = {0.f, 0.f};
#pragma unroll
        for (int ni = 0; ni < 6; ++ni) {
          int col = 8 * ni + 2 * t;
          float2 flo = *reinterpret_cast<const float2 *>(st.xf + (r0 + g) * KP + col);
          float2 fhi = *reinterpret_cast<const float2 *>(st.xf + (r0 + g + 8) * KP + col);
          float2 alo = unpack2(av[ni][0]), ahi = unpack2(av[ni][1]);
          float fv[4] = {flo.x, flo.y, fhi.x, fhi.y}, aa[4] = {alo.x, alo.y, ahi.x, ahi.y};
          float da[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            df[ni][e] += dx[ni][e] * aa[e];
            da[e] = dx[ni][e] * fv[e];
            s1[e >> 1] += df[ni][e];
            s2[e >> 1] += df[ni][e] * fv[e];
          }
          *reinterpret_cast<uint32_t *>(st.des + (r0 + g) * EP + col) = pack2(da[0], da[1]);
          *reinterpret_cast<uint32_t *>(st.des + (r0 + g + 8) * EP + col) = pack2(da[2], da[3]);
          *reinterpret_cast<uint32_t *>(st.des + (r0 + g) * EP + F + col) = pack2(dx[ni][0], dx[ni][1]);
          *reinterpret_cast<uint32_t *>(st.des + (r0 + g + 8) * EP + F + col) = pack2(dx[ni][2], dx[ni][3]);
        }
        // LayerNorm (no affine) backward: dv = (df - mean(df) - f mean(df f)) / sigma
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          s1[h] += __shfl_xor_sync(GS_FULL, s1[h], 1); s1[h] += __shfl_xor_sync(GS_FULL, s1[h], 2);
          s2[h] += __shfl_xor_sync(GS_FULL, s2[h], 1); s2[h] += __shfl_xor_sync(GS_FULL, s2[h], 2);
          s1[h] *= (1.0f / F); s2[h] *= (1.0f / F);
        }
        float rinv[2] = {st.inv[r0 + g] * GRAD_UNSCALE, st.inv[r0 + g + 8] * GRAD_UNSCALE};
        int64_t rrow[2] = {st.row[r0 + g], st.row[r0 + g + 8]};
#pragma unroll
        for (int ni = 0; ni < 6; ++ni) {
          int col = 8 * ni + 2 * t;
          float2 flo = *reinterpret_cast<const float2 *>(st.xf + (r0 + g) * KP + col);
          float2 fhi = *reinterpret_cast<const float2 *>(st.xf + (r0 + g + 8) * KP + col);
          float fv[4] = {flo.x, flo.y, fhi.x, fhi.y}, dv[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) dv[e] = (df[ni][e] - s1[e >> 1] - fv[e] * s2[e >> 1]) * rinv[e >> 1];
          if (ni < PF / 8) {
#pragma unroll
            for (int h = 0; h < 2; ++h) {
              if (rrow[h] < 0) continue;
              float2 *dst = reinterpret_cast<float2 *>(d_feature + rrow[h] * PF + col);
              float2 v = make_float2(dv[2 * h], dv[2 * h + 1]);
              if (accumulate) { float2 old = *dst; v.x += old.x; v.y += old.y; }
              *dst = v;
            }
          } else {
            float c0 = dv[0] + dv[2], c1 = dv[1] + dv[3];  // column sums over the tile's rows -> d glo
#pragma unroll
            for (int o = 4; o < 32; o <<= 1) { c0 += __shfl_xor_sync(GS_FULL, c0, o); c1 += __shfl_xor_sync(GS_FULL, c1, o); }
            if (g == 0) { atomicAdd(acc.glo + col - PF, c0); atomicAdd(acc.glo + col + 1 - PF, c1); }
          }
        }
      }
      __syncthreads();
      dw_coop(acc.w1d, acc.b1d, 2 * H, F, stages[0].dhs, TP, warp, stages[0].xh, KP, 7, lane);
      dw_coop(acc.wsh, acc.bsh, 2 * F, S, stages[0].des, EP, warp, stages[0].yh, KP, 5, lane);
      if (warp < 2) dw_coop(acc.wsh, acc.bsh, 2 * F, S, stages[0].des, EP, CM_WARPS + warp, stages[0].yh, KP, 5, lane);
      if (warp == 3) dw_coop(acc.w2d, acc.b2d, 4, H, stages[0].dos[1], OP, 0, stages[0].gs, HP, 5, lane);
      __syncthreads();
    }
    __syncthreads();
  }
  __syncthreads();
  // one flush of the CTA's accumulators
  const float *srcs[10] = {acc.w1b, acc.b1b, acc.w2b, acc.b2b, acc.wsh, acc.bsh, acc.w1d, acc.b1d, acc.w2d, acc.b2d};
  const int sizes[10] = {2 * H * F, 2 * H, 4 * H, 4, 2 * F * S, 2 * F, 2 * H * F, 2 * H, 4 * H, 4};
#pragma unroll
  for (int k = 0; k < 10; ++k)
    for (int i = threadIdx.x; i < sizes[k]; i += CM_THREADS) {
      float v = srcs[k][i] * GRAD_UNSCALE;
      if (v != 0.f) atomicAdd(gd.w[k] + i, v);
    }
  if (threadIdx.x < GF && acc.glo[threadIdx.x] != 0.f) atomicAdd(d_glo + threadIdx.x, acc.glo[threadIdx.x]);  // dv is unscaled
}

size_t color_bwd_smem_bytes() {
  auto al = [](size_t x) { return (x + 15) & ~(size_t)15; };
  return al(sizeof(Weights)) + al(sizeof(WeightsT)) + al(sizeof(GradAcc)) + CM_WARPS * sizeof(StageB) +
         CMB_CHUNK * sizeof(uint16_t) + GF * sizeof(float) + 16;
}

}  // namespace

extern "C" int gs_color_mlp_fwd(const gs_color_mlp *h_model, int64_t n, const float *in_view, int in_view_stride,
                                const float *feature, const float *position, const float *h_cam_pos, const float *glo,
                                float specular_weight, float *rgb, float *specular, void *stream) {
  int rc = color_check(h_model);
  if (rc != GS_OK) return rc;
  GS_REQUIRE(feature && position && h_cam_pos && glo && rgb, "null argument");
  if (n <= 0) return GS_OK;
  ColorDev m;
  for (int k = 0; k < 10; ++k) m.w[k] = h_model->weights[k];
  static bool configured = false;
  size_t smem = color_smem_bytes();
  if (!configured) {
    GS_CUDA(cudaFuncSetAttribute(k_color_fwd, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    configured = true;
  }
  int64_t chunks = (n + (int64_t)FW_WARPS * FW_SLICE - 1) / ((int64_t)FW_WARPS * FW_SLICE);
  int grid = (int)(chunks < 148 ? chunks : 148);  // persistent: one CTA of 12 warps per SM (212 KB of shared memory)
  k_color_fwd<<<grid, FW_THREADS, smem, (cudaStream_t)stream>>>(m, n, in_view, in_view_stride, feature, position, h_cam_pos[0],
                                                                h_cam_pos[1], h_cam_pos[2], glo, specular_weight, rgb, specular);
  GS_LAUNCH_CHECK();
  return GS_OK;
}

extern "C" int gs_color_mlp_bwd(const gs_color_mlp *h_model, const gs_color_mlp *h_grads, int64_t n, const float *row_w,
                                int row_w_stride, const float *feature, const float *position, const float *h_cam_pos,
                                const float *glo, float specular_weight, const float *d_rgb, int d_rgb_stride,
                                float w_specular_reg, const float *reg_sums, float *d_feature, float *d_glo, int accumulate,
                                void *stream) {
  int rc = color_check(h_model);
  if (rc != GS_OK) return rc;
  rc = color_check(h_grads);
  if (rc != GS_OK) return rc;
  GS_REQUIRE(row_w && feature && position && h_cam_pos && glo && d_rgb && d_feature && d_glo, "null argument");
  GS_REQUIRE(d_rgb_stride >= 3, "d_rgb stride must be >= 3");
  if (n <= 0) return GS_OK;
  ColorDev m;
  ColorGradDev gd;
  for (int k = 0; k < 10; ++k) { m.w[k] = h_model->weights[k]; gd.w[k] = const_cast<float *>(h_grads->weights[k]); }
  static bool configured = false;
  size_t smem = color_bwd_smem_bytes();
  if (!configured) {
    GS_CUDA(cudaFuncSetAttribute(k_color_bwd, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    configured = true;
  }
  int64_t chunks = (n + CMB_CHUNK - 1) / CMB_CHUNK;
  int grid = (int)(chunks < 148 ? chunks : 148);  // persistent, one CTA per SM (218 KB of shared memory)
  k_color_bwd<<<grid, CM_THREADS, smem, (cudaStream_t)stream>>>(m, gd, n, row_w, row_w_stride, feature, position, h_cam_pos[0],
                                                                h_cam_pos[1], h_cam_pos[2], glo, specular_weight, d_rgb,
                                                                d_rgb_stride, w_specular_reg, reg_sums, d_feature, d_glo, accumulate);
  GS_LAUNCH_CHECK();
  return GS_OK;
}

// Dense Adam for the small replicated state around the colour path (ten weight arrays = 10 088 floats, one GLO row = 32
// floats): torch.optim.Adam's update (reference ColorModel.optimizer, scene/color_model.py:162-170, and the GLO table's
// SparseAdam on the rows an image touched, :37-44) in ONE launch instead of ~10 elementwise ones.
namespace {
__global__ void __launch_bounds__(256) k_adam(int64_t n, float *__restrict__ p, float *__restrict__ g, float *__restrict__ m,
                                              float *__restrict__ v, float lr_over_bc1, float b1, float b2, float inv_bc2,
                                              float eps, int zero_grad) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  float gi = g[i];
  float mi = b1 * m[i] + (1.0f - b1) * gi;
  float vi = b2 * v[i] + (1.0f - b2) * gi * gi;
  m[i] = mi; v[i] = vi;
  p[i] -= lr_over_bc1 * mi / (sqrtf(vi * inv_bc2) + eps);
  if (zero_grad) g[i] = 0.f;
}
}  // namespace

namespace {
// Adam on the rows of a small table that were touched since the last step (the GLO embeddings, reference GLOTable's
// SparseAdam, scene/color_model.py:37-44): per-row step counts and the touched flags live on the device, so neither a
// host-built index list nor a read-back is needed (under DP the flags are all-reduced with the gradients).
__global__ void __launch_bounds__(64) k_adam_rows(int dim, float *__restrict__ p, float *__restrict__ g, float *__restrict__ m,
                                                  float *__restrict__ v, int32_t *__restrict__ steps, float *__restrict__ touched,
                                                  float lr, float b1, float b2, float eps) {
  int r = blockIdx.x;
  if (!(touched[r] > 0.f)) return;
  int t = steps[r] + 1;
  float lr_over_bc1 = lr / (1.0f - powf(b1, (float)t)), inv_bc2 = 1.0f / (1.0f - powf(b2, (float)t));
  for (int k = threadIdx.x; k < dim; k += blockDim.x) {
    int64_t i = (int64_t)r * dim + k;
    float gi = g[i];
    float mi = b1 * m[i] + (1.0f - b1) * gi, vi = b2 * v[i] + (1.0f - b2) * gi * gi;
    m[i] = mi; v[i] = vi;
    p[i] -= lr_over_bc1 * mi / (sqrtf(vi * inv_bc2) + eps);
    g[i] = 0.f;
  }
  __syncthreads();
  if (threadIdx.x == 0) { steps[r] = t; touched[r] = 0.f; }
}
}  // namespace

extern "C" int gs_adam_rows(int rows, int dim, float *param, float *grad, float *exp_avg, float *exp_avg_sq, int32_t *steps,
                            float *touched, float lr, float beta1, float beta2, float eps, void *stream) {
  GS_REQUIRE(param && grad && exp_avg && exp_avg_sq && steps && touched, "null argument");
  if (rows <= 0 || dim <= 0) return GS_OK;
  k_adam_rows<<<rows, 64, 0, (cudaStream_t)stream>>>(dim, param, grad, exp_avg, exp_avg_sq, steps, touched, lr, beta1, beta2, eps);
  GS_LAUNCH_CHECK();
  return GS_OK;
}

extern "C" int gs_adam_step(int64_t n, float *param, float *grad, float *exp_avg, float *exp_avg_sq, float lr, float beta1,
                            float beta2, float eps, int step, int zero_grad, void *stream) {
  GS_REQUIRE(param && grad && exp_avg && exp_avg_sq, "null argument");
  GS_REQUIRE(step >= 1, "step counts from 1");
  if (n <= 0) return GS_OK;
  double bc1 = 1.0 - pow((double)beta1, step), bc2 = 1.0 - pow((double)beta2, step);
  k_adam<<<gs_div_up(n, 256), 256, 0, (cudaStream_t)stream>>>(n, param, grad, exp_avg, exp_avg_sq, (float)(lr / bc1), beta1, beta2,
                                                            (float)(1.0 / bc2), eps, zero_grad);
  GS_LAUNCH_CHECK();
  return GS_OK;
}
