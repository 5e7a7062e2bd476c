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
constexpr int CM_THREADS = 128, CM_WARPS = CM_THREADS / 32, CM_CHUNK = 1024;

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
  float o[2][32 * 8];  // outputs of the two MLPs (4 valid columns)
};

__device__ __forceinline__ float hr(float v) { return __half2float(__float2half_rn(v)); }
__device__ __forceinline__ uint32_t pack2(float a, float b) {
  __half2 h = __floats2half2_rn(a, b);
  return *reinterpret_cast<uint32_t *>(&h);
}
__device__ __forceinline__ float sigmoidf(float v) { return 1.0f / (1.0f + __expf(-v)); }

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
  for (int i = threadIdx.x; i < (int)(sizeof(Weights) / 4); i += CM_THREADS) z[i] = 0u;  // the pads must be zero
  __syncthreads();
  for (int i = threadIdx.x; i < 2 * H * F; i += CM_THREADS) {
    int o = i / F, k = i - o * F;
    W.w1b[o * KP + k] = __float2half_rn(m.w[0][i]);
    W.w1d[o * KP + k] = __float2half_rn(m.w[6][i]);
  }
  for (int i = threadIdx.x; i < 2 * F * S; i += CM_THREADS) {
    int o = i / S, k = i - o * S;
    W.wsh[o * KP + k] = __float2half_rn(m.w[4][i]);
  }
  for (int i = threadIdx.x; i < 4 * H; i += CM_THREADS) {
    int o = i / H, k = i - o * H;
    W.w2b[o * HP + k] = __float2half_rn(m.w[2][i]);
    W.w2d[o * HP + k] = __float2half_rn(m.w[8][i]);
  }
  for (int i = threadIdx.x; i < 2 * H; i += CM_THREADS) { W.b1b[i] = hr(m.w[1][i]); W.b1d[i] = hr(m.w[7][i]); }
  for (int i = threadIdx.x; i < 2 * F; i += CM_THREADS) W.bsh[i] = hr(m.w[5][i]);
  if (threadIdx.x < 4) { W.b2b[threadIdx.x] = hr(m.w[3][threadIdx.x]); W.b2d[threadIdx.x] = hr(m.w[9][threadIdx.x]); }
  __syncthreads();
}

// lane = row: LayerNorm of [feature, glo] and the SH coefficients of the view direction into the warp's staging tiles.
// Returns 1 / sigma of the row (the backward needs it).
__device__ __forceinline__ float stage_row(Stage &st, int lane, bool valid, int64_t i, const float *__restrict__ feature,
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

// GLU layer + output layer of one MLP for one 16-row tile: A = input fragments (3 k-steps of 16).
// Leaves the fp16-rounded outputs (4 valid columns) in st.o[which].
__device__ __forceinline__ void mlp_tile(const __half *__restrict__ w1, const float *__restrict__ b1,
                                         const __half *__restrict__ w2, const float *__restrict__ b2, float *__restrict__ out,
                                         int r0, int g, int t, const uint32_t (&A)[3][4]) {
  uint32_t A2[2][4];
#pragma unroll
  for (int j = 0; j < 2; ++j) {
#pragma unroll
    for (int q = 0; q < 2; ++q) {
      int col = 8 * (2 * j + q) + 2 * t;
      float ca[4] = {b1[col], b1[col + 1], b1[col], b1[col + 1]};
      float cb[4] = {b1[H + col], b1[H + col + 1], b1[H + col], b1[H + col + 1]};
#pragma unroll
      for (int ks = 0; ks < 3; ++ks) {
        uint32_t b0, b1r;
        load_b(b0, b1r, w1, KP, 8 * (2 * j + q), 16 * ks, g, t);
        mma16816(ca, A[ks], b0, b1r);
        load_b(b0, b1r, w1, KP, H + 8 * (2 * j + q), 16 * ks, g, t);
        mma16816(cb, A[ks], b0, b1r);
      }
      // nn.GLU on the fp16 layer output: first half * sigmoid(second half), rounded to fp16
      A2[j][2 * q] = pack2(hr(ca[0]) * sigmoidf(hr(cb[0])), hr(ca[1]) * sigmoidf(hr(cb[1])));
      A2[j][2 * q + 1] = pack2(hr(ca[2]) * sigmoidf(hr(cb[2])), hr(ca[3]) * sigmoidf(hr(cb[3])));
    }
  }
  float co[4] = {b2[2 * t], b2[2 * t + 1], b2[2 * t], b2[2 * t + 1]};  // b2 is zero padded to 8
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    uint32_t b0, b1r;
    load_b(b0, b1r, w2, HP, 0, 16 * j, g, t);
    mma16816(co, A2[j], b0, b1r);
  }
  if (t < 2) {
    *reinterpret_cast<float2 *>(out + (r0 + g) * 8 + 2 * t) = make_float2(hr(co[0]), hr(co[1]));
    *reinterpret_cast<float2 *>(out + (r0 + g + 8) * 8 + 2 * t) = make_float2(hr(co[2]), hr(co[3]));
  }
}

// direction encoder + affine modulation for one 16-row tile: AX = fragments of x = f * a + b, (a, b) = Wsh Y + bsh
__device__ __forceinline__ void affine_tile(const Weights &W, const Stage &st, int r0, int g, int t, uint32_t (&AX)[3][4]) {
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
    }
  }
}

// block compaction of the rows of one chunk that have work (in view / blended), like k_project_bwd
__device__ __forceinline__ int compact_chunk(int64_t base, int64_t n, const float *__restrict__ mask, int mask_stride,
                                             int *s_list, int *s_count) {
  if (threadIdx.x == 0) *s_count = 0;
  __syncthreads();
  for (int k = threadIdx.x; k < CM_CHUNK; k += CM_THREADS) {
    int64_t i = base + k;
    if (i < n && (!mask || mask[(int64_t)mask_stride * i] > 0.f)) s_list[atomicAdd(s_count, 1)] = k;
  }
  __syncthreads();
  return *s_count;
}

__global__ void __launch_bounds__(CM_THREADS) k_color_fwd(ColorDev m, int64_t n, const float *__restrict__ in_view,
                                                          int in_view_stride, const float *__restrict__ feature,
                                                          const float *__restrict__ position, float cx, float cy, float cz,
                                                          const float *__restrict__ glo, float specular_weight,
                                                          float *__restrict__ rgb, float *__restrict__ specular) {
  extern __shared__ __align__(16) unsigned char smem[];
  Weights &W = *reinterpret_cast<Weights *>(smem);
  Stage *stages = reinterpret_cast<Stage *>(smem + ((sizeof(Weights) + 15) & ~(size_t)15));
  int *s_list = reinterpret_cast<int *>(stages + CM_WARPS);
  float *s_glo = reinterpret_cast<float *>(s_list + CM_CHUNK);
  int *s_count = reinterpret_cast<int *>(s_glo + GF);
  load_weights(W, m);
  if (threadIdx.x < GF) s_glo[threadIdx.x] = glo[threadIdx.x];
  __syncthreads();
  const float cam[3] = {cx, cy, cz};
  int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  Stage &st = stages[warp];
  for (int64_t base = (int64_t)blockIdx.x * CM_CHUNK; base < n; base += (int64_t)gridDim.x * CM_CHUNK) {
    int cnt = compact_chunk(base, n, in_view, in_view_stride, s_list, s_count);
    for (int k0 = warp * 32; k0 < cnt; k0 += CM_WARPS * 32) {
      bool valid = k0 + lane < cnt;
      int64_t i = valid ? base + s_list[k0 + lane] : 0;
      stage_row(st, lane, valid, i, feature, position, cam, s_glo);
      __syncwarp();
#pragma unroll 1
      for (int mt = 0; mt < 2; ++mt) {
        int r0 = 16 * mt;
        if (k0 + r0 >= cnt) break;
        uint32_t A[3][4];
#pragma unroll
        for (int ks = 0; ks < 3; ++ks) load_a(A[ks], st.xh, KP, r0, 16 * ks, g, t);
        mlp_tile(W.w1b, W.b1b, W.w2b, W.b2b, st.o[0], r0, g, t, A);
        affine_tile(W, st, r0, g, t, A);
        mlp_tile(W.w1d, W.b1d, W.w2d, W.b2d, st.o[1], r0, g, t, A);
      }
      __syncwarp();
      if (valid) {
        const float *o = st.o[0] + lane * 8, *o2 = st.o[1] + lane * 8;
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
    __syncthreads();
  }
}

size_t color_smem_bytes() {
  return ((sizeof(Weights) + 15) & ~(size_t)15) + CM_WARPS * sizeof(Stage) + CM_CHUNK * sizeof(int) + GF * sizeof(float) + 16;
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
  int64_t chunks = (n + CM_CHUNK - 1) / CM_CHUNK;
  int grid = (int)(chunks < 148 * 2 ? chunks : 148 * 2);  // persistent: 2 CTAs (97 KB of shared memory each) per SM
  k_color_fwd<<<grid, CM_THREADS, smem, (cudaStream_t)stream>>>(m, n, in_view, in_view_stride, feature, position, h_cam_pos[0],
                                                                h_cam_pos[1], h_cam_pos[2], glo, specular_weight, rgb, specular);
  GS_LAUNCH_CHECK();
  return GS_OK;
}
