// raster_common.cuh — helpers shared by the rasteriser kernels (raster.cu: generic gather-staged kernels for any channel
// count; raster_rec.cu: record-staged asynchronous kernels of the training path).
#pragma once
#include "common.cuh"
#include "footprint.cuh"  // det_logf

constexpr int RT = 64;  // threads per tile
constexpr float LOG2E = 1.4426950408889634f;
constexpr float LN2 = 0.6931471805599453f;

__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// ---- warp transpose-reductions: NV per-lane values -> NV totals spread over lanes --------------------
// 12 values, 13 shuffles.  On return lane L (even, !(b2&&b1)) holds the total of value `index`.
__device__ __forceinline__ float multi_reduce12(const float *v, int lane, int &index, bool &valid) {
  bool b4 = lane & 16, b3 = lane & 8, b2 = lane & 4, b1 = lane & 2;
  float a[6];
#pragma unroll
  for (int i = 0; i < 6; ++i) {
    float keep = b4 ? v[6 + i] : v[i], send = b4 ? v[i] : v[6 + i];
    a[i] = keep + __shfl_xor_sync(GS_FULL, send, 16);
  }
  float b[3];
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    float keep = b3 ? a[3 + i] : a[i], send = b3 ? a[i] : a[3 + i];
    b[i] = keep + __shfl_xor_sync(GS_FULL, send, 8);
  }
  float c0, c1;
  {
    float keep = b2 ? b[2] : b[0], send = b2 ? b[0] : b[2];
    c0 = keep + __shfl_xor_sync(GS_FULL, send, 4);
    float keep1 = b2 ? 0.f : b[1], send1 = b2 ? b[1] : 0.f;
    c1 = keep1 + __shfl_xor_sync(GS_FULL, send1, 4);
  }
  float keep = b1 ? c1 : c0, send = b1 ? c0 : c1;
  float d = keep + __shfl_xor_sync(GS_FULL, send, 2);
  d += __shfl_xor_sync(GS_FULL, d, 1);
  index = (b4 ? 6 : 0) + (b3 ? 3 : 0) + (b2 ? 2 : (b1 ? 1 : 0));
  valid = !(b2 && b1) && !(lane & 1);
  return d;
}

// 16 values, 16 shuffles; lane L (even) holds the total of value index = lane >> 1.
__device__ __forceinline__ float multi_reduce16(const float *v, int lane, int &index, bool &valid) {
  bool b4 = lane & 16, b3 = lane & 8, b2 = lane & 4, b1 = lane & 2;
  float a[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    float keep = b4 ? v[8 + i] : v[i], send = b4 ? v[i] : v[8 + i];
    a[i] = keep + __shfl_xor_sync(GS_FULL, send, 16);
  }
  float b[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float keep = b3 ? a[4 + i] : a[i], send = b3 ? a[i] : a[4 + i];
    b[i] = keep + __shfl_xor_sync(GS_FULL, send, 8);
  }
  float c[2];
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    float keep = b2 ? b[2 + i] : b[i], send = b2 ? b[i] : b[2 + i];
    c[i] = keep + __shfl_xor_sync(GS_FULL, send, 4);
  }
  float keep = b1 ? c[1] : c[0], send = b1 ? c[0] : c[1];
  float d = keep + __shfl_xor_sync(GS_FULL, send, 2);
  d += __shfl_xor_sync(GS_FULL, d, 1);
  index = (b4 ? 8 : 0) + (b3 ? 4 : 0) + (b2 ? 2 : 0) + (b1 ? 1 : 0);
  valid = !(lane & 1);
  return d;
}


// ---- staged splat record (GS_REC_FLOATS = 12 floats, 48 B, three 16-byte words) ----------------------------------
//   w0 = {mean.x, mean.y, A', B'}   A' = -log2(e)/2 · conic.a, B' = -log2(e) · conic.b
//   w1 = {C', opacity, pcut, row}   C' = -log2(e)/2 · conic.c, pcut = log2(thr / opacity) (Spec R blend bound)
//   w2 = {r, g, b, ey}              ey = (loosened) y half-extent of the support {alpha >= thr}
// Same arithmetic as the gather staging of raster.cu (stage_splat): both paths take identical per-pixel decisions.
struct SplatRec {
  float4 w0, w1, w2;
};

__device__ __forceinline__ SplatRec make_record(float mx, float my, float ca, float cb, float cc, float op, float thr, int id,
                                                float r, float g, float b) {
  SplatRec R;
  float pcut = (op > 0.f) ? __fmul_rn(det_logf(__fdiv_rn(thr, op)), LOG2E) : 1.0f;
  R.w0 = make_float4(mx, my, __fmul_rn(-0.5f * LOG2E, ca), __fmul_rn(-LOG2E, cb));
  R.w1 = make_float4(__fmul_rn(-0.5f * LOG2E, cc), op, pcut, __int_as_float(id));
  float det = ca * cc - cb * cb;
  float k2 = fmaxf(-2.0f * LN2 * pcut, 0.f);
  float ey = (det > 0.f) ? sqrtf(k2 * ca / det) * 1.001f + 0.01f : 3.0e38f;
  R.w2 = make_float4(r, g, b, ey);
  return R;
}

// ---- asynchronous copy + mbarrier (sm_80+ LDGSTS, completion tracked by an sm_90+ mbarrier: SYNCS) ---------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void cp_async16(uint32_t dst, const void *src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void mbar_init(uint32_t bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
// arrive on the barrier once every cp.async this thread issued so far has landed (counts against the init count)
__device__ __forceinline__ void cp_async_arrive(uint32_t bar) {
  asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("{\n.reg .b64 st;\nmbarrier.arrive.shared::cta.b64 st, [%0];\n}\n" ::"r"(bar) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
      "selp.u32 %0, 1, 0, p;\n"
      "}\n"
      : "=r"(ok)
      : "r"(bar), "r"(parity)
      : "memory");
  return ok != 0u;
}
// wait for the phase of the given parity; a copy that never lands traps (reported as a CUDA error) instead of hanging
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  for (uint32_t spins = 0; !mbar_try_wait(bar, parity); ++spins) {
    if (spins > (1u << 22)) __trap();
  }
}
