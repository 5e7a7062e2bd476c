// reg.cuh — per-row terms of MLPScene.compute_reg / reg_loss (reference scene/mlp_scene.py:246-288, util/misc.py:68-69
// saturate) and their gradients; shared by the standalone pass (controller.cu) and the fused projection backward.
#pragma once
#include "common.cuh"

struct RegTerms {
  float ns, sat, dsat, aspect, inv_d, s2[3];
  int kmax, kmin;
};

__device__ __forceinline__ RegTerms reg_terms(float op, float d, const float *__restrict__ ls) {
  RegTerms t;
  float s[3] = {expf(ls[0]), expf(ls[1]), expf(ls[2])};
  t.inv_d = 1.0f / d;
  float id2 = t.inv_d * t.inv_d;
#pragma unroll
  for (int k = 0; k < 3; ++k) t.s2[k] = s[k] * s[k] * id2;
  t.ns = t.s2[0] + t.s2[1] + t.s2[2];
  float e = 1.0f / expf(4.0f * op), u = 1.0f - e;  // saturate(op, gain 4, k 2) = (1 - 1/exp(4 op))^2
  t.sat = u * u;
  t.dsat = 8.0f * u * e;
  t.kmax = 0; t.kmin = 0;
#pragma unroll
  for (int k = 1; k < 3; ++k) {
    if (s[k] > s[t.kmax]) t.kmax = k;
    if (s[k] < s[t.kmin]) t.kmin = k;
  }
  t.aspect = s[t.kmax] / s[t.kmin];
  return t;
}

// fused form inside gs_project_bwd: weights + the device pointers the row needs (sums == nullptr: off)
struct RegFuse {
  float w_scale, w_opacity, w_aspect;
  const float *sums, *depth;
};

struct RegGrad {
  float d_op, d_depth, d_ls[3];
};

// gradients of Σ_k weight_k · mean_k for one blended row, c = w / m (w: the row's visibility, m: blended rows)
__device__ __forceinline__ RegGrad reg_grad(const RegTerms &t, float c, float w_scale, float w_opacity, float w_aspect) {
  RegGrad g;
  float cn = (w_scale + w_opacity * t.sat) * c;  // d total / d norm_scale
  g.d_op = w_opacity * c * t.ns * t.dsat;
  g.d_depth = cn * (-2.0f * t.ns * t.inv_d);
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    float v = cn * 2.0f * t.s2[k];
    if (t.kmax != t.kmin) {
      if (k == t.kmax) v += w_aspect * c * t.aspect;
      if (k == t.kmin) v -= w_aspect * c * t.aspect;
    }
    g.d_ls[k] = v;
  }
  return g;
}
