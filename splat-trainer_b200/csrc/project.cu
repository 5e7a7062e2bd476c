// project.cu — 3D->2D projection (EWA covariance, conic, opacity) fused with SH colour, and its
// backward.  Serves taichi_splatting.project_to_image / evaluate_sh_at call sites (reference
// scene/mlp_scene.py:415,375; scene/transfer_sh.py:49) — SURVEY §8 a6, a11.
//
// One thread per Gaussian, geometry read as SoA attribute planes (coalesced 12/12/16/4-byte rows),
// SH read as one 16-B-aligned row per in-view Gaussian only (12 x LDG.128 for degree 3), so the
// kernel's DRAM traffic is 44·N + (192 + 48)·V bytes forward.  HBM-bound; no tensor cores.
#include "common.cuh"
#include "footprint.cuh"
#include "reg.cuh"
#include "raster_common.cuh"  // make_record: the rasteriser's staged splat record

// ---- SH basis: reference scene/mlp/rsh.py:110-157 (rsh_cart_3) ------------------------------------
template <int DEG>
__device__ __forceinline__ void sh_basis(float x, float y, float z, float *Y) {
  Y[0] = 0.282094791773878f;
  if (DEG >= 1) {
    Y[1] = -0.48860251190292f * y;
    Y[2] = 0.48860251190292f * z;
    Y[3] = -0.48860251190292f * x;
  }
  if (DEG >= 2) {
    float x2 = x * x, y2 = y * y, z2 = z * z;
    Y[4] = 1.09254843059208f * x * y;
    Y[5] = -1.09254843059208f * y * z;
    Y[6] = 0.94617469575756f * z2 - 0.31539156525252f;
    Y[7] = -1.09254843059208f * x * z;
    Y[8] = 0.54627421529604f * x2 - 0.54627421529604f * y2;
    if (DEG >= 3) {
      Y[9] = -0.590043589926644f * y * (3.0f * x2 - y2);
      Y[10] = 2.89061144264055f * x * y * z;
      Y[11] = 0.304697199642977f * y * (1.5f - 7.5f * z2);
      Y[12] = 1.24392110863372f * z * (1.5f * z2 - 0.5f) - 0.497568443453487f * z;
      Y[13] = 0.304697199642977f * x * (1.5f - 7.5f * z2);
      Y[14] = 1.44530572132028f * z * (x2 - y2);
      Y[15] = -0.590043589926644f * x * (x2 - 3.0f * y2);
    }
  }
}

// g += w * grad Y_k(x,y,z) for the polynomials above
template <int DEG>
__device__ __forceinline__ void sh_basis_grad_acc(float x, float y, float z, const float *w, float &gx, float &gy, float &gz) {
  if (DEG >= 1) {
    const float c1 = 0.48860251190292f;
    gy += -c1 * w[1]; gz += c1 * w[2]; gx += -c1 * w[3];
  }
  if (DEG >= 2) {
    const float c2 = 1.09254843059208f;
    gx += c2 * y * w[4]; gy += c2 * x * w[4];
    gy += -c2 * z * w[5]; gz += -c2 * y * w[5];
    gz += 2.0f * 0.94617469575756f * z * w[6];
    gx += -c2 * z * w[7]; gz += -c2 * x * w[7];
    gx += 2.0f * 0.54627421529604f * x * w[8]; gy += -2.0f * 0.54627421529604f * y * w[8];
  }
  if (DEG >= 3) {
    float x2 = x * x, y2 = y * y, z2 = z * z;
    const float c9 = 0.590043589926644f, c10 = 2.89061144264055f, c11 = 0.304697199642977f, c14 = 1.44530572132028f;
    gx += -c9 * 6.0f * x * y * w[9]; gy += -c9 * (3.0f * x2 - 3.0f * y2) * w[9];
    gx += c10 * y * z * w[10]; gy += c10 * x * z * w[10]; gz += c10 * x * y * w[10];
    gy += c11 * (1.5f - 7.5f * z2) * w[11]; gz += c11 * y * -15.0f * z * w[11];
    gz += (1.24392110863372f * (4.5f * z2 - 0.5f) - 0.497568443453487f) * w[12];
    gx += c11 * (1.5f - 7.5f * z2) * w[13]; gz += c11 * x * -15.0f * z * w[13];
    gx += c14 * 2.0f * x * z * w[14]; gy += c14 * -2.0f * y * z * w[14]; gz += c14 * (x2 - y2) * w[14];
    gx += -c9 * (3.0f * x2 - 3.0f * y2) * w[15]; gy += c9 * 6.0f * x * y * w[15];
  }
}

// load one SH row (3*K floats) into registers; vector loads when the row is 16-B aligned
template <int K>
__device__ __forceinline__ void load_sh_row(const float *__restrict__ row, float *v) {
  if ((3 * K) % 4 == 0) {
    const float4 *r4 = reinterpret_cast<const float4 *>(row);
#pragma unroll
    for (int i = 0; i < 3 * K / 4; ++i) {
      float4 t = __ldg(r4 + i);
      v[4 * i] = t.x; v[4 * i + 1] = t.y; v[4 * i + 2] = t.z; v[4 * i + 3] = t.w;
    }
  } else {
#pragma unroll
    for (int i = 0; i < 3 * K; ++i) v[i] = __ldg(row + i);
  }
}

template <int DEG>
__device__ __forceinline__ void sh_color(const float *__restrict__ row, float px, float py, float pz, const float *cam_pos, float *rgb) {
  constexpr int K = (DEG + 1) * (DEG + 1);
  float dx = px - cam_pos[0], dy = py - cam_pos[1], dz = pz - cam_pos[2];
  float len = sqrtf(dx * dx + dy * dy + dz * dz);
  float inv = len > 0.f ? 1.0f / len : 0.f;
  float Y[16];
  sh_basis<DEG>(dx * inv, dy * inv, dz * inv, Y);
  float v[3 * K];
  load_sh_row<K>(row, v);
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    float a = 0.5f;
#pragma unroll
    for (int k = 0; k < K; ++k) a += Y[k] * v[c * K + k];
    rgb[c] = a;
  }
}

// ---- shared forward math --------------------------------------------------------------------------
struct Proj {
  float x, y, z, inv_z, tx, ty, txc, tyc;
  float qx, qy, qz, qw, qlen;
  float R[9], s[3], Tm[6], A[6];
  float cxx, cxy, cyy, det0, det1, aa, sig;
  bool clamped_x, clamped_y;
  float u, v, ca, cb, cc, op;
};

__device__ __forceinline__ bool project_one(const CamDev &cam, const CfgDev &cfg, float px, float py, float pz,
                                            float ls0, float ls1, float ls2, float4 q, float al, Proj &t) {
  const float *T = cam.T;
  t.x = T[0] * px + T[1] * py + T[2] * pz + T[3];
  t.y = T[4] * px + T[5] * py + T[6] * pz + T[7];
  t.z = T[8] * px + T[9] * py + T[10] * pz + T[11];
  if (!(t.z > cam.near_plane) || !(t.z < cam.far_plane)) return false;
  t.inv_z = 1.0f / t.z;
  t.tx = t.x * t.inv_z; t.ty = t.y * t.inv_z;
  t.clamped_x = (t.tx < -cam.limx) || (t.tx > cam.limx);
  t.clamped_y = (t.ty < -cam.limy) || (t.ty > cam.limy);
  t.txc = fminf(fmaxf(t.tx, -cam.limx), cam.limx);
  t.tyc = fminf(fmaxf(t.ty, -cam.limy), cam.limy);
  float ql = sqrtf(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w);
  if (!(ql > 0.f)) return false;
  t.qlen = ql;
  float iq = 1.0f / ql;
  float x = q.x * iq, y = q.y * iq, z = q.z * iq, w = q.w * iq;
  t.qx = x; t.qy = y; t.qz = z; t.qw = w;
  t.R[0] = 1.f - 2.f * (y * y + z * z); t.R[1] = 2.f * (x * y - w * z); t.R[2] = 2.f * (x * z + w * y);
  t.R[3] = 2.f * (x * y + w * z); t.R[4] = 1.f - 2.f * (x * x + z * z); t.R[5] = 2.f * (y * z - w * x);
  t.R[6] = 2.f * (x * z - w * y); t.R[7] = 2.f * (y * z + w * x); t.R[8] = 1.f - 2.f * (x * x + y * y);
  t.s[0] = expf(ls0); t.s[1] = expf(ls1); t.s[2] = expf(ls2);
  float J00 = cam.fx * t.inv_z, J02 = -cam.fx * t.txc * t.inv_z;
  float J11 = cam.fy * t.inv_z, J12 = -cam.fy * t.tyc * t.inv_z;
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    t.Tm[k] = J00 * T[k] + J02 * T[8 + k];
    t.Tm[3 + k] = J11 * T[4 + k] + J12 * T[8 + k];
  }
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    t.A[k] = (t.Tm[0] * t.R[k] + t.Tm[1] * t.R[3 + k] + t.Tm[2] * t.R[6 + k]) * t.s[k];
    t.A[3 + k] = (t.Tm[3] * t.R[k] + t.Tm[4] * t.R[3 + k] + t.Tm[5] * t.R[6 + k]) * t.s[k];
  }
  float c00 = t.A[0] * t.A[0] + t.A[1] * t.A[1] + t.A[2] * t.A[2];
  float c01 = t.A[0] * t.A[3] + t.A[1] * t.A[4] + t.A[2] * t.A[5];
  float c11 = t.A[3] * t.A[3] + t.A[4] * t.A[4] + t.A[5] * t.A[5];
  t.det0 = c00 * c11 - c01 * c01;
  t.cxx = c00 + cfg.blur; t.cyy = c11 + cfg.blur; t.cxy = c01;
  t.det1 = t.cxx * t.cyy - t.cxy * t.cxy;
  if (!(t.det1 > 1e-12f)) return false;
  t.sig = 1.0f / (1.0f + expf(-al));
  t.aa = 1.0f;
  if (cfg.antialias) {
    float r = t.det0 / t.det1;
    t.aa = r > 0.f ? sqrtf(r) : 0.f;
  }
  t.op = t.sig * t.aa;
  if (!(t.op > cfg.thr)) return false;
  float inv = 1.0f / t.det1;
  t.u = cam.fx * t.tx + cam.cx;
  t.v = cam.fy * t.ty + cam.cy;
  t.ca = t.cyy * inv; t.cb = -t.cxy * inv; t.cc = t.cxx * inv;
  return true;
}

template <int DEG>
__global__ void __launch_bounds__(128) k_project_fwd(CamDev cam, CfgDev cfg, int64_t n, const float *__restrict__ position,
                                                     const float *__restrict__ log_scaling, const float *__restrict__ rotation,
                                                     const float *__restrict__ alpha_logit, const float *__restrict__ sh,
                                                     float *__restrict__ g2d, float *__restrict__ depth,
                                                     float *__restrict__ rgb, float *__restrict__ acc, int acc_stride,
                                                     uint4 *__restrict__ tile_recs, int32_t *__restrict__ tile_counts,
                                                     float *__restrict__ records) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  float px = position[3 * i], py = position[3 * i + 1], pz = position[3 * i + 2];
  Proj t;
  float4 q = reinterpret_cast<const float4 *>(rotation)[i];
  bool ok = project_one(cam, cfg, px, py, pz, log_scaling[3 * i], log_scaling[3 * i + 1], log_scaling[3 * i + 2], q,
                        alpha_logit[i], t);
  int ntiles = 0;
  if (ok) {
    Footprint fp = row_footprint(t.u, t.v, t.ca, t.cb, t.cc, t.op, t.z, cfg.thr, cam.W, cam.H);
    if (tile_recs) {  // same footprint code as the mapper: gs_tile_map(tile_info_in) reuses the outcome
      uint4 rec;
      ntiles = footprint_record(fp, cam.W, cam.H, rec);
      if (ntiles > 0) tile_recs[i] = rec;
    } else {
      ntiles = footprint_count(fp, cam.W, cam.H);
    }
    ok = ntiles > 0;
  }
  if (tile_counts) tile_counts[i] = ntiles;
  float2 *g = reinterpret_cast<float2 *>(g2d + GS_G2D_STRIDE * i);
  if (!ok) {
    g[0] = make_float2(0.f, 0.f); g[1] = make_float2(1.f, 0.f); g[2] = make_float2(1.f, 0.f);
    depth[i] = __int_as_float(0x7f800000);
    if (rgb) { rgb[3 * i] = 0.f; rgb[3 * i + 1] = 0.f; rgb[3 * i + 2] = 0.f; }
    return;
  }
  g[0] = make_float2(t.u, t.v); g[1] = make_float2(t.ca, t.cb); g[2] = make_float2(t.cc, t.op);
  depth[i] = t.z;
  if (sh && rgb) {
    constexpr int K = (DEG + 1) * (DEG + 1);
    float c[3];
    sh_color<DEG>(sh + (int64_t)3 * K * i, px, py, pz, cam.cam_pos, c);
    rgb[3 * i] = c[0]; rgb[3 * i + 1] = c[1]; rgb[3 * i + 2] = c[2];
    if (records) {  // the rasteriser's staged record of this row: written once here, gathered per tile by cp.async
      SplatRec R = make_record(t.u, t.v, t.ca, t.cb, t.cc, t.op, cfg.thr, (int)i, c[0], c[1], c[2]);
      float4 *dst = reinterpret_cast<float4 *>(records + (int64_t)GS_REC_FLOATS * i);
      dst[0] = R.w0; dst[1] = R.w1; dst[2] = R.w2;
    }
  }
  if (acc) {
    float4 *a = reinterpret_cast<float4 *>(acc + (int64_t)acc_stride * i);
    for (int k = 0; k < acc_stride / 4; ++k) a[k] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
}

// one row of the projection / SH backward
template <int DEG>
__device__ __forceinline__ void project_bwd_row(const CamDev &cam, const CfgDev &cfg, int64_t i, const float *__restrict__ position,
                                                const float *__restrict__ log_scaling, const float *__restrict__ rotation,
                                                const float *__restrict__ alpha_logit, const float *__restrict__ sh,
                                                const float *__restrict__ d_g2d, int d_g2d_stride,
                                                const float *__restrict__ d_rgb, int d_rgb_stride,
                                                const float *__restrict__ d_depth, const RegFuse &reg, float row_w,
                                                const float *__restrict__ g2d, float *__restrict__ g_position,
                                                float *__restrict__ g_log_scaling, float *__restrict__ g_rotation,
                                                float *__restrict__ g_alpha_logit, float *__restrict__ g_sh, int accumulate) {
  // every load of the row is issued up front (rows are scattered over GBs of planes: each access is a full DRAM +
  // TLB round trip, and the few rows with work leave no other warps to hide it behind)
  constexpr int KK = (DEG + 1) * (DEG + 1);
  float px = position[3 * i], py = position[3 * i + 1], pz = position[3 * i + 2];
  float ls0 = log_scaling[3 * i], ls1 = log_scaling[3 * i + 1], ls2 = log_scaling[3 * i + 2];
  float al = alpha_logit[i];
  float4 q = reinterpret_cast<const float4 *>(rotation)[i];
  const float *dg = d_g2d + (int64_t)d_g2d_stride * i;
  float d_u = dg[0], d_v = dg[1], d_ca = dg[2], d_cb = dg[3], d_cc = dg[4], d_op = dg[5];
  float dc_pre[3] = {0.f, 0.f, 0.f};
  float shv[3 * KK];
  if (sh && d_rgb) {
    const float *dr = d_rgb + (int64_t)d_rgb_stride * i;
    dc_pre[0] = dr[0]; dc_pre[1] = dr[1]; dc_pre[2] = dr[2];
    load_sh_row<KK>(sh + (int64_t)3 * KK * i, shv);
  }
  float dd = d_depth ? d_depth[i] : 0.f;
  float reg_op = 0.f, reg_z = 1.f, reg_m = 0.f;
  if (reg.sums) { reg_op = g2d[GS_G2D_STRIDE * i + 5]; reg_z = reg.depth[i]; reg_m = reg.sums[3]; }
  Proj t;
  if (!project_one(cam, cfg, px, py, pz, ls0, ls1, ls2, q, al, t)) return;
  // regularisation gradients of this image (same arithmetic as k_reg_grads): they enter through the 2-D opacity,
  // the depth and log_scaling
  float reg_ls[3] = {0.f, 0.f, 0.f};
  if (reg.sums && reg_m > 0.f && row_w > 0.f) {
    const float lsv[3] = {ls0, ls1, ls2};
    RegGrad rg = reg_grad(reg_terms(reg_op, reg_z, lsv), row_w / reg_m, reg.w_scale, reg.w_opacity, reg.w_aspect);
    d_op += rg.d_op;
    dd += rg.d_depth;
    reg_ls[0] = rg.d_ls[0]; reg_ls[1] = rg.d_ls[1]; reg_ls[2] = rg.d_ls[2];
  }
  // conic -> covariance: S = -C G C
  float Ga = d_ca, Gb = 0.5f * d_cb, Gc = d_cc;
  float M00 = t.ca * Ga + t.cb * Gb, M01 = t.ca * Gb + t.cb * Gc;
  float M10 = t.cb * Ga + t.cc * Gb, M11 = t.cb * Gb + t.cc * Gc;
  float g_cxx = -(M00 * t.ca + M01 * t.cb);
  float g_cxy = -2.0f * (M00 * t.cb + M01 * t.cc);
  float g_cyy = -(M10 * t.cb + M11 * t.cc);
  float g_sig = d_op * t.aa;
  if (cfg.antialias && t.aa > 0.f) {
    float g_aa = d_op * t.sig;
    float k = g_aa / (2.0f * t.aa * t.det1);
    float f2 = t.aa * t.aa;
    g_cxx += k * ((t.cyy - cfg.blur) - f2 * t.cyy);
    g_cyy += k * ((t.cxx - cfg.blur) - f2 * t.cxx);
    g_cxy += k * (-2.0f * t.cxy + f2 * 2.0f * t.cxy);
  }
  float gal = g_sig * t.sig * (1.0f - t.sig);
  float gA[6];
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    gA[k] = 2.0f * g_cxx * t.A[k] + g_cxy * t.A[3 + k];
    gA[3 + k] = 2.0f * g_cyy * t.A[3 + k] + g_cxy * t.A[k];
  }
  float gTm[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
  for (int r = 0; r < 2; ++r)
#pragma unroll
    for (int j = 0; j < 3; ++j)
#pragma unroll
      for (int k = 0; k < 3; ++k) gTm[3 * r + j] += gA[3 * r + k] * t.R[3 * j + k] * t.s[k];
  float gR[9], gls[3];
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    float gs = 0.f;
#pragma unroll
    for (int j = 0; j < 3; ++j) {
      float gM = t.Tm[j] * gA[k] + t.Tm[3 + j] * gA[3 + k];
      gR[3 * j + k] = gM * t.s[k];
      gs += gM * t.R[3 * j + k];
    }
    gls[k] = gs * t.s[k] + reg_ls[k];
  }
  float x = t.qx, y = t.qy, z = t.qz, w = t.qw;
  float gq[4];
  gq[0] = 2.f * (y * gR[1] + z * gR[2] + y * gR[3] - 2.f * x * gR[4] - w * gR[5] + z * gR[6] + w * gR[7] - 2.f * x * gR[8]);
  gq[1] = 2.f * (-2.f * y * gR[0] + x * gR[1] + w * gR[2] + x * gR[3] + z * gR[5] - w * gR[6] + z * gR[7] - 2.f * y * gR[8]);
  gq[2] = 2.f * (-2.f * z * gR[0] - w * gR[1] + x * gR[2] + w * gR[3] - 2.f * z * gR[4] + y * gR[5] + x * gR[6] + y * gR[7]);
  gq[3] = 2.f * (-z * gR[1] + y * gR[2] + z * gR[3] - x * gR[5] - y * gR[6] + x * gR[7]);
  float dotq = gq[0] * x + gq[1] * y + gq[2] * z + gq[3] * w;
  float iql = 1.0f / t.qlen;
  gq[0] = (gq[0] - x * dotq) * iql; gq[1] = (gq[1] - y * dotq) * iql;
  gq[2] = (gq[2] - z * dotq) * iql; gq[3] = (gq[3] - w * dotq) * iql;
  const float *T = cam.T;
  float gJ00 = gTm[0] * T[0] + gTm[1] * T[1] + gTm[2] * T[2];
  float gJ02 = gTm[0] * T[8] + gTm[1] * T[9] + gTm[2] * T[10];
  float gJ11 = gTm[3] * T[4] + gTm[4] * T[5] + gTm[5] * T[6];
  float gJ12 = gTm[3] * T[8] + gTm[4] * T[9] + gTm[5] * T[10];
  float iz = t.inv_z, iz2 = iz * iz;
  float g_z = -gJ00 * cam.fx * iz2 - gJ11 * cam.fy * iz2 + gJ02 * cam.fx * t.txc * iz2 + gJ12 * cam.fy * t.tyc * iz2;
  float g_tx = t.clamped_x ? 0.f : -gJ02 * cam.fx * iz;
  float g_ty = t.clamped_y ? 0.f : -gJ12 * cam.fy * iz;
  g_tx += cam.fx * d_u;
  g_ty += cam.fy * d_v;
  float g_x = g_tx * iz, g_y = g_ty * iz;
  g_z += -(g_tx * t.tx + g_ty * t.ty) * iz;
  g_z += dd;
  float gp[3];
#pragma unroll
  for (int k = 0; k < 3; ++k) gp[k] = T[k] * g_x + T[4 + k] * g_y + T[8 + k] * g_z;

  if (sh && d_rgb) {
    constexpr int K = KK;
    const float dc[3] = {dc_pre[0], dc_pre[1], dc_pre[2]};
    float dx = px - cam.cam_pos[0], dy = py - cam.cam_pos[1], dz = pz - cam.cam_pos[2];
    float len = sqrtf(dx * dx + dy * dy + dz * dz);
    float inv = len > 0.f ? 1.0f / len : 0.f;
    float nx = dx * inv, ny = dy * inv, nz = dz * inv;
    float Y[16];
    sh_basis<DEG>(nx, ny, nz, Y);
    const float *v = shv;
    float wk[16];
#pragma unroll
    for (int k = 0; k < K; ++k) wk[k] = v[k] * dc[0] + v[K + k] * dc[1] + v[2 * K + k] * dc[2];
    float gnx = 0.f, gny = 0.f, gnz = 0.f;
    sh_basis_grad_acc<DEG>(nx, ny, nz, wk, gnx, gny, gnz);
    float dot = gnx * nx + gny * ny + gnz * nz;
    gp[0] += (gnx - nx * dot) * inv; gp[1] += (gny - ny * dot) * inv; gp[2] += (gnz - nz * dot) * inv;
    if (g_sh) {
      float *gr = g_sh + (int64_t)3 * K * i;
      if ((3 * K) % 4 == 0) {
        float4 *g4 = reinterpret_cast<float4 *>(gr);
#pragma unroll
        for (int j = 0; j < 3 * K / 4; ++j) {
          float4 o;
          o.x = Y[(4 * j) % K] * dc[(4 * j) / K];
          o.y = Y[(4 * j + 1) % K] * dc[(4 * j + 1) / K];
          o.z = Y[(4 * j + 2) % K] * dc[(4 * j + 2) / K];
          o.w = Y[(4 * j + 3) % K] * dc[(4 * j + 3) / K];
          if (accumulate) { float4 p = g4[j]; o.x += p.x; o.y += p.y; o.z += p.z; o.w += p.w; }
          g4[j] = o;
        }
      } else {
#pragma unroll
        for (int j = 0; j < 3 * K; ++j) {
          float o = Y[j % K] * dc[j / K];
          gr[j] = accumulate ? gr[j] + o : o;
        }
      }
    }
  }
  if (accumulate) {
#pragma unroll
    for (int k = 0; k < 3; ++k) { g_position[3 * i + k] += gp[k]; g_log_scaling[3 * i + k] += gls[k]; }
    float4 *gq4 = reinterpret_cast<float4 *>(g_rotation) + i;
    float4 p = *gq4;
    *gq4 = make_float4(p.x + gq[0], p.y + gq[1], p.z + gq[2], p.w + gq[3]);
    g_alpha_logit[i] += gal;
  } else {
#pragma unroll
    for (int k = 0; k < 3; ++k) { g_position[3 * i + k] = gp[k]; g_log_scaling[3 * i + k] = gls[k]; }
    reinterpret_cast<float4 *>(g_rotation)[i] = make_float4(gq[0], gq[1], gq[2], gq[3]);
    g_alpha_logit[i] = gal;
  }
}

// Block-compacted backward: a block first collects the rows of its slice that have work (blended by a pixel /
// in view) into shared memory — in a trained scene only a few percent of the rows are — and then its threads walk
// that dense list, so warps stay full instead of dragging one active lane each through ~60 scattered accesses.
constexpr int PB_THREADS = 128;
constexpr int PB_ITEMS = 8;
constexpr int PB_ROWS = PB_THREADS * PB_ITEMS;

template <int DEG>
__global__ void __launch_bounds__(PB_THREADS) k_project_bwd(CamDev cam, CfgDev cfg, int64_t n, const float *__restrict__ position,
                                                            const float *__restrict__ log_scaling, const float *__restrict__ rotation,
                                                            const float *__restrict__ alpha_logit, const float *__restrict__ sh,
                                                            const float *__restrict__ g2d, const float *__restrict__ d_g2d,
                                                            int d_g2d_stride, const float *__restrict__ d_rgb, int d_rgb_stride,
                                                            const float *__restrict__ d_depth, const float *__restrict__ row_vis,
                                                            int row_vis_stride, RegFuse reg, float *__restrict__ g_position,
                                                            float *__restrict__ g_log_scaling, float *__restrict__ g_rotation,
                                                            float *__restrict__ g_alpha_logit, float *__restrict__ g_sh, int accumulate) {
  __shared__ int s_list[PB_ROWS];
  __shared__ int s_count;
  if (threadIdx.x == 0) s_count = 0;
  __syncthreads();
  int64_t base = (int64_t)blockIdx.x * PB_ROWS;
  // scan of the CTA's rows: all loads of the "blended by some pixel" plane are issued before the first test (this sweep of
  // n rows is all the kernel does for most CTAs; at the occupancy the row work below allows, dependent loads one after the
  // other left it latency-bound); the flag of the forward culling is only read for the few rows that pass
  float rv[PB_ITEMS];
#pragma unroll
  for (int k = 0; k < PB_ITEMS; ++k) {
    int64_t i = base + k * PB_THREADS + threadIdx.x;
    rv[k] = (i < n) ? (row_vis ? row_vis[(int64_t)row_vis_stride * i] : 1.f) : 0.f;
  }
#pragma unroll
  for (int k = 0; k < PB_ITEMS; ++k) {
    if (!(rv[k] > 0.f)) continue;  // blended by no pixel
    int64_t i = base + k * PB_THREADS + threadIdx.x;
    if (!(g2d[GS_G2D_STRIDE * i + 5] > 0.f)) continue;  // culled in forward
    s_list[atomicAdd(&s_count, 1)] = (int)(i - base);
  }
  __syncthreads();
  int cnt = s_count;
  for (int k = threadIdx.x; k < cnt; k += PB_THREADS) {
    int64_t i = base + s_list[k];
    float row_w = reg.sums ? row_vis[(int64_t)row_vis_stride * i] : 0.f;
    project_bwd_row<DEG>(cam, cfg, i, position, log_scaling, rotation, alpha_logit, sh, d_g2d, d_g2d_stride, d_rgb, d_rgb_stride,
                         d_depth, reg, row_w, g2d, g_position, g_log_scaling, g_rotation, g_alpha_logit, g_sh, accumulate);
  }
}

// ---- evaluate_sh_at -------------------------------------------------------------------------------
struct Vec3 { float v[3]; };

template <int DEG>
__global__ void __launch_bounds__(128) k_sh_eval_fwd(int64_t nv, const int64_t *__restrict__ indexes, const float *__restrict__ sh,
                                                     const float *__restrict__ position, Vec3 cam_pos, float *__restrict__ rgb) {
  int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= nv) return;
  int64_t i = indexes ? indexes[j] : j;
  constexpr int K = (DEG + 1) * (DEG + 1);
  float c[3];
  sh_color<DEG>(sh + (int64_t)3 * K * i, position[3 * i], position[3 * i + 1], position[3 * i + 2], cam_pos.v, c);
  rgb[3 * j] = c[0]; rgb[3 * j + 1] = c[1]; rgb[3 * j + 2] = c[2];
}

// indexes may repeat rows in principle (torch indexing semantics) -> accumulate with atomics
template <int DEG>
__global__ void __launch_bounds__(128) k_sh_eval_bwd(int64_t nv, const int64_t *__restrict__ indexes, const float *__restrict__ sh,
                                                     const float *__restrict__ position, Vec3 cam_pos,
                                                     const float *__restrict__ d_rgb, float *__restrict__ g_sh,
                                                     float *__restrict__ g_position) {
  int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= nv) return;
  int64_t i = indexes ? indexes[j] : j;
  constexpr int K = (DEG + 1) * (DEG + 1);
  float dc[3] = {d_rgb[3 * j], d_rgb[3 * j + 1], d_rgb[3 * j + 2]};
  float dx = position[3 * i] - cam_pos.v[0], dy = position[3 * i + 1] - cam_pos.v[1], dz = position[3 * i + 2] - cam_pos.v[2];
  float len = sqrtf(dx * dx + dy * dy + dz * dz);
  float inv = len > 0.f ? 1.0f / len : 0.f;
  float nx = dx * inv, ny = dy * inv, nz = dz * inv;
  float Y[16];
  sh_basis<DEG>(nx, ny, nz, Y);
  if (g_sh) {
    float *gr = g_sh + (int64_t)3 * K * i;
#pragma unroll
    for (int k = 0; k < 3 * K; ++k) atomicAdd(gr + k, Y[k % K] * dc[k / K]);
  }
  if (g_position) {
    float v[3 * K];
    load_sh_row<K>(sh + (int64_t)3 * K * i, v);
    float wk[16];
#pragma unroll
    for (int k = 0; k < K; ++k) wk[k] = v[k] * dc[0] + v[K + k] * dc[1] + v[2 * K + k] * dc[2];
    float gnx = 0.f, gny = 0.f, gnz = 0.f;
    sh_basis_grad_acc<DEG>(nx, ny, nz, wk, gnx, gny, gnz);
    float dot = gnx * nx + gny * ny + gnz * nz;
    atomicAdd(g_position + 3 * i, (gnx - nx * dot) * inv);
    atomicAdd(g_position + 3 * i + 1, (gny - ny * dot) * inv);
    atomicAdd(g_position + 3 * i + 2, (gnz - nz * dot) * inv);
  }
}

#define GS_DISPATCH_DEG(deg, CALL)                \
  switch (deg) {                                  \
    case 0: { constexpr int D = 0; CALL; } break; \
    case 1: { constexpr int D = 1; CALL; } break; \
    case 2: { constexpr int D = 2; CALL; } break; \
    case 3: { constexpr int D = 3; CALL; } break; \
    default: gs_set_error("sh_degree must be 0..3"); return GS_ERR_UNSUPPORTED; \
  }

extern "C" int gs_project_fwd(const gs_camera *h_cam, const gs_raster_config *h_cfg, int64_t n, const float *position,
                              const float *log_scaling, const float *rotation, const float *alpha_logit, const float *sh,
                              int sh_degree, float *g2d, float *depth, float *rgb, float *acc, int acc_stride,
                              void *tile_info, float *records, void *stream) {
  GS_REQUIRE(h_cam && h_cfg, "null camera/config");
  GS_REQUIRE(h_cfg->tile_size == GS_TILE, "tile_size must be 16");
  GS_REQUIRE(records == nullptr || (sh && rgb), "staged records need the SH colour (use gs_raster_records otherwise)");
  GS_REQUIRE(acc == nullptr || (acc_stride > 0 && acc_stride % 4 == 0), "acc_stride must be a multiple of 4");
  if (n <= 0) return GS_OK;
  CamDev cam = gs_make_cam(h_cam, h_cfg);
  CfgDev cfg = gs_make_cfg(h_cfg);
  cudaStream_t st = (cudaStream_t)stream;
  int blocks = gs_div_up(n, 128);
  TileInfo ti = {nullptr, nullptr, 0};
  if (tile_info) ti = gs_tile_info_carve(tile_info, n, h_cam->width, h_cam->height);
  GS_DISPATCH_DEG(sh_degree, (k_project_fwd<D><<<blocks, 128, 0, st>>>(cam, cfg, n, position, log_scaling, rotation,
                                                                        alpha_logit, sh, g2d, depth, rgb, acc, acc_stride,
                                                                        ti.recs, ti.counts, records)));
  GS_LAUNCH_CHECK();
  return GS_OK;
}

extern "C" int gs_project_bwd(const gs_camera *h_cam, const gs_raster_config *h_cfg, int64_t n, const float *position,
                              const float *log_scaling, const float *rotation, const float *alpha_logit, const float *sh,
                              int sh_degree, const float *g2d, const float *d_g2d, int d_g2d_stride, const float *d_rgb,
                              int d_rgb_stride, const float *d_depth, const float *row_vis, int row_vis_stride,
                              const gs_reg_fuse *h_reg, float *g_position, float *g_log_scaling, float *g_rotation,
                              float *g_alpha_logit, float *g_sh, int accumulate, void *stream) {
  GS_REQUIRE(h_cam && h_cfg, "null camera/config");
  GS_REQUIRE(d_g2d && d_g2d_stride >= 6, "d_g2d stride must be >= 6");
  RegFuse reg = {0.f, 0.f, 0.f, nullptr, nullptr};
  if (h_reg) {
    GS_REQUIRE(h_reg->sums && h_reg->depth && row_vis && g2d, "fused regularisation needs sums, depth, row_vis and g2d");
    reg.w_scale = h_reg->w_scale; reg.w_opacity = h_reg->w_opacity; reg.w_aspect = h_reg->w_aspect;
    reg.sums = h_reg->sums; reg.depth = h_reg->depth;
  }
  if (n <= 0) return GS_OK;
  CamDev cam = gs_make_cam(h_cam, h_cfg);
  CfgDev cfg = gs_make_cfg(h_cfg);
  cudaStream_t st = (cudaStream_t)stream;
  int blocks = gs_div_up(n, PB_ROWS);
  GS_DISPATCH_DEG(sh_degree, (k_project_bwd<D><<<blocks, PB_THREADS, 0, st>>>(cam, cfg, n, position, log_scaling, rotation,
                                                                        alpha_logit, sh, g2d, d_g2d, d_g2d_stride, d_rgb,
                                                                        d_rgb_stride, d_depth, row_vis, row_vis_stride, reg,
                                                                        g_position, g_log_scaling, g_rotation, g_alpha_logit, g_sh, accumulate)));
  GS_LAUNCH_CHECK();
  return GS_OK;
}

extern "C" int gs_sh_eval_fwd(int64_t v, const int64_t *indexes, const float *sh, int sh_degree, const float *position,
                              const float *h_cam_pos, float *rgb, void *stream) {
  if (v <= 0) return GS_OK;
  Vec3 c = {{h_cam_pos[0], h_cam_pos[1], h_cam_pos[2]}};
  cudaStream_t st = (cudaStream_t)stream;
  GS_DISPATCH_DEG(sh_degree, (k_sh_eval_fwd<D><<<gs_div_up(v, 128), 128, 0, st>>>(v, indexes, sh, position, c, rgb)));
  GS_LAUNCH_CHECK();
  return GS_OK;
}

extern "C" int gs_sh_eval_bwd(int64_t v, const int64_t *indexes, const float *sh, int sh_degree, const float *position,
                              const float *h_cam_pos, const float *d_rgb, float *g_sh, float *g_position, void *stream) {
  if (v <= 0) return GS_OK;
  Vec3 c = {{h_cam_pos[0], h_cam_pos[1], h_cam_pos[2]}};
  cudaStream_t st = (cudaStream_t)stream;
  GS_DISPATCH_DEG(sh_degree, (k_sh_eval_bwd<D><<<gs_div_up(v, 128), 128, 0, st>>>(v, indexes, sh, position, c, d_rgb, g_sh, g_position)));
  GS_LAUNCH_CHECK();
  return GS_OK;
}
