// Small fp32 linear-algebra helpers for the per-Gaussian kernels.
//
// Sort keys and tile rectangles must be bit-identical to the reference's, and those are functions of
// fp32 expressions whose rounding depends on evaluation order and on which products nvcc contracts
// into FMAs.  The helpers below therefore keep the *expression shapes* of the arithmetic the reference
// executes (column-major 3x3 products with the term order of glm's operator*, the point transforms of
// HR/cuda_rasterizer/auxiliary.h:83-122), and this translation unit is compiled with the same
// floating-point flags as the reference (-fmad=true, IEEE div/sqrt, no fast-math).
#pragma once
#include <cuda_runtime.h>
#include "../../include/h3dgs_b200.h"

namespace h3dgs {

// Column-major 3x3: c[j][i] is column j, row i.
struct M3 {
  float c[3][3];
};

__device__ __forceinline__ M3 m3(float a00, float a01, float a02, float a10, float a11, float a12,
                                 float a20, float a21, float a22) {
  M3 m;
  m.c[0][0] = a00; m.c[0][1] = a01; m.c[0][2] = a02;
  m.c[1][0] = a10; m.c[1][1] = a11; m.c[1][2] = a12;
  m.c[2][0] = a20; m.c[2][1] = a21; m.c[2][2] = a22;
  return m;
}

__device__ __forceinline__ M3 m3_mul(const M3& A, const M3& B) {
  M3 r;
#pragma unroll
  for (int j = 0; j < 3; ++j) {
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      r.c[j][i] = A.c[0][i] * B.c[j][0] + A.c[1][i] * B.c[j][1] + A.c[2][i] * B.c[j][2];
    }
  }
  return r;
}

__device__ __forceinline__ M3 m3_t(const M3& A) {
  M3 r;
#pragma unroll
  for (int j = 0; j < 3; ++j)
#pragma unroll
    for (int i = 0; i < 3; ++i) r.c[j][i] = A.c[i][j];
  return r;
}

__device__ __forceinline__ float3 xform4x3(const float3& p, const float* __restrict__ m) {
  float3 r = {
      m[0] * p.x + m[4] * p.y + m[8] * p.z + m[12],
      m[1] * p.x + m[5] * p.y + m[9] * p.z + m[13],
      m[2] * p.x + m[6] * p.y + m[10] * p.z + m[14],
  };
  return r;
}

__device__ __forceinline__ float4 xform4x4(const float3& p, const float* __restrict__ m) {
  float4 r = {
      m[0] * p.x + m[4] * p.y + m[8] * p.z + m[12],
      m[1] * p.x + m[5] * p.y + m[9] * p.z + m[13],
      m[2] * p.x + m[6] * p.y + m[10] * p.z + m[14],
      m[3] * p.x + m[7] * p.y + m[11] * p.z + m[15],
  };
  return r;
}

__device__ __forceinline__ float3 xform_vec4x3_t(const float3& p, const float* __restrict__ m) {
  float3 r = {
      m[0] * p.x + m[1] * p.y + m[2] * p.z,
      m[4] * p.x + m[5] * p.y + m[6] * p.z,
      m[8] * p.x + m[9] * p.y + m[10] * p.z,
  };
  return r;
}

// NDC -> pixel centre; the reference evaluates this in double because of its double literals
// (auxiliary.h:53-56) and the rounding back to float decides tile membership, so keep it in FP64.
__device__ __forceinline__ float ndc_to_pix(float v, int S) { return ((v + 1.0) * S - 1.0) * 0.5; }

// Rotation matrix of an (un-normalised) quaternion (w,x,y,z), listed column by column
// (forward.cu:190-201).  Which of two products nvcc folds into an FMA depends on use counts inside
// the optimiser, not on the source, so the contraction pattern is pinned here with intrinsics to the
// one the reference build has in its sm_100a SASS: r*x, x*z, r*z, y*y, z*z are rounded products; the
// products involving y (y*z, r*y, x*y) and x*x are fused.
__device__ __forceinline__ M3 quat_to_m3(float r, float x, float y, float z) {
  const float rx = __fmul_rn(r, x), xz = __fmul_rn(x, z), rz = __fmul_rn(r, z);
  const float yy = __fmul_rn(y, y), zz = __fmul_rn(z, z);
  const float yz_m_rx = __fmaf_rn(y, z, -rx), yz_p_rx = __fmaf_rn(y, z, rx);
  const float xz_p_ry = __fmaf_rn(r, y, xz), xz_m_ry = __fmaf_rn(-r, y, xz);
  const float xy_m_rz = __fmaf_rn(x, y, -rz), xy_p_rz = __fmaf_rn(x, y, rz);
  const float yy_zz = __fadd_rn(yy, zz), xx_zz = __fmaf_rn(x, x, zz), xx_yy = __fmaf_rn(x, x, yy);
  return m3(__fsub_rn(1.f, __fmul_rn(2.f, yy_zz)), __fmul_rn(2.f, xy_m_rz), __fmul_rn(2.f, xz_p_ry),
            __fmul_rn(2.f, xy_p_rz), __fsub_rn(1.f, __fmul_rn(2.f, xx_zz)), __fmul_rn(2.f, yz_m_rx),
            __fmul_rn(2.f, xz_m_ry), __fmul_rn(2.f, yz_p_rx), __fsub_rn(1.f, __fmul_rn(2.f, xx_yy)));
}

// World-space covariance from scale and rotation: Sigma = (S R)^T (S R); upper triangle out.
__device__ __forceinline__ void cov3d_from_scale_rot(float sx, float sy, float sz, float mod, float qr,
                                                     float qx, float qy, float qz, float* cov6) {
  M3 S = m3(1.f, 0.f, 0.f, 0.f, 1.f, 0.f, 0.f, 0.f, 1.f);
  S.c[0][0] = mod * sx;
  S.c[1][1] = mod * sy;
  S.c[2][2] = mod * sz;
  M3 R = quat_to_m3(qr, qx, qy, qz);
  M3 M = m3_mul(S, R);
  M3 Sig = m3_mul(m3_t(M), M);
  cov6[0] = Sig.c[0][0];
  cov6[1] = Sig.c[0][1];
  cov6[2] = Sig.c[0][2];
  cov6[3] = Sig.c[1][1];
  cov6[4] = Sig.c[1][2];
  cov6[5] = Sig.c[2][2];
}

// EWA projection of a 3D covariance (Zwicker et al. 2002, eqs. 29/31), with the 1.3*tanfov frustum
// clamp of the view-space mean (forward.cu:141-176).  Also returns T = W*J for the backward pass.
__device__ __forceinline__ float3 cov2d_ewa(const float3& mean, float focal_x, float focal_y,
                                            float tan_fovx, float tan_fovy, const float* cov6,
                                            const float* __restrict__ view, M3* T_out, float3* t_out,
                                            float* txtz_out, float* tytz_out) {
  float3 t = xform4x3(mean, view);
  const float limx = 1.3f * tan_fovx;
  const float limy = 1.3f * tan_fovy;
  const float txtz = t.x / t.z;
  const float tytz = t.y / t.z;
  t.x = min(limx, max(-limx, txtz)) * t.z;
  t.y = min(limy, max(-limy, tytz)) * t.z;

  M3 J = m3(focal_x / t.z, 0.0f, -(focal_x * t.x) / (t.z * t.z),
            0.0f, focal_y / t.z, -(focal_y * t.y) / (t.z * t.z),
            0, 0, 0);
  M3 W = m3(view[0], view[4], view[8], view[1], view[5], view[9], view[2], view[6], view[10]);
  M3 T = m3_mul(W, J);
  M3 V = m3(cov6[0], cov6[1], cov6[2], cov6[1], cov6[3], cov6[4], cov6[2], cov6[4], cov6[5]);
  M3 cov = m3_mul(m3_mul(m3_t(T), m3_t(V)), T);
  if (T_out) *T_out = T;
  if (t_out) *t_out = t;
  if (txtz_out) *txtz_out = txtz;
  if (tytz_out) *tytz_out = tytz;
  return make_float3(cov.c[0][0], cov.c[0][1], cov.c[1][1]);
}

// Real spherical-harmonics constants (degree <= 3), as in auxiliary.h:34-51 / utils/sh_utils.py.
__device__ constexpr float kSH0 = 0.28209479177387814f;
__device__ constexpr float kSH1 = 0.4886025119029199f;
__device__ constexpr float kSH2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                                      -1.0925484305920792f, 0.5462742152960396f};
__device__ constexpr float kSH3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f,
                                      0.3731763325901154f,  -0.4570457994644658f, 1.445305721320277f,
                                      -0.5900435899266435f};

// The 16 basis values for a unit direction.
__device__ __forceinline__ void sh_basis(int deg, float x, float y, float z, float* b) {
  b[0] = kSH0;
  if (deg > 0) {
    b[1] = -kSH1 * y;
    b[2] = kSH1 * z;
    b[3] = -kSH1 * x;
    if (deg > 1) {
      float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
      b[4] = kSH2[0] * xy;
      b[5] = kSH2[1] * yz;
      b[6] = kSH2[2] * (2.0f * zz - xx - yy);
      b[7] = kSH2[3] * xz;
      b[8] = kSH2[4] * (xx - yy);
      if (deg > 2) {
        b[9] = kSH3[0] * y * (3.0f * xx - yy);
        b[10] = kSH3[1] * xy * z;
        b[11] = kSH3[2] * y * (4.0f * zz - xx - yy);
        b[12] = kSH3[3] * z * (2.0f * zz - 3.0f * xx - 3.0f * yy);
        b[13] = kSH3[4] * x * (4.0f * zz - xx - yy);
        b[14] = kSH3[5] * z * (xx - yy);
        b[15] = kSH3[6] * x * (xx - 3.0f * yy);
      }
    }
  }
}

__device__ __forceinline__ int sh_count(int deg) { return (deg + 1) * (deg + 1); }

// ---- parameter activations of scene/gaussian_model.py:40-48,411-412, applied to gathered rows ----
// Each one restates the arithmetic of the torch CUDA kernel the reference runs over the whole store, so
// that the activated values (and with them radii, tile counts and sort keys) come out bit-identical:
//   torch.exp      -> expf
//   torch.sigmoid  -> 1 / (1 + expf(-x))
//   torch.abs      -> fabsf
//   F.normalize    -> x / max(sqrt(sum x_i^2), 1e-12); the sum is torch's block-x reduction over the four
//                     components: each lane squares its own element, then shuffle-down by 2 and by 1,
//                     i.e. (q0^2 + q2^2) + (q1^2 + q3^2)  (variant 2, the default: measured bit-identical to
//                     F.normalize on B200 with harness/diag_raw.py; variants 0, 1, 3 are not).  Other orders
//                     stay selectable for that diagnostic (H3DGS_NORM_VARIANT).
__device__ __forceinline__ float act_scalar(int act, float x) {
  if (act == H3DGS_ACT_EXP) return expf(x);
  if (act == H3DGS_ACT_SIGMOID) return __fdiv_rn(1.0f, __fadd_rn(1.0f, expf(-x)));
  if (act == H3DGS_ACT_ABS) return fabsf(x);
  return x;
}
// d act(x) / dx given the raw x and the activated y
__device__ __forceinline__ float act_scalar_grad(int act, float x, float y) {
  if (act == H3DGS_ACT_EXP) return y;
  if (act == H3DGS_ACT_SIGMOID) return (1.0f - y) * y;
  if (act == H3DGS_ACT_ABS) return x > 0.f ? 1.f : (x < 0.f ? -1.f : 0.f);
  return 1.0f;
}
__device__ __forceinline__ float quat_norm(int variant, const float4& q) {
  const float a = __fmul_rn(q.x, q.x), b = __fmul_rn(q.y, q.y), c = __fmul_rn(q.z, q.z), d = __fmul_rn(q.w, q.w);
  float s;
  if (variant == 1) s = __fadd_rn(__fadd_rn(__fadd_rn(a, b), c), d);
  else if (variant == 2) s = __fadd_rn(__fadd_rn(a, c), __fadd_rn(b, d));
  else if (variant == 3) s = __fmaf_rn(q.w, q.w, __fmaf_rn(q.z, q.z, __fmaf_rn(q.y, q.y, a)));
  else s = __fadd_rn(__fadd_rn(a, b), __fadd_rn(c, d));
  return sqrtf(s);
}
__device__ __forceinline__ float4 act_rotation(int act, int variant, const float4& q) {
  if (act != H3DGS_ACT_NORMALIZE) return q;
  const float d = fmaxf(quat_norm(variant, q), 1e-12f);
  return make_float4(__fdiv_rn(q.x, d), __fdiv_rn(q.y, d), __fdiv_rn(q.z, d), __fdiv_rn(q.w, d));
}
// gradient of y = q / max(||q||, eps) contracted with g (autograd of F.normalize: div, clamp_min, norm)
__device__ __forceinline__ float4 act_rotation_grad(int act, int variant, const float4& q, const float4& g) {
  if (act != H3DGS_ACT_NORMALIZE) return g;
  const float n = quat_norm(variant, q);
  const float d = fmaxf(n, 1e-12f);
  const float inv = 1.0f / d;
  float4 r = make_float4(g.x * inv, g.y * inv, g.z * inv, g.w * inv);
  if (n > 1e-12f) {
    const float dot = g.x * q.x + g.y * q.y + g.z * q.z + g.w * q.w;
    const float k = dot * inv * inv / n;
    r.x -= q.x * k; r.y -= q.y * k; r.z -= q.z * k; r.w -= q.w * k;
  }
  return r;
}


// Visit the 45 consecutive floats of a `_features_rest` row (180 bytes, only 4-byte aligned: the 16-byte phase
// depends on the row id) with the widest aligned loads: f(i, value) for i = 0..44, i a compile-time constant
// after unrolling.  11 or 12 load instructions per row instead of 45 -- the per-slot kernels are bound by the
// number of memory operations of their scattered rows, not by bytes.
template <int HEAD, typename F>
__device__ __forceinline__ void row45_aligned(const float* __restrict__ q, F&& f) {
  if (HEAD == 1 || HEAD == 3) f(0, __ldg(q));
  if (HEAD == 2) { const float2 x = __ldg(reinterpret_cast<const float2*>(q)); f(0, x.x); f(1, x.y); }
  if (HEAD == 3) { const float2 x = __ldg(reinterpret_cast<const float2*>(q + 1)); f(1, x.x); f(2, x.y); }
  constexpr int Q = (45 - HEAD) / 4;
#pragma unroll
  for (int g = 0; g < Q; ++g) {
    const int i = HEAD + 4 * g;
    const float4 x = __ldg(reinterpret_cast<const float4*>(q + i));
    f(i, x.x); f(i + 1, x.y); f(i + 2, x.z); f(i + 3, x.w);
  }
  constexpr int T = HEAD + 4 * Q;   // 44, 45, 42, 43
  if (T == 44) f(44, __ldg(q + 44));
  if (T == 42) { const float2 x = __ldg(reinterpret_cast<const float2*>(q + 42)); f(42, x.x); f(43, x.y); f(44, __ldg(q + 44)); }
  if (T == 43) { const float2 x = __ldg(reinterpret_cast<const float2*>(q + 43)); f(43, x.x); f(44, x.y); }
}
template <typename F>
__device__ __forceinline__ void row45(const float* __restrict__ q, F&& f) {
  switch (((16 - (int)(reinterpret_cast<size_t>(q) & 15)) >> 2) & 3) {   // floats up to the next 16-byte boundary
    case 0: row45_aligned<0>(q, f); break;
    case 1: row45_aligned<1>(q, f); break;
    case 2: row45_aligned<2>(q, f); break;
    default: row45_aligned<3>(q, f); break;
  }
}

}  // namespace h3dgs
