// Per-slot forward preprocessing: optional fused cut interpolation (child/parent lerp), EWA projection,
// culling, AA compensation, tile rectangle, SH -> RGB, and the packed 64-byte record the blend kernels
// stage into shared memory.
//
// Reference behaviour being reproduced: HR/cuda_rasterizer/forward.cu:218-434 (preprocessCUDA),
// :141-215 (computeCov2D / computeCov3D), :25-138 (SH), HR/cuda_rasterizer/auxiliary.h:53-102, and for the
// fused-cut mode the Python lerp of gaussian_renderer/__init__.py:202-235.
#include "common.cuh"
#include "gsmath.cuh"

namespace h3dgs {

namespace {

__device__ __forceinline__ float lerp_rn(float t, float a, float omt, float b) {
  // torch evaluates interps*a + interps_inv*b as three separately rounded ops
  return __fadd_rn(__fmul_rn(t, a), __fmul_rn(omt, b));
}

__device__ __forceinline__ float3 ld3(const float* __restrict__ p, size_t row) {
  const float* q = p + row * 3;
  return make_float3(__ldg(q), __ldg(q + 1), __ldg(q + 2));
}

// Tile rectangle of a pixel-space box, clamped to the grid (auxiliary.h:70-80).
__device__ __forceinline__ void tile_rect(float px, float py, int ex, int ey, int gx, int gy, int& x0,
                                          int& y0, int& x1, int& y1) {
  x0 = min(gx, max((int)0, (int)((px - ex) / TILE_X)));
  y0 = min(gy, max((int)0, (int)((py - ey) / TILE_Y)));
  x1 = min(gx, max((int)0, (int)((px + ex + TILE_X - 1) / TILE_X)));
  y1 = min(gy, max((int)0, (int)((py + ey + TILE_Y - 1) / TILE_Y)));
}

template <bool FUSED>
__global__ void __launch_bounds__(256, 3) preprocess_fwd_kernel(const PreArgs a) {
  const int t_idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (t_idx >= a.P) return;

  // every slot starts culled
  a.radii[t_idx] = 0;
  a.ws.tiles_touched[t_idx] = 0;

  size_t r = (size_t)t_idx;
  long long p = -1;
  float t = 1.0f, omt = 0.0f;
  bool lerp = false;
  if (FUSED) {
    const int K = a.P - a.skyboxnum;
    if (t_idx >= K) {
      r = (size_t)(a.rows - a.skyboxnum + (t_idx - K));
    } else {
      r = (size_t)__ldg(a.idx + t_idx);
      if (a.pidx != nullptr) {
        p = __ldg(a.pidx + t_idx);
        if (p >= 0) {
          t = __ldg(a.ts + t_idx);
          omt = __fsub_rn(1.0f, t);
          lerp = (omt != 0.0f);  // t == 1 reproduces the child bit for bit
        }
      }
    }
  }
  const size_t pr = (size_t)(p < 0 ? 0 : p);

  float3 p_orig = ld3(a.means, r);
  if (FUSED && lerp) {
    float3 q = ld3(a.means, pr);
    p_orig = make_float3(lerp_rn(t, p_orig.x, omt, q.x), lerp_rn(t, p_orig.y, omt, q.y),
                         lerp_rn(t, p_orig.z, omt, q.z));
  }

  const float* __restrict__ view = a.view;
  const float* __restrict__ proj = a.proj;

  // screen space
  float4 p_hom = xform4x4(p_orig, proj);
  float p_w = 1.0f / (p_hom.w + 0.0000001f);
  float3 p_proj = {p_hom.x * p_w, p_hom.y * p_w, p_hom.z * p_w};
  float3 p_view = xform4x3(p_orig, view);
  if (p_view.z <= 0.2f) return;  // near cull (forward.cu:316)

  float* cov6 = a.ws.cov3D + (size_t)t_idx * 6;
  float c6[6];
  if (a.cov3D_pre == nullptr) {
    float3 sc = ld3(a.scales, r);
    sc = make_float3(act_scalar(a.act_scale, sc.x), act_scalar(a.act_scale, sc.y), act_scalar(a.act_scale, sc.z));
    const float4 q4 = act_rotation(a.act_rot, a.norm_variant, __ldg(reinterpret_cast<const float4*>(a.rots) + r));
    float qr = q4.x, qx = q4.y, qy = q4.z, qz = q4.w;
    if (FUSED && lerp) {
      float3 s2 = ld3(a.scales, pr);
      s2 = make_float3(act_scalar(a.act_scale, s2.x), act_scalar(a.act_scale, s2.y), act_scalar(a.act_scale, s2.z));
      sc = make_float3(lerp_rn(t, sc.x, omt, s2.x), lerp_rn(t, sc.y, omt, s2.y),
                       lerp_rn(t, sc.z, omt, s2.z));
      float4 o = act_rotation(a.act_rot, a.norm_variant, __ldg(reinterpret_cast<const float4*>(a.rots) + pr));
      float dotp = (qr * o.x + qx * o.y) + (qy * o.z + qz * o.w);
      if (dotp < 0.0f) { o.x = -o.x; o.y = -o.y; o.z = -o.z; o.w = -o.w; }
      qr = lerp_rn(t, qr, omt, o.x);
      qx = lerp_rn(t, qx, omt, o.y);
      qy = lerp_rn(t, qy, omt, o.z);
      qz = lerp_rn(t, qz, omt, o.w);
    }
    cov3d_from_scale_rot(sc.x, sc.y, sc.z, a.scale_modifier, qr, qx, qy, qz, c6);
  } else {
    const float* cp = a.cov3D_pre + r * 6;
#pragma unroll
    for (int i = 0; i < 6; ++i) c6[i] = __ldg(cp + i);
    if (FUSED && lerp) {
      const float* cq = a.cov3D_pre + pr * 6;
#pragma unroll
      for (int i = 0; i < 6; ++i) c6[i] = lerp_rn(t, c6[i], omt, __ldg(cq + i));
    }
  }
#pragma unroll
  for (int i = 0; i < 6; ++i) cov6[i] = c6[i];

  float3 cov = cov2d_ewa(p_orig, a.focal_x, a.focal_y, a.tan_fovx, a.tan_fovy, c6, view, nullptr,
                         nullptr, nullptr, nullptr);

  // low-pass dilation + opacity compensation (forward.cu:355-363)
  constexpr float h_var = 0.3f;
  const float det_cov = cov.x * cov.z - cov.y * cov.y;
  cov.x += h_var;
  cov.z += h_var;
  const float det_cov_plus_h_cov = cov.x * cov.z - cov.y * cov.y;
  const float h_convolution_scaling = sqrt(max(0.000025f, det_cov / det_cov_plus_h_cov));

  const float det = det_cov_plus_h_cov;
  if (det == 0.0f) return;
  float det_inv = 1.f / det;
  float3 conic = {cov.z * det_inv, -cov.y * det_inv, cov.x * det_inv};

  float mid = 0.5f * (cov.x + cov.z);
  float lambda1 = mid + sqrt(max(0.1f, mid * mid - det));
  float lambda2 = mid - sqrt(max(0.1f, mid * mid - det));
  float my_radius = ceil(3.f * sqrt(max(lambda1, lambda2)));
  float2 point_image = {ndc_to_pix(p_proj.x, a.W), ndc_to_pix(p_proj.y, a.H)};

  const int ex = (int)ceil(3.f * sqrt(cov.x));
  const int ey = (int)ceil(3.f * sqrt(cov.z));
  int x0, y0, x1, y1;
  tile_rect(point_image.x, point_image.y, ex, ey, a.grid_x, a.grid_y, x0, y0, x1, y1);
  if ((x1 - x0) * (y1 - y0) == 0) return;

  // colour
  float3 rgb;
  unsigned clamped = 0;
  if (a.colors == nullptr) {
    const float3 cam = make_float3(__ldg(a.campos), __ldg(a.campos + 1), __ldg(a.campos + 2));
    float3 dir = make_float3(p_orig.x - cam.x, p_orig.y - cam.y, p_orig.z - cam.z);
    float len = sqrt(dir.x * dir.x + dir.y * dir.y + dir.z * dir.z);
    dir.x = dir.x / len;
    dir.y = dir.y / len;
    dir.z = dir.z / len;
    float b[16];
    sh_basis(a.D, dir.x, dir.y, dir.z, b);
    const int n = sh_count(a.D);
    const float* __restrict__ sh = a.shs + r * (size_t)a.M * 3;
    const float* __restrict__ shp = a.shs + pr * (size_t)a.M * 3;
    float cr = 0.f, cg = 0.f, cb = 0.f;
    const bool vec = (a.M == 16) && ((reinterpret_cast<size_t>(a.shs) & 15) == 0);
    if (a.shs_rest != nullptr) {
      // raw store: coefficient 0 lives in _features_dc [rows,1,3], the others in _features_rest [rows,M-1,3]
      const float* __restrict__ dc = a.shs + r * 3;
      const float* __restrict__ dcp = a.shs + pr * 3;
      const float* __restrict__ rs = a.shs_rest + r * (size_t)(a.M - 1) * 3;
      const float* __restrict__ rsp = a.shs_rest + pr * (size_t)(a.M - 1) * 3;
      float v[48];
#pragma unroll
      for (int i = 0; i < 3; ++i) v[i] = __ldg(dc + i);
#pragma unroll
      for (int i = 3; i < 48; ++i)
        if (i < n * 3) v[i] = __ldg(rs + (i - 3));
      if (FUSED && lerp) {
#pragma unroll
        for (int i = 0; i < 3; ++i) v[i] = lerp_rn(t, v[i], omt, __ldg(dcp + i));
#pragma unroll
        for (int i = 3; i < 48; ++i)
          if (i < n * 3) v[i] = lerp_rn(t, v[i], omt, __ldg(rsp + (i - 3)));
      }
      cr = b[0] * v[0]; cg = b[0] * v[1]; cb = b[0] * v[2];
#pragma unroll
      for (int k = 1; k < 16; ++k) {
        if (k < n) {
          cr = cr + b[k] * v[3 * k];
          cg = cg + b[k] * v[3 * k + 1];
          cb = cb + b[k] * v[3 * k + 2];
        }
      }
    } else if (vec) {
      // 48 floats = 12 float4 per row
      const float4* s4 = reinterpret_cast<const float4*>(sh);
      const float4* p4 = reinterpret_cast<const float4*>(shp);
      float v[48];
#pragma unroll
      for (int i = 0; i < 12; ++i) {
        if (i * 4 < n * 3) {
          float4 x = __ldg(s4 + i);
          v[4 * i] = x.x; v[4 * i + 1] = x.y; v[4 * i + 2] = x.z; v[4 * i + 3] = x.w;
        }
      }
      if (FUSED && lerp) {
#pragma unroll
        for (int i = 0; i < 12; ++i) {
          if (i * 4 < n * 3) {
            float4 y = __ldg(p4 + i);
            v[4 * i] = lerp_rn(t, v[4 * i], omt, y.x);
            v[4 * i + 1] = lerp_rn(t, v[4 * i + 1], omt, y.y);
            v[4 * i + 2] = lerp_rn(t, v[4 * i + 2], omt, y.z);
            v[4 * i + 3] = lerp_rn(t, v[4 * i + 3], omt, y.w);
          }
        }
      }
      cr = b[0] * v[0]; cg = b[0] * v[1]; cb = b[0] * v[2];
#pragma unroll
      for (int k = 1; k < 16; ++k) {
        if (k < n) {
          cr = cr + b[k] * v[3 * k];
          cg = cg + b[k] * v[3 * k + 1];
          cb = cb + b[k] * v[3 * k + 2];
        }
      }
    } else {
      for (int k = 0; k < n; ++k) {
        float s0 = __ldg(sh + 3 * k), s1 = __ldg(sh + 3 * k + 1), s2 = __ldg(sh + 3 * k + 2);
        if (FUSED && lerp) {
          s0 = lerp_rn(t, s0, omt, __ldg(shp + 3 * k));
          s1 = lerp_rn(t, s1, omt, __ldg(shp + 3 * k + 1));
          s2 = lerp_rn(t, s2, omt, __ldg(shp + 3 * k + 2));
        }
        if (k == 0) { cr = b[0] * s0; cg = b[0] * s1; cb = b[0] * s2; }
        else { cr = cr + b[k] * s0; cg = cg + b[k] * s1; cb = cb + b[k] * s2; }
      }
    }
    cr += 0.5f; cg += 0.5f; cb += 0.5f;
    clamped = (cr < 0 ? 1u : 0u) | (cg < 0 ? 2u : 0u) | (cb < 0 ? 4u : 0u);
    rgb = make_float3(max(cr, 0.0f), max(cg, 0.0f), max(cb, 0.0f));
  } else {
    rgb = ld3(a.colors, r);
    if (FUSED && lerp) {
      float3 q = ld3(a.colors, pr);
      rgb = make_float3(lerp_rn(t, rgb.x, omt, q.x), lerp_rn(t, rgb.y, omt, q.y),
                        lerp_rn(t, rgb.z, omt, q.z));
    }
  }

  float opacity = act_scalar(a.act_opac, __ldg(a.opac + r));
  if (FUSED && lerp) opacity = lerp_rn(t, opacity, omt, act_scalar(a.act_opac, __ldg(a.opac + pr)));
  const float op = opacity * h_convolution_scaling;

  // Bound of the quadratic form inside which this Gaussian can reach alpha >= 1/255:
  // o exp(-q/2) >= 1/255  <=>  q = a dx^2 + 2 b dx dy + c dy^2 <= 2 ln(255 o); kept with a safety margin far
  // above the rounding of either side, so the blend kernels' block test (blend_common.cuh) only ever drops
  // pairs the reference would reject at its alpha < 1/255 test.
  float qmax = -1.0f;                       // opacity below 1/255: contributes nowhere
  {
    const float l = 2.0f * logf(255.0f * op);
    if (l > 0.0f) qmax = l * 1.002f + 0.002f;
    else if (!(op < 1.0f / 255.0f)) qmax = __int_as_float(0x7f800000);   // NaN / inf opacity: never cull
  }

  // per-slot interpolation terms of the blend (forward.cu:510-511); slot-indexed arrays
  float bt = 1.0f, binvk = 1.0f;
  if (a.ts != nullptr && a.kids != nullptr && t_idx < a.P - a.skyboxnum) {
    bt = __ldg(a.ts + t_idx);
    binvk = 1.0f / (float)__ldg(a.kids + t_idx);
  }

  Rec rec;
  rec.q0 = make_float4(point_image.x, point_image.y, qmax, 0.0f);
  rec.q1 = make_float4(conic.x, conic.y, conic.z, op);
  rec.q2 = make_float4(rgb.x, rgb.y, rgb.z, 1 / p_view.z);
  rec.q3 = make_float4(bt, binvk, p_view.z, h_convolution_scaling);
  a.ws.rec[t_idx] = rec;

  a.ws.depth[t_idx] = p_view.z;
  a.ws.clamped[t_idx] = (uint8_t)clamped;
  a.ws.tile_rect[t_idx] = make_ushort4((unsigned short)x0, (unsigned short)y0, (unsigned short)x1,
                                       (unsigned short)y1);
  a.radii[t_idx] = (int)my_radius;
  a.ws.tiles_touched[t_idx] = (uint32_t)((y1 - y0) * (x1 - x0));
}

__global__ void mark_visible_kernel(int P, const float* __restrict__ means,
                                    const float* __restrict__ view, const float* __restrict__ proj,
                                    uint8_t* __restrict__ present) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P) return;
  float3 p = ld3(means, (size_t)i);
  float3 v = xform4x3(p, view);
  present[i] = (v.z <= 0.2f) ? 0 : 1;  // rasterizer_impl.cu:54-66 / auxiliary.h:164-189
}

}  // namespace

cudaError_t launch_preprocess_fwd(const PreArgs& a, cudaStream_t s) {
  if (a.P <= 0) return cudaSuccess;
  const int blocks = (a.P + 255) / 256;
  ProfScope ps(K_PREPROCESS_FWD, s);
  if (a.idx != nullptr)
    preprocess_fwd_kernel<true><<<blocks, 256, 0, s>>>(a);
  else
    preprocess_fwd_kernel<false><<<blocks, 256, 0, s>>>(a);
  return cudaGetLastError();
}

cudaError_t launch_mark_visible(int P, const float* means, const float* view, const float* proj,
                                uint8_t* present, cudaStream_t s) {
  if (P <= 0) return cudaSuccess;
  ProfScope ps(K_MARK_VISIBLE, s);
  mark_visible_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, means, view, proj, present);
  return cudaGetLastError();
}

}  // namespace h3dgs
