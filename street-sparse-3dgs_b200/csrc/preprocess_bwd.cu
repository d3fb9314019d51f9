// Per-slot backward preprocessing: from the reduced screen-space gradients (GradRec) to gradients of
// means, SH / colours, opacity, scales, rotations (or the precomputed covariance), including the
// chain through the fused child/parent interpolation.
//
// One kernel replaces computeCov2DCUDA + preprocessCUDA<backward> (HR/cuda_rasterizer/backward.cu:147-495)
// and the caller-side zero fill of the gradient tensors (HR/rasterize_points.cu:168-198): in plain mode
// every output row is written exactly once (zeros for culled slots); in fused-cut mode outputs are
// zeroed by the entry point and visible slots scatter t*g to the child row and (1-t)*g to the parent
// row, which is what autograd does for the Python lerp (gaussian_renderer/__init__.py:202-235).
#include "common.cuh"
#include "gsmath.cuh"

namespace h3dgs {

namespace {

#ifndef PRE_BWD_CTAS
#define PRE_BWD_CTAS 2   // resident CTAs per SM of the per-slot backward kernel (register budget 65536 / (256 x CTAs))
#endif

__device__ __forceinline__ float sq(float x) { return x * x; }

__device__ __forceinline__ float3 ld3(const float* __restrict__ p, size_t row) {
  const float* q = p + row * 3;
  return make_float3(__ldg(q), __ldg(q + 1), __ldg(q + 2));
}
__device__ __forceinline__ float lerp_rn(float t, float a, float omt, float b) {
  return __fadd_rn(__fmul_rn(t, a), __fmul_rn(omt, b));
}

// d(v/|v|)/dv applied to dv (auxiliary.h:132-142)
__device__ __forceinline__ float3 dnormvdv(float3 v, float3 dv) {
  float sum2 = v.x * v.x + v.y * v.y + v.z * v.z;
  float invsum32 = 1.0f / sqrt(sum2 * sum2 * sum2);
  float3 r;
  r.x = ((+sum2 - v.x * v.x) * dv.x - v.y * v.x * dv.y - v.z * v.x * dv.z) * invsum32;
  r.y = (-v.x * v.y * dv.x + (sum2 - v.y * v.y) * dv.y - v.z * v.y * dv.z) * invsum32;
  r.z = (-v.x * v.z * dv.x - v.y * v.z * dv.y + (sum2 - v.z * v.z) * dv.z) * invsum32;
  return r;
}

template <bool FUSED>
__device__ __forceinline__ void put(float* base, size_t row, int width, int k, float v, size_t prow,
                                    bool lerp, float t, float omt, float psign = 1.0f) {
  if (base == nullptr) return;
  if (!FUSED) {
    base[row * width + k] = v;
  } else {
    if (!lerp) {
      atomicAdd(base + row * width + k, v);
    } else {
      atomicAdd(base + row * width + k, t * v);
      atomicAdd(base + prow * width + k, omt * psign * v);
    }
  }
}

// child / parent values already carry their t, (1-t) and activation-derivative factors
template <bool FUSED>
__device__ __forceinline__ void put2(float* base, size_t row, int width, int k, float vc, size_t prow,
                                     bool lerp, float vp) {
  if (base == nullptr) return;
  if (!FUSED) {
    base[row * width + k] = vc;
  } else {
    atomicAdd(base + row * width + k, vc);
    if (lerp) atomicAdd(base + prow * width + k, vp);
  }
}

// Pass 1, one thread per slot: culled slots get their zero rows (plain mode: of every output; fused-cut mode: of
// dL_dmeans2D only, the store rows are zero-filled by the caller), visible slots are listed in slot order.
// ---- vector reductions to global memory (sm_90+) ----
// The scatter of fused-cut mode is bound by the NUMBER of L2 reduction operations (one per lane and
// instruction; 118 scalar ones per interpolated slot), not by bytes: rows are therefore added with the widest
// reduction their alignment allows -- 16 bytes per operation on the 180/192-byte SH rows and the quaternions,
// 8 + 4 bytes on the 12-byte rows.
__device__ __forceinline__ void red1(float* p, float a) {
  asm volatile("red.global.add.f32 [%0], %1;\n" ::"l"(p), "f"(a) : "memory");
}
__device__ __forceinline__ void red2(float* p, float a, float b) {
  asm volatile("red.global.add.v2.f32 [%0], {%1, %2};\n" ::"l"(p), "f"(a), "f"(b) : "memory");
}
__device__ __forceinline__ void red4(float* p, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};\n" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d)
               : "memory");
}
// three consecutive floats at a 4-byte aligned address
__device__ __forceinline__ void red_row3(float* p, float a, float b, float c) {
  if ((reinterpret_cast<size_t>(p) & 7) == 0) { red2(p, a, b); red1(p + 2, c); }
  else { red1(p, a); red2(p + 1, b, c); }
}
// N consecutive floats s * v[0..N) at an address that is 4 * HEAD bytes short of 16-byte alignment
template <int HEAD, int N>
__device__ __forceinline__ void red_row_aligned(float* p, const float (&v)[48], float s) {
#pragma unroll
  for (int i = 0; i < HEAD; ++i) red1(p + i, s * v[i]);
  constexpr int Q = (N - HEAD) / 4;
#pragma unroll
  for (int q = 0; q < Q; ++q) {
    const int i = HEAD + 4 * q;
    red4(p + i, s * v[i], s * v[i + 1], s * v[i + 2], s * v[i + 3]);
  }
#pragma unroll
  for (int i = HEAD + 4 * Q; i < N; ++i) red1(p + i, s * v[i]);
}
template <int N>
__device__ __forceinline__ void red_row(float* p, const float (&v)[48], float s) {
  switch ((16 - (int)(reinterpret_cast<size_t>(p) & 15)) >> 2 & 3) {   // floats up to the next 16-byte boundary
    case 0: red_row_aligned<0, N>(p, v, s); break;
    case 1: red_row_aligned<1, N>(p, v, s); break;
    case 2: red_row_aligned<2, N>(p, v, s); break;
    default: red_row_aligned<3, N>(p, v, s); break;
  }
}

template <bool FUSED>
__global__ void __launch_bounds__(256) list_visible_kernel(const PreBwdArgs b) {
  const PreArgs& a = b.f;
  const int t_idx = blockIdx.x * blockDim.x + threadIdx.x;
  const H3dgsGrads& o = b.out;
  const int M = a.M;
  const bool valid = t_idx < a.P;
  const bool visible = valid && a.radii[t_idx] > 0;
  if (visible) b.vis_list[a.ws.offsets[t_idx] - 1u] = (uint32_t)t_idx;
  if (FUSED && o.row_map != nullptr) {
    // compact gradients: every store row a visible slot adds into gets a compact row.  The first thread to turn a
    // row's map entry from -1 into its (negative) tag owns the row; the owners of a warp take their ids with ONE
    // atomic on the row counter (one per row serialised 1.7 M atomics on a single address) and publish them; ids are
    // only read by the next kernel.  Whole warps walk through here together (no early return above).
    size_t r = 0;
    long long p = -1;
    bool lerp = false;
    if (visible) {
      const int K = a.P - a.skyboxnum;
      if (t_idx >= K) {
        r = (size_t)(a.rows - a.skyboxnum + (t_idx - K));
      } else {
        r = (size_t)__ldg(a.idx + t_idx);
        if (a.pidx != nullptr) {
          p = __ldg(a.pidx + t_idx);
          if (p >= 0) lerp = __fsub_rn(1.0f, __ldg(a.ts + t_idx)) != 0.0f;
        }
      }
    }
    const int tag = -2 - 2 * t_idx;
    const int lane_id = threadIdx.x & 31;
#pragma unroll
    for (int side = 0; side < 2; ++side) {
      const bool want = visible && (side == 0 || lerp);
      const size_t row = side ? (size_t)(p < 0 ? 0 : p) : r;
      const bool won = want && atomicCAS(o.row_map + row, -1, tag - side) == -1;
      const unsigned m = __ballot_sync(0xffffffffu, won);
      if (m == 0) continue;
      int base = 0;
      if (lane_id == __ffs(m) - 1) base = atomicAdd(o.compact_count, __popc(m));
      base = __shfl_sync(0xffffffffu, base, __ffs(m) - 1);
      if (won) {
        const int id = base + __popc(m & ((1u << lane_id) - 1u));
        if (id < o.compact_capacity) {
          o.compact_rows[id] = (long long)row;
          o.row_map[row] = id;
        }
      }
    }
  }
  if (!valid || visible) return;
  if (o.dL_dmeans2D) {
    o.dL_dmeans2D[(size_t)t_idx * 3] = 0.f;
    o.dL_dmeans2D[(size_t)t_idx * 3 + 1] = 0.f;
    o.dL_dmeans2D[(size_t)t_idx * 3 + 2] = 0.f;
  }
  if (!FUSED) {
    const size_t r = (size_t)t_idx;
    if (o.dL_dmeans3D) { o.dL_dmeans3D[r * 3] = 0.f; o.dL_dmeans3D[r * 3 + 1] = 0.f; o.dL_dmeans3D[r * 3 + 2] = 0.f; }
    if (o.dL_dcolors) { o.dL_dcolors[r * 3] = 0.f; o.dL_dcolors[r * 3 + 1] = 0.f; o.dL_dcolors[r * 3 + 2] = 0.f; }
    if (o.dL_dopacity) o.dL_dopacity[r] = 0.f;
    if (o.dL_dscales) { o.dL_dscales[r * 3] = 0.f; o.dL_dscales[r * 3 + 1] = 0.f; o.dL_dscales[r * 3 + 2] = 0.f; }
    if (o.dL_drotations) *reinterpret_cast<float4*>(o.dL_drotations + r * 4) = make_float4(0.f, 0.f, 0.f, 0.f);
    if (o.dL_dcov3D) for (int i = 0; i < 6; ++i) o.dL_dcov3D[r * 6 + i] = 0.f;
    if (o.dL_dsh && o.dL_dsh_rest) {
      for (int i = 0; i < 3; ++i) o.dL_dsh[r * 3 + i] = 0.f;
      for (int i = 0; i < (M - 1) * 3; ++i) o.dL_dsh_rest[r * (size_t)(M - 1) * 3 + i] = 0.f;
    } else if (o.dL_dsh) {
      float* d = o.dL_dsh + r * (size_t)M * 3;
      if (((M * 3) & 3) == 0 && ((reinterpret_cast<size_t>(o.dL_dsh) & 15) == 0)) {
        for (int i = 0; i < M * 3 / 4; ++i) reinterpret_cast<float4*>(d)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
      } else {
        for (int i = 0; i < M * 3; ++i) d[i] = 0.f;
      }
    }
  }
}

// Pass 2, one thread per VISIBLE slot (typically a quarter of the slots of a cut: launching over all slots left
// three lanes in four idle through ~1500 instructions of a latency-bound kernel).  The grid covers P; CTAs
// beyond the visible count, which the forward left in the geometry workspace, exit at once.
template <bool FUSED>
__global__ void __launch_bounds__(256, PRE_BWD_CTAS) preprocess_bwd_kernel(const PreBwdArgs b) {
  const PreArgs& a = b.f;
  const unsigned long long V = *reinterpret_cast<const unsigned long long*>(a.ws.counters + 2);
  const unsigned long long vi = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (vi >= V) return;
  const int t_idx = (int)b.vis_list[vi];
  const H3dgsGrads& o = b.out;
  const int M = a.M;

  // ---- which rows? (same decode as the forward) ----
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
          lerp = (omt != 0.0f);
        }
      }
    }
  }
  const size_t pr = (size_t)(p < 0 ? 0 : p);
  // rows of the OUTPUT tensors: the store rows themselves, or their compact rows (H3dgsGrads.compact_*)
  size_t wr = r, wpr = pr;
  if (FUSED && o.row_map != nullptr) {
    const int cr = o.row_map[r];
    const int cpr = lerp ? o.row_map[pr] : cr;
    if (cr < 0 || cr >= o.compact_capacity || cpr < 0 || cpr >= o.compact_capacity) return;   // capacity exceeded
    wr = (size_t)cr;
    wpr = (size_t)cpr;
  }

  // The kernel is a long chain of dependent arithmetic fed by ~25 scattered rows (child and parent parameters,
  // then the same rows of the gradient tensors): without help every row is a separate DRAM round trip in the
  // middle of the chain.  Ask L2 for all of them now.
  if (FUSED) {
    auto pf = [](const void* q) { asm volatile("prefetch.global.L2 [%0];" ::"l"(q)); };
    auto pf_row = [&](const float* base, size_t row, int width) {
      if (base == nullptr) return;
      const char* q = reinterpret_cast<const char*>(base + row * (size_t)width);
      pf(q);
      if (width > 16) pf(q + width * 4 - 4);   // rows longer than 64 B: both ends (the 180/192-byte SH rows)
      if (width > 32) pf(q + 128);
    };
    const bool split = a.shs_rest != nullptr;
    const int wsh = split ? (M - 1) * 3 : M * 3;
    for (int side = 0; side < (lerp ? 2 : 1); ++side) {
      const size_t row = side ? pr : r;
      const size_t wrow = side ? wpr : wr;
      pf_row(a.means, row, 3);
      pf_row(a.scales, row, 3);
      pf_row(a.rots, row, 4);
      pf_row(a.opac, row, 1);
      pf_row(split ? a.shs_rest : a.shs, row, wsh);
      pf_row(o.dL_dmeans3D, wrow, 3);
      pf_row(o.dL_dscales, wrow, 3);
      pf_row(o.dL_drotations, wrow, 4);
      pf_row(o.dL_dopacity, wrow, 1);
      pf_row(split ? o.dL_dsh_rest : o.dL_dsh, wrow, wsh);
      if (split) { pf_row(a.shs, row, 3); pf_row(o.dL_dsh, wrow, 3); }
    }
  }

  if (FUSED && o.dirty_rows != nullptr) {
    // remember which store rows this render adds into (child, and parent when interpolating)
    // (two entries per visible slot, one warp-aggregated atomic; without a parent the child is listed twice)
    const unsigned m = __activemask();
    const int lane_id = threadIdx.x & 31;
    const int leader = __ffs(m) - 1;
    int base = 0;
    if (lane_id == leader) base = atomicAdd(o.dirty_rows, 2 * __popc(m));
    base = __shfl_sync(m, base, leader);
    const int at = 1 + base + 2 * __popc(m & ((1u << lane_id) - 1u));
    o.dirty_rows[at] = (int32_t)r;
    o.dirty_rows[at + 1] = (int32_t)(lerp ? pr : r);
  }

  const GradRec gr = b.grad[t_idx];
  const float dL_dmean2D_x = gr.v[0], dL_dmean2D_y = gr.v[1];
  const float3 dL_dconic = {gr.v[2], gr.v[3], gr.v[4]};
  const float dL_dopacity_v = gr.v[5];
  float3 dL_dRGB = {gr.v[6], gr.v[7], gr.v[8]};
  const float dL_dinvdepth = gr.v[9];

  if (o.dL_dmeans2D) {
    o.dL_dmeans2D[(size_t)t_idx * 3] = dL_dmean2D_x;
    o.dL_dmeans2D[(size_t)t_idx * 3 + 1] = dL_dmean2D_y;
    o.dL_dmeans2D[(size_t)t_idx * 3 + 2] = 0.f;
  }

  float3 mean = ld3(a.means, r);
  if (FUSED && lerp) {
    float3 q = ld3(a.means, pr);
    mean = make_float3(lerp_rn(t, mean.x, omt, q.x), lerp_rn(t, mean.y, omt, q.y),
                       lerp_rn(t, mean.z, omt, q.z));
  }
  const float* __restrict__ view = a.view;
  const float* __restrict__ proj = a.proj;

  // ---------------- 2D covariance / conic backward (backward.cu:147-326) ----------------
  float c6[6];
#pragma unroll
  for (int i = 0; i < 6; ++i) c6[i] = a.ws.cov3D[(size_t)t_idx * 6 + i];
  M3 T;
  float3 tv;
  float txtz, tytz;
  float3 cov = cov2d_ewa(mean, a.focal_x, a.focal_y, a.tan_fovx, a.tan_fovy, c6, view, &T, &tv, &txtz,
                         &tytz);
  const float limx = 1.3f * a.tan_fovx, limy = 1.3f * a.tan_fovy;
  const float x_grad_mul = txtz < -limx || txtz > limx ? 0 : 1;
  const float y_grad_mul = tytz < -limy || tytz > limy ? 0 : 1;

  float c_xx = cov.x, c_xy = cov.y, c_yy = cov.z;
  constexpr float h_var = 0.3f;
  const float det_cov = c_xx * c_yy - c_xy * c_xy;
  c_xx += h_var;
  c_yy += h_var;
  const float det_cov_plus_h_cov = c_xx * c_yy - c_xy * c_xy;
  const float h_convolution_scaling = sqrt(max(0.000025f, det_cov / det_cov_plus_h_cov));

  const float tc = lerp ? t : 1.0f;   // weight of the child row
  const float o_raw = __ldg(a.opac + r);
  const float o_act = act_scalar(a.act_opac, o_raw);
  float op_raw = 0.f, op_act = 0.f;
  float opacity = o_act;
  if (FUSED && lerp) {
    op_raw = __ldg(a.opac + pr);
    op_act = act_scalar(a.act_opac, op_raw);
    opacity = lerp_rn(t, o_act, omt, op_act);
  }
  const float d_h_convolution_scaling = dL_dopacity_v * opacity;
  const float dL_dopacity_out = dL_dopacity_v * h_convolution_scaling;
  const float d_inside_root = (det_cov / det_cov_plus_h_cov) <= 0.000025f
                                  ? 0.f
                                  : d_h_convolution_scaling / (2 * h_convolution_scaling);

  float dL_dc_xx = 0, dL_dc_xy = 0, dL_dc_yy = 0;
  {
    // gradient of sqrt(det/det') through the dilated entries, evaluated at the dilated values exactly
    // as backward.cu:234-244 does
    const float x = c_xx, y = c_yy, z = c_xy, w = h_var;
    const float denom_f = d_inside_root / sq(w * w + w * (x + y) + x * y - z * z);
    dL_dc_xx = w * (w * y + y * y + z * z) * denom_f;
    dL_dc_yy = w * (w * x + x * x + z * z) * denom_f;
    dL_dc_xy = -2.f * w * z * (w + x + y) * denom_f;
  }

  const float denom = c_xx * c_yy - c_xy * c_xy;
  const float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
  float dcov[6] = {0, 0, 0, 0, 0, 0};
  if (denom2inv != 0) {
    dL_dc_xx += denom2inv * (-c_yy * c_yy * dL_dconic.x + 2 * c_xy * c_yy * dL_dconic.y +
                             (denom - c_xx * c_yy) * dL_dconic.z);
    dL_dc_yy += denom2inv * (-c_xx * c_xx * dL_dconic.z + 2 * c_xx * c_xy * dL_dconic.y +
                             (denom - c_xx * c_yy) * dL_dconic.x);
    dL_dc_xy += denom2inv * 2 *
                (c_xy * c_yy * dL_dconic.x - (denom + 2 * c_xy * c_xy) * dL_dconic.y +
                 c_xx * c_xy * dL_dconic.z);
    dcov[0] = (T.c[0][0] * T.c[0][0] * dL_dc_xx + T.c[0][0] * T.c[1][0] * dL_dc_xy + T.c[1][0] * T.c[1][0] * dL_dc_yy);
    dcov[3] = (T.c[0][1] * T.c[0][1] * dL_dc_xx + T.c[0][1] * T.c[1][1] * dL_dc_xy + T.c[1][1] * T.c[1][1] * dL_dc_yy);
    dcov[5] = (T.c[0][2] * T.c[0][2] * dL_dc_xx + T.c[0][2] * T.c[1][2] * dL_dc_xy + T.c[1][2] * T.c[1][2] * dL_dc_yy);
    dcov[1] = 2 * T.c[0][0] * T.c[0][1] * dL_dc_xx + (T.c[0][0] * T.c[1][1] + T.c[0][1] * T.c[1][0]) * dL_dc_xy + 2 * T.c[1][0] * T.c[1][1] * dL_dc_yy;
    dcov[2] = 2 * T.c[0][0] * T.c[0][2] * dL_dc_xx + (T.c[0][0] * T.c[1][2] + T.c[0][2] * T.c[1][0]) * dL_dc_xy + 2 * T.c[1][0] * T.c[1][2] * dL_dc_yy;
    dcov[4] = 2 * T.c[0][2] * T.c[0][1] * dL_dc_xx + (T.c[0][1] * T.c[1][2] + T.c[0][2] * T.c[1][1]) * dL_dc_xy + 2 * T.c[1][1] * T.c[1][2] * dL_dc_yy;
  }

  // Vrk rows (symmetric)
  const float V00 = c6[0], V01 = c6[1], V02 = c6[2], V11 = c6[3], V12 = c6[4], V22 = c6[5];
  const float a0 = T.c[0][0] * V00 + T.c[0][1] * V01 + T.c[0][2] * V02;
  const float a1 = T.c[0][0] * V01 + T.c[0][1] * V11 + T.c[0][2] * V12;
  const float a2 = T.c[0][0] * V02 + T.c[0][1] * V12 + T.c[0][2] * V22;
  const float b0 = T.c[1][0] * V00 + T.c[1][1] * V01 + T.c[1][2] * V02;
  const float b1 = T.c[1][0] * V01 + T.c[1][1] * V11 + T.c[1][2] * V12;
  const float b2 = T.c[1][0] * V02 + T.c[1][1] * V12 + T.c[1][2] * V22;
  const float dL_dT00 = 2 * a0 * dL_dc_xx + b0 * dL_dc_xy;
  const float dL_dT01 = 2 * a1 * dL_dc_xx + b1 * dL_dc_xy;
  const float dL_dT02 = 2 * a2 * dL_dc_xx + b2 * dL_dc_xy;
  const float dL_dT10 = 2 * b0 * dL_dc_yy + a0 * dL_dc_xy;
  const float dL_dT11 = 2 * b1 * dL_dc_yy + a1 * dL_dc_xy;
  const float dL_dT12 = 2 * b2 * dL_dc_yy + a2 * dL_dc_xy;

  // W columns as in cov2d_ewa: W.c[j][i] = view[4*i + j]
  const float W00 = view[0], W01 = view[4], W02 = view[8];
  const float W10 = view[1], W11 = view[5], W12 = view[9];
  const float W20 = view[2], W21 = view[6], W22 = view[10];
  const float dL_dJ00 = W00 * dL_dT00 + W01 * dL_dT01 + W02 * dL_dT02;
  const float dL_dJ02 = W20 * dL_dT00 + W21 * dL_dT01 + W22 * dL_dT02;
  const float dL_dJ11 = W10 * dL_dT10 + W11 * dL_dT11 + W12 * dL_dT12;
  const float dL_dJ12 = W20 * dL_dT10 + W21 * dL_dT11 + W22 * dL_dT12;

  const float tz = 1.f / tv.z;
  const float tz2 = tz * tz;
  const float tz3 = tz2 * tz;
  const float h_x = a.focal_x, h_y = a.focal_y;
  const float dL_dtx = x_grad_mul * -h_x * tz2 * dL_dJ02;
  const float dL_dty = y_grad_mul * -h_y * tz2 * dL_dJ12;
  float dL_dtz = -h_x * tz2 * dL_dJ00 - h_y * tz2 * dL_dJ11 + (2 * h_x * tv.x) * tz3 * dL_dJ02 +
                 (2 * h_y * tv.y) * tz3 * dL_dJ12;
  if (b.has_invdepth) dL_dtz -= dL_dinvdepth / (tv.z * tv.z);
  float3 dL_dmean = xform_vec4x3_t(make_float3(dL_dtx, dL_dty, dL_dtz), view);

  // ---------------- projection of the mean (backward.cu:433-447) ----------------
  {
    float4 m_hom = xform4x4(mean, proj);
    float m_w = 1.0f / (m_hom.w + 0.0000001f);
    float mul1 = (proj[0] * mean.x + proj[4] * mean.y + proj[8] * mean.z + proj[12]) * m_w * m_w;
    float mul2 = (proj[1] * mean.x + proj[5] * mean.y + proj[9] * mean.z + proj[13]) * m_w * m_w;
    dL_dmean.x += (proj[0] * m_w - proj[3] * mul1) * dL_dmean2D_x + (proj[1] * m_w - proj[3] * mul2) * dL_dmean2D_y;
    dL_dmean.y += (proj[4] * m_w - proj[7] * mul1) * dL_dmean2D_x + (proj[5] * m_w - proj[7] * mul2) * dL_dmean2D_y;
    dL_dmean.z += (proj[8] * m_w - proj[11] * mul1) * dL_dmean2D_x + (proj[9] * m_w - proj[11] * mul2) * dL_dmean2D_y;
  }

  // ---------------- colour (backward.cu:23-142) ----------------
  if (a.colors == nullptr) {
    const unsigned cl = a.ws.clamped[t_idx];
    dL_dRGB.x *= (cl & 1u) ? 0 : 1;
    dL_dRGB.y *= (cl & 2u) ? 0 : 1;
    dL_dRGB.z *= (cl & 4u) ? 0 : 1;
    const float3 cam = make_float3(__ldg(a.campos), __ldg(a.campos + 1), __ldg(a.campos + 2));
    const float3 dir_orig = make_float3(mean.x - cam.x, mean.y - cam.y, mean.z - cam.z);
    const float len = sqrt(dir_orig.x * dir_orig.x + dir_orig.y * dir_orig.y + dir_orig.z * dir_orig.z);
    const float x = dir_orig.x / len, y = dir_orig.y / len, z = dir_orig.z / len;
    float bs[16];
    sh_basis(a.D, x, y, z, bs);
    const int n = sh_count(a.D);
    const bool split = a.shs_rest != nullptr;   // raw store: coefficient 0 in shs [rows,1,3], rest in shs_rest
    const float* __restrict__ sh = split ? a.shs_rest + r * (size_t)(M - 1) * 3 - 3 : a.shs + r * (size_t)M * 3;
    const float* __restrict__ shp = split ? a.shs_rest + pr * (size_t)(M - 1) * 3 - 3 : a.shs + pr * (size_t)M * 3;

    // dL/dsh rows
    if (FUSED && M == 16 && o.dL_dsh != nullptr && (!split || o.dL_dsh_rest != nullptr)) {
      // all 48 products in registers, then wide reductions into the child row (weight t) and the parent row (1-t)
      float v[48];
#pragma unroll
      for (int k = 0; k < 16; ++k) {
        const float w = (k < n) ? bs[k] : 0.f;
        v[3 * k] = w * dL_dRGB.x; v[3 * k + 1] = w * dL_dRGB.y; v[3 * k + 2] = w * dL_dRGB.z;
      }
      for (int side = 0; side < (lerp ? 2 : 1); ++side) {
        const size_t row = side ? wpr : wr;
        const float sc = lerp ? (side ? omt : t) : 1.0f;
        if (split) {
          red_row3(o.dL_dsh + row * 3, sc * v[0], sc * v[1], sc * v[2]);
          float* q = o.dL_dsh_rest + row * 45;
          // rest row = v[3..48): shift the view by three floats
          switch ((16 - (int)(reinterpret_cast<size_t>(q) & 15)) >> 2 & 3) {
            case 0:
#pragma unroll
              for (int i = 0; i < 44; i += 4) red4(q + i, sc * v[3 + i], sc * v[4 + i], sc * v[5 + i], sc * v[6 + i]);
              red1(q + 44, sc * v[47]);
              break;
            case 1:
              red1(q, sc * v[3]);
#pragma unroll
              for (int i = 1; i < 45; i += 4) red4(q + i, sc * v[3 + i], sc * v[4 + i], sc * v[5 + i], sc * v[6 + i]);
              break;
            case 2:
              red2(q, sc * v[3], sc * v[4]);
#pragma unroll
              for (int i = 2; i < 42; i += 4) red4(q + i, sc * v[3 + i], sc * v[4 + i], sc * v[5 + i], sc * v[6 + i]);
              red2(q + 42, sc * v[45], sc * v[46]);
              red1(q + 44, sc * v[47]);
              break;
            default:
              red1(q, sc * v[3]); red2(q + 1, sc * v[4], sc * v[5]);
#pragma unroll
              for (int i = 3; i < 43; i += 4) red4(q + i, sc * v[3 + i], sc * v[4 + i], sc * v[5 + i], sc * v[6 + i]);
              red2(q + 43, sc * v[46], sc * v[47]);
              break;
          }
        } else {
          red_row<48>(o.dL_dsh + row * 48, v, sc);
        }
      }
    } else if (split) {
      put<FUSED>(o.dL_dsh, wr, 3, 0, bs[0] * dL_dRGB.x, wpr, lerp, t, omt);
      put<FUSED>(o.dL_dsh, wr, 3, 1, bs[0] * dL_dRGB.y, wpr, lerp, t, omt);
      put<FUSED>(o.dL_dsh, wr, 3, 2, bs[0] * dL_dRGB.z, wpr, lerp, t, omt);
      const int Wd = (M - 1) * 3;
      for (int k = 1; k < M; ++k) {
        const float w = (k < n) ? bs[k] : 0.f;
        if (!FUSED || k < n) {   // in fused mode the rows are pre-zeroed: skip the zero atomics
          put<FUSED>(o.dL_dsh_rest, wr, Wd, 3 * k - 3, w * dL_dRGB.x, wpr, lerp, t, omt);
          put<FUSED>(o.dL_dsh_rest, wr, Wd, 3 * k - 2, w * dL_dRGB.y, wpr, lerp, t, omt);
          put<FUSED>(o.dL_dsh_rest, wr, Wd, 3 * k - 1, w * dL_dRGB.z, wpr, lerp, t, omt);
        }
      }
    } else {
      for (int k = 0; k < M; ++k) {
        const float w = (k < n) ? bs[k] : 0.f;
        if (!FUSED || k < n) {
          put<FUSED>(o.dL_dsh, wr, M * 3, 3 * k + 0, w * dL_dRGB.x, wpr, lerp, t, omt);
          put<FUSED>(o.dL_dsh, wr, M * 3, 3 * k + 1, w * dL_dRGB.y, wpr, lerp, t, omt);
          put<FUSED>(o.dL_dsh, wr, M * 3, 3 * k + 2, w * dL_dRGB.z, wpr, lerp, t, omt);
        }
      }
    }

    // d colour / d direction
    float3 dx3 = {0, 0, 0}, dy3 = {0, 0, 0}, dz3 = {0, 0, 0};
    // lerped coefficients 1..15: whole rows with aligned vector loads when the layout allows (gsmath.cuh), else
    // coefficient by coefficient
    const bool rows_vec = FUSED && M == 16 && a.D == 3 &&
                          (split || (reinterpret_cast<size_t>(a.shs) & 15) == 0);
    float sv[48];
    if (rows_vec) {
      if (split) {
        row45(a.shs_rest + r * 45, [&](int i, float x) { sv[3 + i] = x; });
        if (lerp) row45(a.shs_rest + pr * 45, [&](int i, float x) { sv[3 + i] = lerp_rn(t, sv[3 + i], omt, x); });
      } else {
        const float4* c4 = reinterpret_cast<const float4*>(a.shs + r * 48);
        const float4* p4 = reinterpret_cast<const float4*>(a.shs + pr * 48);
#pragma unroll
        for (int i = 0; i < 12; ++i) {
          const float4 x = __ldg(c4 + i);
          sv[4 * i] = x.x; sv[4 * i + 1] = x.y; sv[4 * i + 2] = x.z; sv[4 * i + 3] = x.w;
        }
        if (lerp) {
#pragma unroll
          for (int i = 0; i < 12; ++i) {
            const float4 y = __ldg(p4 + i);
            sv[4 * i] = lerp_rn(t, sv[4 * i], omt, y.x);
            sv[4 * i + 1] = lerp_rn(t, sv[4 * i + 1], omt, y.y);
            sv[4 * i + 2] = lerp_rn(t, sv[4 * i + 2], omt, y.z);
            sv[4 * i + 3] = lerp_rn(t, sv[4 * i + 3], omt, y.w);
          }
        }
      }
    }
    auto S = [&](int k) -> float3 {   // k >= 1 only: `sh` is biased by one coefficient in split mode
      if (rows_vec) return make_float3(sv[3 * k], sv[3 * k + 1], sv[3 * k + 2]);
      float3 v = make_float3(__ldg(sh + 3 * k), __ldg(sh + 3 * k + 1), __ldg(sh + 3 * k + 2));
      if (FUSED && lerp) {
        v.x = lerp_rn(t, v.x, omt, __ldg(shp + 3 * k));
        v.y = lerp_rn(t, v.y, omt, __ldg(shp + 3 * k + 1));
        v.z = lerp_rn(t, v.z, omt, __ldg(shp + 3 * k + 2));
      }
      return v;
    };
    auto axpy = [](float3& acc, float s, const float3& v) {
      acc.x += s * v.x; acc.y += s * v.y; acc.z += s * v.z;
    };
    if (a.D > 0) {
      const float3 s1 = S(1), s2 = S(2), s3 = S(3);
      axpy(dx3, -kSH1, s3);
      axpy(dy3, -kSH1, s1);
      axpy(dz3, kSH1, s2);
      if (a.D > 1) {
        const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
        const float3 s4 = S(4), s5 = S(5), s6 = S(6), s7 = S(7), s8 = S(8);
        axpy(dx3, kSH2[0] * y, s4); axpy(dx3, kSH2[2] * 2.f * -x, s6); axpy(dx3, kSH2[3] * z, s7); axpy(dx3, kSH2[4] * 2.f * x, s8);
        axpy(dy3, kSH2[0] * x, s4); axpy(dy3, kSH2[1] * z, s5); axpy(dy3, kSH2[2] * 2.f * -y, s6); axpy(dy3, kSH2[4] * 2.f * -y, s8);
        axpy(dz3, kSH2[1] * y, s5); axpy(dz3, kSH2[2] * 2.f * 2.f * z, s6); axpy(dz3, kSH2[3] * x, s7);
        if (a.D > 2) {
          const float3 s9 = S(9), s10 = S(10), s11 = S(11), s12 = S(12), s13 = S(13), s14 = S(14), s15 = S(15);
          axpy(dx3, kSH3[0] * 3.f * 2.f * xy, s9); axpy(dx3, kSH3[1] * yz, s10); axpy(dx3, kSH3[2] * -2.f * xy, s11);
          axpy(dx3, kSH3[3] * -3.f * 2.f * xz, s12); axpy(dx3, kSH3[4] * (-3.f * xx + 4.f * zz - yy), s13);
          axpy(dx3, kSH3[5] * 2.f * xz, s14); axpy(dx3, kSH3[6] * 3.f * (xx - yy), s15);
          axpy(dy3, kSH3[0] * 3.f * (xx - yy), s9); axpy(dy3, kSH3[1] * xz, s10); axpy(dy3, kSH3[2] * (-3.f * yy + 4.f * zz - xx), s11);
          axpy(dy3, kSH3[3] * -3.f * 2.f * yz, s12); axpy(dy3, kSH3[4] * -2.f * xy, s13);
          axpy(dy3, kSH3[5] * -2.f * yz, s14); axpy(dy3, kSH3[6] * -3.f * 2.f * xy, s15);
          axpy(dz3, kSH3[1] * xy, s10); axpy(dz3, kSH3[2] * 4.f * 2.f * yz, s11); axpy(dz3, kSH3[3] * 3.f * (2.f * zz - xx - yy), s12);
          axpy(dz3, kSH3[4] * 4.f * 2.f * xz, s13); axpy(dz3, kSH3[5] * (xx - yy), s14);
        }
      }
    }
    const float3 dL_ddir = {dx3.x * dL_dRGB.x + dx3.y * dL_dRGB.y + dx3.z * dL_dRGB.z,
                            dy3.x * dL_dRGB.x + dy3.y * dL_dRGB.y + dy3.z * dL_dRGB.z,
                            dz3.x * dL_dRGB.x + dz3.y * dL_dRGB.y + dz3.z * dL_dRGB.z};
    const float3 dm = dnormvdv(dir_orig, dL_ddir);
    dL_dmean.x += dm.x; dL_dmean.y += dm.y; dL_dmean.z += dm.z;
    if (!FUSED && o.dL_dcolors) {
      o.dL_dcolors[r * 3] = 0.f; o.dL_dcolors[r * 3 + 1] = 0.f; o.dL_dcolors[r * 3 + 2] = 0.f;
    }
  } else {
    put<FUSED>(o.dL_dcolors, wr, 3, 0, dL_dRGB.x, wpr, lerp, t, omt);
    put<FUSED>(o.dL_dcolors, wr, 3, 1, dL_dRGB.y, wpr, lerp, t, omt);
    put<FUSED>(o.dL_dcolors, wr, 3, 2, dL_dRGB.z, wpr, lerp, t, omt);
    if (!FUSED && o.dL_dsh && !o.dL_dsh_rest) for (int i = 0; i < M * 3; ++i) o.dL_dsh[r * (size_t)M * 3 + i] = 0.f;
  }

  if (FUSED && o.dL_dmeans3D != nullptr) {
    const float sc = lerp ? t : 1.0f;
    red_row3(o.dL_dmeans3D + wr * 3, sc * dL_dmean.x, sc * dL_dmean.y, sc * dL_dmean.z);
    if (lerp) red_row3(o.dL_dmeans3D + wpr * 3, omt * dL_dmean.x, omt * dL_dmean.y, omt * dL_dmean.z);
  } else {
    put<FUSED>(o.dL_dmeans3D, wr, 3, 0, dL_dmean.x, wpr, lerp, t, omt);
    put<FUSED>(o.dL_dmeans3D, wr, 3, 1, dL_dmean.y, wpr, lerp, t, omt);
    put<FUSED>(o.dL_dmeans3D, wr, 3, 2, dL_dmean.z, wpr, lerp, t, omt);
  }
  put2<FUSED>(o.dL_dopacity, wr, 1, 0, tc * dL_dopacity_out * act_scalar_grad(a.act_opac, o_raw, o_act), wpr,
              lerp, omt * dL_dopacity_out * act_scalar_grad(a.act_opac, op_raw, op_act));

  // ---------------- 3D covariance -> scale / rotation (backward.cu:330-393) ----------------
  if (a.cov3D_pre == nullptr) {
    const float3 s_raw = ld3(a.scales, r);
    const float3 s_act = make_float3(act_scalar(a.act_scale, s_raw.x), act_scalar(a.act_scale, s_raw.y),
                                     act_scalar(a.act_scale, s_raw.z));
    float3 sc = s_act;
    const float4 q_raw = __ldg(reinterpret_cast<const float4*>(a.rots) + r);
    const float4 q4 = act_rotation(a.act_rot, a.norm_variant, q_raw);
    float qr = q4.x, qx = q4.y, qy = q4.z, qz = q4.w;
    float psign = 1.0f;
    float3 sp_raw = {0.f, 0.f, 0.f}, sp_act = {0.f, 0.f, 0.f};
    float4 qp_raw = {0.f, 0.f, 0.f, 0.f};
    if (FUSED && lerp) {
      sp_raw = ld3(a.scales, pr);
      sp_act = make_float3(act_scalar(a.act_scale, sp_raw.x), act_scalar(a.act_scale, sp_raw.y),
                           act_scalar(a.act_scale, sp_raw.z));
      const float3 s2 = sp_act;
      sc = make_float3(lerp_rn(t, sc.x, omt, s2.x), lerp_rn(t, sc.y, omt, s2.y), lerp_rn(t, sc.z, omt, s2.z));
      qp_raw = __ldg(reinterpret_cast<const float4*>(a.rots) + pr);
      float4 op = act_rotation(a.act_rot, a.norm_variant, qp_raw);
      float dotp = (qr * op.x + qx * op.y) + (qy * op.z + qz * op.w);
      if (dotp < 0.0f) { psign = -1.0f; op.x = -op.x; op.y = -op.y; op.z = -op.z; op.w = -op.w; }
      qr = lerp_rn(t, qr, omt, op.x);
      qx = lerp_rn(t, qx, omt, op.y);
      qy = lerp_rn(t, qy, omt, op.z);
      qz = lerp_rn(t, qz, omt, op.w);
    }
    const M3 R = quat_to_m3(qr, qx, qy, qz);
    const float sx = a.scale_modifier * sc.x, sy = a.scale_modifier * sc.y, sz = a.scale_modifier * sc.z;
    M3 S = m3(1.f, 0.f, 0.f, 0.f, 1.f, 0.f, 0.f, 0.f, 1.f);
    S.c[0][0] = sx; S.c[1][1] = sy; S.c[2][2] = sz;
    const M3 Mm = m3_mul(S, R);
    const M3 dSig = m3(dcov[0], 0.5f * dcov[1], 0.5f * dcov[2], 0.5f * dcov[1], dcov[3], 0.5f * dcov[4],
                       0.5f * dcov[2], 0.5f * dcov[4], dcov[5]);
    M3 dM = m3_mul(Mm, dSig);
#pragma unroll
    for (int j = 0; j < 3; ++j)
#pragma unroll
      for (int i = 0; i < 3; ++i) dM.c[j][i] *= 2.0f;
    const M3 Rt = m3_t(R);
    M3 dMt = m3_t(dM);
    // the reference stores the scale gradient without the scale_modifier factor
    const float ds0 = Rt.c[0][0] * dMt.c[0][0] + Rt.c[0][1] * dMt.c[0][1] + Rt.c[0][2] * dMt.c[0][2];
    const float ds1 = Rt.c[1][0] * dMt.c[1][0] + Rt.c[1][1] * dMt.c[1][1] + Rt.c[1][2] * dMt.c[1][2];
    const float ds2 = Rt.c[2][0] * dMt.c[2][0] + Rt.c[2][1] * dMt.c[2][1] + Rt.c[2][2] * dMt.c[2][2];
#pragma unroll
    for (int i = 0; i < 3; ++i) { dMt.c[0][i] *= sx; dMt.c[1][i] *= sy; dMt.c[2][i] *= sz; }
    const float x = qx, y = qy, z = qz, rr = qr;
    float4 dq;
    dq.x = 2 * z * (dMt.c[0][1] - dMt.c[1][0]) + 2 * y * (dMt.c[2][0] - dMt.c[0][2]) + 2 * x * (dMt.c[1][2] - dMt.c[2][1]);
    dq.y = 2 * y * (dMt.c[1][0] + dMt.c[0][1]) + 2 * z * (dMt.c[2][0] + dMt.c[0][2]) + 2 * rr * (dMt.c[1][2] - dMt.c[2][1]) - 4 * x * (dMt.c[2][2] + dMt.c[1][1]);
    dq.z = 2 * x * (dMt.c[1][0] + dMt.c[0][1]) + 2 * rr * (dMt.c[2][0] - dMt.c[0][2]) + 2 * z * (dMt.c[1][2] + dMt.c[2][1]) - 4 * y * (dMt.c[2][2] + dMt.c[0][0]);
    dq.w = 2 * rr * (dMt.c[0][1] - dMt.c[1][0]) + 2 * x * (dMt.c[2][0] + dMt.c[0][2]) + 2 * y * (dMt.c[1][2] + dMt.c[2][1]) - 4 * z * (dMt.c[1][1] + dMt.c[0][0]);

    const float gs0 = tc * ds0 * act_scalar_grad(a.act_scale, s_raw.x, s_act.x);
    const float gs1 = tc * ds1 * act_scalar_grad(a.act_scale, s_raw.y, s_act.y);
    const float gs2 = tc * ds2 * act_scalar_grad(a.act_scale, s_raw.z, s_act.z);
    const float gp0 = omt * ds0 * act_scalar_grad(a.act_scale, sp_raw.x, sp_act.x);
    const float gp1 = omt * ds1 * act_scalar_grad(a.act_scale, sp_raw.y, sp_act.y);
    const float gp2 = omt * ds2 * act_scalar_grad(a.act_scale, sp_raw.z, sp_act.z);
    // dq is in (r, x, y, z) = float4 (x, y, z, w) component order, like the stored quaternion
    const float4 gc = act_rotation_grad(a.act_rot, a.norm_variant, q_raw,
                                        make_float4(tc * dq.x, tc * dq.y, tc * dq.z, tc * dq.w));
    float4 gp = {0.f, 0.f, 0.f, 0.f};
    if (FUSED && lerp) {
      const float w = omt * psign;
      gp = act_rotation_grad(a.act_rot, a.norm_variant, qp_raw, make_float4(w * dq.x, w * dq.y, w * dq.z, w * dq.w));
    }
    if (FUSED) {
      if (o.dL_dscales != nullptr) {
        red_row3(o.dL_dscales + wr * 3, gs0, gs1, gs2);
        if (lerp) red_row3(o.dL_dscales + wpr * 3, gp0, gp1, gp2);
      }
      if (o.dL_drotations != nullptr) {
        red4(o.dL_drotations + wr * 4, gc.x, gc.y, gc.z, gc.w);
        if (lerp) red4(o.dL_drotations + wpr * 4, gp.x, gp.y, gp.z, gp.w);
      }
    } else {
      put2<FUSED>(o.dL_dscales, wr, 3, 0, gs0, wpr, lerp, gp0);
      put2<FUSED>(o.dL_dscales, wr, 3, 1, gs1, wpr, lerp, gp1);
      put2<FUSED>(o.dL_dscales, wr, 3, 2, gs2, wpr, lerp, gp2);
      put2<FUSED>(o.dL_drotations, wr, 4, 0, gc.x, wpr, lerp, gp.x);
      put2<FUSED>(o.dL_drotations, wr, 4, 1, gc.y, wpr, lerp, gp.y);
      put2<FUSED>(o.dL_drotations, wr, 4, 2, gc.z, wpr, lerp, gp.z);
      put2<FUSED>(o.dL_drotations, wr, 4, 3, gc.w, wpr, lerp, gp.w);
    }
    if (!FUSED && o.dL_dcov3D) for (int i = 0; i < 6; ++i) o.dL_dcov3D[r * 6 + i] = dcov[i];
  } else {
    for (int i = 0; i < 6; ++i) put<FUSED>(o.dL_dcov3D, wr, 6, i, dcov[i], wpr, lerp, t, omt);
    if (!FUSED) {
      if (o.dL_dscales) { o.dL_dscales[r * 3] = 0.f; o.dL_dscales[r * 3 + 1] = 0.f; o.dL_dscales[r * 3 + 2] = 0.f; }
      if (o.dL_drotations) *reinterpret_cast<float4*>(o.dL_drotations + r * 4) = make_float4(0.f, 0.f, 0.f, 0.f);
    }
  }
}

// Compact gradients: zero the first U rows of every compact tensor (U lives on the device), and afterwards put the
// row map back to all -1 by visiting the U rows that were claimed.
struct CompactZero {
  float* base[8];
  int width[8];
  int n;
};
__global__ void __launch_bounds__(256) zero_compact_kernel(const CompactZero z, const int* __restrict__ count,
                                                           int capacity) {
  const long long U = min(*count, capacity);
  const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (int k = 0; k < z.n; ++k) {
    // rows [0, U) of a compact tensor are one contiguous span; the tensors come from the caching allocator
    // (512-byte aligned), so the span starts on a 16-byte boundary
    const long long n = U * z.width[k];
    float4* q = reinterpret_cast<float4*>(z.base[k]);
    const long long n4 = n >> 2;
    for (long long i = tid; i < n4; i += stride) q[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (tid < (n & 3)) z.base[k][n4 * 4 + tid] = 0.f;
  }
}
__global__ void __launch_bounds__(256) reset_row_map_kernel(int* __restrict__ row_map,
                                                            const long long* __restrict__ rows,
                                                            const int* __restrict__ count, int capacity) {
  const long long U = min(*count, capacity);
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < U; i += stride) row_map[rows[i]] = -1;
}

}  // namespace

cudaError_t launch_preprocess_bwd(const PreBwdArgs& b, cudaStream_t s) {
  if (b.f.P <= 0) return cudaSuccess;
  const int blocks = (b.f.P + 255) / 256;
  ProfScope ps(K_PREPROCESS_BWD, s);
  if (b.f.idx != nullptr && b.out.row_map != nullptr) {
    const H3dgsGrads& o = b.out;
    list_visible_kernel<true><<<blocks, 256, 0, s>>>(b);
    CompactZero z;
    z.n = 0;
    const int M = b.f.M;
    const bool split = b.f.shs_rest != nullptr;
    auto add = [&](float* p, int w) { if (p != nullptr && w > 0) { z.base[z.n] = p; z.width[z.n] = w; ++z.n; } };
    add(o.dL_dmeans3D, 3); add(o.dL_dsh, split ? 3 : M * 3); add(o.dL_dsh_rest, split ? (M - 1) * 3 : 0);
    add(o.dL_dcolors, 3); add(o.dL_dopacity, 1); add(o.dL_dscales, 3); add(o.dL_drotations, 4); add(o.dL_dcov3D, 6);
    zero_compact_kernel<<<148 * 8, 256, 0, s>>>(z, o.compact_count, o.compact_capacity);
    preprocess_bwd_kernel<true><<<blocks, 256, 0, s>>>(b);
    reset_row_map_kernel<<<148 * 4, 256, 0, s>>>(o.row_map, reinterpret_cast<const long long*>(o.compact_rows),
                                                 o.compact_count, o.compact_capacity);
  } else if (b.f.idx != nullptr) {
    list_visible_kernel<true><<<blocks, 256, 0, s>>>(b);
    preprocess_bwd_kernel<true><<<blocks, 256, 0, s>>>(b);
  } else {
    list_visible_kernel<false><<<blocks, 256, 0, s>>>(b);
    preprocess_bwd_kernel<false><<<blocks, 256, 0, s>>>(b);
  }
  return cudaGetLastError();
}

}  // namespace h3dgs
