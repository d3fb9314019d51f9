// Shared-memory staging of per-slot records for the tile blend kernels, and the per-pair alpha
// evaluation used identically by the forward and backward passes.
#pragma once
#include "common.cuh"

namespace h3dgs {

constexpr int BLEND_THREADS = 256;
constexpr int BLEND_BATCH = 256;   // list entries staged per round (HR/cuda_rasterizer/forward.cu:478-512 uses 256 too)
constexpr int SUB_W = 8;           // each warp owns an 8x4 pixel block of the 16x16 tile
constexpr int SUB_H = 4;

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}

// ---- Ampere-style async copy (LDGSTS), 16 bytes per instruction ----
__device__ __forceinline__ void cp_async16(void* dst_smem, const void* src_gmem) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(smem_u32(dst_smem)), "l"(src_gmem)
               : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;\n" ::"n"(N) : "memory");
}

// ---- bulk async copy engine (TMA, 1-D): one instruction moves a whole record, completion by mbarrier ----
__device__ __forceinline__ void mbar_init(unsigned long long* bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() {
  asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long* bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(unsigned long long* bar, unsigned parity) {
  unsigned ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}\n"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, unsigned parity) {
  unsigned spins = 0;
  while (!mbar_try_wait(bar, parity)) {
    if (++spins > (1u << 26)) __trap();  // never hang the GPU on a lost transaction
  }
}
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, unsigned bytes,
                                         unsigned long long* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n" ::"r"(
          smem_u32(dst_smem)),
      "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}

// vector reduction to global memory (sm_90+): one instruction adds four floats
__device__ __forceinline__ void red_add_v4(float* addr, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};\n" ::"l"(addr), "f"(a), "f"(b), "f"(c),
               "f"(d)
               : "memory");
}
__device__ __forceinline__ void red_add_v2(float* addr, float a, float b) {
  asm volatile("red.global.add.v2.f32 [%0], {%1, %2};\n" ::"l"(addr), "f"(a), "f"(b) : "memory");
}

// MUFU.RCP without the range handling of a full division (callers refine with a Newton step)
__device__ __forceinline__ float __frcp_approx(float x) {
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}

__device__ __forceinline__ float ex2_approx(float x) {
  float r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}
__device__ __forceinline__ float lg2_approx(float x) {
  float r;
  asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}

// power = -0.5 (a dx^2 + c dy^2) - b dx dy with the roundings of the reference build (HR/cuda_rasterizer/forward.cu:529,
// backward.cu:611 as nvcc contracts them for sm_100a: a dx^2 is fused into the rounded c dy^2, the cross term is
// rounded on its own).  Written out because the keep / skip decision at alpha = 1/255 hangs on the last bit, and
// nvcc's own choice of which product to fuse depends on the surrounding code (it fused the other one in
// blend_bwd2.cu: two Gaussians of a 100 k scene came out 10 % different).
__device__ __forceinline__ float blend_power(const float4& con_o, float dx, float dy) {
  const float s = __fmaf_rn(dx, __fmul_rn(dx, con_o.x), __fmul_rn(dy, __fmul_rn(dy, con_o.z)));
  return __fsub_rn(__fmul_rn(s, -0.5f), __fmul_rn(dy, __fmul_rn(dx, con_o.y)));
}

// Can the Gaussian of record (q0, q1) reach alpha >= 1/255 anywhere in the pixel block
// [x0, x0+SUB_W-1] x [y0, y0+SUB_H-1]?  alpha >= 1/255  <=>  a dx^2 + 2 b dx dy + c dy^2 <= 2 ln(255 o) = q0.z
// (preprocess stores that bound with a small safety margin, +inf for "never cull", negative for "never
// visible").  The left side is minimised over the block exactly: the minimiser of a convex quadratic over a
// rectangle that does not contain its centre lies on an edge facing the centre, so it is the smaller of the
// two one-dimensional minima on the vertical line x = nearest x and on the horizontal line y = nearest y.
// One lane runs this for one list entry on behalf of its whole warp, so it costs ~1 instruction per pair.
__device__ __forceinline__ bool block_hit(const float4& q0, const float4& q1, float x0, float y0) {
  const float X0 = x0 - q0.x, X1 = X0 + (float)(SUB_W - 1);
  const float Y0 = y0 - q0.y, Y1 = Y0 + (float)(SUB_H - 1);
  const float dxn = fmaxf(X0, fminf(0.0f, X1));   // nearest block point, per axis (0 when the centre is inside)
  const float dyn = fmaxf(Y0, fminf(0.0f, Y1));
  const float a = q1.x, b = q1.y, c = q1.z;
  const float dys = fmaxf(Y0, fminf(__fdividef(-b * dxn, c), Y1));   // best dy on the line dx = dxn
  const float dxs = fmaxf(X0, fminf(__fdividef(-b * dyn, a), X1));   // best dx on the line dy = dyn
  // q = s + t with s = a dx^2 + c dy^2 >= 0 and the cross term t; the per-pixel evaluation the blend loop (and the
  // reference) performs rounds each product, i.e. it can come out lower than the exact value by a few ulp of
  // s + |t| -- for splats thousands of pixels long and sub-pixel thin that exceeds any fixed margin on q0.z, so
  // the bound is lowered by 4e-7 (s + |t|) before the comparison.
  const float sa = a * dxn * dxn + c * dys * dys, ta = 2.0f * b * dxn * dys;
  const float sb = a * dxs * dxs + c * dyn * dyn, tb = 2.0f * b * dxs * dyn;
  const float qa = fmaf(-4e-7f, sa + fabsf(ta), sa + ta);
  const float qb = fmaf(-4e-7f, sb + fabsf(tb), sb + tb);
  return !(fminf(qa, qb) > q0.z);   // NaN anywhere -> keep the entry, the blend itself decides
}

}  // namespace h3dgs
