// Backward tile alpha blend, two pixels per lane.  Replaces renderCUDA of HR/cuda_rasterizer/backward.cu:498-721;
// same mathematics, work items, snapshots and staging as blend_bwd.cu (read that file's header first).
//
// Why a second formulation: blend_bwd.cu is bound by issue slots, and a third of its ~130 instructions per
// (warp, surviving Gaussian) is the cross-lane reduction of the ten gradient sums (12 shuffles, 12 adds, 18 selects)
// -- a cost per (warp, Gaussian), not per pixel.  Here a warp owns an 8x8 pixel block and every lane two pixels
// (rows r and r + 4 of the block): the two pixels' ten terms are added in registers first and ONE butterfly serves
// 64 pixels.  The two pixel chains of a lane are independent, so the warp also carries twice the instruction-level
// parallelism; a CTA is 4 warps (128 threads) per 16x16 tile.  Culling is per 8x8 block (coarser than the 8x4 blocks
// of blend_bwd.cu: ~19 % more pairs are evaluated, ~40 % fewer instructions are issued).
#include "blend_common.cuh"

namespace h3dgs {

namespace {

constexpr int B2_THREADS = 128;
constexpr int B2_BATCH = 128;      // list entries staged per round (one per thread)
constexpr int B2_SUB = 8;          // 8x8 pixel block per warp
constexpr int ACC_W = 12;
#ifndef B2_MIN_CTAS
#define B2_MIN_CTAS 9   // resident CTAs per SM the register allocation aims at (56 registers, 8 bytes spilled): 1.32 -> 1.27 ms; 10 (51 registers) is slower (1.35)
#endif

struct Bwd2Smem {
  Rec rec[2][B2_BATCH];
  float acc[B2_BATCH][ACC_W];
  int red[B2_THREADS / 32];
};

// block_hit of blend_common.cuh for an 8x8 block (see there for the derivation)
__device__ __forceinline__ bool block_hit8(const float4& q0, const float4& q1, float x0, float y0) {
  const float X0 = x0 - q0.x, X1 = X0 + (float)(B2_SUB - 1);
  const float Y0 = y0 - q0.y, Y1 = Y0 + (float)(B2_SUB - 1);
  const float dxn = fmaxf(X0, fminf(0.0f, X1));
  const float dyn = fmaxf(Y0, fminf(0.0f, Y1));
  const float a = q1.x, b = q1.y, c = q1.z;
  const float dys = fmaxf(Y0, fminf(__fdividef(-b * dxn, c), Y1));
  const float dxs = fmaxf(X0, fminf(__fdividef(-b * dyn, a), X1));
  const float sa = a * dxn * dxn + c * dys * dys, ta = 2.0f * b * dxn * dys;
  const float sb = a * dxs * dxs + c * dyn * dyn, tb = 2.0f * b * dxs * dyn;
  const float qa = fmaf(-4e-7f, sa + fabsf(ta), sa + ta);
  const float qb = fmaf(-4e-7f, sb + fabsf(tb), sb + tb);
  return !(fminf(qa, qb) > q0.z);
}

// The per-pixel part of one surviving list entry, in two phases so that a survivor no lane contributes to can be
// dropped after the first: (A) alpha with the reference's backward arithmetic and the keep / skip decision,
// (B) the walking state (T, B) advanced and the pixel's ten gradient terms ADDED into v[].  Identical, instruction
// for instruction, to the body of blend_bwd.cu's survivor loop.
struct PixA {
  float dx, dy, G, test_alpha, alpha, opac_mult;
  bool contrib;
};

template <bool INTERP>
__device__ __forceinline__ PixA pixel_alpha(const float2 xy, const float4 con_o, const float2 tk, float pixx, float pixy,
                                            bool live, int& cls) {
  PixA r;
  r.dx = xy.x - pixx;
  r.dy = xy.y - pixy;
  const float power = blend_power(con_o, r.dx, r.dy);
  r.G = ex2_approx(power * 1.4426950408889634f);
  r.test_alpha = con_o.w * r.G;
  const float my_alpha = min(0.99f, r.test_alpha);
  float alpha = my_alpha;
  r.opac_mult = 1.0f;
  if (INTERP) {
    if (tk.x != 1.0f) {
      const float om = 1.0f - my_alpha;
      float pw = ex2_approx(tk.y * lg2_approx(om));
      alpha = tk.x * my_alpha + (1.0f - tk.x) * (1.0f - pw);
      if (fabsf(alpha - 1.0f / 255.0f) < 1e-6f) {   // the keep / skip decision must be the reference's
        pw = powf(1.0f - min(0.99f, con_o.w * expf(power)), tk.y);
        alpha = tk.x * min(0.99f, con_o.w * expf(power)) + (1.0f - tk.x) * (1.0f - pw);
      }
      r.opac_mult = tk.x - pw * __frcp_approx(om) * (tk.x - 1.0f) * tk.y;
    } else if (fabsf(alpha - 1.0f / 255.0f) < 2e-8f) {
      alpha = min(0.99f, con_o.w * expf(power));
    }
  } else if (fabsf(alpha - 1.0f / 255.0f) < 2e-8f) {
    alpha = min(0.99f, con_o.w * expf(power));
  }
  r.alpha = alpha;
  r.contrib = live & !(power > 0.0f) & !(alpha < 1.0f / 255.0f);
  cls = !live ? 0 : (power > 0.0f ? 1 : (!r.contrib ? 2 : 3));
  return r;
}

template <bool DEPTH, bool FIRST>
__device__ __forceinline__ void pixel_terms(const PixA& r, const float4 con_o, const float4 q2, float g0, float g1,
                                            float g2, float gD, float Tfb, float& T, float& B, float v[10]) {
  const float a_eff = r.contrib ? r.alpha : 0.f;
  const float om_a = 1.f - a_eff;
  float rcp_om = __frcp_approx(om_a);
  rcp_om = fmaf(fmaf(-om_a, rcp_om, 1.0f), rcp_om, rcp_om);
  T *= rcp_om;
  float u = q2.x * g0 + q2.y * g1 + q2.z * g2;
  if (DEPTH) u += q2.w * gD;
  const float d = u - B;
  float dL_dalpha = fmaf(d, T, Tfb * rcp_om);
  B = fmaf(a_eff, d, B);
  if (r.test_alpha > 0.99f) dL_dalpha = 0.f;
  const float Gc = r.contrib ? r.G : 0.f;
  const float w = a_eff * T;
  const float GdA = Gc * dL_dalpha;
  const float q = con_o.w * GdA;
  const float v0 = q * r.dx, v1 = q * r.dy;
  if (FIRST) {
    v[0] = v0; v[1] = v1; v[2] = v0 * r.dx; v[3] = v0 * r.dy; v[4] = v1 * r.dy; v[5] = r.opac_mult * GdA;
    v[6] = w * g0; v[7] = w * g1; v[8] = w * g2; v[9] = DEPTH ? w * gD : 0.f;
  } else {
    v[0] += v0; v[1] += v1; v[2] = fmaf(v0, r.dx, v[2]); v[3] = fmaf(v0, r.dy, v[3]); v[4] = fmaf(v1, r.dy, v[4]);
    v[5] = fmaf(r.opac_mult, GdA, v[5]);
    v[6] = fmaf(w, g0, v[6]); v[7] = fmaf(w, g1, v[7]); v[8] = fmaf(w, g2, v[8]);
    if (DEPTH) v[9] = fmaf(w, gD, v[9]);
  }
}

template <bool INTERP, bool DEPTH, bool COUNT = false>
__global__ void __launch_bounds__(B2_THREADS, B2_MIN_CTAS) blend_bwd2_kernel(const BlendArgs a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  Bwd2Smem& sm = *reinterpret_cast<Bwd2Smem*>(smem_raw);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  int tile, k = 0;
  if ((int)blockIdx.x < a.nseg) {   // work items that start at a list cut lead the grid (blend_bwd.cu)
    const unsigned x = blockIdx.x + 1;
    tile = (int)a.seg_tile[x];
    const uint2 r = a.ranges[tile];
    k = (int)x - (int)(r.x / BLEND_SEG);
    if (k < 1 || r.x + (unsigned)k * BLEND_SEG >= r.y) return;
  } else {
    const int i = (int)blockIdx.x - a.nseg;
    tile = a.order ? (int)a.order[i] : i;
  }
  const int tx = tile % a.grid_x, ty = tile / a.grid_x;
  const int lx = (warp & 1) * B2_SUB + (lane & 7);          // pixel column inside the tile
  const int ly0 = (warp >> 1) * B2_SUB + (lane >> 3);       // row of this lane's first pixel; the second is 4 below
  const int bx0 = tx * TILE_X + (warp & 1) * B2_SUB, by0 = ty * TILE_Y + (warp >> 1) * B2_SUB;
  const int px = tx * TILE_X + lx;
  const float pixx = (float)px;
  const float fx0 = (float)bx0, fy0 = (float)by0;
  const size_t plane = (size_t)a.W * a.H;

  unsigned long long t_start = 0;
  if (COUNT) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_start));
  const uint2 range = a.ranges[tile];
  const int n = (int)(range.y - range.x);
  const uint32_t* __restrict__ list = a.point_list + range.x;
  constexpr int REC_CHUNKS = INTERP ? 4 : 3;

  int last[2];
  float g0[2], g1[2], g2[2], gD[2], Tfb[2], T[2], B[2], pixy[2];
  size_t pixi[2];
  bool inside[2];
#pragma unroll
  for (int p = 0; p < 2; ++p) {
    const int py = ty * TILE_Y + ly0 + 4 * p;
    pixy[p] = (float)py;
    inside[p] = px < a.W && py < a.H;
    pixi[p] = (size_t)py * a.W + px;
    last[p] = 0;
    g0[p] = g1[p] = g2[p] = gD[p] = 0.f;
    float T_final = 0.f, bgdot = 0.f;
    if (inside[p]) {
      last[p] = (int)a.n_contrib[pixi[p]];
      g0[p] = a.dL_dcolor[pixi[p]];
      g1[p] = a.dL_dcolor[plane + pixi[p]];
      g2[p] = a.dL_dcolor[2 * plane + pixi[p]];
      if (DEPTH) gD[p] = a.dL_dinvdepth[pixi[p]];
      T_final = a.final_T[pixi[p]];
      bgdot = a.bg[0] * g0[p] + a.bg[1] * g1[p] + a.bg[2] * g2[p];
    }
    T[p] = T_final;
    B[p] = 0.f;
    Tfb[p] = -T_final * bgdot;
  }
  int wlast = max(last[0], last[1]);
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) wlast = max(wlast, __shfl_xor_sync(0xffffffffu, wlast, d));
  if (lane == 0) sm.red[warp] = wlast;
  for (int i = tid; i < B2_BATCH * ACC_W; i += B2_THREADS) (&sm.acc[0][0])[i] = 0.f;
  __syncthreads();
  int blast = 0;
#pragma unroll
  for (int w = 0; w < B2_THREADS / 32; ++w) blast = max(blast, sm.red[w]);
  const int seg_lo = k * BLEND_SEG;
  if (seg_lo >= min(blast, n)) return;
  const int nn = min(seg_lo + BLEND_SEG, min(blast, n));
  const int len = nn - seg_lo;
  const int rounds = (len + B2_BATCH - 1) / B2_BATCH;

  auto issue = [&](int b, uint32_t id) {
    const int j = b * B2_BATCH + tid;
    const int s = b & 1;
    if (j < len) {
      const char* src = reinterpret_cast<const char*>(a.rec + id);
      char* dst = reinterpret_cast<char*>(&sm.rec[s][tid]);
#pragma unroll
      for (int c = 0; c < REC_CHUNKS; ++c) cp_async16(dst + 16 * c, src + 16 * c);
    }
    cp_async_commit();
  };

  int red_dst = -1;   // which of the ten sums is complete in this lane after the transposed butterfly (blend_bwd.cu)
  if ((lane & 1) == 0) {
    const int b4 = (lane >> 4) & 1, b3 = (lane >> 3) & 1, b2 = (lane >> 2) & 1, b1 = (lane >> 1) & 1;
    if (b1) {
      if (!b2 && !b3) red_dst = b4 * 5 + 2;
    } else {
      red_dst = b4 * 5 + (b2 ? (b3 ? 4 : 1) : (b3 ? 3 : 0));
    }
    if (!DEPTH && red_dst == 9) red_dst = -1;
  }
  const uint32_t acc_addr = smem_u32(&sm.acc[0][red_dst < 0 ? 0 : red_dst]);

#pragma unroll
  for (int p = 0; p < 2; ++p) {
    if (inside[p] && last[p] > nn) {
      // this pixel's walk started in a deeper item: pick it up at the cut from the forward's snapshot, which is
      // indexed by the FORWARD kernel's thread id of the pixel (8x4 blocks: warp = (row / 4) * 2 + column / 8)
      const int ly = ly0 + 4 * p;
      const int ftid = (((ly >> 2) * 2 + (lx >> 3)) << 5) + ((ly & 3) << 3) + (lx & 7);
      const float* __restrict__ sp = a.snap + ((size_t)(range.x / BLEND_SEG) + (size_t)(k + 1)) * SNAP_FLOATS + ftid;
      const float4 tot = a.accum[pixi[p]];
      T[p] = sp[0];
      float behind = (tot.x - sp[TILE_PIX]) * g0[p] + (tot.y - sp[2 * TILE_PIX]) * g1[p] + (tot.z - sp[3 * TILE_PIX]) * g2[p];
      if (DEPTH) behind += (tot.w - sp[4 * TILE_PIX]) * gD[p];
      B[p] = behind / T[p];
    }
  }
  unsigned cntA = 0, cntB = 0, cntD = 0, cntS = 0;
  const float half_w = 0.5f * (float)a.W, half_h = 0.5f * (float)a.H;

  uint32_t id_cur = (tid < len) ? __ldg(list + (nn - 1 - tid)) : 0u;
  issue(0, id_cur);
  uint32_t id_next = (B2_BATCH + tid < len) ? __ldg(list + (nn - 1 - B2_BATCH - tid)) : 0u;

  for (int b = 0; b < rounds; ++b) {
    const int s = b & 1;
    uint32_t id_after = 0;
    if (b + 1 < rounds) {
      issue(b + 1, id_next);
      const int j2 = (b + 2) * B2_BATCH + tid;
      id_after = (j2 < len) ? __ldg(list + (nn - 1 - j2)) : 0u;
      cp_async_wait<1>();
    } else {
      cp_async_wait<0>();
    }
    __syncthreads();

    const int base = b * B2_BATCH;
    const int cnt = min(B2_BATCH, len - base);
    const int first = max(0, nn - min(wlast, nn) - base);
    const Rec* __restrict__ recs = sm.rec[s];
    for (int c = first & ~31; c < cnt; c += 32) {
      const int j = c + lane;
      bool hit = false;
      if (j >= first && j < cnt) hit = block_hit8(recs[j].q0, recs[j].q1, fx0, fy0);
      unsigned m = __ballot_sync(0xffffffffu, hit);
      while (m) {
        const int jj = c + __ffs(m) - 1;
        m &= m - 1;
        const float2 xy = *reinterpret_cast<const float2*>(&recs[jj].q0);
        const float4 con_o = recs[jj].q1;
        float2 tk = make_float2(1.0f, 1.0f);
        if (INTERP) tk = *reinterpret_cast<const float2*>(&recs[jj].q3);
        const int pos = nn - 1 - (base + jj);
        int c0, c1;
        const PixA A0 = pixel_alpha<INTERP>(xy, con_o, tk, pixx, pixy[0], pos < last[0], c0);
        const PixA A1 = pixel_alpha<INTERP>(xy, con_o, tk, pixx, pixy[1], pos < last[1], c1);
        if (COUNT) {
          cntA += (c0 == 1) + (c1 == 1); cntB += (c0 == 2) + (c1 == 2); cntD += (c0 == 3) + (c1 == 3);
          cntS += 2;
        }
        if (!__any_sync(0xffffffffu, A0.contrib | A1.contrib)) continue;
        const float4 q2 = recs[jj].q2;
        float v[10];
        pixel_terms<DEPTH, true>(A0, con_o, q2, g0[0], g1[0], g2[0], gD[0], Tfb[0], T[0], B[0], v);
        pixel_terms<DEPTH, false>(A1, con_o, q2, g0[1], g1[1], g2[1], gD[1], Tfb[1], T[1], B[1], v);
        const bool h16 = lane & 16, h8 = lane & 8, h4 = lane & 4, h2 = lane & 2;
        float r0 = (h16 ? v[5] : v[0]) + __shfl_xor_sync(0xffffffffu, h16 ? v[0] : v[5], 16);
        float r1 = (h16 ? v[6] : v[1]) + __shfl_xor_sync(0xffffffffu, h16 ? v[1] : v[6], 16);
        float r2 = (h16 ? v[7] : v[2]) + __shfl_xor_sync(0xffffffffu, h16 ? v[2] : v[7], 16);
        float r3 = (h16 ? v[8] : v[3]) + __shfl_xor_sync(0xffffffffu, h16 ? v[3] : v[8], 16);
        float r4 = (h16 ? v[9] : v[4]) + __shfl_xor_sync(0xffffffffu, h16 ? v[4] : v[9], 16);
        float s0 = (h8 ? r3 : r0) + __shfl_xor_sync(0xffffffffu, h8 ? r0 : r3, 8);
        float s1 = (h8 ? r4 : r1) + __shfl_xor_sync(0xffffffffu, h8 ? r1 : r4, 8);
        float s2 = r2 + __shfl_xor_sync(0xffffffffu, r2, 8);
        float t0 = (h4 ? s1 : s0) + __shfl_xor_sync(0xffffffffu, h4 ? s0 : s1, 4);
        float t1 = s2 + __shfl_xor_sync(0xffffffffu, s2, 4);
        float uu = (h2 ? t1 : t0) + __shfl_xor_sync(0xffffffffu, h2 ? t0 : t1, 2);
        uu += __shfl_xor_sync(0xffffffffu, uu, 1);
        if (red_dst >= 0)
          asm volatile("red.shared.add.f32 [%0], %1;" ::"r"(acc_addr + (uint32_t)jj * (ACC_W * 4u)), "f"(uu) : "memory");
      }
    }
    __syncthreads();

    if (tid < cnt) {   // flush: thread t owns list entry base + t
      float4 s0 = *reinterpret_cast<float4*>(&sm.acc[tid][0]);
      float4 s1 = *reinterpret_cast<float4*>(&sm.acc[tid][4]);
      float2 s2 = *reinterpret_cast<float2*>(&sm.acc[tid][8]);
      const bool any = (s0.x != 0.f) | (s0.y != 0.f) | (s0.z != 0.f) | (s0.w != 0.f) | (s1.x != 0.f) |
                       (s1.y != 0.f) | (s1.z != 0.f) | (s1.w != 0.f) | (s2.x != 0.f) | (s2.y != 0.f);
      if (any) {
        const float4 con_o = recs[tid].q1;
        const float mx = -(s0.x * con_o.x + s0.y * con_o.y) * half_w;
        const float my = -(s0.y * con_o.z + s0.x * con_o.y) * half_h;
        float* gr = a.grad[id_cur].v;
        red_add_v4(gr, mx, my, -0.5f * s0.z, -0.5f * s0.w);
        red_add_v4(gr + 4, -0.5f * s1.x, s1.y, s1.z, s1.w);
        red_add_v2(gr + 8, s2.x, s2.y);
        *reinterpret_cast<float4*>(&sm.acc[tid][0]) = make_float4(0.f, 0.f, 0.f, 0.f);
        *reinterpret_cast<float4*>(&sm.acc[tid][4]) = make_float4(0.f, 0.f, 0.f, 0.f);
        *reinterpret_cast<float2*>(&sm.acc[tid][8]) = make_float2(0.f, 0.f);
      }
    }
    id_cur = id_next;
    id_next = id_after;
  }
  if (COUNT && a.stats) {
    if (tid == 0 && k == 0) {
      unsigned long long t_end;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_end));
      a.stats[16 + 4 * tile + 2] = t_start;
      a.stats[16 + 4 * tile + 3] = t_end;
    }
    const unsigned vv[4] = {cntA, cntB, cntD, cntS};
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const unsigned t = __reduce_add_sync(0xffffffffu, vv[q]);
      if (lane == 0 && t) atomicAdd(a.stats + 5 + q, (unsigned long long)t);
    }
  }
}

template <bool INTERP, bool DEPTH>
cudaError_t launch2(const BlendArgs& a, cudaStream_t s) {
  const int tiles = a.grid_x * a.grid_y + a.nseg;
  const size_t smem = sizeof(Bwd2Smem);
  static bool attr_done = false;   // one flag per instantiation
  if (!attr_done) {
    cudaError_t e = cudaFuncSetAttribute(blend_bwd2_kernel<INTERP, DEPTH, false>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(blend_bwd2_kernel<INTERP, DEPTH, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             (int)smem);
    if (e != cudaSuccess) return e;
    attr_done = true;
  }
  ProfScope ps(K_BLEND_BWD, s);
  if (a.stats) blend_bwd2_kernel<INTERP, DEPTH, true><<<tiles, B2_THREADS, smem, s>>>(a);
  else blend_bwd2_kernel<INTERP, DEPTH, false><<<tiles, B2_THREADS, smem, s>>>(a);
  return cudaGetLastError();
}

}  // namespace

cudaError_t launch_blend_bwd2(const BlendArgs& a, cudaStream_t s) {
  const bool depth = a.dL_dinvdepth != nullptr;
  if (a.interp) return depth ? launch2<true, true>(a, s) : launch2<true, false>(a, s);
  return depth ? launch2<false, true>(a, s) : launch2<false, false>(a, s);
}

}  // namespace h3dgs
