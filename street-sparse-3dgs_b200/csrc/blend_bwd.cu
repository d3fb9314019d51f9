// Backward tile alpha blend.  Replaces renderCUDA of HR/cuda_rasterizer/backward.cu:498-721.
//
// Same tiling, staging and per-warp culling as the forward kernel.  Each pixel walks its list back to
// front from its last contributor, rebuilding T by division from the stored final transmittance and
// re-evaluating alpha with the reference's backward arithmetic (expf, and precise powf on interpolated
// slots -- the forward used __powf; that asymmetry is the reference's and is part of its gradients).
// The per-channel recursion "colour accumulated behind me" is carried as one scalar,
//     u_i = c_i.g + invdepth_i gD,   A_i = alpha_{i+1} u_{i+1} + (1 - alpha_{i+1}) A_{i+1},
//     dL/dalpha_i = (u_i - A_i) T_i - T_final/(1-alpha_i) bg.g ,
// which is the same linear recursion as backward.cu:657-691 contracted with the pixel gradient.
//
// Gradient reduction: per (warp, Gaussian) the ten partial sums are reduced with warp shuffles, added
// into a per-CTA shared-memory accumulator, and flushed once per (tile, Gaussian) with three vector
// red.global.add instructions -- instead of the reference's ten global atomics per (pixel, Gaussian).
#include <cstdlib>

#include "blend_common.cuh"

namespace h3dgs {

namespace {

constexpr int ACC_W = 12;  // 10 sums, padded to a multiple of four words

struct BwdSmem {
  Rec rec[2][BLEND_BATCH];
  float acc[BLEND_BATCH][ACC_W];
  unsigned long long bar[2];
  int red[BLEND_THREADS / 32];
};

// COUNT: instrumented pass -- stats[5] live pairs rejected at power > 0, [6] at alpha < 1/255, [7] contributing,
// [8] lane slots issued.
template <bool INTERP, bool DEPTH, bool TMA, bool COUNT = false>
__global__ void __launch_bounds__(BLEND_THREADS) blend_bwd_kernel(const BlendArgs a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  BwdSmem& sm = *reinterpret_cast<BwdSmem*>(smem_raw);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  // Work item = (tile, k): list positions [k SEG, (k+1) SEG) of the tile.  Items with k >= 1 lead the grid, one
  // CTA per possible cut number x = floor(global list position / SEG): the tile that owns list position x SEG
  // is read off the sorted keys, and the item exists if that tile has a cut inside block x.
  int tile, k = 0;
  if ((int)blockIdx.x < a.nseg) {
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
  const int bx0 = tx * TILE_X + (warp & 1) * SUB_W;
  const int by0 = ty * TILE_Y + (warp >> 1) * SUB_H;
  const int px = bx0 + (lane & 7), py = by0 + (lane >> 3);
  const bool inside = px < a.W && py < a.H;
  const float pixx = (float)px, pixy = (float)py;
  const float fx0 = (float)bx0, fy0 = (float)by0;
  const size_t pix = (size_t)py * a.W + px;
  const size_t plane = (size_t)a.W * a.H;

  unsigned long long t_start = 0;
  if (COUNT) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_start));
  const uint2 range = a.ranges[tile];
  const int n = (int)(range.y - range.x);
  const uint32_t* __restrict__ list = a.point_list + range.x;
  constexpr unsigned REC_BYTES = INTERP ? 64u : 48u;

  // per-pixel constants
  int last = 0;
  float g0 = 0.f, g1 = 0.f, g2 = 0.f, gD = 0.f, T_final = 0.f, bgdot = 0.f;
  if (inside) {
    last = (int)a.n_contrib[pix];
    g0 = a.dL_dcolor[pix];
    g1 = a.dL_dcolor[plane + pix];
    g2 = a.dL_dcolor[2 * plane + pix];
    if (DEPTH) gD = a.dL_dinvdepth[pix];
    T_final = a.final_T[pix];
    bgdot = a.bg[0] * g0 + a.bg[1] * g1 + a.bg[2] * g2;
  }
  // how far does this warp / this CTA have to walk?
  int wlast = last;
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) wlast = max(wlast, __shfl_xor_sync(0xffffffffu, wlast, d));
  if (lane == 0) sm.red[warp] = wlast;
  for (int i = tid; i < BLEND_BATCH * ACC_W; i += BLEND_THREADS) (&sm.acc[0][0])[i] = 0.f;
  if (TMA && tid == 0) {
    mbar_init(&sm.bar[0], 1);
    mbar_init(&sm.bar[1], 1);
    mbar_fence_init();
  }
  __syncthreads();
  int blast = 0;
#pragma unroll
  for (int w = 0; w < BLEND_THREADS / 32; ++w) blast = max(blast, sm.red[w]);
  const int seg_lo = k * BLEND_SEG;
  if (seg_lo >= min(blast, n)) return;   // nobody in this tile got that deep
  const int nn = min(seg_lo + BLEND_SEG, min(blast, n));  // one past the deepest list position of this item
  const int len = nn - seg_lo;
  const int rounds = (len + BLEND_BATCH - 1) / BLEND_BATCH;

  auto issue = [&](int b, uint32_t id) {
    const int j = b * BLEND_BATCH + tid;
    const int s = b & 1;
    if (TMA) {
      if (tid == 0) {
        const int cnt = min(BLEND_BATCH, len - b * BLEND_BATCH);
        mbar_expect_tx(&sm.bar[s], (unsigned)cnt * REC_BYTES);
      }
      if (j < len) bulk_g2s(&sm.rec[s][tid], a.rec + id, REC_BYTES, &sm.bar[s]);
    } else {
      if (j < len) {
        const char* src = reinterpret_cast<const char*>(a.rec + id);
        char* dst = reinterpret_cast<char*>(&sm.rec[s][tid]);
        cp_async16(dst, src);
        cp_async16(dst + 16, src + 16);
        cp_async16(dst + 32, src + 32);
        if (INTERP) cp_async16(dst + 48, src + 48);
      }
      cp_async_commit();
    }
  };

  // which of the ten sums ends up complete in this lane after the transposed butterfly:
  // group (lane bit 4) * 5 + (bit1 ? 2 : bit2 ? (bit3 ? 4 : 1) : (bit3 ? 3 : 0)); one depositor per sum
  int red_dst = -1;
  if ((lane & 1) == 0) {
    const int b4 = (lane >> 4) & 1, b3 = (lane >> 3) & 1, b2 = (lane >> 2) & 1, b1 = (lane >> 1) & 1;
    if (b1) {
      if (!b2 && !b3) red_dst = b4 * 5 + 2;
    } else {
      red_dst = b4 * 5 + (b2 ? (b3 ? 4 : 1) : (b3 ? 3 : 0));
    }
    if (!DEPTH && red_dst == 9) red_dst = -1;
  }
  // Walking state of a pixel: T in front of the entry at hand, and B = "what lies behind it, contracted with the
  // pixel gradient" (the reference's accum_rec . dL_dpixel, backward.cu:657-691), updated as B += alpha (u - B).
  // An entry that does not contribute enters with alpha = 0 and leaves both unchanged: no selects.
  float T = T_final, B = 0.f;
  const float Tfb = -T_final * bgdot;
  // shared-memory address of this lane's accumulator column (see the butterfly below)
  const uint32_t acc_addr = smem_u32(&sm.acc[0][red_dst < 0 ? 0 : red_dst]);
  if (inside && last > nn) {
    // this pixel's walk started in a deeper item: pick it up at the cut from the forward's snapshot.  What lies
    // behind the cut, contracted with the pixel gradient, is (total - in front of the cut) / T at the cut.
    const float* __restrict__ sp =
        a.snap + ((size_t)(range.x / BLEND_SEG) + (size_t)(k + 1)) * SNAP_FLOATS + tid;
    const float4 tot = a.accum[pix];
    T = sp[0];
    float behind = (tot.x - sp[TILE_PIX]) * g0 + (tot.y - sp[2 * TILE_PIX]) * g1 + (tot.z - sp[3 * TILE_PIX]) * g2;
    if (DEPTH) behind += (tot.w - sp[4 * TILE_PIX]) * gD;
    B = behind / T;
  }
  unsigned cntA = 0, cntB = 0, cntD = 0, cntS = 0;
  const float half_w = 0.5f * (float)a.W, half_h = 0.5f * (float)a.H;

  // staging slot k of round b holds list position nn - 1 - (b * BATCH + k): back to front
  uint32_t id_cur = (tid < len) ? __ldg(list + (nn - 1 - tid)) : 0u;
  issue(0, id_cur);
  uint32_t id_next = (BLEND_BATCH + tid < len) ? __ldg(list + (nn - 1 - BLEND_BATCH - tid)) : 0u;

  for (int b = 0; b < rounds; ++b) {
    const int s = b & 1;
    uint32_t id_after = 0;
    if (b + 1 < rounds) {
      issue(b + 1, id_next);
      const int j2 = (b + 2) * BLEND_BATCH + tid;
      id_after = (j2 < len) ? __ldg(list + (nn - 1 - j2)) : 0u;
      if (!TMA) cp_async_wait<1>();
    } else {
      if (!TMA) cp_async_wait<0>();
    }
    if (TMA) mbar_wait(&sm.bar[s], (unsigned)((b >> 1) & 1));
    __syncthreads();  // records of round b visible; accumulator zeroed by the previous flush

    const int base = b * BLEND_BATCH;
    const int cnt = min(BLEND_BATCH, len - base);
    // slots whose list position is >= this warp's last contributor come first; skip them
    const int first = max(0, nn - min(wlast, nn) - base);
    const Rec* __restrict__ recs = sm.rec[s];
    for (int c = first & ~31; c < cnt; c += 32) {
      const int j = c + lane;
      bool hit = false;
      if (j >= first && j < cnt) hit = block_hit(recs[j].q0, recs[j].q1, fx0, fy0);
      unsigned m = __ballot_sync(0xffffffffu, hit);
      while (m) {
        const int jj = c + __ffs(m) - 1;
        m &= m - 1;
        // Straight-line evaluation with the decisions folded into one predicate: lanes that do not contribute
        // (behind their last contributor, power > 0, alpha < 1/255) run the same arithmetic on values nobody
        // reads and deposit exact zeros (G, w and dL_dalpha are forced to 0 for them).
        const float2 xy = *reinterpret_cast<const float2*>(&recs[jj].q0);
        const float4 con_o = recs[jj].q1;
        const float dx = xy.x - pixx, dy = xy.y - pixy;
        const float power = blend_power(con_o, dx, dy);
        // G, alpha and T are rebuilt with the MUFU approximations (ex2 / lg2 / rcp, relative error ~1e-7 -- the
        // level at which the reference's own forward (__powf) and backward (pow) already disagree); the keep / skip
        // decision at 1/255 below is the one place where a last-bit difference would change the result visibly,
        // and is redone exactly there.
        const float G = ex2_approx(power * 1.4426950408889634f);
        const float test_alpha = con_o.w * G;
        const float my_alpha = min(0.99f, test_alpha);
        float alpha = my_alpha;
        float opac_mult = 1.0f;
        if (INTERP) {
          const float2 tk = *reinterpret_cast<const float2*>(&recs[jj].q3);
          if (tk.x != 1.0f) {
            // (1-a)^(1/k) (backward.cu:631) and (1-a)^(1/k - 1) = that / (1-a) (backward.cu:714); 1-a is in [0.01, 1]
            const float om = 1.0f - my_alpha;
            float pw = ex2_approx(tk.y * lg2_approx(om));
            alpha = tk.x * my_alpha + (1.0f - tk.x) * (1.0f - pw);
            // the keep / skip decision below must be the reference's: within reach of the 1/255 threshold (the
            // approximation is off by < 2e-7 in alpha) redo it with the reference's expf and powf themselves
            if (fabsf(alpha - 1.0f / 255.0f) < 1e-6f) {
              pw = powf(1.0f - min(0.99f, con_o.w * expf(power)), tk.y);
              alpha = tk.x * min(0.99f, con_o.w * expf(power)) + (1.0f - tk.x) * (1.0f - pw);
            }
            // d alpha / d a = t - (1-a)^(1/k - 1) (t-1) / k
            opac_mult = tk.x - pw * __frcp_approx(om) * (tk.x - 1.0f) * tk.y;
          } else if (fabsf(alpha - 1.0f / 255.0f) < 2e-8f) {   // ex2.approx vs expf: a few ulp of alpha
            alpha = min(0.99f, con_o.w * expf(power));
          }
        } else if (fabsf(alpha - 1.0f / 255.0f) < 2e-8f) {
          alpha = min(0.99f, con_o.w * expf(power));
        }
        const bool contrib = (nn - 1 - (base + jj) < last) & !(power > 0.0f) & !(alpha < 1.0f / 255.0f);
        if (COUNT) {
          if (nn - 1 - (base + jj) < last) {
            if (power > 0.0f) ++cntA; else if (!contrib) ++cntB; else ++cntD;
          }
          ++cntS;
        }
        if (!__any_sync(0xffffffffu, contrib)) continue;
        const float4 q2 = recs[jj].q2;
        // 1 - alpha is in [0.01, 1]: the approximate reciprocal plus one Newton step is exact to an ulp there
        // (and exactly 1 for alpha = 0) and needs none of the range checks of a full-range division
        const float a_eff = contrib ? alpha : 0.f;
        const float om_a = 1.f - a_eff;
        float rcp_om = __frcp_approx(om_a);
        rcp_om = fmaf(fmaf(-om_a, rcp_om, 1.0f), rcp_om, rcp_om);
        T *= rcp_om;
        float u = q2.x * g0 + q2.y * g1 + q2.z * g2;
        if (DEPTH) u += q2.w * gD;
        const float d = u - B;
        float dL_dalpha = fmaf(d, T, Tfb * rcp_om);
        B = fmaf(a_eff, d, B);
        // clamped alpha passes no gradient (backward.cu:619,693); entries that do not contribute carry Gc = w = 0
        if (test_alpha > 0.99f) dL_dalpha = 0.f;
        const float Gc = contrib ? G : 0.f;
        const float w = a_eff * T;
        const float GdA = Gc * dL_dalpha;
        const float q = con_o.w * GdA;
        const float v0 = q * dx;
        const float v1 = q * dy;
        const float v2 = v0 * dx;
        const float v3 = v0 * dy;
        const float v4 = v1 * dy;
        const float v5 = opac_mult * GdA;
        const float v6 = w * g0;
        const float v7 = w * g1;
        const float v8 = w * g2;
        const float v9 = DEPTH ? w * gD : 0.f;
        {

          // Transposed butterfly: at every level a lane keeps half of its values and hands the other half
          // to its partner, so the ten sums cost 5+3+2+1+1 = 12 shuffles instead of 10 x 5.
          const bool h16 = lane & 16, h8 = lane & 8, h4 = lane & 4, h2 = lane & 2;
          float r0 = (h16 ? v5 : v0) + __shfl_xor_sync(0xffffffffu, h16 ? v0 : v5, 16);
          float r1 = (h16 ? v6 : v1) + __shfl_xor_sync(0xffffffffu, h16 ? v1 : v6, 16);
          float r2 = (h16 ? v7 : v2) + __shfl_xor_sync(0xffffffffu, h16 ? v2 : v7, 16);
          float r3 = (h16 ? v8 : v3) + __shfl_xor_sync(0xffffffffu, h16 ? v3 : v8, 16);
          float r4 = (h16 ? v9 : v4) + __shfl_xor_sync(0xffffffffu, h16 ? v4 : v9, 16);
          float s0 = (h8 ? r3 : r0) + __shfl_xor_sync(0xffffffffu, h8 ? r0 : r3, 8);
          float s1 = (h8 ? r4 : r1) + __shfl_xor_sync(0xffffffffu, h8 ? r1 : r4, 8);
          float s2 = r2 + __shfl_xor_sync(0xffffffffu, r2, 8);
          float t0 = (h4 ? s1 : s0) + __shfl_xor_sync(0xffffffffu, h4 ? s0 : s1, 4);
          float t1 = s2 + __shfl_xor_sync(0xffffffffu, s2, 4);
          float u = (h2 ? t1 : t0) + __shfl_xor_sync(0xffffffffu, h2 ? t0 : t1, 2);
          u += __shfl_xor_sync(0xffffffffu, u, 1);
          // lane (b4 b3 b2 b1 0) now holds the complete sum number red_dst (see above); ten lanes deposit
          if (red_dst >= 0)
            asm volatile("red.shared.add.f32 [%0], %1;" ::"r"(acc_addr + (uint32_t)jj * (ACC_W * 4u)), "f"(u) : "memory");
        }
      }
    }
    __syncthreads();  // all partial sums of this round are in the accumulator

    // flush: thread t owns list entry base + t
    if (tid < cnt) {
      float4 s0 = *reinterpret_cast<float4*>(&sm.acc[tid][0]);
      float4 s1 = *reinterpret_cast<float4*>(&sm.acc[tid][4]);
      float2 s2 = *reinterpret_cast<float2*>(&sm.acc[tid][8]);
      const bool any = (s0.x != 0.f) | (s0.y != 0.f) | (s0.z != 0.f) | (s0.w != 0.f) | (s1.x != 0.f) |
                       (s1.y != 0.f) | (s1.z != 0.f) | (s1.w != 0.f) | (s2.x != 0.f) | (s2.y != 0.f);
      if (any) {
        const float4 con_o = recs[tid].q1;
        // dL/dmean2D from dG/ddelta (backward.cu:699-704), dL/dconic (backward.cu:707-709)
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
    if (tid == 0 && k == 0) {   // per-tile times cover the tile's first work item only
      unsigned long long t_end;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_end));
      a.stats[16 + 4 * tile + 2] = t_start;
      a.stats[16 + 4 * tile + 3] = t_end;
    }
    const unsigned v[4] = {cntA, cntB, cntD, cntS};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const unsigned t = __reduce_add_sync(0xffffffffu, v[k]);
      if (lane == 0 && t) atomicAdd(a.stats + 5 + k, (unsigned long long)t);
    }
  }
}

template <bool INTERP, bool DEPTH>
cudaError_t launch2(const BlendArgs& a, bool tma, cudaStream_t s) {
  const int tiles = a.grid_x * a.grid_y + a.nseg;
  const size_t smem = sizeof(BwdSmem);
  cudaError_t e;
  ProfScope ps(K_BLEND_BWD, s);
  if (a.stats) {
    auto k = blend_bwd_kernel<INTERP, DEPTH, false, true>;
    e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    k<<<tiles, BLEND_THREADS, smem, s>>>(a);
  } else if (tma) {
    auto k = blend_bwd_kernel<INTERP, DEPTH, true>;
    e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    k<<<tiles, BLEND_THREADS, smem, s>>>(a);
  } else {
    auto k = blend_bwd_kernel<INTERP, DEPTH, false>;
    e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    k<<<tiles, BLEND_THREADS, smem, s>>>(a);
  }
  return cudaGetLastError();
}

}  // namespace

int blend_stage_mode();

// H3DGS_BWD=1 selects this file's one-pixel-per-lane kernel; the default is blend_bwd2.cu (two pixels per lane,
// one cross-lane reduction per 64 pixels).  The bulk-copy staging mode exists in this kernel only.
static int g_bwd_variant = -1;
static int bwd_variant() {
  if (g_bwd_variant < 0) {
    const char* e = std::getenv("H3DGS_BWD");
    g_bwd_variant = (e && e[0] == '1') ? 1 : 2;
  }
  return g_bwd_variant;
}
int set_bwd_variant(int v) {   // diagnostics (h3dgs_debug_set): returns the previous variant
  const int prev = bwd_variant();
  if (v == 1 || v == 2) g_bwd_variant = v;
  return prev;
}

cudaError_t launch_blend_bwd(const BlendArgs& a, cudaStream_t s) {
  const bool tma = blend_stage_mode() == 1;
  if (!tma && bwd_variant() == 2) return launch_blend_bwd2(a, s);
  const bool depth = a.dL_dinvdepth != nullptr;
  if (a.interp) return depth ? launch2<true, true>(a, tma, s) : launch2<true, false>(a, tma, s);
  return depth ? launch2<false, true>(a, tma, s) : launch2<false, false>(a, tma, s);
}

}  // namespace h3dgs
