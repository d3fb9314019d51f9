// Forward tile alpha blend.  Replaces renderCUDA of HR/cuda_rasterizer/forward.cu:439-584.
//
// One CTA per 16x16 tile, eight warps each owning an 8x4 pixel block.  Per round, 256 list entries are
// staged into shared memory as 64-byte records (double buffered; by the bulk-copy engine or LDGSTS).
// Each warp first tests 32 staged entries at a time against its pixel block (one entry per lane, a
// ballot gives the survivors) and only evaluates survivors per pixel, in list order, so every pixel
// sees exactly the sequence of blends the reference performs: alpha, the 1/255 and T<1e-4 decisions,
// n_contrib and final_T are bit-identical.  Warps retire as soon as their 32 pixels are saturated.
// Every BLEND_SEG entries the per-pixel state is also written out (a few dozen long horizon tiles only): the
// backward pass cuts long lists there into independent work items.
#include "blend_common.cuh"

namespace h3dgs {

namespace {

struct FwdSmem {
  Rec rec[2][BLEND_BATCH];
  unsigned long long bar[2];
};

// COUNT: the untimed instrumented pass of bench.py -- per (live pixel, evaluated entry) pair, how far it got:
// stats[0] rejected at power > 0, [1] at alpha < 1/255, [2] at the T test, [3] blended, [4] lane slots issued.
template <bool INTERP, bool DEPTH, bool TMA, bool COUNT = false>
__global__ void __launch_bounds__(BLEND_THREADS) blend_fwd_kernel(const BlendArgs a) {
  __shared__ FwdSmem sm;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int tile = a.order ? (int)a.order[blockIdx.x] : (int)blockIdx.x;
  const int tx = tile % a.grid_x, ty = tile / a.grid_x;
  const int bx0 = tx * TILE_X + (warp & 1) * SUB_W;
  const int by0 = ty * TILE_Y + (warp >> 1) * SUB_H;
  const int px = bx0 + (lane & 7), py = by0 + (lane >> 3);
  const bool inside = px < a.W && py < a.H;
  const float pixx = (float)px, pixy = (float)py;
  const float fx0 = (float)bx0, fy0 = (float)by0;

  unsigned long long t_start = 0;
  if (COUNT) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_start));
  const uint2 range = a.ranges[tile];
  const int n = (int)(range.y - range.x);
  const int rounds = (n + BLEND_BATCH - 1) / BLEND_BATCH;
  const uint32_t* __restrict__ list = a.point_list + range.x;
  constexpr unsigned REC_BYTES = INTERP ? 64u : 48u;

  if (TMA) {
    if (tid == 0) {
      mbar_init(&sm.bar[0], 1);
      mbar_init(&sm.bar[1], 1);
      mbar_fence_init();
    }
    __syncthreads();
  }

  auto issue = [&](int b, uint32_t id) {
    const int j = b * BLEND_BATCH + tid;
    const int s = b & 1;
    if (TMA) {
      if (tid == 0) {
        const int cnt = min(BLEND_BATCH, n - b * BLEND_BATCH);
        mbar_expect_tx(&sm.bar[s], (unsigned)cnt * REC_BYTES);
      }
      if (j < n) bulk_g2s(&sm.rec[s][tid], a.rec + id, REC_BYTES, &sm.bar[s]);
    } else {
      if (j < n) {
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

  float T = inside ? 1.0f : -1.0f;   // sign bit = "done" (saturated, or outside the image)
  float C0 = 0.f, C1 = 0.f, C2 = 0.f, D = 0.f;
  uint32_t last_contributor = 0;
  unsigned cntA = 0, cntB = 0, cntC = 0, cntD = 0, cntS = 0;

  // software pipeline: ids two rounds ahead, records one round ahead
  uint32_t id_next = 0;
  if (rounds > 0) {
    uint32_t id0 = (tid < n) ? __ldg(list + tid) : 0u;
    issue(0, id0);
    if (rounds > 1) id_next = (BLEND_BATCH + tid < n) ? __ldg(list + BLEND_BATCH + tid) : 0u;
  }

  for (int b = 0; b < rounds; ++b) {
    const int s = b & 1;
    // list cut: leave what every pixel has accumulated in front of it for the backward work item that starts
    // (in walking order: ends) here -- see blend_bwd.cu
    if (a.snap && b > 0 && (b & (BLEND_SEG / BLEND_BATCH - 1)) == 0) {
      const size_t slot = (size_t)(range.x / BLEND_SEG) + (size_t)(b / (BLEND_SEG / BLEND_BATCH));
      float* __restrict__ sp = a.snap + slot * SNAP_FLOATS + tid;
      sp[0] = T;
      sp[TILE_PIX] = C0;
      sp[2 * TILE_PIX] = C1;
      sp[3 * TILE_PIX] = C2;
      if (DEPTH) sp[4 * TILE_PIX] = D;
    }
    if (b + 1 < rounds) {
      issue(b + 1, id_next);
      const int j2 = (b + 2) * BLEND_BATCH + tid;
      id_next = (j2 < n) ? __ldg(list + j2) : 0u;
      if (!TMA) cp_async_wait<1>();
    } else {
      if (!TMA) cp_async_wait<0>();
    }
    if (TMA) mbar_wait(&sm.bar[s], (unsigned)((b >> 1) & 1));
    else __syncthreads();

    const bool warp_live = __any_sync(0xffffffffu, T > 0.0f);
    if (warp_live) {
      const int cnt = min(BLEND_BATCH, n - b * BLEND_BATCH);
      const Rec* __restrict__ recs = sm.rec[s];
      for (int c = 0; c < cnt; c += 32) {
        const int j = c + lane;
        bool hit = false;
        if (j < cnt) hit = block_hit(recs[j].q0, recs[j].q1, fx0, fy0);
        unsigned m = __ballot_sync(0xffffffffu, hit);
        // Straight-line evaluation, decisions folded into predicates: the arithmetic of every lane that
        // blends is the reference's, lanes that do not blend compute values nobody reads.  `eval` is the part
        // that does not depend on the running transmittance; survivors are taken two at a time so that two
        // such chains (LDS -> power -> exp -> alpha) are in flight per warp -- what bounds the long horizon
        // tiles that end up alone on their SM.  `apply` then blends in list order.
        auto eval = [&](int jj, float& alpha, bool& ok, bool& pos) {
          const float2 xy = *reinterpret_cast<const float2*>(&recs[jj].q0);
          const float4 con_o = recs[jj].q1;
          const float dx = xy.x - pixx, dy = xy.y - pixy;
          const float power = blend_power(con_o, dx, dy);
          const float my_alpha = min(0.99f, con_o.w * exp(power));
          alpha = my_alpha;
          if (INTERP) {
            const float2 tk = *reinterpret_cast<const float2*>(&recs[jj].q3);
            if (tk.x != 1.0f) {  // t == 1 gives alpha == my_alpha exactly
              // slot gate (forward.cu:537): skybox slots carry t = 1 in their record already
              const float kidsqrt_alpha = 1.0f - __powf(1.0f - my_alpha, tk.y);
              alpha = tk.x * my_alpha + (1.0f - tk.x) * kidsqrt_alpha;
            }
          }
          ok = !(power > 0.0f) & !(alpha < 1.0f / 255.0f);
          pos = power > 0.0f;
        };
        auto apply = [&](int jj, float alpha, bool ok, bool pos) {
          // a saturated pixel carries -T: its test_T is negative, so it can never take another entry
          const float test_T = T * (1 - alpha);
          const bool take = ok & !(test_T < 0.0001f);
          if (COUNT) {
            if (T > 0.0f) {
              if (pos) ++cntA; else if (!ok) ++cntB; else if (!take) ++cntC; else ++cntD;
            }
            ++cntS;
          }
          if (ok & !take) T = -fabsf(T);
          const float4 q2 = recs[jj].q2;
          if (take) {
            C0 += q2.x * alpha * T;
            C1 += q2.y * alpha * T;
            C2 += q2.z * alpha * T;
            if (DEPTH) D += q2.w * alpha * T;
            T = test_T;
            last_contributor = (uint32_t)(b * BLEND_BATCH + jj + 1);
          }
        };
        while (m) {
          const int j0 = c + __ffs(m) - 1;
          m &= m - 1;
          float a0, a1;
          bool k0, k1, p0, p1;
          if (m) {
            const int j1 = c + __ffs(m) - 1;
            m &= m - 1;
            eval(j0, a0, k0, p0);
            eval(j1, a1, k1, p1);
            apply(j0, a0, k0, p0);
            apply(j1, a1, k1, p1);
          } else {
            eval(j0, a0, k0, p0);
            apply(j0, a0, k0, p0);
          }
        }
        if (!__any_sync(0xffffffffu, T > 0.0f)) break;
      }
    }
    // all warps saturated -> stop; the barrier also frees stage s for the copy issued next round
    if (__syncthreads_and(T < 0.0f ? 1 : 0)) {
      // the bulk copies of round b + 1 are already in flight: a CTA must not exit (and hand its shared memory
      // and mbarrier to the next CTA on this SM) before they have landed
      if (TMA && b + 1 < rounds) mbar_wait(&sm.bar[(b + 1) & 1], (unsigned)(((b + 1) >> 1) & 1));
      break;
    }
  }

  if (COUNT && a.stats) {
    if (tid == 0) {   // per-tile (start, end) in ns of the global timer: load-balance diagnostics
      unsigned long long t_end;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_end));
      a.stats[16 + 4 * tile] = t_start;
      a.stats[16 + 4 * tile + 1] = t_end;
    }
    const unsigned v[5] = {cntA, cntB, cntC, cntD, cntS};
#pragma unroll
    for (int k = 0; k < 5; ++k) {
      const unsigned t = __reduce_add_sync(0xffffffffu, v[k]);
      if (lane == 0 && t) atomicAdd(a.stats + k, (unsigned long long)t);
    }
  }
  if (a.tile_last) {
    const uint32_t wl = __reduce_max_sync(0xffffffffu, last_contributor);
    if (lane == 0 && wl) atomicMax(a.tile_last + tile, wl);
  }
  if (inside) {
    const size_t pix = (size_t)py * a.W + px;
    const size_t plane = (size_t)a.W * a.H;
    T = fabsf(T);
    a.final_T[pix] = T;
    a.n_contrib[pix] = last_contributor;
    a.accum[pix] = make_float4(C0, C1, C2, D);
    a.out_color[pix] = C0 + T * a.bg[0];
    a.out_color[plane + pix] = C1 + T * a.bg[1];
    a.out_color[2 * plane + pix] = C2 + T * a.bg[2];
    if (DEPTH) a.out_invdepth[pix] = D;
  }
}

template <bool INTERP, bool DEPTH>
cudaError_t launch2(const BlendArgs& a, bool tma, cudaStream_t s) {
  const int tiles = a.grid_x * a.grid_y;
  ProfScope ps(K_BLEND_FWD, s);
  if (a.stats)
    blend_fwd_kernel<INTERP, DEPTH, false, true><<<tiles, BLEND_THREADS, 0, s>>>(a);
  else if (tma)
    blend_fwd_kernel<INTERP, DEPTH, true><<<tiles, BLEND_THREADS, 0, s>>>(a);
  else
    blend_fwd_kernel<INTERP, DEPTH, false><<<tiles, BLEND_THREADS, 0, s>>>(a);
  return cudaGetLastError();
}

}  // namespace

int blend_stage_mode();  // api.cu: 0 = LDGSTS, 1 = bulk-copy engine

cudaError_t launch_blend_fwd(const BlendArgs& a, cudaStream_t s) {
  const bool tma = blend_stage_mode() == 1;
  const bool depth = a.out_invdepth != nullptr;
  if (a.interp) return depth ? launch2<true, true>(a, tma, s) : launch2<true, false>(a, tma, s);
  return depth ? launch2<false, true>(a, tma, s) : launch2<false, false>(a, tma, s);
}

}  // namespace h3dgs
