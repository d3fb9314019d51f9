// Single-pass inclusive prefix sum (decoupled look-back) of the per-slot tile counts.
// Replaces cub::DeviceScan::InclusiveSum at HR/cuda_rasterizer/rasterizer_impl.cu:316 and
// GH/runtime_switching.cu:531-533.  8 bytes of HBM traffic per element, one launch.
#include "common.cuh"

namespace h3dgs {

namespace {

constexpr unsigned long long FLAG_AGG = 1ull << 62;
constexpr unsigned long long FLAG_INC = 2ull << 62;
constexpr unsigned long long VAL_MASK = (1ull << 62) - 1;

// element `index` owns outputs [prev, cur): it is the first one reaching into every chunk that starts there
__device__ __forceinline__ void mark_chunks(uint32_t* __restrict__ chunk_first, unsigned long long prev,
                                            unsigned long long cur, uint32_t index) {
  for (unsigned long long c = (prev + EMIT_CHUNK - 1) / EMIT_CHUNK; c * EMIT_CHUNK < cur; ++c)
    chunk_first[c] = index;
}

// MODE 0: scan in[i].   MODE 1: scan (in[i] != 0) -- the rank of every visible slot -- and also add up in[i]
// itself into counters[4..5] (the instance total).   MODE 2: scan in[idx[i]] (tile counts in depth order).
// MODE 3: `in` holds the slots' tile rectangles (ushort4); scan the number of strip records of slot idx[i]: one per
// tile row and STRIP_TILES-wide strip its rectangle reaches into (binning.cu).
template <int MODE>
__global__ void __launch_bounds__(SCAN_THREADS) scan_lookback_kernel(
    const uint32_t* __restrict__ in, const uint32_t* __restrict__ idx, uint32_t* __restrict__ out, int n_host,
    const unsigned long long* __restrict__ n_dev, uint32_t* __restrict__ chunk_first,
    volatile unsigned long long* status, uint32_t* counters) {
  __shared__ uint32_t s_tile;
  // element count from the host, or left on the device by an earlier kernel (then the grid covers n_host and the
  // surplus CTAs leave: look-back only ever reads the tiles in front of a tile)
  const int n = n_dev ? (int)*n_dev : n_host;
  __shared__ unsigned long long s_warp[SCAN_THREADS / 32];
  __shared__ unsigned long long s_prefix;

  if (threadIdx.x == 0) s_tile = atomicAdd(&counters[0], 1u);
  __syncthreads();
  const uint32_t tile = s_tile;
  if ((size_t)tile * SCAN_TILE >= (size_t)n) return;
  const size_t base = (size_t)tile * SCAN_TILE + (size_t)threadIdx.x * SCAN_ITEMS;

  uint32_t v[SCAN_ITEMS];
  const uint32_t* __restrict__ src = (MODE >= 2) ? idx : in;
  if (base + SCAN_ITEMS <= (size_t)n) {
    const uint4* p = reinterpret_cast<const uint4*>(src + base);
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS / 4; ++i) {
      uint4 x = __ldg(p + i);
      v[4 * i] = x.x; v[4 * i + 1] = x.y; v[4 * i + 2] = x.z; v[4 * i + 3] = x.w;
    }
  } else {
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; ++i) v[i] = (base + i < (size_t)n) ? src[base + i] : 0u;
  }
  if (MODE == 2) {
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; ++i) v[i] = (base + i < (size_t)n) ? __ldg(in + v[i]) : 0u;
  }
  if (MODE == 3) {
    const uint2* __restrict__ rects = reinterpret_cast<const uint2*>(in);
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; ++i) {
      uint32_t c = 0;
      if (base + i < (size_t)n) {
        const uint2 r = __ldg(rects + v[i]);   // (x0 | y0 << 16, x1 | y1 << 16), x1 > x0 and y1 > y0 for visible slots
        const uint32_t x0 = r.x & 0xffffu, y0 = r.x >> 16, x1 = r.y & 0xffffu, y1 = r.y >> 16;
        c = (y1 - y0) * (((x1 - 1u) / STRIP_TILES) - (x0 / STRIP_TILES) + 1u);
      }
      v[i] = c;
    }
  }
  if (MODE == 1) {
    unsigned long long raw = 0;
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; ++i) { raw += v[i]; v[i] = v[i] != 0u ? 1u : 0u; }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) raw += __shfl_xor_sync(0xffffffffu, raw, d);
    if ((threadIdx.x & 31) == 0 && raw) atomicAdd(reinterpret_cast<unsigned long long*>(counters + 4), raw);
  }
  unsigned long long tsum = 0;
#pragma unroll
  for (int i = 0; i < SCAN_ITEMS; ++i) tsum += v[i];

  // block exclusive scan of per-thread sums
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  unsigned long long incl = tsum;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    unsigned long long o = __shfl_up_sync(0xffffffffu, incl, d);
    if (lane >= d) incl += o;
  }
  if (lane == 31) s_warp[warp] = incl;
  __syncthreads();
  unsigned long long warp_off = 0, block_sum = 0;
#pragma unroll
  for (int w = 0; w < SCAN_THREADS / 32; ++w) {
    unsigned long long x = s_warp[w];
    if (w < warp) warp_off += x;
    block_sum += x;
  }
  const unsigned long long thread_excl = warp_off + incl - tsum;

  // publish aggregate, then look back for the exclusive prefix of this tile
  if (warp == 0) {
    if (lane == 0) status[tile] = (tile == 0 ? FLAG_INC : FLAG_AGG) | (block_sum & VAL_MASK);
    unsigned long long prefix = 0;
    if (tile > 0) {
      long long look = (long long)tile - 1;
      while (true) {
        long long idx = look - lane;
        unsigned long long w = FLAG_INC;  // out-of-range predecessors act as a zero inclusive prefix
        if (idx >= 0) {
          do { w = status[idx]; } while ((w >> 62) == 0);
        }
        const unsigned inc_mask = __ballot_sync(0xffffffffu, (w >> 62) == 2);
        unsigned long long val = w & VAL_MASK;
        if (inc_mask) {
          const int first = __ffs(inc_mask) - 1;  // nearest predecessor holding an inclusive prefix
          if (lane > first) val = 0;
        }
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) val += __shfl_xor_sync(0xffffffffu, val, d);
        prefix += val;
        if (inc_mask) break;
        look -= 32;
      }
      if (lane == 0) status[tile] = FLAG_INC | ((prefix + block_sum) & VAL_MASK);
    }
    if (lane == 0) s_prefix = prefix;
  }
  __syncthreads();
  unsigned long long run = s_prefix + thread_excl;

  if (base + SCAN_ITEMS <= (size_t)n) {
    uint4* q = reinterpret_cast<uint4*>(out + base);
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS / 4; ++i) {
      uint4 x;
      if (MODE >= 2) {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          mark_chunks(chunk_first, run, run + v[4 * i + k], (uint32_t)(base + 4 * i + k));
          run += v[4 * i + k];
          (&x.x)[k] = (uint32_t)run;
        }
      } else {
        run += v[4 * i]; x.x = (uint32_t)run;
        run += v[4 * i + 1]; x.y = (uint32_t)run;
        run += v[4 * i + 2]; x.z = (uint32_t)run;
        run += v[4 * i + 3]; x.w = (uint32_t)run;
      }
      q[i] = x;
    }
  } else {
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; ++i) {
      if (MODE >= 2 && base + i < (size_t)n) mark_chunks(chunk_first, run, run + v[i], (uint32_t)(base + i));
      run += v[i];
      if (base + i < (size_t)n) out[base + i] = (uint32_t)run;
    }
  }
  // the tile holding element n-1 reports the grand total (as u64: the u32 outputs may have wrapped)
  if ((size_t)tile == ((size_t)n - 1) / SCAN_TILE && threadIdx.x == SCAN_THREADS - 1) {
    *reinterpret_cast<unsigned long long*>(counters + 2) = run;
  }
}

}  // namespace

template <int MODE>
static cudaError_t launch_scan(const uint32_t* in, const uint32_t* idx, uint32_t* out, int n,
                               const unsigned long long* n_dev, uint32_t* chunk_first, unsigned long long* status,
                               uint32_t* counters, cudaStream_t s) {
  if (n <= 0) return cudaSuccess;
  const int tiles = (n + SCAN_TILE - 1) / SCAN_TILE;
  cudaError_t e = cudaMemsetAsync(status, 0, sizeof(unsigned long long) * (size_t)(tiles + 1), s);
  if (e != cudaSuccess) return e;
  e = cudaMemsetAsync(counters, 0, 8 * sizeof(uint32_t), s);
  if (e != cudaSuccess) return e;
  ProfScope ps(K_SCAN, s);
  scan_lookback_kernel<MODE><<<tiles, SCAN_THREADS, 0, s>>>(in, idx, out, n, n_dev, chunk_first, status, counters);
  return cudaGetLastError();
}

cudaError_t launch_scan_inclusive(const uint32_t* in, uint32_t* out, int n, unsigned long long* status,
                                  uint32_t* counters, cudaStream_t s) {
  return launch_scan<0>(in, nullptr, out, n, nullptr, nullptr, status, counters, s);
}
cudaError_t launch_scan_visible(const uint32_t* tiles_touched, uint32_t* rank, int n,
                                unsigned long long* status, uint32_t* counters, cudaStream_t s) {
  return launch_scan<1>(tiles_touched, nullptr, rank, n, nullptr, nullptr, status, counters, s);
}
cudaError_t launch_scan_strips(const ushort4* tile_rect, const uint32_t* order, uint32_t* out, int n_cap,
                               const unsigned long long* n_dev, uint32_t* chunk_first, unsigned long long* status,
                               uint32_t* counters, cudaStream_t s) {
  return launch_scan<3>(reinterpret_cast<const uint32_t*>(tile_rect), order, out, n_cap, n_dev, chunk_first, status,
                        counters, s);
}
cudaError_t launch_scan_counts(const uint32_t* in, uint32_t* out, long long n_cap, const unsigned long long* n_dev,
                               unsigned long long* status, uint32_t* counters, cudaStream_t s) {
  if (n_cap > 0x7fffffffll) return cudaErrorInvalidValue;
  return launch_scan<0>(in, nullptr, out, (int)n_cap, n_dev, nullptr, status, counters, s);
}

}  // namespace h3dgs
