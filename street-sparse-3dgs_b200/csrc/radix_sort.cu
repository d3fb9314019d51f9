// Onesweep least-significant-digit radix sort passes used by the two-level instance sort.
//
// The reference sorts R (tile << 32 | depth) keys with cub::DeviceRadixSort::SortPairs
// (HR/cuda_rasterizer/rasterizer_impl.cu:352-357): 45 key bits at 1920x1080, i.e. 6 (CUB, 8-bit digits)
// sweeps over 12-byte pairs.  The same ordered list is produced here in levels (see binning.cu):
//   1. the V visible slots are sorted by depth (32-bit keys, 4 sweeps over V 8-byte pairs, V << R);
//   2. one record per (slot, tile row, 64-tile strip) is emitted in that order and stably partitioned by strip id:
//      ceil(strip bits / 8) sweeps (1 at 1080p) over ~R / 3 6-byte pairs;
//   3. the records of a strip are expanded into its tiles' lists by bit-matrix transposes (binning.cu).
// Equal keys keep their order in every level, exactly as the stable CUB sort leaves them.
//
// One kernel per digit: a CTA ranks a 6144-element tile in shared memory, resolves its global offsets with a
// decoupled look-back over the preceding tiles' per-digit counts and scatters; every element is read and
// written once per digit.
#include "common.cuh"

namespace h3dgs {

namespace {

constexpr int RADIX = SORT_RADIX;       // 256
constexpr int WARPS = SORT_THREADS / 32;
constexpr uint32_t ST_AGG = 1u << 30;
constexpr uint32_t ST_INC = 2u << 30;
constexpr uint32_t ST_VAL = (1u << 30) - 1;

// per-pass digit histograms of n keys: hist[p][d]
template <typename K>
__global__ void __launch_bounds__(512) digit_hist_kernel(const K* __restrict__ keys, long long n_host,
                                                         const unsigned long long* __restrict__ n_dev, int passes,
                                                         SortShifts sh, uint32_t* __restrict__ hist) {
  const long long n = n_dev ? (long long)*n_dev : n_host;
  __shared__ uint32_t h[SORT_MAX_PASSES * RADIX];
  for (int i = threadIdx.x; i < passes * RADIX; i += blockDim.x) h[i] = 0;
  __syncthreads();
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const uint32_t k = (uint32_t)keys[i];
#pragma unroll
    for (int p = 0; p < SORT_MAX_PASSES; ++p)
      if (p < passes) atomicAdd(&h[p * RADIX + ((k >> sh.shift[p]) & sh.mask[p])], 1u);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < passes * RADIX; i += blockDim.x) {
    const uint32_t c = h[i];
    if (c) atomicAdd(&hist[i], c);
  }
}

// exclusive scan of each pass' histogram, in place (one block of RADIX threads per pass)
__global__ void __launch_bounds__(RADIX) scan_hist_kernel(uint32_t* __restrict__ hist) {
  __shared__ uint32_t s_w[RADIX / 32];
  uint32_t* h = hist + (size_t)blockIdx.x * RADIX;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t c = h[threadIdx.x];
  uint32_t incl = c;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    uint32_t o = __shfl_up_sync(0xffffffffu, incl, d);
    if (lane >= d) incl += o;
  }
  if (lane == 31) s_w[warp] = incl;
  __syncthreads();
  uint32_t off = 0;
  for (int w = 0; w < warp; ++w) off += s_w[w];
  h[threadIdx.x] = off + incl - c;
}

template <typename K>
struct SweepSmem {
  uint32_t cnt[WARPS][RADIX];   // per-warp digit counters -> exclusive warp offsets
  uint32_t vals[SORT_TILE];
  K keys[SORT_TILE];
  uint32_t bin_start[RADIX];    // start of each digit inside the tile
  uint32_t gbase[RADIX];        // global position of tile-local slot p with digit d: gbase[d] + p
  uint32_t warp_sums[RADIX / 32];
  uint32_t tile;
};

// OUT 0: write (key, value).  OUT 1: write value only (last depth pass: only the order is needed).
template <typename K, int OUT>
__global__ void __launch_bounds__(SORT_THREADS, 3) onesweep_kernel(
    const K* __restrict__ kin, K* __restrict__ kout, const uint32_t* __restrict__ vin,
    uint32_t* __restrict__ vout, long long n_host, const unsigned long long* __restrict__ n_dev,
    int shift, uint32_t mask, const uint32_t* __restrict__ digit_base, volatile uint32_t* status,
    uint32_t* counter) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  SweepSmem<K>& sm = *reinterpret_cast<SweepSmem<K>*>(smem_raw);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

  // element count: known to the host (one CTA per tile) or left on the device by a scan (the V visible slots of the
  // depth passes, the strip records of the strip passes: a persistent grid whose CTAs keep taking tiles from the
  // dynamic counter until the elements are covered -- tiles are handed out in order, every CTA is resident, so
  // the look-back always waits on a tile that is being worked on)
  const long long n = n_dev ? (long long)*n_dev : n_host;
  for (;;) {
  if (tid == 0) sm.tile = atomicAdd(counter, 1u);
  for (int i = tid; i < WARPS * RADIX; i += SORT_THREADS) (&sm.cnt[0][0])[i] = 0;
  __syncthreads();
  const uint32_t tile = sm.tile;
  const long long tile_base = (long long)tile * SORT_TILE;
  if (tile_base >= n) return;
  const long long remaining = n - tile_base;
  const int tile_valid = (int)(remaining < SORT_TILE ? remaining : SORT_TILE);

  // warp-striped load: element order in memory == (warp, item, lane) order == ranking order
  K key[SORT_ITEMS];
  const int wbase = warp * (32 * SORT_ITEMS) + lane;
#pragma unroll
  for (int i = 0; i < SORT_ITEMS; ++i) {
    const int local = wbase + i * 32;
    key[i] = (local < tile_valid) ? kin[tile_base + local] : (K)~(K)0;
  }

  uint32_t rank[SORT_ITEMS];
  const uint32_t lt = (1u << lane) - 1u;
#pragma unroll
  for (int i = 0; i < SORT_ITEMS; ++i) {
    const uint32_t d = ((uint32_t)key[i] >> shift) & mask;
    const uint32_t peers = __match_any_sync(0xffffffffu, d);
    const int leader = __ffs(peers) - 1;
    uint32_t prior = 0;
    // one shared-memory atomic by the leader of each digit group (the counters of a warp are its own, so the
    // atomic is only the fused read-modify-write; program order keeps the ranking stable)
    if (lane == leader) prior = atomicAdd(&sm.cnt[warp][d], (uint32_t)__popc(peers));
    prior = __shfl_sync(0xffffffffu, prior, leader);
    rank[i] = prior + __popc(peers & lt);
  }
  __syncthreads();

  // one thread per digit: warp counts -> exclusive warp offsets and the tile total
  uint32_t total = 0, incl = 0;
  if (tid < RADIX) {
    uint32_t run = 0;
#pragma unroll
    for (int w = 0; w < WARPS; ++w) {
      const uint32_t c = sm.cnt[w][tid];
      sm.cnt[w][tid] = run;
      run += c;
    }
    total = run;
    status[(size_t)tile * RADIX + tid] = (tile == 0 ? ST_INC : ST_AGG) | total;
    incl = total;
#pragma unroll
    for (int k = 1; k < 32; k <<= 1) {
      uint32_t o = __shfl_up_sync(0xffffffffu, incl, k);
      if (lane >= k) incl += o;
    }
    if (lane == 31) sm.warp_sums[warp] = incl;
  }
  __syncthreads();
  if (tid < RADIX) {
    uint32_t off = 0;
    for (int w = 0; w < warp; ++w) off += sm.warp_sums[w];
    const uint32_t start = off + incl - total;
    sm.bin_start[tid] = start;
    // decoupled look-back over the preceding tiles
    uint32_t excl = 0;
    if (tile > 0) {
      // Eight predecessors are read at once (independent loads, one L2 round trip) and then consumed in order:
      // with a few hundred tiles resident at the same time the chain of publish -> observe hops, not the
      // bandwidth, is what a pass over a small array costs.
      constexpr int LB = 8;
      long long t = (long long)tile - 1;
      bool done = false;
      while (!done) {
        uint32_t w[LB];
#pragma unroll
        for (int j = 0; j < LB; ++j) {
          const long long tj = t - j;
          w[j] = (tj >= 0) ? status[(size_t)tj * RADIX + tid] : 0u;
        }
#pragma unroll
        for (int j = 0; j < LB; ++j) {
          if (done) break;
          const long long tj = t - j;
          uint32_t x = w[j];
          while ((x >> 30) == 0) x = status[(size_t)tj * RADIX + tid];   // not published yet: spin on this one
          excl += x & ST_VAL;
          if ((x >> 30) == 2 || tj == 0) done = true;
        }
        t -= LB;
      }
      status[(size_t)tile * RADIX + tid] = ST_INC | ((excl + total) & ST_VAL);
    }
    sm.gbase[tid] = digit_base[tid] + excl - start;
  }
  __syncthreads();

  // scatter into tile-local sorted order
#pragma unroll
  for (int i = 0; i < SORT_ITEMS; ++i) {
    const int local = wbase + i * 32;
    const uint32_t d = ((uint32_t)key[i] >> shift) & mask;
    const uint32_t pos = sm.bin_start[d] + sm.cnt[warp][d] + rank[i];
    sm.keys[pos] = key[i];
    sm.vals[pos] = (local < tile_valid) ? vin[tile_base + local] : 0u;
  }
  __syncthreads();

  // coalesced write-out; padding keys (all ones) sort to the very end of the tile
  for (int p = tid; p < tile_valid; p += SORT_THREADS) {
    const K k = sm.keys[p];
    const uint32_t d = ((uint32_t)k >> shift) & mask;
    const uint32_t g = sm.gbase[d] + (uint32_t)p;
    const uint32_t v = sm.vals[p];
    if (OUT == 0) kout[g] = k;
    vout[g] = v;
  }
  if (n_dev == nullptr) return;
  __syncthreads();   // shared memory is reused by the next tile
  }
}

template <typename K, int OUT>
cudaError_t launch_pass(const K* kin, K* kout, const uint32_t* vin, uint32_t* vout, long long n,
                        const unsigned long long* n_dev, int shift,
                        uint32_t mask, const uint32_t* digit_base, uint32_t* status, uint32_t* counter,
                        cudaStream_t s) {
  auto k = onesweep_kernel<K, OUT>;
  const size_t smem = sizeof(SweepSmem<K>);
  static bool attr_done = false;   // one flag per template instantiation
  if (!attr_done) {
    cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    attr_done = true;
  }
  long long tiles = (n + SORT_TILE - 1) / SORT_TILE;
  if (n_dev != nullptr) {   // persistent grid: 3 CTAs per SM are resident (launch bounds), never more CTAs than tiles
    int dev = 0, sms = 148;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (tiles > (long long)sms * 3) tiles = (long long)sms * 3;
  }
  ProfScope ps(K_SORT_PASS, s);
  k<<<(unsigned)tiles, SORT_THREADS, smem, s>>>(kin, kout, vin, vout, n, n_dev, shift, mask, digit_base, status,
                                                counter);
  return cudaGetLastError();
}

template <typename K>
cudaError_t launch_hist(const K* keys, long long n, const unsigned long long* n_dev, int passes,
                        const SortShifts& sh, uint32_t* hist, cudaStream_t s) {
  int blocks = (int)((n + 512 * 16 - 1) / (512 * 16));
  if (blocks > 148 * 4) blocks = 148 * 4;
  if (blocks < 1) blocks = 1;
  { ProfScope ps(K_SORT_HIST, s);
    digit_hist_kernel<K><<<blocks, 512, 0, s>>>(keys, n, n_dev, passes, sh, hist); }
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return e;
  { ProfScope ps(K_SORT_SCAN, s);
    scan_hist_kernel<<<passes, RADIX, 0, s>>>(hist); }
  return cudaGetLastError();
}

template <typename K>
cudaError_t strip_sort(BinWs& b, int64_t n_cap, const unsigned long long* n_dev, const SortPlan& plan,
                       cudaStream_t s) {
  K* keys[2] = {reinterpret_cast<K*>(b.skeys[0]), reinterpret_cast<K*>(b.skeys[1])};
  SortShifts sh;
  for (int p = 0; p < SORT_MAX_PASSES; ++p) {
    sh.shift[p] = p < plan.passes ? plan.shift[p] : 0;
    sh.mask[p] = p < plan.passes ? ((1u << plan.bits[p]) - 1u) : 0u;
  }
  cudaError_t e = launch_hist<K>(keys[0], n_cap, n_dev, plan.passes, sh, b.strip_hist, s);
  if (e != cudaSuccess) return e;
  const size_t tiles = (size_t)((n_cap + SORT_TILE - 1) / SORT_TILE);
  for (int p = 0; p < plan.passes; ++p) {
    const int src = p & 1, dst = src ^ 1;
    e = launch_pass<K, 0>(keys[src], keys[dst], b.svals[src], b.svals[dst], n_cap, n_dev, sh.shift[p], sh.mask[p],
                          b.strip_hist + (size_t)p * RADIX, b.strip_status + (size_t)p * tiles * RADIX,
                          b.counters + p, s);
    if (e != cudaSuccess) return e;
  }
  return cudaSuccess;
}

}  // namespace

// Sort the V compacted (depth bits, slot) pairs in g.dkeys[0] / g.dvals[0]; the digit histograms were
// accumulated by the compaction kernel.  V lives on the device (*v_dev); the grids cover the P slots.  The
// sorted slot order ends up in g.dvals[0] (4 passes).
cudaError_t launch_depth_sort(GeomWs& g, int P, const unsigned long long* v_dev, cudaStream_t s) {
  if (P <= 0) return cudaSuccess;
  { ProfScope ps(K_SORT_SCAN, s);
    scan_hist_kernel<<<SORT_DEPTH_PASSES, RADIX, 0, s>>>(g.depth_hist); }
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return e;
  for (int p = 0; p < SORT_DEPTH_PASSES; ++p) {
    const int src = p & 1, dst = src ^ 1;
    uint32_t* status = g.depth_status + (size_t)p * g.sort_tiles * RADIX;
    if (p + 1 < SORT_DEPTH_PASSES)
      e = launch_pass<uint32_t, 0>(g.dkeys[src], g.dkeys[dst], g.dvals[src], g.dvals[dst], P, v_dev,
                                   8 * p, 0xffu, g.depth_hist + (size_t)p * RADIX, status, g.dcounters + p, s);
    else
      e = launch_pass<uint32_t, 1>(g.dkeys[src], g.dkeys[dst], g.dvals[src], g.dvals[dst], P, v_dev,
                                   8 * p, 0xffu, g.depth_hist + (size_t)p * RADIX, status, g.dcounters + p, s);
    if (e != cudaSuccess) return e;
  }
  static_assert((SORT_DEPTH_PASSES & 1) == 0, "sorted order must land in dvals[0]");
  return cudaSuccess;
}

// Generic stable sort of n (u32 key, u32 value) pairs with the same onesweep passes (hier_build.cu: the three
// coordinate orders of the KD-tree).  ws: sort_pairs_ws_words(n) u32 words, zeroed here; the sorted pairs end up in
// keys[0] / vals[0] (4 passes).
size_t sort_pairs_ws_words(long long n) {
  const size_t tiles = (size_t)((n + SORT_TILE - 1) / SORT_TILE);
  return (size_t)SORT_DEPTH_PASSES * RADIX + (size_t)SORT_DEPTH_PASSES * tiles * RADIX + 8;
}
cudaError_t launch_sort_pairs(uint32_t* const keys[2], uint32_t* const vals[2], long long n, uint32_t* ws,
                              cudaStream_t s) {
  if (n <= 0) return cudaSuccess;
  const size_t tiles = (size_t)((n + SORT_TILE - 1) / SORT_TILE);
  uint32_t* hist = ws;
  uint32_t* status = hist + (size_t)SORT_DEPTH_PASSES * RADIX;
  uint32_t* counters = status + (size_t)SORT_DEPTH_PASSES * tiles * RADIX;
  cudaError_t e = cudaMemsetAsync(ws, 0, sort_pairs_ws_words(n) * sizeof(uint32_t), s);
  if (e != cudaSuccess) return e;
  SortShifts sh;
  for (int p = 0; p < SORT_MAX_PASSES; ++p) { sh.shift[p] = 8 * p; sh.mask[p] = 0xffu; }
  e = launch_hist<uint32_t>(keys[0], n, nullptr, SORT_DEPTH_PASSES, sh, hist, s);
  if (e != cudaSuccess) return e;
  for (int p = 0; p < SORT_DEPTH_PASSES; ++p) {
    const int src = p & 1, dst = src ^ 1;
    e = launch_pass<uint32_t, 0>(keys[src], keys[dst], vals[src], vals[dst], n, nullptr, 8 * p, 0xffu,
                                 hist + (size_t)p * RADIX, status + (size_t)p * tiles * RADIX, counters + p, s);
    if (e != cudaSuccess) return e;
  }
  return cudaSuccess;
}

// Stable partition of the emitted (strip id, slot) records by strip id; *n_dev of them (at most n_cap), sorted
// records in b.skeys[b.sorted] / b.svals[b.sorted].
cudaError_t launch_strip_sort(BinWs& b, int64_t n_cap, const unsigned long long* n_dev, const SortPlan& plan,
                              cudaStream_t s) {
  if (n_cap <= 0) return cudaSuccess;
  if (plan.key_bytes == 2) return strip_sort<uint16_t>(b, n_cap, n_dev, plan, s);
  return strip_sort<uint32_t>(b, n_cap, n_dev, plan, s);
}

}  // namespace h3dgs
