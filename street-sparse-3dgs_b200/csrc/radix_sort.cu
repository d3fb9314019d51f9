// Onesweep least-significant-digit radix sort of (u64 key, u32 value) pairs on the low `total` bits.
// Replaces cub::DeviceRadixSort::SortPairs at HR/cuda_rasterizer/rasterizer_impl.cu:352-357.
//
// One global histogram pass (all digits at once), then one kernel per digit that ranks a 6144-pair
// tile in shared memory, resolves its global offsets with a decoupled look-back over the previous
// tiles' per-digit counts and scatters -- every pair is read and written exactly once per digit.
// Digits are 9 bits wide (45 key bits at 1920x1080 -> 5 passes instead of the 6 an 8-bit digit
// needs).  The sort is stable, so equal (tile, depth) keys keep emission order exactly as with CUB.
#include "common.cuh"

namespace h3dgs {

namespace {

constexpr int RADIX = SORT_RADIX_MAX;
constexpr int WARPS = SORT_THREADS / 32;
constexpr uint32_t ST_AGG = 1u << 30;
constexpr uint32_t ST_INC = 2u << 30;
constexpr uint32_t ST_VAL = (1u << 30) - 1;

struct PlanDev {
  int num_passes;
  int shift[SORT_MAX_PASSES];
  uint32_t mask[SORT_MAX_PASSES];
};

__global__ void __launch_bounds__(512) radix_hist_kernel(const unsigned long long* __restrict__ keys,
                                                         long long n, PlanDev plan,
                                                         uint32_t* __restrict__ hist) {
  __shared__ uint32_t h[SORT_MAX_PASSES * RADIX];
  for (int i = threadIdx.x; i < plan.num_passes * RADIX; i += blockDim.x) h[i] = 0;
  __syncthreads();
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const unsigned long long k = keys[i];
#pragma unroll
    for (int p = 0; p < SORT_MAX_PASSES; ++p) {
      if (p < plan.num_passes)
        atomicAdd(&h[p * RADIX + (uint32_t)((k >> plan.shift[p]) & plan.mask[p])], 1u);
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < plan.num_passes * RADIX; i += blockDim.x) {
    const uint32_t c = h[i];
    if (c) atomicAdd(&hist[i], c);
  }
}

// exclusive scan of each pass' 512-bin histogram, in place (one block per pass)
__global__ void __launch_bounds__(RADIX) radix_scan_hist_kernel(uint32_t* __restrict__ hist) {
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

struct SweepSmem {
  uint32_t cnt[WARPS][RADIX];           // per-warp digit counters -> exclusive warp offsets
  unsigned long long keys[SORT_TILE];
  uint32_t vals[SORT_TILE];
  uint32_t bin_start[RADIX];            // start of each digit inside the tile
  uint32_t gbase[RADIX];                // global position of tile-local slot p with digit d: gbase[d] + p
  uint32_t warp_sums[WARPS];
  uint32_t tile;
};

__global__ void __launch_bounds__(SORT_THREADS) onesweep_kernel(
    const unsigned long long* __restrict__ kin, unsigned long long* __restrict__ kout,
    const uint32_t* __restrict__ vin, uint32_t* __restrict__ vout, long long n, int shift,
    uint32_t mask, const uint32_t* __restrict__ digit_base, volatile uint32_t* status,
    uint32_t* counter) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  SweepSmem& sm = *reinterpret_cast<SweepSmem*>(smem_raw);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

  if (tid == 0) sm.tile = atomicAdd(counter, 1u);
  for (int i = tid; i < WARPS * RADIX; i += SORT_THREADS) (&sm.cnt[0][0])[i] = 0;
  __syncthreads();
  const uint32_t tile = sm.tile;
  const long long tile_base = (long long)tile * SORT_TILE;
  const long long remaining = n - tile_base;
  const int tile_valid = (int)(remaining < SORT_TILE ? remaining : SORT_TILE);

  // warp-striped load: element order in memory == (warp, item, lane) order == ranking order
  unsigned long long key[SORT_ITEMS];
  const int wbase = warp * (32 * SORT_ITEMS) + lane;
#pragma unroll
  for (int i = 0; i < SORT_ITEMS; ++i) {
    const int local = wbase + i * 32;
    key[i] = (local < tile_valid) ? kin[tile_base + local] : ~0ull;
  }

  uint32_t rank[SORT_ITEMS];
  const uint32_t lt = (1u << lane) - 1u;
#pragma unroll
  for (int i = 0; i < SORT_ITEMS; ++i) {
    const uint32_t d = (uint32_t)((key[i] >> shift) & mask);
    const uint32_t peers = __match_any_sync(0xffffffffu, d);
    const int leader = __ffs(peers) - 1;
    uint32_t prior = 0;
    if (lane == leader) {
      prior = sm.cnt[warp][d];
      sm.cnt[warp][d] = prior + __popc(peers);
    }
    prior = __shfl_sync(0xffffffffu, prior, leader);
    rank[i] = prior + __popc(peers & lt);
    __syncwarp();
  }
  __syncthreads();

  // one thread per digit: warp counts -> exclusive warp offsets, tile total, look-back
  {
    const int d = tid;  // SORT_THREADS == RADIX
    uint32_t run = 0;
#pragma unroll
    for (int w = 0; w < WARPS; ++w) {
      const uint32_t c = sm.cnt[w][d];
      sm.cnt[w][d] = run;
      run += c;
    }
    const uint32_t total = run;
    volatile uint32_t* my_status = status + (size_t)tile * RADIX;
    my_status[d] = (tile == 0 ? ST_INC : ST_AGG) | total;

    // exclusive scan of totals over digits -> position of each digit inside the tile
    uint32_t incl = total;
#pragma unroll
    for (int k = 1; k < 32; k <<= 1) {
      uint32_t o = __shfl_up_sync(0xffffffffu, incl, k);
      if (lane >= k) incl += o;
    }
    if (lane == 31) sm.warp_sums[warp] = incl;
    __syncthreads();
    uint32_t off = 0;
    for (int w = 0; w < warp; ++w) off += sm.warp_sums[w];
    const uint32_t start = off + incl - total;
    sm.bin_start[d] = start;

    uint32_t excl = 0;
    if (tile > 0) {
      long long t = (long long)tile - 1;
      while (true) {
        volatile uint32_t* ps = status + (size_t)t * RADIX;
        uint32_t w;
        do { w = ps[d]; } while ((w >> 30) == 0);
        excl += w & ST_VAL;
        if ((w >> 30) == 2 || t == 0) break;
        --t;
      }
      my_status[d] = ST_INC | ((excl + total) & ST_VAL);
    }
    sm.gbase[d] = digit_base[d] + excl - start;
  }
  __syncthreads();

  // scatter into tile-local sorted order
#pragma unroll
  for (int i = 0; i < SORT_ITEMS; ++i) {
    const int local = wbase + i * 32;
    const uint32_t d = (uint32_t)((key[i] >> shift) & mask);
    const uint32_t pos = sm.bin_start[d] + sm.cnt[warp][d] + rank[i];
    sm.keys[pos] = key[i];
    sm.vals[pos] = (local < tile_valid) ? vin[tile_base + local] : 0u;
  }
  __syncthreads();

  // coalesced write-out; padding keys (all ones) sort to the very end of the tile
  for (int p = tid; p < tile_valid; p += SORT_THREADS) {
    const unsigned long long k = sm.keys[p];
    const uint32_t d = (uint32_t)((k >> shift) & mask);
    const uint32_t g = sm.gbase[d] + (uint32_t)p;
    kout[g] = k;
    vout[g] = sm.vals[p];
  }
}

}  // namespace

cudaError_t launch_radix_sort(BinWs& b, int64_t R, const SortPlan& plan, cudaStream_t s) {
  if (R <= 0) return cudaSuccess;
  static_assert(SORT_THREADS == RADIX, "one thread per digit");
  static bool attr_done = false;
  const size_t smem = sizeof(SweepSmem);
  if (!attr_done) {
    cudaError_t e = cudaFuncSetAttribute(onesweep_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)smem);
    if (e != cudaSuccess) return e;
    attr_done = true;
  }
  cudaError_t e;
  { ProfScope ps(K_MEMSET, s); e = cudaMemsetAsync(b.zero_base, 0, b.zero_bytes, s); }
  if (e != cudaSuccess) return e;

  PlanDev pd;
  pd.num_passes = plan.num_passes;
  for (int i = 0; i < SORT_MAX_PASSES; ++i) {
    pd.shift[i] = i < plan.num_passes ? plan.shift[i] : 0;
    pd.mask[i] = i < plan.num_passes ? ((1u << plan.bits[i]) - 1u) : 0u;
  }
  const long long tiles = (R + SORT_TILE - 1) / SORT_TILE;
  int hist_blocks = (int)((R + 512 * 16 - 1) / (512 * 16));
  if (hist_blocks > 148 * 4) hist_blocks = 148 * 4;
  if (hist_blocks < 1) hist_blocks = 1;
  { ProfScope ps(K_SORT_HIST, s);
  radix_hist_kernel<<<hist_blocks, 512, 0, s>>>(b.keys[0], R, pd, b.hist); }
  { ProfScope ps(K_SORT_SCAN, s);
  radix_scan_hist_kernel<<<plan.num_passes, RADIX, 0, s>>>(b.hist); }
  for (int p = 0; p < plan.num_passes; ++p) {
    const int src = p & 1, dst = src ^ 1;
    ProfScope ps(K_SORT_PASS, s);
    onesweep_kernel<<<(unsigned)tiles, SORT_THREADS, smem, s>>>(
        b.keys[src], b.keys[dst], b.vals[src], b.vals[dst], R, pd.shift[p], pd.mask[p],
        b.hist + (size_t)p * RADIX, b.status + (size_t)p * tiles * RADIX, b.counters + p);
  }
  return cudaGetLastError();
}

}  // namespace h3dgs
