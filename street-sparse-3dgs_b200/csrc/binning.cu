// Binning: visible-slot compaction, ordered instance emission and per-tile range detection.
// Replaces duplicateWithKeys and identifyTileRanges (HR/cuda_rasterizer/rasterizer_impl.cu:70-142).
//
// The reference emits (tile << 32 | depth, slot) pairs in slot order and sorts all R of them on 45 bits.
// Here the V visible slots are first put in (depth, slot) order (a 32-bit sort over V << R items,
// radix_sort.cu), instances are emitted in THAT order -- row-major over each slot's tile rectangle, as
// the reference does -- and then stably partitioned by tile id only.  Within a tile the list is therefore
// ordered by (depth bits, slot), which is exactly what the reference's stable sort of the full keys yields.
#include "common.cuh"

namespace h3dgs {

namespace {

constexpr int EMIT_SERIAL_MAX = 8;  // rectangles up to this many tiles are written by their own thread

// slot -> compact position rank-1; also the four 8-bit digit histograms of the depth keys
__global__ void __launch_bounds__(256) compact_visible_kernel(
    int P, const uint32_t* __restrict__ tiles_touched, const uint32_t* __restrict__ rank,
    const float* __restrict__ depth, uint32_t* __restrict__ dkeys, uint32_t* __restrict__ dvals,
    uint32_t* __restrict__ hist) {
  __shared__ uint32_t h[SORT_DEPTH_PASSES * SORT_RADIX];
  for (int i = threadIdx.x; i < SORT_DEPTH_PASSES * SORT_RADIX; i += blockDim.x) h[i] = 0;
  __syncthreads();
  const int stride = gridDim.x * blockDim.x;
  for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < P; idx += stride) {
    if (tiles_touched[idx] == 0) continue;
    const uint32_t pos = rank[idx] - 1u;
    const uint32_t k = __float_as_uint(depth[idx]);
    dkeys[pos] = k;
    dvals[pos] = (uint32_t)idx;
#pragma unroll
    for (int p = 0; p < SORT_DEPTH_PASSES; ++p) atomicAdd(&h[p * SORT_RADIX + ((k >> (8 * p)) & 0xffu)], 1u);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < SORT_DEPTH_PASSES * SORT_RADIX; i += blockDim.x) {
    const uint32_t c = h[i];
    if (c) atomicAdd(&hist[i], c);
  }
}

// one thread per visible slot, in (depth, slot) order
template <typename K>
__global__ void __launch_bounds__(256) emit_instances_kernel(
    long long V, const uint32_t* __restrict__ order, const uint32_t* __restrict__ soffs,
    const ushort4* __restrict__ rect, int grid_x, K* __restrict__ keys, uint32_t* __restrict__ vals) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const int lane = threadIdx.x & 31;
  uint32_t off = 0, slot = 0;
  int x0 = 0, y0 = 0, w = 0, n = 0;
  if (i < V) {
    slot = order[i];
    off = (i == 0) ? 0u : soffs[i - 1];
    const ushort4 r = rect[slot];
    x0 = r.x; y0 = r.y;
    w = (int)r.z - (int)r.x;
    n = w * ((int)r.w - (int)r.y);
  }
  if (n > 0 && n <= EMIT_SERIAL_MAX) {
    int x = 0, y = 0;
    for (int j = 0; j < n; ++j) {
      keys[off + j] = (K)((y0 + y) * grid_x + (x0 + x));
      vals[off + j] = slot;
      if (++x == w) { x = 0; ++y; }
    }
  }
  // wide rectangles: the whole warp writes one rectangle at a time (coalesced, balanced)
  unsigned big = __ballot_sync(0xffffffffu, n > EMIT_SERIAL_MAX);
  while (big) {
    const int src = __ffs(big) - 1;
    big &= big - 1;
    const uint32_t o = __shfl_sync(0xffffffffu, off, src);
    const uint32_t id = __shfl_sync(0xffffffffu, slot, src);
    const int bx0 = __shfl_sync(0xffffffffu, x0, src);
    const int by0 = __shfl_sync(0xffffffffu, y0, src);
    const int bw = __shfl_sync(0xffffffffu, w, src);
    const int bn = __shfl_sync(0xffffffffu, n, src);
    for (int j = lane; j < bn; j += 32) {
      const int y = j / bw, x = j - y * bw;
      keys[o + j] = (K)((by0 + y) * grid_x + (bx0 + x));
      vals[o + j] = id;
    }
  }
}

__global__ void __launch_bounds__(256) tile_ranges_kernel(long long L,
                                                          const unsigned long long* __restrict__ keys,
                                                          uint2* __restrict__ ranges) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= L) return;
  const uint32_t cur = (uint32_t)(keys[i] >> 32);
  if (i == 0) {
    ranges[cur].x = 0;
  } else {
    const uint32_t prev = (uint32_t)(keys[i - 1] >> 32);
    if (cur != prev) {
      ranges[prev].y = (uint32_t)i;
      ranges[cur].x = (uint32_t)i;
    }
    // Kept inside the i > 0 branch as in the reference (rasterizer_impl.cu:139-140): a render with
    // exactly one instance leaves its tile range at (0,0).  Parity over repair; see DESIGN.md.
    if (i == L - 1) ranges[cur].y = (uint32_t)L;
  }
}

}  // namespace

cudaError_t launch_compact_visible(int P, const GeomWs& g, BinWs& b, cudaStream_t s) {
  if (P <= 0) return cudaSuccess;
  int blocks = (P + 256 * 8 - 1) / (256 * 8);
  if (blocks > 148 * 8) blocks = 148 * 8;
  ProfScope ps(K_EMIT, s);
  compact_visible_kernel<<<blocks, 256, 0, s>>>(P, g.tiles_touched, g.offsets, g.depth, b.dkeys[0], b.dvals[0],
                                                b.depth_hist);
  return cudaGetLastError();
}

cudaError_t launch_emit_instances(int64_t V, const GeomWs& g, BinWs& b, int grid_x, int key_bytes,
                                  cudaStream_t s) {
  if (V <= 0) return cudaSuccess;
  const unsigned blocks = (unsigned)((V + 255) / 256);
  ProfScope ps(K_EMIT, s);
  if (key_bytes == 2)
    emit_instances_kernel<uint16_t><<<blocks, 256, 0, s>>>(V, b.dvals[0], b.soffs, g.tile_rect, grid_x,
                                                           reinterpret_cast<uint16_t*>(b.tkeys[0]), b.tvals[0]);
  else
    emit_instances_kernel<uint32_t><<<blocks, 256, 0, s>>>(V, b.dvals[0], b.soffs, g.tile_rect, grid_x,
                                                           reinterpret_cast<uint32_t*>(b.tkeys[0]), b.tvals[0]);
  return cudaGetLastError();
}

cudaError_t launch_tile_ranges(int64_t R, const unsigned long long* keys, uint2* ranges, int tiles,
                               cudaStream_t s) {
  cudaError_t e = cudaMemsetAsync(ranges, 0, sizeof(uint2) * (size_t)tiles, s);
  if (e != cudaSuccess || R <= 0) return e;
  const long long blocks = (R + 255) / 256;
  ProfScope ps(K_TILE_RANGES, s);
  tile_ranges_kernel<<<(unsigned)blocks, 256, 0, s>>>(R, keys, ranges);
  return cudaGetLastError();
}

}  // namespace h3dgs
