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

// Ordered emission, load balanced over OUTPUT positions: CTA c writes instances [c*EMIT_CHUNK, (c+1)*EMIT_CHUNK).
// The gathered scan (scan.cu, MODE 2) recorded for every chunk the first visible slot (in (depth, slot)
// order) that reaches into it; the CTA stages the end offsets of its slots in shared memory and every thread
// locates the slot of its output position by binary search there.  A slot covering 8000 tiles and 8000 slots
// covering one tile each cost the same, and all stores are coalesced.
template <typename K>
__global__ void __launch_bounds__(256) emit_instances_kernel(
    long long R, long long V, const uint32_t* __restrict__ order, const uint32_t* __restrict__ soffs,
    const uint32_t* __restrict__ chunk_first, const ushort4* __restrict__ rect, int grid_x,
    K* __restrict__ keys, uint32_t* __restrict__ vals) {
  __shared__ uint32_t ends[EMIT_CHUNK + 1];
  const long long c = blockIdx.x;
  const long long cs = c * EMIT_CHUNK;
  const long long ce = (cs + EMIT_CHUNK < R) ? cs + EMIT_CHUNK : R;
  const uint32_t lo = chunk_first[c];
  const uint32_t hi = (c + 1 < (long long)gridDim.x) ? chunk_first[c + 1] : (uint32_t)(V - 1);
  const int ns = (int)(hi - lo + 1);   // <= EMIT_CHUNK + 1: every slot has at least one instance
  for (int k = threadIdx.x; k < ns; k += blockDim.x) ends[k] = soffs[lo + k];
  const uint32_t first_start = (lo == 0) ? 0u : soffs[lo - 1];
  __syncthreads();
  for (long long o = cs + threadIdx.x; o < ce; o += blockDim.x) {
    const uint32_t pos = (uint32_t)o;
    int a = 0, b = ns - 1;   // smallest k with ends[k] > pos
    while (a < b) {
      const int m = (a + b) >> 1;
      if (ends[m] > pos) b = m; else a = m + 1;
    }
    const uint32_t start = (a == 0) ? first_start : ends[a - 1];
    const uint32_t slot = __ldg(order + lo + a);
    const ushort4 r = rect[slot];
    const int w = (int)r.z - (int)r.x;
    const int j = (int)(pos - start);
    const int y = j / w, x = j - y * w;
    keys[o] = (K)(((int)r.y + y) * grid_x + ((int)r.x + x));
    vals[o] = slot;
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

// Heavy-first launch order for the blend kernels.  Tile work is wildly uneven in street views (median list
// ~1.7 k entries, horizon tiles 25 k): in blockIdx order the few longest tiles start late and the SMs idle
// 40 % of the kernel waiting for them.  One CTA counting-sorts the tiles into 64 logarithmic buckets
// (4 per octave of the length), longest first; the order inside a bucket does not matter.
__device__ __forceinline__ int work_bucket(uint32_t len) {
  if (len < 4u) return 63;
  const int msb = 31 - __clz(len);
  const int sub = (int)((len >> (msb - 2)) & 3u);
  const int b = msb * 4 + sub - 8;          // len 4..7 -> 0..3, doubling every 4
  return 63 - (b > 63 ? 63 : b);
}
__global__ void __launch_bounds__(1024) tile_order_kernel(const uint2* __restrict__ ranges,
                                                          const uint32_t* __restrict__ lengths, int tiles,
                                                          uint32_t* __restrict__ order) {
  __shared__ uint32_t cnt[64];
  if (threadIdx.x < 64) cnt[threadIdx.x] = 0;
  __syncthreads();
  for (int t = threadIdx.x; t < tiles; t += blockDim.x) {
    const uint32_t len = ranges ? (ranges[t].y - ranges[t].x) : lengths[t];
    atomicAdd(&cnt[work_bucket(len)], 1u);
  }
  __syncthreads();
  if (threadIdx.x < 32) {
    // exclusive scan of the 64 counters by one warp (two per lane)
    const uint32_t a = cnt[2 * threadIdx.x], b = cnt[2 * threadIdx.x + 1];
    uint32_t incl = a + b;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const uint32_t o = __shfl_up_sync(0xffffffffu, incl, d);
      if ((int)threadIdx.x >= d) incl += o;
    }
    cnt[2 * threadIdx.x] = incl - a - b;
    cnt[2 * threadIdx.x + 1] = incl - b;
  }
  __syncthreads();
  for (int t = threadIdx.x; t < tiles; t += blockDim.x) {
    const uint32_t len = ranges ? (ranges[t].y - ranges[t].x) : lengths[t];
    order[atomicAdd(&cnt[work_bucket(len)], 1u)] = (uint32_t)t;
  }
}

}  // namespace

cudaError_t launch_tile_order(const uint2* ranges, const uint32_t* lengths, int tiles, uint32_t* order,
                              cudaStream_t s) {
  if (tiles <= 0) return cudaSuccess;
  ProfScope ps(K_TILE_RANGES, s);
  tile_order_kernel<<<1, 1024, 0, s>>>(ranges, lengths, tiles, order);
  return cudaGetLastError();
}

cudaError_t launch_compact_visible(int P, GeomWs& g, cudaStream_t s) {
  if (P <= 0) return cudaSuccess;
  int blocks = (P + 256 * 8 - 1) / (256 * 8);
  if (blocks > 148 * 8) blocks = 148 * 8;
  ProfScope ps(K_EMIT, s);
  compact_visible_kernel<<<blocks, 256, 0, s>>>(P, g.tiles_touched, g.offsets, g.depth, g.dkeys[0], g.dvals[0],
                                                g.depth_hist);
  return cudaGetLastError();
}

cudaError_t launch_emit_instances(int64_t R, int64_t V, const GeomWs& g, BinWs& b, int grid_x, int key_bytes,
                                  cudaStream_t s) {
  if (V <= 0 || R <= 0) return cudaSuccess;
  const unsigned blocks = (unsigned)((R + EMIT_CHUNK - 1) / EMIT_CHUNK);
  ProfScope ps(K_EMIT, s);
  if (key_bytes == 2)
    emit_instances_kernel<uint16_t><<<blocks, 256, 0, s>>>(R, V, g.dvals[0], g.soffs, g.chunk_first, g.tile_rect,
                                                           grid_x, reinterpret_cast<uint16_t*>(b.tkeys[0]),
                                                           b.tvals[0]);
  else
    emit_instances_kernel<uint32_t><<<blocks, 256, 0, s>>>(R, V, g.dvals[0], g.soffs, g.chunk_first, g.tile_rect,
                                                           grid_x, reinterpret_cast<uint32_t*>(b.tkeys[0]),
                                                           b.tvals[0]);
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
