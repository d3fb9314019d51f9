// Instance emission ((tile, depth) keys + slot values) and per-tile range detection.
// Replaces duplicateWithKeys and identifyTileRanges (HR/cuda_rasterizer/rasterizer_impl.cu:70-142).
// Emission order within a slot is row-major over its tile rectangle and slots are laid out by the
// inclusive scan, so the unsorted arrays are element-for-element those of the reference; the stable
// sort that follows therefore yields identical sorted lists.
#include "common.cuh"

namespace h3dgs {

namespace {

constexpr int EMIT_SERIAL_MAX = 8;  // rectangles up to this many tiles are written by their own thread

__global__ void __launch_bounds__(256) emit_instances_kernel(
    int P, const int* __restrict__ radii, const float* __restrict__ depth,
    const ushort4* __restrict__ rect, const uint32_t* __restrict__ offsets, int grid_x,
    unsigned long long* __restrict__ keys, uint32_t* __restrict__ vals) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  const int lane = threadIdx.x & 31;
  uint32_t off = 0, dbits = 0;
  int x0 = 0, y0 = 0, w = 0, n = 0;
  if (idx < P && radii[idx] > 0) {
    off = (idx == 0) ? 0u : offsets[idx - 1];
    const ushort4 r = rect[idx];
    x0 = r.x; y0 = r.y;
    w = (int)r.z - (int)r.x;
    n = w * ((int)r.w - (int)r.y);
    dbits = __float_as_uint(depth[idx]);
  }
  if (n > 0 && n <= EMIT_SERIAL_MAX) {
    int x = 0, y = 0;
    for (int i = 0; i < n; ++i) {
      unsigned long long key = (unsigned long long)(uint32_t)((y0 + y) * grid_x + (x0 + x));
      key = (key << 32) | dbits;
      keys[off + i] = key;
      vals[off + i] = (uint32_t)idx;
      if (++x == w) { x = 0; ++y; }
    }
  }
  // wide rectangles: the whole warp writes one rectangle at a time (coalesced, balanced)
  unsigned big = __ballot_sync(0xffffffffu, n > EMIT_SERIAL_MAX);
  while (big) {
    const int src = __ffs(big) - 1;
    big &= big - 1;
    const uint32_t o = __shfl_sync(0xffffffffu, off, src);
    const uint32_t db = __shfl_sync(0xffffffffu, dbits, src);
    const int bx0 = __shfl_sync(0xffffffffu, x0, src);
    const int by0 = __shfl_sync(0xffffffffu, y0, src);
    const int bw = __shfl_sync(0xffffffffu, w, src);
    const int bn = __shfl_sync(0xffffffffu, n, src);
    const uint32_t id = (uint32_t)(idx - lane + src);
    for (int i = lane; i < bn; i += 32) {
      const int y = i / bw, x = i - y * bw;
      unsigned long long key = (unsigned long long)(uint32_t)((by0 + y) * grid_x + (bx0 + x));
      key = (key << 32) | db;
      keys[o + i] = key;
      vals[o + i] = id;
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

cudaError_t launch_emit_instances(int P, const int* radii, const GeomWs& g, int grid_x,
                                  unsigned long long* keys, uint32_t* vals, cudaStream_t s) {
  if (P <= 0) return cudaSuccess;
  ProfScope ps(K_EMIT, s);
  emit_instances_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, radii, g.depth, g.tile_rect, g.offsets,
                                                        grid_x, keys, vals);
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
