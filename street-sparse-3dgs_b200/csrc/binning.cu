// Binning: visible-slot compaction, ordered strip-record emission, expansion into per-tile lists and ranges.
// Replaces duplicateWithKeys and identifyTileRanges (HR/cuda_rasterizer/rasterizer_impl.cu:70-142).
//
// The reference emits (tile << 32 | depth, slot) pairs in slot order and sorts all R of them on 45 bits.
// Here the V visible slots are first put in (depth, slot) order (a 32-bit sort over V << R items,
// radix_sort.cu).  In THAT order one record is emitted per (slot, tile row, strip of STRIP_TILES tiles of the row)
// the slot's rectangle reaches into -- about R / 3 records -- and the records are stably partitioned by strip id
// (one 8-bit sweep at 1080p).  A warp then takes EXPAND_CHUNK consecutive records of one strip: every record is
// a run of covered tiles, i.e. a 64-bit mask; 32 masks are transposed across the warp (5 shuffle stages per 32
// tiles), after which lane j holds, for tiles j and j + 32 of the strip, the bit set of records covering them.
// A counting pass, one scan over the ragged (strip, tile, chunk) table and a writing pass put every slot at its
// final place: inside a tile the list is ordered by (depth bits, slot), exactly what the reference's stable sort
// of the full keys yields, and the table's prefix sums ARE the tile ranges.
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

// Ordered emission, load balanced over OUTPUT positions: CTA c writes records [c*EMIT_CHUNK, (c+1)*EMIT_CHUNK).
// The gathered scan (scan.cu, MODE 3) recorded for every chunk the first visible slot (in (depth, slot)
// order) that reaches into it; the CTA stages the end offsets of its slots in shared memory and every thread
// locates the slot of its output position by binary search there.  All stores are coalesced.  The record count
// lives on the device (*n_dev); the grid covers its upper bound R.
template <typename K>
__global__ void __launch_bounds__(256) emit_strips_kernel(
    const unsigned long long* __restrict__ n_dev, long long V, const uint32_t* __restrict__ order,
    const uint32_t* __restrict__ soffs, const uint32_t* __restrict__ chunk_first, const ushort4* __restrict__ rect,
    int strips_per_row, K* __restrict__ keys, uint32_t* __restrict__ vals) {
  __shared__ uint32_t ends[EMIT_CHUNK + 1];
  const long long n = (long long)*n_dev;
  const long long c = blockIdx.x;
  const long long cs = c * EMIT_CHUNK;
  if (cs >= n) return;
  const long long ce = (cs + EMIT_CHUNK < n) ? cs + EMIT_CHUNK : n;
  const uint32_t lo = chunk_first[c];
  const uint32_t hi = (cs + EMIT_CHUNK < n) ? chunk_first[c + 1] : (uint32_t)(V - 1);
  const int ns = (int)(hi - lo + 1);   // <= EMIT_CHUNK + 1: every slot has at least one record
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
    const int sc0 = (int)r.x / STRIP_TILES;
    const int w = ((int)r.z - 1) / STRIP_TILES - sc0 + 1;
    const int j = (int)(pos - start);
    const int y = j / w, x = j - y * w;
    keys[o] = (K)(((int)r.y + y) * strips_per_row + (sc0 + x));
    vals[o] = slot;
  }
}

// Records of each strip in the sorted order: boundaries of the sorted strip ids (strip_range is pre-zeroed, empty
// strips stay (0, 0)).
template <typename K>
__global__ void __launch_bounds__(256) strip_ranges_kernel(const unsigned long long* __restrict__ n_dev,
                                                           const K* __restrict__ keys, uint2* __restrict__ range) {
  const long long n = (long long)*n_dev;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const uint32_t cur = (uint32_t)keys[i];
    if (i == 0) {
      range[cur].x = 0;
    } else {
      const uint32_t prev = (uint32_t)keys[i - 1];
      if (cur != prev) {
        range[prev].y = (uint32_t)i;
        range[cur].x = (uint32_t)i;
      }
    }
    if (i == n - 1) range[cur].y = (uint32_t)n;
  }
}

// One CTA: chunks per strip, their exclusive scan, the strip of every chunk, and the totals the next kernels size
// themselves by (counters[6] = chunks, counters[4..5] = u64 length of the ragged count table).
__global__ void __launch_bounds__(1024) chunk_table_kernel(int strips, const uint2* __restrict__ range,
                                                           uint32_t* __restrict__ nch, uint32_t* __restrict__ cbase,
                                                           uint32_t* __restrict__ chunk_strip,
                                                           uint32_t* __restrict__ counters) {
  __shared__ uint32_t s_warp[32];
  __shared__ uint32_t s_carry;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) s_carry = 0;
  __syncthreads();
  for (int s0 = 0; s0 < strips; s0 += 1024) {
    const int s = s0 + (int)threadIdx.x;
    uint32_t c = 0;
    if (s < strips) {
      const uint2 r = range[s];
      c = (r.y - r.x + EXPAND_CHUNK - 1) / EXPAND_CHUNK;
    }
    uint32_t incl = c;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const uint32_t o = __shfl_up_sync(0xffffffffu, incl, d);
      if (lane >= d) incl += o;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    uint32_t off = s_carry;
    for (int w = 0; w < warp; ++w) off += s_warp[w];
    const uint32_t first = off + incl - c;
    if (s < strips) {
      nch[s] = c;
      cbase[s] = first;
      for (uint32_t k = 0; k < c; ++k) chunk_strip[first + k] = (uint32_t)s;
    }
    __syncthreads();
    if (threadIdx.x == 1023) s_carry = off + incl;
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    counters[6] = s_carry;
    *reinterpret_cast<unsigned long long*>(counters + 4) = (unsigned long long)s_carry * STRIP_TILES;
  }
}

// 32 x 32 bit-matrix transpose across a warp: lane r gives row r, lane c receives column c (bit r = row r's bit c).
__device__ __forceinline__ uint32_t transpose32(uint32_t x, int lane) {
#pragma unroll
  for (int j = 16; j >= 1; j >>= 1) {
    const uint32_t m = j == 16 ? 0x0000ffffu : j == 8 ? 0x00ff00ffu : j == 4 ? 0x0f0f0f0fu : j == 2 ? 0x33333333u
                                                                                                     : 0x55555555u;
    const uint32_t y = __shfl_xor_sync(0xffffffffu, x, j);
    x = (lane & j) ? (((y >> j) & m) | (x & ~m)) : ((x & m) | ((y & m) << j));
  }
  return x;
}

// One warp per chunk of EXPAND_CHUNK sorted records of one strip.  WRITE 0: count the instances every tile of the
// strip receives from the chunk.  WRITE 1: with the scanned table, write the slots to their final places.
// Table layout: entry (strip s, tile j of the strip, chunk k of the strip) at STRIP_TILES * cbase[s] + j * nch[s] + k,
// so that one scan over the table in memory order walks tiles in id order and, inside a tile, chunks in order.
template <int WRITE>
__global__ void __launch_bounds__(256) expand_kernel(const uint32_t* __restrict__ counters,
                                                     const uint32_t* __restrict__ chunk_strip,
                                                     const uint2* __restrict__ range, const uint32_t* __restrict__ nch,
                                                     const uint32_t* __restrict__ cbase,
                                                     const uint32_t* __restrict__ svals, const ushort4* __restrict__ rect,
                                                     int strips_per_row, uint32_t* __restrict__ counts,
                                                     const uint32_t* __restrict__ offs, uint32_t* __restrict__ point_list) {
  __shared__ uint32_t s_slot[8][32];
  // WRITE: the four slots of a 16-byte aligned group of a tile's list are collected here and stored at once (every
  // lane appends to its own tile's list: scalar stores were one 4-byte L2 write per instance)
  __shared__ __align__(16) uint32_t s_quad[WRITE ? 8 : 1][WRITE ? STRIP_TILES : 1][4];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t chunk = blockIdx.x * 8u + (uint32_t)warp;
  if (chunk >= counters[6]) return;
  const uint32_t s = chunk_strip[chunk];
  const uint32_t n = nch[s], k = chunk - cbase[s];
  const uint2 sr = range[s];
  const uint32_t rs = sr.x + k * EXPAND_CHUNK;
  const uint32_t re = min(sr.y, rs + EXPAND_CHUNK);
  const int tile0 = (int)(s % (uint32_t)strips_per_row) * STRIP_TILES;   // first tile column of the strip
  const size_t e0 = (size_t)STRIP_TILES * cbase[s] + (size_t)lane * n + k;
  const size_t e1 = e0 + (size_t)32 * n;
  uint32_t pos0 = 0, pos1 = 0;
  // positions [lo, hi) of a list are written as whole aligned groups of four, the rest one by one
  uint32_t lo0 = 0, hi0 = 0, lo1 = 0, hi1 = 0;
  if (WRITE) {
    const uint32_t end0 = offs[e0], end1 = offs[e1];
    pos0 = end0 - counts[e0];
    pos1 = end1 - counts[e1];
    lo0 = (pos0 + 3u) & ~3u; hi0 = end0 & ~3u;
    lo1 = (pos1 + 3u) & ~3u; hi1 = end1 & ~3u;
  }
  // software pipeline: the records of the next group are in flight while this one is expanded
  uint32_t slot_n = 0;
  uint2 xr_n = make_uint2(0u, 0u);
  if (rs + lane < re) {
    slot_n = __ldg(svals + rs + lane);
    const ushort4 r = rect[slot_n];
    xr_n = make_uint2(r.x, r.z);
  }
  for (uint32_t i = rs; i < re; i += 32) {
    const uint32_t slot = slot_n;
    const uint2 xr = xr_n;
    const bool valid = i + lane < re;
    slot_n = 0;
    xr_n = make_uint2(0u, 0u);
    if (i + 32 + lane < re) {
      slot_n = __ldg(svals + i + 32 + lane);
      const ushort4 r = rect[slot_n];
      xr_n = make_uint2(r.x, r.z);
    }
    // covered tiles of the strip as a 64-bit mask
    unsigned long long mask = 0;
    if (valid) {
      const int a = max((int)xr.x - tile0, 0), b = min((int)xr.y - tile0, STRIP_TILES);   // 0 <= a < b <= 64
      const unsigned long long upto_b = b >= 64 ? ~0ull : ((1ull << b) - 1ull);
      mask = upto_b & ~((1ull << a) - 1ull);
    }
    const uint32_t c0 = transpose32((uint32_t)mask, lane);          // records covering tile lane
    const uint32_t c1 = transpose32((uint32_t)(mask >> 32), lane);  // records covering tile lane + 32
    if (!WRITE) {
      pos0 += __popc(c0);
      pos1 += __popc(c1);
    } else {
      __syncwarp();
      s_slot[warp][lane] = slot;
      __syncwarp();
      uint32_t bts = c0;
      while (bts) {
        const int e = __ffs(bts) - 1;
        bts &= bts - 1;
        const uint32_t v = s_slot[warp][e];
        if (pos0 < lo0 || pos0 >= hi0) {
          point_list[pos0] = v;
        } else {
          s_quad[warp][lane][pos0 & 3u] = v;
          if ((pos0 & 3u) == 3u)
            *reinterpret_cast<uint4*>(point_list + (pos0 - 3u)) = *reinterpret_cast<const uint4*>(s_quad[warp][lane]);
        }
        ++pos0;
      }
      bts = c1;
      while (bts) {
        const int e = __ffs(bts) - 1;
        bts &= bts - 1;
        const uint32_t v = s_slot[warp][e];
        if (pos1 < lo1 || pos1 >= hi1) {
          point_list[pos1] = v;
        } else {
          s_quad[warp][lane + 32][pos1 & 3u] = v;
          if ((pos1 & 3u) == 3u)
            *reinterpret_cast<uint4*>(point_list + (pos1 - 3u)) = *reinterpret_cast<const uint4*>(s_quad[warp][lane + 32]);
        }
        ++pos1;
      }
    }
  }
  if (!WRITE) {
    counts[e0] = pos0;
    counts[e1] = pos1;
  }
}

// Tile ranges straight from the scanned table: the list of tile (y, x) starts at the exclusive prefix of its first
// chunk and ends at the inclusive prefix of its last one.  Empty tiles keep (0, 0) as identifyTileRanges leaves
// them, and so does everything when the render has exactly one instance (rasterizer_impl.cu:139-140 sets the end of
// the last range inside the i > 0 branch; parity over repair).  seg_tile[x] = tile whose list holds
// instance x * BLEND_SEG (the backward blend's work items start there).
__global__ void __launch_bounds__(256) tile_ranges_kernel(int tiles, int grid_x, int strips_per_row, long long R,
                                                          const uint32_t* __restrict__ nch,
                                                          const uint32_t* __restrict__ cbase,
                                                          const uint32_t* __restrict__ counts,
                                                          const uint32_t* __restrict__ offs, uint2* __restrict__ ranges,
                                                          uint32_t* __restrict__ seg_tile) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= tiles) return;
  const int y = t / grid_x, x = t - y * grid_x;
  const int s = y * strips_per_row + x / STRIP_TILES, j = x % STRIP_TILES;
  const uint32_t n = nch[s];
  uint32_t start = 0, end = 0;
  if (n > 0 && R > 1) {
    const size_t e = (size_t)STRIP_TILES * cbase[s] + (size_t)j * n;
    start = offs[e] - counts[e];
    end = offs[e + n - 1];
  }
  if (end == start) start = end = 0;
  ranges[t] = make_uint2(start, end);
  for (uint32_t m = (start + BLEND_SEG - 1) / BLEND_SEG * BLEND_SEG; m < end; m += BLEND_SEG)
    seg_tile[m / BLEND_SEG] = (uint32_t)t;
}

// point_list_keys for h3dgs_debug_view: one CTA per tile
__global__ void __launch_bounds__(256) fill_keys64_kernel(const uint2* __restrict__ ranges,
                                                          const uint32_t* __restrict__ point_list,
                                                          const uint32_t* __restrict__ depth_bits,
                                                          unsigned long long* __restrict__ keys64) {
  const uint2 r = ranges[blockIdx.x];
  for (uint32_t i = r.x + threadIdx.x; i < r.y; i += blockDim.x)
    keys64[i] = ((unsigned long long)blockIdx.x << 32) | (unsigned long long)depth_bits[point_list[i]];
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

cudaError_t launch_emit_strips(int64_t R, int64_t V, const GeomWs& g, BinWs& b, const SortPlan& plan, cudaStream_t s) {
  if (V <= 0 || R <= 0) return cudaSuccess;
  const unsigned blocks = (unsigned)((R + EMIT_CHUNK - 1) / EMIT_CHUNK);   // R bounds the record count
  const unsigned long long* n_dev = reinterpret_cast<const unsigned long long*>(g.gscan_counters + 2);
  ProfScope ps(K_EMIT, s);
  if (plan.key_bytes == 2)
    emit_strips_kernel<uint16_t><<<blocks, 256, 0, s>>>(n_dev, V, g.dvals[0], g.soffs, g.chunk_first, g.tile_rect,
                                                        plan.strips_per_row, reinterpret_cast<uint16_t*>(b.skeys[0]),
                                                        b.svals[0]);
  else
    emit_strips_kernel<uint32_t><<<blocks, 256, 0, s>>>(n_dev, V, g.dvals[0], g.soffs, g.chunk_first, g.tile_rect,
                                                        plan.strips_per_row, reinterpret_cast<uint32_t*>(b.skeys[0]),
                                                        b.svals[0]);
  return cudaGetLastError();
}

// sorted strip records -> point_list, ranges, seg_tile
cudaError_t launch_expand(int64_t R, const GeomWs& g, BinWs& b, const SortPlan& plan, int grid_x, int grid_y,
                          uint2* ranges, cudaStream_t s) {
  if (R <= 0) return cudaSuccess;
  const unsigned long long* n_dev = reinterpret_cast<const unsigned long long*>(g.gscan_counters + 2);
  {
    ProfScope ps(K_TILE_RANGES, s);
    if (plan.key_bytes == 2)
      strip_ranges_kernel<uint16_t><<<148 * 8, 256, 0, s>>>(n_dev, reinterpret_cast<const uint16_t*>(b.skeys[b.sorted]),
                                                            b.strip_range);
    else
      strip_ranges_kernel<uint32_t><<<148 * 8, 256, 0, s>>>(n_dev, reinterpret_cast<const uint32_t*>(b.skeys[b.sorted]),
                                                            b.strip_range);
    chunk_table_kernel<<<1, 1024, 0, s>>>(plan.strips, b.strip_range, b.strip_nch, b.strip_cbase, b.chunk_strip,
                                          b.counters);
  }
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return e;
  const unsigned blocks = (unsigned)((b.max_chunks + 7) / 8);
  {
    ProfScope ps(K_EMIT, s);
    expand_kernel<0><<<blocks, 256, 0, s>>>(b.counters, b.chunk_strip, b.strip_range, b.strip_nch, b.strip_cbase,
                                            b.svals[b.sorted], g.tile_rect, plan.strips_per_row, b.counts, nullptr,
                                            nullptr);
  }
  e = cudaGetLastError();
  if (e != cudaSuccess) return e;
  e = launch_scan_counts(b.counts, b.offs, (long long)(b.max_chunks * STRIP_TILES),
                         reinterpret_cast<const unsigned long long*>(b.counters + 4), b.cscan_status, b.cscan_counters, s);
  if (e != cudaSuccess) return e;
  {
    ProfScope ps(K_EMIT, s);
    expand_kernel<1><<<blocks, 256, 0, s>>>(b.counters, b.chunk_strip, b.strip_range, b.strip_nch, b.strip_cbase,
                                            b.svals[b.sorted], g.tile_rect, plan.strips_per_row, b.counts, b.offs,
                                            b.point_list);
  }
  e = cudaGetLastError();
  if (e != cudaSuccess) return e;
  const int tiles = grid_x * grid_y;
  ProfScope ps(K_TILE_RANGES, s);
  tile_ranges_kernel<<<(tiles + 255) / 256, 256, 0, s>>>(tiles, grid_x, plan.strips_per_row, (long long)R, b.strip_nch,
                                                         b.strip_cbase, b.counts, b.offs, ranges, b.seg_tile);
  return cudaGetLastError();
}

cudaError_t launch_fill_keys64(const uint2* ranges, int tiles, const uint32_t* point_list, const float* depth,
                               unsigned long long* keys64, cudaStream_t s) {
  if (tiles <= 0) return cudaSuccess;
  fill_keys64_kernel<<<tiles, 256, 0, s>>>(ranges, point_list, reinterpret_cast<const uint32_t*>(depth), keys64);
  return cudaGetLastError();
}

}  // namespace h3dgs
