// Shared declarations for the sm_100a kernels behind include/h3dgs_b200.h.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>
#include "../../include/h3dgs_b200.h"

namespace h3dgs {

constexpr int TILE_X = 16;            // HR/cuda_rasterizer/config.h:15-16 (BLOCK_X/BLOCK_Y) -- fixed by parity
constexpr int TILE_Y = 16;
constexpr int TILE_PIX = TILE_X * TILE_Y;

// Per-slot record written by preprocess and staged into shared memory by the blend kernels.
//   q0 = (pixel x, pixel y, 2 ln(255 opacity) with margin: bound of the conic form for alpha >= 1/255, 0)
//   q1 = (conic a, conic b, conic c, opacity * AA-compensation)
//   q2 = (r, g, b, 1/depth)
//   q3 = (t, 1/kids, depth, h_convolution_scaling)
struct __align__(16) Rec {
  float4 q0, q1, q2, q3;
};
static_assert(sizeof(Rec) == 64, "Rec must be 64 bytes");

// Per-slot gradient accumulator filled by blend_bwd (vector red.global.add), consumed by preprocess_bwd.
//   v[0..1] = dL/dmean2D (pixel units already scaled by 0.5*W, 0.5*H)
//   v[2..4] = dL/dconic (xx, xy, yy)      v[5] = dL/d(opacity*h)
//   v[6..8] = dL/drgb                      v[9] = dL/d(1/depth)     v[10..11] = padding
struct __align__(16) GradRec {
  float v[12];
};
static_assert(sizeof(GradRec) == 48, "GradRec must be 48 bytes");

inline __host__ __device__ size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

template <typename T>
inline void carve(char*& p, T*& out, size_t count) {
  p = reinterpret_cast<char*>(align_up(reinterpret_cast<size_t>(p), 256));
  out = reinterpret_cast<T*>(p);
  p += count * sizeof(T);
}

// ---- scan geometry (decoupled look-back) ----
constexpr int SCAN_THREADS = 256;
constexpr int SCAN_ITEMS = 16;
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;

// ---- two-level instance sort geometry (radix_sort.cu / binning.cu) ----
constexpr int SORT_RADIX_BITS = 8;
constexpr int SORT_RADIX = 1 << SORT_RADIX_BITS;
constexpr int SORT_THREADS = 512;
constexpr int SORT_ITEMS = 12;
constexpr int SORT_TILE = SORT_THREADS * SORT_ITEMS;
constexpr int SORT_MAX_PASSES = 4;
constexpr int SORT_DEPTH_PASSES = 4;   // 32 depth bits, 8 per pass
constexpr int EMIT_CHUNK = 2048;       // strip records written per CTA of the emission kernel
constexpr int STRIP_TILES = 64;        // tiles of one tile row that a warp expands together (two per lane)
constexpr int EXPAND_CHUNK = 512;      // strip records per warp of the expansion kernels
constexpr long long MAX_INSTANCES = 1ll << 30;   // a render with more tile instances is rejected

// Everything whose size is known before the render's instance count R is: per-slot state and the depth sort of
// the visible slots (V <= P, sized by P).  The kernels of this part take V from device memory, so the whole
// first stage of the forward can be enqueued without a host round trip.
struct GeomWs {
  Rec* rec;
  float* depth;
  float* cov3D;
  ushort4* tile_rect;
  uint32_t* tiles_touched;
  uint32_t* offsets;                // inclusive rank of the slot among the visible slots
  uint8_t* clamped;
  unsigned long long* scan_status;  // one per scan tile
  uint32_t* counters;               // [0] dynamic tile id, [2..3] V (u64), [4..5] R (u64)
  // depth sort of the visible slots
  uint32_t* dkeys[2];               // depth bits of the visible slots, double buffered
  uint32_t* dvals[2];               // visible slots; dvals[0] ends up in (depth, slot) order
  uint32_t* soffs;                  // inclusive scan of the strip-record counts in that order
  uint32_t* chunk_first;            // [MAX_INSTANCES / EMIT_CHUNK + 2] first position (in that order) reaching into each chunk
  unsigned long long* gscan_status; // look-back words of the gathered scan
  uint32_t* gscan_counters;
  char* zero_base;                  // depth_hist .. dcounters are contiguous: one memset
  size_t zero_bytes;
  uint32_t* depth_hist;             // [SORT_DEPTH_PASSES][SORT_RADIX]
  uint32_t* depth_status;           // [SORT_DEPTH_PASSES][tiles(P)][SORT_RADIX] look-back words
  uint32_t* dcounters;              // [0..3] dynamic tile ids of the depth passes
  size_t sort_tiles;                // ceil(P / SORT_TILE): stride of depth_status per pass
  size_t bytes;
  static GeomWs make(void* base, int P) {
    GeomWs w;
    char* p = reinterpret_cast<char*>(base);
    char* p0 = p;
    size_t n = (size_t)(P > 0 ? P : 1);
    carve(p, w.rec, n);
    carve(p, w.depth, n);
    carve(p, w.cov3D, n * 6);
    carve(p, w.tile_rect, n);
    carve(p, w.tiles_touched, n);
    carve(p, w.offsets, n);
    carve(p, w.clamped, n);
    carve(p, w.scan_status, (n + SCAN_TILE - 1) / SCAN_TILE + 1);
    carve(p, w.counters, 8);
    carve(p, w.dkeys[0], n);
    carve(p, w.dkeys[1], n);
    carve(p, w.dvals[0], n);
    carve(p, w.dvals[1], n);
    carve(p, w.soffs, n);
    carve(p, w.chunk_first, (size_t)(MAX_INSTANCES / EMIT_CHUNK) + 2);
    carve(p, w.gscan_status, (n + SCAN_TILE - 1) / SCAN_TILE + 1);
    carve(p, w.gscan_counters, 8);
    w.sort_tiles = (n + SORT_TILE - 1) / SORT_TILE;
    carve(p, w.depth_hist, (size_t)SORT_DEPTH_PASSES * SORT_RADIX);
    w.zero_base = reinterpret_cast<char*>(w.depth_hist);
    carve(p, w.depth_status, (size_t)SORT_DEPTH_PASSES * w.sort_tiles * SORT_RADIX);
    carve(p, w.dcounters, 8);
    w.zero_bytes = (size_t)(p - w.zero_base);
    w.bytes = (size_t)(p - p0) + 256;
    return w;
  }
};

struct ImageWs {
  uint2* ranges;      // per tile
  float* final_T;     // per pixel
  uint32_t* n_contrib;
  float4* accum;      // (C.r, C.g, C.b, D) without background, per pixel
  uint32_t* order_fwd;  // per tile: launch order of the forward blend, longest list first
  uint32_t* order_bwd;  // per tile: launch order of the backward blend, deepest last contributor first
  uint32_t* tile_last;  // per tile: max over its pixels of n_contrib (how far the backward walks)
  size_t bytes;
  static ImageWs make(void* base, int W, int H) {
    ImageWs w;
    char* p = reinterpret_cast<char*>(base);
    char* p0 = p;
    size_t npix = (size_t)W * H;
    size_t tiles = (size_t)((W + TILE_X - 1) / TILE_X) * ((H + TILE_Y - 1) / TILE_Y);
    carve(p, w.ranges, tiles);
    carve(p, w.final_T, npix);
    carve(p, w.n_contrib, npix);
    carve(p, w.accum, npix);
    carve(p, w.order_fwd, tiles);
    carve(p, w.order_bwd, tiles);
    carve(p, w.tile_last, tiles);
    w.bytes = (size_t)(p - p0) + 256;
    return w;
  }
};

// Backward blend work items: a tile list is cut every BLEND_SEG entries (tile-relative); the forward blend
// leaves a per-pixel snapshot (T, C, D in front of the cut) at every cut it crosses.  Cuts of different tiles
// are more than BLEND_SEG list positions apart, so floor(global position / BLEND_SEG) numbers them uniquely.
constexpr int BLEND_SEG = 1024;
constexpr int SNAP_PLANES = 5;         // T, C.r, C.g, C.b, D -- one float per pixel of the tile each
constexpr int SNAP_FLOATS = SNAP_PLANES * TILE_X * TILE_Y;

struct SortShifts {
  int shift[SORT_MAX_PASSES];
  uint32_t mask[SORT_MAX_PASSES];
};

struct SortPlan {
  int strips_per_row;              // ceil(grid_x / STRIP_TILES)
  int strips;                      // strips_per_row * grid_y
  int passes;                      // ceil(bits of a strip id / 8)
  int shift[SORT_MAX_PASSES];
  int bits[SORT_MAX_PASSES];
  int key_bytes;                   // 2 (strip ids fit 16 bits) or 4
};

// diagnostics (h3dgs_debug_set key 2, api.cu): bit 0 = u32 strip keys even when 16 bits would do, bit 1 = at least
// 16 key bits, i.e. two strip passes -- the paths an image below 65 536 strips never takes
int strip_key_debug_mode();

inline SortPlan make_sort_plan(int W, int H) {
  const int grid_x = (W + TILE_X - 1) / TILE_X, grid_y = (H + TILE_Y - 1) / TILE_Y;
  SortPlan pl;
  pl.strips_per_row = (grid_x + STRIP_TILES - 1) / STRIP_TILES;
  pl.strips = pl.strips_per_row * grid_y;
  int total = 1;
  while (total < 32 && (1ll << total) < (long long)pl.strips) ++total;
  const int dbg = strip_key_debug_mode();
  if ((dbg & 2) && total < 16) total = 16;
  pl.passes = (total + SORT_RADIX_BITS - 1) / SORT_RADIX_BITS;
  const int base = total / pl.passes, extra = total % pl.passes;
  int sft = 0;
  for (int i = 0; i < SORT_MAX_PASSES; ++i) {
    pl.bits[i] = i < pl.passes ? base + (i < extra ? 1 : 0) : 0;
    pl.shift[i] = sft;
    sft += pl.bits[i];
  }
  pl.key_bytes = (total <= 16 && !(dbg & 1)) ? 2 : 4;
  return pl;
}

// Binning workspace: everything sized by the instance count R (allocated once the host has read R).  The number
// of strip records (one per visible slot, tile row and 64-tile strip its rectangle reaches into) is only known on
// the device; it never exceeds R, which sizes their buffers.
struct BinWs {
  void* skeys[2];               // strip ids of the emitted records (u16 or u32), double buffered
  uint32_t* svals[2];           // slot of each record, double buffered
  uint32_t* point_list;         // [R] slots of all instances, by tile, (depth, slot) order inside a tile
  unsigned long long* keys64;   // [R] point_list_keys: (tile << 32) | depth bits -- filled for h3dgs_debug_view only
  float* snap;                  // [R / BLEND_SEG + 2][SNAP_FLOATS] forward snapshots at the list cuts
  uint32_t* seg_tile;           // [R / BLEND_SEG + 2] tile whose list holds instance x * BLEND_SEG
  uint32_t* chunk_strip;        // [max_chunks] strip of each expansion chunk
  uint32_t* counts;             // [STRIP_TILES * max_chunks] instances per (strip, tile of the strip, chunk), ragged
  uint32_t* offs;               // inclusive scan of counts
  unsigned long long* cscan_status;
  uint32_t* cscan_counters;
  char* zero_base;              // strip_hist .. strip_range are contiguous: one memset
  size_t zero_bytes;
  uint32_t* strip_hist;         // [SORT_MAX_PASSES][SORT_RADIX]
  uint32_t* strip_status;       // [passes][tiles(R)][SORT_RADIX]
  uint32_t* counters;           // [0..3]: dynamic tile ids of the strip passes; [4..5] u64 scan length, [6] chunks
  uint2* strip_range;           // [strips] records of each strip in the sorted order
  uint32_t* strip_nch;          // [strips] expansion chunks of each strip
  uint32_t* strip_cbase;        // [strips] first chunk of each strip
  int sorted;                   // index (0/1) of the skeys / svals buffers holding the sorted records
  size_t max_chunks;
  size_t bytes;
  static BinWs make(void* base, int64_t R, int W, int H) {
    BinWs w;
    SortPlan pl = make_sort_plan(W, H);
    char* p = reinterpret_cast<char*>(base);
    char* p0 = p;
    const size_t n = (size_t)(R > 0 ? R : 1);
    const size_t tiles_r = (n + SORT_TILE - 1) / SORT_TILE;
    w.max_chunks = n / EXPAND_CHUNK + (size_t)pl.strips + 1;
    char* tk;
    carve(p, tk, n * pl.key_bytes); w.skeys[0] = tk;
    carve(p, tk, n * pl.key_bytes); w.skeys[1] = tk;
    carve(p, w.svals[0], n);
    carve(p, w.svals[1], n);
    carve(p, w.point_list, n);
    carve(p, w.keys64, n);
    carve(p, w.snap, (n / BLEND_SEG + 2) * (size_t)SNAP_FLOATS);
    carve(p, w.seg_tile, n / BLEND_SEG + 2);
    carve(p, w.chunk_strip, w.max_chunks);
    carve(p, w.counts, w.max_chunks * STRIP_TILES);
    carve(p, w.offs, w.max_chunks * STRIP_TILES);
    carve(p, w.cscan_status, (w.max_chunks * STRIP_TILES + SCAN_TILE - 1) / SCAN_TILE + 1);
    carve(p, w.cscan_counters, 8);
    w.sorted = pl.passes & 1;
    carve(p, w.strip_hist, (size_t)SORT_MAX_PASSES * SORT_RADIX);
    w.zero_base = reinterpret_cast<char*>(w.strip_hist);
    carve(p, w.strip_status, (size_t)pl.passes * tiles_r * SORT_RADIX);
    carve(p, w.counters, 8);
    carve(p, w.strip_range, (size_t)pl.strips);
    w.zero_bytes = (size_t)(p - w.zero_base);
    carve(p, w.strip_nch, (size_t)pl.strips);
    carve(p, w.strip_cbase, (size_t)pl.strips);
    w.bytes = (size_t)(p - p0) + 256;
    return w;
  }
};

struct BwdWs {
  GradRec* grad;
  uint32_t* vis_list;   // visible slots in slot order (compacted for the per-slot backward kernel)
  size_t bytes;
  static BwdWs make(void* base, int P) {
    BwdWs w;
    char* p = reinterpret_cast<char*>(base);
    char* p0 = p;
    carve(p, w.grad, (size_t)(P > 0 ? P : 1));
    carve(p, w.vis_list, (size_t)(P > 0 ? P : 1));
    w.bytes = (size_t)(p - p0) + 256;
    return w;
  }
};

// error plumbing (api.cu)
int fail(int code, const char* what);
int cuda_fail(cudaError_t e, const char* what);
#define H3_CUDA(expr)                                         \
  do {                                                        \
    cudaError_t e__ = (expr);                                 \
    if (e__ != cudaSuccess) return ::h3dgs::cuda_fail(e__, #expr); \
  } while (0)

// ---- instrumentation (api.cu) ----
enum KernelId {
  K_PREPROCESS_FWD = 0, K_SCAN, K_EMIT, K_SORT_HIST, K_SORT_SCAN, K_SORT_PASS, K_TILE_RANGES, K_BLEND_FWD,
  K_BLEND_BWD, K_PREPROCESS_BWD, K_CUT_MARK, K_CUT_PUT, K_CUT_WEIGHTS, K_ADAM, K_MARK_VISIBLE, K_MEMSET,
  K_LOSS, K_HIER_TREE, K_HIER_MERGE, K_DENSIFY, K_COUNT
};
static_assert(K_COUNT == H3DGS_PROF_KERNELS, "kernel id table out of sync with the header");
void prof_begin(int id, cudaStream_t s);
void prof_end(int id, cudaStream_t s);
struct ProfScope {
  int id; cudaStream_t s;
  ProfScope(int id_, cudaStream_t s_) : id(id_), s(s_) { prof_begin(id, s); }
  ~ProfScope() { prof_end(id, s); }
};

// ---- kernel launchers (one per .cu) ----
struct PreArgs {
  int P, rows, D, M, skyboxnum;
  int W, H, grid_x, grid_y;
  float scale_modifier, tan_fovx, tan_fovy, focal_x, focal_y;
  const float* means; const float* shs; const float* colors; const float* opac;
  const float* scales; const float* rots; const float* cov3D_pre;
  const float* shs_rest;                 // raw store: shs = [rows,1,3], shs_rest = [rows,M-1,3]
  int act_scale, act_rot, act_opac;      // H3DGS_ACT_*
  int norm_variant;                      // summation order of ||q||^2, see gsmath.cuh act_rotation
  const int* idx; const int* pidx; const float* ts; const int* kids;
  const float* view; const float* proj; const float* campos;
  int* radii;
  GeomWs ws;
};
cudaError_t launch_preprocess_fwd(const PreArgs& a, cudaStream_t s);
cudaError_t launch_mark_visible(int P, const float* means, const float* view, const float* proj,
                                uint8_t* present, cudaStream_t s);

// inclusive scan of u32 -> u32, total (u64) written to total_out[0..1]
cudaError_t launch_scan_inclusive(const uint32_t* in, uint32_t* out, int n, unsigned long long* status,
                                  uint32_t* counters, cudaStream_t s);

// rank (inclusive) of every visible slot among the visible ones; totals: counters[2..3] = V, counters[4..5] = R
cudaError_t launch_scan_visible(const uint32_t* tiles_touched, uint32_t* rank, int n,
                                unsigned long long* status, uint32_t* counters, cudaStream_t s);
// inclusive scan of the strip-record counts of slots order[i], i < *n_dev (a device u64; at most n_cap); the total
// is left at counters[2..3]
cudaError_t launch_scan_strips(const ushort4* tile_rect, const uint32_t* order, uint32_t* out, int n_cap,
                               const unsigned long long* n_dev, uint32_t* chunk_first, unsigned long long* status,
                               uint32_t* counters, cudaStream_t s);
// inclusive scan of in[0 .. *n_dev) (at most n_cap)
cudaError_t launch_scan_counts(const uint32_t* in, uint32_t* out, long long n_cap, const unsigned long long* n_dev,
                               unsigned long long* status, uint32_t* counters, cudaStream_t s);

// binning.cu: visible slots -> (depth bits, slot) pairs + depth digit histograms; ordered emission of strip
// records; expansion of the sorted records into point_list, ranges and seg_tile
cudaError_t launch_compact_visible(int P, GeomWs& g, cudaStream_t s);
cudaError_t launch_emit_strips(int64_t R, int64_t V, const GeomWs& g, BinWs& b, const SortPlan& plan, cudaStream_t s);
cudaError_t launch_expand(int64_t R, const GeomWs& g, BinWs& b, const SortPlan& plan, int grid_x, int grid_y,
                          uint2* ranges, cudaStream_t s);
cudaError_t launch_fill_keys64(const uint2* ranges, int tiles, const uint32_t* point_list, const float* depth,
                               unsigned long long* keys64, cudaStream_t s);
// order[] = tile ids sorted by descending work (ranges given: list length; else lengths[] itself), 64 log buckets
cudaError_t launch_tile_order(const uint2* ranges, const uint32_t* lengths, int tiles, uint32_t* order,
                              cudaStream_t s);
// radix_sort.cu
// sorts the *v_dev (device u64, at most P) compacted (depth bits, slot) pairs of g.dkeys[0] / g.dvals[0]
cudaError_t launch_depth_sort(GeomWs& g, int P, const unsigned long long* v_dev, cudaStream_t s);
cudaError_t launch_strip_sort(BinWs& b, int64_t n_cap, const unsigned long long* n_dev, const SortPlan& plan,
                              cudaStream_t s);
// stable sort of n (u32 key, u32 value) pairs, result in keys[0] / vals[0]; ws = sort_pairs_ws_words(n) u32 words
size_t sort_pairs_ws_words(long long n);
cudaError_t launch_sort_pairs(uint32_t* const keys[2], uint32_t* const vals[2], long long n, uint32_t* ws,
                              cudaStream_t s);

struct BlendArgs {
  int W, H, grid_x, grid_y, P, skyboxnum;
  int interp;       // ts && kids
  const uint2* ranges; const uint32_t* point_list; const Rec* rec;
  const float* bg;
  float* final_T; uint32_t* n_contrib; float4* accum;
  float* out_color; float* out_invdepth;   // fwd
  const float* dL_dcolor; const float* dL_dinvdepth; GradRec* grad;  // bwd
  const uint32_t* order;   // CTA i processes tile order[i] (heavy tiles first); NULL = identity
  uint32_t* tile_last;     // fwd: per-tile max of n_contrib (pre-zeroed)
  float* snap;             // fwd writes, bwd reads: per-pixel state at the list cuts (BinWs::snap)
  const uint32_t* seg_tile;  // bwd: tile whose list holds instance x * BLEND_SEG (finds the tile of a cut)
  int nseg;                // bwd: work items that start at a cut, (R - 1) / BLEND_SEG; they lead the grid
  unsigned long long* stats;  // instrumented pass (h3dgs_blend_stats): pair-category counters, else NULL
};
cudaError_t launch_blend_fwd(const BlendArgs& a, cudaStream_t s);
cudaError_t launch_blend_bwd(const BlendArgs& a, cudaStream_t s);
cudaError_t launch_blend_bwd2(const BlendArgs& a, cudaStream_t s);   // two pixels per lane (blend_bwd2.cu)

struct PreBwdArgs {
  PreArgs f;
  const GradRec* grad;
  int has_invdepth;
  H3dgsGrads out;
  uint32_t* vis_list;
};
cudaError_t launch_preprocess_bwd(const PreBwdArgs& a, cudaStream_t s);

}  // namespace h3dgs
