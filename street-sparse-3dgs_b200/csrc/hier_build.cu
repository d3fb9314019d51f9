// GPU hierarchy build: the tree stages of the reference's offline GaussianHierarchyCreator
//   KD-tree      GH/PointbasedKdTreeGenerator.cpp:16-73   recursive median split along the longest side of the bounds
//   merge        GH/ClusterMerger.cpp:23-165              bottom-up moment-matched merge of the two children
//   align        GH/rotation_aligner.cpp:14-114           top-down signed-permutation match of child to parent axes
//   flatten      GH/writer.cpp:18-115                     pre-order node table / Gaussian rows / boxes
// as level-synchronous kernels.  The reference runs all four as single-threaded recursions on the host.
//
// Design
//   * The SHAPE of the tree depends on N alone (a node of m Gaussians splits into m/2 and m - m/2), so every node's
//     final id, parent, pre-order row, height and child ids are integers computed top-down without looking at the
//     data: the children of the k-th internal node in pre-order are nodes 2k+1 and 2k+2 (writer.cpp populateRec),
//     k(left) = k + 1, k(right) = k + m/2, row(left) = row + 1, row(right) = row + 2 (m/2).  Kernels write straight
//     into the final node table; nothing is renumbered afterwards.
//   * WHICH Gaussian lands in which leaf: three index lists, each sorted once by one coordinate (onesweep radix sort,
//     ties by Gaussian index), are kept segmented by node.  Per level: segmented min/max of the members' bounds
//     (warp-segmented shuffles, one atomic per warp segment) -> axis per node -> the first m/2 entries of the node's
//     segment in the list of that axis go left -> all three lists are stably partitioned inside their segments
//     (one 3-lane prefix scan of the "goes left" flags).  No per-node sort, no recursion: ceil(log2 N) + 1 levels of
//     bandwidth-bound passes over N indices.
//   * Merge runs level by level from the deepest level up (one thread per node for the 3x3 moments and the
//     eigen-decomposition, twelve threads per node for the 48 SH coefficients), alignment level by level from the
//     root down.  Products and sums that the x86 reference rounds separately are written with __fmul_rn / __fadd_rn
//     so that ptxas cannot contract them.
//   * The symmetric 3x3 eigenproblem is solved by cyclic Jacobi rotations in double precision (the reference uses
//     Eigen's float QL iteration): eigenvalues are then correctly rounded floats, eigenvector signs are free in both
//     (tests/hier_build_util.py compares ellipsoids, and axes after alignment).
#include <math.h>

#include <vector>

#include "common.cuh"

namespace h3dgs {

namespace {

#define H3_CUDA_RET(expr) do { cudaError_t e__ = (expr); if (e__ != cudaSuccess) return e__; } while (0)

constexpr int HB_THREADS = 256;
constexpr int HB_MAX_LEVELS = 40;

__device__ __forceinline__ uint32_t f2ord(float f) {   // order-preserving float -> u32
  const uint32_t b = __float_as_uint(f);
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float ord2f(uint32_t u) {
  return __uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u);
}

// per-node build state next to the final node table
struct HbNodes {
  int* nodes;        // [nn,7] final table (depth, parent, start, count_leafs, count_merged, start_children, count_children)
  int* begin;        // [nn] first position of the node's segment in the three lists
  int* size;         // [nn] Gaussians below the node
  int* krank;        // [nn] internal nodes before this one in pre-order
  uint32_t* box;     // [nn,6] order-encoded min xyz, max xyz
  uint8_t* axis;     // [nn]
  int8_t* level;     // [nn] distance from the root
};

struct HbWs {
  uint32_t* list[2][3];   // index lists per axis, double buffered
  uint32_t* keys[2];      // sort keys (initial sorts only)
  uint32_t* node_of[2];   // node of every Gaussian, double buffered across levels
  uint8_t* side;          // 1 = goes to the second child at this level
  uint32_t* lscan;        // [3,N] exclusive scan of the "goes left" flags per list
  uint32_t* bsum;         // [3, scan blocks] block totals -> block offsets
  uint32_t* sort_ws;
  float* lo;              // [N,3] position - 3 max scale
  float* hi;              // [N,3]
  HbNodes nd;
  int* level_nodes;       // [nn] node ids grouped by level
  int* level_fill;        // [HB_MAX_LEVELS] append cursors
  float* scale;           // [nn,3] scale of every row (leaf: input; merged: sqrt |eigenvalue|), log taken at the end
  float* weight;          // [nn,2] normalised merge weights of the node's two children
  int* row_of_gaussian;   // [N] pre-order row of every input Gaussian
  float* cov;             // [nn,6] when the caller does not ask for the covariances
  int* errflag;
  size_t bytes;
  static HbWs make(void* base, long long N) {
    HbWs w;
    char* p = reinterpret_cast<char*>(base);
    char* p0 = p;
    const size_t n = (size_t)(N > 0 ? N : 1), nn = 2 * n - 1;
    const size_t sblocks = (n + HB_THREADS * 4 - 1) / (HB_THREADS * 4);
    for (int b = 0; b < 2; ++b)
      for (int a = 0; a < 3; ++a) carve(p, w.list[b][a], n);
    carve(p, w.keys[0], n); carve(p, w.keys[1], n);
    carve(p, w.node_of[0], n); carve(p, w.node_of[1], n);
    carve(p, w.side, n);
    carve(p, w.lscan, 3 * n);
    carve(p, w.bsum, 3 * sblocks + 4);
    carve(p, w.sort_ws, sort_pairs_ws_words((long long)n));
    carve(p, w.lo, 3 * n); carve(p, w.hi, 3 * n);
    w.nd.nodes = nullptr;      // the node table itself is an output, set by the driver
    carve(p, w.nd.begin, nn); carve(p, w.nd.size, nn); carve(p, w.nd.krank, nn);
    carve(p, w.nd.box, 6 * nn); carve(p, w.nd.axis, nn); carve(p, w.nd.level, nn);
    carve(p, w.level_nodes, nn); carve(p, w.level_fill, HB_MAX_LEVELS);
    carve(p, w.scale, 3 * nn); carve(p, w.weight, 2 * nn);
    carve(p, w.row_of_gaussian, n); carve(p, w.cov, 6 * nn); carve(p, w.errflag, 4);
    w.bytes = (size_t)(p - p0) + 256;
    return w;
  }
};

// ---------------------------------------------------------------------------------------------------------------
// KD-tree
// ---------------------------------------------------------------------------------------------------------------

// bounds of every Gaussian (PointbasedKdTreeGenerator.cpp:26-31) and the three sort keys
__global__ void __launch_bounds__(HB_THREADS) hb_init_kernel(int N, const float* __restrict__ pos,
                                                            const float* __restrict__ scale, float* __restrict__ lo,
                                                            float* __restrict__ hi, uint32_t* __restrict__ node_of) {
  const int g = blockIdx.x * HB_THREADS + threadIdx.x;
  if (g >= N) return;
  const float r = __fmul_rn(3.0f, fmaxf(fmaxf(scale[3 * g], scale[3 * g + 1]), scale[3 * g + 2]));
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    const float x = pos[3 * g + a];
    lo[3 * g + a] = __fsub_rn(x, r);
    hi[3 * g + a] = __fadd_rn(x, r);
  }
  node_of[g] = 0;
}

__global__ void __launch_bounds__(HB_THREADS) hb_keys_kernel(int N, const float* __restrict__ pos, int a,
                                                            uint32_t* __restrict__ keys, uint32_t* __restrict__ vals) {
  const int g = blockIdx.x * HB_THREADS + threadIdx.x;
  if (g >= N) return;
  float x = pos[3 * g + a];
  if (x == 0.0f) x = 0.0f;   // -0 and +0 compare equal in the reference's comparator: one key for both
  keys[g] = f2ord(x);
  vals[g] = (uint32_t)g;
}

__global__ void hb_root_kernel(int N, HbNodes nd, int* __restrict__ level_nodes, int* __restrict__ level_fill) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  nd.begin[0] = 0; nd.size[0] = N; nd.krank[0] = 0; nd.level[0] = 0;
  int* t = nd.nodes;
  t[1] = -1; t[2] = 0;
  level_nodes[0] = 0;
  for (int i = 0; i < HB_MAX_LEVELS; ++i) level_fill[i] = 0;
}

__global__ void __launch_bounds__(HB_THREADS) hb_box_reset_kernel(long long n6, uint32_t* __restrict__ box) {
  const long long i = (long long)blockIdx.x * HB_THREADS + threadIdx.x;
  if (i >= n6) return;
  box[i] = (i % 6) < 3 ? 0xffffffffu : 0u;
}

// Segmented min / max of the member bounds over the nodes of the current level.  The nodes' members are contiguous in
// list order, so lanes of a warp that share a node form runs: a segmented shuffle reduction leaves each run's result
// in its first lane, which issues the six atomics (a CTA whose 256 entries all belong to one node reduces through
// shared memory first: 6 atomics per CTA at the top levels).
__global__ void __launch_bounds__(HB_THREADS) hb_bounds_kernel(int N, const uint32_t* __restrict__ list,
                                                              const uint32_t* __restrict__ node_of,
                                                              const int8_t* __restrict__ node_level, int level,
                                                              const float* __restrict__ lo, const float* __restrict__ hi,
                                                              uint32_t* __restrict__ box) {
  __shared__ uint32_t s_first, s_last;
  __shared__ float s_red[HB_THREADS / 32][6];
  const int p = blockIdx.x * HB_THREADS + threadIdx.x;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const bool valid = p < N;
  uint32_t node = 0xffffffffu;
  float v[6] = {INFINITY, INFINITY, INFINITY, -INFINITY, -INFINITY, -INFINITY};
  if (valid) {
    const uint32_t g = list[p];
    node = node_of[g];
#pragma unroll
    for (int a = 0; a < 3; ++a) { v[a] = lo[3 * g + a]; v[3 + a] = hi[3 * g + a]; }
    if (node_level[node] != level) node = 0xfffffffeu;   // leaf finished at an earlier level: nothing to add
  }
  if (threadIdx.x == 0) s_first = node;
  if (p == min(N, (int)(blockIdx.x + 1) * HB_THREADS) - 1) s_last = node;
  __syncthreads();
  const bool uniform = s_first == s_last && s_first < 0xfffffffeu;   // segments are contiguous: first == last => all
  if (uniform) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
      for (int a = 0; a < 3; ++a) {
        v[a] = fminf(v[a], __shfl_down_sync(0xffffffffu, v[a], o));
        v[3 + a] = fmaxf(v[3 + a], __shfl_down_sync(0xffffffffu, v[3 + a], o));
      }
    }
    if (lane == 0)
#pragma unroll
      for (int a = 0; a < 6; ++a) s_red[warp][a] = v[a];
    __syncthreads();
    if (threadIdx.x < 6) {
      const int a = threadIdx.x;
      float r = s_red[0][a];
      for (int w = 1; w < HB_THREADS / 32; ++w) r = a < 3 ? fminf(r, s_red[w][a]) : fmaxf(r, s_red[w][a]);
      if (a < 3) atomicMin(box + (size_t)s_first * 6 + a, f2ord(r));
      else atomicMax(box + (size_t)s_first * 6 + a, f2ord(r));
    }
    return;
  }
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t on = __shfl_down_sync(0xffffffffu, node, o);
    const bool take = (lane + o < 32) && on == node;
#pragma unroll
    for (int a = 0; a < 3; ++a) {
      const float l = __shfl_down_sync(0xffffffffu, v[a], o);
      const float h = __shfl_down_sync(0xffffffffu, v[3 + a], o);
      if (take) { v[a] = fminf(v[a], l); v[3 + a] = fmaxf(v[3 + a], h); }
    }
  }
  const uint32_t prev = __shfl_up_sync(0xffffffffu, node, 1);
  const bool head = lane == 0 || prev != node;
  if (head && node < 0xfffffffeu) {
#pragma unroll
    for (int a = 0; a < 3; ++a) {
      atomicMin(box + (size_t)node * 6 + a, f2ord(v[a]));
      atomicMax(box + (size_t)node * 6 + 3 + a, f2ord(v[3 + a]));
    }
  }
}

// One thread per node of the level: split axis (PointbasedKdTreeGenerator.cpp:41-53: strict >, first axis wins
// ties, axis 0 when every extent is <= 0), node-table row, and the two children with their final ids.
__global__ void __launch_bounds__(HB_THREADS) hb_split_kernel(int count, const int* __restrict__ level_list, HbNodes nd,
                                                             int* __restrict__ next_list, int* __restrict__ next_fill,
                                                             int level) {
  const int i = blockIdx.x * HB_THREADS + threadIdx.x;
  if (i >= count) return;
  const int id = level_list[i];
  const int m = nd.size[id], k = nd.krank[id];
  int* row = nd.nodes + (size_t)id * 7;
  row[0] = m > 1 ? 32 - __clz(m - 1) : 0;   // height: the larger half decides, ceil(log2 m)
  row[3] = m == 1;
  row[4] = m > 1;
  row[5] = 2 * k + 1;                        // basenodes.size() when the node is visited (writer.cpp:57)
  row[6] = m > 1 ? 2 : 0;
  if (m <= 1) return;
  const uint32_t* b = nd.box + (size_t)id * 6;
  int axis = 0;
  float greatest = 0.0f;
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    const float d = __fsub_rn(ord2f(b[3 + a]), ord2f(b[a]));
    if (d > greatest) { greatest = d; axis = a; }
  }
  nd.axis[id] = (uint8_t)axis;
  const int half = m / 2, c0 = 2 * k + 1, c1 = c0 + 1;
  const int start = row[2], begin = nd.begin[id];
  nd.begin[c0] = begin;        nd.size[c0] = half;     nd.krank[c0] = k + 1;    nd.level[c0] = (int8_t)(level + 1);
  nd.begin[c1] = begin + half; nd.size[c1] = m - half; nd.krank[c1] = k + half; nd.level[c1] = (int8_t)(level + 1);
  int* r0 = nd.nodes + (size_t)c0 * 7;
  int* r1 = nd.nodes + (size_t)c1 * 7;
  r0[1] = id; r0[2] = start + 1;
  r1[1] = id; r1[2] = start + 2 * half;      // the first child's subtree holds 2 half - 1 rows
  const int slot = atomicAdd(next_fill, 2);
  next_list[slot] = c0;
  next_list[slot + 1] = c1;
}

// side[g] = 1 when g lies in the upper part of its node's segment in the list of the node's split axis.  Every
// Gaussian of a splitting node matches in exactly one of the three lists.
__global__ void __launch_bounds__(HB_THREADS) hb_side_kernel(int N, const uint32_t* __restrict__ l0,
                                                            const uint32_t* __restrict__ l1,
                                                            const uint32_t* __restrict__ l2,
                                                            const uint32_t* __restrict__ node_of, HbNodes nd, int level,
                                                            uint8_t* __restrict__ side) {
  const int p = blockIdx.x * HB_THREADS + threadIdx.x;
  if (p >= N) return;
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    const uint32_t g = (a == 0 ? l0 : a == 1 ? l1 : l2)[p];
    const uint32_t node = node_of[g];
    if (nd.level[node] != level) { if (a == 0) side[g] = 1; continue; }   // finished leaf: stays (flag "not left")
    const int m = nd.size[node];
    if (m <= 1) { if (a == 0) side[g] = 1; continue; }
    if (nd.axis[node] == a) side[g] = (uint8_t)((p - nd.begin[node]) >= m / 2);
  }
}

// 3-lane exclusive scan of the "goes left" flags, three kernels: per-CTA totals, scan of the totals, final values.
constexpr int HB_SCAN_ITEMS = 4;
constexpr int HB_SCAN_TILE = HB_THREADS * HB_SCAN_ITEMS;

__device__ __forceinline__ uint32_t hb_left_flags(const uint32_t* __restrict__ l0, const uint32_t* __restrict__ l1,
                                                  const uint32_t* __restrict__ l2, const uint8_t* __restrict__ side,
                                                  int p) {   // packed counts: 10 bits per list
  return (uint32_t)(side[l0[p]] == 0) | ((uint32_t)(side[l1[p]] == 0) << 10) | ((uint32_t)(side[l2[p]] == 0) << 20);
}

__global__ void __launch_bounds__(HB_THREADS) hb_scan_totals_kernel(int N, const uint32_t* __restrict__ l0,
                                                                   const uint32_t* __restrict__ l1,
                                                                   const uint32_t* __restrict__ l2,
                                                                   const uint8_t* __restrict__ side, int blocks,
                                                                   uint32_t* __restrict__ bsum) {
  __shared__ uint32_t s_w[HB_THREADS / 32][3];
  const int base = blockIdx.x * HB_SCAN_TILE + threadIdx.x * HB_SCAN_ITEMS;
  uint32_t c = 0;
#pragma unroll
  for (int i = 0; i < HB_SCAN_ITEMS; ++i)
    if (base + i < N) c += hb_left_flags(l0, l1, l2, side, base + i);
  uint32_t v[3] = {c & 1023u, (c >> 10) & 1023u, c >> 20};
#pragma unroll
  for (int a = 0; a < 3; ++a) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v[a] += __shfl_down_sync(0xffffffffu, v[a], o);
    if ((threadIdx.x & 31) == 0) s_w[threadIdx.x >> 5][a] = v[a];
  }
  __syncthreads();
  if (threadIdx.x < 3) {
    uint32_t t = 0;
    for (int w = 0; w < HB_THREADS / 32; ++w) t += s_w[w][threadIdx.x];
    bsum[(size_t)threadIdx.x * blocks + blockIdx.x] = t;
  }
}

// exclusive scan of each list's CTA totals, in place (one CTA per list; serial over chunks of 1024)
__global__ void __launch_bounds__(1024) hb_scan_blocks_kernel(int blocks, uint32_t* __restrict__ bsum) {
  __shared__ uint32_t s_w[32];
  __shared__ uint32_t s_carry;
  uint32_t* b = bsum + (size_t)blockIdx.x * blocks;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) s_carry = 0;
  __syncthreads();
  for (int base = 0; base < blocks; base += 1024) {
    const int i = base + threadIdx.x;
    const uint32_t c = i < blocks ? b[i] : 0u;
    uint32_t incl = c;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const uint32_t o = __shfl_up_sync(0xffffffffu, incl, d);
      if (lane >= d) incl += o;
    }
    if (lane == 31) s_w[warp] = incl;
    __syncthreads();
    uint32_t off = s_carry;
    for (int w = 0; w < warp; ++w) off += s_w[w];
    if (i < blocks) b[i] = off + incl - c;
    __syncthreads();
    if (threadIdx.x == 1023) s_carry = off + incl;
    __syncthreads();
  }
}

// Final exclusive scan value of every list position (the scatter needs the value at segment starts too).
__global__ void __launch_bounds__(HB_THREADS) hb_scan_final_kernel(
    int N, const uint32_t* __restrict__ l0, const uint32_t* __restrict__ l1, const uint32_t* __restrict__ l2,
    const uint8_t* __restrict__ side, int blocks, const uint32_t* __restrict__ bsum, uint32_t* __restrict__ lscan) {
  __shared__ uint32_t s_w[HB_THREADS / 32][3];
  const int base = blockIdx.x * HB_SCAN_TILE + threadIdx.x * HB_SCAN_ITEMS;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint32_t f[HB_SCAN_ITEMS];
  uint32_t c = 0;
#pragma unroll
  for (int i = 0; i < HB_SCAN_ITEMS; ++i) {
    f[i] = base + i < N ? hb_left_flags(l0, l1, l2, side, base + i) : 0u;
    c += f[i];
  }
  uint32_t excl[3];
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    const uint32_t mine = (c >> (10 * a)) & 1023u;
    uint32_t incl = mine;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const uint32_t o = __shfl_up_sync(0xffffffffu, incl, d);
      if (lane >= d) incl += o;
    }
    if (lane == 31) s_w[warp][a] = incl;
    excl[a] = incl - mine;
  }
  __syncthreads();
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    uint32_t off = bsum[(size_t)a * blocks + blockIdx.x];
    for (int w = 0; w < warp; ++w) off += s_w[w][a];
    excl[a] += off;
  }
#pragma unroll
  for (int i = 0; i < HB_SCAN_ITEMS; ++i) {
    if (base + i < N) {
#pragma unroll
      for (int a = 0; a < 3; ++a) lscan[(size_t)a * N + base + i] = excl[a];
    }
#pragma unroll
    for (int a = 0; a < 3; ++a) excl[a] += (f[i] >> (10 * a)) & 1u;
  }
}

__global__ void __launch_bounds__(HB_THREADS) hb_scatter_kernel(
    int N, const uint32_t* __restrict__ l0, const uint32_t* __restrict__ l1, const uint32_t* __restrict__ l2,
    uint32_t* __restrict__ o0, uint32_t* __restrict__ o1, uint32_t* __restrict__ o2, const uint8_t* __restrict__ side,
    const uint32_t* __restrict__ node_of, uint32_t* __restrict__ node_next, HbNodes nd, int level,
    const uint32_t* __restrict__ lscan) {
  const int p = blockIdx.x * HB_THREADS + threadIdx.x;
  if (p >= N) return;
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    const uint32_t* l = a == 0 ? l0 : a == 1 ? l1 : l2;
    uint32_t* o = a == 0 ? o0 : a == 1 ? o1 : o2;
    const uint32_t g = l[p];
    const uint32_t node = node_of[g];
    const int m = nd.size[node];
    if (nd.level[node] != level || m <= 1) {   // leaf: keeps its place and its node
      o[p] = g;
      if (a == 0) node_next[g] = node;
      continue;
    }
    const int b = nd.begin[node], half = m / 2;
    const uint32_t* sc = lscan + (size_t)a * N;
    const int left_before = (int)(sc[p] - sc[b]);
    const bool right = side[g] != 0;
    const int dst = right ? b + half + ((p - b) - left_before) : b + left_before;
    o[dst] = g;
    if (a == 0) node_next[g] = (uint32_t)(2 * nd.krank[node] + 1 + (right ? 1 : 0));
  }
}

// ---------------------------------------------------------------------------------------------------------------
// Gaussian rows: leaves, bottom-up merge, top-down alignment
// ---------------------------------------------------------------------------------------------------------------

struct HbRows {     // the output arrays, indexed by pre-order row
  float* pos;       // [nn,3]
  float* shs;       // [nn,48]
  float* opac;      // [nn]
  float* scale;     // [nn,3] plain scales while building (workspace), log_scales at the end (output)
  float* rot;       // [nn,4] (w, x, y, z)
  float* cov;       // [nn,6] moment-matched covariance (output for tests / workspace)
};

__device__ __forceinline__ float hb_surface(float s0, float s1, float s2) {   // ClusterMerger.cpp:16-21
  return __fadd_rn(__fadd_rn(__fmul_rn(s0, s1), __fmul_rn(s0, s2)), __fmul_rn(s1, s2));
}

// common.h:86-102 with every product and sum rounded on its own (x86 build of the reference: no FMA)
__device__ __forceinline__ void hb_matrix_from_quat(const float* q, float R[3][3]) {
  const float s = q[0], x = q[1], y = q[2], z = q[3];
  auto sq2 = [](float a, float b) { return __fmul_rn(2.0f, __fadd_rn(__fmul_rn(a, a), __fmul_rn(b, b))); };
  auto pm = [](float a, float b, float c, float d, bool plus) {
    const float u = __fmul_rn(a, b), v = __fmul_rn(c, d);
    return __fmul_rn(2.0f, plus ? __fadd_rn(u, v) : __fsub_rn(u, v));
  };
  R[0][0] = __fsub_rn(1.0f, sq2(y, z)); R[0][1] = pm(x, y, s, z, false); R[0][2] = pm(x, z, s, y, true);
  R[1][0] = pm(x, y, s, z, true);  R[1][1] = __fsub_rn(1.0f, sq2(x, z)); R[1][2] = pm(y, z, s, x, false);
  R[2][0] = pm(x, z, s, y, false); R[2][1] = pm(y, z, s, x, true);  R[2][2] = __fsub_rn(1.0f, sq2(x, y));
}

// common.h:104-133: T = R diag(scale), covariance = T T^t (upper triangle)
__device__ __forceinline__ void hb_covariance(const float* scale, const float* q, float cov[6]) {
  float R[3][3], T[3][3];
  hb_matrix_from_quat(q, R);
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) T[i][j] = __fmul_rn(R[i][j], scale[j]);
  auto dot3 = [&](int i, int j) {
    return __fadd_rn(__fadd_rn(__fmul_rn(T[i][0], T[j][0]), __fmul_rn(T[i][1], T[j][1])), __fmul_rn(T[i][2], T[j][2]));
  };
  cov[0] = dot3(0, 0); cov[1] = dot3(0, 1); cov[2] = dot3(0, 2);
  cov[3] = dot3(1, 1); cov[4] = dot3(1, 2); cov[5] = dot3(2, 2);
}

// Eigen::Quaternionf(Matrix3f) (Eigen/src/Geometry/Quaternion.h, quaternionbase_assign_impl) -> (w, x, y, z)
__device__ __forceinline__ void hb_quat_from_matrix(const float m[3][3], float q[4]) {
  float t = __fadd_rn(__fadd_rn(m[0][0], m[1][1]), m[2][2]);
  if (t > 0.0f) {
    t = sqrtf(__fadd_rn(t, 1.0f));
    q[0] = __fmul_rn(0.5f, t);
    t = __fdiv_rn(0.5f, t);
    q[1] = __fmul_rn(__fsub_rn(m[2][1], m[1][2]), t);
    q[2] = __fmul_rn(__fsub_rn(m[0][2], m[2][0]), t);
    q[3] = __fmul_rn(__fsub_rn(m[1][0], m[0][1]), t);
  } else {
    int i = 0;
    if (m[1][1] > m[0][0]) i = 1;
    if (m[2][2] > (i == 0 ? m[0][0] : m[1][1])) i = 2;
    const int j = (i + 1) % 3, k = (j + 1) % 3;
    auto at = [&](int r, int c) {
      float v = 0.0f;
#pragma unroll
      for (int rr = 0; rr < 3; ++rr)
#pragma unroll
        for (int cc = 0; cc < 3; ++cc) if (rr == r && cc == c) v = m[rr][cc];
      return v;
    };
    t = sqrtf(__fadd_rn(__fsub_rn(__fsub_rn(at(i, i), at(j, j)), at(k, k)), 1.0f));
    float out[3];
    const float qi = __fmul_rn(0.5f, t);
    t = __fdiv_rn(0.5f, t);
    q[0] = __fmul_rn(__fsub_rn(at(k, j), at(j, k)), t);
    const float qj = __fmul_rn(__fadd_rn(at(j, i), at(i, j)), t);
    const float qk = __fmul_rn(__fadd_rn(at(k, i), at(i, k)), t);
#pragma unroll
    for (int a = 0; a < 3; ++a) out[a] = a == i ? qi : a == j ? qj : qk;
    q[1] = out[0]; q[2] = out[1]; q[3] = out[2];
  }
}

// Symmetric 3x3 eigen-decomposition by cyclic Jacobi rotations in double precision.  Eigenvalues ascending (as
// Eigen's SelfAdjointEigenSolver orders them), eigenvectors in the columns of V.
__device__ void hb_eigh3(const float c[6], float lam[3], float V[3][3]) {
  double a[3][3] = {{(double)c[0], (double)c[1], (double)c[2]},
                    {(double)c[1], (double)c[3], (double)c[4]},
                    {(double)c[2], (double)c[4], (double)c[5]}};
  double v[3][3] = {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}};
  for (int sweep = 0; sweep < 16; ++sweep) {
    const double off = fabs(a[0][1]) + fabs(a[0][2]) + fabs(a[1][2]);
    const double diag = fabs(a[0][0]) + fabs(a[1][1]) + fabs(a[2][2]);
    if (off <= 1e-300 || off <= 1e-17 * diag) break;
#pragma unroll
    for (int pq = 0; pq < 3; ++pq) {
      const int p = pq == 2 ? 1 : 0, q = pq == 0 ? 1 : 2;
      const double apq = a[p][q];
      if (apq == 0.0) continue;
      const double theta = (a[q][q] - a[p][p]) / (2.0 * apq);
      const double t = (theta >= 0 ? 1.0 : -1.0) / (fabs(theta) + sqrt(theta * theta + 1.0));
      const double cs = 1.0 / sqrt(t * t + 1.0), sn = t * cs;
      const double app = a[p][p], aqq = a[q][q];
      a[p][p] = app - t * apq;
      a[q][q] = aqq + t * apq;
      a[p][q] = a[q][p] = 0.0;
      const int r = 3 - p - q;
      const double arp = a[r][p], arq = a[r][q];
      a[r][p] = a[p][r] = cs * arp - sn * arq;
      a[r][q] = a[q][r] = sn * arp + cs * arq;
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        const double vkp = v[k][p], vkq = v[k][q];
        v[k][p] = cs * vkp - sn * vkq;
        v[k][q] = sn * vkp + cs * vkq;
      }
    }
  }
  int o0 = 0, o1 = 1, o2 = 2;   // ascending order of the diagonal
  auto sw = [](int& x, int& y) { const int t = x; x = y; y = t; };
  if (a[o1][o1] < a[o0][o0]) sw(o0, o1);
  if (a[o2][o2] < a[o1][o1]) sw(o1, o2);
  if (a[o1][o1] < a[o0][o0]) sw(o0, o1);
  const int ord[3] = {o0, o1, o2};
#pragma unroll
  for (int j = 0; j < 3; ++j) {
    double l = 0, col[3] = {0, 0, 0};
#pragma unroll
    for (int s = 0; s < 3; ++s)
      if (ord[j] == s) { l = a[s][s]; col[0] = v[0][s]; col[1] = v[1][s]; col[2] = v[2][s]; }
    lam[j] = (float)l;
    V[0][j] = (float)col[0]; V[1][j] = (float)col[1]; V[2][j] = (float)col[2];
  }
}

// leaves: input Gaussian -> its pre-order row (writer.cpp:36-44), covariance from scale and rotation
__global__ void __launch_bounds__(HB_THREADS) hb_leaf_rows_kernel(int N, const uint32_t* __restrict__ node_of,
                                                                 const int* __restrict__ nodes,
                                                                 const float* __restrict__ pos,
                                                                 const float* __restrict__ opac,
                                                                 const float* __restrict__ scale,
                                                                 const float* __restrict__ rot, HbRows out,
                                                                 int* __restrict__ row_of_gaussian,
                                                                 int* __restrict__ leaf_order) {
  const int g = blockIdx.x * HB_THREADS + threadIdx.x;
  if (g >= N) return;
  const uint32_t node = node_of[g];
  const int row = nodes[(size_t)node * 7 + 2];
  row_of_gaussian[g] = row;
  if (leaf_order) leaf_order[row] = g;
  float s[3], q[4], c[6];
#pragma unroll
  for (int a = 0; a < 3; ++a) { out.pos[3 * (size_t)row + a] = pos[3 * (size_t)g + a]; s[a] = scale[3 * (size_t)g + a]; }
#pragma unroll
  for (int a = 0; a < 4; ++a) q[a] = rot[4 * (size_t)g + a];
  hb_covariance(s, q, c);
#pragma unroll
  for (int a = 0; a < 3; ++a) out.scale[3 * (size_t)row + a] = s[a];
#pragma unroll
  for (int a = 0; a < 4; ++a) out.rot[4 * (size_t)row + a] = q[a];
#pragma unroll
  for (int a = 0; a < 6; ++a) out.cov[6 * (size_t)row + a] = c[a];
  out.opac[row] = opac[g];
}

// 48 SH floats of every leaf: twelve threads per Gaussian, one float4 each
__global__ void __launch_bounds__(HB_THREADS) hb_leaf_shs_kernel(long long n12, const int* __restrict__ row_of_gaussian,
                                                                const float4* __restrict__ shs, float4* __restrict__ out) {
  const long long i = (long long)blockIdx.x * HB_THREADS + threadIdx.x;
  if (i >= n12) return;
  const long long g = i / 12;
  const int j = (int)(i - g * 12);
  out[(size_t)row_of_gaussian[g] * 12 + j] = __ldg(shs + i);
}

// ClusterMerger.cpp:44-139 for one internal node: weights, position, covariance moments, eigen-decomposition.
// One thread per node of the level; the SH rows are blended by hb_merge_shs_kernel with the weights left here.
__global__ void __launch_bounds__(128) hb_merge_kernel(int count, const int* __restrict__ level_list, HbNodes nd,
                                                      HbRows R, float* __restrict__ weight, int* __restrict__ errflag) {
  const int i = blockIdx.x * 128 + threadIdx.x;
  if (i >= count) return;
  const int id = level_list[i];
  const int m = nd.size[id];
  if (m <= 1) return;
  const size_t row = (size_t)nd.nodes[(size_t)id * 7 + 2];
  const size_t r[2] = {row + 1, row + 2 * (size_t)(m / 2)};
  float w[2], p[2][3];
#pragma unroll
  for (int c = 0; c < 2; ++c) {
    w[c] = __fmul_rn(R.opac[r[c]], hb_surface(R.scale[3 * r[c]], R.scale[3 * r[c] + 1], R.scale[3 * r[c] + 2]));
#pragma unroll
    for (int a = 0; a < 3; ++a) p[c][a] = R.pos[3 * r[c] + a];
  }
  const float wsum = __fadd_rn(__fadd_rn(0.0f, w[0]), w[1]);
  const float a0 = __fdiv_rn(w[0], wsum), a1 = __fdiv_rn(w[1], wsum);
  float pos[3];
#pragma unroll
  for (int a = 0; a < 3; ++a) pos[a] = __fadd_rn(__fadd_rn(0.0f, __fmul_rn(a0, p[0][a])), __fmul_rn(a1, p[1][a]));
  float cov[6] = {0, 0, 0, 0, 0, 0};
#pragma unroll
  for (int c = 0; c < 2; ++c) {
    const float a = c == 0 ? a0 : a1;
    const float dx = __fsub_rn(p[c][0], pos[0]), dy = __fsub_rn(p[c][1], pos[1]), dz = __fsub_rn(p[c][2], pos[2]);
    const float outer[6] = {__fmul_rn(dx, dx), __fmul_rn(dy, dx), __fmul_rn(dz, dx),
                            __fmul_rn(dy, dy), __fmul_rn(dz, dy), __fmul_rn(dz, dz)};
#pragma unroll
    for (int j = 0; j < 6; ++j) cov[j] = __fadd_rn(cov[j], __fmul_rn(a, __fadd_rn(R.cov[6 * r[c] + j], outer[j])));
  }
  float lam[3], V[3][3];
  hb_eigh3(cov, lam, V);
  if (lam[0] != lam[0] || lam[1] != lam[1] || lam[2] != lam[2]) atomicOr(errflag, 1);   // "Found Nans!" (:85-88)
  float mc[6] = {cov[0], cov[1], cov[2], cov[3], cov[4], cov[5]};
  for (int loops = 0; (lam[0] == 0.0f || lam[1] == 0.0f || lam[2] == 0.0f) && loops < 1000; ++loops) {   // :90-104
    mc[0] = __fadd_rn(mc[0], fmaxf(__fmul_rn(mc[0], 0.0001f), 1.1920929e-07f));
    mc[3] = __fadd_rn(mc[3], fmaxf(__fmul_rn(mc[3], 0.0001f), 1.1920929e-07f));
    mc[5] = __fadd_rn(mc[5], fmaxf(__fmul_rn(mc[5], 0.0001f), 1.1920929e-07f));
    hb_eigh3(mc, lam, V);
  }
  {   // right-handed frame (:106-112)
    const float cx = V[1][0] * V[2][1] - V[2][0] * V[1][1];
    const float cy = V[2][0] * V[0][1] - V[0][0] * V[2][1];
    const float cz = V[0][0] * V[1][1] - V[1][0] * V[0][1];
    if (cx * V[0][2] + cy * V[1][2] + cz * V[2][2] < 0.0f) { V[0][2] = -V[0][2]; V[1][2] = -V[1][2]; V[2][2] = -V[2][2]; }
  }
  const float s0 = sqrtf(fabsf(lam[0])), s1 = sqrtf(fabsf(lam[1])), s2 = sqrtf(fabsf(lam[2]));
  float q[4];
  hb_quat_from_matrix(V, q);
#pragma unroll
  for (int a = 0; a < 3; ++a) R.pos[3 * row + a] = pos[a];
  R.scale[3 * row] = s0; R.scale[3 * row + 1] = s1; R.scale[3 * row + 2] = s2;
#pragma unroll
  for (int a = 0; a < 4; ++a) R.rot[4 * row + a] = q[a];
#pragma unroll
  for (int a = 0; a < 6; ++a) R.cov[6 * row + a] = cov[a];
  R.opac[row] = __fdiv_rn(wsum, hb_surface(s0, s1, s2));
  weight[2 * (size_t)id] = a0;
  weight[2 * (size_t)id + 1] = a1;
}

// clustered.shs += a * g->shs (:60-64): twelve threads per node, one float4 of the 48 coefficients each
__global__ void __launch_bounds__(HB_THREADS) hb_merge_shs_kernel(long long count12, const int* __restrict__ level_list,
                                                                 HbNodes nd, const float* __restrict__ weight,
                                                                 float4* __restrict__ shs) {
  const long long t = (long long)blockIdx.x * HB_THREADS + threadIdx.x;
  if (t >= count12) return;
  const int i = (int)(t / 12), j = (int)(t - (long long)i * 12);
  const int id = level_list[i];
  const int m = nd.size[id];
  if (m <= 1) return;
  const size_t row = (size_t)nd.nodes[(size_t)id * 7 + 2];
  const float a0 = weight[2 * (size_t)id], a1 = weight[2 * (size_t)id + 1];
  const float4 x = shs[(row + 1) * 12 + j], y = shs[(row + 2 * (size_t)(m / 2)) * 12 + j];
  auto mix = [&](float u, float v) { return __fadd_rn(__fadd_rn(0.0f, __fmul_rn(a0, u)), __fmul_rn(a1, v)); };
  shs[row * 12 + j] = make_float4(mix(x.x, y.x), mix(x.y, y.y), mix(x.z, y.z), mix(x.w, y.w));
}

// rotation_aligner.cpp:23-88 for every node of a level below the root: the right-handed signed column permutation of
// the node's rotation that is closest (Frobenius inner product) to its parent's, scales permuted alike.
__global__ void __launch_bounds__(128) hb_align_kernel(int count, const int* __restrict__ level_list, HbNodes nd, HbRows R,
                                                      int* __restrict__ errflag) {
  const int i = blockIdx.x * 128 + threadIdx.x;
  if (i >= count) return;
  const int id = level_list[i];
  const int parent = nd.nodes[(size_t)id * 7 + 1];
  if (parent < 0) return;
  const size_t row = (size_t)nd.nodes[(size_t)id * 7 + 2], prow = (size_t)nd.nodes[(size_t)parent * 7 + 2];
  float qr[4], qm[4], Rref[3][3], Rm[3][3];
#pragma unroll
  for (int a = 0; a < 4; ++a) { qr[a] = R.rot[4 * prow + a]; qm[a] = R.rot[4 * row + a]; }
  hb_matrix_from_quat(qr, Rref);
  hb_matrix_from_quat(qm, Rm);
  float best_score = 0.0f;
  int best_perm = -1, best_state = 0;
  for (int perm = 0; perm < 6; ++perm) {   // (i, j, k) in the reference's loop order
    const int pi = perm >> 1;
    const int pj = (perm & 1) ? (pi == 2 ? 1 : 2) : (pi == 0 ? 1 : 0);
    const int pk = 3 - pi - pj;
    float base[3][3];
#pragma unroll
    for (int r_ = 0; r_ < 3; ++r_) {
      base[r_][0] = pi == 0 ? Rm[r_][0] : pi == 1 ? Rm[r_][1] : Rm[r_][2];
      base[r_][1] = pj == 0 ? Rm[r_][0] : pj == 1 ? Rm[r_][1] : Rm[r_][2];
      base[r_][2] = pk == 0 ? Rm[r_][0] : pk == 1 ? Rm[r_][1] : Rm[r_][2];
    }
    for (int state = 0; state < 8; ++state) {
      float t[3][3];
#pragma unroll
      for (int r_ = 0; r_ < 3; ++r_)
#pragma unroll
        for (int w = 0; w < 3; ++w) t[r_][w] = (state & (1 << w)) ? -base[r_][w] : base[r_][w];
      const float cx = __fsub_rn(__fmul_rn(t[1][0], t[2][1]), __fmul_rn(t[2][0], t[1][1]));
      const float cy = __fsub_rn(__fmul_rn(t[2][0], t[0][1]), __fmul_rn(t[0][0], t[2][1]));
      const float cz = __fsub_rn(__fmul_rn(t[0][0], t[1][1]), __fmul_rn(t[1][0], t[0][1]));
      const float det = __fadd_rn(__fadd_rn(__fmul_rn(cx, t[0][2]), __fmul_rn(cy, t[1][2])), __fmul_rn(cz, t[2][2]));
      if (det < 0.0f) continue;
      float score = 0.0f;
#pragma unroll
      for (int r_ = 0; r_ < 3; ++r_)
#pragma unroll
        for (int c_ = 0; c_ < 3; ++c_) score = __fadd_rn(score, __fmul_rn(t[r_][c_], Rref[r_][c_]));
      if (score > best_score) { best_score = score; best_perm = perm; best_state = state; }
    }
  }
  if (best_perm < 0) { atomicOr(errflag, 2); return; }   // the reference reads an uninitialised matrix here
  const int pi = best_perm >> 1;
  const int pj = (best_perm & 1) ? (pi == 2 ? 1 : 2) : (pi == 0 ? 1 : 0);
  const int pk = 3 - pi - pj;
  const int src[3] = {pi, pj, pk};
  float best[3][3], sc[3], s_in[3] = {R.scale[3 * row], R.scale[3 * row + 1], R.scale[3 * row + 2]};
#pragma unroll
  for (int w = 0; w < 3; ++w) {
    const float sg = (best_state & (1 << w)) ? -1.0f : 1.0f;
#pragma unroll
    for (int r_ = 0; r_ < 3; ++r_) {
      const float v = src[w] == 0 ? Rm[r_][0] : src[w] == 1 ? Rm[r_][1] : Rm[r_][2];
      best[r_][w] = sg < 0 ? -v : v;
    }
    sc[w] = src[w] == 0 ? s_in[0] : src[w] == 1 ? s_in[1] : s_in[2];
  }
  float q[4];
  hb_quat_from_matrix(best, q);
#pragma unroll
  for (int a = 0; a < 4; ++a) R.rot[4 * row + a] = q[a];
#pragma unroll
  for (int a = 0; a < 3; ++a) R.scale[3 * row + a] = sc[a];
}

// boxes decoded into the reference's layout (minn.xyz, w; maxx.xyz, w), w = largest extent for merged nodes
// (ClusterMerger.cpp:161-163), and scales -> log scales (writer.cpp:41,51)
__global__ void __launch_bounds__(HB_THREADS) hb_finish_kernel(long long nn, HbNodes nd, float* __restrict__ boxes,
                                                              const float* __restrict__ scale,
                                                              float* __restrict__ log_scales) {
  const long long id = (long long)blockIdx.x * HB_THREADS + threadIdx.x;
  if (id >= nn) return;
  const uint32_t* b = nd.box + (size_t)id * 6;
  const float lo[3] = {ord2f(b[0]), ord2f(b[1]), ord2f(b[2])}, hi[3] = {ord2f(b[3]), ord2f(b[4]), ord2f(b[5])};
  float w = 0.0f;
  if (nd.size[id] > 1)
    w = fmaxf(fmaxf(__fsub_rn(hi[0], lo[0]), __fsub_rn(hi[1], lo[1])), __fsub_rn(hi[2], lo[2]));
  float4* o = reinterpret_cast<float4*>(boxes) + 2 * id;
  o[0] = make_float4(lo[0], lo[1], lo[2], w);
  o[1] = make_float4(hi[0], hi[1], hi[2], w);
  // row `id` of the Gaussian arrays (rows and nodes are both 0..nn-1: one pass serves both)
#pragma unroll
  for (int a = 0; a < 3; ++a) log_scales[3 * id + a] = logf(scale[3 * id + a]);
}

}  // namespace

size_t hier_build_ws_bytes(long long N) { return HbWs::make(nullptr, N).bytes; }

// nodes per level of the tree over N Gaussians (the shape is a function of N alone)
static std::vector<long long> hb_level_counts(long long N) {
  std::vector<long long> out;
  long long s_lo = N, c_lo = 1, c_hi = 0;   // a level holds c_lo nodes of size s_lo and c_hi nodes of size s_lo + 1
  for (;;) {
    out.push_back(c_lo + c_hi);
    if (s_lo + (c_hi ? 1 : 0) <= 1) break;
    // children of size s: s/2 and s - s/2
    long long cnt[4] = {0, 0, 0, 0};
    const long long base = s_lo / 2;          // smallest child size
    auto add = [&](long long s, long long c) {
      if (c == 0 || s <= 1) return;            // leaves have no children
      cnt[s / 2 - base] += c;
      cnt[s - s / 2 - base] += c;
    };
    add(s_lo, c_lo);
    add(s_lo + 1, c_hi);
    if (cnt[0] + cnt[1] + cnt[2] + cnt[3] == 0) break;
    // sizes at one level differ by at most one; find the populated pair
    long long first = 0;
    while (first < 3 && cnt[first] == 0) ++first;
    s_lo = base + first;
    c_lo = cnt[first];
    c_hi = first + 1 < 4 ? cnt[first + 1] : 0;
  }
  return out;
}

cudaError_t run_hier_build(long long N64, const float* pos, const float* shs, const float* opac, const float* scale,
                           const float* rot, int align, float* o_pos, float* o_shs, float* o_opac, float* o_logs,
                           float* o_rot, int* o_nodes, float* o_boxes, float* o_cov, int* o_leaf_order, void* workspace,
                           int* err_host, cudaStream_t s) {
  *err_host = 0;
  if (N64 <= 0) return cudaSuccess;
  const int N = (int)N64;
  const long long nn = 2 * N64 - 1;
  HbWs w = HbWs::make(workspace, N64);
  w.nd.nodes = o_nodes;
  HbRows R{o_pos, o_shs, o_opac, w.scale, o_rot, o_cov ? o_cov : w.cov};
  const int gN = (N + HB_THREADS - 1) / HB_THREADS;
  const std::vector<long long> counts = hb_level_counts(N64);
  const int levels = (int)counts.size();
  if (levels > HB_MAX_LEVELS) return cudaErrorInvalidValue;
  std::vector<long long> off(levels + 1, 0);
  for (int l = 0; l < levels; ++l) off[l + 1] = off[l] + counts[l];

  cudaError_t e;
  {
    ProfScope ps(K_HIER_TREE, s);
    H3_CUDA_RET(cudaMemsetAsync(w.errflag, 0, 4 * sizeof(int), s));
    H3_CUDA_RET(cudaMemsetAsync(o_nodes, 0, (size_t)nn * 7 * sizeof(int), s));
    hb_init_kernel<<<gN, HB_THREADS, 0, s>>>(N, pos, scale, w.lo, w.hi, w.node_of[0]);
    hb_box_reset_kernel<<<(unsigned)((nn * 6 + HB_THREADS - 1) / HB_THREADS), HB_THREADS, 0, s>>>(nn * 6, w.nd.box);
    hb_root_kernel<<<1, 32, 0, s>>>(N, w.nd, w.level_nodes, w.level_fill);
    // the three coordinate orders; the sort leaves its result in buffer 0, list[1][a] serves as the other value buffer
    for (int a = 0; a < 3; ++a) {
      hb_keys_kernel<<<gN, HB_THREADS, 0, s>>>(N, pos, a, w.keys[0], w.list[0][a]);
      uint32_t* const ks[2] = {w.keys[0], w.keys[1]};
      uint32_t* const vs[2] = {w.list[0][a], w.list[1][a]};
      e = launch_sort_pairs(ks, vs, N, w.sort_ws, s);
      if (e != cudaSuccess) return e;
    }
    const int sblocks = (N + HB_SCAN_TILE - 1) / HB_SCAN_TILE;
    int cur = 0;
    for (int l = 0; l < levels; ++l) {
      const int cnt = (int)counts[l];
      const int* list = w.level_nodes + off[l];
      hb_bounds_kernel<<<gN, HB_THREADS, 0, s>>>(N, w.list[cur][0], w.node_of[cur], w.nd.level, l, w.lo, w.hi, w.nd.box);
      hb_split_kernel<<<(cnt + HB_THREADS - 1) / HB_THREADS, HB_THREADS, 0, s>>>(
          cnt, list, w.nd, l + 1 < levels ? w.level_nodes + off[l + 1] : nullptr, w.level_fill + l + 1, l);
      if (l + 1 == levels) break;
      const uint32_t *l0 = w.list[cur][0], *l1 = w.list[cur][1], *l2 = w.list[cur][2];
      hb_side_kernel<<<gN, HB_THREADS, 0, s>>>(N, l0, l1, l2, w.node_of[cur], w.nd, l, w.side);
      hb_scan_totals_kernel<<<sblocks, HB_THREADS, 0, s>>>(N, l0, l1, l2, w.side, sblocks, w.bsum);
      hb_scan_blocks_kernel<<<3, 1024, 0, s>>>(sblocks, w.bsum);
      hb_scan_final_kernel<<<sblocks, HB_THREADS, 0, s>>>(N, l0, l1, l2, w.side, sblocks, w.bsum, w.lscan);
      hb_scatter_kernel<<<gN, HB_THREADS, 0, s>>>(N, l0, l1, l2, w.list[cur ^ 1][0], w.list[cur ^ 1][1],
                                                  w.list[cur ^ 1][2], w.side, w.node_of[cur], w.node_of[cur ^ 1], w.nd,
                                                  l, w.lscan);
      cur ^= 1;
    }
    hb_leaf_rows_kernel<<<gN, HB_THREADS, 0, s>>>(N, w.node_of[cur], o_nodes, pos, opac, scale, rot, R,
                                                  w.row_of_gaussian, o_leaf_order);
    hb_leaf_shs_kernel<<<(unsigned)((N64 * 12 + HB_THREADS - 1) / HB_THREADS), HB_THREADS, 0, s>>>(
        N64 * 12, w.row_of_gaussian, reinterpret_cast<const float4*>(shs), reinterpret_cast<float4*>(o_shs));
  }
  e = cudaGetLastError();
  if (e != cudaSuccess) return e;
  {
    ProfScope ps(K_HIER_MERGE, s);
    for (int l = levels - 2; l >= 0; --l) {
      const int cnt = (int)counts[l];
      const int* list = w.level_nodes + off[l];
      hb_merge_kernel<<<(cnt + 127) / 128, 128, 0, s>>>(cnt, list, w.nd, R, w.weight, w.errflag);
      hb_merge_shs_kernel<<<(unsigned)(((long long)cnt * 12 + HB_THREADS - 1) / HB_THREADS), HB_THREADS, 0, s>>>(
          (long long)cnt * 12, list, w.nd, w.weight, reinterpret_cast<float4*>(o_shs));
    }
    if (align)
      for (int l = 1; l < levels; ++l) {
        const int cnt = (int)counts[l];
        hb_align_kernel<<<(cnt + 127) / 128, 128, 0, s>>>(cnt, w.level_nodes + off[l], w.nd, R, w.errflag);
      }
    hb_finish_kernel<<<(unsigned)((nn + HB_THREADS - 1) / HB_THREADS), HB_THREADS, 0, s>>>(nn, w.nd, o_boxes, w.scale,
                                                                                          o_logs);
  }
  e = cudaGetLastError();
  if (e != cudaSuccess) return e;
  int flags[4] = {0, 0, 0, 0};
  e = cudaMemcpyAsync(flags, w.errflag, sizeof(flags), cudaMemcpyDeviceToHost, s);
  if (e != cudaSuccess) return e;
  e = cudaStreamSynchronize(s);
  if (e != cudaSuccess) return e;
  *err_host = flags[0];
  return cudaSuccess;
}

}  // namespace h3dgs

extern "C" {

size_t h3dgs_hier_build_workspace_bytes(int64_t num_gaussians) {
  return h3dgs::hier_build_ws_bytes(num_gaussians);
}

int h3dgs_hier_build(int64_t num_gaussians, const float* positions, const float* shs, const float* opacities,
                     const float* scales, const float* rotations, int32_t align, float* out_positions, float* out_shs,
                     float* out_opacities, float* out_log_scales, float* out_rotations, int32_t* out_nodes,
                     float* out_boxes, float* out_covariances, int32_t* out_leaf_order, void* workspace,
                     size_t workspace_bytes, h3dgs_stream_t stream) {
  using namespace h3dgs;
  if (num_gaussians < 0 || num_gaussians > (1ll << 30)) return fail(H3DGS_EINVAL, "h3dgs_hier_build: bad Gaussian count");
  if (num_gaussians == 0) return 0;
  if (!positions || !shs || !opacities || !scales || !rotations || !out_positions || !out_shs || !out_opacities ||
      !out_log_scales || !out_rotations || !out_nodes || !out_boxes || !workspace)
    return fail(H3DGS_EINVAL, "h3dgs_hier_build: NULL argument");
  if (workspace_bytes < hier_build_ws_bytes(num_gaussians)) return fail(H3DGS_ENOSPACE, "h3dgs_hier_build: workspace too small");
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  if (out_leaf_order) {
    cudaError_t e = cudaMemsetAsync(out_leaf_order, 0xff, (size_t)(2 * num_gaussians - 1) * sizeof(int32_t), s);
    if (e != cudaSuccess) return cuda_fail(e, "cudaMemsetAsync(leaf_order)");
  }
  int err = 0;
  cudaError_t e = run_hier_build(num_gaussians, positions, shs, opacities, scales, rotations, align, out_positions,
                                 out_shs, out_opacities, out_log_scales, out_rotations, out_nodes, out_boxes,
                                 out_covariances, out_leaf_order, workspace, &err, s);
  if (e != cudaSuccess) return cuda_fail(e, "h3dgs_hier_build");
  if (err & 1) return fail(H3DGS_EINVAL, "h3dgs_hier_build: Found Nans! (ClusterMerger.cpp:85-88)");
  if (err & 2) return fail(H3DGS_EINVAL, "h3dgs_hier_build: rotation alignment found no candidate with a positive score");
  return 0;
}

}  // extern "C"
