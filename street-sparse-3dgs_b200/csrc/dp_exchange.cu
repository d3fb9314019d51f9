// Gradient exchange of camera-batch data parallelism over NVLink peer memory (SURVEY.md 8e mode B, BASELINE.json
// config 4: "camera batch data-parallel with NCCL grad allreduce").  The reference has no in-process parallelism;
// the semantics are those of an all-reduce(sum) / world of the six per-row gradient tensors followed by
// OurAdam.step(relevant) (train_post.py:171-190), restricted to the rows some rank's render touched.
//
// Every rank's backward leaves COMPACT gradients (H3dgsGrads.compact_*).  Instead of all-reducing N-row tensors or
// a padded union, the exchange is an all-gather THROUGH PEER MEMORY that is summed while it is read:
//   1. dp_pack_kernel        the rows whose gate (opacity gradient) is non-zero -> one packed [U, 59] block + row ids
//                            in this rank's arena (symmetric memory: every peer maps it)
//   2. (device-side barrier of the symmetric-memory handle: torch plumbing, no data)
//   3. dp_accumulate_kernel  once per peer, IN RANK ORDER: streams the peer's block over NVLink (contiguous 236-byte
//                            rows, fully coalesced remote loads) and adds it into a local union buffer; a local
//                            row -> union-slot map assigns slots on first touch.  A peer lists a row at most once, so
//                            one launch never adds twice into one slot, and launches are stream-ordered: every rank
//                            forms the same sums in the same order and the replicas stay bit-identical.
//   4. dp_adam_kernel        average + sparse Adam over the union, restoring the map to -1 on the way.
// Bytes over NVLink per rank and step: (world - 1) x U x 240 B, read once, nothing written remotely.
#include "common.cuh"

namespace h3dgs {

namespace {

struct DpTensors {
  const float* grad[H3DGS_ADAM_MAX_TENSORS];
  int width[H3DGS_ADAM_MAX_TENSORS];
  int start[H3DGS_ADAM_MAX_TENSORS + 1];
  int n;
};

// one warp per 32 compact rows: ballot of the gate, one atomic per warp, then the warp copies the surviving rows
__global__ void __launch_bounds__(256) dp_pack_kernel(const DpTensors t, const long long* __restrict__ compact_rows,
                                                     const int* __restrict__ compact_count,
                                                     const float* __restrict__ gate, int* __restrict__ out_rows,
                                                     float* __restrict__ out_packed, int* __restrict__ out_count,
                                                     int out_capacity, int* __restrict__ overflow) {
  const int W = t.start[t.n];
  const int U = *compact_count;
  const int lane = threadIdx.x & 31;
  const int warps = (gridDim.x * blockDim.x) >> 5;
  for (int base = (((blockIdx.x * blockDim.x) + threadIdx.x) >> 5) * 32; base < U; base += warps * 32) {
    const int i = base + lane;
    const bool keep = i < U && (gate == nullptr || gate[i] != 0.0f);
    const unsigned m = __ballot_sync(0xffffffffu, keep);
    if (m == 0) continue;
    int slot0 = 0;
    if (lane == 0) slot0 = atomicAdd(out_count, __popc(m));
    slot0 = __shfl_sync(0xffffffffu, slot0, 0);
    if (slot0 + __popc(m) > out_capacity) {   // the arena is too small for this render: reported, nothing written
      if (lane == 0) atomicExch(overflow, 1);
      continue;
    }
    if (keep) out_rows[slot0 + __popc(m & ((1u << lane) - 1u))] = (int)compact_rows[i];
    unsigned rest = m;
    int slot = slot0;
    while (rest) {
      const int src = base + __ffs(rest) - 1;
      rest &= rest - 1;
      for (int e = lane; e < W; e += 32) {
        int ti = 0;
#pragma unroll
        for (int k = 1; k < H3DGS_ADAM_MAX_TENSORS; ++k)
          if (k < t.n && e >= t.start[k]) ti = k;
        out_packed[(size_t)slot * W + e] = t.grad[ti][(size_t)src * t.width[ti] + (e - t.start[ti])];
      }
      ++slot;
    }
  }
}

// 32 peer rows per warp iteration: every lane resolves one row's union slot (one warp-aggregated atomic for the new
// ones), then the warp moves the rows four at a time so that eight remote loads per lane are in flight.  peer_* may
// be remote (NVLink) pointers; W <= 64.  owner_mod > 1: only the rows with (row >> 6) % owner_mod == owner_rem are
// taken (sharded optimizer: every rank sums and updates the rows it owns; runs of 64 consecutive rows share an owner so
// that the rows of one 32-byte sector of a 12-byte-per-row tensor are written by one kernel, close together in time).
constexpr int DP_RB = 4;
__global__ void __launch_bounds__(256) dp_accumulate_kernel(const int* __restrict__ peer_rows,
                                                           const float* __restrict__ peer_packed,
                                                           const int* __restrict__ peer_count, int peer_capacity,
                                                           int W, int owner_mod, int owner_rem,
                                                           int* __restrict__ row_map, int* __restrict__ union_rows,
                                                           float* __restrict__ union_packed,
                                                           int* __restrict__ union_count, int union_capacity,
                                                           int* __restrict__ overflow) {
  int U = *peer_count;
  if (U > peer_capacity) U = peer_capacity;   // the peer overflowed its arena (it reports that itself)
  const int lane = threadIdx.x & 31;
  const int warps = (gridDim.x * blockDim.x) >> 5;
  const bool two = lane + 32 < W;
  for (int base = (((blockIdx.x * blockDim.x) + threadIdx.x) >> 5) * 32; base < U; base += warps * 32) {
    const int i = base + lane;
    int slot = -1;
    bool fresh = false, mine = false;
    int row = 0;
    if (i < U) {
      row = peer_rows[i];
      mine = owner_mod <= 1 || ((row >> H3DGS_DP_OWNER_SHIFT) % owner_mod) == owner_rem;
      if (mine) {
        slot = row_map[row];
        fresh = slot < 0;
      }
    }
    const unsigned nm = __ballot_sync(0xffffffffu, fresh);
    if (nm) {
      int s0 = 0;
      if (lane == 0) s0 = atomicAdd(union_count, __popc(nm));
      s0 = __shfl_sync(0xffffffffu, s0, 0);
      if (fresh) {
        slot = s0 + __popc(nm & ((1u << lane) - 1u));
        if (slot < union_capacity) { row_map[row] = slot; union_rows[slot] = row; }
      }
    }
    if (mine && slot >= union_capacity) { atomicExch(overflow, 2); slot = -1; }
    unsigned todo = __ballot_sync(0xffffffffu, slot >= 0);
    while (todo) {
      int lj[DP_RB], sj[DP_RB], fj[DP_RB];
      float a[DP_RB][2], d[DP_RB][2];
#pragma unroll
      for (int r = 0; r < DP_RB; ++r) {
        lj[r] = todo ? __ffs(todo) - 1 : -1;
        if (todo) todo &= todo - 1;
        sj[r] = __shfl_sync(0xffffffffu, slot, lj[r] & 31);
        fj[r] = __shfl_sync(0xffffffffu, (int)fresh, lj[r] & 31);
        if (lj[r] < 0) sj[r] = -1;
      }
#pragma unroll
      for (int r = 0; r < DP_RB; ++r) {
        a[r][0] = a[r][1] = d[r][0] = d[r][1] = 0.0f;
        if (sj[r] < 0) continue;
        const float* src = peer_packed + (size_t)(base + lj[r]) * W;
        const float* dst = union_packed + (size_t)sj[r] * W;
        a[r][0] = src[lane];
        if (two) a[r][1] = src[lane + 32];
        if (!fj[r]) {
          d[r][0] = dst[lane];
          if (two) d[r][1] = dst[lane + 32];
        }
      }
#pragma unroll
      for (int r = 0; r < DP_RB; ++r) {
        if (sj[r] < 0) continue;
        float* dst = union_packed + (size_t)sj[r] * W;
        dst[lane] = fj[r] ? a[r][0] : __fadd_rn(d[r][0], a[r][0]);
        if (two) dst[lane + 32] = fj[r] ? a[r][1] : __fadd_rn(d[r][1], a[r][1]);
      }
    }
  }
}

struct DpAdam {
  H3dgsAdamTensor t[H3DGS_ADAM_MAX_TENSORS];
  int start[H3DGS_ADAM_MAX_TENSORS + 1];
  int n;
};

// Sparse Adam over the union: gradient = union_packed * scale (the 1/world average), arithmetic as sparse_adam.cu.
// A warp takes four rows per iteration, lanes cover a row's W <= 64 floats (two per lane): the 4 x 2 x 4 loads of a
// lane are issued before the first dependent instruction -- the rows are scattered over a multi-GB store, the kernel
// lives on memory-level parallelism.
__device__ __forceinline__ void adam_update(const H3dgsAdamTensor& T, float g, float& m, float& v, float& p) {
  m = __fmaf_rn(T.one_minus_beta1, g, __fmul_rn(m, T.beta1));
  v = __fmaf_rn(__fmul_rn(T.one_minus_beta2, g), g, __fmul_rn(v, T.beta2));
  const float inv_bc2 = 1.0f / T.bias_correction2_sqrt;
  const float denom = __fadd_rn(__fmul_rn(sqrtf(v), inv_bc2), T.eps);
  p = __fmaf_rn(-T.step_size, m / denom, p);
}

__global__ void __launch_bounds__(256) dp_adam_kernel(const DpAdam pk, const int* __restrict__ union_rows,
                                                     const int* __restrict__ union_count,
                                                     const float* __restrict__ union_packed, float scale,
                                                     int* __restrict__ row_map, int union_capacity,
                                                     int* __restrict__ out_rows, float* __restrict__ out_packed,
                                                     int* __restrict__ out_count) {
  const int W = pk.start[pk.n];
  int U = *union_count;
  if (U > union_capacity) U = union_capacity;
  if (out_count != nullptr && blockIdx.x == 0 && threadIdx.x == 0) *out_count = U;
  const int lane = threadIdx.x & 31;
  const int warps = (gridDim.x * blockDim.x) >> 5;
  // the two elements of a row this lane owns, and the tensor each falls into
  int ti[2], col[2];
  bool on[2];
#pragma unroll
  for (int q = 0; q < 2; ++q) {
    const int e = lane + 32 * q;
    on[q] = e < W;
    int t = 0;
#pragma unroll
    for (int i = 1; i < H3DGS_ADAM_MAX_TENSORS; ++i)
      if (i < pk.n && e >= pk.start[i]) t = i;
    ti[q] = t;
    col[q] = e - pk.start[t];
  }
  for (int base = (((blockIdx.x * blockDim.x) + threadIdx.x) >> 5) * DP_RB; base < U; base += warps * DP_RB) {
    int row[DP_RB];
    float g[DP_RB][2], m[DP_RB][2], v[DP_RB][2], p[DP_RB][2];
#pragma unroll
    for (int r = 0; r < DP_RB; ++r) row[r] = base + r < U ? union_rows[base + r] : -1;
#pragma unroll
    for (int r = 0; r < DP_RB; ++r) {
      if (row[r] < 0) continue;
#pragma unroll
      for (int q = 0; q < 2; ++q) {
        if (!on[q]) continue;
        const H3dgsAdamTensor& T = pk.t[ti[q]];
        const size_t off = (size_t)row[r] * T.width + col[q];
        g[r][q] = union_packed[(size_t)(base + r) * W + lane + 32 * q];
        m[r][q] = T.exp_avg[off];
        v[r][q] = T.exp_avg_sq[off];
        p[r][q] = T.param[off];
      }
    }
#pragma unroll
    for (int r = 0; r < DP_RB; ++r) {
      if (row[r] < 0) continue;
      if (lane == 0 && row_map != nullptr) row_map[row[r]] = -1;
      if (lane == 0 && out_rows != nullptr) out_rows[base + r] = row[r];
#pragma unroll
      for (int q = 0; q < 2; ++q) {
        if (!on[q]) continue;
        const H3dgsAdamTensor& T = pk.t[ti[q]];
        const size_t off = (size_t)row[r] * T.width + col[q];
        adam_update(T, __fmul_rn(g[r][q], scale), m[r][q], v[r][q], p[r][q]);
        T.exp_avg[off] = m[r][q];
        T.exp_avg_sq[off] = v[r][q];
        T.param[off] = p[r][q];
        if (out_packed != nullptr) out_packed[(size_t)(base + r) * W + lane + 32 * q] = p[r][q];
      }
    }
  }
}

// Sharded optimizer, last stage: the parameter rows a peer owns and has just updated -> the local replica.
__global__ void __launch_bounds__(256) dp_scatter_params_kernel(const DpAdam pk, const int* __restrict__ peer_rows,
                                                               const float* __restrict__ peer_packed,
                                                               const int* __restrict__ peer_count, int peer_capacity) {
  const int W = pk.start[pk.n];
  int U = *peer_count;
  if (U > peer_capacity) U = peer_capacity;
  const int lane = threadIdx.x & 31;
  const int warps = (gridDim.x * blockDim.x) >> 5;
  int ti[2], col[2];
  bool on[2];
#pragma unroll
  for (int q = 0; q < 2; ++q) {
    const int e = lane + 32 * q;
    on[q] = e < W;
    int t = 0;
#pragma unroll
    for (int i = 1; i < H3DGS_ADAM_MAX_TENSORS; ++i)
      if (i < pk.n && e >= pk.start[i]) t = i;
    ti[q] = t;
    col[q] = e - pk.start[t];
  }
  constexpr int RB = 8;
  for (int base = (((blockIdx.x * blockDim.x) + threadIdx.x) >> 5) * RB; base < U; base += warps * RB) {
    int row[RB];
    float v[RB][2];
#pragma unroll
    for (int r = 0; r < RB; ++r) row[r] = base + r < U ? peer_rows[base + r] : -1;
#pragma unroll
    for (int r = 0; r < RB; ++r) {
      if (row[r] < 0) continue;
#pragma unroll
      for (int q = 0; q < 2; ++q)
        if (on[q]) v[r][q] = peer_packed[(size_t)(base + r) * W + lane + 32 * q];
    }
#pragma unroll
    for (int r = 0; r < RB; ++r) {
      if (row[r] < 0) continue;
#pragma unroll
      for (int q = 0; q < 2; ++q)
        if (on[q]) pk.t[ti[q]].param[(size_t)row[r] * pk.t[ti[q]].width + col[q]] = v[r][q];
    }
  }
}

int grid_for(long long threads) {
  long long blocks = (threads + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  if (blocks < 1) blocks = 1;
  return (int)blocks;
}

}  // namespace

}  // namespace h3dgs

extern "C" {

int h3dgs_dp_pack(const H3dgsRowTensor* grads, int32_t num_tensors, const int64_t* compact_rows,
                  const int32_t* compact_count, int32_t capacity, const float* gate, int32_t* out_rows,
                  float* out_packed, int32_t* out_count, int32_t out_capacity, int32_t* overflow,
                  h3dgs_stream_t stream) {
  using namespace h3dgs;
  if (!grads || num_tensors <= 0 || num_tensors > H3DGS_ADAM_MAX_TENSORS || !compact_rows || !compact_count ||
      !out_rows || !out_packed || !out_count || !overflow || capacity < 0 || out_capacity < 0)
    return fail(H3DGS_EINVAL, "h3dgs_dp_pack: bad argument");
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  DpTensors t;
  t.n = num_tensors;
  int acc = 0;
  for (int i = 0; i < num_tensors; ++i) {
    if (!grads[i].data || grads[i].width <= 0) return fail(H3DGS_EINVAL, "h3dgs_dp_pack: bad gradient tensor");
    t.grad[i] = grads[i].data;
    t.width[i] = grads[i].width;
    t.start[i] = acc;
    acc += grads[i].width;
  }
  for (int i = num_tensors; i <= H3DGS_ADAM_MAX_TENSORS; ++i) t.start[i] = acc;
  H3_CUDA(cudaMemsetAsync(out_count, 0, sizeof(int32_t), s));
  if (capacity == 0) return 0;
  ProfScope ps(K_ADAM, s);
  dp_pack_kernel<<<grid_for(capacity), 256, 0, s>>>(t, reinterpret_cast<const long long*>(compact_rows), compact_count,
                                                     gate, out_rows, out_packed, out_count, out_capacity, overflow);
  H3_CUDA(cudaGetLastError());
  return 0;
}

int h3dgs_dp_accumulate(const int32_t* peer_rows, const float* peer_packed, const int32_t* peer_count,
                        int32_t peer_capacity, int32_t width, int32_t owner_mod, int32_t owner_rem,
                        int32_t* row_map, int32_t* union_rows,
                        float* union_packed, int32_t* union_count, int32_t union_capacity, int32_t* overflow,
                        h3dgs_stream_t stream) {
  using namespace h3dgs;
  if (!peer_rows || !peer_packed || !peer_count || width <= 0 || !row_map || !union_rows || !union_packed ||
      !union_count || !overflow || peer_capacity < 0 || union_capacity < 0 ||
      (owner_mod > 1 && (owner_rem < 0 || owner_rem >= owner_mod)))
    return fail(H3DGS_EINVAL, "h3dgs_dp_accumulate: bad argument");
  if (peer_capacity == 0) return 0;
  if (width > 64) return fail(H3DGS_EINVAL, "h3dgs_dp_accumulate: rows wider than 64 floats are not supported");
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  ProfScope ps(K_ADAM, s);
  dp_accumulate_kernel<<<grid_for((long long)peer_capacity), 256, 0, s>>>(
      peer_rows, peer_packed, peer_count, peer_capacity, width, owner_mod, owner_rem, row_map, union_rows, union_packed, union_count, union_capacity, overflow);
  H3_CUDA(cudaGetLastError());
  return 0;
}

int h3dgs_dp_adam(const H3dgsAdamTensor* tensors, int32_t num_tensors, const int32_t* union_rows,
                  const int32_t* union_count, const float* union_packed, float scale, int32_t* row_map,
                  int32_t union_capacity, int32_t* out_rows, float* out_packed, int32_t* out_count,
                  h3dgs_stream_t stream) {
  using namespace h3dgs;
  if (!tensors || num_tensors <= 0 || num_tensors > H3DGS_ADAM_MAX_TENSORS || !union_rows || !union_count ||
      !union_packed || union_capacity < 0)
    return fail(H3DGS_EINVAL, "h3dgs_dp_adam: bad argument");
  DpAdam pk;
  pk.n = num_tensors;
  int acc = 0;
  for (int i = 0; i < num_tensors; ++i) {
    if (!tensors[i].param || !tensors[i].exp_avg || !tensors[i].exp_avg_sq || tensors[i].width <= 0)
      return fail(H3DGS_EINVAL, "h3dgs_dp_adam: bad tensor");
    pk.t[i] = tensors[i];
    pk.start[i] = acc;
    acc += tensors[i].width;
  }
  for (int i = num_tensors; i <= H3DGS_ADAM_MAX_TENSORS; ++i) pk.start[i] = acc;
  if (union_capacity == 0) return 0;
  if (acc > 64) return fail(H3DGS_EINVAL, "h3dgs_dp_adam: rows wider than 64 floats are not supported");
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  ProfScope ps(K_ADAM, s);
  dp_adam_kernel<<<grid_for((long long)union_capacity * 32 / DP_RB), 256, 0, s>>>(pk, union_rows, union_count, union_packed,
                                                                           scale, row_map, union_capacity, out_rows,
                                                                           out_packed, out_count);
  H3_CUDA(cudaGetLastError());
  return 0;
}

int h3dgs_dp_scatter_params(const H3dgsAdamTensor* tensors, int32_t num_tensors, const int32_t* peer_rows,
                            const float* peer_packed, const int32_t* peer_count, int32_t peer_capacity,
                            h3dgs_stream_t stream) {
  using namespace h3dgs;
  if (!tensors || num_tensors <= 0 || num_tensors > H3DGS_ADAM_MAX_TENSORS || !peer_rows || !peer_packed ||
      !peer_count || peer_capacity < 0)
    return fail(H3DGS_EINVAL, "h3dgs_dp_scatter_params: bad argument");
  DpAdam pk;
  pk.n = num_tensors;
  int acc = 0;
  for (int i = 0; i < num_tensors; ++i) {
    if (!tensors[i].param || tensors[i].width <= 0) return fail(H3DGS_EINVAL, "h3dgs_dp_scatter_params: bad tensor");
    pk.t[i] = tensors[i];
    pk.start[i] = acc;
    acc += tensors[i].width;
  }
  for (int i = num_tensors; i <= H3DGS_ADAM_MAX_TENSORS; ++i) pk.start[i] = acc;
  if (acc > 64) return fail(H3DGS_EINVAL, "h3dgs_dp_scatter_params: rows wider than 64 floats are not supported");
  if (peer_capacity == 0) return 0;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  ProfScope ps(K_ADAM, s);
  dp_scatter_params_kernel<<<grid_for((long long)peer_capacity * 4), 256, 0, s>>>(pk, peer_rows, peer_packed, peer_count,
                                                                                  peer_capacity);
  H3_CUDA(cudaGetLastError());
  return 0;
}

}  // extern "C"
