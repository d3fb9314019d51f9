// Fused sparse Adam: one launch updates the rows listed in `relevant` of every parameter tensor.
// Replaces the gather -> update -> scatter sequence of scene/OurAdam.py:249-337 (about ten torch
// kernels and three row-gathered temporaries per parameter group).  Rows that are not listed keep
// stale moments, exactly as in the reference.  Per element: 4 reads + 3 writes of 4 bytes.
//
// Rounding follows the order of the torch ops being replaced (mul_, add_(alpha), addcmul_, sqrt,
// scalar division as multiplication by the fp32 reciprocal, add_(eps), addcdiv_).
#include "common.cuh"

namespace h3dgs {

namespace {

struct AdamPack {
  H3dgsAdamTensor t[H3DGS_ADAM_MAX_TENSORS];
  int start[H3DGS_ADAM_MAX_TENSORS + 1];  // prefix sums of widths
  int n;
};

__global__ void __launch_bounds__(256) sparse_adam_kernel(const AdamPack pk,
                                                          const long long* __restrict__ relevant,
                                                          long long K) {
  const int total_w = pk.start[pk.n];
  const long long gid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= K * total_w) return;
  const long long k = gid / total_w;
  const int e = (int)(gid - k * total_w);
  int ti = 0;
#pragma unroll
  for (int i = 1; i < H3DGS_ADAM_MAX_TENSORS; ++i)
    if (i < pk.n && e >= pk.start[i]) ti = i;
  const H3dgsAdamTensor& T = pk.t[ti];
  const long long row = relevant ? relevant[k] : k;
  // a sorted list with repeated ids updates each id once (the reference scatters identical copies, OurAdam.py:334-337)
  if (relevant && k > 0 && relevant[k - 1] == row) return;
  const size_t off = (size_t)row * T.width + (e - pk.start[ti]);

  const float g = T.grad[off];
  float m = T.exp_avg[off];
  float v = T.exp_avg_sq[off];
  float p = T.param[off];

  const float omb1 = T.one_minus_beta1;
  const float omb2 = T.one_minus_beta2;
  m = __fmaf_rn(omb1, g, __fmul_rn(m, T.beta1));
  v = __fmaf_rn(__fmul_rn(omb2, g), g, __fmul_rn(v, T.beta2));
  const float inv_bc2 = 1.0f / T.bias_correction2_sqrt;
  const float denom = __fadd_rn(__fmul_rn(sqrtf(v), inv_bc2), T.eps);
  p = __fmaf_rn(-T.step_size, m / denom, p);

  T.exp_avg[off] = m;
  T.exp_avg_sq[off] = v;
  T.param[off] = p;
}

// compact gradients: row i of every grad tensor belongs to store row rows[i]; the row count lives on the device.
// A warp takes four rows per iteration and its lanes cover a row's floats (two per lane, rows of up to 64 floats):
// the 4 x 2 x 4 loads of a lane are issued before the first dependent instruction.  The rows are scattered over a
// multi-GB store, so the kernel lives on memory-level parallelism, not on occupancy alone.
constexpr int ADAM_RB = 4;
__global__ void __launch_bounds__(256) sparse_adam_compact_rows_kernel(const AdamPack pk,
                                                                       const long long* __restrict__ rows,
                                                                       const int* __restrict__ count,
                                                                       const float* __restrict__ gate) {
  const int W = pk.start[pk.n];
  const int U = *count;
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
  for (int base = (((blockIdx.x * blockDim.x) + threadIdx.x) >> 5) * ADAM_RB; base < U; base += warps * ADAM_RB) {
    long long row[ADAM_RB];
    float g[ADAM_RB][2], m[ADAM_RB][2], v[ADAM_RB][2], p[ADAM_RB][2];
#pragma unroll
    for (int r = 0; r < ADAM_RB; ++r) {
      const int k = base + r;
      row[r] = (k < U && (gate == nullptr || gate[k] != 0.0f)) ? rows[k] : -1;
    }
#pragma unroll
    for (int r = 0; r < ADAM_RB; ++r) {
      if (row[r] < 0) continue;
#pragma unroll
      for (int q = 0; q < 2; ++q) {
        if (!on[q]) continue;
        const H3dgsAdamTensor& T = pk.t[ti[q]];
        const size_t off = (size_t)row[r] * T.width + col[q];
        g[r][q] = T.grad[(size_t)(base + r) * T.width + col[q]];
        m[r][q] = T.exp_avg[off];
        v[r][q] = T.exp_avg_sq[off];
        p[r][q] = T.param[off];
      }
    }
#pragma unroll
    for (int r = 0; r < ADAM_RB; ++r) {
      if (row[r] < 0) continue;
#pragma unroll
      for (int q = 0; q < 2; ++q) {
        if (!on[q]) continue;
        const H3dgsAdamTensor& T = pk.t[ti[q]];
        const size_t off = (size_t)row[r] * T.width + col[q];
        float mm = m[r][q], vv = v[r][q], pp = p[r][q];
        const float gg = g[r][q];
        mm = __fmaf_rn(T.one_minus_beta1, gg, __fmul_rn(mm, T.beta1));
        vv = __fmaf_rn(__fmul_rn(T.one_minus_beta2, gg), gg, __fmul_rn(vv, T.beta2));
        const float inv_bc2 = 1.0f / T.bias_correction2_sqrt;
        const float denom = __fadd_rn(__fmul_rn(sqrtf(vv), inv_bc2), T.eps);
        pp = __fmaf_rn(-T.step_size, mm / denom, pp);
        T.exp_avg[off] = mm;
        T.exp_avg_sq[off] = vv;
        T.param[off] = pp;
      }
    }
  }
}

// general form (rows wider than 64 floats): one thread per element
__global__ void __launch_bounds__(256) sparse_adam_compact_kernel(const AdamPack pk,
                                                                  const long long* __restrict__ rows,
                                                                  const int* __restrict__ count,
                                                                  const float* __restrict__ gate) {
  const int total_w = pk.start[pk.n];
  const long long U = *count;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long gid = (long long)blockIdx.x * blockDim.x + threadIdx.x; gid < U * total_w; gid += stride) {
    const long long k = gid / total_w;
    if (gate != nullptr && gate[k] == 0.0f) continue;
    const int e = (int)(gid - k * total_w);
    int ti = 0;
#pragma unroll
    for (int i = 1; i < H3DGS_ADAM_MAX_TENSORS; ++i)
      if (i < pk.n && e >= pk.start[i]) ti = i;
    const H3dgsAdamTensor& T = pk.t[ti];
    const int w = e - pk.start[ti];
    const size_t off = (size_t)rows[k] * T.width + w;
    const float g = T.grad[(size_t)k * T.width + w];
    float m = T.exp_avg[off];
    float v = T.exp_avg_sq[off];
    float p = T.param[off];
    m = __fmaf_rn(T.one_minus_beta1, g, __fmul_rn(m, T.beta1));
    v = __fmaf_rn(__fmul_rn(T.one_minus_beta2, g), g, __fmul_rn(v, T.beta2));
    const float inv_bc2 = 1.0f / T.bias_correction2_sqrt;
    const float denom = __fadd_rn(__fmul_rn(sqrtf(v), inv_bc2), T.eps);
    p = __fmaf_rn(-T.step_size, m / denom, p);
    T.exp_avg[off] = m;
    T.exp_avg_sq[off] = v;
    T.param[off] = p;
  }
}

struct RowPack {
  H3dgsRowTensor t[8];
  int start[9];   // prefix sums of widths
  int n;
};

// one thread per packed float: element e of packed row k lives in tensor ti at row rows[k]
template <bool PACK>
__global__ void __launch_bounds__(256) pack_rows_kernel(const RowPack pk, const long long* __restrict__ rows,
                                                        long long K, float* __restrict__ packed, float scale) {
  const int total_w = pk.start[pk.n];
  const long long gid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= K * total_w) return;
  const long long k = gid / total_w;
  const int e = (int)(gid - k * total_w);
  int ti = 0;
#pragma unroll
  for (int i = 1; i < 8; ++i)
    if (i < pk.n && e >= pk.start[i]) ti = i;
  float* p = pk.t[ti].data + (size_t)rows[k] * pk.t[ti].width + (e - pk.start[ti]);
  if (PACK) packed[gid] = *p;
  else *p = packed[gid] * scale;
}

}  // namespace

cudaError_t run_pack_rows(const H3dgsRowTensor* tensors, int n, const long long* rows, long long K, float* packed,
                          float scale, bool pack, cudaStream_t s) {
  RowPack pk;
  pk.n = n;
  int acc = 0;
  for (int i = 0; i < n; ++i) {
    pk.t[i] = tensors[i];
    pk.start[i] = acc;
    acc += tensors[i].width;
  }
  for (int i = n; i <= 8; ++i) pk.start[i] = acc;
  const long long work = K * acc;
  if (work <= 0) return cudaSuccess;
  const long long blocks = (work + 255) / 256;
  ProfScope ps(K_ADAM, s);
  if (pack) pack_rows_kernel<true><<<(unsigned)blocks, 256, 0, s>>>(pk, rows, K, packed, scale);
  else pack_rows_kernel<false><<<(unsigned)blocks, 256, 0, s>>>(pk, rows, K, packed, scale);
  return cudaGetLastError();
}

cudaError_t run_sparse_adam_compact(const H3dgsAdamTensor* tensors, int n, const long long* rows, const int* count,
                                    int capacity, const float* gate, cudaStream_t s) {
  AdamPack pk;
  pk.n = n;
  int acc = 0;
  for (int i = 0; i < n; ++i) {
    pk.t[i] = tensors[i];
    pk.start[i] = acc;
    acc += tensors[i].width;
  }
  for (int i = n; i <= H3DGS_ADAM_MAX_TENSORS; ++i) pk.start[i] = acc;
  const long long work = (long long)capacity * acc;
  if (work <= 0) return cudaSuccess;
  // the grid covers the capacity; the kernel stops at the device-side row count
  long long blocks = (work + 255) / 256;
  if (blocks > 148 * 64) blocks = 148 * 64;
  ProfScope ps(K_ADAM, s);
  if (acc <= 64) {
    long long b2 = ((long long)capacity * 32 / ADAM_RB + 255) / 256;
    if (b2 > 148 * 16) b2 = 148 * 16;
    if (b2 < 1) b2 = 1;
    sparse_adam_compact_rows_kernel<<<(unsigned)b2, 256, 0, s>>>(pk, rows, count, gate);
  } else {
    sparse_adam_compact_kernel<<<(unsigned)blocks, 256, 0, s>>>(pk, rows, count, gate);
  }
  return cudaGetLastError();
}

cudaError_t run_sparse_adam(const H3dgsAdamTensor* tensors, int n, const long long* relevant,
                            long long num_relevant, long long num_rows, cudaStream_t s) {
  AdamPack pk;
  pk.n = n;
  int acc = 0;
  for (int i = 0; i < n; ++i) {
    pk.t[i] = tensors[i];
    pk.start[i] = acc;
    acc += tensors[i].width;
  }
  for (int i = n; i <= H3DGS_ADAM_MAX_TENSORS; ++i) pk.start[i] = acc;
  pk.start[n] = acc;
  const long long K = relevant ? num_relevant : num_rows;
  const long long work = K * acc;
  if (work <= 0) return cudaSuccess;
  const long long blocks = (work + 255) / 256;
  ProfScope ps(K_ADAM, s);
  sparse_adam_kernel<<<(unsigned)blocks, 256, 0, s>>>(pk, relevant, K);
  return cudaGetLastError();
}

}  // namespace h3dgs
