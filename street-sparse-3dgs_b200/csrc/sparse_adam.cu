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

// compact gradients: row i of every grad tensor belongs to store row rows[i]; the row count lives on the device
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
  sparse_adam_compact_kernel<<<(unsigned)blocks, 256, 0, s>>>(pk, rows, count, gate);
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
