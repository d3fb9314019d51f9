// Measured pipe ceilings for the roofline of the blend kernels ("FP32/MUFU ceilings: measure with an
// FMA / ex2 micro-kernel").  Three dependent-chain micro-kernels, every SM full (8 CTAs x 256 threads per SM, 8
// independent chains per thread), timed with CUDA events on the caller's stream:
//   [0] FFMA, three-register form            -> TFLOP/s (1 FFMA = 2 FLOP): the FP32 pipe as compiled code uses it
//   [1] packed fma.rn.f32x2 (FFMA2, sm_100+) -> TFLOP/s
//   [2] MUFU.EX2 (ex2.approx.ftz.f32)        -> G results/s
//   [3] warp instructions issued per second in kernel [0] (its loop is FFMA only): the issue ceiling seen by a
//       kernel whose hot loop is fma-pipe work, in G warp-instr/s
#include "common.cuh"

namespace h3dgs {
namespace {

constexpr int PK_ITERS = 2048;
constexpr int PK_CHAINS = 8;

__global__ void __launch_bounds__(256) peak_ffma_kernel(float* out, float b, float c) {
  float a[PK_CHAINS];
#pragma unroll
  for (int k = 0; k < PK_CHAINS; ++k) a[k] = (float)(threadIdx.x + k) * 1e-3f;
  for (int i = 0; i < PK_ITERS; ++i) {
#pragma unroll
    for (int k = 0; k < PK_CHAINS; ++k) a[k] = fmaf(a[k], b, c);
  }
  float s = 0.f;
#pragma unroll
  for (int k = 0; k < PK_CHAINS; ++k) s += a[k];
  if (s == 12345.678f) out[0] = s;   // never true: keeps the chains alive
}

__global__ void __launch_bounds__(256) peak_ffma2_kernel(float* out, float b, float c) {
  unsigned long long a[PK_CHAINS];
  unsigned long long bb, cc;
  asm("mov.b64 %0, {%1, %1};" : "=l"(bb) : "f"(b));
  asm("mov.b64 %0, {%1, %1};" : "=l"(cc) : "f"(c));
#pragma unroll
  for (int k = 0; k < PK_CHAINS; ++k) {
    const float x = (float)(threadIdx.x + k) * 1e-3f;
    asm("mov.b64 %0, {%1, %2};" : "=l"(a[k]) : "f"(x), "f"(x + 0.5f));
  }
  for (int i = 0; i < PK_ITERS; ++i) {
#pragma unroll
    for (int k = 0; k < PK_CHAINS; ++k) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(a[k]) : "l"(bb), "l"(cc));
  }
  float s = 0.f;
#pragma unroll
  for (int k = 0; k < PK_CHAINS; ++k) {
    float lo, hi;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(a[k]));
    s += lo + hi;
  }
  if (s == 12345.678f) out[0] = s;
}

__global__ void __launch_bounds__(256) peak_mufu_kernel(float* out, float c) {
  float a[PK_CHAINS];
#pragma unroll
  for (int k = 0; k < PK_CHAINS; ++k) a[k] = (float)(threadIdx.x + k) * 1e-3f + c;
  for (int i = 0; i < PK_ITERS; ++i) {
#pragma unroll
    for (int k = 0; k < PK_CHAINS; ++k) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(a[k]));
  }
  float s = 0.f;
#pragma unroll
  for (int k = 0; k < PK_CHAINS; ++k) s += a[k];
  if (s == 12345.678f) out[0] = s;
}

}  // namespace

// out[4] as documented above; best of `reps` launches each (after one warm-up launch)
cudaError_t run_measure_peaks(double* out, float* scratch, int reps, cudaStream_t s) {
  int dev = 0, sms = 148;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int grid = sms * 8;
  cudaEvent_t a, b;
  if ((e = cudaEventCreate(&a)) != cudaSuccess) return e;
  if ((e = cudaEventCreate(&b)) != cudaSuccess) return e;
  const double lanes = (double)grid * 256.0 * PK_ITERS * PK_CHAINS;   // lane-operations per launch
  double best[3] = {1e30, 1e30, 1e30};
  for (int k = 0; k < 3; ++k) {
    for (int r = 0; r < reps + 1; ++r) {
      cudaEventRecord(a, s);
      if (k == 0) peak_ffma_kernel<<<grid, 256, 0, s>>>(scratch, 0.999f, 1e-3f);
      else if (k == 1) peak_ffma2_kernel<<<grid, 256, 0, s>>>(scratch, 0.999f, 1e-3f);
      else peak_mufu_kernel<<<grid, 256, 0, s>>>(scratch, -20.0f);
      cudaEventRecord(b, s);
      if ((e = cudaEventSynchronize(b)) != cudaSuccess) return e;
      float ms = 0.f;
      cudaEventElapsedTime(&ms, a, b);
      if (r > 0 && ms < best[k]) best[k] = ms;
    }
  }
  cudaEventDestroy(a);
  cudaEventDestroy(b);
  out[0] = lanes * 2.0 / (best[0] * 1e-3) / 1e12;
  out[1] = lanes * 4.0 / (best[1] * 1e-3) / 1e12;
  out[2] = lanes / (best[2] * 1e-3) / 1e9;
  out[3] = lanes / 32.0 / (best[0] * 1e-3) / 1e9;
  return cudaGetLastError();
}

}  // namespace h3dgs
