// Fused photometric loss: L1 + SSIM (11x11 Gaussian window) forward and backward.
// Replaces utils/loss_utils.py:17-63 (l1_loss, ssim / _ssim: five depthwise conv2d passes, ~15 elementwise
// kernels and their autograd) as used by train_single.py:116-141 and train_post.py:132-137.
//
// Forward, one pass over the two images: a 16x16 pixel tile (+5 halo, zero padded like conv2d(padding=5)) of both
// images is staged in shared memory, the five windowed moments (mu1, mu2, E[x^2], E[y^2], E[xy]) are formed
// separably (11 horizontal taps, then 11 vertical taps -- the 2-D window of create_window is the outer product
// of the 1-D one), the SSIM map value and its three partial derivatives with respect to (mu1, E[x^2], E[xy]) are
// computed per pixel, and the tile's sums of SSIM and |x - y| go to two double accumulators.
// Backward: d loss / d x = conv(g dm/dmu1) + 2 x conv(g dm/dE11) + y conv(g dm/dE12) + l1 term, the same zero-padded
// separable convolution applied to the three derivative maps (the window is symmetric).
#include <cstring>

#include "common.cuh"

namespace h3dgs {

namespace {

constexpr int LT = 16;          // tile edge
constexpr int LH = 5;           // halo (window_size // 2)
constexpr int LW = LT + 2 * LH; // 26

struct LossArgs {
  int C, H, W;
  const float* img1;      // rendered image [C,H,W]
  const float* img2;      // ground truth   [C,H,W]
  const float* mask;      // [H,W] multiplied into img1 (train_single.py:118-120 `image *= alpha_mask`), or NULL
  float w[11];            // normalised 1-D Gaussian window (utils/loss_utils.py:23-25)
  float* ssim_map;        // [C,H,W] or NULL
  float* dm_dmu1;         // [C,H,W] or NULL (no gradient wanted)
  float* dm_de11;
  float* dm_de12;
  double* sums;           // [0] += sum of the SSIM map, [1] += sum of |img1 - img2|
  // backward
  const float* dL_dmap;   // [C,H,W] per-pixel upstream gradient of the SSIM map, or NULL: uniform *ssim_scale
  const float* upstream;  // device scalar multiplied into both terms (d total / d loss), or NULL = 1
  float ssim_scale;       // d loss / d map value (uniform case), e.g. -lambda / (C H W)
  float l1_scale;         // d loss / d |x - y|, e.g. (1 - lambda) / (C H W); 0 = no L1 term
  float* dL_dimg1;        // [C,H,W]
};

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
  return v;
}

// Tiling of the two SSIM kernels: a CTA of 256 threads produces a 32 x 32 output tile from a 42 x 42 input patch
// (1.72 loads per pixel; 16 x 16 tiles needed 2.64).  Horizontal pass: a thread makes FOUR neighbouring outputs of a
// row from 14 inputs held in registers (7 shared-memory loads per output and quantity pair instead of 22; the
// squares and the product of an input are formed once, not once per tap).  Vertical pass: a thread makes four
// outputs of a column from 14 rows.  Every output is still the sum over k = 0..10 of w[k] * value[k], accumulated
// in that order with one FMA per tap: bit-identical to the one-output-per-thread form.
constexpr int ST = 32;             // tile edge
constexpr int SW = ST + 2 * LH;    // 42
constexpr int SP = SW + 1;         // row pitch of the input patch in shared memory (odd: conflict-free for both passes)
constexpr int SR = 4;              // outputs per thread and pass
constexpr int S_THREADS = 256;

template <bool GRAD>
__global__ void __launch_bounds__(S_THREADS) ssim_fwd_kernel(const LossArgs a) {
  __shared__ float s1[SW][SP], s2[SW][SP];
  __shared__ float hz[5][SW][ST + 1];   // +1: the four rows a warp writes in the horizontal pass fall into different banks
  __shared__ float red[2][S_THREADS / 32];
  const int tid = threadIdx.x;
  const int x0 = blockIdx.x * ST, y0 = blockIdx.y * ST, ch = blockIdx.z;
  const size_t plane = (size_t)a.H * a.W;
  const float* __restrict__ p1 = a.img1 + ch * plane;
  const float* __restrict__ p2 = a.img2 + ch * plane;

  for (int i = tid; i < SW * SW; i += S_THREADS) {
    const int ly = i / SW, lx = i - ly * SW;
    const int gy = y0 + ly - LH, gx = x0 + lx - LH;
    float v1 = 0.f, v2 = 0.f;
    if (gy >= 0 && gy < a.H && gx >= 0 && gx < a.W) {
      v1 = __ldg(p1 + (size_t)gy * a.W + gx);
      v2 = __ldg(p2 + (size_t)gy * a.W + gx);
      if (a.mask) v1 *= __ldg(a.mask + (size_t)gy * a.W + gx);
    }
    s1[ly][lx] = v1;
    s2[ly][lx] = v2;
  }
  __syncthreads();
  // horizontal: task = (row of the patch, group of four output columns)
  for (int i = tid; i < SW * (ST / SR); i += S_THREADS) {
    const int row = i / (ST / SR), col = (i - row * (ST / SR)) * SR;
    float u[SR + 10], v[SR + 10], uu[SR + 10], vv[SR + 10], uv[SR + 10];
#pragma unroll
    for (int j = 0; j < SR + 10; ++j) {
      u[j] = s1[row][col + j];
      v[j] = s2[row][col + j];
      uu[j] = u[j] * u[j];
      vv[j] = v[j] * v[j];
      uv[j] = u[j] * v[j];
    }
#pragma unroll
    for (int o = 0; o < SR; ++o) {
      float m1 = 0.f, m2 = 0.f, e11 = 0.f, e22 = 0.f, e12 = 0.f;
#pragma unroll
      for (int k = 0; k < 11; ++k) {
        const float wk = a.w[k];
        m1 += wk * u[o + k];
        m2 += wk * v[o + k];
        e11 += wk * uu[o + k];
        e22 += wk * vv[o + k];
        e12 += wk * uv[o + k];
      }
      hz[0][row][col + o] = m1; hz[1][row][col + o] = m2; hz[2][row][col + o] = e11; hz[3][row][col + o] = e22;
      hz[4][row][col + o] = e12;
    }
  }
  __syncthreads();
  // vertical: thread = (column, group of four output rows)
  const int tx = tid & (ST - 1), r0 = (tid / ST) * SR;
  float acc[5][SR];
#pragma unroll
  for (int q = 0; q < 5; ++q) {
    float h[SR + 10];
#pragma unroll
    for (int j = 0; j < SR + 10; ++j) h[j] = hz[q][r0 + j][tx];
#pragma unroll
    for (int o = 0; o < SR; ++o) {
      float t = 0.f;
#pragma unroll
      for (int k = 0; k < 11; ++k) t += a.w[k] * h[o + k];
      acc[q][o] = t;
    }
  }
  const int gx = x0 + tx;
  float sum_map = 0.f, sum_ad = 0.f;
#pragma unroll
  for (int o = 0; o < SR; ++o) {
    const int ty = r0 + o, gy = y0 + ty;
    if (gx < a.W && gy < a.H) {
      // utils/loss_utils.py:43-56
      const float mu1 = acc[0][o], mu2 = acc[1][o], e11 = acc[2][o], e22 = acc[3][o], e12 = acc[4][o];
      const float C1 = 0.01f * 0.01f, C2 = 0.03f * 0.03f;
      const float mu1_sq = mu1 * mu1, mu2_sq = mu2 * mu2, mu1_mu2 = mu1 * mu2;
      const float sigma1_sq = e11 - mu1_sq, sigma2_sq = e22 - mu2_sq, sigma12 = e12 - mu1_mu2;
      const float A = 2.f * mu1_mu2 + C1, B = 2.f * sigma12 + C2;
      const float Cc = mu1_sq + mu2_sq + C1, D = sigma1_sq + sigma2_sq + C2;
      const float inv = 1.f / (Cc * D);
      const float map = (A * B) * inv;
      const size_t oo = ch * plane + (size_t)gy * a.W + gx;
      if (a.ssim_map) a.ssim_map[oo] = map;
      sum_map += map;
      sum_ad += fabsf(s1[ty + LH][tx + LH] - s2[ty + LH][tx + LH]);
      if (GRAD) {
        // with sigma1_sq = E11 - mu1^2 and sigma12 = E12 - mu1 mu2 substituted
        a.dm_dmu1[oo] = 2.f * mu2 * (B - A) * inv - map * (2.f * mu1 * (D - Cc)) * inv;
        a.dm_de11[oo] = -map / D;
        a.dm_de12[oo] = 2.f * A * inv;
      }
    }
  }
  const float sm = warp_sum(sum_map), sa = warp_sum(sum_ad);
  if ((tid & 31) == 0) { red[0][tid >> 5] = sm; red[1][tid >> 5] = sa; }
  __syncthreads();
  if (tid < 2) {
    double t = 0.0;
    for (int i = 0; i < S_THREADS / 32; ++i) t += (double)red[tid][i];
    atomicAdd(a.sums + tid, t);
  }
}

__global__ void __launch_bounds__(S_THREADS) ssim_bwd_kernel(const LossArgs a) {
  __shared__ float s[3][SW][SP];
  __shared__ float hz[3][SW][ST + 1];
  const int tid = threadIdx.x;
  const int x0 = blockIdx.x * ST, y0 = blockIdx.y * ST, ch = blockIdx.z;
  const size_t plane = (size_t)a.H * a.W;
  const size_t base = ch * plane;
  for (int i = tid; i < SW * SW; i += S_THREADS) {
    const int ly = i / SW, lx = i - ly * SW;
    const int gy = y0 + ly - LH, gx = x0 + lx - LH;
    float v0 = 0.f, v1 = 0.f, v2 = 0.f;
    if (gy >= 0 && gy < a.H && gx >= 0 && gx < a.W) {
      const size_t o = base + (size_t)gy * a.W + gx;
      const float g = a.dL_dmap ? __ldg(a.dL_dmap + o) : 1.0f;
      v0 = g * __ldg(a.dm_dmu1 + o);
      v1 = g * __ldg(a.dm_de11 + o);
      v2 = g * __ldg(a.dm_de12 + o);
    }
    s[0][ly][lx] = v0; s[1][ly][lx] = v1; s[2][ly][lx] = v2;
  }
  __syncthreads();
  for (int i = tid; i < SW * (ST / SR); i += S_THREADS) {
    const int row = i / (ST / SR), col = (i - row * (ST / SR)) * SR;
#pragma unroll
    for (int q = 0; q < 3; ++q) {
      float h[SR + 10];
#pragma unroll
      for (int j = 0; j < SR + 10; ++j) h[j] = s[q][row][col + j];
#pragma unroll
      for (int o = 0; o < SR; ++o) {
        float c = 0.f;
#pragma unroll
        for (int k = 0; k < 11; ++k) c += a.w[k] * h[o + k];
        hz[q][row][col + o] = c;
      }
    }
  }
  __syncthreads();
  const int tx = tid & (ST - 1), r0 = (tid / ST) * SR;
  float acc[3][SR];
#pragma unroll
  for (int q = 0; q < 3; ++q) {
    float h[SR + 10];
#pragma unroll
    for (int j = 0; j < SR + 10; ++j) h[j] = hz[q][r0 + j][tx];
#pragma unroll
    for (int o = 0; o < SR; ++o) {
      float c = 0.f;
#pragma unroll
      for (int k = 0; k < 11; ++k) c += a.w[k] * h[o + k];
      acc[q][o] = c;
    }
  }
  const int gx = x0 + tx;
  const float up0 = a.upstream ? __ldg(a.upstream) : 1.0f;
#pragma unroll
  for (int o = 0; o < SR; ++o) {
    const int gy = y0 + r0 + o;
    if (gx >= a.W || gy >= a.H) continue;
    const float c0 = acc[0][o], c1 = acc[1][o], c2 = acc[2][o];
    const size_t oo = base + (size_t)gy * a.W + gx;
    const float m = a.mask ? __ldg(a.mask + (size_t)gy * a.W + gx) : 1.0f;
    const float x = __ldg(a.img1 + oo) * m, y = __ldg(a.img2 + oo);
    const float up = up0 * m;   // d (img1 m) / d img1 = m
    float g = (c0 + 2.f * x * c1 + y * c2) * (a.dL_dmap ? 1.0f : a.ssim_scale);
    if (a.l1_scale != 0.f) {
      const float d = x - y;
      g += a.l1_scale * (d > 0.f ? 1.f : (d < 0.f ? -1.f : 0.f));   // torch.abs backward: sign(x - y)
    }
    a.dL_dimg1[oo] = g * up;
  }
}

// Depth loss of train_single.py:133-139: mean |(pred - target) mask| over n elements.  Forward adds the sum to
// *sum; backward writes upstream * scale * sign((pred - target) mask) * mask.
__global__ void __launch_bounds__(256) masked_l1_fwd_kernel(long long n, const float* __restrict__ pred,
                                                            const float* __restrict__ target,
                                                            const float* __restrict__ mask, double* sum) {
  __shared__ float red[8];
  float acc = 0.f;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    float d = __ldg(pred + i) - __ldg(target + i);
    if (mask) d *= __ldg(mask + i);
    acc += fabsf(d);
  }
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
    for (int i = 0; i < 8; ++i) t += (double)red[i];
    atomicAdd(sum, t);
  }
}

__global__ void __launch_bounds__(256) masked_l1_bwd_kernel(long long n, const float* __restrict__ pred,
                                                            const float* __restrict__ target,
                                                            const float* __restrict__ mask,
                                                            const float* __restrict__ upstream, float scale,
                                                            float* __restrict__ grad) {
  const float up = (upstream ? __ldg(upstream) : 1.0f) * scale;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const float m = mask ? __ldg(mask + i) : 1.0f;
    const float d = (__ldg(pred + i) - __ldg(target + i)) * m;
    grad[i] = up * m * (d > 0.f ? 1.f : (d < 0.f ? -1.f : 0.f));
  }
}

// Exposure affine + clamp of the render glue (gaussian_renderer/__init__.py:115-118, :277-281):
//   out[c] = clamp(sum_k in[k] E[k][c] + E[c][3], 0, 1)        (matmul(image.permute(1,2,0), E[:3,:3]) + E[:3,3])
// exposure: device [3,4] row-major, or NULL for the clamp alone.  One thread per pixel, all three channels.
struct Expo {
  float m[3][3];   // m[k][c] = E[k][c]
  float t[3];      // t[c] = E[c][3]
};
__device__ __forceinline__ Expo load_expo(const float* __restrict__ e) {
  Expo x;
#pragma unroll
  for (int k = 0; k < 3; ++k) {
#pragma unroll
    for (int c = 0; c < 3; ++c) x.m[k][c] = e ? __ldg(e + 4 * k + c) : (k == c ? 1.0f : 0.0f);
    x.t[k] = e ? __ldg(e + 4 * k + 3) : 0.0f;
  }
  return x;
}
// torch.matmul of a [.., 3] row with a 3x3 matrix: fp32 dot product, k ascending, fused multiply-adds (cuBLAS sgemv /
// gemm inner loop); then the offset as a separate add
__device__ __forceinline__ float expo_apply(const Expo& x, int c, float r0, float r1, float r2) {
  return __fadd_rn(fmaf(r2, x.m[2][c], fmaf(r1, x.m[1][c], __fmul_rn(r0, x.m[0][c]))), x.t[c]);
}

__global__ void __launch_bounds__(256) exposure_clamp_fwd_kernel(long long npix, const float* __restrict__ in,
                                                                const float* __restrict__ exposure,
                                                                float* __restrict__ out) {
  const Expo x = load_expo(exposure);
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < npix; i += (long long)gridDim.x * blockDim.x) {
    const float r0 = __ldg(in + i), r1 = __ldg(in + npix + i), r2 = __ldg(in + 2 * npix + i);
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const float v = exposure ? expo_apply(x, c, r0, r1, r2) : (c == 0 ? r0 : c == 1 ? r1 : r2);
      out[c * npix + i] = fminf(fmaxf(v, 0.0f), 1.0f);
    }
  }
}

// grad_in[k] = sum_c gate_c grad_out[c] E[k][c];  grad_E[k][c] += sum_pix gate_c grad_out[c] in[k],
// grad_E[c][3] += sum_pix gate_c grad_out[c];  gate_c = 1 where the unclamped value lies in [0, 1] (torch's clamp
// backward passes the gradient at the bounds themselves)
__global__ void __launch_bounds__(256) exposure_clamp_bwd_kernel(long long npix, const float* __restrict__ in,
                                                                const float* __restrict__ exposure,
                                                                const float* __restrict__ grad_out,
                                                                float* __restrict__ grad_in,
                                                                float* __restrict__ grad_exposure) {
  __shared__ float red[8][12];
  const Expo x = load_expo(exposure);
  float acc[12];
#pragma unroll
  for (int j = 0; j < 12; ++j) acc[j] = 0.0f;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < npix; i += (long long)gridDim.x * blockDim.x) {
    const float r[3] = {__ldg(in + i), __ldg(in + npix + i), __ldg(in + 2 * npix + i)};
    float g[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const float v = exposure ? expo_apply(x, c, r[0], r[1], r[2]) : r[c];
      g[c] = (v >= 0.0f && v <= 1.0f) ? __ldg(grad_out + c * npix + i) : 0.0f;
    }
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      grad_in[k * npix + i] = exposure ? fmaf(g[2], x.m[k][2], fmaf(g[1], x.m[k][1], g[0] * x.m[k][0])) : g[k];
      if (grad_exposure) {
#pragma unroll
        for (int c = 0; c < 3; ++c) acc[4 * k + c] = fmaf(g[c], r[k], acc[4 * k + c]);
        acc[4 * k + 3] += g[k];
      }
    }
  }
  if (grad_exposure == nullptr) return;
#pragma unroll
  for (int j = 0; j < 12; ++j) {
    const float v = warp_sum(acc[j]);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5][j] = v;
  }
  __syncthreads();
  if (threadIdx.x < 12) {
    float t = 0.0f;
    for (int w = 0; w < 8; ++w) t += red[w][threadIdx.x];
    atomicAdd(grad_exposure + threadIdx.x, t);
  }
}

LossArgs make_args(int C, int H, int W, const float* img1, const float* img2, const float* mask,
                   const float* window) {
  LossArgs a;
  std::memset(&a, 0, sizeof(a));
  a.C = C; a.H = H; a.W = W; a.img1 = img1; a.img2 = img2; a.mask = mask;
  for (int i = 0; i < 11; ++i) a.w[i] = window[i];
  return a;
}

}  // namespace

cudaError_t run_ssim_forward(int C, int H, int W, const float* img1, const float* img2, const float* mask,
                             const float* window, float* ssim_map, float* dm_dmu1, float* dm_de11, float* dm_de12, double* sums,
                             cudaStream_t s) {
  LossArgs a = make_args(C, H, W, img1, img2, mask, window);
  a.ssim_map = ssim_map; a.dm_dmu1 = dm_dmu1; a.dm_de11 = dm_de11; a.dm_de12 = dm_de12; a.sums = sums;
  const dim3 grid((W + ST - 1) / ST, (H + ST - 1) / ST, C), block(S_THREADS);
  ProfScope ps(K_LOSS, s);
  if (dm_dmu1) ssim_fwd_kernel<true><<<grid, block, 0, s>>>(a);
  else ssim_fwd_kernel<false><<<grid, block, 0, s>>>(a);
  return cudaGetLastError();
}

cudaError_t run_ssim_backward(int C, int H, int W, const float* img1, const float* img2, const float* mask,
                              const float* window, const float* dm_dmu1, const float* dm_de11, const float* dm_de12,
                              const float* dL_dmap, const float* upstream, float ssim_scale, float l1_scale,
                              float* dL_dimg1, cudaStream_t s) {
  LossArgs a = make_args(C, H, W, img1, img2, mask, window);
  a.dm_dmu1 = const_cast<float*>(dm_dmu1); a.dm_de11 = const_cast<float*>(dm_de11);
  a.dm_de12 = const_cast<float*>(dm_de12);
  a.dL_dmap = dL_dmap; a.upstream = upstream; a.ssim_scale = ssim_scale; a.l1_scale = l1_scale;
  a.dL_dimg1 = dL_dimg1;
  const dim3 grid((W + ST - 1) / ST, (H + ST - 1) / ST, C), block(S_THREADS);
  ProfScope ps(K_LOSS, s);
  ssim_bwd_kernel<<<grid, block, 0, s>>>(a);
  return cudaGetLastError();
}

cudaError_t run_masked_l1(long long n, const float* pred, const float* target, const float* mask, double* sum,
                          const float* upstream, float scale, float* grad, cudaStream_t s) {
  if (n <= 0) return cudaSuccess;
  long long blocks = (n + 256 * 8 - 1) / (256 * 8);
  if (blocks > 148 * 8) blocks = 148 * 8;
  ProfScope ps(K_LOSS, s);
  if (grad) masked_l1_bwd_kernel<<<(unsigned)blocks, 256, 0, s>>>(n, pred, target, mask, upstream, scale, grad);
  else masked_l1_fwd_kernel<<<(unsigned)blocks, 256, 0, s>>>(n, pred, target, mask, sum);
  return cudaGetLastError();
}

cudaError_t run_exposure_clamp(long long npix, const float* in, const float* exposure, float* out,
                               const float* grad_out, float* grad_in, float* grad_exposure, cudaStream_t s) {
  if (npix <= 0) return cudaSuccess;
  long long blocks = (npix + 256 * 4 - 1) / (256 * 4);
  if (blocks > 148 * 8) blocks = 148 * 8;
  ProfScope ps(K_LOSS, s);
  if (grad_in) exposure_clamp_bwd_kernel<<<(unsigned)blocks, 256, 0, s>>>(npix, in, exposure, grad_out, grad_in, grad_exposure);
  else exposure_clamp_fwd_kernel<<<(unsigned)blocks, 256, 0, s>>>(npix, in, exposure, out);
  return cudaGetLastError();
}

}  // namespace h3dgs

extern "C" {

int h3dgs_exposure_clamp_forward(int32_t H, int32_t W, const float* image, const float* exposure, float* out,
                                 h3dgs_stream_t stream) {
  using namespace h3dgs;
  if (H < 0 || W < 0 || !image || !out) return fail(H3DGS_EINVAL, "h3dgs_exposure_clamp_forward: bad argument");
  cudaError_t e = run_exposure_clamp((long long)H * W, image, exposure, out, nullptr, nullptr, nullptr,
                                     reinterpret_cast<cudaStream_t>(stream));
  return e == cudaSuccess ? 0 : cuda_fail(e, "h3dgs_exposure_clamp_forward");
}

int h3dgs_exposure_clamp_backward(int32_t H, int32_t W, const float* image, const float* exposure,
                                  const float* grad_out, float* grad_image, float* grad_exposure,
                                  h3dgs_stream_t stream) {
  using namespace h3dgs;
  if (H < 0 || W < 0 || !image || !grad_out || !grad_image || (grad_exposure && !exposure))
    return fail(H3DGS_EINVAL, "h3dgs_exposure_clamp_backward: bad argument");
  cudaError_t e = run_exposure_clamp((long long)H * W, image, exposure, nullptr, grad_out, grad_image, grad_exposure,
                                     reinterpret_cast<cudaStream_t>(stream));
  return e == cudaSuccess ? 0 : cuda_fail(e, "h3dgs_exposure_clamp_backward");
}

}  // extern "C"
