// Densification statistics, selection and rebuild.  Replaces, for single-chunk training
// (train_single.py:191-197):
//   stats   gaussians.max_radii2D[vis] = max(max_radii2D[vis], radii); add_densification_stats
//           (scene/gaussian_model.py:780-794): three gathers, a norm, two maxima and three scatters per iteration
//   select  the masks of densify_and_clone / densify_and_split / the opacity prune (scene/gaussian_model.py:672-760)
//   build   cat_tensors_to_optimizer + prune_points x 2 (scene/gaussian_model.py:560-670): the reference rebuilds all six
//           parameter tensors and their twelve Adam moment tensors three times (clone: cat; split: cat, then a masked
//           copy; prune: a masked copy), ~60 N-row passes.  Here every surviving, cloned or split row is written once
//           into its final place: [kept rows | clones | first split children | second split children], exactly the row
//           order the reference ends with.
// Row widths are the model's: xyz 3, f_dc 3, f_rest 45 (up to 64 per tensor here), opacity 1, scaling 3, rotation 4.
#include "common.cuh"

namespace h3dgs {

namespace {

constexpr int DN_CLONE = 1, DN_SPLIT = 2, DN_LOW = 4;

__device__ __forceinline__ float act_opacity(float raw, int act) {
  if (act == H3DGS_ACT_SIGMOID) return 1.0f / (1.0f + expf(-raw));
  if (act == H3DGS_ACT_ABS) return fabsf(raw);
  return raw;
}

// one thread per visible row: train_single.py:193 and gaussian_model.py:780-794 (is_depth_only multiplies by 1: no-op)
__global__ void __launch_bounds__(256) densify_stats_kernel(long long K, const long long* __restrict__ rows,
                                                           const int* __restrict__ radii,
                                                           const float* __restrict__ vs_grad, float* __restrict__ accum,
                                                           float* __restrict__ denom, float* __restrict__ max_radii) {
  const long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= K) return;
  const long long r = rows[k];
  if (max_radii && radii) max_radii[r] = fmaxf(max_radii[r], (float)radii[k]);
  const float gx = vs_grad[3 * r], gy = vs_grad[3 * r + 1];
  const float norm = sqrtf(__fadd_rn(__fmul_rn(gx, gx), __fmul_rn(gy, gy)));
  const float a = accum[r];
  // torch.max propagates NaN from either side
  accum[r] = (norm != norm || a != a) ? __int_as_float(0x7fc00000) : fmaxf(norm, a);
  denom[r] += 1.0f;
}

// per-row decisions (gaussian_model.py:676-685, 714-722, 752-755) and the three 0/1 counts the scans run over
__global__ void __launch_bounds__(256) densify_select_kernel(int N, const float* __restrict__ accum,
                                                            const float* __restrict__ max_radii,
                                                            const float* __restrict__ opacity_raw, int opacity_act,
                                                            const float* __restrict__ scaling_raw, float grad_threshold,
                                                            float min_opacity, float dense_extent, int scaffold,
                                                            uint8_t* __restrict__ flags, uint32_t* __restrict__ c_keep,
                                                            uint32_t* __restrict__ c_clone, uint32_t* __restrict__ c_split) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N) return;
  float g = accum[i];
  if (g != g) g = 0.0f;                                        // grads[grads.isnan()] = 0.0
  const float op = act_opacity(opacity_raw[i], opacity_act);
  const float smax = fmaxf(fmaxf(expf(scaling_raw[3 * i]), expf(scaling_raw[3 * i + 1])), expf(scaling_raw[3 * i + 2]));
  const float crit = __fmul_rn(__fmul_rn(g, max_radii[i]), powf(op, 0.2f));
  const bool free_row = i >= scaffold;                         // no densification / pruning of the scaffold
  const bool base = crit >= grad_threshold && op > 0.15f && free_row;
  const bool clone = base && smax <= dense_extent;
  const bool split = base && smax > dense_extent;
  const bool low = op < min_opacity && free_row;
  flags[i] = (uint8_t)((clone ? DN_CLONE : 0) | (split ? DN_SPLIT : 0) | (low ? DN_LOW : 0));
  c_keep[i] = (!split && !low) ? 1u : 0u;
  c_clone[i] = (clone && !low) ? 1u : 0u;
  c_split[i] = (split && !low) ? 1u : 0u;
}

struct DnTensors {
  const float* src[H3DGS_ADAM_MAX_TENSORS];
  const float* m_src[H3DGS_ADAM_MAX_TENSORS];   // exp_avg or NULL
  const float* v_src[H3DGS_ADAM_MAX_TENSORS];   // exp_avg_sq or NULL
  float* dst[H3DGS_ADAM_MAX_TENSORS];
  float* m_dst[H3DGS_ADAM_MAX_TENSORS];
  float* v_dst[H3DGS_ADAM_MAX_TENSORS];
  int width[H3DGS_ADAM_MAX_TENSORS];
  int n;
  int xyz, scaling, rotation;   // tensor indices with special treatment in split children
};

// one warp per old row; lanes cover the floats of a row tensor by tensor
__global__ void __launch_bounds__(256) densify_build_kernel(int N, const DnTensors t, const uint8_t* __restrict__ flags,
                                                           const uint32_t* __restrict__ s_keep,
                                                           const uint32_t* __restrict__ s_clone,
                                                           const uint32_t* __restrict__ s_split, int K, int C, int S,
                                                           const float* __restrict__ samples /*[2S,3]*/) {
  const int i = (int)(((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5);
  const int lane = threadIdx.x & 31;
  if (i >= N) return;
  const int f = flags[i];
  const bool low = f & DN_LOW;
  if (low) return;                                             // the row and everything made from it is pruned
  const bool split = f & DN_SPLIT, clone = f & DN_CLONE;
  if (!split) {
    const size_t d0 = s_keep[i] - 1;                           // inclusive scans
    const size_t d1 = clone ? (size_t)K + s_clone[i] - 1 : 0;
    for (int k = 0; k < t.n; ++k) {
      const int w = t.width[k];
      for (int e = lane; e < w; e += 32) {
        const float p = t.src[k][(size_t)i * w + e];
        t.dst[k][d0 * w + e] = p;
        if (t.m_dst[k]) {
          t.m_dst[k][d0 * w + e] = t.m_src[k][(size_t)i * w + e];
          t.v_dst[k][d0 * w + e] = t.v_src[k][(size_t)i * w + e];
        }
        if (clone) {                                           // densify_and_clone: a copy, fresh moments
          t.dst[k][d1 * w + e] = p;
          if (t.m_dst[k]) { t.m_dst[k][d1 * w + e] = 0.0f; t.v_dst[k][d1 * w + e] = 0.0f; }
        }
      }
    }
    return;
  }
  // densify_and_split (gaussian_model.py:687-697): two children, position = R sample + xyz with R from the normalised
  // raw rotation (utils/general_utils.py:82-103), scaling = log(exp(s) / (0.8 * 2)), everything else copied
  const int rk = (int)s_split[i] - 1;
  const float* q = t.src[t.rotation] + (size_t)i * 4;
  const float q0 = q[0], q1 = q[1], q2 = q[2], q3 = q[3];
  const float norm = sqrtf(q0 * q0 + q1 * q1 + q2 * q2 + q3 * q3);
  const float r = q0 / norm, x = q1 / norm, y = q2 / norm, z = q3 / norm;
  const float R[3][3] = {{1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y)},
                         {2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x)},
                         {2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y)}};
  for (int j = 0; j < 2; ++j) {
    const size_t d = (size_t)K + C + (size_t)j * S + rk;
    const float* sm = samples + ((size_t)j * S + rk) * 3;
    for (int k = 0; k < t.n; ++k) {
      const int w = t.width[k];
      for (int e = lane; e < w; e += 32) {
        float p = t.src[k][(size_t)i * w + e];
        if (k == t.xyz) p = fmaf(R[e][2], sm[2], fmaf(R[e][1], sm[1], R[e][0] * sm[0])) + p;
        else if (k == t.scaling) p = logf(expf(p) * (1.0f / 1.6f));
        t.dst[k][d * w + e] = p;
        if (t.m_dst[k]) { t.m_dst[k][d * w + e] = 0.0f; t.v_dst[k][d * w + e] = 0.0f; }
      }
    }
  }
}

// std of the split samples: get_scaling[selected].repeat(2, 1) (gaussian_model.py:687)
__global__ void __launch_bounds__(256) densify_split_std_kernel(int N, const uint8_t* __restrict__ flags,
                                                               const uint32_t* __restrict__ s_split, int S,
                                                               const float* __restrict__ scaling_raw,
                                                               float* __restrict__ stds /*[2S,3]*/) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N) return;
  const int f = flags[i];
  if (!(f & DN_SPLIT) || (f & DN_LOW)) return;
  const int rk = (int)s_split[i] - 1;
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    const float s = expf(scaling_raw[3 * i + a]);
    stds[(size_t)rk * 3 + a] = s;
    stds[((size_t)S + rk) * 3 + a] = s;
  }
}

struct DnWs {
  uint8_t* flags;
  uint32_t *c_keep, *c_clone, *c_split, *s_keep, *s_clone, *s_split;
  unsigned long long* status;
  uint32_t* counters;
  size_t bytes;
  static DnWs make(void* base, long long N) {
    DnWs w;
    char* p = reinterpret_cast<char*>(base);
    char* p0 = p;
    const size_t n = (size_t)(N > 0 ? N : 1);
    carve(p, w.flags, n);
    carve(p, w.c_keep, n); carve(p, w.c_clone, n); carve(p, w.c_split, n);
    carve(p, w.s_keep, n); carve(p, w.s_clone, n); carve(p, w.s_split, n);
    carve(p, w.status, 3 * ((n + SCAN_TILE - 1) / SCAN_TILE + 1));
    carve(p, w.counters, 3 * 8);
    w.bytes = (size_t)(p - p0) + 256;
    return w;
  }
};

}  // namespace

}  // namespace h3dgs

extern "C" {

int h3dgs_densify_stats(int64_t num_visible, const int64_t* visible_rows, const int32_t* radii,
                        const float* viewspace_grad, float* xyz_gradient_accum, float* denom, float* max_radii2D,
                        h3dgs_stream_t stream) {
  using namespace h3dgs;
  if (num_visible < 0 || (num_visible > 0 && (!visible_rows || !viewspace_grad || !xyz_gradient_accum || !denom)))
    return fail(H3DGS_EINVAL, "h3dgs_densify_stats: bad argument");
  if (num_visible == 0) return 0;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  ProfScope ps(K_DENSIFY, s);
  densify_stats_kernel<<<(unsigned)((num_visible + 255) / 256), 256, 0, s>>>(
      num_visible, reinterpret_cast<const long long*>(visible_rows), radii, viewspace_grad, xyz_gradient_accum, denom,
      max_radii2D);
  H3_CUDA(cudaGetLastError());
  return 0;
}

size_t h3dgs_densify_workspace_bytes(int64_t num_rows) { return h3dgs::DnWs::make(nullptr, num_rows).bytes; }

int h3dgs_densify_select(int32_t num_rows, const float* xyz_gradient_accum, const float* max_radii2D,
                         const float* opacity_raw, int32_t opacity_activation, const float* scaling_raw,
                         float grad_threshold, float min_opacity, float dense_extent, int32_t scaffold_points,
                         void* workspace, size_t workspace_bytes, int32_t* counts_host /*[3]: kept, clones, splits*/,
                         h3dgs_stream_t stream) {
  using namespace h3dgs;
  if (num_rows < 0 || !counts_host) return fail(H3DGS_EINVAL, "h3dgs_densify_select: bad argument");
  counts_host[0] = counts_host[1] = counts_host[2] = 0;
  if (num_rows == 0) return 0;
  if (!xyz_gradient_accum || !max_radii2D || !opacity_raw || !scaling_raw || !workspace)
    return fail(H3DGS_EINVAL, "h3dgs_densify_select: NULL argument");
  if (workspace_bytes < DnWs::make(nullptr, num_rows).bytes) return fail(H3DGS_ENOSPACE, "h3dgs_densify_select: workspace too small");
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  DnWs w = DnWs::make(workspace, num_rows);
  const size_t tiles = ((size_t)num_rows + SCAN_TILE - 1) / SCAN_TILE + 1;
  H3_CUDA(cudaMemsetAsync(w.status, 0, 3 * tiles * sizeof(unsigned long long), s));
  H3_CUDA(cudaMemsetAsync(w.counters, 0, 3 * 8 * sizeof(uint32_t), s));
  {
    ProfScope ps(K_DENSIFY, s);
    densify_select_kernel<<<(num_rows + 255) / 256, 256, 0, s>>>(num_rows, xyz_gradient_accum, max_radii2D, opacity_raw,
                                                                 opacity_activation, scaling_raw, grad_threshold,
                                                                 min_opacity, dense_extent, scaffold_points, w.flags,
                                                                 w.c_keep, w.c_clone, w.c_split);
  }
  H3_CUDA(cudaGetLastError());
  H3_CUDA(launch_scan_inclusive(w.c_keep, w.s_keep, num_rows, w.status, w.counters, s));
  H3_CUDA(launch_scan_inclusive(w.c_clone, w.s_clone, num_rows, w.status + tiles, w.counters + 8, s));
  H3_CUDA(launch_scan_inclusive(w.c_split, w.s_split, num_rows, w.status + 2 * tiles, w.counters + 16, s));
  uint32_t last[3];
  H3_CUDA(cudaMemcpyAsync(&last[0], w.s_keep + num_rows - 1, 4, cudaMemcpyDeviceToHost, s));
  H3_CUDA(cudaMemcpyAsync(&last[1], w.s_clone + num_rows - 1, 4, cudaMemcpyDeviceToHost, s));
  H3_CUDA(cudaMemcpyAsync(&last[2], w.s_split + num_rows - 1, 4, cudaMemcpyDeviceToHost, s));
  H3_CUDA(cudaStreamSynchronize(s));   // the caller allocates the new tensors from these counts
  for (int k = 0; k < 3; ++k) counts_host[k] = (int32_t)last[k];
  return 0;
}

int h3dgs_densify_split_std(int32_t num_rows, const float* scaling_raw, const void* workspace, int32_t num_splits,
                            float* stds, h3dgs_stream_t stream) {
  using namespace h3dgs;
  if (num_rows <= 0 || num_splits <= 0) return 0;
  if (!scaling_raw || !workspace || !stds) return fail(H3DGS_EINVAL, "h3dgs_densify_split_std: NULL argument");
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  DnWs w = DnWs::make(const_cast<void*>(workspace), num_rows);
  ProfScope ps(K_DENSIFY, s);
  densify_split_std_kernel<<<(num_rows + 255) / 256, 256, 0, s>>>(num_rows, w.flags, w.s_split, num_splits, scaling_raw, stds);
  H3_CUDA(cudaGetLastError());
  return 0;
}

int h3dgs_densify_build(int32_t num_rows, const H3dgsDensifyTensor* tensors, int32_t num_tensors, int32_t xyz_index,
                        int32_t scaling_index, int32_t rotation_index, const void* workspace, int32_t kept,
                        int32_t clones, int32_t splits, const float* split_samples, h3dgs_stream_t stream) {
  using namespace h3dgs;
  if (num_rows <= 0) return 0;
  if (!tensors || num_tensors <= 0 || num_tensors > H3DGS_ADAM_MAX_TENSORS || !workspace || kept < 0 || clones < 0 ||
      splits < 0 || (splits > 0 && !split_samples) || xyz_index < 0 || xyz_index >= num_tensors || scaling_index < 0 ||
      scaling_index >= num_tensors || rotation_index < 0 || rotation_index >= num_tensors)
    return fail(H3DGS_EINVAL, "h3dgs_densify_build: bad argument");
  DnTensors t;
  t.n = num_tensors;
  t.xyz = xyz_index; t.scaling = scaling_index; t.rotation = rotation_index;
  for (int k = 0; k < num_tensors; ++k) {
    const H3dgsDensifyTensor& x = tensors[k];
    if (!x.param || !x.new_param || x.width <= 0 || ((x.new_exp_avg != nullptr) != (x.new_exp_avg_sq != nullptr)) ||
        (x.new_exp_avg && (!x.exp_avg || !x.exp_avg_sq)))
      return fail(H3DGS_EINVAL, "h3dgs_densify_build: bad tensor");
    t.src[k] = x.param; t.m_src[k] = x.exp_avg; t.v_src[k] = x.exp_avg_sq;
    t.dst[k] = x.new_param; t.m_dst[k] = x.new_exp_avg; t.v_dst[k] = x.new_exp_avg_sq;
    t.width[k] = x.width;
  }
  if (tensors[xyz_index].width != 3 || tensors[scaling_index].width != 3 || tensors[rotation_index].width != 4)
    return fail(H3DGS_EINVAL, "h3dgs_densify_build: xyz / scaling / rotation must be 3 / 3 / 4 floats wide");
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  DnWs w = DnWs::make(const_cast<void*>(workspace), num_rows);
  ProfScope ps(K_DENSIFY, s);
  const long long threads = (long long)num_rows * 32;
  densify_build_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, s>>>(num_rows, t, w.flags, w.s_keep, w.s_clone,
                                                                        w.s_split, kept, clones, splits, split_samples);
  H3_CUDA(cudaGetLastError());
  return 0;
}

}  // extern "C"
