// Hierarchy cut kernels: granularity cut (expand_to_size) and per-entry interpolation weights.
// Replaces markNodesForSize / putRenderIndices / computeTsIndexed and their drivers
// (GH/runtime_switching.cu:55-80,108-143,402-540).
//
// The projected-size arithmetic (clamp to the box, squared distance with the same product order,
// IEEE sqrt and divide) keeps the reference's expression shapes so that the integer outputs -- which
// hang on a float comparison per node -- come out identical.
#include "common.cuh"

namespace h3dgs {

namespace {

// size of a node as seen from `vp`: box extent over distance to the box; FLT_MAX inside the box
__device__ __forceinline__ float projected_size(const float* __restrict__ boxes, size_t i, float vx,
                                                float vy, float vz) {
  const float4 mn = __ldg(reinterpret_cast<const float4*>(boxes) + 2 * i);
  const float4 mx = __ldg(reinterpret_cast<const float4*>(boxes) + 2 * i + 1);
  bool inside = true;
  inside &= vx >= mn.x && vx <= mx.x;
  inside &= vy >= mn.y && vy <= mx.y;
  inside &= vz >= mn.z && vz <= mx.z;
  if (inside) return 3.402823466e+38f;
  const float cx = max(mn.x, min(mx.x, vx));
  const float cy = max(mn.y, min(mx.y, vy));
  const float cz = max(mn.z, min(mx.z, vz));
  const float dx = vx - cx, dy = vy - cy, dz = vz - cz;
  const float min_dist = sqrt(dx * dx + dy * dy + dz * dz);
  return mn.w / min_dist;
}

__global__ void __launch_bounds__(256) mark_nodes_for_size_kernel(
    int N, const int* __restrict__ nodes, const float* __restrict__ boxes,
    const float* __restrict__ viewpoint, float target_size, uint32_t* __restrict__ counts) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N) return;
  const float vx = __ldg(viewpoint), vy = __ldg(viewpoint + 1), vz = __ldg(viewpoint + 2);
  const int* p = nodes + (size_t)i * 7;
  const int depth = __ldg(p), parent = __ldg(p + 1), count_leafs = __ldg(p + 3),
            count_merged = __ldg(p + 4);
  const float size = projected_size(boxes, (size_t)i, vx, vy, vz);
  int count = 0;
  if (size >= target_size) {
    count = count_leafs;
  } else if (parent != -1) {
    const float parent_size = projected_size(boxes, (size_t)parent, vx, vy, vz);
    if (parent_size >= target_size) {
      count = count_leafs;
      if (depth != 0) count += count_merged;
    }
  }
  counts[i] = (uint32_t)count;
}

__global__ void __launch_bounds__(256) put_render_indices_kernel(
    int N, const int* __restrict__ nodes, const uint32_t* __restrict__ counts,
    const uint32_t* __restrict__ offsets, int* __restrict__ render_indices,
    int* __restrict__ parent_indices, int* __restrict__ nodes_for_render_indices) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N) return;
  const int count = (int)counts[i];
  if (count == 0) return;
  const uint32_t offset = (i == 0) ? 0u : offsets[i - 1];
  const int* p = nodes + (size_t)i * 7;
  const int parent = __ldg(p + 1), start = __ldg(p + 2);
  int parentgaussian = -1;
  if (parent != -1) parentgaussian = __ldg(nodes + (size_t)parent * 7 + 2);
  for (int k = 0; k < count; ++k) {
    render_indices[offset + k] = start + k;
    if (parent_indices) parent_indices[offset + k] = parentgaussian;
    if (nodes_for_render_indices) nodes_for_render_indices[offset + k] = i;
  }
}

__global__ void __launch_bounds__(256) interpolation_weights_kernel(
    int K, const int* __restrict__ indices, const int* __restrict__ nodes,
    const float* __restrict__ boxes, float vx, float vy, float vz, float target_size,
    float* __restrict__ ts, int* __restrict__ kids) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= K) return;
  const int node_id = __ldg(indices + i);
  const int parent = __ldg(nodes + (size_t)node_id * 7 + 1);
  float t;
  int nk = 1;
  if (parent == -1) {
    t = 1.0f;
  } else {
    const float parentsize = projected_size(boxes, (size_t)parent, vx, vy, vz);
    if (parentsize > 2.0f * target_size) {
      t = 1.0f;
    } else {
      const float size = projected_size(boxes, (size_t)node_id, vx, vy, vz);
      const float start = max(0.5f * parentsize, size);
      const float diff = parentsize - start;
      if (diff <= 0) {
        t = 1.0f;
      } else {
        const float tdiff = max(0.0f, target_size - start);
        t = max(1.0f - (tdiff / diff), 0.0f);
      }
    }
    nk = __ldg(nodes + (size_t)parent * 7 + 6);
  }
  ts[i] = t;
  kids[i] = nk;
}

}  // namespace

struct ExpandWs {
  uint32_t* counts;
  uint32_t* offsets;
  unsigned long long* status;
  uint32_t* counters;
  size_t bytes;
  static ExpandWs make(void* base, int N) {
    ExpandWs w;
    char* p = reinterpret_cast<char*>(base);
    char* p0 = p;
    size_t n = (size_t)(N > 0 ? N : 1);
    carve(p, w.counts, n);
    carve(p, w.offsets, n);
    carve(p, w.status, (n + SCAN_TILE - 1) / SCAN_TILE + 1);
    carve(p, w.counters, 8);
    w.bytes = (size_t)(p - p0) + 256;
    return w;
  }
};

size_t expand_ws_bytes(int N) { return ExpandWs::make(nullptr, N).bytes; }

cudaError_t run_expand_to_size(int N, float target_size, const int* nodes, const float* boxes,
                               const float* viewpoint, int* render_indices, int* parent_indices,
                               int* nodes_for_render_indices, void* workspace, int* count_host,
                               cudaStream_t s) {
  *count_host = 0;
  if (N <= 0) return cudaSuccess;
  ExpandWs w = ExpandWs::make(workspace, N);
  const int blocks = (N + 255) / 256;
  { ProfScope ps(K_CUT_MARK, s);
  mark_nodes_for_size_kernel<<<blocks, 256, 0, s>>>(N, nodes, boxes, viewpoint, target_size, w.counts); }
  cudaError_t e = launch_scan_inclusive(w.counts, w.offsets, N, w.status, w.counters, s);
  if (e != cudaSuccess) return e;
  { ProfScope ps(K_CUT_PUT, s);
  put_render_indices_kernel<<<blocks, 256, 0, s>>>(N, nodes, w.counts, w.offsets, render_indices,
                                                   parent_indices, nodes_for_render_indices); }
  unsigned long long total = 0;
  e = cudaMemcpyAsync(&total, w.counters + 2, sizeof(total), cudaMemcpyDeviceToHost, s);
  if (e != cudaSuccess) return e;
  e = cudaStreamSynchronize(s);
  if (e != cudaSuccess) return e;
  *count_host = (int)total;
  return cudaGetLastError();
}

cudaError_t run_interpolation_weights(int K, const int* indices, float target_size, const int* nodes,
                                      const float* boxes, float vx, float vy, float vz, float* ts,
                                      int* kids, cudaStream_t s) {
  if (K <= 0) return cudaSuccess;
  ProfScope ps(K_CUT_WEIGHTS, s);
  interpolation_weights_kernel<<<(K + 255) / 256, 256, 0, s>>>(K, indices, nodes, boxes, vx, vy, vz,
                                                               target_size, ts, kids);
  return cudaGetLastError();
}

}  // namespace h3dgs
