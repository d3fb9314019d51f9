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

// ---- incremental cut: one refinement step of an existing cut (Switching::changeToSizeStep and its kernels,
// GH/runtime_switching.cu:145-400; the viewer moves the cut a level per frame instead of recomputing it) ----

// changeNodesOnce (:145-207): per cut node, stay (1), split into itself + its children (1 + count_children), or
// collapse into the parent (0); nodes whose children are not resident (start_children == -1) are flagged instead
__global__ void __launch_bounds__(256) cut_step_decide_kernel(int N, const int* __restrict__ indices,
                                                             const int* __restrict__ nodes,
                                                             const float* __restrict__ boxes,
                                                             const float* __restrict__ viewpoint, float target_size,
                                                             int* __restrict__ split, uint32_t* __restrict__ counts,
                                                             uint32_t* __restrict__ need) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N) return;
  const float vx = __ldg(viewpoint), vy = __ldg(viewpoint + 1), vz = __ldg(viewpoint + 2);
  const int id = __ldg(indices + i);
  const int* p = nodes + (size_t)id * 7;
  const int depth = __ldg(p), parent = __ldg(p + 1), start_children = __ldg(p + 5), count_children = __ldg(p + 6);
  const float size = projected_size(boxes, (size_t)id, vx, vy, vz);
  uint32_t count = 1, need_child = 0;
  if (size >= target_size) {
    if (depth > 0 && split[id] == 0) {
      if (start_children == -1) {
        need_child = 1;
      } else {
        count += (uint32_t)count_children;
        split[id] = 1;
      }
    }
  } else if (parent != -1) {
    if (projected_size(boxes, (size_t)parent, vx, vy, vz) < target_size) {   // collapse
      split[parent] = 0;
      count = 0;
    }
  }
  counts[i] = count;
  need[i] = need_child;
}

// putNodes (:209-230) + the DeviceSelect::Flagged of the nodes that need their children loaded
__global__ void __launch_bounds__(256) cut_step_put_nodes_kernel(int N, const int* __restrict__ indices,
                                                                const int* __restrict__ nodes,
                                                                const uint32_t* __restrict__ counts,
                                                                const uint32_t* __restrict__ offsets,
                                                                const uint32_t* __restrict__ need,
                                                                const uint32_t* __restrict__ need_offsets,
                                                                int* __restrict__ next_nodes, int max_nodes,
                                                                int* __restrict__ nodes_to_expand) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N) return;
  const int id = __ldg(indices + i);
  if (need[i] && nodes_to_expand) nodes_to_expand[need_offsets[i] - 1u] = id;
  const uint32_t count = counts[i];
  if (count == 0) return;
  const uint32_t off = i == 0 ? 0u : offsets[i - 1];
  if (off + count > (uint32_t)max_nodes) return;   // the caller is told through new_N > max_nodes
  next_nodes[off] = id;
  const int start_children = __ldg(nodes + (size_t)id * 7 + 5);
  for (uint32_t k = 1; k < count; ++k) next_nodes[off + k] = start_children + (int)k - 1;
}

// countRenderIndicesIndexed (:232-246)
__global__ void __launch_bounds__(256) cut_step_count_render_kernel(int N, const int* __restrict__ indices,
                                                                   const int* __restrict__ nodes,
                                                                   const int* __restrict__ split,
                                                                   uint32_t* __restrict__ counts) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N) return;
  const int id = __ldg(indices + i);
  const int* p = nodes + (size_t)id * 7;
  int count = __ldg(p + 3);
  if (__ldg(p) > 0 && split[id] == 0) count += __ldg(p + 4);
  counts[i] = (uint32_t)count;
}

// putRenderIndicesIndexed (:248-281)
__global__ void __launch_bounds__(256) cut_step_put_render_kernel(int N, const int* __restrict__ indices,
                                                                 const int* __restrict__ nodes,
                                                                 const uint32_t* __restrict__ counts,
                                                                 const uint32_t* __restrict__ offsets,
                                                                 int* __restrict__ render_indices,
                                                                 int* __restrict__ parent_indices,
                                                                 int* __restrict__ nodes_of_render_indices) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N) return;
  const int id = __ldg(indices + i);
  const int* p = nodes + (size_t)id * 7;
  const int parent = __ldg(p + 1), start = __ldg(p + 2);
  const int parent_gaussian = parent != -1 ? __ldg(nodes + (size_t)parent * 7 + 2) : -1;
  const uint32_t off = i == 0 ? 0u : offsets[i - 1];
  for (uint32_t k = 0; k < counts[i]; ++k) {
    render_indices[off + k] = start + (int)k;
    if (parent_indices) parent_indices[off + k] = parent_gaussian;
    if (nodes_of_render_indices) nodes_of_render_indices[off + k] = id;
  }
}

}  // namespace

struct CutStepWs {
  uint32_t *counts, *offsets, *need, *need_offsets;
  unsigned long long* status;
  uint32_t* counters;
  size_t bytes;
  static CutStepWs make(void* base, int max_nodes) {
    CutStepWs w;
    char* p = reinterpret_cast<char*>(base);
    char* p0 = p;
    const size_t n = (size_t)(max_nodes > 0 ? max_nodes : 1);
    carve(p, w.counts, n); carve(p, w.offsets, n); carve(p, w.need, n); carve(p, w.need_offsets, n);
    carve(p, w.status, (n + SCAN_TILE - 1) / SCAN_TILE + 1);
    carve(p, w.counters, 8);
    w.bytes = (size_t)(p - p0) + 256;
    return w;
  }
};

size_t cut_step_ws_bytes(int max_nodes) { return CutStepWs::make(nullptr, max_nodes).bytes; }

// -> out_host[0] = success (0 when the new node list would exceed max_nodes: nothing else is written),
//    [1] = new node count, [2] = render entries, [3] = nodes that need their children loaded
cudaError_t run_cut_step(int N, const int* node_indices, int* new_node_indices, int max_nodes, const int* nodes,
                         const float* boxes, const float* viewpoint, float target_size, int* split,
                         int* render_indices, int* parent_indices, int* nodes_of_render_indices, int* nodes_to_expand,
                         void* workspace, int* out_host, cudaStream_t s) {
  out_host[0] = 1; out_host[1] = out_host[2] = out_host[3] = 0;
  if (N <= 0) return cudaSuccess;
  CutStepWs w = CutStepWs::make(workspace, max_nodes);
  const int blocks = (N + 255) / 256;
  { ProfScope ps(K_CUT_MARK, s);
    cut_step_decide_kernel<<<blocks, 256, 0, s>>>(N, node_indices, nodes, boxes, viewpoint, target_size, split,
                                                  w.counts, w.need); }
  cudaError_t e = launch_scan_inclusive(w.need, w.need_offsets, N, w.status, w.counters, s);
  if (e != cudaSuccess) return e;
  e = launch_scan_inclusive(w.counts, w.offsets, N, w.status, w.counters, s);
  if (e != cudaSuccess) return e;
  uint32_t last[2] = {0, 0};
  e = cudaMemcpyAsync(&last[0], w.need_offsets + N - 1, 4, cudaMemcpyDeviceToHost, s);
  if (e != cudaSuccess) return e;
  e = cudaMemcpyAsync(&last[1], w.offsets + N - 1, 4, cudaMemcpyDeviceToHost, s);
  if (e != cudaSuccess) return e;
  e = cudaStreamSynchronize(s);   // the reference synchronises here too (:345-347): the new list may not fit
  if (e != cudaSuccess) return e;
  out_host[3] = (int)last[0];
  out_host[1] = (int)last[1];
  if ((int)last[1] > max_nodes) { out_host[0] = 0; return cudaSuccess; }
  const int newN = (int)last[1];
  { ProfScope ps(K_CUT_PUT, s);
    cut_step_put_nodes_kernel<<<blocks, 256, 0, s>>>(N, node_indices, nodes, w.counts, w.offsets, w.need,
                                                     w.need_offsets, new_node_indices, max_nodes, nodes_to_expand); }
  if (newN == 0) return cudaGetLastError();
  const int rblocks = (newN + 255) / 256;
  { ProfScope ps(K_CUT_MARK, s);
    cut_step_count_render_kernel<<<rblocks, 256, 0, s>>>(newN, new_node_indices, nodes, split, w.counts); }
  e = launch_scan_inclusive(w.counts, w.offsets, newN, w.status, w.counters, s);
  if (e != cudaSuccess) return e;
  { ProfScope ps(K_CUT_PUT, s);
    cut_step_put_render_kernel<<<rblocks, 256, 0, s>>>(newN, new_node_indices, nodes, w.counts, w.offsets,
                                                       render_indices, parent_indices, nodes_of_render_indices); }
  e = cudaMemcpyAsync(&last[0], w.offsets + newN - 1, 4, cudaMemcpyDeviceToHost, s);
  if (e != cudaSuccess) return e;
  e = cudaStreamSynchronize(s);
  if (e != cudaSuccess) return e;
  out_host[2] = (int)last[0];
  return cudaGetLastError();
}

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

extern "C" {

size_t h3dgs_cut_step_workspace_bytes(int32_t max_nodes) { return h3dgs::cut_step_ws_bytes(max_nodes); }

int h3dgs_cut_step(int32_t num_cut_nodes, const int32_t* node_indices, int32_t* new_node_indices, int32_t max_nodes,
                   const int32_t* nodes, const float* boxes, const float* viewpoint_dev, float target_size,
                   int32_t* split, int32_t* render_indices, int32_t* parent_indices,
                   int32_t* nodes_of_render_indices, int32_t* nodes_to_expand, void* workspace,
                   size_t workspace_bytes, int32_t* out_host, h3dgs_stream_t stream) {
  using namespace h3dgs;
  if (num_cut_nodes < 0 || max_nodes < num_cut_nodes || !out_host)
    return fail(H3DGS_EINVAL, "h3dgs_cut_step: bad sizes");
  if (num_cut_nodes > 0 && (!node_indices || !new_node_indices || !nodes || !boxes || !viewpoint_dev || !split ||
                            !render_indices || !workspace))
    return fail(H3DGS_EINVAL, "h3dgs_cut_step: NULL argument");
  if (workspace_bytes < cut_step_ws_bytes(max_nodes)) return fail(H3DGS_ENOSPACE, "h3dgs_cut_step: workspace too small");
  cudaError_t e = run_cut_step(num_cut_nodes, node_indices, new_node_indices, max_nodes, nodes, boxes, viewpoint_dev,
                               target_size, split, render_indices, parent_indices, nodes_of_render_indices,
                               nodes_to_expand, workspace, out_host, reinterpret_cast<cudaStream_t>(stream));
  return e == cudaSuccess ? 0 : cuda_fail(e, "h3dgs_cut_step");
}

}  // extern "C"

// ---- expand_to_size without a caller-side workspace (h3dgs_expand_to_size_static) ----
// For callers that cannot allocate (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc): the
// scratch of the cut (per-node counts, their scan, the scan's look-back words) lives in this file-scope device
// array, which every GPU that loads the library holds until exit.  All calls share it and it keeps the previous
// call's contents: the counts and offsets are fully rewritten for nodes 0..N-1 by every call and the look-back
// words and counters are cleared on the stream before the scan (scan.cu launch_scan), so nothing stale is read --
// but two calls that overlap on the GPU (different streams, different host threads) race on it.
constexpr size_t STATIC_CUT_BYTES =
    2 * sizeof(uint32_t) * (size_t)H3DGS_STATIC_CUT_MAX_NODES                                      // counts, offsets
    + sizeof(unsigned long long) * ((size_t)H3DGS_STATIC_CUT_MAX_NODES / h3dgs::SCAN_TILE + 2)     // look-back words
    + 8 * sizeof(uint32_t) + 5 * 256;                                                              // counters, carve slack
static __device__ __align__(256) unsigned char g_static_cut_scratch[STATIC_CUT_BYTES];   // 8.0 MiB per GPU

namespace h3dgs {
namespace {

// In host code `g_static_cut_scratch` is only a handle; its device address is asked for once per device, with that
// device current.
constexpr int STATIC_MAX_DEVICES = 64;
void* g_static_cut_ptr[STATIC_MAX_DEVICES] = {nullptr};

int static_cut_scratch(void** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return cuda_fail(e, "h3dgs_expand_to_size_static: cudaGetDevice");
  if (dev < 0 || dev >= STATIC_MAX_DEVICES) return fail(H3DGS_EINVAL, "h3dgs_expand_to_size_static: device ordinal above 63");
  void* p = __atomic_load_n(&g_static_cut_ptr[dev], __ATOMIC_ACQUIRE);
  if (p == nullptr) {
    e = cudaGetSymbolAddress(&p, g_static_cut_scratch);
    if (e != cudaSuccess) return cuda_fail(e, "h3dgs_expand_to_size_static: cudaGetSymbolAddress");
    if (p == nullptr || (reinterpret_cast<size_t>(p) & 255))
      return fail(H3DGS_EINVAL, "h3dgs_expand_to_size_static: scratch address is null or misaligned");
    __atomic_store_n(&g_static_cut_ptr[dev], p, __ATOMIC_RELEASE);   // every thread would store the same address
  }
  *out = p;
  return 0;
}

}  // namespace
}  // namespace h3dgs

extern "C" {

int32_t h3dgs_expand_to_size_static_max_nodes(void) { return H3DGS_STATIC_CUT_MAX_NODES; }

int h3dgs_expand_to_size_static(int32_t N, float target_size, const int32_t* nodes, const float* boxes,
                                const float* viewpoint_dev, int32_t* render_indices, int32_t* parent_indices,
                                int32_t* nodes_for_render_indices, int32_t* count_host, h3dgs_stream_t stream) {
  using namespace h3dgs;
  if (N < 0) return fail(H3DGS_EINVAL, "negative node count");
  if (!count_host) return fail(H3DGS_EINVAL, "count_host is required");
  *count_host = 0;
  if (N == 0) return 0;
  if (N > H3DGS_STATIC_CUT_MAX_NODES)
    return fail(H3DGS_ENOSPACE, "h3dgs_expand_to_size_static: more than H3DGS_STATIC_CUT_MAX_NODES (1048576) nodes; "
                                "use h3dgs_expand_to_size with a caller-side workspace");
  if (!nodes || !boxes || !viewpoint_dev || !render_indices)
    return fail(H3DGS_EINVAL, "nodes, boxes, viewpoint and render_indices are required");
  if (reinterpret_cast<size_t>(boxes) & 15) return fail(H3DGS_EINVAL, "boxes must be 16-byte aligned");
  if (expand_ws_bytes(N) > STATIC_CUT_BYTES) return fail(H3DGS_ENOSPACE, "h3dgs_expand_to_size_static: scratch layout exceeds the array");
  void* scratch = nullptr;
  int rc = static_cut_scratch(&scratch);
  if (rc) return rc;
  cudaError_t e = run_expand_to_size(N, target_size, nodes, boxes, viewpoint_dev, render_indices, parent_indices,
                                     nodes_for_render_indices, scratch, count_host, reinterpret_cast<cudaStream_t>(stream));
  return e == cudaSuccess ? 0 : cuda_fail(e, "h3dgs_expand_to_size_static");
}

}  // extern "C"
