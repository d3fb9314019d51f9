// extern "C" entry points of libh3dgs_b200.so (include/h3dgs_b200.h): argument checks, workspace carving,
// kernel sequencing.  No torch, no allocation.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "common.cuh"

namespace h3dgs {

static thread_local char g_err[512] = "";

int fail(int code, const char* what) {
  std::snprintf(g_err, sizeof(g_err), "%s", what);
  return code;
}
int cuda_fail(cudaError_t e, const char* what) {
  std::snprintf(g_err, sizeof(g_err), "CUDA error %d (%s) in %s", (int)e, cudaGetErrorString(e), what);
  return (int)e;
}

// ---- instrumentation ----
namespace {
struct Prof {
  bool on = false;
  long long launches_total = 0;
  double ms[K_COUNT] = {0};
  long long count[K_COUNT] = {0};
  struct Pair { int id; cudaEvent_t a, b; };
  std::vector<Pair> pending;
  std::vector<cudaEvent_t> pool;
  cudaEvent_t cur[K_COUNT] = {nullptr};
  cudaEvent_t get() {
    if (!pool.empty()) { cudaEvent_t e = pool.back(); pool.pop_back(); return e; }
    cudaEvent_t e; cudaEventCreate(&e); return e;
  }
} g_prof;
const char* const kKernelNames[K_COUNT] = {
  "preprocess_fwd", "scan", "emit_instances", "sort_histogram", "sort_scan", "sort_onesweep_pass",
  "tile_ranges", "blend_fwd", "blend_bwd", "preprocess_bwd", "cut_mark_nodes", "cut_put_indices",
  "cut_interp_weights", "sparse_adam", "mark_visible", "memset", "loss", "hier_build_tree", "hier_build_merge",
  "densify"};
}  // namespace

void prof_begin(int id, cudaStream_t s) {
  if (id != K_MEMSET) g_prof.launches_total++;
  if (!g_prof.on) return;
  g_prof.cur[id] = g_prof.get();
  cudaEventRecord(g_prof.cur[id], s);
}
void prof_end(int id, cudaStream_t s) {
  if (!g_prof.on || !g_prof.cur[id]) return;
  cudaEvent_t b = g_prof.get();
  cudaEventRecord(b, s);
  g_prof.pending.push_back({id, g_prof.cur[id], b});
  g_prof.cur[id] = nullptr;
}

// device pointer of the pair-category counters while an instrumented pass is on (h3dgs_blend_stats)
static unsigned long long* g_blend_stats = nullptr;

// H3DGS_STAGE=tma|ldgsts selects how the blend kernels stage records into shared memory.
int set_bwd_variant(int v);   // blend_bwd.cu
int blend_stage_mode() {
  static int mode = -1;
  if (mode < 0) {
    const char* e = std::getenv("H3DGS_STAGE");
    mode = (e && (std::strcmp(e, "tma") == 0 || std::strcmp(e, "bulk") == 0)) ? 1 : 0;
  }
  return mode;
}

// H3DGS_NORM_VARIANT=0..3 selects the summation order of ||q||^2 in the fused F.normalize (parity diagnostic;
// 2, the default, is the order torch's CUDA reduction uses, see gsmath.cuh).
int norm_variant() {
  static int v = -1;
  if (v < 0) {
    const char* e = std::getenv("H3DGS_NORM_VARIANT");
    v = e ? std::atoi(e) : 2;
    if (v < 0 || v > 3) v = 2;
  }
  return v;
}

// Zero fill with plain 16-byte stores (H3DGS_ZERO=kernel) instead of cudaMemsetAsync (default).
__global__ void fill_zero_kernel(float* __restrict__ p, size_t n) {
  const size_t head = (16 - (reinterpret_cast<size_t>(p) & 15)) / 4 % 4;   // floats up to 16-byte alignment
  const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  if (tid < head && tid < n) p[tid] = 0.f;
  if (n <= head) return;
  float4* q = reinterpret_cast<float4*>(p + head);
  const size_t n4 = (n - head) / 4;
  for (size_t i = tid; i < n4; i += stride) q[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  const size_t tail = head + n4 * 4;
  if (tid < n - tail) p[tail + tid] = 0.f;
}
int zero_mode() {
  static int mode = -1;
  if (mode < 0) {
    const char* e = std::getenv("H3DGS_ZERO");
    mode = (e && std::strcmp(e, "kernel") == 0) ? 1 : 0;
  }
  return mode;
}
cudaError_t zero_fill(float* p, size_t n, cudaStream_t s) {
  if (!p || n == 0) return cudaSuccess;
  if (zero_mode() == 0) return cudaMemsetAsync(p, 0, n * sizeof(float), s);
  const size_t want = (n / 4 + 255) / 256;
  const int blocks = (int)(want < 148 * 16 ? (want ? want : 1) : 148 * 16);
  fill_zero_kernel<<<blocks, 256, 0, s>>>(p, n);
  return cudaGetLastError();
}

// side stream for work that overlaps the caller's stream (one per device, created on first use)
struct SideStream {
  cudaStream_t stream = nullptr;
  cudaEvent_t fork = nullptr, join = nullptr;
};
SideStream* side_stream() {
  static SideStream per_device[64];
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return nullptr;
  SideStream& ss = per_device[dev];
  if (!ss.stream) {
    if (cudaStreamCreateWithFlags(&ss.stream, cudaStreamNonBlocking) != cudaSuccess) { ss.stream = nullptr; return nullptr; }
    cudaEventCreateWithFlags(&ss.fork, cudaEventDisableTiming);
    cudaEventCreateWithFlags(&ss.join, cudaEventDisableTiming);
  }
  return &ss;
}
bool overlap_zero() {
  static int v = -1;
  if (v < 0) {
    const char* e = std::getenv("H3DGS_OVERLAP_ZERO");
    v = (e && std::strcmp(e, "0") == 0) ? 0 : 1;
  }
  return v == 1;
}

int g_strip_key_debug = 0;
int strip_key_debug_mode() { return g_strip_key_debug; }

size_t expand_ws_bytes(int N);
cudaError_t run_expand_to_size(int N, float target_size, const int* nodes, const float* boxes,
                               const float* viewpoint, int* render_indices, int* parent_indices,
                               int* nodes_for_render_indices, void* workspace, int* count_host,
                               cudaStream_t s);
cudaError_t run_interpolation_weights(int K, const int* indices, float target_size, const int* nodes,
                                      const float* boxes, float vx, float vy, float vz, float* ts,
                                      int* kids, cudaStream_t s);
cudaError_t run_sparse_adam(const H3dgsAdamTensor* tensors, int n, const long long* relevant,
                            long long num_relevant, long long num_rows, cudaStream_t s);
cudaError_t run_sparse_adam_compact(const H3dgsAdamTensor* tensors, int n, const long long* rows, const int* count,
                                    int capacity, const float* gate, cudaStream_t s);
cudaError_t run_ssim_forward(int C, int H, int W, const float* img1, const float* img2, const float* mask,
                             const float* window, float* ssim_map, float* dm_dmu1, float* dm_de11, float* dm_de12, double* sums,
                             cudaStream_t s);
cudaError_t run_masked_l1(long long n, const float* pred, const float* target, const float* mask, double* sum,
                          const float* upstream, float scale, float* grad, cudaStream_t s);
cudaError_t run_ssim_backward(int C, int H, int W, const float* img1, const float* img2, const float* mask,
                              const float* window, const float* dm_dmu1, const float* dm_de11, const float* dm_de12,
                              const float* dL_dmap, const float* upstream, float ssim_scale, float l1_scale,
                              float* dL_dimg1, cudaStream_t s);
cudaError_t run_measure_peaks(double* out, float* scratch, int reps, cudaStream_t s);
cudaError_t run_pack_rows(const H3dgsRowTensor* tensors, int n, const long long* rows, long long K, float* packed,
                          float scale, bool pack, cudaStream_t s);

namespace {

int check_scene(const H3dgsView* v, const H3dgsGaussians* g) {
  if (!v || !g) return fail(H3DGS_EINVAL, "null view / gaussians");
  if (v->image_width <= 0 || v->image_height <= 0) return fail(H3DGS_EINVAL, "image size must be positive");
  if (!v->viewmatrix || !v->projmatrix || !v->campos || !v->background)
    return fail(H3DGS_EINVAL, "viewmatrix, projmatrix, campos and background are required");
  if (g->P < 0 || g->store_rows < 0) return fail(H3DGS_EINVAL, "negative sizes");
  if (g->P == 0) return 0;
  if (!g->means3D || !g->opacities) return fail(H3DGS_EINVAL, "means3D and opacities are required");
  if ((g->shs == nullptr) == (g->colors_precomp == nullptr))
    return fail(H3DGS_EINVAL, "Please provide excatly one of either SHs or precomputed colors!");
  const bool sr = g->scales != nullptr && g->rotations != nullptr;
  if ((!sr && g->cov3D_precomp == nullptr) || ((g->scales || g->rotations) && g->cov3D_precomp))
    return fail(H3DGS_EINVAL,
                "Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!");
  if (g->shs && (g->sh_degree < 0 || g->sh_degree > 3 || (g->sh_degree + 1) * (g->sh_degree + 1) > g->sh_coeffs))
    return fail(H3DGS_EINVAL, "sh_degree must be 0..3 and fit in sh_coeffs");
  if (g->shs_rest && (!g->shs || g->sh_coeffs < 2)) return fail(H3DGS_EINVAL, "shs_rest needs shs (dc) and sh_coeffs >= 2");
  if ((g->scale_activation != H3DGS_ACT_NONE && g->scale_activation != H3DGS_ACT_EXP) ||
      (g->rotation_activation != H3DGS_ACT_NONE && g->rotation_activation != H3DGS_ACT_NORMALIZE) ||
      (g->opacity_activation != H3DGS_ACT_NONE && g->opacity_activation != H3DGS_ACT_SIGMOID &&
       g->opacity_activation != H3DGS_ACT_ABS))
    return fail(H3DGS_EINVAL, "unsupported activation code");
  if ((g->scale_activation || g->rotation_activation) && g->cov3D_precomp)
    return fail(H3DGS_EINVAL, "scale / rotation activations need scales + rotations");
  if (g->skyboxnum < 0 || g->skyboxnum > g->P) return fail(H3DGS_EINVAL, "skyboxnum out of range");
  if (g->render_indices == nullptr) {
    if (g->P != g->store_rows) return fail(H3DGS_EINVAL, "plain mode needs P == store_rows");
  } else {
    if (g->skyboxnum > g->store_rows) return fail(H3DGS_EINVAL, "skyboxnum exceeds store_rows");
    if (g->parent_indices && !g->ts) return fail(H3DGS_EINVAL, "parent_indices needs interpolation weights");
  }
  if (g->rotations && (reinterpret_cast<size_t>(g->rotations) & 15))
    return fail(H3DGS_EINVAL, "rotations must be 16-byte aligned");
  return 0;
}

PreArgs make_pre(const H3dgsView* v, const H3dgsGaussians* g, int* radii, void* geom) {
  PreArgs a;
  std::memset(&a, 0, sizeof(a));
  a.P = g->P; a.rows = g->store_rows; a.D = g->sh_degree; a.M = g->sh_coeffs; a.skyboxnum = g->skyboxnum;
  a.W = v->image_width; a.H = v->image_height;
  a.grid_x = (a.W + TILE_X - 1) / TILE_X; a.grid_y = (a.H + TILE_Y - 1) / TILE_Y;
  a.scale_modifier = g->scale_modifier;
  a.tan_fovx = v->tan_fovx; a.tan_fovy = v->tan_fovy;
  a.focal_y = a.H / (2.0f * a.tan_fovy);   // rasterizer_impl.cu:241-242
  a.focal_x = a.W / (2.0f * a.tan_fovx);
  a.means = g->means3D; a.shs = g->shs; a.colors = g->colors_precomp; a.opac = g->opacities;
  a.scales = g->scales; a.rots = g->rotations; a.cov3D_pre = g->cov3D_precomp;
  a.shs_rest = g->shs ? g->shs_rest : nullptr;
  a.act_scale = g->scale_activation; a.act_rot = g->rotation_activation; a.act_opac = g->opacity_activation;
  a.norm_variant = norm_variant();
  a.idx = g->render_indices; a.pidx = g->render_indices ? g->parent_indices : nullptr;
  a.ts = g->ts; a.kids = g->kids;
  a.view = v->viewmatrix; a.proj = v->projmatrix; a.campos = v->campos;
  a.radii = radii;
  a.ws = GeomWs::make(geom, g->P);
  return a;
}

int debug_sync(int debug, cudaStream_t s, const char* where) {
  if (!debug) return 0;
  cudaError_t e = cudaStreamSynchronize(s);
  if (e == cudaSuccess) e = cudaGetLastError();
  if (e != cudaSuccess) return cuda_fail(e, where);
  return 0;
}

}  // namespace
}  // namespace h3dgs

using namespace h3dgs;

extern "C" {

int h3dgs_abi_version(void) { return H3DGS_ABI_VERSION; }
const char* h3dgs_last_error(void) { return g_err; }

size_t h3dgs_geom_bytes(int32_t P) { return GeomWs::make(nullptr, P).bytes; }
size_t h3dgs_image_bytes(int32_t W, int32_t H) { return ImageWs::make(nullptr, W, H).bytes; }
size_t h3dgs_binning_bytes(int64_t R, int32_t W, int32_t H) { return BinWs::make(nullptr, R, W, H).bytes; }
size_t h3dgs_backward_workspace_bytes(int32_t P) { return BwdWs::make(nullptr, P).bytes; }
size_t h3dgs_expand_workspace_bytes(int32_t N) { return expand_ws_bytes(N); }

int h3dgs_forward_begin(const H3dgsView* view, const H3dgsGaussians* g, int32_t* radii, void* geom,
                        size_t geom_bytes, int64_t* counts_host, void* counts_event, int debug,
                        h3dgs_stream_t stream) {
  int rc = check_scene(view, g);
  if (rc) return rc;
  if (!counts_host) return fail(H3DGS_EINVAL, "counts_host is required");
  counts_host[0] = 0;
  counts_host[1] = 0;
  cudaStream_t s = (cudaStream_t)stream;
  if (g->P == 0) {
    if (counts_event) H3_CUDA(cudaEventRecord((cudaEvent_t)counts_event, s));
    return 0;
  }
  if (!radii || !geom) return fail(H3DGS_EINVAL, "radii and geom are required");
  if (geom_bytes < h3dgs_geom_bytes(g->P)) return fail(H3DGS_ENOSPACE, "geom workspace too small");
  if (reinterpret_cast<size_t>(geom) & 15) return fail(H3DGS_EINVAL, "geom must be 16-byte aligned");
  PreArgs a = make_pre(view, g, radii, geom);
  GeomWs& gw = a.ws;
  { ProfScope ps(K_MEMSET, s); H3_CUDA(cudaMemsetAsync(gw.zero_base, 0, gw.zero_bytes, s)); }
  H3_CUDA(launch_preprocess_fwd(a, s));
  if ((rc = debug_sync(debug, s, "preprocess_fwd"))) return rc;
  H3_CUDA(launch_scan_visible(gw.tiles_touched, gw.offsets, g->P, gw.scan_status, gw.counters, s));
  // (R, V) leave for the host now; the kernels below take V from the device, so the host can wait for this copy
  // alone (counts_event) while the GPU goes on with the depth sort
  static_assert(sizeof(unsigned long long) == sizeof(int64_t), "counter width");
  H3_CUDA(cudaMemcpyAsync(counts_host + 1, gw.counters + 2, sizeof(int64_t), cudaMemcpyDeviceToHost, s));   // V
  H3_CUDA(cudaMemcpyAsync(counts_host, gw.counters + 4, sizeof(int64_t), cudaMemcpyDeviceToHost, s));       // R
  if (counts_event) H3_CUDA(cudaEventRecord((cudaEvent_t)counts_event, s));
  const unsigned long long* v_dev = reinterpret_cast<const unsigned long long*>(gw.counters + 2);
  // visible slots -> (depth, slot) order -> inclusive scan of their strip-record counts in that order
  H3_CUDA(launch_compact_visible(g->P, gw, s));
  if ((rc = debug_sync(debug, s, "compact_visible"))) return rc;
  H3_CUDA(launch_depth_sort(gw, g->P, v_dev, s));
  if ((rc = debug_sync(debug, s, "depth_sort"))) return rc;
  H3_CUDA(launch_scan_strips(gw.tile_rect, gw.dvals[0], gw.soffs, g->P, v_dev, gw.chunk_first, gw.gscan_status,
                             gw.gscan_counters, s));
  if (!counts_event) {
    H3_CUDA(cudaStreamSynchronize(s));
    if (counts_host[0] >= MAX_INSTANCES) return fail(H3DGS_EINVAL, "more than 2^30 tile instances");
  }
  return 0;
}

int h3dgs_forward_finish(const H3dgsView* view, const H3dgsGaussians* g, void* geom, void* binning,
                         size_t binning_bytes, void* image, int64_t R, int64_t V, float* out_color,
                         float* out_invdepth, int debug, h3dgs_stream_t stream, void* pre_blend_event) {
  int rc = check_scene(view, g);
  if (rc) return rc;
  if (!out_color) return fail(H3DGS_EINVAL, "out_color is required");
  cudaStream_t s = (cudaStream_t)stream;
  const int W = view->image_width, H = view->image_height;
  const size_t plane = (size_t)W * H;
  if (g->P == 0 || R <= 0) {
    // nothing to blend: the reference returns the zero-initialised image here (rasterizer_impl.cu:327-328)
    H3_CUDA(cudaMemsetAsync(out_color, 0, plane * 3 * sizeof(float), s));
    if (out_invdepth) H3_CUDA(cudaMemsetAsync(out_invdepth, 0, plane * sizeof(float), s));
    if (image) {
      ImageWs iw = ImageWs::make(image, W, H);
      H3_CUDA(cudaMemsetAsync(iw.n_contrib, 0, plane * sizeof(uint32_t), s));
    }
    if (pre_blend_event) H3_CUDA(cudaEventRecord((cudaEvent_t)pre_blend_event, s));
    return 0;
  }
  if (!geom || !binning || !image) return fail(H3DGS_EINVAL, "geom, binning and image are required");
  if (R >= MAX_INSTANCES) return fail(H3DGS_EINVAL, "more than 2^30 tile instances");
  if (V <= 0 || V > R || V > g->P) return fail(H3DGS_EINVAL, "num_visible must be in [1, min(P, num_rendered)]");
  if (binning_bytes < h3dgs_binning_bytes(R, W, H)) return fail(H3DGS_ENOSPACE, "binning workspace too small");
  GeomWs gw = GeomWs::make(geom, g->P);
  BinWs bw = BinWs::make(binning, R, W, H);
  ImageWs iw = ImageWs::make(image, W, H);
  const int grid_x = (W + TILE_X - 1) / TILE_X, grid_y = (H + TILE_Y - 1) / TILE_Y;
  const SortPlan plan = make_sort_plan(W, H);

  // ordered emission of strip records -> stable partition by strip -> expansion into the tiles' lists
  { ProfScope ps(K_MEMSET, s); H3_CUDA(cudaMemsetAsync(bw.zero_base, 0, bw.zero_bytes, s)); }
  H3_CUDA(launch_emit_strips(R, V, gw, bw, plan, s));
  if ((rc = debug_sync(debug, s, "emit_strips"))) return rc;
  H3_CUDA(launch_strip_sort(bw, R, reinterpret_cast<const unsigned long long*>(gw.gscan_counters + 2), plan, s));
  if ((rc = debug_sync(debug, s, "strip_sort"))) return rc;
  H3_CUDA(launch_expand(R, gw, bw, plan, grid_x, grid_y, iw.ranges, s));
  if ((rc = debug_sync(debug, s, "expand"))) return rc;
  H3_CUDA(launch_tile_order(iw.ranges, nullptr, grid_x * grid_y, iw.order_fwd, s));
  H3_CUDA(cudaMemsetAsync(iw.tile_last, 0, sizeof(uint32_t) * (size_t)(grid_x * grid_y), s));

  BlendArgs b;
  std::memset(&b, 0, sizeof(b));
  b.W = W; b.H = H; b.grid_x = grid_x; b.grid_y = grid_y; b.P = g->P; b.skyboxnum = g->skyboxnum;
  b.interp = (g->ts != nullptr && g->kids != nullptr) ? 1 : 0;
  b.ranges = iw.ranges; b.point_list = bw.point_list; b.rec = gw.rec; b.bg = view->background;
  b.final_T = iw.final_T; b.n_contrib = iw.n_contrib; b.accum = iw.accum;
  b.out_color = out_color; b.out_invdepth = out_invdepth;
  b.order = iw.order_fwd; b.tile_last = iw.tile_last; b.stats = g_blend_stats;
  b.snap = bw.snap;
  if (pre_blend_event) H3_CUDA(cudaEventRecord((cudaEvent_t)pre_blend_event, s));
  H3_CUDA(launch_blend_fwd(b, s));
  if ((rc = debug_sync(debug, s, "blend_fwd"))) return rc;
  return 0;
}

int h3dgs_backward(const H3dgsView* view, const H3dgsGaussians* g, const int32_t* radii, const void* geom,
                   const void* binning, const void* image, int64_t R, const float* dL_dout_color,
                   const float* dL_dout_invdepth, const H3dgsGrads* grads, void* workspace,
                   size_t workspace_bytes, int debug, h3dgs_stream_t stream) {
  int rc = check_scene(view, g);
  if (rc) return rc;
  if (!grads) return fail(H3DGS_EINVAL, "grads is required");
  cudaStream_t s = (cudaStream_t)stream;
  const int P = g->P, W = view->image_width, H = view->image_height;
  const size_t rows = (size_t)g->store_rows;
  const int M = g->sh_coeffs;
  const bool fused = g->render_indices != nullptr;

  // The N-row zero fill of fused-cut mode is pure HBM write bandwidth (4 GB at 17 M rows) while the blend
  // kernel that follows is issue-bound with ~1 % DRAM utilisation: run the fill on a side stream, forked
  // here and joined before the scatter kernel, so that the two overlap (H3DGS_OVERLAP_ZERO=0 disables).
  const bool compact = grads->row_map != nullptr || grads->compact_rows != nullptr || grads->compact_count != nullptr;
  if (compact) {
    if (!fused) return fail(H3DGS_EINVAL, "compact gradients need the fused-cut mode (render_indices)");
    if (!grads->row_map || !grads->compact_rows || !grads->compact_count || grads->compact_capacity <= 0)
      return fail(H3DGS_EINVAL, "compact gradients need row_map, compact_rows, compact_count and a positive capacity");
    if (grads->dirty_rows) return fail(H3DGS_EINVAL, "dirty_rows and compact gradients exclude each other");
    if (reinterpret_cast<size_t>(grads->dL_dmeans3D) & 15 || reinterpret_cast<size_t>(grads->dL_dsh) & 15 ||
        reinterpret_cast<size_t>(grads->dL_dsh_rest) & 15 || reinterpret_cast<size_t>(grads->dL_dopacity) & 15 ||
        reinterpret_cast<size_t>(grads->dL_dscales) & 15 || reinterpret_cast<size_t>(grads->dL_drotations) & 15 ||
        reinterpret_cast<size_t>(grads->dL_dcolors) & 15 || reinterpret_cast<size_t>(grads->dL_dcov3D) & 15)
      return fail(H3DGS_EINVAL, "compact gradient tensors must be 16-byte aligned");
    H3_CUDA(cudaMemsetAsync(grads->compact_count, 0, sizeof(int32_t), (cudaStream_t)stream));
  }
  const bool busy = !(P == 0 || R <= 0);
  // (not while per-kernel timing is on: overlapped kernels would be charged each other's time)
  SideStream* side = (busy && fused && !compact && !grads->rows_prezeroed && overlap_zero() && !g_prof.on) ? side_stream() : nullptr;
  cudaStream_t zs = s;
  if (side) {
    H3_CUDA(cudaEventRecord(side->fork, s));
    H3_CUDA(cudaStreamWaitEvent(side->stream, side->fork, 0));
    zs = side->stream;
  }
  auto zero = [&](float* p, size_t n) -> cudaError_t {
    return zero_fill(p, n, zs);
  };
  if (grads->dirty_rows) H3_CUDA(cudaMemsetAsync(grads->dirty_rows, 0, sizeof(int32_t), s));
  if (!compact && (P == 0 || R <= 0 || (fused && !grads->rows_prezeroed))) {
    ProfScope ps(K_MEMSET, zs);
    // fused-cut mode scatters into store rows; empty renders produce all-zero gradients
    H3_CUDA(zero(grads->dL_dmeans3D, rows * 3));
    const bool split = g->shs && g->shs_rest;
    H3_CUDA(zero(grads->dL_dsh, rows * (size_t)(split ? 1 : M) * 3));
    if (split) H3_CUDA(zero(grads->dL_dsh_rest, rows * (size_t)(M - 1) * 3));
    H3_CUDA(zero(grads->dL_dcolors, rows * 3));
    H3_CUDA(zero(grads->dL_dopacity, rows));
    H3_CUDA(zero(grads->dL_dscales, rows * 3));
    H3_CUDA(zero(grads->dL_drotations, rows * 4));
    H3_CUDA(zero(grads->dL_dcov3D, rows * 6));
  }
  if (side) H3_CUDA(cudaEventRecord(side->join, side->stream));
  if (P == 0 || R <= 0) {
    H3_CUDA(zero(grads->dL_dmeans2D, (size_t)P * 3));
    return 0;
  }
  if (!radii || !geom || !binning || !image || !dL_dout_color || !workspace)
    return fail(H3DGS_EINVAL, "radii, geom, binning, image, dL_dout_color and workspace are required");
  if (workspace_bytes < h3dgs_backward_workspace_bytes(P)) return fail(H3DGS_ENOSPACE, "backward workspace too small");

  GeomWs gw = GeomWs::make(const_cast<void*>(geom), P);
  BinWs bw = BinWs::make(const_cast<void*>(binning), R, W, H);
  ImageWs iw = ImageWs::make(const_cast<void*>(image), W, H);
  BwdWs ww = BwdWs::make(workspace, P);
  { ProfScope ps(K_MEMSET, s); H3_CUDA(cudaMemsetAsync(ww.grad, 0, sizeof(GradRec) * (size_t)P, s)); }

  BlendArgs b;
  std::memset(&b, 0, sizeof(b));
  b.W = W; b.H = H; b.grid_x = (W + TILE_X - 1) / TILE_X; b.grid_y = (H + TILE_Y - 1) / TILE_Y;
  b.P = P; b.skyboxnum = g->skyboxnum;
  b.interp = (g->ts != nullptr && g->kids != nullptr) ? 1 : 0;
  b.ranges = iw.ranges; b.point_list = bw.point_list; b.rec = gw.rec; b.bg = view->background;
  b.final_T = iw.final_T; b.n_contrib = iw.n_contrib; b.accum = iw.accum;
  b.dL_dcolor = dL_dout_color; b.dL_dinvdepth = dL_dout_invdepth; b.grad = ww.grad;
  H3_CUDA(launch_tile_order(nullptr, iw.tile_last, b.grid_x * b.grid_y, iw.order_bwd, s));
  b.order = iw.order_bwd; b.stats = g_blend_stats;
  b.snap = bw.snap; b.seg_tile = bw.seg_tile; b.nseg = R > 0 ? (int)((R - 1) / BLEND_SEG) : 0;
  H3_CUDA(launch_blend_bwd(b, s));
  if ((rc = debug_sync(debug, s, "blend_bwd"))) return rc;

  if (side) H3_CUDA(cudaStreamWaitEvent(s, side->join, 0));
  PreBwdArgs pb;
  pb.f = make_pre(view, g, const_cast<int*>(radii), const_cast<void*>(geom));
  pb.grad = ww.grad;
  pb.vis_list = ww.vis_list;
  pb.has_invdepth = dL_dout_invdepth != nullptr;
  pb.out = *grads;
  H3_CUDA(launch_preprocess_bwd(pb, s));
  if ((rc = debug_sync(debug, s, "preprocess_bwd"))) return rc;
  return 0;
}

// one warp per listed row
__global__ void __launch_bounds__(256) zero_rows_kernel(const int32_t* __restrict__ dirty, long long max_entries,
                                                        H3dgsRowTensor t0, H3dgsRowTensor t1, H3dgsRowTensor t2,
                                                        H3dgsRowTensor t3, H3dgsRowTensor t4, H3dgsRowTensor t5,
                                                        H3dgsRowTensor t6, H3dgsRowTensor t7) {
  long long n = dirty[0];
  if (n > max_entries) n = max_entries;
  const int lane = threadIdx.x & 31;
  const long long warps = (long long)gridDim.x * (blockDim.x >> 5);
  const H3dgsRowTensor ts[8] = {t0, t1, t2, t3, t4, t5, t6, t7};
  for (long long w = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); w < n; w += warps) {
    const size_t row = (size_t)dirty[1 + w];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      if (ts[k].data == nullptr) continue;
      float* p = ts[k].data + row * (size_t)ts[k].width;
      for (int i = lane; i < ts[k].width; i += 32) p[i] = 0.f;
    }
  }
}

int h3dgs_zero_rows(const int32_t* dirty_rows, int64_t max_entries, const H3dgsRowTensor* tensors,
                    int32_t num_tensors, h3dgs_stream_t stream) {
  if (!dirty_rows || !tensors || num_tensors < 0 || num_tensors > 8 || max_entries < 0)
    return fail(H3DGS_EINVAL, "h3dgs_zero_rows: bad arguments");
  if (max_entries == 0 || num_tensors == 0) return 0;
  H3dgsRowTensor t[8];
  std::memset(t, 0, sizeof(t));
  for (int i = 0; i < num_tensors; ++i) {
    if (tensors[i].data && tensors[i].width <= 0) return fail(H3DGS_EINVAL, "h3dgs_zero_rows: width must be positive");
    t[i] = tensors[i];
  }
  cudaStream_t s = (cudaStream_t)stream;
  long long blocks = (max_entries + 7) / 8;
  if (blocks > 148 * 16) blocks = 148 * 16;
  ProfScope ps(K_MEMSET, s);
  zero_rows_kernel<<<(unsigned)blocks, 256, 0, s>>>(dirty_rows, max_entries, t[0], t[1], t[2], t[3], t[4], t[5],
                                                    t[6], t[7]);
  H3_CUDA(cudaGetLastError());
  return 0;
}

// One launch zero-fills every span: grid-stride over 16-byte stores, spans walked one after the other.
struct SpanPack {
  H3dgsSpan sp[8];
  int n;
};
__global__ void __launch_bounds__(256) fill_spans_kernel(const SpanPack pk) {
  const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (int k = 0; k < pk.n; ++k) {
    float* p = pk.sp[k].data;
    const size_t n = (size_t)pk.sp[k].count;
    const size_t head = ((16 - (reinterpret_cast<size_t>(p) & 15)) / 4) % 4;   // floats up to 16-byte alignment
    if (tid < head && tid < n) p[tid] = 0.f;
    if (n <= head) continue;
    float4* q = reinterpret_cast<float4*>(p + head);
    const size_t n4 = (n - head) / 4;
    const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
    size_t i = tid;
    for (; i + 3 * stride < n4; i += 4 * stride) {
      q[i] = z; q[i + stride] = z; q[i + 2 * stride] = z; q[i + 3 * stride] = z;
    }
    for (; i < n4; i += stride) q[i] = z;
    const size_t tail = head + n4 * 4;
    if (tid < n - tail) p[tail + tid] = 0.f;
  }
}

int h3dgs_zero_fill(const H3dgsSpan* spans, int32_t num_spans, int32_t ctas_per_sm, h3dgs_stream_t stream) {
  if (num_spans < 0 || num_spans > 8 || (num_spans > 0 && !spans)) return fail(H3DGS_EINVAL, "h3dgs_zero_fill: 0..8 spans");
  if (ctas_per_sm < 0 || ctas_per_sm > 8) return fail(H3DGS_EINVAL, "h3dgs_zero_fill: ctas_per_sm must be 0..8");
  SpanPack pk;
  pk.n = 0;
  for (int i = 0; i < num_spans; ++i) {
    if (spans[i].count < 0) return fail(H3DGS_EINVAL, "h3dgs_zero_fill: negative count");
    if (!spans[i].data || spans[i].count == 0) continue;
    if (reinterpret_cast<size_t>(spans[i].data) & 3) return fail(H3DGS_EINVAL, "h3dgs_zero_fill: unaligned span");
    pk.sp[pk.n++] = spans[i];
  }
  if (pk.n == 0) return 0;
  int dev = 0, sms = 148;
  if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  cudaStream_t s = (cudaStream_t)stream;
  g_prof.launches_total++;
  ProfScope ps(K_MEMSET, s);
  fill_spans_kernel<<<sms * (ctas_per_sm ? ctas_per_sm : 2), 256, 0, s>>>(pk);
  H3_CUDA(cudaGetLastError());
  return 0;
}

int h3dgs_mark_visible(int32_t P, const float* means3D, const float* viewmatrix, const float* projmatrix,
                       uint8_t* present, h3dgs_stream_t stream) {
  if (P < 0) return fail(H3DGS_EINVAL, "negative P");
  if (P == 0) return 0;
  if (!means3D || !viewmatrix || !projmatrix || !present) return fail(H3DGS_EINVAL, "null argument");
  H3_CUDA(launch_mark_visible(P, means3D, viewmatrix, projmatrix, present, (cudaStream_t)stream));
  return 0;
}

int h3dgs_debug_view(int32_t P, int32_t W, int32_t H, int64_t R, const void* geom, const void* binning,
                     const void* image, H3dgsDebugView* out) {
  if (!out) return fail(H3DGS_EINVAL, "null out");
  std::memset(out, 0, sizeof(*out));
  if (geom) {
    GeomWs gw = GeomWs::make(const_cast<void*>(geom), P);
    out->depths = gw.depth;
    out->tiles_touched = gw.tiles_touched;
    out->point_offsets = gw.offsets;
    out->records = reinterpret_cast<const float*>(gw.rec);
    out->cov3D = gw.cov3D;
    out->tile_rects = reinterpret_cast<const uint16_t*>(gw.tile_rect);
  }
  if (binning && R > 0) {
    BinWs bw = BinWs::make(const_cast<void*>(binning), R, W, H);
    out->point_list = bw.point_list;
    if (geom && image) {
      // the 64-bit keys are not part of the product path any more: materialise them for the parity tests
      GeomWs gw = GeomWs::make(const_cast<void*>(geom), P);
      ImageWs iw = ImageWs::make(const_cast<void*>(image), W, H);
      const int tiles = ((W + TILE_X - 1) / TILE_X) * ((H + TILE_Y - 1) / TILE_Y);
      H3_CUDA(cudaDeviceSynchronize());
      H3_CUDA(launch_fill_keys64(iw.ranges, tiles, bw.point_list, gw.depth, bw.keys64, (cudaStream_t)0));
      H3_CUDA(cudaDeviceSynchronize());
      out->point_list_keys = reinterpret_cast<const uint64_t*>(bw.keys64);
    }
  }
  if (image) {
    ImageWs iw = ImageWs::make(const_cast<void*>(image), W, H);
    out->ranges = reinterpret_cast<const uint32_t*>(iw.ranges);
    out->final_T = iw.final_T;
    out->n_contrib = iw.n_contrib;
    out->order_fwd = iw.order_fwd;
    out->order_bwd = iw.order_bwd;
    out->tile_last = iw.tile_last;
  }
  return 0;
}

int h3dgs_expand_to_size(int32_t N, float target_size, const int32_t* nodes, const float* boxes,
                         const float* viewpoint_dev, int32_t* render_indices, int32_t* parent_indices,
                         int32_t* nodes_for_render_indices, void* workspace, size_t workspace_bytes,
                         int32_t* count_host, h3dgs_stream_t stream) {
  if (N < 0) return fail(H3DGS_EINVAL, "negative node count");
  if (!count_host) return fail(H3DGS_EINVAL, "count_host is required");
  *count_host = 0;
  if (N == 0) return 0;
  if (!nodes || !boxes || !viewpoint_dev || !render_indices || !workspace)
    return fail(H3DGS_EINVAL, "nodes, boxes, viewpoint, render_indices and workspace are required");
  if (reinterpret_cast<size_t>(boxes) & 15) return fail(H3DGS_EINVAL, "boxes must be 16-byte aligned");
  if (workspace_bytes < expand_ws_bytes(N)) return fail(H3DGS_ENOSPACE, "expand workspace too small");
  H3_CUDA(run_expand_to_size(N, target_size, nodes, boxes, viewpoint_dev, render_indices, parent_indices,
                             nodes_for_render_indices, workspace, count_host, (cudaStream_t)stream));
  return 0;
}

int h3dgs_interpolation_weights(int32_t K, const int32_t* node_indices, float target_size,
                                const int32_t* nodes, const float* boxes, float vx, float vy, float vz,
                                float* ts, int32_t* kids, h3dgs_stream_t stream) {
  if (K < 0) return fail(H3DGS_EINVAL, "negative entry count");
  if (K == 0) return 0;
  if (!node_indices || !nodes || !boxes || !ts || !kids) return fail(H3DGS_EINVAL, "null argument");
  if (reinterpret_cast<size_t>(boxes) & 15) return fail(H3DGS_EINVAL, "boxes must be 16-byte aligned");
  H3_CUDA(run_interpolation_weights(K, node_indices, target_size, nodes, boxes, vx, vy, vz, ts, kids,
                                    (cudaStream_t)stream));
  return 0;
}

int h3dgs_sparse_adam(const H3dgsAdamTensor* tensors, int32_t num_tensors, const int64_t* relevant,
                      int64_t num_relevant, int64_t num_rows, h3dgs_stream_t stream) {
  if (!tensors || num_tensors <= 0 || num_tensors > H3DGS_ADAM_MAX_TENSORS)
    return fail(H3DGS_EINVAL, "1..H3DGS_ADAM_MAX_TENSORS tensors required");
  for (int i = 0; i < num_tensors; ++i) {
    if (!tensors[i].param || !tensors[i].grad || !tensors[i].exp_avg || !tensors[i].exp_avg_sq || tensors[i].width <= 0)
      return fail(H3DGS_EINVAL, "adam tensor with null pointer or non-positive width");
  }
  if (num_relevant < 0 || num_rows < 0) return fail(H3DGS_EINVAL, "negative row count");
  H3_CUDA(run_sparse_adam(tensors, num_tensors, reinterpret_cast<const long long*>(relevant), num_relevant,
                          num_rows, (cudaStream_t)stream));
  return 0;
}

int h3dgs_sparse_adam_compact(const H3dgsAdamTensor* tensors, int32_t num_tensors, const int64_t* compact_rows,
                              const int32_t* compact_count, int32_t capacity, const float* gate,
                              h3dgs_stream_t stream) {
  if (!tensors || num_tensors <= 0 || num_tensors > H3DGS_ADAM_MAX_TENSORS)
    return fail(H3DGS_EINVAL, "1..H3DGS_ADAM_MAX_TENSORS tensors required");
  for (int i = 0; i < num_tensors; ++i) {
    if (!tensors[i].param || !tensors[i].grad || !tensors[i].exp_avg || !tensors[i].exp_avg_sq || tensors[i].width <= 0)
      return fail(H3DGS_EINVAL, "adam tensor with null pointer or non-positive width");
  }
  if (capacity < 0) return fail(H3DGS_EINVAL, "negative capacity");
  if (capacity == 0) return 0;
  if (!compact_rows || !compact_count) return fail(H3DGS_EINVAL, "compact_rows and compact_count are required");
  H3_CUDA(run_sparse_adam_compact(tensors, num_tensors, reinterpret_cast<const long long*>(compact_rows), compact_count,
                                  capacity, gate, (cudaStream_t)stream));
  return 0;
}

static int pack_rows_common(const H3dgsRowTensor* tensors, int32_t n, const int64_t* rows, int64_t K, float* packed,
                            float scale, bool pack, h3dgs_stream_t stream) {
  if (!tensors || n <= 0 || n > 8) return fail(H3DGS_EINVAL, "pack/unpack rows: 1..8 tensors required");
  if (K < 0) return fail(H3DGS_EINVAL, "pack/unpack rows: negative row count");
  if (K == 0) return 0;
  if (!rows || !packed) return fail(H3DGS_EINVAL, "pack/unpack rows: rows and packed are required");
  for (int i = 0; i < n; ++i)
    if (!tensors[i].data || tensors[i].width <= 0) return fail(H3DGS_EINVAL, "pack/unpack rows: null tensor or non-positive width");
  H3_CUDA(run_pack_rows(tensors, n, reinterpret_cast<const long long*>(rows), K, packed, scale, pack,
                        (cudaStream_t)stream));
  return 0;
}
int h3dgs_pack_rows(const H3dgsRowTensor* tensors, int32_t num_tensors, const int64_t* rows, int64_t num_rows,
                    float* packed, h3dgs_stream_t stream) {
  return pack_rows_common(tensors, num_tensors, rows, num_rows, packed, 1.0f, true, stream);
}
int h3dgs_unpack_rows(const H3dgsRowTensor* tensors, int32_t num_tensors, const int64_t* rows, int64_t num_rows,
                      const float* packed, float scale, h3dgs_stream_t stream) {
  return pack_rows_common(tensors, num_tensors, rows, num_rows, const_cast<float*>(packed), scale, false, stream);
}

int h3dgs_ssim_forward(int32_t C, int32_t H, int32_t W, const float* img1, const float* img2, const float* mask,
                       const float* window11, float* ssim_map, float* dm_dmu1, float* dm_de11, float* dm_de12, double* sums,
                       h3dgs_stream_t stream) {
  if (C < 0 || H < 0 || W < 0) return fail(H3DGS_EINVAL, "h3dgs_ssim_forward: negative size");
  if (C == 0 || H == 0 || W == 0) return 0;
  if (C > 65535) return fail(H3DGS_EINVAL, "h3dgs_ssim_forward: more than 65535 channels");
  if (!img1 || !img2 || !window11 || !sums) return fail(H3DGS_EINVAL, "h3dgs_ssim_forward: img1, img2, window and sums are required");
  if ((dm_dmu1 != nullptr) != (dm_de11 != nullptr) || (dm_dmu1 != nullptr) != (dm_de12 != nullptr))
    return fail(H3DGS_EINVAL, "h3dgs_ssim_forward: pass all three derivative maps or none");
  H3_CUDA(run_ssim_forward(C, H, W, img1, img2, mask, window11, ssim_map, dm_dmu1, dm_de11, dm_de12, sums,
                           (cudaStream_t)stream));
  return 0;
}

int h3dgs_ssim_backward(int32_t C, int32_t H, int32_t W, const float* img1, const float* img2, const float* mask,
                        const float* window11, const float* dm_dmu1, const float* dm_de11, const float* dm_de12, const float* dL_dmap,
                        const float* upstream, float ssim_scale, float l1_scale, float* dL_dimg1,
                        h3dgs_stream_t stream) {
  if (C < 0 || H < 0 || W < 0) return fail(H3DGS_EINVAL, "h3dgs_ssim_backward: negative size");
  if (C == 0 || H == 0 || W == 0) return 0;
  if (C > 65535) return fail(H3DGS_EINVAL, "h3dgs_ssim_backward: more than 65535 channels");
  if (!img1 || !img2 || !window11 || !dm_dmu1 || !dm_de11 || !dm_de12 || !dL_dimg1)
    return fail(H3DGS_EINVAL, "h3dgs_ssim_backward: null argument");
  H3_CUDA(run_ssim_backward(C, H, W, img1, img2, mask, window11, dm_dmu1, dm_de11, dm_de12, dL_dmap, upstream, ssim_scale,
                            l1_scale, dL_dimg1, (cudaStream_t)stream));
  return 0;
}

int h3dgs_masked_l1_forward(int64_t n, const float* pred, const float* target, const float* mask, double* sum,
                            h3dgs_stream_t stream) {
  if (n < 0) return fail(H3DGS_EINVAL, "h3dgs_masked_l1_forward: negative size");
  if (n == 0) return 0;
  if (!pred || !target || !sum) return fail(H3DGS_EINVAL, "h3dgs_masked_l1_forward: pred, target and sum are required");
  H3_CUDA(run_masked_l1(n, pred, target, mask, sum, nullptr, 0.f, nullptr, (cudaStream_t)stream));
  return 0;
}

int h3dgs_masked_l1_backward(int64_t n, const float* pred, const float* target, const float* mask,
                             const float* upstream, float scale, float* dL_dpred, h3dgs_stream_t stream) {
  if (n < 0) return fail(H3DGS_EINVAL, "h3dgs_masked_l1_backward: negative size");
  if (n == 0) return 0;
  if (!pred || !target || !dL_dpred)
    return fail(H3DGS_EINVAL, "h3dgs_masked_l1_backward: pred, target and dL_dpred are required");
  H3_CUDA(run_masked_l1(n, pred, target, mask, nullptr, upstream, scale, dL_dpred, (cudaStream_t)stream));
  return 0;
}

int h3dgs_blend_stats(void* counters) {
  g_blend_stats = reinterpret_cast<unsigned long long*>(counters);
  return 0;
}

int h3dgs_measure_peaks(double* out4, void* scratch, int32_t reps, h3dgs_stream_t stream) {
  if (!out4 || !scratch || reps < 1 || reps > 100) return fail(H3DGS_EINVAL, "h3dgs_measure_peaks: out4, scratch and 1..100 reps");
  H3_CUDA(run_measure_peaks(out4, reinterpret_cast<float*>(scratch), reps, (cudaStream_t)stream));
  return 0;
}

int h3dgs_debug_set(int key, int value) {
  if (key == 1) return h3dgs::set_bwd_variant(value);
  if (key == 2) {
    const int prev = h3dgs::g_strip_key_debug;
    h3dgs::g_strip_key_debug = value & 3;
    return prev;
  }
  return -1;
}
int64_t h3dgs_launch_count(void) { return g_prof.launches_total; }
int h3dgs_profile_enable(int enable) {
  const int prev = g_prof.on ? 1 : 0;
  g_prof.on = enable != 0;
  return prev;
}
int h3dgs_profile_read(double* ms, int64_t* launches, int reset) {
  for (auto& p : g_prof.pending) {
    H3_CUDA(cudaEventSynchronize(p.b));
    float t = 0.f;
    H3_CUDA(cudaEventElapsedTime(&t, p.a, p.b));
    g_prof.ms[p.id] += t;
    g_prof.count[p.id] += 1;
    g_prof.pool.push_back(p.a);
    g_prof.pool.push_back(p.b);
  }
  g_prof.pending.clear();
  for (int i = 0; i < K_COUNT; ++i) {
    if (ms) ms[i] = g_prof.ms[i];
    if (launches) launches[i] = g_prof.count[i];
    if (reset) { g_prof.ms[i] = 0; g_prof.count[i] = 0; }
  }
  return 0;
}
const char* h3dgs_profile_name(int id) { return (id >= 0 && id < K_COUNT) ? kKernelNames[id] : ""; }

}  // extern "C"
