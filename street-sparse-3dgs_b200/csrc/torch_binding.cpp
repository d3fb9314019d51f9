// The torch extension of the drop-in packages: the ONLY translation unit that sees torch.  It allocates outputs and
// workspaces with the caching allocator, takes torch's current CUDA stream and calls the C ABI of
// include/h3dgs_b200.h (libh3dgs_b200.so) -- one C++ call per forward and one per backward, where the ctypes
// binding (h3dgs_native.py, kept as the reference binding of the ABI and used by the tests of the ABI itself)
// spends ~100 us of interpreter time per call marshalling pointers.
//
// Interface being replaced: RasterizeGaussiansCUDA / RasterizeGaussiansBackwardCUDA
// (HR/rasterize_points.h:18-73, HR/rasterize_points.cu:35-242).  Same argument order for the reference's part of
// the signature; the trailing arguments (skybox count, raw-store tensors and activation codes, gradient layout) are
// this repository's extensions and default to the reference behaviour.
#include <torch/extension.h>

#include <ATen/cuda/CUDAContext.h>
#include <ATen/cuda/CUDAEvent.h>
#include <c10/cuda/CUDACachingAllocator.h>
#include <c10/cuda/CUDAGuard.h>
#include <c10/cuda/CUDAStream.h>
#include <cuda_runtime.h>

#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <tuple>
#include <vector>

#include "../../include/h3dgs_b200.h"

namespace {

using torch::Tensor;

inline bool absent(const Tensor& t) { return !t.defined() || t.numel() == 0; }

void check_rc(int rc, const char* what) {
  if (rc == 0) return;
  // same exception type the ctypes binding raises from (RuntimeError subclass there, RuntimeError here)
  TORCH_CHECK(false, what, " failed (code ", rc, "): ", h3dgs_last_error());
}

// contiguous fp32 / int32 CUDA tensor or "absent"; keeps the (possibly copied) tensor alive in `keep`
const float* f32(const Tensor& t, const char* name, std::vector<Tensor>& keep) {
  if (absent(t)) return nullptr;
  TORCH_CHECK_TYPE(t.scalar_type() == torch::kFloat32, name, " must be float32, got ", t.scalar_type());
  TORCH_CHECK_VALUE(t.is_cuda(), name, " must be a CUDA tensor");
  keep.push_back(t.contiguous());
  return keep.back().data_ptr<float>();
}
const int32_t* i32(const Tensor& t, const char* name, std::vector<Tensor>& keep) {
  if (absent(t)) return nullptr;
  TORCH_CHECK_TYPE(t.scalar_type() == torch::kInt32, name, " must be int32, got ", t.scalar_type());
  TORCH_CHECK_VALUE(t.is_cuda(), name, " must be a CUDA tensor");
  keep.push_back(t.contiguous());
  return keep.back().data_ptr<int32_t>();
}

struct Scene {
  H3dgsView view;
  H3dgsGaussians g;
  std::vector<Tensor> keep;
  bool fused;
};

// Same checks and the same struct contents as diff_gaussian_rasterization/_C.py::_scene.
void make_scene(Scene& sc, const Tensor& bg, const Tensor& indices, const Tensor& parent_indices, const Tensor& ts,
                const Tensor& kids, const Tensor& means3D, const Tensor& colors, const Tensor& opacity,
                const Tensor& scales, const Tensor& rotations, double scale_modifier, const Tensor& cov3D_precomp,
                const Tensor& viewmatrix, const Tensor& projmatrix, double tan_fovx, double tan_fovy, int64_t H,
                int64_t W, const Tensor& sh, int64_t degree, const Tensor& campos, int64_t skyboxnum,
                const Tensor& sh_rest, int64_t act_scale, int64_t act_rot, int64_t act_opac) {
  if (means3D.ndimension() != 2 || means3D.size(1) != 3) AT_ERROR("means3D must have dimensions (num_points, 3)");
  auto& keep = sc.keep;
  const int64_t rows = means3D.size(0);
  sc.fused = !absent(indices);
  const int64_t P = sc.fused ? indices.size(0) + skyboxnum : rows;
  sc.view.image_width = (int32_t)W;
  sc.view.image_height = (int32_t)H;
  sc.view.tan_fovx = (float)tan_fovx;
  sc.view.tan_fovy = (float)tan_fovy;
  sc.view.viewmatrix = f32(viewmatrix, "viewmatrix", keep);
  sc.view.projmatrix = f32(projmatrix, "projmatrix", keep);
  sc.view.campos = f32(campos, "campos", keep);
  sc.view.background = f32(bg, "bg", keep);
  int64_t M = absent(sh) ? 0 : sh.size(1);
  if (!absent(sh_rest)) {
    TORCH_CHECK(!absent(sh) && sh.size(1) == 1 && sh_rest.size(0) == sh.size(0),
                "sh_rest needs sh of shape (rows, 1, 3) with the same number of rows");
    M = 1 + sh_rest.size(1);
  }
  H3dgsGaussians& g = sc.g;
  std::memset(&g, 0, sizeof(g));
  g.P = (int32_t)P;
  g.store_rows = (int32_t)rows;
  g.sh_degree = (int32_t)degree;
  g.sh_coeffs = (int32_t)M;
  g.skyboxnum = (int32_t)skyboxnum;
  g.scale_modifier = (float)scale_modifier;
  g.means3D = f32(means3D, "means3D", keep);
  g.shs = f32(sh, "sh", keep);
  g.colors_precomp = f32(colors, "colors_precomp", keep);
  g.opacities = f32(opacity, "opacities", keep);
  g.scales = f32(scales, "scales", keep);
  g.rotations = f32(rotations, "rotations", keep);
  g.cov3D_precomp = f32(cov3D_precomp, "cov3D_precomp", keep);
  g.render_indices = sc.fused ? i32(indices, "render_indices", keep) : nullptr;
  g.parent_indices = sc.fused ? i32(parent_indices, "parent_indices", keep) : nullptr;
  g.ts = f32(ts, "interpolation_weights", keep);
  g.kids = i32(kids, "num_node_kids", keep);
  g.shs_rest = absent(sh_rest) ? nullptr : f32(sh_rest, "sh_rest", keep);
  g.scale_activation = (int32_t)act_scale;
  g.rotation_activation = (int32_t)act_rot;
  g.opacity_activation = (int32_t)act_opac;
  if (sc.fused && !absent(parent_indices) && parent_indices.size(0) < indices.size(0))
    TORCH_CHECK(false, "parent_indices is shorter than render_indices");
  if (!absent(ts) && ts.size(0) < P - skyboxnum)
    TORCH_CHECK(false, "interpolation_weights must have at least one entry per rendered Gaussian");
  if (!absent(kids) && kids.size(0) < P - skyboxnum)
    TORCH_CHECK(false, "num_node_kids must have at least one entry per rendered Gaussian");
}

// Pinned host words + event through which (num_rendered, num_visible) reach the host; a small free list so that
// renders in flight on different streams each hold their own.
struct Counts {
  int64_t* host = nullptr;
  cudaEvent_t event = nullptr;
};
std::mutex g_counts_mutex;
std::vector<Counts> g_counts_free;

Counts get_counts() {
  {
    std::lock_guard<std::mutex> lock(g_counts_mutex);
    if (!g_counts_free.empty()) {
      Counts c = g_counts_free.back();
      g_counts_free.pop_back();
      return c;
    }
  }
  Counts c;
  C10_CUDA_CHECK(cudaHostAlloc(reinterpret_cast<void**>(&c.host), 2 * sizeof(int64_t), cudaHostAllocDefault));
  C10_CUDA_CHECK(cudaEventCreateWithFlags(&c.event, cudaEventDisableTiming));
  return c;
}
void put_counts(const Counts& c) {
  std::lock_guard<std::mutex> lock(g_counts_mutex);
  g_counts_free.push_back(c);
}

}  // namespace

namespace {

// Gradient tensors of a fused-cut render, allocated by the FORWARD and zero-filled on a side stream forked right
// before the blend kernel (same scheme as _C.py::_Prezero): the 4 GB fill of a 17 M-row store then runs under the
// issue-bound forward blend instead of in front of the backward kernels.
struct Prezero {
  std::vector<Tensor> tensors;   // means3D, sh, sh_rest, colors, opacity, scales, rot, cov3D (undefined = absent)
  Tensor means2D;
  std::unique_ptr<at::cuda::CUDAEvent> done;
};

struct GradShapes {
  std::vector<int64_t> shape[8];
  bool present[8];
};
// index: 0 means3D, 1 sh, 2 sh_rest, 3 colors, 4 opacity, 5 scales, 6 rot, 7 cov3D
GradShapes grad_shapes(int64_t rows, int64_t M, bool has_sh, bool split, bool has_colors, bool has_scales,
                       bool has_rot, bool has_cov) {
  GradShapes s;
  s.shape[0] = {rows, 3};                      s.present[0] = true;
  s.shape[1] = {rows, split ? 1 : M, 3};       s.present[1] = has_sh;
  s.shape[2] = {rows, M - 1, 3};               s.present[2] = split;
  s.shape[3] = {rows, 3};                      s.present[3] = has_colors;
  s.shape[4] = {rows, 1};                      s.present[4] = true;
  s.shape[5] = {rows, 3};                      s.present[5] = has_scales;
  s.shape[6] = {rows, 4};                      s.present[6] = has_rot;
  s.shape[7] = {rows, 6};                      s.present[7] = has_cov;
  return s;
}

using ForwardResult = std::tuple<int64_t, Tensor, Tensor, Tensor, Tensor, Tensor, Tensor, int64_t,
                                 std::shared_ptr<Prezero>>;

// -> (num_rendered, color[3,H,W], radii[P], geomBuffer, binningBuffer, imgBuffer, invdepth, num_visible, prezero)
ForwardResult rasterize_forward(const Tensor& bg, const Tensor& indices, const Tensor& parent_indices,
                                const Tensor& ts, const Tensor& kids, const Tensor& means3D, const Tensor& colors,
                                const Tensor& opacity, const Tensor& scales, const Tensor& rotations,
                                double scale_modifier, const Tensor& cov3D_precomp, const Tensor& viewmatrix,
                                const Tensor& projmatrix, double tan_fovx, double tan_fovy, int64_t image_height,
                                int64_t image_width, const Tensor& sh, int64_t degree, const Tensor& campos,
                                bool prefiltered, bool debug, bool do_depth, int64_t skyboxnum, const Tensor& sh_rest,
                                int64_t act_scale, int64_t act_rot, int64_t act_opac, int64_t prezero_means2D_rows) {
  (void)prefiltered;
  Scene sc;
  make_scene(sc, bg, indices, parent_indices, ts, kids, means3D, colors, opacity, scales, rotations, scale_modifier,
             cov3D_precomp, viewmatrix, projmatrix, tan_fovx, tan_fovy, image_height, image_width, sh, degree, campos,
             skyboxnum, sh_rest, act_scale, act_rot, act_opac);
  const c10::cuda::CUDAGuard guard(means3D.device());
  const auto stream = c10::cuda::getCurrentCUDAStream();
  const int64_t P = sc.g.P, H = image_height, W = image_width;
  const auto bytes = means3D.options().dtype(torch::kUInt8);
  const auto f32o = means3D.options().dtype(torch::kFloat32);
  Tensor radii = torch::empty({P}, means3D.options().dtype(torch::kInt32));
  Tensor geom = torch::empty({(int64_t)h3dgs_geom_bytes((int32_t)P)}, bytes);
  Counts counts = get_counts();
  check_rc(h3dgs_forward_begin(&sc.view, &sc.g, P ? radii.data_ptr<int32_t>() : nullptr, geom.data_ptr(),
                               (size_t)geom.numel(), counts.host, counts.event, debug ? 1 : 0, stream.stream()),
           "h3dgs_forward_begin");
  // everything that does not need R is allocated while the preprocess kernels run
  Tensor img = torch::empty({(int64_t)h3dgs_image_bytes((int32_t)W, (int32_t)H)}, bytes);
  Tensor out_color = torch::empty({3, H, W}, f32o);
  Tensor out_invdepth = torch::empty({do_depth ? 1 : 0, H, W}, f32o);
  // wait for the counts alone: the depth sort of the visible slots is already queued behind them
  C10_CUDA_CHECK(cudaEventSynchronize(counts.event));
  const int64_t R = counts.host[0], V = counts.host[1];
  put_counts(counts);
  TORCH_CHECK(R < (int64_t(1) << 30), "more than 2^30 tile instances");
  Tensor binning = torch::empty({R > 0 ? (int64_t)h3dgs_binning_bytes(R, (int32_t)W, (int32_t)H) : 0}, bytes);

  std::shared_ptr<Prezero> pre;
  std::unique_ptr<at::cuda::CUDAEvent> fork;
  if (prezero_means2D_rows >= 0 && sc.fused && R > 0 && P > 0) {
    // allocated BEFORE the fork point is recorded: whatever used these blocks last was enqueued earlier
    pre = std::make_shared<Prezero>();
    const bool split = sc.g.shs_rest != nullptr;
    GradShapes gs = grad_shapes(sc.g.store_rows, sc.g.sh_coeffs, sc.g.shs != nullptr, split,
                                sc.g.colors_precomp != nullptr, sc.g.scales != nullptr, sc.g.rotations != nullptr,
                                sc.g.cov3D_precomp != nullptr);
    pre->tensors.resize(8);
    for (int i = 0; i < 8; ++i)
      if (gs.present[i]) pre->tensors[i] = torch::empty(gs.shape[i], f32o);
    pre->means2D = torch::empty({std::max<int64_t>(prezero_means2D_rows, P), 3}, f32o);
    fork = std::make_unique<at::cuda::CUDAEvent>();
    fork->record(stream);   // materialises the event; the library re-records it in front of the blend kernel
  }
  check_rc(h3dgs_forward_finish(&sc.view, &sc.g, geom.data_ptr(), R > 0 ? binning.data_ptr() : nullptr,
                                (size_t)binning.numel(), img.data_ptr(), R, V, out_color.data_ptr<float>(),
                                do_depth ? out_invdepth.data_ptr<float>() : nullptr, debug ? 1 : 0, stream.stream(),
                                fork ? (void*)fork->event() : nullptr),
           "h3dgs_forward_finish");
  if (pre) {
    auto side = c10::cuda::getStreamFromPool(false, stream.device_index());
    fork->block(side);
    H3dgsSpan spans[9];
    int n = 0;
    for (auto& t : pre->tensors)
      if (t.defined() && t.numel()) spans[n++] = H3dgsSpan{t.data_ptr<float>(), t.numel()};
    if (pre->means2D.size(0) > P)   // rows of a store-sized means2D beyond the rendered slots get no gradient
      spans[n++] = H3dgsSpan{pre->means2D.data_ptr<float>() + P * 3, (pre->means2D.size(0) - P) * 3};
    for (int i = 0; i < n; i += 8)
      check_rc(h3dgs_zero_fill(spans + i, std::min(8, n - i), 0, side.stream()), "h3dgs_zero_fill");
    for (auto& t : pre->tensors)
      if (t.defined()) c10::cuda::CUDACachingAllocator::recordStream(t.storage().data_ptr(), side);
    c10::cuda::CUDACachingAllocator::recordStream(pre->means2D.storage().data_ptr(), side);
    pre->done = std::make_unique<at::cuda::CUDAEvent>();
    pre->done->record(side);
  }
  return ForwardResult(R, out_color, radii, geom, binning, img, out_invdepth, V, pre);
}

}  // namespace

namespace {

// int32 [rows] scratch of the compact-gradient mode, all -1 between calls (the library restores it); one per device
std::mutex g_map_mutex;
std::map<int, Tensor> g_row_maps;
Tensor row_map(const torch::TensorOptions& like, int device_index, int64_t rows) {
  std::lock_guard<std::mutex> lock(g_map_mutex);
  auto it = g_row_maps.find(device_index);
  if (it == g_row_maps.end() || it->second.size(0) != rows) {
    g_row_maps.erase(device_index);
    it = g_row_maps.emplace(device_index, torch::full({rows}, -1, like.dtype(torch::kInt32))).first;
  }
  return it->second;
}

// -> [dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dsh_rest, dL_dscales, dL_drotations,
//     compact_rows, compact_count]; the reference's eight gradients in its order (HR/rasterize_points.cu:138-242)
//     with dL_dsh_rest inserted and the two compact-mode tensors appended (undefined when not used).
// layout: 0 = dense (reference), 1 = compact (H3dgsGrads.compact_*; the per-row tensors are [2 * num_visible, ...]).
std::vector<Tensor> rasterize_backward(const Tensor& bg, const Tensor& indices, const Tensor& parent_indices,
                                       const Tensor& ts, const Tensor& kids, const Tensor& means3D,
                                       const Tensor& radii, const Tensor& colors, const Tensor& opacities,
                                       const Tensor& scales, const Tensor& rotations, double scale_modifier,
                                       const Tensor& cov3D_precomp, const Tensor& viewmatrix,
                                       const Tensor& projmatrix, double tan_fovx, double tan_fovy,
                                       const Tensor& dL_dout_color, const Tensor& dL_dout_invdepth, const Tensor& sh,
                                       int64_t degree, const Tensor& campos, const Tensor& geomBuffer, int64_t R,
                                       const Tensor& binningBuffer, const Tensor& imageBuffer, bool debug,
                                       int64_t skyboxnum, const Tensor& sh_rest, int64_t act_scale, int64_t act_rot,
                                       int64_t act_opac, std::shared_ptr<Prezero> prezero, int64_t layout,
                                       int64_t num_visible) {
  const int64_t H = dL_dout_color.size(1), W = dL_dout_color.size(2);
  Scene sc;
  make_scene(sc, bg, indices, parent_indices, ts, kids, means3D, colors, opacities, scales, rotations, scale_modifier,
             cov3D_precomp, viewmatrix, projmatrix, tan_fovx, tan_fovy, H, W, sh, degree, campos, skyboxnum, sh_rest,
             act_scale, act_rot, act_opac);
  const c10::cuda::CUDAGuard guard(means3D.device());
  const auto stream = c10::cuda::getCurrentCUDAStream();
  const int64_t P = sc.g.P, rows = sc.g.store_rows;
  const bool split = sc.g.shs_rest != nullptr;
  const auto f32o = means3D.options().dtype(torch::kFloat32);
  const bool compact = layout == 1;
  TORCH_CHECK(!compact || sc.fused, "sparse gradients need a hierarchy-cut render (render_indices)");
  if (prezero && !(sc.fused && P > 0 && R > 0 && prezero->done)) prezero.reset();
  if (compact) prezero.reset();

  const int64_t grad_rows = compact ? std::max<int64_t>(2 * num_visible, 1) : rows;
  GradShapes gs = grad_shapes(grad_rows, sc.g.sh_coeffs, sc.g.shs != nullptr, split, sc.g.colors_precomp != nullptr,
                              sc.g.scales != nullptr, sc.g.rotations != nullptr, sc.g.cov3D_precomp != nullptr);
  std::vector<Tensor> t(8);
  Tensor dL_dmeans2D;
  if (prezero) {
    prezero->done->block(stream);
    t = prezero->tensors;
    dL_dmeans2D = prezero->means2D;
  } else {
    for (int i = 0; i < 8; ++i)
      if (gs.present[i]) t[i] = torch::empty(gs.shape[i], f32o);
    dL_dmeans2D = torch::empty({P, 3}, f32o);
  }
  auto ptr = [](const Tensor& x) -> float* { return (x.defined() && x.numel()) ? x.data_ptr<float>() : nullptr; };
  H3dgsGrads grads;
  std::memset(&grads, 0, sizeof(grads));
  grads.dL_dmeans2D = ptr(dL_dmeans2D);
  grads.dL_dmeans3D = ptr(t[0]);
  grads.dL_dsh = ptr(t[1]);
  grads.dL_dsh_rest = ptr(t[2]);
  grads.dL_dcolors = ptr(t[3]);
  grads.dL_dopacity = ptr(t[4]);
  grads.dL_dscales = ptr(t[5]);
  grads.dL_drotations = ptr(t[6]);
  grads.dL_dcov3D = ptr(t[7]);
  grads.rows_prezeroed = prezero ? 1 : 0;
  Tensor compact_rows, compact_count, map;
  if (compact) {
    compact_rows = torch::empty({grad_rows}, means3D.options().dtype(torch::kInt64));
    compact_count = torch::empty({1}, means3D.options().dtype(torch::kInt32));
    map = row_map(means3D.options(), stream.device_index(), rows);
    grads.compact_rows = compact_rows.data_ptr<int64_t>();
    grads.compact_count = compact_count.data_ptr<int32_t>();
    grads.row_map = map.data_ptr<int32_t>();
    grads.compact_capacity = (int32_t)grad_rows;
  }
  Tensor ws = torch::empty({(int64_t)h3dgs_backward_workspace_bytes((int32_t)P)}, means3D.options().dtype(torch::kUInt8));
  std::vector<Tensor> keep;
  const float* dcol = f32(dL_dout_color, "dL_dout_color", keep);
  const float* ddep = f32(dL_dout_invdepth, "dL_dout_invdepth", keep);
  auto raw = [](const Tensor& x) -> const void* { return (x.defined() && x.numel()) ? x.data_ptr() : nullptr; };
  check_rc(h3dgs_backward(&sc.view, &sc.g, absent(radii) ? nullptr : radii.data_ptr<int32_t>(), raw(geomBuffer),
                          raw(binningBuffer), raw(imageBuffer), R, dcol, ddep, &grads, ws.data_ptr(),
                          (size_t)ws.numel(), debug ? 1 : 0, stream.stream()),
           "h3dgs_backward");
  // gradients of absent optional inputs come back empty, as from the ctypes binding
  auto or_empty = [&](const Tensor& x, std::vector<int64_t> shape) { return x.defined() ? x : torch::empty(shape, f32o); };
  return {dL_dmeans2D, or_empty(t[3], {0, 3}), t[4], t[0], or_empty(t[7], {0, 6}), or_empty(t[1], {0, 0, 3}),
          t[2].defined() ? t[2] : Tensor(), or_empty(t[5], {0, 3}), or_empty(t[6], {0, 4}), compact_rows, compact_count};
}

}  // namespace

PYBIND11_MODULE(TORCH_EXTENSION_NAME, m) {
  m.doc() = "torch extension over the C ABI of include/h3dgs_b200.h (forward / backward of the rasterizer)";
  py::class_<Prezero, std::shared_ptr<Prezero>>(m, "Prezero")
      .def_property_readonly("ready", [](const Prezero& p) { return p.done != nullptr; });
  m.def("rasterize_forward", &rasterize_forward);
  m.def("rasterize_backward", &rasterize_backward);
  m.def("row_map", [](const Tensor& like, int64_t rows) { return row_map(like.options(), like.get_device(), rows); });
  m.def("abi_version", []() { return h3dgs_abi_version(); });
}
