/*
 * h3dgs_b200.h -- C ABI of the B200-native hierarchical Gaussian rasterizer, hierarchy cut and
 * sparse-Adam step.  Plain pointers and sizes only: no torch, no C++ types.
 *
 * Every entry point names the reference interface it replaces (paths relative to the reference
 * checkout; HR = submodules/hierarchy-rasterizer, GH = submodules/gaussianhierarchy).
 *
 * Conventions
 *   - all device pointers are contiguous fp32 / int32 unless stated otherwise;
 *   - a NULL pointer means "absent" exactly where the reference accepts an empty tensor
 *     (HR/rasterize_points.cu:109-120);
 *   - every function returns 0 on success, a negative H3DGS_E* code on a usage error, or a positive
 *     cudaError_t value; h3dgs_last_error() returns a static description of the last failure
 *     of the calling thread;
 *   - nothing here allocates device memory: workspaces are sized by the *_bytes() queries and owned
 *     by the caller (mirrors required<T>(P), HR/cuda_rasterizer/rasterizer_impl.h:68-74);
 *   - all kernels are enqueued on the stream that is passed in (the reference used legacy stream 0,
 *     HR/rasterize_points.cu:102-133).
 */
#ifndef H3DGS_B200_H_
#define H3DGS_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define H3DGS_ABI_VERSION 8

#if defined(__GNUC__)
#define H3DGS_API __attribute__((visibility("default")))
#else
#define H3DGS_API
#endif

#define H3DGS_EINVAL    (-1)  /* bad argument (NULL where required, negative size, ...) */
#define H3DGS_ENOSPACE  (-2)  /* a caller-provided workspace is too small */
#define H3DGS_EIO       (-3)  /* file could not be opened / short read */
#define H3DGS_ELOSSY    (-4)  /* write_hierarchy: node counters exceed the fp16 container ("Would lose information!") */

typedef void* h3dgs_stream_t; /* a cudaStream_t */

H3DGS_API int h3dgs_abi_version(void);
H3DGS_API const char* h3dgs_last_error(void);

/* ------------------------------------------------------------------------------------------------
 * Rasterizer.  Replaces CudaRasterizer::Rasterizer::{forward,backward,markVisible}
 * (HR/cuda_rasterizer/rasterizer.h:21-105, rasterizer_impl.cu:145-510) as driven by
 * RasterizeGaussiansCUDA / RasterizeGaussiansBackwardCUDA (HR/rasterize_points.cu:35-242).
 * ------------------------------------------------------------------------------------------------ */

/* Camera + image: fields of GaussianRasterizationSettings (HR/diff_gaussian_rasterization/__init__.py:146-163). */
typedef struct H3dgsView {
  int32_t image_width, image_height;
  float tan_fovx, tan_fovy;
  const float* viewmatrix; /* device [16], world->view, transposed (column-major when read linearly) */
  const float* projmatrix; /* device [16], full projection, same convention */
  const float* campos;     /* device [3] */
  const float* background; /* device [3] */
} H3dgsView;

/*
 * Gaussians submitted to one render.
 *
 * Plain mode (render_indices == NULL): P == store_rows, slot i reads row i.
 *
 * Fused-cut mode (render_indices != NULL): the attribute arrays are the whole model store
 * [store_rows, ...]; slot i < K = P - skyboxnum renders row render_indices[i] blended with row
 * parent_indices[i] by ts[i] exactly as gaussian_renderer.render_post does in Python
 * (gaussian_renderer/__init__.py:202-235): x = t*x_child + (1-t)*x_parent for means, scales, SH and
 * opacities, and rotations with the parent sign-flipped when dot(child,parent) < 0;
 * parent_indices[i] == -1 means "no parent".  Slots i >= K are the skybox: store rows
 * store_rows - skyboxnum + (i - K), never interpolated (HR/cuda_rasterizer/forward.cu:267-277 gives
 * the same slot layout).  The same arithmetic is what the reference's native index path does in
 * forward.cu:290-343,425; its backward for that path is incomplete (backward.cu:458-494), so the
 * gradient contract here is the Python one.
 *
 * ts / kids are indexed by slot and may be longer than P (train_post.py:62-63).  When both are
 * non-NULL, slots < P - skyboxnum use alpha = t*a + (1-t)*(1-(1-a)^(1/kids)) (forward.cu:537-546).
 */
typedef struct H3dgsGaussians {
  int32_t P;          /* render slots */
  int32_t store_rows; /* rows of the attribute arrays */
  int32_t sh_degree;  /* active degree D, 0..3 */
  int32_t sh_coeffs;  /* M = coefficients per row in shs (16 at degree 3); 0 when shs == NULL */
  int32_t skyboxnum;
  float scale_modifier;
  const float* means3D;        /* [rows,3] */
  const float* shs;            /* [rows,M,3] or NULL */
  const float* colors_precomp; /* [rows,3] or NULL (exactly one of shs / colors_precomp) */
  const float* opacities;      /* [rows,1] */
  const float* scales;         /* [rows,3] or NULL */
  const float* rotations;      /* [rows,4] or NULL */
  const float* cov3D_precomp;  /* [rows,6] or NULL (exactly one of scales+rotations / cov3D_precomp) */
  const int32_t* render_indices; /* [K] or NULL */
  const int32_t* parent_indices; /* [K] or NULL */
  const float* ts;               /* [>=P] or NULL */
  const int32_t* kids;           /* [>=P] or NULL */
  /* Raw parameter store (ABI 2).  The reference activates the WHOLE store on every render
   * (scene/gaussian_model.py:121-145: exp of _scaling, normalize of _rotation, sigmoid/abs of _opacity,
   * torch.cat of _features_dc/_features_rest) before render_post gathers the cut rows from it
   * (gaussian_renderer/__init__.py:202-235).  With these fields the kernels read the raw parameters and
   * apply the same activations to the rows they gather; the backward then returns gradients of the RAW
   * parameters.  All zero / NULL = the arrays above are already activated (ABI 1 behaviour). */
  const float* shs_rest;        /* NULL, or [rows,M-1,3]: then shs is [rows,1,3] (_features_dc/_features_rest) */
  int32_t scale_activation;     /* H3DGS_ACT_NONE | H3DGS_ACT_EXP */
  int32_t rotation_activation;  /* H3DGS_ACT_NONE | H3DGS_ACT_NORMALIZE  (x / max(||x||_2, 1e-12)) */
  int32_t opacity_activation;   /* H3DGS_ACT_NONE | H3DGS_ACT_SIGMOID | H3DGS_ACT_ABS */
} H3dgsGaussians;

#define H3DGS_ACT_NONE      0
#define H3DGS_ACT_EXP       1
#define H3DGS_ACT_NORMALIZE 2
#define H3DGS_ACT_SIGMOID   3
#define H3DGS_ACT_ABS       4

/* Workspace sizes.  geom: per-slot state (GeometryState, rasterizer_impl.cu:159-175) plus the depth sort of
 * the visible slots -- everything that can be sized before the render's instance count is known;
 * image: per-pixel + per-tile state (ImageState, :177-184);
 * binning: instance keys/values + sort workspace (BinningState, :186-199), sized by num_rendered. */
H3DGS_API size_t h3dgs_geom_bytes(int32_t P);
H3DGS_API size_t h3dgs_image_bytes(int32_t image_width, int32_t image_height);
H3DGS_API size_t h3dgs_binning_bytes(int64_t num_rendered, int32_t image_width, int32_t image_height);

/*
 * Forward, stage 1: preprocess (EWA projection, SH, culling, fused interpolation), tile-count scan, and -- with the
 * visible count V taken from DEVICE memory -- compaction of the visible slots, their depth sort and the scan of
 * their tile counts in depth order.  Replaces FORWARD::preprocess + cub::DeviceScan::InclusiveSum + the
 * num_rendered read-back (rasterizer_impl.cu:275-325) and the first half of the sort.  Writes radii[P] (0 = culled).
 *
 * The instance count R (the reference's num_rendered) sizes the binning workspace, so it must reach the host -- but
 * the reference's way (a blocking copy, rasterizer_impl.cu:324-325) leaves the GPU without work while the host
 * wakes up and queues the rest.  Here the copy of counts_host[0] = R, counts_host[1] = V (host memory, pinned for
 * a truly asynchronous copy) is enqueued right after the scan and counts_event (a cudaEvent_t) is recorded behind
 * it; ~0.2 ms of kernels that do not need R on the host follow in the same call.  The caller waits for
 * counts_event, allocates the binning workspace and calls h3dgs_forward_finish while those kernels run.
 * counts_event == NULL: the call synchronises `stream` itself before it returns.
 */
H3DGS_API int h3dgs_forward_begin(const H3dgsView* view, const H3dgsGaussians* g,
                        int32_t* radii, void* geom, size_t geom_bytes,
                        int64_t* counts_host, void* counts_event, int debug, h3dgs_stream_t stream);

/*
 * Forward, stage 2: ordered instance emission, stable onesweep partition by tile (with stage 1's depth sort: the
 * reference's sort by (tile, depth), bit-identical lists), tile ranges, alpha blend.
 * Replaces duplicateWithKeys + cub::DeviceRadixSort::SortPairs + identifyTileRanges + FORWARD::render
 * (rasterizer_impl.cu:336-390).  out_color is [3,H,W]; out_invdepth is [1,H,W] or NULL.
 * num_rendered == 0 only fills out_color with the background.
 * pre_blend_event (a cudaEvent_t, or NULL) is recorded on `stream` right before the blend kernel: the point
 * after which the forward stops using HBM bandwidth, where a caller can fork bandwidth-only side work
 * (h3dgs_zero_fill of the gradient tensors the backward will add into) onto another stream.
 */
H3DGS_API int h3dgs_forward_finish(const H3dgsView* view, const H3dgsGaussians* g,
                         void* geom, void* binning, size_t binning_bytes, void* image,
                         int64_t num_rendered, int64_t num_visible, float* out_color, float* out_invdepth,
                         int debug, h3dgs_stream_t stream, void* pre_blend_event);

/* Gradient outputs of the backward pass; any pointer may be NULL (that gradient is not produced).
 * Row counts: dL_dmeans2D is per slot [P,3]; every other tensor is per store row [store_rows, ...]
 * (in plain mode the two coincide, which is the reference layout, rasterize_points.cu:168-198).
 * All non-NULL outputs are completely overwritten (no caller-side zero fill needed). */
typedef struct H3dgsGrads {
  float* dL_dmeans2D;   /* [P,3]  (z component is always 0) */
  float* dL_dmeans3D;   /* [rows,3] */
  float* dL_dsh;        /* [rows,M,3] */
  float* dL_dcolors;    /* [rows,3]   (gradient of colors_precomp) */
  float* dL_dopacity;   /* [rows,1] */
  float* dL_dscales;    /* [rows,3] */
  float* dL_drotations; /* [rows,4] */
  float* dL_dcov3D;     /* [rows,6]   (gradient of cov3D_precomp) */
  float* dL_dsh_rest;   /* [rows,M-1,3] when g->shs_rest is given (then dL_dsh is [rows,1,3]); else NULL */
  /* Fused-cut mode only (ABI 4), both optional.  The per-store-row outputs are N-row tensors of which a
   * render touches only the rows of its cut; zero-filling all N rows (4 GB at 17 M rows) costs more than
   * the gradient kernel itself.  A caller that recycles its gradient tensors can instead
   *   - set rows_prezeroed = 1: the per-row outputs already hold zeros, skip the zero fill, and
   *   - pass dirty_rows (device int32 [1 + 2*P]): [0] receives the number n of store rows this call added
   *     into, [1..n] their ids (duplicates possible) -- feed it to h3dgs_zero_rows before the next reuse. */
  int32_t* dirty_rows;
  int32_t rows_prezeroed;
  /* Compact gradients (fused-cut mode only, ABI 7; all NULL / 0 = the dense layout above).  A render adds into the
   * rows of its cut only -- a few % of the store -- yet the dense layout costs a zero fill of every row (4 GB at
   * 17 M rows) and the optimizer then scans all N rows to find the touched ones again (train_single.py:225-229:
   * `opacity.grad != 0`, nonzero).  With these fields the per-store-row pointers above address COMPACT tensors
   * [compact_capacity, width] instead: every store row this render touches (the child and parent rows of its
   * visible slots, its visible skybox rows) gets one compact row; compact_rows[i] receives the store row of
   * compact row i and *compact_count their number U (U <= 2 * num_visible, which is therefore always a
   * sufficient capacity).  Compact rows [0, U) are completely written (zero where a tensor gets no contribution);
   * rows beyond U are left untouched.  row_map is scratch: int32 [store_rows], all -1 on entry, all -1 again when
   * the call has finished.  h3dgs_sparse_adam_compact consumes the result without a host round trip. */
  int64_t* compact_rows;
  int32_t* compact_count;
  int32_t* row_map;
  int32_t compact_capacity;
} H3dgsGrads;

H3DGS_API size_t h3dgs_backward_workspace_bytes(int32_t P);

/* Zero the rows listed in dirty_rows (layout above) of up to 8 row-major fp32 tensors [rows,width]. */
typedef struct H3dgsRowTensor {
  float* data;
  int32_t width;
} H3dgsRowTensor;
H3DGS_API int h3dgs_zero_rows(const int32_t* dirty_rows, int64_t max_entries, const H3dgsRowTensor* tensors,
                    int32_t num_tensors, h3dgs_stream_t stream);

/* Zero-fill up to 8 fp32 arrays (data, count floats) with ONE launch of a deliberately small grid
 * (ctas_per_sm resident CTAs per SM, 1..8; 0 = 2), so that the fill can share the SMs with an issue-bound
 * kernel running on another stream (the blend kernels use ~1 % of the HBM bandwidth).  This is the zero
 * fill of the reference's backward (torch::zeros x 9, rasterize_points.cu:168-198) taken off the critical
 * path; pass the filled tensors to h3dgs_backward with rows_prezeroed = 1. */
typedef struct H3dgsSpan {
  float* data;
  int64_t count;
} H3dgsSpan;
H3DGS_API int h3dgs_zero_fill(const H3dgsSpan* spans, int32_t num_spans, int32_t ctas_per_sm,
                    h3dgs_stream_t stream);

/*
 * Backward.  Replaces BACKWARD::render + computeCov2DCUDA + BACKWARD::preprocess
 * (rasterizer_impl.cu:397-510; HR/cuda_rasterizer/backward.cu:147-721).
 * dL_dout_color is [3,H,W]; dL_dout_invdepth is [1,H,W] or NULL.  geom/binning/image are the
 * buffers the forward filled for the same (view, g).
 */
H3DGS_API int h3dgs_backward(const H3dgsView* view, const H3dgsGaussians* g, const int32_t* radii,
                   const void* geom, const void* binning, const void* image, int64_t num_rendered,
                   const float* dL_dout_color, const float* dL_dout_invdepth,
                   const H3dgsGrads* grads, void* workspace, size_t workspace_bytes,
                   int debug, h3dgs_stream_t stream);

/* Near-plane visibility mask; replaces Rasterizer::markVisible / checkFrustum
 * (rasterizer_impl.cu:54-66,145-157).  present is uint8 [P] (torch.bool storage). */
H3DGS_API int h3dgs_mark_visible(int32_t P, const float* means3D, const float* viewmatrix,
                       const float* projmatrix, uint8_t* present, h3dgs_stream_t stream);

/* Parity / debugging view into the opaque buffers (all device pointers into the given buffers).
 * keys/values are the sorted instance list (BinningState::point_list_keys / point_list),
 * ranges is uint2 per tile (ImageState::ranges). */
typedef struct H3dgsDebugView {
  const uint64_t* point_list_keys;
  const uint32_t* point_list;
  const uint32_t* ranges;        /* [tiles,2] */
  const float* final_T;          /* [H*W] */
  const uint32_t* n_contrib;     /* [H*W] */
  const float* depths;           /* [P] */
  const uint32_t* tiles_touched; /* [P] */
  const uint32_t* point_offsets; /* [P] inclusive rank of the slot among the visible slots */
  const float* records;          /* [P,16] packed per-slot record, see DESIGN.md */
  const float* cov3D;            /* [P,6] */
  const uint16_t* tile_rects;    /* [P,4] x0,y0,x1,y1 in tiles */
  const uint32_t* order_fwd;     /* [tiles] launch order of the forward blend (heaviest tile first) */
  const uint32_t* order_bwd;     /* [tiles] launch order of the backward blend (valid after a backward) */
  const uint32_t* tile_last;     /* [tiles] max n_contrib over the tile's pixels */
} H3dgsDebugView;
H3DGS_API int h3dgs_debug_view(int32_t P, int32_t image_width, int32_t image_height, int64_t num_rendered,
                     const void* geom, const void* binning, const void* image, H3dgsDebugView* out);

/* ------------------------------------------------------------------------------------------------
 * Hierarchy cut.  Replaces Switching::expandToSize / getTsIndexed (GH/runtime_switching.cu:402-540)
 * and Traversal::expandToTarget (GH/traversal.cpp:15-38) as bound by GH/torch/torch_interface.cpp:69-120.
 * nodes is int32 [N,7] = (depth,parent,start,count_leafs,count_merged,start_children,count_children)
 * (GH/types.h:51-60); boxes is fp32 [N,2,4] (GH/types.h:43-47).
 * ------------------------------------------------------------------------------------------------ */
H3DGS_API size_t h3dgs_expand_workspace_bytes(int32_t num_nodes);

/* viewpoint is a DEVICE pointer to 3 floats (as in the reference, torch_interface.cpp:92).
 * Writes render_indices / parent_indices / nodes_for_render_indices (each may be NULL except
 * render_indices) and, after synchronising `stream`, the number of entries to *count_host. */
H3DGS_API int h3dgs_expand_to_size(int32_t num_nodes, float target_size, const int32_t* nodes,
                         const float* boxes, const float* viewpoint_dev,
                         int32_t* render_indices, int32_t* parent_indices,
                         int32_t* nodes_for_render_indices,
                         void* workspace, size_t workspace_bytes,
                         int32_t* count_host, h3dgs_stream_t stream);

/* h3dgs_expand_to_size without a workspace argument and without any allocation.  NO COUNTERPART IN THE REFERENCE:
 * added for callers that cannot call cudaMalloc (code under CUDA graph
 * capture, tight serving loops, harnesses that forbid it).  The scratch (two u32 per node + scan look-back words,
 * 8.0 MiB) is a file-scope static __device__ array in the library, resident on every GPU that loads it until exit;
 * the call rejects num_nodes > H3DGS_STATIC_CUT_MAX_NODES with H3DGS_ENOSPACE.  All calls share that array: calls
 * that overlap on the GPU (two streams, two host threads) race on it -- serialise them.  Same outputs as
 * h3dgs_expand_to_size, same synchronisation of `stream` before *count_host is written. */
#define H3DGS_STATIC_CUT_MAX_NODES (1 << 20)
H3DGS_API int32_t h3dgs_expand_to_size_static_max_nodes(void);
H3DGS_API int h3dgs_expand_to_size_static(int32_t num_nodes, float target_size, const int32_t* nodes,
                         const float* boxes, const float* viewpoint_dev,
                         int32_t* render_indices, int32_t* parent_indices,
                         int32_t* nodes_for_render_indices,
                         int32_t* count_host, h3dgs_stream_t stream);

/* Viewpoint by value (the reference reads it on the host, torch_interface.cpp:116-117). */
H3DGS_API int h3dgs_interpolation_weights(int32_t num_entries, const int32_t* node_indices, float target_size,
                                const int32_t* nodes, const float* boxes,
                                float vx, float vy, float vz,
                                float* ts, int32_t* kids, h3dgs_stream_t stream);

/* Incremental cut: ONE refinement step of an existing cut, as the reference's streaming viewer
 * applies per frame instead of recomputing the cut.  Replaces Switching::changeToSizeStep and its kernels
 * changeNodesOnce / putNodes / countRenderIndicesIndexed / putRenderIndicesIndexed with their CUB scan and select
 * (GH/runtime_switching.cu:145-400).  node_indices [num_cut_nodes] is the current node list (start from {0}, the
 * root, and split all zero); split is int32 [num_nodes], updated in place.  A node larger than target_size that is
 * not split yet is joined by its children; a node whose PARENT is smaller than target_size drops out (the parent,
 * still in the list, renders again).  Nodes with start_children == -1 (children not resident) are listed in
 * nodes_to_expand (may be NULL) instead.  Writes the new node list (capacity max_nodes) and its render / parent /
 * node-of-render index arrays.  Synchronises `stream` (the new sizes must reach the host, as in the reference).
 * out_host[4] = {success, new node count, render entries, nodes needing expansion}; success == 0 when the new list
 * does not fit max_nodes (then only split[] and the counts have been touched, exactly like the reference). */
H3DGS_API size_t h3dgs_cut_step_workspace_bytes(int32_t max_nodes);
H3DGS_API int h3dgs_cut_step(int32_t num_cut_nodes, const int32_t* node_indices, int32_t* new_node_indices,
                   int32_t max_nodes, const int32_t* nodes, const float* boxes, const float* viewpoint_dev,
                   float target_size, int32_t* split, int32_t* render_indices, int32_t* parent_indices,
                   int32_t* nodes_of_render_indices, int32_t* nodes_to_expand, void* workspace,
                   size_t workspace_bytes, int32_t* out_host, h3dgs_stream_t stream);

/* Host function on HOST nodes: DFS pre-order leaf/merged cut.  Call with out == NULL to get the
 * count; returns the number of indices (>= 0) or a negative error. */
H3DGS_API int64_t h3dgs_expand_to_target(int32_t num_nodes, const int32_t* nodes_host, int32_t target,
                               int32_t* out_host, int64_t capacity);

/* ------------------------------------------------------------------------------------------------
 * .hier files.  Replaces HierarchyLoader::load / HierarchyWriter::write
 * (GH/hierarchy_loader.cpp:26-128, GH/hierarchy_writer.cpp:27-118).  Host pointers.
 * ------------------------------------------------------------------------------------------------ */
H3DGS_API int h3dgs_hier_read_header(const char* path, int64_t* num_gaussians, int64_t* num_nodes, int* compressed);
H3DGS_API int h3dgs_hier_read(const char* path, float* pos /*[P,3]*/, float* shs /*[P,48]*/, float* alphas /*[P]*/,
                    float* scales /*[P,3]*/, float* rot /*[P,4]*/, int32_t* nodes /*[N,7]*/,
                    float* boxes /*[N,8]*/);
H3DGS_API int h3dgs_hier_write(const char* path, int64_t num_gaussians, int64_t num_nodes,
                     const float* pos, const float* shs, const float* alphas, const float* log_scales,
                     const float* rot, const int32_t* nodes, const float* boxes, int compressed);

/* ------------------------------------------------------------------------------------------------
 * Sparse Adam.  Replaces scene/OurAdam.py:249-337 (_single_tensor_adam: gather rows `relevant`,
 * update moments and parameter, scatter back) for all parameter tensors of one step in one launch.
 * step_size = lr / (1 - beta1^step) and bias_correction2_sqrt = sqrt(1 - beta2^step) are computed
 * by the caller in double precision exactly as the reference does on the host.
 * ------------------------------------------------------------------------------------------------ */
typedef struct H3dgsAdamTensor {
  float* param;        /* [rows,width] */
  const float* grad;   /* [rows,width] */
  float* exp_avg;      /* [rows,width] */
  float* exp_avg_sq;   /* [rows,width] */
  int32_t width;       /* floats per row */
  float step_size;
  float bias_correction2_sqrt;
  float beta1, beta2, eps;
  float one_minus_beta1, one_minus_beta2; /* (1 - beta) rounded from the host's double, as torch passes it */
} H3dgsAdamTensor;

/* Gradient exchange of camera-batch data parallelism (mode B): gather the listed rows of up to 8
 * row-major fp32 tensors into ONE packed [num_rows, sum of widths] buffer (the all-reduce payload), and the
 * inverse scatter, which multiplies by `scale` (1/world for an average).  rows: device int64 [num_rows]. */
H3DGS_API int h3dgs_pack_rows(const H3dgsRowTensor* tensors, int32_t num_tensors, const int64_t* rows,
                    int64_t num_rows, float* packed, h3dgs_stream_t stream);
H3DGS_API int h3dgs_unpack_rows(const H3dgsRowTensor* tensors, int32_t num_tensors, const int64_t* rows,
                    int64_t num_rows, const float* packed, float scale, h3dgs_stream_t stream);

#define H3DGS_ADAM_MAX_TENSORS 8
/* relevant: device int64 [num_relevant] row ids, or NULL for a dense update of rows [0,num_rows). */
H3DGS_API int h3dgs_sparse_adam(const H3dgsAdamTensor* tensors, int32_t num_tensors,
                      const int64_t* relevant, int64_t num_relevant, int64_t num_rows,
                      h3dgs_stream_t stream);

/* Sparse Adam on COMPACT gradients (H3dgsGrads.compact_*): tensors[k].grad is [capacity, width] and its row i
 * belongs to store row compact_rows[i], i < *compact_count (a device word: no host read-back).  gate (device,
 * [capacity], or NULL): rows whose gate value is exactly 0 are skipped -- with the compact opacity gradient this is
 * the reference's `relevant = (opacity.grad != 0).nonzero()` (train_single.py:225-229) evaluated on the device. */
H3DGS_API int h3dgs_sparse_adam_compact(const H3dgsAdamTensor* tensors, int32_t num_tensors,
                              const int64_t* compact_rows, const int32_t* compact_count, int32_t capacity,
                              const float* gate, h3dgs_stream_t stream);

/* Camera-batch data parallelism over NVLink peer memory (SURVEY.md 8e mode B): an all-gather of every rank's
 * gated compact gradient rows that is summed while it is read, then ONE sparse Adam step over the union -- the
 * semantics of all-reduce(sum)/world of the per-row gradients followed by OurAdam.step(relevant)
 * (train_post.py:171-190), without N-row tensors and without a host round trip.  All pointers are device pointers;
 * peer_* may point into ANOTHER GPU's memory (symmetric memory / CUDA IPC mapping): the kernels only load from them.
 *
 * h3dgs_dp_pack: compact gradients (H3dgsGrads.compact_*; grads[k].data is [capacity, width]) -> the rows whose gate
 *   is non-zero (gate NULL: all rows), packed as [count, sum of widths] floats + int32 store rows; *out_count is
 *   reset by the call; *overflow is set to 1 when out_capacity is too small (the block is then invalid).
 * h3dgs_dp_accumulate: add one rank's block into the local union; row_map is int32 [store rows], all -1 on entry to
 *   the first call of a step.  Call once per rank, in rank order, on one stream: every replica then forms identical
 *   sums.  *union_count must be zero before the first call of a step.
 * h3dgs_dp_adam: param/moment update of the union rows with gradient = union_packed * scale (tensors[k].grad is
 *   ignored), and row_map restored to -1 for those rows.
 * Sharded optimizer (ZeRO-1 over rows; the replicated update costs every rank the whole union, 3.6 M scattered rows
 * for 8 cameras): h3dgs_dp_accumulate with owner_mod = world, owner_rem = rank takes only the rows this rank owns
 * ((row >> H3DGS_DP_OWNER_SHIFT) % world == rank: runs of 64 consecutive rows share an owner, so that neighbouring
 * rows -- which share DRAM sectors in the narrow tensors -- are written together); h3dgs_dp_adam with out_* non-NULL also leaves the updated parameter rows as a packed block
 * (out_rows / out_packed / *out_count, in symmetric memory); h3dgs_dp_scatter_params copies a peer's block into the
 * local parameter tensors.  Moments of a row are then valid on its owner only. */
#define H3DGS_DP_OWNER_SHIFT 6
H3DGS_API int h3dgs_dp_pack(const H3dgsRowTensor* grads, int32_t num_tensors, const int64_t* compact_rows,
                  const int32_t* compact_count, int32_t capacity, const float* gate, int32_t* out_rows,
                  float* out_packed, int32_t* out_count, int32_t out_capacity, int32_t* overflow,
                  h3dgs_stream_t stream);
H3DGS_API int h3dgs_dp_accumulate(const int32_t* peer_rows, const float* peer_packed, const int32_t* peer_count,
                        int32_t peer_capacity, int32_t width, int32_t owner_mod, int32_t owner_rem,
                        int32_t* row_map, int32_t* union_rows,
                        float* union_packed, int32_t* union_count, int32_t union_capacity, int32_t* overflow,
                        h3dgs_stream_t stream);
H3DGS_API int h3dgs_dp_adam(const H3dgsAdamTensor* tensors, int32_t num_tensors, const int32_t* union_rows,
                  const int32_t* union_count, const float* union_packed, float scale, int32_t* row_map,
                  int32_t union_capacity, int32_t* out_rows, float* out_packed, int32_t* out_count,
                  h3dgs_stream_t stream);
H3DGS_API int h3dgs_dp_scatter_params(const H3dgsAdamTensor* tensors, int32_t num_tensors, const int32_t* peer_rows,
                            const float* peer_packed, const int32_t* peer_count, int32_t peer_capacity,
                            h3dgs_stream_t stream);

/* ------------------------------------------------------------------------------------------------
 * Densification: statistics every iteration, selection + rebuild every densification interval
 * (train_single.py:191-197, scene/gaussian_model.py:560-760,780-794).
 *
 * h3dgs_densify_stats: for the k-th visible row r = visible_rows[k]: max_radii2D[r] = max(max_radii2D[r], radii[k]),
 *   xyz_gradient_accum[r] = max(||viewspace_grad[r, :2]||, xyz_gradient_accum[r]), denom[r] += 1.  viewspace_grad is
 *   [rows,3]; radii / max_radii2D may be NULL (statistics only).
 * h3dgs_densify_select: clone / split / prune decisions of densify_and_prune for every row and the prefix sums of the
 *   three categories; synchronises `stream` and returns the sizes (rows kept, clones, splits) -- the caller
 *   allocates [kept + clones + 2 splits]-row tensors.  dense_extent = percent_dense * scene_extent.  Rows below
 *   scaffold_points are never cloned, split or pruned.  min_opacity must not exceed 0.15 (the selection's own bound).
 * h3dgs_densify_split_std: stds [2 splits, 3] of the split samples (get_scaling[selected].repeat(2, 1)); the caller
 *   draws torch.normal(0, stds) -- the same call the reference makes, so the random stream is consumed identically.
 * h3dgs_densify_build: writes every tensor's new rows in the reference's final order [kept | clones | split children,
 *   first copies | second copies]; moments of kept rows are carried over, of new rows zero.
 * ------------------------------------------------------------------------------------------------ */
typedef struct H3dgsDensifyTensor {
  const float* param;       /* [rows,width] */
  const float* exp_avg;     /* [rows,width] or NULL (no optimizer state yet) */
  const float* exp_avg_sq;
  float* new_param;         /* [kept + clones + 2 splits, width] */
  float* new_exp_avg;       /* same shape, or NULL */
  float* new_exp_avg_sq;
  int32_t width;
} H3dgsDensifyTensor;
H3DGS_API int h3dgs_densify_stats(int64_t num_visible, const int64_t* visible_rows, const int32_t* radii,
                        const float* viewspace_grad, float* xyz_gradient_accum, float* denom, float* max_radii2D,
                        h3dgs_stream_t stream);
H3DGS_API size_t h3dgs_densify_workspace_bytes(int64_t num_rows);
H3DGS_API int h3dgs_densify_select(int32_t num_rows, const float* xyz_gradient_accum, const float* max_radii2D,
                         const float* opacity_raw, int32_t opacity_activation, const float* scaling_raw,
                         float grad_threshold, float min_opacity, float dense_extent, int32_t scaffold_points,
                         void* workspace, size_t workspace_bytes, int32_t* counts_host, h3dgs_stream_t stream);
H3DGS_API int h3dgs_densify_split_std(int32_t num_rows, const float* scaling_raw, const void* workspace,
                            int32_t num_splits, float* stds, h3dgs_stream_t stream);
H3DGS_API int h3dgs_densify_build(int32_t num_rows, const H3dgsDensifyTensor* tensors, int32_t num_tensors,
                        int32_t xyz_index, int32_t scaling_index, int32_t rotation_index, const void* workspace,
                        int32_t kept, int32_t clones, int32_t splits, const float* split_samples,
                        h3dgs_stream_t stream);

/* ------------------------------------------------------------------------------------------------
 * Hierarchy build.  Replaces the tree stages of the offline GaussianHierarchyCreator
 * (GH/mainHierarchyCreator.cpp:130-141): PointbasedKdTreeGenerator::generate (GH/PointbasedKdTreeGenerator.cpp:16-73),
 * ClusterMerger::merge (GH/ClusterMerger.cpp:23-165), RotationAligner::align (GH/rotation_aligner.cpp:14-114) and
 * Writer::makeHierarchy (GH/writer.cpp:18-115), which the reference runs as single-threaded host recursions.
 *
 * Input: N ACTIVATED Gaussians on the device (what Loader::loadPly leaves in memory, GH/loader.cpp:102-158):
 * positions [N,3], shs [N,48], opacities [N], scales [N,3] (not log), rotations [N,4] (w,x,y,z).
 * Output (device, 2N-1 rows each, pre-order as HierarchyWriter::write expects them): Gaussian rows with LOG scales,
 * nodes int32 [2N-1,7], boxes fp32 [2N-1,2,4].  out_covariances [2N-1,6] (the moment-matched covariance every merge
 * accumulated) and out_leaf_order [2N-1] (input index of a leaf row, -1 for merged rows) may be NULL.
 * align != 0 runs the rotation alignment.  Synchronises `stream` before returning (the error flags of the
 * merge -- the reference's "Found Nans!" exception -- are read back).
 * Ties between equal coordinates at a median are broken by Gaussian index (std::nth_element leaves them in an
 * implementation-defined order).
 * ------------------------------------------------------------------------------------------------ */
H3DGS_API size_t h3dgs_hier_build_workspace_bytes(int64_t num_gaussians);
H3DGS_API int h3dgs_hier_build(int64_t num_gaussians, const float* positions, const float* shs, const float* opacities,
                     const float* scales, const float* rotations, int32_t align,
                     float* out_positions, float* out_shs, float* out_opacities, float* out_log_scales,
                     float* out_rotations, int32_t* out_nodes, float* out_boxes,
                     float* out_covariances, int32_t* out_leaf_order,
                     void* workspace, size_t workspace_bytes, h3dgs_stream_t stream);

/* ------------------------------------------------------------------------------------------------
 * Instrumentation (no reference counterpart).  Launch counting is always on; per-kernel timing wraps
 * every kernel in a pair of CUDA events on the launching stream while enabled.
 * ------------------------------------------------------------------------------------------------ */
#define H3DGS_PROF_KERNELS 20
/*
 * Fused photometric loss.  Replaces utils/loss_utils.py:17-63 (l1_loss, ssim / _ssim: five
 * depthwise conv2d, ~15 elementwise kernels and their autograd) as used by train_single.py:116-141 and
 * train_post.py:132-137.  Images are [C,H,W] fp32; window11 is the HOST array of the 11 normalised Gaussian
 * weights (loss_utils.py:23-25, sigma 1.5); convolution borders are zero padded like conv2d(padding=5).
 * Forward: adds the sum of the SSIM map to sums[0] and the sum of |img1 - img2| to sums[1] (device doubles,
 * zeroed by the caller); optionally writes the map and, when a backward will follow, the three per-pixel partial
 * derivatives of the map with respect to the windowed moments (mu1, E[img1^2], E[img1 img2]).
 * Backward: dL_dimg1 = upstream * (conv(g dm_dmu1) + 2 img1 conv(g dm_de11) + img2 conv(g dm_de12)
 *                                 + l1_scale * sign(img1 - img2)),
 * g = dL_dmap per pixel, or the uniform ssim_scale when dL_dmap is NULL; upstream is a DEVICE scalar or NULL (= 1).
 * For loss = (1 - l) * mean|x - y| + l * (1 - mean(ssim)): ssim_scale = -l / (C H W), l1_scale = (1 - l) / (C H W).
 * mask ([H,W], device, or NULL): the alpha mask of train_single.py:118-120 / train_post.py:134-136 -- every formula
 * above is evaluated on img1 * mask (broadcast over the channels) and dL_dimg1 carries the extra factor mask.
 */
H3DGS_API int h3dgs_ssim_forward(int32_t C, int32_t H, int32_t W, const float* img1, const float* img2,
                       const float* mask, const float* window11, float* ssim_map, float* dm_dmu1, float* dm_de11, float* dm_de12,
                       double* sums, h3dgs_stream_t stream);
H3DGS_API int h3dgs_ssim_backward(int32_t C, int32_t H, int32_t W, const float* img1, const float* img2,
                        const float* mask, const float* window11, const float* dm_dmu1, const float* dm_de11, const float* dm_de12,
                        const float* dL_dmap, const float* upstream, float ssim_scale, float l1_scale,
                        float* dL_dimg1, h3dgs_stream_t stream);

/*
 * Depth loss of train_single.py:133-139, `torch.abs((invDepth - mono_invdepth) * depth_mask).mean()`, over n
 * elements (mask NULL = all ones).  Forward adds sum |(pred - target) mask| to *sum (device double, zeroed by the
 * caller).  Backward: dL_dpred = upstream * scale * mask * sign((pred - target) mask); upstream is a DEVICE scalar or
 * NULL (= 1); for the mean, scale = 1 / n.
 */
/* Exposure affine + clamp of the render glue (gaussian_renderer/__init__.py:115-118, 277-281):
 * out = clamp(matmul(image.permute(1,2,0), E[:3,:3]).permute(2,0,1) + E[:3,3,None,None], 0, 1), image [3,H,W],
 * exposure E device [3,4] row-major or NULL (clamp alone).  Backward: grad_image [3,H,W]; grad_exposure device [12]
 * is ADDED into (caller zeroes it) or NULL.  Replaces permute / matmul / permute / add / clamp and their autograd. */
H3DGS_API int h3dgs_exposure_clamp_forward(int32_t H, int32_t W, const float* image, const float* exposure, float* out,
                                 h3dgs_stream_t stream);
H3DGS_API int h3dgs_exposure_clamp_backward(int32_t H, int32_t W, const float* image, const float* exposure,
                                  const float* grad_out, float* grad_image, float* grad_exposure,
                                  h3dgs_stream_t stream);
H3DGS_API int h3dgs_masked_l1_forward(int64_t n, const float* pred, const float* target, const float* mask, double* sum,
                            h3dgs_stream_t stream);
H3DGS_API int h3dgs_masked_l1_backward(int64_t n, const float* pred, const float* target, const float* mask,
                             const float* upstream, float scale, float* dL_dpred, h3dgs_stream_t stream);

/* Instrumented pass for the FLOP accounting of the blend kernels: while `counters` (device,
 * (16 + 4 x tiles) x uint64, zeroed by the caller) is set, the blend kernels run a counting variant and add, per evaluated
 * (live pixel, list entry) pair: [0] forward pairs rejected at power > 0, [1] at alpha < 1/255, [2] at the T
 * test, [3] blended, [4] forward lane slots issued; [5] backward pairs rejected at power > 0, [6] at
 * alpha < 1/255, [7] contributing, [8] backward lane slots issued; from word 16 on, four words per tile: global-timer ns
 * at which the tile's forward CTA started / ended and its backward CTA started / ended (load-balance diagnostics,
 * harness/diag_balance.py).  NULL switches it off.  Never timed. */
H3DGS_API int h3dgs_blend_stats(void* counters);
/* Measured pipe ceilings of the device the blend kernels' roofline is quoted against (SURVEY.md 8d asks for an FMA /
 * ex2 micro-kernel instead of a datasheet figure).  Runs three dependent-chain micro-kernels that fill every SM, best
 * of `reps` launches each, synchronising `stream`:  out[0] FP32 TFLOP/s with three-register FFMA (2 FLOP each),
 * out[1] the same with packed fma.rn.f32x2 (FFMA2), out[2] MUFU.EX2 results in G/s, out[3] warp instructions issued
 * per second (G/s) by the FFMA kernel.  scratch: device pointer to at least one float. */
H3DGS_API int h3dgs_measure_peaks(double* out4, void* scratch, int32_t reps, h3dgs_stream_t stream);
/* Diagnostics switch: key 1 = backward blend variant (1 = one pixel per lane, 2 = two pixels per lane, the default;
 * env H3DGS_BWD); key 2 = strip-key layout of the binning sort for tests (bit 0: u32 keys, bit 1: two sort passes --
 * what an image with more than 65 536 strips would use; buffers sized under one setting must not be used under
 * another); returns the previous value, -1 for an unknown key. */
H3DGS_API int h3dgs_debug_set(int key, int value);
H3DGS_API int64_t h3dgs_launch_count(void);          /* kernels launched by this library so far (process-wide) */
H3DGS_API int h3dgs_profile_enable(int enable);      /* returns the previous setting */
/* Synchronises pending events and adds their durations; ms[i] / launches[i] are per kernel id
 * 0..H3DGS_PROF_KERNELS-1 (see h3dgs_profile_name).  reset != 0 clears the accumulators afterwards. */
H3DGS_API int h3dgs_profile_read(double* ms, int64_t* launches, int reset);
H3DGS_API const char* h3dgs_profile_name(int kernel_id);

#ifdef __cplusplus
}
#endif
#endif /* H3DGS_B200_H_ */
