// TEST INFRASTRUCTURE ONLY -- command-line driver around the reference's own Switching::changeToSizeStep
// (GH/runtime_switching.cu:283-400), which has no Python binding.  Compiled by oracle/build_ref.py with nvcc from the
// reference source where it lies (nothing is copied); runs on the GPU box as the checker of h3dgs_cut_step.
//
//   ref_switch_tool <in.blob> <out.blob>
//   in : int32 Nn, maxN, views, steps;  int32 nodes[Nn*7];  float boxes[Nn*8];  per view: float vp[3], float target
//   out: per (view, step): int32 success, newN, newR, need;  then (success only) int32 node_list[newN],
//        render[newR], parent[newR], nodes_of[newR];  int32 to_expand[need]
// The cut starts as {root} with split all zero and is carried from step to step and from view to view.
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <vector>

#include <cuda_runtime.h>

#include "runtime_switching.h"

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { std::fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e_)); return 1; } } while (0)

int main(int argc, char** argv) {
  if (argc != 3) { std::fprintf(stderr, "usage: ref_switch_tool <in.blob> <out.blob>\n"); return 2; }
  std::ifstream f(argv[1], std::ios_base::binary);
  if (!f.good()) { std::fprintf(stderr, "cannot open %s\n", argv[1]); return 1; }
  int hdr[4];
  f.read(reinterpret_cast<char*>(hdr), sizeof(hdr));
  const int Nn = hdr[0], maxN = hdr[1], views = hdr[2], steps = hdr[3];
  std::vector<int> nodes((size_t)Nn * 7);
  std::vector<float> boxes((size_t)Nn * 8), cams((size_t)views * 4);
  f.read(reinterpret_cast<char*>(nodes.data()), nodes.size() * 4);
  f.read(reinterpret_cast<char*>(boxes.data()), boxes.size() * 4);
  f.read(reinterpret_cast<char*>(cams.data()), cams.size() * 4);
  int *d_nodes, *d_idx[2], *d_split, *d_render, *d_parent, *d_nodesof, *d_expand, *d_src, *d_dst, *d_num;
  float *d_boxes, *d_vp;
  char* d_c;
  CK(cudaMalloc(&d_nodes, nodes.size() * 4)); CK(cudaMalloc(&d_boxes, boxes.size() * 4)); CK(cudaMalloc(&d_vp, 16));
  for (int k = 0; k < 2; ++k) CK(cudaMalloc(&d_idx[k], (size_t)maxN * 4));
  CK(cudaMalloc(&d_split, (size_t)Nn * 4)); CK(cudaMalloc(&d_render, (size_t)Nn * 4));
  CK(cudaMalloc(&d_parent, (size_t)Nn * 4)); CK(cudaMalloc(&d_nodesof, (size_t)Nn * 4));
  CK(cudaMalloc(&d_expand, (size_t)maxN * 4)); CK(cudaMalloc(&d_src, (size_t)maxN * 4));
  CK(cudaMalloc(&d_dst, (size_t)maxN * 4)); CK(cudaMalloc(&d_num, 16)); CK(cudaMalloc(&d_c, (size_t)maxN));
  CK(cudaMemcpy(d_nodes, nodes.data(), nodes.size() * 4, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(d_boxes, boxes.data(), boxes.size() * 4, cudaMemcpyHostToDevice));
  CK(cudaMemset(d_split, 0, (size_t)Nn * 4));
  CK(cudaMemset(d_idx[0], 0, (size_t)maxN * 4));   // {root}
  char* scratch = nullptr;
  size_t scratch_size = 0;
  int cur = 0, N = 1;
  std::ofstream o(argv[2], std::ios_base::binary);
  std::vector<int> buf;
  auto dump = [&](const int* dev, int n) {
    buf.resize(n > 0 ? n : 1);
    if (n > 0) cudaMemcpy(buf.data(), dev, (size_t)n * 4, cudaMemcpyDeviceToHost);
    o.write(reinterpret_cast<const char*>(buf.data()), (size_t)(n > 0 ? n : 0) * 4);
  };
  for (int v = 0; v < views; ++v) {
    CK(cudaMemcpy(d_vp, &cams[(size_t)v * 4], 12, cudaMemcpyHostToDevice));
    for (int s = 0; s < steps; ++s) {
      int ok = 0, newN = 0, newR = 0, need = 0;
      Switching::changeToSizeStep(cams[(size_t)v * 4 + 3], N, d_idx[cur], d_idx[cur ^ 1], d_nodes, d_boxes, d_vp, 0.f, 0.f,
                                  1.f, d_split, d_render, d_parent, d_nodesof, d_expand, nullptr, scratch, scratch_size,
                                  d_src, d_dst, d_c, d_num, maxN, ok, &newN, &newR, &need, nullptr);
      CK(cudaDeviceSynchronize());
      const int rec[4] = {ok, newN, ok ? newR : 0, need};
      o.write(reinterpret_cast<const char*>(rec), sizeof(rec));
      if (ok) {
        cur ^= 1;
        N = newN;
        dump(d_idx[cur], newN);
        dump(d_render, newR); dump(d_parent, newR); dump(d_nodesof, newR);
      }
      dump(d_expand, need);
    }
  }
  return 0;
}
