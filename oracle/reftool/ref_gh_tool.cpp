// TEST INFRASTRUCTURE ONLY -- a command-line driver around the reference's own host C++ classes, compiled by
// oracle/build_ref.py from the sources where they lie under /root/reference/submodules/gaussianhierarchy
// (hierarchy_writer.cpp, hierarchy_loader.cpp, ...; nothing is copied).  The reference's Python binding of
// write_hierarchy crashes in this image, so golden .hier files are produced by calling HierarchyWriter::write
// (GH/hierarchy_writer.cpp:27-118) and HierarchyLoader::load (GH/hierarchy_loader.cpp:26-128) directly.
//
//   ref_gh_tool hier-write <in.blob> <out.hier> <compressed 0|1>
//   ref_gh_tool hier-load  <in.hier> <out.blob>
//   ref_gh_tool build      <in.gblob> <out.blob> <align 0|1>   [prints "seconds kd merge align" on stdout]
//        the reference Creator's tree stages in its order (GH/mainHierarchyCreator.cpp:130-141):
//        PointbasedKdTreeGenerator::generate, ClusterMerger::merge, RotationAligner::align (optional), then
//        Writer::makeHierarchy (GH/writer.cpp:94-115) to flatten the tree.  The appearance filter (needs the COLMAP
//        source directory) is not run.  gblob: int32 N, then ACTIVATED Gaussians float pos[N*3], shs[N*48],
//        opacity[N], scale[N*3], rotation[N*4] (what Loader::loadBin leaves in memory, GH/loader.cpp:146-158).
//        out.blob holds scales as Writer emits them (log).
//
// blob: int32 P, int32 N, then float pos[P*3], shs[P*48], opacities[P], log_scales[P*3], rotations[P*4],
//       int32 nodes[N*7], float boxes[N*8]   (tests/golden/make_golden_hier.py writes / reads it)
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <stdexcept>
#include <string>
#include <vector>

#include <chrono>

#include "ClusterMerger.h"
#include "PointbasedKdTreeGenerator.h"
#include "hierarchy_loader.h"
#include "hierarchy_writer.h"
#include "rotation_aligner.h"
#include "writer.h"

namespace {

template <typename T>
void rd(std::ifstream& f, T* p, size_t n) { f.read(reinterpret_cast<char*>(p), sizeof(T) * n); }
template <typename T>
void wr(std::ofstream& f, const T* p, size_t n) { f.write(reinterpret_cast<const char*>(p), sizeof(T) * n); }

int hier_write(const char* in, const char* out, bool compressed) {
  std::ifstream f(in, std::ios_base::binary);
  if (!f.good()) throw std::runtime_error("cannot open blob");
  int P = 0, N = 0;
  rd(f, &P, 1); rd(f, &N, 1);
  std::vector<Eigen::Vector3f> pos(P), scales(P);
  std::vector<SHs> shs(P);
  std::vector<float> op(P);
  std::vector<Eigen::Vector4f> rot(P);
  std::vector<Node> nodes(N);
  std::vector<Box> boxes(N);
  static_assert(sizeof(Eigen::Vector3f) == 12 && sizeof(SHs) == 192 && sizeof(Node) == 28 && sizeof(Box) == 32, "layout");
  rd(f, pos.data(), P); rd(f, shs.data(), P); rd(f, op.data(), P); rd(f, scales.data(), P); rd(f, rot.data(), P);
  rd(f, nodes.data(), N); rd(f, boxes.data(), N);
  HierarchyWriter w;
  w.write(out, P, N, pos.data(), shs.data(), op.data(), scales.data(), rot.data(), nodes.data(), boxes.data(), compressed);
  return 0;
}

int hier_load(const char* in, const char* out) {
  std::vector<Eigen::Vector3f> pos, scales;
  std::vector<SHs> shs;
  std::vector<float> op;
  std::vector<Eigen::Vector4f> rot;
  std::vector<Node> nodes;
  std::vector<Box> boxes;
  HierarchyLoader::load(in, pos, shs, op, scales, rot, nodes, boxes);
  std::ofstream f(out, std::ios_base::binary);
  int P = (int)pos.size(), N = (int)nodes.size();
  wr(f, &P, 1); wr(f, &N, 1);
  wr(f, pos.data(), P); wr(f, shs.data(), P); wr(f, op.data(), P); wr(f, scales.data(), P); wr(f, rot.data(), P);
  wr(f, nodes.data(), N); wr(f, boxes.data(), N);
  return 0;
}

int build(const char* in, const char* out, bool align) {
  std::ifstream f(in, std::ios_base::binary);
  if (!f.good()) throw std::runtime_error("cannot open gblob");
  int N = 0;
  rd(f, &N, 1);
  std::vector<Eigen::Vector3f> pos(N), scales(N);
  std::vector<SHs> shs(N);
  std::vector<float> op(N);
  std::vector<Eigen::Vector4f> rot(N);
  rd(f, pos.data(), N); rd(f, shs.data(), N); rd(f, op.data(), N); rd(f, scales.data(), N); rd(f, rot.data(), N);
  std::vector<Gaussian> gaussians(N);
  for (int i = 0; i < N; ++i) {
    Gaussian& g = gaussians[i];
    g.position = pos[i]; g.shs = shs[i]; g.opacity = op[i]; g.scale = scales[i]; g.rotation = rot[i];
    computeCovariance(g.scale, g.rotation, g.covariance);
  }
  using clk = std::chrono::steady_clock;
  auto t0 = clk::now();
  PointbasedKdTreeGenerator generator;
  ExplicitTreeNode* root = generator.generate(gaussians);
  auto t1 = clk::now();
  ClusterMerger merger;
  merger.merge(root, gaussians);
  auto t2 = clk::now();
  if (align) RotationAligner::align(root, gaussians);
  auto t3 = clk::now();
  std::vector<Eigen::Vector3f> o_pos, o_scales;
  std::vector<Eigen::Vector4f> o_rot;
  std::vector<float> o_op;
  std::vector<SHs> o_shs;
  std::vector<Node> nodes;
  std::vector<Box> boxes;
  Writer::makeHierarchy(gaussians, root, o_pos, o_rot, o_scales, o_op, o_shs, nodes, boxes);
  std::ofstream o(out, std::ios_base::binary);
  int P = (int)o_pos.size(), NN = (int)nodes.size();
  wr(o, &P, 1); wr(o, &NN, 1);
  wr(o, o_pos.data(), P); wr(o, o_shs.data(), P); wr(o, o_op.data(), P); wr(o, o_scales.data(), P); wr(o, o_rot.data(), P);
  wr(o, nodes.data(), NN); wr(o, boxes.data(), NN);
  auto sec = [](clk::time_point a, clk::time_point b) { return std::chrono::duration<double>(b - a).count(); };
  std::printf("seconds %.6f %.6f %.6f\n", sec(t0, t1), sec(t1, t2), sec(t2, t3));
  return 0;
}

}  // namespace

int main(int argc, char** argv) {
  try {
    if (argc == 5 && !std::strcmp(argv[1], "hier-write")) return hier_write(argv[2], argv[3], std::atoi(argv[4]) != 0);
    if (argc == 4 && !std::strcmp(argv[1], "hier-load")) return hier_load(argv[2], argv[3]);
    if (argc == 5 && !std::strcmp(argv[1], "build")) return build(argv[2], argv[3], std::atoi(argv[4]) != 0);
    std::fprintf(stderr, "usage: ref_gh_tool hier-write <blob> <hier> <0|1> | hier-load <hier> <blob>\n");
    return 2;
  } catch (const std::exception& e) {
    std::fprintf(stderr, "ref_gh_tool: %s\n", e.what());
    return 1;
  }
}
