// .hier reader / writer (host code).  File layout as produced by the reference's HierarchyWriter and
// consumed by HierarchyLoader (GH/hierarchy_writer.cpp:27-118, GH/hierarchy_loader.cpp:26-128):
//
//   int32 P            >= 0: fp32 payload;  < 0: |P| Gaussians, fp16 payload ("compressed")
//   pos      [P,3] f32
//   rot      [P,4] f32|f16      (w,x,y,z)
//   scales   [P,3] f32|f16      (log scales)
//   opacity  [P]   f32|f16      (activated)
//   shs      [P,48] f32|f16     (coefficient-major, rgb-minor)
//   int32 N
//   nodes    fp32 form: [N,7] int32 (depth,parent,start,count_leafs,count_merged,start_children,count_children)
//            fp16 form: [N] {int32 parent, start, start_children; int16 depth, count_children, count_leafs, count_merged}
//   boxes    [N,2,4] f32|f16
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <vector>

#include "../../include/h3dgs_b200.h"

namespace h3dgs {
int fail(int code, const char* what);
}

namespace {

inline uint16_t f32_to_f16_rne(float f) {
  uint32_t x;
  std::memcpy(&x, &f, 4);
  const uint32_t sign = (x >> 16) & 0x8000u;
  x &= 0x7fffffffu;
  if (x >= 0x7f800000u) return (uint16_t)(sign | 0x7c00u | (x > 0x7f800000u ? 0x200u | ((x >> 13) & 0x3ffu) : 0u));
  if (x >= 0x477ff000u) return (uint16_t)(sign | 0x7c00u);  // rounds to >= 65520 -> inf
  if (x < 0x33000001u) return (uint16_t)sign;               // <= 2^-25 -> 0 (tie to even)
  int exp = (int)(x >> 23) - 127;
  uint32_t man = (x & 0x7fffffu) | 0x800000u;
  int shift;
  uint32_t hexp;
  if (exp < -14) {  // subnormal half
    shift = 13 + (-14 - exp);
    hexp = 0;
  } else {
    shift = 13;
    hexp = (uint32_t)(exp + 15);
  }
  uint32_t q = man >> shift;
  const uint32_t rem = man & ((1u << shift) - 1u);
  const uint32_t half = 1u << (shift - 1);
  if (rem > half || (rem == half && (q & 1u))) q++;
  uint32_t h;
  if (hexp == 0) h = q;  // may carry into the exponent: that is the correct normal number
  else h = ((hexp - 1) << 10) + q;  // q includes the implicit bit (0x400); carry handled by the add
  return (uint16_t)(sign | h);
}

inline float f16_to_f32(uint16_t h) {
  const uint32_t sign = (uint32_t)(h & 0x8000u) << 16;
  uint32_t exp = (h >> 10) & 0x1fu;
  uint32_t man = h & 0x3ffu;
  uint32_t x;
  if (exp == 0) {
    if (man == 0) {
      x = sign;
    } else {
      int e = -1;
      do { man <<= 1; ++e; } while (!(man & 0x400u));
      x = sign | ((uint32_t)(127 - 15 - e) << 23) | ((man & 0x3ffu) << 13);
    }
  } else if (exp == 31) {
    x = sign | 0x7f800000u | (man << 13);
  } else {
    x = sign | ((exp + 112) << 23) | (man << 13);
  }
  float f;
  std::memcpy(&f, &x, 4);
  return f;
}

struct HalfNodeRec {
  int32_t parent, start, start_children;
  int16_t dccc[4];
};
static_assert(sizeof(HalfNodeRec) == 20, "packed half node is 20 bytes");

struct File {
  FILE* f;
  explicit File(const char* path, const char* mode) : f(std::fopen(path, mode)) {}
  ~File() { if (f) std::fclose(f); }
};

bool read_exact(FILE* f, void* dst, size_t bytes) { return bytes == 0 || std::fread(dst, 1, bytes, f) == bytes; }
bool write_exact(FILE* f, const void* src, size_t bytes) { return bytes == 0 || std::fwrite(src, 1, bytes, f) == bytes; }

bool read_halfs(FILE* f, float* dst, size_t count) {
  std::vector<uint16_t> buf(count);
  if (!read_exact(f, buf.data(), count * 2)) return false;
  for (size_t i = 0; i < count; ++i) dst[i] = f16_to_f32(buf[i]);
  return true;
}
bool write_halfs(FILE* f, const float* src, size_t count) {
  std::vector<uint16_t> buf(count);
  for (size_t i = 0; i < count; ++i) buf[i] = f32_to_f16_rne(src[i]);
  return write_exact(f, buf.data(), count * 2);
}

}  // namespace

extern "C" int h3dgs_hier_read_header(const char* path, int64_t* num_gaussians, int64_t* num_nodes,
                                      int* compressed) {
  if (!path || !num_gaussians || !num_nodes) return h3dgs::fail(H3DGS_EINVAL, "hier_read_header: null argument");
  File fp(path, "rb");
  if (!fp.f) return h3dgs::fail(H3DGS_EIO, "File not found!");
  int32_t P;
  if (!read_exact(fp.f, &P, 4)) return h3dgs::fail(H3DGS_EIO, "hier: short read (header)");
  const bool comp = P < 0;
  const int64_t n = comp ? -(int64_t)P : (int64_t)P;
  const int64_t per = comp ? (12 + 8 + 6 + 2 + 96) : (12 + 16 + 12 + 4 + 192);
  if (std::fseek(fp.f, (long)(4 + n * per), SEEK_SET) != 0) return h3dgs::fail(H3DGS_EIO, "hier: seek failed");
  int32_t N;
  if (!read_exact(fp.f, &N, 4)) return h3dgs::fail(H3DGS_EIO, "hier: short read (node count)");
  *num_gaussians = n;
  *num_nodes = N;
  if (compressed) *compressed = comp ? 1 : 0;
  return 0;
}

extern "C" int h3dgs_hier_read(const char* path, float* pos, float* shs, float* alphas, float* scales,
                               float* rot, int32_t* nodes, float* boxes) {
  if (!path) return h3dgs::fail(H3DGS_EINVAL, "hier_read: null path");
  File fp(path, "rb");
  if (!fp.f) return h3dgs::fail(H3DGS_EIO, "File not found!");
  int32_t P;
  if (!read_exact(fp.f, &P, 4)) return h3dgs::fail(H3DGS_EIO, "hier: short read (header)");
  bool ok = true;
  if (P >= 0) {
    const size_t n = (size_t)P;
    ok = ok && read_exact(fp.f, pos, n * 12);
    ok = ok && read_exact(fp.f, rot, n * 16);
    ok = ok && read_exact(fp.f, scales, n * 12);
    ok = ok && read_exact(fp.f, alphas, n * 4);
    ok = ok && read_exact(fp.f, shs, n * 192);
    int32_t N = 0;
    ok = ok && read_exact(fp.f, &N, 4);
    ok = ok && read_exact(fp.f, nodes, (size_t)N * 28);
    ok = ok && read_exact(fp.f, boxes, (size_t)N * 32);
  } else {
    const size_t n = (size_t)(-(int64_t)P);
    ok = ok && read_exact(fp.f, pos, n * 12);
    ok = ok && read_halfs(fp.f, rot, n * 4);
    ok = ok && read_halfs(fp.f, scales, n * 3);
    ok = ok && read_halfs(fp.f, alphas, n);
    ok = ok && read_halfs(fp.f, shs, n * 48);
    int32_t N = 0;
    ok = ok && read_exact(fp.f, &N, 4);
    if (ok) {
      std::vector<HalfNodeRec> hn((size_t)N);
      ok = read_exact(fp.f, hn.data(), (size_t)N * sizeof(HalfNodeRec));
      for (size_t i = 0; ok && i < (size_t)N; ++i) {
        int32_t* d = nodes + i * 7;
        d[0] = hn[i].dccc[0];       // depth
        d[1] = hn[i].parent;
        d[2] = hn[i].start;
        d[3] = hn[i].dccc[2];       // count_leafs
        d[4] = hn[i].dccc[3];       // count_merged
        d[5] = hn[i].start_children;
        d[6] = hn[i].dccc[1];       // count_children
      }
      ok = ok && read_halfs(fp.f, boxes, (size_t)N * 8);
    }
  }
  if (!ok) return h3dgs::fail(H3DGS_EIO, "hier: short read");
  return 0;
}

extern "C" int h3dgs_hier_write(const char* path, int64_t num_gaussians, int64_t num_nodes,
                                const float* pos, const float* shs, const float* alphas,
                                const float* log_scales, const float* rot, const int32_t* nodes,
                                const float* boxes, int compressed) {
  if (!path || num_gaussians < 0 || num_nodes < 0) return h3dgs::fail(H3DGS_EINVAL, "hier_write: bad argument");
  const size_t n = (size_t)num_gaussians, N = (size_t)num_nodes;
  if (compressed) {
    for (size_t i = 0; i < N; ++i) {
      const int32_t* s = nodes + i * 7;
      if (s[0] > 32000 || s[6] > 32000 || s[3] > 32000 || s[4] > 32000)
        return h3dgs::fail(H3DGS_ELOSSY, "Would lose information!");
    }
  }
  File fp(path, "wb");
  if (!fp.f) return h3dgs::fail(H3DGS_EIO, "File not created!");
  bool ok = true;
  if (!compressed) {
    const int32_t P = (int32_t)num_gaussians, NN = (int32_t)num_nodes;
    ok = ok && write_exact(fp.f, &P, 4);
    ok = ok && write_exact(fp.f, pos, n * 12);
    ok = ok && write_exact(fp.f, rot, n * 16);
    ok = ok && write_exact(fp.f, log_scales, n * 12);
    ok = ok && write_exact(fp.f, alphas, n * 4);
    ok = ok && write_exact(fp.f, shs, n * 192);
    ok = ok && write_exact(fp.f, &NN, 4);
    ok = ok && write_exact(fp.f, nodes, N * 28);
    ok = ok && write_exact(fp.f, boxes, N * 32);
  } else {
    const int32_t P = (int32_t)(-num_gaussians), NN = (int32_t)num_nodes;
    ok = ok && write_exact(fp.f, &P, 4);
    ok = ok && write_exact(fp.f, pos, n * 12);
    ok = ok && write_halfs(fp.f, rot, n * 4);
    ok = ok && write_halfs(fp.f, log_scales, n * 3);
    ok = ok && write_halfs(fp.f, alphas, n);
    ok = ok && write_halfs(fp.f, shs, n * 48);
    ok = ok && write_exact(fp.f, &NN, 4);
    std::vector<HalfNodeRec> hn(N);
    for (size_t i = 0; i < N; ++i) {
      const int32_t* s = nodes + i * 7;
      hn[i].parent = s[1];
      hn[i].start = s[2];
      hn[i].start_children = s[5];
      hn[i].dccc[0] = (int16_t)s[0];
      hn[i].dccc[1] = (int16_t)s[6];
      hn[i].dccc[2] = (int16_t)s[3];
      hn[i].dccc[3] = (int16_t)s[4];
    }
    ok = ok && write_exact(fp.f, hn.data(), N * sizeof(HalfNodeRec));
    ok = ok && write_halfs(fp.f, boxes, N * 8);
  }
  if (!ok) return h3dgs::fail(H3DGS_EIO, "hier: short write");
  return 0;
}

extern "C" int64_t h3dgs_expand_to_target(int32_t num_nodes, const int32_t* nodes, int32_t target,
                                          int32_t* out, int64_t capacity) {
  // DFS pre-order (GH/traversal.cpp:15-38), iterative: children are pushed in reverse so that they
  // pop in order.
  if (num_nodes <= 0) return 0;
  if (!nodes) return h3dgs::fail(H3DGS_EINVAL, "expand_to_target: null nodes");
  std::vector<int32_t> stack;
  stack.push_back(0);
  int64_t count = 0;
  while (!stack.empty()) {
    const int32_t id = stack.back();
    stack.pop_back();
    if (id < 0 || id >= num_nodes) return h3dgs::fail(H3DGS_EINVAL, "expand_to_target: node id out of range");
    const int32_t* nd = nodes + (size_t)id * 7;
    const int32_t depth = nd[0], start = nd[2], leafs = nd[3], merged = nd[4], sc = nd[5], cc = nd[6];
    for (int32_t i = 0; i < leafs; ++i) {
      if (out) {
        if (count >= capacity) return h3dgs::fail(H3DGS_ENOSPACE, "expand_to_target: output too small");
        out[count] = start + i;
      }
      ++count;
    }
    if (depth <= target) {
      for (int32_t i = 0; i < merged; ++i) {
        if (out) {
          if (count >= capacity) return h3dgs::fail(H3DGS_ENOSPACE, "expand_to_target: output too small");
          out[count] = start + leafs + i;
        }
        ++count;
      }
    } else {
      for (int32_t i = cc - 1; i >= 0; --i) stack.push_back(sc + i);
    }
  }
  return count;
}
