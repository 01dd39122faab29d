// csrc/rt_bvh.cu — host-built bounding-volume hierarchy over the sphere list (optional mode).
//
// The reference has no acceleration structure: hittable_list::hit tests every sphere
// (src/cpu/hittable_list.h:28-43).  This is the north-star's "optional BVH built on the host"
// for scenes where N x rays is out of reach (BASELINE.json configs[3], 100 k spheres).  The
// traversal (rt_wave.cu, bvh_trace) must return exactly the sphere the reference's linear scan
// returns, so the builder only has to be CONSERVATIVE: every sphere lies inside its leaf box,
// every child box inside its parent, with f32 bounds rounded outwards.
//
// Build: top-down, binned surface-area heuristic (16 bins on the longest centroid axis), leaves
// of <= kLeafMax spheres, nodes stored depth-first with the two children adjacent.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <vector>

#include "rt_kernels.h"

namespace rtb {

namespace {

struct Box {
  double lo[3], hi[3];
  void reset() { for (int k = 0; k < 3; ++k) { lo[k] = INFINITY; hi[k] = -INFINITY; } }
  void grow(const Box &b) { for (int k = 0; k < 3; ++k) { lo[k] = std::min(lo[k], b.lo[k]); hi[k] = std::max(hi[k], b.hi[k]); } }
  void grow_point(const double p[3]) { for (int k = 0; k < 3; ++k) { lo[k] = std::min(lo[k], p[k]); hi[k] = std::max(hi[k], p[k]); } }
  double area() const {
    double d[3] = {hi[0] - lo[0], hi[1] - lo[1], hi[2] - lo[2]};
    if (d[0] < 0) return 0;
    return 2.0 * (d[0] * d[1] + d[1] * d[2] + d[2] * d[0]);
  }
};

float down(double v) { float f = (float)v; return (double)f > v ? std::nextafter(f, -INFINITY) : f; }
float up(double v) { float f = (float)v; return (double)f < v ? std::nextafter(f, INFINITY) : f; }

struct Builder {
  const double4 *sph;  // {cx, cy, cz, signed radius}
  std::vector<Box> box;
  std::vector<uint32_t> &prim;
  std::vector<BvhNode> &nodes;
  static constexpr int kBins = 16;

  Builder(const double4 *s, uint32_t n, std::vector<uint32_t> &p, std::vector<BvhNode> &nd)
      : sph(s), box(n), prim(p), nodes(nd) {
    for (uint32_t i = 0; i < n; ++i) {
      // a relative margin on top of the exact extent: the traversal tests boxes in f32
      const double r = std::fabs(s[i].w);
      const double c[3] = {s[i].x, s[i].y, s[i].z};
      for (int k = 0; k < 3; ++k) {
        const double m = 1e-6 * (std::fabs(c[k]) + r) + 1e-9;
        box[i].lo[k] = c[k] - r - m;
        box[i].hi[k] = c[k] + r + m;
      }
    }
  }

  void set_bounds(BvhNode &nd, const Box &b) {
    nd.lo[0] = down(b.lo[0]); nd.lo[1] = down(b.lo[1]); nd.lo[2] = down(b.lo[2]);
    nd.hi[0] = up(b.hi[0]); nd.hi[1] = up(b.hi[1]); nd.hi[2] = up(b.hi[2]);
  }

  // builds the subtree over prim[first, first+count) into nodes[at]
  void build(uint32_t at, uint32_t first, uint32_t count, int depth) {
    Box b, cb;
    b.reset(); cb.reset();
    for (uint32_t i = first; i < first + count; ++i) {
      b.grow(box[prim[i]]);
      const double c[3] = {sph[prim[i]].x, sph[prim[i]].y, sph[prim[i]].z};
      cb.grow_point(c);
    }
    set_bounds(nodes[at], b);
    auto make_leaf = [&]() { nodes[at].first = first; nodes[at].count = count; };
    if (count <= kBvhLeafMax || depth >= (int)kBvhMaxDepth - 1) { make_leaf(); return; }
    int axis = 0;
    double ext[3] = {cb.hi[0] - cb.lo[0], cb.hi[1] - cb.lo[1], cb.hi[2] - cb.lo[2]};
    if (ext[1] > ext[axis]) axis = 1;
    if (ext[2] > ext[axis]) axis = 2;
    uint32_t mid = first + count / 2;
    auto centre = [&](uint32_t p) { return axis == 0 ? sph[p].x : axis == 1 ? sph[p].y : sph[p].z; };
    if (ext[axis] > 0) {
      // binned SAH along the longest centroid axis
      Box bb[kBins]; uint32_t bc[kBins] = {};
      for (auto &x : bb) x.reset();
      const double scale = kBins / ext[axis];
      auto bin_of = [&](uint32_t p) {
        int k = (int)((centre(p) - cb.lo[axis]) * scale);
        return k < 0 ? 0 : k >= kBins ? kBins - 1 : k;
      };
      for (uint32_t i = first; i < first + count; ++i) { int k = bin_of(prim[i]); bb[k].grow(box[prim[i]]); bc[k]++; }
      double right_area[kBins]; uint32_t right_cnt[kBins];
      Box acc; acc.reset(); uint32_t n = 0;
      for (int k = kBins - 1; k > 0; --k) { acc.grow(bb[k]); n += bc[k]; right_area[k] = acc.area(); right_cnt[k] = n; }
      acc.reset(); n = 0;
      double best = INFINITY; int best_k = -1;
      for (int k = 0; k < kBins - 1; ++k) {
        acc.grow(bb[k]); n += bc[k];
        if (n == 0 || right_cnt[k + 1] == 0) continue;
        const double cost = acc.area() * n + right_area[k + 1] * right_cnt[k + 1];
        if (cost < best) { best = cost; best_k = k; }
      }
      if (best_k >= 0) {
        auto it = std::partition(prim.begin() + first, prim.begin() + first + count,
                                 [&](uint32_t p) { return bin_of(p) <= best_k; });
        mid = (uint32_t)(it - prim.begin());
      }
    }
    if (mid == first || mid == first + count) {  // degenerate split: median by index order
      mid = first + count / 2;
      std::nth_element(prim.begin() + first, prim.begin() + mid, prim.begin() + first + count,
                       [&](uint32_t a, uint32_t c) { return centre(a) < centre(c) || (centre(a) == centre(c) && a < c); });
    }
    const uint32_t left = (uint32_t)nodes.size();
    nodes.push_back(BvhNode{});
    nodes.push_back(BvhNode{});
    nodes[at].first = left;  // children at left, left + 1
    nodes[at].count = 0;
    build(left, first, mid - first, depth + 1);
    build(left + 1, mid, first + count - mid, depth + 1);
  }
};

}  // namespace

void build_bvh(const double4 *spheres, uint32_t n, std::vector<BvhNode> &nodes,
               std::vector<uint32_t> &prim) {
  nodes.clear();
  prim.resize(n);
  for (uint32_t i = 0; i < n; ++i) prim[i] = i;
  nodes.reserve(2 * (size_t)n / 2 + 2);
  nodes.push_back(BvhNode{});
  if (n == 0) {
    nodes[0].lo[0] = nodes[0].lo[1] = nodes[0].lo[2] = 1.f;
    nodes[0].hi[0] = nodes[0].hi[1] = nodes[0].hi[2] = -1.f;  // empty box: never entered
    nodes[0].first = 0; nodes[0].count = 0;
    return;
  }
  Builder b(spheres, n, prim, nodes);
  b.build(0, 0, n, 0);
  // leaves visit their spheres in ascending index order (cache-friendlier; the closest-hit rule
  // itself is order-independent, see bvh_trace)
  for (auto &nd : nodes)
    if (nd.count) std::sort(prim.begin() + nd.first, prim.begin() + nd.first + nd.count);
}

}  // namespace rtb
