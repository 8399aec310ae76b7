// crt_bvh.cpp -- binned-SAH BVH2 builder (host side of the scene-upload module).
//
// New capability: the reference's intersect() is a linear loop (ComputeShader.wgsl:503-518); the
// 1M / 10M-triangle configs need an accelerator that returns the same answer.  The tree only decides
// which primitives get tested; the tests themselves and the tie rule live in crt_device.cuh, and the
// boxes are padded so that culling is conservative with respect to the f32 rounding of those tests
// (DESIGN.md "Conservative culling").  Tree shape therefore affects speed, never results.
#include "crt_bvh.h"

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstring>
#include <future>
#include <limits>

namespace crt {
namespace {

constexpr int kBins = 32;
constexpr uint32_t kLeafMax = 8;         // leaf encoding holds count-1 in 3 bits
constexpr uint32_t kForceMedianDepth = 36;
constexpr float kTraversalCost = 1.0f;   // relative to one triangle test

struct Item { float lo[3], hi[3]; uint32_t id; };

struct Box {
  float lo[3], hi[3];
  void reset() { for (int a = 0; a < 3; a++) { lo[a] = std::numeric_limits<float>::infinity(); hi[a] = -lo[a]; } }
  void grow(const float* l, const float* h) { for (int a = 0; a < 3; a++) { lo[a] = std::min(lo[a], l[a]); hi[a] = std::max(hi[a], h[a]); } }
  void grow(const Box& b) { grow(b.lo, b.hi); }
  float half_area() const {
    float dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
    if (dx < 0.0f) return 0.0f;
    return dx * dy + dy * dz + dz * dx;
  }
};

struct BNode { Box box; uint32_t left = 0, right = 0, first = 0, count = 0; };

struct Builder {
  std::vector<Item>& items;
  std::vector<BNode> pool;
  std::atomic<uint32_t> next{1};
  std::atomic<uint32_t> max_depth{0};
  int par_depth;

  Builder(std::vector<Item>& it, int threads) : items(it), pool(2 * it.size() + 2) {
    par_depth = 0;
    while ((1 << par_depth) < threads * 4) par_depth++;
  }

  void note_depth(uint32_t d) {
    uint32_t cur = max_depth.load(std::memory_order_relaxed);
    while (d > cur && !max_depth.compare_exchange_weak(cur, d, std::memory_order_relaxed)) {}
  }

  void build(uint32_t me, uint32_t first, uint32_t count, uint32_t depth) {
    BNode& n = pool[me];
    Box box, cbox;
    box.reset(); cbox.reset();
    for (uint32_t i = first; i < first + count; i++) {
      const Item& it = items[i];
      box.grow(it.lo, it.hi);
      float c[3] = {0.5f * (it.lo[0] + it.hi[0]), 0.5f * (it.lo[1] + it.hi[1]), 0.5f * (it.lo[2] + it.hi[2])};
      cbox.grow(c, c);
    }
    n.box = box; n.first = first; n.count = 0;
    note_depth(depth);
    if (count <= 2) { n.count = count; return; }

    uint32_t mid = 0;
    bool have_split = false;
    if (depth < kForceMedianDepth) {
      float best_cost = std::numeric_limits<float>::infinity();
      int best_axis = -1, best_bin = -1;
      for (int ax = 0; ax < 3; ax++) {
        const float c0 = cbox.lo[ax], ext = cbox.hi[ax] - cbox.lo[ax];
        if (!(ext > 0.0f)) continue;
        const float scale = (float)kBins / ext;
        Box bb[kBins]; uint32_t bc[kBins];
        for (int b = 0; b < kBins; b++) { bb[b].reset(); bc[b] = 0; }
        for (uint32_t i = first; i < first + count; i++) {
          const Item& it = items[i];
          int b = (int)((0.5f * (it.lo[ax] + it.hi[ax]) - c0) * scale);
          b = b < 0 ? 0 : (b >= kBins ? kBins - 1 : b);
          bb[b].grow(it.lo, it.hi); bc[b]++;
        }
        float right_area[kBins]; uint32_t right_cnt[kBins];
        Box acc; acc.reset(); uint32_t cnt = 0;
        for (int b = kBins - 1; b > 0; b--) { acc.grow(bb[b]); cnt += bc[b]; right_area[b] = acc.half_area(); right_cnt[b] = cnt; }
        acc.reset(); cnt = 0;
        for (int b = 0; b < kBins - 1; b++) {
          acc.grow(bb[b]); cnt += bc[b];
          if (cnt == 0 || right_cnt[b + 1] == 0) continue;
          float cost = acc.half_area() * (float)cnt + right_area[b + 1] * (float)right_cnt[b + 1];
          if (cost < best_cost) { best_cost = cost; best_axis = ax; best_bin = b; }
        }
      }
      if (best_axis >= 0) {
        const float pa = box.half_area();
        const float split_cost = kTraversalCost + (pa > 0.0f ? best_cost / pa : 0.0f);
        if (count <= kLeafMax && (float)count <= split_cost) { n.count = count; return; }
        const float c0 = cbox.lo[best_axis], scale = (float)kBins / (cbox.hi[best_axis] - cbox.lo[best_axis]);
        const int ax = best_axis, bin = best_bin;
        Item* b = items.data() + first;
        Item* m = std::partition(b, b + count, [=](const Item& it) {
          int k = (int)((0.5f * (it.lo[ax] + it.hi[ax]) - c0) * scale);
          k = k < 0 ? 0 : (k >= kBins ? kBins - 1 : k);
          return k <= bin;
        });
        mid = (uint32_t)(m - b);
        have_split = mid > 0 && mid < count;
      }
    }
    if (!have_split) {
      if (count <= kLeafMax && depth < kForceMedianDepth) { n.count = count; return; }
      if (count <= 4) { n.count = count; return; }
      // object median along the widest centroid axis (also the fallback for coincident centroids)
      int ax = 0;
      float e0 = cbox.hi[0] - cbox.lo[0], e1 = cbox.hi[1] - cbox.lo[1], e2 = cbox.hi[2] - cbox.lo[2];
      if (e1 > e0) { ax = 1; e0 = e1; }
      if (e2 > e0) ax = 2;
      mid = count / 2;
      Item* b = items.data() + first;
      std::nth_element(b, b + mid, b + count, [ax](const Item& x, const Item& y) {
        float cx = x.lo[ax] + x.hi[ax], cy = y.lo[ax] + y.hi[ax];
        return cx < cy || (cx == cy && x.id < y.id);
      });
    }
    const uint32_t l = next.fetch_add(2, std::memory_order_relaxed);
    n.left = l; n.right = l + 1;
    if ((int)depth < par_depth && count > 8192) {
      auto fut = std::async(std::launch::async, [=] { build(l, first, mid, depth + 1); });
      build(l + 1, first + mid, count - mid, depth + 1);
      fut.get();
    } else {
      build(l, first, mid, depth + 1);
      build(l + 1, first + mid, count - mid, depth + 1);
    }
  }
};

inline int32_t child_code(const std::vector<BNode>& pool, uint32_t c, const std::vector<int32_t>& inner_index) {
  const BNode& n = pool[c];
  if (n.count) return ~(int32_t)((n.first << 3) | (n.count - 1));
  return inner_index[c];
}

}  // namespace

void build_bvh(const uint8_t* prims, const std::vector<uint32_t>& tri_pos, int n_threads, BvhResult& out) {
  out.nodes.clear(); out.tri_src.clear(); out.n_nodes = 0; out.max_depth = 0; out.pad = 0.0f;
  const size_t n = tri_pos.size();
  if (!n) return;
  std::vector<Item> items(n);
  Box scene; scene.reset();
  for (size_t i = 0; i < n; i++) {
    const float* d1 = (const float*)(prims + (size_t)tri_pos[i] * 80 + 16);
    const float* d2 = (const float*)(prims + (size_t)tri_pos[i] * 80 + 32);
    const float* d3 = (const float*)(prims + (size_t)tri_pos[i] * 80 + 48);
    Item& it = items[i];
    for (int a = 0; a < 3; a++) {
      const float v0 = d1[a], v1 = d1[a] + d2[a], v2 = d1[a] + d3[a];
      it.lo[a] = std::min(v0, std::min(v1, v2));
      it.hi[a] = std::max(v0, std::max(v1, v2));
    }
    it.id = tri_pos[i];
    scene.grow(it.lo, it.hi);
  }
  const float ext = std::max(scene.hi[0] - scene.lo[0], std::max(scene.hi[1] - scene.lo[1], scene.hi[2] - scene.lo[2]));
  const float pad = std::max(ext, 1.0f) * (1.0f / 32768.0f);
  out.pad = pad;
  for (int a = 0; a < 3; a++) { out.lo[a] = scene.lo[a] - pad; out.hi[a] = scene.hi[a] + pad; }

  Builder b(items, n_threads < 1 ? 1 : n_threads);
  b.build(0, 0, (uint32_t)n, 0);
  out.max_depth = b.max_depth.load();

  out.tri_src.resize(n);
  for (size_t i = 0; i < n; i++) out.tri_src[i] = items[i].id;

  // flatten: inner nodes in depth-first order (left subtree right after its parent)
  const std::vector<BNode>& pool = b.pool;
  std::vector<int32_t> inner_index(pool.size(), -1);
  std::vector<uint32_t> order;
  if (pool[0].count) {
    // a lone leaf: wrap it so the root is an inner node; the second child is an empty box
    out.nodes.assign(16, 0.0f);
    float* q = out.nodes.data();
    const BNode& r = pool[0];
    q[0] = r.box.lo[0] - pad; q[1] = r.box.hi[0] + pad; q[2] = r.box.lo[1] - pad; q[3] = r.box.hi[1] + pad;
    q[4] = 1e30f; q[5] = -1e30f; q[6] = 1e30f; q[7] = -1e30f;
    q[8] = r.box.lo[2] - pad; q[9] = r.box.hi[2] + pad; q[10] = 1e30f; q[11] = -1e30f;
    int32_t c0 = ~(int32_t)((r.first << 3) | (r.count - 1));
    std::memcpy(&q[12], &c0, 4); std::memcpy(&q[13], &c0, 4);
    out.n_nodes = 1;
    return;
  }
  std::vector<uint32_t> stack{0};
  while (!stack.empty()) {
    uint32_t c = stack.back(); stack.pop_back();
    inner_index[c] = (int32_t)order.size();
    order.push_back(c);
    const BNode& nd = pool[c];
    if (!pool[nd.right].count) stack.push_back(nd.right);
    if (!pool[nd.left].count) stack.push_back(nd.left);
  }
  out.n_nodes = (uint32_t)order.size();
  out.nodes.resize((size_t)out.n_nodes * 16);
  for (size_t k = 0; k < order.size(); k++) {
    const BNode& nd = pool[order[k]];
    const BNode& c0 = pool[nd.left];
    const BNode& c1 = pool[nd.right];
    float* q = out.nodes.data() + k * 16;
    q[0] = c0.box.lo[0] - pad; q[1] = c0.box.hi[0] + pad; q[2] = c0.box.lo[1] - pad; q[3] = c0.box.hi[1] + pad;
    q[4] = c1.box.lo[0] - pad; q[5] = c1.box.hi[0] + pad; q[6] = c1.box.lo[1] - pad; q[7] = c1.box.hi[1] + pad;
    q[8] = c0.box.lo[2] - pad; q[9] = c0.box.hi[2] + pad; q[10] = c1.box.lo[2] - pad; q[11] = c1.box.hi[2] + pad;
    int32_t a = child_code(pool, nd.left, inner_index), bcode = child_code(pool, nd.right, inner_index);
    std::memcpy(&q[12], &a, 4); std::memcpy(&q[13], &bcode, 4);
    q[14] = 0.0f; q[15] = 0.0f;
  }
}

}  // namespace crt
