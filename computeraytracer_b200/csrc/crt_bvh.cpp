// crt_bvh.cpp -- binned-SAH BVH2 builder (host side of the scene-upload module).
//
// New capability: the reference's intersect() is a linear loop (ComputeShader.wgsl:503-518); the
// 1M / 10M-triangle configs need an accelerator that returns the same answer.  The tree only decides
// which primitives get tested; the tests themselves and the tie rule live in crt_device.cuh, and the
// boxes are padded so that culling is conservative with respect to the f32 rounding of those tests
// (DESIGN.md "Conservative culling").  Tree shape therefore affects speed, never results.
#include "crt_bvh.h"
#include "crt_bvh_ploc.h"

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <future>
#include <limits>
#include <numeric>

namespace crt {
namespace {

constexpr int kBinsDefault = 32, kBinsMax = 256;   // SAH bins per axis (BvhOptions::bins; 128 bins + exact slots: wf_traverse -1.7 %, profiles/r02a_traverse_schedule_ab.txt)
constexpr uint32_t kLeafMax = 3;         // a wide node's meta byte holds the leaf's triangle count in 3 unary bits
constexpr uint32_t kBvh2LeafMax = 1;     // the binary tree is split down to single triangles; the collapse decides the leaves
constexpr float kNodeCost = 1.0f;        // SAH weights of the collapse (Ylitie et al. 2017, section 4.1); the triangle weight
// (BvhOptions::prim_cost, default 0.75; measured 0.3 .. 1.5 on mesh1m and soup10m: flat above 0.75; 0.3 is 1 % / 5 % slower)
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

struct BNode { Box box; uint32_t left = 0, right = 0, first = 0, count = 0, total = 0; };   // count != 0: BVH2 leaf; total: triangles below

struct Builder {
  std::vector<Item>& items;
  std::vector<BNode> pool;
  std::atomic<uint32_t> next{1};
  std::atomic<uint32_t> max_depth{0};
  int par_depth;
  int kBins = kBinsDefault;

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
    n.box = box; n.first = first; n.count = 0; n.total = count;
    note_depth(depth);
    if (count <= kBvh2LeafMax) { n.count = count; return; }

    uint32_t mid = 0;
    bool have_split = false;
    if (depth < kForceMedianDepth) {
      float best_cost = std::numeric_limits<float>::infinity();
      int best_axis = -1, best_bin = -1;
      for (int ax = 0; ax < 3; ax++) {
        const float c0 = cbox.lo[ax], ext = cbox.hi[ax] - cbox.lo[ax];
        if (!(ext > 0.0f)) continue;
        const float scale = (float)kBins / ext;
        Box bb[kBinsMax]; uint32_t bc[kBinsMax];
        for (int b = 0; b < kBins; b++) { bb[b].reset(); bc[b] = 0; }
        for (uint32_t i = first; i < first + count; i++) {
          const Item& it = items[i];
          int b = (int)((0.5f * (it.lo[ax] + it.hi[ax]) - c0) * scale);
          b = b < 0 ? 0 : (b >= kBins ? kBins - 1 : b);
          bb[b].grow(it.lo, it.hi); bc[b]++;
        }
        float right_area[kBinsMax]; uint32_t right_cnt[kBinsMax];
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
        if (count <= kBvh2LeafMax && (float)count <= split_cost) { n.count = count; return; }
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
      if (count <= kBvh2LeafMax && depth < kForceMedianDepth) { n.count = count; return; }
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

// ---- collapse to an 8-wide BVH with quantised child boxes (node layout: crt_device.cuh, "Wide8")
struct WideChild { uint32_t bvh2; float c[3]; };

inline uint32_t f2u(float f) { uint32_t u; std::memcpy(&u, &f, 4); return u; }

// exponent e of the quantisation step, 2^e >= ext / 255, clamped so that both 2^e and 2^(e+15) are normal floats (the node
// stores the biased exponent of 2^15 * step: the device builds that factor with one shift, crt_device.cuh wide_node)
inline int scale_exp(float ext) {
  int e = -126;
  if (ext > 0.0f) {
    e = (int)std::ceil(std::log2((double)ext / 255.0));
    while (std::ldexp(255.0, e) < (double)ext) e++;   // guard against log2 rounding
    if (e < -126) e = -126;
    if (e > 112) e = 112;
  }
  return e;
}

}  // namespace

BvhOptions bvh_options(uint32_t bvh_tuning, float prim_cost) {
  BvhOptions o;
  const int bins = (int)(bvh_tuning & 0x1FFu);
  if (bins >= 4 && bins <= kBinsMax) o.bins = bins;
  o.optimal_slots = (bvh_tuning & 0x200u) != 0u;
  if (prim_cost > 0.0f) o.prim_cost = prim_cost;
  return o;
}

void build_bvh(const uint8_t* prims, const std::vector<uint32_t>& tri_pos, int n_threads, const BvhOptions& opt, BvhResult& out) {
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
  // The padding has to cover what f32 does to the triangles themselves, not only the extent of the mesh: a vertex v0 + e1 and
  // the difference o - v0 are rounded to an ulp of the COORDINATE, so a mesh of extent 2 around x = 2000 needs a pad that
  // scales with 2000 (ext / 32768 would be below one ulp there and vanish in lo - pad).
  float maxabs = 0.0f;
  for (int a = 0; a < 3; a++) maxabs = std::max(maxabs, std::max(std::fabs(scene.lo[a]), std::fabs(scene.hi[a])));
  const float pad = std::max(std::max(ext, maxabs), 1.0f) * (1.0f / 32768.0f);
  out.pad = pad;
  for (int a = 0; a < 3; a++) { out.lo[a] = scene.lo[a] - pad; out.hi[a] = scene.hi[a] + pad; }

  Builder b(items, n_threads < 1 ? 1 : n_threads);
  b.kBins = opt.bins;
  b.build(0, 0, (uint32_t)n, 0);
  const std::vector<BNode>& pool = b.pool;

  const bool optimal_slots = opt.optimal_slots;
  // ---- which BVH2 subtrees become wide nodes, which become leaves: the SAH-optimal collapse of Ylitie et al. (2017,
  // section 4.1) by dynamic programming.  cost[n][i-1] = cheapest way to represent the subtree of n as at most i
  // children of a wide node (i = 1: one child, either a leaf of <= 3 triangles or a wide node of its own).
  const uint32_t n_bvh2 = b.next.load();
  const float kPrimCost = opt.prim_cost;
  std::vector<float> cost((size_t)n_bvh2 * 7);
  std::vector<uint8_t> dec((size_t)n_bvh2 * 8);   // [0..6]: split for i = 1..7 (0 = "same as i-1"; for i = 1: 0 = leaf, else split of the wide node); 
  const float inf = std::numeric_limits<float>::infinity();
  for (uint32_t k = n_bvh2; k-- > 0;) {           // children are allocated after their parent: reverse order is post-order
    const BNode& nd = pool[k];
    float* c = &cost[(size_t)k * 7];
    uint8_t* dk = &dec[(size_t)k * 8];
    const float area = nd.box.half_area();
    const float leaf = nd.total <= kLeafMax ? area * (float)nd.total * kPrimCost : inf;
    if (nd.count) { for (int i = 0; i < 7; i++) { c[i] = leaf; dk[i] = 0; } continue; }
    const float* cl = &cost[(size_t)nd.left * 7];
    const float* cr = &cost[(size_t)nd.right * 7];
    float best = inf; int best_k = 1;
    for (int j = 1; j <= 7; j++) { const float v = cl[j - 1] + cr[7 - j]; if (v < best) { best = v; best_k = j; } }   // j + (8-j) = 8 slots
    const float internal = best + area * kNodeCost;
    if (leaf <= internal) { c[0] = leaf; dk[0] = 0; } else { c[0] = internal; dk[0] = (uint8_t)best_k; }
    for (int i = 2; i <= 7; i++) {
      float v = c[i - 2]; int vk = 0;
      for (int j = 1; j < i; j++) { const float u = cl[j - 1] + cr[i - j - 1]; if (u < v) { v = u; vk = j; } }
      c[i - 1] = v; dk[i - 1] = (uint8_t)vk;
    }
  }
  // the children of BVH2 subtree n when it may occupy up to i slots
  struct Collect {
    const std::vector<BNode>& pool; const std::vector<uint8_t>& dec;
    void run(uint32_t n, int i, uint32_t* out, int& nc) const {
      while (i > 1 && dec[(size_t)n * 8 + (i - 1)] == 0) i--;
      if (i == 1 || pool[n].count) { out[nc++] = n; return; }
      const int j = dec[(size_t)n * 8 + (i - 1)];
      run(pool[n].left, j, out, nc);
      run(pool[n].right, i - j, out, nc);
    }
  } collect{pool, dec};
  auto is_leaf_child = [&](uint32_t n) { return pool[n].count != 0 || dec[(size_t)n * 8] == 0; };

  // Wide nodes are created breadth-first: the inner children of a node get consecutive indices (child_base + rank in
  // slot order), its leaf children's triangles consecutive slots from tri_base (offset + count <= 24).
  std::vector<uint32_t> wide_src{0};      // wide node -> BVH2 node it was collapsed from
  std::vector<uint32_t> wide_depth{1};
  out.tri_src.reserve(n);
  out.nodes.reserve((n / 4 + 16) * 4 * CRT_NODE_STRIDE);
  for (size_t k = 0; k < wide_src.size(); k++) {
    const BNode& top = pool[wide_src[k]];
    uint32_t ch[8]; int nc = 0;
    if (top.count || (k == 0 && dec[0] == 0)) {
      ch[nc++] = wide_src[k];                              // the whole scene is one leaf (<= 3 triangles)
    } else {
      const int j = dec[(size_t)wide_src[k] * 8];          // the split the wide node of this subtree was costed with
      collect.run(top.left, j, ch, nc);
      collect.run(top.right, 8 - j, ch, nc);
    }
    // node frame: origin = lower corner of the children's padded boxes, one power-of-two scale per axis
    Box nb; nb.reset();
    for (int i = 0; i < nc; i++) nb.grow(pool[ch[i]].box);
    float p[3]; uint32_t eb[3]; double sc[3];   // eb: biased exponent of 2^15 * step
    for (int a = 0; a < 3; a++) {
      p[a] = nb.lo[a] - pad;
      const int e = scale_exp((nb.hi[a] + pad) - p[a] + 2.0f * pad);   // slack for the float rounding of hi + pad - p
      eb[a] = (uint32_t)(e + 15 + 127);
      sc[a] = std::ldexp(1.0, e);
    }
    // slots: a child's slot bits say on which side of the node centre it lies (bit set = upper side), so that
    // slot ^ (7 ^ ray octant) orders the children near to far; greedy assignment by centroid offset
    float cen[3] = {0.5f * (nb.lo[0] + nb.hi[0]), 0.5f * (nb.lo[1] + nb.hi[1]), 0.5f * (nb.lo[2] + nb.hi[2])};
    int slot_of[8]; bool slot_used[8] = {false, false, false, false, false, false, false, false};
    bool child_done[8] = {false, false, false, false, false, false, false, false};
    for (int round = 0; round < nc; round++) {
      int bi = -1, bs = -1; float bv = -std::numeric_limits<float>::infinity();
      for (int i = 0; i < nc; i++) {
        if (child_done[i]) continue;
        const Box& cb = pool[ch[i]].box;
        const float off[3] = {0.5f * (cb.lo[0] + cb.hi[0]) - cen[0], 0.5f * (cb.lo[1] + cb.hi[1]) - cen[1], 0.5f * (cb.lo[2] + cb.hi[2]) - cen[2]};
        for (int sl = 0; sl < 8; sl++) {
          if (slot_used[sl]) continue;
          const float v = ((sl & 1) ? off[0] : -off[0]) + ((sl & 2) ? off[1] : -off[1]) + ((sl & 4) ? off[2] : -off[2]);
          if (v > bv) { bv = v; bi = i; bs = sl; }
        }
      }
      slot_of[bi] = bs; slot_used[bs] = true; child_done[bi] = true;
    }
    if (optimal_slots) {
      // tuning (BvhOptions::optimal_slots): the assignment that maximises the summed score exactly, by dynamic programming over the
      // sets of used slots (child k takes a free slot of a set with k slots used)
      float score[8][8];
      for (int i = 0; i < nc; i++) {
        const Box& cb = pool[ch[i]].box;
        const float off[3] = {0.5f * (cb.lo[0] + cb.hi[0]) - cen[0], 0.5f * (cb.lo[1] + cb.hi[1]) - cen[1], 0.5f * (cb.lo[2] + cb.hi[2]) - cen[2]};
        for (int sl = 0; sl < 8; sl++) score[i][sl] = ((sl & 1) ? off[0] : -off[0]) + ((sl & 2) ? off[1] : -off[1]) + ((sl & 4) ? off[2] : -off[2]);
      }
      float dp[256]; int8_t from[256];
      for (int m = 0; m < 256; m++) { dp[m] = -std::numeric_limits<float>::infinity(); from[m] = -1; }
      dp[0] = 0.0f;
      for (int m = 0; m < 256; m++) {
        const int k = __builtin_popcount((unsigned)m);
        if (k >= nc || dp[m] == -std::numeric_limits<float>::infinity()) continue;
        for (int sl = 0; sl < 8; sl++) {
          if (m & (1 << sl)) continue;
          const float v = dp[m] + score[k][sl];
          if (v > dp[m | (1 << sl)]) { dp[m | (1 << sl)] = v; from[m | (1 << sl)] = (int8_t)sl; }
        }
      }
      int bm = -1; float bv = -std::numeric_limits<float>::infinity();
      for (int m = 0; m < 256; m++) if (__builtin_popcount((unsigned)m) == nc && dp[m] > bv) { bv = dp[m]; bm = m; }
      for (int k = nc - 1, m = bm; k >= 0; k--) { slot_of[k] = from[m]; m &= ~(1 << from[m]); }
    }
    int child_in[8]; for (int sl = 0; sl < 8; sl++) child_in[sl] = -1;
    for (int i = 0; i < nc; i++) child_in[slot_of[i]] = i;

    uint32_t w[20];
    for (int i = 0; i < 20; i++) w[i] = 0;
    uint8_t meta[8] = {0, 0, 0, 0, 0, 0, 0, 0}, qlo[3][8], qhi[3][8];
    for (int a = 0; a < 3; a++) for (int sl = 0; sl < 8; sl++) { qlo[a][sl] = 255; qhi[a][sl] = 0; }
    uint32_t imask = 0;
    const uint32_t child_base = (uint32_t)wide_src.size(), tri_base = (uint32_t)out.tri_src.size();
    for (int sl = 0; sl < 8; sl++) {
      const int i = child_in[sl];
      if (i < 0) continue;
      const BNode& c = pool[ch[i]];
      for (int a = 0; a < 3; a++) {   // conservative in exact arithmetic: p + qlo * 2^e <= lo - pad, p + qhi * 2^e >= hi + pad
        const double lo = (double)(c.box.lo[a] - pad) - (double)p[a], hi = (double)(c.box.hi[a] + pad) - (double)p[a];
        double ql = std::floor(lo / sc[a]), qh = std::ceil(hi / sc[a]);
        ql = ql < 0.0 ? 0.0 : (ql > 255.0 ? 255.0 : ql);
        qh = qh < 0.0 ? 0.0 : (qh > 255.0 ? 255.0 : qh);
        qlo[a][sl] = (uint8_t)ql; qhi[a][sl] = (uint8_t)qh;
      }
      if (is_leaf_child(ch[i])) {
        const uint32_t off = (uint32_t)out.tri_src.size() - tri_base;
        for (uint32_t t = 0; t < c.total; t++) out.tri_src.push_back(items[c.first + t].id);   // a subtree's items are contiguous
        meta[sl] = (uint8_t)((((1u << c.total) - 1u) << 5) | off);
      } else {
        imask |= 1u << sl;
        meta[sl] = (uint8_t)((1u << 5) | (24u + (uint32_t)sl));
        wide_src.push_back(ch[i]);
        wide_depth.push_back(wide_depth[k] + 1);
      }
    }
    w[0] = f2u(p[0]); w[1] = f2u(p[1]); w[2] = f2u(p[2]);
    w[3] = eb[0] | (eb[1] << 8) | (eb[2] << 16) | (imask << 24);
    w[4] = child_base; w[5] = tri_base;
    std::memcpy(&w[6], meta, 8);
    std::memcpy(&w[8], qlo[0], 8); std::memcpy(&w[10], qlo[1], 8); std::memcpy(&w[12], qlo[2], 8);
    std::memcpy(&w[14], qhi[0], 8); std::memcpy(&w[16], qhi[1], 8); std::memcpy(&w[18], qhi[2], 8);
    out.nodes.insert(out.nodes.end(), w, w + 20);
    for (int pad_w = 20; pad_w < 4 * CRT_NODE_STRIDE; pad_w++) out.nodes.push_back(0u);
    out.max_depth = std::max(out.max_depth, wide_depth[k]);
  }
  out.n_nodes = (uint32_t)wide_src.size();
}

// The device build's algorithm (crt_bvh_ploc.h: Morton sort, PLOC with the collapse's dynamic programme, level-by-level
// emission) driven by plain loops on the host.  The CUDA kernels of crt_bvh_device.cu run the same per-element functions
// with one thread per index; this driver exists so that the CPU test-suite can check that logic against the builder
// invariants without a GPU (crt_host_build_bvh with CRT_BVH_BUILDER_DEVICE).  Single-threaded, not a product path.
void build_bvh_ploc_host(const uint8_t* prims, const std::vector<uint32_t>& tri_pos, const BvhOptions& opt, BvhResult& out) {
  using namespace ploc;
  out.nodes.clear(); out.tri_src.clear(); out.n_nodes = 0; out.max_depth = 0; out.pad = 0.0f;
  const uint32_t n = (uint32_t)tri_pos.size();
  if (!n) return;
  std::vector<Box3> boxes_in(n);
  float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
  float clo[3] = {INFINITY, INFINITY, INFINITY}, chi[3] = {-INFINITY, -INFINITY, -INFINITY};
  for (uint32_t i = 0; i < n; i++) {
    boxes_in[i] = tri_box(prims, tri_pos[i]);
    for (int a = 0; a < 3; a++) {
      lo[a] = std::min(lo[a], boxes_in[i].lo[a]); hi[a] = std::max(hi[a], boxes_in[i].hi[a]);
      const float c = 0.5f * (boxes_in[i].lo[a] + boxes_in[i].hi[a]);
      clo[a] = std::min(clo[a], c); chi[a] = std::max(chi[a], c);
    }
  }
  float ext = 0.0f, maxabs = 0.0f;
  for (int a = 0; a < 3; a++) { ext = std::max(ext, hi[a] - lo[a]); maxabs = std::max(maxabs, std::max(std::fabs(lo[a]), std::fabs(hi[a]))); }
  const float pad = std::max(std::max(ext, maxabs), 1.0f) * (1.0f / 32768.0f);
  out.pad = pad;
  for (int a = 0; a < 3; a++) { out.lo[a] = lo[a] - pad; out.hi[a] = hi[a] + pad; }
  std::vector<uint64_t> keys(n);
  std::vector<uint32_t> sorted(n);
  for (uint32_t i = 0; i < n; i++) keys[i] = morton63(boxes_in[i], clo, chi);
  std::iota(sorted.begin(), sorted.end(), 0u);
  std::stable_sort(sorted.begin(), sorted.end(), [&](uint32_t a, uint32_t b) { return keys[a] < keys[b]; });   // the radix sort is stable too

  const size_t n2 = 2 * (size_t)n;
  std::vector<Box3> box(n2); std::vector<uint32_t> left(n2), right(n2), total(n2); std::vector<float> cost(n2 * 7); std::vector<uint8_t> dec(n2 * 8);
  Tree t; t.n_leaves = n; t.box = box.data(); t.left = left.data(); t.right = right.data(); t.total = total.data(); t.cost = cost.data(); t.dec = dec.data();
  std::vector<uint32_t> leaf_tri(n), cl_a(n), cl_b(n), nn(n);
  std::vector<uint64_t> word(std::max<size_t>(n, 1)), scan(std::max<size_t>(n, 1));
  for (uint32_t k = 0; k < n; k++) { box[k] = boxes_in[sorted[k]]; total[k] = 1; dp_node(t, k, opt.prim_cost); leaf_tri[k] = tri_pos[sorted[k]]; cl_a[k] = k; }
  uint32_t n_cl = n, n_nodes2 = n;
  while (n_cl > 1) {
    for (uint32_t i = 0; i < n_cl; i++) nn[i] = find_nn(t, cl_a.data(), n_cl, i);
    uint64_t acc = 0;
    for (uint32_t i = 0; i < n_cl; i++) { word[i] = merge_word(nn.data(), i); scan[i] = acc; acc += word[i]; }
    for (uint32_t i = 0; i < n_cl; i++) apply_merge(t, cl_a.data(), nn.data(), word.data(), scan.data(), i, n_nodes2, opt.prim_cost, cl_b.data());
    n_cl = (uint32_t)(acc & 0xFFFFFFFFull); n_nodes2 += (uint32_t)(acc >> 32);
    cl_a.swap(cl_b);
  }
  const uint32_t root = n_nodes2 - 1u;
  std::vector<uint32_t> frontier_a{root}, frontier_b;
  EmitParams ep; ep.pad = pad; ep.exact_slots = opt.optimal_slots ? 1u : 0u;
  out.tri_src.assign(n, 0u);
  uint32_t level_base = 0, tri_before = 0, depth = 0;
  while (!frontier_a.empty()) {
    const uint32_t n_f = (uint32_t)frontier_a.size();
    std::vector<Collected> col(n_f);
    word.assign(n_f, 0); scan.assign(n_f, 0);
    uint64_t acc = 0;
    for (uint32_t f = 0; f < n_f; f++) { word[f] = emit_collect(t, frontier_a[f], ep, col[f]); scan[f] = acc; acc += word[f]; }
    frontier_b.assign((size_t)(acc & 0xFFFFFFFFull), 0u);
    out.nodes.resize((size_t)(level_base + n_f) * 4 * CRT_NODE_STRIDE);
    for (uint32_t f = 0; f < n_f; f++) {
      const uint32_t inner_off = (uint32_t)(scan[f] & 0xFFFFFFFFull), tri_off = (uint32_t)(scan[f] >> 32);
      emit_node(t, col[f], ep, level_base + f, level_base + n_f + inner_off, tri_before + tri_off, leaf_tri.data(), out.nodes.data(), frontier_b.data(), inner_off,
                out.tri_src.data());
    }
    level_base += n_f; tri_before += (uint32_t)(acc >> 32);
    frontier_a.swap(frontier_b);
    depth++;
  }
  out.n_nodes = level_base; out.max_depth = depth;
}

}  // namespace crt
