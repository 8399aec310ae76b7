// crt_bvh_ploc.h -- the per-element steps of the parallel BVH build, written once for both compilers.
//
// The device build of the scene-upload module (crt_bvh_device.cu) is:
//   1. triangle boxes + scene bounds            (tri_box)
//   2. 63-bit Morton codes of the box centres, radix sort                               (morton63)
//   3. PLOC -- parallel locally-ordered clustering (Meister & Bittner, "Parallel locally-ordered clustering for bounding
//      volume hierarchy construction", TVCG 2018): every cluster looks at its 2 x kRadius neighbours in Morton order for
//      the one whose union with it has the smallest area; mutual nearest neighbours merge into a new BVH2 node; the
//      cluster list is compacted; repeat until one cluster is left.  Agglomerative, so its trees are close to
//      full-sweep SAH quality, and every step is a flat loop over the clusters.
//      A node is complete when it is created (its children exist already), so the dynamic programme of the Wide8
//      collapse (Ylitie, Karras, Laine 2017, section 4.1; the same recurrence as the host builder, crt_bvh.cpp) is
//      evaluated right there (dp_node): no bottom-up pass with atomics.
//   4. top-down emission of the 8-wide nodes, one level of the wide tree per step: collect each wide node's children from
//      the DP's decisions, assign slots, count inner children and leaf triangles (emit_collect); exclusive scan; write
//      the node, the next level's frontier and the leaf-ordered triangle list (emit_node).  Scans instead of atomics:
//      the result is deterministic and numbered breadth-first exactly like the host builder's.
// Every function here is a pure per-index step over plain arrays; the CUDA kernels call them with one thread per index
// and a CPU driver (crt_bvh.cpp: build_bvh_ploc_host) calls them in loops.  That driver is what the CPU test-suite
// checks against the builder invariants (tests/test_bvh_builder.py), so the logic is validated without a GPU.
//
// Like everything in the builder, this decides which triangles get tested, never which hit wins: boxes are padded
// (DESIGN.md "Conservative culling") and the quantised planes are rounded outward in exact arithmetic.
#pragma once
#include <stdint.h>
#include <string.h>
#include <math.h>

#ifdef __CUDACC__
#define CRT_HD __host__ __device__ __forceinline__
#else
#define CRT_HD inline
#endif

#ifndef CRT_NODE_STRIDE
#define CRT_NODE_STRIDE 5
#endif

namespace crt {
namespace ploc {

// Search radius in the Morton-ordered cluster list.  Measured with tools/bvh_eval.py (the library's own tree walked on the CPU
// in wf_traverse's order; fetched bytes per queued ray, host SAH build = 798 on the 1M mesh / 2761 on a 1M soup):
//   radius   1     2     3     4     8     12    24    48
//   mesh1m   780   817   802   808   822   813   860   835
//   soup 1M  2856  2839  2854  2929  3083  3212  3201  3452
// A wide window lets far-apart clusters with a small joint box pair up early and leaves their Morton neighbours to merge
// into long thin boxes later; with the nearest neighbours only, the merges follow the space-filling curve's locality and
// the collapse's dynamic programme does the rest.  It is also the cheapest search.
#ifndef CRT_PLOC_RADIUS
#define CRT_PLOC_RADIUS 1
#endif
constexpr int kRadius = CRT_PLOC_RADIUS;
constexpr uint32_t kLeafMax = 3;       // a wide node's meta byte holds the leaf's triangle count in 3 unary bits
constexpr float kNodeCost = 1.0f;
constexpr uint32_t kNone = 0xFFFFFFFFu;

struct Box3 { float lo[3], hi[3]; };

CRT_HD float half_area(const Box3& b) {
  const float dx = b.hi[0] - b.lo[0], dy = b.hi[1] - b.lo[1], dz = b.hi[2] - b.lo[2];
  if (dx < 0.0f) return 0.0f;
  return dx * dy + dy * dz + dz * dx;
}
CRT_HD Box3 merge(const Box3& a, const Box3& b) {
  Box3 r;
  for (int k = 0; k < 3; k++) { r.lo[k] = a.lo[k] < b.lo[k] ? a.lo[k] : b.lo[k]; r.hi[k] = a.hi[k] > b.hi[k] ? a.hi[k] : b.hi[k]; }
  return r;
}

// box of the triangle at position `pos` of binding 8 (80-byte records: data1 = v0 @16, data2 = e1 @32, data3 = e2 @48), with the
// vertices evaluated in f32 exactly as the host builder does
CRT_HD Box3 tri_box(const uint8_t* prims, uint32_t pos) {
  float d1[3], d2[3], d3[3];
  memcpy(d1, prims + (size_t)pos * 80 + 16, 12);
  memcpy(d2, prims + (size_t)pos * 80 + 32, 12);
  memcpy(d3, prims + (size_t)pos * 80 + 48, 12);
  Box3 b;
  for (int a = 0; a < 3; a++) {
    const float v0 = d1[a], v1 = d1[a] + d2[a], v2 = d1[a] + d3[a];
    b.lo[a] = fminf(v0, fminf(v1, v2));
    b.hi[a] = fmaxf(v0, fmaxf(v1, v2));
  }
  return b;
}

CRT_HD uint64_t expand21(uint64_t v) {   // 21 bits -> every third bit
  v &= 0x1FFFFFull;
  v = (v | (v << 32)) & 0x1F00000000FFFFull;
  v = (v | (v << 16)) & 0x1F0000FF0000FFull;
  v = (v | (v << 8)) & 0x100F00F00F00F00Full;
  v = (v | (v << 4)) & 0x10C30C30C30C30C3ull;
  v = (v | (v << 2)) & 0x1249249249249249ull;
  return v;
}

// 63-bit Morton code of a box centre inside the centroid bounds [clo, chi]
CRT_HD uint64_t morton63(const Box3& b, const float* clo, const float* chi) {
  uint64_t q[3];
  for (int a = 0; a < 3; a++) {
    const float c = 0.5f * (b.lo[a] + b.hi[a]);
    const float ext = chi[a] - clo[a];
    double u = ext > 0.0f ? ((double)c - (double)clo[a]) / (double)ext : 0.0;
    u = u < 0.0 ? 0.0 : (u > 1.0 ? 1.0 : u);
    uint64_t v = (uint64_t)(u * 2097151.0);
    q[a] = v > 2097151ull ? 2097151ull : v;
  }
  return (expand21(q[0]) << 2) | (expand21(q[1]) << 1) | expand21(q[2]);
}

// BVH2 in creation order: nodes 0 .. n_leaves-1 are the triangles in Morton order, the inner nodes follow, children before
// parents, the root last (2 n_leaves - 2).
struct Tree {
  uint32_t n_leaves;
  Box3* box;
  uint32_t* left;     // inner nodes only (index - n_leaves is NOT applied: arrays are sized 2n)
  uint32_t* right;
  uint32_t* total;    // triangles below
  float* cost;        // [7 per node] cost[k][i-1]: cheapest way to represent subtree k as at most i children of a wide node
  uint8_t* dec;       // [8 per node] the split behind cost[k][i-1] (0 = "same as i-1"; i = 1: 0 = leaf, else split of the wide node)
};

// the collapse's dynamic programme for node k (crt_bvh.cpp has the same recurrence with the same weights)
CRT_HD void dp_node(const Tree& t, uint32_t k, float prim_cost) {
  float* c = t.cost + (size_t)k * 7;
  uint8_t* dk = t.dec + (size_t)k * 8;
  const float area = half_area(t.box[k]);
  const float inf = INFINITY;
  const uint32_t total = t.total[k];
  const float leaf = total <= kLeafMax ? area * (float)total * prim_cost : inf;
  if (k < t.n_leaves) { for (int i = 0; i < 7; i++) { c[i] = leaf; dk[i] = 0; } dk[7] = 0; return; }
  const float* cl = t.cost + (size_t)t.left[k] * 7;
  const float* cr = t.cost + (size_t)t.right[k] * 7;
  float best = inf; int best_k = 1;
  for (int j = 1; j <= 7; j++) { const float v = cl[j - 1] + cr[7 - j]; if (v < best) { best = v; best_k = j; } }   // j + (8-j) = 8 slots
  const float internal = best + area * kNodeCost;
  if (leaf <= internal) { c[0] = leaf; dk[0] = 0; } else { c[0] = internal; dk[0] = (uint8_t)best_k; }
  for (int i = 2; i <= 7; i++) {
    float v = c[i - 2]; int vk = 0;
    for (int j = 1; j < i; j++) { const float u = cl[j - 1] + cr[i - j - 1]; if (u < v) { v = u; vk = j; } }
    c[i - 1] = v; dk[i - 1] = (uint8_t)vk;
  }
  dk[7] = 0;
}

// ---- PLOC, one iteration over the cluster list cl[0 .. n_cl)
// step 1: nearest neighbour of cluster i within the radius (smallest union area; ties to the lower position)
CRT_HD uint32_t find_nn(const Tree& t, const uint32_t* cl, uint32_t n_cl, uint32_t i) {
  const Box3 bi = t.box[cl[i]];
  const uint32_t j0 = i > (uint32_t)kRadius ? i - kRadius : 0u;
  const uint32_t j1 = i + kRadius + 1u < n_cl ? i + kRadius + 1u : n_cl;
  float best = INFINITY; uint32_t bj = kNone;
  for (uint32_t j = j0; j < j1; j++) {
    if (j == i) continue;
    const float a = half_area(merge(bi, t.box[cl[j]]));
    if (a < best || bj == kNone) { best = a; bj = j; }
  }
  return bj;
}
// step 2: what becomes of cluster i -- bit 0: it stays in the list (as itself or as the new node), bit 32: it is the left
// partner of a merge and creates a node.  The exclusive scan of these words gives the output position (low half) and the
// ordinal of the new node (high half).
CRT_HD uint64_t merge_word(const uint32_t* nn, uint32_t i) {
  const uint32_t j = nn[i];
  if (j != kNone && nn[j] == i) return i < j ? ((1ull << 32) | 1ull) : 0ull;
  return 1ull;
}
// step 3: write cluster i's successor; a merge creates node `first_new + ordinal`
CRT_HD void apply_merge(const Tree& t, const uint32_t* cl, const uint32_t* nn, const uint64_t* word, const uint64_t* scan, uint32_t i,
                        uint32_t first_new, float prim_cost, uint32_t* cl_out) {
  const uint64_t w = word[i];
  if (!(w & 1ull)) return;
  const uint32_t pos = (uint32_t)(scan[i] & 0xFFFFFFFFull);
  if (w >> 32) {
    const uint32_t id = first_new + (uint32_t)(scan[i] >> 32);
    const uint32_t a = cl[i], b = cl[nn[i]];
    t.left[id] = a; t.right[id] = b;
    t.box[id] = merge(t.box[a], t.box[b]);
    t.total[id] = t.total[a] + t.total[b];
    dp_node(t, id, prim_cost);
    cl_out[pos] = id;
  } else {
    cl_out[pos] = cl[i];
  }
}

// ---- emission of the wide tree, one level per step
struct EmitParams {
  float pad;
  uint32_t exact_slots;
};

struct Collected {        // what emit_collect leaves for emit_node: the children of one wide node by slot
  uint32_t child[8];      // BVH2 node in slot sl, or kNone
};

CRT_HD bool is_leaf_child(const Tree& t, uint32_t n) { return n < t.n_leaves || t.dec[(size_t)n * 8] == 0; }

// the children of BVH2 subtree `n` when it may occupy up to `i` slots (the host builder's Collect::run, without recursion)
CRT_HD void collect(const Tree& t, uint32_t n, int i, uint32_t* out, int& nc) {
  uint32_t sn[16]; int si[16]; int sp = 0;
  sn[sp] = n; si[sp] = i; sp++;
  while (sp) {
    --sp; n = sn[sp]; i = si[sp];
    while (i > 1 && t.dec[(size_t)n * 8 + (i - 1)] == 0) i--;
    if (i == 1 || n < t.n_leaves) { out[nc++] = n; continue; }
    const int j = t.dec[(size_t)n * 8 + (i - 1)];
    sn[sp] = t.right[n]; si[sp] = i - j; sp++;   // right is pushed first: the left subtree's children come out first
    sn[sp] = t.left[n]; si[sp] = j; sp++;
  }
}

// A child's slot bits say on which side of the node centre it lies (bit set = upper side), so that slot ^ (7 ^ ray octant)
// orders the children near to far: greedy assignment by centroid offset, or the exact one (dynamic programme over slot sets).
CRT_HD void assign_slots(const Tree& t, const uint32_t* ch, int nc, const Box3& nb, bool exact, int* slot_of) {
  const float cen[3] = {0.5f * (nb.lo[0] + nb.hi[0]), 0.5f * (nb.lo[1] + nb.hi[1]), 0.5f * (nb.lo[2] + nb.hi[2])};
  float score[8][8];
  for (int i = 0; i < nc; i++) {
    const Box3& cb = t.box[ch[i]];
    const float off[3] = {0.5f * (cb.lo[0] + cb.hi[0]) - cen[0], 0.5f * (cb.lo[1] + cb.hi[1]) - cen[1], 0.5f * (cb.lo[2] + cb.hi[2]) - cen[2]};
    for (int sl = 0; sl < 8; sl++) score[i][sl] = ((sl & 1) ? off[0] : -off[0]) + ((sl & 2) ? off[1] : -off[1]) + ((sl & 4) ? off[2] : -off[2]);
  }
  if (!exact) {
    bool slot_used[8] = {false, false, false, false, false, false, false, false};
    bool child_done[8] = {false, false, false, false, false, false, false, false};
    for (int round = 0; round < nc; round++) {
      int bi = -1, bs = -1; float bv = -INFINITY;
      for (int i = 0; i < nc; i++) {
        if (child_done[i]) continue;
        for (int sl = 0; sl < 8; sl++) {
          if (slot_used[sl]) continue;
          if (score[i][sl] > bv || bi < 0) { bv = score[i][sl]; bi = i; bs = sl; }
        }
      }
      slot_of[bi] = bs; slot_used[bs] = true; child_done[bi] = true;
    }
    return;
  }
  float dp[256]; signed char from[256];
  for (int m = 0; m < 256; m++) { dp[m] = -INFINITY; from[m] = -1; }
  dp[0] = 0.0f;
  for (int m = 0; m < 256; m++) {
    int k = 0; for (int b = 0; b < 8; b++) k += (m >> b) & 1;
    if (k >= nc || (m != 0 && from[m] < 0)) continue;
    for (int sl = 0; sl < 8; sl++) {
      if (m & (1 << sl)) continue;
      const float v = dp[m] + score[k][sl];
      if (v > dp[m | (1 << sl)] || from[m | (1 << sl)] < 0) { dp[m | (1 << sl)] = v; from[m | (1 << sl)] = (signed char)sl; }
    }
  }
  int bm = -1; float bv = -INFINITY;
  for (int m = 0; m < 256; m++) {
    int k = 0; for (int b = 0; b < 8; b++) k += (m >> b) & 1;
    if (k == nc && from[m] >= 0 && (dp[m] > bv || bm < 0)) { bv = dp[m]; bm = m; }
  }
  for (int k = nc - 1, m = bm; k >= 0; k--) { slot_of[k] = from[m]; m &= ~(1 << from[m]); }
}

// step 1 of a level: the children of the wide node made from BVH2 node `src`; returns inner children (low half) and leaf
// triangles (high half) for the scan
CRT_HD uint64_t emit_collect(const Tree& t, uint32_t src, const EmitParams& ep, Collected& out) {
  uint32_t ch[8]; int nc = 0;
  if (src < t.n_leaves || t.dec[(size_t)src * 8] == 0) {
    ch[nc++] = src;                                    // the whole scene is one leaf (<= 3 triangles): only the root can be
  } else {
    const int j = t.dec[(size_t)src * 8];              // the split the wide node of this subtree was costed with
    collect(t, t.left[src], j, ch, nc);
    collect(t, t.right[src], 8 - j, ch, nc);
  }
  Box3 nb = t.box[ch[0]];
  for (int i = 1; i < nc; i++) nb = merge(nb, t.box[ch[i]]);
  int slot_of[8];
  assign_slots(t, ch, nc, nb, ep.exact_slots != 0u, slot_of);
  for (int sl = 0; sl < 8; sl++) out.child[sl] = kNone;
  uint32_t inner = 0, tris = 0;
  for (int i = 0; i < nc; i++) {
    out.child[slot_of[i]] = ch[i];
    if (is_leaf_child(t, ch[i])) tris += t.total[ch[i]]; else inner++;
  }
  return (uint64_t)inner | ((uint64_t)tris << 32);
}

// exponent e of the quantisation step, 2^e >= ext / 255, clamped so that both 2^e and 2^(e+15) are normal floats
CRT_HD int scale_exp(float ext) {
  int e = -126;
  if (ext > 0.0f) {
    e = (int)ilogb((double)ext / 255.0);
    while (ldexp(255.0, e) < (double)ext) e++;
    while (e > -126 && ldexp(255.0, e - 1) >= (double)ext) e--;
    if (e < -126) e = -126;
    if (e > 112) e = 112;
  }
  return e;
}

CRT_HD uint32_t f2u(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }

// step 2 of a level: write wide node `node_index` (layout: crt_device.cuh "Wide8"), its inner children's frontier entries
// (they become nodes child_base + rank in slot order) and its leaf triangles' entries of the leaf-ordered list.
//   leaf_tri[k] = triangle (ordinal in the caller's list) at BVH2 leaf k; tri_out[slot] = that ordinal
CRT_HD void emit_node(const Tree& t, const Collected& col, const EmitParams& ep, uint32_t node_index, uint32_t child_base, uint32_t tri_base,
                      const uint32_t* leaf_tri, uint32_t* nodes, uint32_t* next_frontier, uint32_t next_frontier_base, uint32_t* tri_out) {
  const float pad = ep.pad;
  Box3 nb; bool first = true;
  for (int sl = 0; sl < 8; sl++) {
    if (col.child[sl] == kNone) continue;
    nb = first ? t.box[col.child[sl]] : merge(nb, t.box[col.child[sl]]);
    first = false;
  }
  // node frame: origin = lower corner of the children's padded boxes, one power-of-two scale per axis
  float p[3]; uint32_t eb[3]; double sc[3];   // eb: biased exponent of 2^15 * step
  for (int a = 0; a < 3; a++) {
    p[a] = nb.lo[a] - pad;
    const int e = scale_exp((nb.hi[a] + pad) - p[a] + 2.0f * pad);   // slack for the float rounding of hi + pad - p
    eb[a] = (uint32_t)(e + 15 + 127);
    sc[a] = ldexp(1.0, e);
  }
  uint32_t w[20];
  for (int i = 0; i < 20; i++) w[i] = 0;
  uint8_t meta[8] = {0, 0, 0, 0, 0, 0, 0, 0}, qlo[3][8], qhi[3][8];
  for (int a = 0; a < 3; a++) for (int sl = 0; sl < 8; sl++) { qlo[a][sl] = 255; qhi[a][sl] = 0; }
  uint32_t imask = 0, n_inner = 0, tri_off = 0;
  for (int sl = 0; sl < 8; sl++) {
    const uint32_t c = col.child[sl];
    if (c == kNone) continue;
    const Box3& cb = t.box[c];
    for (int a = 0; a < 3; a++) {   // conservative in exact arithmetic: p + qlo * 2^e <= lo - pad, p + qhi * 2^e >= hi + pad
      const double lo = (double)(cb.lo[a] - pad) - (double)p[a], hi = (double)(cb.hi[a] + pad) - (double)p[a];
      double ql = floor(lo / sc[a]), qh = ceil(hi / sc[a]);
      ql = ql < 0.0 ? 0.0 : (ql > 255.0 ? 255.0 : ql);
      qh = qh < 0.0 ? 0.0 : (qh > 255.0 ? 255.0 : qh);
      qlo[a][sl] = (uint8_t)ql; qhi[a][sl] = (uint8_t)qh;
    }
    if (is_leaf_child(t, c)) {
      // the (<= 3) triangles below c, left to right
      uint32_t st[8]; int sp = 0; st[sp++] = c;
      const uint32_t cnt = t.total[c];
      uint32_t k = 0;
      while (sp) {
        const uint32_t x = st[--sp];
        if (x < t.n_leaves) { tri_out[tri_base + tri_off + k] = leaf_tri[x]; k++; }
        else { st[sp++] = t.right[x]; st[sp++] = t.left[x]; }
      }
      meta[sl] = (uint8_t)((((1u << cnt) - 1u) << 5) | tri_off);
      tri_off += cnt;
    } else {
      imask |= 1u << sl;
      meta[sl] = (uint8_t)((1u << 5) | (24u + (uint32_t)sl));
      next_frontier[next_frontier_base + n_inner] = c;
      n_inner++;
    }
  }
  w[0] = f2u(p[0]); w[1] = f2u(p[1]); w[2] = f2u(p[2]);
  w[3] = eb[0] | (eb[1] << 8) | (eb[2] << 16) | (imask << 24);
  w[4] = child_base; w[5] = tri_base;
  memcpy(&w[6], meta, 8);
  memcpy(&w[8], qlo[0], 8); memcpy(&w[10], qlo[1], 8); memcpy(&w[12], qlo[2], 8);
  memcpy(&w[14], qhi[0], 8); memcpy(&w[16], qhi[1], 8); memcpy(&w[18], qhi[2], 8);
  uint32_t* o = nodes + (size_t)node_index * 4 * CRT_NODE_STRIDE;
  for (int i = 0; i < 20; i++) o[i] = w[i];
  for (int i = 20; i < 4 * CRT_NODE_STRIDE; i++) o[i] = 0u;
}

}  // namespace ploc
}  // namespace crt
