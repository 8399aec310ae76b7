// crt_bvh_device.cu -- the scene-upload module's BVH build on the GPU: Morton sort, PLOC clustering with the Wide8
// collapse's dynamic programme evaluated as nodes are created, level-by-level emission of the 8-wide nodes.
// The per-element logic is crt_bvh_ploc.h (shared with the CPU driver the test-suite runs); this file is the kernels
// around it, the scans / sort (CUB) and the host loop.  Round 1 built the tree on the host cores: 0.6 s at 1M triangles,
// 5.7 s at 10M, re-paid on every scene edit (crt_set_binding(8)).
#include "crt_bvh_device.h"

#include <algorithm>
#include <cub/cub.cuh>
#include <vector>

#include "crt_bvh.h"
#include "crt_bvh_ploc.h"
#include "crt_device.cuh"
#include "crt_kernels.h"

namespace {

using namespace crt::ploc;

struct Bounds {              // scene bounds as order-preserving integers (atomicMin / atomicMax), then centroid bounds
  int lo[3], hi[3], clo[3], chi[3];
  int maxabs;
};

__device__ __forceinline__ int f2ord(float f) { const int i = __float_as_int(f); return i >= 0 ? i : i ^ 0x7FFFFFFF; }
__host__ __device__ __forceinline__ float ord2f(int i) { const int j = i >= 0 ? i : i ^ 0x7FFFFFFF; float f; memcpy(&f, &j, 4); return f; }

__global__ void bounds_init_kernel(Bounds* b) {
  for (int a = 0; a < 3; a++) { b->lo[a] = b->clo[a] = 0x7FFFFFFF; b->hi[a] = b->chi[a] = (int)0x80000000; }
  b->maxabs = 0;
}

// leaf boxes (in input order) + scene and centroid bounds
__global__ void tri_boxes_kernel(const uint8_t* __restrict__ prims, const uint32_t* __restrict__ tri_pos, uint32_t n, Box3* __restrict__ boxes, Bounds* bounds) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
  float clo[3] = {INFINITY, INFINITY, INFINITY}, chi[3] = {-INFINITY, -INFINITY, -INFINITY};
  if (i < n) {
    const Box3 b = tri_box(prims, tri_pos[i]);
    boxes[i] = b;
    for (int a = 0; a < 3; a++) { lo[a] = b.lo[a]; hi[a] = b.hi[a]; clo[a] = chi[a] = 0.5f * (b.lo[a] + b.hi[a]); }
  }
  for (int a = 0; a < 3; a++) {
    for (int o = 16; o > 0; o >>= 1) {
      lo[a] = fminf(lo[a], __shfl_xor_sync(0xffffffffu, lo[a], o)); hi[a] = fmaxf(hi[a], __shfl_xor_sync(0xffffffffu, hi[a], o));
      clo[a] = fminf(clo[a], __shfl_xor_sync(0xffffffffu, clo[a], o)); chi[a] = fmaxf(chi[a], __shfl_xor_sync(0xffffffffu, chi[a], o));
    }
  }
  if ((threadIdx.x & 31u) == 0 && lo[0] <= hi[0]) {
    for (int a = 0; a < 3; a++) {
      atomicMin(&bounds->lo[a], f2ord(lo[a])); atomicMax(&bounds->hi[a], f2ord(hi[a]));
      atomicMin(&bounds->clo[a], f2ord(clo[a])); atomicMax(&bounds->chi[a], f2ord(chi[a]));
      atomicMax(&bounds->maxabs, __float_as_int(fmaxf(fabsf(lo[a]), fabsf(hi[a]))));   // non-negative floats order like ints
    }
  }
}

__global__ void morton_kernel(const Box3* __restrict__ boxes, uint32_t n, const Bounds* __restrict__ bounds, uint64_t* __restrict__ keys, uint32_t* __restrict__ vals) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  float clo[3], chi[3];
  for (int a = 0; a < 3; a++) { clo[a] = ord2f(bounds->clo[a]); chi[a] = ord2f(bounds->chi[a]); }
  keys[i] = morton63(boxes[i], clo, chi);
  vals[i] = i;
}

// BVH2 leaves in Morton order: leaf k = triangle sorted[k]
__global__ void leaves_kernel(Tree t, const Box3* __restrict__ boxes_in, const uint32_t* __restrict__ sorted, const uint32_t* __restrict__ tri_pos,
                              uint32_t* __restrict__ leaf_tri, uint32_t* __restrict__ cl, float prim_cost) {
  const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= t.n_leaves) return;
  const uint32_t src = sorted[k];
  t.box[k] = boxes_in[src];
  t.total[k] = 1u;
  dp_node(t, k, prim_cost);
  leaf_tri[k] = tri_pos[src];
  cl[k] = k;
}

__global__ void ploc_nn_kernel(Tree t, const uint32_t* __restrict__ cl, uint32_t n_cl, uint32_t* __restrict__ nn) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n_cl) nn[i] = find_nn(t, cl, n_cl, i);
}
__global__ void ploc_word_kernel(const uint32_t* __restrict__ nn, uint32_t n_cl, uint64_t* __restrict__ word) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n_cl) word[i] = merge_word(nn, i);
}
// (the scan is exclusive: the totals are the last element's scan + word)
__global__ void ploc_apply_kernel(Tree t, const uint32_t* __restrict__ cl, const uint32_t* __restrict__ nn, const uint64_t* __restrict__ word,
                                  const uint64_t* __restrict__ scan, uint32_t n_cl, uint32_t first_new, float prim_cost, uint32_t* __restrict__ cl_out,
                                  uint64_t* __restrict__ totals) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_cl) return;
  apply_merge(t, cl, nn, word, scan, i, first_new, prim_cost, cl_out);
  if (i == n_cl - 1u) *totals = scan[i] + word[i];
}

__global__ void emit_collect_kernel(Tree t, const uint32_t* __restrict__ frontier, uint32_t n_f, EmitParams ep, Collected* __restrict__ col, uint64_t* __restrict__ word) {
  const uint32_t f = blockIdx.x * blockDim.x + threadIdx.x;
  if (f >= n_f) return;
  Collected c;
  word[f] = emit_collect(t, frontier[f], ep, c);
  col[f] = c;
}
__global__ void emit_node_kernel(Tree t, const Collected* __restrict__ col, const uint64_t* __restrict__ word, const uint64_t* __restrict__ scan, uint32_t n_f,
                                 EmitParams ep, uint32_t level_base, uint32_t tri_before, const uint32_t* __restrict__ leaf_tri, uint32_t* __restrict__ nodes,
                                 uint32_t* __restrict__ next_frontier, uint32_t* __restrict__ tri_out, uint64_t* __restrict__ totals) {
  const uint32_t f = blockIdx.x * blockDim.x + threadIdx.x;
  if (f >= n_f) return;
  const uint32_t inner_off = (uint32_t)(scan[f] & 0xFFFFFFFFull), tri_off = (uint32_t)(scan[f] >> 32);
  emit_node(t, col[f], ep, level_base + f, level_base + n_f + inner_off, tri_before + tri_off, leaf_tri, nodes, next_frontier, inner_off, tri_out);
  if (f == n_f - 1u) *totals = scan[f] + word[f];
}

struct Tmp {   // device temporaries, freed on every way out
  std::vector<void*> ptrs;
  ~Tmp() { for (void* p : ptrs) cudaFree(p); }
  template <typename T> cudaError_t alloc(T** p, size_t n) { const cudaError_t e = cudaMalloc((void**)p, (n ? n : 1) * sizeof(T)); if (e == cudaSuccess) ptrs.push_back(*p); return e; }
};

}  // namespace

namespace crt {

#define BV(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { if (msg) *msg = #call; return e_; } } while (0)

cudaError_t build_bvh_device(const void* prims_raw, const uint32_t* tri_pos, uint32_t n, uint32_t bvh_tuning, float prim_cost_in,
                             cudaStream_t stream, DeviceBvh* out, std::string* msg) {
  const BvhOptions opt = bvh_options(bvh_tuning, prim_cost_in);
  const float prim_cost = opt.prim_cost;
  cudaEvent_t ev0, ev1;
  BV(cudaEventCreate(&ev0)); BV(cudaEventCreate(&ev1));
  struct EvGuard { cudaEvent_t a, b; ~EvGuard() { cudaEventDestroy(a); cudaEventDestroy(b); } } evg{ev0, ev1};
  BV(cudaEventRecord(ev0, stream));
  Tmp tmp;
  const uint32_t B = 256;
  auto grid = [&](uint32_t k) { return (k + B - 1) / B; };

  // ---- 1. boxes and bounds
  Box3* boxes_in; Bounds* d_bounds;
  BV(tmp.alloc(&boxes_in, n)); BV(tmp.alloc(&d_bounds, 1));
  bounds_init_kernel<<<1, 1, 0, stream>>>(d_bounds);
  tri_boxes_kernel<<<grid(n), B, 0, stream>>>((const uint8_t*)prims_raw, tri_pos, n, boxes_in, d_bounds);
  BV(cudaGetLastError());
  // ---- 2. Morton codes, sort
  uint64_t *keys, *keys_sorted; uint32_t *vals, *sorted;
  BV(tmp.alloc(&keys, n)); BV(tmp.alloc(&keys_sorted, n)); BV(tmp.alloc(&vals, n)); BV(tmp.alloc(&sorted, n));
  morton_kernel<<<grid(n), B, 0, stream>>>(boxes_in, n, d_bounds, keys, vals);
  BV(cudaGetLastError());
  size_t cub_bytes = 0, scan_bytes = 0;
  BV(cub::DeviceRadixSort::SortPairs(nullptr, cub_bytes, keys, keys_sorted, vals, sorted, (int)n, 0, 63, stream));
  uint64_t *word, *scan, *d_totals;
  BV(tmp.alloc(&word, n)); BV(tmp.alloc(&scan, n)); BV(tmp.alloc(&d_totals, 1));
  BV(cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, word, scan, (int)n, stream));
  uint8_t* cub_tmp;
  BV(tmp.alloc(&cub_tmp, std::max(cub_bytes, scan_bytes)));
  size_t cb = std::max(cub_bytes, scan_bytes);
  BV(cub::DeviceRadixSort::SortPairs(cub_tmp, cb, keys, keys_sorted, vals, sorted, (int)n, 0, 63, stream));
  // ---- 3. PLOC
  Tree t;
  t.n_leaves = n;
  const size_t n2 = 2 * (size_t)n;
  BV(tmp.alloc(&t.box, n2)); BV(tmp.alloc(&t.left, n2)); BV(tmp.alloc(&t.right, n2)); BV(tmp.alloc(&t.total, n2));
  BV(tmp.alloc(&t.cost, n2 * 7)); BV(tmp.alloc(&t.dec, n2 * 8));
  uint32_t *leaf_tri, *cl_a, *cl_b, *nn;
  BV(tmp.alloc(&leaf_tri, n)); BV(tmp.alloc(&cl_a, n)); BV(tmp.alloc(&cl_b, n)); BV(tmp.alloc(&nn, n));
  leaves_kernel<<<grid(n), B, 0, stream>>>(t, boxes_in, sorted, tri_pos, leaf_tri, cl_a, prim_cost);
  BV(cudaGetLastError());
  uint64_t* h_totals;
  BV(cudaMallocHost(&h_totals, 8));
  struct HostGuard { void* p; ~HostGuard() { cudaFreeHost(p); } } hg{h_totals};
  uint32_t n_cl = n, n_nodes2 = n, iters = 0;
  while (n_cl > 1) {
    ploc_nn_kernel<<<grid(n_cl), B, 0, stream>>>(t, cl_a, n_cl, nn);
    ploc_word_kernel<<<grid(n_cl), B, 0, stream>>>(nn, n_cl, word);
    cb = std::max(cub_bytes, scan_bytes);
    BV(cub::DeviceScan::ExclusiveSum(cub_tmp, cb, word, scan, (int)n_cl, stream));
    ploc_apply_kernel<<<grid(n_cl), B, 0, stream>>>(t, cl_a, nn, word, scan, n_cl, n_nodes2, prim_cost, cl_b, d_totals);
    BV(cudaGetLastError());
    BV(cudaMemcpyAsync(h_totals, d_totals, 8, cudaMemcpyDeviceToHost, stream));
    BV(cudaStreamSynchronize(stream));
    const uint32_t kept = (uint32_t)(*h_totals & 0xFFFFFFFFull), merged = (uint32_t)(*h_totals >> 32);
    if (merged == 0 || kept != n_cl - merged) { if (msg) *msg = "PLOC made no progress"; return cudaErrorUnknown; }
    n_cl = kept; n_nodes2 += merged;
    std::swap(cl_a, cl_b);
    iters++;
  }
  const uint32_t root = n_nodes2 - 1u;   // == 2n - 2 (n == 1: leaf 0)
  // ---- bounds and padding (crt_bvh.cpp has the reasoning)
  Bounds hb;
  BV(cudaMemcpyAsync(&hb, d_bounds, sizeof(hb), cudaMemcpyDeviceToHost, stream));
  BV(cudaStreamSynchronize(stream));
  float lo[3], hi[3], ext = 0.0f, maxabs;
  memcpy(&maxabs, &hb.maxabs, 4);
  for (int a = 0; a < 3; a++) { lo[a] = ord2f(hb.lo[a]); hi[a] = ord2f(hb.hi[a]); ext = std::max(ext, hi[a] - lo[a]); }
  const float pad = std::max(std::max(ext, maxabs), 1.0f) * (1.0f / 32768.0f);
  out->pad = pad;
  for (int a = 0; a < 3; a++) { out->lo[a] = lo[a] - pad; out->hi[a] = hi[a] + pad; }
  // ---- 4. emission, one level of the wide tree per step.  A wide node has at least two children unless it is the root of a
  // scene of <= 3 triangles, so there are fewer wide nodes than triangles.
  const size_t max_nodes = std::max<size_t>(n, 1);
  const size_t node_bytes = (max_nodes * 16 * CRT_NODE_STRIDE + 255) & ~(size_t)255, tri_bytes = (size_t)n * 48;
  uint32_t *frontier_a, *frontier_b, *tri_src; Collected* col; uint32_t* nodes_tmp;
  BV(tmp.alloc(&frontier_a, max_nodes)); BV(tmp.alloc(&frontier_b, max_nodes)); BV(tmp.alloc(&tri_src, n)); BV(tmp.alloc(&col, max_nodes));
  BV(tmp.alloc(&nodes_tmp, max_nodes * 4 * CRT_NODE_STRIDE));
  BV(cudaMemcpyAsync(frontier_a, &root, 4, cudaMemcpyHostToDevice, stream));
  EmitParams ep; ep.pad = pad; ep.exact_slots = opt.optimal_slots ? 1u : 0u;
  uint32_t n_f = 1, level_base = 0, tri_before = 0, depth = 0;
  while (n_f) {
    if ((size_t)level_base + n_f > max_nodes) { if (msg) *msg = "wide node count exceeds the bound"; return cudaErrorUnknown; }
    emit_collect_kernel<<<grid(n_f), B, 0, stream>>>(t, frontier_a, n_f, ep, col, word);
    cb = std::max(cub_bytes, scan_bytes);
    BV(cub::DeviceScan::ExclusiveSum(cub_tmp, cb, word, scan, (int)n_f, stream));
    emit_node_kernel<<<grid(n_f), B, 0, stream>>>(t, col, word, scan, n_f, ep, level_base, tri_before, leaf_tri, nodes_tmp, frontier_b, tri_src, d_totals);
    BV(cudaGetLastError());
    BV(cudaMemcpyAsync(h_totals, d_totals, 8, cudaMemcpyDeviceToHost, stream));
    BV(cudaStreamSynchronize(stream));
    level_base += n_f;
    n_f = (uint32_t)(*h_totals & 0xFFFFFFFFull);
    tri_before += (uint32_t)(*h_totals >> 32);
    std::swap(frontier_a, frontier_b);
    depth++;
  }
  if (tri_before != n) { if (msg) *msg = "emission lost triangles"; return cudaErrorUnknown; }
  const uint32_t n_nodes = level_base;
  // ---- the resident copy: nodes, then the triangles in leaf order
  const size_t nb = ((size_t)n_nodes * 16 * CRT_NODE_STRIDE + 255) & ~(size_t)255;
  uint8_t* res = nullptr;
  BV(cudaMalloc((void**)&res, nb + tri_bytes));
  cudaError_t e = cudaMemcpyAsync(res, nodes_tmp, (size_t)n_nodes * 16 * CRT_NODE_STRIDE, cudaMemcpyDeviceToDevice, stream);
  if (e == cudaSuccess) e = launch_repack_tri(prims_raw, tri_src, n, (float4*)(res + nb), stream);
  if (e == cudaSuccess) e = cudaEventRecord(ev1, stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(stream);
  if (e != cudaSuccess) { cudaFree(res); if (msg) *msg = "final copy"; return e; }
  float ms = 0.0f;
  cudaEventElapsedTime(&ms, ev0, ev1);
  out->nodes = (uint4*)res; out->tri = (float4*)(res + nb); out->bytes = nb + tri_bytes;
  out->n_nodes = n_nodes; out->max_depth = depth; out->build_ms = ms; out->ploc_iterations = iters;
  (void)node_bytes;
  return cudaSuccess;
}

}  // namespace crt
