// crt_bvh.h -- host-side BVH construction for the scene-upload module.
#pragma once
#ifndef CRT_NODE_STRIDE
#define CRT_NODE_STRIDE 5   // 16-byte words between consecutive Wide8 nodes (5 = packed; 6 = 96-byte stride, tuning)
#endif
#include <stdint.h>
#include <vector>

namespace crt {

struct BvhResult {
  std::vector<uint32_t> nodes;     // 20 words (5 x 16 bytes) per 8-wide node, device layout (crt_device.cuh, "Wide8")
  std::vector<uint32_t> tri_src;   // triangle slot (leaf order) -> position in the primitives binding
  uint32_t n_nodes = 0;
  uint32_t max_depth = 0;          // depth of the wide tree
  float pad = 0.0f;
  double build_ms = 0.0;
  float lo[3] = {0, 0, 0}, hi[3] = {0, 0, 0};   // padded bounds of all triangles
};

// Builder settings (crt_params.bvh_tuning / bvh_prim_cost; tools/bvh_eval.py).  None of them can change a result.
struct BvhOptions {
  int bins = 32;                // SAH bins per axis (4 .. 256)
  bool optimal_slots = false;   // exact child-slot assignment (dynamic programme over slot sets) instead of the greedy one
  float prim_cost = 0.75f;      // SAH weight of a triangle test relative to a node visit in the Wide8 collapse
};
BvhOptions bvh_options(uint32_t bvh_tuning, float prim_cost);

// prims: binding-8 bytes (80-byte records); tri_pos: positions of the category-2 records.
// Deterministic: the result does not depend on the number of threads.
void build_bvh(const uint8_t* prims, const std::vector<uint32_t>& tri_pos, int n_threads, const BvhOptions& opt, BvhResult& out);

// The device build's algorithm run by loops on the host (validation of crt_bvh_ploc.h without a GPU; single-threaded).
void build_bvh_ploc_host(const uint8_t* prims, const std::vector<uint32_t>& tri_pos, const BvhOptions& opt, BvhResult& out);

}  // namespace crt
