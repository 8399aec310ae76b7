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
  float lo[3] = {0, 0, 0}, hi[3] = {0, 0, 0};   // padded bounds of all triangles
};

// prims: binding-8 bytes (80-byte records); tri_pos: positions of the category-2 records.
// Deterministic: the result does not depend on the number of threads.
void build_bvh(const uint8_t* prims, const std::vector<uint32_t>& tri_pos, int n_threads, BvhResult& out);

}  // namespace crt
