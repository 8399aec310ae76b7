// crt_bvh_device.h -- BVH construction on the GPU (scene-upload module; crt_bvh_device.cu, steps in crt_bvh_ploc.h)
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <string>

namespace crt {

struct DeviceBvh {
  uint4* nodes = nullptr;     // one allocation: the Wide8 nodes, then (256-byte aligned) the leaf-ordered triangles
  float4* tri = nullptr;      // points into the same allocation
  size_t bytes = 0;
  uint32_t n_nodes = 0, max_depth = 0;
  float lo[3] = {0, 0, 0}, hi[3] = {0, 0, 0}, pad = 0.0f;   // padded bounds of all triangles
  double build_ms = 0.0;      // device time of the build (CUDA events), triangle repack included
  uint32_t ploc_iterations = 0;
};

// prims_raw: binding 8 on the device (80-byte records); tri_pos: positions of its n_tri category-2 records (device).
// Synchronous.  On success the caller owns out->nodes (cudaFree).  Deterministic: scans, no atomics in the layout.
cudaError_t build_bvh_device(const void* prims_raw, const uint32_t* tri_pos, uint32_t n_tri, uint32_t bvh_tuning, float prim_cost,
                             cudaStream_t stream, DeviceBvh* out, std::string* msg);

}  // namespace crt
