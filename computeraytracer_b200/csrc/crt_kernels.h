// crt_kernels.h -- host-callable launchers of crt_kernels.cu
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

struct DevScene;
struct LinPrim;
struct LightRec;

#define CRT_MAX_GROUP 16
struct CombinePeers {        // the members' accumulators (binding 1) as seen from member 0 (peer pointers over NVLink)
  const float4* acc[CRT_MAX_GROUP];
  uint32_t n;
  uint32_t tiles;            // 1: CRT_SHARD_TILES (take each pixel from its owner), 0: CRT_SHARD_SPLIT_SPP (sum in rank order)
};

namespace crt {
cudaError_t launch_repack_lin(const void* prims, const uint32_t* lin_src, uint32_t n, LinPrim* out, cudaStream_t s);
cudaError_t launch_repack_tri(const void* prims, const uint32_t* tri_src, uint32_t n, float4* out, cudaStream_t s);
cudaError_t launch_repack_lights(const void* lights, uint32_t n, LightRec* out, cudaStream_t s);
cudaError_t launch_tea_tab(uint32_t W, uint32_t H, uint32_t* out, cudaStream_t s);
cudaError_t launch_camera_setup(const float* cam, float vph, float* out12, cudaStream_t s);
size_t megakernel_tables_bytes(const DevScene& sc);
cudaError_t megakernel_config(bool stats, size_t smem, int* blocks_per_sm);
cudaError_t launch_megakernel(const DevScene& sc, bool stats, int grid, size_t smem, bool tables_in_smem, float4* acc, uint32_t* fb,
                              uint32_t sample_start, uint32_t n_frames, uint32_t n_units, uint32_t* work_counter,
                              unsigned long long* counters, cudaStream_t s);
cudaError_t launch_trace_rays(const DevScene& sc, const float4* rays, const uint32_t* exclude, uint32_t n, uint32_t* hit_idx, float* t,
                              unsigned long long* counters, cudaStream_t s);
cudaError_t launch_resolve(const float4* acc, uint32_t* fb, uint32_t n, uint32_t sample, float exposure, cudaStream_t s);
cudaError_t launch_pack_tiles(const float4* acc, float4* out, uint32_t n_units, uint32_t rank, uint32_t world, uint32_t tiles_x, uint32_t W,
                              uint32_t H, cudaStream_t s);
cudaError_t launch_unpack_tiles(const float4* in, float4* acc, uint32_t n_units, uint32_t rank, uint32_t world, uint32_t tiles_x, uint32_t W,
                                uint32_t H, cudaStream_t s);
cudaError_t launch_combine_peers(const CombinePeers& peers, float4* comb, uint32_t* fb, uint32_t W, uint32_t H, uint32_t tiles_x,
                                 uint32_t sample, float exposure, cudaStream_t s);
cudaError_t launch_fp32_peak(float* out, int grid, uint32_t iters, cudaStream_t s);   // 64 FFMA per thread and iteration
cudaError_t launch_detmath(const float* x, uint32_t n, float* sn, float* cs, float* ex, float* pw, cudaStream_t s);
}  // namespace crt
