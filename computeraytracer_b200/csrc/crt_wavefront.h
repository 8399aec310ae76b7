// crt_wavefront.h -- device buffers and launchers of the wavefront path (crt_wavefront.cu)
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

struct DevScene;

#define CRT_WF_MAX_ITERS 8192   // per-iteration counters are indexed by iteration and never reset inside a dispatch

// Path state, one entry per slot (= one owned pixel), structure of arrays of 16-byte words.
struct WfState {
  float4* S0;          // ray origin, last_bounce_pdf
  float4* S1;          // ray direction, etaScale
  float4* S2;          // accumulated radiance (4 wavelengths)
  float4* S3;          // beta
  uint4* S4;           // RNG state
  uint4* S5;           // exclude, depth | flags << 16, first wavelength, current sample ordinal
  float4* S6;          // shadow phase: surface normal, reflectance | light << 12
  float4* H;           // hit of the current ray: t, primitive index, reference, query flags
  uint32_t* list[2];   // live slots of this / the next iteration
  uint32_t* bvh_list;  // slots whose ray must traverse the BVH this iteration
  uint32_t* n_active;  // [CRT_WF_MAX_ITERS + 4]; entry CRT_WF_MAX_ITERS + 2 is the chunk's work counter
  uint32_t* n_bvh;     // [CRT_WF_MAX_ITERS]
  uint32_t* n_fetch;   // [CRT_WF_MAX_ITERS]
  uint32_t tables_in_smem;
};

// One chunk of frames rendered as a wavefront.
struct WfChunk {
  uint32_t n_units;     // pixels (8x4-tile order) this rank owns
  uint32_t n_slots;     // size of the path-slot pool working on the chunk
  uint32_t n_samples;   // samples of the chunk this rank renders (ordinals 0 .. n_samples-1)
  uint32_t s_first;     // frame number of ordinal 0
  uint32_t s_stride;    // frame number step between ordinals (world size under split-spp, else 1)
  uint32_t n_pixels;    // W * H: stride of the staging buffer
  float4* stage;        // [n_samples][n_pixels]: XYZ of each finished path, summed in order by wf_accumulate
};

namespace crt {
size_t wf_tables_bytes(const DevScene& sc);
cudaError_t wf_config(bool stats, size_t smem, int* logic_blocks_per_sm, int* trav_blocks_per_sm);
cudaError_t launch_wf_logic(const DevScene& sc, const WfState& w, const WfChunk& ck, bool stats, int grid, size_t smem, uint32_t it,
                            unsigned long long* counters, cudaStream_t s);
cudaError_t launch_wf_accumulate(const DevScene& sc, const WfChunk& ck, float4* acc, uint32_t* fb, uint32_t sample_after, cudaStream_t s);
cudaError_t launch_wf_traverse(const DevScene& sc, const WfState& w, bool stats, int grid, uint32_t it, unsigned long long* counters,
                               cudaStream_t s);
}  // namespace crt
