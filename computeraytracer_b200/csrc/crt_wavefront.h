// crt_wavefront.h -- device buffers and launchers of the wavefront path (crt_wavefront.cu)
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

struct DevScene;

#define CRT_WF_MAX_ITERS 8192   // per-iteration counters are indexed by iteration and never reset inside a dispatch

// Queue lengths of one iteration.  Indexed by iteration, so nothing is reset inside a dispatch.  Every counter that takes
// atomics from all warps of a kernel sits in a 128-byte line of its own: with all of them in one 32-byte sector the atomics of
// the BVH queue, the shadow queue and the fetch cursor went through one L2 atomic unit (r02x A/B: wf_logic 143.1 -> 141.7 ms with
// 128-byte blocks, 140.1 with 1 KB blocks -- whose 33 MB memset per chunk is not worth it).
#ifndef WF_ITER_SPREAD
#define WF_ITER_SPREAD 32   // 32-bit words per block
#endif
struct WfIter {
  uint32_t n_bounce;   // slots whose extension ray hit a surface: wf_shade
  uint32_t n_finish;   // slots whose path is over (miss, or only an NEE answer was outstanding): wf_regen finishes them
  uint32_t n_regen;    // slots freed by wf_shade in this iteration: wf_regen starts new paths on them
  uint32_t pad_a[WF_ITER_SPREAD - 3];
  uint32_t n_bvh;      // extension rays queued for the BVH in this iteration
  uint32_t pad_b[WF_ITER_SPREAD - 1];
  uint32_t n_fetch;    // wf_traverse's fetch cursor into the BVH queue
  uint32_t pad_c[WF_ITER_SPREAD - 1];
  uint32_t n_shadow;   // NEE shadow rays queued for the BVH in this iteration (traversed after the extension rays)
  uint32_t pad_d[WF_ITER_SPREAD - 1];
};

#define CRT_WF_SLOW_RAYS 64
#define CRT_WF_REC_WORDS 10     // 16-byte words per slot record (160 bytes = 5 DRAM sectors); layout in crt_wavefront.cu
#define WF_REC_WORDS CRT_WF_REC_WORDS

// One chunk of frames rendered as a wavefront.
struct WfChunk {
  uint32_t n_units;     // pixels (8x4-tile order) this rank owns
  uint32_t n_slots;     // size of the path-slot pool working on the chunk
  uint32_t n_samples;   // samples of the chunk this rank renders (ordinals 0 .. n_samples-1)
  uint32_t s_first;     // frame number of ordinal 0
  uint32_t s_stride;    // frame number step between ordinals (world size under split-spp, else 1)
  uint32_t n_pixels;    // W * H
  float4* stage;        // [n_samples][n_units]: XYZ of each finished path, indexed by work item, summed in order by wf_accumulate
};

// Device-resident control block of the running chunk: written by wf_begin, advanced by wf_next.  The iteration kernels read
// the chunk from here, and -- when their `it` argument is CRT_WF_IT_DEVICE -- the iteration number too, so that the whole
// iteration loop can run as a CUDA-graph WHILE node without the host (crt_wavefront.cu, wf_graph_build).
struct WfCtl {
  uint32_t it;          // current iteration
  uint32_t error;       // 1: the per-iteration counters ran out (CRT_WF_MAX_ITERS)
  uint32_t iters_total; // iterations run since the context was created (diagnostics)
  uint32_t launches_total; // kernels the graph loop has launched
  WfChunk ck;
};
#define CRT_WF_IT_DEVICE 0xFFFFFFFFu

struct WfState {
  WfCtl* ctl;          // control block (device memory)
  float4* rec;         // [slots * WF_REC_WORDS] path state, one contiguous record per slot
  uint32_t* bounce[2]; // shade queue of this / the next iteration
  uint32_t* finish[2]; // finish queue of this / the next iteration
  uint32_t* regen;     // slots freed by wf_shade in the current iteration
  uint32_t* bvh_list;  // [slots] slots whose extension ray must traverse the BVH
  uint32_t* shadow_list; // [slots] slots whose NEE shadow ray must (kept apart so that warps hold rays of one kind)
  uint8_t* cls;        // [slots] class of the slot's next pass (live mode 2): 0 bounce, 1 finish, 2 retired
  WfIter* iter;        // [CRT_WF_MAX_ITERS + 2]
  uint32_t* work;      // next unassigned (sample, pixel) work item of the chunk
  uint32_t tables_in_smem;
  uint32_t slow_min;   // diagnostics (CRT_SLOW_RAY_NODES, stats builds): log rays that visit at least this many BVH nodes; 0 = off
  float4* slow;        // [1 + 3 * CRT_WF_SLOW_RAYS]: word 0 = count, then (origin, iteration), (direction, nodes), (exclude, anyhit, slot, triangles)
};


namespace crt {
size_t wf_tables_bytes(const DevScene& sc);
// resident blocks per SM of wf_logic / wf_shade / wf_regen / wf_traverse
cudaError_t wf_config(bool stats, size_t smem, int* logic_blocks_per_sm, int* shade_blocks_per_sm, int* regen_blocks_per_sm,
                      int* trav_blocks_per_sm);
// start of a chunk: the control block takes the chunk, iteration 0
cudaError_t launch_wf_begin(const WfState& w, const WfChunk& ck, cudaStream_t s);
// live_mode: 0 one live list, 1 routed lists (pushed by wf_logic and wf_traverse), 2 class bytes + wf_partition (crt_wavefront.cu)
// it: the iteration, or CRT_WF_IT_DEVICE (read it from the control block)
cudaError_t launch_wf_logic(const DevScene& sc, const WfState& w, bool stats, uint32_t live_mode, int grid, size_t smem, uint32_t it,
                            unsigned long long* counters, cudaStream_t s);
cudaError_t launch_wf_shade(const DevScene& sc, const WfState& w, const WfChunk& ck, bool stats, uint32_t live_mode, int grid, size_t smem, uint32_t it,
                            unsigned long long* counters, cudaStream_t s);
cudaError_t launch_wf_regen(const DevScene& sc, const WfState& w, const WfChunk& ck, bool stats, uint32_t live_mode, int grid, size_t smem, uint32_t it,
                            unsigned long long* counters, cudaStream_t s);
cudaError_t launch_wf_traverse(const DevScene& sc, const WfState& w, bool stats, bool route, bool fix_class, int grid, uint32_t it,
                               unsigned long long* counters, cudaStream_t s);
cudaError_t launch_wf_partition(const WfState& w, int grid, uint32_t it, cudaStream_t s);
// the end of a chunk: every slot still live at iteration `it` runs its path to the end in one thread (it: as above)
cudaError_t launch_wf_tail(const DevScene& sc, const WfState& w, bool stats, uint32_t live_mode, int grid, size_t smem, uint32_t it,
                           unsigned long long* counters, cudaStream_t s);
cudaError_t launch_wf_accumulate(const DevScene& sc, const WfChunk& ck, float4* acc, uint32_t* fb, uint32_t sample_after, cudaStream_t s);
// The iteration loop of one chunk as a CUDA graph: WHILE (live slots) { wf_logic; wf_traverse; wf_partition; wf_next }.  The graph
// depends on the scene (a kernel parameter), the buffers and the launch shapes, not on the chunk: build once, launch per chunk.
cudaError_t wf_graph_build(const DevScene& sc, const WfState& w, bool stats, uint32_t logic_mode, uint32_t live_mode, int grid_l, int grid_t,
                           int grid_p, size_t smem, unsigned long long* counters, uint32_t kernels_per_iter, uint32_t tail_max, cudaGraph_t* graph, cudaGraphExec_t* exec);
}  // namespace crt
