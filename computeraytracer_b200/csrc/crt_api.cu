// crt_api.cu -- the C ABI of include/crt.h: context, bindings, scene-upload module, dispatch.
//
// Replaces the host half of the reference's WebGPU contract (src/main.js): buffer creation and
// upload (MJ:139-393), bind group (MJ:465-529), the per-frame dispatch pair (MJ:597-611).
// There is no CPU fallback: without a usable sm_100 device crt_create fails.
#include <cuda_runtime.h>

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <string>
#include <utility>
#include <thread>
#include <vector>

#include "../../include/crt.h"
#include "crt_bvh.h"
#include "crt_device.cuh"
#include "crt_kernels.h"
#include "crt_wavefront.h"

#define CRT_VERSION_STRING "libcrt 0.1 (sm_100a)"

struct crt_ctx {
  int device = 0;
  int sm_count = 0;
  size_t total_mem = 0;
  cudaStream_t stream = nullptr;
  cudaStream_t own_stream = nullptr;
  uint64_t launches = 0;   // hot-path kernel launches since the last crt_reset_counters
  cudaEvent_t ev_start = nullptr, ev_stop = nullptr;
  bool timed = false;
  std::string err;

  // host copies of the reference-layout bindings
  std::vector<uint8_t> b_camera, b_cie, b_spectra, b_lights, b_prims;
  bool have[9] = {false, false, false, false, false, false, false, false, false};
  bool scene_dirty = true, tables_dirty = true, camera_dirty = true;

  crt_params params;
  uint32_t rank = 0, world = 1, shard_mode = CRT_SHARD_NONE;
  uint32_t sample = 0;   // binding 2
  uint32_t W = 0, H = 0;

  // device state
  float4* d_acc = nullptr;       // binding 1
  uint32_t* d_fb = nullptr;      // binding 0
  uint32_t* d_tea = nullptr;     // per-pixel tea(x, y*100) of the RNG seed (CS:98)
  size_t acc_pixels = 0;
  float* d_spectra = nullptr; float* d_cie = nullptr;
  void* d_prims_raw = nullptr;   // binding 8 as uploaded
  LinPrim* d_lin = nullptr; LightRec* d_lights = nullptr;
  float4* d_tri = nullptr; uint4* d_nodes = nullptr;   // one allocation: nodes, then triangles (d_tri points into it)
  size_t scene_bytes = 0;
  uint32_t* d_work = nullptr;
  unsigned long long* d_counters = nullptr;
  DevScene scene;
  uint64_t stats[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  std::vector<uint8_t> pending_acc;  // binding 1 set before the size is known

  // wavefront path (crt_wavefront.cu)
  WfState wf;
  void* wf_mem = nullptr;            // one allocation behind all WfState arrays
  uint32_t wf_slots = 0;
  uint32_t* wf_host_count = nullptr; // pinned: live-slot count read back at check points
  void* wf_stage = nullptr;          // per-sample XYZ staging of the current chunk
  size_t wf_stage_bytes = 0;
  uint64_t wf_last_work = 0;
  uint32_t wf_last_iters = 0;        // iterations the previous chunk needed (predicts the next check point)
  uint64_t wf_iterations = 0;

  // CRT_FLAG_KERNEL_TIMES: events recorded between consecutive hot-path launches
  std::vector<cudaEvent_t> ev_free;
  std::vector<std::pair<cudaEvent_t, int>> ev_seq;   // (event, class of the kernel launched right after it; -1 = end of a run)
  double k_ms[CRT_NUM_KERNEL_CLASSES] = {};
  uint64_t k_launches[CRT_NUM_KERNEL_CLASSES] = {};
};

namespace {

int fail(crt_ctx* c, int code, const char* fmt, ...) {
  if (c) {
    char buf[512];
    va_list ap; va_start(ap, fmt); vsnprintf(buf, sizeof(buf), fmt, ap); va_end(ap);
    c->err = buf;
  }
  return code;
}

#define CU(call)                                                                                         \
  do {                                                                                                   \
    cudaError_t e_ = (call);                                                                             \
    if (e_ != cudaSuccess)                                                                               \
      return fail(ctx, e_ == cudaErrorMemoryAllocation ? CRT_ERR_OOM : CRT_ERR_CUDA, "%s: %s (%s:%d)", #call, \
                  cudaGetErrorString(e_), __FILE__, __LINE__);                                           \
  } while (0)

// CRT_FLAG_KERNEL_TIMES: an event on the stream before the next launch of class `cls` (-1 closes a run)
cudaError_t mark(crt_ctx* ctx, int cls) {
  if (!(ctx->params.flags & CRT_FLAG_KERNEL_TIMES)) return cudaSuccess;
  if (cls < 0 && (ctx->ev_seq.empty() || ctx->ev_seq.back().second < 0)) return cudaSuccess;
  cudaEvent_t e;
  if (!ctx->ev_free.empty()) { e = ctx->ev_free.back(); ctx->ev_free.pop_back(); }
  else { cudaError_t rc = cudaEventCreate(&e); if (rc != cudaSuccess) return rc; }
  ctx->ev_seq.emplace_back(e, cls);
  return cudaEventRecord(e, ctx->stream);
}

// fold the recorded events into k_ms / k_launches; synchronises
cudaError_t drain_marks(crt_ctx* ctx) {
  if (ctx->ev_seq.empty()) return cudaSuccess;
  cudaError_t rc = cudaStreamSynchronize(ctx->stream);
  if (rc != cudaSuccess) return rc;
  for (size_t i = 0; i + 1 < ctx->ev_seq.size(); i++) {
    const int cls = ctx->ev_seq[i].second;
    if (cls < 0) continue;
    float ms = 0.0f;
    if ((rc = cudaEventElapsedTime(&ms, ctx->ev_seq[i].first, ctx->ev_seq[i + 1].first)) != cudaSuccess) return rc;
    ctx->k_ms[cls] += ms; ctx->k_launches[cls]++;
  }
  for (auto& e : ctx->ev_seq) ctx->ev_free.push_back(e.first);
  ctx->ev_seq.clear();
  return cudaSuccess;
}

template <typename T>
void dfree(T*& p) { if (p) { cudaFree(p); p = nullptr; } }

int ensure_frame_buffers(crt_ctx* ctx) {
  const size_t n = (size_t)ctx->W * ctx->H;
  if (n == ctx->acc_pixels && ctx->d_acc) return CRT_OK;
  dfree(ctx->d_acc); dfree(ctx->d_fb); dfree(ctx->d_tea);
  ctx->acc_pixels = 0;
  if (n == 0) return fail(ctx, CRT_ERR_SIZE, "camera viewport is %ux%u", ctx->W, ctx->H);
  CU(cudaMalloc(&ctx->d_acc, n * sizeof(float4)));
  CU(cudaMalloc(&ctx->d_fb, n * sizeof(uint32_t)));
  CU(cudaMalloc(&ctx->d_tea, n * sizeof(uint32_t)));
  CU(cudaMemsetAsync(ctx->d_acc, 0, n * sizeof(float4), ctx->stream));   // WebGPU zero-initialises, MJ:298-302
  CU(cudaMemsetAsync(ctx->d_fb, 0, n * sizeof(uint32_t), ctx->stream));
  ctx->acc_pixels = n;
  if (ctx->pending_acc.size() == n * 16) {
    CU(cudaMemcpyAsync(ctx->d_acc, ctx->pending_acc.data(), n * 16, cudaMemcpyHostToDevice, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
  }
  ctx->pending_acc.clear();
  return CRT_OK;
}

// camera binding -> W, H and the hoisted camera frame (CS:85, CS:470-487)
int upload_camera(crt_ctx* ctx) {
  const float* cam = (const float*)ctx->b_camera.data();
  // u32(f32) of CS:85: truncate, saturate
  auto f2u = [](float f) -> uint32_t { if (!(f > 0.0f)) return 0u; if (f >= 4294967296.0f) return 0xFFFFFFFFu; return (uint32_t)f; };
  const uint32_t W = f2u(cam[11]), H = f2u(cam[12]);
  if (W == 0 || H == 0 || (uint64_t)W * H > (1ull << 28)) return fail(ctx, CRT_ERR_SIZE, "camera viewport %ux%u out of range", W, H);
  ctx->W = W; ctx->H = H;
  int rc = ensure_frame_buffers(ctx);
  if (rc) return rc;
  float* d_cam = nullptr; float* d_out = nullptr;
  CU(cudaMalloc(&d_cam, 64));
  CU(cudaMalloc(&d_out, 48));
  CU(cudaMemcpyAsync(d_cam, cam, 64, cudaMemcpyHostToDevice, ctx->stream));
  const float vph = 2.0f * (float)tan((double)(cam[13] / 2.0f));   // tan pin: double libm, rounded once (CS:479)
  CU(crt::launch_camera_setup(d_cam, vph, d_out, ctx->stream));
  float out[12];
  CU(cudaMemcpyAsync(out, d_out, 48, cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  cudaFree(d_cam); cudaFree(d_out);
  DevScene& s = ctx->scene;
  memcpy(s.eye, out, 12); memcpy(s.hor, out + 3, 12); memcpy(s.ver, out + 6, 12); memcpy(s.llc, out + 9, 12);
  s.W = W; s.H = H;
  s.tiles_x = (W + 7) / 8; s.tiles_y = (H + 3) / 4;
  CU(crt::launch_tea_tab(W, H, ctx->d_tea, ctx->stream));   // the table is indexed x + y * W, so every camera change rebuilds it
  s.tea_tab = ctx->d_tea;
  ctx->camera_dirty = false;
  return CRT_OK;
}

int upload_tables(crt_ctx* ctx) {
  if (ctx->b_spectra.size() % (CRT_LAMBDA_RANGE * 4) != 0 || ctx->b_spectra.empty())
    return fail(ctx, CRT_ERR_SIZE, "binding 6 must be a multiple of %u bytes", CRT_LAMBDA_RANGE * 4);
  const uint32_t n_spectra = (uint32_t)(ctx->b_spectra.size() / (CRT_LAMBDA_RANGE * 4));
  if (n_spectra > 4096) return fail(ctx, CRT_ERR_SIZE, "at most 4096 spectra are supported (got %u)", n_spectra);
  dfree(ctx->d_spectra); dfree(ctx->d_cie);
  CU(cudaMalloc(&ctx->d_spectra, ctx->b_spectra.size()));
  CU(cudaMalloc(&ctx->d_cie, ctx->b_cie.size()));
  CU(cudaMemcpyAsync(ctx->d_spectra, ctx->b_spectra.data(), ctx->b_spectra.size(), cudaMemcpyHostToDevice, ctx->stream));
  CU(cudaMemcpyAsync(ctx->d_cie, ctx->b_cie.data(), ctx->b_cie.size(), cudaMemcpyHostToDevice, ctx->stream));
  ctx->scene.spectra = ctx->d_spectra; ctx->scene.cie = ctx->d_cie; ctx->scene.n_spectra = n_spectra;
  ctx->tables_dirty = false;
  return CRT_OK;
}

// The scene-upload module: binding 7/8 bytes -> linear list (patches, spheres), triangle array in BVH
// leaf order, BVH nodes, light records.  All repacking arithmetic runs on the device so it follows the
// same numerics contract as the kernels.
int upload_scene(crt_ctx* ctx) {
  const auto t0 = std::chrono::steady_clock::now();
  const size_t n_prims = ctx->b_prims.size() / 80;
  const size_t n_lights = ctx->b_lights.size() / 80;
  if (n_lights == 0) return fail(ctx, CRT_ERR_SIZE, "binding 7 holds no light (the shader samples one per diffuse bounce, CS:383)");
  if (n_lights > 4096) return fail(ctx, CRT_ERR_SIZE, "at most 4096 lights are supported");
  if (n_prims >= (1u << 28)) return fail(ctx, CRT_ERR_SIZE, "too many primitives");
  const uint8_t* prims = ctx->b_prims.data();
  std::vector<uint32_t> lin_pos, tri_pos;
  for (size_t i = 0; i < n_prims; i++) {
    uint32_t cat; memcpy(&cat, prims + i * 80, 4);
    if (cat == 2u) tri_pos.push_back((uint32_t)i);
    else if (cat <= 1u) lin_pos.push_back((uint32_t)i);
    // other categories match no branch of ray_intersection (CS:525, CS:584) and can never be hit
  }
  crt::BvhResult bvh;
  const int threads = (int)std::max(1u, std::thread::hardware_concurrency());
  crt::build_bvh(prims, tri_pos, threads, bvh);
  if (bvh.max_depth + 2 > CRT_BVH_STACK) return fail(ctx, CRT_ERR_SIZE, "BVH depth %u exceeds the traversal stack", bvh.max_depth);

  dfree(ctx->d_prims_raw); dfree(ctx->d_lin); dfree(ctx->d_lights); ctx->d_tri = nullptr; dfree(ctx->d_nodes); ctx->scene_bytes = 0;
  DevScene& s = ctx->scene;
  s.n_lin = (uint32_t)lin_pos.size(); s.n_tri = (uint32_t)tri_pos.size(); s.n_nodes = bvh.n_nodes; s.n_lights = (uint32_t)n_lights;
  s.lin_g = nullptr; s.tri = nullptr; s.nodes = nullptr; s.lights_g = nullptr;
  for (int a = 0; a < 3; a++) { s.bvh_lo[a] = bvh.lo[a]; s.bvh_hi[a] = bvh.hi[a]; }

  void* d_lights_raw = nullptr; uint32_t* d_src = nullptr;
  if (n_prims) {
    CU(cudaMalloc(&ctx->d_prims_raw, n_prims * 80));
    CU(cudaMemcpyAsync(ctx->d_prims_raw, prims, n_prims * 80, cudaMemcpyHostToDevice, ctx->stream));
  }
  const size_t n_src = std::max(lin_pos.size(), bvh.tri_src.size());
  if (n_src) CU(cudaMalloc(&d_src, n_src * 4));
  if (s.n_lin) {
    CU(cudaMalloc(&ctx->d_lin, (size_t)s.n_lin * sizeof(LinPrim)));
    CU(cudaMemcpyAsync(d_src, lin_pos.data(), lin_pos.size() * 4, cudaMemcpyHostToDevice, ctx->stream));
    CU(crt::launch_repack_lin(ctx->d_prims_raw, d_src, s.n_lin, ctx->d_lin, ctx->stream));
    const uint32_t nc = s.n_lin < CRT_LIN_PARAM_MAX ? s.n_lin : CRT_LIN_PARAM_MAX;
    CU(cudaMemcpyAsync(s.lin_c, ctx->d_lin, nc * sizeof(LinPrim), cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    s.lin_g = ctx->d_lin;
  }
  if (s.n_tri) {
    // one allocation, nodes first
    const size_t node_bytes = (bvh.nodes.size() * 4 + 255) & ~(size_t)255, tri_bytes = (size_t)s.n_tri * 48;
    CU(cudaMalloc(&ctx->d_nodes, node_bytes + tri_bytes));
    ctx->d_tri = (float4*)((uint8_t*)ctx->d_nodes + node_bytes);
    ctx->scene_bytes = node_bytes + tri_bytes;
    CU(cudaMemcpyAsync(d_src, bvh.tri_src.data(), bvh.tri_src.size() * 4, cudaMemcpyHostToDevice, ctx->stream));
    CU(crt::launch_repack_tri(ctx->d_prims_raw, d_src, s.n_tri, ctx->d_tri, ctx->stream));
    CU(cudaMemcpyAsync(ctx->d_nodes, bvh.nodes.data(), bvh.nodes.size() * 4, cudaMemcpyHostToDevice, ctx->stream));
    s.tri = ctx->d_tri; s.nodes = ctx->d_nodes;
  }
  CU(cudaMalloc(&d_lights_raw, n_lights * 80));
  CU(cudaMalloc(&ctx->d_lights, n_lights * sizeof(LightRec)));
  CU(cudaMemcpyAsync(d_lights_raw, ctx->b_lights.data(), n_lights * 80, cudaMemcpyHostToDevice, ctx->stream));
  CU(crt::launch_repack_lights(d_lights_raw, (uint32_t)n_lights, ctx->d_lights, ctx->stream));
  const uint32_t nl = n_lights < CRT_LIGHT_PARAM_MAX ? (uint32_t)n_lights : CRT_LIGHT_PARAM_MAX;
  CU(cudaMemcpyAsync(s.lights_c, ctx->d_lights, nl * sizeof(LightRec), cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  s.lights_g = ctx->d_lights;
  cudaFree(d_lights_raw);
  if (d_src) cudaFree(d_src);
  dfree(ctx->d_prims_raw);   // the SoA layouts are the resident copy

  // does any light's primitive (data4.w, CS:386) live in the BVH?  then shadow rays need closest-hit
  s.shadow_closest = 0;
  for (size_t l = 0; l < n_lights; l++) {
    uint32_t want; memcpy(&want, ctx->b_lights.data() + l * 80 + 76, 4);
    for (uint32_t p : tri_pos) {
      uint32_t idx; memcpy(&idx, prims + (size_t)p * 80 + 76, 4);
      if (idx == want) { s.shadow_closest = 1; break; }
    }
    if (s.shadow_closest) break;
  }
  const auto t1 = std::chrono::steady_clock::now();
  ctx->stats[0] = n_prims; ctx->stats[1] = s.n_lin; ctx->stats[2] = s.n_tri; ctx->stats[3] = s.n_nodes;
  ctx->stats[4] = (uint64_t)s.n_nodes * 16 * CRT_NODE_STRIDE; ctx->stats[5] = (uint64_t)s.n_tri * 48;
  ctx->stats[6] = (uint64_t)std::chrono::duration_cast<std::chrono::microseconds>(t1 - t0).count();
  ctx->stats[7] = bvh.max_depth;
  ctx->scene_dirty = false;
  return CRT_OK;
}

int prepare(crt_ctx* ctx) {
  for (int b = 4; b <= 8; b++)
    if (!ctx->have[b]) return fail(ctx, CRT_ERR_STATE, "binding %d has not been set", b);
  int rc;
  if (ctx->camera_dirty && (rc = upload_camera(ctx))) return rc;
  if (ctx->tables_dirty && (rc = upload_tables(ctx))) return rc;
  if (ctx->scene_dirty && (rc = upload_scene(ctx))) return rc;
  DevScene& s = ctx->scene;
  s.max_depth = ctx->params.max_depth; s.grid_size = ctx->params.grid_size ? ctx->params.grid_size : 1;
  s.t_min = ctx->params.t_min; s.exposure = ctx->params.exposure;
  s.rank = ctx->rank; s.world = ctx->world; s.shard_mode = ctx->shard_mode;
  return CRT_OK;
}

#ifndef CRT_WF_LIVE_BVH
#define CRT_WF_LIVE_BVH 2      // live mode for scenes with triangles (mesh1m: 351.5 -> 341.4 ms per 256-spp step; 1 is 40 % slower:
#endif                        // lists filled by wf_traverse lose the slot order and with it the sequential record stream)
#ifndef CRT_WF_LIVE_LINEAR
#define CRT_WF_LIVE_LINEAR 1   // live mode for linear-list scenes (reference default scene 1080p x 64 frames: 64.8 -> 55.0 ms)
#endif
constexpr size_t kWfCounterBytes = (size_t)(CRT_WF_MAX_ITERS + 2) * sizeof(WfIter) + 64;   // iteration records + the work counter

int ensure_wavefront(crt_ctx* ctx, uint32_t n_slots) {
  if (ctx->wf_mem && ctx->wf_slots >= n_slots) return CRT_OK;
  if (ctx->wf_mem) { cudaFree(ctx->wf_mem); ctx->wf_mem = nullptr; }
  // one record per slot; shade and finish queues of two iterations, the regen queue, the BVH queue (2 entries per slot)
  const size_t slow_bytes = (size_t)(1 + 3 * CRT_WF_SLOW_RAYS) * 16;
  const size_t cls_bytes = ((size_t)n_slots + 255) & ~(size_t)255;
  const size_t total = (size_t)(CRT_WF_REC_WORDS * 16 + 7 * 4) * n_slots + kWfCounterBytes + 256 + slow_bytes + cls_bytes;
  CU(cudaMalloc(&ctx->wf_mem, total));
  uint8_t* p = (uint8_t*)ctx->wf_mem;
  WfState& w = ctx->wf;
  w.rec = (float4*)p; p += (size_t)CRT_WF_REC_WORDS * 16 * n_slots;
  w.bounce[0] = (uint32_t*)p; p += 4 * (size_t)n_slots;
  w.bounce[1] = (uint32_t*)p; p += 4 * (size_t)n_slots;
  w.finish[0] = (uint32_t*)p; p += 4 * (size_t)n_slots;
  w.finish[1] = (uint32_t*)p; p += 4 * (size_t)n_slots;
  w.regen = (uint32_t*)p; p += 4 * (size_t)n_slots;
  w.bvh_list = (uint32_t*)p; p += 4 * (size_t)n_slots;
  w.shadow_list = (uint32_t*)p; p += 4 * (size_t)n_slots;
  w.iter = (WfIter*)p; p += (size_t)(CRT_WF_MAX_ITERS + 2) * sizeof(WfIter);
  w.work = (uint32_t*)p; p += 256;
  w.slow = (float4*)p; p += slow_bytes;
  w.cls = (uint8_t*)p;
  CU(cudaMemset(w.slow, 0, slow_bytes));
  {
    const char* e = getenv("CRT_SLOW_RAY_NODES");   // diagnostics: log rays that visit at least this many nodes (stats mode)
    w.slow_min = e ? (uint32_t)strtoul(e, nullptr, 10) : 0u;
  }
  ctx->wf_slots = n_slots;
  if (!ctx->wf_host_count) CU(cudaMallocHost(&ctx->wf_host_count, 8));
  return CRT_OK;
}

// n frames as a wavefront, in chunks.  Iteration 0 starts a path on every slot (wf_regen); every later iteration is
// wf_shade (one bounce for the slots with a surface hit), wf_regen (finish the slots whose path is over, start new paths on
// all free slots) and, when the scene has triangles, wf_traverse (the BVH queue both of them filled).  A chunk's loop ends
// when a check point reads back empty shade and finish queues; wf_accumulate then adds the chunk's samples to binding 1 in
// frame order and rebuilds binding 0.
int dispatch_wavefront(crt_ctx* ctx, uint32_t n, uint32_t n_units, bool stats) {
  const DevScene& s = ctx->scene;
  size_t smem = crt::wf_tables_bytes(s);
  ctx->wf.tables_in_smem = smem <= 48 * 1024 ? 1u : 0u;
  if (!ctx->wf.tables_in_smem) smem = 0;
  int bl = 0, bs = 0, br = 0, bt = 0;
  CU(crt::wf_config(stats, smem, &bl, &bs, &br, &bt));
  if (bl < 1 || bs < 1 || br < 1 || bt < 1) return fail(ctx, CRT_ERR_CUDA, "wavefront kernels do not fit on an SM");
  const int grid_l = ctx->sm_count * bl, grid_s = ctx->sm_count * bs, grid_r = ctx->sm_count * br, grid_t = ctx->sm_count * bt;
  const bool split = (ctx->params.flags & CRT_FLAG_WF_SPLIT) != 0;   // three-kernel form (wf_shade / wf_regen) instead of wf_logic
  // a path lives for at most max_depth+1 extension rays plus one pass for a trailing NEE answer
  const uint64_t worst = (uint64_t)ctx->params.max_depth + 3ull;
  if (worst + 2 > CRT_WF_MAX_ITERS) return fail(ctx, CRT_ERR_ARG, "max_depth %u is too deep for the wavefront kernel; use CRT_KERNEL_MEGA", ctx->params.max_depth);
  // frames per chunk: each chunk ends in a tail of iterations with few live paths (~4 ms on mesh1m), so chunks are as long
  // as the staging buffer allows (measured: 64 -> 128 -> 256 frames per chunk = 2568 -> 2638 -> 2669 Mrays/s)
  uint32_t chunk = ctx->params.frames_per_launch ? ctx->params.frames_per_launch : 256u;
  // path slots in flight, measured per 64-spp step of mesh1m: 8M -> 114 ms, 16M -> 107 ms, 32M -> 104 ms; 64M is another 1-2 %
  const uint32_t pool = ctx->params.wf_slots ? ctx->params.wf_slots : (64u << 20);
  const size_t n_pixels = ctx->acc_pixels;
  // guards: staging <= a quarter of the device's memory (45 GB of a B200's 180), work items of a chunk < 2^31
  const uint64_t stage_cap = std::max<uint64_t>(ctx->total_mem / 4, 1ull << 30);
  while (chunk > 1 && ((uint64_t)n_pixels * chunk * 16ull > stage_cap || (uint64_t)n_units * chunk >= (1ull << 31))) chunk--;
  const bool split_spp = ctx->shard_mode == CRT_SHARD_SPLIT_SPP && ctx->world > 1;
  // under split-spp a rank renders every world-th frame: size the chunk by the frames this rank renders, not by the
  // frame numbers it spans (8 ranks with 64-frame chunks ran 8 samples per chunk: the pool half empty, 8x the tails)
  if (split_spp) chunk = (uint32_t)std::min<uint64_t>((uint64_t)chunk * ctx->world, 1u << 20);
  // how wf_logic's live slots reach the next iteration (crt_wavefront.cu, wf_emit): 0 one list, 1 routed lists, 2 class
  // bytes + wf_partition.  In 1 and 2 wf_logic runs over the finish list and then the bounce list, so the lanes of a warp
  // are in one phase.
  uint32_t live_mode = s.n_tri ? CRT_WF_LIVE_BVH : CRT_WF_LIVE_LINEAR;
  if (const char* e = getenv("CRT_WF_ROUTED")) live_mode = (uint32_t)atoi(e) % 3u;   // tuning override
  if (split && live_mode == 0) live_mode = 1;   // the split kernels always work over routed lists: filled as rays finish (1) or by wf_partition (2)
  // routed forms: paths that end in a bounce warp are finished by the next iteration's finish warps (mesh1m: 339.8 ->
  // 334.5 ms per step; the reference default scene: 56.0 -> 58.0 ms, so linear-list scenes finish in place)
  uint32_t defer = s.n_tri ? 1u : 0u;
  if (const char* e = getenv("CRT_WF_DEFER")) defer = atoi(e) != 0;   // tuning override
  const uint32_t logic_mode = live_mode | ((live_mode && defer) ? 0x100u : 0u);
  uint32_t left = n;
  while (left) {
    const uint32_t k = left < chunk ? left : chunk;
    WfChunk ck;
    ck.n_units = n_units; ck.n_pixels = (uint32_t)n_pixels;
    if (split_spp) {
      uint32_t first = ctx->sample + 1u;
      first += (ctx->rank + ctx->world - ((first - 1u) % ctx->world)) % ctx->world;
      const uint32_t end = ctx->sample + k;
      ck.s_first = first; ck.s_stride = ctx->world;
      ck.n_samples = first <= end ? (end - first) / ctx->world + 1u : 0u;
    } else {
      ck.s_first = ctx->sample + 1u; ck.s_stride = 1u; ck.n_samples = k;
    }
    const uint64_t total_work = (uint64_t)n_units * ck.n_samples;
    ck.n_slots = (uint32_t)std::min<uint64_t>(std::max<uint64_t>(total_work, 32), pool);
    // worst case: every slot runs ceil(total/slots) paths of `worst` iterations one after the other
    const uint64_t iters_worst = worst * ((total_work + ck.n_slots - 1) / ck.n_slots + 1) + 2;
    if (iters_worst > CRT_WF_MAX_ITERS && k > 1) {   // shrink the chunk so the per-iteration counters cannot run out
      chunk = std::max(1u, k / 2);
      continue;
    }
    const size_t stage_bytes = (size_t)std::max(1u, ck.n_samples) * n_pixels * 16;
    if (stage_bytes > ctx->wf_stage_bytes) {
      if (ctx->wf_stage) { cudaFree(ctx->wf_stage); ctx->wf_stage = nullptr; ctx->wf_stage_bytes = 0; }
      CU(cudaMalloc(&ctx->wf_stage, stage_bytes));
      ctx->wf_stage_bytes = stage_bytes;
    }
    ck.stage = (float4*)ctx->wf_stage;
    if (ck.n_samples && n_units) {
      int rc = ensure_wavefront(ctx, ck.n_slots);
      if (rc) return rc;
      CU(cudaMemsetAsync(ctx->wf.iter, 0, kWfCounterBytes, ctx->stream));
      uint32_t it = 0;
      uint32_t next_check = (ctx->wf_last_iters > 4 && ctx->wf_last_work == total_work) ? ctx->wf_last_iters - 1 : (uint32_t)(2 * (total_work / ck.n_slots) + 2);
      for (;;) {
        if (!split) {
          CU(mark(ctx, CRT_KC_WF_LOGIC));
          CU(crt::launch_wf_logic(s, ctx->wf, ck, stats, logic_mode, grid_l, smem, it, ctx->d_counters, ctx->stream));
          ctx->launches++;
        } else {
          if (it) {
            CU(mark(ctx, CRT_KC_WF_SHADE));
            CU(crt::launch_wf_shade(s, ctx->wf, ck, stats, live_mode, grid_s, smem, it, ctx->d_counters, ctx->stream));
            ctx->launches++;
          }
          CU(mark(ctx, CRT_KC_WF_REGEN));
          CU(crt::launch_wf_regen(s, ctx->wf, ck, stats, live_mode, grid_r, smem, it, ctx->d_counters, ctx->stream));
          ctx->launches++;
        }
        if (s.n_tri) {
          CU(mark(ctx, CRT_KC_WF_TRAVERSE));
          CU(crt::launch_wf_traverse(s, ctx->wf, stats, live_mode == 1, live_mode == 2, grid_t, it, ctx->d_counters, ctx->stream));
          ctx->launches++;
        }
        if (live_mode == 2) {
          CU(mark(ctx, CRT_KC_WF_PARTITION));
          CU(crt::launch_wf_partition(ctx->wf, ck.n_slots, ctx->sm_count * 8, it, ctx->stream));
          ctx->launches++;
        }
        it++;
        if (it >= next_check) {
          CU(mark(ctx, -1));
          CU(cudaMemcpyAsync(ctx->wf_host_count, &ctx->wf.iter[it], 8, cudaMemcpyDeviceToHost, ctx->stream));   // n_bounce, n_finish
          CU(cudaStreamSynchronize(ctx->stream));
          if (ctx->wf_host_count[0] == 0 && ctx->wf_host_count[1] == 0) break;
          next_check = it + 4;
        }
        if (it + 2 >= CRT_WF_MAX_ITERS) return fail(ctx, CRT_ERR_STATE, "wavefront iteration limit reached");
      }
      ctx->wf_last_iters = it;
      ctx->wf_last_work = total_work;
      ctx->wf_iterations += it;
    }
    ctx->sample += k;   // UpdateVariables.wgsl:3-8, once per frame
    CU(mark(ctx, CRT_KC_WF_ACCUMULATE));
    CU(crt::launch_wf_accumulate(s, ck, ctx->d_acc, ctx->d_fb, ctx->sample, ctx->stream));
    CU(mark(ctx, -1));
    ctx->launches++;
    left -= k;
  }
  return CRT_OK;
}

size_t binding_bytes(const crt_ctx* ctx, uint32_t binding) {
  switch (binding) {
    case 0: return (size_t)ctx->W * ctx->H * 4;
    case 1: return (size_t)ctx->W * ctx->H * 16;
    case 2: return 4;
    default: return 0;
  }
}

}  // namespace

extern "C" {

const char* crt_version(void) { return CRT_VERSION_STRING; }

void crt_default_params(crt_params* p) {
  if (!p) return;
  memset(p, 0, sizeof(*p));
  p->max_depth = 100;   // CS:14
  p->grid_size = 16;    // CS:15
  p->t_min = 0.001f;    // CS:720
  p->exposure = 2.2f;   // CS:112
  p->kernel = CRT_KERNEL_AUTO;
  p->frames_per_launch = 0;
}

int crt_create(int device, crt_ctx** out) {
  if (!out) return CRT_ERR_ARG;
  *out = nullptr;
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0 || device < 0 || device >= n) return CRT_ERR_NODEVICE;
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return CRT_ERR_NODEVICE;
  if (prop.major != 10) return CRT_ERR_NODEVICE;   // the kernels are sm_100a only
  crt_ctx* ctx = new (std::nothrow) crt_ctx();
  if (!ctx) return CRT_ERR_OOM;
  ctx->device = device;
  ctx->sm_count = prop.multiProcessorCount;
  ctx->total_mem = prop.totalGlobalMem;
  crt_default_params(&ctx->params);
  memset(&ctx->scene, 0, sizeof(ctx->scene));
  memset(&ctx->wf, 0, sizeof(ctx->wf));
  if (cudaSetDevice(device) != cudaSuccess || cudaStreamCreateWithFlags(&ctx->own_stream, cudaStreamNonBlocking) != cudaSuccess ||
      cudaEventCreate(&ctx->ev_start) != cudaSuccess || cudaEventCreate(&ctx->ev_stop) != cudaSuccess ||
      cudaMalloc(&ctx->d_work, 4) != cudaSuccess || cudaMalloc(&ctx->d_counters, CRT_NUM_COUNTERS * 8) != cudaSuccess ||
      cudaMemset(ctx->d_counters, 0, CRT_NUM_COUNTERS * 8) != cudaSuccess) {
    crt_destroy(ctx);
    return CRT_ERR_CUDA;
  }
  ctx->stream = ctx->own_stream;
  *out = ctx;
  return CRT_OK;
}

void crt_destroy(crt_ctx* ctx) {
  if (!ctx) return;
  cudaSetDevice(ctx->device);
  if (ctx->stream) cudaStreamSynchronize(ctx->stream);
  dfree(ctx->d_acc); dfree(ctx->d_fb); dfree(ctx->d_tea); dfree(ctx->d_spectra); dfree(ctx->d_cie); dfree(ctx->d_prims_raw);
  dfree(ctx->d_lin); dfree(ctx->d_lights); dfree(ctx->d_nodes); dfree(ctx->d_work); dfree(ctx->d_counters);
  if (ctx->wf_mem) cudaFree(ctx->wf_mem);
  if (ctx->wf_stage) cudaFree(ctx->wf_stage);
  if (ctx->wf_host_count) cudaFreeHost(ctx->wf_host_count);
  if (ctx->ev_start) cudaEventDestroy(ctx->ev_start);
  if (ctx->ev_stop) cudaEventDestroy(ctx->ev_stop);
  for (auto& e : ctx->ev_seq) cudaEventDestroy(e.first);
  for (auto& e : ctx->ev_free) cudaEventDestroy(e);
  if (ctx->own_stream) cudaStreamDestroy(ctx->own_stream);
  delete ctx;
}

const char* crt_last_error(crt_ctx* ctx) { return ctx ? ctx->err.c_str() : "null context"; }

int crt_set_binding(crt_ctx* ctx, uint32_t binding, const void* bytes, size_t nbytes) {
  if (!ctx) return CRT_ERR_ARG;
  if (!bytes && nbytes) return fail(ctx, CRT_ERR_ARG, "null bytes");
  CU(cudaSetDevice(ctx->device));
  const uint8_t* p = (const uint8_t*)bytes;
  switch (binding) {
    case 1: {   // resume: restore a saved accumulator
      if (ctx->have[4] && ctx->camera_dirty) { int rc = upload_camera(ctx); if (rc) return rc; }
      if (!ctx->acc_pixels) { ctx->pending_acc.assign(p, p + nbytes); return CRT_OK; }
      if (nbytes != ctx->acc_pixels * 16) return fail(ctx, CRT_ERR_SIZE, "binding 1 is %zu bytes, got %zu", ctx->acc_pixels * 16, nbytes);
      CU(cudaMemcpyAsync(ctx->d_acc, p, nbytes, cudaMemcpyHostToDevice, ctx->stream));
      CU(cudaStreamSynchronize(ctx->stream));
      return CRT_OK;
    }
    case 2:
      if (nbytes != 4) return fail(ctx, CRT_ERR_SIZE, "binding 2 is 4 bytes, got %zu", nbytes);
      memcpy(&ctx->sample, p, 4);
      return CRT_OK;
    case 3:   // bound by the reference, read by no live shader code (CS:4): accepted and ignored
      if (nbytes % 64) return fail(ctx, CRT_ERR_SIZE, "binding 3 has a 64-byte stride, got %zu bytes", nbytes);
      ctx->have[3] = true;
      return CRT_OK;
    case 4:
      if (nbytes != 64) return fail(ctx, CRT_ERR_SIZE, "binding 4 (camera) is 64 bytes, got %zu", nbytes);
      ctx->b_camera.assign(p, p + nbytes); ctx->have[4] = true; ctx->camera_dirty = true;
      return CRT_OK;
    case 5:
      if (nbytes != 3 * 471 * 4) return fail(ctx, CRT_ERR_SIZE, "binding 5 (CIE) is %d bytes, got %zu", 3 * 471 * 4, nbytes);
      ctx->b_cie.assign(p, p + nbytes); ctx->have[5] = true; ctx->tables_dirty = true;
      return CRT_OK;
    case 6:
      if (nbytes == 0 || nbytes % (CRT_LAMBDA_RANGE * 4)) return fail(ctx, CRT_ERR_SIZE, "binding 6 (spectra) has a %u-byte stride, got %zu bytes", CRT_LAMBDA_RANGE * 4, nbytes);
      ctx->b_spectra.assign(p, p + nbytes); ctx->have[6] = true; ctx->tables_dirty = true;
      return CRT_OK;
    case 7:
      if (nbytes == 0 || nbytes % 80) return fail(ctx, CRT_ERR_SIZE, "binding 7 (lights) has an 80-byte stride, got %zu bytes", nbytes);
      ctx->b_lights.assign(p, p + nbytes); ctx->have[7] = true; ctx->scene_dirty = true;
      return CRT_OK;
    case 8:
      if (nbytes % 80) return fail(ctx, CRT_ERR_SIZE, "binding 8 (primitives) has an 80-byte stride, got %zu bytes", nbytes);
      ctx->b_prims.assign(p, p + nbytes); ctx->have[8] = true; ctx->scene_dirty = true;
      return CRT_OK;
    default:
      return fail(ctx, CRT_ERR_BINDING, "binding %u cannot be set", binding);
  }
}

int crt_reset(crt_ctx* ctx) {
  if (!ctx) return CRT_ERR_ARG;
  CU(cudaSetDevice(ctx->device));
  ctx->sample = 0;
  ctx->pending_acc.clear();
  if (ctx->d_acc) {
    CU(cudaMemsetAsync(ctx->d_acc, 0, ctx->acc_pixels * sizeof(float4), ctx->stream));
    CU(cudaMemsetAsync(ctx->d_fb, 0, ctx->acc_pixels * sizeof(uint32_t), ctx->stream));
  }
  return CRT_OK;
}

int crt_set_params(crt_ctx* ctx, const crt_params* p) {
  if (!ctx || !p) return CRT_ERR_ARG;
  if (p->grid_size == 0) return fail(ctx, CRT_ERR_ARG, "grid_size must be >= 1 (sample %% grid_size, CS:497)");
  if (p->kernel > CRT_KERNEL_WAVEFRONT) return fail(ctx, CRT_ERR_ARG, "unknown kernel %u", p->kernel);
  ctx->params = *p;
  return CRT_OK;
}

int crt_get_params(crt_ctx* ctx, crt_params* p) {
  if (!ctx || !p) return CRT_ERR_ARG;
  *p = ctx->params;
  return CRT_OK;
}

int crt_set_shard(crt_ctx* ctx, uint32_t rank, uint32_t world, uint32_t mode) {
  if (!ctx) return CRT_ERR_ARG;
  if (world == 0 || rank >= world || mode > CRT_SHARD_SPLIT_SPP) return fail(ctx, CRT_ERR_ARG, "bad shard (%u of %u, mode %u)", rank, world, mode);
  ctx->rank = rank; ctx->world = world; ctx->shard_mode = world > 1 ? mode : CRT_SHARD_NONE;
  return CRT_OK;
}

int crt_dispatch_frames(crt_ctx* ctx, uint32_t n) {
  if (!ctx) return CRT_ERR_ARG;
  CU(cudaSetDevice(ctx->device));
  int rc = prepare(ctx);
  if (rc) return rc;
  if (n > 0xFFFFFFFFu - ctx->sample) return fail(ctx, CRT_ERR_ARG, "frame counter would overflow");
  const DevScene& s = ctx->scene;
  const bool stats = (ctx->params.flags & CRT_FLAG_STATS) != 0;
  uint32_t n_tiles = s.tiles_x * s.tiles_y;
  if (ctx->shard_mode == CRT_SHARD_TILES && ctx->world > 1)
    n_tiles = (n_tiles > ctx->rank) ? (n_tiles - ctx->rank + ctx->world - 1) / ctx->world : 0;
  const uint32_t n_units = n_tiles * 32u;
  // AUTO: the wavefront form when there is a BVH (ncu: the megakernel runs 4.75 of 32 lanes there) and, for linear-list
  // scenes, whenever the dispatch has two frames or more.  Measured on the reference's default scene at 1080p
  // (profiles/r01r_auto_threshold.txt), ms per dispatch of 1 / 2 / 4 / 64 frames: megakernel 2.2 / 4.4 / 8.3 / 125,
  // wavefront 2.8 / 3.8 / 5.6 / 64 -- the wavefront pays ~2 ms per dispatch for the thin iterations of the paths that loop
  // to max_depth inside the glass sphere and half the megakernel's time per frame after that.  Both forms give the same bits.
  uint32_t kernel = ctx->params.kernel;
  if (kernel == CRT_KERNEL_AUTO)
    kernel = (ctx->params.max_depth <= 1000 && (s.n_tri > 0 || n >= 2)) ? CRT_KERNEL_WAVEFRONT : CRT_KERNEL_MEGA;
  // a light that is itself a triangle needs closest-hit shadow rays: only the megakernel carries that path
  if (s.shadow_closest) kernel = CRT_KERNEL_MEGA;
  CU(cudaEventRecord(ctx->ev_start, ctx->stream));
  if (kernel == CRT_KERNEL_WAVEFRONT) {
    if (n) { rc = dispatch_wavefront(ctx, n, n_units, stats); if (rc) return rc; }
  } else {
    size_t smem = crt::megakernel_tables_bytes(s);
    bool tables_in_smem = smem <= 48 * 1024;
    if (!tables_in_smem) smem = 0;
    int bps = 0;
    CU(crt::megakernel_config(stats, smem, &bps));
    if (bps < 1) return fail(ctx, CRT_ERR_CUDA, "megakernel does not fit on an SM");
    const int grid = ctx->sm_count * bps;
    uint32_t chunk = ctx->params.frames_per_launch ? ctx->params.frames_per_launch : 8u;
    uint32_t left = n;
    while (left) {
      const uint32_t k = left < chunk ? left : chunk;
      CU(cudaMemsetAsync(ctx->d_work, 0, 4, ctx->stream));
      CU(mark(ctx, CRT_KC_MEGA));
      CU(crt::launch_megakernel(s, stats, grid, smem, tables_in_smem, ctx->d_acc, ctx->d_fb, ctx->sample, k, n_units, ctx->d_work,
                                ctx->d_counters, ctx->stream));
      ctx->sample += k;   // UpdateVariables.wgsl:3-8, once per frame
      ctx->launches++;
      left -= k;
    }
    CU(mark(ctx, -1));
  }
  if (ctx->ev_seq.size() > 32768) CU(drain_marks(ctx));
  CU(cudaEventRecord(ctx->ev_stop, ctx->stream));
  ctx->timed = true;
  return CRT_OK;
}

int crt_synchronize(crt_ctx* ctx) {
  if (!ctx) return CRT_ERR_ARG;
  CU(cudaSetDevice(ctx->device));
  CU(cudaStreamSynchronize(ctx->stream));
  return CRT_OK;
}

int crt_last_dispatch_ms(crt_ctx* ctx, float* ms) {
  if (!ctx || !ms) return CRT_ERR_ARG;
  if (!ctx->timed) return fail(ctx, CRT_ERR_STATE, "no dispatch has been timed yet");
  CU(cudaSetDevice(ctx->device));
  CU(cudaEventSynchronize(ctx->ev_stop));
  CU(cudaEventElapsedTime(ms, ctx->ev_start, ctx->ev_stop));
  return CRT_OK;
}

int crt_read_binding(crt_ctx* ctx, uint32_t binding, void* dst, size_t nbytes) {
  if (!ctx || !dst) return CRT_ERR_ARG;
  if (binding > 2) return fail(ctx, CRT_ERR_BINDING, "binding %u cannot be read back", binding);
  CU(cudaSetDevice(ctx->device));
  if (binding == 2) {
    if (nbytes != 4) return fail(ctx, CRT_ERR_SIZE, "binding 2 is 4 bytes, got %zu", nbytes);
    CU(cudaStreamSynchronize(ctx->stream));
    memcpy(dst, &ctx->sample, 4);
    return CRT_OK;
  }
  if (ctx->have[4] && ctx->camera_dirty) { int rc = upload_camera(ctx); if (rc) return rc; }
  if (!ctx->d_acc) return fail(ctx, CRT_ERR_STATE, "no frame buffers yet: set binding 4 (camera) first");
  const size_t want = binding_bytes(ctx, binding);
  if (nbytes != want) return fail(ctx, CRT_ERR_SIZE, "binding %u is %zu bytes, got %zu", binding, want, nbytes);
  CU(cudaMemcpyAsync(dst, binding == 0 ? (const void*)ctx->d_fb : (const void*)ctx->d_acc, nbytes, cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  return CRT_OK;
}

int crt_resolve_framebuffer(crt_ctx* ctx) {
  if (!ctx) return CRT_ERR_ARG;
  CU(cudaSetDevice(ctx->device));
  if (!ctx->d_acc) return fail(ctx, CRT_ERR_STATE, "no frame buffers yet");
  CU(crt::launch_resolve(ctx->d_acc, ctx->d_fb, (uint32_t)ctx->acc_pixels, ctx->sample, ctx->params.exposure, ctx->stream));
  return CRT_OK;
}

int crt_device_ptr(crt_ctx* ctx, uint32_t binding, void** dptr, size_t* nbytes) {
  if (!ctx || !dptr) return CRT_ERR_ARG;
  if (binding > 1) return fail(ctx, CRT_ERR_BINDING, "binding %u has no device buffer", binding);
  CU(cudaSetDevice(ctx->device));
  if (ctx->have[4] && ctx->camera_dirty) { int rc = upload_camera(ctx); if (rc) return rc; }
  if (!ctx->d_acc) return fail(ctx, CRT_ERR_STATE, "no frame buffers yet: set binding 4 (camera) first");
  *dptr = binding == 0 ? (void*)ctx->d_fb : (void*)ctx->d_acc;
  if (nbytes) *nbytes = binding_bytes(ctx, binding);
  return CRT_OK;
}

int crt_set_stream(crt_ctx* ctx, void* cuda_stream) {
  if (!ctx) return CRT_ERR_ARG;
  CU(cudaSetDevice(ctx->device));
  CU(cudaStreamSynchronize(ctx->stream));
  ctx->stream = cuda_stream ? (cudaStream_t)cuda_stream : ctx->own_stream;
  return CRT_OK;
}

int crt_stream(crt_ctx* ctx, void** cuda_stream) {
  if (!ctx || !cuda_stream) return CRT_ERR_ARG;
  *cuda_stream = (void*)ctx->stream;
  return CRT_OK;
}

int crt_trace_rays(crt_ctx* ctx, const float* rays, const uint32_t* exclude, size_t n, uint32_t* hit_idx, float* t) {
  if (!ctx || (!rays && n) || (!hit_idx && n) || (!t && n)) return CRT_ERR_ARG;
  if (n > 0x7FFFFFFFu) return fail(ctx, CRT_ERR_SIZE, "too many rays");
  CU(cudaSetDevice(ctx->device));
  int rc = prepare(ctx);
  if (rc) return rc;
  if (!n) return CRT_OK;
  float4* d_rays = nullptr; uint32_t* d_ex = nullptr; uint32_t* d_idx = nullptr; float* d_t = nullptr;
  CU(cudaMalloc(&d_rays, n * 32));
  CU(cudaMalloc(&d_idx, n * 4));
  CU(cudaMalloc(&d_t, n * 4));
  CU(cudaMemcpyAsync(d_rays, rays, n * 32, cudaMemcpyHostToDevice, ctx->stream));
  if (exclude) {
    CU(cudaMalloc(&d_ex, n * 4));
    CU(cudaMemcpyAsync(d_ex, exclude, n * 4, cudaMemcpyHostToDevice, ctx->stream));
  }
  CU(crt::launch_trace_rays(ctx->scene, d_rays, d_ex, (uint32_t)n, d_idx, d_t, ctx->d_counters, ctx->stream));
  CU(cudaMemcpyAsync(hit_idx, d_idx, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaMemcpyAsync(t, d_t, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  cudaFree(d_rays); cudaFree(d_idx); cudaFree(d_t); if (d_ex) cudaFree(d_ex);
  return CRT_OK;
}

int crt_get_counters(crt_ctx* ctx, uint64_t* out, size_t n) {
  if (!ctx || !out) return CRT_ERR_ARG;
  if (n > CRT_NUM_COUNTERS) n = CRT_NUM_COUNTERS;
  CU(cudaSetDevice(ctx->device));
  unsigned long long tmp[CRT_NUM_COUNTERS];
  CU(cudaMemcpyAsync(tmp, ctx->d_counters, sizeof(tmp), cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  tmp[CRT_CNT_LAUNCHES] = ctx->launches;
  tmp[CRT_CNT_WF_ITERATIONS] = ctx->wf_iterations;
  for (size_t i = 0; i < n; i++) out[i] = tmp[i];
  if (ctx->wf_mem && ctx->wf.slow_min) {   // diagnostics: print and clear the slow-ray log (CRT_SLOW_RAY_NODES)
    static float4 log[1 + 3 * CRT_WF_SLOW_RAYS];
    CU(cudaMemcpy(log, ctx->wf.slow, sizeof(log), cudaMemcpyDeviceToHost));
    uint32_t cnt; memcpy(&cnt, &log[0].x, 4);
    fprintf(stderr, "[libcrt] %u rays visited >= %u BVH nodes\n", cnt, ctx->wf.slow_min);
    for (uint32_t k = 0; k < cnt && k < CRT_WF_SLOW_RAYS; k++) {
      const float4 a = log[1 + 3 * k], b = log[2 + 3 * k], c = log[3 + 3 * k];
      uint32_t it, nodes, excl, any, slot, tris;
      memcpy(&it, &a.w, 4); memcpy(&nodes, &b.w, 4); memcpy(&excl, &c.x, 4); memcpy(&any, &c.y, 4); memcpy(&slot, &c.z, 4); memcpy(&tris, &c.w, 4);
      fprintf(stderr, "[libcrt] slow ray: o=(%.9g, %.9g, %.9g) d=(%.9g, %.9g, %.9g) exclude=%u anyhit=%u iteration=%u nodes=%u triangles=%u slot=%u\n",
              a.x, a.y, a.z, b.x, b.y, b.z, excl, any, it, nodes, tris, slot);
    }
    CU(cudaMemset(ctx->wf.slow, 0, 16));
  }
  return CRT_OK;
}

int crt_reset_counters(crt_ctx* ctx) {
  if (!ctx) return CRT_ERR_ARG;
  CU(cudaSetDevice(ctx->device));
  CU(cudaMemsetAsync(ctx->d_counters, 0, CRT_NUM_COUNTERS * 8, ctx->stream));
  ctx->launches = 0;
  ctx->wf_iterations = 0;
  CU(drain_marks(ctx));
  for (int k = 0; k < CRT_NUM_KERNEL_CLASSES; k++) { ctx->k_ms[k] = 0.0; ctx->k_launches[k] = 0; }
  return CRT_OK;
}

int crt_get_kernel_times(crt_ctx* ctx, double* ms, uint64_t* launches, size_t n) {
  if (!ctx || !ms || !launches) return CRT_ERR_ARG;
  CU(cudaSetDevice(ctx->device));
  CU(drain_marks(ctx));
  for (size_t k = 0; k < n && k < CRT_NUM_KERNEL_CLASSES; k++) { ms[k] = ctx->k_ms[k]; launches[k] = ctx->k_launches[k]; }
  return CRT_OK;
}

int crt_scene_stats(crt_ctx* ctx, uint64_t* out, size_t n) {
  if (!ctx || !out) return CRT_ERR_ARG;
  CU(cudaSetDevice(ctx->device));
  int rc = prepare(ctx);
  if (rc) return rc;
  for (size_t i = 0; i < n && i < 8; i++) out[i] = ctx->stats[i];
  return CRT_OK;
}

// Test hook, host only: the BVH builder of the scene-upload module without a device (tests/test_bvh_builder.py).
CRT_API int crt_host_build_bvh(const void* prims, size_t nbytes, int threads, uint32_t* nodes, size_t node_words, uint32_t* tri_src,
                               size_t tri_cap, uint64_t* info, float* bounds) {
  if ((!prims && nbytes) || nbytes % 80 || !info) return CRT_ERR_ARG;
  const uint8_t* pb = (const uint8_t*)prims;
  std::vector<uint32_t> tri_pos;
  for (size_t i = 0; i < nbytes / 80; i++) {
    uint32_t cat; memcpy(&cat, pb + i * 80, 4);
    if (cat == 2u) tri_pos.push_back((uint32_t)i);
  }
  crt::BvhResult bvh;
  crt::build_bvh(pb, tri_pos, threads, bvh);
  info[0] = bvh.n_nodes; info[1] = bvh.tri_src.size(); info[2] = bvh.max_depth; info[3] = 4 * CRT_NODE_STRIDE;
  if (bounds) { for (int a = 0; a < 3; a++) { bounds[a] = bvh.lo[a]; bounds[3 + a] = bvh.hi[a]; } bounds[6] = bvh.pad; }
  if (!nodes) return CRT_OK;
  if (node_words < bvh.nodes.size() || tri_cap < bvh.tri_src.size() || (!tri_src && !bvh.tri_src.empty())) return CRT_ERR_SIZE;
  if (!bvh.nodes.empty()) memcpy(nodes, bvh.nodes.data(), bvh.nodes.size() * 4);
  if (!bvh.tri_src.empty()) memcpy(tri_src, bvh.tri_src.data(), bvh.tri_src.size() * 4);
  return CRT_OK;
}

// Test hook (not in crt.h's reference mapping): the CUDA build of crt_detmath.h on n inputs.
CRT_API int crt_test_detmath(crt_ctx* ctx, const float* x, size_t n, float* sn, float* cs, float* ex, float* pw) {
  if (!ctx || !x || !sn || !cs || !ex || !pw) return CRT_ERR_ARG;
  CU(cudaSetDevice(ctx->device));
  float* d = nullptr;
  CU(cudaMalloc(&d, n * 4 * 5));
  CU(cudaMemcpyAsync(d, x, n * 4, cudaMemcpyHostToDevice, ctx->stream));
  CU(crt::launch_detmath(d, (uint32_t)n, d + n, d + 2 * n, d + 3 * n, d + 4 * n, ctx->stream));
  CU(cudaMemcpyAsync(sn, d + n, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaMemcpyAsync(cs, d + 2 * n, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaMemcpyAsync(ex, d + 3 * n, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaMemcpyAsync(pw, d + 4 * n, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  cudaFree(d);
  return CRT_OK;
}

}  // extern "C"
