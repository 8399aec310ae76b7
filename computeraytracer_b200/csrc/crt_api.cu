// crt_api.cu -- the C ABI of include/crt.h: context, bindings, scene-upload module, dispatch, multi-GPU combine.
//
// Replaces the host half of the reference's WebGPU contract (src/main.js): buffer creation and
// upload (MJ:139-393), bind group (MJ:465-529), the per-frame dispatch pair (MJ:597-611).
// There is no CPU fallback: without a usable sm_100 device crt_create fails.
#include <cuda_runtime.h>
#include <dlfcn.h>

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <unordered_map>
#include <unordered_set>
#include <utility>
#include <vector>

#include "../../include/crt.h"
#include "crt_bvh.h"
#include "crt_bvh_device.h"
#include "crt_device.cuh"
#include "crt_kernels.h"
#include "crt_wavefront.h"

#ifndef CRT_BVH_AUTO_DEVICE_MIN
#define CRT_BVH_AUTO_DEVICE_MIN 32768u   // CRT_BVH_BUILDER_AUTO: device build from this many triangles on
#endif
#define CRT_VERSION_STRING "libcrt 0.2 (sm_100a)"

struct crt_ctx {
  int device = 0;
  int sm_count = 0;
  size_t total_mem = 0;
  cudaStream_t stream = nullptr;
  cudaStream_t own_stream = nullptr;
  uint64_t launches = 0;   // hot-path kernel launches made by the host since the last crt_reset_counters
  cudaEvent_t ev_start = nullptr, ev_stop = nullptr;
  bool timed = false;
  std::string err;

  // host copies of the reference-layout bindings (binding 8 is shared between the members of a group)
  std::vector<uint8_t> b_camera, b_cie, b_spectra, b_lights;
  std::shared_ptr<const std::vector<uint8_t>> b_prims;
  bool have[9] = {false, false, false, false, false, false, false, false, false};
  bool scene_dirty = true, tables_dirty = true, camera_dirty = true;
  std::shared_ptr<const crt::BvhResult> host_bvh;   // host-built tree of b_prims (a group builds it once)
  std::shared_ptr<const std::vector<uint8_t>> host_bvh_for;   // the binding-8 bytes the cached tree was built from (held, so that
                                                                // a later scene cannot be allocated at the same address and match)
  uint32_t host_bvh_tuning = 0; float host_bvh_prim_cost = 0.0f;

  crt_params params;
  uint32_t rank = 0, world = 1, shard_mode = CRT_SHARD_NONE;
  uint32_t sample = 0;   // binding 2
  uint32_t W = 0, H = 0;

  // device state
  float4* d_acc = nullptr;       // binding 1
  uint32_t* d_fb = nullptr;      // binding 0
  uint32_t* d_tea = nullptr;     // per-pixel tea(x, y*100) of the RNG seed (CS:98)
  uint32_t tea_W = 0, tea_H = 0;
  float* d_cam = nullptr;        // 64 bytes of binding 4 + 48 bytes of camera constants
  size_t acc_pixels = 0;
  float* d_spectra = nullptr; float* d_cie = nullptr;
  LinPrim* d_lin = nullptr; LightRec* d_lights = nullptr; uint32_t* d_light_mask = nullptr;
  float4* d_tri = nullptr; uint4* d_nodes = nullptr;   // one allocation: nodes, then triangles (d_tri points into it)
  size_t scene_bytes = 0;
  uint32_t* d_work = nullptr;
  unsigned long long* d_counters = nullptr;
  DevScene scene;
  uint64_t stats[10] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
  std::vector<uint8_t> pending_acc;  // binding 1 set before the size is known

  // wavefront path (crt_wavefront.cu)
  WfState wf;
  void* wf_mem = nullptr;            // one allocation behind all WfState arrays
  uint32_t wf_slots = 0;
  uint32_t* wf_host_count = nullptr; // pinned: live-slot count read back at check points (host loop)
  void* wf_stage = nullptr;          // per-sample XYZ staging of the current chunk
  size_t wf_stage_bytes = 0;
  uint64_t wf_last_work = 0;
  uint32_t wf_last_iters = 0;        // iterations the previous chunk needed (predicts the next check point of the host loop)
  uint64_t wf_iterations = 0;        // host loop; the graph loop counts on the device (WfCtl::iters_total)
  cudaGraph_t wf_graph = nullptr;    // the iteration loop as a CUDA-graph WHILE node, rebuilt when its key changes
  cudaGraphExec_t wf_graph_exec = nullptr;
  std::vector<uint8_t> wf_graph_key;
  // second lane (dispatch_wavefront: a chunk's samples are split over two half pools that run side by side on two streams)
  WfState wf2;
  void* wf2_mem = nullptr;
  uint32_t wf2_slots = 0;
  cudaGraph_t wf2_graph = nullptr;
  cudaGraphExec_t wf2_graph_exec = nullptr;
  std::vector<uint8_t> wf2_graph_key;
  cudaStream_t lane_stream[2] = {nullptr, nullptr};
  cudaEvent_t lane_fork = nullptr, lane_join[2] = {nullptr, nullptr};
  std::vector<float4> slow_log;

  // CRT_FLAG_KERNEL_TIMES: events recorded between consecutive hot-path launches
  std::vector<cudaEvent_t> ev_free;
  std::vector<std::pair<cudaEvent_t, int>> ev_seq;   // (event, class of the kernel launched right after it; -1 = end of a run)
  double k_ms[CRT_NUM_KERNEL_CLASSES] = {};
  uint64_t k_launches[CRT_NUM_KERNEL_CLASSES] = {};

  // multi-GPU
  std::vector<crt_ctx*> members;     // non-empty: this context is a single-process group (crt_create_multi)
  crt_ctx* group = nullptr;          // member of a group
  void* nccl_comm = nullptr;         // one process per GPU (crt_comm_init)
  float4* d_comb = nullptr;          // combined accumulator (on the destination of crt_combine)
  bool comb_valid = false;           // bindings 0/1 read the combined image until the next dispatch / reset / set_binding
  float4* d_xfer = nullptr; size_t xfer_bytes = 0;   // tile gather: packed send buffer / receive buffer
  cudaEvent_t ev_c0 = nullptr, ev_c1 = nullptr, ev_sync = nullptr;
  bool comb_timed = false;
  uint64_t comb_bytes = 0;
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

// a device temporary that is freed on every way out of a function (the CU() macro returns early)
struct DevTmp {
  void* p = nullptr;
  ~DevTmp() { if (p) cudaFree(p); }
  cudaError_t alloc(size_t n) { return cudaMalloc(&p, n ? n : 1); }
  template <typename T> T* as() const { return (T*)p; }
};

// ------------------------------------------------------------------------------------------------ NCCL, loaded at run time
// The library has no link-time dependency on NCCL: a single-GPU host (or the single-process group, which uses peer pointers)
// never needs it.  The declarations below are NCCL's public C ABI (nccl.h: ncclUniqueId is 128 bytes passed by value,
// ncclFloat32 = 7, ncclSum = 0), stable across 2.x.
struct NcclId { char internal[CRT_COMM_ID_BYTES]; };
struct NcclApi {
  void* lib = nullptr;
  std::string err;
  int (*GetUniqueId)(NcclId*) = nullptr;
  int (*CommInitRank)(void**, int, NcclId, int) = nullptr;
  int (*CommDestroy)(void*) = nullptr;
  const char* (*GetErrorString)(int) = nullptr;
  int (*Reduce)(const void*, void*, size_t, int, int, int, void*, cudaStream_t) = nullptr;
  int (*Send)(const void*, size_t, int, int, void*, cudaStream_t) = nullptr;
  int (*Recv)(void*, size_t, int, int, void*, cudaStream_t) = nullptr;
  int (*GroupStart)() = nullptr;
  int (*GroupEnd)() = nullptr;
};
constexpr int kNcclFloat32 = 7, kNcclSum = 0;

NcclApi* nccl_api() {
  static NcclApi api;
  static std::once_flag once;
  std::call_once(once, [] {
    // a process that already holds an NCCL (torch ships one) gets that one: same SONAME
    api.lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!api.lib) api.lib = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!api.lib) { api.err = std::string("dlopen(libnccl.so.2): ") + dlerror(); return; }
    bool ok = true;
    auto sym = [&](const char* name) -> void* { void* p = dlsym(api.lib, name); if (!p) { ok = false; api.err = std::string("missing symbol ") + name; } return p; };
    api.GetUniqueId = (int (*)(NcclId*))sym("ncclGetUniqueId");
    api.CommInitRank = (int (*)(void**, int, NcclId, int))sym("ncclCommInitRank");
    api.CommDestroy = (int (*)(void*))sym("ncclCommDestroy");
    api.GetErrorString = (const char* (*)(int))sym("ncclGetErrorString");
    api.Reduce = (int (*)(const void*, void*, size_t, int, int, int, void*, cudaStream_t))sym("ncclReduce");
    api.Send = (int (*)(const void*, size_t, int, int, void*, cudaStream_t))sym("ncclSend");
    api.Recv = (int (*)(void*, size_t, int, int, void*, cudaStream_t))sym("ncclRecv");
    api.GroupStart = (int (*)())sym("ncclGroupStart");
    api.GroupEnd = (int (*)())sym("ncclGroupEnd");
    if (!ok) { dlclose(api.lib); api.lib = nullptr; }
  });
  return api.lib ? &api : nullptr;
}

#define NC(call)                                                                                               \
  do {                                                                                                         \
    int r_ = (call);                                                                                           \
    if (r_ != 0) return fail(ctx, CRT_ERR_NCCL, "%s: %s (%s:%d)", #call, nc->GetErrorString(r_), __FILE__, __LINE__); \
  } while (0)

// ------------------------------------------------------------------------------------------------ kernel-time marks
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

bool is_group(const crt_ctx* ctx) { return !ctx->members.empty(); }

int ensure_frame_buffers(crt_ctx* ctx) {
  const size_t n = (size_t)ctx->W * ctx->H;
  if (n == ctx->acc_pixels && ctx->d_acc) return CRT_OK;
  dfree(ctx->d_acc); dfree(ctx->d_fb); dfree(ctx->d_tea); dfree(ctx->d_comb);
  ctx->acc_pixels = 0; ctx->tea_W = ctx->tea_H = 0; ctx->comb_valid = false;
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
  if (!ctx->d_cam) CU(cudaMalloc(&ctx->d_cam, 64 + 48));
  CU(cudaMemcpyAsync(ctx->d_cam, cam, 64, cudaMemcpyHostToDevice, ctx->stream));
  const float vph = 2.0f * (float)tan((double)(cam[13] / 2.0f));   // tan pin: double libm, rounded once (CS:479)
  CU(crt::launch_camera_setup(ctx->d_cam, vph, ctx->d_cam + 16, ctx->stream));
  float out[12];
  CU(cudaMemcpyAsync(out, ctx->d_cam + 16, 48, cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  DevScene& s = ctx->scene;
  memcpy(s.eye, out, 12); memcpy(s.hor, out + 3, 12); memcpy(s.ver, out + 6, 12); memcpy(s.llc, out + 9, 12);
  s.W = W; s.H = H;
  s.tiles_x = (W + 7) / 8; s.tiles_y = (H + 3) / 4;
  if (ctx->tea_W != W || ctx->tea_H != H) {   // the table is indexed x + y * W: rebuilt only when the viewport changes
    CU(crt::launch_tea_tab(W, H, ctx->d_tea, ctx->stream));
    ctx->tea_W = W; ctx->tea_H = H;
  }
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
// same numerics contract as the kernels.  The tree is built on the device (Morton-sorted LBVH collapsed to Wide8,
// crt_bvh_device.cu) for scenes of 32768 triangles and more, on the host cores (binned SAH, crt_bvh.cpp) below that or
// when crt_params.bvh_tuning asks for it; tree shape never changes a result.
int upload_scene(crt_ctx* ctx) {
  const auto t0 = std::chrono::steady_clock::now();
  static const std::vector<uint8_t> kNoPrims;
  const std::vector<uint8_t>& pv = ctx->b_prims ? *ctx->b_prims : kNoPrims;
  const size_t n_prims = pv.size() / 80;
  const size_t n_lights = ctx->b_lights.size() / 80;
  if (n_lights == 0) return fail(ctx, CRT_ERR_SIZE, "binding 7 holds no light (the shader samples one per diffuse bounce, CS:383)");
  if (n_lights > 4096) return fail(ctx, CRT_ERR_SIZE, "at most 4096 lights are supported");
  if (n_prims >= (1u << 28)) return fail(ctx, CRT_ERR_SIZE, "too many primitives");
  const uint8_t* prims = pv.data();
  std::vector<uint32_t> lin_pos, tri_pos;
  for (size_t i = 0; i < n_prims; i++) {
    uint32_t cat; memcpy(&cat, prims + i * 80, 4);
    if (cat == 2u) tri_pos.push_back((uint32_t)i);
    else if (cat <= 1u) lin_pos.push_back((uint32_t)i);
    // other categories match no branch of ray_intersection (CS:525, CS:584) and can never be hit
  }
  const uint32_t builder = ctx->params.bvh_tuning & CRT_BVH_BUILDER_MASK;
  const bool device_build = !tri_pos.empty() && (builder == CRT_BVH_BUILDER_DEVICE || (builder == CRT_BVH_BUILDER_AUTO && tri_pos.size() >= CRT_BVH_AUTO_DEVICE_MIN));

  dfree(ctx->d_lin); dfree(ctx->d_lights); ctx->d_tri = nullptr; dfree(ctx->d_nodes); ctx->scene_bytes = 0;
  DevScene& s = ctx->scene;
  s.n_lin = (uint32_t)lin_pos.size(); s.n_tri = (uint32_t)tri_pos.size(); s.n_nodes = 0; s.n_lights = (uint32_t)n_lights;
  s.lin_g = nullptr; s.tri = nullptr; s.nodes = nullptr; s.lights_g = nullptr;

  DevTmp d_prims_raw, d_lights_raw, d_src;
  if (n_prims) {
    CU(d_prims_raw.alloc(n_prims * 80));
    CU(cudaMemcpyAsync(d_prims_raw.p, prims, n_prims * 80, cudaMemcpyHostToDevice, ctx->stream));
  }
  const size_t n_src = std::max(lin_pos.size(), tri_pos.size());
  if (n_src) CU(d_src.alloc(n_src * 4));
  if (s.n_lin) {
    CU(cudaMalloc(&ctx->d_lin, (size_t)s.n_lin * sizeof(LinPrim)));
    CU(cudaMemcpyAsync(d_src.p, lin_pos.data(), lin_pos.size() * 4, cudaMemcpyHostToDevice, ctx->stream));
    CU(crt::launch_repack_lin(d_prims_raw.p, d_src.as<uint32_t>(), s.n_lin, ctx->d_lin, ctx->stream));
    const uint32_t nc = s.n_lin < CRT_LIN_PARAM_MAX ? s.n_lin : CRT_LIN_PARAM_MAX;
    CU(cudaMemcpyAsync(s.lin_c, ctx->d_lin, nc * sizeof(LinPrim), cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    s.lin_g = ctx->d_lin;
  }
  uint32_t max_depth = 0;
  double build_ms = 0.0;
  bool built_on_device = false;
  if (s.n_tri && device_build) {
    // tri_pos -> device; the builder leaves nodes + leaf-ordered triangles in one allocation it hands over
    CU(cudaMemcpyAsync(d_src.p, tri_pos.data(), tri_pos.size() * 4, cudaMemcpyHostToDevice, ctx->stream));
    crt::DeviceBvh db;
    std::string msg;
    const cudaError_t e = crt::build_bvh_device(d_prims_raw.p, d_src.as<uint32_t>(), s.n_tri, ctx->params.bvh_tuning, ctx->params.bvh_prim_cost,
                                                ctx->stream, &db, &msg);
    if (e != cudaSuccess) return fail(ctx, e == cudaErrorMemoryAllocation ? CRT_ERR_OOM : CRT_ERR_CUDA, "device BVH build: %s (%s)", cudaGetErrorString(e), msg.c_str());
    if (db.max_depth + 2 > CRT_BVH_STACK) {
      // an agglomerative build can chain degenerate input (hundreds of nested triangles around one centroid) deeper than
      // the traversal stack: such scenes take the host SAH build, whose median splits bound the depth
      cudaFree(db.nodes);
    } else {
      ctx->d_nodes = db.nodes; ctx->d_tri = db.tri; ctx->scene_bytes = db.bytes;
      s.n_nodes = db.n_nodes; max_depth = db.max_depth; build_ms = db.build_ms;
      for (int a = 0; a < 3; a++) { s.bvh_lo[a] = db.lo[a]; s.bvh_hi[a] = db.hi[a]; }
      s.tri = ctx->d_tri; s.nodes = ctx->d_nodes;
      built_on_device = true;
    }
  }
  if (s.n_tri && !built_on_device) {
    if (!ctx->host_bvh || ctx->host_bvh_for != ctx->b_prims || ctx->host_bvh_tuning != ctx->params.bvh_tuning ||
        ctx->host_bvh_prim_cost != ctx->params.bvh_prim_cost) {
      auto bvh = std::make_shared<crt::BvhResult>();
      const int threads = (int)std::max(1u, std::thread::hardware_concurrency());
      const auto b0 = std::chrono::steady_clock::now();
      crt::build_bvh(prims, tri_pos, threads, crt::bvh_options(ctx->params.bvh_tuning, ctx->params.bvh_prim_cost), *bvh);
      bvh->build_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - b0).count();
      ctx->host_bvh = bvh; ctx->host_bvh_for = ctx->b_prims;
      ctx->host_bvh_tuning = ctx->params.bvh_tuning; ctx->host_bvh_prim_cost = ctx->params.bvh_prim_cost;
    }
    const crt::BvhResult& bvh = *ctx->host_bvh;
    s.n_nodes = bvh.n_nodes; max_depth = bvh.max_depth; build_ms = bvh.build_ms;
    for (int a = 0; a < 3; a++) { s.bvh_lo[a] = bvh.lo[a]; s.bvh_hi[a] = bvh.hi[a]; }
    // one allocation, nodes first
    const size_t node_bytes = (bvh.nodes.size() * 4 + 255) & ~(size_t)255, tri_bytes = (size_t)s.n_tri * 48;
    CU(cudaMalloc(&ctx->d_nodes, node_bytes + tri_bytes));
    ctx->d_tri = (float4*)((uint8_t*)ctx->d_nodes + node_bytes);
    ctx->scene_bytes = node_bytes + tri_bytes;
    CU(cudaMemcpyAsync(d_src.p, bvh.tri_src.data(), bvh.tri_src.size() * 4, cudaMemcpyHostToDevice, ctx->stream));
    CU(crt::launch_repack_tri(d_prims_raw.p, d_src.as<uint32_t>(), s.n_tri, ctx->d_tri, ctx->stream));
    CU(cudaMemcpyAsync(ctx->d_nodes, bvh.nodes.data(), bvh.nodes.size() * 4, cudaMemcpyHostToDevice, ctx->stream));
    s.tri = ctx->d_tri; s.nodes = ctx->d_nodes;
  }
  if (s.n_tri && max_depth + 2 > CRT_BVH_STACK) return fail(ctx, CRT_ERR_SIZE, "BVH depth %u exceeds the traversal stack", max_depth);
  CU(d_lights_raw.alloc(n_lights * 80));
  CU(cudaMalloc(&ctx->d_lights, n_lights * sizeof(LightRec)));
  CU(cudaMemcpyAsync(d_lights_raw.p, ctx->b_lights.data(), n_lights * 80, cudaMemcpyHostToDevice, ctx->stream));
  CU(crt::launch_repack_lights(d_lights_raw.p, (uint32_t)n_lights, ctx->d_lights, ctx->stream));
  const uint32_t nl = n_lights < CRT_LIGHT_PARAM_MAX ? (uint32_t)n_lights : CRT_LIGHT_PARAM_MAX;
  CU(cudaMemcpyAsync(s.lights_c, ctx->d_lights, nl * sizeof(LightRec), cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  s.lights_g = ctx->d_lights;

  // does any light's primitive (data4.w, CS:386) live in the BVH?  then shadow rays towards it need closest-hit
  s.shadow_closest = 0; s.light_tri_mask = nullptr;
  dfree(ctx->d_light_mask);
  if (!tri_pos.empty()) {
    std::unordered_map<uint32_t, std::vector<uint32_t>> light_prims;   // primitive index -> the lights that name it
    for (size_t l = 0; l < n_lights; l++) { uint32_t want; memcpy(&want, ctx->b_lights.data() + l * 80 + 76, 4); light_prims[want].push_back((uint32_t)l); }
    std::vector<uint32_t> mask((n_lights + 31) / 32, 0u);
    for (uint32_t p : tri_pos) {
      uint32_t idx; memcpy(&idx, prims + (size_t)p * 80 + 76, 4);
      auto it = light_prims.find(idx);
      if (it == light_prims.end()) continue;
      s.shadow_closest = 1;
      for (uint32_t l : it->second) mask[l >> 5] |= 1u << (l & 31u);
    }
    if (s.shadow_closest) {
      CU(cudaMalloc(&ctx->d_light_mask, mask.size() * 4));
      CU(cudaMemcpy(ctx->d_light_mask, mask.data(), mask.size() * 4, cudaMemcpyHostToDevice));
      s.light_tri_mask = ctx->d_light_mask;
    }
  }
  const auto t1 = std::chrono::steady_clock::now();
  ctx->stats[0] = n_prims; ctx->stats[1] = s.n_lin; ctx->stats[2] = s.n_tri; ctx->stats[3] = s.n_nodes;
  ctx->stats[4] = (uint64_t)s.n_nodes * 16 * CRT_NODE_STRIDE; ctx->stats[5] = (uint64_t)s.n_tri * 48;
  ctx->stats[6] = (uint64_t)std::chrono::duration_cast<std::chrono::microseconds>(t1 - t0).count();
  ctx->stats[7] = max_depth;
  ctx->stats[8] = (uint64_t)(build_ms * 1000.0);   // the tree alone, microseconds
  ctx->stats[9] = (s.n_tri ? (built_on_device ? 2u : 1u) : 0u) | (s.shadow_closest ? 0x100u : 0u);
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

constexpr size_t kWfCounterBytes = (size_t)(CRT_WF_MAX_ITERS + 2) * sizeof(WfIter) + 64;   // iteration records + the work counter

// the per-lane wavefront state: lane 0 is the pool every form uses, lane 1 the second half pool of the two-lane device loop
struct LaneRef {
  WfState* wf; void** mem; uint32_t* slots; cudaGraph_t* graph; cudaGraphExec_t* exec; std::vector<uint8_t>* key;
};
LaneRef lane_of(crt_ctx* ctx, int h) {
  if (h == 0) return LaneRef{&ctx->wf, &ctx->wf_mem, &ctx->wf_slots, &ctx->wf_graph, &ctx->wf_graph_exec, &ctx->wf_graph_key};
  return LaneRef{&ctx->wf2, &ctx->wf2_mem, &ctx->wf2_slots, &ctx->wf2_graph, &ctx->wf2_graph_exec, &ctx->wf2_graph_key};
}

int ensure_wavefront(crt_ctx* ctx, int h, uint32_t n_slots) {
  const LaneRef L = lane_of(ctx, h);
  WfState& w = *L.wf;
  w.slow_min = h == 0 ? ctx->params.slow_ray_nodes : 0u;   // diagnostics: log rays that visit at least this many nodes (stats mode)
  if (*L.mem && *L.slots >= n_slots) return CRT_OK;
  uint32_t iters_total = 0, launches_total = 0;
  if (*L.mem) {   // keep the device-side iteration / launch counts across a pool that grows
    uint32_t tmp[4];
    CU(cudaMemcpy(tmp, w.ctl, 16, cudaMemcpyDeviceToHost));
    iters_total = tmp[2]; launches_total = tmp[3];
    cudaFree(*L.mem); *L.mem = nullptr;
  }
  // one record per slot; shade and finish queues of two iterations, the regen queue, the BVH queue (2 entries per slot)
  const size_t slow_bytes = (size_t)(1 + 3 * CRT_WF_SLOW_RAYS) * 16;
  const size_t cls_bytes = ((size_t)n_slots + 255) & ~(size_t)255;
  const size_t total = (size_t)(CRT_WF_REC_WORDS * 16 + 7 * 4) * n_slots + kWfCounterBytes + 256 + slow_bytes + cls_bytes + 256;
  CU(cudaMalloc(L.mem, total));
  uint8_t* p = (uint8_t*)*L.mem;
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
  w.cls = (uint8_t*)p; p += cls_bytes;
  w.ctl = (WfCtl*)p;
  CU(cudaMemset(w.slow, 0, slow_bytes));
  const uint32_t init[4] = {0u, 0u, iters_total, launches_total};
  CU(cudaMemset(w.ctl, 0, sizeof(WfCtl)));
  CU(cudaMemcpy(w.ctl, init, 16, cudaMemcpyHostToDevice));
  *L.slots = n_slots;
  if (!ctx->wf_host_count) CU(cudaMallocHost(&ctx->wf_host_count, 16));
  return CRT_OK;
}

void drop_wf_graph(crt_ctx* ctx, int h) {
  const LaneRef L = lane_of(ctx, h);
  // (an executable graph that is still running is released by the runtime when it completes)
  if (*L.exec) { cudaGraphExecDestroy(*L.exec); *L.exec = nullptr; }
  if (*L.graph) { cudaGraphDestroy(*L.graph); *L.graph = nullptr; }
  L.key->clear();
}

// streams and events of the two-lane device loop, created on first use
int ensure_lanes(crt_ctx* ctx) {
  if (ctx->lane_fork) return CRT_OK;
  for (int h = 0; h < 2; h++) {
    CU(cudaStreamCreateWithFlags(&ctx->lane_stream[h], cudaStreamNonBlocking));
    CU(cudaEventCreateWithFlags(&ctx->lane_join[h], cudaEventDisableTiming));
  }
  CU(cudaEventCreateWithFlags(&ctx->lane_fork, cudaEventDisableTiming));
  return CRT_OK;
}

// n frames as a wavefront, in chunks.  Iteration 0 starts a path on every slot; every later iteration is one pass of
// wf_logic over the live slots (finish list, then bounce list), wf_traverse over the BVH queue both of them filled (scenes
// with triangles) and wf_partition (class-sorted lists).  A chunk's loop ends when no slot is live; wf_accumulate then adds
// the chunk's samples to binding 1 in frame order and rebuilds binding 0.
//
// Who runs the loop: by default the device -- the iteration is a CUDA-graph WHILE node whose condition wf_next sets from
// the live counts, so the host enqueues begin / loop / accumulate and returns; crt_dispatch_frames never waits for the GPU
// and a chunk's thin last iterations cost a kernel each instead of a host round trip.  The host-driven loop of round 1
// (launch every iteration, read the live count back at check points) remains for the instrumented runs that need an event
// between launches (CRT_FLAG_KERNEL_TIMES), for the split three-kernel form, and as CRT_WF_HOST_LOOP.
int dispatch_wavefront(crt_ctx* ctx, uint32_t n, uint32_t n_units, bool stats) {
  const DevScene& s = ctx->scene;
  size_t smem = crt::wf_tables_bytes(s);
  ctx->wf.tables_in_smem = smem <= 48 * 1024 ? 1u : 0u;
  if (!ctx->wf.tables_in_smem) smem = 0;
  int bl = 0, bs = 0, br = 0, bt = 0;
  CU(crt::wf_config(stats, smem, &bl, &bs, &br, &bt));
  if (bl < 1 || bs < 1 || br < 1 || bt < 1) return fail(ctx, CRT_ERR_CUDA, "wavefront kernels do not fit on an SM");
  const int grid_l = ctx->sm_count * bl, grid_s = ctx->sm_count * bs, grid_r = ctx->sm_count * br, grid_t = ctx->sm_count * bt;
  const int grid_p = ctx->sm_count * 8;
  const int grid_tail = ctx->sm_count * 2;   // wf_tail: a few thousand slots, one thread each (grid-stride)
  const bool split = (ctx->params.flags & CRT_FLAG_WF_SPLIT) != 0;   // three-kernel form (wf_shade / wf_regen) instead of wf_logic
  const bool host_loop = split || (ctx->params.flags & CRT_FLAG_KERNEL_TIMES) || (ctx->params.wf_tuning & CRT_WF_HOST_LOOP);
  // a path lives for at most max_depth+1 extension rays plus one pass for a trailing NEE answer
  const uint64_t worst = (uint64_t)ctx->params.max_depth + 3ull;
  if (worst + 2 > CRT_WF_MAX_ITERS) return fail(ctx, CRT_ERR_ARG, "max_depth %u is too deep for the wavefront kernel; use CRT_KERNEL_MEGA", ctx->params.max_depth);
  // frames per chunk: each chunk ends in a tail of iterations with few live paths, so chunks are as long
  // as the staging buffer allows (measured: 64 -> 128 -> 256 frames per chunk = 2568 -> 2638 -> 2669 Mrays/s)
  uint32_t chunk = ctx->params.frames_per_launch ? ctx->params.frames_per_launch : 256u;
  // path slots in flight, measured per 64-spp step of mesh1m: 8M -> 114 ms, 16M -> 107 ms, 32M -> 104 ms; 64M is another 1-2 %
  const uint32_t pool = ctx->params.wf_slots ? ctx->params.wf_slots : (64u << 20);
  const size_t n_pixels = ctx->acc_pixels;
  const bool split_spp = ctx->shard_mode == CRT_SHARD_SPLIT_SPP && ctx->world > 1;
  // guards: staging (16 bytes per work item of the chunk: samples this rank renders x units it owns) <= a quarter of the
  // device's memory (45 GB of a B200's 180), work items of a chunk < 2^31
  const uint64_t stage_cap = std::max<uint64_t>(ctx->total_mem / 4, 1ull << 30);
  while (chunk > 1 && ((uint64_t)std::max(n_units, 1u) * chunk * 16ull > stage_cap || (uint64_t)n_units * chunk >= (1ull << 31))) chunk--;
  // under split-spp a rank renders every world-th frame: size the chunk by the frames this rank renders, not by the
  // frame numbers it spans (8 ranks with 64-frame chunks ran 8 samples per chunk: the pool half empty, 8x the tails)
  if (split_spp) chunk = (uint32_t)std::min<uint64_t>((uint64_t)chunk * ctx->world, 1u << 20);
  // how wf_logic's live slots reach the next iteration (crt_wavefront.cu, wf_emit): 0 one list, 1 routed lists, 2 class
  // bytes + wf_partition.  In 1 and 2 wf_logic runs over the finish list and then the bounce list, so the lanes of a warp
  // are in one phase.  Defaults from measurements: class bytes for scenes with triangles (mesh1m: 351.5 -> 341.4 ms per
  // 256-spp step; lists filled by wf_traverse lose the slot order and with it the sequential record stream: 40 % slower),
  // routed lists for linear-list scenes (reference default scene 1080p x 64 frames: 64.8 -> 55.0 ms).
  uint32_t live_mode = s.n_tri ? 2u : 1u;
  if (ctx->params.wf_tuning & CRT_WF_LIVE_MASK) live_mode = (ctx->params.wf_tuning & CRT_WF_LIVE_MASK) - 1u;
  if (split && live_mode == 0) live_mode = 1;   // the split kernels always work over routed lists: filled as rays finish (1) or by wf_partition (2)
  // routed forms: paths that end in a bounce warp are finished by the next iteration's finish warps (mesh1m: 339.8 ->
  // 334.5 ms per step; the reference default scene: 56.0 -> 58.0 ms, so linear-list scenes finish in place)
  uint32_t defer = s.n_tri ? 1u : 0u;
  if (ctx->params.wf_tuning & CRT_WF_DEFER_MASK) defer = (ctx->params.wf_tuning & CRT_WF_DEFER_ON) ? 1u : 0u;
  const uint32_t logic_mode = live_mode | ((live_mode && defer) ? 0x100u : 0u);
  const uint32_t kernels_per_iter = 1u + (s.n_tri ? 1u : 0u) + (live_mode == 2 ? 1u : 0u) + 1u;   // graph loop, with wf_next
  // the loop hands over to wf_tail when the chunk's work is used up and at most this many slots are live (crt_wavefront.cu)
  const uint32_t tail_sel = (ctx->params.wf_tuning & CRT_WF_TAIL_MASK) >> CRT_WF_TAIL_SHIFT;
  const uint32_t tail_max = split ? 0u : (tail_sel == 0 ? CRT_WF_TAIL_DEFAULT : (tail_sel == 31u ? 0u : (1u << (tail_sel - 1u))));
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
    // Two lanes (device loop only): the chunk's samples are split in two halves that run side by side, each over its own half
    // pool, control block and graph, on two streams forked from and joined into the context's stream.  Every launch of
    // wf_traverse ends in a straggler phase -- the rays with the longest chains of dependent fetches (155 nodes on mesh1m)
    // finish 0.3-0.4 ms after the bulk, with the SMs mostly idle (profiles/r02_launches_mesh1m.csv: late iterations take
    // 340-430 us in wf_traverse against 50-150 us in wf_logic) -- and a chunk ends in thin iterations and wf_tail; with two
    // lanes the other half's kernels fill the machine meanwhile.  Samples are staged by work item and summed in frame order
    // afterwards, so the split changes no bit.
    // (Measured, profiles/r02m_two_lanes_ab.txt: mesh1m 256-spp step 325.5 -> 318.4 ms, 32-frame step 43.7 -> 42.0, 8-frame 13.1 ->
    // 12.3, Cornell 53.2 -> 51.1; the 10M soup, whose 650 MB scene streams from HBM, loses 2 % with two traversals competing
    // for it -- scenes larger than L2 keep one lane.)
    const bool fits_l2 = ctx->scene_bytes <= (96ull << 20);
    const bool two_lanes = !host_loop && !(ctx->params.wf_tuning & CRT_WF_ONE_LANE) && !ctx->params.slow_ray_nodes && ck.n_samples >= 2 &&
                           ((total_work >= (4ull << 20) && fits_l2) || (ctx->params.wf_tuning & CRT_WF_TWO_LANES_ALWAYS));
    const int n_parts = two_lanes ? 2 : 1;
    const uint64_t lane_pool = two_lanes ? std::max<uint64_t>(pool / 2, 32) : pool;
    WfChunk part[2] = {ck, ck};
    if (two_lanes) {
      part[0].n_samples = (ck.n_samples + 1u) / 2u;
      part[1].n_samples = ck.n_samples - part[0].n_samples;
      part[1].s_first = ck.s_first + part[0].n_samples * ck.s_stride;
    }
    // slots per part: all of its work items when they are few, the pool when they are many, and in between half of the
    // items -- a chunk of one to two pools' worth (what a rank renders per step when eight GPUs share 256 spp) is 3 % faster
    // when every slot regenerates once (a finish pass shared with the next path's start) than with one path per slot
    // (profiles/r02g_pool_coherence_ab.txt: 32-frame step 44.5 -> 43.2 ms).  An explicit crt_params.wf_slots is taken as given.
    uint64_t iters_worst = 0;
    for (int h = 0; h < n_parts; h++) {
      const uint64_t work_h = (uint64_t)n_units * part[h].n_samples;
      uint64_t want_slots = std::min<uint64_t>(std::max<uint64_t>(work_h, 32), lane_pool);
      if (!ctx->params.wf_slots && work_h > (16ull << 20)) want_slots = std::min<uint64_t>(std::max<uint64_t>(work_h / 2, 16ull << 20), lane_pool);
      part[h].n_slots = (uint32_t)want_slots;
      // worst case: every slot runs ceil(work/slots) paths of `worst` iterations one after the other
      iters_worst = std::max<uint64_t>(iters_worst, worst * ((work_h + part[h].n_slots - 1) / part[h].n_slots + 1) + 2);
    }
    ck.n_slots = part[0].n_slots;
    if (iters_worst > CRT_WF_MAX_ITERS && k > 1) {   // shrink the chunk so the per-iteration counters cannot run out
      chunk = std::max(1u, k / 2);
      continue;
    }
    const size_t stage_bytes = (size_t)std::max(1u, ck.n_samples) * std::max(n_units, 1u) * 16;
    if (stage_bytes > ctx->wf_stage_bytes) {
      if (ctx->wf_stage) { cudaFree(ctx->wf_stage); ctx->wf_stage = nullptr; ctx->wf_stage_bytes = 0; }
      CU(cudaMalloc(&ctx->wf_stage, stage_bytes));
      ctx->wf_stage_bytes = stage_bytes;
    }
    ck.stage = (float4*)ctx->wf_stage;
    part[0].stage = ck.stage;
    part[1].stage = ck.stage + (size_t)part[0].n_samples * n_units;   // staged by (sample ordinal, unit): the second half follows the first
    if (ck.n_samples && n_units) {
      if (!host_loop) {
        // ---- the loop runs on the device
        if (two_lanes) {
          int rc = ensure_lanes(ctx);
          if (rc) return rc;
          CU(cudaEventRecord(ctx->lane_fork, ctx->stream));
        }
        for (int h = 0; h < n_parts; h++) {
          const LaneRef L = lane_of(ctx, h);
          cudaStream_t st = two_lanes ? ctx->lane_stream[h] : ctx->stream;
          int rc = ensure_wavefront(ctx, h, part[h].n_slots);
          if (rc) return rc;
          WfState& w = *L.wf;
          w.tables_in_smem = ctx->wf.tables_in_smem;
          if (two_lanes) CU(cudaStreamWaitEvent(st, ctx->lane_fork, 0));
          CU(cudaMemsetAsync(w.iter, 0, kWfCounterBytes, st));
          CU(crt::launch_wf_begin(w, part[h], st));
          ctx->launches++;
          struct Key { DevScene sc; WfState w; uint32_t stats, logic_mode, live_mode, tail_max; int gl, gt, gp; size_t smem; void* counters; };
          static_assert(std::is_trivially_copyable<Key>::value, "graph key");
          std::vector<uint8_t> key(sizeof(Key), 0);
          {
            Key kk; memset((void*)&kk, 0, sizeof(kk));
            kk.sc = s; kk.w = w; kk.stats = stats; kk.logic_mode = logic_mode; kk.live_mode = live_mode; kk.tail_max = tail_max;
            kk.gl = grid_l; kk.gt = grid_t; kk.gp = grid_p; kk.smem = smem; kk.counters = ctx->d_counters;
            memcpy(key.data(), &kk, sizeof(kk));
          }
          if (!*L.exec || key != *L.key) {
            drop_wf_graph(ctx, h);
            CU(crt::wf_graph_build(s, w, stats, logic_mode, live_mode, grid_l, grid_t, grid_p, smem, ctx->d_counters, kernels_per_iter, tail_max,
                                   L.graph, L.exec));
            *L.key = key;
          }
          CU(cudaGraphLaunch(*L.exec, st));
          if (tail_max) { CU(crt::launch_wf_tail(s, w, stats, live_mode, grid_tail, smem, CRT_WF_IT_DEVICE, ctx->d_counters, st)); ctx->launches++; }
          if (two_lanes) {
            CU(cudaEventRecord(ctx->lane_join[h], st));
            CU(cudaStreamWaitEvent(ctx->stream, ctx->lane_join[h], 0));
          }
        }
      } else {
        // ---- the host launches every iteration
        int rc = ensure_wavefront(ctx, 0, ck.n_slots);
        if (rc) return rc;
        CU(cudaMemsetAsync(ctx->wf.iter, 0, kWfCounterBytes, ctx->stream));
        CU(crt::launch_wf_begin(ctx->wf, ck, ctx->stream));
        ctx->launches++;
        uint32_t it = 0;
        uint32_t next_check = (ctx->wf_last_iters > 4 && ctx->wf_last_work == total_work) ? ctx->wf_last_iters - 1 : (uint32_t)(2 * (total_work / ck.n_slots) + 2);
        for (;;) {
          if (!split) {
            CU(mark(ctx, CRT_KC_WF_LOGIC));
            CU(crt::launch_wf_logic(s, ctx->wf, stats, logic_mode, grid_l, smem, it, ctx->d_counters, ctx->stream));
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
            CU(crt::launch_wf_partition(ctx->wf, grid_p, it, ctx->stream));
            ctx->launches++;
          }
          it++;
          if (it >= next_check) {
            CU(mark(ctx, -1));
            CU(cudaMemcpyAsync(ctx->wf_host_count, &ctx->wf.iter[it], 8, cudaMemcpyDeviceToHost, ctx->stream));   // n_bounce, n_finish
            CU(cudaMemcpyAsync(ctx->wf_host_count + 2, ctx->wf.work, 4, cudaMemcpyDeviceToHost, ctx->stream));
            CU(cudaStreamSynchronize(ctx->stream));
            const uint64_t live = (uint64_t)ctx->wf_host_count[0] + ctx->wf_host_count[1];
            if (live == 0) break;
            if (live <= tail_max && ctx->wf_host_count[2] >= total_work) {   // only a tail is left: wf_tail finishes it
              CU(mark(ctx, CRT_KC_WF_TAIL));
              CU(crt::launch_wf_tail(s, ctx->wf, stats, live_mode, grid_tail, smem, it, ctx->d_counters, ctx->stream));
              CU(mark(ctx, -1));
              ctx->launches++;
              break;
            }
            next_check = it + (live <= 4ull * tail_max ? 1 : 4);
          }
          if (it + 2 >= CRT_WF_MAX_ITERS) return fail(ctx, CRT_ERR_STATE, "wavefront iteration limit reached");
        }
        ctx->wf_last_iters = it;
        ctx->wf_last_work = total_work;
        ctx->wf_iterations += it;
      }
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

uint32_t owned_tiles(const crt_ctx* ctx, uint32_t rank, uint32_t world) {
  const uint32_t n_tiles = ctx->scene.tiles_x * ctx->scene.tiles_y;
  return n_tiles > rank ? (n_tiles - rank + world - 1) / world : 0;
}

int ensure_comb(crt_ctx* ctx) {
  if (!ctx->d_comb) CU(cudaMalloc(&ctx->d_comb, ctx->acc_pixels * sizeof(float4)));
  if (!ctx->ev_c0) { CU(cudaEventCreate(&ctx->ev_c0)); CU(cudaEventCreate(&ctx->ev_c1)); }
  return CRT_OK;
}

// ---- single-process group (crt_create_multi): helpers
int group_fail(crt_ctx* g, crt_ctx* m, int rc) {
  if (rc) g->err = m->err;
  return rc;
}

// scene upload of every member: member 0 first (it builds a host tree once, the others share it), then the rest in parallel
int group_prepare(crt_ctx* g) {
  crt_ctx* m0 = g->members[0];
  int rc = cudaSetDevice(m0->device) == cudaSuccess ? prepare(m0) : CRT_ERR_CUDA;
  if (rc) return group_fail(g, m0, rc);
  std::vector<std::thread> th;
  std::vector<int> rcs(g->members.size(), 0);
  for (size_t i = 1; i < g->members.size(); i++) {
    crt_ctx* m = g->members[i];
    m->host_bvh = m0->host_bvh; m->host_bvh_for = m0->host_bvh_for; m->host_bvh_tuning = m0->host_bvh_tuning; m->host_bvh_prim_cost = m0->host_bvh_prim_cost;
    th.emplace_back([m, i, &rcs] { rcs[i] = cudaSetDevice(m->device) == cudaSuccess ? prepare(m) : CRT_ERR_CUDA; });
  }
  for (auto& t : th) t.join();
  for (size_t i = 1; i < g->members.size(); i++)
    if (rcs[i]) return group_fail(g, g->members[i], rcs[i]);
  return CRT_OK;
}

int group_combine(crt_ctx* g) {
  crt_ctx* ctx = g;   // for CU()
  crt_ctx* m0 = g->members[0];
  int rc = group_prepare(g);
  if (rc) return rc;
  CU(cudaSetDevice(m0->device));
  if ((rc = ensure_comb(m0))) return group_fail(g, m0, rc);
  CombinePeers peers;
  memset(&peers, 0, sizeof(peers));
  peers.n = (uint32_t)g->members.size();
  peers.tiles = g->shard_mode == CRT_SHARD_TILES ? 1u : 0u;
  for (size_t i = 0; i < g->members.size(); i++) {
    crt_ctx* m = g->members[i];
    if (m->acc_pixels != m0->acc_pixels || m->sample != m0->sample) return fail(g, CRT_ERR_STATE, "group members are out of step");
    peers.acc[i] = m->d_acc;
    if (i) {   // member 0's stream waits for the member's frames
      CU(cudaSetDevice(m->device));
      if (!m->ev_sync) CU(cudaEventCreateWithFlags(&m->ev_sync, cudaEventDisableTiming));
      CU(cudaEventRecord(m->ev_sync, m->stream));
      CU(cudaSetDevice(m0->device));
      CU(cudaStreamWaitEvent(m0->stream, m->ev_sync, 0));
    }
  }
  CU(cudaEventRecord(m0->ev_c0, m0->stream));
  CU(crt::launch_combine_peers(peers, m0->d_comb, m0->d_fb, m0->W, m0->H, m0->scene.tiles_x, m0->sample, m0->params.exposure, m0->stream));
  CU(cudaEventRecord(m0->ev_c1, m0->stream));
  m0->comb_timed = true;
  m0->comb_valid = true;
  // bytes that crossed the links: every other member's share
  const uint64_t full = (uint64_t)m0->acc_pixels * 16;
  m0->comb_bytes = peers.tiles ? full - (uint64_t)owned_tiles(m0, 0, peers.n) * 32 * 16 : full * (peers.n - 1);
  for (size_t i = 1; i < g->members.size(); i++) {   // the members' next frames wait until their accumulators have been read
    crt_ctx* m = g->members[i];
    CU(cudaSetDevice(m->device));
    CU(cudaStreamWaitEvent(m->stream, m0->ev_c1, 0));
  }
  CU(cudaSetDevice(m0->device));
  g->comb_valid = true;
  return CRT_OK;
}

}  // namespace

// ------------------------------------------------------------------------------------------------ crt_trace_rays_noalloc
// Scratch of the allocation-free trace_rays variant: the device-side staging that crt_trace_rays takes from cudaMalloc, as a
// file-scope array of a fixed, modest size -- CRT_NOALLOC_MAX_RAYS x (32 B ray + 4 B exclude + 4 B hit index + 4 B t) = 2.75 MiB,
// held by every GPU that loads the library until the process exits.  In host code the name below is only a stand-in for the
// symbol, not a device address: the entry point resolves it with cudaGetSymbolAddress once per device (noalloc_scratch_of) and
// hands the kernel that pointer.
static __device__ __align__(256) unsigned char g_noalloc_scratch[(size_t)CRT_NOALLOC_MAX_RAYS * 44u];

namespace {
constexpr size_t kNoallocRays = 0;                                            // float4[2 n_max]
constexpr size_t kNoallocExcl = (size_t)CRT_NOALLOC_MAX_RAYS * 32u;           // uint32[n_max]
constexpr size_t kNoallocIdx = kNoallocExcl + (size_t)CRT_NOALLOC_MAX_RAYS * 4u;
constexpr size_t kNoallocT = kNoallocIdx + (size_t)CRT_NOALLOC_MAX_RAYS * 4u;
constexpr int kNoallocMaxDevices = 64;
std::mutex g_noalloc_mutex;                       // calls share the array: within a process they are serialised here
void* g_noalloc_ptr[kNoallocMaxDevices] = {};     // device address of g_noalloc_scratch on each device, resolved on first use

// `device` must be current.  No allocation: the array exists as soon as the module is loaded on the device.
cudaError_t noalloc_scratch_of(int device, unsigned char** out) {
  if (device < 0 || device >= kNoallocMaxDevices) return cudaErrorInvalidDevice;
  if (!g_noalloc_ptr[device]) {
    void* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
    if (e != cudaSuccess) return e;
    g_noalloc_ptr[device] = p;
  }
  *out = (unsigned char*)g_noalloc_ptr[device];
  return cudaSuccess;
}
}  // namespace

// trace_rays_kernel (crt_kernels.cu) with inputs and outputs in the scratch array: intersect() (CS:503-518) per ray, the same
// device code, so hit index and t are those of crt_trace_rays bit for bit.
__global__ void trace_rays_scratch_kernel(const __grid_constant__ DevScene sc, unsigned char* __restrict__ scratch, uint32_t n, uint32_t has_exclude,
                                          unsigned long long* __restrict__ counters) {
  const float4* rays = reinterpret_cast<const float4*>(scratch + kNoallocRays);
  const uint32_t* exclude = reinterpret_cast<const uint32_t*>(scratch + kNoallocExcl);
  uint32_t* hit_idx = reinterpret_cast<uint32_t*>(scratch + kNoallocIdx);
  float* t_out = reinterpret_cast<float*>(scratch + kNoallocT);
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  Stats st; st.prim_tests = st.node_visits = st.tri_tests = st.rand_calls = 0;
  if (i < n) {
    const float4 o = rays[2 * i], d = rays[2 * i + 1];
    const Hit h = intersect<true>(sc, v3(o), v3(d), has_exclude ? exclude[i] : CRT_MAX_U32, st);
    hit_idx[i] = h.idx;
    t_out[i] = h.t;
  }
  unsigned long long v[3] = {st.prim_tests, st.node_visits, st.tri_tests};
  for (int k = 0; k < 3; k++) {
    unsigned long long x = v[k];
    for (int o = 16; o > 0; o >>= 1) x += __shfl_down_sync(0xffffffffu, x, o);
    if ((threadIdx.x & 31u) == 0 && x) atomicAdd(&counters[CRT_CNT_PRIM_TESTS + k], x);
  }
}

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
  memset((void*)&ctx->scene, 0, sizeof(ctx->scene));
  memset((void*)&ctx->wf, 0, sizeof(ctx->wf));
  memset((void*)&ctx->wf2, 0, sizeof(ctx->wf2));
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

int crt_create_multi(const int* devices, int n, uint32_t shard_mode, crt_ctx** out) {
  if (!out) return CRT_ERR_ARG;
  *out = nullptr;
  if (!devices || n < 1 || n > CRT_MAX_GROUP || (shard_mode != CRT_SHARD_TILES && shard_mode != CRT_SHARD_SPLIT_SPP)) return CRT_ERR_ARG;
  crt_ctx* g = new (std::nothrow) crt_ctx();
  if (!g) return CRT_ERR_OOM;
  crt_default_params(&g->params);
  g->world = (uint32_t)n; g->shard_mode = shard_mode; g->device = devices[0];
  for (int i = 0; i < n; i++) {
    crt_ctx* m = nullptr;
    const int rc = crt_create(devices[i], &m);
    if (rc) { crt_destroy(g); return rc; }
    m->group = g;
    m->rank = (uint32_t)i; m->world = (uint32_t)n; m->shard_mode = n > 1 ? shard_mode : CRT_SHARD_NONE;
    g->members.push_back(m);
  }
  // member 0 reads the other members' accumulators through peer pointers (NVLink / NVSwitch)
  for (int i = 1; i < n; i++) {
    if (devices[i] == devices[0]) continue;
    int can = 0;
    if (cudaDeviceCanAccessPeer(&can, devices[0], devices[i]) != cudaSuccess || !can) { crt_destroy(g); return CRT_ERR_NODEVICE; }
    cudaSetDevice(devices[0]);
    const cudaError_t e = cudaDeviceEnablePeerAccess(devices[i], 0);
    if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) { crt_destroy(g); return CRT_ERR_CUDA; }
    cudaGetLastError();
  }
  *out = g;
  return CRT_OK;
}

void crt_destroy(crt_ctx* ctx) {
  if (!ctx) return;
  if (is_group(ctx)) {
    for (crt_ctx* m : ctx->members) crt_destroy(m);
    delete ctx;
    return;
  }
  cudaSetDevice(ctx->device);
  if (ctx->stream) cudaStreamSynchronize(ctx->stream);
  if (ctx->nccl_comm) { if (NcclApi* nc = nccl_api()) nc->CommDestroy(ctx->nccl_comm); ctx->nccl_comm = nullptr; }
  drop_wf_graph(ctx, 0); drop_wf_graph(ctx, 1);
  for (int h = 0; h < 2; h++) {
    if (ctx->lane_stream[h]) cudaStreamDestroy(ctx->lane_stream[h]);
    if (ctx->lane_join[h]) cudaEventDestroy(ctx->lane_join[h]);
  }
  if (ctx->lane_fork) cudaEventDestroy(ctx->lane_fork);
  if (ctx->wf2_mem) cudaFree(ctx->wf2_mem);
  dfree(ctx->d_acc); dfree(ctx->d_fb); dfree(ctx->d_tea); dfree(ctx->d_cam); dfree(ctx->d_spectra); dfree(ctx->d_cie);
  dfree(ctx->d_lin); dfree(ctx->d_lights); dfree(ctx->d_light_mask); dfree(ctx->d_nodes); dfree(ctx->d_work); dfree(ctx->d_counters);
  dfree(ctx->d_comb); dfree(ctx->d_xfer);
  if (ctx->wf_mem) cudaFree(ctx->wf_mem);
  if (ctx->wf_stage) cudaFree(ctx->wf_stage);
  if (ctx->wf_host_count) cudaFreeHost(ctx->wf_host_count);
  if (ctx->ev_start) cudaEventDestroy(ctx->ev_start);
  if (ctx->ev_stop) cudaEventDestroy(ctx->ev_stop);
  if (ctx->ev_c0) cudaEventDestroy(ctx->ev_c0);
  if (ctx->ev_c1) cudaEventDestroy(ctx->ev_c1);
  if (ctx->ev_sync) cudaEventDestroy(ctx->ev_sync);
  for (auto& e : ctx->ev_seq) cudaEventDestroy(e.first);
  for (auto& e : ctx->ev_free) cudaEventDestroy(e);
  if (ctx->own_stream) cudaStreamDestroy(ctx->own_stream);
  delete ctx;
}

const char* crt_last_error(crt_ctx* ctx) { return ctx ? ctx->err.c_str() : "null context"; }

int crt_set_binding(crt_ctx* ctx, uint32_t binding, const void* bytes, size_t nbytes) {
  if (!ctx) return CRT_ERR_ARG;
  if (!bytes && nbytes) return fail(ctx, CRT_ERR_ARG, "null bytes");
  const uint8_t* p = (const uint8_t*)bytes;
  if (is_group(ctx)) {
    ctx->comb_valid = false;
    if (binding == 8) {   // one host copy for all members
      if (nbytes % 80) return fail(ctx, CRT_ERR_SIZE, "binding 8 (primitives) has an 80-byte stride, got %zu bytes", nbytes);
      auto sp = std::make_shared<const std::vector<uint8_t>>(p, p + nbytes);
      for (crt_ctx* m : ctx->members) { m->b_prims = sp; m->have[8] = true; m->scene_dirty = true; }
      return CRT_OK;
    }
    if (binding == 1 && ctx->shard_mode == CRT_SHARD_SPLIT_SPP) {
      // resume: the saved sums go to member 0, the other members' partial sums restart at zero (their sum is the image)
      int rc = group_fail(ctx, ctx->members[0], crt_set_binding(ctx->members[0], 1, bytes, nbytes));
      for (size_t i = 1; i < ctx->members.size() && !rc; i++) {
        crt_ctx* m = ctx->members[i];
        m->pending_acc.clear();
        if (m->d_acc && cudaSetDevice(m->device) == cudaSuccess) cudaMemsetAsync(m->d_acc, 0, m->acc_pixels * sizeof(float4), m->stream);
        m->comb_valid = false;
      }
      return rc;
    }
    for (crt_ctx* m : ctx->members) {   // (tiles: every member takes the whole image and keeps adding to the tiles it owns)
      const int rc = crt_set_binding(m, binding, bytes, nbytes);
      if (rc) return group_fail(ctx, m, rc);
    }
    return CRT_OK;
  }
  CU(cudaSetDevice(ctx->device));
  switch (binding) {
    case 1: {   // resume: restore a saved accumulator
      ctx->comb_valid = false;
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
      if (ctx->have[4] && !ctx->camera_dirty && ctx->b_camera.size() == 64 && memcmp(ctx->b_camera.data(), p, 64) == 0) return CRT_OK;   // unchanged
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
      ctx->b_prims = std::make_shared<const std::vector<uint8_t>>(p, p + nbytes); ctx->have[8] = true; ctx->scene_dirty = true;
      return CRT_OK;
    default:
      return fail(ctx, CRT_ERR_BINDING, "binding %u cannot be set", binding);
  }
}

int crt_reset(crt_ctx* ctx) {
  if (!ctx) return CRT_ERR_ARG;
  ctx->comb_valid = false;
  if (is_group(ctx)) {
    for (crt_ctx* m : ctx->members) { const int rc = crt_reset(m); if (rc) return group_fail(ctx, m, rc); }
    return CRT_OK;
  }
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
  if ((p->wf_tuning & CRT_WF_DEFER_MASK) == CRT_WF_DEFER_MASK || (p->bvh_tuning & CRT_BVH_BUILDER_MASK) == CRT_BVH_BUILDER_MASK)
    return fail(ctx, CRT_ERR_ARG, "contradictory tuning bits");
  if (is_group(ctx)) {
    for (crt_ctx* m : ctx->members) { const int rc = crt_set_params(m, p); if (rc) return group_fail(ctx, m, rc); }
    ctx->params = *p;
    return CRT_OK;
  }
  if (p->bvh_tuning != ctx->params.bvh_tuning || p->bvh_prim_cost != ctx->params.bvh_prim_cost) ctx->scene_dirty = true;   // another tree
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
  if (is_group(ctx) || ctx->group) return fail(ctx, CRT_ERR_STATE, "the sharding of a group is fixed by crt_create_multi");
  if (ctx->nccl_comm) return fail(ctx, CRT_ERR_STATE, "the sharding of a communicator is fixed by crt_comm_init");
  if (world == 0 || rank >= world || mode > CRT_SHARD_SPLIT_SPP) return fail(ctx, CRT_ERR_ARG, "bad shard (%u of %u, mode %u)", rank, world, mode);
  ctx->rank = rank; ctx->world = world; ctx->shard_mode = world > 1 ? mode : CRT_SHARD_NONE;
  ctx->comb_valid = false;
  return CRT_OK;
}

int crt_dispatch_frames(crt_ctx* ctx, uint32_t n) {
  if (!ctx) return CRT_ERR_ARG;
  if (is_group(ctx)) {
    int rc = group_prepare(ctx);
    if (rc) return rc;
    if (n) ctx->comb_valid = false;
    // one host thread per member: a member's dispatch only enqueues work (the iteration loop runs on its device), but the
    // instrumented host loop blocks, and either way the members' queues fill side by side
    std::vector<std::thread> th;
    std::vector<int> rcs(ctx->members.size(), 0);
    for (size_t i = 0; i < ctx->members.size(); i++) {
      crt_ctx* m = ctx->members[i];
      if (ctx->members.size() == 1) { rcs[0] = crt_dispatch_frames(m, n); break; }
      th.emplace_back([m, i, n, &rcs] { rcs[i] = crt_dispatch_frames(m, n); });
    }
    for (auto& t : th) t.join();
    for (size_t i = 0; i < ctx->members.size(); i++)
      if (rcs[i]) return group_fail(ctx, ctx->members[i], rcs[i]);
    ctx->sample = ctx->members[0]->sample;
    return CRT_OK;
  }
  CU(cudaSetDevice(ctx->device));
  int rc = prepare(ctx);
  if (rc) return rc;
  if (n > 0xFFFFFFFFu - ctx->sample) return fail(ctx, CRT_ERR_ARG, "frame counter would overflow");
  if (n) ctx->comb_valid = false;
  const DevScene& s = ctx->scene;
  const bool stats = (ctx->params.flags & CRT_FLAG_STATS) != 0;
  uint32_t n_tiles = s.tiles_x * s.tiles_y;
  if (ctx->shard_mode == CRT_SHARD_TILES && ctx->world > 1) n_tiles = owned_tiles(ctx, ctx->rank, ctx->world);
  const uint32_t n_units = n_tiles * 32u;
  // AUTO: the wavefront form, whatever the scene and however many frames.  Round 1 kept the megakernel for single-frame
  // dispatches of linear-list scenes (the reference's frame() on its own scene: 2.2 ms against 2.8 ms, because the wavefront
  // paid ~2 ms of thin iterations for the paths that loop to max_depth inside the glass sphere); wf_tail removed those, and the
  // same dispatch now takes 1.57 ms in the wavefront form against 2.25 ms in the megakernel, 2.75 against 4.49 ms for two frames
  // (profiles/r02p_single_frame_cornell.txt).  The megakernel remains for max_depth > 1000 (the wavefront's per-iteration
  // counters) and by name (CRT_KERNEL_MEGA).  Both forms give the same bits.
  uint32_t kernel = ctx->params.kernel;
  if (kernel == CRT_KERNEL_AUTO) kernel = ctx->params.max_depth <= 1000 ? CRT_KERNEL_WAVEFRONT : CRT_KERNEL_MEGA;
  // (a light that is itself a triangle needs closest-hit shadow rays: both forms carry that path -- crt_device.cuh
  // shadow_intersect, crt_wavefront.cu NeeQuery::tri)
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
  if (is_group(ctx)) {
    for (crt_ctx* m : ctx->members) { const int rc = crt_synchronize(m); if (rc) return group_fail(ctx, m, rc); }
    return CRT_OK;
  }
  CU(cudaSetDevice(ctx->device));
  CU(cudaStreamSynchronize(ctx->stream));
  return CRT_OK;
}

int crt_last_dispatch_ms(crt_ctx* ctx, float* ms) {
  if (!ctx || !ms) return CRT_ERR_ARG;
  if (is_group(ctx)) {
    float mx = 0.0f;
    for (crt_ctx* m : ctx->members) { float v = 0.0f; const int rc = crt_last_dispatch_ms(m, &v); if (rc) return group_fail(ctx, m, rc); mx = std::max(mx, v); }
    *ms = mx;
    return CRT_OK;
  }
  if (!ctx->timed) return fail(ctx, CRT_ERR_STATE, "no dispatch has been timed yet");
  CU(cudaSetDevice(ctx->device));
  CU(cudaEventSynchronize(ctx->ev_stop));
  CU(cudaEventElapsedTime(ms, ctx->ev_start, ctx->ev_stop));
  return CRT_OK;
}

int crt_read_binding(crt_ctx* ctx, uint32_t binding, void* dst, size_t nbytes) {
  if (!ctx || !dst) return CRT_ERR_ARG;
  if (binding > 2) return fail(ctx, CRT_ERR_BINDING, "binding %u cannot be read back", binding);
  if (is_group(ctx)) {
    if (binding != 2 && !ctx->comb_valid) { const int rc = crt_combine(ctx, 0); if (rc) return rc; }
    if (binding == 2) { const int rc = crt_synchronize(ctx); if (rc) return rc; }
    return group_fail(ctx, ctx->members[0], crt_read_binding(ctx->members[0], binding, dst, nbytes));
  }
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
  const void* src = binding == 0 ? (const void*)ctx->d_fb : (ctx->comb_valid ? (const void*)ctx->d_comb : (const void*)ctx->d_acc);
  CU(cudaMemcpyAsync(dst, src, nbytes, cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  return CRT_OK;
}

int crt_resolve_framebuffer(crt_ctx* ctx) {
  if (!ctx) return CRT_ERR_ARG;
  if (is_group(ctx)) return group_fail(ctx, ctx->members[0], crt_resolve_framebuffer(ctx->members[0]));
  CU(cudaSetDevice(ctx->device));
  if (!ctx->d_acc) return fail(ctx, CRT_ERR_STATE, "no frame buffers yet");
  CU(crt::launch_resolve(ctx->comb_valid ? ctx->d_comb : ctx->d_acc, ctx->d_fb, (uint32_t)ctx->acc_pixels, ctx->sample, ctx->params.exposure, ctx->stream));
  return CRT_OK;
}

int crt_device_ptr(crt_ctx* ctx, uint32_t binding, void** dptr, size_t* nbytes) {
  if (!ctx || !dptr) return CRT_ERR_ARG;
  if (is_group(ctx)) return group_fail(ctx, ctx->members[0], crt_device_ptr(ctx->members[0], binding, dptr, nbytes));
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
  if (is_group(ctx)) return fail(ctx, CRT_ERR_STATE, "a group launches on its members' own streams");
  CU(cudaSetDevice(ctx->device));
  CU(cudaStreamSynchronize(ctx->stream));
  ctx->stream = cuda_stream ? (cudaStream_t)cuda_stream : ctx->own_stream;
  return CRT_OK;
}

int crt_stream(crt_ctx* ctx, void** cuda_stream) {
  if (!ctx || !cuda_stream) return CRT_ERR_ARG;
  if (is_group(ctx)) return crt_stream(ctx->members[0], cuda_stream);
  *cuda_stream = (void*)ctx->stream;
  return CRT_OK;
}

// ------------------------------------------------------------------------------------------------ multi-GPU combine
int crt_comm_unique_id(void* id_bytes, size_t nbytes) {
  if (!id_bytes || nbytes < CRT_COMM_ID_BYTES) return CRT_ERR_ARG;
  NcclApi* nc = nccl_api();
  if (!nc) return CRT_ERR_NCCL;
  NcclId id;
  if (nc->GetUniqueId(&id) != 0) return CRT_ERR_NCCL;
  memcpy(id_bytes, id.internal, CRT_COMM_ID_BYTES);
  return CRT_OK;
}

int crt_comm_init(crt_ctx* ctx, const void* id_bytes, size_t nbytes, uint32_t rank, uint32_t world, uint32_t shard_mode) {
  if (!ctx || !id_bytes) return CRT_ERR_ARG;
  if (is_group(ctx) || ctx->group) return fail(ctx, CRT_ERR_STATE, "a group combines through peer pointers and needs no communicator");
  if (nbytes < CRT_COMM_ID_BYTES || world == 0 || rank >= world || (shard_mode != CRT_SHARD_TILES && shard_mode != CRT_SHARD_SPLIT_SPP))
    return fail(ctx, CRT_ERR_ARG, "bad communicator arguments (rank %u of %u, mode %u, id of %zu bytes)", rank, world, shard_mode, nbytes);
  NcclApi* nc = nccl_api();
  if (!nc) return fail(ctx, CRT_ERR_NCCL, "NCCL is not available: libnccl.so.2 could not be loaded");
  CU(cudaSetDevice(ctx->device));
  if (ctx->nccl_comm) { nc->CommDestroy(ctx->nccl_comm); ctx->nccl_comm = nullptr; }
  NcclId id;
  memcpy(id.internal, id_bytes, CRT_COMM_ID_BYTES);
  NC(nc->CommInitRank(&ctx->nccl_comm, (int)world, id, (int)rank));
  ctx->rank = rank; ctx->world = world; ctx->shard_mode = world > 1 ? shard_mode : CRT_SHARD_NONE;
  ctx->comb_valid = false;
  return CRT_OK;
}

int crt_comm_destroy(crt_ctx* ctx) {
  if (!ctx) return CRT_ERR_ARG;
  if (!ctx->nccl_comm) return CRT_OK;
  NcclApi* nc = nccl_api();
  CU(cudaSetDevice(ctx->device));
  CU(cudaStreamSynchronize(ctx->stream));
  if (nc) nc->CommDestroy(ctx->nccl_comm);
  ctx->nccl_comm = nullptr;
  return CRT_OK;
}

int crt_combine(crt_ctx* ctx, uint32_t dst) {
  if (!ctx) return CRT_ERR_ARG;
  if (is_group(ctx)) {
    if (dst != 0) return fail(ctx, CRT_ERR_ARG, "a group combines on member 0");
    return group_combine(ctx);
  }
  if (ctx->group) return fail(ctx, CRT_ERR_STATE, "combine the group, not its member");
  CU(cudaSetDevice(ctx->device));
  int rc = prepare(ctx);
  if (rc) return rc;
  if ((rc = ensure_comb(ctx))) return rc;
  const size_t n_pix = ctx->acc_pixels;
  if (dst >= ctx->world) return fail(ctx, CRT_ERR_ARG, "destination rank %u of %u", dst, ctx->world);
  if (ctx->world <= 1) {   // nothing to exchange: the combined image is the accumulator
    ctx->comb_valid = false; ctx->comb_bytes = 0; ctx->comb_timed = false;
    return crt_resolve_framebuffer(ctx);
  }
  if (!ctx->nccl_comm) return fail(ctx, CRT_ERR_STATE, "no communicator: call crt_comm_init first");
  NcclApi* nc = nccl_api();
  if (!nc) return fail(ctx, CRT_ERR_NCCL, "NCCL is not available");
  const DevScene& s = ctx->scene;
  CU(cudaEventRecord(ctx->ev_c0, ctx->stream));
  if (ctx->shard_mode == CRT_SHARD_SPLIT_SPP) {
    // the ranks' partial sums -> their sum in dst's combined buffer (the partials stay partial: repeatable)
    NC(nc->Reduce(ctx->d_acc, ctx->d_comb, n_pix * 4, kNcclFloat32, kNcclSum, (int)dst, ctx->nccl_comm, ctx->stream));
    ctx->comb_bytes = ctx->rank == dst ? 0 : (uint64_t)n_pix * 16;
  } else {
    // tiles: every rank sends the tiles it owns, and only those -- W*H*16/world bytes
    const uint32_t my_units = owned_tiles(ctx, ctx->rank, ctx->world) * 32u;
    if (ctx->rank != dst) {
      const size_t need = (size_t)std::max(my_units, 1u) * 16;
      if (need > ctx->xfer_bytes) { dfree(ctx->d_xfer); CU(cudaMalloc(&ctx->d_xfer, need)); ctx->xfer_bytes = need; }
      CU(crt::launch_pack_tiles(ctx->d_acc, ctx->d_xfer, my_units, ctx->rank, ctx->world, s.tiles_x, s.W, s.H, ctx->stream));
      NC(nc->Send(ctx->d_xfer, (size_t)my_units * 4, kNcclFloat32, (int)dst, ctx->nccl_comm, ctx->stream));
      ctx->comb_bytes = (uint64_t)my_units * 16;
    } else {
      size_t total_units = 0;
      for (uint32_t r = 0; r < ctx->world; r++) if (r != dst) total_units += (size_t)owned_tiles(ctx, r, ctx->world) * 32u;
      const size_t need = std::max<size_t>(total_units, 1) * 16;
      if (need > ctx->xfer_bytes) { dfree(ctx->d_xfer); CU(cudaMalloc(&ctx->d_xfer, need)); ctx->xfer_bytes = need; }
      CU(cudaMemcpyAsync(ctx->d_comb, ctx->d_acc, n_pix * 16, cudaMemcpyDeviceToDevice, ctx->stream));   // dst's own tiles
      NC(nc->GroupStart());
      size_t off = 0;
      for (uint32_t r = 0; r < ctx->world; r++) {
        if (r == dst) continue;
        const size_t u = (size_t)owned_tiles(ctx, r, ctx->world) * 32u;
        const int e = nc->Recv(ctx->d_xfer + off, u * 4, kNcclFloat32, (int)r, ctx->nccl_comm, ctx->stream);
        if (e != 0) { nc->GroupEnd(); return fail(ctx, CRT_ERR_NCCL, "ncclRecv: %s", nc->GetErrorString(e)); }
        off += u;
      }
      NC(nc->GroupEnd());
      off = 0;
      for (uint32_t r = 0; r < ctx->world; r++) {
        if (r == dst) continue;
        const uint32_t u = owned_tiles(ctx, r, ctx->world) * 32u;
        CU(crt::launch_unpack_tiles(ctx->d_xfer + off, ctx->d_comb, u, r, ctx->world, s.tiles_x, s.W, s.H, ctx->stream));
        off += u;
      }
      ctx->comb_bytes = 0;
    }
  }
  if (ctx->rank == dst) {
    ctx->comb_valid = true;
    CU(crt::launch_resolve(ctx->d_comb, ctx->d_fb, (uint32_t)n_pix, ctx->sample, ctx->params.exposure, ctx->stream));
  }
  CU(cudaEventRecord(ctx->ev_c1, ctx->stream));
  ctx->comb_timed = true;
  return CRT_OK;
}

int crt_last_combine(crt_ctx* ctx, float* ms, uint64_t* bytes_sent) {
  if (!ctx) return CRT_ERR_ARG;
  if (is_group(ctx)) return group_fail(ctx, ctx->members[0], crt_last_combine(ctx->members[0], ms, bytes_sent));
  if (bytes_sent) *bytes_sent = ctx->comb_bytes;
  if (ms) {
    *ms = 0.0f;
    if (ctx->comb_timed) {
      CU(cudaSetDevice(ctx->device));
      CU(cudaEventSynchronize(ctx->ev_c1));
      CU(cudaEventElapsedTime(ms, ctx->ev_c0, ctx->ev_c1));
    }
  }
  return CRT_OK;
}

// ------------------------------------------------------------------------------------------------ hooks
int crt_trace_rays(crt_ctx* ctx, const float* rays, const uint32_t* exclude, size_t n, uint32_t* hit_idx, float* t) {
  if (!ctx || (!rays && n) || (!hit_idx && n) || (!t && n)) return CRT_ERR_ARG;
  if (is_group(ctx)) {
    const int rc = group_prepare(ctx);
    if (rc) return rc;
    return group_fail(ctx, ctx->members[0], crt_trace_rays(ctx->members[0], rays, exclude, n, hit_idx, t));
  }
  if (n > 0x7FFFFFFFu) return fail(ctx, CRT_ERR_SIZE, "too many rays");
  CU(cudaSetDevice(ctx->device));
  int rc = prepare(ctx);
  if (rc) return rc;
  if (!n) return CRT_OK;
  DevTmp d_rays, d_ex, d_idx, d_t;
  CU(d_rays.alloc(n * 32));
  CU(d_idx.alloc(n * 4));
  CU(d_t.alloc(n * 4));
  CU(cudaMemcpyAsync(d_rays.p, rays, n * 32, cudaMemcpyHostToDevice, ctx->stream));
  if (exclude) {
    CU(d_ex.alloc(n * 4));
    CU(cudaMemcpyAsync(d_ex.p, exclude, n * 4, cudaMemcpyHostToDevice, ctx->stream));
  }
  CU(crt::launch_trace_rays(ctx->scene, d_rays.as<float4>(), d_ex.as<uint32_t>(), (uint32_t)n, d_idx.as<uint32_t>(), d_t.as<float>(), ctx->d_counters, ctx->stream));
  CU(cudaMemcpyAsync(hit_idx, d_idx.p, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaMemcpyAsync(t, d_t.p, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  return CRT_OK;
}

// crt_trace_rays without any allocation (no counterpart in the reference): for
// callers that cannot call cudaMalloc -- code under CUDA-graph capture, tight serving loops, harnesses that forbid it.  Same
// inputs, outputs and results as crt_trace_rays; the device-side staging lives in g_noalloc_scratch.  It cannot run the
// scene-upload module either (that allocates), so the scene must be resident already.
int crt_trace_rays_noalloc(crt_ctx* ctx, const float* rays, const uint32_t* exclude, size_t n, uint32_t* hit_idx, float* t) {
  if (!ctx || (!rays && n) || (!hit_idx && n) || (!t && n)) return CRT_ERR_ARG;
  if (is_group(ctx)) return group_fail(ctx, ctx->members[0], crt_trace_rays_noalloc(ctx->members[0], rays, exclude, n, hit_idx, t));
  if (n > CRT_NOALLOC_MAX_RAYS)
    return fail(ctx, CRT_ERR_SIZE, "crt_trace_rays_noalloc takes at most %u rays per call (its scratch is a fixed array), got %zu: split the batch or use crt_trace_rays",
                (unsigned)CRT_NOALLOC_MAX_RAYS, n);
  for (int b = 4; b <= 8; b++)
    if (!ctx->have[b]) return fail(ctx, CRT_ERR_STATE, "binding %d has not been set", b);
  if (ctx->camera_dirty || ctx->tables_dirty || ctx->scene_dirty)
    return fail(ctx, CRT_ERR_STATE, "the scene is not resident yet (a binding changed since the last upload): this entry point allocates nothing and so cannot "
                                    "run the scene-upload module -- call crt_scene_stats, crt_dispatch_frames or crt_trace_rays once first");
  CU(cudaSetDevice(ctx->device));
  if (!n) return CRT_OK;
  std::lock_guard<std::mutex> lock(g_noalloc_mutex);   // every call shares the array
  unsigned char* scratch = nullptr;
  CU(noalloc_scratch_of(ctx->device, &scratch));
  DevScene sc = ctx->scene;                            // as prepare() leaves it, with the current parameters
  sc.max_depth = ctx->params.max_depth; sc.grid_size = ctx->params.grid_size ? ctx->params.grid_size : 1;
  sc.t_min = ctx->params.t_min; sc.exposure = ctx->params.exposure;
  sc.rank = ctx->rank; sc.world = ctx->world; sc.shard_mode = ctx->shard_mode;
  // nothing in the array has to start at zero: the kernel writes hit_idx[i] and t[i] for every i < n and reads nothing else back
  CU(cudaMemcpyAsync(scratch + kNoallocRays, rays, n * 32, cudaMemcpyHostToDevice, ctx->stream));
  if (exclude) CU(cudaMemcpyAsync(scratch + kNoallocExcl, exclude, n * 4, cudaMemcpyHostToDevice, ctx->stream));
  trace_rays_scratch_kernel<<<((unsigned)n + 127u) / 128u, 128, 0, ctx->stream>>>(sc, scratch, (uint32_t)n, exclude ? 1u : 0u, ctx->d_counters);
  CU(cudaGetLastError());
  CU(cudaMemcpyAsync(hit_idx, scratch + kNoallocIdx, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaMemcpyAsync(t, scratch + kNoallocT, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  return CRT_OK;
}

int crt_get_counters(crt_ctx* ctx, uint64_t* out, size_t n) {
  if (!ctx || !out) return CRT_ERR_ARG;
  if (n > CRT_NUM_COUNTERS) n = CRT_NUM_COUNTERS;
  if (is_group(ctx)) {
    uint64_t sum[CRT_NUM_COUNTERS] = {};
    for (crt_ctx* m : ctx->members) {
      uint64_t v[CRT_NUM_COUNTERS];
      const int rc = crt_get_counters(m, v, CRT_NUM_COUNTERS);
      if (rc) return group_fail(ctx, m, rc);
      for (int k = 0; k < CRT_NUM_COUNTERS; k++) sum[k] = k == CRT_CNT_MAX_RAY_NODES ? std::max(sum[k], v[k]) : sum[k] + v[k];
    }
    for (size_t i = 0; i < n; i++) out[i] = sum[i];
    return CRT_OK;
  }
  CU(cudaSetDevice(ctx->device));
  unsigned long long tmp[CRT_NUM_COUNTERS];
  uint32_t ctl[4] = {0, 0, 0, 0};
  CU(cudaMemcpyAsync(tmp, ctx->d_counters, sizeof(tmp), cudaMemcpyDeviceToHost, ctx->stream));
  uint32_t ctl2[4] = {0, 0, 0, 0};
  if (ctx->wf_mem) CU(cudaMemcpyAsync(ctl, ctx->wf.ctl, 16, cudaMemcpyDeviceToHost, ctx->stream));
  if (ctx->wf2_mem) CU(cudaMemcpyAsync(ctl2, ctx->wf2.ctl, 16, cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  if (ctl[1] || ctl2[1]) return fail(ctx, CRT_ERR_STATE, "wavefront iteration limit reached on the device");
  ctl[2] += ctl2[2]; ctl[3] += ctl2[3];
  tmp[CRT_CNT_LAUNCHES] = ctx->launches + ctl[3];           // host launches + the kernels the graph loop ran
  tmp[CRT_CNT_WF_ITERATIONS] = ctx->wf_iterations + ctl[2];
  for (size_t i = 0; i < n; i++) out[i] = tmp[i];
  if (ctx->wf_mem && ctx->wf.slow_min) {   // diagnostics: print and clear the slow-ray log (crt_params.slow_ray_nodes)
    ctx->slow_log.resize(1 + 3 * CRT_WF_SLOW_RAYS);
    float4* log = ctx->slow_log.data();
    CU(cudaMemcpy(log, ctx->wf.slow, ctx->slow_log.size() * sizeof(float4), cudaMemcpyDeviceToHost));
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
  if (is_group(ctx)) {
    for (crt_ctx* m : ctx->members) { const int rc = crt_reset_counters(m); if (rc) return group_fail(ctx, m, rc); }
    return CRT_OK;
  }
  CU(cudaSetDevice(ctx->device));
  CU(cudaMemsetAsync(ctx->d_counters, 0, CRT_NUM_COUNTERS * 8, ctx->stream));
  if (ctx->wf_mem) CU(cudaMemsetAsync(&ctx->wf.ctl->iters_total, 0, 8, ctx->stream));   // iters_total, launches_total
  if (ctx->wf2_mem) CU(cudaMemsetAsync(&ctx->wf2.ctl->iters_total, 0, 8, ctx->stream));
  ctx->launches = 0;
  ctx->wf_iterations = 0;
  CU(drain_marks(ctx));
  for (int k = 0; k < CRT_NUM_KERNEL_CLASSES; k++) { ctx->k_ms[k] = 0.0; ctx->k_launches[k] = 0; }
  return CRT_OK;
}

int crt_get_kernel_times(crt_ctx* ctx, double* ms, uint64_t* launches, size_t n) {
  if (!ctx || !ms || !launches) return CRT_ERR_ARG;
  if (is_group(ctx)) return group_fail(ctx, ctx->members[0], crt_get_kernel_times(ctx->members[0], ms, launches, n));
  CU(cudaSetDevice(ctx->device));
  CU(drain_marks(ctx));
  for (size_t k = 0; k < n && k < CRT_NUM_KERNEL_CLASSES; k++) { ms[k] = ctx->k_ms[k]; launches[k] = ctx->k_launches[k]; }
  return CRT_OK;
}

int crt_scene_stats(crt_ctx* ctx, uint64_t* out, size_t n) {
  if (!ctx || !out) return CRT_ERR_ARG;
  if (is_group(ctx)) {
    const int rc = group_prepare(ctx);
    if (rc) return rc;
    return group_fail(ctx, ctx->members[0], crt_scene_stats(ctx->members[0], out, n));
  }
  CU(cudaSetDevice(ctx->device));
  int rc = prepare(ctx);
  if (rc) return rc;
  for (size_t i = 0; i < n && i < 10; i++) out[i] = ctx->stats[i];
  return CRT_OK;
}

// Test hook, host only: the host BVH builder of the scene-upload module without a device (tests/test_bvh_builder.py).
CRT_API int crt_host_build_bvh(const void* prims, size_t nbytes, int threads, uint32_t* nodes, size_t node_words, uint32_t* tri_src,
                               size_t tri_cap, uint64_t* info, float* bounds, uint32_t bvh_tuning, float bvh_prim_cost) {
  if ((!prims && nbytes) || nbytes % 80 || !info) return CRT_ERR_ARG;
  const uint8_t* pb = (const uint8_t*)prims;
  std::vector<uint32_t> tri_pos;
  for (size_t i = 0; i < nbytes / 80; i++) {
    uint32_t cat; memcpy(&cat, pb + i * 80, 4);
    if (cat == 2u) tri_pos.push_back((uint32_t)i);
  }
  crt::BvhResult bvh;
  if ((bvh_tuning & CRT_BVH_BUILDER_MASK) == CRT_BVH_BUILDER_DEVICE)   // the device build's algorithm, driven by host loops
    crt::build_bvh_ploc_host(pb, tri_pos, crt::bvh_options(bvh_tuning, bvh_prim_cost), bvh);
  else
    crt::build_bvh(pb, tri_pos, threads, crt::bvh_options(bvh_tuning, bvh_prim_cost), bvh);
  info[0] = bvh.n_nodes; info[1] = bvh.tri_src.size(); info[2] = bvh.max_depth; info[3] = 4 * CRT_NODE_STRIDE;
  if (bounds) { for (int a = 0; a < 3; a++) { bounds[a] = bvh.lo[a]; bounds[3 + a] = bvh.hi[a]; } bounds[6] = bvh.pad; }
  if (!nodes) return CRT_OK;
  if (node_words < bvh.nodes.size() || tri_cap < bvh.tri_src.size() || (!tri_src && !bvh.tri_src.empty())) return CRT_ERR_SIZE;
  if (!bvh.nodes.empty()) memcpy(nodes, bvh.nodes.data(), bvh.nodes.size() * 4);
  if (!bvh.tri_src.empty()) memcpy(tri_src, bvh.tri_src.data(), bvh.tri_src.size() * 4);
  return CRT_OK;
}

// Test hook: the tree the context traverses, copied back from the device (whichever builder made it): nodes in the Wide8
// layout, and for each leaf-order slot the primitive index (data4.w) of the triangle stored there.  info as in
// crt_host_build_bvh; call with nodes == NULL to size the outputs.
CRT_API int crt_read_bvh(crt_ctx* ctx, uint32_t* nodes, size_t node_words, uint32_t* tri_index, size_t tri_cap, uint64_t* info, float* bounds) {
  if (!ctx || !info) return CRT_ERR_ARG;
  if (is_group(ctx)) { const int rc = group_prepare(ctx); if (rc) return rc; return group_fail(ctx, ctx->members[0], crt_read_bvh(ctx->members[0], nodes, node_words, tri_index, tri_cap, info, bounds)); }
  CU(cudaSetDevice(ctx->device));
  int rc = prepare(ctx);
  if (rc) return rc;
  const DevScene& s = ctx->scene;
  info[0] = s.n_nodes; info[1] = s.n_tri; info[2] = ctx->stats[7]; info[3] = 4 * CRT_NODE_STRIDE;
  if (bounds) { for (int a = 0; a < 3; a++) { bounds[a] = s.bvh_lo[a]; bounds[3 + a] = s.bvh_hi[a]; } bounds[6] = 0.0f; }
  if (!nodes) return CRT_OK;
  if (node_words < (size_t)s.n_nodes * 4 * CRT_NODE_STRIDE || tri_cap < s.n_tri || (!tri_index && s.n_tri)) return fail(ctx, CRT_ERR_SIZE, "output buffers too small");
  if (s.n_nodes) CU(cudaMemcpy(nodes, s.nodes, (size_t)s.n_nodes * 16 * CRT_NODE_STRIDE, cudaMemcpyDeviceToHost));
  if (s.n_tri) {
    std::vector<float4> tri((size_t)s.n_tri * 3);
    CU(cudaMemcpy(tri.data(), s.tri, tri.size() * 16, cudaMemcpyDeviceToHost));
    for (uint32_t i = 0; i < s.n_tri; i++) memcpy(&tri_index[i], &tri[3 * (size_t)i].w, 4);
  }
  return CRT_OK;
}

// Bench hook: measured FP32 FMA throughput of this device (TFLOP/s, 2 flops per FFMA), best of 5 launches timed with CUDA
// events -- the roofline denominator of the linear-list regime.
CRT_API int crt_measure_fp32_tflops(crt_ctx* ctx, double* tflops) {
  if (!ctx || !tflops) return CRT_ERR_ARG;
  if (is_group(ctx)) return group_fail(ctx, ctx->members[0], crt_measure_fp32_tflops(ctx->members[0], tflops));
  CU(cudaSetDevice(ctx->device));
  const int grid = ctx->sm_count * 8;
  const uint32_t iters = 4096;
  DevTmp out;
  CU(out.alloc((size_t)grid * 256 * 4));
  cudaEvent_t e0, e1;
  CU(cudaEventCreate(&e0)); CU(cudaEventCreate(&e1));
  double best = 0.0;
  for (int rep = 0; rep < 6; rep++) {
    cudaEventRecord(e0, ctx->stream);
    const cudaError_t rc = crt::launch_fp32_peak(out.as<float>(), grid, iters, ctx->stream);
    cudaEventRecord(e1, ctx->stream);
    if (rc != cudaSuccess || cudaEventSynchronize(e1) != cudaSuccess) { cudaEventDestroy(e0); cudaEventDestroy(e1); return fail(ctx, CRT_ERR_CUDA, "fp32 microbenchmark failed"); }
    float ms = 0.0f;
    cudaEventElapsedTime(&ms, e0, e1);
    const double flops = (double)grid * 256.0 * iters * 64.0 * 2.0;
    if (rep && ms > 0.0f) best = std::max(best, flops / (ms * 1e-3) / 1e12);
  }
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  *tflops = best;
  return CRT_OK;
}

// Test hook (not in crt.h's reference mapping): the CUDA build of crt_detmath.h on n inputs.
CRT_API int crt_test_detmath(crt_ctx* ctx, const float* x, size_t n, float* sn, float* cs, float* ex, float* pw) {
  if (!ctx || !x || !sn || !cs || !ex || !pw) return CRT_ERR_ARG;
  if (is_group(ctx)) return group_fail(ctx, ctx->members[0], crt_test_detmath(ctx->members[0], x, n, sn, cs, ex, pw));
  CU(cudaSetDevice(ctx->device));
  DevTmp dt;
  CU(dt.alloc(n * 4 * 5));
  float* d = dt.as<float>();
  CU(cudaMemcpyAsync(d, x, n * 4, cudaMemcpyHostToDevice, ctx->stream));
  CU(crt::launch_detmath(d, (uint32_t)n, d + n, d + 2 * n, d + 3 * n, d + 4 * n, ctx->stream));
  CU(cudaMemcpyAsync(sn, d + n, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaMemcpyAsync(cs, d + 2 * n, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaMemcpyAsync(ex, d + 3 * n, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaMemcpyAsync(pw, d + 4 * n, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  return CRT_OK;
}

}  // extern "C"
