/*
 * crt.h -- C ABI of libcrt, the B200 (sm_100a) implementation of ComputeRayTracer's
 * compute-shader path-tracing hot path.
 *
 * The reference has no FFI: the seam its hot path sits behind is the WebGPU bind-group +
 * dispatch contract between src/main.js and src/shaders/ComputeShader.wgsl.  Each entry
 * point below replaces one piece of that contract and cites it (paths relative to the
 * reference repo; CS = src/shaders/ComputeShader.wgsl, MJ = src/main.js).
 *
 * Conventions: plain pointers and sizes only; every call returns an int status (CRT_OK = 0);
 * nothing throws or aborts across the ABI; a context is used by one host thread at a time.
 * Byte layouts of the bindings are the reference's (CS:1-9, CS:22-47, written by MJ:139-393):
 *
 *   binding 0  framebuffer   W*H*4   RGBA8 unorm, row-major, top-left origin   (CS:1,  MJ:25-40)   out
 *   binding 1  accumulator   W*H*16  vec3<f32> XYZ sum in a 16-byte stride     (CS:2,  MJ:298-302) in/out
 *   binding 2  sample        4       u32 frame counter                         (CS:3,  MJ:304-311) in/out
 *   binding 3  planar patches 64*n   bound by the reference, read by no live code (CS:4, MJ:139-209) ignored
 *   binding 4  camera        64      eye@0 lookat@16 up@32 W@44 H@48 fov@52    (CS:5,22-29, MJ:313-332)
 *   binding 5  CIE           3*471*4 f32 X|Y|Z colour-matching tables, 360..830 nm (CS:6, MJ:380-393)
 *   binding 6  spectra       301*4*k f32, k spectra sampled at 400..700 nm     (CS:7,  MJ:334-378)
 *   binding 7  lights        80*m    Primitive records of the emitters         (CS:8,  MJ:255-296)
 *   binding 8  primitives    80*N    Primitive records (CS:41-47, MJ:211-246): category u32@0
 *                                    (0 patch, 1 sphere, 2 triangle [extension]), data1 vec3@16,
 *                                    data2 vec3@32, data3 vec3@48, data4 uvec4@64 =
 *                                    (emission idx, reflectance idx, material, primitive index)
 */
#ifndef CRT_H
#define CRT_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define CRT_API __attribute__((visibility("default")))
#else
#define CRT_API
#endif

typedef struct crt_ctx crt_ctx;

enum crt_status {
  CRT_OK = 0,
  CRT_ERR_ARG = 1,      /* null pointer / bad enum                                     */
  CRT_ERR_BINDING = 2,  /* binding number not valid for this call                      */
  CRT_ERR_SIZE = 3,     /* byte count does not match the binding's layout              */
  CRT_ERR_STATE = 4,    /* a required binding (4..8) has not been set yet              */
  CRT_ERR_CUDA = 5,     /* CUDA runtime error; text in crt_last_error                  */
  CRT_ERR_OOM = 6,      /* host or device allocation failed                            */
  CRT_ERR_NODEVICE = 7, /* no CUDA device / wrong architecture: there is no CPU fallback */
  CRT_ERR_NCCL = 8      /* NCCL missing (libnccl.so.2 not loadable) or an NCCL call failed; text in crt_last_error */
};

enum crt_shard_mode {
  CRT_SHARD_NONE = 0,
  CRT_SHARD_TILES = 1,     /* this rank owns the 8x4-pixel tiles with tile_id % world == rank */
  CRT_SHARD_SPLIT_SPP = 2  /* this rank renders the frames with (sample-1) % world == rank     */
};

enum crt_kernel {
  CRT_KERNEL_AUTO = 0,
  CRT_KERNEL_MEGA = 1,      /* persistent-thread megakernel with per-lane path regeneration */
  CRT_KERNEL_WAVEFRONT = 2  /* raygen / extend / shade queues with compaction                */
};

/* The reference hard-codes these as WGSL consts and literals; defaults reproduce them. */
typedef struct crt_params {
  uint32_t max_depth;   /* MAXDEPTH, CS:14 (100)                                          */
  uint32_t grid_size;   /* GRID_SIZE, CS:15 (16): jitter cell = sample % grid_size        */
  float    t_min;       /* create_intersection_context, CS:720 (0.001)                    */
  float    exposure;    /* tone_map argument, CS:112 (2.2)                                */
  uint32_t kernel;      /* enum crt_kernel; AUTO picks from the scene         */
  uint32_t frames_per_launch; /* frames per launch (megakernel: samples a lane integrates in registers) or per
                                 wavefront chunk; 0 = library default (wavefront: up to 256, bounded by the
                                 staging buffer); results never depend on it                              */
  uint32_t flags;       /* CRT_FLAG_* bits (instrumentation; results never depend on them)     */
  uint32_t wf_slots;    /* wavefront kernel: path slots in flight (0 = library default, 64 Mi);
                           results never depend on it                                           */
  /* Tuning words (all 0 = library defaults; none of them can change a result -- tests/test_gpu_parity.py runs the
     non-default settings against the oracle).  They replace what round 1 read from the process environment. */
  uint32_t wf_tuning;      /* CRT_WF_* bits below                                                   */
  uint32_t bvh_tuning;     /* bits 0-8: SAH bins per axis (0 = 32); CRT_BVH_EXACT_SLOTS; CRT_BVH_BUILDER_* */
  float    bvh_prim_cost;  /* SAH weight of a triangle test in the Wide8 collapse (0 = 0.75)        */
  uint32_t slow_ray_nodes; /* diagnostics, with CRT_FLAG_STATS: log rays visiting >= this many nodes */
} crt_params;

/* wf_tuning: how wf_logic's live slots reach the next iteration ... */
#define CRT_WF_LIVE_DEFAULT 0u      /* linear-list scenes: routed lists; scenes with triangles: class bytes + wf_partition */
#define CRT_WF_LIVE_ONE_LIST 1u
#define CRT_WF_LIVE_ROUTED 2u
#define CRT_WF_LIVE_CLASS 3u
#define CRT_WF_LIVE_MASK 3u
/* ... whether a path that ends in a bounce warp is finished by the next iteration's finish warps ... */
#define CRT_WF_DEFER_DEFAULT 0u     /* on for scenes with triangles */
#define CRT_WF_DEFER_OFF 4u
#define CRT_WF_DEFER_ON 8u
#define CRT_WF_DEFER_MASK 12u
/* ... and who runs the iteration loop: by default a CUDA-graph WHILE node on the device (crt_dispatch_frames never waits for
 * the GPU); with this bit the host launches every iteration and reads the live count back at check points (the round-1 form;
 * CRT_FLAG_KERNEL_TIMES and CRT_FLAG_WF_SPLIT imply it). */
#define CRT_WF_HOST_LOOP 16u
/* ... whether the device loop splits a chunk's samples over two half pools on two streams (default; hides the straggler phase
 * at the end of every wf_traverse launch and the thin end of a chunk behind the other half's kernels) or keeps one (this bit) */
#define CRT_WF_ONE_LANE 32u
#define CRT_WF_TWO_LANES_ALWAYS 64u   /* two lanes also for chunks below 4 Mi work items (tests) */
/* ... and when the iteration loop hands the chunk's last live slots to wf_tail (one thread runs a path to its end): bits
 * 8-12 = 0: the default (at most CRT_WF_TAIL_DEFAULT live slots and no work item left), v = 1 .. 30: at most 2^(v-1) slots,
 * 31: never (iterate until no slot is live, the round-1 form). */
#define CRT_WF_TAIL_SHIFT 8
#define CRT_WF_TAIL_MASK (31u << CRT_WF_TAIL_SHIFT)
#define CRT_WF_TAIL_OFF (31u << CRT_WF_TAIL_SHIFT)
#define CRT_WF_TAIL_DEFAULT 131072u
/* bvh_tuning */
#define CRT_BVH_BINS_MASK 0x1FFu
#define CRT_BVH_EXACT_SLOTS 0x200u
#define CRT_BVH_BUILDER_AUTO 0u        /* device LBVH build when the scene has >= 32768 triangles, host SAH below */
#define CRT_BVH_BUILDER_HOST_SAH 0x400u /* binned-SAH BVH2 on the host cores + SAH-optimal Wide8 collapse (best tree, slow build) */
#define CRT_BVH_BUILDER_DEVICE 0x800u   /* Morton-sorted LBVH + Wide8 collapse on the GPU (fast build) */
#define CRT_BVH_BUILDER_MASK 0xC00u

#define CRT_FLAG_STATS 1u        /* detailed device counters (separate, slower kernel variants)   */
#define CRT_FLAG_KERNEL_TIMES 2u /* CUDA events between hot-path launches: crt_get_kernel_times   */
#define CRT_FLAG_WF_SPLIT 4u     /* wavefront: wf_shade + wf_regen over compacted queues instead of
                                    the single wf_logic kernel (same results; tuning/evidence)   */

/* Kernel classes of crt_get_kernel_times. */
enum crt_kernel_class {
  CRT_KC_WF_LOGIC = 0,      /* wf_logic: shade + finish + regenerate, one pass per live slot */
  CRT_KC_WF_TRAVERSE = 1,   /* wf_traverse: BVH traversal + triangle tests                   */
  CRT_KC_WF_ACCUMULATE = 2, /* wf_accumulate: ordered sample sum + epilogue                  */
  CRT_KC_MEGA = 3,          /* pt_megakernel                                                 */
  CRT_KC_WF_REGEN = 4,      /* wf_regen: finish ended paths, start new ones on free slots    */
  CRT_KC_WF_SHADE = 5,      /* wf_shade: one bounce per slot of the shade queue (split form) */
  CRT_KC_WF_PARTITION = 6,  /* wf_partition: bounce and finish lists from the slots' class bytes */
  CRT_KC_WF_TAIL = 7,       /* wf_tail: the chunk's last live slots, each path run to its end by one thread */
  CRT_NUM_KERNEL_CLASSES = 8
};

#define CRT_NUM_COUNTERS 16
enum crt_counter {
  CRT_CNT_PATHS = 0,        /* pixel-samples started                                      */
  CRT_CNT_RAYS_EXT = 1,     /* intersect() calls from path_trace, CS:135 (primary+bounce) */
  CRT_CNT_RAYS_SHADOW = 2,  /* intersect() calls from shadow_intersect, CS:699            */
  CRT_CNT_PRIM_TESTS = 3,   /* ray_intersection() calls on patches/spheres, CS:514        */
  CRT_CNT_NODE_VISITS = 4,  /* BVH nodes fetched                                          */
  CRT_CNT_TRI_TESTS = 5,    /* triangle tests                                             */
  CRT_CNT_RAND = 6,         /* rand() calls, CS:893                                       */
  CRT_CNT_TERM_MISS = 7,
  CRT_CNT_TERM_RR = 8,
  CRT_CNT_TERM_LIGHT = 9,
  CRT_CNT_TERM_DEPTH = 10,
  CRT_CNT_DEPTH_SUM = 11,   /* sum over paths of depth at termination                     */
  CRT_CNT_LAUNCHES = 12,    /* hot-path kernel launches (counted on the host)             */
  CRT_CNT_WF_ITERATIONS = 13, /* wavefront iterations (one ray per live path each)        */
  CRT_CNT_BVH_RAYS = 14,    /* wavefront: rays queued for BVH traversal (the rest are settled
                               by the linear list and the root-box test)                  */
  CRT_CNT_MAX_RAY_NODES = 15  /* wavefront, CRT_FLAG_STATS: most BVH nodes visited by one ray (a maximum,
                               not a sum)                                                   */
};

/* requestAdapter/requestDevice + pipeline creation (MJ:8-9, MJ:535-543).  Fails with
 * CRT_ERR_NODEVICE when `device` is not a usable sm_100 GPU. */
CRT_API int crt_create(int device, crt_ctx** out);
CRT_API void crt_destroy(crt_ctx* ctx);

/* createBuffer(mappedAtCreation) + typed-array writes + unmap (MJ:139-393) and the bind group
 * (MJ:465-529).  binding in 1..8; bytes are copied (caller keeps ownership).  Setting 4 or 8
 * re-runs the scene-upload module (SoA repack, BVH build).  Setting 1/2 restores a saved
 * accumulator / frame counter (resume).  Binding 3 is accepted and ignored. */
CRT_API int crt_set_binding(crt_ctx* ctx, uint32_t binding, const void* bytes, size_t nbytes);

/* WebGPU zero-initialises the accumulator (MJ:298-302) and main.js writes sample = 0
 * (MJ:304-311): this restores that state. */
CRT_API int crt_reset(crt_ctx* ctx);

CRT_API void crt_default_params(crt_params* p);
CRT_API int crt_set_params(crt_ctx* ctx, const crt_params* p);
CRT_API int crt_get_params(crt_ctx* ctx, crt_params* p);

/* ---- Multi-GPU (new; the reference is single-device: one adapter, one device, MJ:8-9).  The scene is replicated, pixels
 * (CRT_SHARD_TILES) or frames (CRT_SHARD_SPLIT_SPP) are sharded, and nothing is exchanged while rendering; the one exchange
 * step is crt_combine at readback.  Two ways to run it:
 *
 *  (a) one process, n GPUs -- what a single Node.js process does: crt_create_multi returns ONE context that owns a member
 *      context per device.  Every call below works on it: bindings/params/reset are replicated, crt_dispatch_frames runs all
 *      members, crt_read_binding(0|1) combines first and returns the combined image of member 0, counters are summed.
 *      crt_combine on such a context is one fused kernel on member 0 that reads the other members' accumulators through
 *      NVLink peer pointers, sums them in rank order (deterministic) and writes the combined accumulator and the tone-mapped
 *      framebuffer -- no staging copy, no separate resolve pass.
 *  (b) one process per GPU (torchrun, MPI): each process creates its own context with crt_create and joins a communicator
 *      with crt_comm_init; the host layer only has to carry the 128-byte id from rank 0 to the others.  crt_combine is then
 *      ncclReduce(SUM, f32, W*H*4) to `dst` for split-spp, and for tiles a grouped ncclSend/ncclRecv of the OWNED tiles only
 *      (W*H*16/world bytes per rank), unpacked on `dst`.  NCCL is loaded at run time (dlopen of libnccl.so.2).
 *
 * crt_combine is repeatable: the ranks' partial sums stay partial (split-spp reduces into a separate combined buffer on
 * `dst`; a tile gather overwrites, never adds), so frames / combine / frames / combine gives the same image as one render.
 * After crt_combine, bindings 0 and 1 read on `dst` are the combined image until the next dispatch, reset or set_binding. */
CRT_API int crt_set_shard(crt_ctx* ctx, uint32_t rank, uint32_t world, uint32_t mode);
CRT_API int crt_create_multi(const int* devices, int n, uint32_t shard_mode, crt_ctx** out);
#define CRT_COMM_ID_BYTES 128
CRT_API int crt_comm_unique_id(void* id_bytes, size_t nbytes);   /* rank 0: ncclGetUniqueId */
CRT_API int crt_comm_init(crt_ctx* ctx, const void* id_bytes, size_t nbytes, uint32_t rank, uint32_t world, uint32_t shard_mode);
CRT_API int crt_comm_destroy(crt_ctx* ctx);
/* Collective (every rank calls it with the same dst); asynchronous on the context's stream.  Device time of the most recent
 * combine (ms, CUDA events around it) and the bytes this rank sent: crt_last_combine. */
CRT_API int crt_combine(crt_ctx* ctx, uint32_t dst);
CRT_API int crt_last_combine(crt_ctx* ctx, float* ms, uint64_t* bytes_sent);

/* frame() (MJ:597-611): n times { UpdateVariables.main: sample++ (UpdateVariables.wgsl:3-8);
 * ComputeShader.main over ceil(W/8) x ceil(H/8) workgroups: one more sample per pixel }.
 * Asynchronous on the context's stream. */
CRT_API int crt_dispatch_frames(crt_ctx* ctx, uint32_t n);
CRT_API int crt_synchronize(crt_ctx* ctx);

/* Device time (ms, CUDA events on the context's stream) of the most recent
 * crt_dispatch_frames call; synchronises. */
CRT_API int crt_last_dispatch_ms(crt_ctx* ctx, float* ms);

/* Readback (new API surface: the reference never maps a buffer; MJ:301 only hints COPY_SRC).
 * binding 0, 1 or 2; synchronises.  nbytes must equal the binding's size. */
CRT_API int crt_read_binding(crt_ctx* ctx, uint32_t binding, void* dst, size_t nbytes);

/* Recompute binding 0 from binding 1 and the frame counter (the epilogue of CS:109-115). Used
 * after a multi-GPU combine has changed the accumulator. */
CRT_API int crt_resolve_framebuffer(crt_ctx* ctx);

/* Raw device addresses of bindings 0/1 so the host layer can hand them to NCCL
 * (reduce / gather over NVLink) without a copy.  Also the CUDA stream the context launches on. */
CRT_API int crt_device_ptr(crt_ctx* ctx, uint32_t binding, void** dptr, size_t* nbytes);
CRT_API int crt_stream(crt_ctx* ctx, void** cuda_stream);
/* Launch on a caller-owned CUDA stream (NULL: back to the context's own), e.g. the host layer's
 * current stream so that its events bracket the kernels. */
CRT_API int crt_set_stream(crt_ctx* ctx, void* cuda_stream);

/* Test hook: closest hit of intersect() (CS:503-518) for n caller-supplied rays.
 * rays: n x 8 floats (origin.xyz, unused, direction.xyz, unused).  exclude: n primitive
 * indices (0xFFFFFFFF = none) or NULL.  hit_idx[i] = primitive index or 0xFFFFFFFF, t[i] = the
 * context's final t_max (CS:577; 2139095040.0 on a miss, CS:12). */
CRT_API int crt_trace_rays(crt_ctx* ctx, const float* rays, const uint32_t* exclude, size_t n,
                           uint32_t* hit_idx, float* t);

/* crt_trace_rays for callers that cannot allocate.  NO COUNTERPART IN THE REFERENCE: added
 * for code under CUDA-graph capture, tight serving loops and harnesses that forbid cudaMalloc.  Same arguments, same results (hit
 * index and t bit for bit) as crt_trace_rays, but the call makes no cudaMalloc / cudaFree and calls no allocating library routine:
 * the device-side staging of rays and results is a file-scope static __device__ array inside the library, sized for
 * CRT_NOALLOC_MAX_RAYS rays (2.75 MiB, held by every GPU that loads the library until the process exits).
 *   - n > CRT_NOALLOC_MAX_RAYS is rejected with CRT_ERR_SIZE (split the batch, or use crt_trace_rays);
 *   - the scene must be resident already (the scene-upload module allocates): CRT_ERR_STATE if a binding has been set since the last
 *     crt_dispatch_frames / crt_scene_stats / crt_trace_rays;
 *   - every call on a device shares the one array, which keeps the previous call's contents; nothing in it has to start at zero
 *     (each call overwrites what it reads back).  Calls are serialised within a process and synchronise before returning; calls
 *     that would overlap on the GPU some other way (several processes each hold their own copy, so that is not one) would race. */
#define CRT_NOALLOC_MAX_RAYS 65536u
CRT_API int crt_trace_rays_noalloc(crt_ctx* ctx, const float* rays, const uint32_t* exclude, size_t n,
                                   uint32_t* hit_idx, float* t);

/* Device-side counters accumulated since the last crt_reset_counters (enum crt_counter). */
CRT_API int crt_get_counters(crt_ctx* ctx, uint64_t* out, size_t n);
CRT_API int crt_reset_counters(crt_ctx* ctx);

/* Device time per kernel class (enum crt_kernel_class) since the last crt_reset_counters, measured with
 * CUDA events recorded on the context's stream between consecutive hot-path launches while
 * CRT_FLAG_KERNEL_TIMES is set.  ms[k] = summed duration, launches[k] = launches timed.  Synchronises. */
CRT_API int crt_get_kernel_times(crt_ctx* ctx, double* ms, uint64_t* launches, size_t n);

/* Scene-upload statistics: out[0]=N primitives, [1]=patches+spheres (linear list),
 * [2]=triangles, [3]=BVH nodes, [4]=BVH bytes, [5]=triangle bytes, [6]=scene upload microseconds (repack + tree),
 * [7]=max BVH depth, [8]=microseconds of the tree build alone, [9]=builder (0 no triangles, 1 host SAH, 2 device) |
 * 0x100 when a light is a triangle (shadow rays need closest-hit: megakernel only). */
CRT_API int crt_scene_stats(crt_ctx* ctx, uint64_t* out, size_t n);

/* Test hook: the CUDA build of crt_detmath.h (sin, cos, exp, pow(x, 1/2.4)) on n inputs, compared bit
 * for bit with the oracle's gcc build of the same header. */
CRT_API int crt_test_detmath(crt_ctx* ctx, const float* x, size_t n, float* sin_out, float* cos_out,
                             float* exp_out, float* pow_out);

/* Test hook, host only (needs no device and no context): the scene-upload module's BVH builder
 * (csrc/crt_bvh.cpp) on binding-8 bytes.  Call once with nodes == NULL to size the outputs:
 * info[0] = wide nodes, info[1] = triangles, info[2] = depth of the wide tree, info[3] = 32-bit words
 * per node (the device layout of csrc/crt_device.cuh, "Wide8").  bounds = lo[3], hi[3] of the padded
 * root box, then the padding.  tri_src[slot] = position in binding 8 of the triangle at leaf-order
 * slot `slot`.  The result does not depend on `threads`. */
CRT_API int crt_host_build_bvh(const void* prims, size_t nbytes, int threads, uint32_t* nodes, size_t node_words,
                               uint32_t* tri_src, size_t tri_cap, uint64_t* info, float* bounds, uint32_t bvh_tuning, float bvh_prim_cost);

/* Test hook: the tree the context traverses, copied back from the device whichever builder made it: nodes in the Wide8
 * layout and, per leaf-order slot, the primitive index (data4.w) of the triangle stored there.  info as above; nodes == NULL
 * sizes the outputs. */
CRT_API int crt_read_bvh(crt_ctx* ctx, uint32_t* nodes, size_t node_words, uint32_t* tri_index, size_t tri_cap, uint64_t* info, float* bounds);

/* Bench hook: measured FP32 FMA throughput of the device in TFLOP/s (a microbenchmark of independent FFMA chains): the
 * roofline denominator of the linear-list regime, which is FP32/issue-bound. */
CRT_API int crt_measure_fp32_tflops(crt_ctx* ctx, double* tflops);

CRT_API const char* crt_last_error(crt_ctx* ctx);
CRT_API const char* crt_version(void);

#ifdef __cplusplus
}
#endif
#endif /* CRT_H */
