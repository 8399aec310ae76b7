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
  CRT_ERR_NODEVICE = 7  /* no CUDA device / wrong architecture: there is no CPU fallback */
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
  uint32_t reserved[4];
} crt_params;

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
  CRT_NUM_KERNEL_CLASSES = 7
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

/* Multi-GPU (new; the reference is single-device).  One context per rank/GPU. */
CRT_API int crt_set_shard(crt_ctx* ctx, uint32_t rank, uint32_t world, uint32_t mode);

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

/* Device-side counters accumulated since the last crt_reset_counters (enum crt_counter). */
CRT_API int crt_get_counters(crt_ctx* ctx, uint64_t* out, size_t n);
CRT_API int crt_reset_counters(crt_ctx* ctx);

/* Device time per kernel class (enum crt_kernel_class) since the last crt_reset_counters, measured with
 * CUDA events recorded on the context's stream between consecutive hot-path launches while
 * CRT_FLAG_KERNEL_TIMES is set.  ms[k] = summed duration, launches[k] = launches timed.  Synchronises. */
CRT_API int crt_get_kernel_times(crt_ctx* ctx, double* ms, uint64_t* launches, size_t n);

/* Scene-upload statistics: out[0]=N primitives, [1]=patches+spheres (linear list),
 * [2]=triangles, [3]=BVH nodes, [4]=BVH bytes, [5]=triangle bytes, [6]=build microseconds,
 * [7]=max BVH depth. */
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
                               uint32_t* tri_src, size_t tri_cap, uint64_t* info, float* bounds);

CRT_API const char* crt_last_error(crt_ctx* ctx);
CRT_API const char* crt_version(void);

#ifdef __cplusplus
}
#endif
#endif /* CRT_H */
