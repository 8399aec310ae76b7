// crt_kernels.cu -- CUDA kernels of libcrt (sm_100a).
//
//   pt_megakernel      the hot path: ComputeShader.wgsl main() (CS:77-117) for n frames, as a
//                      persistent-thread kernel with per-lane path regeneration
//   trace_rays_kernel  test hook: intersect() (CS:503-518) on caller-supplied rays
//   resolve_kernel     epilogue of main (CS:109-115) over the whole accumulator
//   repack_* kernels   the scene-upload module: reference-layout buffers -> SoA device layouts
//
// Build flags matter for parity: -fmad=false -prec-div=true -prec-sqrt=true -ftz=false.
#include "crt_device.cuh"
#include "crt_path.cuh"
#include "crt_kernels.h"

#define CRT_BLOCK 128

// ------------------------------------------------------------------------------------------------ upload module
// 80-byte Primitive record (CS:41-47; written by main.js:211-246)
struct RawPrim {
  uint32_t category, pad0[3];
  float d1[3]; uint32_t pad1;
  float d2[3]; uint32_t pad2;
  float d3[3]; uint32_t pad3;
  uint32_t d4[4];
};

__device__ __forceinline__ uint32_t make_meta(const RawPrim& p) {
  uint32_t mat = p.d4[2] > 3u ? 3u : p.d4[2];          // anything that is not diffuse/light/glass takes no branch
  uint32_t emis = p.d4[0] > 0xFFFu ? 0xFFFu : p.d4[0];
  uint32_t refl = p.d4[1] > 0xFFFu ? 0xFFFu : p.d4[1];
  return (p.category & 3u) | (mat << 2) | (emis << 4) | (refl << 16);
}

// linear list: the patches and spheres of binding 8, in index order; lin_src[i] = position in prims[]
__global__ void repack_lin_kernel(const RawPrim* __restrict__ prims, const uint32_t* __restrict__ lin_src, uint32_t n, LinPrim* __restrict__ out) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const RawPrim p = prims[lin_src[i]];
  LinPrim q;
  const float idxf = __uint_as_float(p.d4[3]);
  const float metaf = __uint_as_float(make_meta(p));
  if (p.category == 0u) {
    V3 e1 = v3(p.d2), e2 = v3(p.d3);
    V3 n_ = normalize(cross(e1, e2));                  // CS:536
    q.A = make_float4(n_.x, n_.y, n_.z, idxf);
    q.B = make_float4(p.d1[0], p.d1[1], p.d1[2], metaf);
    q.C = make_float4(e1.x, e1.y, e1.z, dot(e1, e1));  // CS:563
    q.D = make_float4(e2.x, e2.y, e2.z, dot(e2, e2));  // CS:564
  } else {
    float r = p.d2[0];                                 // CS:593-594
    q.A = make_float4(p.d1[0], p.d1[1], p.d1[2], idxf);
    q.B = make_float4(r * r, r, 0.0f, metaf);
    q.C = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    q.D = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
  }
  out[i] = q;
}

// triangles in BVH leaf order: tri_src[slot] = position in prims[]
__global__ void repack_tri_kernel(const RawPrim* __restrict__ prims, const uint32_t* __restrict__ tri_src, uint32_t n, float4* __restrict__ out) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const RawPrim p = prims[tri_src[i]];
  out[3 * i + 0] = make_float4(p.d1[0], p.d1[1], p.d1[2], __uint_as_float(p.d4[3]));
  out[3 * i + 1] = make_float4(p.d2[0], p.d2[1], p.d2[2], __uint_as_float(make_meta(p)));
  out[3 * i + 2] = make_float4(p.d3[0], p.d3[1], p.d3[2], 0.0f);
}

__global__ void repack_lights_kernel(const RawPrim* __restrict__ lights, uint32_t n, LightRec* __restrict__ out) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const RawPrim p = lights[i];
  float light_area = length(v3(p.d2)) * length(v3(p.d3));   // CS:363
  LightRec q;
  q.d1 = make_float4(p.d1[0], p.d1[1], p.d1[2], __uint_as_float(p.d4[3]));
  q.d2 = make_float4(p.d2[0], p.d2[1], p.d2[2], __uint_as_float(p.d4[0]));
  q.d3 = make_float4(p.d3[0], p.d3[1], p.d3[2], 1.0f / light_area);  // CS:364
  out[i] = q;
}

// camera_basis + the pixel-invariant part of camera_ray, CS:470-487.  vph = 2*tan(fov/2) comes from the
// host (double libm tan rounded once to f32: the tan pin of the numerics contract).
__global__ void camera_setup_kernel(const float* __restrict__ cam, float vph, float* __restrict__ out12) {
  V3 eye = v3(cam[0], cam[1], cam[2]), lookat = v3(cam[4], cam[5], cam[6]), up = v3(cam[8], cam[9], cam[10]);
  V3 w = normalize(eye - lookat);
  V3 u = normalize(cross(up, w));
  V3 v = cross(w, u);
  float aspect = cam[11] / cam[12];
  float vpw = aspect * vph;
  V3 hor = u * vpw, ver = v * vph;
  V3 llc = ((eye - hor / 2.0f) - ver / 2.0f) - w;
  out12[0] = eye.x; out12[1] = eye.y; out12[2] = eye.z;
  out12[3] = hor.x; out12[4] = hor.y; out12[5] = hor.z;
  out12[6] = ver.x; out12[7] = ver.y; out12[8] = ver.z;
  out12[9] = llc.x; out12[10] = llc.y; out12[11] = llc.z;
}

// seed.w of CS:98, tea(x, y*100): 16 rounds that depend on the pixel only -- hashed once per camera instead of per path
__global__ void tea_tab_kernel(uint32_t W, uint32_t H, uint32_t* __restrict__ out) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= W * H) return;
  out[i] = tea(i % W, (i / W) * 100u);
}

// ------------------------------------------------------------------------------------------------ the hot path
// (PathState, path_begin, path_consume, path_step: crt_path.cuh)

// lanes of the warp keep tracing: every trip through path_step has all live lanes inside intersect().
template <bool STATS>
__global__ void __launch_bounds__(CRT_BLOCK) pt_megakernel(const __grid_constant__ DevScene sc, float4* __restrict__ acc, uint32_t* __restrict__ fb,
                                                           uint32_t sample_start, uint32_t n_frames, uint32_t n_units,
                                                           uint32_t* __restrict__ work_counter, unsigned long long* __restrict__ counters,
                                                           uint32_t tables_in_smem) {
  extern __shared__ float s_tables[];
  const float* spectra = sc.spectra;
  const float* cie = sc.cie;
  if (tables_in_smem) {
    const uint32_t ns = sc.n_spectra * CRT_LAMBDA_RANGE;
    for (uint32_t i = threadIdx.x; i < ns; i += blockDim.x) s_tables[i] = sc.spectra[i];
    for (uint32_t i = threadIdx.x; i < 3u * 471u; i += blockDim.x) s_tables[ns + i] = sc.cie[i];
    __syncthreads();
    spectra = s_tables;
    cie = s_tables + ns;
  }
  const uint32_t lane = threadIdx.x & 31u;
  const uint32_t s_end = sample_start + n_frames;
  const bool split = (sc.shard_mode == CRT_SHARD_SPLIT_SPP && sc.world > 1);
  const bool tiles = (sc.shard_mode == CRT_SHARD_TILES && sc.world > 1);

  PathState ps;
  Rng rng;
  Stats st; st.prim_tests = st.node_visits = st.tri_tests = st.rand_calls = 0;
  uint32_t n_paths = 0, n_ext = 0, n_shadow = 0, depth_sum = 0;
  uint32_t term_cnt[4] = {0, 0, 0, 0};
  bool alive = false, have_pixel = false, exhausted = false;
  uint32_t px = 0, py = 0, pix = 0, s_cur = 0;
  float4 a = make_float4(0.0f, 0.0f, 0.0f, 0.0f);

  for (;;) {
    // ---- lanes between paths: next sample of the same pixel, or retire the pixel
    if (!alive && have_pixel) {
      uint32_t s = s_cur + 1;
      if (split) s += (sc.rank + sc.world - ((s - 1u) % sc.world)) % sc.world;
      if (s <= s_end) {
        s_cur = s;
        path_begin<STATS, false>(sc, px, py, s, ps, rng, st);
        n_paths++;
        alive = true;
      } else {
        acc[pix] = a;                                                          // CS:108
        fb[pix] = resolve_pixel(a.x, a.y, a.z, s_end, sc.exposure);            // CS:109-115
        have_pixel = false;
      }
    }
    // ---- lanes without a pixel fetch one; the whole warp takes part in the ballot
    for (;;) {
      const bool need = !alive && !have_pixel && !exhausted;
      const uint32_t mask = __ballot_sync(0xffffffffu, need);
      if (mask == 0u) break;
      uint32_t base = 0;
      const uint32_t leader = __ffs(mask) - 1u;
      if (lane == leader) base = atomicAdd(work_counter, (uint32_t)__popc(mask));
      base = __shfl_sync(0xffffffffu, base, leader);
      if (need) {
        const uint32_t unit = base + __popc(mask & ((1u << lane) - 1u));
        if (unit >= n_units) {
          exhausted = true;
        } else {
          uint32_t tile = unit >> 5;
          const uint32_t l = unit & 31u;
          if (tiles) tile = tile * sc.world + sc.rank;
          px = (tile % sc.tiles_x) * 8u + (l & 7u);
          py = (tile / sc.tiles_x) * 4u + (l >> 3);
          if (px < sc.W && py < sc.H) {
            pix = px + py * sc.W;                                              // CS:107
            a = acc[pix];
            uint32_t s = sample_start + 1u;
            if (split) s += (sc.rank + sc.world - ((s - 1u) % sc.world)) % sc.world;
            have_pixel = true;
            if (s <= s_end) {
              s_cur = s;
              path_begin<STATS, false>(sc, px, py, s, ps, rng, st);
              n_paths++;
              alive = true;
            } else {
              s_cur = s_end;   // nothing to render for this rank: retire on the next trip
            }
          }
        }
      }
    }
    if (__ballot_sync(0xffffffffu, alive || have_pixel) == 0u) break;
    // ---- one intersect() per live lane
    if (alive) {
      uint32_t term = 0;
      if (!path_step<STATS>(sc, spectra, ps, rng, n_ext, n_shadow, st, &term)) {
        uint32_t lam[4];
        wavelengths(ps.lam0, lam);
        const V3 xyz = spectral_to_xyz(cie, ps.L, lam);                        // CS:105
        a.x += xyz.x; a.y += xyz.y; a.z += xyz.z;                              // CS:108
        alive = false;
        if (STATS) { term_cnt[term - TERM_MISS]++; depth_sum += ps.depth; }
      }
    }
  }

  // ---- counters: one atomic per warp and counter
  unsigned long long v[CRT_NUM_COUNTERS];
#pragma unroll
  for (int i = 0; i < CRT_NUM_COUNTERS; i++) v[i] = 0;
  v[CRT_CNT_PATHS] = n_paths; v[CRT_CNT_RAYS_EXT] = n_ext; v[CRT_CNT_RAYS_SHADOW] = n_shadow;
  if (STATS) {
    v[CRT_CNT_PRIM_TESTS] = st.prim_tests; v[CRT_CNT_NODE_VISITS] = st.node_visits; v[CRT_CNT_TRI_TESTS] = st.tri_tests;
    v[CRT_CNT_RAND] = st.rand_calls;
    v[CRT_CNT_TERM_MISS] = term_cnt[0]; v[CRT_CNT_TERM_RR] = term_cnt[1]; v[CRT_CNT_TERM_LIGHT] = term_cnt[2];
    v[CRT_CNT_TERM_DEPTH] = term_cnt[3]; v[CRT_CNT_DEPTH_SUM] = depth_sum;
  }
#pragma unroll
  for (int i = 0; i < 12; i++) {
    if (!STATS && i > CRT_CNT_RAYS_SHADOW) break;
    unsigned long long x = v[i];
    for (int o = 16; o > 0; o >>= 1) x += __shfl_down_sync(0xffffffffu, x, o);
    if (lane == 0 && x) atomicAdd(&counters[i], x);
  }
}

// ------------------------------------------------------------------------------------------------ test hook + resolve
__global__ void trace_rays_kernel(const __grid_constant__ DevScene sc, const float4* __restrict__ rays, const uint32_t* __restrict__ exclude,
                                  uint32_t n, uint32_t* __restrict__ hit_idx, float* __restrict__ t_out, unsigned long long* __restrict__ counters) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  Stats st; st.prim_tests = st.node_visits = st.tri_tests = st.rand_calls = 0;
  if (i < n) {
    const float4 o = rays[2 * i], d = rays[2 * i + 1];
    const Hit h = intersect<true>(sc, v3(o), v3(d), exclude ? exclude[i] : CRT_MAX_U32, st);
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

__global__ void resolve_kernel(const float4* __restrict__ acc, uint32_t* __restrict__ fb, uint32_t n, uint32_t sample, float exposure) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float4 a = acc[i];
  fb[i] = resolve_pixel(a.x, a.y, a.z, sample, exposure);
}

// ------------------------------------------------------------------------------------------------ multi-GPU combine
// Tile gather (one process per GPU, NCCL): a rank packs the accumulator entries of the tiles it owns -- W*H*16/world bytes --
// into a contiguous send buffer in unit order; the destination unpacks each peer's buffer into the combined accumulator.
__global__ void pack_tiles_kernel(const float4* __restrict__ acc, float4* __restrict__ out, uint32_t n_units, uint32_t rank, uint32_t world,
                                  uint32_t tiles_x, uint32_t W, uint32_t H) {
  const uint32_t unit = blockIdx.x * blockDim.x + threadIdx.x;
  if (unit >= n_units) return;
  uint32_t px, py;
  const bool in = unit_pixel_of(unit, rank, world, true, tiles_x, W, H, px, py);
  out[unit] = in ? acc[px + py * W] : make_float4(0.0f, 0.0f, 0.0f, 0.0f);
}

__global__ void unpack_tiles_kernel(const float4* __restrict__ in, float4* __restrict__ acc, uint32_t n_units, uint32_t rank, uint32_t world,
                                    uint32_t tiles_x, uint32_t W, uint32_t H) {
  const uint32_t unit = blockIdx.x * blockDim.x + threadIdx.x;
  if (unit >= n_units) return;
  uint32_t px, py;
  if (unit_pixel_of(unit, rank, world, true, tiles_x, W, H, px, py)) acc[px + py * W] = in[unit];
}

// Fused combine of a single-process group (crt_create_multi): member 0 reads the members' accumulators through NVLink peer
// pointers, sums the split-spp partials in rank order (or, for tiles, takes each pixel from the one member that owns it: W*H*16
// bytes cross the links in total, W*H*16/world per member), and writes the combined accumulator and the tone-mapped
// framebuffer (CS:109-115) in the same pass.
__global__ void combine_peers_kernel(const CombinePeers peers, float4* __restrict__ comb, uint32_t* __restrict__ fb, uint32_t W, uint32_t H,
                                     uint32_t tiles_x, uint32_t sample, float exposure) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= W * H) return;
  float4 a;
  if (peers.tiles) {
    const uint32_t px = i % W, py = i / W;
    const uint32_t tile = (py >> 2) * tiles_x + (px >> 3);
    a = peers.acc[tile % peers.n][i];
  } else {
    a = peers.acc[0][i];
    for (uint32_t r = 1; r < peers.n; r++) {
      const float4 b = peers.acc[r][i];
      a.x += b.x; a.y += b.y; a.z += b.z;
    }
  }
  comb[i] = a;
  fb[i] = resolve_pixel(a.x, a.y, a.z, sample, exposure);
}

// FP32 pipe microbenchmark ("measure an FP32 FMA microbenchmark in the bench run"): 8 independent FFMA chains
// per thread (explicit fmaf: the library is built with -fmad=false), 2 flops per FFMA.  The measured rate is the yardstick for
// the linear-list regime (configs 1-2), which is FP32/issue-bound, not HBM-bound.
__global__ void __launch_bounds__(256) fp32_peak_kernel(float* __restrict__ out, uint32_t iters, float a, float b) {
  float x0 = threadIdx.x, x1 = x0 + 1.0f, x2 = x0 + 2.0f, x3 = x0 + 3.0f, x4 = x0 + 4.0f, x5 = x0 + 5.0f, x6 = x0 + 6.0f, x7 = x0 + 7.0f;
  for (uint32_t i = 0; i < iters; i++) {
#pragma unroll
    for (int u = 0; u < 8; u++) {
      x0 = fmaf(x0, a, b); x1 = fmaf(x1, a, b); x2 = fmaf(x2, a, b); x3 = fmaf(x3, a, b);
      x4 = fmaf(x4, a, b); x5 = fmaf(x5, a, b); x6 = fmaf(x6, a, b); x7 = fmaf(x7, a, b);
    }
  }
  const float s = ((x0 + x1) + (x2 + x3)) + ((x4 + x5) + (x6 + x7));
  if (s == 123.456f) out[blockIdx.x * blockDim.x + threadIdx.x] = s;   // never true: keeps the chains alive
}

// detmath probe: the CUDA build of include/crt_detmath.h, compared bit for bit with the oracle's build
__global__ void detmath_kernel(const float* __restrict__ x, uint32_t n, float* s, float* c, float* e, float* p) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  crt_sincosf(x[i], &s[i], &c[i]);
  e[i] = crt_expf(x[i]);
  p[i] = crt_powf(x[i], 0.41666666f);
}

// ------------------------------------------------------------------------------------------------ host launchers
namespace crt {

cudaError_t launch_repack_lin(const void* prims, const uint32_t* lin_src, uint32_t n, LinPrim* out, cudaStream_t s) {
  if (n) repack_lin_kernel<<<(n + 127) / 128, 128, 0, s>>>((const RawPrim*)prims, lin_src, n, out);
  return cudaGetLastError();
}
cudaError_t launch_repack_tri(const void* prims, const uint32_t* tri_src, uint32_t n, float4* out, cudaStream_t s) {
  if (n) repack_tri_kernel<<<(n + 127) / 128, 128, 0, s>>>((const RawPrim*)prims, tri_src, n, out);
  return cudaGetLastError();
}
cudaError_t launch_repack_lights(const void* lights, uint32_t n, LightRec* out, cudaStream_t s) {
  if (n) repack_lights_kernel<<<(n + 127) / 128, 128, 0, s>>>((const RawPrim*)lights, n, out);
  return cudaGetLastError();
}
cudaError_t launch_tea_tab(uint32_t W, uint32_t H, uint32_t* out, cudaStream_t s) {
  tea_tab_kernel<<<(W * H + 255) / 256, 256, 0, s>>>(W, H, out);
  return cudaGetLastError();
}
cudaError_t launch_camera_setup(const float* cam, float vph, float* out12, cudaStream_t s) {
  camera_setup_kernel<<<1, 1, 0, s>>>(cam, vph, out12);
  return cudaGetLastError();
}

size_t megakernel_tables_bytes(const DevScene& sc) { return ((size_t)sc.n_spectra * CRT_LAMBDA_RANGE + 3u * 471u) * sizeof(float); }

cudaError_t megakernel_config(bool stats, size_t smem, int* blocks_per_sm) {
  cudaError_t e;
  if (stats) {
    e = cudaFuncSetAttribute(pt_megakernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    return cudaOccupancyMaxActiveBlocksPerMultiprocessor(blocks_per_sm, pt_megakernel<true>, CRT_BLOCK, smem);
  }
  e = cudaFuncSetAttribute(pt_megakernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  return cudaOccupancyMaxActiveBlocksPerMultiprocessor(blocks_per_sm, pt_megakernel<false>, CRT_BLOCK, smem);
}

cudaError_t launch_megakernel(const DevScene& sc, bool stats, int grid, size_t smem, bool tables_in_smem, float4* acc, uint32_t* fb,
                              uint32_t sample_start, uint32_t n_frames, uint32_t n_units, uint32_t* work_counter,
                              unsigned long long* counters, cudaStream_t s) {
  if (stats)
    pt_megakernel<true><<<grid, CRT_BLOCK, smem, s>>>(sc, acc, fb, sample_start, n_frames, n_units, work_counter, counters, tables_in_smem ? 1u : 0u);
  else
    pt_megakernel<false><<<grid, CRT_BLOCK, smem, s>>>(sc, acc, fb, sample_start, n_frames, n_units, work_counter, counters, tables_in_smem ? 1u : 0u);
  return cudaGetLastError();
}

cudaError_t launch_trace_rays(const DevScene& sc, const float4* rays, const uint32_t* exclude, uint32_t n, uint32_t* hit_idx, float* t,
                              unsigned long long* counters, cudaStream_t s) {
  if (n) trace_rays_kernel<<<(n + 127) / 128, 128, 0, s>>>(sc, rays, exclude, n, hit_idx, t, counters);
  return cudaGetLastError();
}

cudaError_t launch_resolve(const float4* acc, uint32_t* fb, uint32_t n, uint32_t sample, float exposure, cudaStream_t s) {
  if (n) resolve_kernel<<<(n + 255) / 256, 256, 0, s>>>(acc, fb, n, sample, exposure);
  return cudaGetLastError();
}

cudaError_t launch_pack_tiles(const float4* acc, float4* out, uint32_t n_units, uint32_t rank, uint32_t world, uint32_t tiles_x, uint32_t W,
                              uint32_t H, cudaStream_t s) {
  if (n_units) pack_tiles_kernel<<<(n_units + 255) / 256, 256, 0, s>>>(acc, out, n_units, rank, world, tiles_x, W, H);
  return cudaGetLastError();
}
cudaError_t launch_unpack_tiles(const float4* in, float4* acc, uint32_t n_units, uint32_t rank, uint32_t world, uint32_t tiles_x, uint32_t W,
                                uint32_t H, cudaStream_t s) {
  if (n_units) unpack_tiles_kernel<<<(n_units + 255) / 256, 256, 0, s>>>(in, acc, n_units, rank, world, tiles_x, W, H);
  return cudaGetLastError();
}
cudaError_t launch_combine_peers(const CombinePeers& peers, float4* comb, uint32_t* fb, uint32_t W, uint32_t H, uint32_t tiles_x,
                                 uint32_t sample, float exposure, cudaStream_t s) {
  combine_peers_kernel<<<(W * H + 255) / 256, 256, 0, s>>>(peers, comb, fb, W, H, tiles_x, sample, exposure);
  return cudaGetLastError();
}

cudaError_t launch_fp32_peak(float* out, int grid, uint32_t iters, cudaStream_t s) {
  fp32_peak_kernel<<<grid, 256, 0, s>>>(out, iters, 0.999f, 0.001f);
  return cudaGetLastError();
}

cudaError_t launch_detmath(const float* x, uint32_t n, float* sn, float* cs, float* ex, float* pw, cudaStream_t s) {
  if (n) detmath_kernel<<<(n + 255) / 256, 256, 0, s>>>(x, n, sn, cs, ex, pw);
  return cudaGetLastError();
}

}  // namespace crt
