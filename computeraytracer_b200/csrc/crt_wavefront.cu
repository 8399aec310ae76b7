// crt_wavefront.cu -- wavefront form of the hot path (sm_100a), chosen for scenes with a BVH.
//
// ncu evidence for the choice (profiles/r01a_megakernel_mesh1m_*.txt): in the megakernel only 4.75 of 32
// lanes are active per instruction on the 1M-triangle scene -- a warp's lanes are in different phases and
// only a few of them need the BVH at any moment -- while DRAM is idle (0.3 %).  Here the same per-path
// state machine (crt_path.cuh) is run one ray at a time over compacted queues:
//
//   wf_logic     one thread per live path slot: consume the hit of the previous ray (shade, NEE, bounce,
//                Russian roulette), start the next sample when the path ends, scan the linear list (patches,
//                spheres) for the new ray, and append the slot to the BVH queue only if the ray can reach
//                the triangles' bounds;
//   wf_traverse  persistent warps over the BVH queue; a lane that finishes a ray fetches the next one
//                (one warp-aggregated atomic per refill), so every trip of the loop has all lanes stepping.
//
// A pool of path slots works through the chunk's (sample, pixel) items: a slot whose path ends takes the next
// item from a global counter (one atomic per warp), so the queues stay full until the chunk runs out of work
// and several samples of a pixel are in flight at once.  A finished path stores its XYZ in a staging buffer
// indexed by (sample ordinal, pixel); wf_accumulate then
// adds the samples to the accumulator (binding 1) in frame order, as the reference's frames do (CS:108),
// so results are bit-identical to the megakernel and to n single-frame dispatches.  Queues and counters
// live in HBM; per-iteration counters are indexed by iteration, so nothing is reset inside the loop.
#include "crt_device.cuh"
#include "crt_path.cuh"
#include "crt_wavefront.h"

#define WF_BLOCK 256
#define WF_FLAG_SPECULAR 1u
#define WF_FLAG_TRANSMISSION 2u
#define WF_FLAG_SHADOW 4u
#define WF_QUERY_ANYHIT 1u

// unit -> pixel: units enumerate this rank's 8x4 tiles, 32 pixels each (the megakernel's unit order)
__device__ __forceinline__ bool unit_pixel(const DevScene& sc, uint32_t unit, uint32_t& px, uint32_t& py) {
  uint32_t tile = unit >> 5;
  const uint32_t l = unit & 31u;
  if (sc.shard_mode == CRT_SHARD_TILES && sc.world > 1) tile = tile * sc.world + sc.rank;
  px = (tile % sc.tiles_x) * 8u + (l & 7u);
  py = (tile / sc.tiles_x) * 4u + (l >> 3);
  return px < sc.W && py < sc.H;
}

__device__ __forceinline__ void wf_load(const WfState& w, uint32_t slot, PathState& ps, Rng& rng, uint32_t& s_cur) {
  const float4 a = w.S0[slot], b = w.S1[slot];
  ps.o = v3(a); ps.pdf = a.w;
  ps.d = v3(b); ps.etaScale = b.w;
  ps.L = w.S2[slot];
  ps.beta = w.S3[slot];
  const uint4 r = w.S4[slot];
  rng.x = r.x; rng.y = r.y; rng.z = r.z; rng.w = r.w;
  const uint4 m = w.S5[slot];
  ps.exclude = m.x;
  ps.depth = m.y & 0xFFFFu;
  const uint32_t fl = m.y >> 16;
  ps.specular = (fl & WF_FLAG_SPECULAR) != 0; ps.in_transmission = (fl & WF_FLAG_TRANSMISSION) != 0;
  ps.shadow_phase = (fl & WF_FLAG_SHADOW) != 0;
  ps.lam0 = m.z;
  s_cur = m.w;
  if (ps.shadow_phase) {
    const float4 n = w.S6[slot];
    ps.n = v3(n);
    const uint32_t rl = __float_as_uint(n.w);
    ps.refl = rl & 0xFFFu; ps.light = rl >> 12;
  }
}

__device__ __forceinline__ void wf_store(const WfState& w, uint32_t slot, const PathState& ps, const Rng& rng, uint32_t s_cur) {
  w.S0[slot] = make_float4(ps.o.x, ps.o.y, ps.o.z, ps.pdf);
  w.S1[slot] = make_float4(ps.d.x, ps.d.y, ps.d.z, ps.etaScale);
  w.S2[slot] = ps.L;
  w.S3[slot] = ps.beta;
  w.S4[slot] = make_uint4(rng.x, rng.y, rng.z, rng.w);
  const uint32_t fl = (ps.specular ? WF_FLAG_SPECULAR : 0u) | (ps.in_transmission ? WF_FLAG_TRANSMISSION : 0u) |
                      (ps.shadow_phase ? WF_FLAG_SHADOW : 0u);
  w.S5[slot] = make_uint4(ps.exclude, (ps.depth & 0xFFFFu) | (fl << 16), ps.lam0, s_cur);
  if (ps.shadow_phase) w.S6[slot] = make_float4(ps.n.x, ps.n.y, ps.n.z, __uint_as_float((ps.refl & 0xFFFu) | (ps.light << 12)));
}

// append `slot` to a queue with one atomic per warp
__device__ __forceinline__ void wf_push(bool pred, uint32_t slot, uint32_t* __restrict__ list, uint32_t* __restrict__ counter) {
  const uint32_t mask = __ballot_sync(0xffffffffu, pred);   // callers arrive with whole warps
  if (!pred) return;
  const uint32_t lane = threadIdx.x & 31u;
  const uint32_t leader = __ffs(mask) - 1u;
  uint32_t base = 0;
  if (lane == leader) base = atomicAdd(counter, (uint32_t)__popc(mask));
  base = __shfl_sync(mask, base, leader);
  list[base + __popc(mask & ((1u << lane) - 1u))] = slot;
}

template <bool STATS>
__global__ void __launch_bounds__(WF_BLOCK) wf_logic(const __grid_constant__ DevScene sc, const WfState w, const WfChunk ck, uint32_t it,
                                                     unsigned long long* __restrict__ counters) {
  extern __shared__ float s_tables[];
  const float* spectra = sc.spectra;
  const float* cie = sc.cie;
  if (w.tables_in_smem) {
    const uint32_t ns = sc.n_spectra * CRT_LAMBDA_RANGE;
    for (uint32_t i = threadIdx.x; i < ns; i += blockDim.x) s_tables[i] = sc.spectra[i];
    for (uint32_t i = threadIdx.x; i < 3u * 471u; i += blockDim.x) s_tables[ns + i] = sc.cie[i];
    __syncthreads();
    spectra = s_tables;
    cie = s_tables + ns;
  }
  const uint32_t n_cur = it == 0 ? ck.n_slots : w.n_active[it];
  const uint32_t* __restrict__ cur = w.list[it & 1u];
  uint32_t* __restrict__ next = w.list[(it + 1u) & 1u];
  uint32_t* work = &w.n_active[CRT_WF_MAX_ITERS + 2];   // next unassigned work item of the chunk
  const uint32_t total_work = ck.n_units * ck.n_samples; // item = ordinal * n_units + unit (sample-major)
  const uint32_t lane = threadIdx.x & 31u;

  Stats st; st.prim_tests = st.node_visits = st.tri_tests = st.rand_calls = 0;
  uint32_t n_paths = 0, n_ext = 0, n_shadow = 0, depth_sum = 0, n_bvh_rays = 0;
  uint32_t term_cnt[4] = {0, 0, 0, 0};

  // whole warps iterate together so the work fetch and the queue appends can use warp-wide ballots
  const uint32_t stride = gridDim.x * blockDim.x;
  const uint32_t n_round = (n_cur + 31u) & ~31u;
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n_round; i += stride) {
    const bool in_range = i < n_cur;
    bool alive = false, need = false, want_bvh = false;
    uint32_t slot = 0, wid = 0, px = 0, py = 0;
    PathState ps;
    Rng rng;
    // ---- consume the hit of the slot's current ray
    if (in_range) {
      slot = it == 0 ? i : cur[i];
      need = true;
      if (it != 0) {
        wf_load(w, slot, ps, rng, wid);
        const uint32_t j = wid / ck.n_units;
        unit_pixel(sc, wid - j * ck.n_units, px, py);
        const float4 hv = w.H[slot];
        Hit h; h.t = hv.x; h.idx = __float_as_uint(hv.y); h.ref = __float_as_uint(hv.z);
        uint32_t term = 0;
        alive = path_consume<STATS>(sc, spectra, ps, rng, h, st, &term);
        need = !alive;
        if (!alive) {
          uint32_t lam[4];
          wavelengths(ps.lam0, lam);
          const V3 xyz = spectral_to_xyz(cie, ps.L, lam);                      // CS:105
          const uint32_t pix = px + py * sc.W;                                // CS:107
          ck.stage[(size_t)j * ck.n_pixels + pix] = make_float4(xyz.x, xyz.y, xyz.z, 0.0f);   // added in order by wf_accumulate
          if (STATS) { term_cnt[term - TERM_MISS]++; depth_sum += ps.depth; }
        }
      }
    }
    // ---- slots whose path ended (or fresh slots) take the next (sample, pixel) work item: one atomic per warp
    for (;;) {
      const uint32_t mask = __ballot_sync(0xffffffffu, need);
      if (mask == 0u) break;
      uint32_t base = 0;
      const uint32_t leader = __ffs(mask) - 1u;
      if (lane == leader) base = atomicAdd(work, (uint32_t)__popc(mask));
      base = __shfl_sync(0xffffffffu, base, leader);
      if (need) {
        wid = base + __popc(mask & ((1u << lane) - 1u));
        if (wid >= total_work) {
          need = false;   // the chunk has no work left: this slot retires
        } else {
          const uint32_t j = wid / ck.n_units;
          if (unit_pixel(sc, wid - j * ck.n_units, px, py)) {   // edge tiles have units outside the image
            path_begin<STATS>(sc, px, py, ck.s_first + j * ck.s_stride, ps, rng, st);
            n_paths++;
            alive = true;
            need = false;
          }
        }
      }
    }
    // ---- the new ray: linear list here, BVH (if it can matter) in wf_traverse
    if (alive) {
      Hit h = scan_linear<STATS>(sc, ps.o, ps.d, ps.exclude, st);
      uint32_t query = 0;
      if (!ps.shadow_phase) {
        n_ext++;
        want_bvh = hits_root_box(sc, ps.o, ps.d, h.t);
      } else {
        n_shadow++;
        if (sc.shadow_closest) {
          want_bvh = hits_root_box(sc, ps.o, ps.d, h.t);
        } else {
          // only "is the light the closest hit?" matters (crt_device.cuh shadow_intersect)
          const uint32_t include = __float_as_uint(light_at(sc, ps.light).d1.w);
          want_bvh = h.idx == include && hits_root_box(sc, ps.o, ps.d, h.t);
          query = WF_QUERY_ANYHIT;
        }
      }
      wf_store(w, slot, ps, rng, wid);
      w.H[slot] = make_float4(h.t, __uint_as_float(h.idx), __uint_as_float(h.ref), __uint_as_float(query));
    }
    wf_push(alive, slot, next, &w.n_active[it + 1u]);
    wf_push(want_bvh, slot, w.bvh_list, &w.n_bvh[it]);
    n_bvh_rays += want_bvh ? 1u : 0u;
  }

  // ---- counters: one atomic per warp and counter
  unsigned long long v[12];
#pragma unroll
  for (int k = 0; k < 12; k++) v[k] = 0;
  v[CRT_CNT_PATHS] = n_paths; v[CRT_CNT_RAYS_EXT] = n_ext; v[CRT_CNT_RAYS_SHADOW] = n_shadow;
  if (STATS) {
    v[CRT_CNT_PRIM_TESTS] = st.prim_tests; v[CRT_CNT_RAND] = st.rand_calls;
    v[CRT_CNT_TERM_MISS] = term_cnt[0]; v[CRT_CNT_TERM_RR] = term_cnt[1]; v[CRT_CNT_TERM_LIGHT] = term_cnt[2];
    v[CRT_CNT_TERM_DEPTH] = term_cnt[3]; v[CRT_CNT_DEPTH_SUM] = depth_sum;
  }
#pragma unroll
  for (int k = 0; k < 12; k++) {
    if (!STATS && k > CRT_CNT_RAYS_SHADOW) break;
    unsigned long long x = v[k];
    for (int o = 16; o > 0; o >>= 1) x += __shfl_down_sync(0xffffffffu, x, o);
    if (lane == 0 && x) atomicAdd(&counters[k], x);
  }
  {
    unsigned long long x = n_bvh_rays;
    for (int o = 16; o > 0; o >>= 1) x += __shfl_down_sync(0xffffffffu, x, o);
    if (lane == 0 && x) atomicAdd(&counters[CRT_CNT_BVH_RAYS], x);
  }
}

// Persistent traversal of the BVH queue of iteration `it`.  Each lane owns one ray at a time; when its
// traversal ends it writes the hit back and takes the next queue entry, so the warp never waits for its
// slowest ray.  While-while form: the warp keeps stepping inner nodes as long as enough lanes have one;
// a lane that reaches a leaf parks there until the warp switches to the leaf phase, so the (long, rare)
// triangle tests do not run at one or two lanes on every trip.  Arithmetic per node and per triangle is
// trav_inner / trav_leaf (crt_device.cuh), the same code the megakernel and the trace-rays hook run.
#define WF_INNER_MIN_LANES 12
template <bool STATS>
__global__ void __launch_bounds__(WF_BLOCK) wf_traverse(const __grid_constant__ DevScene sc, const WfState w, uint32_t it,
                                                        unsigned long long* __restrict__ counters) {
  const uint32_t n = w.n_bvh[it];
  if (n == 0) return;
  const uint32_t lane = threadIdx.x & 31u;
  const uint32_t* __restrict__ queue = w.bvh_list;
  uint32_t* fetch = &w.n_fetch[it];
  const float4* __restrict__ nodes = sc.nodes;
  const float4* __restrict__ tris = sc.tri;
  const float t_min = sc.t_min;
  Stats st; st.prim_tests = st.node_visits = st.tri_tests = st.rand_calls = 0;

  int stack[CRT_BVH_STACK];
  int node = 0, sp = 0;
  bool have = false, exhausted = false, anyhit = false;
  uint32_t slot = 0, exclude = CRT_MAX_U32;
  V3 o = v3(0.0f, 0.0f, 0.0f), d = v3(0.0f, 0.0f, 1.0f);
  float idx = 0.0f, idy = 0.0f, idz = 0.0f;
  Hit h; h.t = 0.0f; h.idx = CRT_MAX_U32; h.ref = 0;

  for (;;) {
    // ---- refill the lanes that have no ray
    const bool need = !have && !exhausted;
    const uint32_t mask = __ballot_sync(0xffffffffu, need);
    if (mask) {
      uint32_t base = 0;
      const uint32_t leader = __ffs(mask) - 1u;
      if (lane == leader) base = atomicAdd(fetch, (uint32_t)__popc(mask));
      base = __shfl_sync(0xffffffffu, base, leader);
      if (need) {
        const uint32_t q = base + __popc(mask & ((1u << lane) - 1u));
        if (q >= n) {
          exhausted = true;
        } else {
          slot = queue[q];
          const float4 a = w.S0[slot], b = w.S1[slot], hv = w.H[slot];
          o = v3(a); d = v3(b);
          exclude = w.S5[slot].x;
          h.t = hv.x; h.idx = __float_as_uint(hv.y); h.ref = __float_as_uint(hv.z);
          anyhit = (__float_as_uint(hv.w) & WF_QUERY_ANYHIT) != 0u;
          idx = 1.0f / d.x; idy = 1.0f / d.y; idz = 1.0f / d.z;
          node = 0; sp = 0;
          have = true;
        }
      }
    }
    if (__ballot_sync(0xffffffffu, have) == 0u) break;
    // ---- inner-node phase: lanes at a leaf (node < 0) or without a ray wait
    bool done = false;
    for (;;) {
      const bool inner = have && !done && node >= 0;
      const uint32_t m_in = __ballot_sync(0xffffffffu, inner);
      if (__popc(m_in) < WF_INNER_MIN_LANES) {
        // too few lanes left here: go test the parked leaves / refill -- unless nobody is parked either
        const uint32_t m_leaf = __ballot_sync(0xffffffffu, have && !done && node < 0);
        const uint32_t m_idle = __ballot_sync(0xffffffffu, (!have && !exhausted) || done);
        if (m_in == 0u || m_leaf != 0u || m_idle != 0u) break;
      }
      if (inner) {
        if (!trav_inner<STATS>(nodes, o, idx, idy, idz, t_min, h, node, sp, stack, st)) done = true;
      }
    }
    // ---- leaf phase
    if (have && !done && node < 0) {
      if (!trav_leaf<STATS>(tris, o, d, exclude, t_min, anyhit, h, node, sp, stack, st)) done = true;
    }
    if (done) {
      w.H[slot] = make_float4(h.t, __uint_as_float(h.idx), __uint_as_float(h.ref), 0.0f);
      have = false;
    }
  }
  if (STATS) {
    unsigned long long v[2] = {st.node_visits, st.tri_tests};
    for (int k = 0; k < 2; k++) {
      unsigned long long x = v[k];
      for (int s = 16; s > 0; s >>= 1) x += __shfl_down_sync(0xffffffffu, x, s);
      if (lane == 0 && x) atomicAdd(&counters[CRT_CNT_NODE_VISITS + k], x);
    }
  }
}

// acc += sample_1, += sample_2, ... in frame order (CS:108 over the chunk's frames), then the epilogue of
// main (CS:109-115) with the frame counter after the chunk.
__global__ void wf_accumulate(const __grid_constant__ DevScene sc, const WfChunk ck, float4* __restrict__ acc, uint32_t* __restrict__ fb,
                              uint32_t sample_after) {
  const uint32_t unit = blockIdx.x * blockDim.x + threadIdx.x;
  if (unit >= ck.n_units) return;
  uint32_t px, py;
  if (!unit_pixel(sc, unit, px, py)) return;
  const uint32_t pix = px + py * sc.W;
  float4 a = acc[pix];
  for (uint32_t j = 0; j < ck.n_samples; j++) {
    const float4 v = ck.stage[(size_t)j * ck.n_pixels + pix];
    a.x += v.x; a.y += v.y; a.z += v.z;
  }
  acc[pix] = a;
  fb[pix] = resolve_pixel(a.x, a.y, a.z, sample_after, sc.exposure);
}

namespace crt {

size_t wf_tables_bytes(const DevScene& sc) { return ((size_t)sc.n_spectra * CRT_LAMBDA_RANGE + 3u * 471u) * sizeof(float); }

cudaError_t wf_config(bool stats, size_t smem, int* logic_blocks_per_sm, int* trav_blocks_per_sm) {
  cudaError_t e;
  if (stats) {
    if ((e = cudaFuncSetAttribute(wf_logic<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)) != cudaSuccess) return e;
    if ((e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(logic_blocks_per_sm, wf_logic<true>, WF_BLOCK, smem)) != cudaSuccess) return e;
    return cudaOccupancyMaxActiveBlocksPerMultiprocessor(trav_blocks_per_sm, wf_traverse<true>, WF_BLOCK, 0);
  }
  if ((e = cudaFuncSetAttribute(wf_logic<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)) != cudaSuccess) return e;
  if ((e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(logic_blocks_per_sm, wf_logic<false>, WF_BLOCK, smem)) != cudaSuccess) return e;
  return cudaOccupancyMaxActiveBlocksPerMultiprocessor(trav_blocks_per_sm, wf_traverse<false>, WF_BLOCK, 0);
}

cudaError_t launch_wf_logic(const DevScene& sc, const WfState& w, const WfChunk& ck, bool stats, int grid, size_t smem, uint32_t it,
                            unsigned long long* counters, cudaStream_t s) {
  if (stats) wf_logic<true><<<grid, WF_BLOCK, smem, s>>>(sc, w, ck, it, counters);
  else wf_logic<false><<<grid, WF_BLOCK, smem, s>>>(sc, w, ck, it, counters);
  return cudaGetLastError();
}

cudaError_t launch_wf_traverse(const DevScene& sc, const WfState& w, bool stats, int grid, uint32_t it, unsigned long long* counters,
                               cudaStream_t s) {
  if (stats) wf_traverse<true><<<grid, WF_BLOCK, 0, s>>>(sc, w, it, counters);
  else wf_traverse<false><<<grid, WF_BLOCK, 0, s>>>(sc, w, it, counters);
  return cudaGetLastError();
}

cudaError_t launch_wf_accumulate(const DevScene& sc, const WfChunk& ck, float4* acc, uint32_t* fb, uint32_t sample_after, cudaStream_t s) {
  if (ck.n_units) wf_accumulate<<<(ck.n_units + 255) / 256, 256, 0, s>>>(sc, ck, acc, fb, sample_after);
  return cudaGetLastError();
}

}  // namespace crt
