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

#ifndef WF_BLOCK
#define WF_BLOCK 256
#endif
#ifndef WF_LOGIC_BLOCK
#define WF_LOGIC_BLOCK WF_BLOCK
#endif
#ifndef WF_PREFETCH
#define WF_PREFETCH 11       // wf_logic<.., PF = true>, bit k: 32-byte sector k (A..E) of the record a thread handles in its NEXT trip is sent to L2
#endif
#ifndef WF_LOGIC_MIN_BLOCKS
#define WF_LOGIC_MIN_BLOCKS 3   // 3 x 256 threads per SM = 85 registers at most (ptxas picks 106 without the bound)
#endif

// One slot's record: WF_REC_WORDS 16-byte words, contiguous, so that every 32-byte DRAM sector a kernel touches is used
// in full (the first version kept one array per word: each access pulled a half-used sector from a different DRAM page,
// and the logic kernels ran at the resulting DRAM traffic, not at their instruction count).  Words that are used together
// share a sector: wf_traverse reads sectors A and B for an extension ray.
enum {
  WF_R_S0 = 0, WF_R_S1 = 1,   // sector A: ray origin + last_bounce_pdf, ray direction + etaScale
  WF_R_H = 2, WF_R_S5 = 3,    // sector B: hit of the extension ray; exclude, depth|flags|wavelength, sampled light's primitive, work item
  WF_R_S6 = 4, WF_R_S7 = 5,   // sector C: pending NEE shadow ray (direction, light distance) and contribution
  WF_R_S3 = 6, WF_R_S4 = 7,   // sector D: beta, RNG state
  WF_R_S2 = 8,                // sector E: accumulated radiance L -- on its own, because a pass touches it
                              // only when it adds to L or finishes the path (WF_FLAG_LZERO: never written yet, reads as 0)
  WF_R_S8 = 9                 // sector E, second word: only in scenes where a light is a triangle (sc.shadow_closest) -- what the
                              // deferred NEE contribution towards such a light is rebuilt from (NeeQuery::aux)
};
// Record accessors.  (Measured and rejected: sending the record stream through L2 with an evict-first policy -- wf_traverse
// and the next wf_logic find the records in L2 today, profiles/r01t_static_l2_tri2_variants.txt.)
__device__ __forceinline__ float4* rec_ptr(const WfState& w, uint32_t slot, int word) { return &w.rec[(size_t)slot * WF_REC_WORDS + word]; }
__device__ __forceinline__ float4 rec_ld(const WfState& w, uint32_t slot, int word) { return *rec_ptr(w, slot, word); }
__device__ __forceinline__ uint4 rec_ldu(const WfState& w, uint32_t slot, int word) { return *reinterpret_cast<const uint4*>(rec_ptr(w, slot, word)); }
__device__ __forceinline__ void rec_st(const WfState& w, uint32_t slot, int word, float4 v) { *rec_ptr(w, slot, word) = v; }
__device__ __forceinline__ void rec_stu(const WfState& w, uint32_t slot, int word, uint4 v) { *reinterpret_cast<uint4*>(rec_ptr(w, slot, word)) = v; }
#define WF_FLAG_SPECULAR 1u
#define WF_FLAG_TRANSMISSION 2u
#define WF_FLAG_PENDING 4u            // an NEE contribution (S7) waits for the any-hit answer of its shadow ray (S6)
#define WF_FLAG_DEAD 8u               // Russian roulette ended the path; only the pending contribution is outstanding
#define WF_FLAG_LZERO 0x800u          // (bit 31 of the packed word) L is still 0 and its sector has not been written for this path

__device__ __forceinline__ void wf_load(const WfState& w, uint32_t slot, PathState& ps, Rng& rng, uint32_t& wid, uint32_t& flags) {
  const float4 a = rec_ld(w, slot, WF_R_S0), b = rec_ld(w, slot, WF_R_S1);
  ps.o = v3(a); ps.pdf = a.w;
  ps.d = v3(b); ps.etaScale = b.w;
  ps.L = make_float4(0.0f, 0.0f, 0.0f, 0.0f);   // not loaded here: wf_load_L
  ps.beta = rec_ld(w, slot, WF_R_S3);
  const uint4 r = rec_ldu(w, slot, WF_R_S4);
  rng.x = r.x; rng.y = r.y; rng.z = r.z; rng.w = r.w;
  const uint4 m = rec_ldu(w, slot, WF_R_S5);
  ps.exclude = m.x;
  ps.depth = m.y & 0xFFFFu;
  flags = ((m.y >> 16) & 0xFu) | ((m.y >> 20) & WF_FLAG_LZERO);
  ps.lam0 = (m.y >> 20) & 0x7FFu;
  ps.specular = (flags & WF_FLAG_SPECULAR) != 0; ps.in_transmission = (flags & WF_FLAG_TRANSMISSION) != 0;
  ps.shadow_phase = false;
  wid = m.w;
}

__device__ __forceinline__ float4 wf_load_L(const WfState& w, uint32_t slot, uint32_t flags) {
  return (flags & WF_FLAG_LZERO) ? make_float4(0.0f, 0.0f, 0.0f, 0.0f) : rec_ld(w, slot, WF_R_S2);
}

// everything but L (the caller stores L when it changed it, and then clears WF_FLAG_LZERO)
__device__ __forceinline__ void wf_store(const WfState& w, uint32_t slot, const PathState& ps, const Rng& rng, uint32_t wid, uint32_t flags,
                                         uint32_t light_prim) {
  rec_st(w, slot, WF_R_S0, make_float4(ps.o.x, ps.o.y, ps.o.z, ps.pdf));
  rec_st(w, slot, WF_R_S1, make_float4(ps.d.x, ps.d.y, ps.d.z, ps.etaScale));
  rec_st(w, slot, WF_R_S3, ps.beta);
  rec_stu(w, slot, WF_R_S4, make_uint4(rng.x, rng.y, rng.z, rng.w));
  const uint32_t fl = (flags & (WF_FLAG_PENDING | WF_FLAG_DEAD)) | (ps.specular ? WF_FLAG_SPECULAR : 0u) | (ps.in_transmission ? WF_FLAG_TRANSMISSION : 0u);
  rec_stu(w, slot, WF_R_S5, make_uint4(ps.exclude, (ps.depth & 0xFFFFu) | (fl << 16) | (ps.lam0 << 20) | ((flags & WF_FLAG_LZERO) << 20), light_prim, wid));
}

// Appending to a global queue costs one atomic per warp, in two steps so that several queues can be served with the
// atomics in flight together: wf_reserve issues the atomic (leader lane), wf_commit waits for it and writes the entries.
// Callers arrive with whole warps.
struct PushTicket { uint32_t mask, base; };

__device__ __forceinline__ PushTicket wf_reserve(bool pred, uint32_t* __restrict__ counter) {
  PushTicket t;
  t.mask = __ballot_sync(0xffffffffu, pred);
  t.base = 0;
  if (t.mask != 0u && (threadIdx.x & 31u) == (uint32_t)__ffs(t.mask) - 1u) t.base = atomicAdd(counter, (uint32_t)__popc(t.mask));
  return t;
}

__device__ __forceinline__ void wf_commit(const PushTicket& t, bool pred, uint32_t entry, uint32_t* __restrict__ list) {
  if (t.mask == 0u) return;   // warp-uniform
  const uint32_t base = __shfl_sync(0xffffffffu, t.base, __ffs(t.mask) - 1);
  if (pred) list[base + __popc(t.mask & ((1u << (threadIdx.x & 31u)) - 1u))] = entry;
}

__device__ __forceinline__ void wf_push(bool pred, uint32_t entry, uint32_t* __restrict__ list, uint32_t* __restrict__ counter) {
  const PushTicket t = wf_reserve(pred, counter);
  wf_commit(t, pred, entry, list);
}

// The shadow ray of next-event estimation that still needs the BVH: the light is the closest hit of the linear list at
// distance t, the contribution c = (BRDF * Le * w / pdf_l) * beta is final unless a triangle occludes the segment.
// When the sampled light's primitive is itself a triangle, the linear list cannot have found it: the query is a CLOSEST-hit
// traversal that starts from the linear list's hit (t < 0 in the record marks it: bound -t), and the contribution can only be
// computed once the hit on the light is known (CS:393 needs its normal and position).  Then c carries beta as it was at the
// bounce and aux = (cos_theta, reflectance index | light << 12, -, index of the linear list's hit); wf_tri_contribution
// rebuilds BRDF, Le, the MIS weight and the product in the reference's order when the answer is there.
// What an occluded query adds is (BRDF * 0) * beta (CS:407 then CS:187), not c * 0: the two differ when Le is not finite (a light
// of zero area -- a sphere listed as a light -- has an infinite pdf) while BRDF and beta are.  A channel of c that is not finite
// and would be 0 when occluded is stored as a NaN with a payload that arithmetic never produces (it returns 0x7FFFFFFF); the
// payload says what the visible value was.  nee_resolve undoes it on either outcome.
#define WF_NEE_NAN_ZERO 0x7FC0D0D0u   // visible: NaN, occluded: 0
#define WF_NEE_PINF_ZERO 0x7FC0D0D1u  // visible: +inf, occluded: 0
#define WF_NEE_NINF_ZERO 0x7FC0D0D2u  // visible: -inf, occluded: 0
__device__ __forceinline__ bool nee_nonfinite(float x) { return (__float_as_uint(x) & 0x7F800000u) == 0x7F800000u; }
__device__ __forceinline__ float nee_encode1(float c, float brdf, float beta) {
  if (!nee_nonfinite(c)) return c;
  const float c0 = (brdf * 0.0f) * beta;
  if (c0 != c0) return c;   // NaN when occluded too: c * 0 says the same
  return __uint_as_float(c != c ? WF_NEE_NAN_ZERO : (c > 0.0f ? WF_NEE_PINF_ZERO : WF_NEE_NINF_ZERO));
}
// (the rare paths are calls: inlined they cost wf_logic registers it does not have)
__device__ __noinline__ float4 nee_encode_slow(float4 c, float4 brdf, float4 beta) {
  return make_float4(nee_encode1(c.x, brdf.x, beta.x), nee_encode1(c.y, brdf.y, beta.y), nee_encode1(c.z, brdf.z, beta.z), nee_encode1(c.w, brdf.w, beta.w));
}
#ifndef WF_NEE_EXACT
#define WF_NEE_EXACT 1   // A/B switch: 0 = c * 0 for an occluded query, as before (wrong bits when Le is not finite)
#endif
__device__ __forceinline__ float4 nee_encode(float4 c, float4 brdf, float4 beta) {
  if (!WF_NEE_EXACT) return c;
  if (!nee_nonfinite((c.x + c.y) + (c.z + c.w))) return c;   // all channels finite (a sum that overflows takes the call, which leaves finite channels alone)
  return nee_encode_slow(c, brdf, beta);
}
__device__ __forceinline__ float nee_resolve1(float c, bool occluded) {
  const uint32_t b = __float_as_uint(c);
  if (b - WF_NEE_NAN_ZERO <= 2u) {
    if (occluded) return 0.0f;
    return b == WF_NEE_NAN_ZERO ? __uint_as_float(0x7FFFFFFFu) : __uint_as_float(b == WF_NEE_PINF_ZERO ? 0x7F800000u : 0xFF800000u);
  }
  return occluded ? c * 0.0f : c;
}
__device__ __noinline__ float4 nee_resolve_slow(float4 c, bool occluded) {
  return make_float4(nee_resolve1(c.x, occluded), nee_resolve1(c.y, occluded), nee_resolve1(c.z, occluded), nee_resolve1(c.w, occluded));
}
// the contribution of a (non-triangle-light) query once its answer is known
__device__ __forceinline__ float4 nee_resolve(float4 c, bool occluded) {
  if (WF_NEE_EXACT && nee_nonfinite((c.x + c.y) + (c.z + c.w))) return nee_resolve_slow(c, occluded);
  if (occluded) { c.x *= 0.0f; c.y *= 0.0f; c.z *= 0.0f; c.w *= 0.0f; }
  return c;
}

struct NeeQuery {
  bool pending;
  bool tri;
  V3 d;
  float t;
  uint32_t light_prim;
  float4 c;
  float4 aux;
};

// CS:388-400 + CS:187 for a shadow ray whose closest hit turned out to be the sampled triangle light: origin o, direction d,
// hit (t, light_prim, ref) from the traversal; beta / cos_theta / reflectance / light as saved at the bounce.
__device__ __forceinline__ float4 wf_tri_contribution(const DevScene& sc, const float* __restrict__ spectra, V3 o, V3 d, float t, uint32_t light_prim,
                                                      uint32_t ref, float4 beta, float4 aux, uint32_t lam0, bool visible) {
  uint32_t lam[4];
  wavelengths(lam0, lam);
  const float cos_theta = aux.x;
  const uint32_t packed = __float_as_uint(aux.y);
  const float4 refl = sample_spectrum(spectra, sc.n_spectra, packed & 0xFFFu, lam);
  const float4 BRDF = make_float4(refl.x / CRT_PI_F, refl.y / CRT_PI_F, refl.z / CRT_PI_F, refl.w / CRT_PI_F);  // CS:184
  float4 le = make_float4(0.0f, 0.0f, 0.0f, 0.0f);                                                             // CS:407
  if (visible) {
    const LightRec& light = light_at(sc, (packed >> 12) & 0xFFFu);
    const float4 spec = sample_spectrum(spectra, sc.n_spectra, __float_as_uint(light.d2.w), lam);
    le = make_float4(spec.x * cos_theta, spec.y * cos_theta, spec.z * cos_theta, spec.w * cos_theta);
    Hit hs; hs.t = t; hs.idx = light_prim; hs.ref = ref;
    const Surface sh = surface_of(sc, hs, o, d);
    const float pdf_l = compute_light_pdf(sc, CRT_META_EMIS(sh.meta), sh.normal, sh.position, o, d);            // CS:393
    const float pdf_b = crt_div(cos_theta, CRT_PI_F);
    const float w = power_heuristic(pdf_l, pdf_b);
    le = make_float4(crt_div(le.x * w, pdf_l), crt_div(le.y * w, pdf_l), crt_div(le.z * w, pdf_l), crt_div(le.w * w, pdf_l));  // CS:400
  }
  return make_float4((BRDF.x * le.x) * beta.x, (BRDF.y * le.y) * beta.y, (BRDF.z * le.z) * beta.z, (BRDF.w * le.w) * beta.w);    // CS:187
}

// One bounce of path_trace's loop body (CS:133-293), entered with the hit `h` of the current extension ray and left with
// the next extension ray in ps.o / ps.d / ps.exclude.  Unlike path_consume (crt_path.cuh), next-event estimation is not a
// separate phase: nothing after the shadow ray depends on its outcome except the sum L += BRDF * Le * beta (CS:187) --
// the light sample, the hemisphere sample, beta and Russian roulette draw the same random numbers either way -- so the
// bounce is computed in one go and the shadow ray travels with the extension ray.  The linear list answers the shadow
// query here (CS:697-705: is the closest hit the sampled light?); if it is, and the segment can reach the triangles'
// bounds, the any-hit BVH query is handed to the caller in `nee` and the contribution is added when its answer arrives,
// before anything else is added to L, so the order of the float additions is the reference's.
// Returns 0: the path ended; 1: it continues with an extension ray; 2: Russian roulette ended it while nee.pending
// (the slot must live on until the contribution has been added).
template <bool STATS, bool TRI, bool BVH = true>   // BVH = false: the scene has no triangles, so no NEE query is ever left pending
__device__ __forceinline__ int wf_bounce(const DevScene& sc, const float* __restrict__ spectra, PathState& ps, Rng& rng, const Hit& h,
                                         Stats& st, uint32_t* term, uint32_t& n_shadow, NeeQuery& nee) {
  uint32_t lam[4];
  wavelengths(ps.lam0, lam);
  nee.pending = false; nee.tri = false;
  if (h.idx == CRT_MAX_U32) { if (STATS) term[0] = TERM_MISS; return 0; }      // CS:141-144
  ps.exclude = h.idx;                                                         // CS:146
  const Surface s = surface_of(sc, h, ps.o, ps.d);
  const uint32_t material = CRT_META_MAT(s.meta);
  if (material == CRT_MAT_LIGHT) {                                            // CS:149-164
    const float4 le = sample_spectrum(spectra, sc.n_spectra, CRT_META_EMIS(s.meta), lam);
    if (ps.depth == 0 || ps.specular) {
      ps.L.x += ps.beta.x * le.x; ps.L.y += ps.beta.y * le.y; ps.L.z += ps.beta.z * le.z; ps.L.w += ps.beta.w * le.w;
    } else {
      const float pdf_l = compute_light_pdf(sc, CRT_META_EMIS(s.meta), s.normal, s.position, ps.o, ps.d);
      const float w = power_heuristic(ps.pdf, pdf_l);
      ps.L.x += (w * le.x) * ps.beta.x; ps.L.y += (w * le.y) * ps.beta.y;
      ps.L.z += (w * le.z) * ps.beta.z; ps.L.w += (w * le.w) * ps.beta.w;
    }
    if (STATS) term[0] = TERM_LIGHT;
    return 0;
  }
  if (ps.depth >= sc.max_depth) { if (STATS) term[0] = TERM_DEPTH; return 0; }  // CS:167-170
  if (ps.in_transmission) {                                                   // CS:173-179
    const float distance = length(ps.o - s.position);
    const float4 ext = sample_spectrum(spectra, sc.n_spectra, sc.n_spectra - 1, lam);
    ps.beta.x *= crt_expf(-ext.x * distance); ps.beta.y *= crt_expf(-ext.y * distance);
    ps.beta.z *= crt_expf(-ext.z * distance); ps.beta.w *= crt_expf(-ext.w * distance);
  }
  if (material == CRT_MAT_DIFFUSE) {                                          // CS:182-204
    // sample_lights + sample_light, CS:341-355
    const float u0 = rnd(rng);
    uint32_t li = f2u_sat((float)sc.n_lights * u0);
    if (li >= sc.n_lights) li = sc.n_lights - 1;
    const LightRec& light = light_at(sc, li);
    const float lu = rnd(rng), lv = rnd(rng);
    const V3 point = madd(madd(v3(light.d1), lu, v3(light.d2)), lv, v3(light.d3));
    const V3 light_dir = normalize(point - s.position);                       // CS:385
    // compute_light_radiance, CS:379-408
    n_shadow++;
    const uint32_t include = __float_as_uint(light.d1.w);
    const Hit hs = scan_linear<STATS>(sc, s.position, light_dir, ps.exclude, st);   // shadow_intersect, CS:697-705 (linear list)
    const float4 refl = sample_spectrum(spectra, sc.n_spectra, CRT_META_REFL(s.meta), lam);
    const float4 BRDF = make_float4(refl.x / CRT_PI_F, refl.y / CRT_PI_F, refl.z / CRT_PI_F, refl.w / CRT_PI_F);  // CS:184
    const float cos_theta = fmaxf(0.0f, dot(s.normal, light_dir));             // CS:388
    const float4 spec = sample_spectrum(spectra, sc.n_spectra, __float_as_uint(light.d2.w), lam);
    float4 le = make_float4(spec.x * cos_theta, spec.y * cos_theta, spec.z * cos_theta, spec.w * cos_theta);
    bool add_now = true;
    if (hs.idx == include) {
      const Surface sh = surface_of(sc, hs, s.position, light_dir);
      const float pdf_l = compute_light_pdf(sc, CRT_META_EMIS(sh.meta), sh.normal, sh.position, s.position, light_dir);  // CS:393
      const float pdf_b = crt_div(cos_theta, CRT_PI_F);
      const float w = power_heuristic(pdf_l, pdf_b);
      le = make_float4(crt_div(le.x * w, pdf_l), crt_div(le.y * w, pdf_l), crt_div(le.z * w, pdf_l), crt_div(le.w * w, pdf_l));  // CS:400
      if (BVH && hits_root_box(sc, s.position, light_dir, hs.t)) {   // a triangle may still occlude: any-hit query bounded by the light
        add_now = false;
        nee.pending = true; nee.d = light_dir; nee.t = hs.t; nee.light_prim = include;
      }
    } else {
      le = make_float4(0.0f, 0.0f, 0.0f, 0.0f);                                // CS:407
      if (TRI && sc.shadow_closest && ((sc.light_tri_mask[li >> 5] >> (li & 31u)) & 1u) && hits_root_box(sc, s.position, light_dir, hs.t)) {
        // the sampled light is a triangle: only a closest-hit traversal (bounded by the linear list's hit) can find it
        add_now = false;
        nee.pending = true; nee.tri = true; nee.d = light_dir; nee.t = -hs.t; nee.light_prim = include;
        nee.aux = make_float4(cos_theta, __uint_as_float((CRT_META_REFL(s.meta) & 0xFFFu) | ((li & 0xFFFu) << 12) | 0x80000000u), 0.0f,
                              __uint_as_float(hs.idx));
      }
    }
    const float4 c = make_float4((BRDF.x * le.x) * ps.beta.x, (BRDF.y * le.y) * ps.beta.y, (BRDF.z * le.z) * ps.beta.z,
                                 (BRDF.w * le.w) * ps.beta.w);                 // CS:187
    if (add_now) { ps.L.x += c.x; ps.L.y += c.y; ps.L.z += c.z; ps.L.w += c.w; }
    else nee.c = nee.tri ? ps.beta : nee_encode(c, BRDF, ps.beta);   // (a triangle light's contribution is computed when its hit is known: beta as it is now)
    const float u = rnd(rng), v = rnd(rng);
    if (STATS) st.rand_calls += 5;
    float pdf;
    const V3 nd = cosine_hemisphere(s.normal, u, v, pdf);                      // CS:189
    ps.pdf = pdf;
    const float cos_b = fabsf(dot(s.normal, nd));                              // CS:190
    ps.beta.x *= (BRDF.x * cos_b) / pdf; ps.beta.y *= (BRDF.y * cos_b) / pdf;  // CS:192
    ps.beta.z *= (BRDF.z * cos_b) / pdf; ps.beta.w *= (BRDF.w * cos_b) / pdf;
    ps.o = s.position; ps.d = nd;
    ps.specular = false;
  } else if (material == CRT_MAT_GLASS) {                                     // CS:208-276
    const float eta1 = 1.0f, eta2 = 1.5f;
    float eta = eta1 / eta2;
    const float cos_theta = dot(s.normal, ps.d);
    const float reflected = fresnel_s(ps.d, s.normal, eta1, eta2);
    const float pr = reflected, pt = 1.0f - reflected;
    const float u = rnd(rng);
    if (STATS) st.rand_calls += 1;
    V3 current_normal = s.normal;
    if (cos_theta > 0.0f) { eta = 1.0f / eta; current_normal = -current_normal; }
    V3 nd;
    if (u < pr / (pr + pt)) {
      nd = reflect(ps.d, current_normal);
    } else {
      nd = normalize(refract(ps.d, current_normal, eta));
      const float e2 = eta * eta;
      ps.beta.x *= e2; ps.beta.y *= e2; ps.beta.z *= e2; ps.beta.w *= e2;
      ps.etaScale /= e2;
      ps.in_transmission = !ps.in_transmission;
    }
    ps.o = s.position; ps.d = nd;
    ps.specular = true; ps.exclude = CRT_MAX_U32;
  }
  // Russian roulette, CS:279-289
  const float rx = ps.beta.x * ps.etaScale, ry = ps.beta.y * ps.etaScale, rz = ps.beta.z * ps.etaScale;
  const float mb = fmaxf(rx, fmaxf(ry, rz));
  if (ps.depth > 1 && mb < 1.0f) {
    const float q = fmaxf(0.0f, 1.0f - mb);
    if (STATS) st.rand_calls += 1;
    if (rnd(rng) < q) { if (STATS) term[0] = TERM_RR; return nee.pending ? 2 : 0; }
    const float k = 1.0f - q;
    ps.beta.x /= k; ps.beta.y /= k; ps.beta.z /= k; ps.beta.w /= k;
  }
  ps.depth++;                                                                  // CS:292
  return 1;
}

// end of a path: spectral_to_xyz (CS:105) into the staging buffer at (sample ordinal, pixel); wf_accumulate adds in order
__device__ __forceinline__ void wf_finish(const DevScene& /*sc*/, const WfChunk& ck, const float* __restrict__ cie, float4 L, uint32_t lam0, uint32_t wid) {
  uint32_t lam[4];
  wavelengths(lam0, lam);
  const V3 xyz = spectral_to_xyz(cie, L, lam);
  // staged by work item = (sample ordinal, unit): a rank that owns 1/world of the tiles stages 1/world of the bytes, and a
  // warp's 32 units (one 8x4 tile) are 512 contiguous bytes for wf_accumulate
  ck.stage[wid] = make_float4(xyz.x, xyz.y, xyz.z, 0.0f);   // CS:107
}

// the pending NEE contribution of a slot, as the reference would have added it (CS:187 with CS:400 or CS:407)
template <bool TRI>
__device__ __forceinline__ float4 wf_pending(const DevScene& sc, const float* __restrict__ spectra, const WfState& w, uint32_t slot) {
  const float4 q = rec_ld(w, slot, WF_R_S6);
  const bool occluded = q.w < 0.0f;   // wf_traverse marks an occluded segment by a negative length
  float4 c = rec_ld(w, slot, WF_R_S7);
  if (TRI && sc.shadow_closest) {
    const float4 aux = rec_ld(w, slot, WF_R_S8);
    if (__float_as_uint(aux.y) & 0x80000000u) {   // the query went to a triangle light: c is beta, the contribution is built now
      const float4 a = rec_ld(w, slot, WF_R_S0);
      const uint4 m = rec_ldu(w, slot, WF_R_S5);
      return wf_tri_contribution(sc, spectra, v3(a), v3(q), q.w, m.z, __float_as_uint(aux.z), c, aux, (m.y >> 20) & 0x7FFu, !occluded);
    }
  }
  return nee_resolve(c, occluded);   // CS:407: Le = 0, the sum still happens
}
template <bool TRI>
__device__ __forceinline__ void wf_add_pending(const DevScene& sc, const float* __restrict__ spectra, const WfState& w, uint32_t slot, float4& L) {
  const float4 c = wf_pending<TRI>(sc, spectra, w, slot);
  L.x += c.x; L.y += c.y; L.z += c.z; L.w += c.w;
}

struct LogicCounts {
  uint32_t n_paths, n_ext, n_shadow, depth_sum, n_bvh_rays;
  uint32_t term_cnt[4];
};

__device__ __forceinline__ void wf_counts_init(LogicCounts& lc, Stats& st) {
  st.prim_tests = st.node_visits = st.tri_tests = st.rand_calls = 0;
  lc.n_paths = lc.n_ext = lc.n_shadow = lc.depth_sum = lc.n_bvh_rays = 0;
  lc.term_cnt[0] = lc.term_cnt[1] = lc.term_cnt[2] = lc.term_cnt[3] = 0;
}

// one atomic per warp and counter
template <bool STATS>
__device__ __forceinline__ void wf_counts_flush(const LogicCounts& lc, const Stats& st, unsigned long long* __restrict__ counters) {
  const uint32_t lane = threadIdx.x & 31u;
  unsigned long long v[12];
#pragma unroll
  for (int k = 0; k < 12; k++) v[k] = 0;
  v[CRT_CNT_PATHS] = lc.n_paths; v[CRT_CNT_RAYS_EXT] = lc.n_ext; v[CRT_CNT_RAYS_SHADOW] = lc.n_shadow;
  if (STATS) {
    v[CRT_CNT_PRIM_TESTS] = st.prim_tests; v[CRT_CNT_RAND] = st.rand_calls;
    v[CRT_CNT_TERM_MISS] = lc.term_cnt[0]; v[CRT_CNT_TERM_RR] = lc.term_cnt[1]; v[CRT_CNT_TERM_LIGHT] = lc.term_cnt[2];
    v[CRT_CNT_TERM_DEPTH] = lc.term_cnt[3]; v[CRT_CNT_DEPTH_SUM] = lc.depth_sum;
  }
#pragma unroll
  for (int k = 0; k < 12; k++) {
    if (!STATS && k > CRT_CNT_RAYS_SHADOW) break;
    unsigned long long x = v[k];
    for (int o = 16; o > 0; o >>= 1) x += __shfl_down_sync(0xffffffffu, x, o);
    if (lane == 0 && x) atomicAdd(&counters[k], x);
  }
  unsigned long long x = lc.n_bvh_rays;
  for (int o = 16; o > 0; o >>= 1) x += __shfl_down_sync(0xffffffffu, x, o);
  if (lane == 0 && x) atomicAdd(&counters[CRT_CNT_BVH_RAYS], x);
}

// A live slot leaves a logic kernel.  Its new extension ray meets the linear list here; if it can reach the triangles'
// bounds it is queued for the BVH (wf_traverse then routes the slot by the final hit), otherwise the hit is final and the
// slot goes straight to the next iteration's shade queue (surface hit) or finish queue (miss).  A slot that only waits for
// an NEE answer goes to the finish queue.  Called by whole warps.
// `mode` says how the live slots reach the next iteration: WF_LIVE_LIST = one live list in processing order (bounce[]);
// WF_LIVE_ROUTED = pushed here (and by wf_traverse) to the bounce or finish list by the slot's final hit; WF_LIVE_CLASS =
// no list at all: the slot's class byte is written (w.cls: 0 bounce, 1 finish, 2 retired; wf_traverse corrects it when a
// triangle turns a miss into a hit) and wf_partition builds both lists in slot order after the traversal.
enum { WF_LIVE_LIST = 0, WF_LIVE_ROUTED = 1, WF_LIVE_CLASS = 2, WF_LIVE_DEFER = 0x100 };
template <bool STATS, bool TRI>
__device__ __forceinline__ void wf_emit(const DevScene& sc, const WfState& w, uint32_t it, uint32_t mode, bool valid, bool alive, bool has_ext, uint32_t slot,
                                        const PathState& ps, const Rng& rng, uint32_t wid, uint32_t flags, const NeeQuery& nee,
                                        Stats& st, LogicCounts& lc) {
  bool want_bvh = false, to_bounce = false, lin_hit = false;
  if (has_ext) {
    const Hit h = scan_linear<STATS>(sc, ps.o, ps.d, ps.exclude, st);
    lc.n_ext++;
    want_bvh = hits_root_box(sc, ps.o, ps.d, h.t);
    lin_hit = h.idx != CRT_MAX_U32;
    to_bounce = !want_bvh && lin_hit;
    rec_st(w, slot, WF_R_H, make_float4(h.t, __uint_as_float(h.idx), __uint_as_float(h.ref), 0.0f));
  }
  if (mode == WF_LIVE_LIST) to_bounce = alive;   // unrouted single-kernel form: one live list
  bool to_finish = alive && !want_bvh && !to_bounce;
  const bool shadow = alive && nee.pending;
  if (mode == WF_LIVE_CLASS) {
    // class of the slot's next pass as far as the linear list knows it: a hit there stays a hit whatever the BVH adds
    if (valid) w.cls[slot] = !alive ? (uint8_t)2 : (lin_hit ? (uint8_t)0 : (uint8_t)1);
    to_bounce = false; to_finish = false;
  }
  if (alive) {
    wf_store(w, slot, ps, rng, wid, flags, nee.light_prim);
    if (shadow) {
      rec_st(w, slot, WF_R_S6, make_float4(nee.d.x, nee.d.y, nee.d.z, nee.t));
      rec_st(w, slot, WF_R_S7, nee.c);
      if (TRI && sc.shadow_closest) rec_st(w, slot, WF_R_S8, nee.tri ? nee.aux : make_float4(0.0f, 0.0f, 0.0f, 0.0f));
    }
  }
  const uint32_t nx = (it + 1u) & 1u;
  // the queue atomics are issued together and waited for together
  const PushTicket tb = wf_reserve(to_bounce, &w.iter[it + 1u].n_bounce);
  const PushTicket tf = wf_reserve(to_finish, &w.iter[it + 1u].n_finish);
  const PushTicket tq = wf_reserve(want_bvh, &w.iter[it].n_bvh);
  const PushTicket ts = wf_reserve(shadow, &w.iter[it].n_shadow);
  wf_commit(tb, to_bounce, slot, nx ? w.bounce[1] : w.bounce[0]);
  wf_commit(tf, to_finish, slot, nx ? w.finish[1] : w.finish[0]);
  wf_commit(tq, want_bvh, slot, w.bvh_list);
  wf_commit(ts, shadow, slot, w.shadow_list);
  lc.n_bvh_rays += (want_bvh ? 1u : 0u) + (shadow ? 1u : 0u);
}

__device__ __forceinline__ void wf_tables(const DevScene& sc, const WfState& w, float* s_tables, const float*& spectra, const float*& cie) {
  spectra = sc.spectra;
  cie = sc.cie;
  if (w.tables_in_smem) {
    const uint32_t ns = sc.n_spectra * CRT_LAMBDA_RANGE;
    for (uint32_t i = threadIdx.x; i < ns; i += blockDim.x) s_tables[i] = sc.spectra[i];
    for (uint32_t i = threadIdx.x; i < 3u * 471u; i += blockDim.x) s_tables[ns + i] = sc.cie[i];
    __syncthreads();
    spectra = s_tables;
    cie = s_tables + ns;
  }
}

// wf_shade: one bounce (wf_bounce) for every slot of the iteration's shade queue -- all lanes of a warp shade.  A path that
// ends here (light hit, depth limit, Russian roulette with nothing outstanding) is finished and its slot handed to
// wf_regen through the regen queue.
#ifndef WF_SHADE_MIN_BLOCKS
#define WF_SHADE_MIN_BLOCKS 1
#endif
#ifndef WF_REGEN_MIN_BLOCKS
#define WF_REGEN_MIN_BLOCKS 1
#endif
#ifndef WF_TRAV_MIN_BLOCKS
#define WF_TRAV_MIN_BLOCKS 4
#endif
template <bool STATS>
__global__ void __launch_bounds__(WF_BLOCK, WF_SHADE_MIN_BLOCKS) wf_shade(const __grid_constant__ DevScene sc, const WfState w, const WfChunk ck, uint32_t it,
                                                     uint32_t live_mode, unsigned long long* __restrict__ counters) {
  extern __shared__ float s_tables[];
  const float *spectra, *cie;
  wf_tables(sc, w, s_tables, spectra, cie);
  const uint32_t n_cur = w.iter[it].n_bounce;
  const uint32_t* __restrict__ cur = (it & 1u) ? w.bounce[1] : w.bounce[0];
  Stats st;
  LogicCounts lc;
  wf_counts_init(lc, st);
  const uint32_t stride = gridDim.x * blockDim.x;
  const uint32_t n_round = (n_cur + 31u) & ~31u;
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n_round; i += stride) {
    bool alive = false, has_ext = false, ended = false;
    uint32_t slot = 0, wid = 0, flags = 0;
    PathState ps;
    Rng rng;
    NeeQuery nee; nee.pending = false; nee.tri = false; nee.light_prim = 0;
    if (i < n_cur) {
      slot = cur[i];
      wf_load(w, slot, ps, rng, wid, flags);
      ps.L = wf_load_L(w, slot, flags);
      if (flags & WF_FLAG_PENDING) wf_add_pending<true>(sc, spectra, w, slot, ps.L);   // before anything else is added to L
      const float4 hv = rec_ld(w, slot, WF_R_H);
      Hit h; h.t = hv.x; h.idx = __float_as_uint(hv.y); h.ref = __float_as_uint(hv.z);
      uint32_t term = 0;
      const int r = wf_bounce<STATS, true>(sc, spectra, ps, rng, h, st, &term, lc.n_shadow, nee);
      if (STATS && r != 1) { lc.term_cnt[term - TERM_MISS]++; lc.depth_sum += ps.depth; }
      flags = (nee.pending ? WF_FLAG_PENDING : 0u) | (r == 2 ? WF_FLAG_DEAD : 0u);
      alive = r != 0;
      has_ext = r == 1;
      ended = r == 0;
      if (ended) wf_finish(sc, ck, cie, ps.L, ps.lam0, wid);
      else rec_st(w, slot, WF_R_S2, ps.L);   // (the new flags carry no WF_FLAG_LZERO)
    }
    const PushTicket tr = wf_reserve(ended, &w.iter[it].n_regen);   // in flight while wf_emit scans the linear list
    wf_emit<STATS, true>(sc, w, it, live_mode, i < n_cur, alive, has_ext, slot, ps, rng, wid, flags, nee, st, lc);
    wf_commit(tr, ended, slot, w.regen);
  }
  wf_counts_flush<STATS>(lc, st, counters);
}

// wf_regen: the free slots of the iteration take the next (sample, pixel) work items (one atomic per warp) and start
// their paths: seed, camera ray, wavelengths (CS:98-101).  Free slots are: every slot in iteration 0; afterwards the
// iteration's finish queue (extension ray missed, CS:141-144, or only an NEE answer was outstanding) -- finished here,
// which is cheap -- followed by the slots wf_shade freed.  A slot retires when the chunk has no work left.
template <bool STATS>
__global__ void __launch_bounds__(WF_BLOCK, WF_REGEN_MIN_BLOCKS) wf_regen(const __grid_constant__ DevScene sc, const WfState w, const WfChunk ck, uint32_t it,
                                                     uint32_t live_mode, unsigned long long* __restrict__ counters) {
  extern __shared__ float s_tables[];
  const float *spectra, *cie;
  wf_tables(sc, w, s_tables, spectra, cie);
  const uint32_t n_fin = it == 0 ? 0u : w.iter[it].n_finish;
  const uint32_t n_cur = it == 0 ? ck.n_slots : n_fin + w.iter[it].n_regen;
  const uint32_t* __restrict__ fin = (it & 1u) ? w.finish[1] : w.finish[0];
  const uint32_t total_work = ck.n_units * ck.n_samples;   // item = ordinal * n_units + unit (sample-major)
  const uint32_t lane = threadIdx.x & 31u;
  Stats st;
  LogicCounts lc;
  wf_counts_init(lc, st);
  const uint32_t stride = gridDim.x * blockDim.x;
  const uint32_t n_round = (n_cur + 31u) & ~31u;
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n_round; i += stride) {
    bool need = i < n_cur, alive = false;
    uint32_t slot = 0, wid = 0;
    // the work-item atomic of the first fetch is in flight while the ended paths are finished
    uint32_t mask = __ballot_sync(0xffffffffu, need);
    uint32_t first = 0;
    if (mask != 0u && lane == (uint32_t)__ffs(mask) - 1u) first = atomicAdd(w.work, (uint32_t)__popc(mask));
    if (need) {
      if (it == 0) {
        slot = i;
      } else if (i < n_fin) {
        slot = fin[i];
        const uint4 m = rec_ldu(w, slot, WF_R_S5);
        const uint32_t flags = ((m.y >> 16) & 0xFu) | ((m.y >> 20) & WF_FLAG_LZERO);
        float4 L = wf_load_L(w, slot, flags);
        if (flags & WF_FLAG_PENDING) wf_add_pending<true>(sc, spectra, w, slot, L);
        wf_finish(sc, ck, cie, L, (m.y >> 20) & 0x7FFu, m.w);
        if (STATS && !(flags & WF_FLAG_DEAD)) { lc.term_cnt[0]++; lc.depth_sum += m.y & 0xFFFFu; }
      } else {
        slot = w.regen[i - n_fin];
      }
    }
    PathState ps;
    Rng rng;
    NeeQuery nee; nee.pending = false; nee.tri = false; nee.light_prim = 0;
    while (mask != 0u) {
      first = __shfl_sync(0xffffffffu, first, __ffs(mask) - 1);
      if (need) {
        wid = first + __popc(mask & ((1u << lane) - 1u));
        if (wid >= total_work) {
          need = false;   // the chunk has no work left: this slot retires
        } else {
          const uint32_t j = wid / ck.n_units;
          uint32_t px, py;
          if (unit_pixel(sc, wid - j * ck.n_units, px, py)) {   // edge tiles have units outside the image
            path_begin<STATS, true>(sc, px, py, ck.s_first + j * ck.s_stride, ps, rng, st);
            lc.n_paths++;
            alive = true;
            need = false;
          }
        }
      }
      mask = __ballot_sync(0xffffffffu, need);   // lanes that drew a unit outside the image fetch again
      if (mask != 0u && lane == (uint32_t)__ffs(mask) - 1u) first = atomicAdd(w.work, (uint32_t)__popc(mask));
    }
    wf_emit<STATS, true>(sc, w, it, live_mode, i < n_cur, alive, alive, slot, ps, rng, wid, WF_FLAG_LZERO, nee, st, lc);
  }
  wf_counts_flush<STATS>(lc, st, counters);
}

// wf_logic: the single-kernel form of wf_shade + wf_regen.  One thread per live slot: add the pending NEE answer, shade the
// hit (wf_bounce), and if the path ended finish it and start the slot's next path right here.  Lanes of a warp are in
// different phases (ncu: ~17 of 32 active), but these kernels wait on memory and on dependent arithmetic rather than on
// issue slots, and the fused form makes one pass over a slot's record per bounce instead of up to two.
template <bool STATS, bool TRI, bool PF>   // TRI: the scene has a triangle light (sc.shadow_closest); PF: it has a BVH (record prefetch, below)
__global__ void __launch_bounds__(WF_LOGIC_BLOCK, WF_LOGIC_MIN_BLOCKS) wf_logic(const __grid_constant__ DevScene sc, const WfState w, uint32_t it_arg,
                                                     uint32_t routed, unsigned long long* __restrict__ counters) {
  extern __shared__ float s_tables[];
  __shared__ WfChunk s_ck;   // the chunk (written by wf_begin) is read from shared memory where it is used: in registers it
  __shared__ uint32_t s_it;  // cost wf_logic 44 bytes of spills, as a kernel parameter it would pin the graph loop to one chunk
  if (threadIdx.x == 0) { s_ck = w.ctl->ck; s_it = it_arg == CRT_WF_IT_DEVICE ? w.ctl->it : it_arg; }   // graph loop: the iteration lives on the device
  __syncthreads();
  const float *spectra, *cie;
  wf_tables(sc, w, s_tables, spectra, cie);
  const WfChunk& ck = s_ck;
#ifdef WF_IT_PARAM   // (A/B switch, host loop only: the iteration number straight from the kernel parameter)
  const uint32_t it = it_arg;
#else
  const uint32_t it = s_it;
#endif
  // Routed forms (`routed` = WF_LIVE_ROUTED or WF_LIVE_CLASS): the live slots arrive in two lists -- the finish list (extension ray missed, or only an NEE
  // answer was outstanding: these lanes finish the path and start the next one) padded to whole warps, then the bounce
  // list (surface hit: these lanes shade) -- so the lanes of a warp are in the same phase of this kernel.  wf_emit and
  // wf_traverse sort the slots into the two lists by their final hit (WF_LIVE_ROUTED), or wf_partition builds them in slot
  // order from the class bytes (WF_LIVE_CLASS).  Unrouted (WF_LIVE_LIST): one live list (bounce), phases mixed.
  const bool defer = (routed & WF_LIVE_DEFER) != 0u;
  routed &= 0xFFu;
  const uint32_t n_fin = (routed && it != 0) ? w.iter[it].n_finish : 0u;
  const uint32_t n_fin_pad = (n_fin + 31u) & ~31u;
  const uint32_t n_cur = it == 0 ? ck.n_slots : n_fin_pad + w.iter[it].n_bounce;
  const uint32_t* __restrict__ cur = (it & 1u) ? w.bounce[1] : w.bounce[0];
  const uint32_t* __restrict__ fin = (it & 1u) ? w.finish[1] : w.finish[0];
  const uint32_t total_work = ck.n_units * ck.n_samples;   // item = ordinal * n_units + unit (sample-major)
  const uint32_t lane = threadIdx.x & 31u;
  Stats st;
  LogicCounts lc;
  wf_counts_init(lc, st);
  const uint32_t stride = gridDim.x * blockDim.x;
  const uint32_t n_round = (n_cur + 31u) & ~31u;
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n_round; i += stride) {
    bool alive = false, has_ext = false, need = false;
    uint32_t slot = 0, wid = 0, flags = 0;
    PathState ps;
    Rng rng;
    NeeQuery nee; nee.pending = false; nee.tri = false; nee.light_prim = 0;
    const bool valid = i < n_cur && (i >= n_fin_pad || i < n_fin);   // not a padding lane of the finish list
    // PF: the slot this thread handles in its NEXT trip -- its list entry is fetched now, its record is prefetched further down
    const uint32_t i_next = i + stride;
    const bool pf = PF && it != 0 && i_next < n_cur && (i_next >= n_fin_pad || i_next < n_fin);
    uint32_t slot_next = 0;
    if (PF) { if (pf) slot_next = i_next < n_fin_pad ? fin[i_next] : cur[i_next - n_fin_pad]; }
    if (valid) {
      need = true;
      if (it == 0) {
        slot = i;
      } else {
        slot = i < n_fin_pad ? fin[i] : cur[i - n_fin_pad];
        // L stays in memory unless this pass adds to it or finishes the path.  With an NEE answer outstanding L is
        // needed for sure: it is loaded with the rest of the record and the sums run in place, the outstanding
        // contribution first (the reference's order).  Otherwise ps.L starts at 0 and collects what wf_bounce adds (at
        // most one term, CS:152 / CS:156 / CS:187; 0 + c = c), and L is fetched afterwards only if that term is not all
        // zeros (L is never -0, so L + 0 = L bit for bit; a NaN term is not "all zeros") or the path ended.
        wf_load(w, slot, ps, rng, wid, flags);
        const bool had_pending = PF && (flags & WF_FLAG_PENDING) != 0u;   // (without triangles no query is ever left pending)
        const uint32_t lzero = flags & WF_FLAG_LZERO;
        if (had_pending) { ps.L = wf_load_L(w, slot, lzero); wf_add_pending<TRI>(sc, spectra, w, slot, ps.L); }
        int r = 0;
        if (!(flags & WF_FLAG_DEAD)) {
          const float4 hv = rec_ld(w, slot, WF_R_H);
          Hit h; h.t = hv.x; h.idx = __float_as_uint(hv.y); h.ref = __float_as_uint(hv.z);
          uint32_t term = 0;
          r = wf_bounce<STATS, TRI, PF>(sc, spectra, ps, rng, h, st, &term, lc.n_shadow, nee);
          if (STATS && r != 1) { lc.term_cnt[term - TERM_MISS]++; lc.depth_sum += ps.depth; }
          // routed forms with WF_LIVE_DEFER: a path that ends in a warp of the bounce list (light hit, depth limit, Russian
          // roulette) is not finished here by the few lanes it happens to -- the slot goes to the finish list as a dead
          // path, and the next iteration finishes it and starts its next path in a warp whose lanes all do that
          if (defer && r == 0 && i >= n_fin_pad) r = 2;
        }
        flags = (nee.pending ? WF_FLAG_PENDING : 0u) | (r == 2 ? WF_FLAG_DEAD : 0u) | lzero;
        alive = r != 0;
        has_ext = r == 1;
        need = !alive;
        bool store_L = had_pending;
        if (!had_pending) {
          const float4 dl = ps.L;
          const bool has_dl = !(dl.x == 0.0f && dl.y == 0.0f && dl.z == 0.0f && dl.w == 0.0f);
          if (has_dl || !alive) {
            ps.L = wf_load_L(w, slot, lzero);
            if (has_dl) { ps.L.x += dl.x; ps.L.y += dl.y; ps.L.z += dl.z; ps.L.w += dl.w; store_L = true; }
          }
        }
        if (!alive) wf_finish(sc, ck, cie, ps.L, ps.lam0, wid);
        else if (store_L) { rec_st(w, slot, WF_R_S2, ps.L); flags &= ~WF_FLAG_LZERO; }
      }
    }
    if (PF) {
      // Records live in a pool far larger than L2 and a pass finds them in DRAM; the list gives the address a whole trip ahead.
      // Sectors A, B, D (what every pass reads).  r02t/r02u A/B, mesh1m: wf_logic 150.8 -> 142.7 ms; C and E on top change
      // nothing, L1 instead of L2 neither.  Only in the instantiation for scenes with a BVH: in the reference's own scene the pass
      // is bound by its arithmetic and the extra registers cost 2.5 % even when the prefetch is skipped at run time.
      if (pf) {
        const char* p = reinterpret_cast<const char*>(rec_ptr(w, slot_next, 0));
#pragma unroll
        for (int k = 0; k < 5; k++)
          if ((WF_PREFETCH >> k) & 1) asm volatile("prefetch.global.L2 [%0];" ::"l"(p + 32 * k));
      }
    }
    // ---- slots whose path ended (or fresh slots) take the next (sample, pixel) work item: one atomic per warp.  (Dealing a
    // share of the items to the slots statically, without the atomic, changed nothing: profiles/r01t_static_l2_tri2_variants.txt.)
    for (;;) {
      const uint32_t mask = __ballot_sync(0xffffffffu, need);
      if (mask == 0u) break;
      uint32_t first = 0;
      if (lane == (uint32_t)__ffs(mask) - 1u) first = atomicAdd(w.work, (uint32_t)__popc(mask));
      first = __shfl_sync(0xffffffffu, first, __ffs(mask) - 1);
      if (need) {
        wid = first + __popc(mask & ((1u << lane) - 1u));
        if (wid >= total_work) {
          need = false;   // the chunk has no work left: this slot retires
        } else {
          const uint32_t j = wid / ck.n_units;
          uint32_t px, py;
          if (unit_pixel(sc, wid - j * ck.n_units, px, py)) {   // edge tiles have units outside the image
            path_begin<STATS, true>(sc, px, py, ck.s_first + j * ck.s_stride, ps, rng, st);
            lc.n_paths++;
            alive = true; has_ext = true; flags = WF_FLAG_LZERO; nee.pending = false;
            need = false;
          }
        }
      }
    }
    wf_emit<STATS, TRI>(sc, w, it, routed, valid, alive, has_ext, slot, ps, rng, wid, flags, nee, st, lc);
  }
  wf_counts_flush<STATS>(lc, st, counters);
}

// wf_tail: the end of a chunk.  When the chunk has no work item left and only a few slots are still live, every further
// iteration is three launches over a handful of warps -- the GPU idles through ~60 us per iteration for the 20-30
// iterations the longest paths still need (Russian roulette has a long tail; a path in the glass sphere can go to
// max_depth).  That tail was ~4 ms per chunk: 1 % of a 256-frame step, but 10 % of the 32 frames a rank renders per step when
// eight GPUs share it, and more than half of a single-frame dispatch.  Instead the loop stops at that point (wf_next) and
// this kernel runs each remaining slot's path to its end in one thread, megakernel style: the pending NEE answer and the
// next closest hit are found inline (the same scan_linear / traverse / wf_bounce code, the same order of additions to L),
// so the bits are those of the iterations it replaces.  Few lanes per warp are active here (that is why the megakernel
// form loses on whole frames), which does not matter for the few thousand paths of a tail.
template <bool STATS>
__global__ void __launch_bounds__(WF_BLOCK) wf_tail(const __grid_constant__ DevScene sc, const WfState w, uint32_t it_arg, uint32_t routed,
                                                    unsigned long long* __restrict__ counters) {
  extern __shared__ float s_tables[];
  const float *spectra, *cie;
  wf_tables(sc, w, s_tables, spectra, cie);
  const WfChunk ck = w.ctl->ck;
  const uint32_t it = it_arg == CRT_WF_IT_DEVICE ? w.ctl->it : it_arg;
  const uint32_t n_fin = (routed && it != 0) ? w.iter[it].n_finish : 0u;
  const uint32_t n_cur = it == 0 ? 0u : n_fin + w.iter[it].n_bounce;   // (iteration 0 has no live list: nothing has started)
  const uint32_t* __restrict__ cur = (it & 1u) ? w.bounce[1] : w.bounce[0];
  const uint32_t* __restrict__ fin = (it & 1u) ? w.finish[1] : w.finish[0];
  Stats st;
  LogicCounts lc;
  wf_counts_init(lc, st);
  const uint32_t stride = gridDim.x * blockDim.x;
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n_cur; i += stride) {
    const uint32_t slot = i < n_fin ? fin[i] : cur[i - n_fin];
    PathState ps;
    Rng rng;
    uint32_t wid, flags;
    wf_load(w, slot, ps, rng, wid, flags);
    ps.L = wf_load_L(w, slot, flags);
    if (flags & WF_FLAG_PENDING) wf_add_pending<true>(sc, spectra, w, slot, ps.L);   // the outstanding NEE answer first (the reference's order)
    if (!(flags & WF_FLAG_DEAD)) {
      const float4 hv = rec_ld(w, slot, WF_R_H);
      Hit h; h.t = hv.x; h.idx = __float_as_uint(hv.y); h.ref = __float_as_uint(hv.z);
      for (;;) {
        NeeQuery nee; nee.pending = false; nee.tri = false; nee.light_prim = 0;
        uint32_t term = 0;
        const int r = wf_bounce<STATS, true>(sc, spectra, ps, rng, h, st, &term, lc.n_shadow, nee);
        if (STATS && r != 1) { lc.term_cnt[term - TERM_MISS]++; lc.depth_sum += ps.depth; }
        if (nee.pending) {
          // what wf_traverse answers for the shadow queue: does a triangle occlude the segment to the light?  (origin = the new
          // ray's origin, exclude = the surface just hit)
          float4 c = nee.c;
          if (nee.tri) {   // the light is a triangle: closest hit from the linear list's hit on, then the contribution (wf_pending)
            Hit hs; hs.t = -nee.t; hs.idx = __float_as_uint(nee.aux.w); hs.ref = 0;
            traverse<false, STATS>(sc, ps.o, nee.d, ps.exclude, sc.t_min, hs, st);
            c = wf_tri_contribution(sc, spectra, ps.o, nee.d, hs.t, nee.light_prim, hs.ref, nee.c, nee.aux, ps.lam0, hs.idx == nee.light_prim);
          } else {
            Hit hs; hs.t = nee.t; hs.idx = nee.light_prim; hs.ref = 0;
            traverse<true, STATS>(sc, ps.o, nee.d, ps.exclude, sc.t_min, hs, st);
            c = nee_resolve(c, hs.idx != nee.light_prim);   // CS:407, as wf_pending does
          }
          ps.L.x += c.x; ps.L.y += c.y; ps.L.z += c.z; ps.L.w += c.w;
          lc.n_bvh_rays++;
        }
        if (r != 1) break;
        h = scan_linear<STATS>(sc, ps.o, ps.d, ps.exclude, st);   // the next extension ray: linear list, then the triangles
        lc.n_ext++;
        if (hits_root_box(sc, ps.o, ps.d, h.t)) { traverse<false, STATS>(sc, ps.o, ps.d, ps.exclude, sc.t_min, h, st); lc.n_bvh_rays++; }
      }
    } else if (STATS) {
      // (a path Russian roulette ended earlier was counted when it ended)
    }
    wf_finish(sc, ck, cie, ps.L, ps.lam0, wid);
    if (routed == WF_LIVE_CLASS) w.cls[slot] = (uint8_t)2;
  }
  wf_counts_flush<STATS>(lc, st, counters);
  if (STATS) {   // BVH counters of the inline traversals
    unsigned long long v[2] = {st.node_visits, st.tri_tests};
    for (int k = 0; k < 2; k++) {
      unsigned long long x = v[k];
      for (int o = 16; o > 0; o >>= 1) x += __shfl_down_sync(0xffffffffu, x, o);
      if ((threadIdx.x & 31u) == 0 && x) atomicAdd(&counters[CRT_CNT_NODE_VISITS + k], x);
    }
  }
}

// wf_traverse collects the slots of finished extension rays per warp in shared memory and appends them to the next
// iteration's queues 32 at a time: one atomic (and one wait for it) per 32 rays instead of one per completion event.
struct DoneBuf {
  uint32_t* buf;   // 64 entries of this warp
  uint32_t n;      // warp-uniform fill
};

__device__ __forceinline__ void done_flush(DoneBuf& b, uint32_t count, uint32_t* __restrict__ list, uint32_t* __restrict__ counter) {
  const uint32_t lane = threadIdx.x & 31u;
  __syncwarp();
  uint32_t base = 0;
  if (lane == 0) base = atomicAdd(counter, count);
  base = __shfl_sync(0xffffffffu, base, 0);
  if (lane < count) list[base + lane] = b.buf[lane];
  const uint32_t rest = b.n - count;
  const uint32_t moved = lane < rest ? b.buf[count + lane] : 0u;
  __syncwarp();
  if (lane < rest) b.buf[lane] = moved;
  b.n = rest;
}

__device__ __forceinline__ void done_add(DoneBuf& b, bool pred, uint32_t slot, uint32_t* __restrict__ list, uint32_t* __restrict__ counter) {
  const uint32_t mask = __ballot_sync(0xffffffffu, pred);
  if (mask == 0u) return;
  if (pred) b.buf[b.n + __popc(mask & ((1u << (threadIdx.x & 31u)) - 1u))] = slot;
  b.n += __popc(mask);
  if (b.n >= 32u) done_flush(b, 32u, list, counter);
}

// Persistent traversal of the BVH queue of iteration `it`.  Each lane owns one ray at a time; when its
// traversal ends it writes the hit back and takes the next queue entry, so the warp never waits for its
// slowest ray.  Every trip of the loop is one triangle test for the lanes that have triangles pending, then one node
// step (fetch an 8-wide node, test its child boxes) for the lanes whose triangle group is empty by then.  (Measured:
// triangles first is 4 % faster than nodes first; batching the triangle tests until 6..24 lanes have some is 35-85 %
// slower -- the kernel lives on overlapping the node and triangle fetches of different lanes; letting a lane keep walking nodes
// while it holds one postponed triangle group until 4 / 8 / 12 lanes hold triangles is 3-4 % slower although it needs fewer
// instructions, and taking 32 queue entries per atomic with a refill threshold of 2 / 4 is 4 % slower / equal
// (profiles/r02a_traverse_schedule_ab.txt: the kernel waits for its fetches, not for issue slots); prefetch.global.L1 of the
// next node or triangle changes nothing; one merged fetch phase per trip -- triangle lanes and node lanes load through the
// same registers and the warp waits once -- is equal on mesh1m and 5 % slower on the 10M soup.)  Arithmetic per node and per triangle is wide_node / wide_tri
// (crt_device.cuh), the same code the megakernel and the trace-rays hook run.
#ifndef WF_SSTACK
#define WF_SSTACK 8          // entries of a lane's traversal stack kept in shared memory, the rest in local memory (0 -> 8: wf_traverse -1.5 % on mesh1m, -2.6 % on soup10m; 5 = 8)
#endif
#ifndef WF_REFILL_MIN
#define WF_REFILL_MIN 8
#endif
template <bool STATS, bool ROUTE>
__global__ void __launch_bounds__(WF_BLOCK, WF_TRAV_MIN_BLOCKS) wf_traverse(const __grid_constant__ DevScene sc, const WfState w, uint32_t it_arg,
                                                        uint32_t fix_class, unsigned long long* __restrict__ counters) {
  __shared__ uint32_t s_done[WF_BLOCK / 32][2][64];
  const uint32_t it = it_arg == CRT_WF_IT_DEVICE ? w.ctl->it : it_arg;
  const uint32_t n_ext = w.iter[it].n_bvh;
  const uint32_t n = n_ext + w.iter[it].n_shadow;   // one fetch cursor over both queues: extension rays first
  if (n == 0) return;
  DoneBuf db_b, db_f;
  db_b.buf = s_done[threadIdx.x >> 5][0]; db_b.n = 0;
  db_f.buf = s_done[threadIdx.x >> 5][1]; db_f.n = 0;
  uint32_t* const list_b = ((it + 1u) & 1u) ? w.bounce[1] : w.bounce[0];
  uint32_t* const list_f = ((it + 1u) & 1u) ? w.finish[1] : w.finish[0];
  const uint32_t lane = threadIdx.x & 31u;
  const uint32_t* __restrict__ queue = w.bvh_list;
  uint32_t* fetch = &w.iter[it].n_fetch;
  const uint4* __restrict__ nodes = sc.nodes;
  const float4* __restrict__ tris = sc.tri;
  const float t_min = sc.t_min;
  const uint32_t one = wide_one(it);
  Stats st; st.prim_tests = st.node_visits = st.tri_tests = st.rand_calls = 0;

#if WF_SSTACK
  __shared__ uint2 s_stack[WF_SSTACK][WF_BLOCK];   // [entry][thread]: a lane's column, conflict-free whatever the lanes' depths
  uint2 stack[CRT_BVH_STACK - WF_SSTACK];
#define WF_PUSH(g) do { if (sp < WF_SSTACK) s_stack[sp][threadIdx.x] = (g); else stack[sp - WF_SSTACK] = (g); sp++; } while (0)
#define WF_POP(g) do { --sp; if (sp < WF_SSTACK) (g) = s_stack[sp][threadIdx.x]; else (g) = stack[sp - WF_SSTACK]; } while (0)
#else
  uint2 stack[CRT_BVH_STACK];
#define WF_PUSH(g) do { stack[sp++] = (g); } while (0)
#define WF_POP(g) do { (g) = stack[--sp]; } while (0)
#endif
  int sp = 0;
  uint2 G = make_uint2(0u, 0u), T = make_uint2(0u, 0u);   // node group and triangle group of the lane's ray (crt_device.cuh)
  bool have = false, exhausted = false, anyhit = false, shadow = false, lin_miss = false;   // shadow: the entry is an NEE query; anyhit: its mode
  uint32_t slot = 0, exclude = CRT_MAX_U32, light_prim = 0;
  V3 d = v3(0.0f, 0.0f, 1.0f);
  WideRay r = wide_ray(v3(0.0f, 0.0f, 0.0f), d);
  Hit h; h.t = 0.0f; h.idx = CRT_MAX_U32; h.ref = 0;
  uint32_t ray_nodes = 0, ray_tris = 0, max_nodes = 0;   // STATS: per-ray visit counts

  for (;;) {
    // ---- refill the lanes that have no ray
    // (a refill costs the whole warp an atomic round trip plus the loads of the new rays, so it waits until
    //  WF_REFILL_MIN lanes are idle -- or until nobody has a ray)
    const bool need = !have && !exhausted;
    uint32_t mask = __ballot_sync(0xffffffffu, need);
    if (__popc(mask) < WF_REFILL_MIN && __ballot_sync(0xffffffffu, have) != 0u) mask = 0u;
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
          shadow = q >= n_ext;
          anyhit = shadow;
          slot = shadow ? w.shadow_list[q - n_ext] : queue[q];
          const float4 a = rec_ld(w, slot, WF_R_S0);
          const uint4 m = rec_ldu(w, slot, WF_R_S5);
          exclude = m.x;
          if (shadow) {
            // the NEE shadow ray: the light (linear list) is the closest hit so far at distance b.w; any triangle accepted
            // under the tie rule within that bound occludes it (crt_device.cuh shadow_intersect)
            const float4 b = rec_ld(w, slot, WF_R_S6);
            d = v3(b);
            light_prim = m.z;
            h.t = b.w; h.idx = light_prim; h.ref = 0;
            if (b.w < 0.0f) {
              // ... unless the light is itself a triangle (sc.shadow_closest scenes): then this is a closest-hit query that
              // starts from the linear list's hit (distance -b.w, its index in the aux word) and is visible only if the
              // traversal ends on the light
              anyhit = false;
              h.t = -b.w; h.idx = __float_as_uint(rec_ptr(w, slot, WF_R_S8)->w);
            }
          } else {
            const float4 b = rec_ld(w, slot, WF_R_S1), hv = rec_ld(w, slot, WF_R_H);
            d = v3(b);
            h.t = hv.x; h.idx = __float_as_uint(hv.y); h.ref = __float_as_uint(hv.z);
            lin_miss = h.idx == CRT_MAX_U32;
          }
          r = wide_ray(v3(a), d);
          G = make_uint2(0u, r.valid ? 0x80000000u : 0u);   // the root as the only child of a virtual group
          T = make_uint2(0u, 0u);
          sp = 0;
          have = true;
          ray_nodes = ray_tris = 0;
        }
      }
    }
    if (__ballot_sync(0xffffffffu, have) == 0u) break;
    // ---- triangle step: one triangle for every lane that has some left
    bool done = false;
    if (have && T.y != 0u) {
      if (STATS) ray_tris++;
      if (wide_tri<STATS>(tris, T, r.o, d, exclude, t_min, anyhit, h, st)) done = true;
    }
    // ---- node step, for the lanes whose triangle group is (now) empty: take the nearest unvisited child of the current
    //      node group (or of the group on top of the stack), fetch it and test its 8 child boxes
    if (have && !done && T.y == 0u) {
      if (!(G.y & CRT_WIDE_HITS)) {
        if (sp == 0) done = true;
        else WF_POP(G);
      }
      if (!done) {
        const uint32_t node = wide_pop_child(G, r.octinv);
        if (G.y & CRT_WIDE_HITS) WF_PUSH(G);
        wide_node<STATS>(nodes, node, r, one, t_min, h.t, G, T, st);
        if (STATS) ray_nodes++;
      }
    }
    // ---- finished rays: a shadow ray leaves its answer in S6.w; an extension ray leaves its final hit and its slot goes
    //      to the next iteration's shade queue (surface) or finish queue (miss)
    const bool fin_ext = done && !shadow;
    if (STATS && done) {
      max_nodes = max(max_nodes, ray_nodes);
      if (w.slow_min && ray_nodes >= w.slow_min) {
        const uint32_t k = atomicAdd(reinterpret_cast<uint32_t*>(w.slow), 1u);
        if (k < CRT_WF_SLOW_RAYS) {
          w.slow[1 + 3 * k] = make_float4(r.o.x, r.o.y, r.o.z, __uint_as_float(it));
          w.slow[2 + 3 * k] = make_float4(d.x, d.y, d.z, __uint_as_float(ray_nodes));
          w.slow[3 + 3 * k] = make_float4(__uint_as_float(exclude), __uint_as_float(anyhit ? 1u : 0u), __uint_as_float(slot), __uint_as_float(ray_tris));
        }
      }
    }
    if (done) {
      if (shadow) {
        if (anyhit) {
          if (h.idx != light_prim) rec_ptr(w, slot, WF_R_S6)->w = -1.0f;   // occluded
        } else if (h.idx == light_prim) {   // closest-hit query towards a triangle light: visible, leave the hit (else the length stays negative)
          rec_ptr(w, slot, WF_R_S6)->w = h.t;
          rec_ptr(w, slot, WF_R_S8)->z = __uint_as_float(h.ref);
        }
      } else {
        rec_st(w, slot, WF_R_H, make_float4(h.t, __uint_as_float(h.idx), __uint_as_float(h.ref), 0.0f));
        if (fix_class && lin_miss && h.idx != CRT_MAX_U32) w.cls[slot] = (uint8_t)0;   // WF_LIVE_CLASS: a triangle turned the miss into a hit
      }
      have = false;
    }
    if (ROUTE && __ballot_sync(0xffffffffu, fin_ext) != 0u) {   // routed forms: the slot joins the next bounce or finish queue by its final hit
      done_add(db_b, fin_ext && h.idx != CRT_MAX_U32, slot, list_b, &w.iter[it + 1u].n_bounce);
      done_add(db_f, fin_ext && h.idx == CRT_MAX_U32, slot, list_f, &w.iter[it + 1u].n_finish);
    }
  }
  if (ROUTE) {
    if (db_b.n) done_flush(db_b, db_b.n, list_b, &w.iter[it + 1u].n_bounce);
    if (db_f.n) done_flush(db_f, db_f.n, list_f, &w.iter[it + 1u].n_finish);
  }
  if (STATS) {
    unsigned long long v[2] = {st.node_visits, st.tri_tests};
    for (int k = 0; k < 2; k++) {
      unsigned long long x = v[k];
      for (int s = 16; s > 0; s >>= 1) x += __shfl_down_sync(0xffffffffu, x, s);
      if (lane == 0 && x) atomicAdd(&counters[CRT_CNT_NODE_VISITS + k], x);
    }
    for (int s = 16; s > 0; s >>= 1) max_nodes = max(max_nodes, __shfl_down_sync(0xffffffffu, max_nodes, s));
    if (lane == 0) atomicMax(&counters[CRT_CNT_MAX_RAY_NODES], (unsigned long long)max_nodes);
  }
}

// WF_LIVE_CLASS: the bounce and finish lists of iteration it + 1 from the class bytes, in slot order.  A block takes tiles
// of 256 x 16 consecutive slots (a thread reads its 16 class bytes as one 16-byte word), ranks its entries with a block
// scan and reserves the tile's share of both lists with one atomic per list -- per warp and 32 slots the two counters,
// each a single address, would be the whole cost of the pass (measured: 1.1 ms per launch against 0.06 ms of traffic).
// Tiles reach the lists in arrival order, which is close enough to slot order for wf_logic's record stream to stay
// sequential in memory.
#define WF_PART_PER_THREAD 16
__global__ void __launch_bounds__(256) wf_partition(const WfState w, uint32_t it_arg) {
  const uint32_t it = it_arg == CRT_WF_IT_DEVICE ? w.ctl->it : it_arg;
  const uint32_t n_slots = w.ctl->ck.n_slots;
  __shared__ uint32_t s_warp[2][8];
  __shared__ uint32_t s_base[2], s_tot[2];
  __shared__ uint32_t s_out[2][256 * WF_PART_PER_THREAD];   // the tile's entries by class, then copied out coalesced
  uint32_t* const list_b = ((it + 1u) & 1u) ? w.bounce[1] : w.bounce[0];
  uint32_t* const list_f = ((it + 1u) & 1u) ? w.finish[1] : w.finish[0];
  const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
  const uint32_t tile = 256u * WF_PART_PER_THREAD;
  const uint32_t n_tiles = (n_slots + tile - 1u) / tile;
  for (uint32_t t = blockIdx.x; t < n_tiles; t += gridDim.x) {
    const uint32_t first = t * tile + threadIdx.x * WF_PART_PER_THREAD;
    uint4 c4 = make_uint4(0x02020202u, 0x02020202u, 0x02020202u, 0x02020202u);   // class 2: not live
    if (first + WF_PART_PER_THREAD <= n_slots) {
      c4 = *reinterpret_cast<const uint4*>(w.cls + first);   // w.cls is 256-byte aligned, first a multiple of 16
    } else if (first < n_slots) {
      uint32_t v[4] = {0x02020202u, 0x02020202u, 0x02020202u, 0x02020202u};
      for (uint32_t k = 0; first + k < n_slots; k++) v[k >> 2] = (v[k >> 2] & ~(0xFFu << (8u * (k & 3u)))) | ((uint32_t)w.cls[first + k] << (8u * (k & 3u)));
      c4 = make_uint4(v[0], v[1], v[2], v[3]);
    }
    const uint32_t cw[4] = {c4.x, c4.y, c4.z, c4.w};
    uint32_t n0 = 0, n1 = 0;   // entries of class 0 (bounce) and class 1 (finish) among this thread's 16
#pragma unroll
    for (int k = 0; k < 4; k++) {
      const uint32_t is1 = cw[k] & 0x01010101u, is2 = (cw[k] >> 1) & 0x01010101u;
      n1 += __popc(is1);
      n0 += 4u - __popc(is1 | is2);
    }
    // exclusive ranks within the block: warp scan, then the 8 warp totals
    uint32_t x0 = n0, x1 = n1;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t y0 = __shfl_up_sync(0xffffffffu, x0, o), y1 = __shfl_up_sync(0xffffffffu, x1, o);
      if (lane >= (uint32_t)o) { x0 += y0; x1 += y1; }
    }
    if (lane == 31u) { s_warp[0][warp] = x0; s_warp[1][warp] = x1; }
    __syncthreads();
    if (threadIdx.x < 2u) {
      uint32_t tot = 0;
      for (int k = 0; k < 8; k++) { const uint32_t v = s_warp[threadIdx.x][k]; s_warp[threadIdx.x][k] = tot; tot += v; }
      s_tot[threadIdx.x] = tot;
      s_base[threadIdx.x] = tot ? atomicAdd(threadIdx.x ? &w.iter[it + 1u].n_finish : &w.iter[it + 1u].n_bounce, tot) : 0u;
    }
    __syncthreads();
    uint32_t p0 = s_warp[0][warp] + x0 - n0, p1 = s_warp[1][warp] + x1 - n1;
#pragma unroll
    for (int k = 0; k < WF_PART_PER_THREAD; k++) {
      const uint32_t c = (cw[k >> 2] >> (8 * (k & 3))) & 0xFFu;
      if (c == 0u) s_out[0][p0++] = first + k;
      else if (c == 1u) s_out[1][p1++] = first + k;
    }
    __syncthreads();
    for (uint32_t j = threadIdx.x; j < s_tot[0]; j += 256u) list_b[s_base[0] + j] = s_out[0][j];
    for (uint32_t j = threadIdx.x; j < s_tot[1]; j += 256u) list_f[s_base[1] + j] = s_out[1][j];
    __syncthreads();   // the shared arrays are reused by the next tile
  }
}

// Start of a chunk: the control block takes the chunk and iteration 0.  (The per-iteration counters and the work counter are
// cleared by a memset on the same stream.)
__global__ void wf_begin(const WfState w, const WfChunk ck) {
  w.ctl->ck = ck;
  w.ctl->it = 0u;
  w.ctl->error = 0u;
}

// End of an iteration of the graph loop: advance the iteration and tell the WHILE node whether any slot is still live (the
// lists of iteration it + 1 are complete: wf_logic / wf_traverse pushed them or wf_partition built them).
__global__ void wf_next(const WfState w, cudaGraphConditionalHandle loop, uint32_t kernels_per_iter, uint32_t tail_max) {
  const uint32_t it = w.ctl->it + 1u;
  w.ctl->it = it;
  w.ctl->iters_total++;
  w.ctl->launches_total += kernels_per_iter;
  const uint32_t live = w.iter[it].n_bounce + w.iter[it].n_finish;
  bool more = live != 0u;
  // no work item left and only a tail of live slots: wf_tail finishes them (launched after the loop; it finds nothing when the
  // loop ran to the end)
  if (live <= tail_max && *w.work >= w.ctl->ck.n_units * w.ctl->ck.n_samples) more = false;
  if (more && it + 2u >= CRT_WF_MAX_ITERS) { more = false; w.ctl->error = 1u; }   // cannot happen: the host sizes chunks for the worst case
  cudaGraphSetConditional(loop, more ? 1u : 0u);
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
    const float4 v = ck.stage[(size_t)j * ck.n_units + unit];
    a.x += v.x; a.y += v.y; a.z += v.z;
  }
  acc[pix] = a;
  fb[pix] = resolve_pixel(a.x, a.y, a.z, sample_after, sc.exposure);
}

namespace crt {

size_t wf_tables_bytes(const DevScene& sc) { return ((size_t)sc.n_spectra * CRT_LAMBDA_RANGE + 3u * 471u) * sizeof(float); }

cudaError_t wf_config(bool stats, size_t smem, int* logic_blocks_per_sm, int* shade_blocks_per_sm, int* regen_blocks_per_sm,
                      int* trav_blocks_per_sm) {
  cudaError_t e;
#define CRT_WF_CFG(S)                                                                                                              \
  if ((e = cudaFuncSetAttribute(wf_logic<S, false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)) != cudaSuccess) return e;    \
  if ((e = cudaFuncSetAttribute(wf_logic<S, true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)) != cudaSuccess) return e;    \
  if ((e = cudaFuncSetAttribute(wf_logic<S, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)) != cudaSuccess) return e;    \
  if ((e = cudaFuncSetAttribute(wf_logic<S, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)) != cudaSuccess) return e;    \
  if ((e = cudaFuncSetAttribute(wf_shade<S>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)) != cudaSuccess) return e;    \
  if ((e = cudaFuncSetAttribute(wf_regen<S>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)) != cudaSuccess) return e;    \
  if ((e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(logic_blocks_per_sm, wf_logic<S, true, true>, WF_LOGIC_BLOCK, smem)) != cudaSuccess) return e; \
  if ((e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(shade_blocks_per_sm, wf_shade<S>, WF_BLOCK, smem)) != cudaSuccess) return e; \
  if ((e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(regen_blocks_per_sm, wf_regen<S>, WF_BLOCK, smem)) != cudaSuccess) return e; \
  if ((e = cudaFuncSetAttribute(wf_tail<S>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)) != cudaSuccess) return e;     \
  return cudaOccupancyMaxActiveBlocksPerMultiprocessor(trav_blocks_per_sm, wf_traverse<S, true>, WF_BLOCK, 0);
  if (stats) { CRT_WF_CFG(true) }
  CRT_WF_CFG(false)
#undef CRT_WF_CFG
}

cudaError_t launch_wf_begin(const WfState& w, const WfChunk& ck, cudaStream_t s) {
  wf_begin<<<1, 1, 0, s>>>(w, ck);
  return cudaGetLastError();
}

// The instantiation of wf_logic a scene runs: with the triangle-light code only if a light is a triangle, with the record
// prefetch only if the scene has a BVH.
typedef void (*WfLogicFn)(const DevScene, const WfState, uint32_t, uint32_t, unsigned long long*);
template <bool S>
static WfLogicFn wf_logic_pick(const DevScene& sc) {
  const bool pf = sc.n_tri != 0u;   // (hits_root_box, the only source of pending NEE queries, says no without triangles)
  if (sc.shadow_closest) return pf ? wf_logic<S, true, true> : wf_logic<S, true, false>;
  return pf ? wf_logic<S, false, true> : wf_logic<S, false, false>;
}
static WfLogicFn wf_logic_fn(const DevScene& sc, bool stats) { return stats ? wf_logic_pick<true>(sc) : wf_logic_pick<false>(sc); }

cudaError_t launch_wf_logic(const DevScene& sc, const WfState& w, bool stats, uint32_t live_mode, int grid, size_t smem, uint32_t it,
                            unsigned long long* counters, cudaStream_t s) {
  wf_logic_fn(sc, stats)<<<grid, WF_LOGIC_BLOCK, smem, s>>>(sc, w, it, live_mode, counters);
  return cudaGetLastError();
}

cudaError_t launch_wf_shade(const DevScene& sc, const WfState& w, const WfChunk& ck, bool stats, uint32_t live_mode, int grid, size_t smem, uint32_t it,
                            unsigned long long* counters, cudaStream_t s) {
  if (stats) wf_shade<true><<<grid, WF_BLOCK, smem, s>>>(sc, w, ck, it, live_mode, counters);
  else wf_shade<false><<<grid, WF_BLOCK, smem, s>>>(sc, w, ck, it, live_mode, counters);
  return cudaGetLastError();
}

cudaError_t launch_wf_regen(const DevScene& sc, const WfState& w, const WfChunk& ck, bool stats, uint32_t live_mode, int grid, size_t smem, uint32_t it,
                            unsigned long long* counters, cudaStream_t s) {
  if (stats) wf_regen<true><<<grid, WF_BLOCK, smem, s>>>(sc, w, ck, it, live_mode, counters);
  else wf_regen<false><<<grid, WF_BLOCK, smem, s>>>(sc, w, ck, it, live_mode, counters);
  return cudaGetLastError();
}

cudaError_t launch_wf_traverse(const DevScene& sc, const WfState& w, bool stats, bool route, bool fix_class, int grid, uint32_t it,
                               unsigned long long* counters, cudaStream_t s) {
  const uint32_t fc = fix_class ? 1u : 0u;
  if (stats) {
    if (route) wf_traverse<true, true><<<grid, WF_BLOCK, 0, s>>>(sc, w, it, fc, counters);
    else wf_traverse<true, false><<<grid, WF_BLOCK, 0, s>>>(sc, w, it, fc, counters);
  } else {
    if (route) wf_traverse<false, true><<<grid, WF_BLOCK, 0, s>>>(sc, w, it, fc, counters);
    else wf_traverse<false, false><<<grid, WF_BLOCK, 0, s>>>(sc, w, it, fc, counters);
  }
  return cudaGetLastError();
}

cudaError_t launch_wf_tail(const DevScene& sc, const WfState& w, bool stats, uint32_t live_mode, int grid, size_t smem, uint32_t it,
                           unsigned long long* counters, cudaStream_t s) {
  if (stats) wf_tail<true><<<grid, WF_BLOCK, smem, s>>>(sc, w, it, live_mode, counters);
  else wf_tail<false><<<grid, WF_BLOCK, smem, s>>>(sc, w, it, live_mode, counters);
  return cudaGetLastError();
}

cudaError_t launch_wf_partition(const WfState& w, int grid, uint32_t it, cudaStream_t s) {
  wf_partition<<<grid, 256, 0, s>>>(w, it);
  return cudaGetLastError();
}

// WHILE (live slots) { wf_logic; [wf_traverse;] [wf_partition;] wf_next } as one CUDA graph: the host enqueues a chunk with one
// launch and never waits for it; the loop ends on the device (cudaGraphSetConditional in wf_next).
cudaError_t wf_graph_build(const DevScene& sc, const WfState& w, bool stats, uint32_t logic_mode, uint32_t live_mode, int grid_l, int grid_t,
                           int grid_p, size_t smem, unsigned long long* counters, uint32_t kernels_per_iter, uint32_t tail_max, cudaGraph_t* graph, cudaGraphExec_t* exec) {
  cudaError_t e;
  cudaGraph_t g = nullptr;
  if ((e = cudaGraphCreate(&g, 0)) != cudaSuccess) return e;
  cudaGraphConditionalHandle loop;
  if ((e = cudaGraphConditionalHandleCreate(&loop, g, 1u, cudaGraphCondAssignDefault)) != cudaSuccess) { cudaGraphDestroy(g); return e; }
  cudaGraphNodeParams cp = {};
  cp.type = cudaGraphNodeTypeConditional;
  cp.conditional.handle = loop;
  cp.conditional.type = cudaGraphCondTypeWhile;
  cp.conditional.size = 1;
  cudaGraphNode_t cond;
  if ((e = cudaGraphAddNode(&cond, g, nullptr, 0, &cp)) != cudaSuccess) { cudaGraphDestroy(g); return e; }
  cudaGraph_t body = cp.conditional.phGraph_out[0];
  uint32_t it = CRT_WF_IT_DEVICE;
  cudaGraphNode_t prev = nullptr, node = nullptr;
  auto add = [&](void* func, dim3 grid, dim3 block, size_t shmem, void** args) -> cudaError_t {
    cudaKernelNodeParams kp = {};
    kp.func = func; kp.gridDim = grid; kp.blockDim = block; kp.sharedMemBytes = (unsigned)shmem; kp.kernelParams = args; kp.extra = nullptr;
    const cudaError_t rc = cudaGraphAddKernelNode(&node, body, prev ? &prev : nullptr, prev ? 1 : 0, &kp);
    prev = node;
    return rc;
  };
  {
    void* args[] = {(void*)&sc, (void*)&w, &it, &logic_mode, &counters};
    void* f = (void*)wf_logic_fn(sc, stats);
    e = add(f, dim3(grid_l), dim3(WF_LOGIC_BLOCK), smem, args);
    if (e != cudaSuccess) { cudaGraphDestroy(g); return e; }
  }
  if (sc.n_tri) {
    uint32_t fix_class = live_mode == 2 ? 1u : 0u;
    void* args[] = {(void*)&sc, (void*)&w, &it, &fix_class, &counters};
    void* f = live_mode == 1 ? (stats ? (void*)wf_traverse<true, true> : (void*)wf_traverse<false, true>)
                             : (stats ? (void*)wf_traverse<true, false> : (void*)wf_traverse<false, false>);
    e = add(f, dim3(grid_t), dim3(WF_BLOCK), 0, args);
    if (e != cudaSuccess) { cudaGraphDestroy(g); return e; }
  }
  if (live_mode == 2) {
    void* args[] = {(void*)&w, &it};
    e = add((void*)wf_partition, dim3(grid_p), dim3(256), 0, args);
    if (e != cudaSuccess) { cudaGraphDestroy(g); return e; }
  }
  {
    void* args[] = {(void*)&w, &loop, &kernels_per_iter, &tail_max};
    e = add((void*)wf_next, dim3(1), dim3(1), 0, args);
    if (e != cudaSuccess) { cudaGraphDestroy(g); return e; }
  }
  cudaGraphExec_t x = nullptr;
  if ((e = cudaGraphInstantiate(&x, g, 0)) != cudaSuccess) { cudaGraphDestroy(g); return e; }
  *graph = g; *exec = x;
  return cudaSuccess;
}

cudaError_t launch_wf_accumulate(const DevScene& sc, const WfChunk& ck, float4* acc, uint32_t* fb, uint32_t sample_after, cudaStream_t s) {
  if (ck.n_units) wf_accumulate<<<(ck.n_units + 255) / 256, 256, 0, s>>>(sc, ck, acc, fb, sample_after);
  return cudaGetLastError();
}

}  // namespace crt
