// crt_path.cuh -- the per-path state machine of path_trace (CS:119-295), cut at every intersect() call so that
// a path can be advanced one ray at a time: by the megakernel (intersect inline) or by the wavefront kernels
// (intersect in a separate, compacted pass).
#pragma once
#include "crt_device.cuh"

struct PathState {
  V3 o, d;            // current ray: extension ray, or the shadow ray while phase == 1
  float4 L, beta;     // accumulated_radiance, beta (CS:123-124)
  float pdf;          // last_bounce_pdf (CS:125)
  float etaScale;     // CS:130
  uint32_t exclude;   // CS:126
  uint32_t depth;     // CS:122
  uint32_t lam0;      // first wavelength index; the other three follow (CS:321)
  bool specular, in_transmission, shadow_phase;
  // saved across the shadow ray of next-event estimation
  V3 n;               // surface normal at the diffuse hit
  uint32_t refl;      // reflectance spectrum of the diffuse hit
  uint32_t light;     // sampled light
};

enum { TERM_MISS = CRT_CNT_TERM_MISS, TERM_RR = CRT_CNT_TERM_RR, TERM_LIGHT = CRT_CNT_TERM_LIGHT, TERM_DEPTH = CRT_CNT_TERM_DEPTH };

// Everything that follows one intersect() call, up to the point where the next ray is known: `h` is the
// result for the current ray (extension ray, or the NEE shadow ray while ps.shadow_phase).  Returns false
// when the path has terminated (its radiance is complete in ps.L); otherwise ps.o/ps.d/ps.exclude describe
// the next ray.  Shared by the megakernel and the wavefront kernels.
template <bool STATS>
__device__ __forceinline__ bool path_consume(const DevScene& sc, const float* __restrict__ spectra, PathState& ps, Rng& rng,
                                             const Hit& h, Stats& st, uint32_t* term) {
  uint32_t lam[4];
  wavelengths(ps.lam0, lam);
  if (!ps.shadow_phase) {
    // h = intersect(ray, exclude), CS:135
    if (h.idx == CRT_MAX_U32) { if (STATS) term[0] = TERM_MISS; return false; }  // CS:141-144
    ps.exclude = h.idx;                                                       // CS:146
    const Surface s = surface_of(sc, h, ps.o, ps.d);
    const uint32_t material = CRT_META_MAT(s.meta);
    if (material == CRT_MAT_LIGHT) {                                          // CS:149-164
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
      return false;
    }
    if (ps.depth >= sc.max_depth) { if (STATS) term[0] = TERM_DEPTH; return false; }  // CS:167-170
    if (ps.in_transmission) {                                                 // CS:173-179
      const float distance = length(ps.o - s.position);
      const float4 ext = sample_spectrum(spectra, sc.n_spectra, sc.n_spectra - 1, lam);
      ps.beta.x *= crt_expf(-ext.x * distance); ps.beta.y *= crt_expf(-ext.y * distance);
      ps.beta.z *= crt_expf(-ext.z * distance); ps.beta.w *= crt_expf(-ext.w * distance);
    }
    if (material == CRT_MAT_DIFFUSE) {                                        // CS:182-204, first half
      // sample_lights + sample_light, CS:341-355
      const float u0 = rnd(rng);
      uint32_t li = f2u_sat((float)sc.n_lights * u0);
      if (li >= sc.n_lights) li = sc.n_lights - 1;
      const LightRec& light = light_at(sc, li);
      const float u = rnd(rng), v = rnd(rng);
      if (STATS) st.rand_calls += 3;
      const V3 point = madd(madd(v3(light.d1), u, v3(light.d2)), v, v3(light.d3));
      const V3 light_dir = normalize(point - s.position);                     // CS:385
      ps.n = s.normal; ps.refl = CRT_META_REFL(s.meta); ps.light = li;
      ps.o = s.position; ps.d = light_dir;
      ps.shadow_phase = true;
      return true;                                                            // next step: the shadow ray
    }
    if (material == CRT_MAT_GLASS) {                                          // CS:208-276
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
  } else {
    // second half of the diffuse branch: the shadow ray of compute_light_radiance, CS:379-408
    // h = shadow_intersect(ray, include, exclude), CS:386, 697-705
    const LightRec& light = light_at(sc, ps.light);
    const uint32_t include = __float_as_uint(light.d1.w);
    const float4 refl = sample_spectrum(spectra, sc.n_spectra, ps.refl, lam);
    const float4 BRDF = make_float4(refl.x / CRT_PI_F, refl.y / CRT_PI_F, refl.z / CRT_PI_F, refl.w / CRT_PI_F);  // CS:184
    const float cos_theta = fmaxf(0.0f, dot(ps.n, ps.d));                      // CS:388
    const float4 spec = sample_spectrum(spectra, sc.n_spectra, __float_as_uint(light.d2.w), lam);
    float4 le = make_float4(spec.x * cos_theta, spec.y * cos_theta, spec.z * cos_theta, spec.w * cos_theta);
    if (h.idx == include) {
      const Surface sh = surface_of(sc, h, ps.o, ps.d);
      const float pdf_l = compute_light_pdf(sc, CRT_META_EMIS(sh.meta), sh.normal, sh.position, ps.o, ps.d);  // CS:393
      const float pdf_b = crt_div(cos_theta, CRT_PI_F);
      const float w = power_heuristic(pdf_l, pdf_b);
      le = make_float4(crt_div(le.x * w, pdf_l), crt_div(le.y * w, pdf_l), crt_div(le.z * w, pdf_l), crt_div(le.w * w, pdf_l));  // CS:400
    } else {
      le = make_float4(0.0f, 0.0f, 0.0f, 0.0f);                                // CS:407
    }
    ps.L.x += (BRDF.x * le.x) * ps.beta.x; ps.L.y += (BRDF.y * le.y) * ps.beta.y;   // CS:187
    ps.L.z += (BRDF.z * le.z) * ps.beta.z; ps.L.w += (BRDF.w * le.w) * ps.beta.w;
    const float u = rnd(rng), v = rnd(rng);
    if (STATS) st.rand_calls += 2;
    float pdf;
    const V3 nd = cosine_hemisphere(ps.n, u, v, pdf);                          // CS:189
    ps.pdf = pdf;
    const float cos_b = fabsf(dot(ps.n, nd));                                  // CS:190
    ps.beta.x *= (BRDF.x * cos_b) / pdf; ps.beta.y *= (BRDF.y * cos_b) / pdf;  // CS:192
    ps.beta.z *= (BRDF.z * cos_b) / pdf; ps.beta.w *= (BRDF.w * cos_b) / pdf;
    ps.d = nd;                                                                 // origin is already the hit point
    ps.specular = false;
    ps.shadow_phase = false;
  }
  // Russian roulette, CS:279-289
  const float rx = ps.beta.x * ps.etaScale, ry = ps.beta.y * ps.etaScale, rz = ps.beta.z * ps.etaScale;
  const float mb = fmaxf(rx, fmaxf(ry, rz));
  if (ps.depth > 1 && mb < 1.0f) {
    const float q = fmaxf(0.0f, 1.0f - mb);
    if (STATS) st.rand_calls += 1;
    if (rnd(rng) < q) { if (STATS) term[0] = TERM_RR; return false; }
    const float k = 1.0f - q;
    ps.beta.x /= k; ps.beta.y /= k; ps.beta.z /= k; ps.beta.w /= k;
  }
  ps.depth++;                                                                  // CS:292
  return true;
}

// seed + camera_ray + sample_wavelengths + path_trace's initial state, CS:98-101, 121-131
// TEA_TAB: seed lane w from the per-pixel table (wavefront kernels, which wait on memory anyway) or hashed here (the
// megakernel is FP32/issue-bound on linear-list scenes: the dependent load cost it 5 %).
template <bool STATS, bool TEA_TAB>
__device__ __forceinline__ void path_begin(const DevScene& sc, uint32_t px, uint32_t py, uint32_t sample, PathState& ps, Rng& rng, Stats& st) {
  rng.x = py; rng.y = px * 100u; rng.z = sample;
  rng.w = TEA_TAB ? __ldg(&sc.tea_tab[px + py * sc.W]) : tea(px, py * 100u);   // CS:98
  const float G = (float)sc.grid_size;
  const float cell = (float)(sample % sc.grid_size);
  const float s = ((float)px + (cell + rnd(rng)) / G) / (float)sc.W;            // CS:497
  const float t = ((float)sc.H - (float)py + (cell + rnd(rng)) / G) / (float)sc.H;  // CS:498
  const V3 eye = v3(sc.eye);
  const V3 p = madd(madd(v3(sc.llc), s, v3(sc.hor)), t, v3(sc.ver));
  ps.o = eye;
  ps.d = normalize(p - eye);                                                   // CS:493
  ps.lam0 = f2u_sat(301.0f * rnd(rng));                                        // CS:317-319
  if (STATS) st.rand_calls += 3;
  ps.L = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
  ps.beta = make_float4(1.0f, 1.0f, 1.0f, 1.0f);
  ps.pdf = 1.0f; ps.etaScale = 1.0f;
  ps.exclude = CRT_MAX_U32; ps.depth = 0;
  ps.specular = false; ps.in_transmission = false; ps.shadow_phase = false;
}

// Frames sample_start+1 .. sample_start+n_frames (UpdateVariables.wgsl:3-8 increments before each dispatch).
//
// Work unit = one pixel with all of this launch's frames; a lane owns its pixel for the whole launch, so
// the accumulator is read once, the samples are added in frame order in a register (bit-identical to
// n separate dispatches) and accumulator + framebuffer are written once.  A lane whose path ends starts
// its next sample (or fetches its next pixel from a global counter, one atomic per warp) while the other

// intersect for the current ray of `ps` (extension or shadow), then path_consume
template <bool STATS>
__device__ __forceinline__ bool path_step(const DevScene& sc, const float* __restrict__ spectra, PathState& ps, Rng& rng,
                                          uint32_t& n_ext, uint32_t& n_shadow, Stats& st, uint32_t* term) {
  Hit h;
  if (!ps.shadow_phase) {
    n_ext++;
    h = intersect<STATS>(sc, ps.o, ps.d, ps.exclude, st);
  } else {
    n_shadow++;
    const uint32_t include = __float_as_uint(light_at(sc, ps.light).d1.w);
    h = shadow_intersect<STATS>(sc, ps.o, ps.d, include, ps.exclude, st);
  }
  return path_consume<STATS>(sc, spectra, ps, rng, h, st, term);
}
