/*
 * crt_oracle.c -- CPU restatement of ComputeRayTracer's path-tracing compute shader.
 *
 * TEST INFRASTRUCTURE, NOT PRODUCT.  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may build, load or call this file.  libcrt never does.
 *
 * PARITY STATUS: UNPINNED BY THE REFERENCE.  The reference (Meryx/ComputeRayTracer) has no
 * tests, golden vectors or fixtures for this path (package.json:9 is a stub) and its WGSL
 * cannot be executed on this box (no graphics stack).  The pins this oracle is checked against
 * are the survey-derived known answers of SURVEY.md Appendix B (tests/test_oracle_kat.py),
 * which follow from the integer/f32 semantics of the cited shader lines.
 *
 * What it follows (CS = src/shaders/ComputeShader.wgsl of the reference):
 *   main CS:77-117, path_trace CS:119-295, power_heuristic CS:297-302, ray_at CS:304-307,
 *   sample_spectrum CS:310-313, sample_wavelengths CS:315-322, sample_CIE* CS:324-338,
 *   sample_lights/sample_light CS:341-355, compute_light_pdf CS:357-377,
 *   compute_light_radiance CS:379-408, spectral_to_xyz CS:411-426, xyz_to_linear_rgb CS:428-434,
 *   tone_map CS:436-439, gamma_correct CS:441-467, camera_* CS:470-500, intersect CS:503-518,
 *   ray_intersection CS:520-632, shadow_intersect CS:697-705, create_intersection_context
 *   CS:709-721, cosine_weighted_sample_hemisphere CS:751-774, fresnel_s CS:814-837,
 *   tea/pcg4d/rand CS:865-899, and src/shaders/UpdateVariables.wgsl:3-8 (sample++).
 *
 * Implementation-defined WGSL behaviour is pinned as DESIGN.md "Numerics contract" states
 * (SURVEY.md Appendix A with explicit fmaf in dot/cross/ray_at and friends).  Category 2
 * (triangle, Moller-Trumbore) is this project's extension and has no reference counterpart.
 *
 * Build: gcc -O2 -ffp-contract=off -mfma -fno-math-errno -shared -fPIC -pthread (oracle/Makefile).
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <stdio.h>
#include <math.h>
#include <pthread.h>

#include "../include/crt.h"
#include "../include/crt_detmath.h"

#define ORC_API __attribute__((visibility("default")))

/* ------------------------------------------------------------------ constants (CS:11-20) */
#define PI_F 3.14159265359f
#define INFINITY_F 2139095040.0f /* f32(0x7F800000 as an integer), CS:12 -- not +inf */
#define MAX_U32 0xFFFFFFFFu
#define LAMBDA_RANGE 301u
enum { MAT_DIFFUSE = 0, MAT_LIGHT = 1, MAT_GLASS = 2 };

typedef struct { float x, y, z; } v3;
typedef struct { float x, y, z, w; } v4;

/* 80-byte Primitive record, CS:41-47 / MJ:211-246 */
typedef struct {
  uint32_t category; uint32_t pad0[3];
  float d1[3]; uint32_t pad1;
  float d2[3]; uint32_t pad2;
  float d3[3]; uint32_t pad3;
  uint32_t d4[4];
} prim_t;

/* ------------------------------------------------------------------ vector pins */
static inline v3 V3(float x, float y, float z) { v3 r = {x, y, z}; return r; }
static inline v3 ld3(const float* p) { return V3(p[0], p[1], p[2]); }
static inline v3 add3(v3 a, v3 b) { return V3(a.x + b.x, a.y + b.y, a.z + b.z); }
static inline v3 sub3(v3 a, v3 b) { return V3(a.x - b.x, a.y - b.y, a.z - b.z); }
static inline v3 neg3(v3 a) { return V3(-a.x, -a.y, -a.z); }
static inline v3 scale3(v3 a, float s) { return V3(a.x * s, a.y * s, a.z * s); }
static inline v3 div3(v3 a, float s) { return V3(a.x / s, a.y / s, a.z / s); }
static inline float dot3(v3 a, v3 b) { return fmaf(a.z, b.z, fmaf(a.y, b.y, a.x * b.x)); }
static inline v3 cross3(v3 a, v3 b) {
  return V3(fmaf(a.y, b.z, -(a.z * b.y)), fmaf(a.z, b.x, -(a.x * b.z)), fmaf(a.x, b.y, -(a.y * b.x)));
}
static inline float length3(v3 a) { return sqrtf(dot3(a, a)); }
static inline v3 normalize3(v3 a) { return div3(a, length3(a)); }
/* a + s*b, one fmaf per lane (ray_at CS:306 and friends) */
static inline v3 madd3(v3 a, float s, v3 b) { return V3(fmaf(s, b.x, a.x), fmaf(s, b.y, a.y), fmaf(s, b.z, a.z)); }
static inline uint32_t f2u_sat(float f) { /* WGSL u32(f32): truncate, saturate, NaN -> 0 */
  if (!(f > 0.0f)) return 0u;
  if (f >= 4294967296.0f) return MAX_U32;
  return (uint32_t)f;
}

/* ------------------------------------------------------------------ RNG (CS:865-899) */
ORC_API uint32_t orc_tea(uint32_t val0, uint32_t val1) {
  uint32_t v0 = val0, v1 = val1, s0 = 0;
  for (int n = 0; n < 16; n++) {
    s0 += 0x9e3779b9u;
    v0 += ((v1 << 4) + 0xa341316cu) ^ (v1 + s0) ^ ((v1 >> 5) + 0xc8013ea4u);
    v1 += ((v0 << 4) + 0xad90777du) ^ (v0 + s0) ^ ((v0 >> 5) + 0x7e95761eu);
  }
  return v0;
}

ORC_API void orc_pcg4d(uint32_t s[4]) {
  for (int i = 0; i < 4; i++) s[i] = s[i] * 1664525u + 1013904223u;
  s[0] += s[1] * s[3]; s[1] += s[2] * s[0]; s[2] += s[0] * s[1]; s[3] += s[1] * s[2];
  for (int i = 0; i < 4; i++) s[i] ^= s[i] >> 16;
  s[0] += s[1] * s[3]; s[1] += s[2] * s[0]; s[2] += s[0] * s[1]; s[3] += s[1] * s[2];
}

ORC_API void orc_seed(uint32_t px, uint32_t py, uint32_t sample, uint32_t out[4]) { /* CS:98 */
  out[0] = py; out[1] = px * 100u; out[2] = sample; out[3] = orc_tea(px, py * 100u);
}

/* ------------------------------------------------------------------ scene */
typedef struct bvh_node { float lo[3], hi[3]; uint32_t left, right, first, count; } bvh_node;

typedef struct orc_scene {
  const prim_t* prims; uint32_t n_prims;
  const prim_t* lights; uint32_t n_lights;
  const float* cie;            /* 3 x 471 */
  const float* spectra; uint32_t n_spectra; /* n x 301 */
  crt_params prm;
  /* camera constants (pixel-invariant part of camera_ray, CS:478-487) */
  uint32_t W, H;
  v3 eye, hor, ver, llc;
  /* optional BVH over category-2 primitives (oracle-internal accelerator, validated against
   * the linear scan by tests/test_oracle_bvh.py) */
  int use_bvh;
  bvh_node* nodes; uint32_t n_nodes;
  uint32_t* tri_ids; uint32_t n_tris;   /* primitive indices (positions in prims[]) of triangles, BVH order */
  uint32_t* lin_ids; uint32_t n_lin;    /* positions of non-triangle primitives, ascending */
} orc_scene;

typedef struct {
  uint64_t c[CRT_NUM_COUNTERS];
} orc_counters;

typedef struct { v3 o, d; } ray_t;

typedef struct {
  /* ShapeIntersection CS:54-60 */
  v3 position, normal; uint32_t emission_index, reflectance_index, material;
  /* IntersectionContext CS:62-70 */
  float t_min, t_max; uint32_t index, exclude; int hit; v3 ray_origin, ray_direction;
} isect_t;

typedef struct { uint32_t s[4]; orc_counters* cnt; } rng_t;

static inline float rnd(rng_t* r) { /* CS:893-897 */
  orc_pcg4d(r->s);
  r->cnt->c[CRT_CNT_RAND]++;
  return (float)(r->s[0] & 0x00ffffffu) / 16777216.0f;
}

/* ------------------------------------------------------------------ ray_intersection CS:520-632 */
static inline void ray_intersection(const prim_t* p, ray_t ray, isect_t* c, orc_counters* cnt) {
  uint32_t category = p->category;
  if (category == 0) {
    cnt->c[CRT_CNT_PRIM_TESTS]++;
    uint32_t index = p->d4[3];
    if (c->exclude == index) return;
    v3 edge1 = ld3(p->d2), edge2 = ld3(p->d3);
    v3 normal = normalize3(cross3(edge1, edge2));
    v3 direction = ray.d;
    float ndotd = dot3(normal, direction);
    if (ndotd > 0.0f) normal = neg3(normal);
    ndotd = dot3(normal, direction);
    if (fabsf(ndotd) < 0.0001f) return;
    v3 origin = ray.o;
    v3 oo = sub3(ld3(p->d1), origin);
    float t = dot3(normal, oo) / ndotd;
    if (t < c->t_min || t > c->t_max) return;
    v3 P = madd3(origin, t, direction);
    v3 m = sub3(P, ld3(p->d1));
    float u = dot3(m, edge1) / dot3(edge1, edge1);
    float v = dot3(m, edge2) / dot3(edge2, edge2);
    if (u < 0.0f || u > 1.0f || v < 0.0f || v > 1.0f) return;
    c->position = P; c->normal = normal;
    c->emission_index = p->d4[0]; c->reflectance_index = p->d4[1]; c->material = p->d4[2];
    c->t_max = t; c->index = index; c->hit = 1; c->ray_origin = origin; c->ray_direction = direction;
    return;
  }
  if (category == 1) {
    cnt->c[CRT_CNT_PRIM_TESTS]++;
    uint32_t index = p->d4[3];
    if (c->exclude == index) return;
    v3 center = ld3(p->d1);
    float radius = p->d2[0];
    float radius_squared = radius * radius;
    v3 origin = ray.o, direction = ray.d;
    v3 co = sub3(origin, center);
    float a = dot3(direction, direction);
    float b = 2.0f * dot3(direction, co);
    float cc = dot3(co, co) - radius_squared;
    float disc = fmaf(b, b, -((4.0f * a) * cc));
    if (disc <= 0.0f) return;
    float sq = sqrtf(disc);
    float t = (-b - sq) / (2.0f * a);
    if (t < c->t_min || t > c->t_max) {
      t = (-b + sq) / (2.0f * a);
      if (t < c->t_min || t > c->t_max) return;
    }
    v3 P = madd3(origin, t, direction);
    v3 normal = normalize3(sub3(P, center));
    c->position = P; c->normal = normal;
    c->emission_index = p->d4[0]; c->reflectance_index = p->d4[1]; c->material = p->d4[2];
    c->t_max = t; c->index = index; c->hit = 1; c->ray_origin = origin; c->ray_direction = direction;
    return;
  }
  if (category == 2) {
    /* EXTENSION (no reference counterpart): triangle (v0=data1, e1=data2, e2=data3),
     * Moller-Trumbore, written accept-if-good so NaN never hits; geometric normal faces the ray
     * like a patch's (CS:541-544).  See DESIGN.md "Triangle primitive". */
    cnt->c[CRT_CNT_TRI_TESTS]++;
    uint32_t index = p->d4[3];
    if (c->exclude == index) return;
    v3 e1 = ld3(p->d2), e2 = ld3(p->d3);
    v3 d = ray.d;
    v3 pvec = cross3(d, e2);
    float det = dot3(e1, pvec);
    if (!(fabsf(det) >= 1e-8f)) return;
    float inv = 1.0f / det;
    v3 tvec = sub3(ray.o, ld3(p->d1));
    float u = dot3(tvec, pvec) * inv;
    if (!(u >= 0.0f && u <= 1.0f)) return;
    v3 qvec = cross3(tvec, e1);
    float v = dot3(d, qvec) * inv;
    if (!(v >= 0.0f && u + v <= 1.0f)) return;
    float t = dot3(e2, qvec) * inv;
    if (!(t >= c->t_min && t <= c->t_max)) return;
    v3 normal = normalize3(cross3(e1, e2));
    if (dot3(normal, d) > 0.0f) normal = neg3(normal);
    c->position = madd3(ray.o, t, d); c->normal = normal;
    c->emission_index = p->d4[0]; c->reflectance_index = p->d4[1]; c->material = p->d4[2];
    c->t_max = t; c->index = index; c->hit = 1; c->ray_origin = ray.o; c->ray_direction = d;
    return;
  }
}

static inline void ctx_init(isect_t* c, uint32_t exclude, float t_min) { /* CS:709-721 + zero-init CS:505 */
  memset(c, 0, sizeof(*c));
  c->t_min = t_min; c->t_max = INFINITY_F; c->index = MAX_U32; c->exclude = exclude; c->hit = 0;
}

/* ------------------------------------------------------------------ oracle BVH (accelerator) */
/* Equal-t ties must go to the highest primitive index (CS:557 accepts t == t_max, so a later
 * primitive replaces an earlier one).  The linear scan gets that from its order; an unordered
 * traversal needs it explicitly. */
static inline void tri_test_unordered(const prim_t* p, ray_t ray, isect_t* c, orc_counters* cnt) {
  isect_t trial = *c;
  ray_intersection(p, ray, &trial, cnt);
  if (trial.index != c->index || trial.t_max != c->t_max) {
    /* accepted: trial.t_max <= c->t_max */
    if (trial.t_max < c->t_max || c->index == MAX_U32 || trial.index > c->index) *c = trial;
  }
}

static void bvh_traverse(const orc_scene* sc, ray_t ray, isect_t* c, orc_counters* cnt) {
  if (!sc->n_nodes) return;
  /* NaN / inf rays: box tests are meaningless; triangles reject NaN by construction, so the
   * answer is "no triangle hit" -- but stay literal and test them all. */
  float idx = 1.0f / ray.d.x, idy = 1.0f / ray.d.y, idz = 1.0f / ray.d.z;
  uint32_t stack[128]; int sp = 0; stack[sp++] = 0;
  while (sp) {
    const bvh_node* n = &sc->nodes[stack[--sp]];
    cnt->c[CRT_CNT_NODE_VISITS]++;
    float t0 = (n->lo[0] - ray.o.x) * idx, t1 = (n->hi[0] - ray.o.x) * idx;
    float tn = fminf(t0, t1), tf = fmaxf(t0, t1);
    t0 = (n->lo[1] - ray.o.y) * idy; t1 = (n->hi[1] - ray.o.y) * idy;
    tn = fmaxf(tn, fminf(t0, t1)); tf = fminf(tf, fmaxf(t0, t1));
    t0 = (n->lo[2] - ray.o.z) * idz; t1 = (n->hi[2] - ray.o.z) * idz;
    tn = fmaxf(tn, fminf(t0, t1)); tf = fminf(tf, fmaxf(t0, t1));
    /* conservative: written so NaN passes */
    if (tn > tf || tf < c->t_min || tn > c->t_max) continue;
    if (n->count) {
      for (uint32_t i = 0; i < n->count; i++) tri_test_unordered(&sc->prims[sc->tri_ids[n->first + i]], ray, c, cnt);
    } else {
      if (sp + 2 > 128) { fprintf(stderr, "oracle bvh stack overflow\n"); abort(); }
      stack[sp++] = n->left; stack[sp++] = n->right;
    }
  }
}

typedef struct { float c[3]; float lo[3], hi[3]; uint32_t id; } build_item;
static int g_axis;
static int cmp_item(const void* a, const void* b) {
  const build_item* x = (const build_item*)a; const build_item* y = (const build_item*)b;
  if (x->c[g_axis] < y->c[g_axis]) return -1;
  if (x->c[g_axis] > y->c[g_axis]) return 1;
  return (x->id > y->id) - (x->id < y->id);
}
static uint32_t bvh_build_rec(orc_scene* sc, build_item* it, uint32_t first, uint32_t count, float pad) {
  uint32_t me = sc->n_nodes++;
  bvh_node* n = &sc->nodes[me];
  float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
  float clo[3] = {INFINITY, INFINITY, INFINITY}, chi[3] = {-INFINITY, -INFINITY, -INFINITY};
  for (uint32_t i = first; i < first + count; i++)
    for (int a = 0; a < 3; a++) {
      lo[a] = fminf(lo[a], it[i].lo[a]); hi[a] = fmaxf(hi[a], it[i].hi[a]);
      clo[a] = fminf(clo[a], it[i].c[a]); chi[a] = fmaxf(chi[a], it[i].c[a]);
    }
  for (int a = 0; a < 3; a++) { n->lo[a] = lo[a] - pad; n->hi[a] = hi[a] + pad; }
  n->left = n->right = 0; n->first = first; n->count = 0;
  if (count <= 4) { n->count = count; return me; }
  int ax = 0; float ext = chi[0] - clo[0];
  if (chi[1] - clo[1] > ext) { ax = 1; ext = chi[1] - clo[1]; }
  if (chi[2] - clo[2] > ext) { ax = 2; ext = chi[2] - clo[2]; }
  g_axis = ax;
  qsort(it + first, count, sizeof(build_item), cmp_item);
  uint32_t half = count / 2;
  uint32_t l = bvh_build_rec(sc, it, first, half, pad);
  uint32_t r = bvh_build_rec(sc, it, first + half, count - half, pad);
  n = &sc->nodes[me];
  n->left = l; n->right = r;
  return me;
}

static void scene_build_accel(orc_scene* sc) {
  uint32_t nt = 0, nl = 0;
  for (uint32_t i = 0; i < sc->n_prims; i++) { if (sc->prims[i].category == 2) nt++; else nl++; }
  sc->n_tris = nt; sc->n_lin = nl;
  sc->lin_ids = (uint32_t*)malloc(sizeof(uint32_t) * (nl ? nl : 1));
  sc->tri_ids = (uint32_t*)malloc(sizeof(uint32_t) * (nt ? nt : 1));
  build_item* it = (build_item*)malloc(sizeof(build_item) * (nt ? nt : 1));
  uint32_t a = 0, b = 0;
  float slo[3] = {INFINITY, INFINITY, INFINITY}, shi[3] = {-INFINITY, -INFINITY, -INFINITY};
  for (uint32_t i = 0; i < sc->n_prims; i++) {
    const prim_t* p = &sc->prims[i];
    if (p->category != 2) { sc->lin_ids[a++] = i; continue; }
    build_item* q = &it[b++];
    q->id = i;
    for (int k = 0; k < 3; k++) {
      float v0 = p->d1[k], v1 = p->d1[k] + p->d2[k], v2 = p->d1[k] + p->d3[k];
      q->lo[k] = fminf(v0, fminf(v1, v2)); q->hi[k] = fmaxf(v0, fmaxf(v1, v2));
      q->c[k] = 0.5f * (q->lo[k] + q->hi[k]);
      slo[k] = fminf(slo[k], q->lo[k]); shi[k] = fmaxf(shi[k], q->hi[k]);
    }
  }
  sc->n_nodes = 0; sc->nodes = NULL;
  if (nt) {
    /* box padding: covers the rounding of t, of the hit point and of the slab test */
    float ext = fmaxf(shi[0] - slo[0], fmaxf(shi[1] - slo[1], shi[2] - slo[2]));
    float maxabs = 0.0f;   /* f32 rounds vertices and o - v0 to an ulp of the coordinate: the pad scales with it too */
    for (int k = 0; k < 3; k++) maxabs = fmaxf(maxabs, fmaxf(fabsf(slo[k]), fabsf(shi[k])));
    float pad = fmaxf(fmaxf(ext, maxabs), 1.0f) * (1.0f / 16384.0f);
    sc->nodes = (bvh_node*)malloc(sizeof(bvh_node) * (2 * (size_t)nt));
    bvh_build_rec(sc, it, 0, nt, pad);
    for (uint32_t i = 0; i < nt; i++) sc->tri_ids[i] = it[i].id;
  }
  free(it);
}

/* ------------------------------------------------------------------ intersect CS:503-518 */
static void intersect(const orc_scene* sc, ray_t ray, uint32_t exclude, isect_t* out, orc_counters* cnt) {
  ctx_init(out, exclude, sc->prm.t_min);
  if (!sc->use_bvh) {
    for (uint32_t i = 0; i < sc->n_prims; i++) ray_intersection(&sc->prims[i], ray, out, cnt);
    return;
  }
  /* accelerated: same answer, different order (tests prove it against the loop above) */
  for (uint32_t i = 0; i < sc->n_lin; i++) ray_intersection(&sc->prims[sc->lin_ids[i]], ray, out, cnt);
  /* a patch/sphere with a higher index than a triangle wins an exact tie; tri_test_unordered's
   * rule handles both directions because it compares primitive indices */
  bvh_traverse(sc, ray, out, cnt);
}

static void shadow_intersect(const orc_scene* sc, ray_t ray, uint32_t include, uint32_t exclude, isect_t* out,
                             orc_counters* cnt) { /* CS:697-705 */
  intersect(sc, ray, exclude, out, cnt);
  if (out->index != include) out->hit = 0;
}

/* ------------------------------------------------------------------ spectra, lights */
static inline v4 sample_spectrum(const orc_scene* sc, uint32_t index, const uint32_t lam[4]) { /* CS:310-313 */
  if (index >= sc->n_spectra) index = sc->n_spectra - 1; /* OOB pin: clamp */
  const float* s = sc->spectra + (size_t)index * LAMBDA_RANGE;
  v4 r = {s[lam[0]], s[lam[1]], s[lam[2]], s[lam[3]]};
  return r;
}

static inline float power_heuristic(float f_pdf, float g_pdf) { /* CS:297-302 with nf = ng = 1 */
  float f = 1.0f * f_pdf, g = 1.0f * g_pdf;
  return (f * f) / (f * f + g * g);
}

static float compute_light_pdf(const orc_scene* sc, const isect_t* is) { /* CS:357-377 */
  uint32_t li = is->emission_index;
  if (li >= sc->n_lights) li = sc->n_lights - 1; /* CS:362 is out of bounds in the default scene: clamp pin */
  const prim_t* light = &sc->lights[li];
  float light_area = length3(ld3(light->d2)) * length3(ld3(light->d3));
  float light_area_pdf = 1.0f / light_area;
  float abs_cos_theta = fmaxf(0.00001f, fabsf(dot3(is->normal, neg3(is->ray_direction))));
  float distance = length3(sub3(is->position, is->ray_origin));
  float distance_squared = distance * distance; /* pow(distance, 2) pin */
  float geometric_term = abs_cos_theta / distance_squared;
  float light_solid_angle_pdf = light_area_pdf / geometric_term;
  float number_of_lights = (float)sc->n_lights;
  float light_selection_pdf = 1.0f / number_of_lights;
  return light_selection_pdf * light_solid_angle_pdf;
}

static v4 compute_light_radiance(const orc_scene* sc, const isect_t* is, const uint32_t lam[4], rng_t* rng,
                                 orc_counters* cnt) { /* CS:379-408 */
  /* sample_lights CS:341-347 */
  float u0 = rnd(rng);
  uint32_t li = f2u_sat((float)sc->n_lights * u0);
  if (li >= sc->n_lights) li = sc->n_lights - 1;
  const prim_t* light = &sc->lights[li];
  /* sample_light CS:349-355 */
  float u = rnd(rng), v = rnd(rng);
  v3 point_on_light = madd3(madd3(ld3(light->d1), u, ld3(light->d2)), v, ld3(light->d3));
  v3 light_dir = normalize3(sub3(point_on_light, is->position));
  ray_t sray = {is->position, light_dir};
  isect_t sh;
  cnt->c[CRT_CNT_RAYS_SHADOW]++;
  shadow_intersect(sc, sray, light->d4[3], is->index, &sh, cnt);
  float cos_theta = fmaxf(0.0f, dot3(is->normal, light_dir));
  v4 spec = sample_spectrum(sc, light->d4[0], lam);
  v4 le = {spec.x * cos_theta, spec.y * cos_theta, spec.z * cos_theta, spec.w * cos_theta};
  v4 zero = {0, 0, 0, 0};
  if (sh.hit) {
    float pdf_l = compute_light_pdf(sc, &sh);
    float pdf_b = cos_theta / PI_F;
    float w = power_heuristic(pdf_l, pdf_b);
    v4 t = {le.x * w / pdf_l, le.y * w / pdf_l, le.z * w / pdf_l, le.w * w / pdf_l};
    return t;
  }
  return zero;
}

static v3 cosine_weighted_sample_hemisphere(v3 normal, float* pdf, rng_t* rng) { /* CS:751-774 */
  float u = rnd(rng), v = rnd(rng);
  float r = sqrtf(u);
  float theta = (2.0f * PI_F) * v;
  float sn, cs;
  crt_sincosf(theta, &sn, &cs);
  float x = r * cs, y = r * sn;
  float z = sqrtf(fmaxf(0.0f, 1.0f - u));
  v3 up = (fabsf(normal.z) < 0.999f) ? V3(0, 0, 1) : V3(1, 0, 0);
  v3 tangent = normalize3(cross3(up, normal));
  v3 bitangent = cross3(normal, tangent);
  v3 dir = V3(fmaf(normal.x, z, fmaf(bitangent.x, y, tangent.x * x)),
              fmaf(normal.y, z, fmaf(bitangent.y, y, tangent.y * x)),
              fmaf(normal.z, z, fmaf(bitangent.z, y, tangent.z * x)));
  *pdf = z / PI_F;
  return dir;
}

static float fresnel_s(v3 ray_dir, v3 normal, float eta1, float eta2) { /* CS:814-837 */
  float cosi = fminf(fmaxf(dot3(ray_dir, normal), -1.0f), 1.0f);
  float eta = eta1 / eta2;
  if (cosi > 0.0f) eta = eta2 / eta1;
  float sint2 = (eta * eta) * (1.0f - cosi * cosi);
  if (sint2 > 1.0f) return 1.0f;
  float cost = sqrtf(1.0f - sint2);
  cosi = fabsf(cosi);
  float Rs = ((eta1 * cosi) - (eta2 * cost)) / ((eta1 * cosi) + (eta2 * cost));
  float Rp = ((eta2 * cosi) - (eta1 * cost)) / ((eta2 * cosi) + (eta1 * cost));
  return (Rs * Rs + Rp * Rp) / 2.0f;
}

static inline v3 reflect3(v3 e1, v3 e2) { /* WGSL reflect: e1 - 2*dot(e2,e1)*e2 */
  float k = 2.0f * dot3(e2, e1);
  return madd3(e1, -k, e2);
}
static inline v3 refract3(v3 e1, v3 e2, float e3) { /* WGSL refract */
  float d = dot3(e2, e1);
  float k = 1.0f - (e3 * e3) * (1.0f - d * d);
  if (k < 0.0f) return V3(0, 0, 0);
  float s = fmaf(e3, d, sqrtf(k));
  return madd3(scale3(e1, e3), -s, e2);
}

/* ------------------------------------------------------------------ path_trace CS:119-295 */
static v4 path_trace(const orc_scene* sc, ray_t ray, const uint32_t lam[4], rng_t* rng, orc_counters* cnt) {
  uint32_t depth = 0;
  v4 L = {0, 0, 0, 0};
  v4 beta = {1, 1, 1, 1};
  float last_bounce_pdf = 1.0f;
  uint32_t exclude = MAX_U32;
  int specular_bounce = 0;
  float etaScale = 1.0f;
  int inTransmission = 0;
  for (;;) {
    isect_t is;
    cnt->c[CRT_CNT_RAYS_EXT]++;
    intersect(sc, ray, exclude, &is, cnt);
    if (!is.hit) { cnt->c[CRT_CNT_TERM_MISS]++; break; }
    exclude = is.index;
    if (is.material == MAT_LIGHT) {
      v4 le = sample_spectrum(sc, is.emission_index, lam);
      if (depth == 0 || specular_bounce) {
        L.x += beta.x * le.x; L.y += beta.y * le.y; L.z += beta.z * le.z; L.w += beta.w * le.w;
      } else {
        float pdf_l = compute_light_pdf(sc, &is);
        float w = power_heuristic(last_bounce_pdf, pdf_l);
        L.x += (w * le.x) * beta.x; L.y += (w * le.y) * beta.y; L.z += (w * le.z) * beta.z; L.w += (w * le.w) * beta.w;
      }
      cnt->c[CRT_CNT_TERM_LIGHT]++;
      break;
    }
    if (depth >= sc->prm.max_depth) { cnt->c[CRT_CNT_TERM_DEPTH]++; break; }
    if (inTransmission) {
      float distance = length3(sub3(is.ray_origin, is.position));
      v4 ext = sample_spectrum(sc, sc->n_spectra - 1, lam);
      beta.x *= crt_expf(-ext.x * distance); beta.y *= crt_expf(-ext.y * distance);
      beta.z *= crt_expf(-ext.z * distance); beta.w *= crt_expf(-ext.w * distance);
    }
    if (is.material == MAT_DIFFUSE) {
      v4 refl = sample_spectrum(sc, is.reflectance_index, lam);
      v4 BRDF = {refl.x / PI_F, refl.y / PI_F, refl.z / PI_F, refl.w / PI_F};
      v4 le = compute_light_radiance(sc, &is, lam, rng, cnt);
      L.x += (BRDF.x * le.x) * beta.x; L.y += (BRDF.y * le.y) * beta.y;
      L.z += (BRDF.z * le.z) * beta.z; L.w += (BRDF.w * le.w) * beta.w;
      v3 nd = cosine_weighted_sample_hemisphere(is.normal, &last_bounce_pdf, rng);
      float cos_theta = fabsf(dot3(is.normal, nd));
      beta.x *= (BRDF.x * cos_theta) / last_bounce_pdf; beta.y *= (BRDF.y * cos_theta) / last_bounce_pdf;
      beta.z *= (BRDF.z * cos_theta) / last_bounce_pdf; beta.w *= (BRDF.w * cos_theta) / last_bounce_pdf;
      ray.o = is.position; ray.d = nd;
      specular_bounce = 0;
    }
    if (is.material == MAT_GLASS) {
      const float eta1 = 1.0f, eta2 = 1.5f;
      float eta = eta1 / eta2;
      float cos_theta = dot3(is.normal, ray.d);
      float reflected = fresnel_s(ray.d, is.normal, eta1, eta2);
      float pr = reflected;
      float pt = 1.0f - reflected;
      float u = rnd(rng);
      v3 current_direction = ray.d;
      v3 current_normal = is.normal;
      if (cos_theta > 0.0f) { eta = 1.0f / eta; current_normal = neg3(current_normal); }
      v3 new_direction;
      if (u < pr / (pr + pt)) {
        new_direction = reflect3(current_direction, current_normal);
        ray.o = is.position; specular_bounce = 1; exclude = MAX_U32;
      } else {
        specular_bounce = 1; exclude = MAX_U32;
        new_direction = normalize3(refract3(current_direction, current_normal, eta));
        ray.o = is.position;
        float e2 = eta * eta;
        beta.x *= e2; beta.y *= e2; beta.z *= e2; beta.w *= e2;
        etaScale /= e2;
        inTransmission = !inTransmission;
      }
      ray.d = new_direction;
    }
    float rx = beta.x * etaScale, ry = beta.y * etaScale, rz = beta.z * etaScale;
    float mb = fmaxf(rx, fmaxf(ry, rz));
    if (depth > 1 && mb < 1.0f) {
      float q = fmaxf(0.0f, 1.0f - mb);
      if (rnd(rng) < q) { cnt->c[CRT_CNT_TERM_RR]++; break; }
      float k = 1.0f - q;
      beta.x /= k; beta.y /= k; beta.z /= k; beta.w /= k;
    }
    depth++;
  }
  cnt->c[CRT_CNT_DEPTH_SUM] += depth;
  return L;
}

/* ------------------------------------------------------------------ colour CS:411-467 */
static v3 spectral_to_xyz(const orc_scene* sc, v4 L, const uint32_t lam[4]) {
  v3 r;
  float* o = &r.x;
  for (int k = 0; k < 3; k++) {
    const float* c = sc->cie + 471 * k + 40;
    float s = fmaf(c[lam[3]], L.w, fmaf(c[lam[2]], L.z, fmaf(c[lam[1]], L.y, c[lam[0]] * L.x)));
    o[k] = (s * 300.0f) / 427.42758f;
  }
  return r;
}
static inline float srgb_r(float c) { return c < 0.0031308f ? c * 12.92f : 1.055f * crt_powf(c, 0.41666666f) - 0.055f; }
static inline float srgb_g(float c) { return c < 0.0031308f ? c * (12.92f * c) : 1.055f * crt_powf(c, 0.41666666f) - 0.055f; }
static inline float srgb_b(float c) { return c < 0.0031308f ? 12.92f * c : 1.055f * crt_powf(c, 0.41666666f) - 0.055f; }
static inline uint8_t unorm8(float c) { /* textureStore rgba8unorm pin */
  if (c != c) return 0;
  c = fminf(fmaxf(c, 0.0f), 1.0f);
  return (uint8_t)floorf(c * 255.0f + 0.5f);
}
static void resolve_pixel(const orc_scene* sc, const float* acc, uint32_t sample, uint8_t* px) { /* CS:109-115 */
  float fs = (float)sample;
  float X = acc[0] / fs, Y = acc[1] / fs, Z = acc[2] / fs;
  float r = fmaf(-0.4985314f, Z, fmaf(-1.5371385f, Y, 3.2404542f * X));
  float g = fmaf(0.0415560f, Z, fmaf(1.8760108f, Y, -0.9692660f * X));
  float b = fmaf(1.0572252f, Z, fmaf(-0.2040259f, Y, 0.0556434f * X));
  float e = sc->prm.exposure;
  r = 1.0f - crt_expf(-r * e); g = 1.0f - crt_expf(-g * e); b = 1.0f - crt_expf(-b * e);
  px[0] = unorm8(srgb_r(r)); px[1] = unorm8(srgb_g(g)); px[2] = unorm8(srgb_b(b)); px[3] = 255;
}

/* ------------------------------------------------------------------ camera CS:470-500 */
static void camera_setup(orc_scene* sc, const float* cam) {
  v3 eye = V3(cam[0], cam[1], cam[2]), lookat = V3(cam[4], cam[5], cam[6]), up = V3(cam[8], cam[9], cam[10]);
  float vw = cam[11], vh = cam[12], fl = cam[13];
  sc->W = f2u_sat(vw); sc->H = f2u_sat(vh);
  v3 w = normalize3(sub3(eye, lookat));
  v3 u = normalize3(cross3(up, w));
  v3 v = cross3(w, u);
  float aspect = vw / vh;
  float vph = 2.0f * (float)tan((double)(fl / 2.0f)); /* tan pin: double libm, rounded once */
  float vpw = aspect * vph;
  v3 hor = scale3(u, vpw), ver = scale3(v, vph);
  sc->eye = eye; sc->hor = hor; sc->ver = ver;
  sc->llc = sub3(sub3(sub3(eye, div3(hor, 2.0f)), div3(ver, 2.0f)), w);
}

static ray_t camera_ray(const orc_scene* sc, uint32_t px, uint32_t py, uint32_t sample, rng_t* rng, float* s_out,
                        float* t_out) {
  float G = (float)sc->prm.grid_size;
  float cell = (float)(sample % sc->prm.grid_size);
  float s = ((float)px + (cell + rnd(rng)) / G) / (float)sc->W;
  float t = ((float)sc->H - (float)py + (cell + rnd(rng)) / G) / (float)sc->H;
  v3 p = madd3(madd3(sc->llc, s, sc->hor), t, sc->ver);
  ray_t r = {sc->eye, normalize3(sub3(p, sc->eye))};
  if (s_out) *s_out = s;
  if (t_out) *t_out = t;
  return r;
}

/* ------------------------------------------------------------------ one pixel-sample, CS:77-117 */
static void pixel_sample(const orc_scene* sc, uint32_t px, uint32_t py, uint32_t sample, float* acc4, orc_counters* cnt) {
  rng_t rng; rng.cnt = cnt;
  orc_seed(px, py, sample, rng.s);
  cnt->c[CRT_CNT_PATHS]++;
  ray_t ray = camera_ray(sc, px, py, sample, &rng, NULL, NULL);
  /* sample_wavelengths CS:315-322 */
  float u = rnd(&rng);
  uint32_t l0 = f2u_sat(301.0f * u);
  uint32_t lam[4] = {l0, (l0 + 4) % LAMBDA_RANGE, (l0 + 8) % LAMBDA_RANGE, (l0 + 12) % LAMBDA_RANGE};
  v4 L = path_trace(sc, ray, lam, &rng, cnt);
  v3 xyz = spectral_to_xyz(sc, L, lam);
  acc4[0] += xyz.x; acc4[1] += xyz.y; acc4[2] += xyz.z;
}

/* ------------------------------------------------------------------ public API */
ORC_API orc_scene* orc_scene_create(const void* prims, uint32_t n_prims, const void* lights, uint32_t n_lights,
                                    const float* camera16, const float* cie, const float* spectra, uint32_t n_spectra,
                                    const crt_params* prm, int use_bvh) {
  orc_scene* sc = (orc_scene*)calloc(1, sizeof(orc_scene));
  sc->prims = (const prim_t*)prims; sc->n_prims = n_prims;
  sc->lights = (const prim_t*)lights; sc->n_lights = n_lights;
  sc->cie = cie; sc->spectra = spectra; sc->n_spectra = n_spectra;
  sc->prm = *prm;
  camera_setup(sc, camera16);
  sc->use_bvh = use_bvh;
  if (use_bvh) scene_build_accel(sc);
  return sc;
}
ORC_API void orc_scene_destroy(orc_scene* sc) {
  if (!sc) return;
  free(sc->nodes); free(sc->tri_ids); free(sc->lin_ids); free(sc);
}
ORC_API void orc_scene_camera_constants(const orc_scene* sc, float out[12]) {
  const v3* v[4] = {&sc->eye, &sc->hor, &sc->ver, &sc->llc};
  for (int i = 0; i < 4; i++) { out[3 * i] = v[i]->x; out[3 * i + 1] = v[i]->y; out[3 * i + 2] = v[i]->z; }
}

typedef struct {
  const orc_scene* sc; float* acc; uint8_t* fb; uint32_t sample_start, n_frames;
  uint32_t rank, world, mode;
  volatile uint32_t* next_row; orc_counters cnt;
} job_t;

static int owns_pixel(const job_t* j, uint32_t x, uint32_t y) {
  if (j->mode != CRT_SHARD_TILES || j->world <= 1) return 1;
  uint32_t tiles_x = (j->sc->W + 7) / 8;
  uint32_t tile = (y / 4) * tiles_x + x / 8;
  return tile % j->world == j->rank;
}

static void* render_worker(void* arg) {
  job_t* j = (job_t*)arg;
  const orc_scene* sc = j->sc;
  for (;;) {
    uint32_t y = __sync_fetch_and_add(j->next_row, 1);
    if (y >= sc->H) break;
    for (uint32_t x = 0; x < sc->W; x++) {
      if (!owns_pixel(j, x, y)) continue;
      float* a = j->acc + 4 * ((size_t)x + (size_t)y * sc->W);
      uint32_t last = j->sample_start;
      for (uint32_t f = 1; f <= j->n_frames; f++) {
        uint32_t sample = j->sample_start + f; /* sample++ happens before the dispatch: first frame sees 1 */
        last = sample;
        if (j->mode == CRT_SHARD_SPLIT_SPP && j->world > 1 && (sample - 1) % j->world != j->rank) continue;
        pixel_sample(sc, x, y, sample, a, &j->cnt);
      }
      if (j->fb && j->n_frames) resolve_pixel(sc, a, last, j->fb + 4 * ((size_t)x + (size_t)y * sc->W));
    }
  }
  return NULL;
}

/* n_frames x { sample++ ; one sample per pixel }.  acc: W*H*4 f32 in/out.  fb: W*H*4 u8 out
 * (may be NULL).  counters: CRT_NUM_COUNTERS u64, added to (may be NULL). */
ORC_API int orc_render(const orc_scene* sc, uint32_t sample_start, uint32_t n_frames, float* acc, uint8_t* fb,
                       uint64_t* counters, int n_threads, uint32_t rank, uint32_t world, uint32_t mode) {
  if (n_threads < 1) n_threads = 1;
  if (n_threads > 256) n_threads = 256;
  volatile uint32_t next_row = 0;
  job_t* jobs = (job_t*)calloc((size_t)n_threads, sizeof(job_t));
  pthread_t* th = (pthread_t*)calloc((size_t)n_threads, sizeof(pthread_t));
  for (int i = 0; i < n_threads; i++) {
    jobs[i].sc = sc; jobs[i].acc = acc; jobs[i].fb = fb; jobs[i].sample_start = sample_start; jobs[i].n_frames = n_frames;
    jobs[i].rank = rank; jobs[i].world = world ? world : 1; jobs[i].mode = mode; jobs[i].next_row = &next_row;
    if (i) pthread_create(&th[i], NULL, render_worker, &jobs[i]);
  }
  render_worker(&jobs[0]);
  for (int i = 1; i < n_threads; i++) pthread_join(th[i], NULL);
  if (counters)
    for (int i = 0; i < n_threads; i++)
      for (int k = 0; k < CRT_NUM_COUNTERS; k++) counters[k] += jobs[i].cnt.c[k];
  free(jobs); free(th);
  return 0;
}

ORC_API int orc_resolve(const orc_scene* sc, const float* acc, uint32_t sample, uint8_t* fb) {
  for (size_t i = 0; i < (size_t)sc->W * sc->H; i++) resolve_pixel(sc, acc + 4 * i, sample, fb + 4 * i);
  return 0;
}

/* closest hit for caller-supplied rays; rays n x 8 floats (o.xyz, -, d.xyz, -) */
ORC_API int orc_trace_rays(const orc_scene* sc, const float* rays, const uint32_t* exclude, size_t n, uint32_t* hit_idx,
                           float* t, uint64_t* counters) {
  orc_counters cnt; memset(&cnt, 0, sizeof(cnt));
  for (size_t i = 0; i < n; i++) {
    ray_t r = {V3(rays[8 * i], rays[8 * i + 1], rays[8 * i + 2]), V3(rays[8 * i + 4], rays[8 * i + 5], rays[8 * i + 6])};
    isect_t is;
    intersect(sc, r, exclude ? exclude[i] : MAX_U32, &is, &cnt);
    hit_idx[i] = is.index; t[i] = is.t_max;
  }
  if (counters) for (int k = 0; k < CRT_NUM_COUNTERS; k++) counters[k] += cnt.c[k];
  return 0;
}

/* primary-ray probe for the Appendix B.3 pins: out8 = s, t, dir.xyz, hit index (as float bits), t_max, 0 */
ORC_API int orc_primary_probe(const orc_scene* sc, uint32_t px, uint32_t py, uint32_t sample, float* out8) {
  orc_counters cnt; memset(&cnt, 0, sizeof(cnt));
  rng_t rng; rng.cnt = &cnt;
  orc_seed(px, py, sample, rng.s);
  float s, t;
  ray_t ray = camera_ray(sc, px, py, sample, &rng, &s, &t);
  isect_t is;
  intersect(sc, ray, MAX_U32, &is, &cnt);
  out8[0] = s; out8[1] = t; out8[2] = ray.d.x; out8[3] = ray.d.y; out8[4] = ray.d.z;
  out8[5] = crt_u2f(is.index); out8[6] = is.t_max; out8[7] = 0.0f;
  return 0;
}

/* hit-index image of the primary rays of frame `sample` (B.3 histogram) */
ORC_API int orc_primary_hits(const orc_scene* sc, uint32_t sample, uint32_t* hit_idx, float* t_out) {
  orc_counters cnt; memset(&cnt, 0, sizeof(cnt));
  for (uint32_t y = 0; y < sc->H; y++)
    for (uint32_t x = 0; x < sc->W; x++) {
      rng_t rng; rng.cnt = &cnt;
      orc_seed(x, y, sample, rng.s);
      ray_t ray = camera_ray(sc, x, y, sample, &rng, NULL, NULL);
      isect_t is;
      intersect(sc, ray, MAX_U32, &is, &cnt);
      hit_idx[x + (size_t)y * sc->W] = is.index;
      if (t_out) t_out[x + (size_t)y * sc->W] = is.t_max;
    }
  return 0;
}

/* detmath probes so tests can compare the CUDA build of the same header bit for bit */
ORC_API void orc_detmath(const float* x, size_t n, float* sin_out, float* cos_out, float* exp_out, float* pow_out) {
  for (size_t i = 0; i < n; i++) {
    crt_sincosf(x[i], &sin_out[i], &cos_out[i]);
    exp_out[i] = crt_expf(x[i]);
    pow_out[i] = crt_powf(x[i], 0.41666666f);
  }
}
