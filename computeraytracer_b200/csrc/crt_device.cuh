// crt_device.cuh -- device-side building blocks of the path tracer (sm_100a).
//
// Everything here follows the live code of the reference's src/shaders/ComputeShader.wgsl
// ("CS:" citations) under the numerics contract of DESIGN.md: f32 throughout, no implicit
// contraction (-fmad=false), explicit fmaf in dot/cross/ray_at, IEEE div/sqrt, transcendental
// functions from include/crt_detmath.h.  Ray-invariant quantities (unit normals, e.e, r^2,
// camera frame) are hoisted to scene upload; the hoists are bit-identical to evaluating them
// per ray, which the parity tests against the CPU oracle (which does not hoist) prove.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/crt.h"
#include "../../include/crt_detmath.h"

#define CRT_PI_F 3.14159265359f         // CS:11
#define CRT_INFINITY_F 2139095040.0f    // CS:12: f32(0x7F800000 as an integer), not +inf
#define CRT_MAX_U32 0xFFFFFFFFu         // CS:13
#define CRT_LAMBDA_RANGE 301u           // CS:16-17, 318
#define CRT_MAT_DIFFUSE 0u              // CS:18
#define CRT_MAT_LIGHT 1u                // CS:19
#define CRT_MAT_GLASS 2u                // CS:20

#define CRT_LIN_PARAM_MAX 64            // patches/spheres carried in kernel-parameter (constant-bank) space
#define CRT_LIGHT_PARAM_MAX 8
#define CRT_TRI_REF 0x80000000u         // Hit::ref flag: slot in the triangle array
#define CRT_BVH_STACK 32            // node groups: at most one per level of the wide tree

// meta word: category | material << 2 | emission << 4 | reflectance << 16
#define CRT_META_CAT(m) ((m)&3u)
#define CRT_META_MAT(m) (((m) >> 2) & 3u)
#define CRT_META_EMIS(m) (((m) >> 4) & 0xFFFu)
#define CRT_META_REFL(m) (((m) >> 16) & 0xFFFu)

struct LinPrim {   // one patch or sphere of the linear list, 64 bytes
  float4 A;        // patch: unit normal of cross(e1,e2) | sphere: centre       ; w = bits(primitive index)
  float4 B;        // patch: origin                      | sphere: (r*r, r, 0)  ; w = bits(meta)
  float4 C;        // patch: edge1, w = dot(e1,e1)
  float4 D;        // patch: edge2, w = dot(e2,e2)
};

struct LightRec {  // one entry of binding 7 with compute_light_pdf's invariants hoisted (CS:363-364)
  float4 d1;       // origin;  w = bits(data4.w) (primitive index of the light)
  float4 d2;       // edge1;   w = bits(data4.x) (emission spectrum)
  float4 d3;       // edge2;   w = light_area_pdf = 1 / (|d2| * |d3|)
};

struct DevScene {
  // camera (pixel-invariant part of camera_ray, CS:478-487)
  float eye[3], hor[3], ver[3], llc[3];
  uint32_t W, H;
  // params (crt_params)
  uint32_t max_depth, grid_size;
  float t_min, exposure;
  // tables
  const float* spectra;    // n_spectra x 301
  const float* cie;        // 3 x 471
  uint32_t n_spectra;
  // lights
  uint32_t n_lights;
  const LightRec* lights_g;       // all lights (global memory)
  // linear list
  uint32_t n_lin;
  const LinPrim* lin_g;           // all entries (global); the first CRT_LIN_PARAM_MAX are mirrored below
  // triangles + BVH
  uint32_t n_tri, n_nodes;
  const float4* tri;              // 3 x float4 per triangle, BVH leaf order
  const uint4* nodes;             // 5 x 16 bytes per 8-wide node ("Wide8" below)
  uint32_t shadow_closest;        // 1: some light's primitive is a triangle -> shadow rays towards it need closest-hit
  const uint32_t* light_tri_mask; // shadow_closest scenes: bit li set = light li's primitive (data4.w) is a triangle
  float bvh_lo[3], bvh_hi[3];     // padded bounds of all triangles (root box)
  // sharding
  uint32_t rank, world, shard_mode;
  uint32_t tiles_x, tiles_y;
  const uint32_t* tea_tab;        // tea(x, y * 100) per pixel (CS:98): depends on the pixel only, hashed once per camera
  LightRec lights_c[CRT_LIGHT_PARAM_MAX];
  LinPrim lin_c[CRT_LIN_PARAM_MAX];
};

struct Stats {
  uint32_t prim_tests, node_visits, tri_tests, rand_calls;
};

// ------------------------------------------------------------------------------------------------ vectors
struct V3 { float x, y, z; };
__device__ __forceinline__ V3 v3(float x, float y, float z) { V3 r; r.x = x; r.y = y; r.z = z; return r; }
__device__ __forceinline__ V3 v3(const float4& a) { return v3(a.x, a.y, a.z); }
__device__ __forceinline__ V3 v3(const float* p) { return v3(p[0], p[1], p[2]); }
__device__ __forceinline__ V3 operator+(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ V3 operator-(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ V3 operator-(V3 a) { return v3(-a.x, -a.y, -a.z); }
__device__ __forceinline__ V3 operator*(V3 a, float s) { return v3(a.x * s, a.y * s, a.z * s); }
// a / b, IEEE-rounded like the plain operator (the build sets -prec-div=true).  ptxas's division sends a zero numerator to
// its slow path, a subroutine of ~40 instructions that the whole warp waits for; zeros are common here (components of
// axis-aligned normals, clamped cosines, the XYZ of a path that found no light: ncu counted 10 slow-path calls per warp
// and bounce, 12 % of wf_logic's instructions).  0 / b is a zero with the sign of a ^ b for every b that is neither zero
// nor NaN, so those lanes divide 1 / b on the fast path and the zero is selected afterwards.
#ifndef CRT_DIV_SHORTCUT
#define CRT_DIV_SHORTCUT 1
#endif
__device__ __forceinline__ float crt_div(float a, float b) {
#if CRT_DIV_SHORTCUT
  const bool z = a == 0.0f && fabsf(b) > 0.0f;   // false for b = 0 and b = NaN: those take the plain division
  const float q = (z ? 1.0f : a) / b;
  return z ? __uint_as_float((__float_as_uint(a) ^ __float_as_uint(b)) & 0x80000000u) : q;
#else
  return a / b;
#endif
}
__device__ __forceinline__ V3 operator/(V3 a, float s) { return v3(crt_div(a.x, s), crt_div(a.y, s), crt_div(a.z, s)); }
__device__ __forceinline__ float dot(V3 a, V3 b) { return fmaf(a.z, b.z, fmaf(a.y, b.y, a.x * b.x)); }
__device__ __forceinline__ V3 cross(V3 a, V3 b) {
  return v3(fmaf(a.y, b.z, -(a.z * b.y)), fmaf(a.z, b.x, -(a.x * b.z)), fmaf(a.x, b.y, -(a.y * b.x)));
}
__device__ __forceinline__ float length(V3 a) { return sqrtf(dot(a, a)); }
__device__ __forceinline__ V3 normalize(V3 a) { return a / length(a); }
__device__ __forceinline__ V3 madd(V3 a, float s, V3 b) { return v3(fmaf(s, b.x, a.x), fmaf(s, b.y, a.y), fmaf(s, b.z, a.z)); }
__device__ __forceinline__ uint32_t f2u_sat(float f) { return __float2uint_rz(f); }  // WGSL u32(f32): saturating, NaN -> 0

// ------------------------------------------------------------------------------------------------ RNG, CS:865-899
__device__ __forceinline__ uint32_t tea(uint32_t v0, uint32_t v1) {
  uint32_t s0 = 0;
#pragma unroll
  for (int n = 0; n < 16; n++) {
    s0 += 0x9e3779b9u;
    v0 += ((v1 << 4) + 0xa341316cu) ^ (v1 + s0) ^ ((v1 >> 5) + 0xc8013ea4u);
    v1 += ((v0 << 4) + 0xad90777du) ^ (v0 + s0) ^ ((v0 >> 5) + 0x7e95761eu);
  }
  return v0;
}

struct Rng { uint32_t x, y, z, w; };

__device__ __forceinline__ float rnd(Rng& s) {
  s.x = s.x * 1664525u + 1013904223u; s.y = s.y * 1664525u + 1013904223u;
  s.z = s.z * 1664525u + 1013904223u; s.w = s.w * 1664525u + 1013904223u;
  s.x += s.y * s.w; s.y += s.z * s.x; s.z += s.x * s.y; s.w += s.y * s.z;
  s.x ^= s.x >> 16; s.y ^= s.y >> 16; s.z ^= s.z >> 16; s.w ^= s.w >> 16;
  s.x += s.y * s.w; s.y += s.z * s.x; s.z += s.x * s.y; s.w += s.y * s.z;
  return (float)(s.x & 0x00ffffffu) / 16777216.0f;
}

// ------------------------------------------------------------------------------------------------ work units
// unit -> pixel: units enumerate a rank's 8x4-pixel tiles, 32 pixels each; under CRT_SHARD_TILES rank r owns the tiles with
// tile_id % world == r (interleaved, because cost per tile is very uneven: sky against glass silhouettes)
__device__ __forceinline__ bool unit_pixel_of(uint32_t unit, uint32_t rank, uint32_t world, bool tiles_sharded, uint32_t tiles_x, uint32_t W, uint32_t H,
                                              uint32_t& px, uint32_t& py) {
  uint32_t tile = unit >> 5;
  const uint32_t l = unit & 31u;
  if (tiles_sharded) tile = tile * world + rank;
  px = (tile % tiles_x) * 8u + (l & 7u);
  py = (tile / tiles_x) * 4u + (l >> 3);
  return px < W && py < H;
}
__device__ __forceinline__ bool unit_pixel(const DevScene& sc, uint32_t unit, uint32_t& px, uint32_t& py) {
  return unit_pixel_of(unit, sc.rank, sc.world, sc.shard_mode == CRT_SHARD_TILES && sc.world > 1, sc.tiles_x, sc.W, sc.H, px, py);
}

// ------------------------------------------------------------------------------------------------ intersection
struct Hit {
  float t;        // context.t_max after the scan (CS:577); CRT_INFINITY_F when nothing was accepted
  uint32_t idx;   // context.index: primitive index of the winner, CRT_MAX_U32 on a miss
  uint32_t ref;   // where the winner lives: linear-list slot, or CRT_TRI_REF | triangle slot
};

// `q = a / b; q < 0 || q > 1` (IEEE, round to nearest) without the division where the answer is the same bit for bit.
// For 0 < b < 1e18:  a >= 0 gives q >= 0, and q > 1 exactly when a > b (a > b means a >= b + ulp, whose quotient rounds to at
// least 1 + 2^-23; a <= b gives q <= 1; +-0 and NaN numerators fall out right: NaN compares false both ways, like the NaN
// quotient);  a <= -1e-18 gives |q| >= 1e-36, a normal negative number, so q < 0.  What is left -- a in (-1e-18, 0), whose
// quotient can underflow to -0 (which is NOT < 0), and b that is 0, negative, NaN or huge -- takes the real division.  The
// patch test (CS:563-566) uses u and v for nothing but this comparison; two of its three divisions go away for every
// ray that reaches it (ncu r01: IEEE divisions are 12 % of wf_logic's instructions).
__device__ __forceinline__ bool outside_unit(float a, float b) {
#ifndef CRT_UV_DIVIDE   // (A/B switch: the plain divisions)
  if (b > 0.0f && b < 1e18f && !(a < 0.0f && a > -1e-18f)) return a < 0.0f || a > b;
#endif
  const float q = a / b;
  return q < 0.0f || q > 1.0f;
}

// ray_intersection for one linear-list entry, CS:520-632, sequential semantics (called in index order).
// (The early exits stay: evaluating the patch test without them -- same tests, predicated -- made wf_logic 26 % slower;
// whole warps leave at the exits often enough.)
__device__ __forceinline__ void test_lin(const LinPrim& p, uint32_t slot, V3 o, V3 d, uint32_t exclude, float t_min, Hit& h) {
  uint32_t index = __float_as_uint(p.A.w);
  if (exclude == index) return;                      // CS:529 / CS:588
  uint32_t meta = __float_as_uint(p.B.w);
  if (CRT_META_CAT(meta) == 0u) {
    V3 n = v3(p.A);
    float ndotd = dot(n, d);
    // CS:541-546: flipping the normal negates n.d exactly, so |n.d| and the quotient below are unchanged
    if (fabsf(ndotd) < 0.0001f) return;
    V3 p0 = v3(p.B);
    float t = crt_div(dot(n, p0 - o), ndotd);        // CS:554
    if (t < t_min || t > h.t) return;                // CS:557
    V3 m = madd(o, t, d) - p0;                       // CS:561-562
    // CS:563-566: u = dot(m, e1) / dot(e1, e1), v likewise, rejected outside [0, 1] -- u and v are used for nothing else
    if (outside_unit(dot(m, v3(p.C)), p.C.w) || outside_unit(dot(m, v3(p.D)), p.D.w)) return;
    h.t = t; h.idx = index; h.ref = slot;
  } else {
    V3 co = o - v3(p.A);                             // CS:597
    float a = dot(d, d);
    float b = 2.0f * dot(d, co);
    float c = dot(co, co) - p.B.x;
    float disc = fmaf(b, b, -((4.0f * a) * c));      // CS:601
    if (disc <= 0.0f) return;
    float sq = sqrtf(disc);
    float t = (-b - sq) / (2.0f * a);                // CS:606
    if (t < t_min || t > h.t) {
      t = (-b + sq) / (2.0f * a);                    // CS:611
      if (t < t_min || t > h.t) return;
    }
    h.t = t; h.idx = index; h.ref = slot;
  }
}

// Triangle extension (category 2): Moller-Trumbore on (v0, e1, e2), accept-if-good so NaN never hits.
// Returns true and t when the ray hits inside [t_min, t_max].
__device__ __forceinline__ bool test_tri(float4 T0, float4 T1, float4 T2, V3 o, V3 d, float t_min, float t_max, float& t_out) {
  V3 e1 = v3(T1), e2 = v3(T2);
  V3 pvec = cross(d, e2);
  float det = dot(e1, pvec);
  if (!(fabsf(det) >= 1e-8f)) return false;
  float inv = 1.0f / det;
  V3 tvec = o - v3(T0);
  float u = dot(tvec, pvec) * inv;
  if (!(u >= 0.0f && u <= 1.0f)) return false;
  V3 qvec = cross(tvec, e1);
  float v = dot(d, qvec) * inv;
  if (!(v >= 0.0f && u + v <= 1.0f)) return false;
  float t = dot(e2, qvec) * inv;
  if (!(t >= t_min && t <= t_max)) return false;
  t_out = t;
  return true;
}

// "Wide8": an 8-wide BVH over the triangles with quantised child boxes (after Ylitie, Karras, Laine, "Efficient incoherent
// ray traversal on GPUs through compressed wide BVHs", HPG 2017), 80 bytes = 5 x 16-byte words per node:
//   w0 = (p.x, p.y, p.z, ex | ey << 8 | ez << 16 | imask << 24)     node origin, per-axis step s = 2^(e-127-15), inner-child mask
//   w1 = (child_base, tri_base, meta[0..3], meta[4..7])
//   w2 = (qlo.x[0..3], qlo.x[4..7], qlo.y[0..3], qlo.y[4..7])       child box planes: p + q * s, one byte per child
//   w3 = (qlo.z[0..3], qlo.z[4..7], qhi.x[0..3], qhi.x[4..7])
//   w4 = (qhi.y[0..3], qhi.y[4..7], qhi.z[0..3], qhi.z[4..7])
// meta[slot]: 0 = empty; inner child: 1 << 5 | (24 + slot), stored at node child_base + rank among the inner children;
// leaf: unary triangle count (1, 3, 7) << 5 | offset of its first triangle from tri_base (offset + count <= 24).
// A child's slot says on which side of the node centre it lies, so slot ^ (7 ^ ray octant) orders children near to far.
// The quantised planes are conservative in exact arithmetic around boxes that are already padded (crt_bvh.cpp), so culling
// with (tnear <= tfar), tfar clamped to the best t so far, never drops a primitive the linear scan would accept
// (DESIGN.md "Conservative culling"); equal-t ties go to the higher primitive index, which is what the reference's in-order
// scan with `t > t_max` rejection yields.
#ifndef CRT_NODE_STRIDE
#define CRT_NODE_STRIDE 5
#endif
#define CRT_WIDE_HITS 0xFF000000u   // group word: bits 24..31 = inner children still to visit (priority order), 0..7 = imask

struct WideRay {     // per-ray constants of the traversal
  V3 o;
  float idx, idy, idz;
  uint32_t octinv;   // 7 ^ octant of the direction
  bool valid;        // false for a ray with a NaN component: test_tri accepts nothing for it, so there is nothing to traverse
};

__device__ __forceinline__ WideRay wide_ray(V3 o, V3 d) {
  WideRay r;
  r.o = o;
  // |1/d| is clamped to 1e18: with an infinite reciprocal the plane distances q * (s/d) + (p-o)/d would be inf - inf = NaN,
  // the axis would drop out of the slab test (still conservative) and an axis-parallel ray would sweep the whole tree.
  // A finite bound keeps the signs right: both planes on one side of the origin -> miss, origin between them -> all t
  // (the planes are at least the padding away from any accepted hit, and padding * 1e18 exceeds every t of a scene);
  // 1e18 leaves room for the factor 2^15 s of wide_node below FLT_MAX.
  r.idx = fminf(fmaxf(1.0f / d.x, -1e18f), 1e18f);
  r.idy = fminf(fmaxf(1.0f / d.y, -1e18f), 1e18f);
  r.idz = fminf(fmaxf(1.0f / d.z, -1e18f), 1e18f);
  const uint32_t oct = (d.x < 0.0f ? 1u : 0u) | (d.y < 0.0f ? 2u : 0u) | (d.z < 0.0f ? 4u : 0u);
  r.octinv = 7u ^ oct;
  r.valid = o.x == o.x && o.y == o.y && o.z == o.z && d.x == d.x && d.y == d.y && d.z == d.z;
  return r;
}

// byte i of an 8-byte plane array as the float f = 1 + q * 2^-15: the byte is dropped into bits 8..15 of the mantissa of
// 1.0f by one PRMT, and t = (p + q*s - o)/d is evaluated as f * (2^15 s/d) + ((p - o)/d - 2^15 s/d) by one FFMA.  (Going
// through 8388608 + q and subtracting costs an FADD per plane -- 48 per node -- and the integer conversion runs on the
// slow pipe.)  `one` is 0x3F800000 in a register the compiler cannot see through (wide_one): with a literal there, ptxas
// puts the literal in PRMT's third operand and re-materialises the selector in a register before every PRMT.
// Rounding: b' = b - a' loses at most 2^-24 |a'| = 2^-9 quantisation steps; a step is at most 2/255 of the node's extent,
// so the planes move by less than 2^-16 of the scene's extent, half of the padding the boxes were built with
// (crt_bvh.cpp: pad = extent / 32768), which leaves the culling conservative (DESIGN.md "Conservative culling").
__device__ __forceinline__ uint32_t wide_one(uint32_t runtime_below_2_31) { return 0x3F800000u | (runtime_below_2_31 >> 31); }

__device__ __forceinline__ float wide_q(uint32_t lo4, uint32_t hi4, int i, uint32_t one) {
  const uint32_t w = i < 4 ? lo4 : hi4;
  const uint32_t sel = 0x7604u | ((uint32_t)(i & 3) << 4);   // result bytes (3..0) = (0x3F, 0x80, w.byte[i & 3], 0x00)
  return __uint_as_float(__byte_perm(w, one, sel));
}

// Visit one node: slab tests of its 8 children against [t_min, t_max]; returns the node group G (inner children hit, in
// near-to-far priority bits, plus imask) and the triangle group T (one bit per triangle of the leaf children hit).
// The reciprocal direction is finite (wide_ray), so its sign says which of a child's two planes per axis is the near one.
template <bool STATS>
__device__ __forceinline__ void wide_node_eval(const uint4 w0, const uint4 w1, const uint4 w2, const uint4 w3, const uint4 w4, const WideRay& r,
                                               uint32_t one, float t_min, float t_max, uint2& G, uint2& T, Stats& st) {
  if (STATS) st.node_visits++;
  const uint32_t imask = w0.w >> 24;
  // t = ((p + q * s) - o) / d  evaluated as  (1 + q 2^-15) * (2^15 s / d) + ((p - o) / d - 2^15 s / d); the node stores the
  // exponent of 2^15 s
  const float ax = __uint_as_float((w0.w & 0xFFu) << 23) * r.idx;
  const float ay = __uint_as_float(((w0.w >> 8) & 0xFFu) << 23) * r.idy;
  const float az = __uint_as_float(((w0.w >> 16) & 0xFFu) << 23) * r.idz;
  const float bx = (__uint_as_float(w0.x) - r.o.x) * r.idx - ax;
  const float by = (__uint_as_float(w0.y) - r.o.y) * r.idy - ay;
  const float bz = (__uint_as_float(w0.z) - r.o.z) * r.idz - az;
  // near / far plane arrays per axis (qlo: w2.xy, w2.zw, w3.xy; qhi: w3.zw, w4.xy, w4.zw)
  const bool nx = r.idx < 0.0f, ny = r.idy < 0.0f, nz = r.idz < 0.0f;
  const uint32_t xn0 = nx ? w3.z : w2.x, xn1 = nx ? w3.w : w2.y, xf0 = nx ? w2.x : w3.z, xf1 = nx ? w2.y : w3.w;
  const uint32_t yn0 = ny ? w4.x : w2.z, yn1 = ny ? w4.y : w2.w, yf0 = ny ? w2.z : w4.x, yf1 = ny ? w2.w : w4.y;
  const uint32_t zn0 = nz ? w4.z : w3.x, zn1 = nz ? w4.w : w3.y, zf0 = nz ? w3.x : w4.z, zf1 = nz ? w3.y : w4.w;
  uint32_t hitmask = 0;
#pragma unroll
  for (int half = 0; half < 2; half++) {
    // four meta bytes at a time: bit position and bit pattern each child contributes to the hit mask.  An inner child
    // (meta = 001 11sss) goes to bit 24 + (slot ^ octinv): near children get the high bits; a leaf puts its unary
    // triangle count at its triangle offset; an empty child contributes no bits.
    const uint32_t meta4 = half ? w1.w : w1.z;
    const uint32_t is_inner4 = (meta4 & (meta4 << 1)) & 0x10101010u;
    const uint32_t inner_mask4 = (is_inner4 >> 4) * 0xFFu;                    // 0xFF in the bytes of inner children
    const uint32_t bit_index4 = (meta4 ^ ((r.octinv * 0x01010101u) & inner_mask4)) & 0x1F1F1F1Fu;
    const uint32_t child_bits4 = (meta4 >> 5) & 0x07070707u;
#pragma unroll
    for (int j = 0; j < 4; j++) {
      const int i = half * 4 + j;
      const float tnx = fmaf(wide_q(xn0, xn1, i, one), ax, bx), tfx = fmaf(wide_q(xf0, xf1, i, one), ax, bx);
      const float tny = fmaf(wide_q(yn0, yn1, i, one), ay, by), tfy = fmaf(wide_q(yf0, yf1, i, one), ay, by);
      const float tnz = fmaf(wide_q(zn0, zn1, i, one), az, bz), tfz = fmaf(wide_q(zf0, zf1, i, one), az, bz);
      const float tn = fmaxf(fmaxf(tnx, tny), fmaxf(tnz, t_min));
      const float tf = fminf(fminf(tfx, tfy), fminf(tfz, t_max));
      if (tn <= tf) hitmask |= ((child_bits4 >> (8 * j)) & 0xFFu) << ((bit_index4 >> (8 * j)) & 0xFFu);
    }
  }
  G = make_uint2(w1.x, (hitmask & CRT_WIDE_HITS) | imask);
  T = make_uint2(w1.y, hitmask & ~CRT_WIDE_HITS);
}

template <bool STATS>
__device__ __forceinline__ void wide_node(const uint4* __restrict__ nodes, uint32_t node, const WideRay& r, uint32_t one, float t_min,
                                          float t_max, uint2& G, uint2& T, Stats& st) {
  const uint4 w0 = __ldg(&nodes[CRT_NODE_STRIDE * node + 0]);
  const uint4 w1 = __ldg(&nodes[CRT_NODE_STRIDE * node + 1]);
  const uint4 w2 = __ldg(&nodes[CRT_NODE_STRIDE * node + 2]);
  const uint4 w3 = __ldg(&nodes[CRT_NODE_STRIDE * node + 3]);
  const uint4 w4 = __ldg(&nodes[CRT_NODE_STRIDE * node + 4]);
  wide_node_eval<STATS>(w0, w1, w2, w3, w4, r, one, t_min, t_max, G, T, st);
}

// Take the nearest unvisited inner child out of a node group; returns its node index.
__device__ __forceinline__ uint32_t wide_pop_child(uint2& G, uint32_t octinv) {
  const uint32_t b = 31u - (uint32_t)__clz(G.y);
  G.y &= ~(1u << b);
  const uint32_t slot = (b - 24u) ^ octinv;
  return G.x + __popc(G.y & 0xFFu & ((1u << slot) - 1u));
}

template <bool STATS>
__device__ __forceinline__ bool wide_tri_eval(const float4 T0, const float4 T1, const float4 T2, uint32_t slot, V3 o, V3 d, uint32_t exclude,
                                              float t_min, bool anyhit, Hit& h, Stats& st);

// Test one triangle of a triangle group under the tie rule.  Returns true when an any-hit query has found its occluder.
template <bool STATS>
__device__ __forceinline__ bool wide_tri(const float4* __restrict__ tris, uint2& T, V3 o, V3 d, uint32_t exclude, float t_min, bool anyhit,
                                         Hit& h, Stats& st) {
  const uint32_t k = (uint32_t)__ffs(T.y) - 1u;
  T.y &= T.y - 1u;
  const uint32_t slot = T.x + k;
  const float4 T0 = __ldg(&tris[3 * slot + 0]);
  const float4 T1 = __ldg(&tris[3 * slot + 1]);
  const float4 T2 = __ldg(&tris[3 * slot + 2]);
  return wide_tri_eval<STATS>(T0, T1, T2, slot, o, d, exclude, t_min, anyhit, h, st);
}

// The test of wide_tri on a triangle that has been fetched already (slot = its position in the leaf-ordered array).
template <bool STATS>
__device__ __forceinline__ bool wide_tri_eval(const float4 T0, const float4 T1, const float4 T2, uint32_t slot, V3 o, V3 d, uint32_t exclude,
                                              float t_min, bool anyhit, Hit& h, Stats& st) {
  if (STATS) st.tri_tests++;
  const uint32_t index = __float_as_uint(T0.w);
  if (index == exclude) return false;
  float t;
  if (test_tri(T0, T1, T2, o, d, t_min, h.t, t)) {
    if (t < h.t || h.idx == CRT_MAX_U32 || index > h.idx) {
      h.t = t; h.idx = index; h.ref = CRT_TRI_REF | slot;
      return anyhit;
    }
  }
  return false;
}

template <bool ANYHIT, bool STATS>
__device__ __forceinline__ void traverse(const DevScene& sc, V3 o, V3 d, uint32_t exclude, float t_min, Hit& h, Stats& st) {
  if (sc.n_tri == 0) return;
  const WideRay r = wide_ray(o, d);
  if (!r.valid) return;
  uint2 stack[CRT_BVH_STACK];
  int sp = 0;
  uint2 G = make_uint2(0u, 0x80000000u);   // the root as the only child of a virtual group
  const uint32_t one = wide_one(sc.n_tri);
  for (;;) {
    uint2 T = make_uint2(0u, 0u);
    if (G.y & CRT_WIDE_HITS) {
      const uint32_t node = wide_pop_child(G, r.octinv);
      if (G.y & CRT_WIDE_HITS) stack[sp++] = G;
      wide_node<STATS>(sc.nodes, node, r, one, t_min, h.t, G, T, st);
    }
    while (T.y)
      if (wide_tri<STATS>(sc.tri, T, o, d, exclude, t_min, ANYHIT, h, st)) return;
    if (!(G.y & CRT_WIDE_HITS)) {
      if (sp == 0) return;
      G = stack[--sp];
    }
  }
}

// The patches and spheres, in index order with the reference's sequential semantics (CS:508-515).
template <bool STATS>
__device__ __forceinline__ Hit scan_linear(const DevScene& sc, V3 o, V3 d, uint32_t exclude, Stats& st) {
  Hit h; h.t = CRT_INFINITY_F; h.idx = CRT_MAX_U32; h.ref = 0;
  const uint32_t n_c = sc.n_lin < CRT_LIN_PARAM_MAX ? sc.n_lin : CRT_LIN_PARAM_MAX;
  for (uint32_t i = 0; i < n_c; i++) test_lin(sc.lin_c[i], i, o, d, exclude, sc.t_min, h);
  for (uint32_t i = CRT_LIN_PARAM_MAX; i < sc.n_lin; i++) test_lin(sc.lin_g[i], i, o, d, exclude, sc.t_min, h);
  if (STATS) st.prim_tests += sc.n_lin;
  return h;
}

#ifndef CRT_FAST_ROOT_RCP
#define CRT_FAST_ROOT_RCP 1
#endif
// Can the ray reach any triangle at all within [t_min, best]?  Same conservative slab test as a node's.
__device__ __forceinline__ bool hits_root_box(const DevScene& sc, V3 o, V3 d, float best) {
  if (sc.n_tri == 0) return false;
#if CRT_FAST_ROOT_RCP
  // approximate reciprocals (1 ulp): the bounds are padded by 2^-15 of the scene's extent, the test only has to be conservative
  float idx, idy, idz;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(idx) : "f"(d.x));
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(idy) : "f"(d.y));
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(idz) : "f"(d.z));
#else
  const float idx = 1.0f / d.x, idy = 1.0f / d.y, idz = 1.0f / d.z;
#endif
  float a0 = (sc.bvh_lo[0] - o.x) * idx, a1 = (sc.bvh_hi[0] - o.x) * idx;
  float b0 = (sc.bvh_lo[1] - o.y) * idy, b1 = (sc.bvh_hi[1] - o.y) * idy;
  float c0 = (sc.bvh_lo[2] - o.z) * idz, c1 = (sc.bvh_hi[2] - o.z) * idz;
  float tn = fmaxf(fmaxf(fminf(a0, a1), fminf(b0, b1)), fmaxf(fminf(c0, c1), sc.t_min));
  float tf = fminf(fminf(fmaxf(a0, a1), fmaxf(b0, b1)), fminf(fmaxf(c0, c1), best));
  return tn <= tf;
}

// intersect(), CS:503-518: closest accepted t over all primitives, ties to the highest index.
template <bool STATS>
__device__ __forceinline__ Hit intersect(const DevScene& sc, V3 o, V3 d, uint32_t exclude, Stats& st) {
  Hit h = scan_linear<STATS>(sc, o, d, exclude, st);
  traverse<false, STATS>(sc, o, d, exclude, sc.t_min, h, st);
  return h;
}

// shadow_intersect(), CS:697-705: the reference runs the full closest-hit scan and then keeps the hit
// only if the winner is `include`.  The outcome (winner == include or not) is all the caller uses, so
// once the linear list has produced the light at distance t, any triangle accepted at <= t (under the
// tie rule) settles it: an any-hit traversal bounded by t gives the identical boolean.
template <bool STATS>
__device__ __forceinline__ Hit shadow_intersect(const DevScene& sc, V3 o, V3 d, uint32_t include, uint32_t exclude, Stats& st) {
  Hit h = scan_linear<STATS>(sc, o, d, exclude, st);
  if (sc.shadow_closest) {
    traverse<false, STATS>(sc, o, d, exclude, sc.t_min, h, st);
  } else if (h.idx == include) {
    traverse<true, STATS>(sc, o, d, exclude, sc.t_min, h, st);
  }
  return h;
}

// The fields of ShapeIntersection (CS:54-60) for the winner, rebuilt after the scan.
struct Surface {
  V3 position, normal;
  uint32_t meta;
};

__device__ __forceinline__ Surface surface_of(const DevScene& sc, const Hit& h, V3 o, V3 d) {
  Surface s;
  s.position = madd(o, h.t, d);                       // ray_at, CS:304-307
  if (h.ref & CRT_TRI_REF) {
    const uint32_t slot = h.ref & ~CRT_TRI_REF;
    const float4 T1 = __ldg(&sc.tri[3 * slot + 1]);
    const float4 T2 = __ldg(&sc.tri[3 * slot + 2]);
    V3 n = normalize(cross(v3(T1), v3(T2)));
    if (dot(n, d) > 0.0f) n = -n;
    s.normal = n;
    s.meta = __float_as_uint(T1.w);
  } else {
    const LinPrim& p = (h.ref < CRT_LIN_PARAM_MAX) ? sc.lin_c[h.ref] : sc.lin_g[h.ref];
    s.meta = __float_as_uint(p.B.w);
    if (CRT_META_CAT(s.meta) == 0u) {
      V3 n = v3(p.A);
      if (dot(n, d) > 0.0f) n = -n;                   // CS:541-544
      s.normal = n;
    } else {
      s.normal = normalize(s.position - v3(p.A));     // CS:618 (outward, never flipped)
    }
  }
  return s;
}

// ------------------------------------------------------------------------------------------------ shading helpers
__device__ __forceinline__ void wavelengths(uint32_t l0, uint32_t lam[4]) {   // CS:321
  lam[0] = l0; lam[1] = (l0 + 4u) % CRT_LAMBDA_RANGE; lam[2] = (l0 + 8u) % CRT_LAMBDA_RANGE; lam[3] = (l0 + 12u) % CRT_LAMBDA_RANGE;
}

__device__ __forceinline__ float4 sample_spectrum(const float* __restrict__ spectra, uint32_t n_spectra, uint32_t index, const uint32_t lam[4]) {
  if (index >= n_spectra) index = n_spectra - 1;      // out-of-bounds pin: clamp
  const float* s = spectra + index * CRT_LAMBDA_RANGE;
  return make_float4(s[lam[0]], s[lam[1]], s[lam[2]], s[lam[3]]);
}

__device__ __forceinline__ float power_heuristic(float f_pdf, float g_pdf) {  // CS:297-302, nf = ng = 1
  float f = 1.0f * f_pdf, g = 1.0f * g_pdf;
  return (f * f) / (f * f + g * g);
}

__device__ __forceinline__ const LightRec& light_at(const DevScene& sc, uint32_t li) {
  if (li >= sc.n_lights) li = sc.n_lights - 1;         // CS:362 reads out of bounds in the default scene: clamp pin
  return (li < CRT_LIGHT_PARAM_MAX) ? sc.lights_c[li] : sc.lights_g[li];
}

// compute_light_pdf, CS:357-377
__device__ __forceinline__ float compute_light_pdf(const DevScene& sc, uint32_t emission_index, V3 normal, V3 position, V3 ray_origin, V3 ray_direction) {
  const LightRec& light = light_at(sc, emission_index);
  float light_area_pdf = light.d3.w;
  float abs_cos_theta = fmaxf(0.00001f, fabsf(dot(normal, -ray_direction)));
  float distance = length(position - ray_origin);
  float distance_squared = distance * distance;
  float geometric_term = abs_cos_theta / distance_squared;
  float light_solid_angle_pdf = light_area_pdf / geometric_term;
  float light_selection_pdf = 1.0f / (float)sc.n_lights;
  return light_selection_pdf * light_solid_angle_pdf;
}

// cosine_weighted_sample_hemisphere, CS:751-774 (u, v drawn by the caller in that order)
__device__ __forceinline__ V3 cosine_hemisphere(V3 normal, float u, float v, float& pdf) {
  float r = sqrtf(u);
  float theta = (2.0f * CRT_PI_F) * v;
  float sn, cs;
  crt_sincosf(theta, &sn, &cs);
  float x = r * cs, y = r * sn;
  float z = sqrtf(fmaxf(0.0f, 1.0f - u));
  V3 up = (fabsf(normal.z) < 0.999f) ? v3(0.0f, 0.0f, 1.0f) : v3(1.0f, 0.0f, 0.0f);
  V3 tangent = normalize(cross(up, normal));
  V3 bitangent = cross(normal, tangent);
  pdf = z / CRT_PI_F;
  return v3(fmaf(normal.x, z, fmaf(bitangent.x, y, tangent.x * x)),
            fmaf(normal.y, z, fmaf(bitangent.y, y, tangent.y * x)),
            fmaf(normal.z, z, fmaf(bitangent.z, y, tangent.z * x)));
}

// fresnel_s, CS:814-837
__device__ __forceinline__ float fresnel_s(V3 ray_dir, V3 normal, float eta1, float eta2) {
  float cosi = fminf(fmaxf(dot(ray_dir, normal), -1.0f), 1.0f);
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

__device__ __forceinline__ V3 reflect(V3 e1, V3 e2) {   // WGSL reflect
  float k = 2.0f * dot(e2, e1);
  return madd(e1, -k, e2);
}
__device__ __forceinline__ V3 refract(V3 e1, V3 e2, float e3) {   // WGSL refract
  float dd = dot(e2, e1);
  float k = 1.0f - (e3 * e3) * (1.0f - dd * dd);
  if (k < 0.0f) return v3(0.0f, 0.0f, 0.0f);
  float s = fmaf(e3, dd, sqrtf(k));
  return madd(e1 * e3, -s, e2);
}

// spectral_to_xyz, CS:411-426
__device__ __forceinline__ V3 spectral_to_xyz(const float* __restrict__ cie, float4 L, const uint32_t lam[4]) {
  float o[3];
#pragma unroll
  for (int k = 0; k < 3; k++) {
    const float* c = cie + 471 * k + 40;
    float s = fmaf(c[lam[3]], L.w, fmaf(c[lam[2]], L.z, fmaf(c[lam[1]], L.y, c[lam[0]] * L.x)));
    o[k] = crt_div(s * 300.0f, 427.42758f);
  }
  return v3(o[0], o[1], o[2]);
}

// epilogue of main, CS:109-115: average, XYZ -> linear sRGB, tone map, gamma (with the reference's green
// quirk, CS:453), rgba8unorm store
__device__ __forceinline__ uint32_t unorm8(float c) {
  if (c != c) return 0u;
  c = fminf(fmaxf(c, 0.0f), 1.0f);
  return (uint32_t)floorf(c * 255.0f + 0.5f);
}
__device__ __forceinline__ uint32_t resolve_pixel(float X, float Y, float Z, uint32_t sample, float exposure) {
  float fs = (float)sample;
  X = X / fs; Y = Y / fs; Z = Z / fs;
  float r = fmaf(-0.4985314f, Z, fmaf(-1.5371385f, Y, 3.2404542f * X));
  float g = fmaf(0.0415560f, Z, fmaf(1.8760108f, Y, -0.9692660f * X));
  float b = fmaf(1.0572252f, Z, fmaf(-0.2040259f, Y, 0.0556434f * X));
  r = 1.0f - crt_expf(-r * exposure); g = 1.0f - crt_expf(-g * exposure); b = 1.0f - crt_expf(-b * exposure);
  r = r < 0.0031308f ? r * 12.92f : 1.055f * crt_powf(r, 0.41666666f) - 0.055f;
  g = g < 0.0031308f ? g * (12.92f * g) : 1.055f * crt_powf(g, 0.41666666f) - 0.055f;
  b = b < 0.0031308f ? 12.92f * b : 1.055f * crt_powf(b, 0.41666666f) - 0.055f;
  return unorm8(r) | (unorm8(g) << 8) | (unorm8(b) << 16) | 0xFF000000u;
}
