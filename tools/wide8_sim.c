/* wide8_sim.c -- CPU walk of the Wide8 tree in the order wf_traverse uses (csrc/crt_device.cuh: nearest unvisited child of
 * the current node group by ray octant, a node's leaf triangles before its inner children, closest-hit shrinking t_max,
 * any-hit for shadow rays), counting nodes fetched and triangles tested per ray.  A planning tool for the builder
 * (tools/bvh_eval.py): tree shape changes speed only, so candidate builders are compared here by visit counts before
 * they get GPU time.  Arithmetic is plain float, not the kernels' bit-exact contract -- counts, not results.
 *
 *   gcc -O2 -shared -fPIC -o wide8_sim.so tools/wide8_sim.c -lm
 */
#include <math.h>
#include <stdint.h>
#include <string.h>

#define WORDS 20
#define HITS 0xFF000000u

typedef struct {
  uint32_t x, y;
} u2;

static inline float qf(const uint32_t* w, int word, int i) { return (float)((w[word + (i >> 2)] >> (8 * (i & 3))) & 0xFFu); }

/* rays: n x 8 floats (o.xyz, tmax, d.xyz, any-hit flag); out_t, out_slot: closest hit (slot = -1: none);
 * counts: n x 2 uint32 (nodes, triangles).  tris: 9 floats per leaf-order slot (v0, e1, e2). */
static void trace_sorted(const uint32_t* nodes, uint32_t n_nodes, const float* tris, const float* rays, uint32_t n, float tmin, float* out_t,
                         int32_t* out_slot, uint32_t* counts, int by_octant);

/* order = 0: the kernels' order (children by ray octant); 1: children by their actual entry distance, with the
 * subtrees that can no longer hold a closer hit skipped on the way back -- the bound any child ordering could reach;
 * 2: the kernels' octant order, but with that skip on the way back (what remembering each child's entry distance buys);
 * 3: the kernels' order and stack of node groups, each group remembering the smallest entry distance among the children it
 *    still holds: a popped group whose nearest remaining child lies beyond the current hit is dropped whole */
void wide8_trace(const uint32_t* nodes, uint32_t n_nodes, const float* tris, const float* rays, uint32_t n, float tmin, float* out_t,
                 int32_t* out_slot, uint32_t* counts, int order) {
  if (order == 1 || order == 2) { trace_sorted(nodes, n_nodes, tris, rays, n, tmin, out_t, out_slot, counts, order == 2); return; }
  for (uint32_t r = 0; r < n; r++) {
    const float* R = rays + 8 * (size_t)r;
    const float o[3] = {R[0], R[1], R[2]}, d[3] = {R[4], R[5], R[6]};
    float tmax = R[3];
    const int anyhit = R[7] != 0.0f;
    float id[3];
    for (int a = 0; a < 3; a++) {
      float v = 1.0f / d[a];
      if (v > 1e18f) v = 1e18f;
      if (v < -1e18f) v = -1e18f;
      id[a] = v;
    }
    const uint32_t oct = (d[0] < 0 ? 1u : 0u) | (d[1] < 0 ? 2u : 0u) | (d[2] < 0 ? 4u : 0u), octinv = 7u ^ oct;
    uint32_t nn = 0, nt = 0;
    int32_t best = -1;
    u2 stack[64];
    float stack_t[64];
    float G_t = tmin;   /* smallest entry distance among the children the current group still holds (order 3) */
    int sp = 0;
    u2 G = {0u, n_nodes ? 0x80000000u : 0u};   /* the root as the only child of a virtual group */
    int done = 0;
    while (!done) {
      if (!(G.y & HITS)) {
        if (sp == 0) break;
        G = stack[--sp]; G_t = stack_t[sp];
        if (order == 3 && G_t > tmax) { G.y = 0; continue; }
      }
      const uint32_t b = 31u - (uint32_t)__builtin_clz(G.y);   /* nearest unvisited inner child of the group */
      G.y &= ~(1u << b);
      const uint32_t slot = (b - 24u) ^ octinv;
      const uint32_t node = G.x + (uint32_t)__builtin_popcount(G.y & 0xFFu & ((1u << slot) - 1u));
      if (G.y & HITS) { stack[sp] = G; stack_t[sp++] = G_t; }
      const uint32_t* w = nodes + (size_t)node * WORDS;
      nn++;
      float p[3], st[3];
      memcpy(p, w, 12);
      for (int a = 0; a < 3; a++) st[a] = ldexpf(1.0f, (int)((w[3] >> (8 * a)) & 0xFFu) - 127 - 15);
      const uint32_t imask = w[3] >> 24;
      uint32_t hit_inner = 0, tri_mask = 0;
      float child_t[8];
      for (int i = 0; i < 8; i++) {
        const uint32_t meta = (w[6 + (i >> 2)] >> (8 * (i & 3))) & 0xFFu;
        if (!meta) continue;
        float tn = tmin, tf = tmax;
        for (int a = 0; a < 3; a++) {
          const float lo = p[a] + qf(w, 8 + 2 * a, i) * st[a], hi = p[a] + qf(w, 14 + 2 * a, i) * st[a];
          float t0 = (lo - o[a]) * id[a], t1 = (hi - o[a]) * id[a];
          if (t0 > t1) { const float x = t0; t0 = t1; t1 = x; }
          if (t0 > tn) tn = t0;
          if (t1 < tf) tf = t1;
        }
        if (!(tn <= tf)) continue;
        if ((meta >> 5) == 1u && (meta & 31u) >= 24u) { hit_inner |= 1u << (24u + ((uint32_t)i ^ octinv)); child_t[(uint32_t)i ^ octinv] = tn; }
        else tri_mask |= (meta >> 5) << (meta & 31u);
      }
      G.x = w[4];
      G.y = hit_inner | imask;
      G_t = INFINITY;   /* the nearest-by-octant child is taken next without a test: the minimum is over the others */
      if (hit_inner) {
        const uint32_t first = 31u - (uint32_t)__builtin_clz(hit_inner);
        for (uint32_t k = 24; k < 32; k++) if (((hit_inner >> k) & 1u) && k != first && child_t[k - 24] < G_t) G_t = child_t[k - 24];
      }
      while (tri_mask) {   /* the node's triangles first */
        const uint32_t k = (uint32_t)__builtin_ctz(tri_mask);
        tri_mask &= tri_mask - 1u;
        const uint32_t s = w[5] + k;
        const float* T = tris + 9 * (size_t)s;
        nt++;
        const float* e1 = T + 3; const float* e2 = T + 6;
        const float pv[3] = {d[1] * e2[2] - d[2] * e2[1], d[2] * e2[0] - d[0] * e2[2], d[0] * e2[1] - d[1] * e2[0]};
        const float det = e1[0] * pv[0] + e1[1] * pv[1] + e1[2] * pv[2];
        if (!(fabsf(det) >= 1e-8f)) continue;
        const float inv = 1.0f / det;
        const float tv[3] = {o[0] - T[0], o[1] - T[1], o[2] - T[2]};
        const float u = (tv[0] * pv[0] + tv[1] * pv[1] + tv[2] * pv[2]) * inv;
        if (!(u >= 0.0f && u <= 1.0f)) continue;
        const float qv[3] = {tv[1] * e1[2] - tv[2] * e1[1], tv[2] * e1[0] - tv[0] * e1[2], tv[0] * e1[1] - tv[1] * e1[0]};
        const float v = (d[0] * qv[0] + d[1] * qv[1] + d[2] * qv[2]) * inv;
        if (!(v >= 0.0f && u + v <= 1.0f)) continue;
        const float t = (e2[0] * qv[0] + e2[1] * qv[1] + e2[2] * qv[2]) * inv;
        if (!(t >= tmin && t <= tmax)) continue;
        tmax = t; best = (int32_t)s;
        if (anyhit) { done = 1; break; }
      }
    }
    out_t[r] = tmax; out_slot[r] = best;
    counts[2 * (size_t)r] = nn; counts[2 * (size_t)r + 1] = nt;
  }
}

static int tri_test(const float* T, const float* o, const float* d, float tmin, float tmax, float* t_out) {
  const float* e1 = T + 3; const float* e2 = T + 6;
  const float pv[3] = {d[1] * e2[2] - d[2] * e2[1], d[2] * e2[0] - d[0] * e2[2], d[0] * e2[1] - d[1] * e2[0]};
  const float det = e1[0] * pv[0] + e1[1] * pv[1] + e1[2] * pv[2];
  if (!(fabsf(det) >= 1e-8f)) return 0;
  const float inv = 1.0f / det;
  const float tv[3] = {o[0] - T[0], o[1] - T[1], o[2] - T[2]};
  const float u = (tv[0] * pv[0] + tv[1] * pv[1] + tv[2] * pv[2]) * inv;
  if (!(u >= 0.0f && u <= 1.0f)) return 0;
  const float qv[3] = {tv[1] * e1[2] - tv[2] * e1[1], tv[2] * e1[0] - tv[0] * e1[2], tv[0] * e1[1] - tv[1] * e1[0]};
  const float v = (d[0] * qv[0] + d[1] * qv[1] + d[2] * qv[2]) * inv;
  if (!(v >= 0.0f && u + v <= 1.0f)) return 0;
  const float t = (e2[0] * qv[0] + e2[1] * qv[1] + e2[2] * qv[2]) * inv;
  if (!(t >= tmin && t <= tmax)) return 0;
  *t_out = t;
  return 1;
}

static void trace_sorted(const uint32_t* nodes, uint32_t n_nodes, const float* tris, const float* rays, uint32_t n, float tmin, float* out_t,
                         int32_t* out_slot, uint32_t* counts, int by_octant) {
  for (uint32_t r = 0; r < n; r++) {
    const float* R = rays + 8 * (size_t)r;
    const float o[3] = {R[0], R[1], R[2]}, d[3] = {R[4], R[5], R[6]};
    float tmax = R[3];
    const int anyhit = R[7] != 0.0f;
    const uint32_t octinv = 7u ^ ((d[0] < 0 ? 1u : 0u) | (d[1] < 0 ? 2u : 0u) | (d[2] < 0 ? 4u : 0u));
    float id[3];
    for (int a = 0; a < 3; a++) { float v = 1.0f / d[a]; if (v > 1e18f) v = 1e18f; if (v < -1e18f) v = -1e18f; id[a] = v; }
    uint32_t nn = 0, nt = 0;
    int32_t best = -1;
    struct { uint32_t node; float tn, key; } stack[512];
    int sp = 0, done = 0;
    if (n_nodes) { stack[0].node = 0; stack[0].tn = tmin; sp = 1; }
    while (sp && !done) {
      sp--;
      if (stack[sp].tn > tmax) continue;   /* a closer hit was found since this child was pushed */
      const uint32_t* w = nodes + (size_t)stack[sp].node * WORDS;
      nn++;
      float p[3], st[3];
      memcpy(p, w, 12);
      for (int a = 0; a < 3; a++) st[a] = ldexpf(1.0f, (int)((w[3] >> (8 * a)) & 0xFFu) - 127 - 15);
      const uint32_t imask = w[3] >> 24;
      uint32_t tri_mask = 0;
      const int base = sp;
      for (int i = 0; i < 8; i++) {
        const uint32_t meta = (w[6 + (i >> 2)] >> (8 * (i & 3))) & 0xFFu;
        if (!meta) continue;
        float tn = tmin, tf = tmax;
        for (int a = 0; a < 3; a++) {
          const float lo = p[a] + qf(w, 8 + 2 * a, i) * st[a], hi = p[a] + qf(w, 14 + 2 * a, i) * st[a];
          float t0 = (lo - o[a]) * id[a], t1 = (hi - o[a]) * id[a];
          if (t0 > t1) { const float x = t0; t0 = t1; t1 = x; }
          if (t0 > tn) tn = t0;
          if (t1 < tf) tf = t1;
        }
        if (!(tn <= tf)) continue;
        if ((meta >> 5) == 1u && (meta & 31u) >= 24u) {
          int k = sp++;   /* insertion sort: farthest first, so the nearest is popped first */
          const float key = by_octant ? -(float)((uint32_t)i ^ octinv) : tn;   /* smaller key = visited earlier */
          while (k > base && stack[k - 1].key < key) { stack[k] = stack[k - 1]; k--; }
          stack[k].key = key;
          stack[k].node = w[4] + (uint32_t)__builtin_popcount(imask & ((1u << i) - 1u));
          stack[k].tn = tn;
        } else {
          tri_mask |= (meta >> 5) << (meta & 31u);
        }
      }
      while (tri_mask) {
        const uint32_t k = (uint32_t)__builtin_ctz(tri_mask);
        tri_mask &= tri_mask - 1u;
        const uint32_t s = w[5] + k;
        nt++;
        float t;
        if (tri_test(tris + 9 * (size_t)s, o, d, tmin, tmax, &t)) { tmax = t; best = (int32_t)s; if (anyhit) { done = 1; break; } }
      }
    }
    out_t[r] = tmax; out_slot[r] = best;
    counts[2 * (size_t)r] = nn; counts[2 * (size_t)r + 1] = nt;
  }
}

/* ---------------------------------------------------------------------------------------------------------------------
 * Warp-level model of wf_traverse's loop: 32 lanes, each owning one ray at a time, refilled from the queue when at least
 * `refill_min` lanes are idle (or none has a ray).  A trip is a triangle step (one triangle for every lane that has some
 * pending) followed by a node step (for the lanes whose triangle group is empty by then).  Counts warp-level steps and
 * the lanes active in them -- the kernel is issue-bound at about 265 instructions per node step and 80 per triangle
 * step, so steps x instructions per queued ray is what a scheduling variant has to lower.
 *   mode 0: the kernel's schedule.
 *   mode 1: a lane keeps stepping through nodes while it holds one postponed triangle group; the triangle step runs only
 *           when at least `tri_min` lanes hold triangles, or a lane holds a group and has just found a second one, or no
 *           lane has node work (postponed triangles are tested without the closer hit they might have provided, so some
 *           extra nodes are visited).
 * out[0..5] = node steps, lanes in node steps, triangle steps, lanes in triangle steps, nodes visited, triangles tested. */
typedef struct {
  int have, anyhit;
  float o[3], d[3], id[3], tmax;
  uint32_t octinv;
  u2 G, stack[64];
  int sp;
  uint32_t T_base, T_mask;      /* pending triangle group */
  uint32_t P_base, P_mask;      /* postponed triangle group being tested (mode 1) */
  uint32_t Q_base[8], Q_mask[8]; int nq;   /* further postponed groups (mode 1 with depth > 1) */
  int finished_nodes;           /* traversal has no node left; only triangles remain */
  uint32_t idx; int hit;        /* position in the queue; some triangle was accepted */
} Lane;

static void lane_start(Lane* L, const float* R, uint32_t n_nodes) {
  L->have = 1; L->anyhit = R[7] != 0.0f;
  for (int a = 0; a < 3; a++) {
    L->o[a] = R[a]; L->d[a] = R[4 + a];
    float v = 1.0f / L->d[a]; if (v > 1e18f) v = 1e18f; if (v < -1e18f) v = -1e18f; L->id[a] = v;
  }
  L->tmax = R[3];
  L->octinv = 7u ^ ((L->d[0] < 0 ? 1u : 0u) | (L->d[1] < 0 ? 2u : 0u) | (L->d[2] < 0 ? 4u : 0u));
  L->G.x = 0; L->G.y = n_nodes ? 0x80000000u : 0u;
  L->sp = 0; L->T_mask = L->P_mask = 0; L->finished_nodes = 0; L->nq = 0; L->hit = 0;
}

/* one node step; returns 0 when there was no node left */
static int lane_node(Lane* L, const uint32_t* nodes, float tmin) {
  if (!(L->G.y & HITS)) {
    if (L->sp == 0) { L->finished_nodes = 1; return 0; }
    L->G = L->stack[--L->sp];
  }
  const uint32_t b = 31u - (uint32_t)__builtin_clz(L->G.y);
  L->G.y &= ~(1u << b);
  const uint32_t slot = (b - 24u) ^ L->octinv;
  const uint32_t node = L->G.x + (uint32_t)__builtin_popcount(L->G.y & 0xFFu & ((1u << slot) - 1u));
  if (L->G.y & HITS) L->stack[L->sp++] = L->G;
  const uint32_t* w = nodes + (size_t)node * WORDS;
  float p[3], st[3];
  memcpy(p, w, 12);
  for (int a = 0; a < 3; a++) st[a] = ldexpf(1.0f, (int)((w[3] >> (8 * a)) & 0xFFu) - 127 - 15);
  uint32_t hit_inner = 0, tri_mask = 0;
  for (int i = 0; i < 8; i++) {
    const uint32_t meta = (w[6 + (i >> 2)] >> (8 * (i & 3))) & 0xFFu;
    if (!meta) continue;
    float tn = tmin, tf = L->tmax;
    for (int a = 0; a < 3; a++) {
      const float lo = p[a] + qf(w, 8 + 2 * a, i) * st[a], hi = p[a] + qf(w, 14 + 2 * a, i) * st[a];
      float t0 = (lo - L->o[a]) * L->id[a], t1 = (hi - L->o[a]) * L->id[a];
      if (t0 > t1) { const float x = t0; t0 = t1; t1 = x; }
      if (t0 > tn) tn = t0;
      if (t1 < tf) tf = t1;
    }
    if (!(tn <= tf)) continue;
    if ((meta >> 5) == 1u && (meta & 31u) >= 24u) hit_inner |= 1u << (24u + ((uint32_t)i ^ L->octinv));
    else tri_mask |= (meta >> 5) << (meta & 31u);
  }
  L->G.x = w[4]; L->G.y = hit_inner | (w[3] >> 24);
  L->T_base = w[5]; L->T_mask = tri_mask;
  return 1;
}

/* one triangle of the pending group (the postponed one first); returns 1 when an any-hit ray found its occluder */
static int lane_tri(Lane* L, const float* tris, float tmin) {
  uint32_t* mask = L->P_mask ? &L->P_mask : &L->T_mask;
  const uint32_t base = L->P_mask ? L->P_base : L->T_base;
  const uint32_t k = (uint32_t)__builtin_ctz(*mask);
  *mask &= *mask - 1u;
  float t;
  const int hit = tri_test(tris + 9 * (size_t)(base + k), L->o, L->d, tmin, L->tmax, &t);
  if (hit) { L->tmax = t; L->hit = 1; }
  if (!L->P_mask && L->nq) { L->nq--; L->P_base = L->Q_base[L->nq]; L->P_mask = L->Q_mask[L->nq]; }   /* next postponed group */
  return hit && L->anyhit;
}

void wide8_warp_sim(const uint32_t* nodes, uint32_t n_nodes, const float* tris, const float* rays, uint32_t n, float tmin, int mode,
                    int refill_min, int tri_min, int depth, uint64_t* out, float* res_t, int32_t* res_hit) {
#define LANE_DONE(L) do { (L)->have = 0; if (res_t) { res_t[(L)->idx] = (L)->tmax; res_hit[(L)->idx] = (L)->hit; } } while (0)
  /* depth: postponed groups a lane may hold (mode 1); it must stop walking when it holds `depth` and finds another */
  static Lane lane[32];
  for (int l = 0; l < 32; l++) lane[l].have = 0;
  for (int k = 0; k < 6; k++) out[k] = 0;
  uint32_t next = 0;
  for (;;) {
    int idle = 0, busy = 0;
    for (int l = 0; l < 32; l++) { if (lane[l].have) busy++; else idle++; }
    if (next < n && (idle >= refill_min || busy == 0))
      for (int l = 0; l < 32 && next < n; l++) if (!lane[l].have) { lane[l].idx = next; lane_start(&lane[l], rays + 8 * (size_t)next++, n_nodes); }
    busy = 0;
    for (int l = 0; l < 32; l++) busy += lane[l].have;
    if (!busy) break;
    /* ---- triangle step */
    int want = 0, must = 0, node_work = 0;
    for (int l = 0; l < 32; l++) {
      Lane* L = &lane[l];
      if (!L->have) continue;
      const int tri = (L->T_mask | L->P_mask) != 0;
      want += tri;
      if (mode == 0) continue;
      const int full = L->P_mask && L->T_mask && L->nq >= depth - 1;
      if (full || (tri && L->finished_nodes)) must = 1;
      if (!L->finished_nodes && !full) node_work++;
    }
    const int run_tri = mode == 0 ? want > 0 : (want > 0 && (want >= tri_min || must || node_work == 0));
    if (run_tri) {
      out[2]++;
      for (int l = 0; l < 32; l++) {
        Lane* L = &lane[l];
        if (!L->have || !(L->T_mask | L->P_mask)) continue;
        out[3]++; out[5]++;
        if (lane_tri(L, tris, tmin)) LANE_DONE(L);
      }
    }
    /* ---- node step */
    int stepped = 0;
    for (int l = 0; l < 32; l++) {
      Lane* L = &lane[l];
      if (!L->have) continue;
      if (mode == 0) { if (L->T_mask) continue; }
      else {
        if (L->P_mask && L->T_mask && L->nq >= depth - 1) continue;   /* no room to postpone another group: wait for the triangle step */
        if (L->T_mask) {                                   /* postpone, keep walking */
          if (!L->P_mask) { L->P_base = L->T_base; L->P_mask = L->T_mask; }
          else { L->Q_base[L->nq] = L->T_base; L->Q_mask[L->nq] = L->T_mask; L->nq++; }
          L->T_mask = 0;
        }
        if (L->finished_nodes) { if (!L->P_mask) LANE_DONE(L); continue; }
      }
      if (lane_node(L, nodes, tmin)) { stepped++; out[4]++; }
      else if (!(L->T_mask | L->P_mask)) LANE_DONE(L);      /* no node left and no triangle pending: the ray is done */
    }
    if (stepped) { out[0]++; out[1] += (uint64_t)stepped; }
  }
}
