/* srt_oracle.c — TEST INFRASTRUCTURE, not product code.
 *
 * Plain-C CPU restatement of the reference's path-tracing path over the flattened scene of
 * include/srt.h.  Every function cites the reference file:line it follows; the arithmetic keeps the
 * reference's evaluation order (Eigen's dot = a0*b0 + (a1*b1 + a2*b2), Redux.h:101-115; cross,
 * OrthoMethods.h:47-49; vec3.h helpers) and must be compiled WITHOUT FMA contraction or fast-math
 * (oracle/Makefile: -O2 -ffp-contract=off).
 *
 * Parity status: PINNED — tests/test_oracle_vs_reference.py checks this file against the compiled
 * reference (oracle/_ref/libref_oracle.so, built from /root/reference) bit-for-bit on hit records,
 * box decisions, scatter results with injected random numbers, texture lookups, camera rays and the
 * resolve; and tests/golden/ holds vectors generated from the reference for boxes without it.
 *
 * The one thing that is NOT the reference's: random numbers.  The reference draws from one global
 * std::mt19937 in program order (globals.h:30-35) and samples disks/balls by rejection
 * (vec3.h:62-70,88-95); the product uses a per-(pixel, sample) Philox4x32-10 stream with direct
 * sampling of the same distributions (DESIGN.md "random numbers"), and so does so_render here, so
 * that a CUDA frame can be compared pixel by pixel.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this library.
 */
#include <float.h>
#include <math.h>
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "srt.h"

typedef struct { float x, y, z; } v3;

static inline v3 V(float x, float y, float z) { v3 r = {x, y, z}; return r; }
static inline v3 vadd(v3 a, v3 b) { return V(a.x + b.x, a.y + b.y, a.z + b.z); }
static inline v3 vsub(v3 a, v3 b) { return V(a.x - b.x, a.y - b.y, a.z - b.z); }
static inline v3 vscale(float s, v3 a) { return V(s * a.x, s * a.y, s * a.z); }
static inline v3 vneg(v3 a) { return V(-a.x, -a.y, -a.z); }
/* Eigen redux order for 3-vectors: a0*b0 + (a1*b1 + a2*b2) */
static inline float vdot(v3 a, v3 b) { return a.x * b.x + (a.y * b.y + a.z * b.z); }
/* OrthoMethods.h:47-49 */
static inline v3 vcross(v3 a, v3 b) { return V(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
/* vec3.h:29-31 */
static inline float length_squared(v3 v) { return v.x * v.x + v.y * v.y + v.z * v.z; }
/* vec3.h:33-35 */
static inline float vlength(v3 v) { return sqrtf(length_squared(v)); }
/* vec3.h:54-60 */
static inline v3 unit_vector(v3 v) {
  float len = vlength(v);
  if (len != 0) return V(v.x / len, v.y / len, v.z / len);
  return v;
}
/* globals.h:17-24 */
static inline float clampf(float x, float lo, float hi) { if (x < lo) return lo; if (x > hi) return hi; return x; }
static inline v3 ld3(const float* p) { return V(p[0], p[1], p[2]); }
static inline void st3(float* p, v3 v) { p[0] = v.x; p[1] = v.y; p[2] = v.z; }

static const float PI_F = 3.1415926535897932385f; /* globals.h:15 */

/* ------------------------------------------------------------------------------------------ scene */
typedef struct so_scene {
  int64_t n_tri, n_sph, n_nodes;
  float *v0, *v1, *v2, *uv0, *uv1, *uv2; int32_t* tri_mat;
  float *n0, *n1, *n2;   /* optional per-vertex normals (srt.h: tri_n0), NULL = the reference's flat shading */
  float *tg0, *tg1, *tg2; /* optional per-vertex tangents xyz + handedness (srt.h: tri_tan0), NULL = the per-face basis */
  float *c0, *c1, *t0, *t1, *r; int32_t* sph_mat;
  srt_bvh_node* nodes;
  int32_t n_mat, n_tex;
  srt_material* mats;
  srt_texture* texs; /* data pointers own copies */
} so_scene;

static void* dup_mem(const void* p, size_t n) { void* q = malloc(n ? n : 1); if (n) memcpy(q, p, n); return q; }

so_scene* so_scene_create(const srt_scene_desc* d) {
  so_scene* s = (so_scene*)calloc(1, sizeof *s);
  s->n_tri = d->n_tri; s->n_sph = d->n_sph; s->n_nodes = d->n_nodes; s->n_mat = d->n_mat; s->n_tex = d->n_tex;
  size_t t3 = (size_t)d->n_tri * 3 * 4, t2 = (size_t)d->n_tri * 2 * 4;
  s->v0 = dup_mem(d->tri_v0, t3); s->v1 = dup_mem(d->tri_v1, t3); s->v2 = dup_mem(d->tri_v2, t3);
  s->uv0 = dup_mem(d->tri_uv0, t2); s->uv1 = dup_mem(d->tri_uv1, t2); s->uv2 = dup_mem(d->tri_uv2, t2);
  s->tri_mat = dup_mem(d->tri_mat, (size_t)d->n_tri * 4);
  if (d->n_tri > 0 && d->tri_n0 && d->tri_n1 && d->tri_n2) { s->n0 = dup_mem(d->tri_n0, t3); s->n1 = dup_mem(d->tri_n1, t3); s->n2 = dup_mem(d->tri_n2, t3); }
  if (d->n_tri > 0 && d->tri_tan0 && d->tri_tan1 && d->tri_tan2) { size_t t4 = (size_t)d->n_tri * 4 * 4; s->tg0 = dup_mem(d->tri_tan0, t4); s->tg1 = dup_mem(d->tri_tan1, t4); s->tg2 = dup_mem(d->tri_tan2, t4); }
  size_t s3 = (size_t)d->n_sph * 3 * 4, s1 = (size_t)d->n_sph * 4;
  s->c0 = dup_mem(d->sph_c0, s3); s->c1 = dup_mem(d->sph_c1, s3);
  s->t0 = dup_mem(d->sph_t0, s1); s->t1 = dup_mem(d->sph_t1, s1); s->r = dup_mem(d->sph_r, s1);
  s->sph_mat = dup_mem(d->sph_mat, s1);
  s->nodes = dup_mem(d->nodes, (size_t)d->n_nodes * sizeof(srt_bvh_node));
  s->mats = dup_mem(d->materials, (size_t)d->n_mat * sizeof(srt_material));
  s->texs = dup_mem(d->textures, (size_t)d->n_tex * sizeof(srt_texture));
  for (int i = 0; i < s->n_tex; i++)
    if (s->texs[i].kind == SRT_TEX_IMAGE && s->texs[i].data) {
      /* +2 zero bytes: imagePNG::value reads px[1], px[2] after the last texel of a 1-bpp map
         (texture.h:145-147; out of bounds in the reference) — pinned to 0 here and on the GPU */
      size_t n = (size_t)s->texs[i].width * s->texs[i].height * s->texs[i].bpp;
      uint8_t* q = (uint8_t*)calloc(n + 2, 1); memcpy(q, s->texs[i].data, n); s->texs[i].data = q;
    }
  return s;
}

void so_scene_destroy(so_scene* s) {
  if (!s) return;
  free(s->n0); free(s->n1); free(s->n2);
  free(s->tg0); free(s->tg1); free(s->tg2);
  free(s->v0); free(s->v1); free(s->v2); free(s->uv0); free(s->uv1); free(s->uv2); free(s->tri_mat);
  free(s->c0); free(s->c1); free(s->t0); free(s->t1); free(s->r); free(s->sph_mat); free(s->nodes); free(s->mats);
  for (int i = 0; i < s->n_tex; i++) if (s->texs[i].kind == SRT_TEX_IMAGE && s->texs[i].data) free((void*)s->texs[i].data);
  free(s->texs); free(s);
}

/* ------------------------------------------------------------------------------------------ record */
typedef struct { v3 o, dir; float time; } ray_t;
typedef struct { v3 p, normal, tangent, bitangent; float u, v, t; int front_face, mat, prim; } rec_t;
typedef struct { uint64_t box, leaf; } counters_t;

/* ray.h:15-17 */
static inline v3 ray_at(const ray_t* r, float t) { return V(r->o.x + t * r->dir.x, r->o.y + t * r->dir.y, r->o.z + t * r->dir.z); }

/* hittable.h:18-21 */
static inline void set_face_normal(rec_t* rec, const ray_t* r, v3 outward) {
  rec->front_face = vdot(r->dir, outward) < 0;
  rec->normal = rec->front_face ? outward : vneg(outward);
}

/* aabb.h:11-27 (the unitVector at :12 is dead code) */
static int aabb_hit(const float* bmin, const float* bmax, const ray_t* r, float tMin, float tMax) {
  const float o[3] = {r->o.x, r->o.y, r->o.z}, d[3] = {r->dir.x, r->dir.y, r->dir.z};
  for (int axis = 0; axis < 3; axis++) {
    float a = (bmin[axis] - o[axis]) / d[axis];
    float b = (bmax[axis] - o[axis]) / d[axis];
    float t0 = fminf(a, b);
    float t1 = fmaxf(a, b);
    tMin = fmaxf(t0, tMin);
    tMax = fminf(t1, tMax);
    if (tMax <= tMin) return 0;
  }
  return 1;
}

/* sphere.h:47-52 */
static v3 sphere_center(const so_scene* s, int64_t j, float time) {
  v3 c0 = ld3(s->c0 + 3 * j), c1 = ld3(s->c1 + 3 * j);
  if (c0.x != c1.x || c0.y != c1.y || c0.z != c1.z) {
    float f = (time - s->t0[j]) / (s->t1[j] - s->t0[j]);
    return vadd(c0, vscale(f, vsub(c1, c0)));
  }
  return c0;
}

/* sphere.h:96-106; -vec3f::UnitZ() is (-0,-0,-1) */
static void sphere_tangent_basis(v3 n, v3* tangent, v3* bitangent) {
  v3 b;
  float ndoty = n.x * 0.0f + (n.y * 1.0f + n.z * 0.0f);
  if (1.0f - fabsf(ndoty) < FLT_EPSILON) b = V(-0.0f, -0.0f, -1.0f);
  else b = V(0.0f, 1.0f, 0.0f);
  *tangent = unit_vector(vcross(b, n));
  *bitangent = unit_vector(vcross(n, *tangent));
}

/* sphere.h:54-83 (+ getSphereUV :32-38) */
static int sphere_hit(const so_scene* s, int64_t j, const ray_t* r, float tMin, float tMax, rec_t* rec) {
  v3 ctr = sphere_center(s, j, r->time);
  float radius = s->r[j];
  v3 oc = vsub(r->o, ctr);
  float a = length_squared(r->dir);
  float halfB = vdot(oc, r->dir);
  float c = length_squared(oc) - radius * radius;
  float discriminant = halfB * halfB - a * c;
  if (discriminant < 0.0f) return 0;
  float sqrtd = sqrtf(discriminant);
  float root = (-halfB - sqrtd) / a;
  if (root < tMin || root > tMax) {
    root = (-halfB + sqrtd) / a;
    if (root < tMin || root > tMax) return 0;
  }
  rec->t = root;
  rec->p = ray_at(r, rec->t);
  v3 outward = unit_vector(vsub(rec->p, ctr));
  set_face_normal(rec, r, outward);
  float theta = acosf(-outward.y);
  float phi = atan2f(-outward.z, outward.x) + PI_F;
  rec->u = phi / (2.0f * PI_F);
  rec->v = theta / PI_F;
  rec->mat = s->sph_mat[j];
  sphere_tangent_basis(outward, &rec->tangent, &rec->bitangent);
  return 1;
}

/* model.h:214-235 */
static void triangle_tangent_basis(const so_scene* s, int64_t i, v3* tangent, v3* bitangent) {
  v3 p0 = ld3(s->v0 + 3 * i), p1 = ld3(s->v1 + 3 * i), p2 = ld3(s->v2 + 3 * i);
  v3 edge0 = vsub(p1, p0), edge1 = vsub(p2, p0);
  float du0x = s->uv1[2 * i] - s->uv0[2 * i], du0y = s->uv1[2 * i + 1] - s->uv0[2 * i + 1];
  float du1x = s->uv2[2 * i] - s->uv0[2 * i], du1y = s->uv2[2 * i + 1] - s->uv0[2 * i + 1];
  float f = (du0x * du1y - du1x * du0y);
  if (f == 0) f += FLT_EPSILON;
  f = 1.0f / f;
  v3 t, b;
  t.x = f * (du1y * edge0.x - du0y * edge1.x);
  t.y = f * (du1y * edge0.y - du0y * edge1.y);
  t.z = f * (du1y * edge0.z - du0y * edge1.z);
  *tangent = unit_vector(t);
  b.x = f * (-du1x * edge0.x + du0x * edge1.x);
  b.y = f * (-du1x * edge0.y + du0x * edge1.y);
  b.z = f * (-du1x * edge0.z + du0x * edge1.z);
  *bitangent = unit_vector(b);
}

/* model.h:104-181: plane + three inside-edge tests, back-face culled, NO tMax test */
static int triangle_hit(const so_scene* s, int64_t i, const ray_t* r, float tMin, float tMax, rec_t* rec) {
  (void)tMax;
  v3 sv0 = ld3(s->v0 + 3 * i), sv1 = ld3(s->v1 + 3 * i), sv2 = ld3(s->v2 + 3 * i);
  v3 normal = vcross(vsub(sv1, sv0), vsub(sv2, sv0));           /* getNormal, model.h:276-283 */
  float NdotDir = vdot(normal, r->dir);
  if (fabsf(NdotDir) < FLT_EPSILON) return 0;
  if (vdot(r->dir, normal) > 0) return 0;
  float d = -vdot(normal, sv0);
  float t = -(vdot(normal, r->o) + d) / NdotDir;
  if (t < tMin) return 0;
  v3 p = V(r->o.x + t * r->dir.x, r->o.y + t * r->dir.y, r->o.z + t * r->dir.z);
  if (vdot(normal, vcross(vsub(sv1, sv0), vsub(p, sv0))) < 0) return 0;
  if (vdot(normal, vcross(vsub(sv2, sv1), vsub(p, sv1))) < 0) return 0;
  if (vdot(normal, vcross(vsub(sv0, sv2), vsub(p, sv2))) < 0) return 0;
  float d0 = vlength(vsub(p, sv0)), d1 = vlength(vsub(p, sv1)), d2 = vlength(vsub(p, sv2));
  float denom = (1.0f / d0) + (1.0f / d1) + (1.0f / d2);
  float r0 = (1.0f / d0) / denom, r1 = (1.0f / d1) / denom, r2 = (1.0f / d2) / denom;
  float u = r0 * s->uv0[2 * i] + r1 * s->uv1[2 * i] + r2 * s->uv2[2 * i];
  float v = 1.0f - (r0 * s->uv0[2 * i + 1] + r1 * s->uv1[2 * i + 1] + r2 * s->uv2[2 * i + 1]);
  v3 outward = unit_vector(normal);
  rec->t = t;
  rec->p = ray_at(r, rec->t);
  set_face_normal(rec, r, outward);
  if (s->n0) {   /* NOT the reference (srt.h: tri_n0): vertex normals blended with the weights of model.h:158-166 */
    const float *a = s->n0 + 3 * i, *b = s->n1 + 3 * i, *c = s->n2 + 3 * i;
    v3 sn = unit_vector(V(r0 * a[0] + r1 * b[0] + r2 * c[0], r0 * a[1] + r1 * b[1] + r2 * c[1], r0 * a[2] + r1 * b[2] + r2 * c[2]));
    if (sn.x == sn.x && sn.y == sn.y && sn.z == sn.z) rec->normal = rec->front_face ? sn : vneg(sn);
  }
  rec->u = u; rec->v = v;
  rec->mat = s->tri_mat[i];
  triangle_tangent_basis(s, i, &rec->tangent, &rec->bitangent);
  if (s->tg0) {  /* NOT the reference (srt.h: tri_tan0): glTF vertex tangents blended with the same weights */
    const float *a = s->tg0 + 4 * i, *b = s->tg1 + 4 * i, *c = s->tg2 + 4 * i;
    if (a[3] != 0.0f || b[3] != 0.0f || c[3] != 0.0f) {
      v3 T = unit_vector(V(r0 * a[0] + r1 * b[0] + r2 * c[0], r0 * a[1] + r1 * b[1] + r2 * c[1], r0 * a[2] + r1 * b[2] + r2 * c[2]));
      v3 Bu = unit_vector(vcross(rec->normal, T));
      v3 Bt = V(a[3] * Bu.x, a[3] * Bu.y, a[3] * Bu.z);
      float chk = (T.x + T.y + T.z) + (Bt.x + Bt.y + Bt.z);
      if (chk == chk && fabsf(chk) <= 3.4e38f) { rec->tangent = T; rec->bitangent = Bt; }
    }
  }
  return 1;
}

static int prim_hit(const so_scene* s, int32_t prim, const ray_t* r, float tMin, float tMax, rec_t* rec, counters_t* c) {
  int h;
  if (c) c->leaf++;
  if (prim < s->n_tri) h = triangle_hit(s, prim, r, tMin, tMax, rec);
  else h = sphere_hit(s, prim - s->n_tri, r, tMin, tMax, rec);
  if (h) rec->prim = prim;
  return h;
}

/* bvh.h:97-105: box, then left, then right with tMax = hitLeft ? rec.t : tMax.  A span-1 node
   (bvh.h:67-69, right == NONE) visits its single child twice, exactly as left == right does. */
static int child_hit(const so_scene* s, int32_t child, const ray_t* r, float tMin, float tMax, rec_t* rec, counters_t* c) {
  if (child < 0) return prim_hit(s, ~child, r, tMin, tMax, rec, c);
  const srt_bvh_node* n = &s->nodes[child];
  if (c) c->box++;
  if (!aabb_hit(n->bmin, n->bmax, r, tMin, tMax)) return 0;
  int32_t right = (n->right == SRT_BVH_NONE) ? n->left : n->right;
  int hl = child_hit(s, n->left, r, tMin, tMax, rec, c);
  int hr = child_hit(s, right, r, tMin, hl ? rec->t : tMax, rec, c);
  return hl || hr;
}

/* hittableList::hit with its single element, the BVH root (hittablelist.h:33-47, main.cpp:146) */
static int world_hit(const so_scene* s, const ray_t* r, float tMin, float tMax, rec_t* rec, counters_t* c) {
  if (s->n_nodes == 0) return 0;
  return child_hit(s, 0, r, tMin, tMax, rec, c);
}

static ray_t to_ray(const srt_ray* q) { ray_t r; r.o = ld3(q->o); r.dir = ld3(q->dir); r.time = q->time; return r; }
static void from_ray(const ray_t* r, srt_ray* q) { st3(q->o, r->o); st3(q->dir, r->dir); q->time = r->time; }

static void rec_to_hit(int hit, const rec_t* rec, srt_hit* h) {
  memset(h, 0, sizeof *h);
  if (!hit) { h->prim = -1; h->mat = -1; return; }
  h->prim = rec->prim; h->mat = rec->mat; h->front_face = rec->front_face; h->t = rec->t;
  st3(h->p, rec->p); st3(h->normal, rec->normal); st3(h->tangent, rec->tangent); st3(h->bitangent, rec->bitangent);
  h->uv[0] = rec->u; h->uv[1] = rec->v;
}
static rec_t hit_to_rec(const srt_hit* h) {
  rec_t rec; memset(&rec, 0, sizeof rec);
  rec.p = ld3(h->p); rec.normal = ld3(h->normal); rec.tangent = ld3(h->tangent); rec.bitangent = ld3(h->bitangent);
  rec.u = h->uv[0]; rec.v = h->uv[1]; rec.t = h->t; rec.front_face = h->front_face; rec.mat = h->mat; rec.prim = h->prim;
  return rec;
}

void so_trace_rays(const so_scene* s, const srt_ray* rays, int64_t n, float tMin, float tMax, srt_hit* out) {
  for (int64_t i = 0; i < n; i++) {
    ray_t r = to_ray(&rays[i]); rec_t rec; memset(&rec, 0, sizeof rec);
    int h = world_hit(s, &r, tMin, tMax, &rec, NULL);
    rec_to_hit(h, &rec, &out[i]);
  }
}

void so_count_tests(const so_scene* s, const srt_ray* rays, int64_t n, float tMin, float tMax, uint64_t* box, uint64_t* leaf) {
  counters_t c = {0, 0};
  for (int64_t i = 0; i < n; i++) {
    ray_t r = to_ray(&rays[i]); rec_t rec; memset(&rec, 0, sizeof rec);
    world_hit(s, &r, tMin, tMax, &rec, &c);
  }
  *box = c.box; *leaf = c.leaf;
}

void so_aabb_kat(const srt_ray* rays, const float* box, const float* tMinMax, int64_t n, int32_t* out) {
  for (int64_t i = 0; i < n; i++) {
    ray_t r = to_ray(&rays[i]);
    out[i] = aabb_hit(box + 6 * i, box + 6 * i + 3, &r, tMinMax[2 * i], tMinMax[2 * i + 1]);
  }
}

/* ------------------------------------------------------------------------------------------ textures */
/* texture.h:26-28 (solidColor), :42-48 (checker), :129-148 (imagePNG / image3bpp :82-101) */
static v3 texture_value(const so_scene* s, int tex, float u, float v, v3 p) {
  const srt_texture* t = &s->texs[tex];
  if (t->kind == SRT_TEX_SOLID) return ld3(t->color);
  if (t->kind == SRT_TEX_CHECKER) {
    float sines = sinf(10.0f * p.x) * sinf(10.0f * p.y) * sinf(10.0f * p.z);
    v3 c = texture_value(s, sines < 0 ? t->odd : t->even, u, v, p);
    return V(c.x * 255.0f, c.y * 255.0f, c.z * 255.0f);
  }
  if (t->data == NULL) return V(1.0f, 0, 1.0f);
  u = clampf(u, 0, 1.0f);
  v = 1.0f - clampf(v, 0, 1.0f);
  /* static_cast<int> of a NaN is INT_MIN on x86 (out-of-bounds read in the reference);
     pinned to texel 0 here and on the GPU (cvt.rzi of NaN = 0) */
  int i = (u != u) ? 0 : (int)(u * t->width);
  int j = (v != v) ? 0 : (int)(v * t->height);
  if (i >= t->width) i = t->width - 1;
  if (j >= t->height) j = t->height - 1;
  const uint8_t* pixel = t->data + (size_t)j * ((size_t)t->bpp * t->width) + (size_t)i * t->bpp;
  return V(pixel[0], pixel[1], pixel[2]);
}

void so_texture_kat(const so_scene* s, int tex, const float* uvp, int64_t n, float* out) {
  for (int64_t i = 0; i < n; i++) {
    const float* q = uvp + 5 * i;
    st3(out + 3 * i, texture_value(s, tex, q[0], q[1], V(q[2], q[3], q[4])));
  }
}

/* ------------------------------------------------------------------------------------------ materials */
/* pbr.h:58-65 */
static float trowbridge_reitz_ndf(float NdotH, float roughness) {
  float alpha = roughness * roughness;
  float alpha2 = alpha * alpha;
  float NdotH2 = NdotH * NdotH;
  float denom = PI_F * powf(NdotH2 * (alpha2 - 1.0f) + 1.0f, 2.0f);
  return alpha2 / denom;
}
/* pbr.h:69-73 */
static float schlick_gaf(float NdotV, float roughness) {
  float k = ((roughness + 1.0f) * (roughness + 1.0f)) / 8.0f;
  return NdotV / (NdotV * (1.0f - k) + k);
}
/* vec3.h:76-78 */
static v3 reflect(v3 v, v3 n) { float s = 2.0f * vdot(v, n); return V(v.x - s * n.x, v.y - s * n.y, v.z - s * n.z); }
/* vec3.h:80-86 */
static v3 refract(v3 uv, v3 n, float eta) {
  float cosTheta = fminf(vdot(n, vneg(uv)), 1.0f);
  v3 perp = V(eta * (uv.x + cosTheta * n.x), eta * (uv.y + cosTheta * n.y), eta * (uv.z + cosTheta * n.z));
  float sp = -sqrtf(fabsf(1.0f - length_squared(perp)));
  return V(perp.x + sp * n.x, perp.y + sp * n.y, perp.z + sp * n.z);
}

/* material::emitted: default (material.h:18-20) = 0; diffuseLight (material.h:148-150) = emit->value */
static v3 material_emitted(const so_scene* s, const srt_material* m, float u, float v, v3 p) {
  if (m->kind == SRT_MAT_LIGHT) return texture_value(s, m->albedo_tex, u, v, p);
  return V(0, 0, 0);
}

/* material::scatter.  rvec = the random vector the sampler produced (a unit vector for PBR,
   a point in the unit ball for METAL), xi = the uniform number DIELECTRIC compares against. */
static int material_scatter(const so_scene* s, const srt_material* m, const ray_t* rIn, const rec_t* rec, v3 rvec,
                            float xi, v3* attenuation, ray_t* out) {
  if (m->kind == SRT_MAT_LIGHT) return 0;                                   /* material.h:144-146 */
  if (m->kind == SRT_MAT_METAL) {                                           /* material.h:91-97 */
    v3 reflected = reflect(unit_vector(rIn->dir), rec->normal);
    out->o = rec->p;
    out->dir = V(reflected.x + m->fuzz * rvec.x, reflected.y + m->fuzz * rvec.y, reflected.z + m->fuzz * rvec.z);
    out->time = rIn->time;
    *attenuation = V(m->albedo[0], m->albedo[1], m->albedo[2]);
    return vdot(out->dir, rec->normal) > 0;
  }
  if (m->kind == SRT_MAT_DIELECTRIC) {                                      /* material.h:108-126,132-136 */
    *attenuation = V(1.0f, 1.0f, 1.0f);
    float ratio = rec->front_face ? (1.0f / m->ir) : m->ir;
    v3 unitDir = unit_vector(rIn->dir);
    float cosTheta = fminf(vdot(rec->normal, vneg(unitDir)), 1.0f);
    float sinTheta = sqrtf(1.0f - cosTheta * cosTheta);
    int cannotRefract = ratio * sinTheta > 1.0f;
    float r0 = (1.0f - ratio) / (1.0f + ratio);
    r0 = r0 * r0;
    float reflectance = r0 + (1.0f - r0) * powf((1.0f - cosTheta), 5.0f);
    v3 dir;
    if (cannotRefract || reflectance > xi) dir = reflect(unitDir, rec->normal);
    else dir = refract(unitDir, rec->normal, ratio);
    out->o = rec->p; out->dir = dir; out->time = rIn->time;
    return 1;
  }
  /* pbrMetallicRoughness::scatter, material.h:156-245 */
  v3 att, normal;
  float mm, rr;
  if (m->albedo_tex >= 0) {
    att = texture_value(s, m->albedo_tex, rec->u, rec->v, rec->p);
    att = V(att.x / 255.0f, att.y / 255.0f, att.z / 255.0f);
  } else att = V(m->albedo[0], m->albedo[1], m->albedo[2]);
  if (m->normal_tex >= 0) {
    v3 n = texture_value(s, m->normal_tex, rec->u, rec->v, rec->p);
    n = V((n.x - 128.0f) / 128.0f, (n.y - 128.0f) / 128.0f, (n.z - 128.0f) / 128.0f);       /* vec3.h:103-110 */
    v3 T = rec->tangent, B = rec->bitangent, N = rec->normal;                                /* columns of the 3x3 */
    normal = unit_vector(V(T.x * n.x + (B.x * n.y + N.x * n.z), T.y * n.x + (B.y * n.y + N.y * n.z),
                           T.z * n.x + (B.z * n.y + N.z * n.z)));
  } else normal = rec->normal;
  if ((m->flags & SRT_MATF_PACKED_MR) && m->metallic_tex >= 0) {   /* glTF metallicRoughnessTexture: NOT the reference (srt.h) */
    v3 mr = texture_value(s, m->metallic_tex, rec->u, rec->v, rec->p);
    mm = clampf(mr.z / 255.0f * m->metalness, 0, 1.0f);
    rr = clampf(mr.y / 255.0f * m->roughness, 0, 1.0f);
  } else {
    if (m->metallic_tex >= 0) mm = clampf(texture_value(s, m->metallic_tex, rec->u, rec->v, rec->p).x / 255.0f, 0, 1.0f);
    else mm = m->metalness;
    if (m->roughness_tex >= 0) rr = clampf(texture_value(s, m->roughness_tex, rec->u, rec->v, rec->p).y / 255.0f, 0, 1.0f);
    else rr = m->roughness;
  }

  v3 sd = vadd(normal, rvec);
  if (((double)fabsf(sd.x) < 1e-8) && ((double)fabsf(sd.y) < 1e-8) && ((double)fabsf(sd.z) < 1e-8)) sd = normal; /* vec3.h:49-52 */
  sd = unit_vector(sd);
  out->o = rec->p; out->dir = sd; out->time = rIn->time;
  v3 viewVec = vneg(unit_vector(rIn->dir));
  v3 halfVec = unit_vector(vadd(sd, viewVec));
  float NdotL = fmaxf(vdot(normal, sd), 0);
  float NdotH = fmaxf(vdot(normal, halfVec), 0);
  float HdotV = fmaxf(vdot(halfVec, viewVec), 0);
  float NdotV = fmaxf(vdot(normal, viewVec), 0);
  float fr[3] = {m->albedo[0], m->albedo[1], m->albedo[2]};
  float D = trowbridge_reitz_ndf(NdotH, rr);
  float power = powf(2.0f, (-5.55473f * HdotV - 6.98316f) * HdotV);                         /* pbr.h:75-81 */
  float G = schlick_gaf(NdotL, rr) * schlick_gaf(NdotV, rr);
  float a3[3] = {att.x, att.y, att.z}, res[3];
  for (int k = 0; k < 3; k++) {
    float F0 = (1.0f - mm) * 0.4f + mm * fr[k];                                              /* vec3.h:97-101, material.h:228 */
    float F = F0 + (1.0f - F0) * power;
    float fd = a3[k] / PI_F;
    fd *= (1.0f - F);
    fd *= (1.0f - mm);
    fd *= fr[k];
    float fs = D * F * G / (4.0f * NdotV * NdotL + FLT_EPSILON);
    res[k] = (fd + fs) * NdotL;
  }
  *attenuation = V(res[0], res[1], res[2]);
  return 1;
}

void so_shade_kat(const so_scene* s, const srt_ray* raysIn, const srt_hit* hits, const float* xi, int64_t n, float* att,
                  float* emit, srt_ray* raysOut, int32_t* scattered) {
  for (int64_t i = 0; i < n; i++) {
    ray_t rIn = to_ray(&raysIn[i]); rec_t rec = hit_to_rec(&hits[i]);
    const srt_material* m = &s->mats[rec.mat];
    const float* q = xi + 4 * i;
    v3 a = V(0, 0, 0); ray_t out; memset(&out, 0, sizeof out);
    st3(emit + 3 * i, material_emitted(s, m, rec.u, rec.v, rec.p));
    scattered[i] = material_scatter(s, m, &rIn, &rec, V(q[0], q[1], q[2]), q[3], &a, &out);
    st3(att + 3 * i, a); from_ray(&out, &raysOut[i]);
  }
}

/* ------------------------------------------------------------------------------------------ random numbers (NEW) */
/* Philox4x32-10 (Salmon et al. 2011): key = seed, counter = (pixel, sample, block, 0) */
void so_philox(uint64_t seed, uint32_t pixel, uint32_t sample, uint32_t block, uint32_t out[4]) {
  uint32_t c0 = pixel, c1 = sample, c2 = block, c3 = 0;
  uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
  for (int round = 0; round < 10; round++) {
    uint64_t p0 = (uint64_t)0xD2511F53u * c0, p1 = (uint64_t)0xCD9E8D57u * c2;
    uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0, n1 = (uint32_t)p1, n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1, n3 = (uint32_t)p0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}
/* 24-bit uniform in [0,1): the granularity of libstdc++'s generate_canonical<float,24> behind globals.h:31 */
static inline float u01(uint32_t x) { return (float)(x >> 8) * (1.0f / 16777216.0f); }

/* uniform point in the unit disk: same distribution as randomInUnitDisk (vec3.h:88-95), sampled directly */
static void sample_disk(float xa, float xb, float* dx, float* dy) {
  float r = sqrtf(xa), phi = (2.0f * PI_F) * xb;
  *dx = r * cosf(phi); *dy = r * sinf(phi);
}
/* uniform direction: same distribution as randomUnitVector (vec3.h:72-74), sampled directly */
static v3 sample_unit_vector(float xa, float xb) {
  float z = 1.0f - 2.0f * xa;
  float rxy = sqrtf(fmaxf(0.0f, 1.0f - z * z)), phi = (2.0f * PI_F) * xb;
  return V(rxy * cosf(phi), rxy * sinf(phi), z);
}
/* uniform point in the unit ball: same distribution as randomInUnitSphere (vec3.h:62-70) */
static v3 sample_unit_ball(float xa, float xb, float xc) {
  v3 d = sample_unit_vector(xa, xb);
  float r = cbrtf(xc);
  return V(r * d.x, r * d.y, r * d.z);
}

/* camera::getRay (camera.h:40-50) with the lens-disk point and the time variate supplied */
static ray_t camera_get_ray(const srt_camera* c, float s, float t, float diskx, float disky, float xiTime) {
  float rdx = c->lens_radius * diskx, rdy = c->lens_radius * disky;
  v3 offset = V(rdx * c->hor[0] + rdy * c->vert[0], rdx * c->hor[1] + rdy * c->vert[1], rdx * c->hor[2] + rdy * c->vert[2]);
  ray_t r;
  r.o = vadd(ld3(c->origin), offset);
  const float* L = c->lleft; const float* Hh = c->horizontal; const float* Vv = c->vertical; const float* O = c->origin;
  r.dir = V(L[0] + s * Hh[0] + t * Vv[0] - O[0] - offset.x, L[1] + s * Hh[1] + t * Vv[1] - O[1] - offset.y,
            L[2] + s * Hh[2] + t * Vv[2] - O[2] - offset.z);
  r.time = c->t0 + (c->t1 - c->t0) * xiTime;                                   /* globals.h:37-39 */
  return r;
}

void so_getray_kat(const srt_camera* c, const float* in, int64_t n, srt_ray* out) {
  for (int64_t i = 0; i < n; i++) {
    const float* q = in + 5 * i;
    ray_t r = camera_get_ray(c, q[0], q[1], q[2], q[3], q[4]);
    from_ray(&r, &out[i]);
  }
}

/* ------------------------------------------------------------------------------------------ integrator */
#define SO_MAX_DEPTH 64

/* rayColor (main.cpp:33-52) for sample `sample` of pixel (x, y): the recursion
   L = emitted + attenuation * L(next) is unwound in the reference's order (innermost first). */
static v3 path_radiance(const so_scene* s, const srt_render_params* P, ray_t r, uint32_t pixel, uint32_t sample,
                        uint64_t* segments) {
  v3 em[SO_MAX_DEPTH], at[SO_MAX_DEPTH];
  v3 tail = V(0, 0, 0);                               /* maxBounce <= 0 -> (0,0,0), main.cpp:36-37 */
  v3 thr = V(1, 1, 1);
  int depth = 0;
  int maxBounce = P->max_bounce < SO_MAX_DEPTH ? P->max_bounce : SO_MAX_DEPTH;
  for (int b = 0; b < maxBounce; b++) {
    rec_t rec; memset(&rec, 0, sizeof rec);
    ++*segments;
    if (!world_hit(s, &r, 0.001f, INFINITY, &rec, NULL)) { tail = ld3(P->background); break; }      /* main.cpp:39-40 */
    const srt_material* m = &s->mats[rec.mat];
    uint32_t x[4];
    so_philox(P->seed, pixel, sample, 2u + (uint32_t)b, x);
    v3 rvec = V(0, 0, 0);
    if (m->kind == SRT_MAT_PBR) rvec = sample_unit_vector(u01(x[0]), u01(x[1]));
    else if (m->kind == SRT_MAT_METAL) rvec = sample_unit_ball(u01(x[0]), u01(x[1]), u01(x[2]));
    v3 emitted = material_emitted(s, m, rec.u, rec.v, rec.p);                                       /* main.cpp:44 */
    v3 attenuation; ray_t scattered;
    if (!material_scatter(s, m, &r, &rec, rvec, u01(x[2]), &attenuation, &scattered)) { tail = emitted; break; }  /* :46-47 */
    if (P->rr_begin > 0 && b + 1 >= P->rr_begin) {    /* Russian roulette: NOT in the reference; off in parity runs */
      thr = V(thr.x * attenuation.x, thr.y * attenuation.y, thr.z * attenuation.z);
      float q = fminf(fmaxf(fmaxf(thr.x, fmaxf(thr.y, thr.z)), 0.05f), 1.0f);
      if (!(u01(x[3]) < q)) { tail = emitted; break; }
      float inv = 1.0f / q;
      attenuation = V(attenuation.x * inv, attenuation.y * inv, attenuation.z * inv);
      thr = V(thr.x * inv, thr.y * inv, thr.z * inv);
    } else thr = V(thr.x * attenuation.x, thr.y * attenuation.y, thr.z * attenuation.z);
    em[depth] = emitted; at[depth] = attenuation; depth++;
    r = scattered;
  }
  v3 L = tail;
  for (int k = depth - 1; k >= 0; k--)                                                              /* main.cpp:49-51 */
    L = V(em[k].x + L.x * at[k].x, em[k].y + L.y * at[k].y, em[k].z + L.z * at[k].z);
  return L;
}

typedef struct {
  const so_scene* s; const srt_camera* cam; const srt_render_params* P; int y0, y1; float* sum; uint64_t paths, segments;
} band_t;

/* main.cpp:181,205-207 (commented-out 4x rotated grid): AngleAxisf(deg2rad(15) + deg2rad(90 s), UnitZ) * (0, 0.8, 0);
   Eigen's matrix for the z axis has m01 = -sin(angle), m11 = cos(angle) (Geometry/AngleAxis.h toRotationMatrix) */
void so_rotated_grid(int s, float* x, float* y) {
  const float pi = 3.1415926535897932385f;                                                         /* globals.h:15 */
  float ang = 15.0f * pi / 180.0f + (90.0f * (float)s) * pi / 180.0f;                               /* globals.h:26-28 */
  *x = -sinf(ang) * 0.8f;
  *y = cosf(ang) * 0.8f;
}

/* main.cpp:200-227 over rows [y0, y1) and global samples [sample_begin, sample_begin + sample_count) */
static void* render_band(void* arg) {
  band_t* b = (band_t*)arg;
  const srt_render_params* P = b->P;
  int W = P->width, H = P->height;
  int n = P->sample_count > 0 ? P->sample_count : P->spp;
  for (int y = b->y0; y < b->y1; y++) {
    for (int x = 0; x < W; x++) {
      v3 pixelColor = V(0, 0, 0);
      uint32_t pixel = (uint32_t)(y * W + x);
      for (int k = 0; k < n; k++) {
        uint32_t sample = (uint32_t)(P->sample_begin + k);
        uint32_t a[4], c[4];
        so_philox(P->seed, pixel, sample, 0, a);
        so_philox(P->seed, pixel, sample, 1, c);
        float ju = u01(a[0]), jv = u01(a[1]);
        if (P->sampler == 1) so_rotated_grid((int)sample, &ju, &jv);                               /* main.cpp:214-215 */
        float u = ((float)x + ju) / (float)(W - 1);                                                 /* main.cpp:210 */
        float v = ((float)(H - y) + jv) / (float)(H - 1);                                           /* main.cpp:211 */
        float dx, dy;
        sample_disk(u01(a[2]), u01(a[3]), &dx, &dy);
        ray_t r = camera_get_ray(b->cam, u, v, dx, dy, u01(c[0]));
        v3 L = path_radiance(b->s, P, r, pixel, sample, &b->segments);
        pixelColor = vadd(pixelColor, L);                                                           /* main.cpp:217 */
        b->paths++;
      }
      st3(b->sum + ((size_t)y * W + x) * 3, pixelColor);
    }
  }
  return NULL;
}

/* sum: [H*W*3]; rows outside [y0,y1) untouched.  nthreads row bands run in parallel (the reference is
   single threaded; per-pixel Philox streams make the result independent of the banding). */
void so_render(const so_scene* s, const srt_camera* cam, const srt_render_params* P, int y0, int y1, int nthreads,
               float* sum, srt_stats* stats) {
  if (nthreads < 1) nthreads = 1;
  if (nthreads > 256) nthreads = 256;
  band_t bands[256]; pthread_t th[256];
  int rows = y1 - y0;
  for (int i = 0; i < nthreads; i++) {
    band_t b = {s, cam, P, y0 + (int)((int64_t)rows * i / nthreads), y0 + (int)((int64_t)rows * (i + 1) / nthreads), sum, 0, 0};
    bands[i] = b;
  }
  if (nthreads == 1) render_band(&bands[0]);
  else {
    for (int i = 0; i < nthreads; i++) pthread_create(&th[i], NULL, render_band, &bands[i]);
    for (int i = 0; i < nthreads; i++) pthread_join(th[i], NULL);
  }
  if (stats) {
    memset(stats, 0, sizeof *stats);
    for (int i = 0; i < nthreads; i++) { stats->paths += bands[i].paths; stats->segments += bands[i].segments; }
  }
}

/* writeColorTarget, color.h:25-41 */
void so_resolve(const float* sum, int W, int H, int spp, uint8_t* rgba) {
  float scale = 1.0f / spp;
  for (size_t i = 0; i < (size_t)W * H; i++) {
    for (int k = 0; k < 3; k++) {
      float c = sqrtf(sum[3 * i + k] * scale);
      rgba[4 * i + k] = (uint8_t)(256 * clampf(c, 0.0f, 0.999f));
    }
    rgba[4 * i + 3] = 255;
  }
}
