// srt_core.cuh — device-side scene layout, exact fp32 helpers and the ordered BVH traversal.
//
// Arithmetic contract: every comparison that decides hit/miss, and the
// values t / p / normal / uv of an accepted hit, are evaluated as the reference evaluates them —
// same operation order, no FMA contraction (this translation unit is compiled with -fmad=false;
// -prec-div / -prec-sqrt stay at their IEEE defaults, denormals stay on).  FMAs appear only where
// written explicitly (fmaf) in the *filtered* box predicate, whose result is proven equal to the
// exact one or else recomputed exactly.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <float.h>

namespace srt {

struct V3 { float x, y, z; };
__device__ __forceinline__ V3 mk(float x, float y, float z) { V3 r; r.x = x; r.y = y; r.z = z; return r; }
__device__ __forceinline__ V3 operator+(V3 a, V3 b) { return mk(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ V3 operator-(V3 a, V3 b) { return mk(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ V3 operator-(V3 a) { return mk(-a.x, -a.y, -a.z); }
__device__ __forceinline__ V3 operator*(float s, V3 a) { return mk(s * a.x, s * a.y, s * a.z); }
// Eigen's 3-vector dot (Redux.h:101-115): a0*b0 + (a1*b1 + a2*b2)
__device__ __forceinline__ float dot(V3 a, V3 b) { return a.x * b.x + (a.y * b.y + a.z * b.z); }
// Eigen's cross (OrthoMethods.h:47-49)
__device__ __forceinline__ V3 cross(V3 a, V3 b) {
  return mk(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
// vec3.h:29-31
__device__ __forceinline__ float length_squared(V3 v) { return v.x * v.x + v.y * v.y + v.z * v.z; }
// vec3.h:33-35
__device__ __forceinline__ float length(V3 v) { return sqrtf(length_squared(v)); }
// vec3.h:54-60
__device__ __forceinline__ V3 unit_vector(V3 v) {
  float len = length(v);
  if (len != 0) return mk(v.x / len, v.y / len, v.z / len);
  return v;
}
// globals.h:17-24
__device__ __forceinline__ float clampf(float x, float lo, float hi) { return x < lo ? lo : (x > hi ? hi : x); }

#define SRT_PI 3.1415926535897932385f  /* globals.h:15 */

struct Ray { V3 o, d; float time; };

// ---- device scene -----------------------------------------------------------------------------
// BVH node, 32 bytes, "threaded" form of the caller's pre-order tree (include/srt.h: srt_bvh_node).
// The reference's traversal order is fixed (bvh.h:97-105: left subtree, then right subtree), so the
// recursion is equivalent to walking the pre-order array with a skip link — no stack:
//   inner node : box hit -> i+1 (its left child is next in pre-order);  box miss -> skip
//   leaf node  : box hit -> test prim a, then prim b;                   then i+1 either way
// lo = (min.x, min.y, min.z, max.x)   hi = (max.y, max.z, w6, w7)
//   inner: w6 = skip index, w7 = NODE_INNER;  leaf: w6 = prim a, w7 = prim b | NODE_NO_B | NODE_NO_BOX
struct Node { float4 lo, hi; };
#define NODE_INNER  ((int)0x80000000)
#define NODE_NO_B   (-1)     /* span-1 node of the reference (bvh.h:67-69): one primitive */
#define NODE_NO_BOX (-2)     /* a primitive hanging directly under an inner node: no box of its own */

// triangle, 64 bytes: positions + the reference's un-normalised normal N = (v1-v0)x(v2-v0)
// (model.h:276-283) and plane term dd = -(N.v0) (model.h:125), precomputed with the same arithmetic
struct Tri { float4 a, b, c, d; };   // a=(v0,N.x) b=(v1,N.y) c=(v2,N.z) d=(dd, mat, -, -)
struct TriUV { float4 a, b; };       // a=(uv0, uv1) b=(uv2, -, -)
struct Sph { float4 a, b, c; };      // a=(c0, r) b=(c1, t0) c=(t1, mat, moving, -)

struct Tex {
  int kind;                 // SRT_TEX_*
  float r, g, b;            // SOLID
  int even, odd;            // CHECKER
  int w, h;                 // IMAGE
  int missing;              // IMAGE: file was missing -> (1,0,1)
  int pad;
  const uchar4* texels;     // IMAGE: w*h texels in the scene arena, row j at texels + j*w (expanded on the device by k_expand_texels)
};

#define MAT_NEEDS_UV  1
#define MAT_NEEDS_TBN 2
#define MAT_PACKED_MR 4   // SRT_MATF_PACKED_MR
struct Mat {
  int kind, albedo_tex, normal_tex, metallic_tex;
  int roughness_tex, flags; float metalness, roughness;
  float albedo[4];
  float fuzz, ir, pad0, pad1;
};

// ---- compact tree (trees that do not fit shared memory; BASELINE config 5) ---------------------------------------
// A second, 16-byte image of every node for the walk in global memory, where the cost of a step is the number of
// 32-byte sectors it pulls from HBM (profiles/r1_c5_extend_summary.md): the box snapped OUTWARD to a 16-bit grid over
// the root box (6 x uint16) + one word.  Two nodes share a sector, and node i+1 - the left child, next in pre-order -
// is the other half of node i's sector every second time.
//   q.x = lo.x | lo.y << 16    q.y = lo.z | hi.x << 16    q.z = hi.y | hi.z << 16
//   q.w : bit 31 set   -> inner node, bits 0-30 = skip index
//         bit 31 clear -> leaf: bits 0-28 = index u of its first primitive in the LEAF-ORDERED primitive array
//                         (CNODE_NO_PRIM = none: the sentinel), bit 29 = a second primitive follows at u + 1,
//                         bit 30 = the node has no box (a primitive hanging directly under an inner node)
// The snapped box only FILTERS: with outer = [lo, hi] cells and inner = [lo + 1, hi - 1] cells,
//   ray misses outer  => misses the true box;   ray hits inner => hits the true box   (slab intervals are nested),
// both decided with a rounding margin; the thin shell in between (a few % of the tests) and every ray the filter
// cannot serve re-evaluates aabb::hit on the full 32-byte node with the reference's arithmetic.
#define CNODE_INNER   0x80000000u
#define CNODE_TWO     0x20000000u
#define CNODE_NO_BOX  0x40000000u
#define CNODE_NO_PRIM 0x1fffffffu
// leaf-ordered primitive record, 64 bytes: a triangle as Tri with d = (dd, mat, original id, 0), or a sphere as
// (Sph.a, Sph.b, Sph.c, d = (-, mat, original id, 1)) - the two primitives of a leaf are neighbours in memory
struct LPrim { float4 a, b, c, d; };

struct Scene {
  const uint4* cnodes;   // compact tree, or nullptr
  const LPrim* lprims;
  float q_org[3], q_cell[3];
  const Node*  nodes;  int n_nodes;
  const Tri*   tris;   int n_tri;
  const TriUV* tri_uv;
  const float4* tri_vt;  // optional per-vertex tangents (xyz, handedness), 3 x float4 per triangle (srt_scene_desc::tri_tan0..2), or nullptr
  const float4* tri_vn;  // optional per-vertex normals, 3 x float4 per triangle (srt_scene_desc::tri_n0..2), or nullptr
  const Sph*   sphs;   int n_sph;
  const Mat*   mats;   int n_mat;
  const Tex*   texs;   int n_tex;
  int filter_ok;       // every box coordinate is finite and small enough for the filtered predicate
  // k_wf_extend2 (srt_extend2.cuh): the tree is nested (every node's box contains its children's), has no box-less
  // primitives and coordinates within 2^40 -> inner nodes may be culled conservatively; centre = the point box
  // coordinates are taken relative to in the walk's 16-byte nodes (median of the leaf boxes' centres)
  int x2_ok;
  float x2_centre[3];
  // the tree k_wf_extend2 / k_wf_path walk: the same leaves in the same order under FEWER inner nodes (srt_scene_create:
  // an inner node whose box is nearly its parent's culls too little to pay for its test and is left out - its children
  // hang under the parent).  Same 32-byte format and threading as `nodes`; leaves are copies of the original leaf nodes.
  const Node* x2_nodes;
  int x2_n;
  const uint4* x2_cnodes;   // the compact image of x2_nodes (trees in HBM), or nullptr; leaf words index the same lprims
};

// ---- aabb::hit, aabb.h:11-27 --------------------------------------------------------------------
// exact form: six true divisions, fminf/fmaxf (NaN -> the other operand, as in C), tMax <= tMin misses
__device__ __forceinline__ bool box_hit_exact(const float4& lo, const float4& hi, const Ray& r, float tMin, float tMax) {
  float a, b, t0, t1;
  a = (lo.x - r.o.x) / r.d.x; b = (lo.w - r.o.x) / r.d.x;
  t0 = fminf(a, b); t1 = fmaxf(a, b); tMin = fmaxf(t0, tMin); tMax = fminf(t1, tMax);
  if (tMax <= tMin) return false;
  a = (lo.y - r.o.y) / r.d.y; b = (hi.x - r.o.y) / r.d.y;
  t0 = fminf(a, b); t1 = fmaxf(a, b); tMin = fmaxf(t0, tMin); tMax = fminf(t1, tMax);
  if (tMax <= tMin) return false;
  a = (lo.z - r.o.z) / r.d.z; b = (hi.y - r.o.z) / r.d.z;
  t0 = fminf(a, b); t1 = fmaxf(a, b); tMin = fmaxf(t0, tMin); tMax = fminf(t1, tMax);
  if (tMax <= tMin) return false;
  return true;
}

// Filtered predicate.  aabb::hit only returns a bool, so a cheaper evaluation may stand in for it
// wherever it provably gives the same answer.  With id = fl(1/d), oid = fl(o*id) the slab bound
// a' = fma(c, id, -oid) differs from the reference's fl(fl(c-o)/d) by at most
// 4.02u|q| + 1.01u|o/d| (u = 2^-24, q the true quotient); max/min over the axes keep that bound.
// The interval test is decided by the filter when |hi - lo| exceeds 2^-20 (M + |lo| + |hi|) + 1e-30,
// M = max|oid| — a 4x margin over the bound — and by box_hit_exact otherwise.  Rays with a zero,
// non-finite or extreme direction component never use the filter (ok == 0).
struct RayAux { float idx, idy, idz, oidx, oidy, oidz, M; int ok; };

__device__ __forceinline__ RayAux make_aux(const Ray& r, int scene_filter_ok) {
  RayAux a;
  a.idx = 1.0f / r.d.x; a.idy = 1.0f / r.d.y; a.idz = 1.0f / r.d.z;
  a.oidx = r.o.x * a.idx; a.oidy = r.o.y * a.idy; a.oidz = r.o.z * a.idz;
  float mi = fmaxf(fabsf(a.idx), fmaxf(fabsf(a.idy), fabsf(a.idz)));
  a.M = fmaxf(fabsf(a.oidx), fmaxf(fabsf(a.oidy), fabsf(a.oidz)));
  // mi <= 2^40 and M <= 2^60 fail for NaN/inf as well; box coordinates are <= 2^40 when scene_filter_ok
  a.ok = scene_filter_ok && (mi <= 1.0995116e12f) && (a.M <= 1.1529215e18f) && (r.time == r.time);
  return a;
}

template <bool COUNT>
__device__ __forceinline__ bool box_hit(const float4& lo, const float4& hi, const Ray& r, const RayAux& x, float tMin,
                                        float tMax, unsigned long long* exact_count) {
  if (x.ok) {
    float ax = fmaf(lo.x, x.idx, -x.oidx), bx = fmaf(lo.w, x.idx, -x.oidx);
    float ay = fmaf(lo.y, x.idy, -x.oidy), by = fmaf(hi.x, x.idy, -x.oidy);
    float az = fmaf(lo.z, x.idz, -x.oidz), bz = fmaf(hi.y, x.idz, -x.oidz);
    float l = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fmaxf(fminf(az, bz), tMin));
    float h = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fminf(fmaxf(az, bz), tMax));
    float slack = fmaf(9.5367432e-7f, x.M + fabsf(l) + fabsf(h), 1e-30f);
    float diff = h - l;
    if (diff > slack) return true;
    if (diff < -slack) return false;
    // tMax = +inf and an unbounded slab give h = inf: diff = inf > slack handled above; NaN falls through
  }
  if (COUNT) ++*exact_count;
  return box_hit_exact(lo, hi, r, tMin, tMax);
}

// ---- triangle::hit decision, model.h:104-154 ---------------------------------------------------------
// plane intersection + three inside-edge tests, back-face culled, NO tMax test.  Returns t on accept.
__device__ __forceinline__ bool tri_hit(const Tri* __restrict__ tris, int i, const Ray& r, float tMin, float& tOut) {
  const float4 A = __ldg(&tris[i].a), B = __ldg(&tris[i].b), C = __ldg(&tris[i].c);
  const float dd = __ldg(&tris[i].d.x);
  V3 N = mk(A.w, B.w, C.w);
  float NdotDir = dot(N, r.d);
  if (fabsf(NdotDir) < FLT_EPSILON) return false;       // model.h:119
  if (NdotDir > 0) return false;                        // model.h:122 (dir.dot(normal): same products, same sum)
  float t = -(dot(N, r.o) + dd) / NdotDir;              // model.h:125-126
  if (t < tMin) return false;                           // model.h:128; NaN t passes, as in the reference
  V3 v0 = mk(A.x, A.y, A.z), v1 = mk(B.x, B.y, B.z), v2 = mk(C.x, C.y, C.z);
  V3 p = mk(r.o.x + t * r.d.x, r.o.y + t * r.d.y, r.o.z + t * r.d.z);
  if (dot(N, cross(v1 - v0, p - v0)) < 0) return false;  // model.h:136-140
  if (dot(N, cross(v2 - v1, p - v1)) < 0) return false;  // model.h:143-147
  if (dot(N, cross(v0 - v2, p - v2)) < 0) return false;  // model.h:150-154
  tOut = t;
  return true;
}

// sphere.h:47-52
__device__ __forceinline__ V3 sph_center(const float4& a, const float4& b, const float4& c, float time) {
  V3 c0 = mk(a.x, a.y, a.z);
  if (c.z != 0.0f) {    // center0 != center1
    V3 c1 = mk(b.x, b.y, b.z);
    float f = (time - b.w) / (c.x - b.w);
    return c0 + f * (c1 - c0);
  }
  return c0;
}

// sphere::hit decision, sphere.h:54-72: nearest root in the INCLUSIVE range [tMin, tMax]
__device__ __forceinline__ bool sph_hit(const Sph* __restrict__ sphs, int j, const Ray& r, float tMin, float tMax, float& tOut) {
  const float4 a = __ldg(&sphs[j].a), b = __ldg(&sphs[j].b), c = __ldg(&sphs[j].c);
  V3 oc = r.o - sph_center(a, b, c, r.time);
  float aa = length_squared(r.d);
  float halfB = dot(oc, r.d);
  float cc = length_squared(oc) - a.w * a.w;
  float disc = halfB * halfB - aa * cc;
  if (disc < 0.0f) return false;
  float sqrtd = sqrtf(disc);
  float root = (-halfB - sqrtd) / aa;
  if (root < tMin || root > tMax) {
    root = (-halfB + sqrtd) / aa;
    if (root < tMin || root > tMax) return false;
  }
  tOut = root;
  return true;
}

struct Counters { unsigned long long nodes, leaves, exact, sph_leaves, tri_hits; };

// ---- compact-node box filter -------------------------------------------------------------------------------------
// Slab bounds of the snapped box: t(q) = ((org + q cell) - o) / d = fma(q, cid, c0) with cid = fl(cell id), id = fl(1/d),
// c0 = fl(fl(org - o) id).  |t_computed - t_real| <= 3u Mq + u|t| (u = 2^-24, Mq = max_k |c0_k| + 65535 |cid_k|); the
// reference's own fl(fl(c - o) / d) is within 2u|t| of the real quotient.  Both decisions keep a margin of
// 2^-19 (Mq + |l| + |h|) + 1e-30 - more than 5x the sum of those bounds.
struct RayAuxC { float cx, cy, cz, kx, ky, kz, Mq; int ok; };

__device__ __forceinline__ RayAuxC make_aux_c(const Scene& sc, const Ray& r) {
  RayAuxC a;
  const float idx = 1.0f / r.d.x, idy = 1.0f / r.d.y, idz = 1.0f / r.d.z;
  a.cx = sc.q_cell[0] * idx; a.cy = sc.q_cell[1] * idy; a.cz = sc.q_cell[2] * idz;
  a.kx = (sc.q_org[0] - r.o.x) * idx; a.ky = (sc.q_org[1] - r.o.y) * idy; a.kz = (sc.q_org[2] - r.o.z) * idz;
  a.Mq = fmaxf(fmaf(65535.0f, fabsf(a.cx), fabsf(a.kx)), fmaxf(fmaf(65535.0f, fabsf(a.cy), fabsf(a.ky)), fmaf(65535.0f, fabsf(a.cz), fabsf(a.kz))));
  const float mi = fmaxf(fabsf(idx), fmaxf(fabsf(idy), fabsf(idz)));
  a.ok = (mi <= 1.0995116e12f) && (a.Mq <= 1.1529215e18f) && (r.time == r.time);    // false for NaN / inf / zero components
  if (!a.ok) a.Mq = __int_as_float(0x7f800000);          // the margin becomes infinite: every test goes to the exact path
  return a;
}

// -> +1 certain hit, -1 certain miss, 0 undecided
__device__ __forceinline__ int cbox_filter(const uint4& q, const RayAuxC& x, float tMin, float tCur) {
  const float lx = __uint2float_rn(q.x & 0xffffu), ly = __uint2float_rn(q.x >> 16), lz = __uint2float_rn(q.y & 0xffffu);
  const float hx = __uint2float_rn(q.y >> 16), hy = __uint2float_rn(q.z & 0xffffu), hz = __uint2float_rn(q.z >> 16);
  const float ax = fmaf(lx, x.cx, x.kx), bx = fmaf(hx, x.cx, x.kx);
  const float ay = fmaf(ly, x.cy, x.ky), by = fmaf(hy, x.cy, x.ky);
  const float az = fmaf(lz, x.cz, x.kz), bz = fmaf(hz, x.cz, x.kz);
  const float nx = fminf(ax, bx), ny = fminf(ay, by), nz = fminf(az, bz);
  const float fx = fmaxf(ax, bx), fy = fmaxf(ay, by), fz = fmaxf(az, bz);
  const float lo = fmaxf(fmaxf(nx, ny), fmaxf(nz, tMin)), ho = fminf(fminf(fx, fy), fminf(fz, tCur));
  const float ex = fabsf(x.cx), ey = fabsf(x.cy), ez = fabsf(x.cz);
  const float li = fmaxf(fmaxf(nx + ex, ny + ey), fmaxf(nz + ez, tMin)), hi = fminf(fminf(fx - ex, fy - ey), fminf(fz - ez, tCur));
  const float slack = fmaf(1.9073486e-6f, x.Mq + fabsf(lo) + fabsf(ho), 1e-30f);
  if ((ho - lo) < -slack) return -1;
  if ((hi - li) > slack) return 1;
  return 0;
}

#ifndef SRT_LPRIM_NO_ALLOCATE
#define SRT_LPRIM_NO_ALLOCATE 1
#endif
__device__ __forceinline__ float4 ldg_stream(const float4* p) {
  float4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  return v;
}
// one primitive of the leaf-ordered array against the ray (same decisions as tri_hit / sph_hit); `orig` = its id
__device__ __forceinline__ bool lprim_hit(const LPrim* __restrict__ lp, unsigned u, const Ray& r, float tMin, float tCur, float& tOut,
                                          int& orig, bool& is_sphere) {
#if SRT_LPRIM_NO_ALLOCATE
  // a leaf's record is used once by one lane: keep it out of the L1, which the tree's hot nodes need (config 5)
  const float4 A = ldg_stream(&lp[u].a), B = ldg_stream(&lp[u].b), C = ldg_stream(&lp[u].c), D = ldg_stream(&lp[u].d);
#else
  const float4 A = __ldg(&lp[u].a), B = __ldg(&lp[u].b), C = __ldg(&lp[u].c), D = __ldg(&lp[u].d);
#endif
  orig = __float_as_int(D.z);
  is_sphere = __float_as_int(D.w) != 0;
  if (is_sphere) {
    V3 oc = r.o - sph_center(A, B, C, r.time);
    float aa = length_squared(r.d);
    float halfB = dot(oc, r.d);
    float cc = length_squared(oc) - A.w * A.w;
    float disc = halfB * halfB - aa * cc;
    if (disc < 0.0f) return false;
    float sqrtd = sqrtf(disc);
    float root = (-halfB - sqrtd) / aa;
    if (root < tMin || root > tCur) {
      root = (-halfB + sqrtd) / aa;
      if (root < tMin || root > tCur) return false;
    }
    tOut = root;
    return true;
  }
  V3 N = mk(A.w, B.w, C.w);
  float NdotDir = dot(N, r.d);
  if (fabsf(NdotDir) < FLT_EPSILON) return false;
  if (NdotDir > 0) return false;
  float t = -(dot(N, r.o) + D.x) / NdotDir;
  if (t < tMin) return false;
  V3 v0 = mk(A.x, A.y, A.z), v1 = mk(B.x, B.y, B.z), v2 = mk(C.x, C.y, C.z);
  V3 p = mk(r.o.x + t * r.d.x, r.o.y + t * r.d.y, r.o.z + t * r.d.z);
  if (dot(N, cross(v1 - v0, p - v0)) < 0) return false;
  if (dot(N, cross(v2 - v1, p - v1)) < 0) return false;
  if (dot(N, cross(v0 - v2, p - v2)) < 0) return false;
  tOut = t;
  return true;
}

// hittableList::hit over the compact tree: the same walk as traverse() below (pre-order, running tCur, every accept
// overwrites), node decisions through cbox_filter with the exact fallback on the full node.
template <bool COUNT>
__device__ __forceinline__ void traverse_compact(const Scene& sc, const Ray& r, float tMin, float tMax, int& prim, float& tHit,
                                                 Counters* cnt) {
  const RayAuxC aux = make_aux_c(sc, r);
  float tCur = tMax;
  prim = -1;
  int i = 0;
  const int n = sc.n_nodes;
  while (i < n) {
    const uint4 q = __ldg(&sc.cnodes[i]);
    if (COUNT) ++cnt->nodes;
    bool hit = true;
    if (!(q.w & CNODE_NO_BOX) || (q.w & CNODE_INNER)) {
      const int f = cbox_filter(q, aux, tMin, tCur);
      if (f == 0) {
        if (COUNT) ++cnt->exact;
        hit = box_hit_exact(__ldg(&sc.nodes[i].lo), __ldg(&sc.nodes[i].hi), r, tMin, tCur);
      } else hit = f > 0;
    }
    if (q.w & CNODE_INNER) {
      i = hit ? i + 1 : (int)(q.w & 0x7fffffffu);
    } else {
      const unsigned u = q.w & CNODE_NO_PRIM;
      if (hit && u != CNODE_NO_PRIM) {
        const int cntp = (q.w & CNODE_TWO) ? 2 : 1;
#pragma unroll 1
        for (int k = 0; k < cntp; k++) {
          float t; int orig; bool sph;
          const bool acc = lprim_hit(sc.lprims, u + k, r, tMin, tCur, t, orig, sph);
          if (COUNT) { ++cnt->leaves; if (sph) ++cnt->sph_leaves; }
          if (acc) { tCur = t; prim = orig; }
        }
      }
      i = i + 1;
    }
  }
  tHit = tCur;
}

// hittableList::hit -> bvhNode::hit (hittablelist.h:33-47, bvh.h:97-105) in the threaded form.
// Rule: walk in pre-order; one running tCur = t of the most recently ACCEPTED leaf
// (initially tMax); boxes and spheres are tested against [tMin, tCur]; triangles ignore tCur; every
// accept overwrites the record.  `nodes` may point to shared memory.
template <bool COUNT>
__device__ __forceinline__ void traverse(const Scene& sc, const Node* nodes, const Ray& r, float tMin, float tMax,
                                         int& prim, float& tHit, Counters* cnt) {
  const RayAux aux = make_aux(r, sc.filter_ok);
  float tCur = tMax;
  prim = -1;
  int i = 0;
  const int n = sc.n_nodes;
  while (i < n) {
    const float4 lo = nodes[i].lo, hi = nodes[i].hi;
    const int w6 = __float_as_int(hi.z), w7 = __float_as_int(hi.w);
    if (COUNT) ++cnt->nodes;
    bool hit = (w7 == NODE_NO_BOX) ? true : box_hit<COUNT>(lo, hi, r, aux, tMin, tCur, COUNT ? &cnt->exact : nullptr);
    if (w7 == NODE_INNER) {
      i = hit ? i + 1 : w6;
    } else {
      if (hit) {
#pragma unroll 1
        for (int k = 0; k < 2; k++) {
          int p = k == 0 ? w6 : w7;
          if (p < 0) break;
          if (COUNT) { ++cnt->leaves; if (p >= sc.n_tri) ++cnt->sph_leaves; }
          float t;
          bool acc = p < sc.n_tri ? tri_hit(sc.tris, p, r, tMin, t) : sph_hit(sc.sphs, p - sc.n_tri, r, tMin, tCur, t);
          if (acc) { tCur = t; prim = p; }
        }
      }
      i = i + 1;
    }
  }
  tHit = tCur;
}

}  // namespace srt
