// srt_shade.cuh — hit record completion, textures, the four materials, Philox stream, samplers, camera.
// Reference: hittable.h:9-22, sphere.h:32-38,73-81,96-106, model.h:156-178,214-235, texture.h,
// material.h, pbr.h, camera.h:40-50.  Compiled with -fmad=false (see srt_core.cuh).
#pragma once
#ifndef SRT_SMALL_CODE
#define SRT_SMALL_CODE 1   /* sinf and philox as out-of-line functions: one copy each in the shade kernels' instruction stream (4.1 k -> 3.2 k
                              instructions, k_wf_shade2c 0.431 -> 0.412 ms per wave on B200); 0 = inlined everywhere */
#endif
#include "srt_core.cuh"
#include "../../include/srt.h"

namespace srt {

#if SRT_SMALL_CODE
__device__ __noinline__ float sinf_once(float v) { return sinf(v); }
#define SRT_SINF sinf_once
#if SRT_SMALL_CODE >= 2
#define SRT_TEXVAL_INLINE __device__ __noinline__
#else
#define SRT_TEXVAL_INLINE __device__ __forceinline__
#endif
#else
#define SRT_TEXVAL_INLINE __device__ __forceinline__
#define SRT_SINF sinf
#endif

struct Rec {
  V3 p, normal, tangent, bitangent;
  float u, v, t;
  int front_face, mat, prim;
};

// unitVector (vec3.h:54-60).  FAST = the render kernels: one MUFU.RSQ and three products instead of three IEEE
// divisions (<= 2 ulp from the reference's value; shading needs statistical parity only, and the 1e-5 bar on
// normals holds a fortiori).  The known-answer entry points (srt_trace_rays, srt_shade_kat) use the exact form.
template <bool FAST>
__device__ __forceinline__ V3 unit_v(V3 v) {
  if (!FAST) return unit_vector(v);
  const float l2 = length_squared(v);
  if (l2 == 0.0f) return v;
  const float inv = rsqrtf(l2);
  return mk(v.x * inv, v.y * inv, v.z * inv);
}

// hittable.h:18-21
__device__ __forceinline__ void set_face_normal(Rec& rec, const Ray& r, V3 outward) {
  rec.front_face = dot(r.d, outward) < 0;
  rec.normal = rec.front_face ? outward : -outward;
}

// The part of triangle::hit after the accept (model.h:156-178) + calcTangentBasis (model.h:214-235).
// Deferred to the one primitive whose record survives the traversal: every field is a function of
// (primitive, ray, t) only.
template <bool FAST>
__device__ __forceinline__ void finish_triangle(const Scene& sc, int i, const Ray& r, float t, bool want_uv, bool want_tbn,
                                                Rec& rec) {
  const float4 A = __ldg(&sc.tris[i].a), B = __ldg(&sc.tris[i].b), C = __ldg(&sc.tris[i].c), D = __ldg(&sc.tris[i].d);
  V3 v0 = mk(A.x, A.y, A.z), v1 = mk(B.x, B.y, B.z), v2 = mk(C.x, C.y, C.z), N = mk(A.w, B.w, C.w);
  rec.t = t;
  rec.p = mk(r.o.x + t * r.d.x, r.o.y + t * r.d.y, r.o.z + t * r.d.z);     // ray.at(t), ray.h:15-17
  set_face_normal(rec, r, unit_v<FAST>(N));
  rec.mat = __float_as_int(D.y);
  rec.prim = i;
  rec.u = rec.v = 0.0f;
  rec.tangent = rec.bitangent = mk(0, 0, 0);
  const bool smooth = sc.tri_vn != nullptr;
  const bool vtan = want_tbn && sc.tri_vt != nullptr;
  bool have_tbn = false;
  if (want_uv || want_tbn || smooth) {
    const float4 ua = __ldg(&sc.tri_uv[i].a), ub = __ldg(&sc.tri_uv[i].b);
    if (want_uv || smooth || vtan) {                                      // model.h:156-169
      V3 p = rec.p;
      float d0 = length(p - v0), d1 = length(p - v1), d2 = length(p - v2);
      float denom = (1.0f / d0) + (1.0f / d1) + (1.0f / d2);
      float r0 = (1.0f / d0) / denom, r1 = (1.0f / d1) / denom, r2 = (1.0f / d2) / denom;
      if (want_uv) {
        rec.u = r0 * ua.x + r1 * ua.z + r2 * ub.x;
        rec.v = 1.0f - (r0 * ua.y + r1 * ua.w + r2 * ub.y);
      }
      if (smooth) {       // not the reference (srt.h: tri_n0): the vertex normals blended with the same weights as the UVs
        const float4 n0 = __ldg(&sc.tri_vn[3 * (size_t)i]), n1 = __ldg(&sc.tri_vn[3 * (size_t)i + 1]), n2 = __ldg(&sc.tri_vn[3 * (size_t)i + 2]);
        const V3 s = unit_v<FAST>(mk(r0 * n0.x + r1 * n1.x + r2 * n2.x, r0 * n0.y + r1 * n1.y + r2 * n2.y, r0 * n0.z + r1 * n1.z + r2 * n2.z));
        if (s.x == s.x && s.y == s.y && s.z == s.z) rec.normal = rec.front_face ? s : -s;   // NaN weights (p on a vertex) keep the face normal
      }
      if (vtan) {         // not the reference (srt.h: tri_tan0): glTF vertex tangents, blended like the normals
        const float4 t0 = __ldg(&sc.tri_vt[3 * (size_t)i]), t1 = __ldg(&sc.tri_vt[3 * (size_t)i + 1]), t2 = __ldg(&sc.tri_vt[3 * (size_t)i + 2]);
        if (t0.w != 0.0f || t1.w != 0.0f || t2.w != 0.0f) {
          const V3 T = unit_v<FAST>(mk(r0 * t0.x + r1 * t1.x + r2 * t2.x, r0 * t0.y + r1 * t1.y + r2 * t2.y, r0 * t0.z + r1 * t1.z + r2 * t2.z));
          const V3 Bu = unit_v<FAST>(cross(rec.normal, T));
          const V3 Bt = mk(t0.w * Bu.x, t0.w * Bu.y, t0.w * Bu.z);
          const float chk = (T.x + T.y + T.z) + (Bt.x + Bt.y + Bt.z);
          if (chk == chk && fabsf(chk) <= 3.4e38f) { rec.tangent = T; rec.bitangent = Bt; have_tbn = true; }   // else: the per-face basis
        }
      }
    }
    if (want_tbn && !have_tbn) {                                          // model.h:214-235
      V3 e0 = v1 - v0, e1 = v2 - v0;
      float du0x = ua.z - ua.x, du0y = ua.w - ua.y, du1x = ub.x - ua.x, du1y = ub.y - ua.y;
      float f = (du0x * du1y - du1x * du0y);
      if (f == 0) f += FLT_EPSILON;
      f = 1.0f / f;
      rec.tangent = unit_v<FAST>(mk(f * (du1y * e0.x - du0y * e1.x), f * (du1y * e0.y - du0y * e1.y),
                                   f * (du1y * e0.z - du0y * e1.z)));
      rec.bitangent = unit_v<FAST>(mk(f * (-du1x * e0.x + du0x * e1.x), f * (-du1x * e0.y + du0x * e1.y),
                                     f * (-du1x * e0.z + du0x * e1.z)));
    }
  }
}

// The part of sphere::hit after the root is chosen (sphere.h:74-81), getSphereUV (:32-38),
// calcTangentBasis (:96-106)
template <bool FAST>
__device__ __forceinline__ void finish_sphere(const Scene& sc, int j, const Ray& r, float t, bool want_uv, bool want_tbn,
                                              Rec& rec) {
  const float4 a = __ldg(&sc.sphs[j].a), b = __ldg(&sc.sphs[j].b), c = __ldg(&sc.sphs[j].c);
  V3 ctr = sph_center(a, b, c, r.time);
  rec.t = t;
  rec.p = mk(r.o.x + t * r.d.x, r.o.y + t * r.d.y, r.o.z + t * r.d.z);
  V3 outward = unit_v<FAST>(rec.p - ctr);
  set_face_normal(rec, r, outward);
  rec.mat = __float_as_int(c.y);
  rec.prim = sc.n_tri + j;
  rec.u = rec.v = 0.0f;
  rec.tangent = rec.bitangent = mk(0, 0, 0);
  if (want_uv) {
    float theta = acosf(-outward.y);
    float phi = atan2f(-outward.z, outward.x) + SRT_PI;
    rec.u = phi / (2.0f * SRT_PI);
    rec.v = theta / SRT_PI;
  }
  if (want_tbn) {
    V3 bb;
    float ndoty = outward.x * 0.0f + (outward.y * 1.0f + outward.z * 0.0f);
    if (1.0f - fabsf(ndoty) < FLT_EPSILON) bb = mk(-0.0f, -0.0f, -1.0f);
    else bb = mk(0.0f, 1.0f, 0.0f);
    rec.tangent = unit_v<FAST>(cross(bb, outward));
    rec.bitangent = unit_v<FAST>(cross(outward, rec.tangent));
  }
}

template <bool FAST>
__device__ __forceinline__ void finish_hit(const Scene& sc, int prim, const Ray& r, float t, bool want_uv, bool want_tbn,
                                           Rec& rec) {
  if (prim < sc.n_tri) finish_triangle<FAST>(sc, prim, r, t, want_uv, want_tbn, rec);
  else finish_sphere<FAST>(sc, prim - sc.n_tri, r, t, want_uv, want_tbn, rec);
}

// ---- texture::value -----------------------------------------------------------------------------
// imagePNG::value / image3bpp::value (texture.h:129-148, :82-101): the reference's own index
// arithmetic, then one 4-byte texel load (read-only path).  The expanded texel
// (i, j) holds (px[0], px[1], px[2]) for px = data + j*bpp*w + i*bpp, which reproduces the
// "channel 1 of a 1-byte map is the next texel" behaviour of material.h:197.
__device__ __forceinline__ V3 image_value(const Tex& t, float u, float v) {
  if (t.missing) return mk(1.0f, 0.0f, 1.0f);                            // texture.h:130-131
  u = clampf(u, 0, 1.0f);
  v = 1.0f - clampf(v, 0, 1.0f);
  int i = (int)(u * t.w);      // cvt.rzi: NaN -> 0 (the reference reads out of bounds there)
  int j = (int)(v * t.h);
  if (i >= t.w) i = t.w - 1;
  if (j >= t.h) j = t.h - 1;
  const uchar4 px = __ldg(t.texels + (size_t)j * (size_t)t.w + (size_t)i);     // 0 <= i < w, 0 <= j < h by the lines above
  return mk((float)px.x, (float)px.y, (float)px.z);
}

SRT_TEXVAL_INLINE V3 texture_value(const Scene& sc, int id, float u, float v, V3 p) {
  const Tex* t = &sc.texs[id];
  int kind = __ldg(&t->kind);
  float scale = 1.0f;
  if (kind == SRT_TEX_CHECKER) {                                          // texture.h:42-48
    float sines = SRT_SINF(10.0f * p.x) * SRT_SINF(10.0f * p.y) * SRT_SINF(10.0f * p.z);
    t = &sc.texs[sines < 0 ? __ldg(&t->odd) : __ldg(&t->even)];
    kind = __ldg(&t->kind);
    scale = 255.0f;
  }
  V3 c;
  if (kind == SRT_TEX_SOLID) c = mk(__ldg(&t->r), __ldg(&t->g), __ldg(&t->b));   // texture.h:26-28
  else c = image_value(*t, u, v);
  return scale == 1.0f ? c : mk(c.x * scale, c.y * scale, c.z * scale);
}

// ---- materials ----------------------------------------------------------------------------------
// vec3.h:76-78
__device__ __forceinline__ V3 reflect(V3 v, V3 n) { float s = 2.0f * dot(v, n); return mk(v.x - s * n.x, v.y - s * n.y, v.z - s * n.z); }
// vec3.h:80-86
__device__ __forceinline__ V3 refract(V3 uv, V3 n, float eta) {
  float cosTheta = fminf(dot(n, -uv), 1.0f);
  V3 perp = mk(eta * (uv.x + cosTheta * n.x), eta * (uv.y + cosTheta * n.y), eta * (uv.z + cosTheta * n.z));
  float sp = -sqrtf(fabsf(1.0f - length_squared(perp)));
  return mk(perp.x + sp * n.x, perp.y + sp * n.y, perp.z + sp * n.z);
}

// material::emitted (material.h:18-20 default, :148-150 diffuseLight)
__device__ __forceinline__ V3 material_emitted(const Scene& sc, const Mat& m, const Rec& rec) {
  if (m.kind == SRT_MAT_LIGHT) return texture_value(sc, m.albedo_tex, rec.u, rec.v, rec.p);
  return mk(0, 0, 0);
}

// material::scatter.  rvec: unit vector (PBR) / point in the unit ball (METAL); xi: DIELECTRIC's draw.
template <bool FAST>
__device__ __forceinline__ bool material_scatter(const Scene& sc, const Mat& m, const Ray& rIn, const Rec& rec, V3 rvec,
                                                 float xi, V3& attenuation, Ray& out) {
  if (m.kind == SRT_MAT_LIGHT) return false;                              // material.h:144-146
  out.o = rec.p;
  out.time = rIn.time;
  if (m.kind == SRT_MAT_METAL) {                                          // material.h:91-97
    V3 reflected = reflect(unit_v<FAST>(rIn.d), rec.normal);
    out.d = mk(reflected.x + m.fuzz * rvec.x, reflected.y + m.fuzz * rvec.y, reflected.z + m.fuzz * rvec.z);
    attenuation = mk(m.albedo[0], m.albedo[1], m.albedo[2]);
    return dot(out.d, rec.normal) > 0;
  }
  if (m.kind == SRT_MAT_DIELECTRIC) {                                     // material.h:108-126,132-136
    attenuation = mk(1.0f, 1.0f, 1.0f);
    float ratio = rec.front_face ? (1.0f / m.ir) : m.ir;
    V3 unitDir = unit_v<FAST>(rIn.d);
    float cosTheta = fminf(dot(rec.normal, -unitDir), 1.0f);
    float sinTheta = sqrtf(1.0f - cosTheta * cosTheta);
    bool cannotRefract = ratio * sinTheta > 1.0f;
    float r0 = (1.0f - ratio) / (1.0f + ratio);
    r0 = r0 * r0;
    float c1 = 1.0f - cosTheta, c2 = c1 * c1;
    float reflectance = r0 + (1.0f - r0) * (c2 * c2 * c1);                // pow(1-cos, 5)
    out.d = (cannotRefract || reflectance > xi) ? reflect(unitDir, rec.normal) : refract(unitDir, rec.normal, ratio);
    return true;
  }
  // pbrMetallicRoughness::scatter, material.h:156-245 (+ pbr.h:58-81)
  V3 att, normal;
  float mm, rr;
  if (m.albedo_tex >= 0) {
    att = texture_value(sc, m.albedo_tex, rec.u, rec.v, rec.p);
    att = mk(att.x / 255.0f, att.y / 255.0f, att.z / 255.0f);
  } else att = mk(m.albedo[0], m.albedo[1], m.albedo[2]);
  if (m.normal_tex >= 0) {
    V3 n = texture_value(sc, m.normal_tex, rec.u, rec.v, rec.p);
    n = mk((n.x - 128.0f) / 128.0f, (n.y - 128.0f) / 128.0f, (n.z - 128.0f) / 128.0f);     // vec3.h:103-110
    const V3 T = rec.tangent, B = rec.bitangent, N = rec.normal;
    normal = unit_v<FAST>(mk(T.x * n.x + (B.x * n.y + N.x * n.z), T.y * n.x + (B.y * n.y + N.y * n.z),
                            T.z * n.x + (B.z * n.y + N.z * n.z)));
  } else normal = rec.normal;
  if ((m.flags & MAT_PACKED_MR) && m.metallic_tex >= 0) {     // glTF metallicRoughnessTexture, not the reference (srt.h)
    const V3 mr = texture_value(sc, m.metallic_tex, rec.u, rec.v, rec.p);
    mm = clampf(mr.z / 255.0f * m.metalness, 0, 1.0f);
    rr = clampf(mr.y / 255.0f * m.roughness, 0, 1.0f);
  } else {
    if (m.metallic_tex >= 0) mm = clampf(texture_value(sc, m.metallic_tex, rec.u, rec.v, rec.p).x / 255.0f, 0, 1.0f);
    else mm = m.metalness;
    if (m.roughness_tex >= 0) rr = clampf(texture_value(sc, m.roughness_tex, rec.u, rec.v, rec.p).y / 255.0f, 0, 1.0f);
    else rr = m.roughness;
  }

  V3 sd = normal + rvec;
  // vec3.h:49-52 compares fabs() < 1e-8 in double; (double)x < 1e-8  <=>  x <= fl32(1e-8)
  if (fabsf(sd.x) <= 1e-8f && fabsf(sd.y) <= 1e-8f && fabsf(sd.z) <= 1e-8f) sd = normal;
  sd = unit_v<FAST>(sd);
  out.d = sd;
  V3 viewVec = -unit_v<FAST>(rIn.d);
  V3 halfVec = unit_v<FAST>(sd + viewVec);
  float NdotL = fmaxf(dot(normal, sd), 0.0f);
  float NdotH = fmaxf(dot(normal, halfVec), 0.0f);
  float HdotV = fmaxf(dot(halfVec, viewVec), 0.0f);
  float NdotV = fmaxf(dot(normal, viewVec), 0.0f);
  float alpha = rr * rr, alpha2 = alpha * alpha;                                             // pbr.h:58-65
  float dq = (NdotH * NdotH) * (alpha2 - 1.0f) + 1.0f;
  float D = alpha2 / (SRT_PI * (dq * dq));
  float power = exp2f((-5.55473f * HdotV - 6.98316f) * HdotV);                               // pbr.h:75-81
  float k = ((rr + 1.0f) * (rr + 1.0f)) / 8.0f;                                              // pbr.h:69-73
  float G = (NdotL / (NdotL * (1.0f - k) + k)) * (NdotV / (NdotV * (1.0f - k) + k));
  float spec_den = 4.0f * NdotV * NdotL + FLT_EPSILON;
  float a3[3] = {att.x, att.y, att.z}, res[3];
#pragma unroll
  for (int c = 0; c < 3; c++) {
    float F0 = (1.0f - mm) * 0.4f + mm * m.albedo[c];                                        // material.h:228
    float F = F0 + (1.0f - F0) * power;
    float fd = a3[c] / SRT_PI;
    fd *= (1.0f - F);
    fd *= (1.0f - mm);
    fd *= m.albedo[c];
    float fs = D * F * G / spec_den;
    res[c] = (fd + fs) * NdotL;
  }
  attenuation = mk(res[0], res[1], res[2]);
  return true;
}

// ---- random numbers (new: per-(pixel, sample) Philox4x32-10; DESIGN.md "random numbers") --------
__device__ __forceinline__ uint4 philox(uint64_t seed, uint32_t pixel, uint32_t sample, uint32_t block) {
  uint32_t c0 = pixel, c1 = sample, c2 = block, c3 = 0;
  uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
#pragma unroll
  for (int round = 0; round < 10; round++) {
    uint32_t h0 = __umulhi(0xD2511F53u, c0), l0 = 0xD2511F53u * c0;
    uint32_t h1 = __umulhi(0xCD9E8D57u, c2), l1 = 0xCD9E8D57u * c2;
    c0 = h1 ^ c1 ^ k0; c1 = l1; c2 = h0 ^ c3 ^ k1; c3 = l0;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  return make_uint4(c0, c1, c2, c3);
}
// one out-of-line copy for the wavefront's shade kernels (three call sites: the kernel is bound by its instruction cache,
// not by this call); the megakernels keep the inlined form (config 2: 7 % faster that way)
__device__ __noinline__ uint4 philox_ool(uint64_t seed, uint32_t pixel, uint32_t sample, uint32_t block) { return philox(seed, pixel, sample, block); }
template <bool SMALL>
__device__ __forceinline__ uint4 philox_sel(uint64_t seed, uint32_t pixel, uint32_t sample, uint32_t block) {
  return (SMALL && SRT_SMALL_CODE) ? philox_ool(seed, pixel, sample, block) : philox(seed, pixel, sample, block);
}
__device__ __forceinline__ float u01(uint32_t x) { return (float)(x >> 8) * (1.0f / 16777216.0f); }

// same distributions as randomInUnitDisk / randomUnitVector / randomInUnitSphere (vec3.h:62-74,88-95),
// sampled directly instead of by rejection (no data-dependent loop inside a warp)
__device__ __forceinline__ void sample_disk(float xa, float xb, float& dx, float& dy) {
  float r = sqrtf(xa), phi = (2.0f * SRT_PI) * xb, s, c;
  __sincosf(phi, &s, &c);         // phi in [0, 2 pi): absolute error 2^-21, no range-reduction slow path
  dx = r * c; dy = r * s;
}
__device__ __forceinline__ V3 sample_unit_vector(float xa, float xb) {
  float z = 1.0f - 2.0f * xa;
  float rxy = sqrtf(fmaxf(0.0f, 1.0f - z * z)), phi = (2.0f * SRT_PI) * xb, s, c;
  __sincosf(phi, &s, &c);
  return mk(rxy * c, rxy * s, z);
}
__device__ __forceinline__ V3 sample_unit_ball(float xa, float xb, float xc) {
  V3 d = sample_unit_vector(xa, xb);
  float r = cbrtf(xc);
  return mk(r * d.x, r * d.y, r * d.z);
}

// camera::getRay, camera.h:40-50
__device__ __forceinline__ Ray camera_get_ray(const srt_camera& c, float s, float t, float diskx, float disky, float xiTime) {
  float rdx = c.lens_radius * diskx, rdy = c.lens_radius * disky;
  V3 offset = mk(rdx * c.hor[0] + rdy * c.vert[0], rdx * c.hor[1] + rdy * c.vert[1], rdx * c.hor[2] + rdy * c.vert[2]);
  Ray r;
  r.o = mk(c.origin[0] + offset.x, c.origin[1] + offset.y, c.origin[2] + offset.z);
  r.d = mk(c.lleft[0] + s * c.horizontal[0] + t * c.vertical[0] - c.origin[0] - offset.x,
           c.lleft[1] + s * c.horizontal[1] + t * c.vertical[1] - c.origin[1] - offset.y,
           c.lleft[2] + s * c.horizontal[2] + t * c.vertical[2] - c.origin[2] - offset.z);
  r.time = c.t0 + (c.t1 - c.t0) * xiTime;
  return r;
}

}  // namespace srt
