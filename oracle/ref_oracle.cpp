// ref_oracle.cpp — TEST INFRASTRUCTURE, not product code.
//
// The reference's own CPU implementation of the path-tracing path, compiled from the headers where
// they lie under /root/reference (nothing is copied): globals.h color.h hittablelist.h sphere.h
// camera.h material.h bvh.h model.h — everything main.cpp:11-18 includes except gl.h (needs a
// generated glad header + GLFW; no graphics stack on this box).  Built by oracle/Makefile into
// oracle/_ref/libref_oracle.so together with the reference's globals.cpp (stb / cgltf bodies).
//
// main.cpp itself cannot be compiled (it includes gl.h and defines main), so its driver parts are
// restated here, behaviour for behaviour: rayColor (main.cpp:33-52), the scene blocks
// (main.cpp:72-76,82-146 and the commented Shirley block main.cpp:92-122), the camera call
// (main.cpp:164-172) and the pixel/sample loop (main.cpp:200-227).
//
// Two deliberate hooks, neither of which changes what the reference computes:
//  * the std::mt19937 inside randomFloat() (globals.h:30-35) is redirected, by a macro that is live
//    only while globals.h is included, to a forwarding engine over one process-wide std::mt19937
//    that this file can reseed, and whose next outputs it can inject (known-answer tests);
//    default-constructed it yields exactly the reference's sequence (seed 5489);
//  * `private`/`protected` are opened while the first-party headers are included so the exporter
//    can read camera.h:52-62, model.h:45-47, texture.h:30-31,150-153 and material.h:152-153.
//
// Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline / --impl reference legs may load
// this library.  It must never be linked into the product (it reuses the reference's class names).

#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <iostream>
#include <limits>
#include <map>
#include <memory>
#include <random>
#include <string>
#include <vector>
#include <algorithm>
#include <unistd.h>

#include "Eigen/Core"
#include "Eigen/Dense"
#include "Eigen/Geometry"

#include "stb_image.h"
#include "stb_image_write.h"
#include "cgltf.h"

#include "../include/srt.h"

// ---------------------------------------------------------------- RNG hook
static std::mt19937         g_engine;   // default seed 5489, same as globals.h:32
static std::deque<uint32_t> g_inject;   // raw 32-bit outputs served before the engine's own
static uint64_t             g_draws = 0;

namespace std {
struct srt_hooked_mt19937 {
  typedef uint32_t result_type;
  static constexpr result_type min() { return std::mt19937::min(); }
  static constexpr result_type max() { return std::mt19937::max(); }
  result_type operator()() {
    ++g_draws;
    if (!g_inject.empty()) { result_type v = g_inject.front(); g_inject.pop_front(); return v; }
    return g_engine();
  }
};
}  // namespace std

#define mt19937 srt_hooked_mt19937
#include "globals.h"
#undef mt19937

#define private public
#define protected public
#include "color.h"
#include "hittablelist.h"
#include "sphere.h"
#include "camera.h"
#include "material.h"
#include "bvh.h"
#include "model.h"
#undef private
#undef protected

// ---------------------------------------------------------------- primitive ids
// hitRecord has no id field (hittable.h:9-17); a forwarding decorator records it on accept.  The
// last accept is the one whose record survives (every accept overwrites), so g_last_id after
// world.hit() is the reference's answer.
static int g_last_id = -1;
static uint64_t g_leaf_tests = 0, g_box_tests = 0;

struct tagged : public hittable {
  shared_ptr<hittable> inner;
  int id;
  tagged(shared_ptr<hittable> h, int i) : inner(h), id(i) {}
  bool hit(const ray& r, float tMin, float tMax, hitRecord& rec) const override {
    ++g_leaf_tests;
    if (inner->hit(r, tMin, tMax, rec)) { g_last_id = id; return true; }
    return false;
  }
  bool boundingBox(float a, float b, aabb& box) const override { return inner->boundingBox(a, b, box); }
  void calcTangentBasis(const vec3f&, vec3f&, vec3f&) const override {}
  int  populateVector(shared_ptr<hittableVector>) const override { return -1; }
};

// texture living in memory instead of a file (imagePNG::value, texture.h:129-148, stays the
// reference's own code)
struct memPNG : public imagePNG {
  memPNG(const uint8_t* src, int w, int h, int bytesPP) : imagePNG(bytesPP) {
    width = w; height = h; bytesPerScanline = bytesPP * w;
    size_t n = (size_t)w * h * bytesPP;
    data = static_cast<uint8_t*>(::operator new(n + 4));   // +4: the reference reads px[1..2] past a 1-bpp map
    memcpy(data, src, n); memset(data + n, 0, 4);
  }
};

struct RefScene {
  hittableList                          world;      // one element: the BVH root (main.cpp:146)
  std::vector<shared_ptr<hittable>>     prims;      // untagged, id order
  std::vector<int>                      primMat;
  std::vector<shared_ptr<material>>     mats;
  std::vector<shared_ptr<texture>>      texs;
  std::vector<shared_ptr<model>>        models;     // keep alive
  std::vector<shared_ptr<mesh>>         meshes;
  int64_t nTri = 0, nSph = 0;
};

static int texIndex(RefScene* s, const shared_ptr<texture>& t) {
  if (!t) return -1;
  for (size_t i = 0; i < s->texs.size(); i++) if (s->texs[i].get() == t.get()) return (int)i;
  // children first so checker's even/odd ids are smaller than its own
  if (auto c = std::dynamic_pointer_cast<checker>(t)) { texIndex(s, c->even); texIndex(s, c->odd); }
  s->texs.push_back(t);
  return (int)s->texs.size() - 1;
}

static int matIndex(RefScene* s, const shared_ptr<material>& m) {
  for (size_t i = 0; i < s->mats.size(); i++) if (s->mats[i].get() == m.get()) return (int)i;
  s->mats.push_back(m);
  if (auto p = std::dynamic_pointer_cast<pbrMetallicRoughness>(m)) {
    texIndex(s, p->albedoMap); texIndex(s, p->normalMap); texIndex(s, p->metallicMap); texIndex(s, p->roughnessMap);
  } else if (auto l = std::dynamic_pointer_cast<diffuseLight>(m)) {
    texIndex(s, l->emit);
  }
  return (int)s->mats.size() - 1;
}

static void addPrim(RefScene* s, hittableList& objects, shared_ptr<hittable> h, shared_ptr<material> m) {
  int id = (int)s->prims.size();
  s->prims.push_back(h);
  s->primMat.push_back(matIndex(s, m));
  objects.add(make_shared<tagged>(h, id));
}

static void addModel(RefScene* s, hittableList& objects, const char* file) {
  auto mdl = model::create(file);
  if (!mdl->init()) { std::cerr << "ref_oracle: cannot load " << file << "\n"; return; }
  s->models.push_back(mdl);
  for (const auto& msh : mdl->meshes)                // main.cpp:82-86
    for (const auto& tri : msh->triangles) { addPrim(s, objects, tri, msh->matPtr); s->nTri++; }
}

static void addSphere(RefScene* s, hittableList& objects, vec3f c0, vec3f c1, float t0, float t1, float r,
                      shared_ptr<material> m) {
  addPrim(s, objects, make_shared<sphere>(c0, c1, t0, t1, r, m), m);
  s->nSph++;
}

static shared_ptr<pbrMetallicRoughness> pinned(shared_ptr<pbrMetallicRoughness> p, float m, float r) {
  // material.h:25-40 leave metalness/roughness/anisotropy uninitialised; pin them
  p->metalness = m; p->roughness = r; p->anisotropy = 0;
  return p;
}

// ---------------------------------------------------------------- native scenes (reference classes)
static void groundAndLight(RefScene* s, hittableList& objects) {
  auto checkerTex = make_shared<checker>(color3f(0.2f, 0.3f, 0.1f), color3f(0.9f, 0.9f, 0.9f));   // main.cpp:89
  addSphere(s, objects, vec3f(0, -1000, 0.0f), vec3f(0, -1000, 0.0f), 0, 1.0f, 1000,
            pinned(make_shared<pbrMetallicRoughness>(checkerTex), 0, 0));                          // main.cpp:90
  auto lightMat = make_shared<diffuseLight>(color3f(250.2f, 220.9f, 110.2f));                     // main.cpp:126
  addSphere(s, objects, vec3f(-7.0f, 4.0f, 6.0f), vec3f(-7.0f, 4.0f, 6.0f), 0, 1.0f, 1.0f, lightMat);  // :127
}

// which: 1 = Shirley sphere scene (main.cpp:92-122 block re-enabled + main.cpp:124-131,143-144),
//        2 = cube.gltf + square.gltf + ground + light,
//        3 = randomScene() default branch (main.cpp:72-76,82-90,126-146),
//        4 = as 3 but the four globe maps point at files that exist here
static RefScene* buildNative(int which, const char* dataDir) {
  // model.h:395-412 hard-codes "../data/": run from a directory whose ../data is the asset dir
  char cwd[4096]; if (!getcwd(cwd, sizeof cwd)) cwd[0] = 0;
  if (chdir(dataDir) != 0) { std::cerr << "ref_oracle: cannot chdir to " << dataDir << "\n"; return nullptr; }

  g_engine = std::mt19937(); g_inject.clear();       // the reference's start state
  RefScene* s = new RefScene;
  hittableList objects;

  if (which == 1) {
    auto checkerTex = make_shared<checker>(color3f(0.2f, 0.3f, 0.1f), color3f(0.9f, 0.9f, 0.9f));
    addSphere(s, objects, vec3f(0, -1000, 0.0f), vec3f(0, -1000, 0.0f), 0, 1.0f, 1000,
              pinned(make_shared<pbrMetallicRoughness>(checkerTex), 0, 0));
    for (int a = -11; a < 11; a++) {                                                  // main.cpp:92-122
      for (int b = -11; b < 11; b++) {
        auto choose_mat = randomFloat();
        vec3f center(a + 0.9*randomFloat(), 0.2f, b + 0.9f*randomFloat());
        if (length(center - vec3f(4, 0.2, 0)) > 0.9) {
          if (choose_mat < 0.8) {
            color3f albedo(randomFloat() * randomFloat(),
                           randomFloat() * randomFloat(),
                           randomFloat() * randomFloat());
            // pbrMetallicRoughness(color3f) is near-black in the reference (albedo/255);
            // the diffuse role is a solidColor scaled to the 0..255 convention, m = 0, r = 0.5
            color3f a255 = albedo * 255.0f;
            auto mat = pinned(make_shared<pbrMetallicRoughness>(make_shared<solidColor>(a255)), 0, 0.5f);
            vec3f center2 = center + vec3f(0, randomFloat(0, 0.5f), 0);
            addSphere(s, objects, center, center2, 0, 1.0f, 0.2f, mat);
          } else if (choose_mat < 0.95f) {
            auto albedo = randomVec3f(0.5f, 1);
            auto fuzz = randomFloat(0, 0.5f);
            addSphere(s, objects, center, center, 0, 1.0f, 0.2f, make_shared<metal>(albedo, fuzz));
          } else {
            addSphere(s, objects, center, center, 0, 1.0f, 0.2f, make_shared<dielectric>(1.5f));
          }
        }
      }
    }
    addSphere(s, objects, vec3f(0, 1, 0), vec3f(0, 1, 0), 0, 1.0f, 1.0f, make_shared<dielectric>(1.5));  // :124-125
    color3f brown(0.4f * 255.0f, 0.2f * 255.0f, 0.1f * 255.0f);
    addSphere(s, objects, vec3f(-4, 1, 0), vec3f(-4, 1, 0), 0, 1.0f, 1.0f,
              pinned(make_shared<pbrMetallicRoughness>(make_shared<solidColor>(brown)), 0, 0.5f));        // :129-130
    addSphere(s, objects, vec3f(4.0f, 1.0f, 0), vec3f(4.0f, 1.0f, 0), 0, 1.0f, 1.0f,
              make_shared<metal>(color3f(0.7, 0.6, 0.5), 0.0));                                          // :143-144
  } else if (which == 2) {
    addModel(s, objects, "../data/cube.gltf");
    addModel(s, objects, "../data/square.gltf");
    groundAndLight(s, objects);
  } else {
    addModel(s, objects, "../data/masterchief2-separate-xf.gltf");                    // main.cpp:74-75
    groundAndLight(s, objects);
    shared_ptr<imagePNG> ironAlbedo, ironNMap, ironMMap, ironRMap;
    if (which == 3) {                                                                 // main.cpp:133-136
      ironAlbedo = make_shared<imagePNG>("../data/rustediron2_basecolor-2x1.png", 3);
      ironNMap = make_shared<imagePNG>("../data/rustediron2_normal-2x1.png", 3);
      ironMMap = make_shared<imagePNG>("../data/rustediron2_metallic-2x1.png", 1);
      ironRMap = make_shared<imagePNG>("../data/rustediron2_roughness-2x1.png", 1);
    } else {
      ironAlbedo = make_shared<imagePNG>("../data/Image_1.png", 3);
      ironNMap = make_shared<imagePNG>("../data/Image_0.png", 3);
      ironMMap = make_shared<imagePNG>("../data/worldisyoursbanner.png", 1);
      ironRMap = make_shared<imagePNG>("../data/worldisyoursbanner.png", 1);
    }
    auto ironMat = make_shared<pbrMetallicRoughness>(ironAlbedo, ironNMap, ironMMap, ironRMap,
                                                     vec4f(1.0f, 1.0f, 1.0f, 1.0f));  // main.cpp:137-139
    addSphere(s, objects, vec3f(-3.0f, 1.0f, 0.0f), vec3f(-3.0f, 1.0f, 0.0f), 0, 1.0f, 1.0f, ironMat);  // :140
    addSphere(s, objects, vec3f(3.0f, 1.0f, 0), vec3f(3.0f, 1.0f, 0), 0, 1.0f, 1.0f,
              make_shared<metal>(color3f(0.7, 0.6, 0.5), 0.0));                       // main.cpp:143-144
  }
  s->world.add(make_shared<bvhNode>(objects, 0, 1));                                  // main.cpp:146
  if (chdir(cwd) != 0) {}
  return s;
}

// ---------------------------------------------------------------- scene from a flattened description
// Rebuilds reference objects (mesh/triangle/sphere/material/texture) from srt_scene_desc and installs
// the given tree through bvhNode's public default ctor + left/right/box (bvh.h:14,28-31), so the very
// same tree is traversed by bvhNode::hit.  Used where /root/reference/data is absent (GPU box).
static RefScene* buildFromDesc(const srt_scene_desc* d) {
  RefScene* s = new RefScene;
  for (int i = 0; i < d->n_tex; i++) {
    const srt_texture& t = d->textures[i];
    if (t.kind == SRT_TEX_SOLID) s->texs.push_back(make_shared<solidColor>(color3f(t.color[0], t.color[1], t.color[2])));
    else if (t.kind == SRT_TEX_CHECKER) s->texs.push_back(make_shared<checker>(s->texs.at(t.even), s->texs.at(t.odd)));
    else if (t.data) s->texs.push_back(make_shared<memPNG>(t.data, t.width, t.height, t.bpp));
    else s->texs.push_back(make_shared<imagePNG>(t.bpp));
  }
  auto tex = [&](int id) { return id < 0 ? shared_ptr<texture>() : s->texs.at(id); };
  for (int i = 0; i < d->n_mat; i++) {
    const srt_material& m = d->materials[i];
    if (m.kind == SRT_MAT_PBR) {
      auto p = make_shared<pbrMetallicRoughness>(tex(m.albedo_tex), tex(m.normal_tex), tex(m.metallic_tex),
                                                 tex(m.roughness_tex),
                                                 vec4f(m.albedo[0], m.albedo[1], m.albedo[2], m.albedo[3]));
      s->mats.push_back(pinned(p, m.metalness, m.roughness));
    } else if (m.kind == SRT_MAT_METAL) {
      s->mats.push_back(make_shared<metal>(color3f(m.albedo[0], m.albedo[1], m.albedo[2]), m.fuzz));
    } else if (m.kind == SRT_MAT_DIELECTRIC) {
      s->mats.push_back(make_shared<dielectric>(m.ir));
    } else {
      s->mats.push_back(make_shared<diffuseLight>(tex(m.albedo_tex)));
    }
  }
  // triangles: consecutive runs of one material become a mesh of <= 21845 triangles (uint16 indices, model.h:46)
  shared_ptr<mesh> cur; int curMat = -2;
  for (int64_t i = 0; i < d->n_tri; i++) {
    if (!cur || d->tri_mat[i] != curMat || cur->positions.size() + 3 > 65535) {
      cur = mesh::create(); curMat = d->tri_mat[i]; cur->matPtr = s->mats.at(curMat); s->meshes.push_back(cur);
    }
    uint16_t base = (uint16_t)cur->positions.size();
    const float* v[3] = {d->tri_v0 + 3 * i, d->tri_v1 + 3 * i, d->tri_v2 + 3 * i};
    const float* uv[3] = {d->tri_uv0 + 2 * i, d->tri_uv1 + 2 * i, d->tri_uv2 + 2 * i};
    for (int k = 0; k < 3; k++) {
      cur->positions.push_back(vec3f(v[k][0], v[k][1], v[k][2]));
      cur->texcoords.push_back(vec2f(uv[k][0], uv[k][1]));
    }
    auto tri = triangle::create(base, base + 1, base + 2, cur);
    cur->triangles.push_back(tri);
    s->prims.push_back(tri); s->primMat.push_back(curMat);
  }
  s->nTri = d->n_tri;
  for (int64_t i = 0; i < d->n_sph; i++) {
    const float* a = d->sph_c0 + 3 * i; const float* b = d->sph_c1 + 3 * i;
    s->prims.push_back(make_shared<sphere>(vec3f(a[0], a[1], a[2]), vec3f(b[0], b[1], b[2]), d->sph_t0[i], d->sph_t1[i],
                                           d->sph_r[i], s->mats.at(d->sph_mat[i])));
    s->primMat.push_back(d->sph_mat[i]);
  }
  s->nSph = d->n_sph;
  std::vector<shared_ptr<hittable>> tags(s->prims.size());
  for (size_t i = 0; i < tags.size(); i++) tags[i] = make_shared<tagged>(s->prims[i], (int)i);
  std::vector<shared_ptr<bvhNode>> nodes(d->n_nodes);
  for (int64_t i = 0; i < d->n_nodes; i++) nodes[i] = make_shared<bvhNode>();
  for (int64_t i = 0; i < d->n_nodes; i++) {
    const srt_bvh_node& n = d->nodes[i];
    nodes[i]->box = aabb(vec3f(n.bmin[0], n.bmin[1], n.bmin[2]), vec3f(n.bmax[0], n.bmax[1], n.bmax[2]));
    nodes[i]->left = n.left >= 0 ? shared_ptr<hittable>(nodes.at(n.left)) : tags.at(~n.left);
    if (n.right == SRT_BVH_NONE) nodes[i]->right = nodes[i]->left;
    else nodes[i]->right = n.right >= 0 ? shared_ptr<hittable>(nodes.at(n.right)) : tags.at(~n.right);
  }
  if (d->n_nodes > 0) s->world.add(nodes[0]);
  return s;
}

// ---------------------------------------------------------------- export (flatten)
static void flattenNode(const hittable* h, std::vector<srt_bvh_node>& out, int32_t* slot) {
  if (auto t = dynamic_cast<const tagged*>(h)) { *slot = ~t->id; return; }
  auto n = dynamic_cast<const bvhNode*>(h);
  int32_t idx = (int32_t)out.size();
  *slot = idx;
  out.emplace_back();
  for (int k = 0; k < 3; k++) { out[idx].bmin[k] = n->box.minimum(k); out[idx].bmax[k] = n->box.maximum(k); }
  int32_t l, r;
  flattenNode(n->left.get(), out, &l);                 // pre-order, left first (bvh.h:112-148)
  if (n->left.get() == n->right.get()) r = SRT_BVH_NONE;
  else flattenNode(n->right.get(), out, &r);
  out[idx].left = l; out[idx].right = r;
}

extern "C" {

void* ref_scene_native(int which, const char* dataDir) { return buildNative(which, dataDir); }
void* ref_scene_from_desc(const srt_scene_desc* d) { return buildFromDesc(d); }
void  ref_scene_free(void* p) { delete static_cast<RefScene*>(p); }

void ref_scene_sizes(void* p, int64_t* out /*n_tri,n_sph,n_nodes,n_mat,n_tex*/) {
  RefScene* s = static_cast<RefScene*>(p);
  std::vector<srt_bvh_node> nodes; int32_t root;
  flattenNode(s->world.objects[0].get(), nodes, &root);
  out[0] = s->nTri; out[1] = s->nSph; out[2] = (int64_t)nodes.size(); out[3] = (int64_t)s->mats.size();
  out[4] = (int64_t)s->texs.size();
}

// texture data pointers point into scene-owned memory (valid until ref_scene_free)
void ref_scene_export(void* p, float* v0, float* v1, float* v2, float* uv0, float* uv1, float* uv2, int32_t* triMat,
                      float* c0, float* c1, float* t0, float* t1, float* rad, int32_t* sphMat,
                      srt_bvh_node* nodesOut, srt_material* mats, srt_texture* texs) {
  RefScene* s = static_cast<RefScene*>(p);
  for (int64_t i = 0; i < s->nTri; i++) {
    auto tri = std::dynamic_pointer_cast<triangle>(s->prims[i]);
    float* vv[3] = {v0 + 3 * i, v1 + 3 * i, v2 + 3 * i}; float* uu[3] = {uv0 + 2 * i, uv1 + 2 * i, uv2 + 2 * i};
    for (int k = 0; k < 3; k++) {
      const vec3f& P = tri->parentMesh->positions[tri->vertices[k]];
      const vec2f& T = tri->parentMesh->texcoords[tri->vertices[k]];
      vv[k][0] = P(0); vv[k][1] = P(1); vv[k][2] = P(2); uu[k][0] = T(0); uu[k][1] = T(1);
    }
    triMat[i] = s->primMat[i];
  }
  for (int64_t j = 0; j < s->nSph; j++) {
    auto sp = std::dynamic_pointer_cast<sphere>(s->prims[s->nTri + j]);
    for (int k = 0; k < 3; k++) { c0[3 * j + k] = sp->center0(k); c1[3 * j + k] = sp->center1(k); }
    t0[j] = sp->t0; t1[j] = sp->t1; rad[j] = sp->radius; sphMat[j] = s->primMat[s->nTri + j];
  }
  std::vector<srt_bvh_node> nodes; int32_t root;
  flattenNode(s->world.objects[0].get(), nodes, &root);
  memcpy(nodesOut, nodes.data(), nodes.size() * sizeof(srt_bvh_node));
  auto tid = [&](const shared_ptr<texture>& t) { return texIndex(s, t); };
  for (size_t i = 0; i < s->mats.size(); i++) {
    srt_material m; memset(&m, 0, sizeof m);
    m.albedo_tex = m.normal_tex = m.metallic_tex = m.roughness_tex = -1;
    if (auto q = std::dynamic_pointer_cast<pbrMetallicRoughness>(s->mats[i])) {
      m.kind = SRT_MAT_PBR; m.albedo_tex = tid(q->albedoMap); m.normal_tex = tid(q->normalMap);
      m.metallic_tex = tid(q->metallicMap); m.roughness_tex = tid(q->roughnessMap);
      for (int k = 0; k < 4; k++) m.albedo[k] = q->albedo(k);
      m.metalness = q->metalness; m.roughness = q->roughness;
    } else if (auto q = std::dynamic_pointer_cast<metal>(s->mats[i])) {
      m.kind = SRT_MAT_METAL; for (int k = 0; k < 3; k++) m.albedo[k] = q->albedo(k); m.fuzz = q->fuzz;
    } else if (auto q = std::dynamic_pointer_cast<dielectric>(s->mats[i])) {
      m.kind = SRT_MAT_DIELECTRIC; m.ir = q->ir;
    } else if (auto q = std::dynamic_pointer_cast<diffuseLight>(s->mats[i])) {
      m.kind = SRT_MAT_LIGHT; m.albedo_tex = tid(q->emit);
    }
    mats[i] = m;
  }
  for (size_t i = 0; i < s->texs.size(); i++) {
    srt_texture t; memset(&t, 0, sizeof t); t.even = t.odd = -1;
    if (auto q = std::dynamic_pointer_cast<solidColor>(s->texs[i])) {
      t.kind = SRT_TEX_SOLID; for (int k = 0; k < 3; k++) t.color[k] = q->colorValue(k);
    } else if (auto q = std::dynamic_pointer_cast<checker>(s->texs[i])) {
      t.kind = SRT_TEX_CHECKER; t.even = tid(q->even); t.odd = tid(q->odd);
    } else if (auto q = std::dynamic_pointer_cast<imagePNG>(s->texs[i])) {
      t.kind = SRT_TEX_IMAGE; t.width = q->width; t.height = q->height; t.bpp = q->bpp; t.data = q->data;
    }
    texs[i] = t;
  }
}

}  // extern "C"

// ---------------------------------------------------------------- the path itself
static uint64_t g_segments = 0;

// main.cpp:33-52, behaviour for behaviour (+ a segment counter before world.hit)
static color3f rayColor(const ray& r, const color3f& background, const hittable& world, int maxBounce) {
  hitRecord record;
  if (maxBounce <= 0) return color3f(0, 0, 0);
  ++g_segments;
  if (!world.hit(r, 0.001f, infinity, record)) return background;
  ray     scattered;
  color3f attenuation;
  color3f emitted = record.matPtr->emitted(record.uv(0), record.uv(1), record.p);
  if (!record.matPtr->scatter(r, record, attenuation, scattered)) return emitted;
  color3f newColor = rayColor(scattered, background, world, maxBounce - 1);
  newColor = color3f(newColor(0) * attenuation(0), newColor(1) * attenuation(1), newColor(2) * attenuation(2));
  return emitted + newColor;
}

static camera makeCamera(const float* c /*eye3 lookAt3 up3 vfov aspect aperture focus t0 t1*/) {
  return camera(vec3f(c[0], c[1], c[2]), vec3f(c[3], c[4], c[5]), vec3f(c[6], c[7], c[8]), c[9], c[10], c[11], c[12],
                c[13], c[14]);
}

static void fillHit(const RefScene* s, bool hit, const hitRecord& rec, srt_hit* h) {
  memset(h, 0, sizeof *h);
  if (!hit) { h->prim = -1; h->mat = -1; return; }
  h->prim = g_last_id; h->mat = s->primMat[g_last_id]; h->front_face = rec.frontFace ? 1 : 0; h->t = rec.t;
  for (int k = 0; k < 3; k++) {
    h->p[k] = rec.p(k); h->normal[k] = rec.normal(k); h->tangent[k] = rec.tangent(k); h->bitangent[k] = rec.bitangent(k);
  }
  h->uv[0] = rec.uv(0); h->uv[1] = rec.uv(1);
}

static hitRecord toRecord(const RefScene* s, const srt_hit& h) {
  hitRecord rec;
  rec.p = vec3f(h.p[0], h.p[1], h.p[2]); rec.normal = vec3f(h.normal[0], h.normal[1], h.normal[2]);
  rec.tangent = vec3f(h.tangent[0], h.tangent[1], h.tangent[2]);
  rec.bitangent = vec3f(h.bitangent[0], h.bitangent[1], h.bitangent[2]);
  rec.uv = vec2f(h.uv[0], h.uv[1]); rec.t = h.t; rec.frontFace = h.front_face != 0; rec.matPtr = s->mats.at(h.mat);
  return rec;
}

// raw 32-bit engine output that makes randomFloat() return exactly xi (xi a multiple of 2^-24 in [0,1)):
// libstdc++ generate_canonical<float,24> over a 32-bit engine is float(x) / 2^32
static uint32_t rawFor(float xi) { return (uint32_t)(xi * 16777216.0f) << 8; }

extern "C" {

void ref_rng_seed(uint32_t seed) { g_engine.seed(seed); g_inject.clear(); }
void ref_rng_reset(void) { g_engine = std::mt19937(); g_inject.clear(); }
void ref_rng_burn(uint64_t n) { g_engine.discard(n); }
float ref_random_float(void) { return randomFloat(); }
uint64_t ref_rng_draws(void) { return g_draws; }

void ref_camera(const float* c, srt_camera* out) {
  camera cam = makeCamera(c);
  for (int k = 0; k < 3; k++) {
    out->origin[k] = cam.origin(k); out->lleft[k] = cam.lleft(k); out->horizontal[k] = cam.horizontal(k);
    out->vertical[k] = cam.vertical(k); out->w[k] = cam.w(k); out->hor[k] = cam.hor(k); out->vert[k] = cam.vert(k);
  }
  out->lens_radius = cam.lensRadius; out->t0 = cam.t0; out->t1 = cam.t1;
}

// hittableList::hit(r, tMin, tMax, rec) on explicit rays (hittablelist.h:33-47 -> bvh.h:97-105)
void ref_trace_rays(void* p, const srt_ray* rays, int64_t n, float tMin, float tMax, srt_hit* out) {
  RefScene* s = static_cast<RefScene*>(p);
  for (int64_t i = 0; i < n; i++) {
    ray r(vec3f(rays[i].o[0], rays[i].o[1], rays[i].o[2]), vec3f(rays[i].dir[0], rays[i].dir[1], rays[i].dir[2]),
          rays[i].time);
    hitRecord rec; g_last_id = -1;
    bool hit = s->world.hit(r, tMin, tMax, rec);
    fillHit(s, hit, rec, &out[i]);
  }
}

// leaf tests (tagged::hit calls) and box tests for a ray set: box tests counted by re-walking the
// public tree with the reference's own aabb::hit and leaf ::hit (same recursion as bvh.h:97-105)
static bool countingHit(const hittable* h, const ray& r, float tMin, float tMax, hitRecord& rec) {
  auto n = dynamic_cast<const bvhNode*>(h);
  if (!n) return h->hit(r, tMin, tMax, rec);
  ++g_box_tests;
  if (!n->box.hit(r, tMin, tMax)) return false;
  bool hl = countingHit(n->left.get(), r, tMin, tMax, rec);
  bool hr = countingHit(n->right.get(), r, tMin, hl ? rec.t : tMax, rec);
  return hl || hr;
}
void ref_count_tests(void* p, const srt_ray* rays, int64_t n, float tMin, float tMax, uint64_t* boxTests,
                     uint64_t* leafTests) {
  RefScene* s = static_cast<RefScene*>(p);
  g_box_tests = 0; g_leaf_tests = 0;
  for (int64_t i = 0; i < n; i++) {
    ray r(vec3f(rays[i].o[0], rays[i].o[1], rays[i].o[2]), vec3f(rays[i].dir[0], rays[i].dir[1], rays[i].dir[2]),
          rays[i].time);
    hitRecord rec;
    countingHit(s->world.objects[0].get(), r, tMin, tMax, rec);
  }
  *boxTests = g_box_tests; *leafTests = g_leaf_tests;
}

// main.cpp:200-227 over rows [y0, y1).  sum: [H*W*3] float (rows outside the band untouched),
// rgba: [H*W*4] or NULL.  Returns the number of path segments (world.hit calls).
uint64_t ref_render(void* p, const float* camArgs, int W, int H, int spp, int maxBounce, const float* bg, int y0,
                    int y1, float* sum, uint8_t* rgba) {
  RefScene* s = static_cast<RefScene*>(p);
  camera mainCamera = makeCamera(camArgs);
  color3f background(bg[0], bg[1], bg[2]);
  g_segments = 0;
  for (int y = y0; y < y1; ++y) {
    for (int x = 0; x < W; ++x) {
      color3f pixelColor(0, 0, 0);
      for (int sidx = 0; sidx < spp; ++sidx) {
        auto u = float(x + randomFloat()) / (W - 1);                                  // main.cpp:210
        auto v = float((H - y) + randomFloat()) / (H - 1);                            // main.cpp:211
        ray r = mainCamera.getRay(u, v);
        pixelColor += rayColor(r, background, s->world, maxBounce);
      }
      if (sum) for (int k = 0; k < 3; k++) sum[((size_t)y * W + x) * 3 + k] = pixelColor(k);
      if (rgba) writeColorTarget(rgba, x, y, W, H, 4, pixelColor, spp);
    }
  }
  return g_segments;
}

// color.h:25-41 on a summed buffer
void ref_resolve(const float* sum, int W, int H, int spp, uint8_t* rgba) {
  for (int y = 0; y < H; y++) for (int x = 0; x < W; x++) {
    const float* c = sum + ((size_t)y * W + x) * 3;
    writeColorTarget(rgba, x, y, W, H, 4, color3f(c[0], c[1], c[2]), spp);
  }
}

// camera::getRay(s, t) (camera.h:40-50) with the disk point and the time draw injected:
// in[i*5] = (s, t, disk.x, disk.y, xiTime), disk components and xiTime multiples of 2^-24 such that
// disk = -1 + 2*xi is inside the unit disk (first rejection-loop iteration accepts).
void ref_getray_kat(const float* camArgs, const float* in, int64_t n, srt_ray* out) {
  camera cam = makeCamera(camArgs);
  for (int64_t i = 0; i < n; i++) {
    const float* q = in + 5 * i;
    g_inject.clear();
    // vec3f(randomFloat(-1,1), randomFloat(-1,1), 0): g++ evaluates ctor arguments right to left -> y first
    g_inject.push_back(rawFor(q[3])); g_inject.push_back(rawFor(q[2])); g_inject.push_back(rawFor(q[4]));
    ray r = cam.getRay(q[0], q[1]);
    for (int k = 0; k < 3; k++) { out[i].o[k] = r.o(k); out[i].dir[k] = r.dir(k); }
    out[i].time = r.time;
  }
  g_inject.clear();
}

// material::emitted + material::scatter with the random vector injected.
// xi[i*4] = (x, y, z, w): for PBR and METAL (x,y,z) are the three randomFloat() values (multiples of
// 2^-24) whose image -1+2*xi lies inside the unit ball, so randomInUnitSphere (vec3.h:62-70) accepts at
// its first iteration; g++ draws them in z, y, x order.  For DIELECTRIC w is the reflectance draw.
void ref_shade_kat(void* p, const srt_ray* raysIn, const srt_hit* hits, const float* xi, int64_t n, float* att,
                   float* emit, srt_ray* raysOut, int32_t* scattered) {
  RefScene* s = static_cast<RefScene*>(p);
  for (int64_t i = 0; i < n; i++) {
    ray rIn(vec3f(raysIn[i].o[0], raysIn[i].o[1], raysIn[i].o[2]),
            vec3f(raysIn[i].dir[0], raysIn[i].dir[1], raysIn[i].dir[2]), raysIn[i].time);
    hitRecord rec = toRecord(s, hits[i]);
    const float* q = xi + 4 * i;
    g_inject.clear();
    const material* m = rec.matPtr.get();
    if (dynamic_cast<const dielectric*>(m)) g_inject.push_back(rawFor(q[3]));
    else { g_inject.push_back(rawFor(q[2])); g_inject.push_back(rawFor(q[1])); g_inject.push_back(rawFor(q[0])); }
    color3f e = m->emitted(rec.uv(0), rec.uv(1), rec.p);
    color3f a(0, 0, 0); ray sc(vec3f(0, 0, 0), vec3f(0, 0, 0), 0);
    bool ok = m->scatter(rIn, rec, a, sc);
    for (int k = 0; k < 3; k++) {
      att[3 * i + k] = a(k); emit[3 * i + k] = e(k); raysOut[i].o[k] = sc.o(k); raysOut[i].dir[k] = sc.dir(k);
    }
    raysOut[i].time = sc.time; scattered[i] = ok ? 1 : 0;
  }
  g_inject.clear();
}

// texture::value(u, v, p); uvp[i*5] = (u, v, p.x, p.y, p.z)
void ref_texture_kat(void* p, int tex, const float* uvp, int64_t n, float* out) {
  RefScene* s = static_cast<RefScene*>(p);
  const texture* t = s->texs.at(tex).get();
  for (int64_t i = 0; i < n; i++) {
    const float* q = uvp + 5 * i;
    color3f c = t->value(q[0], q[1], vec3f(q[2], q[3], q[4]));
    out[3 * i] = c(0); out[3 * i + 1] = c(1); out[3 * i + 2] = c(2);
  }
}

// aabb::hit (aabb.h:11-27) on explicit boxes: box[i*6] = min3 max3; one ray per box
void ref_aabb_kat(const srt_ray* rays, const float* box, const float* tMinMax, int64_t n, int32_t* out) {
  for (int64_t i = 0; i < n; i++) {
    ray r(vec3f(rays[i].o[0], rays[i].o[1], rays[i].o[2]), vec3f(rays[i].dir[0], rays[i].dir[1], rays[i].dir[2]),
          rays[i].time);
    aabb b(vec3f(box[6 * i], box[6 * i + 1], box[6 * i + 2]), vec3f(box[6 * i + 3], box[6 * i + 4], box[6 * i + 5]));
    out[i] = b.hit(r, tMinMax[2 * i], tMinMax[2 * i + 1]) ? 1 : 0;
  }
}

// The in-pixel position of sample s under the 4x rotated grid main.cpp keeps commented out: the statements of
// main.cpp:181,205-207 run as they stand (Eigen's AngleAxisf), out = newSamplePos(0), newSamplePos(1)
void ref_rotated_grid(int s, float* out) {
  const vec3f samplePos(0, 0.8f, 0);
  Eigen::AngleAxisf rotate(deg2rad(15.0f) + deg2rad(90.0f * s), vec3f::UnitZ());
  Eigen::Matrix<float, 3, 3> m = rotate.matrix();                          // mat33f, main.cpp:26
  vec3f       newSamplePos = m * samplePos;
  out[0] = newSamplePos(0);
  out[1] = newSamplePos(1);
}

}  // extern "C"
