// srt_world.cpp — implementation of include/srt_world.hpp (libsrt_host.so): the scene-description classes with the
// reference's constructor signatures, the glTF reader standing in for gltfLoad (model.h:301-460), the flattening into
// srt_scene_desc that replaces hittableVector::build (hittablevector.h:27-31), and b200Device (the headless glDevice).
// Built with -ffp-contract=off: the few arithmetic results computed here (camera, random scene parameters) must equal
// the reference's.
#include "../../include/srt_world.hpp"

#include <dlfcn.h>

#include <cmath>
#include <cstdio>
#include <cstring>
#include <limits>
#include <map>
#include <random>

#include "srt_json.hpp"

extern "C" {
srt_status srt_png_read(const char* path, int32_t req_comp, int32_t* w, int32_t* h, uint8_t** out);
void srt_host_free(void* p);
}

namespace srt {

const float infinity = std::numeric_limits<float>::infinity();
const float epsilon = std::numeric_limits<float>::epsilon();
const float pi = 3.1415926535897932385f;

// ---- globals.h:30-43: one process-wide generator, default seed, draws counted --------------------------------------
namespace {
struct Rng {
  std::mt19937 gen;
  std::uniform_real_distribution<float> dist{0.0f, 1.0f};
  uint64_t draws = 0;
};
Rng& rng() { static Rng r; return r; }
gltfOptions g_gltf;
}  // namespace

float randomFloat() {
  Rng& r = rng();
  const uint64_t before = r.draws;
  struct Counting {                                   // counts raw outputs (generate_canonical<float,24> takes one each)
    Rng& r;
    typedef std::mt19937::result_type result_type;
    static constexpr result_type min() { return std::mt19937::min(); }
    static constexpr result_type max() { return std::mt19937::max(); }
    result_type operator()() { ++r.draws; return r.gen(); }
  } c{r};
  (void)before;
  return r.dist(c);
}
float randomFloat(float mn, float mx) { return mn + (mx - mn) * randomFloat(); }
int randomInt(int mn, int mx) { return static_cast<int>(randomFloat((float)mn, (float)(mx + 1))); }
vec3f randomVec3f() { return vec3f(randomFloat(), randomFloat(), randomFloat()); }
vec3f randomVec3f(float mn, float mx) { return vec3f(randomFloat(mn, mx), randomFloat(mn, mx), randomFloat(mn, mx)); }
uint64_t randomDraws() { return rng().draws; }
void randomReset() { rng() = Rng(); }

float lengthSquared(const vec3f& v) { return v(0) * v(0) + v(1) * v(1) + v(2) * v(2); }
float length(const vec3f& v) { return sqrtf(lengthSquared(v)); }
float distance(const vec3f& u, const vec3f& v) { return sqrtf(lengthSquared(u - v)); }
vec3f unitVector(const vec3f& v) {
  const float len = length(v);
  return len != 0 ? vec3f(v(0) / len, v(1) / len, v(2) / len) : v;
}

// ---- texture.h:112-123 ---------------------------------------------------------------------------------------------
imagePNG::imagePNG(const char* filename, int bytesPP) : bpp(bytesPP) {
  uint8_t* px = nullptr;
  int32_t w = 0, h = 0;
  if (bytesPP >= 1 && bytesPP <= 4 && srt_png_read(filename, bytesPP, &w, &h, &px) == SRT_OK) {
    data.assign(px, px + (size_t)w * h * bytesPP);
    width = w; height = h;
    srt_host_free(px);
  } else {
    fprintf(stderr, "ERROR: Could not load image file '%s'\n", filename);
  }
}

camera::camera(vec3f eye, vec3f lookAt, vec3f up, float vFOV, float aspect, float aperture, float focusDist, float time0, float time1) {
  srt_camera_make(eye.d, lookAt.d, up.d, vFOV, aspect, aperture, focusDist, time0, time1, &cam);
}

// ---- bvh.h:55-95 ---------------------------------------------------------------------------------------------------
namespace {

// SoA view of a list of spheres / triangles in id order (triangles first, then spheres)
struct Soa {
  std::vector<float> v0, v1, v2, uv0, uv1, uv2, c0, c1, t0, t1, r;
  std::vector<const triangle*> tris;
  std::vector<const sphere*> sphs;
  srt_scene_desc desc{};
  void add(const triangle* t, const mesh* m) {
    const vec3f* P[3]; const vec2f* T[3];
    static const vec2f zero;
    for (int k = 0; k < 3; k++) {
      P[k] = &m->positions.at(t->vertices[k]);
      T[k] = t->vertices[k] < m->texcoords.size() ? &m->texcoords[t->vertices[k]] : &zero;
    }
    std::vector<float>* V[3] = {&v0, &v1, &v2};
    std::vector<float>* U[3] = {&uv0, &uv1, &uv2};
    for (int k = 0; k < 3; k++) {
      V[k]->insert(V[k]->end(), P[k]->d, P[k]->d + 3);
      U[k]->insert(U[k]->end(), T[k]->d, T[k]->d + 2);
    }
    tris.push_back(t);
  }
  // a box as a pseudo triangle (v0 = v2 = min corner, v1 = max corner): triangle::boundingBox of it is the box itself, so the
  // builder sorts and unions a nested bvhNode exactly like the reference does through bvhNode::boundingBox (bvh.h:107-110)
  void add_box(const float* bmin, const float* bmax) {
    v0.insert(v0.end(), bmin, bmin + 3); v1.insert(v1.end(), bmax, bmax + 3); v2.insert(v2.end(), bmin, bmin + 3);
    for (auto* U : {&uv0, &uv1, &uv2}) U->insert(U->end(), 2, 0.0f);
    tris.push_back(nullptr);
  }
  void add(const sphere* s) {
    c0.insert(c0.end(), s->center0.d, s->center0.d + 3);
    c1.insert(c1.end(), s->center1.d, s->center1.d + 3);
    t0.push_back(s->t0); t1.push_back(s->t1); r.push_back(s->radius);
    sphs.push_back(s);
  }
  const srt_scene_desc* finish() {
    desc.n_tri = (int64_t)tris.size();
    desc.tri_v0 = v0.data(); desc.tri_v1 = v1.data(); desc.tri_v2 = v2.data();
    desc.tri_uv0 = uv0.data(); desc.tri_uv1 = uv1.data(); desc.tri_uv2 = uv2.data();
    desc.n_sph = (int64_t)sphs.size();
    desc.sph_c0 = c0.data(); desc.sph_c1 = c1.data(); desc.sph_t0 = t0.data(); desc.sph_t1 = t1.data(); desc.sph_r = r.data();
    return &desc;
  }
};

// objects of a span -> Soa + `order` (insertion position -> global primitive id)
bool collect(const std::vector<shared_ptr<hittable>>& objs, size_t start, size_t end, Soa& soa, std::vector<int32_t>& order,
             std::vector<std::shared_ptr<mesh>>& keep, std::string& err) {
  size_t n_tri = 0;
  for (size_t i = start; i < end; i++) if (dynamic_cast<const triangle*>(objs[i].get())) ++n_tri;
  size_t ti = 0, si = 0;
  order.resize(end - start);
  // two passes so that triangle ids precede sphere ids whatever the insertion order
  for (size_t i = start; i < end; i++)
    if (auto t = dynamic_cast<const triangle*>(objs[i].get())) {
      auto m = t->parentMesh.lock();
      if (!m) { err = "triangle whose mesh is gone"; return false; }
      if (t->vertices[0] >= m->positions.size() || t->vertices[1] >= m->positions.size() || t->vertices[2] >= m->positions.size()) {
        err = "triangle index beyond its mesh's positions"; return false;
      }
      keep.push_back(m);
      soa.add(t, m.get());
      order[i - start] = (int32_t)ti++;
    }
  for (size_t i = start; i < end; i++) {
    if (dynamic_cast<const triangle*>(objs[i].get())) continue;
    if (auto s = dynamic_cast<const sphere*>(objs[i].get())) { soa.add(s); order[i - start] = (int32_t)(n_tri + si++); }
    else { err = "bvhNode over an object that is neither a sphere nor a triangle (nested lists / trees are not supported)"; return false; }
  }
  return true;
}

}  // namespace

bvhNode::bvhNode(const std::vector<shared_ptr<hittable>>& srcObjects, size_t start, size_t end, float t0, float t1)
    : time0(t0), time1(t1) {
  if (start > end || end > srcObjects.size()) { error = "bad span"; return; }
  // The span may hold spheres, triangles and other bvhNodes (bvh.h:14: any hittable).  A nested bvhNode takes part in this
  // tree's construction through its root box only - exactly what the reference's comparator and surroundingBox see - and its
  // own nodes are spliced in afterwards where the builder put the box.
  std::vector<const bvhNode*> nested;
  bool plain = true;
  for (size_t i = start; i < end; i++)
    if (auto nb = dynamic_cast<const bvhNode*>(srcObjects[i].get())) {
      if (!nb->error.empty()) { error = "nested bvhNode: " + nb->error; return; }
      if (nb->nodes.empty()) { error = "nested bvhNode without nodes"; return; }
      plain = false;
    }
  Soa soa;
  std::vector<int32_t> order;
  std::vector<std::shared_ptr<mesh>> keep;
  if (plain) {
    if (!collect(srcObjects, start, end, soa, order, keep, error)) return;
    // prims in ID order (triangles, then spheres)
    prims.resize(end - start);
    for (size_t i = start; i < end; i++) prims[order[i - start]] = srcObjects[i];
  } else {
    // builder ids: triangles and nested boxes share the triangle numbering (span order), spheres follow
    size_t n_tb = 0;
    for (size_t i = start; i < end; i++) if (!dynamic_cast<const sphere*>(srcObjects[i].get())) ++n_tb;
    order.resize(end - start);
    size_t ti = 0, si = 0;
    for (size_t i = start; i < end; i++) {
      const hittable* h = srcObjects[i].get();
      if (auto t = dynamic_cast<const triangle*>(h)) {
        auto m = t->parentMesh.lock();
        if (!m) { error = "triangle whose mesh is gone"; return; }
        if (t->vertices[0] >= m->positions.size() || t->vertices[1] >= m->positions.size() || t->vertices[2] >= m->positions.size()) {
          error = "triangle index beyond its mesh's positions"; return;
        }
        keep.push_back(m);
        soa.add(t, m.get());
        order[i - start] = (int32_t)ti++;
      } else if (auto nb = dynamic_cast<const bvhNode*>(h)) {
        soa.add_box(nb->nodes[0].bmin, nb->nodes[0].bmax);
        order[i - start] = (int32_t)ti++;
      } else if (dynamic_cast<const sphere*>(h)) {
        continue;
      } else { error = "bvhNode over an object that is neither a sphere, a triangle nor a bvhNode"; return; }
    }
    for (size_t i = start; i < end; i++)
      if (auto sp = dynamic_cast<const sphere*>(srcObjects[i].get())) { soa.add(sp); order[i - start] = (int32_t)(n_tb + si++); }
  }
  srt_bvh_node* out = nullptr;
  int64_t n = 0;
  uint64_t draws = 0;
  if (srt_bvh_build_ordered(soa.finish(), order.data(), time0, time1, randomDraws(), 0, &out, &n, &draws) != SRT_OK) {
    error = srt_host_last_error();
    return;
  }
  nodes.assign(out, out + n);
  srt_bvh_free(out);
  Rng& r = rng();                       // the constructor's randomInt(0,2) calls, one per node (bvh.h:60)
  r.gen.discard(draws);
  r.draws += draws;
  if (plain) return;

  // ---- splice: final primitive ids (triangles of this span, then the nested trees' triangles, then spheres likewise) and nodes
  std::vector<shared_ptr<hittable>> tri_objs, sph_objs;
  std::vector<int32_t> tri_final(soa.tris.size(), -1);      // builder triangle id -> final id, or -1 - k for nested tree k
  {
    size_t b = 0;
    for (size_t i = start; i < end; i++) {
      const hittable* h = srcObjects[i].get();
      if (dynamic_cast<const triangle*>(h)) { tri_final[b++] = (int32_t)tri_objs.size(); tri_objs.push_back(srcObjects[i]); }
      else if (auto nb = dynamic_cast<const bvhNode*>(h)) { tri_final[b++] = -1 - (int32_t)nested.size(); nested.push_back(nb); }
    }
    for (size_t i = start; i < end; i++) if (dynamic_cast<const sphere*>(srcObjects[i].get())) sph_objs.push_back(srcObjects[i]);
  }
  std::vector<size_t> ntri(nested.size()), tri_base(nested.size()), sph_base(nested.size());
  for (size_t k = 0; k < nested.size(); k++) {
    size_t c = 0;
    for (auto& o : nested[k]->prims) if (dynamic_cast<const triangle*>(o.get())) ++c;      // (prims are triangles first)
    ntri[k] = c;
    tri_base[k] = tri_objs.size();
    tri_objs.insert(tri_objs.end(), nested[k]->prims.begin(), nested[k]->prims.begin() + c);
  }
  const size_t n_tri_total = tri_objs.size();
  for (size_t k = 0; k < nested.size(); k++) {
    sph_base[k] = sph_objs.size();
    sph_objs.insert(sph_objs.end(), nested[k]->prims.begin() + ntri[k], nested[k]->prims.end());
  }
  prims = tri_objs;
  prims.insert(prims.end(), sph_objs.begin(), sph_objs.end());
  const size_t n_tb_builder = soa.tris.size();
  auto final_leaf = [&](int32_t builder_id, int32_t& nested_k) -> int32_t {       // -> final primitive id, or nested_k >= 0
    nested_k = -1;
    if ((size_t)builder_id < n_tb_builder) {
      const int32_t f = tri_final[builder_id];
      if (f < 0) { nested_k = -1 - f; return -1; }
      return f;
    }
    return (int32_t)(n_tri_total + ((size_t)builder_id - n_tb_builder));          // this span's spheres come first among the spheres
  };
  auto append_nested = [&](size_t k) -> int32_t {                                  // copies tree k behind `nodes`, returns its root
    const int32_t off = (int32_t)nodes.size();
    for (const srt_bvh_node& src : nested[k]->nodes) {
      srt_bvh_node nd = src;
      for (int32_t* c : {&nd.left, &nd.right}) {
        if (*c == SRT_BVH_NONE) continue;
        if (*c >= 0) *c += off;
        else {
          const size_t id = (size_t)(~*c);
          const size_t f = id < ntri[k] ? tri_base[k] + id : n_tri_total + sph_base[k] + (id - ntri[k]);
          *c = ~(int32_t)f;
        }
      }
      nodes.push_back(nd);
    }
    return off;
  };
  const size_t n_outer = nodes.size();
  for (size_t i = 0; i < n_outer; i++) {
    int32_t kl = -1, kr = -1;
    const bool single = nodes[i].right == SRT_BVH_NONE;       // a span of one object: the reference sets left = right (bvh.h:67-69)
    if (nodes[i].left < 0) {
      const int32_t f = final_leaf(~nodes[i].left, kl);
      if (kl < 0) nodes[i].left = ~f;
    }
    if (!single && nodes[i].right < 0) {
      const int32_t f = final_leaf(~nodes[i].right, kr);
      if (kr < 0) nodes[i].right = ~f;
    }
    if (kl >= 0) { const int32_t root = append_nested((size_t)kl); nodes[i].left = root; }
    if (single && kl >= 0) { const int32_t root = append_nested((size_t)kl); nodes[i].right = root; }   // hit twice, as the reference does
    if (kr >= 0) { const int32_t root = append_nested((size_t)kr); nodes[i].right = root; }
  }
}

// ---- flattening ------------------------------------------------------------------------------------------------------
namespace {

struct Tables {
  std::vector<srt_material>& mats;
  std::vector<srt_texture>& texs;
  std::vector<shared_ptr<texture>>& owners;
  std::map<const material*, int> mat_id;
  std::map<const texture*, int> tex_id;
  std::string err;

  // ids in discovery order; a checker's children come first so their ids are smaller than its own
  int tex(const shared_ptr<texture>& t) {
    if (!t) return -1;
    auto it = tex_id.find(t.get());
    if (it != tex_id.end()) return it->second;
    srt_texture r;
    memset(&r, 0, sizeof r);
    r.even = r.odd = -1;
    if (auto c = dynamic_cast<const checker*>(t.get())) {
      r.kind = SRT_TEX_CHECKER;
      r.even = tex(c->even); r.odd = tex(c->odd);
      if (r.even < 0 || r.odd < 0) { err = "checker without both textures"; return -1; }
    } else if (auto s = dynamic_cast<const solidColor*>(t.get())) {
      r.kind = SRT_TEX_SOLID;
      for (int k = 0; k < 3; k++) r.color[k] = s->colorValue(k);
    } else if (auto p = dynamic_cast<const imagePNG*>(t.get())) {
      r.kind = SRT_TEX_IMAGE;
      r.width = p->width; r.height = p->height; r.bpp = p->bpp;
      r.data = p->data.empty() ? nullptr : p->data.data();
    } else {
      err = "unknown texture class"; return -1;
    }
    const int id = (int)texs.size();
    texs.push_back(r);
    owners.push_back(t);
    tex_id[t.get()] = id;
    return id;
  }
  int mat(const shared_ptr<material>& m) {
    if (!m) { err = "primitive without a material"; return -1; }
    auto it = mat_id.find(m.get());
    if (it != mat_id.end()) return it->second;
    srt_material r;
    memset(&r, 0, sizeof r);
    r.albedo_tex = r.normal_tex = r.metallic_tex = r.roughness_tex = -1;
    const int id = (int)mats.size();
    mats.emplace_back();
    mat_id[m.get()] = id;
    if (auto p = dynamic_cast<const pbrMetallicRoughness*>(m.get())) {
      r.kind = SRT_MAT_PBR;
      r.albedo_tex = tex(p->albedoMap); r.normal_tex = tex(p->normalMap);
      r.metallic_tex = tex(p->metallicMap); r.roughness_tex = tex(p->roughnessMap);
      for (int k = 0; k < 4; k++) r.albedo[k] = p->albedo(k);
      r.metalness = p->metalness; r.roughness = p->roughness;
      if (p->sampleMetallicRoughnessMap && p->metallicRoughnessMap) {      // not the reference: SRT_MATF_PACKED_MR
        r.metallic_tex = tex(p->metallicRoughnessMap); r.roughness_tex = -1; r.flags |= SRT_MATF_PACKED_MR;
      }
    } else if (auto q = dynamic_cast<const metal*>(m.get())) {
      r.kind = SRT_MAT_METAL;
      for (int k = 0; k < 3; k++) r.albedo[k] = q->albedo(k);
      r.fuzz = q->fuzz;
    } else if (auto d = dynamic_cast<const dielectric*>(m.get())) {
      r.kind = SRT_MAT_DIELECTRIC; r.ir = d->ir;
    } else if (auto l = dynamic_cast<const diffuseLight*>(m.get())) {
      r.kind = SRT_MAT_LIGHT; r.albedo_tex = tex(l->emit);
    } else {
      err = "unknown material class"; return -1;
    }
    mats[id] = r;
    return id;
  }
};

}  // namespace

bool flatScene::build(const hittableList& world) {
  error.clear();
  if (world.objects.size() != 1) { error = "world must hold exactly one bvhNode (main.cpp:146)"; return false; }
  auto root = dynamic_cast<const bvhNode*>(world.objects[0].get());
  if (!root) { error = "world's object is not a bvhNode"; return false; }
  if (!root->error.empty()) { error = root->error; return false; }
  tri_v0.clear(); tri_v1.clear(); tri_v2.clear(); tri_uv0.clear(); tri_uv1.clear(); tri_uv2.clear(); tri_mat.clear();
  sph_c0.clear(); sph_c1.clear(); sph_t0.clear(); sph_t1.clear(); sph_r.clear(); sph_mat.clear();
  materials.clear(); textures.clear(); texture_owners.clear();
  Soa soa;
  std::vector<int32_t> order;
  std::vector<std::shared_ptr<mesh>> keep;
  if (!collect(root->prims, 0, root->prims.size(), soa, order, keep, error)) return false;
  Tables T{materials, textures, texture_owners, {}, {}, {}};
  // material ids in first-use order over the primitives in id order (triangles carry their mesh's material)
  for (const triangle* t : soa.tris) tri_mat.push_back(T.mat(t->parentMesh.lock()->matPtr));
  for (const sphere* s : soa.sphs) sph_mat.push_back(T.mat(s->matPtr));
  if (!T.err.empty()) { error = T.err; return false; }
  tri_v0.swap(soa.v0); tri_v1.swap(soa.v1); tri_v2.swap(soa.v2); tri_uv0.swap(soa.uv0); tri_uv1.swap(soa.uv1); tri_uv2.swap(soa.uv2);
  sph_c0.swap(soa.c0); sph_c1.swap(soa.c1); sph_t0.swap(soa.t0); sph_t1.swap(soa.t1); sph_r.swap(soa.r);
  nodes = root->nodes;
  // vertex normals, when any mesh has them; triangles of meshes without get their face normal three times
  tri_n0.clear(); tri_n1.clear(); tri_n2.clear();
  bool any_normals = false;
  for (const triangle* t : soa.tris) if (!t->parentMesh.lock()->normals.empty()) { any_normals = true; break; }
  if (any_normals) {
    std::vector<float>* N[3] = {&tri_n0, &tri_n1, &tri_n2};
    for (size_t i = 0; i < soa.tris.size(); i++) {
      const triangle* t = soa.tris[i];
      const auto m = t->parentMesh.lock();
      const bool has = t->vertices[0] < m->normals.size() && t->vertices[1] < m->normals.size() && t->vertices[2] < m->normals.size();
      float fn[3] = {0, 0, 0};
      if (!has) {
        const float *a = &tri_v0[3 * i], *b = &tri_v1[3 * i], *c = &tri_v2[3 * i];
        const double e0[3] = {(double)b[0] - a[0], (double)b[1] - a[1], (double)b[2] - a[2]}, e1[3] = {(double)c[0] - a[0], (double)c[1] - a[1], (double)c[2] - a[2]};
        const double n[3] = {e0[1] * e1[2] - e0[2] * e1[1], e0[2] * e1[0] - e0[0] * e1[2], e0[0] * e1[1] - e0[1] * e1[0]};
        const double len = std::sqrt(n[0] * n[0] + n[1] * n[1] + n[2] * n[2]);
        for (int k = 0; k < 3; k++) fn[k] = len > 0 ? (float)(n[k] / len) : 0.0f;
      }
      for (int k = 0; k < 3; k++) {
        const float* src = has ? m->normals[t->vertices[k]].d : fn;
        N[k]->insert(N[k]->end(), src, src + 3);
      }
    }
  }
  // vertex tangents, when any mesh has them; triangles of meshes without get w = 0 (= keep the per-face basis)
  tri_tan0.clear(); tri_tan1.clear(); tri_tan2.clear();
  bool any_tangents = false;
  for (const triangle* t : soa.tris) if (!t->parentMesh.lock()->tangents.empty()) { any_tangents = true; break; }
  if (any_tangents) {
    std::vector<float>* T[3] = {&tri_tan0, &tri_tan1, &tri_tan2};
    static const float none[4] = {0, 0, 0, 0};
    for (const triangle* t : soa.tris) {
      const auto m = t->parentMesh.lock();
      const bool has = t->vertices[0] < m->tangents.size() && t->vertices[1] < m->tangents.size() && t->vertices[2] < m->tangents.size();
      for (int k = 0; k < 3; k++) {
        const float* src = has ? m->tangents[t->vertices[k]].d : none;
        T[k]->insert(T[k]->end(), src, src + 4);
      }
    }
  }
  memset(&d, 0, sizeof d);
  d.n_tri = (int64_t)tri_mat.size();
  d.tri_v0 = tri_v0.data(); d.tri_v1 = tri_v1.data(); d.tri_v2 = tri_v2.data();
  d.tri_uv0 = tri_uv0.data(); d.tri_uv1 = tri_uv1.data(); d.tri_uv2 = tri_uv2.data(); d.tri_mat = tri_mat.data();
  if (any_normals) { d.tri_n0 = tri_n0.data(); d.tri_n1 = tri_n1.data(); d.tri_n2 = tri_n2.data(); }
  if (any_tangents) { d.tri_tan0 = tri_tan0.data(); d.tri_tan1 = tri_tan1.data(); d.tri_tan2 = tri_tan2.data(); }
  d.n_sph = (int64_t)sph_mat.size();
  d.sph_c0 = sph_c0.data(); d.sph_c1 = sph_c1.data(); d.sph_t0 = sph_t0.data(); d.sph_t1 = sph_t1.data(); d.sph_r = sph_r.data();
  d.sph_mat = sph_mat.data();
  d.n_nodes = (int64_t)nodes.size(); d.nodes = nodes.data();
  d.n_mat = (int32_t)materials.size(); d.materials = materials.data();
  d.n_tex = (int32_t)textures.size(); d.textures = textures.data();
  return true;
}

// ---- glTF (model.h:301-460) -------------------------------------------------------------------------------------------
gltfOptions& gltfConfig() { return g_gltf; }
bool model::init() { return gltfLoad(filename, shared_from_this()); }

namespace {

bool read_file(const std::string& path, std::vector<uint8_t>& out) {
  FILE* f = fopen(path.c_str(), "rb");
  if (!f) return false;
  uint8_t tmp[65536];
  size_t n;
  while ((n = fread(tmp, 1, sizeof tmp, f)) > 0) out.insert(out.end(), tmp, tmp + n);
  fclose(f);
  return true;
}

bool base64(const char* s, size_t n, std::vector<uint8_t>& out) {
  uint32_t acc = 0;
  int bits = 0;
  for (size_t i = 0; i < n; i++) {
    const char c = s[i];
    int v;
    if (c >= 'A' && c <= 'Z') v = c - 'A'; else if (c >= 'a' && c <= 'z') v = c - 'a' + 26;
    else if (c >= '0' && c <= '9') v = c - '0' + 52; else if (c == '+' || c == '-') v = 62;
    else if (c == '/' || c == '_') v = 63; else if (c == '=') break; else return false;
    acc = (acc << 6) | (uint32_t)v;
    bits += 6;
    if (bits >= 8) { bits -= 8; out.push_back((uint8_t)(acc >> bits)); }
  }
  return true;
}

std::string dir_of(const std::string& path) {
  const size_t k = path.find_last_of('/');
  return k == std::string::npos ? std::string() : path.substr(0, k + 1);
}

std::string uri_decode(const std::string& s) {
  std::string o;
  for (size_t i = 0; i < s.size(); i++) {
    if (s[i] == '%' && i + 2 < s.size() + 0 && isxdigit((unsigned char)s[i + 1]) && isxdigit((unsigned char)s[i + 2])) {
      o += (char)strtol(s.substr(i + 1, 2).c_str(), nullptr, 16);
      i += 2;
    } else {
      o += s[i];
    }
  }
  return o;
}

struct Mat4 { double m[16]; };     // column major, as glTF stores it
Mat4 identity() { Mat4 r{}; r.m[0] = r.m[5] = r.m[10] = r.m[15] = 1; return r; }
Mat4 mul(const Mat4& a, const Mat4& b) {
  Mat4 r{};
  for (int c = 0; c < 4; c++) for (int rr = 0; rr < 4; rr++) {
    double s = 0;
    for (int k = 0; k < 4; k++) s += a.m[k * 4 + rr] * b.m[c * 4 + k];
    r.m[c * 4 + rr] = s;
  }
  return r;
}
Mat4 node_matrix(const srt_json::Value& n) {
  if (const auto* m = n.get("matrix")) {
    Mat4 r = identity();
    for (size_t i = 0; i < 16 && i < m->size(); i++) r.m[i] = m->at(i).num;
    return r;
  }
  double t[3] = {0, 0, 0}, q[4] = {0, 0, 0, 1}, s[3] = {1, 1, 1};
  if (const auto* v = n.get("translation")) for (size_t i = 0; i < 3 && i < v->size(); i++) t[i] = v->at(i).num;
  if (const auto* v = n.get("rotation")) for (size_t i = 0; i < 4 && i < v->size(); i++) q[i] = v->at(i).num;
  if (const auto* v = n.get("scale")) for (size_t i = 0; i < 3 && i < v->size(); i++) s[i] = v->at(i).num;
  const double x = q[0], y = q[1], z = q[2], w = q[3];
  Mat4 r = identity();
  r.m[0] = (1 - 2 * (y * y + z * z)) * s[0]; r.m[1] = (2 * (x * y + z * w)) * s[0]; r.m[2] = (2 * (x * z - y * w)) * s[0];
  r.m[4] = (2 * (x * y - z * w)) * s[1]; r.m[5] = (1 - 2 * (x * x + z * z)) * s[1]; r.m[6] = (2 * (y * z + x * w)) * s[1];
  r.m[8] = (2 * (x * z + y * w)) * s[2]; r.m[9] = (2 * (y * z - x * w)) * s[2]; r.m[10] = (1 - 2 * (x * x + y * y)) * s[2];
  r.m[12] = t[0]; r.m[13] = t[1]; r.m[14] = t[2];
  return r;
}

struct Gltf {
  srt_json::Value root;
  std::vector<std::vector<uint8_t>> buffers;
  std::string dir;

  // -> pointer to the first element + stride, per the active rules
  bool view(long accessor, size_t elem_bytes, const uint8_t*& ptr, size_t& stride, size_t& count, const srt_json::Value*& acc) const {
    const auto* accs = root.get("accessors");
    const auto* views = root.get("bufferViews");
    if (!accs || !views || accessor < 0 || (size_t)accessor >= accs->size()) return false;
    acc = &accs->at(accessor);
    const long vi = acc->index("bufferView");
    if (vi < 0 || (size_t)vi >= views->size()) return false;
    const auto& bv = views->at(vi);
    const long bi = bv.index("buffer");
    if (bi < 0 || (size_t)bi >= buffers.size()) return false;
    // every number below comes from an untrusted file: reject anything that is not a finite integer in [0, 2^53] before it
    // is cast, and check the range without arithmetic that could wrap
    auto field = [](const srt_json::Value& v, const char* key, size_t& out) -> bool {
      const double d = v.number(key, 0);
      if (!(d >= 0.0) || !(d <= 9007199254740992.0) || d != std::floor(d)) return false;
      out = (size_t)d;
      return true;
    };
    const size_t size = buffers[bi].size();
    size_t off = 0, acc_off = 0, bs = 0, view_len = 0;
    if (!field(bv, "byteOffset", off) || !field(*acc, "byteOffset", acc_off) || !field(bv, "byteStride", bs) ||
        !field(*acc, "count", count) || !field(bv, "byteLength", view_len)) return false;
    if (off > size) return false;
    size_t avail = size - off;                                   // bytes from the view's start to the end of the buffer
    if (bv.get("byteLength")) { if (view_len > avail) return false; if (!g_gltf.referenceQuirks) avail = view_len; }
    stride = elem_bytes;
    if (!g_gltf.referenceQuirks) {
      if (acc_off > avail) return false;
      off += acc_off; avail -= acc_off;
      if (bs) { if (bs < elem_bytes) return false; stride = bs; }
    }
    if (count && (elem_bytes > avail || (count - 1) > (avail - elem_bytes) / stride)) return false;
    ptr = buffers[bi].data() + off;
    return true;
  }
};

std::string image_path(const Gltf& g, const srt_json::Value& texview) {
  const long ti = texview.index("index");
  const auto* texs = g.root.get("textures");
  const auto* imgs = g.root.get("images");
  if (ti < 0 || !texs || (size_t)ti >= texs->size() || !imgs) return std::string();
  const long src = texs->at(ti).index("source");
  if (src < 0 || (size_t)src >= imgs->size()) return std::string();
  const std::string uri = imgs->at(src).string("uri");
  if (uri.empty()) return std::string();
  return g_gltf.textureDir + uri;                    // model.h:393-394: "../data/" + uri
}

}  // namespace

bool gltfLoad(std::string filename, shared_ptr<model> mdl) {
  std::vector<uint8_t> file;
  if (!read_file(filename, file)) return false;
  Gltf g;
  g.dir = dir_of(filename);
  const uint8_t* json = file.data();
  size_t json_len = file.size();
  const uint8_t* bin = nullptr;
  size_t bin_len = 0;
  if (file.size() >= 20 && !memcmp(file.data(), "glTF", 4)) {            // GLB container
    auto le32 = [&](size_t o) { return (uint32_t)file[o] | ((uint32_t)file[o + 1] << 8) | ((uint32_t)file[o + 2] << 16) | ((uint32_t)file[o + 3] << 24); };
    size_t pos = 12;
    json = nullptr;
    while (pos + 8 <= file.size()) {
      const uint32_t len = le32(pos), type = le32(pos + 4);
      if (pos + 8 + len > file.size()) return false;
      if (type == 0x4e4f534a && !json) { json = file.data() + pos + 8; json_len = len; }
      else if (type == 0x004e4942 && !bin) { bin = file.data() + pos + 8; bin_len = len; }
      pos += 8 + ((len + 3) & ~3u);
    }
    if (!json) return false;
  }
  std::string perr;
  srt_json::Parser P((const char*)json, json_len);
  if (!P.parse(g.root, perr)) return false;

  // cgltf_load_buffers
  if (const auto* bufs = g.root.get("buffers")) {
    for (size_t i = 0; i < bufs->size(); i++) {
      std::vector<uint8_t> data;
      const std::string uri = bufs->at(i).string("uri");
      if (uri.empty()) {
        if (i == 0 && bin) data.assign(bin, bin + bin_len); else return false;
      } else if (uri.compare(0, 5, "data:") == 0) {
        const size_t comma = uri.find(',');
        if (comma == std::string::npos || uri.find(";base64") == std::string::npos) return false;
        if (!base64(uri.c_str() + comma + 1, uri.size() - comma - 1, data)) return false;
      } else if (!read_file(g.dir + uri_decode(uri), data)) {
        return false;
      }
      if (data.size() < (size_t)bufs->at(i).number("byteLength", 0)) return false;
      g.buffers.push_back(std::move(data));
    }
  }

  // mesh index -> world matrix (only used when referenceQuirks is off; the reference ignores nodes)
  std::map<long, Mat4> mesh_xf;
  if (!g_gltf.referenceQuirks) {
    const auto* nodes = g.root.get("nodes");
    std::vector<int> parent(nodes ? nodes->size() : 0, -1);
    for (size_t i = 0; nodes && i < nodes->size(); i++)
      if (const auto* ch = nodes->at(i).get("children"))
        for (size_t k = 0; k < ch->size(); k++) { const long c = (long)ch->at(k).num; if (c >= 0 && (size_t)c < parent.size()) parent[c] = (int)i; }
    for (size_t i = 0; nodes && i < nodes->size(); i++) {
      const long mi = nodes->at(i).index("mesh");
      if (mi < 0) continue;
      Mat4 w = node_matrix(nodes->at(i));
      int guard = 0;
      for (int p = parent[i]; p >= 0 && guard < 256; p = parent[p], ++guard) w = mul(node_matrix(nodes->at(p)), w);
      if (!mesh_xf.count(mi)) mesh_xf[mi] = w;
    }
  }

  const auto* meshes = g.root.get("meshes");
  const auto* materials = g.root.get("materials");
  for (size_t meshIndex = 0; meshes && meshIndex < meshes->size(); meshIndex++) {
    const auto* prims = meshes->at(meshIndex).get("primitives");
    for (size_t primIndex = 0; prims && primIndex < prims->size(); primIndex++) {
      const auto& gp = prims->at(primIndex);
      shared_ptr<mesh> newMesh = mesh::create();
      mdl->meshes.push_back(newMesh);
      // model.h:345,361,450 append to meshes[primIndex], not to newMesh (equal for single-mesh files)
      shared_ptr<mesh> target = g_gltf.referenceQuirks ? mdl->meshes.at(primIndex) : newMesh;

      if (const auto* attrs = gp.get("attributes")) {
        for (const auto& kv : attrs->obj) {                       // file order, like cgltf's attribute list
          const std::string& name = kv.first;
          if (name.empty() || name[0] == '_') continue;
          const std::string base = name.substr(0, name.find('_'));
          const uint8_t* ptr; size_t stride, count; const srt_json::Value* acc;
          if (base == "POSITION") {
            if (!g.view((long)kv.second.num, 12, ptr, stride, count, acc)) return false;
            if (acc->string("type") != "VEC3") continue;
            const Mat4* xf = (!g_gltf.referenceQuirks && mesh_xf.count((long)meshIndex)) ? &mesh_xf[(long)meshIndex] : nullptr;
            for (size_t k = 0; k < count; k++) {
              float v[3];
              memcpy(v, ptr + k * stride, 12);
              if (xf) {
                const double x = v[0], y = v[1], z = v[2];
                for (int r = 0; r < 3; r++) v[r] = (float)(xf->m[r] * x + xf->m[4 + r] * y + xf->m[8 + r] * z + xf->m[12 + r]);
              }
              target->positions.push_back(vec3f(v[0], v[1], v[2]));
            }
          } else if (name == "NORMAL" && g_gltf.smoothNormals) {          // never read by the reference (model.h:335-365)
            if (!g.view((long)kv.second.num, 12, ptr, stride, count, acc)) return false;
            if (acc->string("type") != "VEC3") continue;
            const Mat4* xf = (!g_gltf.referenceQuirks && mesh_xf.count((long)meshIndex)) ? &mesh_xf[(long)meshIndex] : nullptr;
            double it[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};                   // inverse transpose of the upper 3x3, row-major
            if (xf) {
              const double a = xf->m[0], b = xf->m[4], c = xf->m[8], d2 = xf->m[1], e = xf->m[5], f = xf->m[9], g2 = xf->m[2], h = xf->m[6], i2 = xf->m[10];
              const double det = a * (e * i2 - f * h) - b * (d2 * i2 - f * g2) + c * (d2 * h - e * g2);
              if (det != 0) {
                const double inv[9] = {(e * i2 - f * h) / det, (c * h - b * i2) / det, (b * f - c * e) / det,
                                       (f * g2 - d2 * i2) / det, (a * i2 - c * g2) / det, (c * d2 - a * f) / det,
                                       (d2 * h - e * g2) / det, (b * g2 - a * h) / det, (a * e - b * d2) / det};
                for (int r = 0; r < 3; r++) for (int q = 0; q < 3; q++) it[3 * r + q] = inv[3 * q + r];
              }
            }
            for (size_t k = 0; k < count; k++) {
              float v[3];
              memcpy(v, ptr + k * stride, 12);
              if (xf) {
                const double x = v[0], y = v[1], z = v[2];
                double o[3];
                for (int r = 0; r < 3; r++) o[r] = it[3 * r] * x + it[3 * r + 1] * y + it[3 * r + 2] * z;
                const double len = std::sqrt(o[0] * o[0] + o[1] * o[1] + o[2] * o[2]);
                for (int r = 0; r < 3; r++) v[r] = (float)(len > 0 ? o[r] / len : o[r]);
              }
              target->normals.push_back(vec3f(v[0], v[1], v[2]));
            }
          } else if (name == "TANGENT" && g_gltf.vertexTangents) {        // never read by the reference (model.h:335-365)
            if (!g.view((long)kv.second.num, 16, ptr, stride, count, acc)) return false;
            if (acc->string("type") != "VEC4" || (long)acc->number("componentType", 5126) != 5126) continue;
            const Mat4* xf = (!g_gltf.referenceQuirks && mesh_xf.count((long)meshIndex)) ? &mesh_xf[(long)meshIndex] : nullptr;
            for (size_t k = 0; k < count; k++) {
              float v[4];
              memcpy(v, ptr + k * stride, 16);
              if (xf) {                                                    // a direction: the upper 3x3, renormalised; w stays
                const double x = v[0], y = v[1], z = v[2];
                double o[3];
                for (int r = 0; r < 3; r++) o[r] = xf->m[r] * x + xf->m[4 + r] * y + xf->m[8 + r] * z;
                const double len = std::sqrt(o[0] * o[0] + o[1] * o[1] + o[2] * o[2]);
                for (int r = 0; r < 3; r++) v[r] = (float)(len > 0 ? o[r] / len : o[r]);
              }
              target->tangents.push_back(vec4f(v[0], v[1], v[2], v[3]));
            }
          } else if (base == "TEXCOORD") {
            if (!g_gltf.referenceQuirks && name != "TEXCOORD_0") continue;
            if (!g.view((long)kv.second.num, 8, ptr, stride, count, acc)) return false;
            if (acc->string("type") != "VEC2") continue;
            if (!g_gltf.referenceQuirks && (long)acc->number("componentType", 5126) != 5126) continue;
            for (size_t k = 0; k < count; k++) {
              float v[2];
              memcpy(v, ptr + k * stride, 8);
              target->texcoords.push_back(vec2f(v[0], v[1]));
            }
          }
        }
      }

      const long mi = gp.index("material");
      if (mi >= 0 && materials && (size_t)mi < materials->size()) {
        const auto& gm = materials->at(mi);
        if (const auto* pbr = gm.get("pbrMetallicRoughness")) {           // has_pbr_metallic_roughness
          std::string textureFile, normalMapFile, mrFile;
          if (const auto* v = pbr->get("baseColorTexture")) textureFile = image_path(g, *v);
          if (const auto* v = gm.get("normalTexture")) normalMapFile = image_path(g, *v);
          if (const auto* v = pbr->get("metallicRoughnessTexture")) mrFile = image_path(g, *v);
          vec4f baseColor(1.0f, 1.0f, 1.0f, 1.0f);                        // cgltf defaults
          if (const auto* f = pbr->get("baseColorFactor"))
            for (size_t k = 0; k < 4 && k < f->size(); k++) baseColor(k) = (float)f->at(k).num;
          const float metallicness = (float)pbr->number("metallicFactor", 1.0);
          const float roughness = (float)pbr->number("roughnessFactor", 1.0);
          shared_ptr<imagePNG> albedoPNG, normalPNG, mrPNG;
          if (!textureFile.empty()) albedoPNG = make_shared<imagePNG>(textureFile.c_str(), 3);
          if (!normalMapFile.empty()) normalPNG = make_shared<imagePNG>(normalMapFile.c_str(), 3);
          if (!mrFile.empty()) mrPNG = make_shared<imagePNG>(mrFile.c_str(), 3);
          auto pm = make_shared<pbrMetallicRoughness>(albedoPNG, normalPNG, mrPNG, baseColor, metallicness, roughness);
          pm->sampleMetallicRoughnessMap = g_gltf.sampleMetallicRoughness && mrPNG != nullptr;
          newMesh->matPtr = pm;
        }
      }

      const long mode = gp.has("mode") ? gp.index("mode") : 4;
      if (mode == 4 && gp.has("indices")) {
        const uint8_t* ptr; size_t stride, count; const srt_json::Value* acc;
        const auto* accs = g.root.get("accessors");
        const long ai = gp.index("indices");
        long ctype = 5123;
        if (!g_gltf.referenceQuirks && accs && ai >= 0 && (size_t)ai < accs->size()) ctype = (long)accs->at(ai).number("componentType", 5123);
        const size_t eb = ctype == 5121 ? 1 : ctype == 5125 ? 4 : 2;
        if (!g.view(ai, eb, ptr, stride, count, acc)) return false;
        auto idx = [&](size_t k) -> uint32_t {
          const uint8_t* q = ptr + k * stride;
          if (eb == 1) return q[0];
          if (eb == 2) { uint16_t v; memcpy(&v, q, 2); return v; }
          uint32_t v; memcpy(&v, q, 4); return v;
        };
        for (size_t k = 0; k + 2 < count; k += 3) target->triangles.push_back(triangle::create(idx(k), idx(k + 1), idx(k + 2), target));
      }
    }
  }
  return true;
}

// ---- b200Device ---------------------------------------------------------------------------------------------------------
namespace {
struct Api {
  srt_status (*group_create)(const int*, int, srt_device_group**);
  void (*group_destroy)(srt_device_group*);
  int32_t (*group_size)(const srt_device_group*);
  srt_device* (*group_device)(srt_device_group*, int32_t);
  srt_status (*scene_group_create)(srt_device_group*, const srt_scene_desc*, srt_scene_group**);
  void (*scene_group_destroy)(srt_scene_group*);
  srt_scene* (*scene_group_scene)(srt_scene_group*, int32_t);
  srt_status (*render_group)(srt_scene_group*, const srt_camera*, const srt_render_params*, float*, uint8_t*, srt_stats*);
  const char* (*last_error)(void);
  srt_status (*film_create)(srt_device*, int32_t, int32_t, srt_film**);
  void (*film_destroy)(srt_film*);
  srt_status (*film_add)(srt_film*, srt_scene*, const srt_camera*, const srt_render_params*, int32_t, srt_stats*);
  int32_t (*film_samples)(const srt_film*);
  srt_status (*film_resolve)(srt_film*, uint8_t*, float*);
  srt_status (*film_load)(srt_film*, const float*, int32_t);
};
bool bind(void* so, Api& a) {
  a.group_create = (decltype(a.group_create))dlsym(so, "srt_device_group_create");
  a.group_destroy = (decltype(a.group_destroy))dlsym(so, "srt_device_group_destroy");
  a.group_size = (decltype(a.group_size))dlsym(so, "srt_device_group_size");
  a.group_device = (decltype(a.group_device))dlsym(so, "srt_device_group_device");
  a.scene_group_create = (decltype(a.scene_group_create))dlsym(so, "srt_scene_group_create");
  a.scene_group_destroy = (decltype(a.scene_group_destroy))dlsym(so, "srt_scene_group_destroy");
  a.scene_group_scene = (decltype(a.scene_group_scene))dlsym(so, "srt_scene_group_scene");
  a.render_group = (decltype(a.render_group))dlsym(so, "srt_render_group");
  a.last_error = (decltype(a.last_error))dlsym(so, "srt_last_error");
  a.film_create = (decltype(a.film_create))dlsym(so, "srt_film_create");
  a.film_destroy = (decltype(a.film_destroy))dlsym(so, "srt_film_destroy");
  a.film_add = (decltype(a.film_add))dlsym(so, "srt_film_add_samples");
  a.film_samples = (decltype(a.film_samples))dlsym(so, "srt_film_samples");
  a.film_resolve = (decltype(a.film_resolve))dlsym(so, "srt_film_resolve");
  a.film_load = (decltype(a.film_load))dlsym(so, "srt_film_load");
  return a.group_create && a.group_destroy && a.group_size && a.group_device && a.scene_group_create && a.scene_group_destroy &&
         a.scene_group_scene && a.render_group && a.last_error && a.film_create &&
         a.film_destroy && a.film_add && a.film_samples && a.film_resolve && a.film_load;
}
}  // namespace

bool b200Device::init(int w, int h, const hittableList& world, int gpu, const char* library_path) {
  return init(w, h, world, gpu < 0 ? std::vector<int>() : std::vector<int>(1, gpu), library_path);
}

bool b200Device::init(int, int, const hittableList& world, const std::vector<int>& gpus, const char* library_path) {
  terminate();
  flatScene flat;
  if (!flat.build(world)) { err = flat.error; return false; }
  const char* path = library_path ? library_path : (getenv("SRT_B200_LIBRARY") ? getenv("SRT_B200_LIBRARY") : "libsrt_b200.so");
  so = dlopen(path, RTLD_NOW | RTLD_LOCAL);
  if (!so) { err = std::string("cannot load the CUDA library (there is no CPU fallback): ") + dlerror(); return false; }
  Api a;
  if (!bind(so, a)) { err = "libsrt_b200.so lacks an entry point of include/srt.h"; return false; }
  // one device per GPU of the list (empty = every visible GPU), one replica of the scene on each (glDevice::init's upload)
  if (a.group_create(gpus.empty() ? nullptr : gpus.data(), (int)gpus.size(), &group) != SRT_OK) { err = a.last_error(); group = nullptr; return false; }
  if (a.scene_group_create(group, &flat.desc(), &scenes) != SRT_OK) { err = a.last_error(); scenes = nullptr; return false; }
  dev = a.group_device(group, 0);                 // progressive films live on the first GPU
  scene = a.scene_group_scene(scenes, 0);
  return true;
}

bool b200Device::rtFrame(void* frameData, int w, int h, const camera& cam, const color3f& background, int numSamples, int maxBounce,
                         uint64_t seed, srt_stats* stats) {
  if (!scene) { err = "rtFrame before init"; return false; }
  Api a;
  bind(so, a);
  srt_render_params p;
  memset(&p, 0, sizeof p);
  p.width = w; p.height = h; p.spp = numSamples; p.max_bounce = maxBounce; p.seed = seed; p.sampler = sampler;
  for (int k = 0; k < 3; k++) p.background[k] = background(k);
  srt_stats st;
  if (a.render_group(scenes, &cam.desc(), &p, nullptr, static_cast<uint8_t*>(frameData), &st) != SRT_OK) { err = a.last_error(); return false; }
  if (stats) *stats = st;
  return true;
}

bool b200Device::rtFrameProgressive(void* frameData, int w, int h, const camera& cam, const color3f& background, int samples,
                                    int maxBounce, uint64_t seed, srt_stats* stats) {
  if (!scene) { err = "rtFrameProgressive before init"; return false; }
  Api a;
  bind(so, a);
  if (film && (film_w != w || film_h != h)) { a.film_destroy(film); film = nullptr; }
  if (!film) {
    if (a.film_create(dev, w, h, &film) != SRT_OK) { err = a.last_error(); film = nullptr; return false; }
    film_w = w; film_h = h;
  }
  if (a.film_samples(film) > 0 &&
      (seed != film_seed || (film_sampler >= 0 && sampler != film_sampler) || (film_bounce >= 0 && maxBounce != film_bounce))) {
    err = "rtFrameProgressive: the film already holds samples of another seed / sampler / maxBounce (resetFilm() starts a new one)";
    return false;
  }
  film_seed = seed; film_sampler = sampler; film_bounce = maxBounce;
  srt_render_params p;
  memset(&p, 0, sizeof p);
  p.width = w; p.height = h; p.max_bounce = maxBounce; p.seed = seed; p.sampler = sampler;
  for (int k = 0; k < 3; k++) p.background[k] = background(k);
  srt_stats st;
  if (a.film_add(film, scene, &cam.desc(), &p, samples, &st) != SRT_OK) { err = a.last_error(); return false; }
  if (stats) *stats = st;
  if (frameData && a.film_resolve(film, static_cast<uint8_t*>(frameData), nullptr) != SRT_OK) { err = a.last_error(); return false; }
  return true;
}

int b200Device::gpuCount() const {
  if (!group) return 0;
  Api a;
  return bind(so, a) ? a.group_size(group) : 0;
}

int b200Device::samplesDone() const {
  if (!film) return 0;
  Api a;
  bind(so, a);
  return a.film_samples(film);
}

void b200Device::resetFilm() {
  if (!film) return;
  Api a;
  if (bind(so, a)) a.film_destroy(film);
  film = nullptr; film_sampler = film_bounce = -1;
}

bool b200Device::saveCheckpoint(const char* path) {
  if (!film) { err = "no progressive render to save"; return false; }
  Api a;
  bind(so, a);
  std::vector<float> sum((size_t)film_w * film_h * 3);
  if (a.film_resolve(film, nullptr, sum.data()) != SRT_OK) { err = a.last_error(); return false; }
  if (srt_checkpoint_write2(path, film_w, film_h, a.film_samples(film), film_seed, film_sampler, film_bounce, sum.data()) != SRT_OK) { err = srt_host_last_error(); return false; }
  return true;
}

bool b200Device::loadCheckpoint(const char* path, int w, int h) {
  if (!scene) { err = "loadCheckpoint before init"; return false; }
  Api a;
  bind(so, a);
  int32_t cw = 0, ch = 0, done = 0, smp = -1, mb = -1;
  uint64_t seed = 0;
  float* sum = nullptr;
  if (srt_checkpoint_read2(path, &cw, &ch, &done, &seed, &smp, &mb, &sum) != SRT_OK) { err = srt_host_last_error(); return false; }
  if (cw != w || ch != h) { srt_host_free(sum); err = "checkpoint is of a different frame size"; return false; }
  if (film && (film_w != w || film_h != h)) { a.film_destroy(film); film = nullptr; }
  if (!film && a.film_create(dev, w, h, &film) != SRT_OK) { srt_host_free(sum); err = a.last_error(); film = nullptr; return false; }
  film_w = w; film_h = h; film_seed = seed; film_sampler = smp; film_bounce = mb;
  const srt_status rc = a.film_load(film, sum, done);
  srt_host_free(sum);
  if (rc != SRT_OK) { err = a.last_error(); return false; }
  return true;
}

void b200Device::terminate() {
  if (so) {
    Api a;
    if (bind(so, a)) {
      if (film) a.film_destroy(film);
      if (scenes) a.scene_group_destroy(scenes);
      if (group) a.group_destroy(group);
    }
    dlclose(so);
  }
  so = nullptr; dev = nullptr; scene = nullptr; film = nullptr; group = nullptr; scenes = nullptr;
}

}  // namespace srt
