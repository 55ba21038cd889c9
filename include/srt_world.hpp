// srt_world.hpp — the HOST side of the drop-in: scene-description classes with the reference's names and constructor
// signatures, which own the flattening into include/srt.h's srt_scene_desc (last paragraph).
//
// A main.cpp-equivalent (examples/main.cpp) keeps the reference's set-up code (main.cpp:54-172) word for word -
// camera(...), make_shared<sphere>(...), make_shared<pbrMetallicRoughness>(...), model::create(...)->init(),
// make_shared<bvhNode>(objects, 0, 1) - and replaces the pixel loop main.cpp:200-227 by b200Device::rtFrame.
//
//   reference class (file:line)                         here
//   vec2f/vec3f/vec4f/color3f  vec3.h:12-18 (Eigen)     small PODs with the (i) accessor / dot / cross the set-up code uses
//   randomFloat/randomInt      globals.h:30-43          same std::mt19937 + uniform_real_distribution<float>, draws counted
//   texture, solidColor, checker, imagePNG, image3bpp   texture.h:13-154   description only (no value()): sampled on the GPU
//   material, pbrMetallicRoughness (9 ctors), metal,    material.h:15-154  description only (no scatter()): shaded on the GPU
//     dielectric, diffuseLight
//   hittable, sphere, triangle, mesh, model             hittable.h:24-32, sphere.h:9-45, model.h:19-102
//   gltfLoad                                            model.h:301-460    own glTF 2.0 / GLB reader (no cgltf needed)
//   hittableList                                        hittablelist.h:14-34
//   bvhNode                                             bvh.h:14-31,55-95  built by srt_bvh_build (same tree, same RNG use)
//   camera                                              camera.h:8-63      srt_camera_make (+ desc() exposing the members)
//   glDevice::init / rtFrame / terminate                gl.h:28-31         b200Device over libsrt_b200.so (headless)
//   writeColorTarget + stbi_write_png                   color.h:25-41, main.cpp:235-237   srt_resolve on the GPU + srt_png_write
//
// The classes DESCRIBE a scene; none of them can intersect or shade on the CPU (there is no CPU fallback).
// Implementation: sexy-raytracer_b200/host/srt_world.cpp (libsrt_host.so).  Plain C++17, no Eigen, no cgltf, no stb.
#ifndef SRT_WORLD_HPP_
#define SRT_WORLD_HPP_

#include <cstdint>
#include <memory>
#include <string>
#include <vector>

#include "srt.h"
#include "srt_host.h"

namespace srt {

using std::make_shared;
using std::shared_ptr;

// ---- vec3.h:12-18 -----------------------------------------------------------------------------------------------
struct vec2f {
  float d[2] = {0, 0};
  vec2f() = default;
  vec2f(float a, float b) : d{a, b} {}
  float& operator()(int i) { return d[i]; }
  float operator()(int i) const { return d[i]; }
};
struct vec3f {
  float d[3] = {0, 0, 0};
  vec3f() = default;
  template <class A, class B, class C>              // Eigen accepts any mix of int / float / double scalars here
  vec3f(A a, B b, C c) : d{(float)a, (float)b, (float)c} {}
  float& operator()(int i) { return d[i]; }
  float operator()(int i) const { return d[i]; }
  static vec3f UnitX() { return vec3f(1.f, 0.f, 0.f); }
  static vec3f UnitY() { return vec3f(0.f, 1.f, 0.f); }
  static vec3f UnitZ() { return vec3f(0.f, 0.f, 1.f); }
  float dot(const vec3f& o) const { return d[0] * o.d[0] + (d[1] * o.d[1] + d[2] * o.d[2]); }   // Eigen's order
  vec3f cross(const vec3f& o) const {
    return vec3f(d[1] * o.d[2] - d[2] * o.d[1], d[2] * o.d[0] - d[0] * o.d[2], d[0] * o.d[1] - d[1] * o.d[0]);
  }
  bool operator==(const vec3f& o) const { return d[0] == o.d[0] && d[1] == o.d[1] && d[2] == o.d[2]; }
  bool operator!=(const vec3f& o) const { return !(*this == o); }
};
struct vec4f {
  float d[4] = {0, 0, 0, 0};
  vec4f() = default;
  vec4f(float a, float b, float c, float e) : d{a, b, c, e} {}
  float& operator()(int i) { return d[i]; }
  float operator()(int i) const { return d[i]; }
};
using color3f = vec3f;
using color4f = vec4f;
inline vec3f operator+(const vec3f& a, const vec3f& b) { return vec3f(a(0) + b(0), a(1) + b(1), a(2) + b(2)); }
inline vec3f operator-(const vec3f& a, const vec3f& b) { return vec3f(a(0) - b(0), a(1) - b(1), a(2) - b(2)); }
inline vec3f operator-(const vec3f& a) { return vec3f(-a(0), -a(1), -a(2)); }
inline vec3f operator*(float s, const vec3f& a) { return vec3f(s * a(0), s * a(1), s * a(2)); }
inline vec3f operator*(const vec3f& a, float s) { return vec3f(a(0) * s, a(1) * s, a(2) * s); }
inline vec3f operator/(const vec3f& a, float s) { return vec3f(a(0) / s, a(1) / s, a(2) / s); }

// ---- globals.h:13-43, vec3.h:29-60 ------------------------------------------------------------------------------
extern const float infinity, epsilon, pi;
inline float clamp(float x, float mn, float mx) { return x < mn ? mn : (x > mx ? mx : x); }
inline float deg2rad(float degrees) { return degrees * 3.1415926535897932385f / 180.0f; }
float randomFloat();                       // the process-wide default-seeded std::mt19937 of globals.h:30-35
float randomFloat(float mn, float mx);
int   randomInt(int mn, int mx);
vec3f randomVec3f();                       // vec3.h:41-47: the three draws are call arguments (compiler's evaluation order)
vec3f randomVec3f(float mn, float mx);
uint64_t randomDraws();                    // generator outputs consumed so far (= srt_bvh_build's rng_skip)
void     randomReset();                    // back to the default seed, zero draws (tests)
float lengthSquared(const vec3f& v);
float length(const vec3f& v);
float distance(const vec3f& u, const vec3f& v);
vec3f unitVector(const vec3f& v);

// ---- texture.h --------------------------------------------------------------------------------------------------
class texture {
 public:
  virtual ~texture() = default;
};
class solidColor : public texture {
 public:
  solidColor() = default;
  solidColor(color3f c) : colorValue(c) {}
  solidColor(float r, float g, float b) : colorValue(r, g, b) {}
  color3f colorValue;
};
class checker : public texture {
 public:
  checker() = default;
  checker(shared_ptr<texture> _even, shared_ptr<texture> _odd) : odd(_odd), even(_even) {}
  checker(color3f c1, color3f c2) : odd(make_shared<solidColor>(c2)), even(make_shared<solidColor>(c1)) {}
  shared_ptr<texture> odd, even;
};
// imagePNG(filename, bytesPP): texture.h:112-123.  A file that cannot be read leaves data empty -> the device returns the
// (1,0,1) fallback of texture.h:130-131.  Decoding and channel conversion follow stb_image's rules (srt_png_read).
class imagePNG : public texture {
 public:
  explicit imagePNG(int bytesPP) : bpp(bytesPP) {}
  imagePNG(const char* filename, int bytesPP);
  std::vector<uint8_t> data;
  int width = 0, height = 0, bpp = 3;
};
class image3bpp : public imagePNG {            // texture.h:54-107: the same lookup with bpp fixed at 3
 public:
  image3bpp() : imagePNG(3) {}
  explicit image3bpp(const char* filename) : imagePNG(filename, 3) {}
};

// ---- material.h -------------------------------------------------------------------------------------------------
class material {
 public:
  virtual ~material() = default;
};
class pbrMetallicRoughness : public material {
 public:
  // the nine constructors of material.h:25-70, in source order.  Members a constructor leaves uninitialised in the
  // reference (metalness/roughness in the first four) are 0 here - the value pinned on both sides.
  pbrMetallicRoughness(const color3f& a) : albedoMap(make_shared<solidColor>(a)) {}
  pbrMetallicRoughness(shared_ptr<texture> aMap) : albedoMap(aMap) {}
  pbrMetallicRoughness(shared_ptr<texture> aMap, vec4f a) : albedoMap(aMap), albedo(a) {}
  pbrMetallicRoughness(shared_ptr<texture> aMap, shared_ptr<texture> nMap) : albedoMap(aMap), normalMap(nMap) {}
  pbrMetallicRoughness(shared_ptr<texture> aMap, shared_ptr<texture> nMap, shared_ptr<texture> mMap, shared_ptr<texture> rMap)
      : albedoMap(aMap), normalMap(nMap), metallicMap(mMap), roughnessMap(rMap) {}
  pbrMetallicRoughness(shared_ptr<texture> aMap, shared_ptr<texture> nMap, shared_ptr<texture> mMap, shared_ptr<texture> rMap, vec4f a)
      : albedoMap(aMap), normalMap(nMap), metallicMap(mMap), roughnessMap(rMap), albedo(a) {}
  pbrMetallicRoughness(shared_ptr<texture> aMap, shared_ptr<texture> nMap, shared_ptr<texture> mrMap, vec4f a)
      : albedoMap(aMap), normalMap(nMap), metallicRoughnessMap(mrMap), albedo(a) {}
  pbrMetallicRoughness(shared_ptr<texture> aMap, shared_ptr<texture> nMap, shared_ptr<texture> mrMap, vec4f a, float m, float r)
      : albedoMap(aMap), normalMap(nMap), metallicRoughnessMap(mrMap), albedo(a), metalness(m), roughness(r) {}
  pbrMetallicRoughness(shared_ptr<texture> aMap, vec4f a, float m, float r) : albedoMap(aMap), albedo(a), metalness(m), roughness(r) {}

  shared_ptr<texture> albedoMap, normalMap;
  shared_ptr<texture> metallicRoughnessMap;       // stored, never sampled (material.h:78) ...
  bool sampleMetallicRoughnessMap = false;        // ... unless this is set (NOT the reference, SRT_MATF_PACKED_MR in srt.h;
                                                  // gltfLoad sets it when gltfOptions::sampleMetallicRoughness is on)
  shared_ptr<texture> metallicMap, roughnessMap;
  vec4f albedo = vec4f(1.0f, 1.0f, 1.0f, 1.0f);
  float metalness = 0, roughness = 0, anisotropy = 0;
};
class metal : public material {
 public:
  metal(const color3f& a, float f) : albedo(a), fuzz(f < 1.0f ? f : 1.0f) {}
  color3f albedo;
  float fuzz;
};
class dielectric : public material {
 public:
  dielectric(float indexRefraction) : ir(indexRefraction) {}
  float ir;
};
class diffuseLight : public material {
 public:
  diffuseLight(shared_ptr<texture> a) : emit(a) {}
  diffuseLight(color3f c) : emit(make_shared<solidColor>(c)) {}
  shared_ptr<texture> emit;
};

// ---- hittables --------------------------------------------------------------------------------------------------
class hittable {
 public:
  virtual ~hittable() = default;
};
class sphere : public hittable {
 public:
  sphere() = default;
  sphere(vec3f c0, vec3f c1, float time0, float time1, float r, shared_ptr<material> m)
      : center0(c0), center1(c1), t0(time0), t1(time1), radius(r), matPtr(m) {}
  vec3f center0, center1;
  float t0 = 0, t1 = 1;
  float radius = 0;
  shared_ptr<material> matPtr;
};
class mesh;
class model;
// model.h:19-50; indices are 32-bit here (the reference's uint16_t caps a mesh at 65 536 vertices, model.h:46)
class triangle : public hittable {
 public:
  static shared_ptr<triangle> create(uint32_t index0, uint32_t index1, uint32_t index2, shared_ptr<mesh> srcMesh) {
    return shared_ptr<triangle>(new triangle(index0, index1, index2, srcMesh));
  }
  uint32_t vertices[3];
  std::weak_ptr<mesh> parentMesh;      // weak: mesh::triangles owns the triangles (the reference leaks this cycle)
 private:
  triangle(uint32_t a, uint32_t b, uint32_t c, shared_ptr<mesh> m) : vertices{a, b, c}, parentMesh(m) {}
};
class mesh : public hittable {          // model.h:52-76
 public:
  static shared_ptr<mesh> create() { return shared_ptr<mesh>(new mesh()); }
  std::vector<shared_ptr<triangle>> triangles;
  std::vector<vec3f> positions;
  std::vector<vec2f> texcoords;
  std::vector<vec3f> normals;           // glTF NORMAL, filled only under gltfOptions::smoothNormals (the reference never reads it)
  std::vector<vec4f> tangents;          // glTF TANGENT (xyz + handedness), filled only under gltfOptions::vertexTangents
  shared_ptr<material> matPtr;
 private:
  mesh() = default;
};
class model : public hittable, public std::enable_shared_from_this<model> {   // model.h:78-102
 public:
  static shared_ptr<model> create(std::string fn) { return shared_ptr<model>(new model(fn)); }
  bool init();                          // gltfLoad(filename, this), model.h:301-460
  std::string filename;
  std::vector<shared_ptr<mesh>> meshes;
 private:
  explicit model(std::string fn) : filename(fn) {}
};

// How gltfLoad behaves.  The defaults ARE the reference's behaviour (every fix is a deviation, so it
// sits behind a flag).
struct gltfOptions {
  std::string textureDir = "../data/";  // model.h:393,401,409 prefix image URIs with this literal
  bool referenceQuirks = true;          // true: read attributes/indices from bufferView.byteOffset ignoring the accessor's
                                        //       byteOffset and the view's byteStride, append to meshes[primIndex]
                                        //       (model.h:345,361,450), ignore node transforms, 16-bit indices only
                                        // false: honour accessor.byteOffset / byteStride, 8/16/32-bit indices, append to the
                                        //       primitive's own mesh, apply node TRS / matrix transforms to positions
  // Two things the reference does not do at all, both off by default and independent of referenceQuirks:
  bool smoothNormals = false;           // read NORMAL into mesh::normals; a scene holding any is shaded with interpolated
                                        // vertex normals (srt_scene_desc::tri_n0..2); meshes without NORMAL keep face normals
  bool vertexTangents = false;          // read TANGENT into mesh::tangents; triangles of such meshes are shaded with the interpolated
                                        // tangent (srt_scene_desc::tri_tan0..2) instead of calcTangentBasis' per-face one (model.h:214-235)
  bool sampleMetallicRoughness = false; // materials with a metallicRoughnessTexture sample it (glTF 2.0: metal = B * factor,
                                        // roughness = G * factor) instead of storing it unused (material.h:78)
};
gltfOptions& gltfConfig();
bool gltfLoad(std::string filename, shared_ptr<model> model);

class hittableList : public hittable {  // hittablelist.h:14-34
 public:
  hittableList() = default;
  hittableList(shared_ptr<hittable> object) { add(object); }
  void clear() { objects.clear(); }
  void add(shared_ptr<hittable> object) { objects.push_back(object); }
  std::vector<shared_ptr<hittable>> objects;
};

// bvh.h:14-31,55-95.  The constructor consumes the process-wide generator exactly like the reference's recursion (one
// randomInt(0,2) per node, pre-order) and produces the same tree, through srt_bvh_build (in place, multi-threaded).
class bvhNode : public hittable {
 public:
  bvhNode(const hittableList& list, float time0, float time1) : bvhNode(list.objects, 0, list.objects.size(), time0, time1) {}
  bvhNode(const std::vector<shared_ptr<hittable>>& srcObjects, size_t start, size_t end, float time0, float time1);
  std::vector<shared_ptr<hittable>> prims;   // the span's objects in insertion order; leaf id = index into this
  std::vector<srt_bvh_node> nodes;           // pre-order, root 0, leaves ~index into prims
  float time0, time1;
  std::string error;                         // non-empty: the span held something that is neither sphere nor triangle
};

class camera {                            // camera.h:8-63
 public:
  camera(vec3f eye, vec3f lookAt, vec3f up, float vFOV, float aspect, float aperture, float focusDist, float time0, float time1);
  const srt_camera& desc() const { return cam; }   // the ten members that are private in the reference (camera.h:52-62)
 private:
  srt_camera cam;
};

// ---- flattening: replaces hittableVector::build (hittablevector.h:27-31, main.cpp:148-149) -----------------------------
// Owns every array srt_scene_desc points at.  `world` must be what main.cpp:146 builds: a list holding one bvhNode.
// Primitive ids: triangles first, then spheres, each in insertion order (for main.cpp's scene = insertion order).
class flatScene {
 public:
  bool build(const hittableList& world);
  const srt_scene_desc& desc() const { return d; }
  std::string error;
  std::vector<float> tri_v0, tri_v1, tri_v2, tri_uv0, tri_uv1, tri_uv2, sph_c0, sph_c1, sph_t0, sph_t1, sph_r;
  std::vector<float> tri_n0, tri_n1, tri_n2;         // empty unless some mesh carries vertex normals
  std::vector<float> tri_tan0, tri_tan1, tri_tan2;   // empty unless some mesh carries vertex tangents (4 floats per triangle vertex)
  std::vector<int32_t> tri_mat, sph_mat;
  std::vector<srt_bvh_node> nodes;
  std::vector<srt_material> materials;
  std::vector<srt_texture> textures;
  std::vector<shared_ptr<texture>> texture_owners;   // keeps image bytes alive
 private:
  srt_scene_desc d{};
};

// ---- the device, in the shape of glDevice (gl.h:28-31), headless ------------------------------------------------------
// init uploads the flattened scene (glDevice::init's SSBO upload); rtFrame renders ONE whole frame of `numSamples` samples
// per pixel into frameData (RGBA8, row 0 = top, what writeColorTarget produces) and returns false on error (the text is in
// lastError()).  libsrt_b200.so is loaded at run time with dlopen so that libsrt_host.so itself needs no CUDA.
class b200Device {
 public:
  ~b200Device() { terminate(); }
  // gpu < 0 (the default): every visible GPU - the samples of every pixel are split across them and summed on the first one
  // (srt_render_group, include/srt.h); gpu >= 0: that GPU alone; or an explicit list.
  bool init(int w, int h, const hittableList& world, int gpu = -1, const char* library_path = nullptr);
  bool init(int w, int h, const hittableList& world, const std::vector<int>& gpus, const char* library_path = nullptr);
  int  gpuCount() const;
  bool rtFrame(void* frameData, int w, int h, const camera& cam, const color3f& background, int numSamples, int maxBounce,
               uint64_t seed = 1, srt_stats* stats = nullptr);
  // Progressive form - what the reference's `while (glDevice.rtFrame(...))` loop is for (main.cpp:195-199): each call adds
  // `samples` more samples per pixel to a film that stays on the device and resolves the frame so far into frameData.
  // saveCheckpoint / loadCheckpoint write / restore the film (fp32 sums + sample count) so a long render can resume.
  bool rtFrameProgressive(void* frameData, int w, int h, const camera& cam, const color3f& background, int samples, int maxBounce,
                          uint64_t seed = 1, srt_stats* stats = nullptr);
  // A film holds ONE sample stream: once it has samples, a call with another seed, sampler or maxBounce is refused (it would
  // continue the global sample indices with different draws); resetFilm() starts over.
  int  samplesDone() const;
  void resetFilm();
  bool saveCheckpoint(const char* path);
  bool loadCheckpoint(const char* path, int w, int h);
  void terminate();
  const std::string& lastError() const { return err; }
  // where a sample lies inside its pixel: 0 = random jitter (main.cpp:210-211), 1 = the 4x rotated grid the reference keeps
  // commented out (main.cpp:205-207,214-215); see srt_render_params::sampler
  int sampler = 0;
 private:
  void* so = nullptr;
  srt_device_group* group = nullptr;
  srt_scene_group* scenes = nullptr;
  srt_device* dev = nullptr;          // borrowed from the group: its first GPU
  srt_scene* scene = nullptr;         // borrowed: the replica on that GPU
  srt_film* film = nullptr;
  int film_w = 0, film_h = 0;
  uint64_t film_seed = 1;
  int film_sampler = -1, film_bounce = -1;      // -1: not known (a version-1 checkpoint)
  std::string err;
};

}  // namespace srt
#endif  // SRT_WORLD_HPP_
