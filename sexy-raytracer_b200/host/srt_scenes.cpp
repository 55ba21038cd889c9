// srt_scenes.cpp — the reference's own scenes written against include/srt_world.hpp exactly as main.cpp writes them
// against the reference's headers (main.cpp:54-154), plus the C entry points that hand their flattened form to callers
// that are not C++ (tests, bench).  The set-up statements mirror main.cpp line for line so that the process-wide random
// stream is consumed identically; the flattened result is compared bit for bit with the compiled reference's own export
// in tests/test_world.py.
#include <cstring>
#include <string>

#include "../../include/srt_world.hpp"

namespace srt_host_detail {
srt_status fail(srt_status s, const std::string& m);
}
using srt_host_detail::fail;
using namespace srt;

namespace {

void addModel(hittableList& objects, std::vector<shared_ptr<model>>& keep, const std::string& file, bool& ok) {
  auto testModel = model::create(file);
  if (!testModel->init()) { ok = false; return; }
  keep.push_back(testModel);
  for (const auto& mesh : testModel->meshes)            // main.cpp:82-86
    for (const auto& tri : mesh->triangles) objects.add(tri);
}

void groundAndLight(hittableList& objects) {
  auto checkerTex = make_shared<checker>(color3f(0.2f, 0.3f, 0.1f), color3f(0.9f, 0.9f, 0.9f));                       // main.cpp:89
  objects.add(make_shared<sphere>(vec3f(0, -1000, 0.0f), vec3f(0, -1000, 0.0f), 0, 1.0f, 1000,
                                  make_shared<pbrMetallicRoughness>(checkerTex)));                                    // main.cpp:90
  auto lightMat = make_shared<diffuseLight>(color3f(250.2f, 220.9f, 110.2f));                                         // main.cpp:126
  objects.add(make_shared<sphere>(vec3f(-7.0f, 4.0f, 6.0f), vec3f(-7.0f, 4.0f, 6.0f), 0, 1.0f, 1.0f, lightMat));      // main.cpp:127
}

// the diffuse role of the Shirley block: see SURVEY 8.0(4) - pbrMetallicRoughness(color3f) is near-black in the reference
// (albedo / 255), so config 1 feeds a solidColor in the 0..255 convention with metalness 0, roughness 0.5
shared_ptr<material> diffuse255(const color3f& albedo) {
  auto m = make_shared<pbrMetallicRoughness>(make_shared<solidColor>(albedo * 255.0f));
  m->metalness = 0; m->roughness = 0.5f;
  return m;
}

}  // namespace

struct srt_host_scene {
  hittableList world;
  std::vector<shared_ptr<model>> models;
  flatScene flat;
  uint64_t draws = 0;
};

namespace {

bool referenceScene(int which, const std::string& dataDir, srt_host_scene& S, std::string& err) {
  hittableList objects;
  bool ok = true;
  const std::string savedDir = gltfConfig().textureDir;
  gltfConfig().textureDir = dataDir;                    // stands for the literal "../data/" of model.h:393
  randomReset();                                        // the generator's state at the start of main()
  if (which == 11 || which == 12) {
    // NOT in the reference's main.cpp: config 1's spheres with bvhNodes nested inside the outer bvhNode's list, which bvh.h:14
    // allows (any hittable).  11: the small spheres form their own tree next to the ground and the three large spheres;
    // 12: as 11, and the outer list holds nothing but one bvhNode over all of that (a span of one object, bvh.h:67-69).
    // The spheres are created first, with config 1's draws, so their parameters equal config 1's.
    hittableList field, top;
    auto checkerTex = make_shared<checker>(color3f(0.2f, 0.3f, 0.1f), color3f(0.9f, 0.9f, 0.9f));
    top.add(make_shared<sphere>(vec3f(0, -1000, 0.0f), vec3f(0, -1000, 0.0f), 0, 1.0f, 1000, make_shared<pbrMetallicRoughness>(checkerTex)));
    for (int a = -11; a < 11; a++) {
      for (int b = -11; b < 11; b++) {
        auto choose_mat = randomFloat();
        vec3f center(a + 0.9 * randomFloat(), 0.2f, b + 0.9f * randomFloat());
        if (length(center - vec3f(4, 0.2, 0)) > 0.9) {
          if (choose_mat < 0.8) {
            color3f albedo(randomFloat() * randomFloat(), randomFloat() * randomFloat(), randomFloat() * randomFloat());
            auto sphere_material = diffuse255(albedo);
            vec3f center2 = center + vec3f(0, randomFloat(0, 0.5f), 0);
            field.add(make_shared<sphere>(center, center2, 0, 1.0f, 0.2f, sphere_material));
          } else if (choose_mat < 0.95f) {
            auto albedo = randomVec3f(0.5f, 1);
            auto fuzz = randomFloat(0, 0.5f);
            field.add(make_shared<sphere>(center, center, 0, 1.0f, 0.2f, make_shared<metal>(albedo, fuzz)));
          } else {
            field.add(make_shared<sphere>(center, center, 0, 1.0f, 0.2f, make_shared<dielectric>(1.5f)));
          }
        }
      }
    }
    top.add(make_shared<bvhNode>(field, 0, 1));
    top.add(make_shared<sphere>(vec3f(0, 1, 0), vec3f(0, 1, 0), 0, 1.0f, 1.0f, make_shared<dielectric>(1.5)));
    top.add(make_shared<sphere>(vec3f(-4, 1, 0), vec3f(-4, 1, 0), 0, 1.0f, 1.0f, diffuse255(color3f(0.4f, 0.2f, 0.1f))));
    top.add(make_shared<sphere>(vec3f(4.0f, 1.0f, 0), vec3f(4.0f, 1.0f, 0), 0, 1.0f, 1.0f, make_shared<metal>(color3f(0.7, 0.6, 0.5), 0.0)));
    if (which == 11) objects = top;
    else objects.add(make_shared<bvhNode>(top, 0, 1));
  } else if (which == 1) {
    auto checkerTex = make_shared<checker>(color3f(0.2f, 0.3f, 0.1f), color3f(0.9f, 0.9f, 0.9f));
    objects.add(make_shared<sphere>(vec3f(0, -1000, 0.0f), vec3f(0, -1000, 0.0f), 0, 1.0f, 1000, make_shared<pbrMetallicRoughness>(checkerTex)));
    for (int a = -11; a < 11; a++) {                    // main.cpp:92-122
      for (int b = -11; b < 11; b++) {
        auto choose_mat = randomFloat();
        vec3f center(a + 0.9 * randomFloat(), 0.2f, b + 0.9f * randomFloat());
        if (length(center - vec3f(4, 0.2, 0)) > 0.9) {
          if (choose_mat < 0.8) {
            color3f albedo(randomFloat() * randomFloat(), randomFloat() * randomFloat(), randomFloat() * randomFloat());
            auto sphere_material = diffuse255(albedo);
            vec3f center2 = center + vec3f(0, randomFloat(0, 0.5f), 0);
            objects.add(make_shared<sphere>(center, center2, 0, 1.0f, 0.2f, sphere_material));
          } else if (choose_mat < 0.95f) {
            auto albedo = randomVec3f(0.5f, 1);
            auto fuzz = randomFloat(0, 0.5f);
            objects.add(make_shared<sphere>(center, center, 0, 1.0f, 0.2f, make_shared<metal>(albedo, fuzz)));
          } else {
            objects.add(make_shared<sphere>(center, center, 0, 1.0f, 0.2f, make_shared<dielectric>(1.5f)));
          }
        }
      }
    }
    objects.add(make_shared<sphere>(vec3f(0, 1, 0), vec3f(0, 1, 0), 0, 1.0f, 1.0f, make_shared<dielectric>(1.5)));            // main.cpp:124-125
    objects.add(make_shared<sphere>(vec3f(-4, 1, 0), vec3f(-4, 1, 0), 0, 1.0f, 1.0f, diffuse255(color3f(0.4f, 0.2f, 0.1f)))); // main.cpp:129-130
    objects.add(make_shared<sphere>(vec3f(4.0f, 1.0f, 0), vec3f(4.0f, 1.0f, 0), 0, 1.0f, 1.0f,
                                    make_shared<metal>(color3f(0.7, 0.6, 0.5), 0.0)));                                        // main.cpp:143-144
  } else if (which == 2) {
    addModel(objects, S.models, dataDir + "cube.gltf", ok);
    addModel(objects, S.models, dataDir + "square.gltf", ok);
    groundAndLight(objects);
  } else if (which == 3 || which == 4) {
    addModel(objects, S.models, dataDir + "masterchief2-separate-xf.gltf", ok);                                               // main.cpp:74-75
    groundAndLight(objects);
    shared_ptr<imagePNG> ironAlbedo, ironNMap, ironMMap, ironRMap;
    if (which == 3) {                                                                                                         // main.cpp:133-136
      ironAlbedo = make_shared<imagePNG>((dataDir + "rustediron2_basecolor-2x1.png").c_str(), 3);
      ironNMap = make_shared<imagePNG>((dataDir + "rustediron2_normal-2x1.png").c_str(), 3);
      ironMMap = make_shared<imagePNG>((dataDir + "rustediron2_metallic-2x1.png").c_str(), 1);
      ironRMap = make_shared<imagePNG>((dataDir + "rustediron2_roughness-2x1.png").c_str(), 1);
    } else {
      ironAlbedo = make_shared<imagePNG>((dataDir + "Image_1.png").c_str(), 3);
      ironNMap = make_shared<imagePNG>((dataDir + "Image_0.png").c_str(), 3);
      ironMMap = make_shared<imagePNG>((dataDir + "worldisyoursbanner.png").c_str(), 1);
      ironRMap = make_shared<imagePNG>((dataDir + "worldisyoursbanner.png").c_str(), 1);
    }
    auto ironMat = make_shared<pbrMetallicRoughness>(ironAlbedo, ironNMap, ironMMap, ironRMap, vec4f(1.0f, 1.0f, 1.0f, 1.0f));   // main.cpp:137-139
    objects.add(make_shared<sphere>(vec3f(-3.0f, 1.0f, 0.0f), vec3f(-3.0f, 1.0f, 0.0f), 0, 1.0f, 1.0f, ironMat));              // main.cpp:140
    auto material3 = make_shared<metal>(color3f(0.7, 0.6, 0.5), 0.0);                                                          // main.cpp:143
    objects.add(make_shared<sphere>(vec3f(3.0f, 1.0f, 0), vec3f(3.0f, 1.0f, 0), 0, 1.0f, 1.0f, material3));                    // main.cpp:144
  } else {
    err = "unknown reference scene";
    ok = false;
  }
  gltfConfig().textureDir = savedDir;
  if (!ok) { if (err.empty()) err = "cannot load a glTF file under " + dataDir; return false; }
  S.world.add(make_shared<bvhNode>(objects, 0, 1));     // main.cpp:146
  S.draws = randomDraws();
  if (!S.flat.build(S.world)) { err = S.flat.error; return false; }
  return true;
}

}  // namespace

namespace srt {
// for examples/main.cpp: randomScene() of main.cpp:54-154 - which = 3: the default branch; 1: the commented Shirley block
hittableList randomScene(const std::string& dataDir, std::vector<shared_ptr<model>>& keepAlive, int which) {
  srt_host_scene S;
  std::string err;
  if (!referenceScene(which, dataDir, S, err)) return hittableList();
  keepAlive = S.models;
  return S.world;
}
}  // namespace srt

extern "C" {

srt_status srt_host_scene_reference(int which, const char* data_dir, srt_host_scene** out) try {
  if (!out || !data_dir) return fail(SRT_ERR_INVALID, "NULL argument");
  *out = nullptr;
  std::string dir(data_dir);
  if (!dir.empty() && dir.back() != '/') dir += '/';
  auto* S = new srt_host_scene;
  std::string err;
  if (!referenceScene(which, dir, *S, err)) { delete S; return fail(SRT_ERR_INVALID, err); }
  *out = S;
  return SRT_OK;
} catch (const std::bad_alloc&) { return fail(SRT_ERR_NOMEM, "out of memory"); } catch (const std::exception& e_) { return fail(SRT_ERR_INVALID, e_.what()); } catch (...) { return fail(SRT_ERR_INVALID, "unknown exception"); }

srt_status srt_host_scene_gltf(const char* gltf_path, const char* texture_dir, int options, srt_host_scene** out) try {
  if (!out || !gltf_path) return fail(SRT_ERR_INVALID, "NULL argument");
  *out = nullptr;
  auto* S = new srt_host_scene;
  const gltfOptions saved = gltfConfig();
  if (texture_dir) gltfConfig().textureDir = texture_dir;
  gltfConfig().referenceQuirks = (options & SRT_GLTF_REFERENCE_QUIRKS) != 0;
  gltfConfig().smoothNormals = (options & SRT_GLTF_SMOOTH_NORMALS) != 0;
  gltfConfig().sampleMetallicRoughness = (options & SRT_GLTF_SAMPLE_MR) != 0;
  gltfConfig().vertexTangents = (options & SRT_GLTF_VERTEX_TANGENTS) != 0;
  hittableList objects;
  bool ok = true;
  addModel(objects, S->models, gltf_path, ok);
  gltfConfig() = saved;
  if (!ok) { delete S; return fail(SRT_ERR_INVALID, std::string("cannot load ") + gltf_path); }
  if (objects.objects.empty()) { delete S; return fail(SRT_ERR_INVALID, "the glTF file holds no triangles"); }
  S->world.add(make_shared<bvhNode>(objects, 0, 1));
  S->draws = randomDraws();
  if (!S->flat.build(S->world)) { const std::string e = S->flat.error; delete S; return fail(SRT_ERR_INVALID, e); }
  *out = S;
  return SRT_OK;
} catch (const std::bad_alloc&) { return fail(SRT_ERR_NOMEM, "out of memory"); } catch (const std::exception& e_) { return fail(SRT_ERR_INVALID, e_.what()); } catch (...) { return fail(SRT_ERR_INVALID, "unknown exception"); }

const srt_scene_desc* srt_host_scene_desc(const srt_host_scene* s) { return s ? &s->flat.desc() : nullptr; }
uint64_t srt_host_scene_draws(const srt_host_scene* s) { return s ? s->draws : 0; }
void srt_host_scene_free(srt_host_scene* s) { delete s; }

}  // extern "C"
