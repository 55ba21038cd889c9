// examples/main.cpp — the reference's main.cpp with the pixel loop handed to the B200.
//
// Everything up to "world" is the reference's own set-up code (main.cpp:156-187) compiled against include/srt_world.hpp
// instead of the reference's headers; the loop main.cpp:200-227 (and the compiled-out glDevice calls around it,
// main.cpp:190-199,228-232) becomes b200Device::init / rtFrame / terminate; the output stage main.cpp:235-237
// (writeColorTarget + stbi_write_png) becomes the device's resolve + srt_png_write.
//
//   g++ -std=c++17 -Iinclude examples/main.cpp -Lsexy-raytracer_b200 -lsrt_host -o sexy-raytracer
//   SRT_B200_LIBRARY=sexy-raytracer_b200/libsrt_b200.so ./sexy-raytracer [data dir | shirley] [height] [samples] [out.png] [passes] [seed]
// passes > 1 renders progressively (the reference's `while (glDevice.rtFrame(...))` loop, main.cpp:195-199): samples/passes
// more samples per pixel per pass into a film on the device, a checkpoint <out.png>.srtf after every pass; an existing
// checkpoint of the same size is resumed.
// "shirley" renders the sphere scene of the commented block main.cpp:92-122 (needs no asset files) with the camera of
// the commented line main.cpp:160 and depth 10 (BASELINE config 1).
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <iostream>
#include <string>
#include <vector>

#include "srt_world.hpp"

using namespace srt;

namespace srt { hittableList randomScene(const std::string& dataDir, std::vector<shared_ptr<model>>& keepAlive, int which); }

int main(int argc, char** argv) {
  const bool  shirley = argc > 1 && std::string(argv[1]) == "shirley";
  const std::string dataDir = argc > 1 ? std::string(argv[1]) + "/" : "../data/";
  // camera (main.cpp:157-172)
  const auto  aspect = 16.0f / 9.0f;
  vec3f       eye = shirley ? vec3f(13.0f, 2.0f, 3.0f) : vec3f(0.0f, 3.0f, 5.0f);
  vec3f       lookAt = shirley ? vec3f(0, 0, 0) : vec3f(0, 2.5f, 0);
  vec3f       vUp(0, 1.0f, 0);
  auto        distToFocus = 10.0f;
  auto        aperture = 0.1f;
  color3f     background(0.53f, 0.81f, 0.92f);
  camera      mainCamera(eye, lookAt, vUp, shirley ? 20.0f : 70.0f, aspect, aperture, distToFocus, 0, 1.0f);

  // image (main.cpp:174-184)
  const int   imageHeight = argc > 2 ? atoi(argv[2]) : 720;
  const int   imageWidth = static_cast<int>(imageHeight * aspect);
  const int   numSamples = argc > 3 ? atoi(argv[3]) : 5000;
  const int   maxBounce = shirley ? 10 : 4;
  const char* outFile = argc > 4 ? argv[4] : "test.png";
  uint8_t*    target = static_cast<uint8_t*>(malloc(sizeof(uint8_t) * 4 * imageWidth * imageHeight));

  // world (main.cpp:187)
  std::vector<shared_ptr<model>> models;
  hittableList world = randomScene(dataDir, models, shirley ? 1 : 3);
  if (world.objects.empty()) { std::cerr << "cannot build the scene from " << dataDir << "\n"; return 1; }
  std::cout << "P3\n" << imageWidth << ' ' << imageHeight << "\n255\n";

  b200Device device;                                            // glDevice glDevice;             main.cpp:191
  if (!device.init(imageWidth, imageHeight, world)) {           // glDevice.init(...)             main.cpp:192
    std::cerr << "device: " << device.lastError() << "\n";
    return 1;
  }
  std::cerr << "rendering on " << device.gpuCount() << " GPU(s)\n";
  srt_stats stats;
  const int passes = argc > 5 ? atoi(argv[5]) : 1;
  const uint64_t seed = argc > 6 ? strtoull(argv[6], nullptr, 10) : 1;
  if (passes <= 1) {
    if (!device.rtFrame(target, imageWidth, imageHeight, mainCamera, background, numSamples, maxBounce, seed, &stats)) {   // main.cpp:197-227
      std::cerr << "device: " << device.lastError() << "\n";
      return 1;
    }
  } else {
    const std::string ckpt = std::string(outFile) + ".srtf";
    if (device.loadCheckpoint(ckpt.c_str(), imageWidth, imageHeight))
      std::cerr << "resuming " << ckpt << " at " << device.samplesDone() << " samples per pixel\n";
    while (device.samplesDone() < numSamples) {                 // while (glDevice.rtFrame(...))  main.cpp:195-199
      const int k = std::min((numSamples + passes - 1) / passes, numSamples - device.samplesDone());
      if (!device.rtFrameProgressive(target, imageWidth, imageHeight, mainCamera, background, k, maxBounce, seed, &stats) ||
          !device.saveCheckpoint(ckpt.c_str())) {
        std::cerr << "device: " << device.lastError() << "\n";
        return 1;
      }
      std::cerr << "\r" << device.samplesDone() << " / " << numSamples << " samples per pixel " << std::flush;
    }
  }
  device.terminate();                                           // main.cpp:231

  if (srt_png_write(outFile, imageWidth, imageHeight, 4, target, 4 * imageWidth) != SRT_OK) {    // main.cpp:237
    std::cerr << srt_host_last_error() << "\n";
    return 1;
  }
  free(target);
  std::cerr << "\nDone. " << stats.segments << " path segments in " << stats.kernel_ms << " ms on the device ("
            << stats.segments / stats.kernel_ms / 1e3 << " M segments/s)\n";
  return 0;
}
