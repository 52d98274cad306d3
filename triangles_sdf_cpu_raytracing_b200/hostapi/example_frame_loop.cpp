// The viewer's frame loop (src/main.cpp:89-91, 196-208) without the window: same classes, same calls.
//   example_frame_loop <scene.grid|.octree|.obj> <W> <H> <normal|lambert|color> <plane 0|1> <out.raw> [orbit steps [pinned]]
// "pinned": FrameBuffer::pin() + Renderer::frameIsCleared, the page-locked framebuffer the kernel stores straight into.
// out.raw = W*H uint32 colour followed by W*H float depth (the FrameBuffer contents after draw).
// RT_DEVICES=0,1,2,... in the environment lists the GPUs (rt_init): the same single renderer.draw call then renders
// every frame on all of them.
// An output name ending in .ppm writes the colour buffer as a binary PPM instead (the headless stand-in for the
// viewer's SDL blit, main.cpp:238-249); with orbit steps one file per frame: name_000.ppm, name_001.ppm, ...
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <memory>
#include <string>
#include <vector>

#include "scenes.hpp"

static bool ends_with(const std::string &s, const char *suf) {
  size_t n = std::strlen(suf);
  return s.size() >= n && s.compare(s.size() - n, n, suf) == 0;
}

// FrameBuffer::color is R,G,B,A in memory (SDL_PIXELFORMAT_ABGR8888, main.cpp:86); memory row 0 is the top of the image
static void write_ppm(const std::string &name, const FrameBuffer &fb, uint32_t W, uint32_t H) {
  std::ofstream out(name, std::ios::binary);
  out << "P6\n" << W << " " << H << "\n255\n";
  std::string row(static_cast<size_t>(W) * 3, '\0');
  const uint32_t *c = fb.color.data();
  for (uint32_t y = 0; y < H; ++y) {
    for (uint32_t x = 0; x < W; ++x) {
      const uint32_t p = c[static_cast<size_t>(y) * W + x];
      row[3 * x] = static_cast<char>(p & 255u);
      row[3 * x + 1] = static_cast<char>((p >> 8) & 255u);
      row[3 * x + 2] = static_cast<char>((p >> 16) & 255u);
    }
    out.write(row.data(), static_cast<std::streamsize>(row.size()));
  }
}

int main(int argc, char **argv) {
  if (argc < 7) {
    std::fprintf(stderr, "usage: %s scene W H normal|lambert|color plane out.raw [orbit steps]\n", argv[0]);
    return 2;
  }
  try {
    const std::string path = argv[1];
    const uint32_t W = std::atoi(argv[2]), H = std::atoi(argv[3]);
    const std::string mode = argv[4];
    const bool plane = std::atoi(argv[5]) != 0;
    const int orbit = argc > 7 ? std::atoi(argv[7]) : 0;
    const bool pinned = argc > 8 && std::strcmp(argv[8], "pinned") == 0;
    const std::string outName = argv[6];
    const bool ppm = ends_with(outName, ".ppm");

    if (const char *list = std::getenv("RT_DEVICES")) {  // before any scene is created
      std::vector<int> devs;
      for (const char *p = list; *p;) {
        devs.push_back(std::atoi(p));
        while (*p && *p != ',') ++p;
        if (*p == ',') ++p;
      }
      if (rt_init(devs.data(), static_cast<int>(devs.size())) != RT_OK) throw std::runtime_error(rt_last_error());
    }
    std::shared_ptr<IScene> pScene;  // main.cpp:167-191
    float planeY = -1.0f;
    if (ends_with(path, ".grid")) {
      auto g = std::make_shared<SDFGrid>();
      loadSDFGrid(*g, path);
      pScene = g;
    } else if (ends_with(path, ".octree")) {
      auto o = std::make_shared<SDFOctree>();
      loadSDFOctree(*o, path);
      pScene = o;
    } else {
      auto b = std::make_shared<BVHBuilder>();
      b->perform(loadAndScale(path));
      planeY = b->modelMinY();
      pScene = b;
    }
    std::shared_ptr<IScene> fullScene =
        std::make_shared<SceneUnion>(pScene, std::make_shared<Plane>(rtmath::float3{0.0f, 1.0f, 0.0f}, planeY));

    FrameBuffer frameBuf;
    frameBuf.resize(W, H);
    Camera camera({0.0f, 0.0f, 2.5f}, {0.0f, 0.0f, 0.0f}, {0.0f, 1.0f, 0.0f});  // main.cpp:89-91
    camera.setLockUp(true);
    Renderer renderer;
    renderer.lightPos = {2.0f, 2.0f, 2.0f};  // main.cpp:60
    renderer.shadingMode = mode == "normal" ? ShadingMode::Normal : mode == "color" ? ShadingMode::Color : ShadingMode::Lambert;
    renderer.enableShadows = plane;
    renderer.enableReflections = plane;
    if (pinned) {
      frameBuf.pin();
      renderer.frameIsCleared = true;  // clear() below runs right before every draw (main.cpp:197)
    }

    float time = 0.0f;
    for (int frame = 0; frame <= orbit; ++frame) {
      // one 64th of a full orbit per frame: rotate() multiplies by the 0.01 sensitivity (camera.hpp:55)
      if (frame > 0) camera.rotate(static_cast<float>(2.0 * 3.14159265358979323846 / 64.0 / 0.01), 0.0f);
      frameBuf.clear();  // main.cpp:197
      auto proj = rtmath::perspectiveMatrix(45.0f, static_cast<float>(W) / static_cast<float>(H), 0.01f, 100.0f);
      auto projInv = rtmath::inverse4x4(proj);  // main.cpp:198-201
      time = renderer.draw(plane ? *fullScene : *pScene, frameBuf, camera, projInv);  // main.cpp:203 / 206
      if (ppm && orbit > 0) {
        char suffix[16];
        std::snprintf(suffix, sizeof suffix, "_%03d.ppm", frame);
        write_ppm(outName.substr(0, outName.size() - 4) + suffix, frameBuf, W, H);
      }
    }
    if (ppm) {
      if (orbit == 0) write_ppm(outName, frameBuf, W, H);
    } else {
      std::ofstream out(outName, std::ios::binary);
      out.write(reinterpret_cast<const char *>(frameBuf.color.data()), static_cast<std::streamsize>(W) * H * 4);
      out.write(reinterpret_cast<const char *>(frameBuf.t.data()), static_cast<std::streamsize>(W) * H * 4);
    }
    std::printf("Render Time: %.3f ms\n", time);
    return 0;
  } catch (const std::exception &e) {
    std::fprintf(stderr, "error: %s\n", e.what());
    return 1;
  }
}
