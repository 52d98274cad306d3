// C++ face of the B200 path with the reference's own names and signatures (src/raytracing.hpp:9-117), so that a
// caller such as the viewer's frame loop (src/main.cpp:196-208) compiles unchanged:
//
//     state.frameBuf.clear();
//     time = renderer.draw(*pScene, state.frameBuf, state.camera, projInv);
//
// Every class here is a thin owner of an `rt_scene*` / host framebuffer; all arithmetic on rays happens in the CUDA
// kernels behind include/rt_b200.h.  There is no CPU fallback: without a device the constructors / draw throw.
#pragma once
#include <algorithm>
#include <cstdint>
#include <limits>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/rt_b200.h"
#include "camera.hpp"
#include "rt_math.hpp"

namespace rthost {
[[noreturn]] inline void raise(const char *what) { throw std::runtime_error(std::string(what) + ": " + rt_last_error()); }
inline void check(int rc, const char *what) {
  if (rc != RT_OK) raise(what);
}
}  // namespace rthost

// LiteImage::Image2D<T> subset used by FrameBuffer (Image2d.h:67-107): row-major, index y*width + x.
// B200 extension: pin() moves the pixels into page-locked memory (rt_pinned_alloc) and keeps them there across resizes.
template <typename T> class Image2D {
public:
  Image2D() = default;
  Image2D(const Image2D &) = delete;  // owns a page-locked block
  Image2D &operator=(const Image2D &) = delete;
  ~Image2D() { release(); }
  void resize(uint32_t w, uint32_t h) {
    const bool wasPinned = m_pinned != nullptr;
    release();
    m_w = w;
    m_h = h;
    if (wasPinned) allocPinned();
    else m_data.resize(count());
  }
  void clear(T v) { std::fill(data(), data() + count(), v); }
  uint32_t width() const { return m_w; }
  uint32_t height() const { return m_h; }
  T *data() { return m_pinned ? m_pinned : m_data.data(); }
  const T *data() const { return m_pinned ? m_pinned : m_data.data(); }
  T &operator[](size_t i) { return data()[i]; }
  const T &operator[](size_t i) const { return data()[i]; }
  void pin() {
    if (m_pinned || count() == 0) return;
    allocPinned();
    std::copy(m_data.begin(), m_data.end(), m_pinned);
    std::vector<T>().swap(m_data);
  }

private:
  size_t count() const { return static_cast<size_t>(m_w) * m_h; }
  void allocPinned() {
    void *p = nullptr;
    rthost::check(rt_pinned_alloc(count() * sizeof(T), &p), "Image2D::pin");
    m_pinned = static_cast<T *>(p);
  }
  void release() {
    if (m_pinned) rt_pinned_free(m_pinned);
    m_pinned = nullptr;
  }
  uint32_t m_w = 0, m_h = 0;
  std::vector<T> m_data;
  T *m_pinned = nullptr;
};

struct FrameBuffer {  // raytracing.hpp:9-20
  Image2D<uint32_t> color;
  Image2D<float> t;
  void resize(uint32_t newWidth, uint32_t newHeight) {
    color.resize(newWidth, newHeight);
    t.resize(newWidth, newHeight);
  }
  void clear() {
    color.clear(0);
    t.clear(std::numeric_limits<float>::infinity());
  }
  // B200 extension: page-locked storage.  Renderer::draw with frameIsCleared then lets the kernel store its tiles
  // straight into the two images; without pin() the frame travels through pipelined copies.
  void pin() {
    color.pin();
    t.pin();
  }
};

struct HitInfo {  // raytracing.hpp:67-73
  bool hitten = false;
  float t = std::numeric_limits<float>::infinity();
  rtmath::float3 normal = {0.0f, 1.0f, 0.0f};
  rtmath::float3 albedo = {1.0f, 1.0f, 1.0f};
  float reflectiveness = 0.0f;
};

class IScene {  // raytracing.hpp:75-81
public:
  virtual HitInfo intersect(const rtmath::float3 &rayPos, const rtmath::float3 &rayDir, float tNear, float tFar) const = 0;
  virtual ~IScene() {}
  // the uploaded scene behind this object (nullptr: not renderable by the CUDA path)
  virtual const rt_scene *deviceScene() const { return nullptr; }
  // ground plane y of SceneUnion(model, Plane({0,1,0}, y)); only SceneUnion overrides it
  virtual bool groundPlane(float &) const { return false; }
};

// a scene that lives on the GPU; IScene::intersect forwards one ray to rt_intersect_host
class DeviceScene : public IScene {
public:
  HitInfo intersect(const rtmath::float3 &rayPos, const rtmath::float3 &rayDir, float tNear, float tFar) const override {
    uint8_t hit = 0;
    float t = 0, n[3] = {0, 1, 0};
    rthost::check(rt_intersect_host(handle(), &rayPos.x, &rayDir.x, 1, tNear, tFar, 0, 0.0f, &hit, &t, n, nullptr),
                  "IScene::intersect");
    HitInfo h;
    h.hitten = hit != 0;
    if (h.hitten) {
      h.t = t;
      h.normal = {n[0], n[1], n[2]};
    }
    return h;
  }
  const rt_scene *deviceScene() const override { return handle(); }

protected:
  const rt_scene *handle() const {
    if (!m_scene) throw std::runtime_error("scene has not been uploaded (load / perform it first)");
    return m_scene.get();
  }
  void adopt(rt_scene *s) { m_scene = std::shared_ptr<rt_scene>(s, rt_scene_destroy); }
  std::shared_ptr<rt_scene> m_scene;
};

// Plane({0,1,0}, offset): the viewer's checkerboard ground (raytracing.hpp:119-186, main.cpp:189). Only this
// orientation exists on the device path.
class Plane final : public IScene {
public:
  explicit Plane(const rtmath::float3 &normal, float offset = 0) : m_normal(normal), m_offset(offset) {
    if (!(normal.x == 0.0f && normal.y == 1.0f && normal.z == 0.0f))
      throw std::runtime_error("Plane: only the ground plane normal {0,1,0} is supported on the device path");
  }
  HitInfo intersect(const rtmath::float3 &, const rtmath::float3 &, float, float) const override {
    throw std::runtime_error("Plane::intersect: render it through SceneUnion + Renderer::draw");
  }
  float offset() const { return m_offset; }

private:
  rtmath::float3 m_normal;
  float m_offset;
};

class SceneUnion : public IScene {  // raytracing.hpp:83-97, for (device model, ground plane)
public:
  SceneUnion(std::shared_ptr<IScene> pFirst, std::shared_ptr<IScene> pSecond) : m_pFirst(pFirst), m_pSecond(pSecond) {
    m_plane = dynamic_cast<const Plane *>(m_pSecond.get());
    if (!m_pFirst->deviceScene() || !m_plane)
      throw std::runtime_error("SceneUnion: supported form is (device scene, Plane), as main.cpp:189-190 builds it");
  }
  HitInfo intersect(const rtmath::float3 &rayPos, const rtmath::float3 &rayDir, float tNear, float tFar) const override {
    uint8_t hit = 0;
    float t = 0, n[3] = {0, 1, 0};
    uint32_t prim = 0;
    rthost::check(rt_intersect_host(m_pFirst->deviceScene(), &rayPos.x, &rayDir.x, 1, tNear, tFar, 1, m_plane->offset(),
                                    &hit, &t, n, &prim),
                  "SceneUnion::intersect");
    HitInfo h;
    h.hitten = hit != 0;
    if (h.hitten) {
      h.t = t;
      h.normal = {n[0], n[1], n[2]};
      if (prim == RT_PRIM_PLANE) h.reflectiveness = 0.3f;
    }
    return h;
  }
  const rt_scene *deviceScene() const override { return m_pFirst->deviceScene(); }
  bool groundPlane(float &y) const override {
    y = m_plane->offset();
    return true;
  }

private:
  std::shared_ptr<IScene> m_pFirst, m_pSecond;
  const Plane *m_plane = nullptr;
};

enum class ShadingMode { Normal, Lambert, Color };  // raytracing.hpp:99

struct Renderer {  // raytracing.hpp:101-117
  rtmath::float3 lightPos;
  bool enableShadows = true;
  bool enableReflections = true;
  ShadingMode shadingMode = ShadingMode::Lambert;
  // B200 extension: when true, draw() treats the framebuffer as freshly cleared (FrameBuffer::clear fused into the
  // launch: nothing is uploaded). The viewer calls clear() right before draw (main.cpp:197), so it can set this.
  bool frameIsCleared = false;

  // Returns milliseconds like the reference (raytracing.cpp:97-101): the device time of the render kernel.
  float draw(const IScene &scene, FrameBuffer &frameBuffer, const Camera &camera, const rtmath::float4x4 projInv) const {
    const rt_scene *dev = scene.deviceScene();
    if (!dev) throw std::runtime_error("Renderer::draw: this IScene has no device scene (no CPU fallback exists)");
    rt_frame f{};
    rtmath::float3 pos = camera.position();
    rtmath::float4x4 viewInv = rtmath::inverse4x4(camera.lookAtMatrix());  // raytracing.cpp:73-75
    f.ray_pos[0] = pos.x; f.ray_pos[1] = pos.y; f.ray_pos[2] = pos.z;
    for (int i = 0; i < 16; ++i) {
      f.view_inv[i] = viewInv.data()[i];
      f.proj_inv[i] = projInv.data()[i];
    }
    f.width = frameBuffer.color.width();
    f.height = frameBuffer.color.height();
    f.shading_mode = static_cast<int>(shadingMode);
    f.light_pos[0] = lightPos.x; f.light_pos[1] = lightPos.y; f.light_pos[2] = lightPos.z;
    f.enable_shadows = enableShadows;
    f.enable_reflections = enableReflections;
    float y = 0.0f;
    f.has_plane = scene.groundPlane(y);
    f.plane_y = y;
    f.clear = frameIsCleared;
    float deviceMs = 0.0f;
    rthost::check(rt_render_host(dev, &f, frameBuffer.color.data(), frameBuffer.t.data(), nullptr, &deviceMs, nullptr, nullptr),
                  "Renderer::draw");
    return deviceMs;
  }
};
