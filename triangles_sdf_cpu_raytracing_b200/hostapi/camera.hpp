// Orbit camera with the reference's public interface (src/camera.hpp:7-61, src/quaternion.hpp:8-70).
// Host-only: it produces the eye position and the view matrix the kernels consume (~35 floats per frame).
// Same formulas in the same order as camera.cpp:5-62 / quaternion.hpp:40-70, so the matrices match the reference's
// bit for bit (tests/test_host_mirror.py compares them with the golden cameras).
#pragma once
#include "rt_math.hpp"

struct Quaternion : rtmath::float4 {
  Quaternion() : rtmath::float4{0.0f, 0.0f, 0.0f, 1.0f} {}
  Quaternion(float x_, float y_, float z_, float w_) : rtmath::float4{x_, y_, z_, w_} {}
  Quaternion conjugate() const { return Quaternion(-x, -y, -z, w); }
  float magnitude() const { return rtmath::length(static_cast<const rtmath::float4 &>(*this)); }
  Quaternion normalized() const {
    float n = magnitude();
    return {x / n, y / n, z / n, w / n};
  }
};

inline Quaternion operator*(const Quaternion &a, const Quaternion &b) {  // quaternion.hpp:40-45
  return Quaternion{a.w * b.x + a.x * b.w + a.y * b.z - a.z * b.y, a.w * b.y - a.x * b.z + a.y * b.w + a.z * b.x,
                    a.w * b.z + a.x * b.y - a.y * b.x + a.z * b.w, a.w * b.w - a.x * b.x - a.y * b.y - a.z * b.z};
}
inline rtmath::float3 rotateVector(const rtmath::float3 &v, const Quaternion &q) {  // quaternion.hpp:47-52
  Quaternion p(v.x, v.y, v.z, 0.0f);
  Quaternion r = q * p * q.conjugate();
  return {r.x, r.y, r.z};
}
inline Quaternion angleAxis(float angle, const rtmath::float3 &axis) {  // quaternion.hpp:54-70
  rtmath::float3 n = rtmath::normalize(axis);
  float half = angle * 0.5f;
  Quaternion q;
  q.w = std::cos(half);
  q.x = n.x * std::sin(half);
  q.y = n.y * std::sin(half);
  q.z = n.z * std::sin(half);
  return q;
}

class Camera {
public:
  Camera() = default;
  Camera(const rtmath::float3 &position, const rtmath::float3 &target,
         const rtmath::float3 &up = rtmath::float3{0.0f, 1.0f, 0.0f})
      : m_position(position), m_target(target) {
    updateOrientation(up);
  }
  void resetPosition(const rtmath::float3 &p) { m_position = p; updateOrientation(up()); }
  void resetTarget(const rtmath::float3 &t) { m_target = t; updateOrientation(up()); }
  void rotate(float dx, float dy) {  // camera.cpp:5-20
    float pitch = dy * m_sensetivity, yaw = dx * m_sensetivity;
    Quaternion qYaw = angleAxis(yaw, up());
    Quaternion qPitch = angleAxis(pitch, right());
    m_orientation = (qPitch * qYaw * m_orientation).normalized();
    updateVectors();
    if (m_lockUp) updateOrientation(m_lockedUp);
  }
  rtmath::float4x4 lookAtMatrix() const { return rtmath::lookAt(m_position, m_target, up()); }
  rtmath::float3 position() const { return m_position; }
  rtmath::float3 target() const { return m_target; }
  rtmath::float3 up() const { return rtmath::normalize(rotateVector({0.0f, 1.0f, 0.0f}, m_orientation)); }
  rtmath::float3 right() const { return rtmath::normalize(rotateVector({1.0f, 0.0f, 0.0f}, m_orientation)); }
  rtmath::float3 forward() const { return rtmath::normalize(target() - position()); }
  float sensetivity() const { return m_sensetivity; }
  void setLockUp(bool to) {
    if (!m_lockUp && to) m_lockedUp = up();
    m_lockUp = to;
  }
  bool isLockedUp() { return m_lockUp; }

private:
  void updateVectors() {  // camera.cpp:22-28
    rtmath::float3 rotatedForward = rtmath::normalize(rotateVector({0.0f, 0.0f, -1.0f}, m_orientation));
    float distance = rtmath::length(m_target - m_position);
    m_position = m_target - rotatedForward * distance;
  }
  void updateOrientation(rtmath::float3 upv) {  // camera.cpp:36-62: rotation matrix -> quaternion (not renormalised:
    auto m = rtmath::lookAt(position(), target(), upv);  // `normalize(m_orientation)` there discards its result)
    float t;
    if (m(2, 2) < 0) {
      if (m(0, 0) > m(1, 1)) {
        t = 1 + m(0, 0) - m(1, 1) - m(2, 2);
        m_orientation = Quaternion(t, m(0, 1) + m(1, 0), m(2, 0) + m(0, 2), m(1, 2) - m(2, 1));
      } else {
        t = 1 - m(0, 0) + m(1, 1) - m(2, 2);
        m_orientation = Quaternion(m(0, 1) + m(1, 0), t, m(1, 2) + m(2, 1), m(2, 0) - m(0, 2));
      }
    } else {
      if (m(0, 0) < -m(1, 1)) {
        t = 1 - m(0, 0) - m(1, 1) + m(2, 2);
        m_orientation = Quaternion(m(2, 0) + m(0, 2), m(1, 2) + m(2, 1), t, m(0, 1) - m(1, 0));
      } else {
        t = 1 + m(0, 0) + m(1, 1) + m(2, 2);
        m_orientation = Quaternion(m(1, 2) - m(2, 1), m(2, 0) - m(0, 2), m(0, 1) - m(1, 0), t);
      }
    }
  }
  rtmath::float3 m_position;
  rtmath::float3 m_target;
  Quaternion m_orientation;
  float m_sensetivity = 0.01f;
  bool m_lockUp = false;
  rtmath::float3 m_lockedUp;
};
