// Minimal host-side vector/matrix types for the C++ face of the B200 path.
//
// The reference uses LiteMath (external/LiteMath/LiteMath.h) for these.  Only the handful of formulas that decide
// the numbers the kernels receive are restated here, in the same operation order so that the matrices are
// bit-identical (this file is compiled with -ffp-contract=off):
//   normalize   LiteMath.h:814      a * (1/length)
//   cross       LiteMath.h:835-840
//   lookAt      LiteMath.h:1780-1805
//   perspectiveMatrix LiteMath.h:1807-1825
//   inverse4x4  LiteMath.h:1580-1664 (cofactor expansion, scaled by 1/det)
// Layouts match LiteMath's (float3 = 12 B, float4 = 16 B, float4x4 = 4 column float4s, operator()(row, col)), so a
// build inside the reference tree can alias them to the LiteMath types (define RT_HOST_USE_LITEMATH).
#pragma once
#include <cmath>
#include <cstdint>

namespace rtmath {

struct float2 { float x = 0, y = 0; };
struct uint3 { uint32_t x = 0, y = 0, z = 0; };
struct float3 {
  float x = 0, y = 0, z = 0;
  float3() = default;
  float3(float a, float b, float c) : x(a), y(b), z(c) {}
  explicit float3(float a) : x(a), y(a), z(a) {}
  float &operator[](int i) { return (&x)[i]; }
  float operator[](int i) const { return (&x)[i]; }
};
struct float4 {
  float x = 0, y = 0, z = 0, w = 0;
  float4() = default;
  float4(float a, float b, float c, float d) : x(a), y(b), z(c), w(d) {}
  float &operator[](int i) { return (&x)[i]; }
  float operator[](int i) const { return (&x)[i]; }
};

inline float3 operator+(float3 a, float3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline float3 operator-(float3 a, float3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline float3 operator*(float3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
inline float3 operator*(float s, float3 a) { return {s * a.x, s * a.y, s * a.z}; }
inline float3 operator/(float3 a, float s) { return {a.x / s, a.y / s, a.z / s}; }
inline float3 operator-(float3 a) { return {-a.x, -a.y, -a.z}; }
inline float dot(float3 a, float3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline float dot(float4 a, float4 b) { return a.x * b.x + a.y * b.y + a.z * b.z + a.w * b.w; }
inline float length(float3 a) { return std::sqrt(dot(a, a)); }
inline float length(float4 a) { return std::sqrt(dot(a, a)); }
inline float3 normalize(float3 a) {
  float lenInv = 1.0f / std::sqrt(dot(a, a));
  return a * lenInv;
}
inline float3 cross(float3 a, float3 b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
inline float3 to_float3(float4 v) { return {v.x, v.y, v.z}; }
inline float4 to_float4(float3 v, float w) { return {v.x, v.y, v.z, w}; }

struct float4x4 {
  float4 m_col[4];
  float4x4() { identity(); }
  void identity() {
    m_col[0] = {1, 0, 0, 0};
    m_col[1] = {0, 1, 0, 0};
    m_col[2] = {0, 0, 1, 0};
    m_col[3] = {0, 0, 0, 1};
  }
  float &operator()(int row, int col) { return m_col[col][row]; }
  float operator()(int row, int col) const { return m_col[col][row]; }
  const float *data() const { return &m_col[0].x; }
};

inline float4x4 lookAt(float3 eye, float3 center, float3 up) {
  float3 z = normalize(float3{eye.x - center.x, eye.y - center.y, eye.z - center.z});
  float3 y = up;
  float3 x = cross(y, z);
  y = cross(z, x);
  x = normalize(x);
  y = normalize(y);
  float4x4 M;
  M.m_col[0] = {x.x, y.x, z.x, 0.0f};
  M.m_col[1] = {x.y, y.y, z.y, 0.0f};
  M.m_col[2] = {x.z, y.z, z.z, 0.0f};
  M.m_col[3] = {-x.x * eye.x - x.y * eye.y - x.z * eye.z, -y.x * eye.x - y.y * eye.y - y.z * eye.z,
                -z.x * eye.x - z.y * eye.y - z.z * eye.z, 1.0f};
  return M;
}

inline float4x4 perspectiveMatrix(float fovy, float aspect, float zNear, float zFar) {
  const float ymax = zNear * tanf(fovy * 3.14159265358979323846f / 360.0f);
  const float xmax = ymax * aspect;
  const float left = -xmax, right = +xmax, bottom = -ymax, top = +ymax;
  const float temp = 2.0f * zNear, temp2 = right - left, temp3 = top - bottom, temp4 = zFar - zNear;
  float4x4 r;
  r.m_col[0] = {temp / temp2, 0.0f, 0.0f, 0.0f};
  r.m_col[1] = {0.0f, temp / temp3, 0.0f, 0.0f};
  r.m_col[2] = {(right + left) / temp2, (top + bottom) / temp3, (-zFar - zNear) / temp4, -1.0f};
  r.m_col[3] = {0.0f, 0.0f, (-temp * zFar) / temp4, 0.0f};
  return r;
}

// Cofactor inverse.  Each output element is (three products summed) minus (three products summed) of 2x2
// sub-determinant pairs; the pair tables below enumerate them in the order LiteMath.h:1580-1664 evaluates them.
inline float4x4 inverse4x4(const float4x4 &a) {
  float4x4 m;
  float p[12];
  // pairs of rows 2,3 (for output rows taken from input rows 1 and 0)
  auto pairs = [&](int r0, int r1) {
    static const int c[12][2] = {{2, 3}, {3, 2}, {1, 3}, {3, 1}, {1, 2}, {2, 1}, {0, 3}, {3, 0}, {0, 2}, {2, 0}, {0, 1}, {1, 0}};
    for (int i = 0; i < 12; ++i) p[i] = a(r0, c[i][0]) * a(r1, c[i][1]);
  };
  // element = (p[i0]*a(r,c0) + p[i1]*a(r,c1) + p[i2]*a(r,c2)) - (p[j0]*a(r,c0') + ...)
  struct Term { int out_r, out_c, src_r; int pi[3], pc[3], ni[3], nc[3]; };
  static const Term first[8] = {
      {0, 0, 1, {0, 3, 4}, {1, 2, 3}, {1, 2, 5}, {1, 2, 3}},   {1, 0, 1, {1, 6, 9}, {0, 2, 3}, {0, 7, 8}, {0, 2, 3}},
      {2, 0, 1, {2, 7, 10}, {0, 1, 3}, {3, 6, 11}, {0, 1, 3}}, {3, 0, 1, {5, 8, 11}, {0, 1, 2}, {4, 9, 10}, {0, 1, 2}},
      {0, 1, 0, {1, 2, 5}, {1, 2, 3}, {0, 3, 4}, {1, 2, 3}},   {1, 1, 0, {0, 7, 8}, {0, 2, 3}, {1, 6, 9}, {0, 2, 3}},
      {2, 1, 0, {3, 6, 11}, {0, 1, 3}, {2, 7, 10}, {0, 1, 3}}, {3, 1, 0, {4, 9, 10}, {0, 1, 2}, {5, 8, 11}, {0, 1, 2}}};
  pairs(2, 3);
  for (const Term &t : first) {
    float v = p[t.pi[0]] * a(t.src_r, t.pc[0]) + p[t.pi[1]] * a(t.src_r, t.pc[1]) + p[t.pi[2]] * a(t.src_r, t.pc[2]);
    v -= p[t.ni[0]] * a(t.src_r, t.nc[0]) + p[t.ni[1]] * a(t.src_r, t.nc[1]) + p[t.ni[2]] * a(t.src_r, t.nc[2]);
    m(t.out_r, t.out_c) = v;
  }
  static const Term second[8] = {
      {0, 2, 3, {0, 3, 4}, {1, 2, 3}, {1, 2, 5}, {1, 2, 3}},   {1, 2, 3, {1, 6, 9}, {0, 2, 3}, {0, 7, 8}, {0, 2, 3}},
      {2, 2, 3, {2, 7, 10}, {0, 1, 3}, {3, 6, 11}, {0, 1, 3}}, {3, 2, 3, {5, 8, 11}, {0, 1, 2}, {4, 9, 10}, {0, 1, 2}},
      {0, 3, 2, {2, 5, 1}, {2, 3, 1}, {4, 0, 3}, {3, 1, 2}},   {1, 3, 2, {8, 0, 7}, {3, 0, 2}, {6, 9, 1}, {2, 3, 0}},
      {2, 3, 2, {6, 11, 3}, {1, 3, 0}, {10, 2, 7}, {3, 0, 1}}, {3, 3, 2, {10, 4, 9}, {2, 0, 1}, {8, 11, 5}, {1, 2, 0}}};
  pairs(0, 1);
  for (const Term &t : second) {
    float v = p[t.pi[0]] * a(t.src_r, t.pc[0]) + p[t.pi[1]] * a(t.src_r, t.pc[1]) + p[t.pi[2]] * a(t.src_r, t.pc[2]);
    v -= p[t.ni[0]] * a(t.src_r, t.nc[0]) + p[t.ni[1]] * a(t.src_r, t.nc[1]) + p[t.ni[2]] * a(t.src_r, t.nc[2]);
    m(t.out_r, t.out_c) = v;
  }
  const float k = 1.0f / (a(0, 0) * m(0, 0) + a(0, 1) * m(1, 0) + a(0, 2) * m(2, 0) + a(0, 3) * m(3, 0));
  for (int c = 0; c < 4; ++c) {
    m.m_col[c].x *= k;
    m.m_col[c].y *= k;
    m.m_col[c].z *= k;
    m.m_col[c].w *= k;
  }
  return m;
}

}  // namespace rtmath
