// Minimal stand-in for <glm/glm.hpp> (glm is not in this image): the handful of types the reference's camera and
// colour-mapping TESTS use, convertible to and from this repo's Vec3 / Mat4 (which replace glm in the headless host
// layer, include/nbody/nbody.hpp).  Column-major mat4 like glm's.  Test infrastructure only (see gtest/gtest.h here).
#pragma once
#include <cmath>
#include <ostream>

#include "nbody/nbody.hpp"

namespace glm {

struct vec3 {
  union {
    struct { float x, y, z; };
    struct { float r, g, b; };
  };
  vec3() : x(0), y(0), z(0) {}
  vec3(float a, float b_, float c) : x(a), y(b_), z(c) {}
  explicit vec3(float s) : x(s), y(s), z(s) {}
  vec3(const nbody::Vec3& v) : x(v.x), y(v.y), z(v.z) {}
  operator nbody::Vec3() const { return nbody::Vec3(x, y, z); }
  vec3 operator+(const vec3& o) const { return {x + o.x, y + o.y, z + o.z}; }
  vec3 operator-(const vec3& o) const { return {x - o.x, y - o.y, z - o.z}; }
  vec3 operator*(float s) const { return {x * s, y * s, z * s}; }
  bool operator==(const vec3& o) const { return x == o.x && y == o.y && z == o.z; }
  bool operator!=(const vec3& o) const { return !(*this == o); }
};
inline std::ostream& operator<<(std::ostream& os, const vec3& v) { return os << "(" << v.x << ", " << v.y << ", " << v.z << ")"; }
inline float length(const vec3& v) { return std::sqrt(v.x * v.x + v.y * v.y + v.z * v.z); }
inline float dot(const vec3& a, const vec3& b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline vec3 normalize(const vec3& v) { const float l = length(v); return {v.x / l, v.y / l, v.z / l}; }

struct vec4 {
  union {
    struct { float x, y, z, w; };
    struct { float r, g, b, a; };
  };
  vec4() : x(0), y(0), z(0), w(0) {}
  vec4(float a_, float b_, float c, float d) : x(a_), y(b_), z(c), w(d) {}
  vec4(const vec3& v, float d) : x(v.x), y(v.y), z(v.z), w(d) {}
  vec4& operator/=(float s) { x /= s; y /= s; z /= s; w /= s; return *this; }
  vec4 operator/(float s) const { return {x / s, y / s, z / s, w / s}; }
  float& operator[](int i) { return (&x)[i]; }
  float operator[](int i) const { return (&x)[i]; }
};

struct mat4 {
  vec4 col[4];
  mat4() = default;
  mat4(const nbody::Mat4& m) {
    for (int c = 0; c < 4; c++)
      for (int r_ = 0; r_ < 4; r_++) col[c][r_] = m.at(c, r_);
  }
  vec4& operator[](int c) { return col[c]; }
  const vec4& operator[](int c) const { return col[c]; }
  vec4 operator*(const vec4& v) const {
    vec4 o;
    for (int r_ = 0; r_ < 4; r_++) o[r_] = col[0][r_] * v.x + col[1][r_] * v.y + col[2][r_] * v.z + col[3][r_] * v.w;
    return o;
  }
};

}  // namespace glm
