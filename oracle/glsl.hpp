// ORACLE — TEST INFRASTRUCTURE ONLY. Not part of the product path.
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
// legs may load anything under oracle/.
//
// Minimal GLSL-like vector/matrix vocabulary written from scratch (glm is not
// available), so the reference's shader math (src/shaders/**) can be restated
// statement-by-statement in C++.  Semantics pinned here:
//   * every operation is IEEE fp32, evaluated left to right, no FMA contraction
//     (build with -ffp-contract=off);
//   * min/max follow the GLSL spec text ("y if y < x, otherwise x");
//   * normalize(v) = v * inversesqrt(dot(v,v)), inversesqrt(x) = 1/sqrt(x);
//   * mix(a,b,t) = a*(1-t) + b*t;
//   * mat4 is column-major, M*v = c0*v.x + c1*v.y + c2*v.z + c3*v.w.
// PARITY UNPINNED: the reference ships no golden vectors.
#pragma once
#include <cmath>
#include <cstdint>

namespace glsl {

constexpr float PI = 3.1415926535897932384626433832795f;

struct vec2 { float x, y; vec2() : x(0), y(0) {} vec2(float a, float b) : x(a), y(b) {} explicit vec2(float a) : x(a), y(a) {} };
struct vec3 { float x, y, z; vec3() : x(0), y(0), z(0) {} vec3(float a, float b, float c) : x(a), y(b), z(c) {} explicit vec3(float a) : x(a), y(a), z(a) {}
  vec3(vec2 v, float c) : x(v.x), y(v.y), z(c) {} vec2 xy() const { return vec2(x, y); } };
struct vec4 { float x, y, z, w; vec4() : x(0), y(0), z(0), w(0) {} vec4(float a, float b, float c, float d) : x(a), y(b), z(c), w(d) {}
  vec4(vec3 v, float d) : x(v.x), y(v.y), z(v.z), w(d) {} vec3 xyz() const { return vec3(x, y, z); } vec2 xy() const { return vec2(x, y); } };
struct ivec2 { int x, y; ivec2() : x(0), y(0) {} ivec2(int a, int b) : x(a), y(b) {} };

inline ivec2 operator+(ivec2 a, ivec2 b) { return ivec2(a.x + b.x, a.y + b.y); }

// GLSL float->int conversion truncates toward zero.
inline ivec2 to_ivec2(vec2 v) { return ivec2((int)v.x, (int)v.y); }
inline vec2 to_vec2(ivec2 v) { return vec2((float)v.x, (float)v.y); }

#define GLSL_BINOP2(op) \
  inline vec2 operator op(vec2 a, vec2 b) { return vec2(a.x op b.x, a.y op b.y); } \
  inline vec2 operator op(vec2 a, float b) { return vec2(a.x op b, a.y op b); } \
  inline vec2 operator op(float a, vec2 b) { return vec2(a op b.x, a op b.y); }
#define GLSL_BINOP3(op) \
  inline vec3 operator op(vec3 a, vec3 b) { return vec3(a.x op b.x, a.y op b.y, a.z op b.z); } \
  inline vec3 operator op(vec3 a, float b) { return vec3(a.x op b, a.y op b, a.z op b); } \
  inline vec3 operator op(float a, vec3 b) { return vec3(a op b.x, a op b.y, a op b.z); }
#define GLSL_BINOP4(op) \
  inline vec4 operator op(vec4 a, vec4 b) { return vec4(a.x op b.x, a.y op b.y, a.z op b.z, a.w op b.w); } \
  inline vec4 operator op(vec4 a, float b) { return vec4(a.x op b, a.y op b, a.z op b, a.w op b); } \
  inline vec4 operator op(float a, vec4 b) { return vec4(a op b.x, a op b.y, a op b.z, a op b.w); }
GLSL_BINOP2(+) GLSL_BINOP2(-) GLSL_BINOP2(*) GLSL_BINOP2(/)
GLSL_BINOP3(+) GLSL_BINOP3(-) GLSL_BINOP3(*) GLSL_BINOP3(/)
GLSL_BINOP4(+) GLSL_BINOP4(-) GLSL_BINOP4(*) GLSL_BINOP4(/)
inline vec2 operator-(vec2 a) { return vec2(-a.x, -a.y); }
inline vec3 operator-(vec3 a) { return vec3(-a.x, -a.y, -a.z); }
inline vec3& operator+=(vec3& a, vec3 b) { a = a + b; return a; }
inline vec3& operator-=(vec3& a, vec3 b) { a = a - b; return a; }
inline vec3& operator*=(vec3& a, float b) { a = a * b; return a; }
inline vec3& operator*=(vec3& a, vec3 b) { a = a * b; return a; }
inline vec3& operator/=(vec3& a, vec3 b) { a = a / b; return a; }
inline vec3& operator/=(vec3& a, float b) { a = a / b; return a; }
inline vec2& operator*=(vec2& a, float b) { a = a * b; return a; }
inline vec2& operator*=(vec2& a, vec2 b) { a = a * b; return a; }
inline vec4& operator/=(vec4& a, float b) { a = a / b; return a; }

// GLSL spec: min(x,y) = y if y < x else x ; max(x,y) = y if x < y else x.
inline float min(float x, float y) { return (y < x) ? y : x; }
inline float max(float x, float y) { return (x < y) ? y : x; }
inline int min(int x, int y) { return (y < x) ? y : x; }
inline int max(int x, int y) { return (x < y) ? y : x; }
inline vec3 min(vec3 a, vec3 b) { return vec3(min(a.x, b.x), min(a.y, b.y), min(a.z, b.z)); }
inline vec3 max(vec3 a, vec3 b) { return vec3(max(a.x, b.x), max(a.y, b.y), max(a.z, b.z)); }
inline float clamp(float x, float lo, float hi) { return min(max(x, lo), hi); }
inline vec3 clamp(vec3 v, vec3 lo, vec3 hi) { return min(max(v, lo), hi); }
inline float abs(float x) { return std::fabs(x); }
inline vec2 abs(vec2 v) { return vec2(std::fabs(v.x), std::fabs(v.y)); }
inline vec4 abs(vec4 v) { return vec4(std::fabs(v.x), std::fabs(v.y), std::fabs(v.z), std::fabs(v.w)); }
inline float floor(float x) { return std::floor(x); }
inline vec2 floor(vec2 v) { return vec2(std::floor(v.x), std::floor(v.y)); }
inline float fract(float x) { return x - std::floor(x); }
inline float sqrt(float x) { return std::sqrt(x); }
inline float inversesqrt(float x) { return 1.0f / std::sqrt(x); }
inline float mix(float a, float b, float t) { return a * (1.0f - t) + b * t; }
inline vec3 mix(vec3 a, vec3 b, float t) { return a * (1.0f - t) + b * t; }
inline vec4 mix(vec4 a, vec4 b, float t) { return a * (1.0f - t) + b * t; }

inline float dot(vec2 a, vec2 b) { return a.x * b.x + a.y * b.y; }
inline float dot(vec3 a, vec3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline float dot(vec4 a, vec4 b) { return a.x * b.x + a.y * b.y + a.z * b.z + a.w * b.w; }
inline float length(vec2 v) { return std::sqrt(dot(v, v)); }
inline float length(vec3 v) { return std::sqrt(dot(v, v)); }
inline vec2 normalize(vec2 v) { return v * inversesqrt(dot(v, v)); }
inline vec3 normalize(vec3 v) { return v * inversesqrt(dot(v, v)); }
inline vec3 cross(vec3 a, vec3 b) { return vec3(a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y); }
inline vec3 reflect(vec3 I, vec3 N) { return I - 2.0f * dot(N, I) * N; }

struct mat4 {
  vec4 c[4];  // columns
  mat4() {}
  explicit mat4(const float* colmajor16) { for (int i = 0; i < 4; i++) c[i] = vec4(colmajor16[4 * i], colmajor16[4 * i + 1], colmajor16[4 * i + 2], colmajor16[4 * i + 3]); }
};
inline vec4 operator*(const mat4& m, vec4 v) { return m.c[0] * v.x + m.c[1] * v.y + m.c[2] * v.z + m.c[3] * v.w; }

}  // namespace glsl
