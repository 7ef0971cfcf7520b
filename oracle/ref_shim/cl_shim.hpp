// cl_shim.hpp — TEST INFRASTRUCTURE (oracle/), never linked into the product.
//
// The minimum of OpenCL C 1.x needed to compile the reference's kernel file
// /root/reference/src/sphFluid.cl *unchanged in its arithmetic* as C++17
// (SURVEY.md §8c).  Only one mechanical rewrite is applied to the text by
// oracle/build_ref.py: vector literals `(float4)( a,b,c,d )` become constructor
// calls `float4( a,b,c,d )`.
//
// Definitions where OpenCL leaves latitude (recorded in DESIGN.md "Oracle"):
//   native_sqrt(x), sqrt(x)  := correctly rounded sqrtf
//   native_divide(a,b), a/b  := IEEE division
//   dot(float4,float4)       := ((x*x + y*y) + z*z) + w*w, no contraction
//   select(a,b,c) (scalar)   := c ? b : a
//   vector `<`               := -1 per true lane, 0 per false lane
//   (int)float               := C truncation
// Build flags: -O2 -ffp-contract=off -fno-fast-math (the reference builds its
// kernels with -cl-opt-disable: no FMA contraction, owOpenCLSolver.cpp:196-201).
#pragma once
#include <cmath>
#include <cstdio>
#include <algorithm>

typedef unsigned int uint;

struct float2 { float x, y; };
struct uint2 { uint x, y; };

struct int4 {
  int x, y, z, w;
  int4() : x(0), y(0), z(0), w(0) {}
  int4(int a, int b, int c, int d) : x(a), y(b), z(c), w(d) {}
};
inline int4 operator+(int4 a, int4 b) { return int4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
inline int4 operator*(int s, int4 a) { return int4(s * a.x, s * a.y, s * a.z, s * a.w); }

struct float4 {
  float x, y, z, w;
  float4() : x(0.f), y(0.f), z(0.f), w(0.f) {}
  float4(float s) : x(s), y(s), z(s), w(s) {}  // OpenCL scalar -> vector widening (sphFluid.cl:1127)
  float4(float a, float b, float c, float d) : x(a), y(b), z(c), w(d) {}
  float4 &operator+=(float4 b) { x += b.x; y += b.y; z += b.z; w += b.w; return *this; }
  float4 &operator*=(float s) { x *= s; y *= s; z *= s; w *= s; return *this; }
  float4 &operator/=(float s) { x /= s; y /= s; z /= s; w /= s; return *this; }
};
inline float4 operator+(float4 a, float4 b) { return float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
inline float4 operator-(float4 a, float4 b) { return float4(a.x - b.x, a.y - b.y, a.z - b.z, a.w - b.w); }
inline float4 operator-(float4 a) { return float4(-a.x, -a.y, -a.z, -a.w); }
inline float4 operator*(float4 a, float s) { return float4(a.x * s, a.y * s, a.z * s, a.w * s); }
inline float4 operator*(float s, float4 a) { return float4(s * a.x, s * a.y, s * a.z, s * a.w); }
inline float4 operator/(float4 a, float s) { return float4(a.x / s, a.y / s, a.z / s, a.w / s); }
inline int4 operator<(float4 a, float s) {
  return int4(a.x < s ? -1 : 0, a.y < s ? -1 : 0, a.z < s ? -1 : 0, a.w < s ? -1 : 0);
}
inline float dot(float4 a, float4 b) { return ((a.x * b.x + a.y * b.y) + a.z * b.z) + a.w * b.w; }

inline float native_sqrt(float v) { return sqrtf(v); }
inline float native_divide(float a, float b) { return a / b; }
template <class T, class C> inline T select(T a, T b, C c) { return c ? b : a; }
using std::max;
using std::min;

// One work-item at a time: the driver sets this before each kernel call.
extern thread_local int cl_shim_global_id;
inline int get_global_id(int) { return cl_shim_global_id; }

#define __kernel
#define __global
#define cl_khr_fp64 1
