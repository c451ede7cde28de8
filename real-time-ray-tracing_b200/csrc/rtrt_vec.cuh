// Device vector maths for the shading code.  Operation order and the compensated dot / cross follow
// src/linearMath.h (FMA :41-43, dop :47-53, TwoProd/TwoSum/InnerProduct :64-91, cross :468, dot :484) so
// that results agree with the CPU oracle to the last few ulps (TU is built with --fmad=false; fused
// multiply-adds appear only where the reference writes FMA).
#pragma once
#include "rtrt_types.h"
#include "bvh_block.cuh"   // sel_min / sel_max
#include <cuda_runtime.h>

namespace rtrt {

#define RTRT_PI_OVER_4 0.7853981633974483096156608458198757210492f
#define RTRT_PI_OVER_2 1.5707963267948966192313216916397514420985f
#define RTRT_TWO_PI 6.2831853071795864769252867665590057683943f
#define RTRT_INV_TWO_PI 0.15915494309f
#define RTRT_M_PI 3.1415926535897932384626422832795028841971f   /* src/linearMath.h:14-16 (float on the reference's platform) */
#define RTRT_SAFE_COSINE_EPSI 1e-5f

#define VD __device__ __forceinline__

VD Float3 f3(float a) { return Float3{a, a, a}; }
VD Float3 f3(float a, float b, float c) { return Float3{a, b, c}; }
VD Float3 operator+(Float3 a, Float3 b) { return Float3{a.x + b.x, a.y + b.y, a.z + b.z}; }
VD Float3 operator-(Float3 a, Float3 b) { return Float3{a.x - b.x, a.y - b.y, a.z - b.z}; }
VD Float3 operator*(Float3 a, Float3 b) { return Float3{a.x * b.x, a.y * b.y, a.z * b.z}; }
VD Float3 operator/(Float3 a, Float3 b) { return Float3{a.x / b.x, a.y / b.y, a.z / b.z}; }
VD Float3 operator*(Float3 a, float s) { return Float3{a.x * s, a.y * s, a.z * s}; }
VD Float3 operator*(float s, Float3 a) { return Float3{a.x * s, a.y * s, a.z * s}; }
VD Float3 operator/(Float3 a, float s) { return Float3{a.x / s, a.y / s, a.z / s}; }
VD Float3 operator+(Float3 a, float s) { return Float3{a.x + s, a.y + s, a.z + s}; }
VD Float3 operator-(Float3 a, float s) { return Float3{a.x - s, a.y - s, a.z - s}; }
VD Float3 operator-(Float3 a) { return Float3{-a.x, -a.y, -a.z}; }
VD Float2 operator+(Float2 a, Float2 b) { return Float2{a.x + b.x, a.y + b.y}; }
VD Float2 operator-(Float2 a, Float2 b) { return Float2{a.x - b.x, a.y - b.y}; }
VD Float2 operator*(Float2 a, Float2 b) { return Float2{a.x * b.x, a.y * b.y}; }
VD Float2 operator/(Float2 a, Float2 b) { return Float2{a.x / b.x, a.y / b.y}; }
VD Float2 operator*(Float2 a, float s) { return Float2{a.x * s, a.y * s}; }
VD Float2 operator+(Float2 a, float s) { return Float2{a.x + s, a.y + s}; }
VD Float2 operator-(Float2 a, float s) { return Float2{a.x - s, a.y - s}; }
VD Float2 operator/(Float2 a, float s) { return Float2{a.x / s, a.y / s}; }
VD Float4 operator*(Float4 a, float s) { return Float4{a.x * s, a.y * s, a.z * s, a.w * s}; }
VD Float4 operator+(Float4 a, Float4 b) { return Float4{a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w}; }
VD Float4 operator/(Float4 a, float s) { return Float4{a.x / s, a.y / s, a.z / s, a.w / s}; }

VD float dopf(float a, float b, float c, float d) {
    const float cd = c * d;
    const float err = __fmaf_rn(-c, d, cd);
    const float r = __fmaf_rn(a, b, -cd);
    return r + err;
}
struct CFloat { float v, err; };
VD CFloat two_prod(float a, float b) { const float ab = a * b; return CFloat{ab, __fmaf_rn(a, b, -ab)}; }
VD CFloat two_sum(float a, float b) { const float s = a + b, d = s - a; return CFloat{s, (a - (s - d)) + (b - d)}; }
VD float dot(Float3 a, Float3 b) {
    const CFloat zz = two_prod(a.z, b.z), yy = two_prod(a.y, b.y);
    const CFloat s1 = two_sum(yy.v, zz.v);
    const CFloat t1{s1.v, yy.err + (zz.err + s1.err)};
    const CFloat xx = two_prod(a.x, b.x);
    const CFloat s0 = two_sum(xx.v, t1.v);
    return s0.v + (xx.err + (t1.err + s0.err));
}
VD Float3 cross(Float3 a, Float3 b) { return Float3{dopf(a.y, b.z, a.z, b.y), dopf(a.z, b.x, a.x, b.z), dopf(a.x, b.y, a.y, b.x)}; }
VD Float3 normalize(Float3 v) { const float n = sqrtf(v.x * v.x + v.y * v.y + v.z * v.z); return Float3{v.x / n, v.y / n, v.z / n}; }
VD Float3 abs3(Float3 v) { return Float3{fabsf(v.x), fabsf(v.y), fabsf(v.z)}; }
VD float getmax3(Float3 v) { return sel_max(sel_max(v.x, v.y), v.z); }
VD Float3 max3f(Float3 a, Float3 b) { return Float3{sel_max(a.x, b.x), sel_max(a.y, b.y), sel_max(a.z, b.z)}; }
VD float clampf(float a, float lo = 0.0f, float hi = 1.0f) { return a < lo ? lo : a > hi ? hi : a; }
VD int clampi(int a, int lo, int hi) { return a < lo ? lo : a > hi ? hi : a; }
VD float max1f(float a, float b) { return (a < b) ? b : a; }          // src/linearMath.h:247
VD bool isnan3(Float3 v) { return v.x != v.x || v.y != v.y || v.z != v.z; }
VD Float3 xyz(Float4 v) { return Float3{v.x, v.y, v.z}; }

// column-major Mat3 with compensated row products (src/linearMath.h:506-539)
struct Mat3 { float m00, m10, m20, m01, m11, m21, m02, m12, m22; };
VD Mat3 mat3_cols(Float3 a, Float3 b, Float3 c) { return Mat3{a.x, a.y, a.z, b.x, b.y, b.z, c.x, c.y, c.z}; }
VD Mat3 transpose(Mat3 m) { Mat3 r = m; r.m01 = m.m10; r.m10 = m.m01; r.m20 = m.m02; r.m02 = m.m20; r.m21 = m.m12; r.m12 = m.m21; return r; }
VD Float3 mul(const Mat3& m, Float3 v) {
    return Float3{dot(Float3{m.m00, m.m01, m.m02}, v), dot(Float3{m.m10, m.m11, m.m12}, v), dot(Float3{m.m20, m.m21, m.m22}, v)};
}

// src/linearMath.h:725-738
VD void localize_sample(Float3 n, Float3& u, Float3& v) {
    Float3 w = f3(1, 0, 0);
    if (fabsf(n.x) > 0.707f) w = f3(0, 1, 0);
    u = cross(n, w);
    v = cross(n, u);
}

}  // namespace rtrt
