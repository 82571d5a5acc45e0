// pt_math.h -- float3 arithmetic shared by the CUDA kernels and (compiled by g++) the CPU restatement.
// Mirrors the operator set of the reference's vec3 (vec3.h:11-199): plain fp32, no fast-math.
#pragma once
#include <math.h>
#include <stdint.h>

#if defined(__CUDACC__)
#define PT_HD __host__ __device__ __forceinline__
#define PT_D __device__ __forceinline__
// rarely executed, large (sinf / cosf / the Perlin loop): kept out of line so that the hot code of the kernels stays
// small and contiguous -- they are instruction-cache bound (ncu: no_instruction 13-29 % of stall samples)
#define PT_HD_COLD static __host__ __device__ __noinline__
#else
#define PT_HD inline
#define PT_HD_COLD static inline
#endif

#define PT_PI_F 3.14159265358979323846f
#define PT_FLT_MAX 3.402823466e+38f

namespace pt {

struct V3 {
    float x, y, z;
};

PT_HD V3 v3(float x, float y, float z)
{
    V3 r;
    r.x = x;
    r.y = y;
    r.z = z;
    return r;
}
PT_HD V3 operator+(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
PT_HD V3 operator-(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
PT_HD V3 operator-(V3 a) { return v3(-a.x, -a.y, -a.z); }
PT_HD V3 operator*(V3 a, V3 b) { return v3(a.x * b.x, a.y * b.y, a.z * b.z); }
PT_HD V3 operator*(float t, V3 a) { return v3(t * a.x, t * a.y, t * a.z); }
PT_HD V3 operator*(V3 a, float t) { return v3(t * a.x, t * a.y, t * a.z); }
// a / b, IEEE.  On the GPU an fp32 division is a short FFMA sequence guarded by FCHK, which sends every ZERO numerator
// to a ~50-instruction subroutine -- and zeros are everywhere here: normals of axis-aligned surfaces, the plane
// distance of a ray that starts on the plane.  (+-0) / b is +-0 with the xor of the signs for any non-zero, non-NaN
// b, so those lanes skip the division (ncu: 16 % of k_nee's instructions were that subroutine).
PT_HD float fdiv(float a, float b)
{
#if defined(__CUDA_ARCH__) && defined(PT_FAST_DIV)
    return __fdividef(a, b); // A/B: approximate division (2 ulp), no FCHK slow path
#elif defined(__CUDA_ARCH__)
    // branch-free: such a lane divides b by b (fast path) and the quotient is replaced by the signed zero
    const bool zero = a == 0.f && b == b && b != 0.f;
    const float q = (zero ? b : a) / b;
    return zero ? __int_as_float((__float_as_int(a) ^ __float_as_int(b)) & (int)0x80000000) : q;
#else
    return a / b;
#endif
}
PT_HD V3 operator/(V3 a, float t) { return v3(fdiv(a.x, t), fdiv(a.y, t), fdiv(a.z, t)); }
PT_HD V3 &operator+=(V3 &a, V3 b)
{
    a.x += b.x;
    a.y += b.y;
    a.z += b.z;
    return a;
}
PT_HD float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
PT_HD V3 cross(V3 a, V3 b) { return v3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
PT_HD float length(V3 a) { return sqrtf(a.x * a.x + a.y * a.y + a.z * a.z); }
PT_HD float squared_length(V3 a) { return a.x * a.x + a.y * a.y + a.z * a.z; }
// a / t as one reciprocal and three products.  The reference's vec3 divides component-wise (vec3.h:124-127); the
// hit routine keeps that (operator/) because its records are compared bit for bit, the shading code -- compared
// statistically -- uses this form: an IEEE division costs ~10 instructions on the GPU.
PT_HD V3 div_fast(V3 a, float t)
{
    float k = 1.0f / t;
    return v3(a.x * k, a.y * k, a.z * k);
}
PT_HD V3 unit_vector(V3 a) { return div_fast(a, length(a)); } // vec3.h:196-199
// Eigen 3.2.10 Vector3f::normalized(): v / sqrt(x*x + (y*y + z*z)) -- its 3-element sum splits 1 + 2
// (Core/Redux.h redux_novec_unroller).  Used where the reference normalises through Eigen (transform3.h:60-63).
PT_HD V3 eigen_normalized(V3 a)
{
    float yz = a.y * a.y + a.z * a.z;
    return a / sqrtf(a.x * a.x + yz);
}
PT_HD bool is_nan(float x) { return !(x == x); }
PT_HD bool is_nan(V3 a) { return is_nan(a.x) || is_nan(a.y) || is_nan(a.z); }
PT_HD float get(V3 a, int i) { return i == 0 ? a.x : (i == 1 ? a.y : a.z); }

// 3x4 affine stored row major: m[r*4+c]
PT_HD V3 xform_point(const float *m, V3 p)
{
    return v3(m[0] * p.x + m[1] * p.y + m[2] * p.z + m[3], m[4] * p.x + m[5] * p.y + m[6] * p.z + m[7],
              m[8] * p.x + m[9] * p.y + m[10] * p.z + m[11]);
}
PT_HD V3 xform_vector(const float *m, V3 v)
{
    return v3(m[0] * v.x + m[1] * v.y + m[2] * v.z, m[4] * v.x + m[5] * v.y + m[6] * v.z,
              m[8] * v.x + m[9] * v.y + m[10] * v.z);
}
// transpose(linear part) * v : with m = the INVERSE transform this is the un-normalised normal transform
// (transform3.h:60-63: (L^-1)^T n).
PT_HD V3 xform_vector_transposed(const float *m, V3 v)
{
    return v3(m[0] * v.x + m[4] * v.y + m[8] * v.z, m[1] * v.x + m[5] * v.y + m[9] * v.z,
              m[2] * v.x + m[6] * v.y + m[10] * v.z);
}

// onb::build_from_w + onb::local (helpers.h:112-136)
struct Onb {
    V3 u, v, w;
};
PT_HD Onb onb_from_w(V3 n)
{
    Onb o;
    o.w = unit_vector(n);
    V3 a = fabsf(o.w.x) > 0.9f ? v3(0.f, 1.f, 0.f) : v3(1.f, 0.f, 0.f);
    o.v = unit_vector(cross(o.w, a));
    o.u = cross(o.w, o.v);
    return o;
}
PT_HD V3 onb_local(const Onb &o, V3 a) { return a.x * o.u + a.y * o.v + a.z * o.w; }

} // namespace pt
