// Device math of the forward tracers: float32 vectors with GLSL semantics and the
// deterministic transcendental functions of the arithmetic contract.
//
// Arithmetic contract: IEEE binary32, round to nearest even, no contraction (the library is
// compiled with -fmad=false), fused multiply-add only where fmaf() is written. + - * / sqrt
// and fma are correctly rounded on the GPU (-prec-div=true -prec-sqrt=true, the nvcc
// defaults), so a fixed operation sequence gives fixed bits. sin/cos/exp/log/pow are fixed
// polynomial sequences instead of the CUDA intrinsics; GLSL only specifies its built-ins to
// a few ulp, so any 1-2 ulp implementation is faithful to the reference shaders
// (sh/ = src/theia/shader/ of the reference).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

#define TDEV __device__ __forceinline__

namespace theia {

// ---------------------------------------------------------------- bit casts / classification
TDEV uint32_t f2u(float f) { return __float_as_uint(f); }
TDEV float u2f(uint32_t u) { return __uint_as_float(u); }
TDEV int32_t f2i(float f) { return __float_as_int(f); }
TDEV float i2f(int32_t i) { return __int_as_float(i); }

#define T_INF (__uint_as_float(0x7F800000u))
#define T_NAN (__uint_as_float(0x7FC00000u))

TDEV bool isinf_(float x) { return (f2u(x) & 0x7FFFFFFFu) == 0x7F800000u; }
TDEV bool isnan_(float x) { return x != x; }

// GLSL min/max/clamp with a fixed NaN behaviour: the first operand is kept unless the
// comparison holds.
TDEV float min_(float a, float b) { return b < a ? b : a; }
TDEV float max_(float a, float b) { return a < b ? b : a; }
TDEV float clamp_(float x, float lo, float hi) { return min_(max_(x, lo), hi); }
TDEV float abs_(float x) { return u2f(f2u(x) & 0x7FFFFFFFu); }
TDEV float sign_(float x) { return x > 0.0f ? 1.0f : (x < 0.0f ? -1.0f : 0.0f); }
TDEV float mix_(float a, float b, float t) { return a * (1.0f - t) + b * t; }
// sh/math.glsl:18-20
TDEV float signBit(float f) { return u2f((f2u(f) & 0x80000000u) | 0x3F800000u); }

// ---------------------------------------------------------------- deterministic math
// sin/cos: Cody-Waite reduction by pi/2 in three fma steps, degree 7/8 minimax kernels.
// Experiment switch THEIA_MATH_OUTLINE: one out-of-line copy of exp / log / sincos per kernel.
#ifdef THEIA_MATH_OUTLINE
#define TMATH static __device__ __noinline__
#else
#define TMATH TDEV
#endif
TDEV void dm_sincos_inl(float x, float* s_out, float* c_out) {
    const float TWO_OVER_PI = 0.636619772367581343f;
    const float MAGIC = 12582912.0f;  // 1.5 * 2^23
    const float PIO2_HI = 1.57079637050628662109375f;
    const float PIO2_MID = -4.37113900018624283e-8f;
    const float PIO2_LO = -1.71512449329647e-15f;
    float t = fmaf(x, TWO_OVER_PI, MAGIC);
    int32_t q = f2i(t);
    float j = t - MAGIC;
    float r = fmaf(-j, PIO2_HI, x);
    r = fmaf(-j, PIO2_MID, r);
    r = fmaf(-j, PIO2_LO, r);
    float z = r * r;
    float sp = fmaf(-1.9515295891e-4f, z, 8.3321608736e-3f);
    sp = fmaf(sp, z, -1.6666654611e-1f);
    float sn = fmaf(sp * z, r, r);
    float cp = fmaf(2.443315711809948e-5f, z, -1.388731625493765e-3f);
    cp = fmaf(cp, z, 4.166664568298827e-2f);
    float cs = fmaf(cp * z, z, fmaf(-0.5f, z, 1.0f));
    float s = (q & 1) ? cs : sn;
    float c = (q & 1) ? sn : cs;
    if (q & 2) s = -s;
    if ((q + 1) & 2) c = -c;
    *s_out = s;
    *c_out = c;
}

// exp: n = round(x log2 e), two step reduction, degree 6 polynomial, two step scaling.
#ifdef THEIA_MATH_OUTLINE
static __device__ __noinline__ float2 dm_sincos2(float x) {
    float2 r;
    dm_sincos_inl(x, &r.x, &r.y);
    return r;
}
TDEV void dm_sincos(float x, float* s_out, float* c_out) {
    float2 r = dm_sincos2(x);
    *s_out = r.x;
    *c_out = r.y;
}
#else
TDEV void dm_sincos(float x, float* s_out, float* c_out) { dm_sincos_inl(x, s_out, c_out); }
#endif

TMATH float dm_exp(float x) {
    if (isnan_(x)) return x;
    if (x > 88.7228394f) return T_INF;
    if (x < -103.972084f) return 0.0f;
    const float LOG2E = 1.44269504088896341f;
    const float MAGIC = 12582912.0f;
    const float LN2_HI = 0.693359375f;
    const float LN2_LO = -2.12194440e-4f;
    float t = fmaf(x, LOG2E, MAGIC);
    float n = t - MAGIC;
    float r = fmaf(-n, LN2_HI, x);
    r = fmaf(-n, LN2_LO, r);
    float p = fmaf(1.9875691500e-4f, r, 1.3981999507e-3f);
    p = fmaf(p, r, 8.3334519073e-3f);
    p = fmaf(p, r, 4.1665795894e-2f);
    p = fmaf(p, r, 1.6666665459e-1f);
    p = fmaf(p, r, 5.0000001201e-1f);
    p = fmaf(p * r, r, r) + 1.0f;
    int32_t ni = (int32_t)n;
    int32_t n1 = ni >> 1;
    int32_t n2 = ni - n1;
    float s1 = u2f((uint32_t)(n1 + 127) << 23);
    float s2 = u2f((uint32_t)(n2 + 127) << 23);
    return (p * s1) * s2;
}

// tanh(x) = 1 - 2 / (exp(2x) + 1) on |x| (odd extension); absolute error ~1e-7, which is what
// its only user, the kernel CDF of sh/response.histogram.kernel.glsl:62-70, needs.
TDEV float dm_tanh(float x) {
    if (x != x) return x;
    float a = abs_(x);
    float r = a > 10.0f ? 1.0f : 1.0f - 2.0f / (dm_exp(2.0f * a) + 1.0f);
    return x < 0.0f ? -r : r;
}
// atan: Cephes atanf scheme (reduce by tan(3 pi / 8), tan(pi / 8); degree 4 odd polynomial in x^2)
TDEV float dm_atan(float x) {
    if (x != x) return x;
    float a = abs_(x), y = 0.0f;
    if (a > 2.414213562373095f) { y = 1.5707963267948966f; a = -(1.0f / a); }
    else if (a > 0.4142135623730950f) { y = 0.7853981633974483f; a = (a - 1.0f) / (a + 1.0f); }
    float z = a * a;
    float p = fmaf(8.05374449538e-2f, z, -1.38776856032e-1f);
    p = fmaf(p, z, 1.99777106478e-1f);
    p = fmaf(p, z, -3.33329491539e-1f);
    y += fmaf(p * z, a, a);
    return x < 0.0f ? -y : y;
}

// log: mantissa/exponent split at sqrt(1/2), degree 9 polynomial.
TMATH float dm_log(float x) {
    if (isnan_(x)) return x;
    if (x < 0.0f) return T_NAN;
    if (x == 0.0f) return -T_INF;
    if (isinf_(x)) return x;
    int32_t e = 0;
    if (x < 1.17549435e-38f) { x *= 8388608.0f; e = -23; }
    uint32_t ix = f2u(x);
    e += (int32_t)(ix >> 23) - 126;
    float m = u2f((ix & 0x007FFFFFu) | 0x3F000000u);
    if (m < 0.707106781186547524f) { e -= 1; m = m + m - 1.0f; }
    else { m = m - 1.0f; }
    float z = m * m;
    float p = fmaf(7.0376836292e-2f, m, -1.1514610310e-1f);
    p = fmaf(p, m, 1.1676998740e-1f);
    p = fmaf(p, m, -1.2420140846e-1f);
    p = fmaf(p, m, 1.4249322787e-1f);
    p = fmaf(p, m, -1.6668057665e-1f);
    p = fmaf(p, m, 2.0000714765e-1f);
    p = fmaf(p, m, -2.4999993993e-1f);
    p = fmaf(p, m, 3.3333331174e-1f);
    float fe = (float)e;
    float y = (m * z) * p;
    y = fmaf(-2.12194440e-4f, fe, y);
    y = fmaf(-0.5f, z, y);
    float r = m + y;
    r = fmaf(0.693359375f, fe, r);
    return r;
}

// pow(x, y), x >= 0: exp(y log x) with log x carried as an unevaluated sum of two floats (H + L, ~2^-40 relative),
// so that the result stays within ~2 ulp for |y log x| up to the range of exp. Float operations only (fma, +, -, *,
// /): the first form went through double precision, which cost the cascade / muon light sources a third of their
// generate kernel on a GPU with few FP64 units.
static __device__ __noinline__ float dm_pow(float x, float y) {
    if (isnan_(x) || isnan_(y)) return T_NAN;
    if (y == 0.0f) return 1.0f;
    if (x < 0.0f) return T_NAN;
    if (x == 0.0f) return y > 0.0f ? 0.0f : T_INF;
    if (isinf_(x)) return y > 0.0f ? T_INF : 0.0f;
    if (x == 1.0f) return 1.0f;
    if (isinf_(y)) {
        bool big = x > 1.0f;
        return ((y > 0.0f) == big) ? T_INF : 0.0f;
    }
    // ---- log x = e ln 2 + log m, m in (sqrt(1/2), sqrt(2)], log m = 2 atanh(s), s = (m - 1) / (m + 1)
    int32_t e = 0;
    if (x < 1.17549435e-38f) { x *= 8388608.0f; e = -23; }
    uint32_t ix = f2u(x);
    e += (int32_t)(ix >> 23) - 127;
    float m = u2f((ix & 0x007FFFFFu) | 0x3F800000u);
    if (m > 1.41421354f) { m *= 0.5f; e += 1; }
    float f = m - 1.0f;                 // exact
    float g = m + 1.0f;                 // rounded: m + 1 = g + gl exactly
    float gl = m - (g - 1.0f);
    float sh = f / g;
    float r = fmaf(-sh, g, f);          // f - sh g, exact
    r = fmaf(-sh, gl, r);
    float sl = r / g;                   // s = sh + sl
    float z = sh * sh;
    float p = fmaf(1.5384615385e-1f, z, 1.8181818182e-1f);  // 2/13, 2/11, 2/9, 2/7, 2/5, 2/3
    p = fmaf(p, z, 2.2222222222e-1f);
    p = fmaf(p, z, 2.8571428571e-1f);
    p = fmaf(p, z, 4.0e-1f);
    p = fmaf(p, z, 6.6666666667e-1f);
    float lh = sh + sh;
    float ll = fmaf(sh * z, p, (sl + sl) * fmaf(z, 1.0f + z, 1.0f));  // d/ds 2 atanh(s) = 2 / (1 - s^2)
    float fe = (float)e;
    float a = fe * 0.693359375f;        // exact (ln 2 split as in dm_log)
    float H = a + lh;                   // two-sum
    float bv = H - a;
    float L = (a - (H - bv)) + (lh - bv);
    L += fmaf(fe, -2.12194440e-4f, ll);
    float H2 = H + L;
    L = L - (H2 - H);
    H = H2;
    // ---- t = y (H + L) as th + tl
    float th = y * H;
    float tl = fmaf(y, H, -th) + y * L;
    if (th > 89.0f) return T_INF;
    if (th < -104.0f) return 0.0f;
    float q = dm_exp(th);
    return fmaf(q, tl, q);
}

// ---------------------------------------------------------------- vectors
struct vec2 { float x, y; };
struct vec3 { float x, y, z; };
struct vec4 { float x, y, z, w; };

TDEV vec3 v3(float x, float y, float z) { return vec3{x, y, z}; }
TDEV vec3 v3(float s) { return vec3{s, s, s}; }
TDEV vec3 v3(const float* p) { return vec3{p[0], p[1], p[2]}; }
TDEV vec4 v4(const float* p) { return vec4{p[0], p[1], p[2], p[3]}; }
TDEV vec3 operator+(vec3 a, vec3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
TDEV vec3 operator-(vec3 a, vec3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
TDEV vec3 operator-(vec3 a) { return {-a.x, -a.y, -a.z}; }
TDEV vec3 operator*(vec3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
TDEV vec3 operator*(float s, vec3 a) { return {a.x * s, a.y * s, a.z * s}; }
TDEV vec3 operator/(vec3 a, float s) { return {a.x / s, a.y / s, a.z / s}; }
// a + s*b with one fma per component
TDEV vec3 fma3(float s, vec3 b, vec3 a) { return {fmaf(s, b.x, a.x), fmaf(s, b.y, a.y), fmaf(s, b.z, a.z)}; }
// dot: fma chain z(y(x))
TDEV float dot(vec3 a, vec3 b) { return fmaf(a.z, b.z, fmaf(a.y, b.y, a.x * b.x)); }
TDEV vec3 cross(vec3 a, vec3 b) {
    return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}
TDEV float length(vec3 a) { return sqrtf(dot(a, a)); }
TDEV float distance(vec3 a, vec3 b) { return length(a - b); }
TDEV vec3 normalize(vec3 a) {
    float inv = 1.0f / sqrtf(dot(a, a));
    return a * inv;
}
TDEV vec4 operator/(vec4 a, float s) { return {a.x / s, a.y / s, a.z / s, a.w / s}; }

// column major 3x3 (GLSL mat3(c0, c1, c2))
struct mat3 { vec3 c0, c1, c2; };
TDEV mat3 mat3_diag(float s) { return {{s, 0, 0}, {0, s, 0}, {0, 0, s}}; }
TDEV vec3 operator*(const mat3& m, vec3 v) {
    return {fmaf(m.c2.x, v.z, fmaf(m.c1.x, v.y, m.c0.x * v.x)),
            fmaf(m.c2.y, v.z, fmaf(m.c1.y, v.y, m.c0.y * v.x)),
            fmaf(m.c2.z, v.z, fmaf(m.c1.z, v.y, m.c0.z * v.x))};
}
// v * M = transpose(M) * v
TDEV vec3 mulT(vec3 v, const mat3& m) { return {dot(v, m.c0), dot(v, m.c1), dot(v, m.c2)}; }

#define T_PI 3.141592653589793238f
#define T_TWO_PI 6.283185307179586477f
#define T_PI_OVER_FOUR 0.7853981633974483096f
#define T_PI_OVER_TWO 1.570796326794896619f
#define T_INV_4PI 0.0795774715459476679f
#define T_SPEED_OF_LIGHT 0.299792458f
#define T_INV_SPEED_OF_LIGHT 3.335640951981521f

// sh/math.glsl:24-29 (precise): ab - cd with one fma error term
TDEV float prodDiff(float a, float b, float c, float d) {
    float cd = c * d;
    float diff = fmaf(a, b, -cd);
    float err = fmaf(-c, d, cd);
    return diff + err;
}
// sh/math.glsl:33-39
TDEV vec3 crosser(vec3 v, vec3 w) {
    return {prodDiff(v.y, w.z, v.z, w.y), prodDiff(v.z, w.x, v.x, w.z), prodDiff(v.x, w.y, v.y, w.x)};
}
// sh/math.glsl:45-55
TDEV mat3 createLocalCOSY(vec3 vz) {
    float s = signBit(vz.z);
    float a = -1.0f / (s + vz.z);
    float b = vz.x * vz.y * a;
    vec3 vx = {1.0f + s * vz.x * vz.x * a, s * b, -s * vz.x};
    vec3 vy = {b, s + vz.y * vz.y * a, -vz.y};
    vx = normalize(vx);
    vy = normalize(vy);
    return {vx, vy, vz};
}
// sh/math.glsl:58-63
TDEV vec3 perpendicularTo(vec3 v) {
    float s = signBit(v.z);
    float a = -1.0f / (s + v.z);
    float b = v.x * v.y * a;
    return {b, s + v.y * v.y * a, -v.y};
}
// sh/math.glsl:66-78
TDEV vec3 perpendicularTo(vec3 a, vec3 b) {
    vec3 c = crosser(a, b);
    float len = length(c);
    if (len < 1e-5f) return perpendicularTo(a);
    return c / len;
}

}  // namespace theia
