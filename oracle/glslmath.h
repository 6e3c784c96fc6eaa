// ORACLE - TEST INFRASTRUCTURE ONLY. Never included, linked or called by the product path.
//
// GLSL-flavoured float32 vector helpers and the deterministic transcendental functions used
// by the CPU restatement of theia's forward tracing shaders.
//
// Arithmetic contract (DESIGN.md "Arithmetic contract"):
//  * IEEE binary32, round to nearest even, no contraction (compile with -ffp-contract=off),
//    fused multiply-add only where fmaf() is written.
//  * + - * / sqrt fma are correctly rounded on both the CPU and the GPU, so the same
//    operation sequence gives the same bits. sin/cos/exp/log/pow are NOT taken from libm
//    (glibc and CUDA differ by an ulp): they are restated here as fixed polynomial
//    sequences; the CUDA side carries the same sequences (theia_b200/csrc/detmath.cuh).
//    GLSL built-ins are only specified to a few ulp (Vulkan spec, "Precision and Operation
//    of SPIR-V Instructions"), so any 1-2 ulp implementation is a faithful restatement.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>

namespace orc {

// ---------------------------------------------------------------- bit casts
static inline uint32_t f2u(float f) { uint32_t u; std::memcpy(&u, &f, 4); return u; }
static inline float u2f(uint32_t u) { float f; std::memcpy(&f, &u, 4); return f; }
static inline int32_t f2i(float f) { int32_t u; std::memcpy(&u, &f, 4); return u; }
static inline float i2f(int32_t u) { float f; std::memcpy(&f, &u, 4); return f; }
static inline uint64_t d2u(double f) { uint64_t u; std::memcpy(&u, &f, 8); return u; }
static inline double u2d(uint64_t u) { double f; std::memcpy(&f, &u, 8); return f; }

static const float F_INF = u2f(0x7F800000u);
static const float F_NAN = u2f(0x7FC00000u);

static inline bool isinf_(float x) { return (f2u(x) & 0x7FFFFFFFu) == 0x7F800000u; }
static inline bool isnan_(float x) { return x != x; }

// GLSL min/max/clamp/abs/sign/mix/fract with a fixed NaN behaviour (second operand wins
// the comparison only if the comparison is true).
static inline float min_(float a, float b) { return b < a ? b : a; }
static inline float max_(float a, float b) { return a < b ? b : a; }
static inline float clamp_(float x, float lo, float hi) { return min_(max_(x, lo), hi); }
static inline float abs_(float x) { return u2f(f2u(x) & 0x7FFFFFFFu); }
static inline float sign_(float x) { return x > 0.0f ? 1.0f : (x < 0.0f ? -1.0f : 0.0f); }
static inline float mix_(float a, float b, float t) { return a * (1.0f - t) + b * t; }
static inline float fract_(float x) { return x - std::floor(x); }
// sh/math.glsl:18-20
static inline float signBit(float f) { return u2f((f2u(f) & 0x80000000u) | 0x3F800000u); }

// ---------------------------------------------------------------- deterministic math
// sin/cos: Cody-Waite reduction by pi/2 with three fma steps, Cephes sinf/cosf kernels.
static inline void dm_sincos(float x, float* s_out, float* c_out) {
    const float TWO_OVER_PI = 0.636619772367581343f;
    const float MAGIC = 12582912.0f;  // 1.5 * 2^23
    const float PIO2_HI = 1.57079637050628662109375f;       // float(pi/2)
    const float PIO2_MID = -4.37113900018624283e-8f;        // float(pi/2 - hi)
    const float PIO2_LO = -1.71512449329647e-15f;           // float(pi/2 - hi - mid)
    float t = fmaf(x, TWO_OVER_PI, MAGIC);
    int32_t q = f2i(t);  // low bits hold the integer
    float j = t - MAGIC;
    float r = fmaf(-j, PIO2_HI, x);
    r = fmaf(-j, PIO2_MID, r);
    r = fmaf(-j, PIO2_LO, r);
    float z = r * r;
    // sin(r), |r| <= pi/4
    float sp = fmaf(-1.9515295891e-4f, z, 8.3321608736e-3f);
    sp = fmaf(sp, z, -1.6666654611e-1f);
    float sn = fmaf(sp * z, r, r);
    // cos(r)
    float cp = fmaf(2.443315711809948e-5f, z, -1.388731625493765e-3f);
    cp = fmaf(cp, z, 4.166664568298827e-2f);
    float cs = fmaf(cp * z, z, fmaf(-0.5f, z, 1.0f));
    // quadrant
    float s = (q & 1) ? cs : sn;
    float c = (q & 1) ? sn : cs;
    if (q & 2) s = -s;
    if ((q + 1) & 2) c = -c;
    *s_out = s;
    *c_out = c;
}
static inline float dm_sin(float x) { float s, c; dm_sincos(x, &s, &c); return s; }
static inline float dm_cos(float x) { float s, c; dm_sincos(x, &s, &c); return c; }

// exp: n = round(x*log2e), r = x - n*ln2 (two steps), degree 6 polynomial, two step scaling.
static inline float dm_exp(float x) {
    if (isnan_(x)) return x;
    if (x > 88.7228394f) return F_INF;
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
    int32_t ni = (int32_t)n;  // exact, |n| <= 150
    int32_t n1 = ni >> 1;     // floor(n/2)
    int32_t n2 = ni - n1;
    float s1 = u2f((uint32_t)(n1 + 127) << 23);
    float s2 = u2f((uint32_t)(n2 + 127) << 23);
    return (p * s1) * s2;
}

// tanh(x) = 1 - 2 / (exp(2x) + 1) on |x| (odd extension); absolute error ~1e-7, which is what
// its only user, the kernel CDF of sh/response.histogram.kernel.glsl:62-70, needs.
static inline float dm_tanh(float x) {
    if (x != x) return x;
    float a = abs_(x);
    float r = a > 10.0f ? 1.0f : 1.0f - 2.0f / (dm_exp(2.0f * a) + 1.0f);
    return x < 0.0f ? -r : r;
}
// atan: Cephes atanf scheme (reduce by tan(3 pi / 8), tan(pi / 8); degree 4 odd polynomial in x^2)
static inline float dm_atan(float x) {
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

// log: Cephes logf. Denormals are scaled by 2^23 first.
static inline float dm_log(float x) {
    if (isnan_(x)) return x;
    if (x < 0.0f) return F_NAN;
    if (x == 0.0f) return -F_INF;
    if (isinf_(x)) return x;
    int32_t e = 0;
    if (x < 1.17549435e-38f) { x *= 8388608.0f; e = -23; }
    uint32_t ix = f2u(x);
    e += (int32_t)(ix >> 23) - 126;                     // x = m * 2^e, m in [0.5, 1)
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
static inline float dm_pow(float x, float y) {
    if (isnan_(x) || isnan_(y)) return F_NAN;
    if (y == 0.0f) return 1.0f;
    if (x < 0.0f) return F_NAN;
    if (x == 0.0f) return y > 0.0f ? 0.0f : F_INF;
    if (isinf_(x)) return y > 0.0f ? F_INF : 0.0f;
    if (x == 1.0f) return 1.0f;
    if (isinf_(y)) {
        bool big = x > 1.0f;
        return ((y > 0.0f) == big) ? F_INF : 0.0f;
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
    if (th > 89.0f) return F_INF;
    if (th < -104.0f) return 0.0f;
    float q = dm_exp(th);
    return fmaf(q, tl, q);
}

// ---------------------------------------------------------------- vectors
struct vec2 { float x, y; };
struct vec3 { float x, y, z; };
struct vec4 { float x, y, z, w; };

static inline vec3 v3(float x, float y, float z) { return vec3{x, y, z}; }
static inline vec3 v3(float s) { return vec3{s, s, s}; }
static inline vec3 v3(const float* p) { return vec3{p[0], p[1], p[2]}; }
static inline vec4 v4(const float* p) { return vec4{p[0], p[1], p[2], p[3]}; }
static inline vec3 operator+(vec3 a, vec3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
static inline vec3 operator-(vec3 a, vec3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
static inline vec3 operator-(vec3 a) { return {-a.x, -a.y, -a.z}; }
static inline vec3 operator*(vec3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
static inline vec3 operator*(float s, vec3 a) { return {a.x * s, a.y * s, a.z * s}; }
static inline vec3 operator*(vec3 a, vec3 b) { return {a.x * b.x, a.y * b.y, a.z * b.z}; }
static inline vec3 operator/(vec3 a, float s) { return {a.x / s, a.y / s, a.z / s}; }
// a + s*b with one fma per component
static inline vec3 fma3(float s, vec3 b, vec3 a) {
    return {fmaf(s, b.x, a.x), fmaf(s, b.y, a.y), fmaf(s, b.z, a.z)};
}
// dot: fma chain z(y(x))
static inline float dot(vec3 a, vec3 b) { return fmaf(a.z, b.z, fmaf(a.y, b.y, a.x * b.x)); }
// GLSL cross without contraction (used on "precise" values)
static inline vec3 cross(vec3 a, vec3 b) {
    return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}
static inline float length(vec3 a) { return std::sqrt(dot(a, a)); }
static inline float distance(vec3 a, vec3 b) { return length(a - b); }
static inline vec3 normalize(vec3 a) {
    float inv = 1.0f / std::sqrt(dot(a, a));
    return a * inv;
}
static inline vec4 operator*(vec4 a, float s) { return {a.x * s, a.y * s, a.z * s, a.w * s}; }
static inline vec4 operator/(vec4 a, float s) { return {a.x / s, a.y / s, a.z / s, a.w / s}; }

// column major 3x3 (GLSL mat3(c0,c1,c2))
struct mat3 { vec3 c0, c1, c2; };
static inline mat3 mat3_diag(float s) { return {{s, 0, 0}, {0, s, 0}, {0, 0, s}}; }
// M * v
static inline vec3 operator*(const mat3& m, vec3 v) {
    return {fmaf(m.c2.x, v.z, fmaf(m.c1.x, v.y, m.c0.x * v.x)),
            fmaf(m.c2.y, v.z, fmaf(m.c1.y, v.y, m.c0.y * v.x)),
            fmaf(m.c2.z, v.z, fmaf(m.c1.z, v.y, m.c0.z * v.x))};
}
// v * M = transpose(M) * v
static inline vec3 mulT(vec3 v, const mat3& m) { return {dot(v, m.c0), dot(v, m.c1), dot(v, m.c2)}; }

// Mueller matrices on this path only ever have the block structures below, but the oracle
// keeps the full 4x4 (column major) product to stay literal.
struct mat4 { float m[4][4]; };  // m[col][row]
static inline mat4 mat4_identity() {
    mat4 r{};
    for (int i = 0; i < 4; ++i) r.m[i][i] = 1.0f;
    return r;
}
static inline vec4 operator*(const mat4& a, vec4 v) {
    float in[4] = {v.x, v.y, v.z, v.w};
    float out[4];
    for (int r = 0; r < 4; ++r) {
        float acc = a.m[0][r] * in[0];
        acc = fmaf(a.m[1][r], in[1], acc);
        acc = fmaf(a.m[2][r], in[2], acc);
        acc = fmaf(a.m[3][r], in[3], acc);
        out[r] = acc;
    }
    return {out[0], out[1], out[2], out[3]};
}

// ---------------------------------------------------------------- sh/math.glsl
#define ORC_PI 3.141592653589793238f
#define ORC_TWO_PI 6.283185307179586477f
#define ORC_INV_4PI 0.0795774715459476679f
#define ORC_INV_2PI 0.1591549430918953357f
#define ORC_PI_OVER_FOUR 0.7853981633974483096f
#define ORC_PI_OVER_TWO 1.570796326794896619f

// sh/math.glsl:24-29 (precise): ab - cd with an fma error term
static inline float prodDiff(float a, float b, float c, float d) {
    float cd = c * d;
    float diff = fmaf(a, b, -cd);
    float err = fmaf(-c, d, cd);
    return diff + err;
}
// sh/math.glsl:33-39
static inline vec3 crosser(vec3 v, vec3 w) {
    return {prodDiff(v.y, w.z, v.z, w.y), prodDiff(v.z, w.x, v.x, w.z), prodDiff(v.x, w.y, v.y, w.x)};
}
// sh/math.glsl:45-55
static inline mat3 createLocalCOSY(vec3 vz) {
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
static inline vec3 perpendicularTo(vec3 v) {
    float s = signBit(v.z);
    float a = -1.0f / (s + v.z);
    float b = v.x * v.y * a;
    return {b, s + v.y * v.y * a, -v.y};
}
// sh/math.glsl:66-78
static inline vec3 perpendicularTo(vec3 a, vec3 b) {
    vec3 c = crosser(a, b);
    float len = length(c);
    if (len < 1e-5f) return perpendicularTo(a);
    return c / len;
}

}  // namespace orc
