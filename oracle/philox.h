// ORACLE - TEST INFRASTRUCTURE ONLY. Never included, linked or called by the product path.
//
// CPU restatement of theia's Philox4x32-10 usage (sh/random.philox.glsl, sh/random.util.glsl).
#pragma once
#include "glslmath.h"
#include "../include/theia_b200.h"

namespace orc {

// Plain Philox4x32-10 block function (Salmon et al., SC'11), the arithmetic of
// sh/random.philox.glsl:31-49.
static inline void philox4x32_10(const uint32_t ctr[4], const uint32_t key_in[2], uint32_t out[4]) {
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u;
    uint32_t x = ctr[0], y = ctr[1], z = ctr[2], w = ctr[3];
    uint32_t k0 = key_in[0], k1 = key_in[1];
    for (int i = 0; i < 10; ++i) {
        uint64_t p0 = (uint64_t)M0 * x;
        uint64_t p1 = (uint64_t)M1 * z;
        uint32_t hi0 = (uint32_t)(p0 >> 32), lo0 = (uint32_t)p0;
        uint32_t hi1 = (uint32_t)(p1 >> 32), lo1 = (uint32_t)p1;
        uint32_t nx = hi1 ^ y ^ k0, ny = lo1, nz = hi0 ^ w ^ k1, nw = lo0;
        x = nx; y = ny; z = nz; w = nw;
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    out[0] = x; out[1] = y; out[2] = z; out[3] = w;
}

// uaddCarry
static inline uint32_t addc(uint32_t a, uint32_t b, uint32_t* carry) {
    uint64_t s = (uint64_t)a + b;
    *carry = (uint32_t)(s >> 32);
    return (uint32_t)s;
}

// sh/random.util.glsl:8-14
static inline float normalizeUint(uint32_t i) {
    const float ONE_MINUS_EPSILON = u2f(0x3F7FFFFFu);
    const float EPSILON = u2f(0x2F800000u);
    return min_(ONE_MINUS_EPSILON, (float)i * EPSILON);
}

// sh/random.philox.glsl:15-52: counter = baseCount + (i << 2) with carry chain and end-around
// carry into word 0, key = key + stream with end-around carry.
static inline void philox_theia(const theia_philox_params_t& p, uint32_t stream, uint32_t i, uint32_t out[4]) {
    uint32_t carry = i << 2;
    uint32_t c[4] = {p.offset[0], p.offset[1], p.offset[2], p.offset[3]};
    c[0] = addc(c[0], carry, &carry);
    c[1] = addc(c[1], carry, &carry);
    c[2] = addc(c[2], carry, &carry);
    c[3] = addc(c[3], carry, &carry);
    c[0] += carry;
    uint32_t k[2] = {p.key[0], p.key[1]};
    k[0] = addc(k[0], stream, &carry);
    k[1] = addc(k[1], carry, &carry);
    k[0] += carry;
    philox4x32_10(c, k, out);
}

// sh/random.philox.glsl:54-77. philoxBufferCount is never written, so `count != 0` forces a
// fresh evaluation for every draw with i != 0 mod 2^30; for i << 2 == 0 the draw index is
// 0 mod 4 and `idx == 0` forces it as well: every draw is one full Philox evaluation.
static inline float random_s(const theia_philox_params_t& p, uint32_t stream, uint32_t i) {
    uint32_t r[4];
    philox_theia(p, stream, i, r);
    return normalizeUint(r[i & 3u]);
}

// RNG handle carried by every sampling function: (params, idx = stream, dim = draw counter)
struct Rng {
    const theia_philox_params_t* p;
    uint32_t idx;
    uint32_t dim;
    float next() { float r = random_s(*p, idx, dim); dim += 1; return r; }
    vec2 next2() {
        vec2 r = {random_s(*p, idx, dim), random_s(*p, idx, dim + 1)};
        dim += 2;
        return r;
    }
};

}  // namespace orc
