// Philox4x32-10 with theia's counter mapping, counter for counter
// (sh/random.philox.glsl:15-77, sh/random.util.glsl:8-14 of the reference).
//
// Draw i of stream s is element (i & 3) of Philox4x32-10(ctr = offset + 4 i, key = key + s):
// the shader multiplies the draw index by four and its result cache never hits, so every
// scalar draw is one full block evaluation. The additions are 128 / 64 bit adds with an
// end-around carry into word 0.
#pragma once
#include "detmath.cuh"
#include "../../include/theia_b200.h"

namespace theia {

// One Philox4x32-10 block; returns only the requested word (the compiler drops the dead
// half of the last round).
TDEV void philox4x32_10(uint32_t x, uint32_t y, uint32_t z, uint32_t w, uint32_t k0, uint32_t k1,
                        uint32_t out[4]) {
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u;
#pragma unroll
    for (int i = 0; i < 10; ++i) {
        uint32_t hi0 = __umulhi(M0, x), lo0 = M0 * x;
        uint32_t hi1 = __umulhi(M1, z), lo1 = M1 * z;
        uint32_t nx = hi1 ^ y ^ k0;
        uint32_t nz = hi0 ^ w ^ k1;
        x = nx; y = lo1; z = nz; w = lo0;
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    out[0] = x; out[1] = y; out[2] = z; out[3] = w;
}

// sh/random.util.glsl:8-14
TDEV float normalizeUint(uint32_t i) {
    const float ONE_MINUS_EPSILON = u2f(0x3F7FFFFFu);
    const float EPSILON = u2f(0x2F800000u);
    return min_(ONE_MINUS_EPSILON, (float)i * EPSILON);
}

TDEV uint32_t addc(uint32_t a, uint32_t b, uint32_t& carry) {
    uint32_t s = a + b;
    carry = s < a ? 1u : 0u;
    return s;
}

struct PhiloxKey { uint32_t k0, k1; };

// key = key + stream with end-around carry (sh/random.philox.glsl:25-28); constant per path
TDEV PhiloxKey philoxStreamKey(const theia_philox_params_t& p, uint32_t stream) {
    uint32_t carry;
    uint32_t k0 = addc(p.key[0], stream, carry);
    uint32_t k1 = addc(p.key[1], carry, carry);
    k0 += carry;
    return {k0, k1};
}

// sh/random.philox.glsl:54-77: one draw
TDEV float random_s(const theia_philox_params_t& p, PhiloxKey key, uint32_t i) {
    uint32_t carry = i << 2;
    uint32_t c0 = addc(p.offset[0], carry, carry);
    uint32_t c1 = addc(p.offset[1], carry, carry);
    uint32_t c2 = addc(p.offset[2], carry, carry);
    uint32_t c3 = addc(p.offset[3], carry, carry);
    c0 += carry;
    uint32_t r[4];
    philox4x32_10(c0, c1, c2, c3, key.k0, key.k1, r);
    uint32_t sel = i & 3u;
    uint32_t v = sel == 0 ? r[0] : (sel == 1 ? r[1] : (sel == 2 ? r[2] : r[3]));
    return normalizeUint(v);
}

// Experiment switch: one out-of-line copy of the Philox block per kernel instead of one per draw site.
#ifdef THEIA_PHILOX_OUTLINE
static __device__ __noinline__ float randomCall(const theia_philox_params_t* p, uint32_t k0, uint32_t k1, uint32_t i) {
    return random_s(*p, PhiloxKey{k0, k1}, i);
}
#define THEIA_DRAW(p, key, i) randomCall(p, (key).k0, (key).k1, i)
#else
#define THEIA_DRAW(p, key, i) random_s(*(p), key, i)
#endif

// RNG handle carried by every sampling function: stream key + running draw counter "dim"
struct Rng {
    const theia_philox_params_t* p;
    PhiloxKey key;
    uint32_t idx;
    uint32_t dim;
    TDEV float next() { float r = THEIA_DRAW(p, key, dim); dim += 1; return r; }
    TDEV vec2 next2() {
        vec2 r = {THEIA_DRAW(p, key, dim), THEIA_DRAW(p, key, dim + 1)};
        dim += 2;
        return r;
    }
};

TDEV Rng makeRng(const theia_philox_params_t* p, uint32_t idx) {
    return Rng{p, philoxStreamKey(*p, idx), idx, 0u};
}

}  // namespace theia
