// ORACLE - TEST INFRASTRUCTURE ONLY. Never included, linked or called by the product path.
//
// CPU restatement of the device math of theia's forward tracers. Every function cites the
// GLSL it follows (sh/ = /root/reference/src/theia/shader/). "Device addresses" are host
// pointers here.
#pragma once
#include "glslmath.h"
#include "philox.h"

namespace orc {

#define ORC_SPEED_OF_LIGHT 0.299792458f
#define ORC_INV_SPEED_OF_LIGHT 3.335640951981521f

// ------------------------------------------------------------------ sh/lookup.glsl:14-29
static inline float lookUp(uint64_t table, float u, float nullValue) {
    if (table == 0) return nullValue;
    const int32_t* t = (const int32_t*)(uintptr_t)table;
    int nx = t[0];
    const float* samples = (const float*)(t + 1);
    u = clamp_(u, 0.0f, 1.0f);
    u *= (float)(nx - 1);
    float fl = std::floor(u);
    int lo = (int)fl;
    int hi = (int)std::ceil(u);
    float l = u - fl;
    return samples[lo] * (1.0f - l) + samples[hi] * l;
}

// ------------------------------------------------------------------ sh/material.glsl
#pragma pack(push, 4)
struct MediumGLSL {  // sh/material.glsl:11-34, material.py:122-138 (88 bytes, 8 aligned)
    float lambda_min, lambda_max;
    uint64_t n, vg, mu_a, mu_s, log_phase, phase_sampling, phase_m12, phase_m22, phase_m33, phase_m34;
};
struct MaterialGLSL {  // sh/material.glsl:92-97, material.py:587-595
    uint64_t inside, outside;
    uint32_t flagsInwards, flagsOutwards;
};
#pragma pack(pop)
static_assert(sizeof(MediumGLSL) == 88, "Medium layout");
static_assert(sizeof(MaterialGLSL) == 24, "Material layout");

enum : uint32_t {
    MATERIAL_BLACK_BODY_BIT = 0x01, MATERIAL_DETECTOR_BIT = 0x02, MATERIAL_LIGHT_SOURCE_BIT = 0x04,
    MATERIAL_NO_REFLECT_FWD_BIT = 0x08, MATERIAL_NO_REFLECT_BWD_BIT = 0x10,
    MATERIAL_NO_TRANSMIT_FWD_BIT = 0x20, MATERIAL_NO_TRANSMIT_BWD_BIT = 0x40,
    MATERIAL_VOLUME_BORDER_BIT = 0x80
};

struct MediumConstants { float n, vg, mu_s, mu_e; };

static inline const MediumGLSL* asMedium(uint64_t p) { return (const MediumGLSL*)(uintptr_t)p; }

// sh/material.glsl:37-39
static inline float normalize_lambda(const MediumGLSL* m, float lambda) {
    return clamp_((lambda - m->lambda_min) / (m->lambda_max - m->lambda_min), 0.0f, 1.0f);
}
// sh/material.glsl:50-74
static inline MediumConstants lookUpMedium(uint64_t medium, float lambda) {
    if (medium == 0) return {1.0f, ORC_SPEED_OF_LIGHT, 0.0f, 0.0f};
    const MediumGLSL* m = asMedium(medium);
    float u = normalize_lambda(m, lambda);
    float mu_a = lookUp(m->mu_a, u, 0.0f);
    float mu_s = lookUp(m->mu_s, u, 0.0f);
    float mu_e = mu_a + mu_s;
    return {lookUp(m->n, u, 1.0f), lookUp(m->vg, u, ORC_SPEED_OF_LIGHT), mu_s, mu_e};
}

// ------------------------------------------------------------------ sh/ray.glsl
// RayState + polarization state (PolarizedForwardRay, sh/ray.glsl:22-36,163-170). The
// unpolarized ray simply ignores stokes/polRef.
struct Ray {
    vec3 position, direction;
    uint64_t medium;
    float wavelength, time, lin_contrib, log_contrib;
    MediumConstants constants;
    vec4 stokes;
    vec3 polRef;
};

struct SourceRay {  // sh/lightsource.common.glsl:12-26
    vec3 position, direction;
    vec4 stokes;
    vec3 polRef;
    float startTime, contrib;
};
struct WavelengthSample { float wavelength, contrib; };  // sh/wavelengthsource.common.glsl:4-7

// sh/ray.glsl:136-143
// sh/ray.glsl:131-133
static inline float getContrib(const Ray& r) { return r.lin_contrib * dm_exp(r.log_contrib); }
static inline bool isRayBad(const Ray& r) {
    vec3 p = r.position, d = r.direction;
    return isnan_(p.x) || isnan_(p.y) || isnan_(p.z) || isinf_(p.x) || isinf_(p.y) || isinf_(p.z) ||
           isnan_(d.x) || isnan_(d.y) || isnan_(d.z) || isinf_(d.x) || isinf_(d.y) || isinf_(d.z) ||
           !(length(d) > 0.0f);
}

// sh/lightsource.common.glsl:59-107 (polRef = first column of createLocalCOSY)
static inline SourceRay createSourceRay(vec3 pos, vec3 dir, float startTime, float contrib, bool polarized) {
    SourceRay s{};
    s.position = pos; s.direction = dir; s.startTime = startTime; s.contrib = contrib;
    s.stokes = {1.0f, 0.0f, 0.0f, 0.0f};
    s.polRef = polarized ? createLocalCOSY(dir).c0 : v3(0.0f);
    return s;
}
static inline SourceRay createSourceRay(vec3 pos, vec3 dir, vec4 stokes, vec3 polRef, float startTime, float contrib) {
    return SourceRay{pos, dir, stokes, polRef, startTime, contrib};
}
// sh/ray.glsl:59-79, 188-211
static inline Ray createRay(const SourceRay& s, uint64_t medium, MediumConstants c, WavelengthSample ph) {
    Ray r{};
    r.position = s.position; r.direction = s.direction; r.medium = medium;
    r.wavelength = ph.wavelength; r.time = s.startTime;
    r.lin_contrib = ph.contrib * s.contrib; r.log_contrib = 0.0f; r.constants = c;
    r.stokes = s.stokes; r.polRef = s.polRef;
    return r;
}

// ------------------------------------------------------------------ sh/ray.propagate.glsl
typedef theia_propagate_params_t PropagateParams;

// :32-49
static inline float sampleScatterLength(const Ray& ray, const PropagateParams& p, float u) {
    float dist = p.maxDist;
    float sampleCoef = p.scatterCoefficient >= 0.0f ? p.scatterCoefficient : ray.constants.mu_s;
    bool canScatter = ray.constants.mu_s > 0.0f;
    if (canScatter && sampleCoef != 0.0f) dist = -dm_log(1.0f - u) / sampleCoef;
    return dist;
}
// :70-80
static inline int updateRay(Ray& ray, float dist, const PropagateParams& p) {
    ray.log_contrib = fmaf(-ray.constants.mu_e, dist, ray.log_contrib);
    ray.time += dist / ray.constants.vg;
    return ray.time <= p.maxTime ? THEIA_RESULT_SUCCESS : THEIA_RESULT_RAY_DECAYED;
}
// :101-130
static inline void updateRayIS(Ray& ray, float dist, const PropagateParams& p, bool hit) {
    bool canScatter = ray.constants.mu_s > 0.0f;
    if (!canScatter) return;
    float sampleCoef = p.scatterCoefficient >= 0.0f ? p.scatterCoefficient : ray.constants.mu_s;
    ray.log_contrib = fmaf(sampleCoef, dist, ray.log_contrib);
    if (!hit) ray.lin_contrib /= sampleCoef;
}
// :153-166
static inline int propagateRay(Ray& ray, float dist, const PropagateParams& p) {
    ray.position = fma3(dist, ray.direction, ray.position);
    vec3 q = ray.position;
    bool outside = q.x < p.lowerBBoxCorner[0] || q.y < p.lowerBBoxCorner[1] || q.z < p.lowerBBoxCorner[2] ||
                   q.x > p.upperBBoxCorner[0] || q.y > p.upperBBoxCorner[1] || q.z > p.upperBBoxCorner[2];
    int result = updateRay(ray, dist, p);
    return outside ? THEIA_RESULT_RAY_LOST : result;
}

// ------------------------------------------------------------------ sh/polarization.glsl
// All reference frame rotations are the block matrix [[1,0,0,0],[0,c,-s,0],[0,s,c,0],[0,0,0,1]]
// (column major constructor of :38-43): applied to a Stokes vector it gives
// (I, cQ - sU, sQ + cU, V).
static inline mat4 rotMatrix(float c, float s) {
    mat4 r = mat4_identity();
    r.m[1][1] = c; r.m[1][2] = s;
    r.m[2][1] = -s; r.m[2][2] = c;
    return r;
}
// :22-35
static inline mat4 alignPolRef(vec3 rayDir, vec3 oldRef, vec3 newRef) {
    float cos_phi = dot(oldRef, newRef);
    float sin_phi = dot(crosser(oldRef, newRef), rayDir);
    float c = 2.0f * cos_phi * cos_phi - 1.0f;
    float s = 2.0f * cos_phi * sin_phi;
    return rotMatrix(c, s);
}
// :39-69
static inline mat4 rotatePolRef(vec3 dir, vec3 ref, vec3 nw, vec3* new_ref) {
    vec3 nr = crosser(dir, nw);
    float len = length(nr);
    if (len > 1.0e-7f) {
        nr = nr / len;
    } else {
        *new_ref = ref;
        return mat4_identity();
    }
    *new_ref = nr;
    float cos_phi = dot(ref, nr);
    float sin_phi = dot(crosser(ref, nr), dir);
    float c = 2.0f * cos_phi * cos_phi - 1.0f;
    float s = 2.0f * cos_phi * sin_phi;
    return rotMatrix(c, s);
}
// :89-107
static inline mat4 lookUpPhaseMatrix(uint64_t medium, float cos_theta) {
    if (medium == 0) return mat4_identity();
    const MediumGLSL* m = asMedium(medium);
    float t = 0.5f * (cos_theta + 1.0f);
    float m12 = lookUp(m->phase_m12, t, 0.0f);
    float m22 = lookUp(m->phase_m22, t, 0.0f);
    float m33 = lookUp(m->phase_m33, t, 0.0f);
    float m34 = lookUp(m->phase_m34, t, 0.0f);
    mat4 r{};
    r.m[0][0] = 1.0f; r.m[0][1] = m12;
    r.m[1][0] = m12;  r.m[1][1] = m22;
    r.m[2][2] = m33;  r.m[2][3] = -m34;
    r.m[3][2] = m34;  r.m[3][3] = m33;
    return r;
}
// :109-120
static inline mat4 polarizerMatrix(float p, float s) {
    float att = p * p + s * s;
    float m12 = (p * p - s * s) / att;
    float m33 = (2.0f * p * s) / att;
    mat4 r{};
    r.m[0][0] = 1.0f; r.m[0][1] = m12;
    r.m[1][0] = m12;  r.m[1][1] = 1.0f;
    r.m[2][2] = m33;
    r.m[3][3] = m33;
    return r;
}

// sh/ray.propagate.glsl:193-202
static inline void alignRayToHit(Ray& ray, vec3 hitNormal, bool polarized) {
    if (!polarized) return;
    vec3 polRef;
    mat4 rotate = rotatePolRef(ray.direction, ray.polRef, hitNormal, &polRef);
    ray.stokes = rotate * ray.stokes;
    ray.polRef = polRef;
}
// sh/ray.propagate.glsl:237-264
static inline int propagateRayToHit(Ray& ray, vec3 hitPos, vec3 hitNormal, const PropagateParams& p, bool polarized) {
    alignRayToHit(ray, hitNormal, polarized);
    float dist = distance(ray.position, hitPos);
    ray.position = hitPos;
    return updateRay(ray, dist, p);
}

// ------------------------------------------------------------------ sh/scatter.volume.glsl
// :7-26
static inline vec3 scatterDir(vec3 prevDir, float cos_theta, float phi) {
    prevDir = normalize(prevDir);
    float sin_theta = std::sqrt(max_(1.0f - cos_theta * cos_theta, 0.0f));
    float sp, cp;
    dm_sincos(phi, &sp, &cp);
    vec3 local = {sin_theta * cp, sin_theta * sp, cos_theta};
    local = normalize(local);
    mat3 trafo = createLocalCOSY(prevDir);
    return normalize(trafo * local);
}
// :28-46
static inline float sampleScatterDir(uint64_t medium, vec3 inDir, vec2 rng, float* cos_theta, float* phi) {
    const MediumGLSL* m = asMedium(medium);
    *phi = rng.x * ORC_TWO_PI;
    if (m != nullptr && m->phase_sampling != 0) {
        float ct = lookUp(m->phase_sampling, rng.y, 0.0f);
        ct = clamp_(ct, -1.0f, 1.0f);
        *cos_theta = ct;
        return dm_exp(lookUp(m->log_phase, 0.5f * (ct + 1.0f), 0.0f));
    }
    *cos_theta = 2.0f * rng.y - 1.0f;
    return ORC_INV_4PI;
}
// :48-53
static inline vec3 scatter(uint64_t medium, vec3 inDir, vec2 rng, float* p) {
    float cos_theta, phi;
    *p = sampleScatterDir(medium, inDir, rng, &cos_theta, &phi);
    return scatterDir(inDir, cos_theta, phi);
}
// :55-65
static inline float scatterProb(uint64_t medium, vec3 inDir, vec3 scDir) {
    const MediumGLSL* m = asMedium(medium);
    if (m == nullptr || m->log_phase == 0) return ORC_INV_4PI;
    float cos_theta = dot(inDir, scDir);
    float log_p = lookUp(m->log_phase, 0.5f * (cos_theta + 1.0f), 0.0f);
    return dm_exp(log_p);
}

// ------------------------------------------------------------------ sh/ray.scatter.glsl
// :46-54 polarized part
static inline void scatterPolRay_impl(Ray& ray, vec3 dir, float cos_theta) {
    vec3 polRef;
    mat4 phase = lookUpPhaseMatrix(ray.medium, cos_theta);
    mat4 rotate = rotatePolRef(ray.direction, ray.polRef, dir, &polRef);
    ray.stokes = phase * (rotate * ray.stokes);
    ray.polRef = polRef;
}
// :12-15, 72-76: importance sampled direction, only mu_s applied
static inline void scatterRayIS(Ray& ray, vec3 dir, bool polarized) {
    if (polarized) {
        float cos_theta = dot(ray.direction, dir);
        scatterPolRay_impl(ray, dir, cos_theta);
    }
    ray.direction = dir;
    ray.lin_contrib *= ray.constants.mu_s;
}
// :30-38, 94-102: sample new direction from the phase function
static inline void scatterRay(Ray& ray, vec2 u, bool polarized) {
    float cos_theta, phi;
    sampleScatterDir(ray.medium, ray.direction, u, &cos_theta, &phi);
    vec3 newDir = scatterDir(ray.direction, cos_theta, phi);
    if (polarized) scatterPolRay_impl(ray, newDir, cos_theta);
    ray.direction = newDir;
    ray.lin_contrib *= ray.constants.mu_s;
}

// ------------------------------------------------------------------ sh/response.common.glsl
struct HitItem {
    vec3 position, direction, normal;
    vec4 stokes;
    vec3 polRef;
    float wavelength, time, contrib;
    int32_t objectId;
};

// sh/ray.response.glsl:18-56 (polarized) and :63-84 (unpolarized)
static inline HitItem createHit(const Ray& ray, vec3 objHitPos, vec3 objHitNormal, int32_t objectId,
                                const mat3& worldToObj, bool polarized) {
    float contrib = ray.lin_contrib * dm_exp(ray.log_contrib);
    vec3 objHitDir = normalize(worldToObj * ray.direction);
    HitItem h{};
    h.position = objHitPos; h.direction = objHitDir; h.normal = objHitNormal;
    h.wavelength = ray.wavelength; h.time = ray.time; h.objectId = objectId;
    if (polarized) {
        vec3 hitPolRef = perpendicularTo(objHitDir, objHitNormal);
        vec3 objPolRef = normalize(worldToObj * ray.polRef);
        mat4 mueller = alignPolRef(objHitDir, objPolRef, hitPolRef);
        vec4 stokes = mueller * ray.stokes;
        contrib *= stokes.x;
        stokes = stokes / stokes.x;
        h.stokes = stokes;
        h.polRef = hitPolRef;
    } else {
        h.stokes = {1.0f, 0.0f, 0.0f, 0.0f};
        h.polRef = v3(0.0f);
    }
    h.contrib = contrib;
    return h;
}

// ------------------------------------------------------------------ sh/util.sample.glsl
// :9-16 (note x = sin(theta) sin(phi), y = sin(theta) cos(phi))
static inline vec3 sphericalToCartessian(float phi, float cos_theta) {
    float sin_theta = std::sqrt(max_(1.0f - cos_theta * cos_theta, 0.0f));
    float sp, cp;
    dm_sincos(phi, &sp, &cp);
    return {sin_theta * sp, sin_theta * cp, cos_theta};
}
// :22-26
static inline vec3 sampleDirectionCone(float cos_opening, vec2 rng) {
    float phi = ORC_TWO_PI * rng.x;
    float cos_theta = (1.0f - rng.y) + cos_opening * rng.y;
    return sphericalToCartessian(phi, cos_theta);
}
// :84-88
static inline vec3 sampleUnitSphere(vec2 rng) {
    float phi = ORC_TWO_PI * rng.x;
    float cos_theta = 2.0f * rng.y - 1.0f;
    return sphericalToCartessian(phi, cos_theta);
}

// ------------------------------------------------------------------ sh/math.sphere.glsl:14-39
static inline vec2 intersectSphere(vec3 center, float radius, vec3 origin, vec3 direction) {
    vec3 f = origin - center;
    float b2 = dot(f, direction);
    float r2 = radius * radius;
    vec3 fd = f - b2 * direction;
    float discr = r2 - dot(fd, fd);
    vec2 result = {F_INF, F_INF};
    if (discr >= 0.0f) {
        float c = dot(f, f) - r2;
        float q = -b2 - signBit(b2) * std::sqrt(discr);
        result = {c / q, q};
        if (result.x > result.y) result = {result.y, result.x};
    }
    return result;
}

}  // namespace orc
