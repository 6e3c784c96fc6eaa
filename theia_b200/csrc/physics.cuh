// Ray state, media, propagation, volume scattering, polarization and hit creation of the
// forward tracers. Each function cites the reference shader it implements
// (sh/ = src/theia/shader/). Device addresses inside Medium / Material records are raw
// 64 bit pointers into the caller's MaterialStore blob (src/theia/material.py:884-1117).
#pragma once
#include "philox.cuh"

namespace theia {

// ------------------------------------------------------------------ sh/lookup.glsl:14-29
// Table1D = { int32 nx; float samples[nx]; }. Read through the read-only path: the tables
// of a medium (10 x 4 KiB) stay resident in L1.
TDEV float lookUp(uint64_t table, float u, float nullValue) {
    if (table == 0) return nullValue;
    const int32_t* t = reinterpret_cast<const int32_t*>(table);
    int nx = __ldg(t);
    const float* samples = reinterpret_cast<const float*>(t + 1);
    u = clamp_(u, 0.0f, 1.0f);
    u *= (float)(nx - 1);
    float fl = floorf(u);
    int lo = (int)fl;
    int hi = (int)ceilf(u);
    float l = u - fl;
    return __ldg(samples + lo) * (1.0f - l) + __ldg(samples + hi) * l;
}

// ------------------------------------------------------------------ sh/material.glsl
#pragma pack(push, 4)
struct MediumRec {  // sh/material.glsl:11-34, src/theia/material.py:122-138
    float lambda_min, lambda_max;
    uint64_t n, vg, mu_a, mu_s, log_phase, phase_sampling, phase_m12, phase_m22, phase_m33, phase_m34;
};
struct MaterialRec {  // sh/material.glsl:92-97, src/theia/material.py:587-595
    uint64_t inside, outside;
    uint32_t flagsInwards, flagsOutwards;
};
#pragma pack(pop)
static_assert(sizeof(MediumRec) == 88, "Medium layout");
static_assert(sizeof(MaterialRec) == 24, "Material layout");

enum : uint32_t {  // sh/material.glsl:79-90
    MATERIAL_BLACK_BODY_BIT = 0x01, MATERIAL_DETECTOR_BIT = 0x02, MATERIAL_LIGHT_SOURCE_BIT = 0x04,
    MATERIAL_NO_REFLECT_FWD_BIT = 0x08, MATERIAL_NO_REFLECT_BWD_BIT = 0x10,
    MATERIAL_NO_TRANSMIT_FWD_BIT = 0x20, MATERIAL_NO_TRANSMIT_BWD_BIT = 0x40,
    MATERIAL_VOLUME_BORDER_BIT = 0x80
};

struct MediumConstants { float n, vg, mu_s, mu_e; };

// Records are 4 byte aligned inside the blob: read them word by word.
TDEV float ldf(uint64_t addr, uint32_t byteOffset) {
    return __ldg(reinterpret_cast<const float*>(addr + byteOffset));
}
TDEV uint32_t ldu(uint64_t addr, uint32_t byteOffset) {
    return __ldg(reinterpret_cast<const uint32_t*>(addr + byteOffset));
}
TDEV uint64_t ldp(uint64_t addr, uint32_t byteOffset) {
    return (uint64_t)ldu(addr, byteOffset) | ((uint64_t)ldu(addr, byteOffset + 4) << 32);
}
enum : uint32_t {  // byte offsets inside MediumRec
    MED_N = 8, MED_VG = 16, MED_MU_A = 24, MED_MU_S = 32, MED_LOG_PHASE = 40, MED_PHASE_SAMPLING = 48,
    MED_M12 = 56, MED_M22 = 64, MED_M33 = 72, MED_M34 = 80
};

// sh/material.glsl:37-39
TDEV float normalize_lambda(uint64_t medium, float lambda) {
    float lmin = ldf(medium, 0), lmax = ldf(medium, 4);
    return clamp_((lambda - lmin) / (lmax - lmin), 0.0f, 1.0f);
}
// sh/material.glsl:50-74
TDEV MediumConstants lookUpMedium(uint64_t medium, float lambda) {
    if (medium == 0) return {1.0f, T_SPEED_OF_LIGHT, 0.0f, 0.0f};
    float u = normalize_lambda(medium, lambda);
    float mu_a = lookUp(ldp(medium, MED_MU_A), u, 0.0f);
    float mu_s = lookUp(ldp(medium, MED_MU_S), u, 0.0f);
    float mu_e = mu_a + mu_s;
    return {lookUp(ldp(medium, MED_N), u, 1.0f), lookUp(ldp(medium, MED_VG), u, T_SPEED_OF_LIGHT), mu_s, mu_e};
}

// ------------------------------------------------------------------ sh/ray.glsl
// RayState (+ PolarizedForwardRay, sh/ray.glsl:22-36, 163-170). Unpolarized kernels never
// touch stokes / polRef, so those registers are dead there.
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
struct WavelengthSample { float wavelength, contrib; };

// sh/ray.glsl:136-143
// sh/ray.glsl:131-133
TDEV float getContrib(const Ray& r) { return r.lin_contrib * dm_exp(r.log_contrib); }
TDEV bool isRayBad(const Ray& r) {
    vec3 p = r.position, d = r.direction;
    return isnan_(p.x) || isnan_(p.y) || isnan_(p.z) || isinf_(p.x) || isinf_(p.y) || isinf_(p.z) ||
           isnan_(d.x) || isnan_(d.y) || isnan_(d.z) || isinf_(d.x) || isinf_(d.y) || isinf_(d.z) ||
           !(length(d) > 0.0f);
}

// sh/lightsource.common.glsl:59-107
template <bool POL>
TDEV SourceRay createSourceRay(vec3 pos, vec3 dir, float startTime, float contrib) {
    SourceRay s;
    s.position = pos; s.direction = dir; s.startTime = startTime; s.contrib = contrib;
    s.stokes = {1.0f, 0.0f, 0.0f, 0.0f};
    if (POL) s.polRef = createLocalCOSY(dir).c0;
    else s.polRef = v3(0.0f);
    return s;
}
TDEV SourceRay createSourceRay(vec3 pos, vec3 dir, vec4 stokes, vec3 polRef, float startTime, float contrib) {
    return SourceRay{pos, dir, stokes, polRef, startTime, contrib};
}
// sh/ray.glsl:59-79, 188-211
TDEV Ray createRay(const SourceRay& s, uint64_t medium, MediumConstants c, WavelengthSample ph) {
    Ray r;
    r.position = s.position; r.direction = s.direction; r.medium = medium;
    r.wavelength = ph.wavelength; r.time = s.startTime;
    r.lin_contrib = ph.contrib * s.contrib; r.log_contrib = 0.0f; r.constants = c;
    r.stokes = s.stokes; r.polRef = s.polRef;
    return r;
}

// ------------------------------------------------------------------ sh/ray.propagate.glsl
typedef theia_propagate_params_t PropagateParams;

// :32-49
TDEV float sampleScatterLength(const Ray& ray, const PropagateParams& p, float u) {
    float dist = p.maxDist;
    float sampleCoef = p.scatterCoefficient >= 0.0f ? p.scatterCoefficient : ray.constants.mu_s;
    bool canScatter = ray.constants.mu_s > 0.0f;
    if (canScatter && sampleCoef != 0.0f) dist = -dm_log(1.0f - u) / sampleCoef;
    return dist;
}
// :70-80
TDEV int updateRay(Ray& ray, float dist, const PropagateParams& p) {
    ray.log_contrib = fmaf(-ray.constants.mu_e, dist, ray.log_contrib);
    ray.time += dist / ray.constants.vg;
    return ray.time <= p.maxTime ? THEIA_RESULT_SUCCESS : THEIA_RESULT_RAY_DECAYED;
}
// :101-130
TDEV void updateRayIS(Ray& ray, float dist, const PropagateParams& p, bool hit) {
    bool canScatter = ray.constants.mu_s > 0.0f;
    if (!canScatter) return;
    float sampleCoef = p.scatterCoefficient >= 0.0f ? p.scatterCoefficient : ray.constants.mu_s;
    ray.log_contrib = fmaf(sampleCoef, dist, ray.log_contrib);
    if (!hit) ray.lin_contrib /= sampleCoef;
}
// :153-166
TDEV int propagateRay(Ray& ray, float dist, const PropagateParams& p) {
    ray.position = fma3(dist, ray.direction, ray.position);
    vec3 q = ray.position;
    bool outside = q.x < p.lowerBBoxCorner[0] || q.y < p.lowerBBoxCorner[1] || q.z < p.lowerBBoxCorner[2] ||
                   q.x > p.upperBBoxCorner[0] || q.y > p.upperBBoxCorner[1] || q.z > p.upperBBoxCorner[2];
    int result = updateRay(ray, dist, p);
    return outside ? THEIA_RESULT_RAY_LOST : result;
}

// ------------------------------------------------------------------ sh/polarization.glsl
// 4x4 Mueller matrices, column major like the GLSL constructors: m[col][row]. The products
// are evaluated literally (including the structural zeros) to keep the bits of the
// reference's operation order; only polarized kernels pay for it.
struct mat4 { float m[4][4]; };
TDEV mat4 mat4_zero() {
    mat4 r;
#pragma unroll
    for (int c = 0; c < 4; ++c)
#pragma unroll
        for (int k = 0; k < 4; ++k) r.m[c][k] = 0.0f;
    return r;
}
TDEV mat4 mat4_identity() {
    mat4 r = mat4_zero();
    r.m[0][0] = 1.0f; r.m[1][1] = 1.0f; r.m[2][2] = 1.0f; r.m[3][3] = 1.0f;
    return r;
}
TDEV vec4 operator*(const mat4& a, vec4 v) {
    float in[4] = {v.x, v.y, v.z, v.w};
    float out[4];
#pragma unroll
    for (int r = 0; r < 4; ++r) {
        float acc = a.m[0][r] * in[0];
        acc = fmaf(a.m[1][r], in[1], acc);
        acc = fmaf(a.m[2][r], in[2], acc);
        acc = fmaf(a.m[3][r], in[3], acc);
        out[r] = acc;
    }
    return {out[0], out[1], out[2], out[3]};
}
// rotation of the reference frame: [[1,0,0,0],[0,c,-s,0],[0,s,c,0],[0,0,0,1]] (:38-43)
TDEV mat4 rotMatrix(float c, float s) {
    mat4 r = mat4_identity();
    r.m[1][1] = c; r.m[1][2] = s;
    r.m[2][1] = -s; r.m[2][2] = c;
    return r;
}
// :22-35
TDEV mat4 alignPolRef(vec3 rayDir, vec3 oldRef, vec3 newRef) {
    float cos_phi = dot(oldRef, newRef);
    float sin_phi = dot(crosser(oldRef, newRef), rayDir);
    float c = 2.0f * cos_phi * cos_phi - 1.0f;
    float s = 2.0f * cos_phi * sin_phi;
    return rotMatrix(c, s);
}
// :39-69
TDEV mat4 rotatePolRef(vec3 dir, vec3 ref, vec3 nw, vec3* new_ref) {
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
TDEV mat4 lookUpPhaseMatrix(uint64_t medium, float cos_theta) {
    if (medium == 0) return mat4_identity();
    float t = 0.5f * (cos_theta + 1.0f);
    float m12 = lookUp(ldp(medium, MED_M12), t, 0.0f);
    float m22 = lookUp(ldp(medium, MED_M22), t, 0.0f);
    float m33 = lookUp(ldp(medium, MED_M33), t, 0.0f);
    float m34 = lookUp(ldp(medium, MED_M34), t, 0.0f);
    mat4 r = mat4_zero();
    r.m[0][0] = 1.0f; r.m[0][1] = m12;
    r.m[1][0] = m12;  r.m[1][1] = m22;
    r.m[2][2] = m33;  r.m[2][3] = -m34;
    r.m[3][2] = m34;  r.m[3][3] = m33;
    return r;
}
// :109-120
TDEV mat4 polarizerMatrix(float p, float s) {
    float att = p * p + s * s;
    float m12 = (p * p - s * s) / att;
    float m33 = (2.0f * p * s) / att;
    mat4 r = mat4_zero();
    r.m[0][0] = 1.0f; r.m[0][1] = m12;
    r.m[1][0] = m12;  r.m[1][1] = 1.0f;
    r.m[2][2] = m33;
    r.m[3][3] = m33;
    return r;
}

// sh/ray.propagate.glsl:193-202
template <bool POL>
TDEV void alignRayToHit(Ray& ray, vec3 hitNormal) {
    if (!POL) return;
    vec3 polRef;
    mat4 rotate = rotatePolRef(ray.direction, ray.polRef, hitNormal, &polRef);
    ray.stokes = rotate * ray.stokes;
    ray.polRef = polRef;
}
// sh/ray.propagate.glsl:237-264
template <bool POL>
TDEV int propagateRayToHit(Ray& ray, vec3 hitPos, vec3 hitNormal, const PropagateParams& p) {
    alignRayToHit<POL>(ray, hitNormal);
    float dist = distance(ray.position, hitPos);
    ray.position = hitPos;
    return updateRay(ray, dist, p);
}

// ------------------------------------------------------------------ sh/scatter.volume.glsl
// :7-26
TDEV vec3 scatterDir(vec3 prevDir, float cos_theta, float phi) {
    prevDir = normalize(prevDir);
    float sin_theta = sqrtf(max_(1.0f - cos_theta * cos_theta, 0.0f));
    float sp, cp;
    dm_sincos(phi, &sp, &cp);
    vec3 local = {sin_theta * cp, sin_theta * sp, cos_theta};
    local = normalize(local);
    mat3 trafo = createLocalCOSY(prevDir);
    return normalize(trafo * local);
}
// :28-46
TDEV float sampleScatterDir(uint64_t medium, vec2 rng, float* cos_theta, float* phi) {
    *phi = rng.x * T_TWO_PI;
    uint64_t sampling = medium != 0 ? ldp(medium, MED_PHASE_SAMPLING) : 0;
    if (sampling != 0) {
        float ct = lookUp(sampling, rng.y, 0.0f);
        ct = clamp_(ct, -1.0f, 1.0f);
        *cos_theta = ct;
        return dm_exp(lookUp(ldp(medium, MED_LOG_PHASE), 0.5f * (ct + 1.0f), 0.0f));
    }
    *cos_theta = 2.0f * rng.y - 1.0f;
    return T_INV_4PI;
}
// :48-53
TDEV vec3 scatter(uint64_t medium, vec3 inDir, vec2 rng, float* p) {
    float cos_theta, phi;
    *p = sampleScatterDir(medium, rng, &cos_theta, &phi);
    return scatterDir(inDir, cos_theta, phi);
}
// :55-65
TDEV float scatterProb(uint64_t medium, vec3 inDir, vec3 scDir) {
    uint64_t logPhase = medium != 0 ? ldp(medium, MED_LOG_PHASE) : 0;
    if (logPhase == 0) return T_INV_4PI;
    float cos_theta = dot(inDir, scDir);
    float log_p = lookUp(logPhase, 0.5f * (cos_theta + 1.0f), 0.0f);
    return dm_exp(log_p);
}

// ------------------------------------------------------------------ sh/ray.scatter.glsl
// :46-54
TDEV void scatterPolRay_impl(Ray& ray, vec3 dir, float cos_theta) {
    vec3 polRef;
    mat4 phase = lookUpPhaseMatrix(ray.medium, cos_theta);
    mat4 rotate = rotatePolRef(ray.direction, ray.polRef, dir, &polRef);
    ray.stokes = phase * (rotate * ray.stokes);
    ray.polRef = polRef;
}
// :12-15, 72-76: direction chosen by the caller (importance sampled), only mu_s applied
template <bool POL>
TDEV void scatterRayIS(Ray& ray, vec3 dir) {
    if (POL) {
        float cos_theta = dot(ray.direction, dir);
        scatterPolRay_impl(ray, dir, cos_theta);
    }
    ray.direction = dir;
    ray.lin_contrib *= ray.constants.mu_s;
}
// :30-38, 94-102: new direction sampled from the phase function
template <bool POL>
TDEV void scatterRay(Ray& ray, vec2 u) {
    float cos_theta, phi;
    sampleScatterDir(ray.medium, u, &cos_theta, &phi);
    vec3 newDir = scatterDir(ray.direction, cos_theta, phi);
    if (POL) scatterPolRay_impl(ray, newDir, cos_theta);
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

// sh/ray.response.glsl:18-56 (polarized), :63-84 (unpolarized)
template <bool POL>
TDEV HitItem createHit(const Ray& ray, vec3 objHitPos, vec3 objHitNormal, int32_t objectId, const mat3& worldToObj) {
    float contrib = ray.lin_contrib * dm_exp(ray.log_contrib);
    vec3 objHitDir = normalize(worldToObj * ray.direction);
    HitItem h;
    h.position = objHitPos; h.direction = objHitDir; h.normal = objHitNormal;
    h.wavelength = ray.wavelength; h.time = ray.time; h.objectId = objectId;
    if (POL) {
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
// :9-16 (x = sin(theta) sin(phi), y = sin(theta) cos(phi))
TDEV vec3 sphericalToCartessian(float phi, float cos_theta) {
    float sin_theta = sqrtf(max_(1.0f - cos_theta * cos_theta, 0.0f));
    float sp, cp;
    dm_sincos(phi, &sp, &cp);
    return {sin_theta * sp, sin_theta * cp, cos_theta};
}
// :22-26
TDEV vec3 sampleDirectionCone(float cos_opening, vec2 rng) {
    float phi = T_TWO_PI * rng.x;
    float cos_theta = (1.0f - rng.y) + cos_opening * rng.y;
    return sphericalToCartessian(phi, cos_theta);
}
// :84-88
TDEV vec3 sampleUnitSphere(vec2 rng) {
    float phi = T_TWO_PI * rng.x;
    float cos_theta = 2.0f * rng.y - 1.0f;
    return sphericalToCartessian(phi, cos_theta);
}

// ------------------------------------------------------------------ sh/math.sphere.glsl:14-39
TDEV vec2 intersectSphere(vec3 center, float radius, vec3 origin, vec3 direction) {
    vec3 f = origin - center;
    float b2 = dot(f, direction);
    float r2 = radius * radius;
    vec3 fd = f - b2 * direction;
    float discr = r2 - dot(fd, fd);
    vec2 result = {T_INF, T_INF};
    if (discr >= 0.0f) {
        float c = dot(f, f) - r2;
        float q = -b2 - signBit(b2) * sqrtf(discr);
        result = {c / q, q};
        if (result.x > result.y) result = {result.y, result.x};
    }
    return result;
}

}  // namespace theia
