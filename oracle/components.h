// ORACLE - TEST INFRASTRUCTURE ONLY. Never included, linked or called by the product path.
//
// CPU restatement of the component snippets: light sources, wavelength sources, targets,
// target guides (sh/lightsource.*.glsl, sh/wavelengthsource.*.glsl, sh/target*.glsl).
#pragma once
#include "physics.h"

namespace orc {

// ------------------------------------------------------------------ wavelength sources
// sh/wavelengthsource.const.glsl:10-12, sh/wavelengthsource.uniform.glsl:12-16
// (first = RNG stream of the first path of the launch: host queues are indexed by idx - first)
static inline WavelengthSample sampleWavelength(const theia_component_t& c, Rng& rng, uint32_t first = 0) {
    if (c.kind == THEIA_WAVELENGTH_CONST) {
        const auto* p = (const theia_wavelength_const_t*)c.params;
        return {p->lambda, 1.0f};
    }
    if (c.kind == THEIA_WAVELENGTH_HOST) {  // sh/wavelengthsource.host.glsl:10-13
        const auto* p = (const theia_wavelength_host_t*)c.params;
        const float* q = (const float*)(uintptr_t)p->queue;
        const uint32_t k = rng.idx - first;
        if (k >= p->capacity) return {F_NAN, 0.0f};
        return {q[k], q[p->capacity + k]};
    }
    if (c.kind == THEIA_WAVELENGTH_FUNCTION) {  // sh/wavelengthsource.function.glsl:13-16
        const auto* p = (const theia_wavelength_function_t*)c.params;
        return {lookUp(p->table, rng.next(), 0.0f), p->contrib};
    }
    const auto* p = (const theia_wavelength_uniform_t*)c.params;
    float u = rng.next();
    float lam = mix_(p->lam_min, p->lam_max, u);
    return {lam, p->contrib};
}

// ------------------------------------------------------------------ sh/random.gamma.glsl:9-29
static inline float sampleGamma(float alpha, Rng& rng) {
    float scale = 1.0f;
    if (alpha < 1.0f) {
        scale = dm_pow(rng.next(), 1.0f / alpha);
        alpha += 1.0f;
    }
    float a = std::sqrt(2.0f * alpha - 1.0f);
    float b = alpha - dm_log(4.0f);
    float c = alpha + 1.0f / alpha;
    // the GLSL loop is unbounded; the oracle bounds it to keep broken inputs from hanging
    for (int it = 0; it < 1000000; ++it) {
        vec2 u = rng.next2();
        float V = a * dm_log(u.x / (1.0f - u.x));
        float X = alpha * dm_exp(V);
        if (b + c * V - X >= dm_log(u.x * u.x * u.y)) return scale * X;
    }
    return F_NAN;
}

// ------------------------------------------------------------------ sh/lightsource.particles.common.glsl
// :52-61
static inline float frank_tamm(float lambda, float n) {
    lambda *= 1e-3f;
    float sin2_theta = 1.0f - 1.0f / (n * n);
    float result = ORC_TWO_PI * 7.2973525693f / (lambda * lambda) * sin2_theta;
    return max_(result, 0.0f);
}
// :68-97
static inline float particle_sampleEmissionAngle(float n, float a, float b, float u) {
    float cos_chev = 1.0f / n;
    float int_lower = 1.0f - dm_exp(-b * dm_pow(1.0f - cos_chev, a));
    float int_upper = 1.0f - dm_exp(-b * dm_pow(1.0f + cos_chev, a));
    u = u * (int_upper + int_lower) - int_lower;
    float x = dm_pow(-dm_log(1.0f - abs_(u)) / b, 1.0f / a);
    return cos_chev - sign_(u) * x;
}

// GLSL uint(floor(x)) for x >= 0 (NaN and negative values: 0)
static inline uint32_t floorToUint(float x) { return x >= 1.0f ? (uint32_t)std::fmin(std::floor(x), 4294967040.0f) : 0u; }
// sh/lightsource.cherenkov.common.glsl:13-17 (energy) and :31-35 (photon count)
static inline float frank_tamm_cherenkov(float n, float lambda, bool photonCount) {
    lambda *= 1e-3f;
    float result = photonCount ? 7.2973525693f / (lambda * lambda) * (1.0f - (1.0f / (n * n)))
                               : 9.04756408986352f / (lambda * lambda * lambda) * (1.0f - (1.0f / (n * n)));
    return max_(result, 0.0f);
}
// sh/lightsource.cherenkov.simple.glsl:18-52 and sh/lightsource.cherenkov.track.glsl:23-66
static inline SourceRay cherenkovSample(const theia_component_t& c, float wavelength, float n, float u, float v, bool polarized) {
    const bool photonCount = (c.flags & THEIA_LIGHT_FLAG_PHOTON_COUNT) != 0;
    vec3 pos, particleDir;
    float time, contrib;
    if (c.kind == THEIA_LIGHT_CHERENKOV) {
        const auto* p = (const theia_light_cherenkov_t*)c.params;
        vec3 s = v3(p->trackStart), e = v3(p->trackEnd);
        pos = {mix_(s.x, e.x, u), mix_(s.y, e.y, u), mix_(s.z, e.z, u)};
        time = mix_(p->startTime, p->endTime, u);
        contrib = ORC_TWO_PI * frank_tamm_cherenkov(n, wavelength, photonCount) * p->trackDist;
        particleDir = v3(p->trackDir);
    } else {
        const auto* p = (const theia_light_cherenkov_track_t*)c.params;
        const uint32_t* track = (const uint32_t*)(uintptr_t)p->track;
        const uint32_t trackLength = track[0];
        u *= (float)trackLength;
        uint32_t vertexIdx = std::min(floorToUint(u), trackLength - 1u);
        u = u - std::floor(u);
        const float* a = (const float*)(track + 1) + 4 * (size_t)vertexIdx;
        vec3 s = {a[0], a[1], a[2]}, e = {a[4], a[5], a[6]};
        pos = {mix_(s.x, e.x, u), mix_(s.y, e.y, u), mix_(s.z, e.z, u)};
        time = mix_(a[3], a[7], u);
        particleDir = normalize(e - s);
        float segmentLength = distance(s, e);
        contrib = ORC_TWO_PI * (float)trackLength * segmentLength;
        contrib *= frank_tamm_cherenkov(n, wavelength, photonCount);
    }
    float cos_theta = 1.0f / n;
    float sin_theta = std::sqrt(max_(1.0f - cos_theta * cos_theta, 0.0f));
    float phi = ORC_TWO_PI * v;
    float sp, cp;
    dm_sincos(phi, &sp, &cp);
    vec3 rayDir = createLocalCOSY(particleDir) * vec3{sin_theta * cp, sin_theta * sp, cos_theta};
    if (polarized) {
        vec3 polRef = normalize(cross(rayDir, particleDir));
        return createSourceRay(pos, rayDir, vec4{1.0f, 1.0f, 0.0f, 0.0f}, polRef, time, contrib);
    }
    return createSourceRay(pos, rayDir, time, contrib, false);
}

// ------------------------------------------------------------------ light sources (forward)
static inline SourceRay sampleLight(const theia_component_t& c, float wavelength, MediumConstants medium,
                                    Rng& rng, bool polarized, uint32_t first = 0) {
    switch (c.kind) {
    case THEIA_LIGHT_SPHERICAL: {  // sh/lightsource.spherical.glsl:16-34
        const auto* p = (const theia_light_spherical_t*)c.params;
        vec3 rayDir = sampleUnitSphere(rng.next2());
        float v = rng.next();
        float startTime = mix_(p->t_min, p->t_max, v);
        return createSourceRay(v3(p->position), rayDir, startTime, p->contribFwd, polarized);
    }
    case THEIA_LIGHT_CONE: {  // sh/lightsource.cone.glsl:22-72
        const auto* p = (const theia_light_cone_t*)c.params;
        vec2 u = rng.next2();
        float phi = ORC_TWO_PI * u.x;
        float cos_theta = (1.0f - u.y) + p->cosOpeningAngle * u.y;
        float sin_theta = std::sqrt(max_(1.0f - cos_theta * cos_theta, 0.0f));
        float sp, cp;
        dm_sincos(phi, &sp, &cp);
        vec3 localDir = {sin_theta * cp, sin_theta * sp, cos_theta};
        mat3 trafo = createLocalCOSY(normalize(v3(p->direction)));
        vec3 rayDir = trafo * localDir;
        float v = rng.next();
        float startTime = mix_(p->t_min, p->t_max, v);
        if (c.flags & THEIA_LIGHT_FLAG_POLARIZED) {
            vec3 polRef = v3(p->polRef);
            polRef = polRef - dot(polRef, rayDir) * rayDir;
            polRef = normalize(polRef);
            return createSourceRay(v3(p->position), rayDir, v4(p->stokes), polRef, startTime, p->contribFwd);
        }
        return createSourceRay(v3(p->position), rayDir, startTime, p->contribFwd, polarized);
    }
    case THEIA_LIGHT_PENCIL: {  // sh/lightsource.pencil.glsl:18-35
        const auto* p = (const theia_light_pencil_t*)c.params;
        float u = rng.next();
        float startTime = mix_(p->t_min, p->t_max, u);
        return createSourceRay(v3(p->position), v3(p->direction), v4(p->stokes), v3(p->polRef), startTime, p->budget);
    }
    case THEIA_LIGHT_CASCADE: {  // sh/lightsource.particles.cascade.glsl:40-69
        const auto* p = (const theia_light_cascade_t*)c.params;
        float z = sampleGamma(p->a_long, rng) * p->b_long;
        vec3 dir = v3(p->direction);
        vec3 rayPos = fma3(z, dir, v3(p->startPosition));
        float time = fmaf(z, ORC_INV_SPEED_OF_LIGHT, p->startTime);
        vec2 u = rng.next2();
        float phi = ORC_TWO_PI * u.x;
        float cos_theta = particle_sampleEmissionAngle(medium.n, p->a_angular, p->b_angular, u.y);
        vec3 rayDir = createLocalCOSY(dir) * sphericalToCartessian(phi, cos_theta);
        float contrib = p->effectiveLength;
        if (!(c.flags & THEIA_LIGHT_FLAG_FRANK_TAMM_IS)) contrib *= frank_tamm(wavelength, medium.n);
        return createSourceRay(rayPos, rayDir, time, contrib, polarized);
    }
    case THEIA_LIGHT_MUON: {  // sh/lightsource.particles.muon.glsl:28-59
        const auto* p = (const theia_light_muon_t*)c.params;
        float u = rng.next();
        vec3 s = v3(p->startPosition), e = v3(p->endPosition);
        vec3 position = {mix_(s.x, e.x, u), mix_(s.y, e.y, u), mix_(s.z, e.z, u)};
        float startTime = mix_(p->startTime, p->endTime, u);
        float contrib = distance(s, e);
        contrib *= p->energyScale;
        vec2 v = rng.next2();
        float phi = ORC_TWO_PI * v.x;
        float cos_theta = particle_sampleEmissionAngle(medium.n, p->a_angular, p->b_angular, v.y);
        vec3 trackDir = normalize(e - s);
        vec3 rayDir = createLocalCOSY(trackDir) * sphericalToCartessian(phi, cos_theta);
        if (!(c.flags & THEIA_LIGHT_FLAG_FRANK_TAMM_IS)) contrib *= frank_tamm(wavelength, medium.n);
        return createSourceRay(position, rayDir, startTime, contrib, polarized);
    }
    case THEIA_LIGHT_CHERENKOV:
    case THEIA_LIGHT_CHERENKOV_TRACK: {
        float u = rng.next();
        float v = rng.next();
        return cherenkovSample(c, wavelength, medium.n, u, v, polarized);
    }
    case THEIA_LIGHT_HOST: {  // sh/lightsource.host.glsl:10-17, LOAD_SAMPLE of sh/lightsource.queue.glsl:40-56
        const auto* p = (const theia_light_host_t*)c.params;
        const float* q = (const float*)(uintptr_t)p->queue;
        const size_t C = p->capacity;
        const uint32_t k = rng.idx - first;
        if (k >= p->capacity) break;
        const float* e = q + k;
        vec3 pos = {e[0], e[C], e[2 * C]};
        vec3 dir = {e[3 * C], e[4 * C], e[5 * C]};
        if (p->polarized) {
            vec4 stokes = {e[6 * C], e[7 * C], e[8 * C], e[9 * C]};
            vec3 polRef = {e[10 * C], e[11 * C], e[12 * C]};
            return createSourceRay(pos, dir, stokes, polRef, e[13 * C], e[14 * C]);
        }
        return createSourceRay(pos, dir, e[6 * C], e[7 * C], polarized);
    }
    }
    SourceRay bad{};  // sh/lightsource.common.glsl:28-37
    bad.position = v3(F_NAN); bad.direction = v3(F_NAN);
    bad.startTime = F_NAN; bad.contrib = F_NAN;
    return bad;
}

// ------------------------------------------------------------------ sh/target.common.glsl
struct TargetSample {
    vec3 position, normal;
    float dist;
    vec3 objPosition, objNormal;
    float prob;
    bool valid;
    vec3 offset;
    mat3 worldToObj;
};
static inline TargetSample createTargetSampleMiss() {  // :77-91
    TargetSample t{};
    t.dist = F_INF; t.prob = 0.0f; t.valid = false; t.worldToObj = mat3_diag(1.0f);
    return t;
}

// GLSL column major mat3 of a parameter block
static inline mat3 mat3Of(const float m[9]) { return {{m[0], m[1], m[2]}, {m[3], m[4], m[5]}, {m[6], m[7], m[8]}}; }
static inline mat3 transposeOf(const float m[9]) { return {{m[0], m[3], m[6]}, {m[1], m[4], m[7]}, {m[2], m[5], m[8]}}; }

// sh/util.sample.glsl:31-52 (concentric mapping)
static inline vec3 sampleUnitDisk(vec2 rng) {
    rng = {2.0f * rng.x - 1.0f, 2.0f * rng.y - 1.0f};
    if (rng.x == 0.0f && rng.y == 0.0f) return v3(0.0f);
    float phi, r;
    if (abs_(rng.x) > abs_(rng.y)) {
        r = rng.x;
        phi = ORC_PI_OVER_FOUR * (rng.y / rng.x);
    } else {
        r = rng.y;
        phi = ORC_PI_OVER_TWO - ORC_PI_OVER_FOUR * (rng.x / rng.y);
    }
    float sp, cp;
    dm_sincos(phi, &sp, &cp);
    return {r * cp, r * sp, 0.0f};
}
static inline float length2(float x, float y) { return std::sqrt(fmaf(y, y, x * x)); }

// the plane shapes (flat, disk) as (objToWorld, centre, normal, prob) + an inside test
struct PlaneShape { const float* objToWorld; vec3 center, normal; float prob; bool flat; float width, length, radius; };
static inline PlaneShape planeShape(bool flat, const void* params) {
    PlaneShape s{};
    s.flat = flat;
    if (flat) {
        const auto* p = (const theia_target_flat_t*)params;
        s.objToWorld = p->objToWorld; s.center = v3(p->position); s.normal = v3(p->normal); s.prob = p->prob;
        s.width = p->width; s.length = p->length;
    } else {
        const auto* p = (const theia_target_disk_t*)params;
        s.objToWorld = p->objToWorld; s.center = v3(p->position); s.normal = v3(p->normal); s.prob = p->prob;
        s.radius = p->radius;
    }
    return s;
}
static inline vec3 planeLocalSample(const PlaneShape& s, Rng& rng) {
    if (s.flat) {  // sh/target.flat.glsl:17-20, sh/target_guide.flat.glsl:17-20
        vec2 u = rng.next2();
        return {s.width * (u.x - 0.5f), s.length * (u.y - 0.5f), 0.0f};
    }
    return s.radius * sampleUnitDisk(rng.next2());  // sh/target.disk.glsl:17, sh/target_guide.disk.glsl:16
}
static inline bool planeInside(const PlaneShape& s, vec3 localPos) {
    if (s.flat) return 2.0f * abs_(localPos.x) <= s.width && 2.0f * abs_(localPos.y) <= s.length;
    return length2(localPos.x, localPos.y) <= s.radius;
}
struct PlaneHit { vec3 localPos, pos, normal; float t, side; };
// sh/target.flat.glsl:46-62, sh/target.disk.glsl:42-57, sh/target_guide.flat.glsl:28-36, sh/target_guide.disk.glsl:23-31
static inline PlaneHit planeIntersect(const PlaneShape& s, vec3 observer, vec3 direction) {
    mat3 worldToObj = transposeOf(s.objToWorld);
    vec3 localObs = worldToObj * (observer - s.center);
    vec3 localDir = worldToObj * direction;
    PlaneHit h;
    h.t = -localObs.z / localDir.z;
    h.localPos = fma3(h.t, localDir, localObs);
    h.pos = mat3Of(s.objToWorld) * h.localPos + s.center;
    h.side = signBit(localObs.z);
    h.normal = s.normal * h.side;
    return h;
}
static inline bool isPlaneTarget(uint32_t kind) { return kind == THEIA_TARGET_FLAT || kind == THEIA_TARGET_DISK; }

// sh/target.sphere.glsl:18-43 and sh/target.sphere.inner.glsl:18-36; sh/target.flat.glsl:15-44,
// sh/target.disk.glsl:15-40
static inline TargetSample sampleTarget(const theia_component_t& c, vec3 observer, Rng& rng) {
    if (isPlaneTarget(c.kind)) {
        PlaneShape s = planeShape(c.kind == THEIA_TARGET_FLAT, c.params);
        vec3 localPos = planeLocalSample(s, rng);
        vec3 pos = mat3Of(s.objToWorld) * localPos + s.center;
        float side = sign_(dot(s.normal, observer - pos));
        vec3 normal = s.normal * side;
        TargetSample t{};
        t.position = pos; t.normal = normal; t.dist = distance(observer, pos);
        t.objPosition = localPos; t.objNormal = {0.0f, 0.0f, side};
        t.prob = s.prob; t.valid = dot(normal, normal) != 0.0f;
        t.worldToObj = transposeOf(s.objToWorld);
        t.offset = -(t.worldToObj * s.center);
        return t;
    }
    const auto* p = (const theia_target_sphere_t*)c.params;
    vec3 center = v3(p->position);
    TargetSample t{};
    t.offset = v3(p->invPos);
    t.worldToObj = mat3_diag(p->invRad);
    if (c.kind == THEIA_TARGET_SPHERE) {
        vec3 dir = normalize(observer - center);
        float d = distance(observer, center);
        float cosOpening = p->radius / d;
        vec3 normal = createLocalCOSY(dir) * sampleDirectionCone(cosOpening, rng.next2());
        vec3 pos = normal * p->radius + center;
        float prob = p->prob / (1.0f - cosOpening);
        bool valid = !isinf_(prob);
        if (!valid) prob = 0.0f;
        t.position = pos; t.normal = normal; t.dist = distance(pos, observer);
        t.objPosition = normal; t.objNormal = normal; t.prob = prob; t.valid = valid;
    } else {
        vec3 normal = sampleUnitSphere(rng.next2());
        vec3 pos = p->radius * normal + center;
        t.position = pos; t.normal = -normal; t.dist = distance(observer, pos);
        t.objPosition = normal; t.objNormal = -normal; t.prob = p->prob; t.valid = true;
    }
    return t;
}
// sh/target.sphere.glsl:45-78 and sh/target.sphere.inner.glsl:38-56
static inline TargetSample intersectTarget(const theia_component_t& c, vec3 observer, vec3 direction) {
    if (isPlaneTarget(c.kind)) {  // sh/target.flat.glsl:46-74, sh/target.disk.glsl:42-68
        PlaneShape s = planeShape(c.kind == THEIA_TARGET_FLAT, c.params);
        PlaneHit h = planeIntersect(s, observer, direction);
        bool valid = h.t > 0.0f && planeInside(s, h.localPos);
        TargetSample r{};
        r.position = h.pos; r.normal = h.normal; r.dist = h.t;
        r.objPosition = h.localPos; r.objNormal = {0.0f, 0.0f, h.side};
        r.prob = s.prob * (valid ? 1.0f : 0.0f); r.valid = valid;
        r.worldToObj = transposeOf(s.objToWorld);
        r.offset = -(r.worldToObj * s.center);
        return r;
    }
    const auto* p = (const theia_target_sphere_t*)c.params;
    vec3 center = v3(p->position);
    vec2 ts = intersectSphere(center, p->radius, observer, direction);
    if (c.kind == THEIA_TARGET_SPHERE) {
        float t = ts.x;
        bool hit = (t > 0.0f) && !isinf_(t);
        if (!hit) return createTargetSampleMiss();
        vec3 pos = fma3(t, direction, observer);
        vec3 nrm = normalize(pos - center);
        float d = distance(observer, center);
        float cosOpening = p->radius / d;
        float prob = p->prob / (1.0f - cosOpening);
        bool valid = !isinf_(prob);
        if (!valid) prob = 0.0f;
        TargetSample r{};
        r.position = pos; r.normal = nrm; r.dist = t; r.objPosition = nrm; r.objNormal = nrm;
        r.prob = prob; r.valid = valid; r.offset = v3(p->invPos); r.worldToObj = mat3_diag(p->invRad);
        return r;
    }
    float t = ts.y;
    bool hit = (t > 0.0f) && !isinf_(t);
    vec3 pos = fma3(t, direction, observer);
    vec3 nrm = normalize(center - pos);
    TargetSample r{};
    r.position = pos; r.normal = nrm; r.dist = t; r.objPosition = -nrm; r.objNormal = nrm;
    r.prob = p->prob; r.valid = hit; r.offset = v3(p->invPos); r.worldToObj = mat3_diag(p->invRad);
    return r;
}
// sh/target.sphere.glsl:80-83, sh/target.sphere.inner.glsl:58-60
static inline bool isOccludedByTarget(const theia_component_t& c, vec3 position) {
    if (isPlaneTarget(c.kind)) return false;  // sh/target.flat.glsl:76-79, sh/target.disk.glsl:70-73
    const auto* p = (const theia_target_sphere_t*)c.params;
    float d = distance(position, v3(p->position));
    return c.kind == THEIA_TARGET_SPHERE ? d <= p->radius : d >= p->radius;
}

// ------------------------------------------------------------------ sh/target_guide.*.glsl
struct TargetGuideSample { vec3 dir; float dist, prob; };

// sh/target_guide.common.glsl:11-33 (from a sampled point) and :35-52 (from a direction)
static inline TargetGuideSample guideSampleFromPoint(vec3 observer, vec3 targetPos, vec3 targetNrm, float targetProb) {
    vec3 dir = targetPos - observer;
    float d2 = dot(dir, dir);
    dir = normalize(dir);
    float cosNormal = dot(dir, targetNrm);
    float prob = targetProb * d2 / abs_(cosNormal);
    if (isinf_(prob)) prob = 0.0f;
    prob *= (cosNormal < 0.0f) ? 1.0f : 0.0f;
    return {dir, distance(targetPos, observer), prob};
}
static inline TargetGuideSample guideSampleFromDir(vec3 dir, float dist, vec3 targetNrm, float targetProb) {
    float cosNormal = dot(dir, targetNrm);
    float prob = targetProb * dist * dist / abs_(cosNormal);
    if (isinf_(prob)) prob = 0.0f;
    prob *= (cosNormal < 0.0f) ? 1.0f : 0.0f;
    return {dir, dist, prob};
}
static inline bool isPlaneGuide(uint32_t kind) { return kind == THEIA_GUIDE_FLAT || kind == THEIA_GUIDE_DISK; }

static inline void guideCone(const theia_guide_sphere_t* p, vec3 observer, float* d, vec3* viewDir,
                             float* cosMin, float* prob) {
    vec3 center = v3(p->position);
    *d = distance(center, observer);
    *viewDir = normalize(center - observer);
    float sinMax = p->radius / *d;
    float sin2Max = sinMax * sinMax;
    float cm = 1.0f - std::sqrt(max_(1.0f - sin2Max, 0.0f));
    if (sin2Max < 0.00068523f) cm = 0.5f * sin2Max;
    *cosMin = cm;
    float pr = 1.0f / (ORC_TWO_PI * cm);
    pr *= (*d > p->radius) ? 1.0f : 0.0f;
    *prob = pr;
}
// :15-51
static inline TargetGuideSample sampleTargetGuide(const theia_component_t& c, vec3 observer, Rng& rng) {
    if (isPlaneGuide(c.kind)) {  // sh/target_guide.flat.glsl:15-26, sh/target_guide.disk.glsl:14-21
        PlaneShape s = planeShape(c.kind == THEIA_GUIDE_FLAT, c.params);
        vec3 localPos = planeLocalSample(s, rng);
        vec3 pos = mat3Of(s.objToWorld) * localPos + s.center;
        return guideSampleFromPoint(observer, pos, s.normal, s.prob);
    }
    const auto* p = (const theia_guide_sphere_t*)c.params;
    float d, cosMin, prob;
    vec3 viewDir;
    guideCone(p, observer, &d, &viewDir, &cosMin, &prob);
    float cosTheta = 1.0f - cosMin * rng.next();
    float sinTheta = std::sqrt(max_(1.0f - cosTheta * cosTheta, 0.0f));
    float phi = ORC_TWO_PI * rng.next();
    float sp, cp;
    dm_sincos(phi, &sp, &cp);
    vec3 dir = createLocalCOSY(viewDir) * vec3{sinTheta * sp, sinTheta * cp, cosTheta};
    return {dir, d + p->radius, prob};
}
// :53-81
static inline TargetGuideSample evalTargetGuide(const theia_component_t& c, vec3 observer, vec3 direction) {
    if (isPlaneGuide(c.kind)) {  // sh/target_guide.flat.glsl:28-47, sh/target_guide.disk.glsl:23-43
        PlaneShape s = planeShape(c.kind == THEIA_GUIDE_FLAT, c.params);
        PlaneHit h = planeIntersect(s, observer, direction);
        bool valid = h.t > 0.0f && !isinf_(h.t) && !(h.t != h.t) && planeInside(s, h.localPos);
        return guideSampleFromDir(direction, h.t, s.normal, s.prob * (valid ? 1.0f : 0.0f));
    }
    const auto* p = (const theia_guide_sphere_t*)c.params;
    float d, cosMin, prob;
    vec3 viewDir;
    guideCone(p, observer, &d, &viewDir, &cosMin, &prob);
    float cosDir = dot(viewDir, direction);
    prob *= (cosMin >= 1.0f - cosDir) ? 1.0f : 0.0f;
    return {direction, d + p->radius, prob};
}

}  // namespace orc
