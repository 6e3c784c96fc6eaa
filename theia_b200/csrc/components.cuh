// Component snippets of the reference as device functions: wavelength sources, light
// sources (forward sampling), volume-tracer targets and scene-tracer target guides
// (sh/wavelengthsource.*.glsl, sh/lightsource.*.glsl, sh/random.gamma.glsl, sh/target.*.glsl,
// sh/target_guide.*.glsl). The byte-identical parameter blocks of include/theia_b200.h are
// copied into the kernel's constant parameter space by the launcher.
#pragma once
#include "physics.cuh"

namespace theia {

union LightBlock {
    theia_light_spherical_t spherical;
    theia_light_cone_t cone;
    theia_light_pencil_t pencil;
    theia_light_cascade_t cascade;
    theia_light_muon_t muon;
    theia_light_host_t host;
    theia_light_cherenkov_t cherenkov;
    theia_light_cherenkov_track_t track;
};
union WavelengthBlock {
    theia_wavelength_const_t constant;
    theia_wavelength_uniform_t uniform;
    theia_wavelength_host_t host;
    theia_wavelength_function_t function;
};
// first = RNG stream of the first path of the launch: host queues are indexed by idx - first
struct LightSel { uint32_t kind, flags, first; LightBlock p; };
struct WavelengthSel { uint32_t kind, first; WavelengthBlock p; };
union TargetBlock {
    theia_target_sphere_t sphere;
    theia_target_flat_t flat;
    theia_target_disk_t disk;
};
union GuideBlock {
    theia_guide_sphere_t sphere;
    theia_target_flat_t flat;
    theia_target_disk_t disk;
};
struct TargetSel { uint32_t kind; TargetBlock p; };
struct GuideSel { uint32_t kind; GuideBlock p; };

// ------------------------------------------------------------------ wavelength sources
// sh/wavelengthsource.const.glsl:10-12, sh/wavelengthsource.uniform.glsl:12-16
// FULL = false (here and below): the caller guarantees that none of the components that were added late
// (host / function / Cherenkov sources, plane targets, responses beyond record and histogram) is in
// use, and their code is left out: the volume kernel runs at the limit of the instruction cache.
template <bool FULL = true>
TDEV WavelengthSample sampleWavelength(const WavelengthSel& c, Rng& rng) {
    if (c.kind == THEIA_WAVELENGTH_CONST) return {c.p.constant.lambda, 1.0f};
    if (!FULL) {
        float u0 = rng.next();
        return {mix_(c.p.uniform.lam_min, c.p.uniform.lam_max, u0), c.p.uniform.contrib};
    }
    if (c.kind == THEIA_WAVELENGTH_HOST) {  // sh/wavelengthsource.host.glsl:10-13
        const float* q = reinterpret_cast<const float*>(c.p.host.queue);
        const uint32_t k = rng.idx - c.first, C = c.p.host.capacity;
        if (k >= C) return {T_NAN, 0.0f};
        return {__ldg(q + k), __ldg(q + C + k)};
    }
    if (c.kind == THEIA_WAVELENGTH_FUNCTION)  // sh/wavelengthsource.function.glsl:13-16
        return {lookUp(c.p.function.table, rng.next(), 0.0f), c.p.function.contrib};
    float u = rng.next();
    float lam = mix_(c.p.uniform.lam_min, c.p.uniform.lam_max, u);
    return {lam, c.p.uniform.contrib};
}

// ------------------------------------------------------------------ sh/random.gamma.glsl:9-29
// Cheng's GA rejection sampler: data dependent number of draws (1 + 2k).
static __device__ __noinline__ float sampleGamma(float alpha, Rng& rng) {
    float scale = 1.0f;
    if (alpha < 1.0f) {
        scale = dm_pow(rng.next(), 1.0f / alpha);
        alpha += 1.0f;
    }
    float a = sqrtf(2.0f * alpha - 1.0f);
    float b = alpha - dm_log(4.0f);
    float c = alpha + 1.0f / alpha;
    for (int it = 0; it < 1000000; ++it) {
        vec2 u = rng.next2();
        float V = a * dm_log(u.x / (1.0f - u.x));
        float X = alpha * dm_exp(V);
        if (b + c * V - X >= dm_log(u.x * u.x * u.y)) return scale * X;
    }
    return T_NAN;
}

// ------------------------------------------------------------------ sh/lightsource.particles.common.glsl
// :52-61
TDEV float frank_tamm(float lambda, float n) {
    lambda *= 1e-3f;
    float sin2_theta = 1.0f - 1.0f / (n * n);
    float result = T_TWO_PI * 7.2973525693f / (lambda * lambda) * sin2_theta;
    return max_(result, 0.0f);
}
// :68-97
static __device__ __noinline__ float particle_sampleEmissionAngle(float n, float a, float b, float u) {
    float cos_chev = 1.0f / n;
    float int_lower = 1.0f - dm_exp(-b * dm_pow(1.0f - cos_chev, a));
    float int_upper = 1.0f - dm_exp(-b * dm_pow(1.0f + cos_chev, a));
    u = u * (int_upper + int_lower) - int_lower;
    float x = dm_pow(-dm_log(1.0f - abs_(u)) / b, 1.0f / a);
    return cos_chev - sign_(u) * x;
}

// One trial of the rejection loop of sampleGamma; constants as in sampleGamma.
struct GammaConsts { float alpha, a, b, c; };
TDEV GammaConsts gammaConsts(float alpha) {  // alpha >= 1 (after the rescaling step)
    return {alpha, sqrtf(2.0f * alpha - 1.0f), alpha - dm_log(4.0f), alpha + 1.0f / alpha};
}
TDEV bool gammaTrial(const GammaConsts& g, Rng& rng, float* X_out) {
    vec2 u = rng.next2();
    float V = g.a * dm_log(u.x / (1.0f - u.x));
    float X = g.alpha * dm_exp(V);
    *X_out = X;
    return g.b + g.c * V - X >= dm_log(u.x * u.x * u.y);
}

// Gamma variate sampled ahead of sampleLight (wavefront generate kernel): the value and the
// draw counter after the rejection loop.
struct PreGamma { bool valid; float value; uint32_t dimAfter; };

// sh/lightsource.host.glsl:10-17, LOAD_SAMPLE of sh/lightsource.queue.glsl:40-56 (out of line: see
// commitHitOther in launch.cuh)
template <bool POL>
static __device__ __noinline__ SourceRay hostLightSample(const theia_light_host_t& p, uint32_t k) {
    const float* e = reinterpret_cast<const float*>(p.queue) + k;
    const size_t C = p.capacity;
    vec3 pos = {__ldg(e), __ldg(e + C), __ldg(e + 2 * C)};
    vec3 dir = {__ldg(e + 3 * C), __ldg(e + 4 * C), __ldg(e + 5 * C)};
    if (p.polarized) {
        vec4 stokes = {__ldg(e + 6 * C), __ldg(e + 7 * C), __ldg(e + 8 * C), __ldg(e + 9 * C)};
        vec3 polRef = {__ldg(e + 10 * C), __ldg(e + 11 * C), __ldg(e + 12 * C)};
        return createSourceRay(pos, dir, stokes, polRef, __ldg(e + 13 * C), __ldg(e + 14 * C));
    }
    return createSourceRay<POL>(pos, dir, __ldg(e + 6 * C), __ldg(e + 7 * C));
}

// GLSL uint(floor(x)) for x >= 0 (NaN and negative values: 0)
TDEV uint32_t floorToUint(float x) { return x >= 1.0f ? (uint32_t)fminf(floorf(x), 4294967040.0f) : 0u; }
// sh/lightsource.cherenkov.common.glsl:13-17 (energy) and :31-35 (photon count)
TDEV float frank_tamm_cherenkov(float n, float lambda, bool photonCount) {
    lambda *= 1e-3f;
    float result = photonCount ? 7.2973525693f / (lambda * lambda) * (1.0f - (1.0f / (n * n)))
                               : 9.04756408986352f / (lambda * lambda * lambda) * (1.0f - (1.0f / (n * n)));
    return max_(result, 0.0f);
}
// Cherenkov cone around a particle moving with beta = 1 from `start` to `end`; u, v are the two
// draws. sh/lightsource.cherenkov.simple.glsl:18-52 (one segment, trackLength = 1) and
// sh/lightsource.cherenkov.track.glsl:23-66 (segment picked by u). Out of line, see commitHitOther.
template <bool POL>
static __device__ __noinline__ SourceRay cherenkovSample(const LightSel& c, float wavelength, float n, float u, float v) {
    const bool photonCount = (c.flags & THEIA_LIGHT_FLAG_PHOTON_COUNT) != 0;
    vec3 pos, particleDir;
    float time, contrib;
    if (c.kind == THEIA_LIGHT_CHERENKOV) {
        const theia_light_cherenkov_t& p = c.p.cherenkov;
        vec3 s = v3(p.trackStart), e = v3(p.trackEnd);
        pos = {mix_(s.x, e.x, u), mix_(s.y, e.y, u), mix_(s.z, e.z, u)};
        time = mix_(p.startTime, p.endTime, u);
        contrib = T_TWO_PI * frank_tamm_cherenkov(n, wavelength, photonCount) * p.trackDist;
        particleDir = v3(p.trackDir);
    } else {
        const uint32_t* track = reinterpret_cast<const uint32_t*>(c.p.track.track);
        const uint32_t trackLength = __ldg(track);
        const float4* vertices = reinterpret_cast<const float4*>(track + 1);  // 4 byte aligned only
        u *= (float)trackLength;
        uint32_t vertexIdx = min(floorToUint(u), trackLength - 1u);
        u = u - floorf(u);
        const float* a = reinterpret_cast<const float*>(vertices) + 4 * (size_t)vertexIdx;
        vec3 s = {__ldg(a), __ldg(a + 1), __ldg(a + 2)}, e = {__ldg(a + 4), __ldg(a + 5), __ldg(a + 6)};
        pos = {mix_(s.x, e.x, u), mix_(s.y, e.y, u), mix_(s.z, e.z, u)};
        time = mix_(__ldg(a + 3), __ldg(a + 7), u);
        particleDir = normalize(e - s);
        float segmentLength = distance(s, e);
        contrib = T_TWO_PI * (float)trackLength * segmentLength;
        contrib *= frank_tamm_cherenkov(n, wavelength, photonCount);
    }
    float cos_theta = 1.0f / n;
    float sin_theta = sqrtf(max_(1.0f - cos_theta * cos_theta, 0.0f));
    float phi = T_TWO_PI * v;
    float sp, cp;
    dm_sincos(phi, &sp, &cp);
    vec3 rayDir = createLocalCOSY(particleDir) * vec3{sin_theta * cp, sin_theta * sp, cos_theta};
    if (POL) {
        vec3 polRef = normalize(cross(rayDir, particleDir));
        return createSourceRay(pos, rayDir, vec4{1.0f, 1.0f, 0.0f, 0.0f}, polRef, time, contrib);
    }
    return createSourceRay<false>(pos, rayDir, time, contrib);
}

// ------------------------------------------------------------------ light sources (forward)
// KIND >= 0: the kind of the light source is known at compile time (the kernels of the BASELINE configurations are
// instantiated per light source; the others decide at run time)
template <bool POL, bool FULL = true, int KIND = -1>
TDEV SourceRay sampleLight(const LightSel& c, float wavelength, MediumConstants medium, Rng& rng,
                           PreGamma pre = PreGamma{false, 0.0f, 0u}) {
    switch (KIND >= 0 ? (uint32_t)KIND : (uint32_t)c.kind) {
    case THEIA_LIGHT_SPHERICAL: {  // sh/lightsource.spherical.glsl:16-34
        const theia_light_spherical_t& p = c.p.spherical;
        vec3 rayDir = sampleUnitSphere(rng.next2());
        float v = rng.next();
        float startTime = mix_(p.t_min, p.t_max, v);
        return createSourceRay<POL>(v3(p.position), rayDir, startTime, p.contribFwd);
    }
    case THEIA_LIGHT_CONE: {  // sh/lightsource.cone.glsl:22-72
        const theia_light_cone_t& p = c.p.cone;
        vec2 u = rng.next2();
        float phi = T_TWO_PI * u.x;
        float cos_theta = (1.0f - u.y) + p.cosOpeningAngle * u.y;
        float sin_theta = sqrtf(max_(1.0f - cos_theta * cos_theta, 0.0f));
        float sp, cp;
        dm_sincos(phi, &sp, &cp);
        vec3 localDir = {sin_theta * cp, sin_theta * sp, cos_theta};
        mat3 trafo = createLocalCOSY(normalize(v3(p.direction)));
        vec3 rayDir = trafo * localDir;
        float v = rng.next();
        float startTime = mix_(p.t_min, p.t_max, v);
        if (c.flags & THEIA_LIGHT_FLAG_POLARIZED) {
            vec3 polRef = v3(p.polRef);
            polRef = polRef - dot(polRef, rayDir) * rayDir;
            polRef = normalize(polRef);
            return createSourceRay(v3(p.position), rayDir, v4(p.stokes), polRef, startTime, p.contribFwd);
        }
        return createSourceRay<POL>(v3(p.position), rayDir, startTime, p.contribFwd);
    }
    case THEIA_LIGHT_PENCIL: {  // sh/lightsource.pencil.glsl:18-35
        const theia_light_pencil_t& p = c.p.pencil;
        float u = rng.next();
        float startTime = mix_(p.t_min, p.t_max, u);
        return createSourceRay(v3(p.position), v3(p.direction), v4(p.stokes), v3(p.polRef), startTime, p.budget);
    }
    case THEIA_LIGHT_CASCADE: {  // sh/lightsource.particles.cascade.glsl:40-69
        const theia_light_cascade_t& p = c.p.cascade;
        float gamma;
        if (pre.valid) { gamma = pre.value; rng.dim = pre.dimAfter; }
        else gamma = sampleGamma(p.a_long, rng);
        float z = gamma * p.b_long;
        vec3 dir = v3(p.direction);
        vec3 rayPos = fma3(z, dir, v3(p.startPosition));
        float time = fmaf(z, T_INV_SPEED_OF_LIGHT, p.startTime);
        vec2 u = rng.next2();
        float phi = T_TWO_PI * u.x;
        float cos_theta = particle_sampleEmissionAngle(medium.n, p.a_angular, p.b_angular, u.y);
        vec3 rayDir = createLocalCOSY(dir) * sphericalToCartessian(phi, cos_theta);
        float contrib = p.effectiveLength;
        if (!(c.flags & THEIA_LIGHT_FLAG_FRANK_TAMM_IS)) contrib *= frank_tamm(wavelength, medium.n);
        return createSourceRay<POL>(rayPos, rayDir, time, contrib);
    }
    case THEIA_LIGHT_MUON: {  // sh/lightsource.particles.muon.glsl:28-59
        const theia_light_muon_t& p = c.p.muon;
        float u = rng.next();
        vec3 s = v3(p.startPosition), e = v3(p.endPosition);
        vec3 position = {mix_(s.x, e.x, u), mix_(s.y, e.y, u), mix_(s.z, e.z, u)};
        float startTime = mix_(p.startTime, p.endTime, u);
        float contrib = distance(s, e);
        contrib *= p.energyScale;
        vec2 v = rng.next2();
        float phi = T_TWO_PI * v.x;
        float cos_theta = particle_sampleEmissionAngle(medium.n, p.a_angular, p.b_angular, v.y);
        vec3 trackDir = normalize(e - s);
        vec3 rayDir = createLocalCOSY(trackDir) * sphericalToCartessian(phi, cos_theta);
        if (!(c.flags & THEIA_LIGHT_FLAG_FRANK_TAMM_IS)) contrib *= frank_tamm(wavelength, medium.n);
        return createSourceRay<POL>(position, rayDir, startTime, contrib);
    }
    case THEIA_LIGHT_CHERENKOV:
    case THEIA_LIGHT_CHERENKOV_TRACK: {
        if (!FULL) break;
        float u = rng.next();
        float v = rng.next();
        return cherenkovSample<POL>(c, wavelength, medium.n, u, v);
    }
    case THEIA_LIGHT_HOST: {
        if (!FULL) break;
        const uint32_t k = rng.idx - c.first;
        if (k >= c.p.host.capacity) break;
        return hostLightSample<POL>(c.p.host, k);
    }
    }
    SourceRay bad;  // sh/lightsource.common.glsl:28-37
    bad.position = v3(T_NAN); bad.direction = v3(T_NAN);
    bad.stokes = {1.0f, 0.0f, 0.0f, 0.0f}; bad.polRef = v3(0.0f);
    bad.startTime = T_NAN; bad.contrib = T_NAN;
    return bad;
}

// sh/tracer.volume.forward.glsl:223-229, sh/tracer.scene.forward.glsl:50-56
template <bool POL, bool FULL = true, int KIND = -1>
TDEV Ray sampleRay(const WavelengthSel& wl, const LightSel& light, uint64_t medium, Rng& rng,
                   PreGamma pre = PreGamma{false, 0.0f, 0u}) {
    WavelengthSample photon = sampleWavelength<FULL>(wl, rng);
    MediumConstants constants = lookUpMedium(medium, photon.wavelength);
    SourceRay lightRay = sampleLight<POL, FULL, KIND>(light, photon.wavelength, constants, rng, pre);
    return createRay(lightRay, medium, constants, photon);
}

// ------------------------------------------------------------------ sh/target.common.glsl
// The world -> object transformation of a sample (TargetSample.offset / .worldToObj of the
// reference) depends on the parameter block only: targetWorldToObj() rebuilds it where a hit is
// created instead of carrying a matrix per sample.
struct TargetSample {
    vec3 position, normal;
    float dist;
    vec3 objPosition, objNormal;
    float prob;
    bool valid;
};
TDEV TargetSample createTargetSampleMiss() {  // :77-91
    TargetSample t;
    t.position = v3(0.0f); t.normal = v3(0.0f); t.objPosition = v3(0.0f); t.objNormal = v3(0.0f);
    t.dist = T_INF; t.prob = 0.0f; t.valid = false;
    return t;
}
// GLSL column major mat3 of a parameter block; transposeOf = transpose(objToWorld)
TDEV mat3 mat3Of(const float m[9]) { return {{m[0], m[1], m[2]}, {m[3], m[4], m[5]}, {m[6], m[7], m[8]}}; }
TDEV mat3 transposeOf(const float m[9]) { return {{m[0], m[3], m[6]}, {m[1], m[4], m[7]}, {m[2], m[5], m[8]}}; }
template <bool FULL = true>
TDEV mat3 targetWorldToObj(const TargetSel& c) {
    if (!FULL) return mat3_diag(c.p.sphere.invRad);
    switch (c.kind) {
    case THEIA_TARGET_FLAT: return transposeOf(c.p.flat.objToWorld);
    case THEIA_TARGET_DISK: return transposeOf(c.p.disk.objToWorld);
    default: return mat3_diag(c.p.sphere.invRad);
    }
}

// sh/util.sample.glsl:31-52 (concentric mapping)
TDEV vec3 sampleUnitDisk(vec2 rng) {
    rng = {2.0f * rng.x - 1.0f, 2.0f * rng.y - 1.0f};
    if (rng.x == 0.0f && rng.y == 0.0f) return v3(0.0f);
    float phi, r;
    if (abs_(rng.x) > abs_(rng.y)) {
        r = rng.x;
        phi = T_PI_OVER_FOUR * (rng.y / rng.x);
    } else {
        r = rng.y;
        phi = T_PI_OVER_TWO - T_PI_OVER_FOUR * (rng.x / rng.y);
    }
    float sp, cp;
    dm_sincos(phi, &sp, &cp);
    return {r * cp, r * sp, 0.0f};
}
TDEV float length2(float x, float y) { return sqrtf(fmaf(y, y, x * x)); }

// Plane shapes share everything but the local sample / the inside test:
// sh/target.flat.glsl:15-44 and sh/target.disk.glsl:15-40 (sample), :46-74 and :42-68 (intersect)
TDEV TargetSample planeSample(vec3 localPos, const float objToWorld[9], vec3 center, vec3 nrm, float prob, vec3 observer) {
    vec3 pos = mat3Of(objToWorld) * localPos + center;
    float side = sign_(dot(nrm, observer - pos));
    vec3 normal = nrm * side;
    TargetSample t;
    t.position = pos; t.normal = normal; t.dist = distance(observer, pos);
    t.objPosition = localPos; t.objNormal = {0.0f, 0.0f, side};
    t.prob = prob; t.valid = dot(normal, normal) != 0.0f;
    return t;
}
struct PlaneHit { vec3 localPos, pos, normal; float t, side; };
TDEV PlaneHit planeIntersect(const float objToWorld[9], vec3 center, vec3 nrm, vec3 observer, vec3 direction) {
    mat3 worldToObj = transposeOf(objToWorld);
    vec3 localObs = worldToObj * (observer - center);
    vec3 localDir = worldToObj * direction;
    PlaneHit h;
    h.t = -localObs.z / localDir.z;
    h.localPos = fma3(h.t, localDir, localObs);
    h.pos = mat3Of(objToWorld) * h.localPos + center;
    h.side = signBit(localObs.z);
    h.normal = nrm * h.side;
    return h;
}

// flat / disk targets, out of line (see commitHitOther in launch.cuh): sh/target.flat.glsl:15-44,
// sh/target.disk.glsl:15-40 (sample), sh/target.flat.glsl:46-74, sh/target.disk.glsl:42-68 (intersect)
static __device__ __noinline__ TargetSample planeTargetSample(const TargetSel& c, vec3 observer, vec2 u) {
    if (c.kind == THEIA_TARGET_FLAT) {
        const theia_target_flat_t& p = c.p.flat;
        vec3 localPos = {p.width * (u.x - 0.5f), p.length * (u.y - 0.5f), 0.0f};
        return planeSample(localPos, p.objToWorld, v3(p.position), v3(p.normal), p.prob, observer);
    }
    const theia_target_disk_t& p = c.p.disk;
    vec3 localPos = p.radius * sampleUnitDisk(u);
    return planeSample(localPos, p.objToWorld, v3(p.position), v3(p.normal), p.prob, observer);
}
static __device__ __noinline__ TargetSample planeTargetIntersect(const TargetSel& c, vec3 observer, vec3 direction) {
    TargetSample r;
    const bool flat = c.kind == THEIA_TARGET_FLAT;
    PlaneHit h = flat ? planeIntersect(c.p.flat.objToWorld, v3(c.p.flat.position), v3(c.p.flat.normal), observer, direction)
                      : planeIntersect(c.p.disk.objToWorld, v3(c.p.disk.position), v3(c.p.disk.normal), observer, direction);
    bool valid = h.t > 0.0f;
    if (flat) valid = valid && 2.0f * abs_(h.localPos.x) <= c.p.flat.width && 2.0f * abs_(h.localPos.y) <= c.p.flat.length;
    else valid = valid && length2(h.localPos.x, h.localPos.y) <= c.p.disk.radius;
    r.position = h.pos; r.normal = h.normal; r.dist = h.t;
    r.objPosition = h.localPos; r.objNormal = {0.0f, 0.0f, h.side};
    r.prob = (flat ? c.p.flat.prob : c.p.disk.prob) * (valid ? 1.0f : 0.0f); r.valid = valid;
    return r;
}
TDEV bool isPlaneTarget(uint32_t kind) { return kind == THEIA_TARGET_FLAT || kind == THEIA_TARGET_DISK; }

// sh/target.sphere.glsl:18-43, sh/target.sphere.inner.glsl:18-36 (every shape draws one random2D)
template <bool FULL = true>
TDEV TargetSample sampleTarget(const TargetSel& c, vec3 observer, Rng& rng) {
    const vec2 u = rng.next2();
    if (FULL && isPlaneTarget(c.kind)) return planeTargetSample(c, observer, u);
    const theia_target_sphere_t& p = c.p.sphere;
    vec3 center = v3(p.position);
    TargetSample t;
    if (c.kind == THEIA_TARGET_SPHERE) {
        vec3 dir = normalize(observer - center);
        float d = distance(observer, center);
        float cosOpening = p.radius / d;
        vec3 normal = createLocalCOSY(dir) * sampleDirectionCone(cosOpening, u);
        vec3 pos = normal * p.radius + center;
        float prob = p.prob / (1.0f - cosOpening);
        bool valid = !isinf_(prob);
        if (!valid) prob = 0.0f;
        t.position = pos; t.normal = normal; t.dist = distance(pos, observer);
        t.objPosition = normal; t.objNormal = normal; t.prob = prob; t.valid = valid;
    } else {
        vec3 normal = sampleUnitSphere(u);
        vec3 pos = p.radius * normal + center;
        t.position = pos; t.normal = -normal; t.dist = distance(observer, pos);
        t.objPosition = normal; t.objNormal = -normal; t.prob = p.prob; t.valid = true;
    }
    return t;
}
// sh/target.sphere.glsl:45-78, sh/target.sphere.inner.glsl:38-56, flat / disk as above
template <bool FULL = true>
TDEV TargetSample intersectTarget(const TargetSel& c, vec3 observer, vec3 direction) {
    if (FULL && isPlaneTarget(c.kind)) return planeTargetIntersect(c, observer, direction);
    TargetSample r;
    const theia_target_sphere_t& p = c.p.sphere;
    vec3 center = v3(p.position);
    vec2 ts = intersectSphere(center, p.radius, observer, direction);
    if (c.kind == THEIA_TARGET_SPHERE) {
        float t = ts.x;
        bool hit = (t > 0.0f) && !isinf_(t);
        if (!hit) return createTargetSampleMiss();
        vec3 pos = fma3(t, direction, observer);
        vec3 nrm = normalize(pos - center);
        float d = distance(observer, center);
        float cosOpening = p.radius / d;
        float prob = p.prob / (1.0f - cosOpening);
        bool valid = !isinf_(prob);
        if (!valid) prob = 0.0f;
        r.position = pos; r.normal = nrm; r.dist = t; r.objPosition = nrm; r.objNormal = nrm;
        r.prob = prob; r.valid = valid;
        return r;
    }
    float t = ts.y;
    bool hit = (t > 0.0f) && !isinf_(t);
    vec3 pos = fma3(t, direction, observer);
    vec3 nrm = normalize(center - pos);
    r.position = pos; r.normal = nrm; r.dist = t; r.objPosition = -nrm; r.objNormal = nrm;
    r.prob = p.prob; r.valid = hit;
    return r;
}
// sh/target.sphere.glsl:80-83, sh/target.sphere.inner.glsl:58-60, sh/target.flat.glsl:76-79,
// sh/target.disk.glsl:70-73 (planes never occlude)
template <bool FULL = true>
TDEV bool isOccludedByTarget(const TargetSel& c, vec3 position) {
    if (FULL && isPlaneTarget(c.kind)) return false;
    float d = distance(position, v3(c.p.sphere.position));
    return c.kind == THEIA_TARGET_SPHERE ? d <= c.p.sphere.radius : d >= c.p.sphere.radius;
}

// ------------------------------------------------------------------ sh/target_guide.*.glsl
struct TargetGuideSample { vec3 dir; float dist, prob; };

// sh/target_guide.common.glsl:11-33 (from a sampled point) and :35-52 (from a direction)
TDEV TargetGuideSample guideSampleFromPoint(vec3 observer, vec3 targetPos, vec3 targetNrm, float targetProb) {
    vec3 dir = targetPos - observer;
    float d2 = dot(dir, dir);
    dir = normalize(dir);
    float cosNormal = dot(dir, targetNrm);
    float prob = targetProb * d2 / abs_(cosNormal);
    if (isinf_(prob)) prob = 0.0f;
    prob *= (cosNormal < 0.0f) ? 1.0f : 0.0f;
    return {dir, distance(targetPos, observer), prob};
}
TDEV TargetGuideSample guideSampleFromDir(vec3 dir, float dist, vec3 targetNrm, float targetProb) {
    float cosNormal = dot(dir, targetNrm);
    float prob = targetProb * dist * dist / abs_(cosNormal);
    if (isinf_(prob)) prob = 0.0f;
    prob *= (cosNormal < 0.0f) ? 1.0f : 0.0f;
    return {dir, dist, prob};
}

TDEV void guideCone(const theia_guide_sphere_t& p, vec3 observer, float* d, vec3* viewDir, float* cosMin,
                    float* prob) {
    vec3 center = v3(p.position);
    *d = distance(center, observer);
    *viewDir = normalize(center - observer);
    float sinMax = p.radius / *d;
    float sin2Max = sinMax * sinMax;
    float cm = 1.0f - sqrtf(max_(1.0f - sin2Max, 0.0f));
    if (sin2Max < 0.00068523f) cm = 0.5f * sin2Max;
    *cosMin = cm;
    float pr = 1.0f / (T_TWO_PI * cm);
    pr *= (*d > p.radius) ? 1.0f : 0.0f;
    *prob = pr;
}
// flat / disk guides, out of line: sh/target_guide.flat.glsl:15-26, sh/target_guide.disk.glsl:14-21 (sample),
// sh/target_guide.flat.glsl:28-47, sh/target_guide.disk.glsl:23-43 (eval)
static __device__ __noinline__ TargetGuideSample planeGuideSample(const GuideSel& c, vec3 observer, vec2 u) {
    if (c.kind == THEIA_GUIDE_FLAT) {
        const theia_target_flat_t& p = c.p.flat;
        vec3 localPos = {p.width * (u.x - 0.5f), p.length * (u.y - 0.5f), 0.0f};
        vec3 pos = mat3Of(p.objToWorld) * localPos + v3(p.position);
        return guideSampleFromPoint(observer, pos, v3(p.normal), p.prob);
    }
    const theia_target_disk_t& p = c.p.disk;
    vec3 localPos = p.radius * sampleUnitDisk(u);
    vec3 pos = mat3Of(p.objToWorld) * localPos + v3(p.position);
    return guideSampleFromPoint(observer, pos, v3(p.normal), p.prob);
}
static __device__ __noinline__ TargetGuideSample planeGuideEval(const GuideSel& c, vec3 observer, vec3 direction) {
    const bool flat = c.kind == THEIA_GUIDE_FLAT;
    PlaneHit h = flat ? planeIntersect(c.p.flat.objToWorld, v3(c.p.flat.position), v3(c.p.flat.normal), observer, direction)
                      : planeIntersect(c.p.disk.objToWorld, v3(c.p.disk.position), v3(c.p.disk.normal), observer, direction);
    bool valid = h.t > 0.0f && !isinf_(h.t) && !(h.t != h.t);
    if (flat) valid = valid && 2.0f * abs_(h.localPos.x) <= c.p.flat.width && 2.0f * abs_(h.localPos.y) <= c.p.flat.length;
    else valid = valid && length2(h.localPos.x, h.localPos.y) <= c.p.disk.radius;
    float prob = (flat ? c.p.flat.prob : c.p.disk.prob) * (valid ? 1.0f : 0.0f);
    return guideSampleFromDir(direction, h.t, flat ? v3(c.p.flat.normal) : v3(c.p.disk.normal), prob);
}
TDEV bool isPlaneGuide(uint32_t kind) { return kind == THEIA_GUIDE_FLAT || kind == THEIA_GUIDE_DISK; }

// sh/target_guide.sphere.glsl:15-51 (every shape draws two numbers: random2D in the plane shapes, two
// calls of random in the sphere, which address the same two counters)
template <bool FULL = true>
TDEV TargetGuideSample sampleTargetGuide(const GuideSel& c, vec3 observer, Rng& rng) {
    const vec2 u = rng.next2();
    if (FULL && isPlaneGuide(c.kind)) return planeGuideSample(c, observer, u);
    float d, cosMin, prob;
    vec3 viewDir;
    guideCone(c.p.sphere, observer, &d, &viewDir, &cosMin, &prob);
    float cosTheta = 1.0f - cosMin * u.x;
    float sinTheta = sqrtf(max_(1.0f - cosTheta * cosTheta, 0.0f));
    float phi = T_TWO_PI * u.y;
    float sp, cp;
    dm_sincos(phi, &sp, &cp);
    vec3 dir = createLocalCOSY(viewDir) * vec3{sinTheta * sp, sinTheta * cp, cosTheta};
    return {dir, d + c.p.sphere.radius, prob};
}
// sh/target_guide.sphere.glsl:53-81
template <bool FULL = true>
TDEV TargetGuideSample evalTargetGuide(const GuideSel& c, vec3 observer, vec3 direction) {
    if (FULL && isPlaneGuide(c.kind)) return planeGuideEval(c, observer, direction);
    float d, cosMin, prob;
    vec3 viewDir;
    guideCone(c.p.sphere, observer, &d, &viewDir, &cosMin, &prob);
    float cosDir = dot(viewDir, direction);
    prob *= (cosMin >= 1.0f - cosDir) ? 1.0f : 0.0f;
    return {direction, d + c.p.sphere.radius, prob};
}

}  // namespace theia
