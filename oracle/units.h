// ORACLE - TEST INFRASTRUCTURE ONLY. Never included, linked or called by the product path.
//
// Unit seam of the oracle: one function of the path for n inputs, the CPU mirror of
// theia_unit_eval (include/theia_b200.h). Each unit restates the body of one of the reference's
// test programs (tests/shader/*.test.glsl) or sampler stages around the oracle's own functions,
// so that the numpy pins of the reference's unit tests (tests/test_scatter.py, test_polarization.py,
// test_ray.py, test_target.py, test_target_guide.py, test_light.py, test_math.py, test_random.py)
// can be ported against them (tests/test_reference_units.py).
#pragma once
#include "tracer.h"

namespace orc {

static inline uint64_t mediumOf(const float* lohi) {
    uint32_t lo, hi;
    std::memcpy(&lo, lohi, 4);
    std::memcpy(&hi, lohi + 1, 4);
    return (uint64_t)lo | ((uint64_t)hi << 32);
}
// sh/polarization.glsl:7-17
static inline mat4 rotatePolRefPhi(float phi) { return rotMatrix(dm_cos(2.0f * phi), dm_sin(2.0f * phi)); }
static inline mat4 matmul(const mat4& a, const mat4& b) {  // column major: (a b)[c][r] = sum_k a[k][r] b[c][k]
    mat4 o{};
    for (int c = 0; c < 4; ++c)
        for (int r = 0; r < 4; ++r) {
            float acc = a.m[0][r] * b.m[c][0];
            for (int k = 1; k < 4; ++k) acc = fmaf(a.m[k][r], b.m[c][k], acc);
            o.m[c][r] = acc;
        }
    return o;
}
static inline void put3(float* o, vec3 v) { o[0] = v.x; o[1] = v.y; o[2] = v.z; }
static inline void put4(float* o, vec4 v) { o[0] = v.x; o[1] = v.y; o[2] = v.z; o[3] = v.w; }
static inline void putU(float* o, uint32_t v) { std::memcpy(o, &v, 4); }

static inline int unitWidths(uint32_t unit, uint32_t polarized, uint32_t* in, uint32_t* out) {
    (void)polarized;
    switch (unit) {
    case THEIA_UNIT_SCATTER_DIR: *in = 5; *out = 3; return 0;
    case THEIA_UNIT_REFLECTANCE: *in = 7; *out = 3; return 0;
    case THEIA_UNIT_VOLUME_SCATTER: *in = 5; *out = 4; return 0;
    case THEIA_UNIT_SCATTER_PROB: *in = 8; *out = 1; return 0;
    case THEIA_UNIT_ROTATE_PHI: *in = 5; *out = 4; return 0;
    case THEIA_UNIT_ALIGN_POLREF: *in = 0; *out = 16; return 0;
    case THEIA_UNIT_ROTATE_DIR: *in = 12; *out = 7; return 0;
    case THEIA_UNIT_PHASE_MATRIX: *in = 5; *out = 4; return 0;
    case THEIA_UNIT_SURFACE: *in = 0; *out = 33; return 0;
    case THEIA_UNIT_SIGN_BIT: *in = 1; *out = 1; return 0;
    case THEIA_UNIT_GAMMA: *in = 0; *out = 1; return 0;
    case THEIA_UNIT_EMISSION_ANGLE: *in = 1; *out = 1; return 0;
    case THEIA_UNIT_TARGET_SAMPLE: *in = 0; *out = 29; return 0;
    case THEIA_UNIT_GUIDE_SAMPLE: *in = 0; *out = 13; return 0;
    case THEIA_UNIT_LIGHT_SAMPLE: *in = 0; *out = 0; return 0;
    case THEIA_UNIT_MEDIUM: *in = 3; *out = 12; return 0;
    }
    return -1;
}

// tests/shader/ray.surface.test.glsl (FORWARD): a ray from the light source meets a surface with
// normal `normal` at its own position; the same ray is reflected and transmitted
static inline void unitSurface(const theia_unit_args_t& A, Rng& rng, float* o) {
    const bool pol = A.polarized != 0;
    const vec3 normal = {A.params[0], A.params[1], A.params[2]};
    const float lamMin = A.params[3], lamMax = A.params[4];
    const MaterialGLSL* mat = (const MaterialGLSL*)(uintptr_t)A.material;
    float u = rng.next();
    float lam = lamMin + (lamMax - lamMin) * u;
    MediumConstants consts = lookUpMedium(0, lam);
    SourceRay source = sampleLight(A.light, lam, consts, rng, pol, A.first);
    bool inward = dot(source.direction, normal) < 0.0f;
    uint64_t med = inward ? mat->outside : mat->inside;
    consts = lookUpMedium(med, lam);
    Ray ray = createRay(source, med, consts, WavelengthSample{lam, 1.0f});
    alignRayToHit(ray, normal, pol);
    inward = dot(ray.direction, normal) <= 0.0f;
    SurfaceHit hit{};
    hit.valid = true; hit.material = mat; hit.inward = inward; hit.customId = 0; hit.flags = 0;
    hit.worldPos = ray.position; hit.rayNrm = inward ? normal : -normal;
    hit.objPos = ray.position; hit.objNrm = normal; hit.objDir = ray.direction;
    hit.worldToObj = mat3_diag(1.0f);
    Reflectance surface = fresnelReflect(ray, hit);
    Ray refl = ray, trans = ray;
    reflectRay(refl, hit, surface, pol);
    transmitRay(trans, hit, surface, pol);
    o[0] = ray.wavelength; put3(o + 1, ray.direction); o[4] = ray.lin_contrib * dm_exp(ray.log_contrib);
    put3(o + 5, refl.position); put3(o + 8, refl.direction); o[11] = refl.lin_contrib * dm_exp(refl.log_contrib);
    // (the reference multiplies the transmitted contribution with the reflected ray's exp(log_contrib): same value)
    put3(o + 12, trans.position); put3(o + 15, trans.direction); o[18] = trans.lin_contrib * dm_exp(refl.log_contrib);
    put3(o + 19, refl.polRef); put3(o + 22, trans.polRef);
    put4(o + 25, refl.stokes); put4(o + 29, trans.stokes);
}

// sh/target.sample.glsl:16-48 (checkSample / checkIntersect)
static inline void packChecked(vec3 observer, const TargetSample& ts, float* o, const vec3* direction) {
    float expD = distance(ts.position, observer);
    vec3 expPos = ts.worldToObj * ts.position + ts.offset;
    vec3 expNrm = normalize(ts.worldToObj * ts.normal);
    int32_t error = 0;
    error += (int32_t)(abs_(expD - ts.dist) > 1e-5f);
    error += (int32_t)(length(expPos - ts.objPosition) > 1e-5f) << 1;
    error += (int32_t)(length(expNrm - ts.objNormal) > 1e-5f) << 2;
    if (direction) {
        vec3 sampleDir = normalize(ts.position - observer);
        error += (int32_t)(length(sampleDir - *direction) > 1e-5f) << 3;
    }
    put3(o, ts.position); put3(o + 3, ts.normal); o[6] = ts.prob; putU(o + 7, ts.valid ? 1u : 0u); putU(o + 8, (uint32_t)error);
}
static inline vec3 sampleObserver(const float* p, Rng& rng) {
    float x = mix_(p[0], p[3], rng.next());
    float y = mix_(p[1], p[4], rng.next());
    float z = mix_(p[2], p[5], rng.next());
    return {x, y, z};
}

// SAVE_SAMPLE / SAVE_PHOTON (sh/lightsource.queue.glsl, sh/wavelengthsource.queue.glsl): count word, then
// one array of `cap` entries per component
static inline void saveLightSample(uint32_t* queue, uint32_t cap, bool pol, uint32_t k, const SourceRay& r) {
    float* f = (float*)(queue + 1);
    int c = 0;
    auto put = [&](float v) { f[(size_t)c * cap + k] = v; ++c; };
    put(r.position.x); put(r.position.y); put(r.position.z);
    put(r.direction.x); put(r.direction.y); put(r.direction.z);
    if (pol) {
        put(r.stokes.x); put(r.stokes.y); put(r.stokes.z); put(r.stokes.w);
        put(r.polRef.x); put(r.polRef.y); put(r.polRef.z);
    }
    put(r.startTime); put(r.contrib);
}

static inline int unitEval(const theia_unit_args_t& A) {
    uint32_t wi = 0, wo = 0;
    if (A.struct_size != sizeof(theia_unit_args_t) || unitWidths(A.unit, A.polarized, &wi, &wo)) return -1;
    if (A.unit == THEIA_UNIT_LIGHT_SAMPLE) {
        if (!A.light_queue || !A.wavelength_queue || A.n > A.capacity) return -2;
        const bool pol = A.polarized != 0;
#pragma omp parallel for schedule(static)
        for (int64_t i = 0; i < (int64_t)A.n; ++i) {  // sh/lightsource.sample.glsl:48-73
            Rng rng{&A.rng, A.first + (uint32_t)i, A.base_count};
            WavelengthSample photon = sampleWavelength(A.wavelength, rng, A.first);
            SourceRay ray = sampleLight(A.light, photon.wavelength, lookUpMedium(A.medium, photon.wavelength), rng, pol, A.first);
            saveLightSample(A.light_queue, A.capacity, pol, (uint32_t)i, ray);
            float* w = (float*)(A.wavelength_queue + 1);
            w[i] = photon.wavelength;
            w[(size_t)A.capacity + i] = photon.contrib;
        }
        A.light_queue[0] = A.n;
        A.wavelength_queue[0] = A.n;
        return 0;
    }
    if (!A.out || (wi && !A.in)) return -2;
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < (int64_t)A.n; ++i) {
        const float* q = A.in ? A.in + (size_t)wi * i : nullptr;
        float* o = A.out + (size_t)wo * i;
        Rng rng{&A.rng, A.first + (uint32_t)i, A.base_count};
        switch (A.unit) {
        case THEIA_UNIT_SCATTER_DIR: put3(o, scatterDir(v3(q), q[3], q[4])); break;
        case THEIA_UNIT_REFLECTANCE: {  // tests/shader/scatter.reflectance.test.glsl
            const MaterialGLSL* mat = (const MaterialGLSL*)(uintptr_t)A.material;
            vec3 dir = v3(q), nrm = v3(q + 3);
            float lam = q[6];
            bool inward = dot(dir, nrm) <= 0.0f;
            uint64_t med = inward ? mat->outside : mat->inside;
            Ray ray{};
            ray.position = v3(0.0f); ray.direction = dir; ray.wavelength = lam; ray.time = 0.0f;
            ray.lin_contrib = 1.0f; ray.log_contrib = 0.0f; ray.constants = lookUpMedium(med, lam);
            SurfaceHit hit{};
            hit.valid = true; hit.material = mat; hit.inward = inward;
            hit.rayNrm = inward ? nrm : -nrm; hit.objNrm = nrm; hit.objDir = dir; hit.worldToObj = mat3_diag(1.0f);
            Reflectance s = fresnelReflect(ray, hit);
            o[0] = s.r_s; o[1] = s.r_p; o[2] = s.n_tr;
            break;
        }
        case THEIA_UNIT_VOLUME_SCATTER: {  // tests/shader/scatter.volume.scatter.test.glsl
            float prob;
            vec3 d = scatter(mediumOf(q + 3), v3(q), rng.next2(), &prob);
            put3(o, d); o[3] = prob;
            break;
        }
        case THEIA_UNIT_SCATTER_PROB: o[0] = scatterProb(mediumOf(q + 6), v3(q), v3(q + 3)); break;
        case THEIA_UNIT_ROTATE_PHI: put4(o, rotatePolRefPhi(q[4]) * vec4{q[0], q[1], q[2], q[3]}); break;
        case THEIA_UNIT_ALIGN_POLREF: {  // tests/shader/polarization.alignPolRef.test.glsl
            vec3 dirIn = sampleUnitSphere(rng.next2());
            vec3 dirOut = sampleUnitSphere(rng.next2());
            vec3 polRef = createLocalCOSY(dirIn).c0, newPolRef;
            mat4 rot = rotatePolRef(dirIn, polRef, dirOut, &newPolRef);
            mat4 unrot = alignPolRef(dirIn, newPolRef, polRef);
            mat4 r = matmul(unrot, rot);
            for (int c = 0; c < 4; ++c) for (int k = 0; k < 4; ++k) o[4 * c + k] = r.m[c][k];
            break;
        }
        case THEIA_UNIT_ROTATE_DIR: {  // tests/shader/polarization.rotate.dir.test.glsl
            vec3 dirIn = v3(q + 4), ref = v3(q + 7), nref;
            vec3 dir = scatterDir(dirIn, dm_cos(q[11]), q[10]);
            vec4 st = rotatePolRef(dirIn, ref, dir, &nref) * vec4{q[0], q[1], q[2], q[3]};
            put4(o, st); put3(o + 4, nref);
            break;
        }
        case THEIA_UNIT_PHASE_MATRIX: put4(o, lookUpPhaseMatrix(A.medium, q[4]) * vec4{q[0], q[1], q[2], q[3]}); break;
        case THEIA_UNIT_SURFACE: unitSurface(A, rng, o); break;
        case THEIA_UNIT_SIGN_BIT: o[0] = signBit(q[0]); break;
        case THEIA_UNIT_GAMMA: o[0] = sampleGamma(A.params[0], rng) / A.params[1]; break;
        case THEIA_UNIT_EMISSION_ANGLE: o[0] = particle_sampleEmissionAngle(A.params[0], A.params[1], A.params[2], q[0]); break;
        case THEIA_UNIT_MEDIUM: {  // tests/shader/medium.test.glsl, tests/shader/material.test.glsl
            uint64_t med = A.medium;
            o[10] = o[11] = 0.0f;
            if (A.material != 0) {
                const MaterialGLSL* mat = reinterpret_cast<const MaterialGLSL*>(A.material);
                med = A.params[0] == 0.0f ? mat->inside : mat->outside;
                o[10] = u2f(mat->flagsInwards); o[11] = u2f(mat->flagsOutwards);
            }
            const MediumConstants c = lookUpMedium(med, q[0]);
            const MediumGLSL* m = med ? asMedium(med) : nullptr;
            const float t = 0.5f * (q[1] + 1.0f);
            o[0] = c.n; o[1] = c.vg; o[2] = c.mu_s; o[3] = c.mu_e;
            o[4] = lookUp(m ? m->log_phase : 0, t, 0.0f);
            o[5] = lookUp(m ? m->phase_sampling : 0, q[2], 0.0f);
            o[6] = lookUp(m ? m->phase_m12 : 0, t, 0.0f);
            o[7] = lookUp(m ? m->phase_m22 : 0, t, 0.0f);
            o[8] = lookUp(m ? m->phase_m33 : 0, t, 0.0f);
            o[9] = lookUp(m ? m->phase_m34 : 0, t, 0.0f);
            break;
        }
        case THEIA_UNIT_TARGET_SAMPLE: {  // sh/target.sample.glsl:62-86
            vec3 observer = sampleObserver(A.params, rng);
            vec3 direction = sampleUnitSphere(rng.next2());
            put3(o, observer); put3(o + 3, direction);
            packChecked(observer, sampleTarget(A.target, observer, rng), o + 6, nullptr);
            packChecked(observer, intersectTarget(A.target, observer, direction), o + 15, &direction);
            putU(o + 24, isOccludedByTarget(A.target, observer) ? 1u : 0u);
            for (int k = 25; k < 29; ++k) o[k] = 0.0f;
            break;
        }
        case THEIA_UNIT_GUIDE_SAMPLE: {  // sh/target_guide.sample.glsl:24-44
            vec3 observer = sampleObserver(A.params, rng);
            vec3 direction = sampleUnitSphere(rng.next2());
            TargetGuideSample s = sampleTargetGuide(A.guide, observer, rng);
            TargetGuideSample e = evalTargetGuide(A.guide, observer, direction);
            put3(o, observer);
            put3(o + 3, s.dir); o[6] = s.dist; o[7] = s.prob;
            put3(o + 8, e.dir); o[11] = e.dist; o[12] = e.prob;
            break;
        }
        }
    }
    return 0;
}

}  // namespace orc
