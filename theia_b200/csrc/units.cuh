// Component / shader unit seam (theia_unit_eval, include/theia_b200.h): one function of the path
// evaluated for n independent inputs by one thread each. The units are the bodies of the
// reference's test programs (tests/shader/<unit>.test.glsl) and sampler stages
// (sh/lightsource.sample.glsl, sh/target.sample.glsl, sh/target_guide.sample.glsl) around the
// functions the tracing kernels use, so the reference's unit tests can be run against the device
// code itself (tests/test_reference_units.py).
#pragma once
#include "surface.cuh"
#include "launch.cuh"

namespace theia {

struct UnitDev {
    uint32_t unit, n, polarized, first, base_count, in_width, out_width, capacity;
    const float* in;
    float* out;
    theia_philox_params_t rng;
    uint64_t medium, material;
    float params[8];
    LightSel light;
    WavelengthSel wavelength;
    TargetSel target;
    GuideSel guide;
    uint32_t* light_queue;
    uint32_t* wavelength_queue;
};

__host__ __device__ inline int unitWidths(uint32_t unit, uint32_t* in, uint32_t* out) {
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

TDEV uint64_t unitMediumOf(const float* lohi) { return (uint64_t)f2u(lohi[0]) | ((uint64_t)f2u(lohi[1]) << 32); }
// sh/polarization.glsl:7-17
TDEV mat4 rotatePolRefPhi(float phi) {
    float s, c;
    dm_sincos(2.0f * phi, &s, &c);
    return rotMatrix(c, s);
}
TDEV mat4 unitMatmul(const mat4& a, const mat4& b) {  // column major: (a b)[c][r] = sum_k a[k][r] b[c][k]
    mat4 o = mat4_zero();
    for (int c = 0; c < 4; ++c)
        for (int r = 0; r < 4; ++r) {
            float acc = a.m[0][r] * b.m[c][0];
            for (int k = 1; k < 4; ++k) acc = fmaf(a.m[k][r], b.m[c][k], acc);
            o.m[c][r] = acc;
        }
    return o;
}
TDEV void put3(float* o, vec3 v) { o[0] = v.x; o[1] = v.y; o[2] = v.z; }
TDEV void put4(float* o, vec4 v) { o[0] = v.x; o[1] = v.y; o[2] = v.z; o[3] = v.w; }

// TargetSample.offset of the reference (sh/target.common.glsl): objPosition = worldToObj * position + offset
TDEV vec3 targetOffset(const TargetSel& c) {
    switch (c.kind) {
    case THEIA_TARGET_FLAT: return -(transposeOf(c.p.flat.objToWorld) * v3(c.p.flat.position));
    case THEIA_TARGET_DISK: return -(transposeOf(c.p.disk.objToWorld) * v3(c.p.disk.position));
    default: return v3(c.p.sphere.invPos);
    }
}

// tests/shader/ray.surface.test.glsl (FORWARD): a ray from the light source meets a surface with
// normal `normal` at its own position; the same ray is reflected and transmitted
template <bool POL>
TDEV void unitSurface(const UnitDev& A, Rng& rng, float* o) {
    const vec3 normal = {A.params[0], A.params[1], A.params[2]};
    const float lamMin = A.params[3], lamMax = A.params[4];
    float u = rng.next();
    float lam = lamMin + (lamMax - lamMin) * u;
    MediumConstants consts = lookUpMedium(0, lam);
    SourceRay source = sampleLight<POL, true>(A.light, lam, consts, rng);
    bool inward = dot(source.direction, normal) < 0.0f;
    uint64_t med = ldp(A.material, inward ? MAT_OUTSIDE : MAT_INSIDE);
    consts = lookUpMedium(med, lam);
    Ray ray = createRay(source, med, consts, WavelengthSample{lam, 1.0f});
    alignRayToHit<POL>(ray, normal);
    inward = dot(ray.direction, normal) <= 0.0f;
    SurfaceHit hit;
    hit.valid = true; hit.material = A.material; hit.inward = inward; hit.customId = 0; hit.flags = 0;
    hit.worldPos = ray.position; hit.rayNrm = inward ? normal : -normal;
    hit.objPos = ray.position; hit.objNrm = normal;
    hit.worldToObj = mat3_diag(1.0f);
    Reflectance surface = fresnelReflect(ray, hit);
    Ray refl = ray, trans = ray;
    reflectRay<POL>(refl, hit, surface);
    transmitRay<POL>(trans, hit, surface);
    o[0] = ray.wavelength; put3(o + 1, ray.direction); o[4] = ray.lin_contrib * dm_exp(ray.log_contrib);
    put3(o + 5, refl.position); put3(o + 8, refl.direction); o[11] = refl.lin_contrib * dm_exp(refl.log_contrib);
    put3(o + 12, trans.position); put3(o + 15, trans.direction); o[18] = trans.lin_contrib * dm_exp(refl.log_contrib);
    put3(o + 19, refl.polRef); put3(o + 22, trans.polRef);
    put4(o + 25, refl.stokes); put4(o + 29, trans.stokes);
}

// sh/target.sample.glsl:16-48 (checkSample / checkIntersect)
TDEV void packChecked(const TargetSel& c, vec3 observer, const TargetSample& ts, float* o, const vec3* direction) {
    const mat3 w2o = targetWorldToObj<true>(c);
    float expD = distance(ts.position, observer);
    vec3 expPos = w2o * ts.position + targetOffset(c);
    vec3 expNrm = normalize(w2o * ts.normal);
    int32_t error = 0;
    error += (int32_t)(abs_(expD - ts.dist) > 1e-5f);
    error += (int32_t)(length(expPos - ts.objPosition) > 1e-5f) << 1;
    error += (int32_t)(length(expNrm - ts.objNormal) > 1e-5f) << 2;
    if (direction) {
        vec3 sampleDir = normalize(ts.position - observer);
        error += (int32_t)(length(sampleDir - *direction) > 1e-5f) << 3;
    }
    put3(o, ts.position); put3(o + 3, ts.normal); o[6] = ts.prob; o[7] = u2f(ts.valid ? 1u : 0u); o[8] = u2f((uint32_t)error);
}
TDEV vec3 sampleObserver(const float* p, Rng& rng) {
    float x = mix_(p[0], p[3], rng.next());
    float y = mix_(p[1], p[4], rng.next());
    float z = mix_(p[2], p[5], rng.next());
    return {x, y, z};
}

template <bool POL>
__global__ void __launch_bounds__(128) unitEvalKernel(const __grid_constant__ UnitDev A) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= A.n) return;
    Rng rng = makeRng(&A.rng, A.first + i);
    rng.dim = A.base_count;
    if (A.unit == THEIA_UNIT_LIGHT_SAMPLE) {  // sh/lightsource.sample.glsl:48-73
        WavelengthSample photon = sampleWavelength<true>(A.wavelength, rng);
        SourceRay ray = sampleLight<POL, true>(A.light, photon.wavelength, lookUpMedium(A.medium, photon.wavelength), rng);
        float* f = reinterpret_cast<float*>(A.light_queue + 1) + i;
        const size_t C = A.capacity;
        f[0] = ray.position.x; f[C] = ray.position.y; f[2 * C] = ray.position.z;
        f[3 * C] = ray.direction.x; f[4 * C] = ray.direction.y; f[5 * C] = ray.direction.z;
        f += 6 * C;
        if (POL) {
            f[0] = ray.stokes.x; f[C] = ray.stokes.y; f[2 * C] = ray.stokes.z; f[3 * C] = ray.stokes.w;
            f[4 * C] = ray.polRef.x; f[5 * C] = ray.polRef.y; f[6 * C] = ray.polRef.z;
            f += 7 * C;
        }
        f[0] = ray.startTime; f[C] = ray.contrib;
        float* w = reinterpret_cast<float*>(A.wavelength_queue + 1) + i;
        w[0] = photon.wavelength; w[C] = photon.contrib;
        if (i == 0) { A.light_queue[0] = A.n; A.wavelength_queue[0] = A.n; }
        return;
    }
    const float* q = A.in ? A.in + (size_t)A.in_width * i : nullptr;
    float* o = A.out + (size_t)A.out_width * i;
    switch (A.unit) {
    case THEIA_UNIT_SCATTER_DIR: put3(o, scatterDir(v3(q), q[3], q[4])); break;
    case THEIA_UNIT_REFLECTANCE: {  // tests/shader/scatter.reflectance.test.glsl
        vec3 dir = v3(q), nrm = v3(q + 3);
        float lam = q[6];
        bool inward = dot(dir, nrm) <= 0.0f;
        uint64_t med = ldp(A.material, inward ? MAT_OUTSIDE : MAT_INSIDE);
        Ray ray;
        ray.position = v3(0.0f); ray.direction = dir; ray.medium = med; ray.wavelength = lam; ray.time = 0.0f;
        ray.lin_contrib = 1.0f; ray.log_contrib = 0.0f; ray.constants = lookUpMedium(med, lam);
        ray.stokes = {1.0f, 0.0f, 0.0f, 0.0f}; ray.polRef = v3(0.0f);
        SurfaceHit hit;
        hit.valid = true; hit.material = A.material; hit.inward = inward; hit.customId = 0; hit.flags = 0;
        hit.worldPos = v3(0.0f); hit.rayNrm = inward ? nrm : -nrm; hit.objPos = v3(0.0f); hit.objNrm = nrm;
        hit.worldToObj = mat3_diag(1.0f);
        Reflectance s = fresnelReflect(ray, hit);
        o[0] = s.r_s; o[1] = s.r_p; o[2] = s.n_tr;
        break;
    }
    case THEIA_UNIT_VOLUME_SCATTER: {  // tests/shader/scatter.volume.scatter.test.glsl
        float prob;
        vec3 d = scatter(unitMediumOf(q + 3), v3(q), rng.next2(), &prob);
        put3(o, d); o[3] = prob;
        break;
    }
    case THEIA_UNIT_SCATTER_PROB: o[0] = scatterProb(unitMediumOf(q + 6), v3(q), v3(q + 3)); break;
    case THEIA_UNIT_ROTATE_PHI: put4(o, rotatePolRefPhi(q[4]) * vec4{q[0], q[1], q[2], q[3]}); break;
    case THEIA_UNIT_ALIGN_POLREF: {  // tests/shader/polarization.alignPolRef.test.glsl
        vec3 dirIn = sampleUnitSphere(rng.next2());
        vec3 dirOut = sampleUnitSphere(rng.next2());
        vec3 polRef = createLocalCOSY(dirIn).c0, newPolRef;
        mat4 rot = rotatePolRef(dirIn, polRef, dirOut, &newPolRef);
        mat4 unrot = alignPolRef(dirIn, newPolRef, polRef);
        mat4 r = unitMatmul(unrot, rot);
        for (int c = 0; c < 4; ++c) for (int k = 0; k < 4; ++k) o[4 * c + k] = r.m[c][k];
        break;
    }
    case THEIA_UNIT_ROTATE_DIR: {  // tests/shader/polarization.rotate.dir.test.glsl
        vec3 dirIn = v3(q + 4), ref = v3(q + 7), nref;
        float st, ct;
        dm_sincos(q[11], &st, &ct);
        vec3 dir = scatterDir(dirIn, ct, q[10]);
        vec4 sv = rotatePolRef(dirIn, ref, dir, &nref) * vec4{q[0], q[1], q[2], q[3]};
        put4(o, sv); put3(o + 4, nref);
        break;
    }
    case THEIA_UNIT_PHASE_MATRIX: put4(o, lookUpPhaseMatrix(A.medium, q[4]) * vec4{q[0], q[1], q[2], q[3]}); break;
    case THEIA_UNIT_SURFACE: unitSurface<POL>(A, rng, o); break;
    case THEIA_UNIT_SIGN_BIT: o[0] = signBit(q[0]); break;
    case THEIA_UNIT_GAMMA: o[0] = sampleGamma(A.params[0], rng) / A.params[1]; break;
    case THEIA_UNIT_EMISSION_ANGLE: o[0] = particle_sampleEmissionAngle(A.params[0], A.params[1], A.params[2], q[0]); break;
    case THEIA_UNIT_MEDIUM: {  // tests/shader/medium.test.glsl, tests/shader/material.test.glsl
        // the medium itself, or with a material its inside (params[0] == 0) / outside medium and the flag words
        uint64_t med = A.medium;
        o[10] = o[11] = 0.0f;
        if (A.material != 0) {
            med = ldp(A.material, A.params[0] == 0.0f ? MAT_INSIDE : MAT_OUTSIDE);
            o[10] = u2f(ldu(A.material, MAT_FLAGS_IN)); o[11] = u2f(ldu(A.material, MAT_FLAGS_OUT));
        }
        const MediumConstants c = lookUpMedium(med, q[0]);
        const float t = 0.5f * (q[1] + 1.0f);
        o[0] = c.n; o[1] = c.vg; o[2] = c.mu_s; o[3] = c.mu_e;
        o[4] = lookUp(med ? ldp(med, MED_LOG_PHASE) : 0, t, 0.0f);
        o[5] = lookUp(med ? ldp(med, MED_PHASE_SAMPLING) : 0, q[2], 0.0f);
        o[6] = lookUp(med ? ldp(med, MED_M12) : 0, t, 0.0f);
        o[7] = lookUp(med ? ldp(med, MED_M22) : 0, t, 0.0f);
        o[8] = lookUp(med ? ldp(med, MED_M33) : 0, t, 0.0f);
        o[9] = lookUp(med ? ldp(med, MED_M34) : 0, t, 0.0f);
        break;
    }
    case THEIA_UNIT_TARGET_SAMPLE: {  // sh/target.sample.glsl:62-86
        vec3 observer = sampleObserver(A.params, rng);
        vec3 direction = sampleUnitSphere(rng.next2());
        put3(o, observer); put3(o + 3, direction);
        packChecked(A.target, observer, sampleTarget<true>(A.target, observer, rng), o + 6, nullptr);
        packChecked(A.target, observer, intersectTarget<true>(A.target, observer, direction), o + 15, &direction);
        o[24] = u2f(isOccludedByTarget<true>(A.target, observer) ? 1u : 0u);
        for (int k = 25; k < 29; ++k) o[k] = 0.0f;
        break;
    }
    case THEIA_UNIT_GUIDE_SAMPLE: {  // sh/target_guide.sample.glsl:24-44
        vec3 observer = sampleObserver(A.params, rng);
        vec3 direction = sampleUnitSphere(rng.next2());
        TargetGuideSample s = sampleTargetGuide<true>(A.guide, observer, rng);
        TargetGuideSample e = evalTargetGuide<true>(A.guide, observer, direction);
        put3(o, observer);
        put3(o + 3, s.dir); o[6] = s.dist; o[7] = s.prob;
        put3(o + 8, e.dir); o[11] = e.dist; o[12] = e.prob;
        break;
    }
    default: break;
    }
}

}  // namespace theia
