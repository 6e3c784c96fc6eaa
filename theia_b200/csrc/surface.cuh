// Surface interaction of the scene tracer: ray query result -> SurfaceHit, Fresnel
// coefficients, self-intersection offsets, border crossing, reflection and transmission
// (sh/scene.intersect.glsl, sh/scatter.surface.glsl, sh/ray.surface.glsl of the reference).
#pragma once
#include "physics.cuh"
#include "traverse.cuh"

namespace theia {

// sh/scene.types.glsl:28-57
struct SurfaceHit {
    bool valid;
    uint64_t material;  // address of the Material record
    bool inward;
    int32_t customId;
    uint32_t flags;
    vec3 worldPos, rayNrm;
    vec3 objPos, objNrm;
    mat3 worldToObj;
};

enum : uint32_t { MAT_INSIDE = 0, MAT_OUTSIDE = 8, MAT_FLAGS_IN = 16, MAT_FLAGS_OUT = 20 };

TDEV vec3 ldVertex(uint64_t vertices, int32_t index, uint32_t offset) {
    const float* p = reinterpret_cast<const float*>(vertices) + 6 * (size_t)index + offset;
    return {__ldg(p), __ldg(p + 1), __ldg(p + 2)};
}

TDEV void ldVertex6(uint64_t vertices, int32_t index, vec3& pos, vec3& nrm) {
    const float2* p = reinterpret_cast<const float2*>(reinterpret_cast<const float*>(vertices) + 6 * (size_t)index);
    const float2 a = __ldg(p), b = __ldg(p + 1), c = __ldg(p + 2);
    pos = {a.x, a.y, b.x};
    nrm = {b.y, c.x, c.y};
}

// sh/scene.intersect.glsl:23-99
TDEV int processRayQuery(const SceneView& sc, const Ray& ray, const RayHit& q, SurfaceHit& hit) {
    hit.valid = q.valid;
    hit.worldPos = v3(0.0f);
    if (!hit.valid) return THEIA_RESULT_RAY_MISSED;
    const InstanceRec* in = sc.instances + q.slot;
    hit.customId = __ldg(&in->customIndex);
    const uint64_t vertices = __ldg(&in->vertices);
    const int32_t* index = reinterpret_cast<const int32_t*>(__ldg(&in->indices)) + 3 * (size_t)q.prim;
    const int32_t i0 = __ldg(index), i1 = __ldg(index + 1), i2 = __ldg(index + 2);
    // a vertex = {position, normal} = 24 bytes, 8 byte aligned: three 64 bit loads each
    vec3 v0, v1, v2, n0, nn1, nn2;
    ldVertex6(vertices, i0, v0, n0);
    ldVertex6(vertices, i1, v1, nn1);
    ldVertex6(vertices, i2, v2, nn2);
    vec3 e1 = v1 - v0, e2 = v2 - v0;
    float bx = q.u, by = q.v;
    // precise: v0 + fma(vec3(b.x), e1, b.y * e2)
    hit.objPos = {v0.x + fmaf(bx, e1.x, by * e2.x), v0.y + fmaf(bx, e1.y, by * e2.y), v0.z + fmaf(bx, e1.z, by * e2.z)};
    vec3 objNrm = cross(e1, e2);
    vec3 d1 = nn1 - n0, d2 = nn2 - n0;
    vec3 intNrm = {n0.x + fmaf(bx, d1.x, by * d2.x), n0.y + fmaf(bx, d1.y, by * d2.y), n0.z + fmaf(bx, d1.z, by * d2.z)};
    objNrm = objNrm * signBit(dot(objNrm, intNrm));
    hit.objNrm = normalize(objNrm);
    // world -> object (columns of the 3x4 row major matrix)
    const float4* wq = reinterpret_cast<const float4*>(in->w2o);
    float4 w0 = __ldg(wq), w1 = __ldg(wq + 1), w2 = __ldg(wq + 2);
    hit.worldToObj = mat3{{w0.x, w1.x, w2.x}, {w0.y, w1.y, w2.y}, {w0.z, w1.z, w2.z}};
    vec3 objDir = normalize(hit.worldToObj * ray.direction);
    hit.inward = dot(objDir, hit.objNrm) <= 0.0f;
    hit.material = __ldg(&in->material);
    hit.flags = ldu(hit.material, hit.inward ? MAT_FLAGS_IN : MAT_FLAGS_OUT);
    uint64_t medium = ldp(hit.material, hit.inward ? MAT_OUTSIDE : MAT_INSIDE);
    if (ray.medium != medium) return THEIA_ERROR_MEDIA_MISMATCH;
    vec3 worldNrm = normalize(mulT(hit.objNrm, hit.worldToObj));
    hit.rayNrm = worldNrm * (2.0f * (hit.inward ? 1.0f : 0.0f) - 1.0f);
    const float4* mq = reinterpret_cast<const float4*>(in->o2w);
    float4 m0 = __ldg(mq), m1 = __ldg(mq + 1), m2 = __ldg(mq + 2);
    hit.worldPos.x = m0.w + fmaf(m0.x, hit.objPos.x, fmaf(m0.y, hit.objPos.y, m0.z * hit.objPos.z));
    hit.worldPos.y = m1.w + fmaf(m1.x, hit.objPos.x, fmaf(m1.y, hit.objPos.y, m1.z * hit.objPos.z));
    hit.worldPos.z = m2.w + fmaf(m2.x, hit.objPos.x, fmaf(m2.y, hit.objPos.y, m2.z * hit.objPos.z));
    return THEIA_RESULT_SUCCESS;
}

// sh/scatter.surface.glsl:14-51
struct Reflectance { float n_in, n_tr, r_s, r_p; };
TDEV Reflectance fresnelReflect(const Ray& ray, const SurfaceHit& hit) {
    float cos_i = clamp_(dot(ray.direction, hit.rayNrm), -1.0f, 1.0f);
    float sin_i = sqrtf(max_(1.0f - cos_i * cos_i, 0.0f));
    uint64_t otherMed = ldp(hit.material, hit.inward ? MAT_INSIDE : MAT_OUTSIDE);
    float n_i = ray.constants.n;
    float n_t = 1.0f;
    if (otherMed != 0) {
        float u_t = normalize_lambda(otherMed, ray.wavelength);
        n_t = lookUp(ldp(otherMed, MED_N), u_t, 1.0f);
    }
    float sin_t = sin_i * n_i / n_t;
    float cos_t = sqrtf(max_(1.0f - sin_t * sin_t, 0.0f));
    cos_i = abs_(cos_i);
    float r_s = (n_i * cos_i - n_t * cos_t) / (n_i * cos_i + n_t * cos_t);
    float r_p = (n_t * cos_i - n_i * cos_t) / (n_t * cos_i + n_i * cos_t);
    return {n_i, n_t, r_s, r_p};
}

// sh/ray.surface.glsl:22-36 (Waechter & Binder, Ray Tracing Gems ch. 6)
TDEV float offsetComp(float p, float n) {
    int32_t of_i = (int32_t)(256.0f * n);
    float p_i = i2f(f2i(p) + ((p < 0.0f) ? -of_i : of_i));
    return abs_(p) < (1.0f / 32.0f) ? fmaf(1.0f / 65536.0f, n, p) : p_i;
}
TDEV vec3 offsetRay(vec3 p, vec3 n) { return {offsetComp(p.x, n.x), offsetComp(p.y, n.y), offsetComp(p.z, n.z)}; }

// sh/ray.surface.glsl:44-56
TDEV void crossBorder(Ray& ray, const SurfaceHit& hit) {
    ray.position = offsetRay(hit.worldPos, -hit.rayNrm);
    uint64_t medium = ldp(hit.material, hit.inward ? MAT_INSIDE : MAT_OUTSIDE);
    ray.medium = medium;
    ray.constants = lookUpMedium(medium, ray.wavelength);
}
// GLSL reflect / refract
TDEV vec3 reflect_(vec3 I, vec3 N) { return I - (2.0f * dot(N, I)) * N; }
TDEV vec3 refract_(vec3 I, vec3 N, float eta) {
    float d = dot(N, I);
    float k = 1.0f - eta * eta * (1.0f - d * d);
    if (k < 0.0f) return v3(0.0f);
    return eta * I - (eta * d + sqrtf(k)) * N;
}
// sh/ray.surface.glsl:66-74, 165-177
template <bool POL>
TDEV void reflectRayIS(Ray& ray, const SurfaceHit& hit, const Reflectance& r) {
    ray.position = offsetRay(hit.worldPos, hit.rayNrm);
    ray.direction = normalize(reflect_(ray.direction, hit.rayNrm));
    if (POL) ray.stokes = polarizerMatrix(r.r_p, r.r_s) * ray.stokes;
}
// :82-94, 178-185
template <bool POL>
TDEV void reflectRay(Ray& ray, const SurfaceHit& hit, const Reflectance& r) {
    ray.position = offsetRay(hit.worldPos, hit.rayNrm);
    ray.direction = normalize(reflect_(ray.direction, hit.rayNrm));
    float s = r.r_s, p = r.r_p;
    ray.lin_contrib *= 0.5f * (s * s + p * p);
    if (POL) ray.stokes = polarizerMatrix(r.r_p, r.r_s) * ray.stokes;
}
TDEV void applyTransmitStokes(Ray& ray, const Reflectance& r) {
    float eta = r.n_in / r.n_tr;
    float t_s = r.r_s + 1.0f;
    float t_p = (r.r_p + 1.0f) * eta;
    ray.stokes = polarizerMatrix(t_p, t_s) * ray.stokes;
}
// :104-114, 209-226
template <bool POL>
TDEV void transmitRayIS(Ray& ray, const SurfaceHit& hit, const Reflectance& r) {
    crossBorder(ray, hit);
    float eta = r.n_in / r.n_tr;
    ray.direction = normalize(refract_(ray.direction, hit.rayNrm, eta));
    if (POL) applyTransmitStokes(ray, r);
}
// :124-136, 227-238
template <bool POL>
TDEV void transmitRay(Ray& ray, const SurfaceHit& hit, const Reflectance& r) {
    crossBorder(ray, hit);
    float eta = r.n_in / r.n_tr;
    ray.direction = normalize(refract_(ray.direction, hit.rayNrm, eta));
    float s = r.r_s, p = r.r_p;
    ray.lin_contrib *= 1.0f - 0.5f * (s * s + p * p);
    if (POL) applyTransmitStokes(ray, r);
}

}  // namespace theia
