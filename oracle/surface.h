// ORACLE - TEST INFRASTRUCTURE ONLY. Never included, linked or called by the product path.
//
// CPU restatement of the surface interaction code: sh/scene.intersect.glsl (processRayQuery),
// sh/scatter.surface.glsl (fresnelReflect), sh/ray.surface.glsl (offsetRay, crossBorder,
// reflect/transmit).
#pragma once
#include "scene.h"

namespace orc {

// sh/scene.types.glsl:28-57
struct SurfaceHit {
    bool valid;
    const MaterialGLSL* material;
    bool inward;
    int32_t customId;
    uint32_t flags;
    vec3 worldPos, rayNrm;
    vec3 objPos, objNrm, objDir;
    mat3 worldToObj;
};

// sh/scene.intersect.glsl:23-99
static inline int processRayQuery(const Scene& sc, const Ray& ray, const RayHit& q, SurfaceHit& hit) {
    hit = SurfaceHit{};
    hit.valid = q.valid;
    if (!hit.valid) return THEIA_RESULT_RAY_MISSED;
    const Instance& in = sc.instances[q.instance];
    hit.customId = in.customIndex;
    const float* verts = (const float*)(uintptr_t)in.geom.vertices;
    const int32_t* index = (const int32_t*)(uintptr_t)in.geom.indices + 3 * q.prim;
    const float* p0 = verts + 6 * index[0];
    const float* p1 = verts + 6 * index[1];
    const float* p2 = verts + 6 * index[2];
    vec3 v0 = v3(p0), v1 = v3(p1), v2 = v3(p2);
    vec3 n0 = v3(p0 + 3), nn1 = v3(p1 + 3), nn2 = v3(p2 + 3);
    vec3 e1 = v1 - v0, e2 = v2 - v0;
    float bx = q.u, by = q.v;
    // v0 + fma(vec3(b.x), e1, b.y * e2)  (precise)
    hit.objPos = {v0.x + fmaf(bx, e1.x, by * e2.x), v0.y + fmaf(bx, e1.y, by * e2.y), v0.z + fmaf(bx, e1.z, by * e2.z)};
    hit.objNrm = cross(e1, e2);
    vec3 d1 = nn1 - n0, d2 = nn2 - n0;
    vec3 intNrm = {n0.x + fmaf(bx, d1.x, by * d2.x), n0.y + fmaf(bx, d1.y, by * d2.y), n0.z + fmaf(bx, d1.z, by * d2.z)};
    hit.objNrm = hit.objNrm * signBit(dot(hit.objNrm, intNrm));
    hit.objNrm = normalize(hit.objNrm);
    // world -> object (columns of the 3x4 row major matrix)
    const float* w = in.w2o;
    hit.worldToObj = mat3{{w[0], w[4], w[8]}, {w[1], w[5], w[9]}, {w[2], w[6], w[10]}};
    hit.objDir = normalize(hit.worldToObj * ray.direction);
    hit.inward = dot(hit.objDir, hit.objNrm) <= 0.0f;
    hit.material = (const MaterialGLSL*)(uintptr_t)in.geom.material;
    hit.flags = hit.inward ? hit.material->flagsInwards : hit.material->flagsOutwards;
    uint64_t medium = hit.inward ? hit.material->outside : hit.material->inside;
    if (ray.medium != medium) return THEIA_ERROR_MEDIA_MISMATCH;
    vec3 worldNrm = normalize(mulT(hit.objNrm, hit.worldToObj));
    hit.rayNrm = worldNrm * (2.0f * (hit.inward ? 1.0f : 0.0f) - 1.0f);
    const float* m = in.o2w;
    hit.worldPos.x = m[3] + fmaf(m[0], hit.objPos.x, fmaf(m[1], hit.objPos.y, m[2] * hit.objPos.z));
    hit.worldPos.y = m[7] + fmaf(m[4], hit.objPos.x, fmaf(m[5], hit.objPos.y, m[6] * hit.objPos.z));
    hit.worldPos.z = m[11] + fmaf(m[8], hit.objPos.x, fmaf(m[9], hit.objPos.y, m[10] * hit.objPos.z));
    return THEIA_RESULT_SUCCESS;
}

// sh/scatter.surface.glsl:14-51
struct Reflectance { float n_in, n_tr, r_s, r_p; };
static inline Reflectance fresnelReflect(const Ray& ray, const SurfaceHit& hit) {
    float cos_i = clamp_(dot(ray.direction, hit.rayNrm), -1.0f, 1.0f);
    float sin_i = std::sqrt(max_(1.0f - cos_i * cos_i, 0.0f));
    uint64_t otherMed = hit.inward ? hit.material->inside : hit.material->outside;
    float n_i = ray.constants.n;
    float n_t = 1.0f;
    if (otherMed != 0) {
        const MediumGLSL* om = asMedium(otherMed);
        float u_t = normalize_lambda(om, ray.wavelength);
        n_t = lookUp(om->n, u_t, 1.0f);
    }
    float sin_t = sin_i * n_i / n_t;
    float cos_t = std::sqrt(max_(1.0f - sin_t * sin_t, 0.0f));
    cos_i = abs_(cos_i);
    float r_s = (n_i * cos_i - n_t * cos_t) / (n_i * cos_i + n_t * cos_t);
    float r_p = (n_t * cos_i - n_i * cos_t) / (n_t * cos_i + n_i * cos_t);
    return {n_i, n_t, r_s, r_p};
}

// sh/ray.surface.glsl:22-36
static inline float offsetComp(float p, float n) {
    int32_t of_i = (int32_t)(256.0f * n);
    float p_i = i2f(f2i(p) + ((p < 0.0f) ? -of_i : of_i));
    return abs_(p) < (1.0f / 32.0f) ? fmaf(1.0f / 65536.0f, n, p) : p_i;
}
static inline vec3 offsetRay(vec3 p, vec3 n) { return {offsetComp(p.x, n.x), offsetComp(p.y, n.y), offsetComp(p.z, n.z)}; }

// sh/ray.surface.glsl:44-56
static inline void crossBorder(Ray& ray, const SurfaceHit& hit) {
    ray.position = offsetRay(hit.worldPos, -hit.rayNrm);
    uint64_t medium = hit.inward ? hit.material->inside : hit.material->outside;
    ray.medium = medium;
    ray.constants = lookUpMedium(medium, ray.wavelength);
}
// GLSL reflect / refract
static inline vec3 reflect_(vec3 I, vec3 N) { return I - (2.0f * dot(N, I)) * N; }
static inline vec3 refract_(vec3 I, vec3 N, float eta) {
    float d = dot(N, I);
    float k = 1.0f - eta * eta * (1.0f - d * d);
    if (k < 0.0f) return v3(0.0f);
    return eta * I - (eta * d + std::sqrt(k)) * N;
}
// sh/ray.surface.glsl:66-74, 165-177
static inline void reflectRayIS(Ray& ray, const SurfaceHit& hit, const Reflectance& r, bool polarized) {
    ray.position = offsetRay(hit.worldPos, hit.rayNrm);
    ray.direction = normalize(reflect_(ray.direction, hit.rayNrm));
    if (polarized) ray.stokes = polarizerMatrix(r.r_p, r.r_s) * ray.stokes;
}
// :82-94, 178-185
static inline void reflectRay(Ray& ray, const SurfaceHit& hit, const Reflectance& r, bool polarized) {
    ray.position = offsetRay(hit.worldPos, hit.rayNrm);
    ray.direction = normalize(reflect_(ray.direction, hit.rayNrm));
    float s = r.r_s, p = r.r_p;
    ray.lin_contrib *= 0.5f * (s * s + p * p);
    if (polarized) ray.stokes = polarizerMatrix(r.r_p, r.r_s) * ray.stokes;
}
static inline void applyTransmitStokes(Ray& ray, const Reflectance& r) {
    float eta = r.n_in / r.n_tr;
    float t_s = r.r_s + 1.0f;
    float t_p = (r.r_p + 1.0f) * eta;
    ray.stokes = polarizerMatrix(t_p, t_s) * ray.stokes;
}
// :104-114, 209-226
static inline void transmitRayIS(Ray& ray, const SurfaceHit& hit, const Reflectance& r, bool polarized) {
    crossBorder(ray, hit);
    float eta = r.n_in / r.n_tr;
    ray.direction = normalize(refract_(ray.direction, hit.rayNrm, eta));
    if (polarized) applyTransmitStokes(ray, r);
}
// :124-136, 227-238
static inline void transmitRay(Ray& ray, const SurfaceHit& hit, const Reflectance& r, bool polarized) {
    crossBorder(ray, hit);
    float eta = r.n_in / r.n_tr;
    ray.direction = normalize(refract_(ray.direction, hit.rayNrm, eta));
    float s = r.r_s, p = r.r_p;
    ray.lin_contrib *= 1.0f - 0.5f * (s * s + p * p);
    if (polarized) applyTransmitStokes(ray, r);
}

}  // namespace orc
