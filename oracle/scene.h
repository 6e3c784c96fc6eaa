// ORACLE - TEST INFRASTRUCTURE ONLY. Never included, linked or called by the product path.
//
// Ray query of the oracle. In the reference this arithmetic lives in the Vulkan driver
// (GL_EXT_ray_query against a TLAS, sh/scene.traverse.glsl:310-323); it is NOT in the
// reference sources. The oracle restates the semantics of the Vulkan ray query:
//   closest committed opaque triangle hit with t in (0, tmax), no culling, cull mask 0xFF,
//   barycentrics (u, v) weighting v1 and v2, instance id = TLAS build order,
//   custom index = detectorId, float32 object<->world 3x4 matrices.
// Triangle test: watertight algorithm of Woop, Benthin, Wald (JCGT 2013) in object space
// with an unnormalised object space direction (so t is the world space parameter) and a
// double precision fallback for exactly-zero edge functions. Ties in t are broken towards
// the lower (instance id, primitive id). With this rule the result does not depend on the
// acceleration structure, which is what lets the GPU's wide BVH be checked bit-exactly.
#pragma once
#include <algorithm>
#include <vector>
#include "physics.h"

namespace orc {

struct RayHit {
    bool valid;
    int32_t instance, prim;
    float t, u, v;
};

struct AABB {
    float lo[3], hi[3];
    void reset() { for (int i = 0; i < 3; ++i) { lo[i] = F_INF; hi[i] = -F_INF; } }
    void grow(const float* p) { for (int i = 0; i < 3; ++i) { lo[i] = min_(lo[i], p[i]); hi[i] = max_(hi[i], p[i]); } }
    void grow(const AABB& b) { for (int i = 0; i < 3; ++i) { lo[i] = min_(lo[i], b.lo[i]); hi[i] = max_(hi[i], b.hi[i]); } }
    void pad() {
        for (int i = 0; i < 3; ++i) {
            float m = max_(abs_(lo[i]), abs_(hi[i]));
            float e = 4e-6f * m + 1e-30f;
            lo[i] -= e; hi[i] += e;
        }
    }
};

// Generic binary BVH over boxed primitives (median split along the widest centroid axis).
struct BVH {
    struct Node { AABB box; int32_t left, right, first, count; };  // leaf: count > 0
    std::vector<Node> nodes;
    std::vector<int32_t> order;

    void build(const std::vector<AABB>& boxes) {
        int n = (int)boxes.size();
        order.resize(n);
        for (int i = 0; i < n; ++i) order[i] = i;
        nodes.clear();
        nodes.reserve(2 * n + 1);
        if (n > 0) buildRec(boxes, 0, n);
    }
    int buildRec(const std::vector<AABB>& boxes, int first, int count) {
        int id = (int)nodes.size();
        nodes.push_back(Node{});
        AABB box, cbox;
        box.reset(); cbox.reset();
        for (int i = first; i < first + count; ++i) {
            const AABB& b = boxes[order[i]];
            box.grow(b);
            float c[3] = {0.5f * (b.lo[0] + b.hi[0]), 0.5f * (b.lo[1] + b.hi[1]), 0.5f * (b.lo[2] + b.hi[2])};
            cbox.grow(c);
        }
        box.pad();
        nodes[id].box = box;
        if (count <= 4) {
            nodes[id].first = first; nodes[id].count = count; nodes[id].left = nodes[id].right = -1;
            return id;
        }
        int axis = 0;
        float ext = cbox.hi[0] - cbox.lo[0];
        for (int a = 1; a < 3; ++a) if (cbox.hi[a] - cbox.lo[a] > ext) { ext = cbox.hi[a] - cbox.lo[a]; axis = a; }
        int mid = first + count / 2;
        std::nth_element(order.begin() + first, order.begin() + mid, order.begin() + first + count,
                         [&](int a, int b) {
                             float ca = boxes[a].lo[axis] + boxes[a].hi[axis];
                             float cb = boxes[b].lo[axis] + boxes[b].hi[axis];
                             return ca < cb || (ca == cb && a < b);
                         });
        int l = buildRec(boxes, first, mid - first);
        int r = buildRec(boxes, mid, first + count - mid);
        nodes[id].left = l; nodes[id].right = r; nodes[id].count = 0; nodes[id].first = 0;
        return id;
    }
};

// conservative slab test on [0, tmax]
static inline bool hitBox(const AABB& b, const float o[3], const float inv[3], float tmax) {
    float tn = 0.0f, tf = tmax;
    for (int a = 0; a < 3; ++a) {
        float t0 = (b.lo[a] - o[a]) * inv[a];
        float t1 = (b.hi[a] - o[a]) * inv[a];
        if (t0 != t0 || t1 != t1) continue;  // 0 * inf: origin on the slab plane, keep
        float lo = min_(t0, t1), hi = max_(t0, t1);
        tn = max_(tn, lo);
        tf = min_(tf, hi);
    }
    return tn <= tf * 1.0000005f + 1e-30f;
}

// Per (ray, instance) setup of the watertight test. The permutation (kx, ky, kz) and shear
// (Sx, Sy, Sz) of the paper are folded into three vectors and offsets: the sheared
// coordinates of a vertex v are x' = v . mx + cx, y' = v . my + cy, z' = v . mz + cz with
// mx = e_kx - Sx e_kz, my = e_ky - Sy e_kz, mz = Sz e_kz, c = -(o . m), each one fma chain.
struct WoopRay {
    float o[3], d[3];
    float mx[3], my[3], mz[3];
    float cx, cy, cz;
};
static inline float dot3(const float a[3], const float b[3]) { return fmaf(a[2], b[2], fmaf(a[1], b[1], a[0] * b[0])); }
static inline void woopSetup(WoopRay& r) {
    float ax = abs_(r.d[0]), ay = abs_(r.d[1]), az = abs_(r.d[2]);
    int kz = 0;
    float m = ax;
    if (ay > m) { kz = 1; m = ay; }
    if (az > m) { kz = 2; m = az; }
    int kx = kz + 1; if (kx == 3) kx = 0;
    int ky = kx + 1; if (ky == 3) ky = 0;
    if (r.d[kz] < 0.0f) { int t = kx; kx = ky; ky = t; }
    const float Sz = 1.0f / r.d[kz];
    const float Sx = r.d[kx] * Sz;
    const float Sy = r.d[ky] * Sz;
    for (int k = 0; k < 3; ++k) {
        r.mx[k] = k == kz ? -Sx : (k == kx ? 1.0f : 0.0f);
        r.my[k] = k == kz ? -Sy : (k == ky ? 1.0f : 0.0f);
        r.mz[k] = k == kz ? Sz : 0.0f;
    }
    r.cx = -dot3(r.o, r.mx);
    r.cy = -dot3(r.o, r.my);
    r.cz = -dot3(r.o, r.mz);
}
static inline float shear(const float* v, const float m[3], float c) {
    return fmaf(v[2], m[2], fmaf(v[1], m[1], fmaf(v[0], m[0], c)));
}
// returns true and (t,u,v) if the triangle is hit with t in (0, tmax)
static inline bool woopIntersect(const WoopRay& r, const float* v0, const float* v1, const float* v2,
                                 float tmax, float* t_out, float* u_out, float* v_out) {
    float Ax = shear(v0, r.mx, r.cx), Ay = shear(v0, r.my, r.cy);
    float Bx = shear(v1, r.mx, r.cx), By = shear(v1, r.my, r.cy);
    float Cx = shear(v2, r.mx, r.cx), Cy = shear(v2, r.my, r.cy);
    float U = Cx * By - Cy * Bx;
    float V = Ax * Cy - Ay * Cx;
    float W = Bx * Ay - By * Ax;
    if (U == 0.0f || V == 0.0f || W == 0.0f) {
        double CxBy = (double)Cx * (double)By, CyBx = (double)Cy * (double)Bx;
        U = (float)(CxBy - CyBx);
        double AxCy = (double)Ax * (double)Cy, AyCx = (double)Ay * (double)Cx;
        V = (float)(AxCy - AyCx);
        double BxAy = (double)Bx * (double)Ay, ByAx = (double)By * (double)Ax;
        W = (float)(BxAy - ByAx);
    }
    if ((U < 0.0f || V < 0.0f || W < 0.0f) && (U > 0.0f || V > 0.0f || W > 0.0f)) return false;
    float det = U + V + W;
    if (det == 0.0f) return false;
    float Az = shear(v0, r.mz, r.cz), Bz = shear(v1, r.mz, r.cz), Cz = shear(v2, r.mz, r.cz);
    float T = fmaf(W, Cz, fmaf(V, Bz, U * Az));
    if (det < 0.0f) {
        if (!(T < 0.0f) || !(T > tmax * det)) return false;
    } else {
        if (!(T > 0.0f) || !(T < tmax * det)) return false;
    }
    float rcp = 1.0f / det;
    *t_out = T * rcp;
    *u_out = V * rcp;
    *v_out = W * rcp;
    return true;
}

struct Mesh {
    const float* verts;    // [nVerts][6]
    const int32_t* idx;    // [nTris][3]
    uint32_t nTris;
    BVH bvh;
};
struct Instance {
    float o2w[12], w2o[12];
    int mesh;
    theia_geometry_t geom;
    int32_t customIndex;
};
struct Scene {
    std::vector<Mesh> meshes;
    std::vector<Instance> instances;
    BVH tlas;
    bool bruteForce = false;
    // traversal statistics (filled when stats != nullptr)
};

static inline void transformPoint(const float m[12], const float p[3], float out[3]) {
    for (int i = 0; i < 3; ++i)
        out[i] = fmaf(m[4 * i + 2], p[2], fmaf(m[4 * i + 1], p[1], fmaf(m[4 * i + 0], p[0], m[4 * i + 3])));
}
static inline void transformVector(const float m[12], const float p[3], float out[3]) {
    for (int i = 0; i < 3; ++i)
        out[i] = fmaf(m[4 * i + 2], p[2], fmaf(m[4 * i + 1], p[1], m[4 * i + 0] * p[0]));
}

static inline Scene* sceneCreate(const theia_instance_desc_t* desc, uint32_t n, const uint32_t* nVerts,
                                 const uint32_t* nTris, bool bruteForce) {
    Scene* s = new Scene();
    s->bruteForce = bruteForce;
    std::vector<AABB> ibox(n);
    for (uint32_t i = 0; i < n; ++i) {
        const float* verts = (const float*)(uintptr_t)desc[i].geometry.vertices;
        const int32_t* idx = (const int32_t*)(uintptr_t)desc[i].geometry.indices;
        int mesh = -1;
        for (size_t m = 0; m < s->meshes.size(); ++m)
            if (s->meshes[m].verts == verts && s->meshes[m].idx == idx) mesh = (int)m;
        if (mesh < 0) {
            Mesh me;
            me.verts = verts; me.idx = idx; me.nTris = nTris[i];
            std::vector<AABB> boxes(nTris[i]);
            for (uint32_t t = 0; t < nTris[i]; ++t) {
                boxes[t].reset();
                for (int k = 0; k < 3; ++k) boxes[t].grow(verts + 6 * idx[3 * t + k]);
            }
            me.bvh.build(boxes);
            s->meshes.push_back(std::move(me));
            mesh = (int)s->meshes.size() - 1;
        }
        Instance in;
        std::memcpy(in.o2w, desc[i].obj_to_world, 48);
        std::memcpy(in.w2o, desc[i].world_to_obj, 48);
        in.mesh = mesh; in.geom = desc[i].geometry; in.customIndex = desc[i].custom_index;
        s->instances.push_back(in);
        // world box from the transformed vertices
        ibox[i].reset();
        for (uint32_t v = 0; v < nVerts[i]; ++v) {
            float w[3];
            transformPoint(in.o2w, verts + 6 * v, w);
            ibox[i].grow(w);
        }
        ibox[i].pad();
    }
    s->tlas.build(ibox);
    return s;
}

static inline void intersectInstance(const Scene& sc, int inst, const float o[3], const float d[3], float tmax,
                                     RayHit& best, uint64_t* stats) {
    const Instance& in = sc.instances[inst];
    const Mesh& me = sc.meshes[in.mesh];
    WoopRay r;
    transformPoint(in.w2o, o, r.o);
    transformVector(in.w2o, d, r.d);
    woopSetup(r);
    if (stats) stats[2]++;
    auto test = [&](int prim) {
        const int32_t* ix = me.idx + 3 * prim;
        float t, u, v;
        if (stats) stats[1]++;
        if (!woopIntersect(r, me.verts + 6 * ix[0], me.verts + 6 * ix[1], me.verts + 6 * ix[2], tmax, &t, &u, &v))
            return;
        bool better = !best.valid || t < best.t ||
                      (t == best.t && (inst < best.instance || (inst == best.instance && prim < best.prim)));
        if (better) best = RayHit{true, inst, prim, t, u, v};
    };
    if (sc.bruteForce) {
        for (uint32_t p = 0; p < me.nTris; ++p) test((int)p);
        return;
    }
    float inv[3] = {1.0f / r.d[0], 1.0f / r.d[1], 1.0f / r.d[2]};
    int stack[64], sp = 0;
    if (me.bvh.nodes.empty()) return;
    stack[sp++] = 0;
    while (sp > 0) {
        const BVH::Node& nd = me.bvh.nodes[stack[--sp]];
        if (stats) stats[0]++;
        float lim = best.valid ? best.t : tmax;
        if (!hitBox(nd.box, r.o, inv, lim)) continue;
        if (nd.count > 0) {
            for (int i = nd.first; i < nd.first + nd.count; ++i) test(me.bvh.order[i]);
        } else {
            stack[sp++] = nd.left;
            stack[sp++] = nd.right;
        }
    }
}

// The ray query: closest hit over all instances.
static inline RayHit closestHit(const Scene& sc, vec3 origin, vec3 dir, float tmax, uint64_t* stats = nullptr) {
    RayHit best{false, -1, -1, 0.0f, 0.0f, 0.0f};
    float o[3] = {origin.x, origin.y, origin.z}, d[3] = {dir.x, dir.y, dir.z};
    if (stats) stats[3]++;
    if (sc.bruteForce) {
        for (size_t i = 0; i < sc.instances.size(); ++i) intersectInstance(sc, (int)i, o, d, tmax, best, stats);
        return best;
    }
    if (sc.tlas.nodes.empty()) return best;
    float inv[3] = {1.0f / d[0], 1.0f / d[1], 1.0f / d[2]};
    int stack[64], sp = 0;
    stack[sp++] = 0;
    while (sp > 0) {
        const BVH::Node& nd = sc.tlas.nodes[stack[--sp]];
        if (stats) stats[0]++;
        float lim = best.valid ? best.t : tmax;
        if (!hitBox(nd.box, o, inv, lim)) continue;
        if (nd.count > 0) {
            for (int i = nd.first; i < nd.first + nd.count; ++i)
                intersectInstance(sc, sc.tlas.order[i], o, d, tmax, best, stats);
        } else {
            stack[sp++] = nd.left;
            stack[sp++] = nd.right;
        }
    }
    return best;
}

}  // namespace orc
