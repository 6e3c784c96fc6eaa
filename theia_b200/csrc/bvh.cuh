// Acceleration structure format and software ray query.
//
// The reference traces rays with GL_EXT_ray_query against a Vulkan TLAS built by the driver
// (sh/scene.traverse.glsl:310-323, src/theia/scene.py:550, 668); B200 has no RT cores, so
// the structure and its traversal are part of this library:
//   * two levels: one BLAS per mesh (object space), one TLAS over instances (world space);
//   * 8-wide nodes of 80 bytes with child boxes quantised to 8 bit per axis on a
//     power-of-two grid anchored at the node's lower corner, read with five 128 bit loads;
//     children are stored sorted along the node's widest axis so that slots can be visited
//     near-to-far by scanning the hit mask upwards or downwards;
//   * leaf triangles are stored pre-gathered (3 x 16 bytes, primitive id in the padding).
// Ray query semantics (Vulkan): closest opaque triangle with t in (0, tmax), no culling,
// barycentrics (u, v) weighting v1 and v2. Triangle test: watertight test of Woop, Benthin
// and Wald (JCGT 2013) in object space with the unnormalised direction (t is the world
// parameter), double precision fallback for zero edge functions; ties in t go to the lower
// (instance, primitive). The result is therefore independent of the tree, and the box tests
// only have to be conservative: quantised boxes are rounded outwards and every slab
// interval is widened by a few ulp of its operands.
#pragma once
#include "detmath.cuh"

namespace theia {

struct alignas(16) WideNode {  // 80 bytes
    float px, py, pz;          // origin of the quantisation grid (lower corner)
    uint8_t ex, ey, ez;        // biased float exponent of the grid step per axis
    uint8_t imask;             // bit i set: child i is an inner node
    uint32_t childBase;        // index of the first inner child (inner children are contiguous, slot order)
    uint32_t primBase;         // index of the first primitive of the leaf children (contiguous, slot order)
    uint8_t meta[8];           // 0: empty; inner: 1 + rank among inner children; leaf: count | offset << 3
    uint8_t qlox[8], qloy[8], qloz[8];
    uint8_t qhix[8], qhiy[8], qhiz[8];
};
static_assert(sizeof(WideNode) == 80, "WideNode layout");

struct alignas(16) TriRec {  // 48 bytes
    float v0x, v0y, v0z;
    uint32_t prim;
    float v1x, v1y, v1z;
    uint32_t pad1;
    float v2x, v2y, v2z;
    uint32_t pad2;
};
static_assert(sizeof(TriRec) == 48, "TriRec layout");

#define THEIA_BLAS_MAGIC 0x534C4254u  // "TBLS"
#define THEIA_TLAS_MAGIC 0x534C5454u  // "TTLS"
#ifndef THEIA_LEAF_MAX
#define THEIA_LEAF_MAX 4
#endif
#define THEIA_STACK_SIZE 128  // traversal stack entries per ray

struct alignas(16) BlasHeader {  // 64 bytes, followed by nodes then triangles
    uint32_t magic, nNodes, nTris, orderAxis;
    uint64_t nodesOffset, trisOffset;
    float lo[3], hi[3];  // object space bounds
    uint32_t pad[2];
};
static_assert(sizeof(BlasHeader) == 64, "BlasHeader layout");

struct alignas(16) InstanceRec {  // 144 bytes
    float w2o[12];                // 3x4 row major world -> object
    const WideNode* blasNodes;
    const TriRec* blasTris;
    float o2w[12];                // 3x4 row major object -> world
    uint64_t vertices, indices, material;  // Scene.GLSLGeometry (scene.py:628-635)
    int32_t customIndex;
    uint32_t pad;  // instance id (the records are stored in leaf order)
};
static_assert(sizeof(InstanceRec) == 144, "InstanceRec layout");

struct alignas(16) TlasHeader {  // 64 bytes, followed by nodes then instance records
    uint32_t magic, nNodes, nInstances, pad0;
    uint64_t nodesOffset, instancesOffset;
    float lo[3], hi[3];
    uint32_t pad[2];
};
static_assert(sizeof(TlasHeader) == 64, "TlasHeader layout");

struct RayHit {
    bool valid;
    int32_t instance, prim;  // instance id (TLAS build order), primitive id inside the mesh
    int32_t slot;            // position of the instance record inside the TLAS (leaf order)
    float t, u, v;
};
struct TraversalStats { uint32_t nodes, tris, instances; };

// Node fetch: five 128 bit loads through the read-only path.
struct NodeRegs { uint4 a, b, c, d, e; };
TDEV NodeRegs loadNode(const WideNode* n) {
    const uint4* p = reinterpret_cast<const uint4*>(n);
    NodeRegs r;
    r.a = __ldg(p); r.b = __ldg(p + 1); r.c = __ldg(p + 2); r.d = __ldg(p + 3); r.e = __ldg(p + 4);
    return r;
}
TDEV NodeRegs loadNodeShared(const WideNode* n) {
    const uint4* p = reinterpret_cast<const uint4*>(n);
    NodeRegs r;
    r.a = p[0]; r.b = p[1]; r.c = p[2]; r.d = p[3]; r.e = p[4];
    return r;
}

// Byte i (0..7) of the eight child meta bytes of a node: one byte permute with the slot as its selector.
TDEV uint32_t metaByte(const NodeRegs& n, int i) { return __byte_perm(n.b.z, n.b.w, (uint32_t)i) & 0xFFu; }
// t[i] for a run time i out of eight registers (seven selects; an indexed array would be spilled to local memory)
TDEV float pick8(const float t[8], int i) {
    const bool b0 = (i & 1) != 0, b1 = (i & 2) != 0, b2 = (i & 4) != 0;
    const float a0 = b0 ? t[1] : t[0], a1 = b0 ? t[3] : t[2], a2 = b0 ? t[5] : t[4], a3 = b0 ? t[7] : t[6];
    const float c0 = b1 ? a1 : a0, c1 = b1 ? a3 : a2;
    return b2 ? c1 : c0;
}

// Rays and results stream through once: keep them out of L1, which holds the nodes and triangles that are read again.
#ifndef THEIA_STREAM_RAYS
#define THEIA_STREAM_RAYS 1
#endif
TDEV float4 loadStream(const float4* p) {
#if THEIA_STREAM_RAYS
    float4 v;
    asm("ld.global.nc.L1::no_allocate.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
    return v;
#else
    return __ldg(p);
#endif
}
TDEV void storeStream(uint4* p, uint4 v) {
#if THEIA_STREAM_RAYS
    asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1, %2, %3, %4};" :: "l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
#else
    *p = v;
#endif
}

// (feeds the conservative box tests only: the 1 ulp of rcp.approx is inside their slack)
TDEV float safeInv(float d) {
    float inv;
    asm("rcp.approx.f32 %0, %1;" : "=f"(inv) : "f"(d));
    if (!(abs_(inv) <= 1e30f)) inv = copysignf(1e30f, d);
    return inv;
}
TDEV uint32_t byteOf(uint32_t w, int i) { return (w >> (8 * i)) & 0xFFu; }
// 32768 + (byte i of w) as a float, built by ONE byte permute: bytes {0, b, 0, 0x47} are the
// float 2^15 * (1 + b * 2^-15). (I2F.U8 runs on the quarter rate conversion unit, and 48 of
// them per node made the traversal bound by it.) The caller folds the 2^15 into its constant.
#define THEIA_BYTE_BIAS 32768.0f
static __constant__ uint32_t g_byteMagic = 0x47000000u;
TDEV float biasedByte(uint32_t w, int i, uint32_t magic /* 0x47000000 in a register */) {
    return __uint_as_float(__byte_perm(w, magic, 0x7504u | ((uint32_t)i << 4)));
}

// Ray in the space of the structure being traversed.
struct SlabRay {
    float ox, oy, oz;
    float ix, iy, iz;  // safe inverse direction
};

// Tests the eight quantised child boxes of a node against [0, lim]; returns the hit mask and
// the entry distances. Each axis interval is widened by ~5 ulp of its operands.
TDEV uint32_t testChildren(const NodeRegs& n, const SlabRay& r, float lim, float tmin[8]) {
    // slack per axis: ~5 ulp of the operands (EPS) plus the rounding of the biased constant
    // below, 2^-24 * 2^15 |a| = 2^-9 |a|, covered twice over by 2^-8 |a| (0.4 % of a grid cell)
    const float EPS = 6e-7f, SLACK_A = 256.0f * 6e-7f + 0.00390625f;
    float px = u2f(n.a.x), py = u2f(n.a.y), pz = u2f(n.a.z);
    float sx = u2f(byteOf(n.a.w, 0) << 23), sy = u2f(byteOf(n.a.w, 1) << 23), sz = u2f(byteOf(n.a.w, 2) << 23);
    float ax = sx * r.ix, ay = sy * r.iy, az = sz * r.iz;
    float bx = (px - r.ox) * r.ix, by = (py - r.oy) * r.iy, bz = (pz - r.oz) * r.iz;
    float ex = fmaf(SLACK_A, fabsf(ax), EPS * fabsf(bx));
    float ey = fmaf(SLACK_A, fabsf(ay), EPS * fabsf(by));
    float ez = fmaf(SLACK_A, fabsf(az), EPS * fabsf(bz));
    // plane distance = q * a + b with the byte q read as 32768 + q: subtract 32768 a from b once
    float cx = fmaf(-THEIA_BYTE_BIAS, ax, bx), cy = fmaf(-THEIA_BYTE_BIAS, ay, by), cz = fmaf(-THEIA_BYTE_BIAS, az, bz);
    float bxl = cx - ex, bxh = cx + ex, byl = cy - ey, byh = cy + ey, bzl = cz - ez, bzh = cz + ez;
    // read from constant memory so that the permute takes it as its constant bank operand and
    // the selector as its immediate (as a literal it would claim the immediate slot and every
    // selector would need a register of its own)
    const uint32_t magic = g_byteMagic;
    // per axis: words holding the near / far plane bytes of children 0-3 and 4-7
    // layout: b = {childBase, primBase, meta0-3, meta4-7}, c = {qlox, qlox, qloy, qloy},
    //         d = {qloz, qloz, qhix, qhix}, e = {qhiy, qhiy, qhiz, qhiz}
    const bool nx = r.ix < 0.0f, ny = r.iy < 0.0f, nz = r.iz < 0.0f;
    uint32_t xn0 = nx ? n.d.z : n.c.x, xn1 = nx ? n.d.w : n.c.y, xf0 = nx ? n.c.x : n.d.z, xf1 = nx ? n.c.y : n.d.w;
    uint32_t yn0 = ny ? n.e.x : n.c.z, yn1 = ny ? n.e.y : n.c.w, yf0 = ny ? n.c.z : n.e.x, yf1 = ny ? n.c.w : n.e.y;
    uint32_t zn0 = nz ? n.e.z : n.d.x, zn1 = nz ? n.e.w : n.d.y, zf0 = nz ? n.d.x : n.e.z, zf1 = nz ? n.d.y : n.e.w;
    uint32_t mask = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int j = i & 3;
        uint32_t meta = byteOf(i < 4 ? n.b.z : n.b.w, j);
        float tnx = fmaf(biasedByte(i < 4 ? xn0 : xn1, j, magic), ax, bxl);
        float tfx = fmaf(biasedByte(i < 4 ? xf0 : xf1, j, magic), ax, bxh);
        float tny = fmaf(biasedByte(i < 4 ? yn0 : yn1, j, magic), ay, byl);
        float tfy = fmaf(biasedByte(i < 4 ? yf0 : yf1, j, magic), ay, byh);
        float tnz = fmaf(biasedByte(i < 4 ? zn0 : zn1, j, magic), az, bzl);
        float tfz = fmaf(biasedByte(i < 4 ? zf0 : zf1, j, magic), az, bzh);
        float tn = fmaxf(fmaxf(tnx, tny), fmaxf(tnz, 0.0f));
        float tf = fminf(fminf(tfx, tfy), fminf(tfz, lim));
        tmin[i] = tn;
        if (meta != 0 && tn <= tf) mask |= 1u << i;
    }
    return mask;
}

// Per (ray, instance) setup of the watertight test (Woop, Benthin, Wald, JCGT 2013). The
// permutation (kx, ky, kz) and the shear (Sx, Sy, Sz) of the paper are folded into three
// vectors and offsets, so that the sheared coordinates of a vertex v are plain fma chains
//   x' = v . mx + cx,  y' = v . my + cy,  z' = v . mz + cz
// with mx = e_kx - Sx e_kz, my = e_ky - Sy e_kz, mz = Sz e_kz and c = -(o . m): no per
// triangle selects by axis. Shared vertices get identical sheared coordinates in every
// triangle, which is what makes the edge tests watertight.
struct WoopRay {
    float o[3], d[3];  // object space ray (setup input)
    float mx[3], my[3], mz[3];
    float cx, cy, cz;
};
TDEV float pick(const float v[3], int k) { return k == 0 ? v[0] : (k == 1 ? v[1] : v[2]); }
TDEV float dot3(const float a[3], const float b[3]) { return fmaf(a[2], b[2], fmaf(a[1], b[1], a[0] * b[0])); }
TDEV void woopSetup(WoopRay& r) {
    float ax = abs_(r.d[0]), ay = abs_(r.d[1]), az = abs_(r.d[2]);
    int kz = 0;
    float m = ax;
    if (ay > m) { kz = 1; m = ay; }
    if (az > m) { kz = 2; m = az; }
    int kx = kz + 1; if (kx == 3) kx = 0;
    int ky = kx + 1; if (ky == 3) ky = 0;
    if (pick(r.d, kz) < 0.0f) { int t = kx; kx = ky; ky = t; }
    const float Sz = 1.0f / pick(r.d, kz);
    const float Sx = pick(r.d, kx) * Sz;
    const float Sy = pick(r.d, ky) * Sz;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        r.mx[k] = k == kz ? -Sx : (k == kx ? 1.0f : 0.0f);
        r.my[k] = k == kz ? -Sy : (k == ky ? 1.0f : 0.0f);
        r.mz[k] = k == kz ? Sz : 0.0f;
    }
    r.cx = -dot3(r.o, r.mx);
    r.cy = -dot3(r.o, r.my);
    r.cz = -dot3(r.o, r.mz);
}
TDEV float shear(const float v[3], const float m[3], float c) { return fmaf(v[2], m[2], fmaf(v[1], m[1], fmaf(v[0], m[0], c))); }
TDEV bool woopIntersect(const WoopRay& r, const float v0[3], const float v1[3], const float v2[3], float tmax,
                        float* t_out, float* u_out, float* v_out) {
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

TDEV void transformPoint(const float m[12], const float p[3], float out[3]) {
#pragma unroll
    for (int i = 0; i < 3; ++i)
        out[i] = fmaf(m[4 * i + 2], p[2], fmaf(m[4 * i + 1], p[1], fmaf(m[4 * i + 0], p[0], m[4 * i + 3])));
}
TDEV void transformVector(const float m[12], const float p[3], float out[3]) {
#pragma unroll
    for (int i = 0; i < 3; ++i) out[i] = fmaf(m[4 * i + 2], p[2], fmaf(m[4 * i + 1], p[1], m[4 * i + 0] * p[0]));
}

}  // namespace theia
