// Two level traversal of the wide BVH: the software ray query (see bvh.cuh).
#pragma once
#include "bvh.cuh"

namespace theia {

#ifndef THEIA_TOP_NODES
#define THEIA_TOP_NODES 72  // TLAS nodes staged in shared memory per CTA (breadth first prefix)
#endif

enum : uint32_t { ENTRY_TLAS = 0u, ENTRY_INST = 1u << 30, ENTRY_BLAS = 2u << 30, ENTRY_TYPE = 3u << 30 };

struct SceneView {
    const WideNode* nodes;        // TLAS nodes (global)
    const WideNode* topNodes;     // shared memory copy of the first nTop nodes, or null
    uint32_t nTop;
    const InstanceRec* instances;
};

TDEV SceneView makeSceneView(const void* tlas, const WideNode* sharedTop, uint32_t nTop) {
    const TlasHeader* h = reinterpret_cast<const TlasHeader*>(tlas);
    const char* base = reinterpret_cast<const char*>(tlas);
    SceneView v;
    v.nodes = reinterpret_cast<const WideNode*>(base + h->nodesOffset);
    v.instances = reinterpret_cast<const InstanceRec*>(base + h->instancesOffset);
    v.topNodes = sharedTop;
    v.nTop = sharedTop ? min(nTop, h->nNodes) : 0;
    return v;
}
// Cooperative copy of the top of the TLAS into shared memory (call with the whole CTA).
TDEV uint32_t stageTopNodes(const void* tlas, WideNode* sharedTop, uint32_t cap = THEIA_TOP_NODES) {
    const TlasHeader* h = reinterpret_cast<const TlasHeader*>(tlas);
    const uint4* src = reinterpret_cast<const uint4*>(reinterpret_cast<const char*>(tlas) + h->nodesOffset);
    uint32_t n = min(h->nNodes, cap);
    uint4* dst = reinterpret_cast<uint4*>(sharedTop);
    for (uint32_t i = threadIdx.x; i < n * 5; i += blockDim.x) dst[i] = __ldg(src + i);
    return n;
}

// Slot of the one instance of the target, or ~0 (several, none, or no filter): the first pass of a filtered ray
// query starts at that instance instead of the root of the TLAS.
TDEV uint32_t singleTargetSlot(const float* __restrict__ targetBox, int32_t targetId) {
    if (targetId < 0 || targetBox == nullptr || __ldg(targetBox + 6) != 1.0f) return 0xFFFFFFFFu;
    return f2u(__ldg(targetBox + 7));
}

// Closest hit over all instances. Per node: triangles of hit leaf children are tested right
// away, inner children (and instances) are pushed far-to-near with their entry distance, so
// entries that fall behind a closer hit are dropped when popped.
// targetId >= 0 (shadow rays, see wfTraceQueuedKernel): two passes, first the instances with this custom id alone,
// then, if one was hit, the others up to that hit; exact whenever the closest hit is on such an instance.
template <bool STATS>
TDEV RayHit closestHit(const SceneView& sc, vec3 origin, vec3 dir, float tmax, TraversalStats* st, int32_t targetId = -1,
                       uint32_t onlySlot = 0xFFFFFFFFu /* slot of the target's one instance: the first pass starts there */) {
    RayHit best;
    best.valid = false; best.instance = -1; best.prim = -1; best.slot = -1; best.t = 0.0f; best.u = 0.0f; best.v = 0.0f;

    uint32_t stackKey[THEIA_STACK_SIZE];
    float stackT[THEIA_STACK_SIZE];
    int sp = 0;
    stackKey[0] = onlySlot != 0xFFFFFFFFu ? (ENTRY_INST | onlySlot) : (ENTRY_TLAS | 0u);
    stackT[0] = 0.0f;
    sp = 1;

    const SlabRay world = {origin.x, origin.y, origin.z, safeInv(dir.x), safeInv(dir.y), safeInv(dir.z)};
    SlabRay obj = world;
    WoopRay wr;
    int32_t curInst = -1, curSlot = -1;
    const WideNode* blasNodes = nullptr;
    const TriRec* blasTris = nullptr;

    bool second = false;
    while (true) {
        if (sp == 0) {
            if (targetId < 0 || second || !best.valid) break;
            second = true;  // a hit on the target: is anything of the rest of the scene in front of it?
            stackKey[0] = ENTRY_TLAS | 0u;
            stackT[0] = 0.0f;
            sp = 1;
        }
        --sp;
        const uint32_t key = stackKey[sp];
        const float tEntry = stackT[sp];
        const float lim = best.valid ? best.t : tmax;
        if (tEntry > lim) continue;
        const uint32_t type = key & ENTRY_TYPE;
        const uint32_t index = key & ~ENTRY_TYPE;

        if (type == ENTRY_INST) {
            // enter the instance: object space ray, watertight setup
            const InstanceRec* in = sc.instances + index;
            if (targetId >= 0 && ((__ldg(&in->customIndex) == targetId) == second)) continue;
            const float4* q = reinterpret_cast<const float4*>(in);
            float4 r0 = __ldg(q), r1 = __ldg(q + 1), r2 = __ldg(q + 2);
            const float m[12] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w, r2.x, r2.y, r2.z, r2.w};
            const float o[3] = {origin.x, origin.y, origin.z}, d[3] = {dir.x, dir.y, dir.z};
            transformPoint(m, o, wr.o);
            transformVector(m, d, wr.d);
            woopSetup(wr);
            obj = {wr.o[0], wr.o[1], wr.o[2], safeInv(wr.d[0]), safeInv(wr.d[1]), safeInv(wr.d[2])};
            const uint4 ptrs = __ldg(reinterpret_cast<const uint4*>(in) + 3);
            blasNodes = reinterpret_cast<const WideNode*>((uint64_t)ptrs.x | ((uint64_t)ptrs.y << 32));
            blasTris = reinterpret_cast<const TriRec*>((uint64_t)ptrs.z | ((uint64_t)ptrs.w << 32));
            curSlot = (int32_t)index;
            curInst = (int32_t)__ldg(&in->pad);
            if (STATS) st->instances++;
            stackKey[sp] = ENTRY_BLAS | 0u;
            stackT[sp] = tEntry;
            ++sp;
            continue;
        }

        const bool inBlas = type == ENTRY_BLAS;
        NodeRegs n;
        if (inBlas) n = loadNode(blasNodes + index);
        else if (index < sc.nTop) n = loadNodeShared(sc.topNodes + index);
        else n = loadNode(sc.nodes + index);
        if (STATS) st->nodes++;

        float tmin[8];
        uint32_t mask = testChildren(n, inBlas ? obj : world, lim, tmin);
        if (mask == 0) continue;
        const uint32_t imask = n.a.w >> 24;

        // ---- leaves first (BLAS: triangles now; TLAS: instances are pushed like inner nodes), nearest
        // leaf first (slot order along the node's widest axis), and a leaf whose box starts behind the
        // closest hit so far is skipped
        uint32_t leafMask = mask & ~imask;
        if (inBlas && leafMask) {
            const uint32_t axisL = (n.b.x >> 30);
            const float idirL = axisL == 0 ? obj.ix : (axisL == 1 ? obj.iy : obj.iz);
            const bool fwdL = idirL >= 0.0f;
            while (leafMask) {
                const int i = fwdL ? (__ffs(leafMask) - 1) : (31 - __clz(leafMask));
                leafMask &= ~(1u << i);
                if (best.valid && tmin[i] > best.t) continue;
                const uint32_t meta = byteOf(i < 4 ? n.b.z : n.b.w, i & 3);
                const uint32_t count = meta & 7u, first = n.b.y + (meta >> 3);
                for (uint32_t k = 0; k < count; ++k) {
                    const float4* tp = reinterpret_cast<const float4*>(blasTris + first + k);
                    float4 a = __ldg(tp), b = __ldg(tp + 1), c = __ldg(tp + 2);
                    const float v0[3] = {a.x, a.y, a.z}, v1[3] = {b.x, b.y, b.z}, v2[3] = {c.x, c.y, c.z};
                    const int32_t prim = (int32_t)f2u(a.w);
                    float t, u, v;
                    if (STATS) st->tris++;
                    if (!woopIntersect(wr, v0, v1, v2, tmax, &t, &u, &v)) continue;
                    bool better = !best.valid || t < best.t ||
                                  (t == best.t && (curInst < best.instance || (curInst == best.instance && prim < best.prim)));
                    if (better) { best.valid = true; best.instance = curInst; best.slot = curSlot; best.prim = prim; best.t = t; best.u = u; best.v = v; }
                }
            }
        }
        // ---- push inner children (and TLAS leaves = instances), farthest first.
        // Children are stored sorted along the node's widest axis: scanning the mask downwards
        // for a positive direction pushes far slots first.
        uint32_t pushMask = inBlas ? (mask & imask) : mask;
        const float lim2 = best.valid ? best.t : tmax;
        const uint32_t axis = (n.b.x >> 30);  // widest axis of the node, kept in the top bits of childBase
        const uint32_t childBase = n.b.x & 0x3FFFFFFFu;
        const SlabRay& rr = inBlas ? obj : world;
        const float idirAxis = axis == 0 ? rr.ix : (axis == 1 ? rr.iy : rr.iz);
        const bool forward = idirAxis >= 0.0f;  // near slots have low indices
        while (pushMask) {
            // forward: far = high slot -> take the highest set bit first
            const int i = forward ? (31 - __clz(pushMask)) : (__ffs(pushMask) - 1);
            pushMask &= ~(1u << i);
            if (tmin[i] > lim2) continue;
            uint32_t k;
            if ((imask >> i) & 1u) {
                const uint32_t meta = byteOf(i < 4 ? n.b.z : n.b.w, i & 3);
                k = (inBlas ? ENTRY_BLAS : ENTRY_TLAS) | (childBase + meta - 1u);
            } else {
                const uint32_t meta = byteOf(i < 4 ? n.b.z : n.b.w, i & 3);
                k = ENTRY_INST | (n.b.y + (meta >> 3));
            }
            if (sp < THEIA_STACK_SIZE) {
                stackKey[sp] = k;
                stackT[sp] = tmin[i];
                ++sp;
            }
        }
    }
    return best;
}

}  // namespace theia
