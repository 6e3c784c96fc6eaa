// Ray queries with a warp level queue of triangle tests ("queued" trace kernel).
//
// The two older forms of the ray query kernel (wavefront.cuh) keep a ray's triangle tests in the
// lane that owns the ray, so a warp runs them at 3-7 of 32 lanes (profiles/r1i_wftrace_c2_lines.txt:
// 30 % of the instructions of the kernel), and a lane that met a leaf holds up - or drops out of -
// the node tests of the others. Here the two kinds of work are decoupled:
//
//   * a lane owns a ray and walks its stack: per round ONE node step (pop, slab test of the eight
//     children, push). Leaves are not tested by the owner: they go as items (owner lane, number of
//     triangles, first triangle) into a ring shared by the warp, and the lane carries on with its stack;
//   * whenever the ring holds a full warp of items, they are expanded into triangle slots and all 32 lanes
//     test one triangle each, whichever ray it belongs to: the owner's sheared ray (watertight test,
//     bvh.cuh) lives in shared memory,
//     the result goes back with one 64 bit shared-memory atomicMin per hit on
//     (t bits << 32 | instance * keyStride + primitive), which is exactly the closest-hit rule of
//     the ray query (ties in t to the lower (instance, primitive)), in any order;
//   * a lane whose stack ran empty waits for its queued items, writes the result and takes the next
//     ray of the warp's chunk (rays are handed out in chunks of <= 32 from a global cursor).
//
// The closest hit does not depend on the order of the tests, so deferring them changes no result;
// the only cost is that `lim` (the closest hit so far) is a little stale for the node tests.
// A lane may only enter the next instance once its items of the current one are done (the sheared
// ray is per instance); such lanes, and lanes waiting for their last items, trigger a flush of a
// partially filled ring once there are `flushWaiting` of them.
// One warp per CTA (the queue sits at a fixed shared memory address), 28 CTAs per SM; the bottom of every lane's stack
// lives in shared memory, too. Shadow rays are traced against the instances of the target first (targetId).
#pragma once
#include "traverse.cuh"
#include "launch.cuh"

namespace theia {

#define WQ_CAP 512u  // ring entries per warp (power of two): < 32 left by a flush + at most 8 per lane and node step
#ifndef WQ_WARPS
// Warps per CTA of the queued kernel. One: the warp's queue sits at a fixed shared memory address, so none of its many
// accesses needs address arithmetic or a base register (C3 shadow trace 1.159 -> 1.048 ms, C2 0.405 -> 0.367 ms against
// four warps per CTA at the same number of warps per SM); the top of the TLAS is not staged in shared memory then.
#define WQ_WARPS 1u
#endif
#define WQ_TOP_NODES (WQ_WARPS == 1u ? 0u : (uint32_t)THEIA_TOP_NODES)  // TLAS nodes staged in shared memory per CTA
// Entries of a lane's traversal stack that live in shared memory (the rest in local memory). The stack is read when a
// lane runs out of hit children, with the node fetch waiting on it, and local memory mostly missed L1 there (77 % of
// the sectors: the L1 is churned by nodes and triangles).
#ifndef WQ_SSTACK
#define WQ_SSTACK 6
#endif

struct WarpQueue {
    unsigned long long best[32];  // (t bits << 32) | tie key; low word 0xFFFFFFFF: no hit (t = tmax)
    float bu[32], bv[32];
    uint32_t bslot[32], bprim[32];
    float woop[12][32];           // mx[3], my[3], mz[3], cx, cy, cz of the owner's current instance
    float tmax[32];
    const TriRec* tris[32];
    uint32_t keyBase[32];         // instance id * keyStride
    uint32_t curSlot[32];
    uint32_t items[WQ_CAP];       // leaves: owner lane << 27 | (triangles - 1) << 25 | first triangle in the owner's BLAS
    uint8_t expand[128];          // triangle slot of a pass -> item lane << 2 | triangle within the leaf
    uint32_t tail;                // ring position behind the last item (free running; the head is warp uniform, in registers)
    uint32_t pad_;
    unsigned long long sstack[WQ_SSTACK > 0 ? WQ_SSTACK : 1][32];  // bottom of the lanes' stacks
};

// Tests the triangles of up to 32 queued leaves (n of them from `head`): the leaves are expanded into
// triangle slots so that every lane tests one triangle per pass, whichever leaf and ray it belongs to.
// Warp collective.
TDEV void wqProcess(WarpQueue& q, uint32_t head, uint32_t n) {
    const uint32_t lane = laneId();
    const uint32_t item = lane < n ? q.items[(head + lane) & (WQ_CAP - 1u)] : 0u;
    const uint32_t cnt = lane < n ? ((item >> 25) & 3u) + 1u : 0u;
    uint32_t incl = cnt;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t o = __shfl_up_sync(FULL_MASK, incl, d);
        if (lane >= (uint32_t)d) incl += o;
    }
    const uint32_t total = __shfl_sync(FULL_MASK, incl, 31);
    for (uint32_t k = 0; k < cnt; ++k) q.expand[incl - cnt + k] = (uint8_t)((lane << 2) | k);
    __syncwarp();
    for (uint32_t base = 0; base < total; base += 32u) {
        const uint32_t s = base + lane;
        const uint32_t e = s < total ? q.expand[s] : 0u;
        const uint32_t it = __shfl_sync(FULL_MASK, item, e >> 2);
        bool hit = false;
        unsigned long long key = 0ull;
        uint32_t owner = it >> 27, prim = 0;
        float t = 0.0f, u = 0.0f, v = 0.0f;
        if (s < total) {
            const float4* tp = reinterpret_cast<const float4*>(q.tris[owner] + ((it & 0x01FFFFFFu) + (e & 3u)));
            const float4 a = __ldg(tp), b = __ldg(tp + 1), c = __ldg(tp + 2);
            WoopRay wr;
            wr.mx[0] = q.woop[0][owner]; wr.mx[1] = q.woop[1][owner]; wr.mx[2] = q.woop[2][owner];
            wr.my[0] = q.woop[3][owner]; wr.my[1] = q.woop[4][owner]; wr.my[2] = q.woop[5][owner];
            wr.mz[0] = q.woop[6][owner]; wr.mz[1] = q.woop[7][owner]; wr.mz[2] = q.woop[8][owner];
            wr.cx = q.woop[9][owner]; wr.cy = q.woop[10][owner]; wr.cz = q.woop[11][owner];
            const float v0[3] = {a.x, a.y, a.z}, v1[3] = {b.x, b.y, b.z}, v2[3] = {c.x, c.y, c.z};
            prim = f2u(a.w);
            hit = woopIntersect(wr, v0, v1, v2, q.tmax[owner], &t, &u, &v);
            if (hit) {
                key = ((unsigned long long)f2u(t) << 32) | (unsigned long long)(q.keyBase[owner] + prim);
                atomicMin(&q.best[owner], key);
            }
        }
        __syncwarp();
        if (hit && q.best[owner] == key) {
            q.bu[owner] = u; q.bv[owner] = v; q.bslot[owner] = q.curSlot[owner]; q.bprim[owner] = prim;
        }
        __syncwarp();
    }
}

#ifndef THEIA_QUEUED_CTAS
#define THEIA_QUEUED_CTAS (28u / WQ_WARPS)  // 28 warps per SM (72 registers)
#endif
enum : uint32_t { WT_NONE = 0u, WT_NODE = 1u, WT_INST = 2u };
// rays[2 r], rays[2 r + 1]: ray r < *countPtr; results[r] = (instance slot | ~0, primitive, u, v), or with indirect != 0
// results[k], k = the bits of rays[2 r + 1].w (shadow rays).
// targetId >= 0 (shadow rays: only a closest hit on an instance with this custom id matters, processShadowRay): a ray
// is first traced against those instances alone and, if it hits one, against the others up to that hit, which may
// replace it; the result is the closest hit of the scene whenever that is on such an instance, and unspecified
// (a miss, or a hit elsewhere) when it is not.
// keyStride (TlasHeader::pad[0]) == 0: (instance, primitive) does not fit the tie key: one thread per ray instead.
// thresholds = flushWaiting | instLanes << 8 | refillLanes << 16: a kind of work runs once that many lanes want it
// (or nothing else is left to do): run for a few lanes it costs the warp as many issue slots as for all of them.
__global__ void __launch_bounds__(32 * WQ_WARPS, THEIA_QUEUED_CTAS)
wfTraceQueuedKernel(const void* __restrict__ tlas, const float4* __restrict__ rays, uint4* __restrict__ results,
                    const uint32_t* __restrict__ countPtr, uint32_t* cursor, uint32_t indirect, uint32_t thresholds,
                    int32_t targetId, const float* __restrict__ targetBox) {
#if WQ_WARPS == 1
    WideNode* sharedTop = nullptr;
    const uint32_t nTop = 0u;
#else
    __shared__ float4 sharedRaw[WQ_TOP_NODES * 5];
    WideNode* sharedTop = reinterpret_cast<WideNode*>(sharedRaw);
    const uint32_t nTop = stageTopNodes(tlas, sharedTop, WQ_TOP_NODES);
#endif
    __shared__ WarpQueue queues[WQ_WARPS];
    const uint32_t lane = laneId();
    WarpQueue& q = queues[threadIdx.x >> 5];
    if (lane == 0) q.tail = 0u;
    __syncthreads();
    const SceneView sc = makeSceneView(tlas, sharedTop, nTop);
    const uint32_t keyStride = reinterpret_cast<const TlasHeader*>(tlas)->pad[0];
    const uint32_t total = *countPtr;
    if (keyStride == 0u) {
        for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
            const float4 a = loadStream(rays + 2 * (size_t)i), b = loadStream(rays + 2 * (size_t)i + 1);
            const uint32_t r = indirect ? f2u(b.w) : i;
            const RayHit h = closestHit<false>(sc, vec3{a.x, a.y, a.z}, vec3{b.x, b.y, b.z}, a.w, nullptr, targetId,
                                               singleTargetSlot(targetBox, targetId));
            storeStream(results + r, make_uint4(h.valid ? (uint32_t)h.slot : 0xFFFFFFFFu, (uint32_t)h.prim, f2u(h.u), f2u(h.v)));
        }
        return;
    }
    const uint32_t flushWaiting = thresholds & 0xFFu, instLanes = (thresholds >> 8) & 0xFFu, refillLanes = (thresholds >> 16) & 0xFFu;
    const uint32_t nWarps = gridDim.x * WQ_WARPS;
    // rays per grab of the warp: a full warp while there is plenty, less for short queues (tail balance)
    uint32_t chunk = total / (2u * nWarps);
    chunk = chunk > 32u ? 32u : (chunk < 1u ? 1u : chunk);

    // entry distance bits << 32 | key: one load / store per entry; entries 0 .. WQ_SSTACK - 1 in shared memory
    unsigned long long stack[THEIA_STACK_SIZE - WQ_SSTACK];
    int sp = 0;
    bool have = false;
    uint32_t task = WT_NONE, payload = 0u;  // the entry this lane works on next (popped, not yet done)
    bool taskBlas = false;
    uint32_t rid = 0, pendingEnd = 0u;
    uint32_t wNext = 0u, wEnd = 0u;  // the warp's chunk of rays (uniform)
    bool exhausted = false;          // the cursor ran past the last ray (uniform)
    float ox = 0, oy = 0, oz = 0, dx = 0, dy = 0, dz = 0;
    SlabRay world = {0, 0, 0, 0, 0, 0}, obj = {0, 0, 0, 0, 0, 0};
    const WideNode* blasNodes = nullptr;
    uint32_t head = 0u;  // ring position of the oldest item (uniform)
    bool idleRound = true;  // no node work in the last round (uniform): everything that waits may go ahead
    const bool filtered = targetId >= 0;
    const uint32_t onlySlot = singleTargetSlot(targetBox, targetId);  // the target's one instance: first passes start there
    bool second = false;  // filtered: the lane's ray is on its second pass (instances that do not belong to the target)
    // an instance entry the lane's pass leaves out?
    auto leftOut = [&](uint32_t slot) { return filtered && ((__ldg(&sc.instances[slot].customIndex) == targetId) == second); };

    while (true) {
        // ---- results and refill
        bool clear = (int32_t)(head - pendingEnd) >= 0;
        if (have && task == WT_NONE && sp == 0 && clear && filtered && !second && (uint32_t)q.best[lane] != 0xFFFFFFFFu) {
            second = true;  // a hit on the target: is anything of the rest of the scene in front of it?
            task = WT_NODE; payload = 0u; taskBlas = false;
        }
        if (have && task == WT_NONE && sp == 0 && clear) {
            const unsigned long long b = q.best[lane];
            const bool valid = (uint32_t)b != 0xFFFFFFFFu;
            storeStream(results + rid, make_uint4(valid ? q.bslot[lane] : 0xFFFFFFFFu, valid ? q.bprim[lane] : 0xFFFFFFFFu,
                                      f2u(valid ? q.bu[lane] : 0.0f), f2u(valid ? q.bv[lane] : 0.0f)));
            have = false;
        }
        const uint32_t needMask = __ballot_sync(FULL_MASK, !have);
        if (!exhausted && needMask && ((uint32_t)__popc(needMask) >= refillLanes || idleRound)) {
            if (wNext == wEnd) {
                uint32_t base = 0u;
                if (lane == 0) base = atomicAdd(cursor, chunk);
                base = __shfl_sync(FULL_MASK, base, 0);
                if (base >= total) { exhausted = true; wNext = wEnd = 0u; }
                else { wNext = base; wEnd = min(base + chunk, total); }
            }
            const uint32_t i = wNext + (uint32_t)__popc(needMask & laneMaskLt());
            if (!have && i < wEnd) {
                const float4 a = loadStream(rays + 2 * (size_t)i), b = loadStream(rays + 2 * (size_t)i + 1);
                rid = indirect ? f2u(b.w) : i;
                ox = a.x; oy = a.y; oz = a.z; dx = b.x; dy = b.y; dz = b.z;
                world = {ox, oy, oz, safeInv(dx), safeInv(dy), safeInv(dz)};
                q.best[lane] = ((unsigned long long)f2u(a.w) << 32) | 0xFFFFFFFFull;
                q.tmax[lane] = a.w;
                sp = 0;
                task = WT_NODE; payload = 0u; taskBlas = false;  // root of the TLAS
                if (onlySlot != 0xFFFFFFFFu) { task = WT_INST; payload = onlySlot; }
                second = false;
                have = true;
                pendingEnd = head;
                clear = true;
            }
            wNext = min(wEnd, wNext + (uint32_t)__popc(needMask));
        }
        if (exhausted && __all_sync(FULL_MASK, !have)) break;

        // ---- next entry of the stack that is still in front of the closest hit
        const float lim = u2f((uint32_t)(q.best[lane] >> 32));
        if (task == WT_NONE) {
            while (sp > 0) {
                --sp;
                const unsigned long long e = sp < WQ_SSTACK ? q.sstack[sp][lane] : stack[sp - WQ_SSTACK];
                if (u2f((uint32_t)(e >> 32)) > lim) continue;
                const uint32_t key = (uint32_t)e;
                payload = key & 0x3FFFFFFFu;
                const uint32_t type = key >> 30;
                if (type == WE_INST && leftOut(payload)) continue;
                task = type == WE_INST ? WT_INST : WT_NODE;
                taskBlas = type == WE_BLAS;
                break;
            }
        }
        // ---- instance entry: object space ray, sheared ray of the triangle tests (needs the lane's queued
        // triangles of the previous instance done)
        const uint32_t nodeMask0 = __ballot_sync(FULL_MASK, task == WT_NODE);
        const bool wantInst = task == WT_INST && clear;
        const uint32_t instMask = __ballot_sync(FULL_MASK, wantInst);
        if (instMask && ((uint32_t)__popc(instMask) >= instLanes || nodeMask0 == 0u)) {
            if (wantInst) {
                const InstanceRec* in = sc.instances + payload;
                const float4* qp = reinterpret_cast<const float4*>(in);
                const float4 r0 = __ldg(qp), r1 = __ldg(qp + 1), r2 = __ldg(qp + 2);
                const float m[12] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w, r2.x, r2.y, r2.z, r2.w};
                const float o[3] = {ox, oy, oz}, d[3] = {dx, dy, dz};
                WoopRay wr;
                transformPoint(m, o, wr.o);
                transformVector(m, d, wr.d);
                woopSetup(wr);
                obj = {wr.o[0], wr.o[1], wr.o[2], safeInv(wr.d[0]), safeInv(wr.d[1]), safeInv(wr.d[2])};
                const uint4 ptrs = __ldg(reinterpret_cast<const uint4*>(in) + 3);
                blasNodes = reinterpret_cast<const WideNode*>((uint64_t)ptrs.x | ((uint64_t)ptrs.y << 32));
                q.tris[lane] = reinterpret_cast<const TriRec*>((uint64_t)ptrs.z | ((uint64_t)ptrs.w << 32));
                q.woop[0][lane] = wr.mx[0]; q.woop[1][lane] = wr.mx[1]; q.woop[2][lane] = wr.mx[2];
                q.woop[3][lane] = wr.my[0]; q.woop[4][lane] = wr.my[1]; q.woop[5][lane] = wr.my[2];
                q.woop[6][lane] = wr.mz[0]; q.woop[7][lane] = wr.mz[1]; q.woop[8][lane] = wr.mz[2];
                q.woop[9][lane] = wr.cx; q.woop[10][lane] = wr.cy; q.woop[11][lane] = wr.cz;
                q.keyBase[lane] = __ldg(&in->pad) * keyStride;
                q.curSlot[lane] = payload;
                task = WT_NODE;  // straight on to the root of the BLAS
                taskBlas = true;
                payload = 0u;
            }
        }
        // ---- node step: slab test of the eight children; inner children (TLAS: instances) go to the stack,
        // leaves to the ring of triangle tests
        const uint32_t nodeMask = __ballot_sync(FULL_MASK, task == WT_NODE);
        idleRound = nodeMask == 0u;
        if (task == WT_NODE) {
            task = WT_NONE;
            const bool inBlas = taskBlas;
            const WideNode* np = inBlas ? blasNodes + payload
                                        : (payload < sc.nTop ? sc.topNodes + payload : sc.nodes + payload);
#if WQ_WARPS == 1 && !defined(WQ_GENERIC_NODE_LOAD)
            const NodeRegs n = loadNode(np);  // always global memory: the read-only path
#else
            const NodeRegs n = loadNodeShared(np);
#endif
            float tmin[8];
            uint32_t mask = testChildren(n, inBlas ? obj : world, lim, tmin);
            if (mask) {
                const uint32_t imask = n.a.w >> 24;
                const uint32_t axis = n.b.x >> 30, childBase = n.b.x & 0x3FFFFFFFu;
                const SlabRay& rr = inBlas ? obj : world;
                const float idirAxis = axis == 0 ? rr.ix : (axis == 1 ? rr.iy : rr.iz);
                const bool forward = idirAxis >= 0.0f;
                uint32_t leafMask = inBlas ? (mask & ~imask) : 0u;
                mask &= ~leafMask;
                if (leafMask) {
                    uint32_t pos = atomicAdd(&q.tail, (uint32_t)__popc(leafMask));
                    pendingEnd = pos + (uint32_t)__popc(leafMask);
                    while (leafMask) {
                        const int i = __ffs(leafMask) - 1;
                        leafMask &= leafMask - 1u;
                        const uint32_t meta = metaByte(n, i);
                        q.items[pos & (WQ_CAP - 1u)] = (lane << 27) | (((meta & 7u) - 1u) << 25) | (n.b.y + (meta >> 3));
                        ++pos;
                    }
                }
                if (mask) {
                    // inner children / instances that were hit. The nearest one in slot order along the node's axis is
                    // the lane's next entry and stays in registers; the others go to the stack, nearer ones on top. The
                    // loop runs as often as the lane of the warp with the most of them needs (two or three times as a
                    // rule), so an iteration has to be short: the entry distance comes out of the eight registers by
                    // selects (an array indexed by i would live in local memory), the meta byte by one byte permute.
                    const uint32_t innerKey = ((inBlas ? WE_BLAS : WE_TLAS) << 30) | (childBase - 1u);
                    const uint32_t instKey = (WE_INST << 30) | n.b.y;
                    const int i0 = forward ? (__ffs(mask) - 1) : (31 - __clz(mask));
                    const uint32_t meta0 = metaByte(n, i0);
                    if ((imask >> i0) & 1u) { payload = childBase + meta0 - 1u; task = WT_NODE; taskBlas = inBlas; }
                    else { payload = n.b.y + (meta0 >> 3); task = leftOut(payload) ? WT_NONE : WT_INST; taskBlas = false; }
                    uint32_t rest = mask & ~(1u << i0);
                    // (the stack cannot overflow: theia_tlas_build bounds the depth of the trees to its size)
                    while (rest) {
                        const int i = forward ? (31 - __clz(rest)) : (__ffs(rest) - 1);  // farthest slot first
                        rest &= ~(1u << i);
                        const uint32_t meta = metaByte(n, i);
                        const uint32_t k = ((imask >> i) & 1u) ? innerKey + meta : instKey + (meta >> 3);
                        const unsigned long long e = ((unsigned long long)f2u(pick8(tmin, i)) << 32) | k;
                        if (sp < WQ_SSTACK) q.sstack[sp][lane] = e;
                        else stack[sp - WQ_SSTACK] = e;
                        ++sp;
                    }
                }
            }
        }

        // ---- triangle tests: full warps of leaves; everything once enough lanes wait for theirs
        __syncwarp();
        uint32_t nItems = q.tail - head;
        if (nItems) {
            const bool waiting = have && (int32_t)(head - pendingEnd) < 0 && (task == WT_INST || (task == WT_NONE && sp == 0));
            const uint32_t waitMask = __ballot_sync(FULL_MASK, waiting);
            const bool flushAll = (uint32_t)__popc(waitMask) >= flushWaiting || idleRound;
            while (nItems >= 32u || (flushAll && nItems > 0u)) {
                const uint32_t m = nItems < 32u ? nItems : 32u;
                wqProcess(q, head, m);
                head += m;
                nItems -= m;
            }
        }
    }
}

}  // namespace theia
