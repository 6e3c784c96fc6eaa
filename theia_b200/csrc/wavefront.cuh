// Wavefront form of the SceneForwardTracer (sh/tracer.scene.forward.glsl +
// sh/scene.traverse.glsl of the reference).
//
// The path state lives in HBM as structure-of-arrays and each path segment is split into kernels.
// Live paths are kept DENSE: a kernel that keeps a path alive allocates the next free index of
// the output buffer (one atomic per warp: ballot + prefix popcount) and writes state and ray
// query there, so every kernel reads and writes contiguous indices 0 .. count-1 (coalesced,
// every sector fully used) and there is no queue of slot numbers to gather through. The path
// number (RNG stream, hit tags, track index) travels as one more state word. State and ray
// buffers ping-pong between segments:
//
//   generate : sample wavelength + light, create the ray, prepare the first ray query
//   per segment:
//     trace  : ray queries only (BVH traversal + triangle tests). Persistent warps; a lane
//              that finishes its ray immediately fetches the next one (warp-aggregated
//              counter), so lanes never wait for the longest traversal of their warp.
//     shade  : ray query result -> SurfaceHit, propagation, surface interaction (Fresnel
//              reflect / transmit / detect) or volume scatter, responses, event callbacks,
//              then the preparation of the next ray query, appended to the other buffer.
//     with a target guide, volume scatter events emit two shadow rays instead (phase sample,
//     guide sample); a second trace + shade pass turns them into responses, performs the
//     deferred phase function scatter and prepares the next query.
//
// Per path the sequence of arithmetic operations and of random draws is exactly that of the
// monolithic kernel (trace_scene.cuh) and of the reference shader, so results are identical;
// only the order in which hits reach the response differs.
#pragma once
#include "trace_scene.cuh"
#include <cooperative_groups.h>

namespace theia {

// ------------------------------------------------------------------ workspace
enum : uint32_t {
    WF_PX, WF_PY, WF_PZ, WF_DX, WF_DY, WF_DZ, WF_MED_LO, WF_MED_HI, WF_LAMBDA, WF_TIME, WF_LIN, WF_LOG,
    WF_N, WF_VG, WF_MUS, WF_MUE, WF_DIM, WF_FLAGS, WF_PATH, WF_NFIELDS_UNPOL,
    WF_S0 = WF_NFIELDS_UNPOL, WF_S1, WF_S2, WF_S3, WF_R0, WF_R1, WF_R2, WF_NFIELDS_POL
};
// WF_FLAGS: bits 0-15 segment number "it", bit 16 allowResponse, bit 17 mis_target
enum : uint32_t { WF_ALLOW = 1u << 16, WF_MIS_TARGET = 1u << 17 };
// Counters: one slot of four words per segment generation s = 0 .. pathLength, cleared ONCE per batch (a memset), so
// no launch in between has to reset anything: [0] live entries produced by generation s (s = 0: the generate kernel),
// [1] paths of generation s waiting for shadow rays, [2] / [3] work cursors of the primary / shadow ray query launch,
// [4] shadow rays of generation s that have to be traced (see targetBox).
enum : uint32_t { WF_CNT_QUEUE = 0, WF_CNT_SHADOW = 1, WF_CNT_CURSOR = 2, WF_CNT_SHADOW_CURSOR = 3, WF_CNT_SHADOW_RAYS = 4,
                  WF_SLOT = 8 };

struct Workspace {
    uint32_t capacity;      // paths per batch
    uint32_t* counters;     // [pathLength + 1][WF_SLOT], see above
    // the counters of the launch at hand (set by the host per launch): entries to read, entries produced, shadow entries
    const uint32_t* cntIn;
    uint32_t* cntOut;
    uint32_t* cntShadow;
    uint32_t* cntShadowRays;
    uint32_t* state[2];     // ping-pong, [nFields][capacity] words, dense
    float4* ray[2];         // ping-pong, [capacity][2]: (origin, tmax), (direction, sampledDist)
    uint4* result;          // [capacity]: instance slot (0xFFFFFFFF = miss), primitive, u, v
    uint32_t* shadowState;  // [nFields][capacity]: paths with a pending pair of shadow rays, dense
    // the two shadow rays of a parked path: plane [0] phase function sample, plane [1] guide sample
    float4* shadowDir;      // [2][capacity]: (direction, weight)
    uint4* shadowResult;    // [2][capacity], written for the rays that were traced (WF_LIVE0 / WF_LIVE1 of the path)
    // Only a shadow ray whose closest hit is a target creates a response (processShadowRay), so one that misses the
    // box around all instances the target id selects is not traced at all. The others, dense:
    float4* shadowLive;     // [2 capacity][2]: (origin, tmax), (direction, plane * capacity + slot as bits)
    float* targetBox;       // [8]: lo.xyz, hi.xyz of that box (padded), computed per batch by wfTargetBoxKernel
    uint32_t cullShadow;    // 0: every shadow ray is traced (option wf_shadow_cull, for A/B runs and tests)
};

static inline size_t wfAlign(size_t v) { return (v + 255) & ~size_t(255); }

// Carves the workspace out of a caller supplied buffer (base may be null to get the size).
static inline size_t layoutWorkspace(void* base, uint32_t capacity, bool polarized, bool guide, uint32_t pathLength, Workspace* ws) {
    size_t off = 0;
    char* b = (char*)base;
    auto take = [&](size_t bytes) { void* p = b ? b + off : nullptr; off += wfAlign(bytes); return p; };
    const size_t C = capacity > 0 ? capacity : 1;
    const size_t nFields = polarized ? WF_NFIELDS_POL : WF_NFIELDS_UNPOL;
    Workspace w;
    w.capacity = capacity;
    w.counters = (uint32_t*)take(4 * WF_SLOT * ((size_t)pathLength + 1));
    w.cntIn = nullptr; w.cntOut = nullptr; w.cntShadow = nullptr; w.cntShadowRays = nullptr;
    w.state[0] = (uint32_t*)take(4 * nFields * C);
    w.state[1] = (uint32_t*)take(4 * nFields * C);
    w.ray[0] = (float4*)take(32 * C);
    w.ray[1] = (float4*)take(32 * C);
    w.result = (uint4*)take(16 * C);
    if (guide) {
        w.shadowState = (uint32_t*)take(4 * nFields * C);
        w.shadowDir = (float4*)take(32 * C);
        w.shadowResult = (uint4*)take(32 * C);
        w.shadowLive = (float4*)take(64 * C);
        w.targetBox = (float*)take(32);
    } else {
        w.shadowState = nullptr; w.shadowDir = nullptr; w.shadowResult = nullptr; w.shadowLive = nullptr; w.targetBox = nullptr;
    }
    w.cullShadow = 1u;
    if (ws) *ws = w;
    return off;
}

// ------------------------------------------------------------------ state access
template <bool POL>
TDEV void storeState(uint32_t* state, size_t C, uint32_t slot, const Ray& r, uint32_t dim, uint32_t flags, uint32_t path) {
    uint32_t* s = state + slot;
    s[WF_PX * C] = f2u(r.position.x); s[WF_PY * C] = f2u(r.position.y); s[WF_PZ * C] = f2u(r.position.z);
    s[WF_DX * C] = f2u(r.direction.x); s[WF_DY * C] = f2u(r.direction.y); s[WF_DZ * C] = f2u(r.direction.z);
    s[WF_MED_LO * C] = (uint32_t)r.medium; s[WF_MED_HI * C] = (uint32_t)(r.medium >> 32);
    s[WF_LAMBDA * C] = f2u(r.wavelength); s[WF_TIME * C] = f2u(r.time);
    s[WF_LIN * C] = f2u(r.lin_contrib); s[WF_LOG * C] = f2u(r.log_contrib);
    s[WF_N * C] = f2u(r.constants.n); s[WF_VG * C] = f2u(r.constants.vg);
    s[WF_MUS * C] = f2u(r.constants.mu_s); s[WF_MUE * C] = f2u(r.constants.mu_e);
    s[WF_DIM * C] = dim; s[WF_FLAGS * C] = flags; s[WF_PATH * C] = path;
    if (POL) {
        s[WF_S0 * C] = f2u(r.stokes.x); s[WF_S1 * C] = f2u(r.stokes.y); s[WF_S2 * C] = f2u(r.stokes.z);
        s[WF_S3 * C] = f2u(r.stokes.w);
        s[WF_R0 * C] = f2u(r.polRef.x); s[WF_R1 * C] = f2u(r.polRef.y); s[WF_R2 * C] = f2u(r.polRef.z);
    }
}
template <bool POL>
TDEV void loadState(const uint32_t* state, size_t C, uint32_t slot, Ray& r, uint32_t& dim, uint32_t& flags, uint32_t& path) {
    const uint32_t* s = state + slot;
    r.position = {u2f(s[WF_PX * C]), u2f(s[WF_PY * C]), u2f(s[WF_PZ * C])};
    r.direction = {u2f(s[WF_DX * C]), u2f(s[WF_DY * C]), u2f(s[WF_DZ * C])};
    r.medium = (uint64_t)s[WF_MED_LO * C] | ((uint64_t)s[WF_MED_HI * C] << 32);
    r.wavelength = u2f(s[WF_LAMBDA * C]); r.time = u2f(s[WF_TIME * C]);
    r.lin_contrib = u2f(s[WF_LIN * C]); r.log_contrib = u2f(s[WF_LOG * C]);
    r.constants = {u2f(s[WF_N * C]), u2f(s[WF_VG * C]), u2f(s[WF_MUS * C]), u2f(s[WF_MUE * C])};
    dim = s[WF_DIM * C]; flags = s[WF_FLAGS * C]; path = s[WF_PATH * C];
    if (POL) {
        r.stokes = {u2f(s[WF_S0 * C]), u2f(s[WF_S1 * C]), u2f(s[WF_S2 * C]), u2f(s[WF_S3 * C])};
        r.polRef = {u2f(s[WF_R0 * C]), u2f(s[WF_R1 * C]), u2f(s[WF_R2 * C])};
    } else {
        r.stokes = {1.0f, 0.0f, 0.0f, 0.0f};
        r.polRef = v3(0.0f);
    }
}

// Hands the lanes with `push` consecutive indices of a dense buffer: one atomic per warp.
// MUST be called by all 32 lanes; the value returned to the other lanes is meaningless.
TDEV uint32_t denseAlloc(uint32_t* counter, bool push) {
    uint32_t m = __ballot_sync(FULL_MASK, push);
    if (m == 0) return 0u;
    uint32_t leader = __ffs(m) - 1, base = 0;
    if (laneId() == leader) base = atomicAdd(counter, (uint32_t)__popc(m));
    base = __shfl_sync(FULL_MASK, base, leader);
    return base + __popc(m & laneMaskLt());
}

// ------------------------------------------------------------------ shadow ray culling
// Can a ray reach the box (lo = box[0..2], hi = box[3..5])? Errs on the side of yes: the box is padded by what the
// coordinates can be off (wfTargetBoxKernel), and the distances get 2e-5 of slack for origins far away from it (the
// ray queries carry the same kind of error: their object space rays are rounded, too).
TDEV bool rayMayHitBox(const float* __restrict__ box, vec3 o, vec3 d, float tmax) {
    const float lx = __ldg(box), ly = __ldg(box + 1), lz = __ldg(box + 2);
    const float hx = __ldg(box + 3), hy = __ldg(box + 4), hz = __ldg(box + 5);
    if (!(lx <= hx)) return false;  // no instance belongs to the target
    const float ix = safeInv(d.x), iy = safeInv(d.y), iz = safeInv(d.z);
    const float x0 = (lx - o.x) * ix, x1 = (hx - o.x) * ix;
    const float y0 = (ly - o.y) * iy, y1 = (hy - o.y) * iy;
    const float z0 = (lz - o.z) * iz, z1 = (hz - o.z) * iz;
    const float tn = fmaxf(fmaxf(fminf(x0, x1), fminf(y0, y1)), fmaxf(fminf(z0, z1), 0.0f));
    const float tf = fminf(fminf(fmaxf(x0, x1), fmaxf(y0, y1)), fminf(fmaxf(z0, z1), tmax));
    return tn <= fmaf(2e-5f, tf, tf);
}

// World box around every instance a shadow ray may create a response on: custom id == targetId (any id if
// targetId < 0) and a material with the detector flag on one of its sides (processShadowRay,
// sh/scene.traverse.glsl:165-170). One CTA; box[6] = number of instances with that custom id, box[7] = (as bits) the
// slot of the instance record if there is exactly one (the ray queries start there: singleTargetSlot).
__global__ void __launch_bounds__(256) wfTargetBoxKernel(const void* __restrict__ tlas, int32_t targetId, float* __restrict__ box) {
    const TlasHeader* h = reinterpret_cast<const TlasHeader*>(tlas);
    const InstanceRec* inst = reinterpret_cast<const InstanceRec*>(reinterpret_cast<const char*>(tlas) + h->instancesOffset);
    float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    uint32_t count = 0, slots = 0;
    for (uint32_t i = threadIdx.x; i < h->nInstances; i += blockDim.x) {
        const InstanceRec& in = inst[i];
        if (targetId >= 0 && in.customIndex != targetId) continue;
        ++count; slots += i;
        if (in.material != 0) {
            const MaterialRec* m = reinterpret_cast<const MaterialRec*>(in.material);
            if (((m->flagsInwards | m->flagsOutwards) & MATERIAL_DETECTOR_BIT) == 0) continue;
        }
        const BlasHeader* b = reinterpret_cast<const BlasHeader*>(reinterpret_cast<const char*>(in.blasNodes) - sizeof(BlasHeader));
        float objMag = 0.0f, scale = 0.0f;
        for (int a = 0; a < 3; ++a) {
            objMag = fmaxf(objMag, fmaxf(fabsf(b->lo[a]), fabsf(b->hi[a])));
            scale = fmaxf(scale, fabsf(in.o2w[4 * a]) + fabsf(in.o2w[4 * a + 1]) + fabsf(in.o2w[4 * a + 2]));
        }
        float l[3] = {INFINITY, INFINITY, INFINITY}, u[3] = {-INFINITY, -INFINITY, -INFINITY};
        for (int c = 0; c < 8; ++c) {
            const float p[3] = {(c & 1) ? b->hi[0] : b->lo[0], (c & 2) ? b->hi[1] : b->lo[1], (c & 4) ? b->hi[2] : b->lo[2]};
            float w[3];
            transformPoint(in.o2w, p, w);
            for (int a = 0; a < 3; ++a) { l[a] = fminf(l[a], w[a]); u[a] = fmaxf(u[a], w[a]); }
        }
        for (int a = 0; a < 3; ++a) {
            // 1e-4 of everything that sets the size of a rounding error: world and object coordinates, extent
            const float pad = 1e-4f * (fmaxf(fabsf(l[a]), fabsf(u[a])) + (u[a] - l[a]) + scale * objMag) + 1e-30f;
            lo[a] = fminf(lo[a], l[a] - pad);
            hi[a] = fmaxf(hi[a], u[a] + pad);
        }
    }
    __shared__ float red[8][6];
    __shared__ uint32_t redCount[8], redSlots[8];
    for (int a = 0; a < 3; ++a)
        for (int d = 16; d > 0; d >>= 1) {
            lo[a] = fminf(lo[a], __shfl_xor_sync(FULL_MASK, lo[a], d));
            hi[a] = fmaxf(hi[a], __shfl_xor_sync(FULL_MASK, hi[a], d));
        }
    for (int d = 16; d > 0; d >>= 1) { count += __shfl_xor_sync(FULL_MASK, count, d); slots += __shfl_xor_sync(FULL_MASK, slots, d); }
    if (laneId() == 0) {
        for (int a = 0; a < 3; ++a) { red[threadIdx.x >> 5][a] = lo[a]; red[threadIdx.x >> 5][3 + a] = hi[a]; }
        redCount[threadIdx.x >> 5] = count; redSlots[threadIdx.x >> 5] = slots;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t total = 0, slot = 0;
        for (uint32_t w = 0; w < blockDim.x / 32; ++w) {
            total += redCount[w]; slot += redSlots[w];
            for (int a = 0; a < 3; ++a) { lo[a] = fminf(lo[a], red[w][a]); hi[a] = fmaxf(hi[a], red[w][3 + a]); }
        }
        for (int a = 0; a < 3; ++a) { box[a] = lo[a]; box[3 + a] = hi[a]; }
        box[6] = (float)total;
        box[7] = u2f(total == 1u ? slot : 0xFFFFFFFFu);
    }
}

// ------------------------------------------------------------------ trace kernel
// Ray queries of a dense buffer: rays[2 r], rays[2 r + 1] hold ray r < *countPtr; results[r] receives (instance slot | ~0,
// primitive, u, v). With indirect != 0 (shadow rays) the result goes to results[k] instead, k = the bits of the fourth
// component of rays[2 r + 1]. targetId >= 0: two passes per ray, the target's instances first (see wfTraceQueuedKernel).
//
// Persistent warps, three phases per round so that the lanes of a warp do the same kind of
// work at the same time (the traversal steps differ a lot in cost, and a mixed warp would
// execute them one after the other):
//   F  lanes whose stack ran empty write their result and fetch the next ray (private chunks
//      of WF_FETCH rays from a global cursor, one atomic per round and warp)
//   N  every lane pops nodes / instances until it meets a leaf or runs out of stack
//   L  lanes with a pending leaf test its triangles
// so a lane never waits for the longest traversal of its warp, only for the end of a phase.
#define WF_FETCH 4u
enum : uint32_t { WE_TLAS = 0u, WE_INST = 1u, WE_BLAS = 2u, WE_LEAF = 3u };

#ifndef THEIA_PHASED_CTAS
#define THEIA_PHASED_CTAS 5
#endif
__global__ void __launch_bounds__(128, THEIA_PHASED_CTAS) wfTraceKernel(const void* __restrict__ tlas, const float4* __restrict__ rays,
                                                         uint4* __restrict__ results,
                                                         const uint32_t* __restrict__ countPtr, uint32_t* cursor,
                                                         uint32_t indirect, uint32_t minNodeLanes, int32_t targetId,
                                                         const float* __restrict__ targetBox) {
    __shared__ float4 sharedRaw[THEIA_TOP_NODES * 5];
    WideNode* sharedTop = reinterpret_cast<WideNode*>(sharedRaw);
    const uint32_t nTop = stageTopNodes(tlas, sharedTop);
    __syncthreads();
    const SceneView sc = makeSceneView(tlas, sharedTop, nTop);
    const uint32_t total = *countPtr;
    // rays a lane takes per fetch: 4 while there is plenty of work (fewer atomics), 1 once the
    // queue is short, so that the tail is spread over all lanes instead of queued up in a few
    const uint32_t lanes = gridDim.x * blockDim.x;
    const uint32_t fetch = total / lanes >= 4u * WF_FETCH ? WF_FETCH : 1u;

    uint32_t stackKey[THEIA_STACK_SIZE];
    float stackT[THEIA_STACK_SIZE];
    int sp = 0;
    bool have = false, done = false;
    uint32_t chunkNext = 0, chunkEnd = 0, rid = 0;
    uint32_t pending = 0xFFFFFFFFu;  // leaf waiting for phase L
    float ox = 0, oy = 0, oz = 0, dx = 0, dy = 0, dz = 0, tmax = 0;
    SlabRay world = {0, 0, 0, 0, 0, 0}, obj = {0, 0, 0, 0, 0, 0};
    WoopRay wr;
    wr.cx = wr.cy = wr.cz = 0.0f;
    wr.mx[0] = wr.mx[1] = wr.mx[2] = wr.my[0] = wr.my[1] = wr.my[2] = wr.mz[0] = wr.mz[1] = wr.mz[2] = 0.0f;
    wr.o[0] = wr.o[1] = wr.o[2] = wr.d[0] = wr.d[1] = wr.d[2] = 0.0f;
    int32_t curInst = -1, curSlot = -1;
    const WideNode* blasNodes = nullptr;
    const TriRec* blasTris = nullptr;
    bool bestValid = false;
    int32_t bestInst = -1, bestSlot = -1, bestPrim = -1;
    float bestT = 0.0f, bestU = 0.0f, bestV = 0.0f;

    const uint32_t onlySlot = singleTargetSlot(targetBox, targetId);
    bool second = false;  // targetId >= 0: the ray is on its second pass (instances that do not belong to the target)
    while (true) {
        // ---- phase F
        if (sp == 0 && have && targetId >= 0 && !second && bestValid) {
            second = true;  // a hit on the target: is anything of the rest of the scene in front of it?
            stackKey[0] = WE_TLAS << 30;
            stackT[0] = 0.0f;
            sp = 1;
        }
        const bool idle = sp == 0 && !done;
        if (idle && have) {
            storeStream(results + rid, make_uint4(bestValid ? (uint32_t)bestSlot : 0xFFFFFFFFu, (uint32_t)bestPrim, f2u(bestU), f2u(bestV)));
            have = false;
        }
        const uint32_t needChunk = __ballot_sync(FULL_MASK, idle && chunkNext == chunkEnd);
        if (needChunk) {
            const uint32_t leader = __ffs(needChunk) - 1;
            uint32_t base = 0;
            if (laneId() == leader) base = atomicAdd(cursor, fetch * (uint32_t)__popc(needChunk));
            base = __shfl_sync(FULL_MASK, base, leader);
            if (idle && chunkNext == chunkEnd) {
                chunkNext = base + fetch * (uint32_t)__popc(needChunk & laneMaskLt());
                chunkEnd = chunkNext + fetch;
            }
        }
        if (idle) {
            const uint32_t i = chunkNext++;
            if (i >= total) {
                done = true;
            } else {
                const float4 a = loadStream(rays + 2 * (size_t)i), b = loadStream(rays + 2 * (size_t)i + 1);
                rid = indirect ? f2u(b.w) : i;
                ox = a.x; oy = a.y; oz = a.z; tmax = a.w; dx = b.x; dy = b.y; dz = b.z;
                world = {ox, oy, oz, safeInv(dx), safeInv(dy), safeInv(dz)};
                bestValid = false; bestInst = -1; bestSlot = -1; bestPrim = -1; bestT = 0.0f; bestU = 0.0f; bestV = 0.0f;
                stackKey[0] = onlySlot != 0xFFFFFFFFu ? ((WE_INST << 30) | onlySlot) : (WE_TLAS << 30);
                stackT[0] = 0.0f;
                sp = 1;
                second = false;
                have = true;
            }
        }
        if (__all_sync(FULL_MASK, done)) break;

        // ---- phase N (left early once only a few lanes are still looking for a leaf: they
        // resume in the next round instead of holding up the triangle tests of the others;
        // at most half of the lanes that entered may be dropped this way, or a warp with few
        // busy lanes would leave after every node)
        const uint32_t nodeLimit = min(minNodeLanes, ((uint32_t)__popc(__ballot_sync(FULL_MASK, sp > 0)) + 1u) >> 1);
        bool firstNode = true;  // the first iteration always runs: progress does not rest on what __activemask reports
        while (sp > 0) {
            if (!firstNode && (uint32_t)__popc(__activemask()) < nodeLimit) break;
            firstNode = false;
            // pop until an entry that is still in front of the closest hit
            const float lim = bestValid ? bestT : tmax;
            uint32_t key;
            float tEntry;
            do {
                --sp;
                key = stackKey[sp];
                tEntry = stackT[sp];
            } while (tEntry > lim && sp > 0);
            if (tEntry > lim) break;
            const uint32_t type = key >> 30, payload = key & 0x3FFFFFFFu;
            if (type == WE_LEAF) { pending = payload; break; }
            if (type == WE_INST) {
                const InstanceRec* in = sc.instances + payload;
                if (targetId >= 0 && ((__ldg(&in->customIndex) == targetId) == second)) continue;
                const float4* q = reinterpret_cast<const float4*>(in);
                const float4 r0 = __ldg(q), r1 = __ldg(q + 1), r2 = __ldg(q + 2);
                const float m[12] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w, r2.x, r2.y, r2.z, r2.w};
                const float o[3] = {ox, oy, oz}, d[3] = {dx, dy, dz};
                transformPoint(m, o, wr.o);
                transformVector(m, d, wr.d);
                woopSetup(wr);
                obj = {wr.o[0], wr.o[1], wr.o[2], safeInv(wr.d[0]), safeInv(wr.d[1]), safeInv(wr.d[2])};
                const uint4 ptrs = __ldg(reinterpret_cast<const uint4*>(in) + 3);
                blasNodes = reinterpret_cast<const WideNode*>((uint64_t)ptrs.x | ((uint64_t)ptrs.y << 32));
                blasTris = reinterpret_cast<const TriRec*>((uint64_t)ptrs.z | ((uint64_t)ptrs.w << 32));
                curSlot = (int32_t)payload;
                curInst = (int32_t)__ldg(&in->pad);
                stackKey[sp] = WE_BLAS << 30;
                stackT[sp] = tEntry;
                ++sp;
                continue;
            }
            const bool inBlas = type == WE_BLAS;
            // one load sequence for all three homes of a node (generic address: the top of the
            // TLAS lives in shared memory)
            const WideNode* np = inBlas ? blasNodes + payload
                                        : (payload < sc.nTop ? sc.topNodes + payload : sc.nodes + payload);
            const NodeRegs n = loadNodeShared(np);
            float tmin[8];
            uint32_t mask = testChildren(n, inBlas ? obj : world, lim, tmin);
            if (mask == 0) continue;
            const uint32_t imask = n.a.w >> 24;
            const uint32_t axis = n.b.x >> 30, childBase = n.b.x & 0x3FFFFFFFu;
            const SlabRay& rr = inBlas ? obj : world;
            const float idirAxis = axis == 0 ? rr.ix : (axis == 1 ? rr.iy : rr.iz);
            const bool forward = idirAxis >= 0.0f;
            while (mask) {
                const int i = forward ? (31 - __clz(mask)) : (__ffs(mask) - 1);  // farthest slot first
                mask &= ~(1u << i);
                const uint32_t meta = byteOf(i < 4 ? n.b.z : n.b.w, i & 3);
                uint32_t k;
                if ((imask >> i) & 1u) k = ((inBlas ? WE_BLAS : WE_TLAS) << 30) | (childBase + meta - 1u);
                else if (inBlas) k = (WE_LEAF << 30) | ((n.b.y + (meta >> 3)) << 2) | ((meta & 7u) - 1u);
                else k = (WE_INST << 30) | (n.b.y + (meta >> 3));
                if (sp < THEIA_STACK_SIZE) {
                    stackKey[sp] = k;
                    stackT[sp] = tmin[i];
                    ++sp;
                }
            }
        }

        // ---- phase L (reconverge first: the compiler would otherwise run it per exit path of N)
        __syncwarp();
        if (pending != 0xFFFFFFFFu) {
            const uint32_t first = pending >> 2, count = (pending & 3u) + 1u;
            pending = 0xFFFFFFFFu;
            for (uint32_t k = 0; k < count; ++k) {
                const float4* tp = reinterpret_cast<const float4*>(blasTris + first + k);
                const float4 a = __ldg(tp), b = __ldg(tp + 1), c = __ldg(tp + 2);
                const float v0[3] = {a.x, a.y, a.z}, v1[3] = {b.x, b.y, b.z}, v2[3] = {c.x, c.y, c.z};
                const int32_t prim = (int32_t)f2u(a.w);
                float t, u, v;
                if (!woopIntersect(wr, v0, v1, v2, tmax, &t, &u, &v)) continue;
                const bool better = !bestValid || t < bestT ||
                                    (t == bestT && (curInst < bestInst || (curInst == bestInst && prim < bestPrim)));
                if (better) { bestValid = true; bestInst = curInst; bestSlot = curSlot; bestPrim = prim; bestT = t; bestU = u; bestV = v; }
            }
        }
    }
}

// The same ray queries with one thread per ray and the single-loop traversal of traverse.cuh
// (leaf triangles are tested inside the node iteration). No refill, no phases: a warp lasts as
// long as its longest ray, but every iteration does node and triangle work for all lanes that
// have any. Faster where all rays of a launch cost about the same (C2: -20 %), slower where
// trivial and heavy rays mix (shadow rays of C3: +60 %, C5: +18 %), so api.cu picks the kernel
// per tracer and ray class from timings of the first batches. (A third form, persistent lanes
// around this traversal step with Aila & Laine's refill threshold, lay between the two on
// every workload for every threshold and was dropped.)
#ifndef THEIA_SIMPLE_CTAS
#define THEIA_SIMPLE_CTAS 6
#endif
#ifndef THEIA_SIMPLE_THREADS
#define THEIA_SIMPLE_THREADS 128
#endif
__global__ void __launch_bounds__(THEIA_SIMPLE_THREADS, THEIA_SIMPLE_CTAS) wfTraceSimpleKernel(const void* __restrict__ tlas, const float4* __restrict__ rays,
                                                               uint4* __restrict__ results,
                                                               const uint32_t* __restrict__ countPtr, uint32_t indirect,
                                                               int32_t targetId, const float* __restrict__ targetBox) {
    __shared__ float4 sharedRaw[THEIA_TOP_NODES * 5];
    WideNode* sharedTop = reinterpret_cast<WideNode*>(sharedRaw);
    const uint32_t total = *countPtr;
    if (blockIdx.x * blockDim.x >= total) return;
    const uint32_t nTop = stageTopNodes(tlas, sharedTop);
    __syncthreads();
    const SceneView sc = makeSceneView(tlas, sharedTop, nTop);
    const uint32_t onlySlot = singleTargetSlot(targetBox, targetId);
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        const float4 a = loadStream(rays + 2 * (size_t)i), b = loadStream(rays + 2 * (size_t)i + 1);
        const uint32_t rid = indirect ? f2u(b.w) : i;
        const RayHit h = closestHit<false>(sc, vec3{a.x, a.y, a.z}, vec3{b.x, b.y, b.z}, a.w, nullptr, targetId, onlySlot);
        storeStream(results + rid, make_uint4(h.valid ? (uint32_t)h.slot : 0xFFFFFFFFu, (uint32_t)h.prim, f2u(h.u), f2u(h.v)));
    }
}

// ------------------------------------------------------------------ shading side
// Resident CTAs per SM of the shading kernels. They wait on chains of dependent loads, so more warps
// are worth the spills of a smaller register budget: wfShadeKernel at 3 / 4 / 5 / 6 / 7 / 8 CTAs per SM
// takes 0.237 / 0.201 / 0.195 / 0.191 / 0.182 / 0.182 ms per launch on C2 (profiles/r1h_shade_occupancy_ab.txt).
#ifndef THEIA_SHADE_CTAS
#define THEIA_SHADE_CTAS 7
#endif
#ifndef THEIA_SHADOW_CTAS
#define THEIA_SHADOW_CTAS 5  // 4 / 5 / 6 / 7 / 8: 0.366 / 0.322 / 0.321 / 0.324 / 0.318 ms per launch on C3 (profiles/r2_ab_shade_ctas.txt)
#endif
#ifndef THEIA_GEN_CTAS
#define THEIA_GEN_CTAS 4
#endif
enum : uint32_t { WF_SEQ1 = 1u << 18 };  // one response was already committed in this event
enum : uint32_t { WF_LIVE0 = 1u << 19, WF_LIVE1 = 1u << 20 };  // parked path: its phase / guide shadow ray is traced

// Compile time variant key of the shading kernels (the reference compiles one shader per configuration,
// src/theia/util.py:102-115, src/theia/trace.py:666-685; here the configurations of BASELINE.json get their own
// instantiations and everything else runs the generic ones). The kernels are 110-160 KB of SASS when every
// component kind is a run time switch and starve on instruction fetch (profiles/r1i_wfshade_c2.txt).
//   VK_FULL    all components, everything decided at run time (responses / guides / sources added late, photon tracers)
//   VK_RTGUIDE lean components, guide decided at run time (generate kernel)
//   VK_GUIDE   a sphere target guide is set        VK_HIST  the response is the plain histogram
//   VK_NOCB    no callback
enum : uint32_t { VK_FULL = 1u, VK_RTGUIDE = 2u, VK_GUIDE = 4u, VK_HIST = 8u, VK_NOCB = 16u };
template <uint32_t V>
TDEV bool variantGuide(const DeviceLaunch& P) {
    if (V & (VK_FULL | VK_RTGUIDE)) return P.guide.kind != THEIA_GUIDE_NONE;
    return (V & VK_GUIDE) != 0;
}

// First half of trace() (sh/scene.traverse.glsl:280-323): free path sample, extension towards
// the target guide; writes the ray query to `rayOut` (the entry the path was given in the
// output buffer). The health check of :285 (isRayBad) is done by the caller, which needs its
// outcome to allocate the entry.
template <bool POL, bool FULL>
TDEV void prepareQuery(const DeviceLaunch& P, float4* rayOut, const Ray& ray, Rng& rng, bool allowResponse, bool GUIDE,
                       bool& misTarget) {
    misTarget = false;
    vec3 dir = normalize(ray.direction);
    float dist = sampleScatterLength(ray, P.propagation, rng.next());
    float sampledDist = dist;
    if (GUIDE) {
        TargetGuideSample guideSample = evalTargetGuide<FULL>(P.guide, ray.position, dir);
        bool mis = allowResponse && guideSample.prob > 0.0f;
        if (mis) {
            mis = guideSample.dist > dist;
            dist = max_(guideSample.dist, dist);
        }
        misTarget = mis;
    }
    rayOut[0] = make_float4(ray.position.x, ray.position.y, ray.position.z, dist);
    rayOut[1] = make_float4(dir.x, dir.y, dir.z, sampledDist);
}

// End of a segment (sh/tracer.scene.forward.glsl:87-104): event callback, stop conditions,
// then the next ray query, appended to the output buffer. MUST be called by all 32 lanes.
template <bool POL, uint32_t V>
TDEV void finishSegment(const DeviceLaunch& P, const Workspace& ws, SinkShared* sh, bool active, uint32_t path, Ray& ray,
                        Rng& rng, int result, uint32_t it, bool GUIDE, uint32_t next, uint32_t& nSegments) {
    constexpr bool FULL = (V & VK_FULL) != 0, NOCB = (V & VK_NOCB) != 0;
    if (FULL && P.photon && active) result = photonRoulette<POL>(ray, rng, result);
    onEvent<POL, NOCB>(P.callback, sh, active, ray, result, path, it);
    bool cont = active && result >= 0;
    bool maxIter = false, allow = true;
    if (cont) {
        allow = GUIDE ? (result != THEIA_RESULT_RAY_SCATTERED) : true;
        if (it == P.path_length) { maxIter = true; cont = false; }
        else {
            if (FULL && P.photon) photonRunBoundary(P, rng, it);
            it++;
        }
    }
    onEvent<POL, NOCB>(P.callback, sh, maxIter, ray, THEIA_RESULT_MAX_ITER, path, P.path_length + 1);
    if (cont) nSegments++;
    const bool ok = cont && !isRayBad(ray);
    onEvent<POL, NOCB>(P.callback, sh, cont && !ok, ray, THEIA_ERROR_RAY_BAD, path, it);
    const uint32_t j = denseAlloc(ws.cntOut, ok);
    if (ok) {
        bool mis;
        prepareQuery<POL, FULL>(P, ws.ray[next] + 2 * (size_t)j, ray, rng, allow, GUIDE, mis);
        storeState<POL>(ws.state[next], ws.capacity, j, ray, rng.dim, it | (allow ? WF_ALLOW : 0u) | (mis ? WF_MIS_TARGET : 0u), path);
    }
}

TDEV uint32_t wfWarpBase(uint32_t& stride) {
    const uint32_t warpsPerCta = blockDim.x >> 5;
    stride = gridDim.x * warpsPerCta * 32u;
    return (blockIdx.x * warpsPerCta + (threadIdx.x >> 5)) * 32u;
}

// Cascade light: the gamma variate of the shower depth comes from a rejection loop with a
// data dependent number of trials (sh/random.gamma.glsl:9-29). Run per path inside sampleLight
// a warp would wait for its unluckiest lane; here the warp runs the trials of its path range
// with refill (a lane that accepted takes the next path), parking value and draw counter at
// the path's index of state buffer 1 (unused until the first shade). Same draws, same
// arithmetic as sampleGamma.
TDEV void pregenerateGamma(const DeviceLaunch& P, const Workspace& ws, uint32_t first, uint32_t end) {
    const float alpha0 = P.light.p.cascade.a_long;
    const bool rescale = alpha0 < 1.0f;
    const GammaConsts g = gammaConsts(rescale ? alpha0 + 1.0f : alpha0);
    // draws of sampleWavelength ahead of the light source: the const and host sources take none
    // (sh/wavelengthsource.const.glsl, sh/wavelengthsource.host.glsl), uniform and function one
    const uint32_t skip = (P.wavelength.kind == THEIA_WAVELENGTH_CONST || P.wavelength.kind == THEIA_WAVELENGTH_HOST) ? 0u : 1u;
    const size_t C = ws.capacity;
    uint32_t next = first;
    bool busy = false;
    uint32_t k = 0;
    float scale = 1.0f;
    Rng rng = makeRng(&P.rng, 0);
    for (uint32_t guard = 0;; ++guard) {
        if (guard == 100000000u) {
            // a parameter set whose rejection loop never accepts: the paths still waiting get a NaN depth and a defined
            // draw counter, so they end with ERROR_CODE_RAY_BAD instead of reading stale workspace contents
            if (busy) { ws.state[1][WF_PX * C + k] = f2u(T_NAN); ws.state[1][WF_DIM * C + k] = rng.dim; }
            for (uint32_t j = next + laneId(); j < end; j += 32u) { ws.state[1][WF_PX * C + j] = f2u(T_NAN); ws.state[1][WF_DIM * C + j] = skip; }
            break;
        }
        uint32_t idle = __ballot_sync(FULL_MASK, !busy);
        if (idle) {
            uint32_t avail = end - next, rank = __popc(idle & laneMaskLt());
            if (!busy && rank < avail) {
                busy = true;
                k = next + rank;
                rng = makeRng(&P.rng, P.first_path + k);
                rng.dim = skip;
                scale = 1.0f;
                if (rescale) scale = dm_pow(rng.next(), 1.0f / alpha0);
            }
            next += min((uint32_t)__popc(idle), avail);
        }
        if (!__any_sync(FULL_MASK, busy)) break;
        if (busy) {
            float X;
            if (gammaTrial(g, rng, &X)) {
                ws.state[1][WF_PX * C + k] = f2u(scale * X);
                ws.state[1][WF_DIM * C + k] = rng.dim;
                busy = false;
            }
        }
    }
}

// generate: sh/tracer.scene.forward.glsl:50-66. Each warp owns a contiguous range of paths.
// NOCB: no callback; LIGHT >= 0: kind of the light source, both known at compile time (lean launches only)
template <bool POL, bool FULL, bool NOCB = false, int LIGHT = -1>
__global__ void __launch_bounds__(128, THEIA_GEN_CTAS) wfGenerateKernel(const __grid_constant__ DeviceLaunch P, const Workspace ws) {
    __shared__ SinkShared sh;
    sinkInit(P, &sh, nullptr);
    const bool GUIDE = P.guide.kind != THEIA_GUIDE_NONE;
    const bool PREGAMMA = LIGHT >= 0 ? LIGHT == (int)THEIA_LIGHT_CASCADE : P.light.kind == THEIA_LIGHT_CASCADE;
    const uint32_t warpsPerCta = blockDim.x >> 5, nWarps = gridDim.x * warpsPerCta;
    const uint32_t warp = blockIdx.x * warpsPerCta + (threadIdx.x >> 5);
    const uint32_t perWarp = ((P.batch_size + nWarps - 1) / nWarps + 31u) & ~31u;
    const uint32_t first = min(warp * perWarp, P.batch_size), end = min(first + perWarp, P.batch_size);
    if (PREGAMMA) {
        pregenerateGamma(P, ws, first, end);
        __syncwarp();
    }
    uint32_t nSegments = 0, nPaths = 0;
    for (uint32_t base = first; base < end; base += 32u) {
        const uint32_t k = base + laneId();
        const bool active = k < end;
        Ray ray;
        Rng rng = makeRng(&P.rng, P.first_path + k);
        const bool allow = P.disable_direct_lighting == 0;
        if (active) {
            PreGamma pre = {false, 0.0f, 0u};
            if (PREGAMMA) pre = {true, u2f(ws.state[1][WF_PX * (size_t)ws.capacity + k]), ws.state[1][WF_DIM * (size_t)ws.capacity + k]};
            ray = sampleRay<POL, FULL, LIGHT>(P.wavelength, P.light, P.medium, rng, pre);
            if (FULL && P.photon) { ray.lin_contrib = 1.0f; ray.log_contrib = 0.0f; }  // sh/tracer.scene.photon.init.glsl:50-54
            nPaths++;
            nSegments++;
        }
        onEvent<POL, NOCB>(P.callback, &sh, active, ray, THEIA_RESULT_RAY_CREATED, k, 0);
        const bool ok = active && !isRayBad(ray);
        onEvent<POL, NOCB>(P.callback, &sh, active && !ok, ray, THEIA_ERROR_RAY_BAD, k, 1);
        const uint32_t j = denseAlloc(ws.cntOut, ok);
        if (ok) {
            bool mis;
            prepareQuery<POL, FULL>(P, ws.ray[0] + 2 * (size_t)j, ray, rng, allow, GUIDE, mis);
            storeState<POL>(ws.state[0], ws.capacity, j, ray, rng.dim, 1u | (allow ? WF_ALLOW : 0u) | (mis ? WF_MIS_TARGET : 0u), k);
        }
    }
    sinkFinish(P, &sh, nullptr, nSegments, 0, 0, nPaths);
}

TDEV RayHit resultToHit(uint4 r) {
    RayHit q;
    q.valid = r.x != 0xFFFFFFFFu;
    q.slot = (int32_t)r.x; q.instance = -1; q.prim = (int32_t)r.y; q.t = 0.0f; q.u = u2f(r.z); q.v = u2f(r.w);
    return q;
}

// shade: second half of trace() (sh/scene.traverse.glsl:324-352) + processInteraction (:361-397)
template <bool POL, uint32_t V>
__global__ void __launch_bounds__(128, THEIA_SHADE_CTAS) wfShadeKernel(const __grid_constant__ DeviceLaunch P, const Workspace ws,
                                                         uint32_t cur) {
    constexpr bool FULL = (V & VK_FULL) != 0, HIST = (V & VK_HIST) != 0;
    extern __shared__ float sharedHist[];
    __shared__ SinkShared sh;
    sinkInit(P, &sh, sharedHist);
    float* hist = sinkHist(P, sharedHist);
    const SceneView sc = makeSceneView(P.tlas, nullptr, 0);
    const bool GUIDE = variantGuide<V>(P);
    const uint32_t count = *ws.cntIn;
    const uint32_t next = cur ^ 1u;
    const uint32_t* stateIn = ws.state[cur];
    const float4* rayIn = ws.ray[cur];
    uint32_t nSegments = 0, nResponses = 0, nShadow = 0, stride;

    for (uint32_t base = wfWarpBase(stride); base < count; base += stride) {
        const uint32_t i = base + laneId();
        const bool active = i < count;
        Ray ray;
        uint32_t dim = 0, flags = 1u, path = 0;
        // everything the segment needs is requested in one go: issued after the first use of
        // the state, the query and result loads would add a second memory latency
        float queryDist = 0.0f, sampledDist = 0.0f;
        uint4 queryResult4 = make_uint4(0xFFFFFFFFu, 0u, 0u, 0u);
        if (active) {
            queryDist = rayIn[2 * (size_t)i].w;
            sampledDist = rayIn[2 * (size_t)i + 1].w;
            queryResult4 = ws.result[i];
            loadState<POL>(stateIn, ws.capacity, i, ray, dim, flags, path);
        }
        const uint32_t idx = P.first_path + path;
        Rng rng = makeRng(&P.rng, idx);
        rng.dim = dim;
        const uint32_t it = flags & 0xFFFFu;
        const bool allowResponse = (flags & WF_ALLOW) != 0, misTarget = (flags & WF_MIS_TARGET) != 0;
        const uint32_t tagEvent = it - 1;
        uint32_t hitSeq = 0;
        HitItem item;
        bool has = false;
        SurfaceHit hit;
        hit.valid = false;
        int result = 0;
        if (active) {
            const RayHit q = resultToHit(queryResult4);
            float dist = queryDist;
            int queryResult = processRayQuery(sc, ray, q, hit);
            if (hit.valid) dist = distance(ray.position, hit.worldPos);
            if (GUIDE && misTarget && hit.valid && dist > sampledDist && queryResult >= 0) {
                has = processShadowRay<POL>(P, ray, hit, item);
                hit.valid = false;
                dist = sampledDist;
            }
            int propResult = propagateRay(ray, dist, P.propagation);
            updateRayIS(ray, dist, P.propagation, hit.valid);
            result = queryResult <= THEIA_ERROR_MAX_VALUE ? queryResult : propResult;
        }
        has = commitHit<FULL, HIST>(P.response, hist, has, item, idx, tagEvent | (hitSeq << 16), nResponses, responseDraw<FULL>(P.response, rng, has, item));
        hitSeq += has ? 1u : 0u;
        has = false;

        const bool interact = active && result >= 0;
        const bool last = (FULL && P.photon) ? false : it == P.path_length;  // the photon tracers always scatter
        bool wantMIS = false;
        float uResponse = 2.0f;
        if (interact && hit.valid) {
            // processHit (:73-154)
            result = propagateRayToHit<POL>(ray, hit.worldPos, hit.rayNrm, P.propagation);
            if (result >= 0) {
                result = THEIA_RESULT_RAY_HIT;
                bool isAbs = (hit.flags & MATERIAL_BLACK_BODY_BIT) != 0;
                bool isTarget = (hit.flags & MATERIAL_DETECTOR_BIT) != 0;
                bool volBorder = (hit.flags & MATERIAL_VOLUME_BORDER_BIT) != 0;
                bool canTransmit = (hit.flags & MATERIAL_NO_TRANSMIT_BWD_BIT) == 0;  // sic (:96-97)
                bool canReflect = (hit.flags & MATERIAL_NO_REFLECT_BWD_BIT) == 0;
                Reflectance surface = fresnelReflect(ray, hit);
                if (allowResponse && isTarget && (P.objectId < 0 || hit.customId == P.objectId)) {
                    sceneCreateResponse<POL>(P, ray, hit, surface, isAbs, item);
                    has = true;
                    uResponse = responseDraw<FULL>(P.response, rng, true, item);  // before the reflect / transmit draw
                    result = THEIA_RESULT_RAY_DETECTED;
                }
                if (isAbs) {
                    result = THEIA_RESULT_RAY_ABSORBED;
                } else if (P.disable_volume_border == 0 && volBorder) {
                    crossBorder(ray, hit);
                    result = THEIA_RESULT_VOLUME_HIT;
                } else if (P.disable_transmission == 0) {
                    float r = 0.5f * (surface.r_s * surface.r_s + surface.r_p * surface.r_p);
                    if (canReflect && canTransmit) {
                        if (rng.next() < r) reflectRayIS<POL>(ray, hit, surface);
                        else transmitRayIS<POL>(ray, hit, surface);
                    } else if (canReflect) {
                        reflectRay<POL>(ray, hit, surface);
                    } else if (canTransmit) {
                        transmitRay<POL>(ray, hit, surface);
                    } else {
                        result = THEIA_RESULT_RAY_ABSORBED;
                    }
                } else {
                    if (canReflect) reflectRay<POL>(ray, hit, surface);
                    else result = THEIA_RESULT_RAY_ABSORBED;
                }
            }
        } else if (interact) {
            result = THEIA_RESULT_RAY_SCATTERED;
            wantMIS = GUIDE && !last;
        }
        has = commitHit<FULL, HIST>(P.response, hist, has, item, idx, tagEvent | (hitSeq << 16), nResponses, uResponse);

        // paths that wait for shadow rays are parked in the shadow buffer (warp collective)
        uint32_t js = 0u;
        if (GUIDE) js = denseAlloc(ws.cntShadow, wantMIS);
        bool live0 = false, live1 = false;
        float4 srOrigin = make_float4(0, 0, 0, 0), srDir0 = make_float4(0, 0, 1, 0), srDir1 = make_float4(0, 0, 1, 0);
        float srDist1 = 0.0f;
        if (wantMIS) {
            // sampleTargetMIS (:234-269): the two shadow rays are traced by the next trace pass
            float pPP;
            vec3 dirPhase = scatter(ray.medium, ray.direction, rng.next2(), &pPP);
            TargetGuideSample targetSample = sampleTargetGuide<FULL>(P.guide, ray.position, rng);
            float pTT = targetSample.prob;
            TargetGuideSample phaseSample = evalTargetGuide<FULL>(P.guide, ray.position, dirPhase);
            float pTP = phaseSample.prob;
            float pPT = scatterProb(ray.medium, ray.direction, targetSample.dir);
            float wTarget = pTT * pPT / (pTT * pTT + pPT * pPT);
            float wPhase = pPP * pPP / (pPP * pPP + pTP * pTP);
            // a shadow ray that cannot reach any instance of the target has no effect whatever it hits: not traced
            live0 = !ws.cullShadow || rayMayHitBox(ws.targetBox, ray.position, dirPhase, phaseSample.dist);
            live1 = !ws.cullShadow || rayMayHitBox(ws.targetBox, ray.position, targetSample.dir, targetSample.dist);
            if (live0) ws.shadowDir[js] = make_float4(dirPhase.x, dirPhase.y, dirPhase.z, wPhase);
            if (live1) ws.shadowDir[(size_t)ws.capacity + js] = make_float4(targetSample.dir.x, targetSample.dir.y, targetSample.dir.z, wTarget);
            srOrigin = make_float4(ray.position.x, ray.position.y, ray.position.z, phaseSample.dist);
            srDir0 = make_float4(dirPhase.x, dirPhase.y, dirPhase.z, u2f(js));
            srDir1 = make_float4(targetSample.dir.x, targetSample.dir.y, targetSample.dir.z, u2f(ws.capacity + js));
            srDist1 = targetSample.dist;
            nShadow += 2;
            storeState<POL>(ws.shadowState, ws.capacity, js, ray, rng.dim,
                            it | (hitSeq ? WF_SEQ1 : 0u) | (live0 ? WF_LIVE0 : 0u) | (live1 ? WF_LIVE1 : 0u), path);
        } else if (interact && !hit.valid && !last) {
            scatterRay<POL>(ray, rng.next2());  // volume scatter (:386-393)
        }
        if (GUIDE) {
            // the rays to trace, dense: first the phase function samples of the warp, then its guide samples (rays of a
            // kind are neighbours: the first kind mostly leaves the scene at once, the second crosses it)
            const uint32_t m0 = __ballot_sync(FULL_MASK, live0), m1 = __ballot_sync(FULL_MASK, live1);
            if (m0 | m1) {
                uint32_t base = 0u;
                if (laneId() == 0) base = atomicAdd(ws.cntShadowRays, (uint32_t)(__popc(m0) + __popc(m1)));
                base = __shfl_sync(FULL_MASK, base, 0);
                if (live0) {
                    float4* r = ws.shadowLive + 2 * (size_t)(base + (uint32_t)__popc(m0 & laneMaskLt()));
                    r[0] = srOrigin; r[1] = srDir0;
                }
                if (live1) {
                    float4* r = ws.shadowLive + 2 * (size_t)(base + (uint32_t)__popc(m0) + (uint32_t)__popc(m1 & laneMaskLt()));
                    r[0] = make_float4(srOrigin.x, srOrigin.y, srOrigin.z, srDist1); r[1] = srDir1;
                }
            }
        }
        finishSegment<POL, V>(P, ws, &sh, active && !wantMIS, path, ray, rng, result, it, GUIDE, next, nSegments);
    }
    sinkFinish(P, &sh, sharedHist, nSegments, nResponses, nShadow, 0);
}

// second shading pass (target guide only): shadow ray results -> responses
// (traceShadowRay, sh/scene.traverse.glsl:185-214), then the deferred volume scatter
template <bool POL, uint32_t V>
__global__ void __launch_bounds__(128, THEIA_SHADOW_CTAS) wfShadeShadowKernel(const __grid_constant__ DeviceLaunch P, const Workspace ws,
                                                               uint32_t cur) {
    constexpr bool FULL = (V & VK_FULL) != 0, HIST = (V & VK_HIST) != 0;
    extern __shared__ float sharedHist[];
    __shared__ SinkShared sh;
    sinkInit(P, &sh, sharedHist);
    float* hist = sinkHist(P, sharedHist);
    const SceneView sc = makeSceneView(P.tlas, nullptr, 0);
    const uint32_t count = *ws.cntShadow;
    const uint32_t next = cur ^ 1u;
    uint32_t nSegments = 0, nResponses = 0, stride;

    for (uint32_t base = wfWarpBase(stride); base < count; base += stride) {
        const uint32_t i = base + laneId();
        const bool active = i < count;
        const uint32_t slot = i;
        Ray ray;
        uint32_t dim = 0, flags = 1u, path = 0;
        // both shadow rays (direction + weight) and their results are requested together with the state
        float4 dirW[2] = {make_float4(0, 0, 1, 0), make_float4(0, 0, 1, 0)};
        uint4 res[2] = {make_uint4(0xFFFFFFFFu, 0u, 0u, 0u), make_uint4(0xFFFFFFFFu, 0u, 0u, 0u)};
        if (active) {
            loadState<POL>(ws.shadowState, ws.capacity, i, ray, dim, flags, path);
            // (a ray that was not traced has no result: it counts as a miss)
            if (flags & WF_LIVE0) { dirW[0] = ws.shadowDir[slot]; res[0] = ws.shadowResult[slot]; }
            if (flags & WF_LIVE1) { dirW[1] = ws.shadowDir[(size_t)ws.capacity + slot]; res[1] = ws.shadowResult[(size_t)ws.capacity + slot]; }
        }
        const uint32_t idx = P.first_path + path;
        Rng rng = makeRng(&P.rng, idx);
        rng.dim = dim;
        const uint32_t it = flags & 0xFFFFu;
        const uint32_t tagEvent = it - 1;
        uint32_t hitSeq = (flags & WF_SEQ1) ? 1u : 0u;
#pragma unroll 1
        for (uint32_t j = 0; j < 2; ++j) {
            HitItem item;
            bool has = false;
            const RayHit q = resultToHit(j == 0 ? res[0] : res[1]);
            if (active && q.valid) {  // (a miss has no effect: processShadowRay, sh/scene.traverse.glsl:168)
                const float4 b = j == 0 ? dirW[0] : dirW[1];
                Ray r2 = ray;
                scatterRayIS<POL>(r2, vec3{b.x, b.y, b.z});
                SurfaceHit hit;
                if (processRayQuery(sc, r2, q, hit) >= 0) {
                    r2.lin_contrib *= b.w;
                    has = processShadowRay<POL>(P, r2, hit, item);
                }
            }
            has = commitHit<FULL, HIST>(P.response, hist, has, item, idx, tagEvent | (hitSeq << 16), nResponses, responseDraw<FULL>(P.response, rng, has, item));
            hitSeq += has ? 1u : 0u;
        }
        if (active) scatterRay<POL>(ray, rng.next2());
        finishSegment<POL, V>(P, ws, &sh, active, path, ray, rng, THEIA_RESULT_RAY_SCATTERED, it, true, next, nSegments);
    }
    sinkFinish(P, &sh, sharedHist, nSegments, nResponses, 0, 0);
}

}  // namespace theia

