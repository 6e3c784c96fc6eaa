// GPU builder of the acceleration structures (replaces the driver-side BLAS / TLAS builds
// behind hp.GeometryStore and hp.AccelerationStructure, src/theia/scene.py:550, 668).
//
// Pipeline (all on the device, caller supplied result + scratch memory):
//   1. primitive boxes + centroid bounds          (one thread per primitive)
//   2. 63 bit Morton codes of the centroids, own stable LSD radix sort (8 passes of 8 bits: radixCount / radixScan /
//      radixScatter below)
//   3. binary radix tree (Karras 2012), one thread per inner node
//   4. bottom-up box refit with per-node arrival counters
//      [THEIA_BVH_BUILDER=sah replaces 3 + 4 by a sweep builder: every node range of the Morton
//      order is split where n_L A_L + n_R A_R is smallest (one warp per node, level by level).
//      Measured on the meshes of C2 / C3 / C5 it does not pay: 6.32 nodes + 8.93 triangles per
//      ray against 6.19 + 8.56 of the radix tree on C2, step times within 2 % either way - the
//      SAH-guided collapse of step 5 already decides which boxes a ray meets - so it stays an
//      option. Results do not depend on the tree.]
//   5. collapse into 8-wide nodes: per wide node the binary subtree is opened greedily at the
//      child with the largest surface area (SAH-guided), children are sorted along the node's
//      widest axis, boxes are quantised outwards to 8 bit; breadth first, deterministic
//      (block-wide prefix sums hand out node / primitive slots)
//   6. leaf triangles are gathered into 48 byte records in leaf order.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <vector>
#include "bvh.cuh"
#include "../../include/theia_b200.h"

namespace theia {

void setError(const char* fmt, ...);  // api.cu
void countLaunches(unsigned n);      // api.cu
void tunerForget(const void* tlas);  // api.cu

// ------------------------------------------------------------------ scratch layout
struct BuildScratch {
    float4* boxLo;      // [n] primitive boxes
    float4* boxHi;      // [n]
    uint64_t* keys;     // [n]
    uint64_t* keysAlt;  // [n]
    uint32_t* vals;     // [n]
    uint32_t* valsAlt;  // [n]
    float4* nodeLo;     // [2n-1] binary tree boxes (inner 0..n-2, leaves n-1..2n-2)
    float4* nodeHi;     // [2n-1]
    int32_t* parent;    // [2n-1]
    int2* child;        // [n-1]
    int2* range;        // [n-1] first, last sorted position
    uint32_t* arrive;   // [n-1]
    uint32_t* wideBin;  // [n] binary node behind each wide node
    int4* work[2];      // [n] each: (node, first, last, -) worklists of the SAH sweep builder
    uint32_t* bounds;   // [8] ordered-int centroid bounds: lo xyz, hi xyz
    uint32_t* info;     // [8] nWide, nPrimOut, depth, overflow
    uint32_t* sortHist;  // [256][sortTiles(n)] digit counts per tile of the radix sort, then their exclusive scan
    size_t total;
};

static size_t alignUp(size_t v) { return (v + 255) & ~size_t(255); }

#define SORT_THREADS 256u
#define SORT_ITEMS 16u  // keys per thread and pass: tiles of 4096 keys
static uint32_t sortTiles(uint32_t n) { return (n + SORT_THREADS * SORT_ITEMS - 1) / (SORT_THREADS * SORT_ITEMS); }

static BuildScratch layoutScratch(void* base, uint32_t n) {
    BuildScratch s;
    size_t off = 0;
    char* b = (char*)base;
    auto take = [&](size_t bytes) { void* p = b ? b + off : nullptr; off += alignUp(bytes); return p; };
    uint32_t m = n > 0 ? n : 1;
    s.boxLo = (float4*)take(sizeof(float4) * m);
    s.boxHi = (float4*)take(sizeof(float4) * m);
    s.keys = (uint64_t*)take(8 * m);
    s.keysAlt = (uint64_t*)take(8 * m);
    s.vals = (uint32_t*)take(4 * m);
    s.valsAlt = (uint32_t*)take(4 * m);
    s.nodeLo = (float4*)take(sizeof(float4) * 2 * m);
    s.nodeHi = (float4*)take(sizeof(float4) * 2 * m);
    s.parent = (int32_t*)take(4 * 2 * m);
    s.child = (int2*)take(sizeof(int2) * m);
    s.range = (int2*)take(sizeof(int2) * m);
    s.arrive = (uint32_t*)take(4 * m);
    s.wideBin = (uint32_t*)take(4 * m);
    s.work[0] = (int4*)take(sizeof(int4) * m);
    s.work[1] = (int4*)take(sizeof(int4) * m);
    s.bounds = (uint32_t*)take(32);
    s.info = (uint32_t*)take(32);
    s.sortHist = (uint32_t*)take(4 * 256 * (size_t)sortTiles(m));
    s.total = off;
    return s;
}

// ------------------------------------------------------------------ helpers
// monotone float <-> uint mapping for atomicMin / atomicMax
__device__ __forceinline__ uint32_t orderedBits(float f) {
    uint32_t u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float orderedFloat(uint32_t u) {
    return __uint_as_float((u & 0x80000000u) ? (u & 0x7FFFFFFFu) : ~u);
}
__device__ __forceinline__ uint64_t expandBits21(uint32_t v) {
    uint64_t x = v & 0x1FFFFFu;
    x = (x | x << 32) & 0x1F00000000FFFFull;
    x = (x | x << 16) & 0x1F0000FF0000FFull;
    x = (x | x << 8) & 0x100F00F00F00F00Full;
    x = (x | x << 4) & 0x10C30C30C30C30C3ull;
    x = (x | x << 2) & 0x1249249249249249ull;
    return x;
}

__global__ void initBuildKernel(BuildScratch s, uint32_t n) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < 3) s.bounds[i] = 0xFFFFFFFFu;
    else if (i < 6) s.bounds[i] = 0u;
    if (i < 8) s.info[i] = 0u;
    if (n > 1 && i < n - 1) s.arrive[i] = 0u;
}

__device__ __forceinline__ void publishBox(const BuildScratch& s, uint32_t i, float lo[3], float hi[3]) {
    s.boxLo[i] = make_float4(lo[0], lo[1], lo[2], 0.0f);
    s.boxHi[i] = make_float4(hi[0], hi[1], hi[2], 0.0f);
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        float c = 0.5f * lo[a] + 0.5f * hi[a];
        atomicMin(s.bounds + a, orderedBits(c));
        atomicMax(s.bounds + 3 + a, orderedBits(c));
    }
}

// 1a. triangle boxes; padded by a few ulp so that flat (axis aligned) triangles get a volume
__global__ void triangleBoxKernel(BuildScratch s, const float* __restrict__ verts, const uint32_t* __restrict__ idx,
                                  uint32_t n) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (int k = 0; k < 3; ++k) {
        const float* p = verts + 6 * (size_t)idx[3 * i + k];
        for (int a = 0; a < 3; ++a) { lo[a] = fminf(lo[a], p[a]); hi[a] = fmaxf(hi[a], p[a]); }
    }
    for (int a = 0; a < 3; ++a) {
        float e = 4e-6f * fmaxf(fabsf(lo[a]), fabsf(hi[a])) + 1e-30f;
        lo[a] -= e; hi[a] += e;
    }
    publishBox(s, i, lo, hi);
}

// 1b. instance boxes: world box of the eight transformed corners of the BLAS bounds
__global__ void instanceBoxKernel(BuildScratch s, const InstanceRec* __restrict__ inst, uint32_t n) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const InstanceRec& in = inst[i];
    const BlasHeader* h = reinterpret_cast<const BlasHeader*>(reinterpret_cast<const char*>(in.blasNodes) - sizeof(BlasHeader));
    float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (int c = 0; c < 8; ++c) {
        float p[3] = {(c & 1) ? h->hi[0] : h->lo[0], (c & 2) ? h->hi[1] : h->lo[1], (c & 4) ? h->hi[2] : h->lo[2]};
        float w[3];
        transformPoint(in.o2w, p, w);
        for (int a = 0; a < 3; ++a) { lo[a] = fminf(lo[a], w[a]); hi[a] = fmaxf(hi[a], w[a]); }
    }
    for (int a = 0; a < 3; ++a) {
        float e = 1e-5f * fmaxf(fabsf(lo[a]), fabsf(hi[a])) + 1e-30f;
        lo[a] -= e; hi[a] += e;
    }
    publishBox(s, i, lo, hi);
}

// 2. Morton codes
__global__ void mortonKernel(BuildScratch s, uint32_t n) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4 lo = s.boxLo[i], hi = s.boxHi[i];
    float c[3] = {0.5f * lo.x + 0.5f * hi.x, 0.5f * lo.y + 0.5f * hi.y, 0.5f * lo.z + 0.5f * hi.z};
    uint64_t key = 0;
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        float bl = orderedFloat(s.bounds[a]), bh = orderedFloat(s.bounds[3 + a]);
        float ext = bh - bl;
        float t = ext > 0.0f ? (c[a] - bl) / ext : 0.0f;
        uint32_t q = (uint32_t)fminf(fmaxf(t * 2097152.0f, 0.0f), 2097151.0f);
        key |= expandBits21(q) << a;
    }
    s.keys[i] = key;
    s.vals[i] = i;
}

// 3. Karras 2012. delta = common prefix length, ties broken by the position
__device__ __forceinline__ int deltaFn(const uint64_t* keys, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    uint64_t a = keys[i], b = keys[j];
    if (a == b) return 64 + __clz((uint32_t)i ^ (uint32_t)j);
    return __clzll((long long)(a ^ b));
}
__global__ void radixTreeKernel(BuildScratch s, const uint64_t* __restrict__ keys, uint32_t n) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (int)n - 1) return;
    int N = (int)n;
    int d = deltaFn(keys, N, i, i + 1) - deltaFn(keys, N, i, i - 1) >= 0 ? 1 : -1;
    int dmin = deltaFn(keys, N, i, i - d);
    int lmax = 2;
    while (deltaFn(keys, N, i, i + lmax * d) > dmin) lmax *= 2;
    int l = 0;
    for (int t = lmax / 2; t >= 1; t /= 2)
        if (deltaFn(keys, N, i, i + (l + t) * d) > dmin) l += t;
    int j = i + l * d;
    int dnode = deltaFn(keys, N, i, j);
    int sp = 0;
    for (int t = (l + 1) / 2;; t = (t + 1) / 2) {
        if (deltaFn(keys, N, i, i + (sp + t) * d) > dnode) sp += t;
        if (t == 1) break;
    }
    int gamma = i + sp * d + min(d, 0);
    int first = min(i, j), last = max(i, j);
    int left = (first == gamma) ? (N - 1 + gamma) : gamma;
    int right = (last == gamma + 1) ? (N - 1 + gamma + 1) : gamma + 1;
    s.child[i] = make_int2(left, right);
    s.range[i] = make_int2(first, last);
    s.parent[left] = i;
    s.parent[right] = i;
    if (i == 0) s.parent[0] = -1;
}

// 4. refit: leaves climb, the second arrival at an inner node computes its box
__global__ void refitKernel(BuildScratch s, const uint32_t* __restrict__ order, uint32_t n) {
    uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    uint32_t prim = order[j];
    int node = (int)(n - 1 + j);
    s.nodeLo[node] = s.boxLo[prim];
    s.nodeHi[node] = s.boxHi[prim];
    if (n == 1) return;
    __threadfence();
    int p = s.parent[node];
    while (p >= 0) {
        if (atomicAdd(s.arrive + p, 1u) == 0u) return;
        __threadfence();
        int2 c = s.child[p];
        float4 al = __ldcg(s.nodeLo + c.x), ah = __ldcg(s.nodeHi + c.x);
        float4 bl = __ldcg(s.nodeLo + c.y), bh = __ldcg(s.nodeHi + c.y);
        s.nodeLo[p] = make_float4(fminf(al.x, bl.x), fminf(al.y, bl.y), fminf(al.z, bl.z), 0.0f);
        s.nodeHi[p] = make_float4(fmaxf(ah.x, bh.x), fmaxf(ah.y, bh.y), fmaxf(ah.z, bh.z), 0.0f);
        __threadfence();
        p = s.parent[p];
    }
}

// 3'. SAH sweep builder. One CTA; a level of the tree at a time; one warp per node of the level.
// Node numbering as in the radix tree (root 0, a node split after position g has the children g
// and g + 1, leaves are n - 1 + position), so everything downstream is unchanged.
struct SweepBox { float lx, ly, lz, hx, hy, hz; };
__device__ __forceinline__ SweepBox emptyBox() { return {INFINITY, INFINITY, INFINITY, -INFINITY, -INFINITY, -INFINITY}; }
__device__ __forceinline__ SweepBox uniteBox(SweepBox a, SweepBox b) {
    return {fminf(a.lx, b.lx), fminf(a.ly, b.ly), fminf(a.lz, b.lz), fmaxf(a.hx, b.hx), fmaxf(a.hy, b.hy), fmaxf(a.hz, b.hz)};
}
__device__ __forceinline__ SweepBox shflUpBox(SweepBox b, int o) {
    return {__shfl_up_sync(0xFFFFFFFFu, b.lx, o), __shfl_up_sync(0xFFFFFFFFu, b.ly, o), __shfl_up_sync(0xFFFFFFFFu, b.lz, o),
            __shfl_up_sync(0xFFFFFFFFu, b.hx, o), __shfl_up_sync(0xFFFFFFFFu, b.hy, o), __shfl_up_sync(0xFFFFFFFFu, b.hz, o)};
}
__device__ __forceinline__ SweepBox shflBox(SweepBox b, int src) {
    return {__shfl_sync(0xFFFFFFFFu, b.lx, src), __shfl_sync(0xFFFFFFFFu, b.ly, src), __shfl_sync(0xFFFFFFFFu, b.lz, src),
            __shfl_sync(0xFFFFFFFFu, b.hx, src), __shfl_sync(0xFFFFFFFFu, b.hy, src), __shfl_sync(0xFFFFFFFFu, b.hz, src)};
}
__device__ __forceinline__ float sweepArea(SweepBox b) {
    float dx = b.hx - b.lx, dy = b.hy - b.ly, dz = b.hz - b.lz;
    return dx * dy + dy * dz + dz * dx;
}
// inclusive scan over the lanes (lane k: union of lanes 0..k), then the carry of earlier chunks
__device__ __forceinline__ SweepBox scanBox(SweepBox b, SweepBox carry) {
    const uint32_t lane = threadIdx.x & 31u;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        SweepBox up = shflUpBox(b, o);
        if (lane >= (uint32_t)o) b = uniteBox(b, up);
    }
    return uniteBox(b, carry);
}

#define SWEEP_THREADS 1024
__global__ void __launch_bounds__(SWEEP_THREADS) sahTreeKernel(BuildScratch s, const uint32_t* __restrict__ order, uint32_t n) {
    __shared__ uint32_t counts[2];
    const int N = (int)n;
    float* rightArea = reinterpret_cast<float*>(s.arrive);  // [n] scratch of the sweeps (ranges of a level are disjoint)
    // leaves of the binary tree = primitive boxes in Morton order
    for (uint32_t j = threadIdx.x; j < n; j += blockDim.x) {
        uint32_t prim = order[j];
        s.nodeLo[n - 1 + j] = s.boxLo[prim];
        s.nodeHi[n - 1 + j] = s.boxHi[prim];
    }
    if (threadIdx.x == 0) {
        counts[0] = N > 1 ? 1u : 0u;
        counts[1] = 0u;
        if (N > 1) s.work[0][0] = make_int4(0, 0, N - 1, 0);
    }
    __syncthreads();
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5, nWarps = blockDim.x >> 5;
    for (uint32_t level = 0;; ++level) {
        const uint32_t cur = level & 1u, nxt = cur ^ 1u;
        const uint32_t count = counts[cur];
        if (count == 0) break;
        for (uint32_t w = warp; w < count; w += nWarps) {
            const int4 job = s.work[cur][w];
            const int id = job.x, f = job.y, l = job.z;
            const int cnt = l - f + 1;
            // ---- suffix sweep: rightArea[p] = area of the union of positions p .. l
            SweepBox carry = emptyBox();
            for (int base = 0; base < cnt; base += 32) {
                const int p = l - (base + (int)lane);
                SweepBox b = emptyBox();
                if (p >= f) {
                    float4 lo = s.nodeLo[N - 1 + p], hi = s.nodeHi[N - 1 + p];
                    b = {lo.x, lo.y, lo.z, hi.x, hi.y, hi.z};
                }
                b = scanBox(b, carry);
                if (p >= f) rightArea[p] = sweepArea(b);
                carry = shflBox(b, 31);
            }
            const SweepBox whole = carry;
            __syncwarp();
            // ---- prefix sweep: cost of splitting after position p. Large ranges may not be split
            // in their outer eighths, which bounds the depth of the tree.
            const int margin = cnt >= 16 ? cnt / 8 : 0;
            const int mid2 = f + l - 1;  // twice the middle split position
            float bestCost = INFINITY;
            int bestOff = 0x7FFFFFFF, bestP = f;
            carry = emptyBox();
            for (int base = 0; base < cnt; base += 32) {
                const int p = f + base + (int)lane;
                SweepBox b = emptyBox();
                if (p <= l) {
                    float4 lo = s.nodeLo[N - 1 + p], hi = s.nodeHi[N - 1 + p];
                    b = {lo.x, lo.y, lo.z, hi.x, hi.y, hi.z};
                }
                b = scanBox(b, carry);
                if (p >= f + margin && p < l - margin) {
                    float cost = sweepArea(b) * (float)(p - f + 1) + rightArea[p + 1] * (float)(l - p);
                    int off = abs(2 * p - mid2);
                    if (!(cost == cost)) cost = INFINITY;
                    if (cost < bestCost || (cost == bestCost && (off < bestOff || (off == bestOff && p < bestP)))) {
                        bestCost = cost; bestOff = off; bestP = p;
                    }
                }
                carry = shflBox(b, 31);
            }
#pragma unroll
            for (int o = 16; o >= 1; o >>= 1) {
                float c2 = __shfl_xor_sync(0xFFFFFFFFu, bestCost, o);
                int o2 = __shfl_xor_sync(0xFFFFFFFFu, bestOff, o), p2 = __shfl_xor_sync(0xFFFFFFFFu, bestP, o);
                if (c2 < bestCost || (c2 == bestCost && (o2 < bestOff || (o2 == bestOff && p2 < bestP)))) {
                    bestCost = c2; bestOff = o2; bestP = p2;
                }
            }
            int g = bestP;
            if (bestOff == 0x7FFFFFFF) g = (f + l - 1) / 2;  // no finite cost anywhere: split in the middle
            if (lane == 0) {
                const int left = (g == f) ? (N - 1 + g) : g;
                const int right = (g + 1 == l) ? (N - 1 + g + 1) : g + 1;
                s.child[id] = make_int2(left, right);
                s.range[id] = make_int2(f, l);
                s.nodeLo[id] = make_float4(whole.lx, whole.ly, whole.lz, 0.0f);
                s.nodeHi[id] = make_float4(whole.hx, whole.hy, whole.hz, 0.0f);
                const uint32_t nNew = (g > f ? 1u : 0u) + (g + 1 < l ? 1u : 0u);
                if (nNew) {
                    uint32_t k = atomicAdd(&counts[nxt], nNew);
                    if (g > f) s.work[nxt][k++] = make_int4(g, f, g, 0);
                    if (g + 1 < l) s.work[nxt][k] = make_int4(g + 1, g + 1, l, 0);
                }
            }
        }
        __syncthreads();
        if (threadIdx.x == 0) counts[cur] = 0u;
        __syncthreads();
    }
}

// ------------------------------------------------------------------ 5. collapse to 8-wide nodes
struct CollapseArgs {
    uint32_t n;
    uint32_t tlas;             // 0: BLAS (triangles), 1: TLAS (instances)
    const uint32_t* order;     // sorted position -> primitive
    WideNode* nodes;           // out
    TriRec* tris;              // out (BLAS)
    const float* verts;        // BLAS input
    const uint32_t* idx;       // BLAS input
    InstanceRec* instOut;      // out (TLAS)
    const InstanceRec* instIn; // TLAS input, original order
    uint32_t* headerCounts;    // out: [nNodes, nPrims, depth]
    float* headerBounds;       // out: lo xyz, hi xyz
};

#define COLLAPSE_THREADS 512

__device__ __forceinline__ float boxArea(float4 lo, float4 hi) {
    float dx = hi.x - lo.x, dy = hi.y - lo.y, dz = hi.z - lo.z;
    return dx * dy + dy * dz + dz * dx;
}

// exclusive block scan of two counters; returns totals through tot0 / tot1
__device__ void blockScan2(uint32_t v0, uint32_t v1, uint32_t& ex0, uint32_t& ex1, uint32_t& tot0, uint32_t& tot1,
                           uint32_t* warpBuf /* [2 * 32] */) {
    const uint32_t lane = threadIdx.x & 31u, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
    uint32_t s0 = v0, s1 = v1;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t a = __shfl_up_sync(0xFFFFFFFFu, s0, o), b = __shfl_up_sync(0xFFFFFFFFu, s1, o);
        if (lane >= (uint32_t)o) { s0 += a; s1 += b; }
    }
    if (lane == 31) { warpBuf[wid] = s0; warpBuf[32 + wid] = s1; }
    __syncthreads();
    if (wid == 0) {
        uint32_t a = lane < nw ? warpBuf[lane] : 0u, b = lane < nw ? warpBuf[32 + lane] : 0u;
        uint32_t ia = a, ib = b;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            uint32_t x = __shfl_up_sync(0xFFFFFFFFu, ia, o), y = __shfl_up_sync(0xFFFFFFFFu, ib, o);
            if (lane >= (uint32_t)o) { ia += x; ib += y; }
        }
        warpBuf[lane] = ia - a;       // exclusive warp offsets
        warpBuf[32 + lane] = ib - b;
        if (lane == 31) { warpBuf[64] = ia; warpBuf[65] = ib; }
    }
    __syncthreads();
    ex0 = warpBuf[wid] + s0 - v0;
    ex1 = warpBuf[32 + wid] + s1 - v1;
    tot0 = warpBuf[64];
    tot1 = warpBuf[65];
    __syncthreads();
}

// smallest biased exponent e with 255 * 2^(e-127) >= extent (checked in double precision)
__device__ __forceinline__ uint32_t gridExponent(float lo, float hi) {
    const double extent = (double)hi - (double)lo;
    float step = (float)(extent / 255.0);
    uint32_t bits = __float_as_uint(step);
    uint32_t e = bits >> 23;
    if ((bits & 0x7FFFFFu) != 0u) e += 1;
    if (e < 1u) e = 1u;
    while (e < 254u && 255.0 * (double)__uint_as_float(e << 23) < extent) e += 1;
    if (e > 254u) e = 254u;
    return e;
}

__global__ void __launch_bounds__(COLLAPSE_THREADS) collapseKernel(BuildScratch s, CollapseArgs A) {
    __shared__ uint32_t warpBuf[66];
    __shared__ uint32_t levelBegin, levelEnd, nodeAlloc, primAlloc, depth;
    const int N = (int)A.n;
    const int LEAF = A.tlas ? 1 : THEIA_LEAF_MAX;
    if (threadIdx.x == 0) {
        levelBegin = 0; levelEnd = 1; nodeAlloc = 1; primAlloc = 0; depth = 0;
        s.wideBin[0] = (N > 1) ? 0u : (uint32_t)(N - 1);  // single primitive: the leaf itself
        float4 lo = s.nodeLo[(N > 1) ? 0 : N - 1], hi = s.nodeHi[(N > 1) ? 0 : N - 1];
        A.headerBounds[0] = lo.x; A.headerBounds[1] = lo.y; A.headerBounds[2] = lo.z;
        A.headerBounds[3] = hi.x; A.headerBounds[4] = hi.y; A.headerBounds[5] = hi.z;
    }
    __syncthreads();

    auto primCount = [&](int node) -> int {
        if (node >= N - 1) return 1;
        int2 r = s.range[node];
        return r.y - r.x + 1;
    };
    auto firstPos = [&](int node) -> int { return node >= N - 1 ? node - (N - 1) : s.range[node].x; };

    while (true) {
        const uint32_t lb = levelBegin, le = levelEnd;
        if (lb >= le) break;
        for (uint32_t start = lb; start < le; start += blockDim.x) {
            const uint32_t w = start + threadIdx.x;
            const bool active = w < le;
            int refs[8];
            int cnt = 0;
            if (active) {
                int b = (int)s.wideBin[w];
                if (primCount(b) <= LEAF) {
                    refs[0] = b; cnt = 1;
                } else {
                    int2 c = s.child[b];
                    refs[0] = c.x; refs[1] = c.y; cnt = 2;
                    while (cnt < 8) {
                        int best = -1;
                        float bestArea = -1.0f;
                        for (int k = 0; k < cnt; ++k) {
                            if (primCount(refs[k]) <= LEAF) continue;
                            float a = boxArea(s.nodeLo[refs[k]], s.nodeHi[refs[k]]);
                            if (a > bestArea) { bestArea = a; best = k; }
                        }
                        if (best < 0) break;
                        int2 c2 = s.child[refs[best]];
                        refs[best] = c2.x;
                        refs[cnt++] = c2.y;
                    }
                }
            }
            // node box = union of the child boxes
            float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
            for (int k = 0; k < cnt; ++k) {
                float4 l = s.nodeLo[refs[k]], h = s.nodeHi[refs[k]];
                lo[0] = fminf(lo[0], l.x); lo[1] = fminf(lo[1], l.y); lo[2] = fminf(lo[2], l.z);
                hi[0] = fmaxf(hi[0], h.x); hi[1] = fmaxf(hi[1], h.y); hi[2] = fmaxf(hi[2], h.z);
            }
            int axis = 0;
            if (cnt > 0) {
                float ex = hi[0] - lo[0], ey = hi[1] - lo[1], ez = hi[2] - lo[2];
                axis = (ey > ex && ey >= ez) ? 1 : ((ez > ex && ez > ey) ? 2 : 0);
                // insertion sort of the children by centroid along the widest axis
                for (int a = 1; a < cnt; ++a) {
                    int r = refs[a];
                    float4 l = s.nodeLo[r], h = s.nodeHi[r];
                    float c = axis == 0 ? l.x + h.x : (axis == 1 ? l.y + h.y : l.z + h.z);
                    int b = a - 1;
                    while (b >= 0) {
                        float4 l2 = s.nodeLo[refs[b]], h2 = s.nodeHi[refs[b]];
                        float c2 = axis == 0 ? l2.x + h2.x : (axis == 1 ? l2.y + h2.y : l2.z + h2.z);
                        if (c2 <= c) break;
                        refs[b + 1] = refs[b];
                        --b;
                    }
                    refs[b + 1] = r;
                }
            }
            uint32_t nInner = 0, nPrims = 0;
            for (int k = 0; k < cnt; ++k) {
                int pc = primCount(refs[k]);
                if (pc <= LEAF) nPrims += pc; else nInner += 1;
            }
            uint32_t exInner, exPrims, totInner, totPrims;
            blockScan2(nInner, nPrims, exInner, exPrims, totInner, totPrims, warpBuf);
            const uint32_t childBase = nodeAlloc + exInner, primBase = primAlloc + exPrims;
            if (active) {
                WideNode node;
                memset(&node, 0, sizeof(node));
                node.px = lo[0]; node.py = lo[1]; node.pz = lo[2];
                const uint32_t e[3] = {gridExponent(lo[0], hi[0]), gridExponent(lo[1], hi[1]), gridExponent(lo[2], hi[2])};
                node.ex = (uint8_t)e[0]; node.ey = (uint8_t)e[1]; node.ez = (uint8_t)e[2];
                node.childBase = childBase | ((uint32_t)axis << 30);
                node.primBase = primBase;
                uint32_t rank = 0, off = 0, imask = 0;
                for (int k = 0; k < cnt; ++k) {
                    const int r = refs[k];
                    const int pc = primCount(r);
                    float4 l = s.nodeLo[r], h = s.nodeHi[r];
                    const float cl[3] = {l.x, l.y, l.z}, ch[3] = {h.x, h.y, h.z};
                    uint8_t ql[3], qh[3];
                    for (int a = 0; a < 3; ++a) {
                        const double step = (double)__uint_as_float(e[a] << 23);
                        double fl = floor(((double)cl[a] - (double)lo[a]) / step);
                        double fh = ceil(((double)ch[a] - (double)lo[a]) / step);
                        ql[a] = (uint8_t)fmin(fmax(fl, 0.0), 255.0);
                        qh[a] = (uint8_t)fmin(fmax(fh, 0.0), 255.0);
                    }
                    node.qlox[k] = ql[0]; node.qloy[k] = ql[1]; node.qloz[k] = ql[2];
                    node.qhix[k] = qh[0]; node.qhiy[k] = qh[1]; node.qhiz[k] = qh[2];
                    if (pc <= LEAF) {
                        node.meta[k] = (uint8_t)((uint32_t)pc | (off << 3));
                        const int fp = firstPos(r);
                        for (int j = 0; j < pc; ++j) {
                            const uint32_t prim = A.order[fp + j];
                            const uint32_t dst = primBase + off + j;
                            if (A.tlas) {
                                InstanceRec rec = A.instIn[prim];
                                rec.pad = prim;  // original instance id
                                A.instOut[dst] = rec;
                            } else {
                                TriRec t;
                                const float* p0 = A.verts + 6 * (size_t)A.idx[3 * prim + 0];
                                const float* p1 = A.verts + 6 * (size_t)A.idx[3 * prim + 1];
                                const float* p2 = A.verts + 6 * (size_t)A.idx[3 * prim + 2];
                                t.v0x = p0[0]; t.v0y = p0[1]; t.v0z = p0[2]; t.prim = prim;
                                t.v1x = p1[0]; t.v1y = p1[1]; t.v1z = p1[2]; t.pad1 = 0;
                                t.v2x = p2[0]; t.v2y = p2[1]; t.v2z = p2[2]; t.pad2 = 0;
                                A.tris[dst] = t;
                            }
                        }
                        off += pc;
                    } else {
                        imask |= 1u << k;
                        node.meta[k] = (uint8_t)(1u + rank);
                        s.wideBin[childBase + rank] = (uint32_t)r;
                        rank += 1;
                    }
                }
                node.imask = (uint8_t)imask;
                A.nodes[w] = node;
            }
            __syncthreads();
            if (threadIdx.x == 0) { nodeAlloc += totInner; primAlloc += totPrims; }
            __syncthreads();
        }
        if (threadIdx.x == 0) { levelBegin = le; levelEnd = nodeAlloc; depth += 1; }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        A.headerCounts[0] = nodeAlloc;
        A.headerCounts[1] = primAlloc;
        A.headerCounts[2] = depth;
    }
}

// ------------------------------------------------------------------ host side
static int checkCuda(cudaError_t e, const char* what) {
    if (e == cudaSuccess) return 0;
    setError("%s: %s", what, cudaGetErrorString(e));
    return -1;
}
#define CUDA_OK(expr) do { if (checkCuda((expr), #expr)) return -1; } while (0)

static uint32_t blocksFor(uint32_t n, uint32_t bs) { return (n + bs - 1) / bs; }

// steps 2-4 for n >= 1 primitives whose boxes are in s.boxLo / s.boxHi; returns the sorted order
// ------------------------------------------------------------------ radix sort of (Morton key, primitive) pairs
// Least significant digit first, 8 bits per pass, stable: per pass (1) every tile of 4096 keys counts its digits,
// (2) one CTA scans the counts in (digit, tile) order, (3) every tile walks its keys again in index order and writes
// each behind the keys of its digit that come before it: in earlier tiles (scan), earlier rounds of the tile (running
// base), earlier warps of the round (per warp counts) and lower lanes of the warp (__match_any_sync).
__global__ void __launch_bounds__(SORT_THREADS) radixCountKernel(const uint64_t* __restrict__ keys, uint32_t n, uint32_t shift,
                                                                 uint32_t* __restrict__ hist, uint32_t nTiles) {
    __shared__ uint32_t counts[256];
    counts[threadIdx.x] = 0u;
    __syncthreads();
    const uint32_t base = blockIdx.x * SORT_THREADS * SORT_ITEMS;
    for (uint32_t r = 0; r < SORT_ITEMS; ++r) {
        const uint32_t i = base + r * SORT_THREADS + threadIdx.x;
        if (i < n) atomicAdd(&counts[(uint32_t)(keys[i] >> shift) & 0xFFu], 1u);
    }
    __syncthreads();
    hist[threadIdx.x * nTiles + blockIdx.x] = counts[threadIdx.x];
}

__global__ void __launch_bounds__(1024) radixScanKernel(uint32_t* __restrict__ hist, uint32_t len) {
    __shared__ uint32_t warpSum[32];
    __shared__ uint32_t carry;
    if (threadIdx.x == 0) carry = 0u;
    __syncthreads();
    for (uint32_t base = 0; base < len; base += 1024u) {
        const uint32_t i = base + threadIdx.x;
        const uint32_t v = i < len ? hist[i] : 0u;
        uint32_t incl = v;
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t o = __shfl_up_sync(0xFFFFFFFFu, incl, d);
            if ((threadIdx.x & 31u) >= (uint32_t)d) incl += o;
        }
        if ((threadIdx.x & 31u) == 31u) warpSum[threadIdx.x >> 5] = incl;
        __syncthreads();
        if (threadIdx.x < 32) {
            uint32_t w = warpSum[threadIdx.x];
            for (int d = 1; d < 32; d <<= 1) {
                const uint32_t o = __shfl_up_sync(0xFFFFFFFFu, w, d);
                if (threadIdx.x >= (uint32_t)d) w += o;
            }
            warpSum[threadIdx.x] = w;  // inclusive over the warps
        }
        __syncthreads();
        const uint32_t before = carry + ((threadIdx.x >> 5) ? warpSum[(threadIdx.x >> 5) - 1] : 0u) + incl - v;
        if (i < len) hist[i] = before;
        __syncthreads();
        if (threadIdx.x == 0) carry += warpSum[31];
        __syncthreads();
    }
}

__global__ void __launch_bounds__(SORT_THREADS) radixScatterKernel(const uint64_t* __restrict__ keysIn, const uint32_t* __restrict__ valsIn,
                                                                   uint64_t* __restrict__ keysOut, uint32_t* __restrict__ valsOut,
                                                                   uint32_t n, uint32_t shift, const uint32_t* __restrict__ hist,
                                                                   uint32_t nTiles) {
    __shared__ uint32_t digitBase[256];
    __shared__ uint32_t warpCount[SORT_THREADS / 32][256];
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    digitBase[threadIdx.x] = hist[threadIdx.x * nTiles + blockIdx.x];
    for (uint32_t w = 0; w < SORT_THREADS / 32; ++w) warpCount[w][threadIdx.x] = 0u;
    __syncthreads();
    const uint32_t base = blockIdx.x * SORT_THREADS * SORT_ITEMS;
    for (uint32_t r = 0; r < SORT_ITEMS; ++r) {
        const uint32_t i = base + r * SORT_THREADS + threadIdx.x;
        const bool valid = i < n;
        uint64_t key = 0;
        uint32_t val = 0, digit = 0x100u + lane;  // lanes without a key match nobody
        if (valid) { key = keysIn[i]; val = valsIn[i]; digit = (uint32_t)(key >> shift) & 0xFFu; }
        const uint32_t peers = __match_any_sync(0xFFFFFFFFu, digit);
        const uint32_t rankInWarp = __popc(peers & ((1u << lane) - 1u));
        if (valid && rankInWarp == 0u) warpCount[warp][digit] = __popc(peers);
        __syncthreads();
        if (valid) {
            uint32_t pos = digitBase[digit] + rankInWarp;
            for (uint32_t w = 0; w < warp; ++w) pos += warpCount[w][digit];
            keysOut[pos] = key;
            valsOut[pos] = val;
        }
        __syncthreads();
        uint32_t sum = 0;
        for (uint32_t w = 0; w < SORT_THREADS / 32; ++w) { sum += warpCount[w][threadIdx.x]; warpCount[w][threadIdx.x] = 0u; }
        digitBase[threadIdx.x] += sum;
        __syncthreads();
    }
}

// Sorts n (key, value) pairs by the low 64 bits of the key, stable; the result is back in s.keys / s.vals.
static int radixSortPairs(BuildScratch& s, uint32_t n, cudaStream_t st) {
    const uint32_t nTiles = sortTiles(n);
    uint64_t *kin = s.keys, *kout = s.keysAlt;
    uint32_t *vin = s.vals, *vout = s.valsAlt;
    for (uint32_t shift = 0; shift < 64; shift += 8) {
        radixCountKernel<<<nTiles, SORT_THREADS, 0, st>>>(kin, n, shift, s.sortHist, nTiles);
        radixScanKernel<<<1, 1024, 0, st>>>(s.sortHist, 256u * nTiles);
        radixScatterKernel<<<nTiles, SORT_THREADS, 0, st>>>(kin, vin, kout, vout, n, shift, s.sortHist, nTiles);
        uint64_t* tk = kin; kin = kout; kout = tk;
        uint32_t* tv = vin; vin = vout; vout = tv;
    }
    CUDA_OK(cudaGetLastError());
    return 0;  // eight passes: the sorted pairs are in s.keys / s.vals again
}

static int buildBinaryTree(BuildScratch& s, uint32_t n, cudaStream_t st, const uint32_t** order) {
    const uint32_t bs = 256;
    mortonKernel<<<blocksFor(n, bs), bs, 0, st>>>(s, n);
    if (radixSortPairs(s, n, st)) return -1;
    const uint64_t* keys = s.keys;
    *order = s.vals;
    static const bool sweep = getenv("THEIA_BVH_BUILDER") && !strcmp(getenv("THEIA_BVH_BUILDER"), "sah");
    if (!sweep) {
        if (n > 1) radixTreeKernel<<<blocksFor(n - 1, bs), bs, 0, st>>>(s, keys, n);
        refitKernel<<<blocksFor(n, bs), bs, 0, st>>>(s, *order, n);
    } else {
        sahTreeKernel<<<1, SWEEP_THREADS, 0, st>>>(s, *order, n);
    }
    CUDA_OK(cudaGetLastError());
    countLaunches(n > 1 ? 27 : 26);  // morton, sort (8 passes of 3 kernels), radix tree, refit
    return 0;
}

}  // namespace theia

using namespace theia;

extern "C" int theia_blas_sizes(uint32_t n_triangles, theia_as_sizes_t* sizes) {
    if (!sizes) { setError("theia_blas_sizes: sizes is null"); return -1; }
    if (n_triangles == 0) { setError("theia_blas_sizes: mesh has no triangles"); return -1; }
    uint64_t n = n_triangles;
    sizes->result_bytes = sizeof(BlasHeader) + n * sizeof(WideNode) + n * sizeof(TriRec);
    sizes->scratch_bytes = layoutScratch(nullptr, n_triangles).total;
    return 0;
}

extern "C" int theia_blas_build(const float* vertices_dev, uint32_t n_vertices, const uint32_t* indices_dev,
                                uint32_t n_triangles, void* result_dev, uint64_t result_bytes, void* scratch_dev,
                                uint64_t scratch_bytes, void* stream) {
    theia_as_sizes_t need;
    if (theia_blas_sizes(n_triangles, &need)) return -1;
    if (!vertices_dev || !indices_dev || !result_dev || !scratch_dev || n_vertices == 0) {
        setError("theia_blas_build: null argument");
        return -1;
    }
    if (result_bytes < need.result_bytes || scratch_bytes < need.scratch_bytes) {
        setError("theia_blas_build: buffers too small (need %llu + %llu bytes)", (unsigned long long)need.result_bytes,
                 (unsigned long long)need.scratch_bytes);
        return -1;
    }
    if (((uintptr_t)result_dev & 15u) || ((uintptr_t)scratch_dev & 15u)) {
        setError("theia_blas_build: buffers must be 16 byte aligned");
        return -1;
    }
    cudaStream_t st = (cudaStream_t)stream;
    const uint32_t n = n_triangles, bs = 256;
    BuildScratch s = layoutScratch(scratch_dev, n);
    char* out = (char*)result_dev;
    BlasHeader h;
    memset(&h, 0, sizeof(h));
    h.magic = THEIA_BLAS_MAGIC;
    h.nodesOffset = sizeof(BlasHeader);
    h.trisOffset = sizeof(BlasHeader) + (uint64_t)n * sizeof(WideNode);
    CUDA_OK(cudaMemcpyAsync(out, &h, sizeof(h), cudaMemcpyHostToDevice, st));
    initBuildKernel<<<blocksFor(n > 8 ? n : 8, bs), bs, 0, st>>>(s, n);
    triangleBoxKernel<<<blocksFor(n, bs), bs, 0, st>>>(s, vertices_dev, indices_dev, n);
    const uint32_t* order = nullptr;
    if (buildBinaryTree(s, n, st, &order)) return -1;
    CollapseArgs A;
    memset(&A, 0, sizeof(A));
    A.n = n; A.tlas = 0; A.order = order;
    A.nodes = (WideNode*)(out + h.nodesOffset);
    A.tris = (TriRec*)(out + h.trisOffset);
    A.verts = vertices_dev; A.idx = indices_dev;
    A.headerCounts = (uint32_t*)(out + offsetof(BlasHeader, nNodes));  // nNodes, nTris, orderAxis := depth
    A.headerBounds = (float*)(out + offsetof(BlasHeader, lo));
    collapseKernel<<<1, COLLAPSE_THREADS, 0, st>>>(s, A);
    CUDA_OK(cudaGetLastError());
    countLaunches(3);  // init, boxes, collapse
    // the host copy of h must outlive the async copy
    CUDA_OK(cudaStreamSynchronize(st));
    return 0;
}

extern "C" int theia_tlas_sizes(uint32_t n_instances, theia_as_sizes_t* sizes) {
    if (!sizes) { setError("theia_tlas_sizes: sizes is null"); return -1; }
    if (n_instances == 0) { setError("theia_tlas_sizes: scene has no instances"); return -1; }
    uint64_t n = n_instances;
    sizes->result_bytes = sizeof(TlasHeader) + n * sizeof(WideNode) + n * sizeof(InstanceRec);
    sizes->scratch_bytes = layoutScratch(nullptr, n_instances).total + 256 + n * sizeof(InstanceRec);
    return 0;
}

extern "C" int theia_tlas_build(const theia_instance_desc_t* instances, uint32_t n_instances, void* result_dev,
                                uint64_t result_bytes, void* scratch_dev, uint64_t scratch_bytes, void* stream) {
    theia_as_sizes_t need;
    if (theia_tlas_sizes(n_instances, &need)) return -1;
    if (!instances || !result_dev || !scratch_dev) { setError("theia_tlas_build: null argument"); return -1; }
    if (result_bytes < need.result_bytes || scratch_bytes < need.scratch_bytes) {
        setError("theia_tlas_build: buffers too small (need %llu + %llu bytes)", (unsigned long long)need.result_bytes,
                 (unsigned long long)need.scratch_bytes);
        return -1;
    }
    if (((uintptr_t)result_dev & 15u) || ((uintptr_t)scratch_dev & 15u)) {
        setError("theia_tlas_build: buffers must be 16 byte aligned");
        return -1;
    }
    cudaStream_t st = (cudaStream_t)stream;
    const uint32_t n = n_instances, bs = 256;
    tunerForget(result_dev);  // what the ray query kernel tuner learnt about this address belongs to another scene
    // BLAS headers (host copies, one per distinct BLAS)
    std::map<const void*, BlasHeader> blas;
    uint32_t maxBlasDepth = 0, maxBlasTris = 0;
    for (uint32_t i = 0; i < n; ++i) {
        const void* b = instances[i].blas;
        if (!b) { setError("theia_tlas_build: instance %u has no BLAS", i); return -1; }
        if (blas.count(b)) continue;
        BlasHeader h;
        CUDA_OK(cudaMemcpyAsync(&h, b, sizeof(h), cudaMemcpyDeviceToHost, st));
        CUDA_OK(cudaStreamSynchronize(st));
        if (h.magic != THEIA_BLAS_MAGIC) { setError("theia_tlas_build: instance %u: not a BLAS", i); return -1; }
        blas[b] = h;
        if (h.orderAxis > maxBlasDepth) maxBlasDepth = h.orderAxis;
        if (h.nTris > maxBlasTris) maxBlasTris = h.nTris;
    }
    std::vector<InstanceRec> recs(n);
    for (uint32_t i = 0; i < n; ++i) {
        const theia_instance_desc_t& d = instances[i];
        const BlasHeader& h = blas[d.blas];
        InstanceRec& r = recs[i];
        memset(&r, 0, sizeof(r));
        memcpy(r.w2o, d.world_to_obj, 48);
        memcpy(r.o2w, d.obj_to_world, 48);
        r.blasNodes = (const WideNode*)((const char*)d.blas + h.nodesOffset);
        r.blasTris = (const TriRec*)((const char*)d.blas + h.trisOffset);
        r.vertices = d.geometry.vertices; r.indices = d.geometry.indices; r.material = d.geometry.material;
        r.customIndex = d.custom_index;
        r.pad = i;
    }
    BuildScratch s = layoutScratch(scratch_dev, n);
    InstanceRec* instIn = (InstanceRec*)((char*)scratch_dev + ((s.total + 255) & ~size_t(255)));
    CUDA_OK(cudaMemcpyAsync(instIn, recs.data(), sizeof(InstanceRec) * n, cudaMemcpyHostToDevice, st));
    char* out = (char*)result_dev;
    TlasHeader h;
    memset(&h, 0, sizeof(h));
    h.magic = THEIA_TLAS_MAGIC;
    h.nodesOffset = sizeof(TlasHeader);
    h.instancesOffset = sizeof(TlasHeader) + (uint64_t)n * sizeof(WideNode);
    // stride of the 32 bit tie key (instance * stride + primitive) of the queued ray query kernel
    // (trace_queue.cuh); 0 = the scene does not fit it and that kernel falls back to one thread per ray
    h.pad[0] = ((uint64_t)n * maxBlasTris < 0xFFFFFFFFull && maxBlasTris < (1u << 25)) ? maxBlasTris : 0u;
    CUDA_OK(cudaMemcpyAsync(out, &h, sizeof(h), cudaMemcpyHostToDevice, st));
    initBuildKernel<<<blocksFor(n > 8 ? n : 8, bs), bs, 0, st>>>(s, n);
    instanceBoxKernel<<<blocksFor(n, bs), bs, 0, st>>>(s, instIn, n);
    const uint32_t* order = nullptr;
    if (buildBinaryTree(s, n, st, &order)) return -1;
    CollapseArgs A;
    memset(&A, 0, sizeof(A));
    A.n = n; A.tlas = 1; A.order = order;
    A.nodes = (WideNode*)(out + h.nodesOffset);
    A.instOut = (InstanceRec*)(out + h.instancesOffset);
    A.instIn = instIn;
    A.headerCounts = (uint32_t*)(out + offsetof(TlasHeader, nNodes));  // nNodes, nInstances, pad0 := depth
    A.headerBounds = (float*)(out + offsetof(TlasHeader, lo));
    collapseKernel<<<1, COLLAPSE_THREADS, 0, st>>>(s, A);
    CUDA_OK(cudaGetLastError());
    countLaunches(3);  // init, boxes, collapse
    TlasHeader back;
    CUDA_OK(cudaMemcpyAsync(&back, out, sizeof(back), cudaMemcpyDeviceToHost, st));
    CUDA_OK(cudaStreamSynchronize(st));
    // traversal stack bound: every popped node pushes at most 8 entries
    const uint32_t depth = back.pad0 + maxBlasDepth;
    if (7u * depth + 2u > THEIA_STACK_SIZE) {
        setError("theia_tlas_build: tree too deep for the traversal stack (%u levels)", depth);
        return -1;
    }
    return 0;
}
