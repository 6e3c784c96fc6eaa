// C-ABI entry points of libtheia_b200.so (include/theia_b200.h): error reporting, launch
// validation and translation into the device launch block, plus the small kernels behind
// the RNG sink, the histogram reducer, the hit replay and the ray query test seam.
#include <cuda_runtime.h>
#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <vector>
#include "trace_volume.cuh"
#include "wavefront.cuh"
#include "trace_queue.cuh"
#include "units.cuh"

namespace theia {

static thread_local char g_error[512] = "";
static std::atomic<unsigned long long> g_launches{0};  // kernels launched by this library
void countLaunches(unsigned n) { g_launches += n; }

// ------------------------------------------------------------------ per kernel timing
// Optional (theia_profile_enable): brackets every launch of the tracing kernels with CUDA
// events on the launching stream so that bench.py can report live per-kernel durations.
// Off by default: the events cost a few microseconds per launch.
enum KernelClass : int { KC_VOLUME, KC_SCENE, KC_GENERATE, KC_TRACE, KC_SHADE, KC_SHADOW_TRACE, KC_SHADE_SHADOW,
                         KC_REDUCE, KC_COUNT };
static const char* const g_classNames[KC_COUNT] = {"volumeForwardKernel", "sceneForwardKernel", "wfGenerateKernel",
                                                   "wfTraceKernel", "wfShadeKernel", "wfTraceKernel(shadow)",
                                                   "wfShadeShadowKernel", "histReduceKernel"};
struct PendingTiming { int cls, dev; cudaEvent_t a, b; };
#define THEIA_MAX_DEVICES 64
// Guards the profiler and the trace kernel tuners: the library may be driven by several host threads (two pipeline
// schedulers) and for several devices (events are only ever reused on the device they were created on).
static std::recursive_mutex g_stateMutex;
static struct {
    bool on = false;
    std::vector<PendingTiming> pending;
    std::vector<cudaEvent_t> pool[THEIA_MAX_DEVICES];
    double ms[KC_COUNT] = {};
    unsigned long long n[KC_COUNT] = {};
} g_prof;
static int currentDeviceSlot() {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0) dev = 0;
    return dev % THEIA_MAX_DEVICES;
}
static cudaEvent_t profEvent(int dev) {
    std::lock_guard<std::recursive_mutex> lock(g_stateMutex);
    cudaEvent_t e = nullptr;
    auto& pool = g_prof.pool[dev];
    if (!pool.empty()) { e = pool.back(); pool.pop_back(); }
    else cudaEventCreate(&e);
    return e;
}
static void profRecycle(int dev, cudaEvent_t e) {
    std::lock_guard<std::recursive_mutex> lock(g_stateMutex);
    g_prof.pool[dev].push_back(e);
}
struct Timed {
    PendingTiming t{};
    cudaStream_t st;
    bool active;
    Timed(int cls, cudaStream_t s) : st(s), active(g_prof.on) {
        if (!active) return;
        t.cls = cls; t.dev = currentDeviceSlot(); t.a = profEvent(t.dev); t.b = profEvent(t.dev);
        cudaEventRecord(t.a, st);
    }
    ~Timed() {
        if (!active) return;
        cudaEventRecord(t.b, st);
        std::lock_guard<std::recursive_mutex> lock(g_stateMutex);
        g_prof.pending.push_back(t);
    }
};
static void profCollect() {
    std::lock_guard<std::recursive_mutex> lock(g_stateMutex);
    for (auto& t : g_prof.pending) {
        float ms = 0.0f;
        cudaEventSynchronize(t.b);
        if (cudaEventElapsedTime(&ms, t.a, t.b) == cudaSuccess) { g_prof.ms[t.cls] += ms; g_prof.n[t.cls] += 1; }
        g_prof.pool[t.dev].push_back(t.a);
        g_prof.pool[t.dev].push_back(t.b);
    }
    g_prof.pending.clear();
}

void setError(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_error, sizeof(g_error), fmt, ap);
    va_end(ap);
}
static int cudaFail(cudaError_t e, const char* what) {
    setError("%s: %s", what, cudaGetErrorString(e));
    return -2;
}
#define CUDA_TRY(expr) do { cudaError_t e_ = (expr); if (e_ != cudaSuccess) return cudaFail(e_, #expr); } while (0)

static uint32_t blocksFor(uint64_t n, uint32_t bs) { return (uint32_t)((n + bs - 1) / bs); }

// ------------------------------------------------------------------ small kernels
// sh/random.sink.glsl:17-39
__global__ void philoxFillKernel(theia_philox_params_t p, uint32_t baseStream, uint32_t baseCount, uint32_t streams,
                                 uint32_t samples, uint32_t dim, float* out) {
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (uint64_t)streams * samples) return;
    uint32_t s = (uint32_t)(i / samples), c = (uint32_t)(i % samples);
    PhiloxKey key = philoxStreamKey(p, baseStream + s);
    if (dim == 1) {
        out[i] = random_s(p, key, baseCount + c);
    } else {
        out[2 * i] = random_s(p, key, baseCount + c);
        out[2 * i + 1] = random_s(p, key, baseCount + c + 1);
    }
}
__global__ void philoxRawKernel(const uint32_t* ctr, const uint32_t* key, uint32_t n, uint32_t* out) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t r[4];
    philox4x32_10(ctr[4 * i], ctr[4 * i + 1], ctr[4 * i + 2], ctr[4 * i + 3], key[2 * i], key[2 * i + 1], r);
    out[4 * i] = r[0]; out[4 * i + 1] = r[1]; out[4 * i + 2] = r[2]; out[4 * i + 3] = r[3];
}
__global__ void mathEvalKernel(uint32_t fn, const float* x, const float* y, uint32_t n, float* out) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float s, c;
    switch (fn) {
    case 0: dm_sincos(x[i], &s, &c); out[i] = s; break;
    case 1: dm_sincos(x[i], &s, &c); out[i] = c; break;
    case 2: out[i] = dm_exp(x[i]); break;
    case 3: out[i] = dm_log(x[i]); break;
    case 4: out[i] = dm_pow(x[i], y[i]); break;
    case 5: out[i] = dm_tanh(x[i]); break;
    default: out[i] = dm_atan(x[i]); break;
    }
}
// sh/estimator.reduce.glsl:17-35: per bin Kahan sum over the partial histograms (precise)
// The order of the sum is the reference's (row 0, 1, 2, ...); only the loads are batched so
// that 16 rows are in flight per thread instead of one dependent load per row.
__global__ void histReduceKernel(float* partials, uint32_t nHist, uint32_t nBins, float norm, float* out, int clear) {
    uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= nBins) return;
    constexpr uint32_t BATCH = 16;
    float result = 0.0f, c = 0.0f;
    for (uint32_t n0 = 0; n0 < nHist; n0 += BATCH) {
        float v[BATCH];
#pragma unroll
        for (uint32_t k = 0; k < BATCH; ++k) {
            float* p = partials + (size_t)(n0 + k) * nBins + b;
            v[k] = (n0 + k < nHist) ? __ldcg(p) : 0.0f;
        }
        if (clear) {
#pragma unroll
            for (uint32_t k = 0; k < BATCH; ++k)
                if (n0 + k < nHist) partials[(size_t)(n0 + k) * nBins + b] = 0.0f;
        }
#pragma unroll
        for (uint32_t k = 0; k < BATCH; ++k) {
            if (n0 + k < nHist) {
                float y = __fsub_rn(v[k], c);
                float t = __fadd_rn(result, y);
                c = __fsub_rn(__fsub_rn(t, result), y);
                result = t;
            }
        }
    }
    out[b] = result * norm;
}
// sh/response.replay.glsl:35-50
// hasRng != 0: hit k draws from stream k, starting at its first number (idx = gl_GlobalInvocationID.x, dim = 0, :37-38)
__global__ void __launch_bounds__(128) hitReplayKernel(const uint32_t* queue, uint32_t capacity, uint32_t polarized,
                                                       ResponseDev r, const __grid_constant__ theia_philox_params_t rngParams,
                                                       uint32_t hasRng) {
    extern __shared__ float sharedHist[];
    if (responseHasHistogram(r.kind) && r.hist_shared)
        for (uint32_t b = threadIdx.x; b < r.n_bins; b += blockDim.x) sharedHist[b] = 0.0f;
    __syncthreads();
    float* hist = nullptr;
    if (responseHasHistogram(r.kind))
        hist = r.hist_shared ? sharedHist : r.hist_partials + (size_t)(blockIdx.x % r.hist_capacity) * r.n_bins;
    uint32_t count = min(queue[0], capacity);
    // grid-stride over warps so that all 32 lanes reach commitHit together
    for (uint32_t base = blockIdx.x * blockDim.x; base < capacity; base += gridDim.x * blockDim.x) {
        uint32_t k = base + threadIdx.x;
        bool has = k < count;
        HitItem item;
        item.contrib = 0.0f; item.time = 0.0f;
        if (has) item = loadHit(queue, capacity, polarized != 0, k);
        // response(hit) without the "contrib > 0" test of the tracers' call sites (sh/response.replay.glsl:46-47)
        float u = 2.0f;
        if (hasRng && has && r.kind == THEIA_RESPONSE_STORE_TIME) {
            Rng rng = makeRng(&rngParams, k);
            u = rng.next();
        }
        commitHitImpl<true, false>(r, hist, has, item, k, 0u, u);
    }
    __syncthreads();
    if (responseHasHistogram(r.kind) && r.hist_shared) {
        const bool exclusive = gridDim.x <= r.hist_capacity;
        float* dst = r.hist_partials + (size_t)(blockIdx.x % r.hist_capacity) * r.n_bins;
        for (uint32_t b = threadIdx.x; b < r.n_bins; b += blockDim.x) {
            float v = sharedHist[b];
            if (exclusive) dst[b] += v;
            else if (v != 0.0f) atomicAdd(dst + b, v);
        }
    }
}
// ray query test seam
__global__ void __launch_bounds__(128) tlasQueryKernel(const void* tlas, const float* rays, uint32_t n, int32_t* out,
                                                       float* outBary, unsigned long long* stats) {
    __shared__ float4 sharedRaw[THEIA_TOP_NODES * 5];
    WideNode* sharedTop = reinterpret_cast<WideNode*>(sharedRaw);
    const uint32_t nTop = stageTopNodes(tlas, sharedTop);
    __syncthreads();
    const SceneView sc = makeSceneView(tlas, sharedTop, nTop);
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float* r = rays + 8 * (size_t)i;
    TraversalStats st = {0, 0, 0};
    RayHit h = closestHit<true>(sc, v3(r), v3(r + 4), r[3], &st);
    out[4 * i + 0] = h.valid ? h.instance : -1;
    out[4 * i + 1] = h.valid ? h.prim : -1;
    out[4 * i + 2] = h.valid ? sc.instances[h.slot].customIndex : -1;
    out[4 * i + 3] = 0;
    outBary[4 * i + 0] = h.valid ? h.t : 0.0f;
    outBary[4 * i + 1] = h.valid ? h.u : 0.0f;
    outBary[4 * i + 2] = h.valid ? h.v : 0.0f;
    outBary[4 * i + 3] = 0.0f;
    if (stats) {
        atomicAdd(stats + 0, (unsigned long long)st.nodes);
        atomicAdd(stats + 1, (unsigned long long)st.tris);
        atomicAdd(stats + 2, (unsigned long long)st.instances);
        atomicAdd(stats + 3, 1ull);
    }
}

// the same seam through the wavefront trace kernel (THEIA_QUERY_WAVEFRONT=1)
__global__ void wfQuerySetupKernel(uint32_t* counters, uint32_t n) {
    if (threadIdx.x == 0) { counters[0] = n; counters[1] = 0u; }
}
__global__ void wfQueryConvertKernel(const void* tlas, const float* rays, const uint4* results, uint32_t n, int32_t* out,
                                     float* outBary) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const SceneView sc = makeSceneView(tlas, nullptr, 0);
    const RayHit h = resultToHit(results[i]);
    out[4 * i + 0] = h.valid ? (int32_t)sc.instances[h.slot].pad : -1;
    out[4 * i + 1] = h.valid ? h.prim : -1;
    out[4 * i + 2] = h.valid ? sc.instances[h.slot].customIndex : -1;
    out[4 * i + 3] = 0;
    float t = 0.0f;
    if (h.valid) {
        // the trace kernel keeps (slot, primitive, u, v); t follows from the hit point like in
        // processRayQuery: recompute it with the triangle test of the query kernel
        const float* r = rays + 8 * (size_t)i;
        TraversalStats st = {0, 0, 0};
        t = closestHit<false>(sc, v3(r), v3(r + 4), r[3], &st).t;
    }
    outBary[4 * i + 0] = t;
    outBary[4 * i + 1] = h.valid ? h.u : 0.0f;
    outBary[4 * i + 2] = h.valid ? h.v : 0.0f;
    outBary[4 * i + 3] = 0.0f;
}

// ------------------------------------------------------------------ measurement aids
// Visit counts of the ray queries of a wavefront launch (theia_trace_stats_enable): walks the same rays with the
// instrumented one-thread-per-ray traversal and sums nodes visited / triangles tested / instances entered / rays.
__global__ void __launch_bounds__(128) wfTraceStatsKernel(const void* tlas, const float4* rays, const uint32_t* countPtr,
                                                          unsigned long long* stats, int32_t targetId, const float* targetBox) {
    const SceneView sc = makeSceneView(tlas, nullptr, 0);
    const uint32_t total = *countPtr;
    const uint32_t onlySlot = singleTargetSlot(targetBox, targetId);
    unsigned long long nodes = 0, tris = 0, inst = 0, n = 0;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        const float4 a = __ldg(rays + 2 * (size_t)i), b = __ldg(rays + 2 * (size_t)i + 1);
        TraversalStats st = {0, 0, 0};
        closestHit<true>(sc, vec3{a.x, a.y, a.z}, vec3{b.x, b.y, b.z}, a.w, &st, targetId, onlySlot);
        nodes += st.nodes; tris += st.tris; inst += st.instances; n += 1;
    }
    for (int o = 16; o > 0; o >>= 1) {
        nodes += __shfl_down_sync(FULL_MASK, nodes, o); tris += __shfl_down_sync(FULL_MASK, tris, o);
        inst += __shfl_down_sync(FULL_MASK, inst, o); n += __shfl_down_sync(FULL_MASK, n, o);
    }
    if (laneId() == 0 && n) { atomicAdd(stats + 0, nodes); atomicAdd(stats + 1, tris); atomicAdd(stats + 2, inst); atomicAdd(stats + 3, n); }
}
static unsigned long long* g_traceStats = nullptr;  // device uint64[2][4]: primary, shadow rays; null = off

// L2 read bandwidth probe (theia_l2_probe): every thread streams 128 bit loads that bypass L1 over a buffer that fits
// the L2 cache, `iters` times; the caller times it with events. The xor of everything read lands in *sink.
__global__ void __launch_bounds__(256) l2ProbeKernel(const uint4* buf, uint64_t n16, uint32_t iters, uint32_t* sink) {
    uint4 acc = make_uint4(0u, 0u, 0u, 0u);
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint32_t it = 0; it < iters; ++it)
        for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += stride) {
            const uint4 v = __ldcg(buf + i);
            acc.x ^= v.x; acc.y ^= v.y; acc.z ^= v.z; acc.w ^= v.w;
        }
    const uint32_t x = acc.x ^ acc.y ^ acc.z ^ acc.w;
    if (x == 0x9E3779B9u) atomicXor(sink, x);  // keeps the loads alive; practically never taken
}

// ------------------------------------------------------------------ launch translation
#define HIST_SHARED_MAX_BINS 8192u

static int fillResponse(const theia_response_t& r, ResponseDev& d) {
    memset(&d, 0, sizeof(d));
    d.kind = r.kind;
    if (r.kind == THEIA_RESPONSE_STORE_TIME) {
        if (r.value_kind != THEIA_VALUE_UNIFORM) { setError("store time response: unsupported value response %u", r.value_kind); return -1; }
        if (!r.value_queue || r.value_capacity == 0) { setError("store time response: missing queue"); return -1; }
        d.value_queue = r.value_queue; d.value_capacity = r.value_capacity; d.store_object_id = r.store_object_id;
        d.tag_path = r.tag_path; d.tag_event = r.tag_event;
    } else if (r.kind == THEIA_RESPONSE_OBJECT_HISTOGRAM) {
        if (r.value_kind != THEIA_VALUE_UNIFORM) { setError("object histogram response: unsupported value response %u", r.value_kind); return -1; }
        if (!r.object_hist || r.n_objects == 0 || r.n_bins == 0) { setError("object histogram response: missing histogram buffer"); return -1; }
        if ((uint64_t)r.n_objects * r.n_bins > 0xFFFFFFFFull) { setError("object histogram response: more than 2^32 bins"); return -1; }
        d.t0 = r.hist.t0; d.binSize = r.hist.binSize; d.n_bins = r.n_bins; d.object_hist = r.object_hist; d.n_objects = r.n_objects;
    } else if (r.kind == THEIA_RESPONSE_HISTOGRAM || r.kind == THEIA_RESPONSE_KERNEL_HISTOGRAM) {
        if (r.kind == THEIA_RESPONSE_KERNEL_HISTOGRAM) {
            if (!(r.kernel_bandwidth > 0.0f) || !(r.kernel_support >= 0.0f)) { setError("kernel histogram response: bad kernel parameters"); return -1; }
            d.kernel_bandwidth = r.kernel_bandwidth; d.kernel_support = r.kernel_support;
        }
        if (r.value_kind != THEIA_VALUE_UNIFORM) { setError("histogram response: unsupported value response %u", r.value_kind); return -1; }
        if (!r.hist_partials || r.hist_capacity == 0 || r.n_bins == 0) { setError("histogram response: missing partial histogram buffer"); return -1; }
        d.t0 = r.hist.t0; d.binSize = r.hist.binSize; d.n_bins = r.n_bins; d.hist_capacity = r.hist_capacity;
        d.hist_partials = r.hist_partials;
        d.hist_shared = r.n_bins <= HIST_SHARED_MAX_BINS ? 1u : 0u;
    } else if (r.kind == THEIA_RESPONSE_RECORD) {
        if (!r.queue || r.queue_capacity == 0) { setError("record response: missing queue"); return -1; }
        d.queue = r.queue; d.queue_capacity = r.queue_capacity; d.queue_polarized = r.queue_polarized;
        d.tag_path = r.tag_path; d.tag_event = r.tag_event;
    } else if (r.kind != THEIA_RESPONSE_EMPTY) {
        setError("unknown response kind %u", r.kind);
        return -1;
    }
    return 0;
}

template <typename T>
static int copyBlock(const theia_component_t& c, T& dst, const char* what) {
    if (!c.params || c.params_size != sizeof(T)) {
        setError("%s: parameter block has %u bytes, expected %zu", what, c.params_size, sizeof(T));
        return -1;
    }
    memcpy(&dst, c.params, sizeof(T));
    return 0;
}

// component description (kind + parameter block as the reference's ctypes structure) -> device selector
static int fillTarget(const theia_component_t& c, TargetSel& d) {
    d.kind = c.kind;
    switch (c.kind) {
    case THEIA_TARGET_SPHERE: case THEIA_TARGET_INNER_SPHERE: return copyBlock(c, d.p.sphere, "sphere target");
    case THEIA_TARGET_FLAT: return copyBlock(c, d.p.flat, "flat target");
    case THEIA_TARGET_DISK: return copyBlock(c, d.p.disk, "disk target");
    default: setError("unsupported target kind %u", c.kind); return -1;
    }
}
static int fillGuide(const theia_component_t& c, GuideSel& d) {
    d.kind = c.kind;
    switch (c.kind) {
    case THEIA_GUIDE_NONE: return 0;
    case THEIA_GUIDE_SPHERE: return copyBlock(c, d.p.sphere, "sphere target guide");
    case THEIA_GUIDE_FLAT: return copyBlock(c, d.p.flat, "flat target guide");
    case THEIA_GUIDE_DISK: return copyBlock(c, d.p.disk, "disk target guide");
    default: setError("unsupported guide kind %u", c.kind); return -1;
    }
}
static int fillLight(const theia_component_t& c, uint32_t first, uint32_t n, LightSel& d) {
    d.kind = c.kind; d.flags = c.flags; d.first = first;
    switch (c.kind) {
    case THEIA_LIGHT_SPHERICAL: if (copyBlock(c, d.p.spherical, "spherical light")) return -1; break;
    case THEIA_LIGHT_CONE: if (copyBlock(c, d.p.cone, "cone light")) return -1; break;
    case THEIA_LIGHT_PENCIL: if (copyBlock(c, d.p.pencil, "pencil light")) return -1; break;
    case THEIA_LIGHT_CASCADE: if (copyBlock(c, d.p.cascade, "cascade light")) return -1; break;
    case THEIA_LIGHT_MUON: if (copyBlock(c, d.p.muon, "muon light")) return -1; break;
    case THEIA_LIGHT_CHERENKOV: if (copyBlock(c, d.p.cherenkov, "cherenkov light")) return -1; break;
    case THEIA_LIGHT_CHERENKOV_TRACK:
        if (copyBlock(c, d.p.track, "cherenkov track light")) return -1;
        if (!d.p.track.track) { setError("cherenkov track light: no particle track"); return -1; }
        break;
    case THEIA_LIGHT_HOST:
        if (copyBlock(c, d.p.host, "host light")) return -1;
        if (!d.p.host.queue || d.p.host.capacity < n) { setError("host light: queue holds %u samples, batch needs %u", d.p.host.capacity, n); return -1; }
        break;
    default: setError("unsupported light source kind %u", c.kind); return -1;
    }
    return 0;
}
static int fillWavelength(const theia_component_t& c, uint32_t first, uint32_t n, WavelengthSel& d) {
    d.kind = c.kind; d.first = first;
    switch (c.kind) {
    case THEIA_WAVELENGTH_CONST: if (copyBlock(c, d.p.constant, "const wavelength")) return -1; break;
    case THEIA_WAVELENGTH_UNIFORM: if (copyBlock(c, d.p.uniform, "uniform wavelength")) return -1; break;
    case THEIA_WAVELENGTH_FUNCTION:
        if (copyBlock(c, d.p.function, "function wavelength")) return -1;
        if (!d.p.function.table) { setError("function wavelength: no table"); return -1; }
        break;
    case THEIA_WAVELENGTH_HOST:
        if (copyBlock(c, d.p.host, "host wavelength")) return -1;
        if (!d.p.host.queue || d.p.host.capacity < n) { setError("host wavelength: queue holds %u samples, batch needs %u", d.p.host.capacity, n); return -1; }
        break;
    default: setError("unsupported wavelength source kind %u", c.kind); return -1;
    }
    return 0;
}

static int translate(const theia_trace_launch_t* L, DeviceLaunch& D) {
    if (!L || L->struct_size != sizeof(theia_trace_launch_t)) { setError("launch: struct_size mismatch (ABI version?)"); return -1; }
    memset(&D, 0, sizeof(D));
    D.disable_mis = L->disable_mis; D.disable_direct_lighting = L->disable_direct_lighting;
    D.disable_transmission = L->disable_transmission; D.disable_volume_border = L->disable_volume_border;
    D.hit_use_transmitted_dir = L->hit_use_transmitted_dir; D.path_length = L->path_length;
    D.batch_size = L->batch_size; D.first_path = L->first_path;
    D.rng = L->rng;
    const bool photon = L->tracer == THEIA_TRACER_VOLUME_PHOTON || L->tracer == THEIA_TRACER_SCENE_PHOTON;
    if (photon) {
        if (L->run_length == 0 || L->path_length == 0 || L->path_length % L->run_length != 0) { setError("photon tracer: path_length must be a positive multiple of run_length"); return -1; }
        if (L->guide.kind != THEIA_GUIDE_NONE) { setError("photon tracer: target guides are not used by the photon tracers"); return -1; }
        D.photon = 1u; D.run_length = L->run_length; D.rng_per_run = L->rng_per_run;
    }
    if (L->tracer == THEIA_TRACER_VOLUME_FORWARD || L->tracer == THEIA_TRACER_VOLUME_PHOTON) {
        if (!L->trace_params || L->trace_params_size != sizeof(theia_volume_trace_params_t)) { setError("volume tracer: bad TraceParams block"); return -1; }
        theia_volume_trace_params_t p;
        memcpy(&p, L->trace_params, sizeof(p));
        D.medium = p.medium; D.objectId = p.objectId; D.propagation = p.propagation;
        if (fillTarget(L->target, D.target)) return -1;
    } else if (L->tracer == THEIA_TRACER_SCENE_FORWARD || L->tracer == THEIA_TRACER_SCENE_PHOTON) {
        if (!L->trace_params || L->trace_params_size != sizeof(theia_scene_trace_params_t)) { setError("scene tracer: bad TraceParams block"); return -1; }
        theia_scene_trace_params_t p;
        memcpy(&p, L->trace_params, sizeof(p));
        D.medium = p.sourceMedium; D.objectId = p.targetId; D.propagation = p.propagation;
        if (!L->tlas) { setError("scene tracer: no acceleration structure"); return -1; }
        if (L->path_length == 0) { setError("scene tracer: path_length must be at least 1"); return -1; }
        D.tlas = L->tlas;
        if (fillGuide(L->guide, D.guide)) return -1;
    } else {
        setError("unknown tracer kind %u", L->tracer);
        return -1;
    }
    if (fillLight(L->light, L->first_path, L->batch_size, D.light) || fillWavelength(L->wavelength, L->first_path, L->batch_size, D.wavelength)) return -1;
    if (fillResponse(L->response, D.response)) return -1;
    D.callback.kind = L->callback.kind; D.callback.track_polarized = L->callback.track_polarized;
    D.callback.stats = (unsigned long long*)L->callback.stats; D.callback.track_buffer = L->callback.track_buffer;
    D.callback.track_count = L->callback.track_count; D.callback.track_length = L->callback.track_length;
    if (L->callback.kind == THEIA_CALLBACK_STAT && !L->callback.stats) { setError("statistic callback without buffer"); return -1; }
    if (L->callback.kind == THEIA_CALLBACK_TRACK && !L->callback.track_buffer) { setError("track callback without buffer"); return -1; }
    D.work_counters = (unsigned long long*)L->work_counters;
    return 0;
}

#define TRACE_THREADS 128u
#define QUEUED_THREADS (32u * WQ_WARPS)  // CTA of wfTraceQueuedKernel

struct KernelPlan { const void* fn; size_t smem; uint32_t grid; };

static int plan(const theia_trace_launch_t* L, const DeviceLaunch& D, KernelPlan& K) {
    const bool pol = L->polarized != 0;
    size_t smem = (responseHasHistogram(D.response.kind) && D.response.hist_shared) ? (size_t)D.response.n_bins * 4 : 0;
    if (L->tracer == THEIA_TRACER_VOLUME_FORWARD) {
        // the lean build of the kernel serves the components of the first round (see FULL in components.cuh)
        const bool lean = D.light.kind <= THEIA_LIGHT_MUON && D.wavelength.kind <= THEIA_WAVELENGTH_UNIFORM &&
                          D.target.kind <= THEIA_TARGET_INNER_SPHERE && D.response.kind <= THEIA_RESPONSE_HISTOGRAM;
        if (lean) {
            const bool hist = D.response.kind == THEIA_RESPONSE_HISTOGRAM, nocb = D.callback.kind == THEIA_CALLBACK_EMPTY;
            static const void* const table[2][4] = {
                {(const void*)volumeForwardKernel<false, false, false, false>, (const void*)volumeForwardKernel<false, false, false, true>,
                 (const void*)volumeForwardKernel<false, false, true, false>, (const void*)volumeForwardKernel<false, false, true, true>},
                {(const void*)volumeForwardKernel<true, false, false, false>, (const void*)volumeForwardKernel<true, false, false, true>,
                 (const void*)volumeForwardKernel<true, false, true, false>, (const void*)volumeForwardKernel<true, false, true, true>}};
            K.fn = table[pol ? 1 : 0][(hist ? 2 : 0) | (nocb ? 1 : 0)];
            if (hist && nocb) {  // and the light source, too, for the shape of the BASELINE configurations
                if (D.light.kind == THEIA_LIGHT_SPHERICAL)
                    K.fn = pol ? (const void*)volumeForwardKernel<true, false, true, true, THEIA_LIGHT_SPHERICAL>
                               : (const void*)volumeForwardKernel<false, false, true, true, THEIA_LIGHT_SPHERICAL>;
                else if (D.light.kind == THEIA_LIGHT_CONE)
                    K.fn = pol ? (const void*)volumeForwardKernel<true, false, true, true, THEIA_LIGHT_CONE>
                               : (const void*)volumeForwardKernel<false, false, true, true, THEIA_LIGHT_CONE>;
            }
        } else {
            K.fn = pol ? (const void*)volumeForwardKernel<true, true> : (const void*)volumeForwardKernel<false, true>;
        }
    } else if (L->tracer == THEIA_TRACER_VOLUME_PHOTON) {
        K.fn = pol ? (const void*)volumePhotonKernel<true> : (const void*)volumePhotonKernel<false>;
    } else {
        K.fn = pol ? (const void*)sceneForwardKernel<true> : (const void*)sceneForwardKernel<false>;
        smem += sizeof(WideNode) * THEIA_TOP_NODES;
    }
    K.smem = smem;
    int dev = 0, sms = 0, perSm = 0;
    CUDA_TRY(cudaGetDevice(&dev));
    CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    if (smem > 48 * 1024) CUDA_TRY(cudaFuncSetAttribute(K.fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, K.fn, (int)TRACE_THREADS, smem));
    if (perSm < 1) { setError("tracer kernel does not fit on an SM"); return -1; }
    // persistent grid: a whole number of resident CTAs per SM, never more CTAs than warps of work
    uint32_t full = (uint32_t)sms * (uint32_t)perSm;
    uint32_t need = blocksFor(D.batch_size, TRACE_THREADS);
    K.grid = need < full ? (need > 0 ? need : 1u) : full;
    return 0;
}

// The wavefront kernels exist in a lean instantiation without the components that were added late
// (see FULL in components.cuh); which one a launch needs:
static bool leanLaunch(const DeviceLaunch& D) {
    return D.light.kind <= THEIA_LIGHT_MUON && D.wavelength.kind <= THEIA_WAVELENGTH_UNIFORM &&
           D.guide.kind <= THEIA_GUIDE_SPHERE && D.response.kind <= THEIA_RESPONSE_HISTOGRAM && D.photon == 0;
}
struct WavefrontKernels { const void *gen, *shade, *shadow; };
// Variant tables of the shading kernels (wavefront.cuh, VK_*): lean launches get the instantiation that knows their
// guide / response / callback at compile time, everything else the generic one.
template <bool POL, uint32_t V>
static const void* shadeFn() { return (const void*)wfShadeKernel<POL, V>; }
template <bool POL, uint32_t V>
static const void* shadowFn() { return (const void*)wfShadeShadowKernel<POL, V>; }
template <bool POL>
static WavefrontKernels wavefrontKernelsOf(const DeviceLaunch& D) {
    if (!leanLaunch(D))
        return {(const void*)wfGenerateKernel<POL, true>, shadeFn<POL, VK_FULL>(), shadowFn<POL, VK_FULL>()};
    const bool guide = D.guide.kind != THEIA_GUIDE_NONE, hist = D.response.kind == THEIA_RESPONSE_HISTOGRAM,
               nocb = D.callback.kind == THEIA_CALLBACK_EMPTY;
    const uint32_t key = (guide ? 4u : 0u) | (hist ? 2u : 0u) | (nocb ? 1u : 0u);
    static const void* const shade[8] = {
        shadeFn<POL, 0u>(), shadeFn<POL, VK_NOCB>(), shadeFn<POL, VK_HIST>(), shadeFn<POL, VK_HIST | VK_NOCB>(),
        shadeFn<POL, VK_GUIDE>(), shadeFn<POL, VK_GUIDE | VK_NOCB>(), shadeFn<POL, VK_GUIDE | VK_HIST>(),
        shadeFn<POL, VK_GUIDE | VK_HIST | VK_NOCB>()};
    static const void* const shadow[4] = {
        shadowFn<POL, VK_GUIDE>(), shadowFn<POL, VK_GUIDE | VK_NOCB>(), shadowFn<POL, VK_GUIDE | VK_HIST>(),
        shadowFn<POL, VK_GUIDE | VK_HIST | VK_NOCB>()};
    const void* gen = nocb ? (const void*)wfGenerateKernel<POL, false, true> : (const void*)wfGenerateKernel<POL, false>;
    if (nocb) {
        if (D.light.kind == THEIA_LIGHT_SPHERICAL) gen = (const void*)wfGenerateKernel<POL, false, true, THEIA_LIGHT_SPHERICAL>;
        else if (D.light.kind == THEIA_LIGHT_CONE) gen = (const void*)wfGenerateKernel<POL, false, true, THEIA_LIGHT_CONE>;
        else if (D.light.kind == THEIA_LIGHT_CASCADE) gen = (const void*)wfGenerateKernel<POL, false, true, THEIA_LIGHT_CASCADE>;
    }
    return {gen, shade[key], shadow[key & 3u]};
}
static WavefrontKernels wavefrontKernels(bool pol, const DeviceLaunch& D) {
    return pol ? wavefrontKernelsOf<true>(D) : wavefrontKernelsOf<false>(D);
}

static int persistentGrid(const void* fn, size_t smem, uint32_t* grid, int threads = (int)TRACE_THREADS) {
    // occupancy queries are not free: remember the answer per (device, kernel, shared memory)
    struct Entry { int dev; const void* fn; size_t smem; uint32_t grid; };
    static thread_local Entry cache[16];
    static thread_local int nCache = 0;
    int dev = 0, sms = 0, perSm = 0;
    CUDA_TRY(cudaGetDevice(&dev));
    for (int i = 0; i < nCache; ++i)
        if (cache[i].dev == dev && cache[i].fn == fn && cache[i].smem == smem) { *grid = cache[i].grid; return 0; }
    CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, fn, threads, smem));
    if (perSm < 1) { setError("wavefront kernel does not fit on an SM"); return -1; }
    *grid = (uint32_t)sms * (uint32_t)perSm;
    if (nCache < 16) cache[nCache++] = Entry{dev, fn, smem, *grid};
    return 0;
}

static size_t workspaceBytes(const theia_trace_launch_t* L) {
    if (L->tracer != THEIA_TRACER_SCENE_FORWARD && L->tracer != THEIA_TRACER_SCENE_PHOTON) return 0;
    return layoutWorkspace(nullptr, L->batch_size, L->polarized != 0, L->guide.kind != THEIA_GUIDE_NONE, L->path_length, nullptr);
}

// ------------------------------------------------------------------ options, trace kernel selection
enum : int { TRACE_PHASED = 0, TRACE_SIMPLE = 1, TRACE_QUEUED = 2 };
static int envInt(const char* name, int fallback) {
    const char* e = getenv(name);
    return e ? atoi(e) : fallback;
}
// per ray class (primary, shadow): -1 = pick by measurement; the environment sets the defaults
static int g_optTraceKernel[2] = {envInt("THEIA_WF_TRACE_KERNEL", -1), envInt("THEIA_WF_SHADOW_TRACE_KERNEL", -1)};
static int g_optMinNodeLanes = envInt("THEIA_WF_MIN_NODE_LANES", 16);
static int g_optShadowCull = envInt("THEIA_WF_SHADOW_CULL", 1);  // 0: trace every shadow ray (A/B runs, tests)
static int g_optFlushWaiting = envInt("THEIA_WF_FLUSH_WAITING", 4 | (8 << 8) | (4 << 16));  // thresholds of the queued kernel: waiting lanes | instance lanes << 8 | refill lanes << 16
static int g_optQueryKernel = envInt("THEIA_QUERY_KERNEL", getenv("THEIA_QUERY_WAVEFRONT") ? 1 : 0);  // theia_tlas_query: 0 direct, 1 phased, 2 queued
// Picks the trace kernel of a (device, scene, batch size) from its first nine batches: batch b runs candidate b mod 3
// (phased, one thread per ray, queued), each trace launch bracketed by events; once the events of a timed batch have
// completed (polled, never waited for) they are summed per ray class, and after the ninth the sums of batches 3-8 are
// compared (the first round only warms up: clocks, caches, lazy module loading). All kernels return identical results,
// so tuning is invisible. A scene is known by the address of its TLAS; theia_tlas_build forgets what was learnt about an
// address it writes a new structure to. All of this state is guarded by g_stateMutex, which launchWavefront holds
// while it uses a tuner.
#define TRACE_CANDIDATES 3
struct TraceTuner {
    int dev;
    const void* tlas;
    uint32_t batch;
    bool guide;
    int choice[2] = {-1, -1};
    int step = 0;
    int candidate = 0;
    bool inFlight = false;
    std::vector<std::pair<cudaEvent_t, cudaEvent_t>> pending[2];
    double ms[2][TRACE_CANDIDATES] = {};
};
static std::vector<TraceTuner*> g_tuners;
static void tunerRelease(TraceTuner* t) {
    for (auto& v : t->pending)
        for (auto& e : v) { profRecycle(t->dev, e.first); profRecycle(t->dev, e.second); }
    delete t;
}
static TraceTuner* tunerFor(int dev, const void* tlas, uint32_t batch, bool guide) {
    for (TraceTuner* t : g_tuners)
        if (t->dev == dev && t->tlas == tlas && t->batch == batch && t->guide == guide) return t;
    if (g_tuners.size() >= 64) {  // scenes come and go: forget the oldest entry (nobody holds it: the mutex is ours)
        tunerRelease(g_tuners.front());
        g_tuners.erase(g_tuners.begin());
    }
    TraceTuner* t = new TraceTuner;
    t->dev = dev; t->tlas = tlas; t->batch = batch; t->guide = guide;
    if (!guide) t->choice[1] = TRACE_PHASED;  // no shadow rays to decide about
    g_tuners.push_back(t);
    return t;
}
// called by theia_tlas_build: a new structure at this address is a new scene
void tunerForget(const void* tlas) {
    std::lock_guard<std::recursive_mutex> lock(g_stateMutex);
    const int dev = currentDeviceSlot();
    for (size_t i = 0; i < g_tuners.size();) {
        if (g_tuners[i]->dev == dev && g_tuners[i]->tlas == tlas) { tunerRelease(g_tuners[i]); g_tuners.erase(g_tuners.begin() + i); }
        else ++i;
    }
}
static void tunerPoll(TraceTuner* t) {
    if (!t->inFlight) return;
    for (int c = 0; c < 2; ++c)
        if (!t->pending[c].empty() && cudaEventQuery(t->pending[c].back().second) != cudaSuccess) return;
    for (int c = 0; c < 2; ++c) {
        for (auto& e : t->pending[c]) {
            float ms = 0.0f;
            if (t->step >= TRACE_CANDIDATES && cudaEventElapsedTime(&ms, e.first, e.second) == cudaSuccess) t->ms[c][t->candidate] += ms;
            profRecycle(t->dev, e.first);
            profRecycle(t->dev, e.second);
        }
        t->pending[c].clear();
    }
    t->inFlight = false;
    if (++t->step >= 3 * TRACE_CANDIDATES)
        for (int c = 0; c < 2; ++c)
            if (t->choice[c] < 0) {
                int best = 0;
                for (int k = 1; k < TRACE_CANDIDATES; ++k)
                    if (t->ms[c][k] < t->ms[c][best]) best = k;
                t->choice[c] = best;
            }
}

// Scene tracer, wavefront form: generate, then per segment trace + shade (+ shadow trace +
// shadow shade with a target guide). All launches are asynchronous on the caller's stream.
static int launchWavefront(const theia_trace_launch_t* L, const DeviceLaunch& D) {
    const bool pol = L->polarized != 0, guide = D.guide.kind != THEIA_GUIDE_NONE;
    cudaStream_t st = (cudaStream_t)L->stream;
    if (((uintptr_t)L->workspace & 255u) != 0) { setError("workspace must be 256 byte aligned"); return -1; }
    // the wavefront state keeps the segment number in 16 bits next to its flags (WF_FLAGS)
    if (D.path_length > 0xFFFFu) { setError("scene tracer (wavefront): path_length %u exceeds 65535", D.path_length); return -1; }
    Workspace ws;
    size_t need = layoutWorkspace(L->workspace, D.batch_size, pol, guide, D.path_length, &ws);
    if (L->workspace_bytes < need) {
        setError("workspace too small: %llu bytes given, %llu needed", (unsigned long long)L->workspace_bytes,
                 (unsigned long long)need);
        return -1;
    }
    const size_t histSmem = (responseHasHistogram(D.response.kind) && D.response.hist_shared) ? (size_t)D.response.n_bins * 4 : 0;
    const WavefrontKernels wk = wavefrontKernels(pol, D);
    const void *genFn = wk.gen, *shadeFn = wk.shade, *shadowFn = wk.shadow;
    uint32_t gGen, gShade, gShadow, gTrace, gQueued;
    if (persistentGrid(genFn, 0, &gGen) || persistentGrid(shadeFn, histSmem, &gShade) ||
        persistentGrid(shadowFn, histSmem, &gShadow) || persistentGrid((const void*)wfTraceKernel, 0, &gTrace) ||
        persistentGrid((const void*)wfTraceQueuedKernel, 0, &gQueued, QUEUED_THREADS))
        return -1;
    // partial histograms: one row per CTA of the shading grids
    uint32_t need128 = blocksFor(D.batch_size, TRACE_THREADS);
    if (gGen > need128) gGen = need128;
    if (gShade > need128) gShade = need128;
    if (gShadow > need128) gShadow = need128;
    if (gTrace > need128) gTrace = need128;
    if (gQueued > need128 * (TRACE_THREADS / QUEUED_THREADS)) gQueued = need128 * (TRACE_THREADS / QUEUED_THREADS);
    const uint32_t minNodeLanes = (uint32_t)g_optMinNodeLanes;
    const uint32_t gSimple = need128 > 0 ? need128 : 1u;

    // ---- which trace kernel (per ray class: 0 primary, 1 shadow): forced by option, or timed
    TraceTuner* tuner = nullptr;
    int kernelOf[2] = {g_optTraceKernel[0], g_optTraceKernel[1]};
    bool timeIt = false;
    // held to the end of this function while a tuner is in use: its events are recorded below
    std::unique_lock<std::recursive_mutex> tunerLock(g_stateMutex, std::defer_lock);
    const int devSlot = currentDeviceSlot();
    if (kernelOf[0] < 0 || kernelOf[1] < 0) {
        tunerLock.lock();
        tuner = tunerFor(devSlot, D.tlas, D.batch_size, guide);
        tunerPoll(tuner);
        for (int c = 0; c < 2; ++c) {
            if (kernelOf[c] >= 0) continue;
            if (tuner->choice[c] >= 0) kernelOf[c] = tuner->choice[c];
            else if (!tuner->inFlight) { kernelOf[c] = tuner->step % TRACE_CANDIDATES; timeIt = true; }
            else kernelOf[c] = TRACE_PHASED;  // a timed batch is still running: do not queue up events
        }
        if (timeIt) { tuner->inFlight = true; tuner->candidate = tuner->step % TRACE_CANDIDATES; }
        else tunerLock.unlock();  // decided (or busy): nothing of the tuner is touched below
    }
    // cls 1: shadow rays, a dense list of up to two per path whose results go where the rays say (indirect)
    auto launchTrace = [&](int cls, const float4* rays, uint4* results, const uint32_t* countPtr, uint32_t* cursor) {
        const uint32_t indirect = cls == 1 ? 1u : 0u, rpe = cls == 1 ? 2u : 1u;
        const int32_t targetId = (cls == 1 && g_optShadowCull) ? D.objectId : -1;  // shadow rays: target first
        const float* targetBox = targetId >= 0 ? ws.targetBox : nullptr;
        if (g_traceStats) {  // instrumented pass (measurement only, outside the timed launches)
            wfTraceStatsKernel<<<gSimple < 1184u ? gSimple : 1184u, TRACE_THREADS, 0, st>>>(D.tlas, rays, countPtr, g_traceStats + 4 * cls, targetId, targetBox);
            countLaunches(1);
        }
        Timed tm(cls == 0 ? KC_TRACE : KC_SHADOW_TRACE, st);
        cudaEvent_t e0 = nullptr, e1 = nullptr;
        const bool timed = timeIt && tuner && tuner->choice[cls] < 0 && g_optTraceKernel[cls] < 0;
        if (timed) { e0 = profEvent(devSlot); e1 = profEvent(devSlot); cudaEventRecord(e0, st); }
        if (kernelOf[cls] == TRACE_QUEUED)
            wfTraceQueuedKernel<<<gQueued, QUEUED_THREADS, 0, st>>>(D.tlas, rays, results, countPtr, cursor, indirect,
                                                                  (uint32_t)g_optFlushWaiting, targetId, targetBox);
        else if (kernelOf[cls] == TRACE_SIMPLE)
            wfTraceSimpleKernel<<<rpe * gSimple * (TRACE_THREADS / THEIA_SIMPLE_THREADS), THEIA_SIMPLE_THREADS, 0, st>>>(D.tlas, rays, results, countPtr, indirect, targetId, targetBox);
        else
            wfTraceKernel<<<gTrace, TRACE_THREADS, 0, st>>>(D.tlas, rays, results, countPtr, cursor, indirect, minNodeLanes,
                                                            targetId, targetBox);
        if (timed) { cudaEventRecord(e1, st); tuner->pending[cls].push_back({e0, e1}); }
    };

    // all counters of the batch are cleared here, once (one slot per segment generation: wavefront.cuh)
    CUDA_TRY(cudaMemsetAsync(ws.counters, 0, 4 * WF_SLOT * ((size_t)D.path_length + 1), st));
    auto slot = [&](uint32_t s) { return ws.counters + (size_t)WF_SLOT * s; };
    ws.cullShadow = g_optShadowCull ? 1u : 0u;
    if (guide && ws.cullShadow) {  // the box a shadow ray has to reach to matter (scene and target id may change per batch)
        wfTargetBoxKernel<<<1, 256, 0, st>>>(D.tlas, D.objectId, ws.targetBox);
        countLaunches(1);
    }
    {
        Timed tm(KC_GENERATE, st);
        ws.cntIn = nullptr; ws.cntOut = slot(0) + WF_CNT_QUEUE; ws.cntShadow = nullptr; ws.cntShadowRays = nullptr;
        void* args[] = {(void*)&D, (void*)&ws};
        CUDA_TRY(cudaLaunchKernel(genFn, dim3(gGen), dim3(TRACE_THREADS), args, 0, st));
    }
    uint32_t cur = 0;
    countLaunches(1);
    for (uint32_t it = 1; it <= D.path_length; ++it) {
        countLaunches((guide && it < D.path_length) ? 4 : 2);
        ws.cntIn = slot(it - 1) + WF_CNT_QUEUE; ws.cntOut = slot(it) + WF_CNT_QUEUE; ws.cntShadow = slot(it) + WF_CNT_SHADOW;
        ws.cntShadowRays = slot(it) + WF_CNT_SHADOW_RAYS;
        launchTrace(0, ws.ray[cur], ws.result, ws.cntIn, slot(it) + WF_CNT_CURSOR);
        {
            Timed tm(KC_SHADE, st);
            void* args[] = {(void*)&D, (void*)&ws, (void*)&cur};
            CUDA_TRY(cudaLaunchKernel(shadeFn, dim3(gShade), dim3(TRACE_THREADS), args, histSmem, st));
        }
        if (guide && it < D.path_length) {
            launchTrace(1, ws.shadowLive, ws.shadowResult, ws.cntShadowRays, slot(it) + WF_CNT_SHADOW_CURSOR);
            Timed tm(KC_SHADE_SHADOW, st);
            void* args[] = {(void*)&D, (void*)&ws, (void*)&cur};
            CUDA_TRY(cudaLaunchKernel(shadowFn, dim3(gShadow), dim3(TRACE_THREADS), args, histSmem, st));
        }
        cur ^= 1u;
    }
    CUDA_TRY(cudaGetLastError());
    return 0;
}

}  // namespace theia

using namespace theia;

extern "C" {

const char* theia_last_error(void) { return g_error; }
int theia_abi_version(void) { return THEIA_ABI_VERSION; }
uint64_t theia_kernel_launches(void) { return g_launches.load(); }

int theia_set_option(const char* name, int value) {
    if (!name) { setError("theia_set_option: null name"); return -1; }
    if (!strcmp(name, "wf_trace_kernel")) { g_optTraceKernel[0] = value; return 0; }
    if (!strcmp(name, "wf_shadow_trace_kernel")) { g_optTraceKernel[1] = value; return 0; }
    if (!strcmp(name, "wf_flush_waiting")) { g_optFlushWaiting = value; return 0; }
    if (!strcmp(name, "wf_shadow_cull")) { g_optShadowCull = value != 0; return 0; }
    if (!strcmp(name, "query_kernel")) { g_optQueryKernel = value; return 0; }
    if (!strcmp(name, "wf_min_node_lanes")) { g_optMinNodeLanes = value < 1 ? 1 : (value > 32 ? 32 : value); return 0; }
    setError("theia_set_option: unknown option '%s'", name);
    return -1;
}
int theia_get_trace_choice(const void* tlas_dev, uint32_t batch_size, int guide, int* primary, int* shadow) {
    std::lock_guard<std::recursive_mutex> lock(g_stateMutex);
    const int dev = currentDeviceSlot();
    for (TraceTuner* t : g_tuners)
        if (t->dev == dev && t->tlas == tlas_dev && t->batch == batch_size && t->guide == (guide != 0)) {
            if (primary) *primary = t->choice[0];
            if (shadow) *shadow = t->choice[1];
            return 0;
        }
    if (primary) *primary = -1;
    if (shadow) *shadow = -1;
    return 0;
}

int theia_profile_enable(int on) {
    std::lock_guard<std::recursive_mutex> lock(g_stateMutex);
    profCollect();
    if (on) { for (int k = 0; k < KC_COUNT; ++k) { g_prof.ms[k] = 0.0; g_prof.n[k] = 0; } }
    g_prof.on = on != 0;
    return 0;
}
int theia_profile_read(double* ms, uint64_t* launches, uint32_t n) {
    if (!ms || !launches) { setError("theia_profile_read: null argument"); return -1; }
    std::lock_guard<std::recursive_mutex> lock(g_stateMutex);
    profCollect();
    for (uint32_t k = 0; k < n; ++k) {
        ms[k] = k < KC_COUNT ? g_prof.ms[k] : 0.0;
        launches[k] = k < KC_COUNT ? g_prof.n[k] : 0;
    }
    return KC_COUNT;
}
const char* theia_profile_class_name(uint32_t cls) { return cls < KC_COUNT ? g_classNames[cls] : ""; }

int theia_device_info(int* sm_count, int* cc_major, int* cc_minor) {
    int dev = 0;
    CUDA_TRY(cudaGetDevice(&dev));
    if (sm_count) CUDA_TRY(cudaDeviceGetAttribute(sm_count, cudaDevAttrMultiProcessorCount, dev));
    if (cc_major) CUDA_TRY(cudaDeviceGetAttribute(cc_major, cudaDevAttrComputeCapabilityMajor, dev));
    if (cc_minor) CUDA_TRY(cudaDeviceGetAttribute(cc_minor, cudaDevAttrComputeCapabilityMinor, dev));
    return 0;
}

int theia_philox_fill(const theia_philox_params_t* params, uint32_t baseStream, uint32_t baseCount, uint32_t streams,
                      uint32_t samples, uint32_t dim, float* out_dev, void* stream) {
    if (!params || !out_dev || (dim != 1 && dim != 2)) { setError("theia_philox_fill: bad argument"); return -1; }
    uint64_t n = (uint64_t)streams * samples;
    if (n == 0) return 0;
    philoxFillKernel<<<blocksFor(n, 256), 256, 0, (cudaStream_t)stream>>>(*params, baseStream, baseCount, streams, samples, dim, out_dev);
    countLaunches(1);
    CUDA_TRY(cudaGetLastError());
    return 0;
}

int theia_philox_raw(const uint32_t* ctr_dev, const uint32_t* key_dev, uint32_t n, uint32_t* out_dev, void* stream) {
    if (!ctr_dev || !key_dev || !out_dev) { setError("theia_philox_raw: null argument"); return -1; }
    if (n == 0) return 0;
    philoxRawKernel<<<blocksFor(n, 256), 256, 0, (cudaStream_t)stream>>>(ctr_dev, key_dev, n, out_dev);
    countLaunches(1);
    CUDA_TRY(cudaGetLastError());
    return 0;
}

int theia_math_eval(uint32_t fn, const float* x_dev, const float* y_dev, uint32_t n, float* out_dev, void* stream) {
    if (fn > 6 || !x_dev || !out_dev || (fn == 4 && !y_dev)) { setError("theia_math_eval: bad argument"); return -1; }
    if (n == 0) return 0;
    mathEvalKernel<<<blocksFor(n, 256), 256, 0, (cudaStream_t)stream>>>(fn, x_dev, y_dev, n, out_dev);
    countLaunches(1);
    CUDA_TRY(cudaGetLastError());
    return 0;
}

int theia_hist_reduce(float* partials_dev, uint32_t n_hist, uint32_t n_bins, float norm, float* out_dev, int clear,
                      void* stream) {
    if (!partials_dev || !out_dev) { setError("theia_hist_reduce: null argument"); return -1; }
    if (n_bins == 0) return 0;
    {
        Timed tm(KC_REDUCE, (cudaStream_t)stream);
        histReduceKernel<<<blocksFor(n_bins, 32), 32, 0, (cudaStream_t)stream>>>(partials_dev, n_hist, n_bins, norm, out_dev, clear);
    }
    countLaunches(1);
    CUDA_TRY(cudaGetLastError());
    return 0;
}

int theia_hit_replay_rng(const uint32_t* queue_dev, uint32_t capacity, uint32_t polarized, const theia_response_t* response,
                         const theia_philox_params_t* rng, uint32_t block_size, void* stream) {
    (void)block_size;
    if (!queue_dev || !response) { setError("theia_hit_replay: null argument"); return -1; }
    ResponseDev r;
    if (fillResponse(*response, r)) return -1;
    if (r.kind == THEIA_RESPONSE_STORE_TIME && !rng) { setError("theia_hit_replay: the store time response needs a random number stream (theia_hit_replay_rng)"); return -1; }
    theia_philox_params_t rp;
    memset(&rp, 0, sizeof(rp));
    if (rng) rp = *rng;
    if (capacity == 0) return 0;
    size_t smem = (responseHasHistogram(r.kind) && r.hist_shared) ? (size_t)r.n_bins * 4 : 0;
    uint32_t grid = blocksFor(capacity, 128);
    if (responseHasHistogram(r.kind) && grid > r.hist_capacity) grid = r.hist_capacity;
    if (grid > 148 * 8) grid = 148 * 8;
    hitReplayKernel<<<grid, 128, smem, (cudaStream_t)stream>>>(queue_dev, capacity, polarized, r, rp, rng ? 1u : 0u);
    countLaunches(1);
    CUDA_TRY(cudaGetLastError());
    return 0;
}
int theia_hit_replay(const uint32_t* queue_dev, uint32_t capacity, uint32_t polarized, const theia_response_t* response,
                     uint32_t block_size, void* stream) {
    return theia_hit_replay_rng(queue_dev, capacity, polarized, response, nullptr, block_size, stream);
}

int theia_tlas_query(const void* tlas_dev, const float* rays_dev, uint32_t n, int32_t* out_dev, float* out_bary_dev,
                     uint64_t* stats_dev, void* stream) {
    if (!tlas_dev || !rays_dev || !out_dev || !out_bary_dev) { setError("theia_tlas_query: null argument"); return -1; }
    if (n == 0) return 0;
    if (g_optQueryKernel > 0) {
        // measurement seam: the ray queries of the wavefront tracer on a caller supplied ray list
        cudaStream_t st = (cudaStream_t)stream;
        uint32_t* counters = nullptr;
        uint4* results = nullptr;
        CUDA_TRY(cudaMallocAsync(&counters, 16, st));
        CUDA_TRY(cudaMallocAsync(&results, 16 * (size_t)n, st));
        uint32_t gTrace = 0;
        const bool queued = g_optQueryKernel == 2;
        if (persistentGrid(queued ? (const void*)wfTraceQueuedKernel : (const void*)wfTraceKernel, 0, &gTrace,
                           queued ? (int)QUEUED_THREADS : (int)TRACE_THREADS)) return -1;
        const uint32_t need = blocksFor(n, queued ? QUEUED_THREADS : TRACE_THREADS);
        if (gTrace > need) gTrace = need;
        wfQuerySetupKernel<<<1, 32, 0, st>>>(counters, n);
        {
            Timed tm(KC_TRACE, st);
            if (queued)
                wfTraceQueuedKernel<<<gTrace, QUEUED_THREADS, 0, st>>>(tlas_dev, reinterpret_cast<const float4*>(rays_dev), results,
                                                                      counters, counters + 1, 0u, (uint32_t)g_optFlushWaiting, -1, nullptr);
            else
                wfTraceKernel<<<gTrace, TRACE_THREADS, 0, st>>>(tlas_dev, reinterpret_cast<const float4*>(rays_dev), results,
                                                                counters, counters + 1, 0u, 16u, -1, nullptr);
        }
        wfQueryConvertKernel<<<blocksFor(n, 128), 128, 0, st>>>(tlas_dev, rays_dev, results, n, out_dev, out_bary_dev);
        CUDA_TRY(cudaFreeAsync(counters, st));
        CUDA_TRY(cudaFreeAsync(results, st));
        countLaunches(3);
        CUDA_TRY(cudaGetLastError());
        return 0;
    }
    {
        Timed tm(KC_SCENE, (cudaStream_t)stream);
        tlasQueryKernel<<<blocksFor(n, 128), 128, 0, (cudaStream_t)stream>>>(tlas_dev, rays_dev, n, out_dev, out_bary_dev,
                                                                             (unsigned long long*)stats_dev);
    }
    countLaunches(1);
    CUDA_TRY(cudaGetLastError());
    return 0;
}

int theia_trace_grid_size(const theia_trace_launch_t* launch, uint32_t* n_blocks) {
    DeviceLaunch D;
    KernelPlan K;
    if (!n_blocks) { setError("theia_trace_grid_size: null argument"); return -1; }
    if (translate(launch, D)) return -1;
    if ((launch->tracer == THEIA_TRACER_SCENE_FORWARD || launch->tracer == THEIA_TRACER_SCENE_PHOTON) && launch->workspace) {
        // wavefront: the shading grids are the ones that write partial histograms
        const bool pol = launch->polarized != 0;
        const size_t smem = (responseHasHistogram(D.response.kind) && D.response.hist_shared) ? (size_t)D.response.n_bins * 4 : 0;
        uint32_t gShade = 0, gShadow = 0;
        const WavefrontKernels wk = wavefrontKernels(pol, D);
        if (persistentGrid(wk.shade, smem, &gShade) || persistentGrid(wk.shadow, smem, &gShadow))
            return -1;
        const uint32_t need = blocksFor(D.batch_size, TRACE_THREADS);
        uint32_t g = gShade > gShadow ? gShade : gShadow;
        *n_blocks = g < need ? g : (need > 0 ? need : 1u);
        return 0;
    }
    if (plan(launch, D, K)) return -1;
    *n_blocks = K.grid;
    return 0;
}

int theia_trace_workspace_size(const theia_trace_launch_t* launch, uint64_t* bytes) {
    if (!launch || !bytes || launch->struct_size != sizeof(theia_trace_launch_t)) {
        setError("theia_trace_workspace_size: bad argument");
        return -1;
    }
    *bytes = workspaceBytes(launch);
    return 0;
}

int theia_trace_forward(const theia_trace_launch_t* launch) {
    DeviceLaunch D;
    KernelPlan K;
    if (translate(launch, D)) return -1;
    if (D.batch_size == 0) return 0;
    const bool sceneTracer = launch->tracer == THEIA_TRACER_SCENE_FORWARD || launch->tracer == THEIA_TRACER_SCENE_PHOTON;
    if (sceneTracer && launch->workspace) return launchWavefront(launch, D);
    if (plan(launch, D, K)) return -1;
    void* args[] = {&D};
    {
        Timed tm(sceneTracer ? KC_SCENE : KC_VOLUME, (cudaStream_t)launch->stream);
        CUDA_TRY(cudaLaunchKernel(K.fn, dim3(K.grid), dim3(TRACE_THREADS), args, K.smem, (cudaStream_t)launch->stream));
    }
    countLaunches(1);
    return 0;
}

int theia_trace_stats_enable(uint64_t* stats_dev) {
    g_traceStats = (unsigned long long*)stats_dev;
    return 0;
}

int theia_l2_probe(const void* buf_dev, uint64_t bytes, uint32_t iters, uint32_t* sink_dev, void* stream) {
    if (!buf_dev || !sink_dev || bytes < 16 || ((uintptr_t)buf_dev & 15u)) { setError("theia_l2_probe: bad argument"); return -1; }
    int dev = 0, sms = 0;
    CUDA_TRY(cudaGetDevice(&dev));
    CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    l2ProbeKernel<<<(uint32_t)sms * 8u, 256, 0, (cudaStream_t)stream>>>((const uint4*)buf_dev, bytes / 16, iters, sink_dev);
    countLaunches(1);
    CUDA_TRY(cudaGetLastError());
    return 0;
}

int theia_unit_widths(uint32_t unit, uint32_t polarized, uint32_t* in_width, uint32_t* out_width) {
    (void)polarized;
    if (!in_width || !out_width || unitWidths(unit, in_width, out_width)) { setError("theia_unit_widths: unknown unit %u", unit); return -1; }
    return 0;
}

int theia_unit_eval(const theia_unit_args_t* args) {
    if (!args || args->struct_size != sizeof(theia_unit_args_t)) { setError("theia_unit_eval: struct_size mismatch (ABI version?)"); return -1; }
    UnitDev A;
    memset(&A, 0, sizeof(A));
    if (unitWidths(args->unit, &A.in_width, &A.out_width)) { setError("theia_unit_eval: unknown unit %u", args->unit); return -1; }
    A.unit = args->unit; A.n = args->n; A.polarized = args->polarized; A.first = args->first; A.base_count = args->base_count;
    A.in = args->in; A.out = args->out; A.rng = args->rng; A.medium = args->medium; A.material = args->material;
    memcpy(A.params, args->params, sizeof(A.params));
    if (A.n == 0) return 0;
    const uint32_t u = args->unit;
    if (u == THEIA_UNIT_LIGHT_SAMPLE) {
        if (!args->light_queue || !args->wavelength_queue || args->n > args->capacity) { setError("light sampler: %u samples do not fit queues of %u", args->n, args->capacity); return -1; }
        A.light_queue = args->light_queue; A.wavelength_queue = args->wavelength_queue; A.capacity = args->capacity;
        if (fillWavelength(args->wavelength, args->first, args->n, A.wavelength)) return -1;
    } else if (!A.out || (A.in_width && !A.in)) { setError("theia_unit_eval: null buffer"); return -1; }
    if ((u == THEIA_UNIT_LIGHT_SAMPLE || u == THEIA_UNIT_SURFACE) && fillLight(args->light, args->first, args->n, A.light)) return -1;
    if (u == THEIA_UNIT_TARGET_SAMPLE && fillTarget(args->target, A.target)) return -1;
    if (u == THEIA_UNIT_GUIDE_SAMPLE && (args->guide.kind == THEIA_GUIDE_NONE || fillGuide(args->guide, A.guide))) { if (args->guide.kind == THEIA_GUIDE_NONE) setError("guide sampler: no target guide"); return -1; }
    if ((u == THEIA_UNIT_REFLECTANCE || u == THEIA_UNIT_SURFACE) && !A.material) { setError("theia_unit_eval: unit %u needs a material", u); return -1; }
    cudaStream_t st = (cudaStream_t)args->stream;
    if (A.polarized) unitEvalKernel<true><<<blocksFor(A.n, 128), 128, 0, st>>>(A);
    else unitEvalKernel<false><<<blocksFor(A.n, 128), 128, 0, st>>>(A);
    countLaunches(1);
    CUDA_TRY(cudaGetLastError());
    return 0;
}

}  // extern "C"

// ------------------------------------------------------------------ ray queries without any allocation
// theia_tlas_query_static: the ray query seam through the queued wavefront kernel (trace_queue.cuh), whose
// scratch memory (ray counter, work cursor, one result record per ray) lives in a file-scope device array
// instead of cudaMallocAsync memory, for callers that must not allocate (CUDA graph capture, serving loops).
// Every GPU that loads the library holds the array (4 MiB + 256 B) until the process exits.
#define THEIA_QUERY_STATIC_MAX_RAYS (1u << 18)
static __device__ __align__(256) unsigned char g_queryScratch[256 + 16ull * THEIA_QUERY_STATIC_MAX_RAYS];

// Device address of g_queryScratch on the current device. In host code the array's name is only a stand-in
// for the symbol, not an address a kernel may use: the address is asked for once per device.
static int queryScratchAddress(unsigned char** out) {
    static std::mutex mutex;
    static unsigned char* perDevice[64] = {};
    int dev = 0;
    CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) { setError("theia_tlas_query_static: device ordinal %d not supported", dev); return -1; }
    std::lock_guard<std::mutex> lock(mutex);
    if (!perDevice[dev]) {
        unsigned char* p = nullptr;
        CUDA_TRY(cudaGetSymbolAddress((void**)&p, g_queryScratch));
        if (!p) { setError("theia_tlas_query_static: no address for the scratch array"); return -1; }
        perDevice[dev] = p;
    }
    *out = perDevice[dev];
    return 0;
}

extern "C" int theia_tlas_query_static(const void* tlas_dev, const float* rays_dev, uint32_t n, int32_t* out_dev,
                                       float* out_bary_dev, void* stream) {
    if (!tlas_dev || !rays_dev || !out_dev || !out_bary_dev) { setError("theia_tlas_query_static: null argument"); return -1; }
    if (n > THEIA_QUERY_STATIC_MAX_RAYS) {
        setError("theia_tlas_query_static: %u rays exceed the static scratch array (at most %u per call)", n,
                 THEIA_QUERY_STATIC_MAX_RAYS);
        return -1;
    }
    if (n == 0) return 0;
    unsigned char* scratch = nullptr;
    if (queryScratchAddress(&scratch)) return -1;
    uint32_t* counters = reinterpret_cast<uint32_t*>(scratch);          // [0] rays, [1] cursor: rewritten by every call
    uint4* results = reinterpret_cast<uint4*>(scratch + 256);           // [n]: every entry < n is written before it is read
    cudaStream_t st = (cudaStream_t)stream;
    uint32_t grid = 0;
    if (persistentGrid((const void*)wfTraceQueuedKernel, 0, &grid, (int)QUEUED_THREADS)) return -1;
    const uint32_t need = blocksFor(n, QUEUED_THREADS);
    if (grid > need) grid = need;
    wfQuerySetupKernel<<<1, 32, 0, st>>>(counters, n);
    wfTraceQueuedKernel<<<grid, QUEUED_THREADS, 0, st>>>(tlas_dev, reinterpret_cast<const float4*>(rays_dev), results, counters,
                                                        counters + 1, 0u, (uint32_t)g_optFlushWaiting, -1, nullptr);
    wfQueryConvertKernel<<<blocksFor(n, 128), 128, 0, st>>>(tlas_dev, rays_dev, results, n, out_dev, out_bary_dev);
    countLaunches(3);
    CUDA_TRY(cudaGetLastError());
    return 0;
}
