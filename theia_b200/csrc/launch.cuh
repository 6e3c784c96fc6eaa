// Device-side launch description (the kernel parameter block) and the per-CTA sinks for
// hits and trace events: warp-aggregated shared-memory histogram, warp-aggregated hit queue
// slots, warp-aggregated event counters, track recorder
// (sh/response.histogram.glsl, sh/response.record.glsl, sh/response.queue.glsl,
// sh/callback.stat.glsl, sh/callback.track.glsl of the reference).
#pragma once
#include "components.cuh"

namespace theia {

#define FULL_MASK 0xFFFFFFFFu

struct ResponseDev {
    uint32_t kind;  // THEIA_RESPONSE_*
    float t0, binSize;
    uint32_t n_bins;
    uint32_t hist_capacity;
    uint32_t hist_shared;  // 1: CTA-local histogram lives in shared memory
    float* hist_partials;
    uint32_t* queue;
    uint32_t queue_capacity;
    uint32_t queue_polarized;
    uint32_t* tag_path;
    uint32_t* tag_event;
    float kernel_bandwidth, kernel_support;  // kernel histogram
    uint32_t* value_queue;                   // store time
    uint32_t value_capacity, store_object_id;
    float* object_hist;                      // object histogram
    uint32_t n_objects;
};
__host__ __device__ inline bool responseHasHistogram(uint32_t kind) { return kind == THEIA_RESPONSE_HISTOGRAM || kind == THEIA_RESPONSE_KERNEL_HISTOGRAM; }
struct CallbackDev {
    uint32_t kind;
    uint32_t track_polarized;
    unsigned long long* stats;
    void* track_buffer;
    uint32_t track_count, track_length;
};

// One launch of a forward tracer, passed as a __grid_constant__ kernel parameter.
struct DeviceLaunch {
    uint32_t disable_mis, disable_direct_lighting, disable_transmission, disable_volume_border;
    uint32_t hit_use_transmitted_dir, path_length, batch_size, first_path;
    uint32_t photon, run_length, rng_per_run;  // photon tracers (absorption roulette, per run draw counters)
    theia_philox_params_t rng;
    // TraceParams of either tracer (trace.py:589-598, 1136-1145)
    uint64_t medium;   // volume: medium; scene: sourceMedium
    int32_t objectId;  // volume: objectId; scene: targetId
    PropagateParams propagation;
    LightSel light;
    WavelengthSel wavelength;
    TargetSel target;
    GuideSel guide;
    ResponseDev response;
    CallbackDev callback;
    const void* tlas;
    unsigned long long* work_counters;  // [segments, responses, shadow rays, paths] or null
};

TDEV uint32_t laneId() { return threadIdx.x & 31u; }
TDEV uint32_t laneMaskLt() {
    uint32_t m;
    asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
    return m;
}

// ------------------------------------------------------------------ CTA-shared sink state
struct SinkShared {
    unsigned long long stats[12];
    unsigned long long work[4];
};

// sh/response.queue.glsl:94-133 (SAVE_HIT): field-major SoA behind the count word
TDEV void saveHit(uint32_t* queue, uint32_t cap, bool qpol, uint32_t k, const HitItem& h) {
    float* f = reinterpret_cast<float*>(queue + 1);
    size_t o = k;
    f[o] = h.position.x; o += cap; f[o] = h.position.y; o += cap; f[o] = h.position.z; o += cap;
    f[o] = h.direction.x; o += cap; f[o] = h.direction.y; o += cap; f[o] = h.direction.z; o += cap;
    f[o] = h.normal.x; o += cap; f[o] = h.normal.y; o += cap; f[o] = h.normal.z; o += cap;
    if (qpol) {
        f[o] = h.stokes.x; o += cap; f[o] = h.stokes.y; o += cap; f[o] = h.stokes.z; o += cap;
        f[o] = h.stokes.w; o += cap;
        f[o] = h.polRef.x; o += cap; f[o] = h.polRef.y; o += cap; f[o] = h.polRef.z; o += cap;
    }
    f[o] = h.wavelength; o += cap; f[o] = h.time; o += cap; f[o] = h.contrib; o += cap;
    reinterpret_cast<int32_t*>(f)[o] = h.objectId;
}
TDEV HitItem loadHit(const uint32_t* queue, uint32_t cap, bool qpol, uint32_t k) {
    const float* f = reinterpret_cast<const float*>(queue + 1);
    size_t o = k;
    HitItem h;
    h.position.x = f[o]; o += cap; h.position.y = f[o]; o += cap; h.position.z = f[o]; o += cap;
    h.direction.x = f[o]; o += cap; h.direction.y = f[o]; o += cap; h.direction.z = f[o]; o += cap;
    h.normal.x = f[o]; o += cap; h.normal.y = f[o]; o += cap; h.normal.z = f[o]; o += cap;
    if (qpol) {
        h.stokes.x = f[o]; o += cap; h.stokes.y = f[o]; o += cap; h.stokes.z = f[o]; o += cap;
        h.stokes.w = f[o]; o += cap;
        h.polRef.x = f[o]; o += cap; h.polRef.y = f[o]; o += cap; h.polRef.z = f[o]; o += cap;
    } else {
        h.stokes = {1.0f, 0.0f, 0.0f, 0.0f};
        h.polRef = v3(0.0f);
    }
    h.wavelength = f[o]; o += cap; h.time = f[o]; o += cap; h.contrib = f[o]; o += cap;
    h.objectId = reinterpret_cast<const int32_t*>(f)[o];
    return h;
}

// sh/response.histogram.glsl:59-68: bin = uint(int(floor((t - t0) / binSize)))
TDEV int32_t floatToInt(float f) {  // GLSL int(): out of range and NaN are undefined there, INT_MIN here
    return (f >= 2147483648.0f || f < -2147483648.0f || f != f) ? INT32_MIN : (int32_t)f;
}
TDEV uint32_t histogramBin(const ResponseDev& r, float time) {
    return (uint32_t)floatToInt(floorf((time - r.t0) / r.binSize));
}
// sh/response.histogram.kernel.glsl:62-70: normal CDF after Vazquez-Leal et al. (2011), eq. 4.3
TDEV float kernelCdf(float x) {
    const float c1 = 7.779374467827938f, c2 = 55.5f, c3 = 0.1257926109373887f;
    return 0.5f + 0.5f * dm_tanh(c1 * x - c2 * dm_atan(c3 * x));
}

// Commits one optional hit per lane. MUST be called by all 32 lanes of the warp.
//  - histogram: lanes hitting the same bin are combined with match.any and one lane issues
//    the shared-memory float atomic (float atomics on shared memory are CAS loops, so
//    same-bin collisions inside a warp are the expensive case, e.g. the direct-light peak).
//  - record: one global atomic per warp hands out a contiguous run of queue slots.
//  - kernel histogram: every lane walks the bins its hit touches (plain shared-memory atomics).
//  - store time: one draw of the path's stream per hit decides whether it is kept; slots like record.
//  - object histogram: lanes with the same (object, bin) are combined, one global float atomic each.
// The store-time response consumes one number of the path's stream per hit; it has to be drawn
// where the reference calls response() (other draws of the path may follow before the warp
// reaches commitHit): responseDraw() at that point, its value handed to commitHit.
static __device__ __noinline__ float randomOutOfLine(const theia_philox_params_t* p, PhiloxKey key, uint32_t dim) {
    return random_s(*p, key, dim);  // a whole Philox block: not worth inlining at every response site
}
template <bool FULL = true>
TDEV float responseDraw(const ResponseDev& r, Rng& rng, bool has, const HitItem& item) {
    if (!FULL) return 2.0f;
    if (r.kind != THEIA_RESPONSE_STORE_TIME || !has || !(item.contrib > 0.0f)) return 2.0f;
    float u = randomOutOfLine(rng.p, rng.key, rng.dim);
    rng.dim += 1;
    return u;
}
// Responses other than record / histogram live out of line: inlined at every call site they made
// the tracing kernels (which already run at the limit of the instruction cache) 30 % larger and
// measurably slower. Scalars only: handing over the hit by reference would pin it to local memory.
static __device__ __noinline__ void commitHitOther(const ResponseDev& r, float* hist, bool has, float time, float contrib,
                                                   int32_t objectId, uint32_t path, uint32_t tag, float u) {
    if (r.kind == THEIA_RESPONSE_KERNEL_HISTOGRAM) {  // sh/response.histogram.kernel.glsl:72-101
        if (!has) return;
        const float value = contrib;  // sh/response.uniform.glsl:4-6
        float t = time - r.t0;
        int32_t firstBin = floatToInt(floorf((t - r.kernel_support) / r.binSize));
        int32_t lastBin = floatToInt(ceilf((t + r.kernel_support) / r.binSize));
        firstBin = max(firstBin, 0);
        lastBin = min(lastBin, (int32_t)r.n_bins) - 1;
        t = (float)firstBin * r.binSize + r.t0 - time;
        const float h = 1.0f / r.kernel_bandwidth;
        float prev_cdf = kernelCdf(t * h);
        for (int32_t i = firstBin; i <= lastBin; ++i) {
            t += r.binSize;
            float cdf = kernelCdf(t * h);
            float w = cdf - prev_cdf;
            prev_cdf = cdf;
            atomicAdd(hist + i, w * value);
        }
    } else if (r.kind == THEIA_RESPONSE_STORE_TIME) {  // sh/response.time.store.glsl:14-26
        bool keep = false;
        if (has) keep = u < contrib;
        uint32_t act = __ballot_sync(FULL_MASK, keep);
        if (act == 0) return;
        uint32_t leader = __ffs(act) - 1;
        uint32_t base = 0;
        if (laneId() == leader) base = atomicAdd(r.value_queue, (uint32_t)__popc(act));
        base = __shfl_sync(FULL_MASK, base, leader);
        if (!keep) return;
        uint32_t slot = base + __popc(act & laneMaskLt());
        if (slot < r.value_capacity) {
            reinterpret_cast<float*>(r.value_queue + 1)[slot] = time;
            if (r.store_object_id) reinterpret_cast<int32_t*>(r.value_queue + 1)[(size_t)r.value_capacity + slot] = objectId;
            if (r.tag_path) r.tag_path[slot] = path;
            if (r.tag_event) r.tag_event[slot] = tag;
        }
    } else if (r.kind == THEIA_RESPONSE_OBJECT_HISTOGRAM) {
        uint32_t key = 0;
        if (has) {
            uint32_t bin = histogramBin(r, time);
            has = bin < r.n_bins && (uint32_t)objectId < r.n_objects;
            key = (uint32_t)objectId * r.n_bins + bin;
        }
        uint32_t act = __ballot_sync(FULL_MASK, has);
        if (act == 0 || !has) return;
        uint32_t peers = __match_any_sync(act, key);
        float sum = 0.0f;
        for (uint32_t m = peers; m; m &= m - 1) sum += __shfl_sync(peers, contrib, __ffs(m) - 1);
        if (laneId() == (uint32_t)(__ffs(peers) - 1)) atomicAdd(r.object_hist + key, sum);
    }
}
// HIST: the launch's response is known to be the plain histogram (compile time variant of the wavefront kernels)
template <bool FULL, bool HIST = false>
TDEV void commitHitImpl(const ResponseDev& r, float* hist, bool has, const HitItem& item, uint32_t path, uint32_t tag,
                        float u) {
    if (HIST || r.kind == THEIA_RESPONSE_HISTOGRAM) {
        uint32_t bin = 0;
        if (has) {
            bin = histogramBin(r, item.time);
            has = bin < r.n_bins;
        }
        uint32_t act = __ballot_sync(FULL_MASK, has);
        if (act == 0 || !has) return;
        float value = item.contrib;  // sh/response.uniform.glsl:4-6
        uint32_t peers = __match_any_sync(act, bin);
        uint32_t leader = __ffs(peers) - 1;
        float sum = 0.0f;
        for (uint32_t m = peers; m; m &= m - 1) sum += __shfl_sync(peers, value, __ffs(m) - 1);
        if (laneId() == leader) atomicAdd(hist + bin, sum);
    } else if (!HIST && r.kind == THEIA_RESPONSE_RECORD) {
        uint32_t act = __ballot_sync(FULL_MASK, has);
        if (act == 0) return;
        uint32_t leader = __ffs(act) - 1;
        uint32_t base = 0;
        if (laneId() == leader) base = atomicAdd(r.queue, (uint32_t)__popc(act));
        base = __shfl_sync(FULL_MASK, base, leader);
        if (!has) return;
        uint32_t slot = base + __popc(act & laneMaskLt());
        if (slot < r.queue_capacity) {
            saveHit(r.queue, r.queue_capacity, r.queue_polarized != 0, slot, item);
            if (r.tag_path) r.tag_path[slot] = path;
            if (r.tag_event) r.tag_event[slot] = tag;
        }
    } else if (!HIST && FULL && r.kind != THEIA_RESPONSE_EMPTY) {
        commitHitOther(r, hist, has, item.time, item.contrib, item.objectId, path, tag, u);
    }
}

// Returns whether the lane's hit reached the response (the call sites number a path's responses
// inside an event with it: test tags).
template <bool FULL = true, bool HIST = false>
TDEV bool commitHit(const ResponseDev& r, float* hist, bool has, const HitItem& item, uint32_t path, uint32_t tag,
                    uint32_t& nResponses, float u) {
    has = has && item.contrib > 0.0f;  // "if (hit.contrib > 0.0) response(hit)" at every call site
    if (has) nResponses++;
    commitHitImpl<FULL, HIST>(r, hist, has, item, path, tag, u);
    return has;
}
// sh/callback.stat.glsl:21-62: counter index of a result code (-1: not counted by name)
TDEV int statIndex(int code) {
    switch (code) {
    case THEIA_RESULT_RAY_CREATED: return 0;
    case THEIA_RESULT_RAY_SCATTERED: return 1;
    case THEIA_RESULT_RAY_HIT: return 2;
    case THEIA_RESULT_RAY_DETECTED: return 3;
    case THEIA_RESULT_VOLUME_HIT: return 4;
    case THEIA_RESULT_RAY_LOST: return 5;
    case THEIA_RESULT_RAY_DECAYED: return 6;
    case THEIA_RESULT_RAY_ABSORBED: return 7;
    case THEIA_RESULT_RAY_MISSED: return 8;
    case THEIA_RESULT_MAX_ITER: return 9;
    case THEIA_ERROR_MEDIA_MISMATCH: return 11;
    default: return -1;
    }
}

// onEvent(ray, code, idx, i). The statistic part is warp aggregated and MUST be called by
// all 32 lanes (has = false for lanes without an event); the track part is a plain store.
// NOCB: the launch is known to have no callback (compile time variant of the wavefront kernels)
template <bool POL, bool NOCB = false>
TDEV void onEvent(const CallbackDev& cb, SinkShared* sh, bool has, const Ray& ray, int code, uint32_t k, uint32_t i) {
    if (NOCB) return;
    if (cb.kind == THEIA_CALLBACK_STAT) {
        uint32_t act = __ballot_sync(FULL_MASK, has);
        if (act == 0 || !has) return;
        uint32_t peers = __match_any_sync(act, code);
        if (laneId() == (uint32_t)(__ffs(peers) - 1)) {
            int s = statIndex(code);
            unsigned long long n = (unsigned long long)__popc(peers);
            if (s >= 0) atomicAdd(&sh->stats[s], n);
            if (code <= THEIA_ERROR_MAX_VALUE) atomicAdd(&sh->stats[10], n);
        }
    } else if (cb.kind == THEIA_CALLBACK_TRACK) {
        if (!has || code == THEIA_RESULT_MAX_ITER) return;
        uint32_t T = cb.track_count, Ln = cb.track_length;
        if (k >= T) return;
        uint32_t* n = reinterpret_cast<uint32_t*>(cb.track_buffer);
        int32_t* codes = reinterpret_cast<int32_t*>(n + T);
        float* base = reinterpret_cast<float*>(codes + T);
        n[k] = i;
        codes[k] = code;
        if (i >= Ln) return;
        size_t plane = (size_t)Ln * T;
        size_t o = (size_t)i * T + k;
        base[0 * plane + o] = ray.position.x;
        base[1 * plane + o] = ray.position.y;
        base[2 * plane + o] = ray.position.z;
        base[3 * plane + o] = ray.time;
        if (cb.track_polarized) {
            vec4 st = POL ? ray.stokes : vec4{1.0f, 0.0f, 0.0f, 0.0f};
            vec3 pr = POL ? ray.polRef : v3(0.0f);
            base[4 * plane + o] = st.x; base[5 * plane + o] = st.y;
            base[6 * plane + o] = st.z; base[7 * plane + o] = st.w;
            base[8 * plane + o] = pr.x; base[9 * plane + o] = pr.y; base[10 * plane + o] = pr.z;
        }
    }
}

// Photon tracers, end of an interaction (sh/tracer.scene.photon.init.glsl:90-106,
// sh/tracer.volume.photon.common.glsl:88-104): the contribution is the survival probability so
// far; play roulette on it and start over with one.
template <bool POL>
TDEV int photonRoulette(Ray& ray, Rng& rng, int result) {
    if (result < 0) return result;
    if (POL) {
        ray.lin_contrib *= ray.stokes.x;
        ray.stokes = ray.stokes / ray.stokes.x;
    }
    if (getContrib(ray) <= rng.next()) return THEIA_RESULT_RAY_ABSORBED;
    ray.lin_contrib = 1.0f;
    ray.log_contrib = 0.0f;
    return result;
}
// The reference runs photon paths in shader dispatches of run_length events and starts the draw
// counter of run r at r * rng_per_run: applied when event `it` (1 based) was the last of a run.
TDEV void photonRunBoundary(const DeviceLaunch& P, Rng& rng, uint32_t it) {
    if (P.run_length != 0 && it % P.run_length == 0) rng.dim = (it / P.run_length) * P.rng_per_run;
}

// CTA prologue / epilogue shared by both tracers.
TDEV void sinkInit(const DeviceLaunch& P, SinkShared* sh, float* hist) {
    for (uint32_t i = threadIdx.x; i < 16; i += blockDim.x) {
        if (i < 12) sh->stats[i] = 0; else sh->work[i - 12] = 0;
    }
    if (hist && responseHasHistogram(P.response.kind) && P.response.hist_shared)
        for (uint32_t b = threadIdx.x; b < P.response.n_bins; b += blockDim.x) hist[b] = 0.0f;
    __syncthreads();
}
TDEV float* sinkHist(const DeviceLaunch& P, float* sharedHist) {
    // n_bins too large for shared memory: accumulate straight into this CTA's partial
    if (!responseHasHistogram(P.response.kind)) return nullptr;
    if (P.response.hist_shared) return sharedHist;
    return P.response.hist_partials + (size_t)(blockIdx.x % P.response.hist_capacity) * P.response.n_bins;
}
TDEV void sinkFinish(const DeviceLaunch& P, SinkShared* sh, const float* hist, uint32_t nSegments,
                     uint32_t nResponses, uint32_t nShadow, uint32_t nPaths) {
    if (P.work_counters) {
        uint32_t v[4] = {nSegments, nResponses, nShadow, nPaths};
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            uint32_t s = __reduce_add_sync(FULL_MASK, v[k]);
            if (laneId() == 0 && s) atomicAdd(&sh->work[k], (unsigned long long)s);
        }
    }
    __syncthreads();
    if (hist && responseHasHistogram(P.response.kind) && P.response.hist_shared) {
        // sh/response.histogram.glsl:35-45: one partial histogram per work group
        const bool exclusive = gridDim.x <= P.response.hist_capacity;
        float* dst = P.response.hist_partials + (size_t)(blockIdx.x % P.response.hist_capacity) * P.response.n_bins;
        for (uint32_t b = threadIdx.x; b < P.response.n_bins; b += blockDim.x) {
            float v = hist[b];
            if (exclusive) dst[b] += v;
            else if (v != 0.0f) atomicAdd(dst + b, v);
        }
    }
    if (threadIdx.x < 12 && P.callback.kind == THEIA_CALLBACK_STAT && P.callback.stats) {
        unsigned long long v = sh->stats[threadIdx.x];
        if (v) atomicAdd(P.callback.stats + threadIdx.x, v);
    }
    if (threadIdx.x < 4 && P.work_counters) {
        unsigned long long v = sh->work[threadIdx.x];
        if (v) atomicAdd(P.work_counters + threadIdx.x, v);
    }
}

}  // namespace theia
