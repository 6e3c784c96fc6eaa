// ORACLE - TEST INFRASTRUCTURE ONLY. Never included, linked or called by the product path.
//
// C interface of the CPU oracle (bound from Python by oracle/__init__.py with ctypes).
// It mirrors include/theia_b200.h with host pointers in place of device pointers, so the
// parity tests hand the same launch description to both sides.
//
// Parity status: the reference itself cannot run here (GLSL + Vulkan through hephaistos, see
// DESIGN.md), so this restatement is pinned against every CPU statement the reference's
// tests hold (numpy restatements, CSV golden tables, analytic energy oracles) and against
// the published Philox4x32-10 known-answer vectors; the ray query arithmetic lives in the
// Vulkan driver: parity of hit triangle ids is UNPINNED (the reference never outputs them).
#include <omp.h>
#include <cstdio>
#include <cstdlib>
#include "tracer.h"
#include "units.h"

using namespace orc;

extern "C" {

int orc_abi_version(void) { return THEIA_ABI_VERSION; }

// sh/random.sink.glsl:17-39
int orc_philox_fill(const theia_philox_params_t* p, uint32_t baseStream, uint32_t baseCount, uint32_t streams,
                    uint32_t samples, uint32_t dim, float* out) {
#pragma omp parallel for schedule(static)
    for (int64_t s = 0; s < (int64_t)streams; ++s) {
        for (uint32_t c = 0; c < samples; ++c) {
            size_t i = (size_t)s * samples + c;
            if (dim == 1) {
                out[i] = random_s(*p, baseStream + (uint32_t)s, baseCount + c);
            } else {
                out[2 * i] = random_s(*p, baseStream + (uint32_t)s, baseCount + c);
                out[2 * i + 1] = random_s(*p, baseStream + (uint32_t)s, baseCount + c + 1);
            }
        }
    }
    return 0;
}

int orc_philox_raw(const uint32_t* ctr, const uint32_t* key, uint32_t n, uint32_t* out) {
    for (uint32_t i = 0; i < n; ++i) philox4x32_10(ctr + 4 * i, key + 2 * i, out + 4 * i);
    return 0;
}

int orc_math_eval(uint32_t fn, const float* x, const float* y, uint32_t n, float* out) {
    for (uint32_t i = 0; i < n; ++i) {
        switch (fn) {
        case 0: out[i] = dm_sin(x[i]); break;
        case 1: out[i] = dm_cos(x[i]); break;
        case 2: out[i] = dm_exp(x[i]); break;
        case 3: out[i] = dm_log(x[i]); break;
        case 4: out[i] = dm_pow(x[i], y[i]); break;
        case 5: out[i] = dm_tanh(x[i]); break;
        case 6: out[i] = dm_atan(x[i]); break;
        default: return -1;
        }
    }
    return 0;
}

// ---- component seams (pins of the target / target guide restatements, tests/test_components.py)
// out[i] = {position xyz, normal xyz, dist, objPosition xyz, objNormal xyz, prob, valid} (15 floats)
static void packTargetSample(const TargetSample& t, float* o) {
    o[0] = t.position.x; o[1] = t.position.y; o[2] = t.position.z;
    o[3] = t.normal.x; o[4] = t.normal.y; o[5] = t.normal.z; o[6] = t.dist;
    o[7] = t.objPosition.x; o[8] = t.objPosition.y; o[9] = t.objPosition.z;
    o[10] = t.objNormal.x; o[11] = t.objNormal.y; o[12] = t.objNormal.z; o[13] = t.prob; o[14] = t.valid ? 1.0f : 0.0f;
}
int orc_target_sample(const theia_component_t* c, const theia_philox_params_t* rng, const float* observer, uint32_t n,
                      float* out) {
    for (uint32_t i = 0; i < n; ++i) {
        Rng r{rng, i, 0u};
        packTargetSample(sampleTarget(*c, v3(observer + 3 * i), r), out + 15 * i);
    }
    return 0;
}
int orc_target_intersect(const theia_component_t* c, const float* observer, const float* dir, uint32_t n, float* out) {
    for (uint32_t i = 0; i < n; ++i)
        packTargetSample(intersectTarget(*c, v3(observer + 3 * i), v3(dir + 3 * i)), out + 15 * i);
    return 0;
}
// out[i] = {dir xyz, dist, prob}
int orc_guide_sample(const theia_component_t* c, const theia_philox_params_t* rng, const float* observer, uint32_t n,
                     float* out) {
    for (uint32_t i = 0; i < n; ++i) {
        Rng r{rng, i, 0u};
        TargetGuideSample g = sampleTargetGuide(*c, v3(observer + 3 * i), r);
        float* o = out + 5 * i;
        o[0] = g.dir.x; o[1] = g.dir.y; o[2] = g.dir.z; o[3] = g.dist; o[4] = g.prob;
    }
    return 0;
}
int orc_guide_eval(const theia_component_t* c, const float* observer, const float* dir, uint32_t n, float* out) {
    for (uint32_t i = 0; i < n; ++i) {
        TargetGuideSample g = evalTargetGuide(*c, v3(observer + 3 * i), v3(dir + 3 * i));
        float* o = out + 5 * i;
        o[0] = g.dir.x; o[1] = g.dir.y; o[2] = g.dir.z; o[3] = g.dist; o[4] = g.prob;
    }
    return 0;
}

// ---- unit seam (mirror of theia_unit_eval / theia_unit_widths, host pointers)
int orc_unit_widths(uint32_t unit, uint32_t polarized, uint32_t* in_width, uint32_t* out_width) {
    return unitWidths(unit, polarized, in_width, out_width);
}
int orc_unit_eval(const theia_unit_args_t* args) { return unitEval(*args); }

// ---- scene
void* orc_scene_create(const theia_instance_desc_t* desc, uint32_t n, const uint32_t* nVerts, const uint32_t* nTris,
                       int bruteForce) {
    return sceneCreate(desc, n, nVerts, nTris, bruteForce != 0);
}
void orc_scene_destroy(void* s) { delete (Scene*)s; }

int orc_tlas_query(const void* scene, const float* rays, uint32_t n, int32_t* out, float* out_bary, uint64_t* stats) {
    const Scene& sc = *(const Scene*)scene;
    uint64_t tot[4] = {0, 0, 0, 0};
#pragma omp parallel
    {
        uint64_t st[4] = {0, 0, 0, 0};
#pragma omp for schedule(dynamic, 1024)
        for (int64_t i = 0; i < (int64_t)n; ++i) {
            const float* r = rays + 8 * i;
            RayHit h = closestHit(sc, v3(r), v3(r + 4), r[3], stats ? st : nullptr);
            out[4 * i + 0] = h.valid ? h.instance : -1;
            out[4 * i + 1] = h.valid ? h.prim : -1;
            out[4 * i + 2] = h.valid ? sc.instances[h.instance].customIndex : -1;
            out[4 * i + 3] = 0;
            out_bary[4 * i + 0] = h.valid ? h.t : 0.0f;
            out_bary[4 * i + 1] = h.valid ? h.u : 0.0f;
            out_bary[4 * i + 2] = h.valid ? h.v : 0.0f;
            out_bary[4 * i + 3] = 0.0f;
        }
#pragma omp critical
        for (int k = 0; k < 4; ++k) tot[k] += st[k];
    }
    if (stats) for (int k = 0; k < 4; ++k) stats[k] = tot[k];
    return 0;
}

// ---- hit queue (sh/response.queue.glsl:6-33): uint32 count, then field-major arrays
static void saveHit(uint32_t* queue, uint32_t cap, bool qpol, uint32_t k, const HitItem& h) {
    float* f = (float*)(queue + 1);
    int c = 0;
    auto put = [&](float v) { f[(size_t)c * cap + k] = v; ++c; };
    put(h.position.x); put(h.position.y); put(h.position.z);
    put(h.direction.x); put(h.direction.y); put(h.direction.z);
    put(h.normal.x); put(h.normal.y); put(h.normal.z);
    if (qpol) {
        put(h.stokes.x); put(h.stokes.y); put(h.stokes.z); put(h.stokes.w);
        put(h.polRef.x); put(h.polRef.y); put(h.polRef.z);
    }
    put(h.wavelength); put(h.time); put(h.contrib);
    ((int32_t*)f)[(size_t)c * cap + k] = h.objectId;
}
static HitItem loadHit(const uint32_t* queue, uint32_t cap, bool qpol, uint32_t k) {
    const float* f = (const float*)(queue + 1);
    int c = 0;
    auto get = [&]() { float v = f[(size_t)c * cap + k]; ++c; return v; };
    HitItem h{};
    h.position.x = get(); h.position.y = get(); h.position.z = get();
    h.direction.x = get(); h.direction.y = get(); h.direction.z = get();
    h.normal.x = get(); h.normal.y = get(); h.normal.z = get();
    if (qpol) {
        h.stokes.x = get(); h.stokes.y = get(); h.stokes.z = get(); h.stokes.w = get();
        h.polRef.x = get(); h.polRef.y = get(); h.polRef.z = get();
    } else {
        h.stokes = {1.0f, 0.0f, 0.0f, 0.0f};
    }
    h.wavelength = get(); h.time = get(); h.contrib = get();
    h.objectId = ((const int32_t*)f)[(size_t)c * cap + k];
    return h;
}

// counters[0] = segments (trace() executions), [1] = responses, [2..5] traversal statistics
int orc_trace_forward(const theia_trace_launch_t* L, int nThreads, uint64_t* counters) {
    if (L->struct_size != sizeof(theia_trace_launch_t)) return -1;
    if ((L->tracer == THEIA_TRACER_VOLUME_PHOTON || L->tracer == THEIA_TRACER_SCENE_PHOTON) &&
        (L->run_length == 0 || L->path_length % L->run_length != 0)) return -4;
    const uint32_t bs = L->block_size ? L->block_size : 128;
    const uint32_t nChunks = (L->batch_size + bs - 1) / bs;
    const bool pol = L->polarized != 0;
    const theia_response_t& R = L->response;
    if (nThreads <= 0) nThreads = omp_get_max_threads();
    std::vector<std::vector<TaggedHit>> chunkHits(R.kind == THEIA_RESPONSE_RECORD ? nChunks : 0);
    std::vector<std::vector<Sink::TimeHit>> chunkTimes(R.kind == THEIA_RESPONSE_STORE_TIME ? nChunks : 0);
    uint64_t stats[12] = {0};
    uint64_t segs = 0, resp = 0, trav[4] = {0, 0, 0, 0};
    if (Sink::hasHistogram(R.kind) && R.hist_capacity == 0) return -2;

#pragma omp parallel num_threads(nThreads)
    {
        uint64_t tstats[12] = {0}, tsegs = 0, tresp = 0, ttrav[4] = {0, 0, 0, 0};
#pragma omp for schedule(dynamic, 4)
        for (int64_t c = 0; c < (int64_t)nChunks; ++c) {
            Sink sink(L);
            uint32_t lo = (uint32_t)c * bs, hi = lo + bs;
            if (hi > L->batch_size) hi = L->batch_size;
            if (L->tracer == THEIA_TRACER_VOLUME_FORWARD || L->tracer == THEIA_TRACER_VOLUME_PHOTON) {
                VolumeTracer t{*L, *(const theia_volume_trace_params_t*)L->trace_params, sink, pol};
                if (L->tracer == THEIA_TRACER_VOLUME_FORWARD) for (uint32_t k = lo; k < hi; ++k) t.traceMain(L->first_path + k);
                else for (uint32_t k = lo; k < hi; ++k) t.tracePhoton(L->first_path + k);
            } else {
                SceneTracer t{*L, *(const theia_scene_trace_params_t*)L->trace_params, *(const Scene*)L->tlas,
                              sink, pol, counters ? ttrav : nullptr};
                if (L->tracer == THEIA_TRACER_SCENE_FORWARD) for (uint32_t k = lo; k < hi; ++k) t.traceMain(L->first_path + k);
                else for (uint32_t k = lo; k < hi; ++k) t.tracePhoton(L->first_path + k);
            }
            if (Sink::hasHistogram(R.kind)) {
                float* dst = R.hist_partials + (size_t)((uint32_t)c % R.hist_capacity) * R.n_bins;
#pragma omp critical(orc_hist)
                for (uint32_t b = 0; b < R.n_bins; ++b) dst[b] += sink.hist[b];
            } else if (R.kind == THEIA_RESPONSE_RECORD) {
                chunkHits[c] = std::move(sink.hits);
            } else if (R.kind == THEIA_RESPONSE_STORE_TIME) {
                chunkTimes[c] = std::move(sink.times);
            }
            for (int k = 0; k < 12; ++k) tstats[k] += sink.stats[k];
            tresp += sink.nResponse;
            tsegs += sink.nSegments;
        }
#pragma omp critical(orc_stats)
        {
            for (int k = 0; k < 12; ++k) stats[k] += tstats[k];
            segs += tsegs; resp += tresp;
            for (int k = 0; k < 4; ++k) trav[k] += ttrav[k];
        }
    }
    if (R.kind == THEIA_RESPONSE_RECORD) {
        uint32_t count = R.queue[0];
        for (auto& v : chunkHits)
            for (auto& h : v) {
                if (count < R.queue_capacity) {
                    saveHit(R.queue, R.queue_capacity, R.queue_polarized != 0, count, h.item);
                    if (R.tag_path) R.tag_path[count] = h.path;
                    if (R.tag_event) R.tag_event[count] = h.event;
                }
                ++count;  // the reference's atomicAdd also runs past the capacity
            }
        R.queue[0] = count;
    }
    if (R.kind == THEIA_RESPONSE_STORE_TIME) {  // sh/response.time.store.glsl:4-10 (count, time[], objectId[])
        uint32_t count = R.value_queue[0];
        float* time = (float*)(R.value_queue + 1);
        int32_t* ids = (int32_t*)(R.value_queue + 1) + R.value_capacity;
        for (auto& v : chunkTimes)
            for (auto& h : v) {
                if (count < R.value_capacity) {
                    time[count] = h.time;
                    if (R.store_object_id) ids[count] = h.objectId;
                    if (R.tag_path) R.tag_path[count] = h.path;
                    if (R.tag_event) R.tag_event[count] = h.event;
                }
                ++count;
            }
        R.value_queue[0] = count;
    }
    if (L->callback.kind == THEIA_CALLBACK_STAT && L->callback.stats)
        for (int k = 0; k < 12; ++k) L->callback.stats[k] += stats[k];
    if (L->work_counters) {  // same meaning as on the device; shadow rays are not counted here
        L->work_counters[0] += segs; L->work_counters[1] += resp; L->work_counters[3] += L->batch_size;
    }
    if (counters) {
        counters[0] = segs; counters[1] = resp;
        for (int k = 0; k < 4; ++k) counters[2 + k] = trav[k];
    }
    return 0;
}

// sh/estimator.reduce.glsl:17-35 (precise Kahan sum)
int orc_hist_reduce(float* partials, uint32_t nHist, uint32_t nBins, float norm, float* out, int clear) {
    for (uint32_t b = 0; b < nBins; ++b) {
        float result = 0.0f, c = 0.0f;
        for (uint32_t n = 0; n < nHist; ++n) {
            float y = partials[(size_t)n * nBins + b] - c;
            float t = result + y;
            c = (t - result) - y;
            result = t;
        }
        out[b] = result * norm;
    }
    if (clear) std::memset(partials, 0, sizeof(float) * (size_t)nHist * nBins);
    return 0;
}

// sh/response.replay.glsl:35-50
// rng may be null ("optional RNG", :8-9): responses that draw (store time) need one. Hit k uses stream k from its
// first draw (idx = gl_GlobalInvocationID.x, dim = 0, :37-38).
int orc_hit_replay_rng(const uint32_t* queue, uint32_t capacity, uint32_t polarized, const theia_response_t* response,
                       const theia_philox_params_t* rng, uint32_t block_size) {
    if (response->kind == THEIA_RESPONSE_STORE_TIME && !rng) return -3;
    theia_trace_launch_t L{};
    L.response = *response;
    uint32_t count = queue[0] < capacity ? queue[0] : capacity;
    uint32_t bs = block_size ? block_size : 128;
    uint32_t nChunks = (capacity + bs - 1) / bs;
    std::vector<TaggedHit> all;
    std::vector<Sink::TimeHit> times;
    for (uint32_t c = 0; c < nChunks; ++c) {
        Sink sink(&L);
        for (uint32_t k = c * bs; k < (c + 1) * bs && k < count; ++k) {
            Rng stream{rng, k, 0u};
            sink.response(loadHit(queue, capacity, polarized != 0, k), stream);
        }
        for (auto& t : sink.times) times.push_back(t);
        if (Sink::hasHistogram(response->kind)) {
            float* dst = response->hist_partials + (size_t)(c % response->hist_capacity) * response->n_bins;
            for (uint32_t b = 0; b < response->n_bins; ++b) dst[b] += sink.hist[b];
        } else if (response->kind == THEIA_RESPONSE_RECORD) {
            for (auto& h : sink.hits) all.push_back(h);
        }
    }
    if (response->kind == THEIA_RESPONSE_RECORD) {
        uint32_t n = response->queue[0];
        for (auto& h : all) {
            if (n < response->queue_capacity) saveHit(response->queue, response->queue_capacity, response->queue_polarized != 0, n, h.item);
            ++n;
        }
        response->queue[0] = n;
    }
    if (response->kind == THEIA_RESPONSE_STORE_TIME) {  // sh/response.time.store.glsl:4-10 (count, time[], objectId[])
        uint32_t n = response->value_queue[0];
        float* time = (float*)(response->value_queue + 1);
        int32_t* ids = (int32_t*)(response->value_queue + 1) + response->value_capacity;
        for (auto& h : times) {
            if (n < response->value_capacity) {
                time[n] = h.time;
                if (response->store_object_id) ids[n] = h.objectId;
            }
            ++n;
        }
        response->value_queue[0] = n;
    }
    return 0;
}
int orc_hit_replay(const uint32_t* queue, uint32_t capacity, uint32_t polarized, const theia_response_t* response,
                   uint32_t block_size) {
    return orc_hit_replay_rng(queue, capacity, polarized, response, nullptr, block_size);
}

}  // extern "C"
