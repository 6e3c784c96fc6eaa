// ORACLE - TEST INFRASTRUCTURE ONLY. Never included, linked or called by the product path.
//
// CPU restatement of the two forward tracer mains and of the response / callback snippets:
// sh/tracer.volume.forward.glsl, sh/tracer.scene.forward.glsl, sh/scene.traverse.glsl,
// sh/response.histogram.glsl, sh/response.record.glsl, sh/callback.stat.glsl,
// sh/callback.track.glsl.
#pragma once
#include <vector>
#include "components.h"
#include "surface.h"

namespace orc {

struct TaggedHit { HitItem item; uint32_t path, event; };

// Per work-group (chunk of block_size paths) sink for hits and events.
struct Sink {
    const theia_trace_launch_t* L;
    std::vector<float> hist;        // local histogram of this chunk
    std::vector<TaggedHit> hits;    // recorded hits of this chunk in path order
    uint64_t stats[12] = {0};
    uint64_t nResponse = 0;
    uint64_t nSegments = 0;  // trace() executions
    uint32_t curEvent = 0;
    uint32_t seq = 0;  // running number of the path's hits inside the current event (test tags)
    struct TimeHit { float time; int32_t objectId; uint32_t path, event; };
    std::vector<TimeHit> times;     // store time response: accepted hits of this chunk in path order

    static bool hasHistogram(uint32_t kind) { return kind == THEIA_RESPONSE_HISTOGRAM || kind == THEIA_RESPONSE_KERNEL_HISTOGRAM; }
    explicit Sink(const theia_trace_launch_t* l) : L(l) {
        if (hasHistogram(l->response.kind)) hist.assign(l->response.n_bins, 0.0f);
    }
    static int32_t floatToInt(float f) { return (f >= 2147483648.0f || f < -2147483648.0f || f != f) ? INT32_MIN : (int32_t)f; }
    // sh/response.histogram.kernel.glsl:62-70: normal CDF after Vazquez-Leal et al. (2011), eq. 4.3
    static float kernelCdf(float x) {
        const float c1 = 7.779374467827938f, c2 = 55.5f, c3 = 0.1257926109373887f;
        return 0.5f + 0.5f * dm_tanh(c1 * x - c2 * dm_atan(c3 * x));
    }
    // sh/response.histogram.glsl:59-68, sh/response.record.glsl:23-26, sh/response.histogram.kernel.glsl:72-101,
    // sh/response.time.store.glsl:14-26
    void response(const HitItem& item, Rng& rng) {
        const uint32_t idx = rng.idx;
        nResponse++;
        const theia_response_t& r = L->response;
        if (r.kind == THEIA_RESPONSE_KERNEL_HISTOGRAM) {
            const float value = item.contrib;  // sh/response.uniform.glsl:4-6
            float t = item.time - r.hist.t0;
            int32_t firstBin = floatToInt(std::floor((t - r.kernel_support) / r.hist.binSize));
            int32_t lastBin = floatToInt(std::ceil((t + r.kernel_support) / r.hist.binSize));
            firstBin = std::max(firstBin, 0);
            lastBin = std::min(lastBin, (int32_t)r.n_bins) - 1;
            t = (float)firstBin * r.hist.binSize + r.hist.t0 - item.time;
            const float h = 1.0f / r.kernel_bandwidth;
            float prev_cdf = kernelCdf(t * h);
            for (int32_t i = firstBin; i <= lastBin; ++i) {
                t += r.hist.binSize;
                float cdf = kernelCdf(t * h);
                float w = cdf - prev_cdf;
                prev_cdf = cdf;
                hist[i] += w * value;
            }
        } else if (r.kind == THEIA_RESPONSE_STORE_TIME) {
            float value = item.contrib;
            if (rng.next() < value) times.push_back(TimeHit{item.time, item.objectId, idx, curEvent | (seq << 16)});
            seq++;
        } else if (r.kind == THEIA_RESPONSE_OBJECT_HISTOGRAM) {
            uint32_t bin = (uint32_t)floatToInt(std::floor((item.time - r.hist.t0) / r.hist.binSize));
            if (bin < r.n_bins && (uint32_t)item.objectId < r.n_objects) {
                float* dst = r.object_hist + (size_t)item.objectId * r.n_bins + bin;
#pragma omp atomic
                *dst += item.contrib;
            }
        } else if (r.kind == THEIA_RESPONSE_HISTOGRAM) {
            float value = item.contrib;  // sh/response.uniform.glsl:4-6
            float fb = std::floor((item.time - r.hist.t0) / r.hist.binSize);
            // uint bin = int(floor(..)); negative values wrap to huge unsigned numbers
            int32_t ib = (fb >= 2147483648.0f || fb < -2147483648.0f || fb != fb) ? INT32_MIN : (int32_t)fb;
            uint32_t bin = (uint32_t)ib;
            if (bin < r.n_bins) hist[bin] += value;
        } else if (r.kind == THEIA_RESPONSE_RECORD) {
            hits.push_back(TaggedHit{item, idx, curEvent | (seq << 16)});
            seq++;
        }
    }
    // sh/callback.stat.glsl:21-62, sh/callback.track.glsl:28-98
    void onEvent(const Ray& ray, int code, uint32_t idx, uint32_t i, bool polarized) {
        curEvent = i;
        seq = 0;
        if (i >= 1 && code != THEIA_RESULT_MAX_ITER) nSegments++;
        const theia_callback_t& cb = L->callback;
        if (cb.kind == THEIA_CALLBACK_STAT) {
            switch (code) {
            case THEIA_RESULT_RAY_CREATED: stats[0]++; break;
            case THEIA_RESULT_RAY_SCATTERED: stats[1]++; break;
            case THEIA_RESULT_RAY_HIT: stats[2]++; break;
            case THEIA_RESULT_RAY_DETECTED: stats[3]++; break;
            case THEIA_RESULT_VOLUME_HIT: stats[4]++; break;
            case THEIA_RESULT_RAY_LOST: stats[5]++; break;
            case THEIA_RESULT_RAY_DECAYED: stats[6]++; break;
            case THEIA_RESULT_RAY_ABSORBED: stats[7]++; break;
            case THEIA_RESULT_RAY_MISSED: stats[8]++; break;
            case THEIA_RESULT_MAX_ITER: stats[9]++; break;
            case THEIA_ERROR_MEDIA_MISMATCH: stats[11]++; break;
            }
            if (code <= THEIA_ERROR_MAX_VALUE) stats[10]++;
        } else if (cb.kind == THEIA_CALLBACK_TRACK) {
            if (code == THEIA_RESULT_MAX_ITER) return;
            uint32_t T = cb.track_count, Ln = cb.track_length;
            uint32_t k = idx - L->first_path;
            if (k >= T) return;
            uint32_t* n = (uint32_t*)cb.track_buffer;
            int32_t* codes = (int32_t*)(n + T);
            float* base = (float*)(codes + T);
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
                vec4 st = polarized ? ray.stokes : vec4{1.0f, 0.0f, 0.0f, 0.0f};
                vec3 pr = polarized ? ray.polRef : v3(0.0f);
                base[4 * plane + o] = st.x; base[5 * plane + o] = st.y;
                base[6 * plane + o] = st.z; base[7 * plane + o] = st.w;
                base[8 * plane + o] = pr.x; base[9 * plane + o] = pr.y; base[10 * plane + o] = pr.z;
            }
        }
    }
};

// sh/tracer.*.forward.glsl sampleRay
static inline Ray sampleRay(const theia_trace_launch_t& L, uint64_t medium, Rng& rng) {
    bool pol = L.polarized != 0;
    WavelengthSample photon = sampleWavelength(L.wavelength, rng, L.first_path);
    MediumConstants constants = lookUpMedium(medium, photon.wavelength);
    SourceRay lightRay = sampleLight(L.light, photon.wavelength, constants, rng, pol, L.first_path);
    return createRay(lightRay, medium, constants, photon);
}

// sh/tracer.scene.photon.init.glsl:90-106, sh/tracer.volume.photon.common.glsl:88-104
static inline int photonRoulette(Ray& ray, Rng& rng, int result, bool pol) {
    if (result < 0) return result;
    if (pol) {
        ray.lin_contrib *= ray.stokes.x;
        ray.stokes = ray.stokes / ray.stokes.x;
    }
    if (getContrib(ray) <= rng.next()) return THEIA_RESULT_RAY_ABSORBED;
    ray.lin_contrib = 1.0f;
    ray.log_contrib = 0.0f;
    return result;
}

// =================================================================== volume tracer
struct VolumeTracer {
    const theia_trace_launch_t& L;
    const theia_volume_trace_params_t& P;
    Sink& sink;
    bool pol;

    // sh/tracer.volume.forward.glsl:51-82
    void createResponse(Ray ray, const TargetSample& hit, vec3 dir, float weight, bool scattered, Rng& rng) {
        if (!hit.valid) return;
        if (scattered) scatterRayIS(ray, dir, pol);
        int code = propagateRayToHit(ray, hit.position, hit.normal, P.propagation, pol);
        if (code < 0) return;
        ray.lin_contrib *= weight;
        HitItem item = createHit(ray, hit.objPosition, hit.objNormal, P.objectId, hit.worldToObj, pol);
        if (item.contrib > 0.0f) sink.response(item, rng);
    }
    // :103-114
    static float jacobian_dAdW(vec3 obs, vec3 pos, vec3 nrm) {
        vec3 dir = pos - obs;
        float d2 = dot(dir, dir);
        dir = normalize(dir);
        float factor = d2 / abs_(dot(dir, nrm));
        if (isinf_(factor)) factor = 0.0f;
        return factor;
    }
    // :116-149
    void sampleTargetMIS(const Ray& ray, Rng& rng) {
        uint64_t med = P.medium;
        vec3 obs = ray.position, dir = ray.direction;
        float pPP;
        vec3 dirPhase = scatter(med, dir, rng.next2(), &pPP);
        TargetSample phaseHit = intersectTarget(L.target, obs, dirPhase);
        TargetSample targetHit = sampleTarget(L.target, obs, rng);
        vec3 dirTarget = normalize(targetHit.position - obs);
        float pTT = targetHit.prob * jacobian_dAdW(obs, targetHit.position, targetHit.normal);
        float pPT = scatterProb(med, dir, dirTarget);
        float pTP = phaseHit.prob * jacobian_dAdW(obs, phaseHit.position, phaseHit.normal);
        float wTarget = pTT * pPT / (pTT * pTT + pPT * pPT);
        float wPhase = pPP * pPP / (pPP * pPP + pTP * pTP);
        createResponse(ray, phaseHit, dirPhase, wPhase, true, rng);
        createResponse(ray, targetHit, dirTarget, wTarget, true, rng);
    }
    // :153-208
    int trace(Ray& ray, Rng& rng, bool allowResponse) {
        float u = rng.next();
        float dist = sampleScatterLength(ray, P.propagation, u);
        TargetSample hit = intersectTarget(L.target, ray.position, ray.direction);
        bool hitValid = hit.valid && (hit.dist <= dist);
        dist = min_(hit.dist, dist);
        int result = propagateRay(ray, dist, P.propagation);
        updateRayIS(ray, dist, P.propagation, hitValid);
        if (result < 0) return result;
        if (hitValid && allowResponse) {
            alignRayToHit(ray, hit.normal, pol);
            HitItem item = createHit(ray, hit.objPosition, hit.objNormal, P.objectId, hit.worldToObj, pol);
            if (item.contrib > 0.0f) sink.response(item, rng);
            return THEIA_RESULT_RAY_DETECTED;
        } else if (hitValid) {
            return THEIA_RESULT_RAY_ABSORBED;
        }
        if (!L.disable_mis) sampleTargetMIS(ray, rng);
        return THEIA_RESULT_RAY_SCATTERED;
    }
    // :231-276
    void traceMain(uint32_t idx) {
        Rng rng{&L.rng, idx, 0};
        const bool DIRECT_LIGHTING = L.disable_mis && !L.disable_direct_lighting;
        const bool ALLOW_RESPONSE = L.disable_mis != 0;
        Ray ray = sampleRay(L, P.medium, rng);
        sink.onEvent(ray, THEIA_RESULT_RAY_CREATED, idx, 0, pol);
        if (isOccludedByTarget(L.target, ray.position)) {
            sink.onEvent(ray, THEIA_ERROR_TRACE_ABORT, idx, 0, pol);
            return;
        }
        if (!L.disable_direct_lighting && !L.disable_mis) {
            TargetSample directHit = intersectTarget(L.target, ray.position, ray.direction);
            createResponse(ray, directHit, ray.direction, 1.0f, false, rng);
        }
        int result = trace(ray, rng, DIRECT_LIGHTING);
        sink.onEvent(ray, result, idx, 1, pol);
        if (result < 0) return;
        for (uint32_t i = 0; i < L.path_length; ++i) {
            vec2 u = rng.next2();
            scatterRay(ray, u, pol);
            int res = trace(ray, rng, ALLOW_RESPONSE);
            sink.onEvent(ray, res, idx, i + 2, pol);
            if (res < 0) return;
        }
        sink.onEvent(ray, THEIA_RESULT_MAX_ITER, idx, L.path_length + 2, pol);
    }
    // VolumePhotonTracer: sh/tracer.volume.photon.common.glsl:32-76
    int tracePhotonStep(Ray& ray, Rng& rng) {
        float u = rng.next();
        float dist = sampleScatterLength(ray, P.propagation, u);
        TargetSample hit = intersectTarget(L.target, ray.position, ray.direction);
        bool hitValid = hit.valid && (hit.dist <= dist);
        dist = min_(hit.dist, dist);
        int result = propagateRay(ray, dist, P.propagation);
        updateRayIS(ray, dist, P.propagation, hitValid);
        if (result < 0) return result;
        if (hitValid) {
            alignRayToHit(ray, hit.normal, pol);
            HitItem item = createHit(ray, hit.objPosition, hit.objNormal, P.objectId, hit.worldToObj, pol);
            if (item.contrib > 0.0f) sink.response(item, rng);
            return THEIA_RESULT_RAY_ABSORBED;
        }
        return THEIA_RESULT_RAY_SCATTERED;
    }
    // sh/tracer.volume.photon.init.glsl:30-62, traceLoop of sh/tracer.volume.photon.common.glsl:78-123 over
    // all runs of sh/tracer.volume.photon.loop.glsl:20-31 (run r starts its draw counter at r * rng_per_run)
    void tracePhoton(uint32_t idx) {
        Rng rng{&L.rng, idx, 0};
        Ray ray = sampleRay(L, P.medium, rng);
        ray.lin_contrib = 1.0f;
        ray.log_contrib = 0.0f;
        sink.onEvent(ray, THEIA_RESULT_RAY_CREATED, idx, 0, pol);
        if (isOccludedByTarget(L.target, ray.position)) {
            sink.onEvent(ray, THEIA_ERROR_TRACE_ABORT, idx, 0, pol);
            return;
        }
        for (uint32_t i = 1; i <= L.path_length; ++i) {
            int result = tracePhotonStep(ray, rng);
            scatterRay(ray, rng.next2(), pol);
            result = photonRoulette(ray, rng, result, pol);
            sink.onEvent(ray, result, idx, i, pol);
            if (result < 0) return;
            if (i < L.path_length && i % L.run_length == 0) rng.dim = (i / L.run_length) * L.rng_per_run;
        }
        sink.onEvent(ray, THEIA_RESULT_MAX_ITER, idx, L.path_length + 1, pol);
    }
};

// =================================================================== scene tracer
struct SceneTracer {
    const theia_trace_launch_t& L;
    const theia_scene_trace_params_t& P;
    const Scene& scene;
    Sink& sink;
    bool pol;
    uint64_t* travStats;

    // sh/scene.traverse.glsl:33-68
    void createResponse(Ray ray, const SurfaceHit& hit, const Reflectance& surface, bool absorb, Rng& rng) {
        if (!absorb) {
            if (L.hit_use_transmitted_dir) {
                transmitRay(ray, hit, surface, pol);
            } else {
                vec3 dir = ray.direction;
                transmitRay(ray, hit, surface, pol);
                ray.direction = dir;
            }
        }
        HitItem item = createHit(ray, hit.objPos, hit.objNrm, hit.customId, hit.worldToObj, pol);
        if (item.contrib > 0.0f) sink.response(item, rng);
    }
    // :73-154
    int processHit(Ray& ray, const SurfaceHit& hit, Rng& rng, bool allowResponse) {
        int result = propagateRayToHit(ray, hit.worldPos, hit.rayNrm, P.propagation, pol);
        if (result < 0) return result;
        result = THEIA_RESULT_RAY_HIT;
        bool isAbs = (hit.flags & MATERIAL_BLACK_BODY_BIT) != 0;
        bool isTarget = (hit.flags & MATERIAL_DETECTOR_BIT) != 0;
        bool volBorder = (hit.flags & MATERIAL_VOLUME_BORDER_BIT) != 0;
        bool canTransmit = (hit.flags & MATERIAL_NO_TRANSMIT_BWD_BIT) == 0;  // sic: BWD bits (:96-97)
        bool canReflect = (hit.flags & MATERIAL_NO_REFLECT_BWD_BIT) == 0;
        Reflectance surface = fresnelReflect(ray, hit);
        if (allowResponse && isTarget && (P.targetId < 0 || hit.customId == P.targetId)) {
            createResponse(ray, hit, surface, isAbs, rng);
            result = THEIA_RESULT_RAY_DETECTED;
        }
        if (isAbs) return THEIA_RESULT_RAY_ABSORBED;
        if (!L.disable_volume_border && volBorder) {
            crossBorder(ray, hit);
            return THEIA_RESULT_VOLUME_HIT;
        }
        if (!L.disable_transmission) {
            float r = 0.5f * (surface.r_s * surface.r_s + surface.r_p * surface.r_p);
            if (canReflect && canTransmit) {
                if (rng.next() < r) reflectRayIS(ray, hit, surface, pol);
                else transmitRayIS(ray, hit, surface, pol);
            } else if (canReflect) {
                reflectRay(ray, hit, surface, pol);
            } else if (canTransmit) {
                transmitRay(ray, hit, surface, pol);
            } else {
                result = THEIA_RESULT_RAY_ABSORBED;
            }
        } else {
            if (canReflect) reflectRay(ray, hit, surface, pol);
            else result = THEIA_RESULT_RAY_ABSORBED;
        }
        return result;
    }
    // :160-180
    void processShadowRay(Ray ray, const SurfaceHit& hit, Rng& rng) {
        bool isTarget = (hit.flags & MATERIAL_DETECTOR_BIT) != 0;
        bool correctId = P.targetId < 0 || P.targetId == hit.customId;
        if (!hit.valid || !correctId || !isTarget) return;
        if (propagateRayToHit(ray, hit.worldPos, hit.rayNrm, P.propagation, pol) < 0) return;
        Reflectance surface = fresnelReflect(ray, hit);
        bool black = (hit.flags & MATERIAL_BLACK_BODY_BIT) != 0;
        createResponse(ray, hit, surface, black, rng);
    }
    // :185-214
    void traceShadowRay(Ray ray, vec3 dir, float dist, Rng& rng, float weight) {
        scatterRayIS(ray, dir, pol);
        RayHit q = closestHit(scene, ray.position, dir, dist, travStats);
        SurfaceHit hit;
        if (processRayQuery(scene, ray, q, hit) < 0) return;
        ray.lin_contrib *= weight;
        processShadowRay(ray, hit, rng);
    }
    // :234-269
    void sampleTargetMIS(const Ray& ray, Rng& rng) {
        uint64_t med = ray.medium;
        vec3 obs = ray.position, dir = ray.direction;
        float pPP;
        vec3 dirPhase = scatter(med, dir, rng.next2(), &pPP);
        TargetGuideSample targetSample = sampleTargetGuide(L.guide, obs, rng);
        float pTT = targetSample.prob;
        TargetGuideSample phaseSample = evalTargetGuide(L.guide, obs, dirPhase);
        float pTP = phaseSample.prob;
        float pPT = scatterProb(med, dir, targetSample.dir);
        float wTarget = pTT * pPT / (pTT * pTT + pPT * pPT);
        float wPhase = pPP * pPP / (pPP * pPP + pTP * pTP);
        traceShadowRay(ray, dirPhase, phaseSample.dist, rng, wPhase);
        traceShadowRay(ray, targetSample.dir, targetSample.dist, rng, wTarget);
    }
    // :280-352
    int trace(Ray& ray, SurfaceHit& hit, Rng& rng, bool allowResponse) {
        hit = SurfaceHit{};
        if (isRayBad(ray)) return THEIA_ERROR_RAY_BAD;
        vec3 dir = normalize(ray.direction);
        float dist = sampleScatterLength(ray, P.propagation, rng.next());
        const bool mis = L.guide.kind != THEIA_GUIDE_NONE;
        float sampledDist = dist;
        bool mis_target = false;
        if (mis) {
            TargetGuideSample guideSample = evalTargetGuide(L.guide, ray.position, dir);
            mis_target = allowResponse && guideSample.prob > 0.0f;
            if (mis_target) {
                mis_target = guideSample.dist > dist;
                dist = max_(guideSample.dist, dist);
            }
        }
        RayHit q = closestHit(scene, ray.position, dir, dist, travStats);
        int queryResult = processRayQuery(scene, ray, q, hit);
        if (hit.valid) dist = distance(ray.position, hit.worldPos);
        if (mis && mis_target && hit.valid && dist > sampledDist && queryResult >= 0) {
            processShadowRay(ray, hit, rng);
            hit.valid = false;
            dist = sampledDist;
        }
        int propResult = propagateRay(ray, dist, P.propagation);
        updateRayIS(ray, dist, P.propagation, hit.valid);
        return queryResult <= THEIA_ERROR_MAX_VALUE ? queryResult : propResult;
    }
    // :361-397
    int processInteraction(Ray& ray, const SurfaceHit& hit, Rng& rng, bool allowResponse, bool last) {
        if (hit.valid) return processHit(ray, hit, rng, allowResponse);
        if (!last) {
            if (L.guide.kind != THEIA_GUIDE_NONE) sampleTargetMIS(ray, rng);
            scatterRay(ray, rng.next2(), pol);
        }
        return THEIA_RESULT_RAY_SCATTERED;
    }
    // sh/tracer.scene.forward.glsl:58-106
    void traceMain(uint32_t idx) {
        Rng rng{&L.rng, idx, 0};
        Ray ray = sampleRay(L, P.sourceMedium, rng);
        sink.onEvent(ray, THEIA_RESULT_RAY_CREATED, idx, 0, pol);
        bool allowResponse = !L.disable_direct_lighting;
        for (uint32_t i = 1; i <= L.path_length; ++i) {
            bool last = i == L.path_length;
            SurfaceHit hit;
            int result = trace(ray, hit, rng, allowResponse);
            if (result >= 0) result = processInteraction(ray, hit, rng, allowResponse, last);
            sink.onEvent(ray, result, idx, i, pol);
            if (result < 0) return;
            if (L.guide.kind != THEIA_GUIDE_NONE) allowResponse = result != THEIA_RESULT_RAY_SCATTERED;
            else allowResponse = true;
        }
        sink.onEvent(ray, THEIA_RESULT_MAX_ITER, idx, L.path_length + 1, pol);
    }
    // ScenePhotonTracer: sh/tracer.scene.photon.init.glsl:43-114 followed by the runs of
    // sh/tracer.scene.photon.loop.glsl:38-99 (run r starts its draw counter at r * rng_per_run;
    // what the queue between two runs carries is exactly the ray as it stands)
    void tracePhoton(uint32_t idx) {
        Rng rng{&L.rng, idx, 0};
        Ray ray = sampleRay(L, P.sourceMedium, rng);
        ray.lin_contrib = 1.0f;
        ray.log_contrib = 0.0f;
        sink.onEvent(ray, THEIA_RESULT_RAY_CREATED, idx, 0, pol);
        for (uint32_t i = 1; i <= L.path_length; ++i) {
            SurfaceHit hit;
            int result = trace(ray, hit, rng, true);
            if (result >= 0) result = processInteraction(ray, hit, rng, true, false);
            result = photonRoulette(ray, rng, result, pol);
            sink.onEvent(ray, result, idx, i, pol);
            if (result < 0) return;
            if (i < L.path_length && i % L.run_length == 0) rng.dim = (i / L.run_length) * L.rng_per_run;
        }
        sink.onEvent(ray, THEIA_RESULT_MAX_ITER, idx, L.path_length + 1, pol);
    }
};

}  // namespace orc
