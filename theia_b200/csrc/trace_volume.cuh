// VolumeForwardTracer kernel (sh/tracer.volume.forward.glsl of the reference).
//
// Execution model: persistent warps. Each warp owns a contiguous range of the batch's path
// indices; a lane whose path has ended is refilled with the warp's next unstarted path
// (ballot + prefix popcount), so live paths stay packed in the 32 lanes without a round
// trip of the path state through HBM. One loop iteration advances every live lane by one
// path segment: [create | scatter] -> trace -> [MIS phase sample] -> [MIS target sample],
// with the (optional) hit of each stage committed at a warp-convergent point so the
// histogram / queue / counter updates can be warp aggregated. A path's random numbers are
// a pure function of (key, offset, path index, draw counter), so its result does not depend
// on which lane or iteration executes it.
#pragma once
#include "launch.cuh"

namespace theia {

// sh/tracer.volume.forward.glsl:103-114
TDEV float jacobian_dAdW(vec3 obs, vec3 pos, vec3 nrm) {
    vec3 dir = pos - obs;
    float d2 = dot(dir, dir);
    dir = normalize(dir);
    float factor = d2 / abs_(dot(dir, nrm));
    if (isinf_(factor)) factor = 0.0f;
    return factor;
}

// sh/tracer.volume.forward.glsl:51-82. Works on a copy of the ray; returns whether a hit item
// was produced (the contrib > 0 test happens in commitHit).
template <bool POL, bool FULL>
TDEV bool volumeCreateResponse(const DeviceLaunch& P, Ray ray, const TargetSample& hit, vec3 dir, float weight,
                               bool scattered, HitItem& item) {
    if (!hit.valid) return false;
    if (scattered) scatterRayIS<POL>(ray, dir);
    int code = propagateRayToHit<POL>(ray, hit.position, hit.normal, P.propagation);
    if (code < 0) return false;
    ray.lin_contrib *= weight;
    item = createHit<POL>(ray, hit.objPosition, hit.objNormal, P.objectId, targetWorldToObj<FULL>(P.target));
    return true;
}

#ifndef THEIA_VOLUME_CTAS
#define THEIA_VOLUME_CTAS 6
#endif
// HIST / NOCB: the response is the plain histogram / there is no callback, known at compile time (lean launches of the
// BASELINE configurations get those instantiations, like the shading kernels of the scene tracer: VK_* in wavefront.cuh)
template <bool POL, bool FULL, bool HIST = false, bool NOCB = false, int LIGHT = -1>
__global__ void __launch_bounds__(128, THEIA_VOLUME_CTAS) volumeForwardKernel(const __grid_constant__ DeviceLaunch P) {
    extern __shared__ float sharedHist[];
    __shared__ SinkShared sh;
    sinkInit(P, &sh, sharedHist);
    float* hist = sinkHist(P, sharedHist);

    const bool MIS = P.disable_mis == 0;
    const bool DIRECT = P.disable_direct_lighting == 0;
    const bool DIRECT_LIGHTING = !MIS && DIRECT;  // :34-38
    const bool ALLOW_RESPONSE = !MIS;             // :44-48
    const uint32_t lastSeg = P.path_length;       // trace() calls are numbered 0..PATH_LENGTH

    // contiguous path range of this warp
    const uint32_t warpsPerCta = blockDim.x >> 5;
    const uint32_t nWarps = gridDim.x * warpsPerCta;
    const uint32_t warp = blockIdx.x * warpsPerCta + (threadIdx.x >> 5);
    const uint32_t perWarp = (P.batch_size + nWarps - 1) / nWarps;
    uint32_t next = min(warp * perWarp, P.batch_size);
    const uint32_t end = min(next + perWarp, P.batch_size);

    bool alive = false;
    uint32_t k = 0;    // path number inside the batch
    uint32_t seg = 0;  // index of the next trace() call of this path
    Ray ray;
    Rng rng;
    uint32_t nSegments = 0, nResponses = 0, nPaths = 0;

    while (true) {
        // ---- refill dead lanes with unstarted paths
        uint32_t dead = __ballot_sync(FULL_MASK, !alive);
        if (dead) {
            uint32_t avail = end - next;
            uint32_t rank = __popc(dead & laneMaskLt());
            if (!alive && rank < avail) {
                alive = true;
                k = next + rank;
                seg = 0;
            }
            next += min((uint32_t)__popc(dead), avail);
        }
        if (!__any_sync(FULL_MASK, alive)) break;

        HitItem item;
        bool has = false;
        const uint32_t idx = P.first_path + k;
        uint32_t hitSeq = 0;  // order of this path's hits inside the current event (test tags)

        // ---- stage 1: create the path (:234-250) or scatter (:262)
        bool created = alive && seg == 0;
        bool aborted = false;
        if (created) {
            rng = makeRng(&P.rng, idx);
            ray = sampleRay<POL, FULL, LIGHT>(P.wavelength, P.light, P.medium, rng);
            nPaths++;
            aborted = isOccludedByTarget<FULL>(P.target, ray.position);
            if (!aborted && MIS && DIRECT) {
                TargetSample directHit = intersectTarget<FULL>(P.target, ray.position, ray.direction);
                has = volumeCreateResponse<POL, FULL>(P, ray, directHit, ray.direction, 1.0f, false, item);
            }
        } else if (alive) {
            scatterRay<POL>(ray, rng.next2());
        }
        onEvent<POL, NOCB>(P.callback, &sh, created, ray, THEIA_RESULT_RAY_CREATED, k, 0);
        onEvent<POL, NOCB>(P.callback, &sh, aborted, ray, THEIA_ERROR_TRACE_ABORT, k, 0);
        if (aborted) alive = false;
        has = commitHit<FULL, HIST>(P.response, hist, has, item, idx, seg | (hitSeq << 16), nResponses, responseDraw<FULL>(P.response, rng, has, item));
        hitSeq += has ? 1u : 0u;

        // ---- stage 2: trace (:153-208)
        has = false;
        int result = 0;
        bool wantMIS = false;
        if (alive) {
            nSegments++;
            const bool allowResponse = seg == 0 ? DIRECT_LIGHTING : ALLOW_RESPONSE;
            float u = rng.next();
            float dist = sampleScatterLength(ray, P.propagation, u);
            TargetSample hit = intersectTarget<FULL>(P.target, ray.position, ray.direction);
            bool hitValid = hit.valid && (hit.dist <= dist);
            dist = min_(hit.dist, dist);
            result = propagateRay(ray, dist, P.propagation);
            updateRayIS(ray, dist, P.propagation, hitValid);
            if (result >= 0) {
                if (hitValid && allowResponse) {
                    alignRayToHit<POL>(ray, hit.normal);
                    item = createHit<POL>(ray, hit.objPosition, hit.objNormal, P.objectId, targetWorldToObj<FULL>(P.target));
                    has = true;
                    result = THEIA_RESULT_RAY_DETECTED;
                } else if (hitValid) {
                    result = THEIA_RESULT_RAY_ABSORBED;
                } else {
                    wantMIS = MIS;
                    result = THEIA_RESULT_RAY_SCATTERED;
                }
            }
        }
        // ---- stage 3: MIS, phase function sample (:116-149)
        vec3 dirTarget;
        TargetSample targetHit;
        float wTarget = 0.0f;
        if (wantMIS) {
            float pPP;
            vec3 dirPhase = scatter(P.medium, ray.direction, rng.next2(), &pPP);
            TargetSample phaseHit = intersectTarget<FULL>(P.target, ray.position, dirPhase);
            targetHit = sampleTarget<FULL>(P.target, ray.position, rng);
            dirTarget = normalize(targetHit.position - ray.position);
            float pTT = targetHit.prob * jacobian_dAdW(ray.position, targetHit.position, targetHit.normal);
            float pPT = scatterProb(P.medium, ray.direction, dirTarget);
            float pTP = phaseHit.prob * jacobian_dAdW(ray.position, phaseHit.position, phaseHit.normal);
            wTarget = pTT * pPT / (pTT * pTT + pPT * pPT);
            float wPhase = pPP * pPP / (pPP * pPP + pTP * pTP);
            has = volumeCreateResponse<POL, FULL>(P, ray, phaseHit, dirPhase, wPhase, true, item);
        }
        has = commitHit<FULL, HIST>(P.response, hist, has, item, idx, seg | (hitSeq << 16), nResponses, responseDraw<FULL>(P.response, rng, has, item));
        hitSeq += has ? 1u : 0u;
        // ---- stage 4: MIS, target sample
        has = false;
        if (wantMIS) has = volumeCreateResponse<POL, FULL>(P, ray, targetHit, dirTarget, wTarget, true, item);
        has = commitHit<FULL, HIST>(P.response, hist, has, item, idx, seg | (hitSeq << 16), nResponses, responseDraw<FULL>(P.response, rng, has, item));

        // ---- events and path bookkeeping (:254-275)
        onEvent<POL, NOCB>(P.callback, &sh, alive, ray, result, k, seg + 1);
        bool maxIter = false;
        if (alive) {
            if (result < 0) {
                alive = false;
            } else if (seg == lastSeg) {
                maxIter = true;
                alive = false;
            } else {
                seg++;
            }
        }
        onEvent<POL, NOCB>(P.callback, &sh, maxIter, ray, THEIA_RESULT_MAX_ITER, k, lastSeg + 2);
    }
    sinkFinish(P, &sh, sharedHist, nSegments, nResponses, 0, nPaths);
}

// ------------------------------------------------------------------ VolumePhotonTracer
// sh/tracer.volume.photon.common.glsl:32-123 + sh/tracer.volume.photon.init.glsl:30-62: single
// photons against the target: every event is trace (hit -> response, photon absorbed), scatter,
// absorption roulette. Same persistent-warp form as the forward tracer: the state of a path
// stays in registers, finished lanes take the next path of the warp's range.
template <bool POL>
__global__ void __launch_bounds__(128, 4) volumePhotonKernel(const __grid_constant__ DeviceLaunch P) {
    extern __shared__ float sharedHist[];
    __shared__ SinkShared sh;
    sinkInit(P, &sh, sharedHist);
    float* hist = sinkHist(P, sharedHist);

    const uint32_t warpsPerCta = blockDim.x >> 5;
    const uint32_t nWarps = gridDim.x * warpsPerCta;
    const uint32_t warp = blockIdx.x * warpsPerCta + (threadIdx.x >> 5);
    const uint32_t perWarp = (P.batch_size + nWarps - 1) / nWarps;
    uint32_t next = min(warp * perWarp, P.batch_size);
    const uint32_t end = min(next + perWarp, P.batch_size);

    bool alive = false;
    uint32_t k = 0;   // path number inside the batch
    uint32_t it = 1;  // event number, 1 .. path_length
    Ray ray;
    Rng rng;
    uint32_t nSegments = 0, nResponses = 0, nPaths = 0;

    while (true) {
        uint32_t dead = __ballot_sync(FULL_MASK, !alive);
        bool created = false;
        if (dead) {
            uint32_t avail = end - next;
            uint32_t rank = __popc(dead & laneMaskLt());
            if (!alive && rank < avail) {
                alive = true;
                created = true;
                k = next + rank;
                it = 1;
            }
            next += min((uint32_t)__popc(dead), avail);
        }
        if (!__any_sync(FULL_MASK, alive)) break;
        const uint32_t idx = P.first_path + k;

        // ---- create the path (init :30-62)
        bool aborted = false;
        if (created) {
            rng = makeRng(&P.rng, idx);
            ray = sampleRay<POL>(P.wavelength, P.light, P.medium, rng);
            ray.lin_contrib = 1.0f;
            ray.log_contrib = 0.0f;
            nPaths++;
            aborted = isOccludedByTarget(P.target, ray.position);
        }
        onEvent<POL>(P.callback, &sh, created, ray, THEIA_RESULT_RAY_CREATED, k, 0);
        onEvent<POL>(P.callback, &sh, aborted, ray, THEIA_ERROR_TRACE_ABORT, k, 0);
        if (aborted) alive = false;

        // ---- trace (common :32-76)
        HitItem item;
        bool has = false;
        int result = 0;
        float uResponse = 2.0f;
        if (alive) {
            nSegments++;
            float u = rng.next();
            float dist = sampleScatterLength(ray, P.propagation, u);
            TargetSample hit = intersectTarget(P.target, ray.position, ray.direction);
            bool hitValid = hit.valid && (hit.dist <= dist);
            dist = min_(hit.dist, dist);
            result = propagateRay(ray, dist, P.propagation);
            updateRayIS(ray, dist, P.propagation, hitValid);
            if (result >= 0) {
                if (hitValid) {
                    alignRayToHit<POL>(ray, hit.normal);
                    item = createHit<POL>(ray, hit.objPosition, hit.objNormal, P.objectId, targetWorldToObj(P.target));
                    has = true;
                    uResponse = responseDraw(P.response, rng, true, item);
                    result = THEIA_RESULT_RAY_ABSORBED;  // never detect the same photon twice
                } else {
                    result = THEIA_RESULT_RAY_SCATTERED;
                }
            }
        }
        commitHit(P.response, hist, has, item, idx, it - 1u, nResponses, uResponse);

        // ---- scatter, roulette, event (common :81-108)
        if (alive) {
            scatterRay<POL>(ray, rng.next2());
            result = photonRoulette<POL>(ray, rng, result);
        }
        onEvent<POL>(P.callback, &sh, alive, ray, result, k, it);
        bool maxIter = false;
        if (alive) {
            if (result < 0) {
                alive = false;
            } else if (it == P.path_length) {
                maxIter = true;
                alive = false;
            } else {
                photonRunBoundary(P, rng, it);
                it++;
            }
        }
        onEvent<POL>(P.callback, &sh, maxIter, ray, THEIA_RESULT_MAX_ITER, k, P.path_length + 1);
    }
    sinkFinish(P, &sh, sharedHist, nSegments, nResponses, 0, nPaths);
}

}  // namespace theia
