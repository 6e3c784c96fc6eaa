// SceneForwardTracer kernel (sh/tracer.scene.forward.glsl + sh/scene.traverse.glsl of the
// reference). Same execution model as the volume tracer (persistent warps, dead lanes are
// refilled with unstarted paths by ballot + prefix popcount); one loop iteration advances
// every live lane by one path segment:
//   trace (free path sample, ray query, propagate) -> interaction (surface hit: Fresnel
//   reflect / transmit / detect; volume: next event estimation + phase function scatter).
// Up to three hits per segment are committed at warp-convergent points.
#pragma once
#include "launch.cuh"
#include "surface.cuh"

namespace theia {

static __device__ __noinline__ RayHit rayQuery(const SceneView& sc, vec3 origin, vec3 dir, float tmax) {
    return closestHit<false>(sc, origin, dir, tmax, nullptr);
}

// sh/scene.traverse.glsl:33-68; works on a copy of the ray
template <bool POL>
TDEV void sceneCreateResponse(const DeviceLaunch& P, Ray ray, const SurfaceHit& hit, const Reflectance& surface,
                              bool absorb, HitItem& item) {
    if (!absorb) {
        if (P.hit_use_transmitted_dir) {
            transmitRay<POL>(ray, hit, surface);
        } else {
            vec3 dir = ray.direction;
            transmitRay<POL>(ray, hit, surface);
            ray.direction = dir;
        }
    }
    item = createHit<POL>(ray, hit.objPos, hit.objNrm, hit.customId, hit.worldToObj);
}

// sh/scene.traverse.glsl:160-180; works on a copy of the ray
template <bool POL>
TDEV bool processShadowRay(const DeviceLaunch& P, Ray ray, const SurfaceHit& hit, HitItem& item) {
    bool isTarget = (hit.flags & MATERIAL_DETECTOR_BIT) != 0;
    bool correctId = P.objectId < 0 || P.objectId == hit.customId;
    if (!hit.valid || !correctId || !isTarget) return false;
    if (propagateRayToHit<POL>(ray, hit.worldPos, hit.rayNrm, P.propagation) < 0) return false;
    Reflectance surface = fresnelReflect(ray, hit);
    bool black = (hit.flags & MATERIAL_BLACK_BODY_BIT) != 0;
    sceneCreateResponse<POL>(P, ray, hit, surface, black, item);
    return true;
}

// sh/scene.traverse.glsl:185-214; works on a copy of the ray
template <bool POL>
TDEV bool traceShadowRay(const DeviceLaunch& P, const SceneView& sc, Ray ray, vec3 dir, float dist, float weight,
                         HitItem& item) {
    scatterRayIS<POL>(ray, dir);
    RayHit q = rayQuery(sc, ray.position, dir, dist);
    SurfaceHit hit;
    if (processRayQuery(sc, ray, q, hit) < 0) return false;
    ray.lin_contrib *= weight;
    return processShadowRay<POL>(P, ray, hit, item);
}

template <bool POL>
__global__ void __launch_bounds__(128, 4) sceneForwardKernel(const __grid_constant__ DeviceLaunch P) {
    extern __shared__ float4 sharedRaw[];
    __shared__ SinkShared sh;
    // dynamic shared memory: [top TLAS nodes][histogram]
    WideNode* sharedTop = reinterpret_cast<WideNode*>(sharedRaw);
    float* sharedHist = reinterpret_cast<float*>(sharedTop + THEIA_TOP_NODES);
    const uint32_t nTop = stageTopNodes(P.tlas, sharedTop);
    sinkInit(P, &sh, sharedHist);  // ends with __syncthreads()
    float* hist = sinkHist(P, sharedHist);
    const SceneView sc = makeSceneView(P.tlas, sharedTop, nTop);

    const bool GUIDE = P.guide.kind != THEIA_GUIDE_NONE;
    const uint32_t PATH_LENGTH = P.path_length;

    const uint32_t warpsPerCta = blockDim.x >> 5;
    const uint32_t nWarps = gridDim.x * warpsPerCta;
    const uint32_t warp = blockIdx.x * warpsPerCta + (threadIdx.x >> 5);
    const uint32_t perWarp = (P.batch_size + nWarps - 1) / nWarps;
    uint32_t next = min(warp * perWarp, P.batch_size);
    const uint32_t end = min(next + perWarp, P.batch_size);

    bool alive = false;
    bool allowResponse = false;
    uint32_t k = 0;   // path number inside the batch
    uint32_t it = 1;  // segment number (event index), 1..PATH_LENGTH
    Ray ray;
    Rng rng;
    uint32_t nSegments = 0, nResponses = 0, nShadow = 0, nPaths = 0;

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

        // ---- create the path (sh/tracer.scene.forward.glsl:59-66)
        if (created) {
            rng = makeRng(&P.rng, idx);
            ray = sampleRay<POL>(P.wavelength, P.light, P.medium, rng);
            if (P.photon) { ray.lin_contrib = 1.0f; ray.log_contrib = 0.0f; }  // sh/tracer.scene.photon.init.glsl:50-54
            allowResponse = P.disable_direct_lighting == 0;
            nPaths++;
        }
        onEvent<POL>(P.callback, &sh, created, ray, THEIA_RESULT_RAY_CREATED, k, 0);

        HitItem item;
        bool has = false;
        uint32_t hitSeq = 0;
        const uint32_t tagEvent = it - 1;

        // ---- trace (sh/scene.traverse.glsl:280-352)
        SurfaceHit hit;
        hit.valid = false;
        int result = 0;
        if (alive) {
            nSegments++;
            if (isRayBad(ray)) {
                result = THEIA_ERROR_RAY_BAD;
            } else {
                vec3 dir = normalize(ray.direction);
                float dist = sampleScatterLength(ray, P.propagation, rng.next());
                float sampledDist = dist;
                bool mis_target = false;
                if (GUIDE) {
                    TargetGuideSample guideSample = evalTargetGuide(P.guide, ray.position, dir);
                    mis_target = allowResponse && guideSample.prob > 0.0f;
                    if (mis_target) {
                        mis_target = guideSample.dist > dist;
                        dist = max_(guideSample.dist, dist);
                    }
                }
                RayHit q = rayQuery(sc, ray.position, dir, dist);
                int queryResult = processRayQuery(sc, ray, q, hit);
                if (hit.valid) dist = distance(ray.position, hit.worldPos);
                if (GUIDE && mis_target && hit.valid && dist > sampledDist && queryResult >= 0) {
                    has = processShadowRay<POL>(P, ray, hit, item);
                    hit.valid = false;
                    dist = sampledDist;
                }
                int propResult = propagateRay(ray, dist, P.propagation);
                updateRayIS(ray, dist, P.propagation, hit.valid);
                result = queryResult <= THEIA_ERROR_MAX_VALUE ? queryResult : propResult;
            }
        }
        has = commitHit(P.response, hist, has, item, idx, tagEvent | (hitSeq << 16), nResponses, responseDraw(P.response, rng, has, item));
        hitSeq += has ? 1u : 0u;
        has = false;

        // ---- interaction (sh/scene.traverse.glsl:361-397)
        const bool interact = alive && result >= 0;
        const bool last = P.photon ? false : it == PATH_LENGTH;  // the photon tracers always scatter
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
                    uResponse = responseDraw(P.response, rng, true, item);  // before the reflect / transmit draw
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
        // next event estimation towards the guide (sampleTargetMIS, :234-269)
        vec3 dirTarget = v3(0.0f);
        float distTarget = 0.0f, wTarget = 0.0f;
        if (wantMIS) {
            float pPP;
            vec3 dirPhase = scatter(ray.medium, ray.direction, rng.next2(), &pPP);
            TargetGuideSample targetSample = sampleTargetGuide(P.guide, ray.position, rng);
            float pTT = targetSample.prob;
            TargetGuideSample phaseSample = evalTargetGuide(P.guide, ray.position, dirPhase);
            float pTP = phaseSample.prob;
            float pPT = scatterProb(ray.medium, ray.direction, targetSample.dir);
            wTarget = pTT * pPT / (pTT * pTT + pPT * pPT);
            float wPhase = pPP * pPP / (pPP * pPP + pTP * pTP);
            dirTarget = targetSample.dir;
            distTarget = targetSample.dist;
            nShadow += 2;
            has = traceShadowRay<POL>(P, sc, ray, dirPhase, phaseSample.dist, wPhase, item);
        }
        has = commitHit(P.response, hist, has, item, idx, tagEvent | (hitSeq << 16), nResponses, wantMIS ? responseDraw(P.response, rng, has, item) : uResponse);
        hitSeq += has ? 1u : 0u;
        has = false;
        if (wantMIS) has = traceShadowRay<POL>(P, sc, ray, dirTarget, distTarget, wTarget, item);
        has = commitHit(P.response, hist, has, item, idx, tagEvent | (hitSeq << 16), nResponses, responseDraw(P.response, rng, has, item));
        // volume scatter (:386-393)
        if (interact && !hit.valid && !last) scatterRay<POL>(ray, rng.next2());

        // ---- events and bookkeeping (sh/tracer.scene.forward.glsl:87-104)
        if (P.photon && alive) result = photonRoulette<POL>(ray, rng, result);
        onEvent<POL>(P.callback, &sh, alive, ray, result, k, it);
        bool maxIter = false;
        if (alive) {
            if (result < 0) {
                alive = false;
            } else {
                allowResponse = GUIDE ? (result != THEIA_RESULT_RAY_SCATTERED) : true;
                if (it == PATH_LENGTH) {
                    maxIter = true;
                    alive = false;
                } else {
                    if (P.photon) photonRunBoundary(P, rng, it);
                    it++;
                }
            }
        }
        onEvent<POL>(P.callback, &sh, maxIter, ray, THEIA_RESULT_MAX_ITER, k, PATH_LENGTH + 1);
    }
    sinkFinish(P, &sh, sharedHist, nSegments, nResponses, nShadow, nPaths);
}

}  // namespace theia
