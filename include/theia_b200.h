/*
 * theia_b200.h - C-ABI of the B200-native forward photon tracing path.
 *
 * This is the drop-in boundary. The reference (tkerscher/theia)
 * has no FFI: its boundary is a source-level plugin API in which Python components own
 * ctypes.Structure parameter blocks (layout(scalar), no padding) and raw 64 bit device
 * addresses, and hephaistos dispatches one compute shader per tracer. Every entry point
 * below replaces one such dispatch / driver call; the citation names the reference
 * interface it replaces (paths relative to the reference root, sh/ = src/theia/shader/).
 *
 * Conventions
 *  - plain C, plain pointers and sizes, no torch types.
 *  - every function returns 0 on success, <0 on error; theia_last_error() returns a
 *    thread-local message for the last failing call.
 *  - "dev" pointers are CUDA device pointers (torch.Tensor.data_ptr()); "host" pointers
 *    are ordinary host memory and are only read during the call.
 *  - the caller owns all memory. The library never allocates device memory and never
 *    keeps a pointer beyond the call, except inside acceleration structures it wrote into
 *    a caller supplied buffer.
 *  - work is enqueued on the caller supplied cudaStream_t (void*), no synchronisation.
 *  - parameter blocks are byte-identical to the reference's ctypes.Structure mirrors.
 */
#ifndef THEIA_B200_H
#define THEIA_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define THEIA_ABI_VERSION 2

/* ------------------------------------------------------------------ result codes */
/* sh/result.glsl:10-28, src/theia/trace.py:308-343 */
enum {
    THEIA_RESULT_SUCCESS = 0,
    THEIA_RESULT_RAY_CREATED = 1,
    THEIA_RESULT_RAY_SCATTERED = 2,
    THEIA_RESULT_RAY_HIT = 3,
    THEIA_RESULT_RAY_DETECTED = 4,
    THEIA_RESULT_VOLUME_HIT = 5,
    THEIA_RESULT_RAY_LOST = -1,
    THEIA_RESULT_RAY_DECAYED = -2,
    THEIA_RESULT_RAY_ABSORBED = -3,
    THEIA_RESULT_RAY_MISSED = -4,
    THEIA_RESULT_MAX_ITER = -5,
    THEIA_ERROR_MAX_VALUE = -10,
    THEIA_ERROR_UNKNOWN = -10,
    THEIA_ERROR_MEDIA_MISMATCH = -11,
    THEIA_ERROR_TRACE_ABORT = -12,
    THEIA_ERROR_RAY_BAD = -13
};

/* ------------------------------------------------------------------ parameter blocks
 * All blocks use scalar layout: 4 byte alignment, no padding (sh/preamble.glsl:20-21). */
#pragma pack(push, 4)

/* src/theia/random.py:202-254, sh/random.philox.glsl:8-11 */
typedef struct theia_philox_params {
    uint32_t key[2];    /* 64 bit key, little endian words */
    uint32_t offset[4]; /* 128 bit counter offset, little endian words */
} theia_philox_params_t;

/* sh/ray.propagate.glsl:18-26 */
typedef struct theia_propagate_params {
    float scatterCoefficient;
    float lowerBBoxCorner[3];
    float upperBBoxCorner[3];
    float maxTime;
    float maxDist;
} theia_propagate_params_t;

/* src/theia/trace.py:589-598, sh/tracer.volume.forward.glsl:18-23 */
typedef struct theia_volume_trace_params {
    uint64_t medium;
    int32_t objectId;
    theia_propagate_params_t propagation;
} theia_volume_trace_params_t;

/* src/theia/trace.py:1136-1145, sh/tracer.scene.forward.glsl:36-41 */
typedef struct theia_scene_trace_params {
    int32_t targetId;
    uint64_t sourceMedium;
    theia_propagate_params_t propagation;
} theia_scene_trace_params_t;

/* light sources -------------------------------------------------------------- */
/* src/theia/light.py:1133-1139, sh/lightsource.spherical.glsl:7-14 */
typedef struct theia_light_spherical {
    float position[3];
    float contribFwd;
    float contribBwd;
    float t_min;
    float t_max;
} theia_light_spherical_t;

/* src/theia/light.py:934-944, sh/lightsource.cone.glsl:6-20 */
typedef struct theia_light_cone {
    float position[3];
    float direction[3];
    float cosOpeningAngle;
    float contribFwd;
    float contribBwd;
    float t_min;
    float t_max;
    float polRef[3];
    float stokes[4];
} theia_light_cone_t;

/* src/theia/light.py:1064-1072, sh/lightsource.pencil.glsl:4-16 */
typedef struct theia_light_pencil {
    float position[3];
    float direction[3];
    float budget;
    float t_min;
    float t_max;
    float stokes[4];
    float polRef[3];
} theia_light_pencil_t;

/* src/theia/light.py:1590-1600, sh/lightsource.particles.cascade.glsl:13-38 */
typedef struct theia_light_cascade {
    float startPosition[3];
    float startTime;
    float direction[3];
    float effectiveLength;
    float a_angular;
    float b_angular;
    float a_long;
    float b_long;
} theia_light_cascade_t;

/* src/theia/light.py:1452-1461, sh/lightsource.particles.muon.glsl:13-26 */
typedef struct theia_light_muon {
    float startPosition[3];
    float startTime;
    float endPosition[3];
    float endTime;
    float energyScale;
    float a_angular;
    float b_angular;
} theia_light_muon_t;

/* src/theia/light.py:1215-1223 (CherenkovLightSource), sh/lightsource.cherenkov.simple.glsl:9-16 */
typedef struct theia_light_cherenkov {
    float trackStart[3];
    float startTime;
    float trackEnd[3];
    float endTime;
    float trackDir[3];
    float trackDist;
} theia_light_cherenkov_t;

/* src/theia/light.py:1377-1378 (CherenkovTrackLightSource), sh/lightsource.cherenkov.track.glsl:9-21:
 * track = dev address of { uint32 trackLength; { float x, y, z, t; } vertices[trackLength + 1]; } */
typedef struct theia_light_cherenkov_track {
    uint64_t track;
} theia_light_cherenkov_track_t;

/* src/theia/light.py:692-766 (HostLightSource), sh/lightsource.host.glsl:6-17 with the queue of
 * sh/lightsource.queue.glsl:11-33: field major float arrays of `capacity` entries each, NO count
 * word - posXYZ, dirXYZ, [stokes IQUV, polRef XYZ when polarized], startTime, contrib.
 * Path k of a launch (RNG stream first_path + k) reads entry k. */
typedef struct theia_light_host {
    uint64_t queue; /* dev: float[8 or 15][capacity] */
    uint32_t capacity;
    uint32_t polarized; /* LIGHT_QUEUE_POLARIZED */
} theia_light_host_t;

/* wavelength sources ---------------------------------------------------------- */
/* src/theia/light.py:87-144 (HostWavelengthSource), sh/wavelengthsource.host.glsl:6-13 with the
 * queue of sh/wavelengthsource.queue.glsl:11-14: float wavelength[capacity], contrib[capacity] */
/* src/theia/light.py:372-373 (FunctionWavelengthSource), sh/wavelengthsource.function.glsl:6-11:
 * table = dev address of a Table1D { int32 nx; float samples[nx]; } holding the inverse CDF */
typedef struct theia_wavelength_function {
    uint64_t table;
    float contrib;
} theia_wavelength_function_t;

typedef struct theia_wavelength_host {
    uint64_t queue; /* dev: float[2][capacity] */
    uint32_t capacity;
    uint32_t _pad;
} theia_wavelength_host_t;

/* src/theia/light.py:272-273, sh/wavelengthsource.const.glsl:6-8 */
typedef struct theia_wavelength_const {
    float lambda;
} theia_wavelength_const_t;

/* src/theia/light.py:308-312, sh/wavelengthsource.uniform.glsl:6-10 */
typedef struct theia_wavelength_uniform {
    float lam_min;
    float lam_max;
    float contrib;
} theia_wavelength_uniform_t;

/* targets / guides ------------------------------------------------------------ */
/* src/theia/target.py:105-112 and 169-176, sh/target.sphere.glsl:8-16,
 * sh/target.sphere.inner.glsl:8-16 (last field: hemisphereProb resp. prob) */
typedef struct theia_target_sphere {
    float position[3];
    float radius;
    float invPos[3];
    float invRad;
    float prob;
} theia_target_sphere_t;

/* src/theia/target.py:262-270, sh/target.flat.glsl:6-13; the same block serves the flat target
 * guide (src/theia/target.py:574-582, sh/target_guide.flat.glsl:6-13). objToWorld is GLSL's
 * column major mat3: element [3 c + r]. */
typedef struct theia_target_flat {
    float width;
    float length; /* "height" in the shader and in the guide */
    float position[3];
    float normal[3];
    float prob;
    float objToWorld[9];
} theia_target_flat_t;

/* src/theia/target.py:363-370, sh/target.disk.glsl:6-12; the same block serves the disk target
 * guide (src/theia/target.py:672-679, sh/target_guide.disk.glsl:6-12) */
typedef struct theia_target_disk {
    float position[3];
    float radius;
    float normal[3];
    float prob;
    float objToWorld[9];
} theia_target_disk_t;

/* src/theia/target.py:497-502, sh/target_guide.sphere.glsl:7-10 */
typedef struct theia_guide_sphere {
    float position[3];
    float radius;
    float prob; /* unused by the device code, kept for layout parity */
} theia_guide_sphere_t;

/* src/theia/response.py:1264-1265, sh/response.histogram.glsl:11-14 */
typedef struct theia_histogram_params {
    float t0;
    float binSize;
} theia_histogram_params_t;

/* src/theia/scene.py:628-635, sh/scene.types.glsl:16-20 */
typedef struct theia_geometry {
    uint64_t vertices; /* dev: float[nVerts][6] position, normal */
    uint64_t indices;  /* dev: int32[nTris][3] */
    uint64_t material; /* dev: theia Material (24 B) or 0 */
} theia_geometry_t;

#pragma pack(pop)

/* ------------------------------------------------------------------ component kinds */
enum { THEIA_LIGHT_SPHERICAL = 1, THEIA_LIGHT_CONE = 2, THEIA_LIGHT_PENCIL = 3,
       THEIA_LIGHT_CASCADE = 4, THEIA_LIGHT_MUON = 5, THEIA_LIGHT_HOST = 6, THEIA_LIGHT_CHERENKOV = 7,
       THEIA_LIGHT_CHERENKOV_TRACK = 8 };
/* light flags (reference: preamble macros of the light source snippets) */
enum { THEIA_LIGHT_FLAG_POLARIZED = 1,     /* LIGHTSOURCE_POLARIZED, light.py:1009-1011 */
       THEIA_LIGHT_FLAG_FRANK_TAMM_IS = 2, /* FRANK_TAMM_IS, light.py:1517-1519 */
       THEIA_LIGHT_FLAG_PHOTON_COUNT = 4   /* FRANK_TAMM_USE_PHOTON_COUNT, light.py:1258-1262 */ };
enum { THEIA_WAVELENGTH_CONST = 1, THEIA_WAVELENGTH_UNIFORM = 2, THEIA_WAVELENGTH_HOST = 3,
       THEIA_WAVELENGTH_FUNCTION = 4 };
enum { THEIA_TARGET_NONE = 0, THEIA_TARGET_SPHERE = 1, THEIA_TARGET_INNER_SPHERE = 2, THEIA_TARGET_FLAT = 3,
       THEIA_TARGET_DISK = 4 };
enum { THEIA_GUIDE_NONE = 0, THEIA_GUIDE_SPHERE = 1, THEIA_GUIDE_FLAT = 2, THEIA_GUIDE_DISK = 3 };
enum { THEIA_RESPONSE_EMPTY = 0, THEIA_RESPONSE_RECORD = 1, THEIA_RESPONSE_HISTOGRAM = 2,
       THEIA_RESPONSE_STORE_TIME = 3,       /* sh/response.time.store.glsl */
       THEIA_RESPONSE_KERNEL_HISTOGRAM = 4, /* sh/response.histogram.kernel.glsl */
       THEIA_RESPONSE_OBJECT_HISTOGRAM = 5  /* B200 extension: one time histogram per objectId */ };
enum { THEIA_VALUE_UNIFORM = 1 };
enum { THEIA_CALLBACK_EMPTY = 0, THEIA_CALLBACK_STAT = 1, THEIA_CALLBACK_TRACK = 2 };
enum { THEIA_TRACER_VOLUME_FORWARD = 1, THEIA_TRACER_SCENE_FORWARD = 2,
       /* single photon tracers with absorption roulette: src/theia/trace.py:2657-2959 and 2370-2654,
        * sh/tracer.volume.photon.*.glsl, sh/tracer.scene.photon.*.glsl. They take the TraceParams
        * block of the matching forward tracer. */
       THEIA_TRACER_VOLUME_PHOTON = 3, THEIA_TRACER_SCENE_PHOTON = 4 };

/* A component = kind + flags + host pointer to its byte-identical parameter block. */
typedef struct theia_component {
    uint32_t kind;
    uint32_t flags;
    const void *params;   /* host */
    uint32_t params_size; /* bytes, checked against the kind */
    uint32_t _pad;
} theia_component_t;

/* Hit response (sh/response.*.glsl). */
typedef struct theia_response {
    uint32_t kind;       /* THEIA_RESPONSE_* */
    uint32_t value_kind; /* THEIA_VALUE_* (histogram only) */
    /* histogram: sh/response.histogram.glsl, src/theia/response.py:1200-1421 */
    theia_histogram_params_t hist;
    uint32_t n_bins;
    uint32_t hist_capacity; /* number of partial histograms hist_partials can hold */
    float *hist_partials;   /* dev: float[hist_capacity][n_bins]; partial i is written by CTA i */
    /* record: sh/response.record.glsl, sh/response.queue.glsl:6-33 (SoA, count first) */
    uint32_t *queue;        /* dev: uint32 count; then field-major float/int arrays */
    uint32_t queue_capacity;
    uint32_t queue_polarized;
    /* B200 extension (test seam): if non-null, the path index and the event number that
     * produced hit k are written to tag_path[k], tag_event[k]. */
    uint32_t *tag_path;
    uint32_t *tag_event;
    /* kernel histogram: sh/response.histogram.kernel.glsl:13-20 (ResponseParams behind t0, binSize),
     * src/theia/response.py:1424-1673. Uses hist, n_bins, hist_capacity, hist_partials like the
     * plain histogram; every hit is spread over the bins within kernel_support of its time with
     * the weights of a Gaussian of width kernel_bandwidth. */
    float kernel_bandwidth;
    float kernel_support;
    /* store time: sh/response.time.store.glsl:4-26, src/theia/response.py:656-797. A hit is kept
     * with probability responseValue(hit) (ONE draw of the path's RNG stream per hit, whatever
     * the outcome). value_queue = uint32 count; float time[value_capacity];
     * int32 objectId[value_capacity] (only if store_object_id). */
    uint32_t *value_queue;
    uint32_t value_capacity;
    uint32_t store_object_id; /* RESPONSE_STORE_OBJECT_ID */
    /* object histogram (B200 extension; what config C5's per module light curves need and the
     * reference can only get from HitRecorder + host binning): hits with 0 <= objectId < n_objects
     * add their value to object_hist[objectId][bin]; hist, n_bins give the binning. Accumulated
     * in place with float atomics; the caller clears and normalises. */
    float *object_hist; /* dev: float[n_objects][n_bins] */
    uint32_t n_objects;
    uint32_t _pad1;
} theia_response_t;

/* Trace event callback (sh/callback.*.glsl, src/theia/trace.py:77-305). */
typedef struct theia_callback {
    uint32_t kind;            /* THEIA_CALLBACK_* */
    uint32_t track_polarized; /* TRACK_POLARIZED */
    uint64_t *stats;          /* dev: 12 x uint64 in the order of trace.py:86-100 */
    void *track_buffer;       /* dev: layout of sh/callback.track.glsl:6-25 */
    uint32_t track_count;     /* TRACK_COUNT */
    uint32_t track_length;    /* TRACK_LENGTH */
} theia_callback_t;

/* One launch of a forward tracer. Replaces Tracer.run(i) -> program.dispatchIndirect
 * (src/theia/trace.py:762-770, 1327-1336, 478-481). */
typedef struct theia_trace_launch {
    uint32_t struct_size; /* sizeof(theia_trace_launch_t) */
    uint32_t tracer;      /* THEIA_TRACER_* */
    /* variant key = the reference's preamble macros (trace.py:666-673, 1225-1235) */
    uint32_t polarized;               /* POLARIZATION */
    uint32_t disable_mis;             /* DISABLE_MIS */
    uint32_t disable_direct_lighting; /* DISABLE_DIRECT_LIGHTING */
    uint32_t disable_transmission;    /* DISABLE_TRANSMISSION (scene) */
    uint32_t disable_volume_border;   /* DISABLE_VOLUME_BORDER (scene) */
    uint32_t hit_use_transmitted_dir; /* HIT_USE_TRANSMITTED_DIR (scene) */
    uint32_t path_length;             /* PATH_LENGTH */
    uint32_t block_size;              /* BLOCK_SIZE; only used to size partial histograms */
    /* DispatchParams (trace.py:354-360) */
    uint32_t batch_size;
    /* B200 extension: global index of the first path of this launch. Path k of the launch
     * uses RNG stream first_path + k, so one batch can be split over several GPUs. */
    uint32_t first_path;
    /* TraceParams: theia_volume_trace_params_t or theia_scene_trace_params_t (host) */
    const void *trace_params;
    uint32_t trace_params_size;
    uint32_t _pad0;
    theia_philox_params_t rng;
    theia_component_t light;
    theia_component_t wavelength;
    theia_component_t target; /* volume tracer */
    theia_component_t guide;  /* scene tracer, THEIA_GUIDE_NONE = no MIS */
    theia_response_t response;
    theia_callback_t callback;
    const void *tlas; /* dev: acceleration structure written by theia_tlas_build (scene) */
    void *stream;     /* cudaStream_t */
    /* B200 extension (measurement): if non-null, uint64[4] on the device that the launch adds
     * to: path segments (trace() executions, the reference's EventStatisticCallback sum),
     * responses (hits), shadow rays, paths started. */
    uint64_t *work_counters;
    /* B200 extension: device scratch memory for the wavefront form of the scene tracer
     * (path state, ray / result buffers, queues; size from theia_trace_workspace_size).
     * NULL selects the single-launch form (every path stays in the registers of one thread),
     * which is meant for small batches. Contents need not be preserved between launches. */
    void *workspace;
    uint64_t workspace_bytes;
    /* Photon tracers: path_length = nRuns * run_length scattering events in total. The reference
     * dispatches one shader per run of run_length events and restarts the draw counter of run r
     * at r * rng_per_run (Push.dim, src/theia/trace.py:2645-2650, 2946-2950); here a path simply
     * continues and only its draw counter is set to that value at every run boundary. */
    uint32_t run_length;  /* nScatteringPerRun (PATH_LENGTH of the photon shaders) */
    uint32_t rng_per_run; /* _nRNGPerRun */
} theia_trace_launch_t;

/* ------------------------------------------------------------------ library */
const char *theia_last_error(void);
int theia_abi_version(void);
/* Number of CUDA kernels this library has launched in this process (measurement aid). */
uint64_t theia_kernel_launches(void);

/* Tuning / test seam (no counterpart in the reference). Options:
 *   "wf_trace_kernel", "wf_shadow_trace_kernel": ray query kernel of the wavefront scene tracer for
 *       primary / shadow rays: -1 (default) = picked per scene from timings of the first nine
 *       batches, 0 = phased persistent kernel, 1 = one thread per ray, 2 = queued (a warp level queue of triangle
 *       tests, trace_queue.cuh). Results do not depend on it.
 *   "wf_flush_waiting": thresholds of the queued kernel (waiting lanes | instance lanes << 8 | refill lanes << 16).
 *   "wf_shadow_cull": 1 (default): shadow rays of the target guide that cannot reach the box around the instances the
 *       target id selects are not traced (they cannot create a response; results are identical); 0: all are traced.
 *   "query_kernel": kernel behind theia_tlas_query: 0 = direct traversal (default), 1 = phased, 2 = queued.
 *   "wf_min_node_lanes": 1..32, node phase threshold of the phased kernel (default 16).
 * theia_get_trace_choice reports what was picked for a scene (-1 = still measuring). */
int theia_set_option(const char *name, int value);
int theia_get_trace_choice(const void *tlas_dev, uint32_t batch_size, int guide, int *primary, int *shadow);

/* Optional per kernel timing for benchmarks (no counterpart in the reference, which times
 * whole pipelines only): while enabled, every launch of a tracing kernel is bracketed by CUDA
 * events on its stream. theia_profile_enable(1) clears the accumulators; theia_profile_read
 * waits for the pending events and returns, per kernel class k < n, the summed duration in
 * milliseconds and the number of launches; its return value is the number of classes.
 * Not thread safe; meant for a single benchmarking thread. */
int theia_profile_enable(int on);
int theia_profile_read(double *ms, uint64_t *launches, uint32_t n);
const char *theia_profile_class_name(uint32_t cls);
/* Measurement aids for the roofline block of bench.py (no counterpart in the reference).
 * theia_trace_stats_enable(stats_dev): while stats_dev != NULL every ray query launch of the wavefront scene tracer is
 *   preceded by an instrumented pass over the same rays that adds, for primary rays to stats_dev[0..3] and for shadow
 *   rays to stats_dev[4..7], {nodes visited, triangles tested, instances entered, rays} of the one-thread-per-ray
 *   traversal of the same structure (the algorithmic visit counts). NULL switches it off. Not thread safe.
 * theia_l2_probe: reads a buffer of `bytes` (choose it well below the L2 size) `iters` times with 128 bit loads that
 *   bypass L1; bytes * iters / elapsed time (CUDA events of the caller) is the L2 read bandwidth of the device. */
int theia_trace_stats_enable(uint64_t *stats_dev);
int theia_l2_probe(const void *buf_dev, uint64_t bytes, uint32_t iters, uint32_t *sink_dev, void *stream);
/* Number of SMs etc. of the current device; used by the host to size buffers. */
int theia_device_info(int *sm_count, int *cc_major, int *cc_minor);

/* ------------------------------------------------------------------ RNG
 * Replaces the RNGBufferSink dispatch (src/theia/random.py:44-199, sh/random.sink.glsl):
 * out[s * samples + c] = random_s(baseStream + s, baseCount + c) for dim == 1,
 * out[2*(s*samples+c) + {0,1}] = random2D_s(...) for dim == 2. */
int theia_philox_fill(const theia_philox_params_t *params, uint32_t baseStream,
                      uint32_t baseCount, uint32_t streams, uint32_t samples, uint32_t dim,
                      float *out_dev, void *stream);
/* Raw Philox4x32-10 block function (known-answer tests): out[4*i..] = philox(ctr[i], key[i]). */
int theia_philox_raw(const uint32_t *ctr_dev, const uint32_t *key_dev, uint32_t n,
                     uint32_t *out_dev, void *stream);

/* ------------------------------------------------------------------ acceleration structures
 * Replace hp.GeometryStore (BLAS build, src/theia/scene.py:550) and
 * hp.AccelerationStructure (TLAS build, src/theia/scene.py:668). Vulkan style two step
 * build: query sizes, caller allocates, library writes the structure into the buffer. */
typedef struct theia_as_sizes {
    uint64_t result_bytes;  /* buffer that will hold the finished structure */
    uint64_t scratch_bytes; /* temporary build memory */
} theia_as_sizes_t;

int theia_blas_sizes(uint32_t n_triangles, theia_as_sizes_t *sizes);
/* vertices_dev: float[nVerts][6], indices_dev: uint32[nTris][3] (scene.py:434-440).
 * Both must stay valid while the BLAS is in use (hits fetch vertices from them). */
int theia_blas_build(const float *vertices_dev, uint32_t n_vertices,
                     const uint32_t *indices_dev, uint32_t n_triangles, void *result_dev,
                     uint64_t result_bytes, void *scratch_dev, uint64_t scratch_bytes,
                     void *stream);

/* One TLAS instance (hp.GeometryInstance + Scene.GLSLGeometry, scene.py:595-604, 652-666). */
typedef struct theia_instance_desc {
    float obj_to_world[12]; /* 3x4 row major */
    float world_to_obj[12]; /* 3x4 row major, inverse of the above */
    const void *blas;       /* dev: result buffer of theia_blas_build */
    theia_geometry_t geometry;
    int32_t custom_index;   /* detectorId */
    uint32_t _pad;
} theia_instance_desc_t;

int theia_tlas_sizes(uint32_t n_instances, theia_as_sizes_t *sizes);
/* instances: HOST array. Instance order = instance id (scene.py:652-668). */
int theia_tlas_build(const theia_instance_desc_t *instances, uint32_t n_instances,
                     void *result_dev, uint64_t result_bytes, void *scratch_dev,
                     uint64_t scratch_bytes, void *stream);

/* Test seam for the ray query: closest hit of n rays against a TLAS.
 * rays_dev: float[n][8] = origin xyz, tmax, direction xyz, unused.
 * out_dev: int32[n][4] = instance id (-1 = miss), primitive id, custom index, 0;
 * out_bary_dev: float[n][4] = t, u, v, 0. Also reports traversal statistics if
 * stats_dev != NULL: uint64[4] = nodes visited, triangles tested, instances entered, rays. */
int theia_tlas_query(const void *tlas_dev, const float *rays_dev, uint32_t n, int32_t *out_dev,
                     float *out_bary_dev, uint64_t *stats_dev, void *stream);

/* The same ray queries for callers that must not allocate (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc). No counterpart in the reference.
 * It runs the queued wavefront ray query kernel with its scratch memory (ray counter, work cursor, one 16 byte
 * result record per ray) in a static device array of the library, and calls no cudaMalloc / cudaFree / allocating
 * routine. Limits: n <= 262144 rays per call (more is rejected with an error; split the rays over several calls);
 * every GPU that loads the library holds the array (4 MiB + 256 B) until the process exits. All calls share the
 * array: calls that overlap on the GPU (different streams, or graphs replayed concurrently) race; calls on one
 * stream are safe, each call rewrites what it reads. Same outputs, bit for bit, as theia_tlas_query (no statistics). */
int theia_tlas_query_static(const void *tlas_dev, const float *rays_dev, uint32_t n, int32_t *out_dev,
                            float *out_bary_dev, void *stream);

/* ------------------------------------------------------------------ tracers */
/* Number of CTAs the launch will use = number of partial histograms it writes. */
int theia_trace_grid_size(const theia_trace_launch_t *launch, uint32_t *n_blocks);
/* Bytes of theia_trace_launch_t.workspace the launch needs to run in wavefront form
 * (0 for the volume tracer, which has no wavefront form). */
int theia_trace_workspace_size(const theia_trace_launch_t *launch, uint64_t *bytes);
/* Runs VolumeForwardTracer / SceneForwardTracer for one batch. */
int theia_trace_forward(const theia_trace_launch_t *launch);

/* ------------------------------------------------------------------ responses */
/* Replaces the HistogramReducer dispatch (sh/estimator.reduce.glsl:17-35,
 * src/theia/response.py:1188-1197): out[b] = norm * kahan_sum_h partial[h][b]; the partials
 * are cleared afterwards when clear != 0 (hp.clearTensor(histIn)). */
int theia_hist_reduce(float *partials_dev, uint32_t n_hist, uint32_t n_bins, float norm,
                      float *out_dev, int clear, void *stream);
/* Replaces the HitReplay dispatch (sh/response.replay.glsl:35-50): feeds the hits of a
 * queue (same SoA layout as theia_response_t.queue) to a response. */
int theia_hit_replay(const uint32_t *queue_dev, uint32_t capacity, uint32_t polarized,
                     const theia_response_t *response, uint32_t block_size, void *stream);
/* The same with the replay stage's optional random number generator (src/theia/response.py:278-422, HitReplay(rng=...)):
 * hit k draws from stream k, first number first (sh/response.replay.glsl:37-38). rng may be NULL; the store time
 * response needs one. */
int theia_hit_replay_rng(const uint32_t *queue_dev, uint32_t capacity, uint32_t polarized,
                         const theia_response_t *response, const theia_philox_params_t *rng,
                         uint32_t block_size, void *stream);

/* ------------------------------------------------------------------ device math test seam
 * Evaluates the library's deterministic float math on the device (fn: 0 sin, 1 cos, 2 exp,
 * 3 log, 4 pow(x,y), 5 tanh, 6 atan). Used by the parity tests only. */
int theia_math_eval(uint32_t fn, const float *x_dev, const float *y_dev, uint32_t n,
                    float *out_dev, void *stream);

/* ------------------------------------------------------------------ component / shader unit seam
 * One function of the path evaluated for n independent inputs: the counterpart of the reference's
 * per-function test programs (tests/shader/<unit>.test.glsl, driven by tests/test_scatter.py,
 * test_polarization.py, test_ray.py, test_math.py, test_random.py, test_light.py:721-836) and of
 * its sampler stages (LightSampler src/theia/light.py:492-690 + sh/lightsource.sample.glsl,
 * theia.testing.TargetSampler / TargetGuideSampler src/theia/testing.py:429-638 +
 * sh/target.sample.glsl, sh/target_guide.sample.glsl).
 * in_dev: float[n][in_width] (uint fields as raw bits), out_dev: float[n][out_width], row major;
 * element i draws from Philox stream first + i starting at draw base_count.
 *
 * unit                      in row                              out row
 * SCATTER_DIR               inDir3, cos_theta, phi              dir3               (sh/scatter.volume.glsl:7-26)
 * REFLECTANCE  [material]   dir3, normal3, wavelength           r_s, r_p, n_tr     (sh/scatter.surface.glsl:21-51)
 * VOLUME_SCATTER [rng]      dir3, medium lo, medium hi          dir3, prob         (sh/scatter.volume.glsl:28-58)
 * SCATTER_PROB              inDir3, outDir3, medium lo, hi      prob               (sh/scatter.volume.glsl:60-65)
 * ROTATE_PHI                stokes4, phi                        stokes4            (sh/polarization.glsl:7-21)
 * ALIGN_POLREF [rng]        -                                   mat4 (column major): alignPolRef o rotatePolRef
 * ROTATE_DIR                stokes4, dir3, ref3, phi, theta     stokes4, ref3      (sh/polarization.glsl:44-75)
 * PHASE_MATRIX [medium]     stokes4, cos_theta                  stokes4            (sh/polarization.glsl:84-104)
 * SURFACE [rng, material, params = normal3, lam_min, lam_max; polarized]
 *                           -                                   wavelength, dir3, contrib, refl pos3, dir3, contrib,
 *                                                               trans pos3, dir3, contrib, refl polRef3, trans polRef3,
 *                                                               refl stokes4, trans stokes4  (tests/shader/ray.surface.test.glsl, FORWARD)
 * SIGN_BIT                  x                                   signBit(x)         (sh/math.glsl:18-20)
 * GAMMA [rng, params = alpha, lambda]   -                       sampleGamma        (sh/random.gamma.glsl:9-33)
 * EMISSION_ANGLE [params = n, a, b]     u                       cos_theta          (sh/lightsource.particles.common.glsl:52-97)
 * MEDIUM [medium | material, params[0] = 0 inside / 1 outside]   wavelength, cos_theta, eta   ->
 *                           n, vg, mu_s, mu_e, log_phase(cos_theta), phase_sampling(eta), m12, m22, m33, m34(cos_theta),
 *                           flagsInwards, flagsOutwards (bits; 0 without material)
 *                           (tests/shader/medium.test.glsl, tests/shader/material.test.glsl)
 * TARGET_SAMPLE [rng, target, params = min3, max3]   -          observer3, direction3, sample {pos3, nrm3, prob, valid, error},
 *                                                               hit {pos3, nrm3, prob, valid, error}, occluded   (29 words)
 * GUIDE_SAMPLE [rng, guide, params = min3, max3]     -          observer3, sample {dir3, dist, prob}, eval {dir3, dist, prob}
 * LIGHT_SAMPLE [rng, wavelength, light, medium; polarized]: writes the reference's queues instead of rows:
 *     light_queue_dev = uint32 count + field major LightSourceQueue (sh/lightsource.queue.glsl) of `capacity`
 *     entries, wavelength_queue_dev likewise (sh/wavelengthsource.queue.glsl); count := n. */
enum theia_unit {
    THEIA_UNIT_SCATTER_DIR = 1, THEIA_UNIT_REFLECTANCE, THEIA_UNIT_VOLUME_SCATTER, THEIA_UNIT_SCATTER_PROB,
    THEIA_UNIT_ROTATE_PHI, THEIA_UNIT_ALIGN_POLREF, THEIA_UNIT_ROTATE_DIR, THEIA_UNIT_PHASE_MATRIX,
    THEIA_UNIT_SURFACE, THEIA_UNIT_SIGN_BIT, THEIA_UNIT_GAMMA, THEIA_UNIT_EMISSION_ANGLE,
    THEIA_UNIT_TARGET_SAMPLE, THEIA_UNIT_GUIDE_SAMPLE, THEIA_UNIT_LIGHT_SAMPLE, THEIA_UNIT_MEDIUM, THEIA_UNIT_COUNT
};
typedef struct theia_unit_args {
    uint32_t struct_size; /* sizeof(theia_unit_args_t) */
    uint32_t unit;
    uint32_t n;
    uint32_t polarized;
    const float *in;      /* dev, may be NULL for units without input row */
    float *out;           /* dev */
    theia_philox_params_t rng;
    uint32_t first;       /* stream / idx of element 0 */
    uint32_t base_count;  /* first draw (dim) */
    uint64_t medium;      /* dev address of a medium (0 = vacuum) */
    uint64_t material;    /* dev address of a material */
    float params[8];
    theia_component_t light, wavelength, target, guide;
    uint32_t *light_queue;      /* dev, LIGHT_SAMPLE only */
    uint32_t *wavelength_queue; /* dev, LIGHT_SAMPLE only */
    uint32_t capacity;          /* of both queues */
    uint32_t _pad;
    void *stream;
} theia_unit_args_t;
/* Row widths (in floats) of a unit; LIGHT_SAMPLE reports 0 / 0. */
int theia_unit_widths(uint32_t unit, uint32_t polarized, uint32_t *in_width, uint32_t *out_width);
int theia_unit_eval(const theia_unit_args_t *args);

#ifdef __cplusplus
}
#endif
#endif /* THEIA_B200_H */
