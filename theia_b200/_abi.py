"""ctypes mirrors of the C-ABI structures and enums of include/theia_b200.h."""

from __future__ import annotations

from ctypes import POINTER, Structure, c_float, c_int32, c_uint32, c_uint64, c_void_p

ABI_VERSION = 2

# component kinds
LIGHT_SPHERICAL, LIGHT_CONE, LIGHT_PENCIL, LIGHT_CASCADE, LIGHT_MUON, LIGHT_HOST = 1, 2, 3, 4, 5, 6
LIGHT_CHERENKOV, LIGHT_CHERENKOV_TRACK = 7, 8
LIGHT_FLAG_POLARIZED, LIGHT_FLAG_FRANK_TAMM_IS, LIGHT_FLAG_PHOTON_COUNT = 1, 2, 4
WAVELENGTH_CONST, WAVELENGTH_UNIFORM, WAVELENGTH_HOST, WAVELENGTH_FUNCTION = 1, 2, 3, 4
TARGET_NONE, TARGET_SPHERE, TARGET_INNER_SPHERE, TARGET_FLAT, TARGET_DISK = 0, 1, 2, 3, 4
GUIDE_NONE, GUIDE_SPHERE, GUIDE_FLAT, GUIDE_DISK = 0, 1, 2, 3
RESPONSE_EMPTY, RESPONSE_RECORD, RESPONSE_HISTOGRAM = 0, 1, 2
RESPONSE_STORE_TIME, RESPONSE_KERNEL_HISTOGRAM, RESPONSE_OBJECT_HISTOGRAM = 3, 4, 5
VALUE_UNIFORM = 1
CALLBACK_EMPTY, CALLBACK_STAT, CALLBACK_TRACK = 0, 1, 2
TRACER_VOLUME_FORWARD, TRACER_SCENE_FORWARD, TRACER_VOLUME_PHOTON, TRACER_SCENE_PHOTON = 1, 2, 3, 4


class PhiloxParams(Structure):
    _pack_ = 4
    _fields_ = [("key", c_uint32 * 2), ("offset", c_uint32 * 4)]


class HistogramParams(Structure):
    _pack_ = 4
    _fields_ = [("t0", c_float), ("binSize", c_float)]


class Geometry(Structure):
    _pack_ = 4
    _fields_ = [("vertices", c_uint64), ("indices", c_uint64), ("material", c_uint64)]


class Component(Structure):
    _fields_ = [
        ("kind", c_uint32),
        ("flags", c_uint32),
        ("params", c_void_p),
        ("params_size", c_uint32),
        ("_pad", c_uint32),
    ]


class Response(Structure):
    _fields_ = [
        ("kind", c_uint32),
        ("value_kind", c_uint32),
        ("hist", HistogramParams),
        ("n_bins", c_uint32),
        ("hist_capacity", c_uint32),
        ("hist_partials", c_void_p),
        ("queue", c_void_p),
        ("queue_capacity", c_uint32),
        ("queue_polarized", c_uint32),
        ("tag_path", c_void_p),
        ("tag_event", c_void_p),
        ("kernel_bandwidth", c_float),
        ("kernel_support", c_float),
        ("value_queue", c_void_p),
        ("value_capacity", c_uint32),
        ("store_object_id", c_uint32),
        ("object_hist", c_void_p),
        ("n_objects", c_uint32),
        ("_pad1", c_uint32),
    ]


class Callback(Structure):
    _fields_ = [
        ("kind", c_uint32),
        ("track_polarized", c_uint32),
        ("stats", c_void_p),
        ("track_buffer", c_void_p),
        ("track_count", c_uint32),
        ("track_length", c_uint32),
    ]


class TraceLaunch(Structure):
    _fields_ = [
        ("struct_size", c_uint32),
        ("tracer", c_uint32),
        ("polarized", c_uint32),
        ("disable_mis", c_uint32),
        ("disable_direct_lighting", c_uint32),
        ("disable_transmission", c_uint32),
        ("disable_volume_border", c_uint32),
        ("hit_use_transmitted_dir", c_uint32),
        ("path_length", c_uint32),
        ("block_size", c_uint32),
        ("batch_size", c_uint32),
        ("first_path", c_uint32),
        ("trace_params", c_void_p),
        ("trace_params_size", c_uint32),
        ("_pad0", c_uint32),
        ("rng", PhiloxParams),
        ("light", Component),
        ("wavelength", Component),
        ("target", Component),
        ("guide", Component),
        ("response", Response),
        ("callback", Callback),
        ("tlas", c_void_p),
        ("stream", c_void_p),
        ("work_counters", c_void_p),
        ("workspace", c_void_p),
        ("workspace_bytes", c_uint64),
        ("run_length", c_uint32),
        ("rng_per_run", c_uint32),
    ]


class ASSizes(Structure):
    _fields_ = [("result_bytes", c_uint64), ("scratch_bytes", c_uint64)]


class InstanceDesc(Structure):
    _fields_ = [
        ("obj_to_world", c_float * 12),
        ("world_to_obj", c_float * 12),
        ("blas", c_void_p),
        ("geometry", Geometry),
        ("custom_index", c_int32),
        ("_pad", c_uint32),
    ]


# unit seam (theia_unit_eval)
(UNIT_SCATTER_DIR, UNIT_REFLECTANCE, UNIT_VOLUME_SCATTER, UNIT_SCATTER_PROB, UNIT_ROTATE_PHI, UNIT_ALIGN_POLREF,
 UNIT_ROTATE_DIR, UNIT_PHASE_MATRIX, UNIT_SURFACE, UNIT_SIGN_BIT, UNIT_GAMMA, UNIT_EMISSION_ANGLE,
 UNIT_TARGET_SAMPLE, UNIT_GUIDE_SAMPLE, UNIT_LIGHT_SAMPLE, UNIT_MEDIUM) = range(1, 17)


class UnitArgs(Structure):
    _fields_ = [
        ("struct_size", c_uint32),
        ("unit", c_uint32),
        ("n", c_uint32),
        ("polarized", c_uint32),
        ("in_", c_void_p),
        ("out", c_void_p),
        ("rng", PhiloxParams),
        ("first", c_uint32),
        ("base_count", c_uint32),
        ("medium", c_uint64),
        ("material", c_uint64),
        ("params", c_float * 8),
        ("light", Component),
        ("wavelength", Component),
        ("target", Component),
        ("guide", Component),
        ("light_queue", c_void_p),
        ("wavelength_queue", c_void_p),
        ("capacity", c_uint32),
        ("_pad", c_uint32),
        ("stream", c_void_p),
    ]
