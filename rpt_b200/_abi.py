"""ctypes declarations of include/rptx.h and the loader of librptx.so.

The library is the product; there is no CPU fallback.  If it is missing or cannot be loaded,
importing the renderer fails loudly.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# RPTX_LIBRARY: developer knob for A/B builds of the same ABI (tools/); the product is librptx.so next to this file
LIB_PATH = os.environ.get("RPTX_LIBRARY") or os.path.join(_HERE, "librptx.so")

SHAPE_SPHERE, SHAPE_PLANE, SHAPE_CUBE, SHAPE_MESH, SHAPE_GROUP, SHAPE_MONOMIAL = range(6)
MAT_LAMBERTIAN, MAT_PHONG, MAT_MIRROR, MAT_TRANSMISSIVE = range(4)
LIGHT_POINT, LIGHT_AMBIENT, LIGHT_DIRECTIONAL, LIGHT_OBJECT = range(4)
MEDIUM_NONE, MEDIUM_HOMOGENEOUS_ISOTROPIC, MEDIUM_COLORED_GLOWING_FOG = range(3)
PHOTON_MAP, PHOTON_POINT_BEAM, PHOTON_BEAM_BEAM = range(3)


class ShapeDesc(C.Structure):
    pass


ShapeDesc._fields_ = [
    ("kind", C.c_int32),
    ("has_transform", C.c_int32),
    ("transform", C.c_double * 16),
    ("mesh_id", C.c_uint32),
    ("n_children", C.c_uint32),
    ("plane_normal", C.c_double * 3),
    ("plane_value", C.c_double),
    ("mono_height", C.c_double),
    ("mono_exp", C.c_double),
    ("children", C.POINTER(ShapeDesc)),
]


class MaterialDesc(C.Structure):
    _fields_ = [
        ("kind", C.c_int32),
        ("_pad", C.c_int32),
        ("albedo", C.c_double * 3),
        ("emittance", C.c_double),
        ("shininess", C.c_double),
        ("ior", C.c_double),
    ]


class LightDesc(C.Structure):
    _fields_ = [
        ("kind", C.c_int32),
        ("_pad", C.c_int32),
        ("color", C.c_double * 3),
        ("vec", C.c_double * 3),
        ("shape", ShapeDesc),
        ("material", MaterialDesc),
    ]


class CameraDesc(C.Structure):
    _fields_ = [
        ("eye", C.c_double * 3),
        ("direction", C.c_double * 3),
        ("up", C.c_double * 3),
        ("fov", C.c_double),
        ("aperture", C.c_double),
        ("focal_distance", C.c_double),
    ]


class Params(C.Structure):
    _fields_ = [
        ("width", C.c_uint32),
        ("height", C.c_uint32),
        ("exposure_value", C.c_double),
        ("filter_radius", C.c_uint32),
        ("max_bounces", C.c_uint32),
        ("num_samples", C.c_uint32),
        ("_pad", C.c_uint32),
        ("gather_size", C.c_uint64),
        ("gather_size_volume", C.c_uint64),
        ("watts", C.c_double),
        ("stepsize", C.c_double),
        ("seed", C.c_uint64),
        ("shard_count", C.c_uint32),
        ("shard_index", C.c_uint32),
    ]


class Stats(C.Structure):
    _fields_ = [
        ("path_samples", C.c_uint64),
        ("extend_rays", C.c_uint64),
        ("shadow_rays", C.c_uint64),
        ("shadow_traced", C.c_uint64),
        ("shadow_occlusion", C.c_uint64),
        ("waves", C.c_uint64),
        ("kernel_launches", C.c_uint64),
        ("device_ms", C.c_double),
        ("trace_ms", C.c_double),
        ("extend_ms", C.c_double),
        ("shade_ms", C.c_double),
        ("connect_ms", C.c_double),
        ("shadow_ms", C.c_double),
        ("photons_emitted", C.c_uint64),
        ("photons_surface", C.c_uint64),
        ("photons_volume", C.c_uint64),
        ("gather_queries", C.c_uint64),
        ("gather_visibility", C.c_uint64),
        ("gather_candidates", C.c_uint64),
        ("path_mode", C.c_uint64),
    ]


P = C.c_void_p
DP = C.POINTER(C.c_double)
U32P = C.POINTER(C.c_uint32)
U8P = C.POINTER(C.c_uint8)
U64P = C.POINTER(C.c_uint64)
PHOTON_ROW_F64 = 13  # RPTX_PHOTON_ROW_BYTES / 8
COMM_ID_BYTES = 128
INTERSECT_STATIC_MAX_RAYS = 65536  # RPTX_INTERSECT_STATIC_MAX_RAYS

# name -> (restype, argtypes): every symbol include/rptx.h declares
PROTOTYPES = {
    "rptx_abi_version": (C.c_int, []),
    "rptx_ctx_create": (C.c_int, [C.c_int, C.POINTER(P)]),
    "rptx_ctx_destroy": (None, [P]),
    "rptx_last_error": (C.c_char_p, [P]),
    "rptx_ctx_trim": (C.c_int, [P]),
    "rptx_comm_unique_id": (C.c_int, [P]),
    "rptx_comm_init": (C.c_int, [P, P, C.c_int, C.c_int]),
    "rptx_comm_info": (C.c_int, [P, C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "rptx_comm_destroy": (None, [P]),
    "rptx_scene_create": (C.c_int, [P, C.POINTER(P)]),
    "rptx_scene_destroy": (None, [P]),
    "rptx_scene_add_mesh": (C.c_int, [P, DP, DP, C.c_size_t, U32P]),
    "rptx_scene_add_object": (C.c_int, [P, C.POINTER(ShapeDesc), C.POINTER(MaterialDesc), U32P]),
    "rptx_scene_add_light": (C.c_int, [P, C.POINTER(LightDesc)]),
    "rptx_scene_set_medium": (C.c_int, [P, C.c_int32, C.c_double, C.c_double]),
    "rptx_scene_set_environment_color": (C.c_int, [P, DP]),
    "rptx_scene_set_environment_hdri": (C.c_int, [P, C.c_uint32, C.c_uint32, DP]),
    "rptx_scene_commit": (C.c_int, [P]),
    "rptx_render_begin": (C.c_int, [P, P, C.POINTER(CameraDesc), C.POINTER(Params), C.POINTER(P)]),
    "rptx_film_destroy": (None, [P]),
    "rptx_render_sample": (C.c_int, [P, C.c_uint32]),
    "rptx_render_sample_trace": (C.c_int, [P, C.c_uint32]),
    "rptx_render_sample_commit": (C.c_int, [P, C.c_uint32]),
    "rptx_render_sample_allreduce": (C.c_int, [P]),
    "rptx_film_batch_sum_device_ptr": (C.c_void_p, [P]),
    "rptx_film_image_rgb8": (C.c_int, [P, U8P]),
    "rptx_film_variance": (C.c_int, [P, DP]),
    "rptx_film_mean": (C.c_int, [P, DP]),
    "rptx_photons_emit": (C.c_int, [P, P, C.POINTER(Params), C.c_uint64, C.c_uint64, C.c_uint64, C.c_int32, C.POINTER(P)]),
    "rptx_photons_create": (C.c_int, [P, C.POINTER(P)]),
    "rptx_photons_destroy": (None, [P]),
    "rptx_photons_append_host": (C.c_int, [P, C.c_int32, P, C.c_size_t]),
    "rptx_photons_append_device": (C.c_int, [P, C.c_int32, P, C.c_size_t]),
    "rptx_photons_allgather": (C.c_int, [P, C.POINTER(P)]),
    "rptx_photons_count": (C.c_int, [P, U64P]),
    "rptx_photons_device_ptr": (C.c_void_p, [P, C.c_int32]),
    "rptx_photons_download": (C.c_int, [P, C.c_int32, P]),
    "rptx_photons_stats": (C.c_int, [P, U64P, U64P, DP]),
    "rptx_photon_map_build": (C.c_int, [P, P, P, C.c_int32, C.POINTER(P)]),
    "rptx_photon_map_build_sharded": (C.c_int, [P, P, P, C.c_int32, C.POINTER(P)]),
    "rptx_photon_map_destroy": (None, [P]),
    "rptx_photon_map_stats": (C.c_int, [P, U64P, DP]),
    "rptx_photon_gather_begin": (C.c_int, [P, P, C.POINTER(CameraDesc), C.POINTER(Params), C.POINTER(P)]),
    "rptx_photon_gather_trace": (C.c_int, [P, P, C.c_uint32]),
    "rptx_photon_render": (C.c_int, [P, P, C.POINTER(CameraDesc), C.POINTER(Params), C.c_uint64, C.c_int32, U8P, DP]),
    "rptx_photon_map_knn": (C.c_int, [P, C.c_int32, DP, C.c_size_t, C.c_uint32, U64P, DP, U32P]),
    "rptx_photon_map_spheres": (C.c_int, [P, DP, DP]),
    "rptx_intersect_batch": (C.c_int, [P, DP, DP, C.c_size_t, C.c_double, DP, U32P, U32P, U32P, DP]),
    "rptx_intersect_batch_device": (C.c_int, [P, P, P, C.c_size_t, C.c_double, P, P, P, P, P]),
    "rptx_stats_get": (C.c_int, [P, C.POINTER(Stats)]),
    "rptx_intersect_batch_static": (C.c_int, [P, DP, DP, C.c_size_t, C.c_double, DP, U32P, U32P, U32P, DP, U64P]),
}

_lib = None


def bind(lib, prototypes=PROTOTYPES, prefix_from="rptx_", prefix_to="rptx_"):
    """Attach restype/argtypes; raises AttributeError if a declared symbol is missing."""
    for name, (res, args) in prototypes.items():
        fn = getattr(lib, name.replace(prefix_from, prefix_to, 1))
        fn.restype = res
        fn.argtypes = args
    return lib


def load():
    """Load librptx.so (once).  No fallback: a missing library is an error."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "or `make -C rpt_b200/csrc`. rpt_b200 has no CPU fallback."
            )
        _lib = bind(C.CDLL(LIB_PATH))
        if _lib.rptx_abi_version() != 3:
            raise ImportError("librptx.so ABI version mismatch")
    return _lib
