"""ctypes binding of include/b200rt.h (struct layouts and prototypes).

The library is built in-tree (rust_raytracer_b200/libb200rt.so, see build.py) and loaded from there; a missing
library is a hard error — there is no Python or CPU fallback for any compute entry point.
"""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("B200RT_LIB") or os.path.join(HERE, "libb200rt.so")  # B200RT_LIB: A/B builds of the same library



def _name_bundled_nccl():
    """Point the library's dlopen of NCCL (b200rt_render_multi's ncclReduce path) at the copy bundled with this interpreter's
    torch wheel, when there is one and the host has not chosen (B200RT_NCCL_LIB).  The loader keeps one object per SONAME: a
    process that maps the SYSTEM libnccl.so.2 first and imports torch later breaks that import if torch was built against a
    newer NCCL.  Nothing is imported or loaded here; a path is named."""
    if os.environ.get("B200RT_NCCL_LIB"):
        return
    try:
        import importlib.util
        spec = importlib.util.find_spec("nvidia.nccl")
        for d in (spec.submodule_search_locations if spec else []):
            cand = os.path.join(d, "lib", "libnccl.so.2")
            if os.path.exists(cand):
                os.environ["B200RT_NCCL_LIB"] = cand
                return
    except Exception:
        pass


_name_bundled_nccl()

OK, ERR_INVALID, ERR_CUDA, ERR_NOMEM, ERR_UNSUPPORTED, ERR_IO, ERR_PARSE = range(7)

PRIM_SPHERE, PRIM_MOVING_SPHERE, PRIM_XY_RECT, PRIM_XZ_RECT, PRIM_YZ_RECT, PRIM_BOX, PRIM_TRIANGLE, PRIM_MEDIUM_SPHERE, PRIM_MEDIUM_BOX = range(9)
MAT_LAMBERTIAN, MAT_METAL, MAT_DIELECTRIC, MAT_DIFFUSE_LIGHT, MAT_ISOTROPIC = range(5)
TEX_CONSTANT, TEX_CHECKER, TEX_IMAGE, TEX_NOISE = range(4)
INTEGRATOR_PATH, INTEGRATOR_MIXTURE, INTEGRATOR_NEE = range(3)
PIPELINE_WAVEFRONT, PIPELINE_MEGAKERNEL = range(2)
BACKGROUND_CONSTANT, BACKGROUND_SKY = range(2)
SCENE_FLAT_LIST = 1
FLAG_LIGHT_SCATTERS = 1
EVAL_TEXTURE, EVAL_SCATTER, EVAL_CAMERA, EVAL_SAMPLERS = range(4)


class Prim(C.Structure):
    _fields_ = [("type", C.c_int32), ("material", C.c_int32), ("instance", C.c_int32), ("aux", C.c_int32), ("p", C.c_float * 12)]


class Material(C.Structure):
    _fields_ = [("type", C.c_int32), ("texture", C.c_int32), ("param", C.c_float), ("pad", C.c_float)]


class Texture(C.Structure):
    _fields_ = [("type", C.c_int32), ("image", C.c_int32), ("scale", C.c_float), ("pad", C.c_float), ("c0", C.c_float * 4), ("c1", C.c_float * 4)]


class Image(C.Structure):
    _fields_ = [("offset", C.c_int32), ("width", C.c_int32), ("height", C.c_int32), ("pad", C.c_int32)]


class Instance(C.Structure):
    _fields_ = [("m", C.c_double * 12)]


class Perlin(C.Structure):
    _fields_ = [("ranvec", (C.c_float * 4) * 256), ("perm_x", C.c_int32 * 256), ("perm_y", C.c_int32 * 256), ("perm_z", C.c_int32 * 256)]


class SceneDesc(C.Structure):
    _fields_ = [
        ("n_prims", C.c_uint32), ("n_tri_normals", C.c_uint32), ("n_materials", C.c_uint32), ("n_textures", C.c_uint32),
        ("n_images", C.c_uint32), ("n_instances", C.c_uint32), ("n_lights", C.c_uint32), ("flags", C.c_uint32),
        ("image_bytes", C.c_uint64),
        ("prims", C.POINTER(Prim)), ("tri_normals", C.POINTER(C.c_float)), ("materials", C.POINTER(Material)),
        ("textures", C.POINTER(Texture)), ("images", C.POINTER(Image)), ("image_rgb8", C.POINTER(C.c_uint8)),
        ("instances", C.POINTER(Instance)), ("light_prims", C.POINTER(C.c_int32)), ("perlin", C.POINTER(Perlin)),
        ("light_scales", C.POINTER(C.c_float)),
    ]


class Camera(C.Structure):
    _fields_ = [("origin", C.c_float * 3), ("lower_left_corner", C.c_float * 3), ("horizontal", C.c_float * 3), ("vertical", C.c_float * 3),
                ("u", C.c_float * 3), ("v", C.c_float * 3), ("w", C.c_float * 3), ("lens_radius", C.c_float), ("time0", C.c_float), ("time1", C.c_float)]


class Params(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("spp_begin", C.c_int32), ("spp_count", C.c_int32), ("max_depth", C.c_int32),
                ("integrator", C.c_int32), ("pipeline", C.c_int32), ("background_mode", C.c_int32), ("t_min", C.c_float),
                ("background", C.c_float * 3), ("flags", C.c_uint32), ("pad", C.c_uint32), ("seed", C.c_uint64)]


class Stats(C.Structure):
    _fields_ = [("paths", C.c_uint64), ("extend_rays", C.c_uint64), ("shadow_rays", C.c_uint64), ("kernel_launches", C.c_uint64),
                ("ms_total", C.c_float), ("ms_generate", C.c_float), ("ms_extend", C.c_float), ("ms_shade", C.c_float), ("ms_shadow", C.c_float),
                ("ms_other", C.c_float), ("wavefront_iterations", C.c_uint32), ("pipelines", C.c_uint32),
                ("ms_generate_sum", C.c_float), ("ms_extend_sum", C.c_float), ("ms_shade_sum", C.c_float), ("ms_shadow_sum", C.c_float)]


class SppmParams(C.Structure):
    _fields_ = [("iterations", C.c_int32), ("photons_per_iter", C.c_int32), ("k_global", C.c_int32), ("k_caustic", C.c_int32),
                ("alpha", C.c_double), ("light_scale", C.c_double)]


class SppmStats(C.Structure):
    _fields_ = [("global_photons", C.c_uint64), ("caustic_photons", C.c_uint64), ("photon_rays", C.c_uint64), ("eye_rays", C.c_uint64),
                ("ms_total", C.c_float), ("ms_photons", C.c_float), ("ms_grid", C.c_float), ("ms_eye", C.c_float), ("ms_final", C.c_float),
                ("pad", C.c_uint32)]


class MultiStats(C.Structure):
    _fields_ = [("paths", C.c_uint64), ("extend_rays", C.c_uint64), ("shadow_rays", C.c_uint64), ("ms_render_max", C.c_float),
                ("ms_reduce", C.c_float), ("ms_wall", C.c_float), ("reduce_path", C.c_int32), ("n_dev", C.c_int32), ("pad", C.c_uint32)]


REDUCE_NONE, REDUCE_NCCL, REDUCE_PEER = 0, 1, 2


class Hit(C.Structure):
    _fields_ = [("t", C.c_float), ("prim", C.c_int32), ("nx", C.c_float), ("ny", C.c_float), ("nz", C.c_float), ("front_face", C.c_int32),
                ("u", C.c_float), ("v", C.c_float)]


VP = C.c_void_p
D3 = C.POINTER(C.c_double)
# name -> (restype, argtypes): every symbol include/b200rt.h declares
PROTOTYPES = {
    "b200rt_abi_version": (C.c_int, []),
    "b200rt_device_count": (C.c_int, [C.POINTER(C.c_int)]),
    "b200rt_ctx_create": (C.c_int, [C.c_int, C.POINTER(VP)]),
    "b200rt_ctx_destroy": (C.c_int, [VP]),
    "b200rt_ctx_set_stream": (C.c_int, [VP, VP]),
    "b200rt_last_error": (C.c_char_p, [VP]),
    "b200rt_global_error": (C.c_char_p, []),
    "b200rt_ctx_device_info": (C.c_int, [VP, C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_int64), C.c_char_p, C.c_int]),
    "b200rt_scene_create": (C.c_int, [VP, C.POINTER(SceneDesc), C.POINTER(VP)]),
    "b200rt_scene_destroy": (C.c_int, [VP]),
    "b200rt_scene_info": (C.c_int, [VP, C.POINTER(C.c_uint64), C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]),
    "b200rt_scene_bvh_quality": (C.c_int, [C.POINTER(SceneDesc), C.POINTER(C.c_float), C.POINTER(C.c_float), C.POINTER(C.c_int32), C.POINTER(C.c_float), C.POINTER(C.c_float)]),
    "b200rt_scene_compile_stats": (C.c_int, [C.POINTER(SceneDesc), C.POINTER(C.c_uint32), C.POINTER(C.c_uint32), C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]),
    "b200rt_debug_eval": (C.c_int, [VP, C.POINTER(Camera), C.POINTER(Params), C.c_int32, C.c_uint64, VP, VP]),
    "b200rt_intersect_batch": (C.c_int, [VP, VP, VP, VP, C.c_uint64, C.c_float, VP]),
    "b200rt_intersect_batch_device": (C.c_int, [VP, VP, VP, VP, C.c_uint64, C.c_float, VP]),
    "b200rt_render_device": (C.c_int, [VP, C.POINTER(Camera), C.POINTER(Params), VP, C.POINTER(Stats)]),
    "b200rt_render_device_static": (C.c_int, [VP, C.POINTER(Camera), C.POINTER(Params), VP, C.POINTER(Stats)]),
    "b200rt_render": (C.c_int, [VP, C.POINTER(Camera), C.POINTER(Params), VP, C.POINTER(Stats)]),
    "b200rt_tonemap_rgb8": (C.c_int, [VP, VP, C.c_int32, C.c_int32, VP]),
    "b200rt_tonemap_rgb8_device": (C.c_int, [VP, VP, C.c_int32, C.c_int32, VP]),
    "b200rt_write_png": (C.c_int, [C.c_char_p, VP, C.c_int32, C.c_int32]),
    "b200rt_render_multi": (C.c_int, [C.c_int32, C.POINTER(VP), C.POINTER(Camera), C.POINTER(Params), VP, C.POINTER(Stats), C.POINTER(MultiStats)]),
    "b200rt_render_sppm": (C.c_int, [VP, C.POINTER(Camera), C.POINTER(Params), VP, VP, VP, VP]),
    "b200rt_render_sppm_multi": (C.c_int, [C.c_int32, C.POINTER(VP), C.POINTER(Camera), C.POINTER(Params), VP, VP, VP, VP]),
    "b200rt_camera_make": (C.c_int, [D3, D3, D3, C.c_double, C.c_double, C.c_double, C.c_double, C.c_double, C.c_double, C.POINTER(Camera)]),
    "b200rt_measure_fp32_peak": (C.c_int, [VP, C.POINTER(C.c_double), C.POINTER(C.c_double)]),
    "b200rt_builder_create": (C.c_int, [C.POINTER(VP)]),
    "b200rt_builder_destroy": (C.c_int, [VP]),
    "b200rt_builder_error": (C.c_char_p, [VP]),
    "b200rt_builder_texture_constant": (C.c_int, [VP, C.c_double, C.c_double, C.c_double]),
    "b200rt_builder_texture_checker": (C.c_int, [VP, D3, D3]),
    "b200rt_builder_texture_image": (C.c_int, [VP, VP, C.c_int32, C.c_int32]),
    "b200rt_builder_texture_noise": (C.c_int, [VP, C.c_double]),
    "b200rt_builder_material": (C.c_int, [VP, C.c_int32, C.c_int32, C.c_double]),
    "b200rt_builder_push_transform": (C.c_int, [VP, D3, D3, D3]),
    "b200rt_builder_pop_transform": (C.c_int, [VP]),
    "b200rt_builder_sphere": (C.c_int, [VP, D3, C.c_double, C.c_int32]),
    "b200rt_builder_moving_sphere": (C.c_int, [VP, D3, D3, C.c_double, C.c_double, C.c_double, C.c_int32]),
    "b200rt_builder_rect": (C.c_int, [VP, C.c_int32, C.c_double, C.c_double, C.c_double, C.c_double, C.c_double, C.c_int32]),
    "b200rt_builder_cube": (C.c_int, [VP, D3, D3, C.c_int32]),
    "b200rt_builder_triangle": (C.c_int, [VP, D3, D3, D3, D3, D3, D3, C.c_int32]),
    "b200rt_builder_mesh_obj": (C.c_int, [VP, C.c_char_p, C.c_int32, C.c_int32, C.c_int32]),
    "b200rt_builder_medium_sphere": (C.c_int, [VP, D3, C.c_double, C.c_double, C.c_int32]),
    "b200rt_builder_medium_cube": (C.c_int, [VP, D3, D3, C.c_double, C.c_int32]),
    "b200rt_builder_mark_light": (C.c_int, [VP, C.c_int32]),
    "b200rt_builder_mark_light_scaled": (C.c_int, [VP, C.c_int32, C.c_double]),
    "b200rt_builder_set_flags": (C.c_int, [VP, C.c_uint32]),
    "b200rt_builder_seed_perlin": (C.c_int, [VP, C.c_uint64]),
    "b200rt_builder_load_json": (C.c_int, [VP, C.c_char_p, C.c_uint64, C.POINTER(Camera)]),
    "b200rt_builder_load_yaml": (C.c_int, [VP, C.c_char_p, C.c_uint64, C.POINTER(Camera)]),
    "b200rt_builder_make_scene": (C.c_int, [VP, C.c_char_p, C.c_uint64, C.c_char_p, VP, C.c_int32, C.c_int32, C.c_double, C.POINTER(Camera), C.POINTER(Params)]),
    "b200rt_builder_finish": (C.c_int, [VP, C.POINTER(SceneDesc)]),
}

_lib = None


def lib():
    """The loaded libb200rt.so; raises if it has not been built (no fallback)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError("%s is missing: build it with `python rust_raytracer_b200/build.py` (there is no CPU fallback)" % LIB_PATH)
        handle = C.CDLL(LIB_PATH)
        for name, (res, args) in PROTOTYPES.items():
            fn = getattr(handle, name)  # AttributeError if the library does not export a declared symbol
            fn.restype = res
            fn.argtypes = args
        if handle.b200rt_abi_version() != 1:
            raise RuntimeError("libb200rt.so ABI version mismatch")
        _lib = handle
    return _lib


def d3(v):
    return (C.c_double * 3)(float(v[0]), float(v[1]), float(v[2]))
