"""ctypes view of the C ABI in include/ptb200.h and of the scene-builder C API.

This module only *binds*: every computation happens in libptb200.so (CUDA).  If the library is
missing the import of the product fails loudly — there is no Python or CPU fallback.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
PKG_ROOT = os.path.dirname(HERE)
REPO_ROOT = os.path.dirname(PKG_ROOT)
PRODUCT_LIB = os.path.join(PKG_ROOT, "libptb200.so")

PTB_FLAG_EXACT = 1
PTB_FLAG_BILINEAR = 2
PTB_FLAG_REJECTION = 4
PTB_FLAG_ORIGIN0 = 8
PTB_FLAG_SPECIALIZE = 16
PTB_FLAG_INTERPRET = 32
PTB_HIT_MISS, PTB_HIT_ENTER, PTB_HIT_EXIT = 0, 1, 2
PTB_COMM_ID_BYTES = 128


class PtbPrim(C.Structure):
    _fields_ = [("p", C.c_float * 4), ("type", C.c_int32), ("material", C.c_int32), ("chain", C.c_int32), ("reserved", C.c_int32)]


class PtbXform(C.Structure):
    _fields_ = [("m", C.c_float * 12), ("inv", C.c_float * 12)]


class PtbChain(C.Structure):
    _fields_ = [("first", C.c_int32), ("count", C.c_int32)]


class PtbTexOp(C.Structure):
    _fields_ = [("kind", C.c_int32), ("arg", C.c_int32)]


class PtbMaterial(C.Structure):
    _fields_ = [("tex", C.c_int32 * 5), ("ior", C.c_float), ("reserved", C.c_int32 * 2)]


class PtbImageDesc(C.Structure):
    _fields_ = [("rgba", C.POINTER(C.c_float)), ("width", C.c_int32), ("height", C.c_int32)]


class PtbSceneDesc(C.Structure):
    _fields_ = [
        ("abi_version", C.c_uint32),
        ("n_prims", C.c_int32), ("n_xforms", C.c_int32), ("n_chains", C.c_int32), ("n_chain_xf", C.c_int32), ("n_ops", C.c_int32),
        ("n_materials", C.c_int32), ("n_texops", C.c_int32), ("n_texdata", C.c_int32), ("n_images", C.c_int32),
        ("prims", C.POINTER(PtbPrim)), ("xforms", C.POINTER(PtbXform)), ("chains", C.POINTER(PtbChain)), ("chain_xf", C.POINTER(C.c_int32)),
        ("ops", C.POINTER(C.c_uint32)), ("materials", C.POINTER(PtbMaterial)), ("texops", C.POINTER(PtbTexOp)), ("texdata", C.POINTER(C.c_float)),
        ("images", C.POINTER(PtbImageDesc)),
    ]


class PtbRenderParams(C.Structure):
    _fields_ = [
        ("width", C.c_int32), ("height", C.c_int32),
        ("x0", C.c_int32), ("y0", C.c_int32), ("x1", C.c_int32), ("y1", C.c_int32),
        ("spp", C.c_int32), ("sample_offset", C.c_int32), ("sample_stride", C.c_int32), ("max_depth", C.c_int32),
        ("screen_width", C.c_float), ("screen_height", C.c_float), ("screen_distance", C.c_float),
        ("seed", C.c_uint64), ("flags", C.c_uint32), ("samples_per_pass", C.c_int32),
    ]


class PtbStats(C.Structure):
    _fields_ = [
        ("paths", C.c_uint64), ("rays", C.c_uint64), ("launches", C.c_uint64), ("queue_bytes", C.c_uint64),
        ("ms_total", C.c_float), ("ms_primary", C.c_float), ("ms_bounce", C.c_float), ("ms_resolve", C.c_float),
        ("rays_primary", C.c_uint64), ("rays_bounce", C.c_uint64), ("ms_trace_primary", C.c_float), ("ms_trace_bounce", C.c_float),
        ("rays_first_bounce", C.c_uint64), ("launches_elided", C.c_uint64), ("ms_shade_primary", C.c_float), ("ms_shade_bounce", C.c_float),
    ]

    def as_dict(self):
        return {name: getattr(self, name) for name, _ in self._fields_}


class PtbAdaptiveParams(C.Structure):
    _fields_ = [("block_size", C.c_int32), ("max_sample_size", C.c_int32), ("min_color_delta", C.c_float), ("reserved", C.c_int32)]


class PtbAdaptiveStats(C.Structure):
    _fields_ = [("pixels_traced", C.c_uint64), ("pixels_interpolated", C.c_uint64), ("squares_interpolated", C.c_uint64), ("levels", C.c_int32), ("reserved", C.c_int32)]


# numpy record layouts of PtbHit / PtbSpan (32 and 48 bytes)
HIT_DTYPE = np.dtype([("t", "<f4"), ("normal", "<f4", 3), ("kind", "<i4"), ("prim", "<i4"), ("material", "<i4"), ("n_spans", "<i4")])
SPAN_DTYPE = np.dtype([("start", "<f4"), ("end", "<f4"), ("start_normal", "<f4", 3), ("end_normal", "<f4", 3), ("start_prim", "<i4"), ("end_prim", "<i4"),
                       ("start_material", "<i4"), ("end_material", "<i4")])
assert HIT_DTYPE.itemsize == 32 and SPAN_DTYPE.itemsize == 48

# every symbol include/ptb200.h declares (checked by tests/test_abi.py)
ABI_SYMBOLS = [
    "ptb200_last_error", "ptb200_abi_version", "ptb200_device_count", "ptb200_trim", "ptb200_scene_create", "ptb200_scene_destroy", "ptb200_image_upload",
    "ptb200_intersect", "ptb200_intersect_device", "ptb200_texture_eval", "ptb200_render", "ptb200_render_adaptive", "ptb200_render_accumulate_device", "ptb200_resolve_device", "ptb200_display8_device", "ptb200_display8",
    "ptb200_render_async", "ptb200_job_done", "ptb200_job_read", "ptb200_job_wait", "ptb200_job_destroy", "ptb200_sample_scatter", "ptb200_fp32_peak", "ptb200_specialize_probe", "ptb200_scene_specialized", "ptb200_tile_request_parse", "ptb200_tile_encode", "ptb200_tile_decode",
    "ptb200_comm_unique_id", "ptb200_comm_init_rank", "ptb200_comm_init_all", "ptb200_comm_info", "ptb200_comm_destroy", "ptb200_scene_create_multi", "ptb200_render_multi",
    "ptb200_multi_rows", "ptb200_multi_shard", "ptb200_vector_eval", "ptb200_rng_words", "ptb200_render_pixels", "ptb200_intersect_device_static", "ptb200_static_scratch_bytes",
]

_vp, _f, _i, _sz = C.c_void_p, C.c_float, C.c_int, C.c_size_t
_fp = C.POINTER(C.c_float)

# scene-builder C API shared by libptb200.so (prefix ptbh_) and oracle/_ref/libptref.so (ref_)
BUILDER_SIGS = {
    "last_error": (C.c_char_p, []),
    "matrix_rotate": (None, [_f, _f, _f, C.c_double, _fp]),
    "matrix_translate": (None, [_f, _f, _f, _fp]),
    "matrix_scale": (None, [_f, _f, _f, _fp]),
    "matrix_concat": (None, [_fp, _fp, _fp]),
    "matrix_inverse": (_i, [_fp, _fp]),
    "matrix_determinant": (_f, [_fp]),
    "matrix_apply": (None, [_fp, _fp, _i, _fp]),
    "image_from_rgba": (_vp, [_fp, _i, _i]),
    "image_load": (_vp, [C.c_char_p]),
    "image_width": (_i, [_vp]),
    "image_height": (_i, [_vp]),
    "image_get_rgba": (None, [_vp, _fp]),
    "image_free": (None, [_vp]),
    "write_hdr": (_i, [_fp, _i, _i, C.c_char_p]),
    "tex_color": (_vp, [_f, _f, _f]),
    "tex_image": (_vp, [_vp]),
    "tex_image_alpha": (_vp, [_vp]),
    "tex_skybox": (_vp, [_vp] * 6),
    "tex_skybox_alpha": (_vp, [_vp] * 6),
    "tex_transformed": (_vp, [_fp, _vp]),
    "tex_multiply": (_vp, [_f, _f, _f, _vp]),
    "tex_log": (_vp, [_vp]),
    "tex_mirrorball": (_vp, [_vp]),
    "tex_spherical": (_vp, [_vp]),
    "tex_transform": (_vp, [_fp, _vp]),
    "tex_duplicate": (_vp, [_vp]),
    "tex_free": (None, [_vp]),
    "material": (_vp, [_vp, _vp, _vp, _vp, _f, _vp]),
    "material_transform": (_vp, [_fp, _vp]),
    "material_index": (_i, [_vp]),
    "material_free": (None, [_vp]),
    "materials_reset": (None, []),
    "obj_sphere": (_vp, [_f, _f, _f, _f, _vp]),
    "obj_plane": (_vp, [_f, _f, _f, _f, _vp]),
    "obj_plane_at": (_vp, [_f] * 6 + [_vp]),
    "obj_union": (_vp, [_vp, _vp]),
    "obj_difference": (_vp, [_vp, _vp]),
    "obj_intersection": (_vp, [_vp, _vp]),
    "obj_transformed": (_vp, [_fp, _vp]),
    "obj_transform": (_vp, [_fp, _vp]),
    "obj_duplicate": (_vp, [_vp]),
    "obj_free": (None, [_vp]),
    "vec_reflect": (None, [_fp, _fp, _fp]),
    "vec_refract": (None, [_fp, _f, _fp, _fp]),
    "vec_refract_strength": (_f, [_fp, _f, _fp]),
    "vec_normalize": (None, [_fp, _fp]),
    "trace_spans": (_i, [_vp, _vp, _vp, _sz, _vp, _i, _vp]),
    "tex_eval": (_i, [_vp, _vp, _sz, _vp, _vp]),
}


def bind_builder(lib, prefix):
    for name, (res, args) in BUILDER_SIGS.items():
        fn = getattr(lib, prefix + name)
        fn.restype, fn.argtypes = res, args
    return lib


class PtbError(RuntimeError):
    def __init__(self, status, message):
        super().__init__("ptb200 status %d: %s" % (status, message))
        self.status = status


_product = None


def product():
    """libptb200.so (CUDA kernels + C ABI + host C++ scene API).  Raises if it was not built."""
    global _product
    if _product is None:
        if not os.path.exists(PRODUCT_LIB):
            raise RuntimeError("libptb200.so is missing: run `python -c 'import __graft_entry__ as g; g.build()'` (there is no CPU fallback)")
        lib = C.CDLL(PRODUCT_LIB)
        lib.ptb200_last_error.restype = C.c_char_p
        lib.ptb200_scene_create.argtypes = [C.POINTER(PtbSceneDesc), C.POINTER(_vp)]
        lib.ptb200_scene_destroy.argtypes = [_vp]
        lib.ptb200_scene_destroy.restype = None
        lib.ptb200_image_upload.argtypes = [_vp, _vp, _i, _i, C.POINTER(_i)]
        lib.ptb200_intersect.argtypes = [_vp, _vp, _vp, _sz, _vp, _vp, _i, C.c_uint32]
        lib.ptb200_intersect_device.argtypes = [_vp, _vp, _vp, _sz, _vp, _vp, _i, C.c_uint32, _vp]
        lib.ptb200_texture_eval.argtypes = [_vp, _i, _vp, _sz, _vp, _vp, C.c_uint32]
        lib.ptb200_render.argtypes = [_vp, C.POINTER(PtbRenderParams), _vp, C.POINTER(PtbStats)]
        lib.ptb200_render_adaptive.argtypes = [_vp, C.POINTER(PtbRenderParams), C.POINTER(PtbAdaptiveParams), _vp, _vp, _vp, C.POINTER(PtbStats), C.POINTER(PtbAdaptiveStats)]
        lib.ptb200_render_accumulate_device.argtypes = [_vp, C.POINTER(PtbRenderParams), _vp, C.POINTER(PtbStats), _vp]
        lib.ptb200_resolve_device.argtypes = [_vp, _vp, _sz, _vp]
        lib.ptb200_display8_device.argtypes = [_vp, _vp, _sz, _i, _vp]
        lib.ptb200_display8.argtypes = [_vp, _sz, _i, _vp]
        lib.ptb200_render_async.argtypes = [_vp, C.POINTER(PtbRenderParams), C.POINTER(_vp)]
        lib.ptb200_job_done.argtypes = [_vp]
        lib.ptb200_job_read.argtypes = [_vp, _vp, _i, _i]
        lib.ptb200_job_wait.argtypes = [_vp, C.POINTER(PtbStats)]
        lib.ptb200_job_destroy.argtypes = [_vp]
        lib.ptb200_job_destroy.restype = None
        lib.ptb200_sample_scatter.argtypes = [_fp, _fp, _f, _sz, C.c_uint64, C.c_uint32, _vp]
        lib.ptb200_specialize_probe.argtypes = [C.POINTER(PtbSceneDesc), _i, C.c_char_p, _sz, C.POINTER(_sz), C.POINTER(_sz)]
        lib.ptb200_scene_specialized.argtypes = [_vp, C.POINTER(C.c_double), C.POINTER(_i), C.POINTER(_i)]
        lib.ptb200_tile_request_parse.argtypes = [C.c_char_p, _i, _i, C.POINTER(_i), C.POINTER(_i), C.POINTER(_i)]
        lib.ptb200_tile_encode.argtypes = [_vp, _i, _i, _i, _i, _i, _vp, _i, C.c_char_p, _sz, C.POINTER(_sz)]
        lib.ptb200_tile_decode.argtypes = [C.c_char_p, _sz, _i, _i, _i, _vp, _vp, C.POINTER(_i), C.POINTER(_sz)]
        lib.ptb200_fp32_peak.argtypes = [C.POINTER(C.c_double), C.POINTER(C.c_double)]
        _ip = C.POINTER(_i)
        lib.ptb200_intersect_device_static.argtypes = [_vp, _vp, _vp, _sz, _vp, _vp, _i, C.c_uint32, _vp]
        lib.ptb200_static_scratch_bytes.restype = _sz
        lib.ptb200_render_pixels.argtypes = [_vp, C.POINTER(PtbRenderParams), _vp, _sz, _vp, C.POINTER(PtbStats)]
        lib.ptb200_vector_eval.argtypes = [_i, _vp, _vp, _vp, _sz, _vp]
        lib.ptb200_rng_words.argtypes = [_i, C.c_uint64, _vp, _vp, _vp, _sz, _vp]
        lib.ptb200_comm_unique_id.argtypes = [_vp]
        lib.ptb200_comm_init_rank.argtypes = [C.c_char_p, _i, _i, C.POINTER(_vp)]
        lib.ptb200_comm_init_all.argtypes = [_ip, _i, C.POINTER(_vp)]
        lib.ptb200_comm_info.argtypes = [_vp, _ip, _ip, _ip]
        lib.ptb200_comm_destroy.argtypes = [_vp]
        lib.ptb200_comm_destroy.restype = None
        lib.ptb200_scene_create_multi.argtypes = [_vp, C.POINTER(PtbSceneDesc), C.POINTER(_vp)]
        lib.ptb200_render_multi.argtypes = [_vp, C.POINTER(_vp), C.POINTER(PtbRenderParams), _vp, C.POINTER(_vp), C.POINTER(PtbStats), _vp]
        lib.ptb200_multi_rows.argtypes = [_i, _i, _i, _ip, _ip]
        lib.ptb200_multi_shard.argtypes = [_i, _i, _i, _ip, _ip, _ip]
        bind_builder(lib, "ptbh_")
        lib.ptbh_flatten.restype = _vp
        lib.ptbh_flatten.argtypes = [_vp]
        lib.ptbh_flatten_texture.restype = _vp
        lib.ptbh_flatten_texture.argtypes = [_vp]
        lib.ptbh_flatten_desc.restype = None
        lib.ptbh_flatten_desc.argtypes = [_vp, C.POINTER(PtbSceneDesc)]
        lib.ptbh_flatten_free.restype = None
        lib.ptbh_flatten_free.argtypes = [_vp]
        lib.ptbh_render.argtypes = [_vp, C.POINTER(PtbRenderParams), _vp, C.POINTER(PtbStats)]
        _product = lib
    return _product


def check(status):
    if status != 0:
        raise PtbError(status, product().ptb200_last_error().decode())
