"""Test-side loaders for the CHECKERS: oracle/libptoracle.so (our CPU restatement) and
oracle/_ref/libptref.so (the unmodified reference).  Only tests, smoke() and bench.py's CPU
baseline may use these."""
import ctypes as C
import os

import numpy as np

from ptb200 import capi
from ptb200.builder import Builder

ROOT = capi.REPO_ROOT
ORACLE_LIB = os.path.join(ROOT, "oracle", "libptoracle.so")
REF_LIB = os.path.join(ROOT, "oracle", "_ref", "libptref.so")
REF_FAST_LIB = os.path.join(ROOT, "oracle", "_ref", "libptref_fast.so")
GOLDEN = os.path.join(ROOT, "tests", "golden")

_vp, _f, _i, _sz, _u = C.c_void_p, C.c_float, C.c_int, C.c_size_t, C.c_uint
_ull = C.POINTER(C.c_ulonglong)
_oracle = None
_ref = {}


def oracle():
    global _oracle
    if _oracle is None:
        lib = C.CDLL(ORACLE_LIB)
        D = C.POINTER(capi.PtbSceneDesc)
        lib.orc_intersect.argtypes = [D, _vp, _vp, _sz, _vp, _vp, _i]
        lib.orc_trace_rays.argtypes = [D, _vp, _vp, _sz, _i, _u, _f, _vp]
        lib.orc_render.argtypes = [D, _i, _i, _i, _i, _i, _i, _i, _i, _i, _f, _f, _f, _u, _vp, _ull]
        lib.orc_render_mt.argtypes = [D, _i, _i, _i, _i, _i, _i, _i, _i, _i, _f, _f, _f, _u, _i, C.c_double, _vp, _ull, _ull, C.POINTER(C.c_double)]
        lib.orc_tex_eval.argtypes = [D, _i, _vp, _sz, _vp, _vp]
        lib.orc_rng_u32.argtypes = [_u, _sz, _vp]
        lib.orc_rng_uniform.argtypes = [_u, _sz, _f, _f, _vp]
        lib.orc_rng_ball.argtypes = [_u, _sz, _f, _f, _vp]
        lib.orc_vec_reflect.argtypes = [_vp, _vp, _vp]
        lib.orc_vec_refract.argtypes = [_vp, _f, _vp, _vp]
        lib.orc_vec_refract_strength.argtypes = [_vp, _f, _vp]
        lib.orc_vec_refract_strength.restype = _f
        _oracle = lib
    return _oracle


def have_ref():
    return os.path.exists(REF_LIB)


def ref(fast=False):
    """The unmodified reference behind the builder C API (prefix ref_)."""
    path = REF_FAST_LIB if fast else REF_LIB
    if path not in _ref:
        lib = C.CDLL(path)
        capi.bind_builder(lib, "ref_")
        lib.ref_trace_rays.argtypes = [_vp, _vp, _vp, _sz, _i, _u, _f, _vp]
        lib.ref_render.argtypes = [_vp, _i, _i, _i, _i, _i, _i, _i, _i, _i, _f, _f, _f, _u, _vp, _ull]
        lib.ref_render_mt.argtypes = [_vp, _i, _i, _i, _i, _i, _i, _i, _i, _i, _f, _f, _f, _u, _i, C.c_double, _vp, _ull, _ull, C.POINTER(C.c_double)]
        lib.ref_rng_u32.argtypes = [_u, _sz, _vp]
        lib.ref_rng_uniform.argtypes = [_u, _sz, _f, _f, _vp]
        lib.ref_rng_ball.argtypes = [_u, _sz, _f, _f, _vp]
        _ref[path] = lib
    return _ref[path]


def ref_builder(fast=False):
    return Builder(ref(fast), "ref_")


# ---- oracle wrappers over a flattened scene ----------------------------------------------------
def orc_intersect(flat, origins, dirs, max_spans=16):
    o = np.ascontiguousarray(origins, np.float32).reshape(-1, 3)
    d = np.ascontiguousarray(dirs, np.float32).reshape(-1, 3)
    hits = np.zeros(len(o), capi.HIT_DTYPE)
    spans = np.zeros((len(o), max_spans), capi.SPAN_DTYPE)
    assert oracle().orc_intersect(C.byref(flat.desc), o.ctypes.data, d.ctypes.data, len(o), hits.ctypes.data, spans.ctypes.data, max_spans) == 0
    return hits, spans


def orc_trace_rays(flat, origins, dirs, depth, seed, strength=1.0):
    o = np.ascontiguousarray(origins, np.float32).reshape(-1, 3)
    d = np.ascontiguousarray(dirs, np.float32).reshape(-1, 3)
    rgb = np.zeros((len(o), 3), np.float32)
    assert oracle().orc_trace_rays(C.byref(flat.desc), o.ctypes.data, d.ctypes.data, len(o), depth, seed, strength, rgb.ctypes.data) == 0
    return rgb


def lattice_args(W, H, x0, y0, step, nx, ny, spp, depth):
    return (W, H, x0, y0, step, nx, ny, spp, depth, float(W), float(H), float(2 * min(W, H)))


def orc_render(flat, W, H, x0, y0, step, nx, ny, spp, depth, seed):
    rgb = np.zeros((ny, nx, 3), np.float32)
    rays = C.c_ulonglong()
    assert oracle().orc_render(C.byref(flat.desc), *lattice_args(W, H, x0, y0, step, nx, ny, spp, depth), seed, rgb.ctypes.data, C.byref(rays)) == 0
    return rgb, rays.value


def orc_render_mt(flat, W, H, x0, y0, step, nx, ny, spp, depth, seed, threads, max_seconds=0.0):
    rgb = np.zeros((ny, nx, 3), np.float32)
    rays, pixels, secs = C.c_ulonglong(), C.c_ulonglong(), C.c_double()
    rows = oracle().orc_render_mt(C.byref(flat.desc), *lattice_args(W, H, x0, y0, step, nx, ny, spp, depth), seed, threads, max_seconds, rgb.ctypes.data,
                                  C.byref(rays), C.byref(pixels), C.byref(secs))
    assert rows >= 0
    return rgb, dict(rows=rows, rays=rays.value, pixels=pixels.value, seconds=secs.value)


def orc_tex_eval(flat, root, positions):
    pos = np.ascontiguousarray(positions, np.float32).reshape(-1, 3)
    rgb = np.zeros((len(pos), 3), np.float32)
    f = np.zeros(len(pos), np.float32)
    assert oracle().orc_tex_eval(C.byref(flat.desc), root, pos.ctypes.data, len(pos), rgb.ctypes.data, f.ctypes.data) == 0
    return rgb, f


# ---- reference wrappers over reference objects ----------------------------------------------------
def ref_trace_rays(lib, world, origins, dirs, depth, seed, strength=1.0):
    o = np.ascontiguousarray(origins, np.float32).reshape(-1, 3)
    d = np.ascontiguousarray(dirs, np.float32).reshape(-1, 3)
    rgb = np.zeros((len(o), 3), np.float32)
    assert lib.ref_trace_rays(world, o.ctypes.data, d.ctypes.data, len(o), depth, seed, strength, rgb.ctypes.data) == 0
    return rgb


def ref_render(lib, world, W, H, x0, y0, step, nx, ny, spp, depth, seed):
    rgb = np.zeros((ny, nx, 3), np.float32)
    rays = C.c_ulonglong()
    assert lib.ref_render(world, *lattice_args(W, H, x0, y0, step, nx, ny, spp, depth), seed, rgb.ctypes.data, C.byref(rays)) == 0
    return rgb, rays.value


def ref_render_mt(lib, world, W, H, x0, y0, step, nx, ny, spp, depth, seed, threads, max_seconds=0.0):
    rgb = np.zeros((ny, nx, 3), np.float32)
    rays, pixels, secs = C.c_ulonglong(), C.c_ulonglong(), C.c_double()
    rows = lib.ref_render_mt(world, *lattice_args(W, H, x0, y0, step, nx, ny, spp, depth), seed, threads, max_seconds, rgb.ctypes.data, C.byref(rays),
                             C.byref(pixels), C.byref(secs))
    assert rows >= 0
    return rgb, dict(rows=rows, rays=rays.value, pixels=pixels.value, seconds=secs.value)
