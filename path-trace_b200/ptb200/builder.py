"""Scene construction through the flat builder C API.

`Builder(lib, prefix)` drives either this repository's C++ scene classes (libptb200.so, prefix
"ptbh_") or — in tests and the CPU-baseline leg of bench.py only — the unmodified reference's
classes (oracle/_ref/libptref.so, prefix "ref_").  Both are compiled from the same
scene_capi.inc, so one Python scene function builds the same scene on both sides.
"""
import ctypes as C
import math

import numpy as np

from . import capi

_F12 = C.c_float * 12
_F3 = C.c_float * 3


def _m(values):
    return _F12(*[float(v) for v in values])


class Builder:
    def __init__(self, lib, prefix):
        self.lib, self.prefix = lib, prefix
        self._keep = []  # numpy buffers handed to C that must outlive the call

    def _fn(self, name):
        return getattr(self.lib, self.prefix + name)

    def error(self):
        return self._fn("last_error")().decode()

    # ---- matrices (12 floats, reference Matrix member order) ---------------------------------
    def rotate(self, axis, angle):
        out = _F12()
        self._fn("matrix_rotate")(axis[0], axis[1], axis[2], float(angle), out)
        return list(out)

    def rotate_x(self, angle):
        return self.rotate((1, 0, 0), angle)

    def translate(self, x, y, z):
        out = _F12()
        self._fn("matrix_translate")(x, y, z, out)
        return list(out)

    def scale(self, x, y=None, z=None):
        y = x if y is None else y
        z = x if z is None else z
        out = _F12()
        self._fn("matrix_scale")(x, y, z, out)
        return list(out)

    def concat(self, a, b):
        out = _F12()
        self._fn("matrix_concat")(_m(a), _m(b), out)
        return list(out)

    def inverse(self, a):
        out = _F12()
        if self._fn("matrix_inverse")(_m(a), out) != 0:
            raise ArithmeticError(self.error())
        return list(out)

    def determinant(self, a):
        return self._fn("matrix_determinant")(_m(a))

    def apply(self, a, v, translate=True):
        out = _F3()
        self._fn("matrix_apply")(_m(a), _F3(*v), 1 if translate else 0, out)
        return list(out)

    # ---- images / textures ---------------------------------------------------------------------
    def image(self, rgba):
        rgba = np.ascontiguousarray(rgba, dtype=np.float32)
        assert rgba.ndim == 3 and rgba.shape[2] == 4
        h = self._fn("image_from_rgba")(rgba.ctypes.data_as(C.POINTER(C.c_float)), rgba.shape[1], rgba.shape[0])
        if not h:
            raise RuntimeError(self.error())
        return h

    def image_load(self, path):
        h = self._fn("image_load")(path.encode())
        if not h:
            raise RuntimeError(self.error())
        return h

    def image_pixels(self, img):
        w, h = self._fn("image_width")(img), self._fn("image_height")(img)
        out = np.zeros((h, w, 4), np.float32)
        self._fn("image_get_rgba")(img, out.ctypes.data_as(C.POINTER(C.c_float)))
        return out

    def write_hdr(self, rgba, path):
        rgba = np.ascontiguousarray(rgba, dtype=np.float32)
        if self._fn("write_hdr")(rgba.ctypes.data_as(C.POINTER(C.c_float)), rgba.shape[1], rgba.shape[0], path.encode()) != 0:
            raise RuntimeError(self.error())

    def color(self, r, g=None, b=None):
        g = r if g is None else g
        b = r if b is None else b
        return self._fn("tex_color")(r, g, b)

    def tex_image(self, img):
        return self._fn("tex_image")(img)

    def tex_image_alpha(self, img):
        return self._fn("tex_image_alpha")(img)

    def tex_skybox(self, faces, alpha=False):
        return self._fn("tex_skybox_alpha" if alpha else "tex_skybox")(*faces)

    def tex_transformed(self, m, t):
        return self._fn("tex_transformed")(_m(m), t)

    def tex_multiply(self, rgb, t):
        return self._fn("tex_multiply")(rgb[0], rgb[1], rgb[2], t)

    def tex_log(self, t):
        return self._fn("tex_log")(t)

    def tex_mirrorball(self, t):
        return self._fn("tex_mirrorball")(t)

    def tex_spherical(self, t):
        return self._fn("tex_spherical")(t)

    def tex_transform(self, m, t):
        return self._fn("tex_transform")(_m(m), t)

    def tex_eval(self, tex, positions):
        """Texture::getColor / getFloat through the C++ API of this side."""
        pos = np.ascontiguousarray(positions, dtype=np.float32).reshape(-1, 3)
        rgb = np.zeros((len(pos), 3), np.float32)
        f = np.zeros(len(pos), np.float32)
        if self._fn("tex_eval")(tex, pos.ctypes.data, len(pos), rgb.ctypes.data, f.ctypes.data) != 0:
            raise RuntimeError(self.error())
        return rgb, f

    # ---- materials (defaults as reference include/material.h:18) -------------------------------
    def material(self, reflect=None, scatter=None, emissive=None, transmit=None, ior=1.0, trc=None):
        reflect = self.color(1) if reflect is None else reflect
        scatter = self.color(1) if scatter is None else scatter
        emissive = self.color(0) if emissive is None else emissive
        transmit = self.color(0) if transmit is None else transmit
        trc = self.color(0) if trc is None else trc
        return self._fn("material")(reflect, scatter, emissive, transmit, ior, trc)

    def material_transform(self, m, mat):
        return self._fn("material_transform")(_m(m), mat)

    def material_index(self, mat):
        return self._fn("material_index")(mat)

    def materials_reset(self):
        self._fn("materials_reset")()

    # ---- objects ---------------------------------------------------------------------------------
    def sphere(self, c, r, mat):
        return self._fn("obj_sphere")(c[0], c[1], c[2], r, mat)

    def plane(self, n, d, mat):
        return self._fn("obj_plane")(n[0], n[1], n[2], d, mat)

    def plane_at(self, n, pos, mat):
        return self._fn("obj_plane_at")(n[0], n[1], n[2], pos[0], pos[1], pos[2], mat)

    def union(self, a, b):
        return self._fn("obj_union")(a, b)

    def difference(self, a, b):
        return self._fn("obj_difference")(a, b)

    def intersection(self, a, b):
        return self._fn("obj_intersection")(a, b)

    def transformed(self, m, o):
        """new TransformedObject(m, o): m maps the world ray into o's space (quirk Q7)."""
        return self._fn("obj_transformed")(_m(m), o)

    def transform(self, m, o):
        """PathTrace::transform(m, o): new object, o is not consumed (quirk Q6 for combinators)."""
        return self._fn("obj_transform")(_m(m), o)

    def duplicate(self, o):
        return self._fn("obj_duplicate")(o)

    def free(self, o):
        self._fn("obj_free")(o)

    def union_array(self, objs):
        """Balanced union tree, the shape of unionArray (reference src/test.cpp:52-64)."""
        n = len(objs)
        if n == 1:
            return objs[0]
        if n == 2:
            return self.union(objs[0], objs[1])
        split = n // 2
        return self.union(self.union_array(objs[:split]), self.union_array(objs[split:]))

    # ---- per-ray span dump through this side's Object::makeSpanIterator() ------------------------
    def trace_spans(self, world, origins, dirs, max_spans=16):
        o = np.ascontiguousarray(origins, dtype=np.float32).reshape(-1, 3)
        d = np.ascontiguousarray(dirs, dtype=np.float32).reshape(-1, 3)
        spans = np.zeros((len(o), max_spans), capi.SPAN_DTYPE)
        counts = np.zeros(len(o), np.int32)
        if self._fn("trace_spans")(world, o.ctypes.data, d.ctypes.data, len(o), spans.ctypes.data, max_spans, counts.ctypes.data) != 0:
            raise RuntimeError(self.error())
        return spans, counts

    # ---- vector helpers -----------------------------------------------------------------------------
    def reflect(self, v, n):
        out = _F3()
        self._fn("vec_reflect")(_F3(*v), _F3(*n), out)
        return list(out)

    def refract(self, v, ior, n):
        out = _F3()
        self._fn("vec_refract")(_F3(*v), ior, _F3(*n), out)
        return list(out)

    def refract_strength(self, v, ior, n):
        return self._fn("vec_refract_strength")(_F3(*v), ior, _F3(*n))


def product_builder():
    return Builder(capi.product(), "ptbh_")
