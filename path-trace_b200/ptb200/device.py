"""Product-side wrappers: flatten an Object graph, put it on the GPU, intersect, render.

Everything here calls libptb200.so through the C ABI; host buffers are numpy arrays, device
buffers (for the resident-input bench path and multi-GPU accumulation) are torch tensors whose
data_ptr() is passed as a plain pointer.
"""
import ctypes as C

import numpy as np

from . import capi
from .capi import PtbRenderParams, PtbSceneDesc, PtbStats, check


class Flat:
    """A flattened scene: owns the host arrays a PtbSceneDesc points into."""

    def __init__(self, world=None, texture=None):
        lib = capi.product()
        self.handle = lib.ptbh_flatten(world) if texture is None else lib.ptbh_flatten_texture(texture)
        if not self.handle:
            raise RuntimeError(lib.ptbh_last_error().decode())
        self.desc = PtbSceneDesc()
        lib.ptbh_flatten_desc(self.handle, C.byref(self.desc))

    def pin_images(self):
        """Moves the image pixels into page-locked host memory (torch) and points the description at the copies:
        ptb200_scene_create then DMAs them from where they lie instead of staging them through its own pinned
        buffer.  What a caller that re-creates scenes every frame would do."""
        import torch

        self._pinned = []
        for i in range(self.desc.n_images):
            im = self.desc.images[i]
            n = 4 * im.width * im.height
            t = torch.empty(n, dtype=torch.float32, pin_memory=True)
            C.memmove(t.data_ptr(), im.rgba, 4 * n)
            self._pinned.append(t)
            self.desc.images[i].rgba = C.cast(t.data_ptr(), type(im.rgba))
        return self

    def close(self):
        if self.handle:
            capi.product().ptbh_flatten_free(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # plain-python views, for tests of the flattener
    def ops(self):
        return [self.desc.ops[i] for i in range(self.desc.n_ops)]

    def prims(self):
        return [self.desc.prims[i] for i in range(self.desc.n_prims)]


def default_params(width, height, spp, depth=16, seed=0xB200, flags=0, samples_per_pass=0, sample_offset=0, sample_stride=1, rect=None):
    """Camera of RenderBlock::calcPixelColor (reference src/test.cpp:450)."""
    p = PtbRenderParams()
    p.width, p.height = width, height
    if rect is not None:
        p.x0, p.y0, p.x1, p.y1 = rect
    p.spp, p.sample_offset, p.sample_stride, p.max_depth = spp, sample_offset, sample_stride, depth
    p.screen_width, p.screen_height, p.screen_distance = float(width), float(height), float(2 * min(width, height))
    p.seed, p.flags, p.samples_per_pass = seed, flags, samples_per_pass
    return p


class DeviceScene:
    """A scene resident on the current CUDA device (ptb200_scene_create)."""

    def __init__(self, flat):
        self.flat = flat  # keeps the host arrays alive while create copies them
        self.handle = C.c_void_p()
        check(capi.product().ptb200_scene_create(C.byref(flat.desc), C.byref(self.handle)))

    def close(self):
        if self.handle:
            capi.product().ptb200_scene_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def intersect(self, origins, dirs, max_spans=0, flags=0):
        """Host-buffer ray batch through ptb200_intersect."""
        o = np.ascontiguousarray(origins, dtype=np.float32).reshape(-1, 3)
        d = np.ascontiguousarray(dirs, dtype=np.float32).reshape(-1, 3)
        n = len(o)
        hits = np.zeros(n, capi.HIT_DTYPE)
        spans = np.zeros((n, max_spans), capi.SPAN_DTYPE) if max_spans > 0 else None
        check(capi.product().ptb200_intersect(self.handle, o.ctypes.data, d.ctypes.data, n, hits.ctypes.data, spans.ctypes.data if spans is not None else None,
                                              max_spans, flags))
        return hits, spans

    def intersect_device(self, origins4, dirs4, hits, flags=0, stream=None):
        """Device-resident ray batch: torch tensors (n,4) float32 in, (n,8) int32/float32 out."""
        n = origins4.shape[0]
        check(capi.product().ptb200_intersect_device(self.handle, origins4.data_ptr(), dirs4.data_ptr(), n, hits.data_ptr(), None, 0, flags, stream))

    def texture_eval(self, root, positions, flags=0):
        pos = np.ascontiguousarray(positions, dtype=np.float32).reshape(-1, 3)
        rgb = np.zeros((len(pos), 3), np.float32)
        f = np.zeros(len(pos), np.float32)
        check(capi.product().ptb200_texture_eval(self.handle, root, pos.ctypes.data, len(pos), rgb.ctypes.data, f.ctypes.data, flags))
        return rgb, f

    def render(self, params, out=None):
        """Synchronous frame render into a host (H, W, 4) float32 array."""
        if out is None:
            out = np.zeros((params.height, params.width, 4), np.float32)
        stats = PtbStats()
        check(capi.product().ptb200_render(self.handle, C.byref(params), out.ctypes.data, C.byref(stats)))
        return out, stats

    def render_pixels(self, params, xy):
        """ptb200_render_pixels: the listed pixels (n x 2 ints: x, y) only -> (n, 4) float32, stats."""
        xy = np.ascontiguousarray(xy, dtype=np.int32).reshape(-1, 2)
        out = np.zeros((len(xy), 4), np.float32)
        stats = PtbStats()
        check(capi.product().ptb200_render_pixels(self.handle, C.byref(params), xy.ctypes.data, len(xy), out.ctypes.data, C.byref(stats)))
        return out, stats

    def render_adaptive(self, params, block_size=32, max_sample_size=None, min_color_delta=0.003, want_lattice=True):
        """ptb200_render_adaptive: the reference's pixel-skipping quadtree (src/test.cpp:423-499).
        Defaults are the reference's constants (src/test.cpp:38,50).  Returns image, source map
        (1 traced, 2 interpolated), traced lattice, PtbStats, PtbAdaptiveStats."""
        ap = capi.PtbAdaptiveParams()
        ap.block_size = block_size
        ap.max_sample_size = params.height // (480 // 4) if max_sample_size is None else max_sample_size
        ap.min_color_delta = min_color_delta
        out = np.zeros((params.height, params.width, 4), np.float32)
        source = np.zeros((params.height, params.width), np.uint8)
        bs = max(1, block_size)  # bad sizes are rejected by the library, not here
        lattice = np.zeros((-(-params.height // bs) * bs + 1, -(-params.width // bs) * bs + 1, 4), np.float32)
        stats, astats = PtbStats(), capi.PtbAdaptiveStats()
        check(capi.product().ptb200_render_adaptive(self.handle, C.byref(params), C.byref(ap), out.ctypes.data, source.ctypes.data, lattice.ctypes.data if want_lattice else None, C.byref(stats), C.byref(astats)))
        return out, source, lattice, stats, astats

    def specialized(self):
        """ptb200_scene_specialized: (kernels ready?, compile ms, passes of the last stats render that ran them, passes);
        self.spec_camera_only: the last render ran them for camera rays only (a large program); self.spec_pending: a
        compile a render call asked for is still under way."""
        ms, a, b = C.c_double(), C.c_int(), C.c_int()
        rc = capi.product().ptb200_scene_specialized(self.handle, C.byref(ms), C.byref(a), C.byref(b))
        if rc < 0:
            check(rc)
        self.spec_camera_only = rc == 2
        self.spec_pending = rc == 0 and ms.value < 0
        return rc >= 1, ms.value, a.value, b.value

    def render_accumulate(self, params, accum, stream=None, want_stats=True):
        """Adds this call's radiance sums into a device (H, W, 4) float32 torch tensor."""
        stats = PtbStats()
        check(capi.product().ptb200_render_accumulate_device(self.handle, C.byref(params), accum.data_ptr(), C.byref(stats) if want_stats else None, stream))
        return stats


def specialize_probe(flat, which=0, compile=True):
    """ptb200_specialize_probe (no GPU needed): the CUDA source generated for a flattened scene and, with
    compile, the size of the sm_100a cubin NVRTC makes of it."""
    lib = capi.product()
    n, cubin = C.c_size_t(), C.c_size_t()
    check(lib.ptb200_specialize_probe(C.byref(flat.desc), which, None, 0, C.byref(n), None))
    buf = C.create_string_buffer(n.value + 1)
    check(lib.ptb200_specialize_probe(C.byref(flat.desc), which, buf, len(buf), C.byref(n), C.byref(cubin) if compile else None))
    return buf.value.decode(), cubin.value


def resolve(accum, rgba, stream=None):
    check(capi.product().ptb200_resolve_device(accum.data_ptr(), rgba.data_ptr(), accum.numel() // 4, stream))


def sample_scatter(normal, mirror, scatter, n, seed=1, flags=0):
    out = np.zeros((n, 3), np.float32)
    f3 = C.c_float * 3
    check(capi.product().ptb200_sample_scatter(f3(*normal), f3(*mirror), scatter, n, seed, flags, out.ctypes.data))
    return out


def fp32_peak():
    fma, mhz = C.c_double(), C.c_double()
    check(capi.product().ptb200_fp32_peak(C.byref(fma), C.byref(mhz)))
    return fma.value, mhz.value
