"""Sample-sharded multi-GPU rendering — Python plumbing around the C ABI.

The work itself is in libptb200.so (csrc/ptb_multi.cu): ptb200_comm_init_rank / _init_all,
ptb200_scene_create_multi (images cross PCIe once, NCCL broadcast over NVLink), ptb200_render_multi (every
rank renders its sample shard, ONE ncclReduceScatter of the accumulation frame by row blocks, every rank
resolves and delivers its own rows).  Rank g of G renders the global camera-sample indices g, g+G, ... of every
pixel; the RNG of the render kernels is keyed by the GLOBAL sample index, so the union over ranks is exactly the
single-GPU sample set and the G-GPU image equals the 1-GPU image up to float summation order.

What lives here: shipping the 128-byte NCCL id between processes with torch.distributed (any backend: the
id is just bytes), and thin wrappers.  The partition rules (which samples, which rows) are the library's own
pure functions ptb200_multi_shard / ptb200_multi_rows, so tests/test_multi_gloo.py can run the same
decomposition on CPU with the gloo backend and world size 2.
"""
import ctypes as C

from . import capi


def shard_samples(spp, rank, world):
    """(count, offset, stride) of the global sample indices rank `rank` renders: offset + k*stride,
    k < count (ptb200_multi_shard).  The shards partition range(spp) for any spp >= 0 (ragged when world
    does not divide spp: the first spp % world ranks take one more)."""
    count, offset, stride = C.c_int(), C.c_int(), C.c_int()
    if capi.product().ptb200_multi_shard(spp, world, rank, C.byref(count), C.byref(offset), C.byref(stride)) != 0:
        raise ValueError("bad spp/rank/world: %r/%r/%r" % (spp, rank, world))
    return count.value, offset.value, stride.value


def rows_of(height, rank, world):
    """[row0, row1) of the frame that rank `rank` resolves and delivers (ptb200_multi_rows)."""
    r0, r1 = C.c_int(), C.c_int()
    if capi.product().ptb200_multi_rows(height, world, rank, C.byref(r0), C.byref(r1)) != 0:
        raise ValueError("bad height/rank/world: %r/%r/%r" % (height, rank, world))
    return r0.value, r1.value


def shard_params(make_params, spp, rank, world):
    """make_params(spp=, sample_offset=, sample_stride=) -> PtbRenderParams for this rank, or None
    when the rank has no samples (spp < world)."""
    count, offset, stride = shard_samples(spp, rank, world)
    if count == 0:
        return None
    return make_params(spp=count, sample_offset=offset, sample_stride=stride)


def exchange_id(dist, make_id):
    """Rank 0 makes the communicator id (bytes), every rank returns it.  Works with any torch.distributed
    backend (gloo on CPU in the tests, nccl under torchrun)."""
    box = [make_id() if dist.get_rank() == 0 else None]
    dist.broadcast_object_list(box, src=0)
    return box[0]


def render_rows_emulated(render_accumulate, resolve_rows, make_params, spp, accum, out, dist):
    """The decomposition of ptb200_render_multi with the device calls replaced by callables and the
    reduce-scatter by an all-reduce + slice (gloo has no reduce-scatter): every rank accumulates its sample
    shard into `accum` (zeroed by the caller), the frames are summed, and every rank resolves ITS rows of
    `out` only.  Used by the CPU tests; the product path is Communicator.render."""
    rank, world = dist.get_rank(), dist.get_world_size()
    params = shard_params(make_params, spp, rank, world)
    stats = render_accumulate(params, accum) if params is not None else None
    dist.all_reduce(accum, op=dist.ReduceOp.SUM)
    r0, r1 = rows_of(accum.shape[0], rank, world)
    if r1 > r0:
        resolve_rows(accum[r0:r1], out[r0:r1])
    return stats


class MultiScene:
    """A scene replicated on this rank's device by ptb200_scene_create_multi."""

    def __init__(self, handle, flat):
        self.handle, self.flat = handle, flat

    def close(self):
        if self.handle:
            capi.product().ptb200_scene_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Communicator:
    """One process per GPU: ptb200_comm_init_rank on the current CUDA device, the id shipped through
    torch.distributed (rank 0 makes it)."""

    def __init__(self, dist):
        lib = capi.product()
        self.rank, self.world = dist.get_rank(), dist.get_world_size()

        def make_id():
            buf = C.create_string_buffer(capi.PTB_COMM_ID_BYTES)
            capi.check(lib.ptb200_comm_unique_id(buf))
            return buf.raw

        ident = exchange_id(dist, make_id)
        self.handle = C.c_void_p()
        capi.check(lib.ptb200_comm_init_rank(ident, self.rank, self.world, C.byref(self.handle)))

    def rows(self, height):
        return rows_of(height, self.rank, self.world)

    def create_scene(self, flat):
        """Collective.  Rank 0 uploads the images, the others receive them over NVLink."""
        out = (C.c_void_p * 1)()
        capi.check(capi.product().ptb200_scene_create_multi(self.handle, C.byref(flat.desc), out))
        return MultiScene(C.c_void_p(out[0]), flat)

    def render(self, scene, params, d_rgba=None, host_rgba=None, want_stats=False, stream=None):
        """Collective: one frame of params.spp samples over all ranks.  d_rgba: device (H, W, 4) torch tensor,
        host_rgba: (H, W, 4) float32 numpy array; this rank's rows of either are written."""
        scenes = (C.c_void_p * 1)(scene.handle)
        dptr = (C.c_void_p * 1)(d_rgba.data_ptr()) if d_rgba is not None else None
        stats = capi.PtbStats() if want_stats else None
        capi.check(capi.product().ptb200_render_multi(self.handle, scenes, C.byref(params), host_rgba.ctypes.data if host_rgba is not None else None, dptr,
                                                      C.byref(stats) if want_stats else None, stream))
        return stats

    def close(self):
        if self.handle:
            capi.product().ptb200_comm_destroy(self.handle)
            self.handle = None


class LocalCommunicator:
    """One process, n devices (ptb200_comm_init_all): what a C++ main calls; here for tests."""

    def __init__(self, devices):
        lib = capi.product()
        self.n = len(devices)
        arr = (C.c_int * self.n)(*devices)
        self.handle = C.c_void_p()
        capi.check(lib.ptb200_comm_init_all(arr, self.n, C.byref(self.handle)))

    def create_scenes(self, flat):
        out = (C.c_void_p * self.n)()
        capi.check(capi.product().ptb200_scene_create_multi(self.handle, C.byref(flat.desc), out))
        return [MultiScene(C.c_void_p(h), flat) for h in out]

    def render(self, scenes, params, host_rgba, want_stats=False):
        handles = (C.c_void_p * self.n)(*[s.handle for s in scenes])
        stats = (capi.PtbStats * self.n)() if want_stats else None
        capi.check(capi.product().ptb200_render_multi(self.handle, handles, C.byref(params), host_rgba.ctypes.data, None, stats, None))
        return list(stats) if want_stats else None

    def close(self):
        if self.handle:
            capi.product().ptb200_comm_destroy(self.handle)
            self.handle = None
