"""Sample-sharded multi-GPU rendering: one process per GPU, rank g of G renders
the global camera-sample indices {g, g+G, g+2G, ...} of every pixel, then ONE sum-reduce of the
accumulation buffers to rank 0 and a resolve (divide by the sample count) there.

The RNG of the render kernels is keyed by the GLOBAL sample index, so the union over ranks is exactly
the single-GPU sample set and the G-GPU image equals the 1-GPU image up to float summation order.
There is no data-path collective besides that reduce (132.7 MB at 3840x2160, once per frame).

Only plumbing lives here (torch.distributed); the rendering is ptb200_render_accumulate_device and
ptb200_resolve_device in libptb200.so.  tests/test_multi_gloo.py runs this logic on CPU with the gloo
backend and world size 2.
"""


def shard_samples(spp, rank, world):
    """(count, offset, stride) of the global sample indices rank `rank` renders: offset + k*stride,
    k < count.  The shards partition range(spp) for any spp >= 0 (ragged when world does not divide
    spp: the first spp % world ranks take one more)."""
    if world <= 0 or not 0 <= rank < world:
        raise ValueError("bad rank/world: %r/%r" % (rank, world))
    if spp < 0:
        raise ValueError("negative sample count")
    count = (spp - rank + world - 1) // world if spp > rank else 0
    return count, rank, world


def shard_params(make_params, spp, rank, world):
    """make_params(spp=, sample_offset=, sample_stride=) -> PtbRenderParams for this rank, or None
    when the rank has no samples (spp < world)."""
    count, offset, stride = shard_samples(spp, rank, world)
    if count == 0:
        return None
    return make_params(spp=count, sample_offset=offset, sample_stride=stride)


def reduce_accum(accum, dist=None, dst=0):
    """Sum-reduce the accumulation buffers (rgb = radiance sums, a = finished samples) to rank dst."""
    if dist is not None and dist.is_initialized() and dist.get_world_size() > 1:
        dist.reduce(accum, dst=dst, op=dist.ReduceOp.SUM)
    return accum


def render_sharded(render_accumulate, resolve, make_params, spp, accum, out, dist=None):
    """One frame: every rank accumulates its sample shard into `accum` (zeroed by the caller),
    the buffers are summed to rank 0, rank 0 resolves into `out`.  `render_accumulate(params, accum)`
    and `resolve(accum, out)` are the device calls (DeviceScene.render_accumulate / device.resolve).
    Returns whatever render_accumulate returned (stats) or None for an idle rank."""
    rank = dist.get_rank() if dist is not None and dist.is_initialized() else 0
    world = dist.get_world_size() if dist is not None and dist.is_initialized() else 1
    params = shard_params(make_params, spp, rank, world)
    stats = render_accumulate(params, accum) if params is not None else None
    reduce_accum(accum, dist)
    if rank == 0:
        resolve(accum, out)
    return stats
