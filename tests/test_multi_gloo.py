"""Host-side logic of the sample-sharded multi-GPU path (ptb200/multi.py) on CPU: gloo backend,
world size 2 (and 3, ragged).  The device calls are replaced by a deterministic stand-in that adds
f(pixel, global sample index) for every sample of the shard — what matters here is that the shards
partition the sample set, that one sum-reduce + resolve reproduces the unsharded frame, and that only
rank 0 resolves."""
import os
import socket

import numpy as np
import pytest

from ptb200 import multi

torch = pytest.importorskip("torch")
import torch.distributed as dist  # noqa: E402
import torch.multiprocessing as mp  # noqa: E402

W, H = 16, 9


def test_shards_partition_the_sample_set():
    for spp in (0, 1, 2, 7, 8, 1024):
        for world in (1, 2, 3, 4, 8):
            seen = []
            for rank in range(world):
                count, offset, stride = multi.shard_samples(spp, rank, world)
                seen += [offset + k * stride for k in range(count)]
            assert sorted(seen) == list(range(spp)), (spp, world)
    with pytest.raises(ValueError):
        multi.shard_samples(8, 2, 2)
    with pytest.raises(ValueError):
        multi.shard_samples(-1, 0, 2)


def contribution(sample):
    """Radiance the stand-in renderer adds to every pixel for one global sample index."""
    pix = torch.arange(W * H, dtype=torch.float32).reshape(H, W)
    rgb = torch.stack([pix * 0.01 + sample, torch.full((H, W), 0.5 * sample), pix % 3 + 1.0], -1)
    return torch.cat([rgb, torch.ones(H, W, 1)], -1)  # alpha = one finished sample


def fake_render_accumulate(params, accum):
    for k in range(params["spp"]):
        accum += contribution(params["sample_offset"] + k * params["sample_stride"])
    return params["spp"]


def fake_resolve(accum, out):
    out[..., :3] = accum[..., :3] / accum[..., 3:4]
    out[..., 3] = 1.0


def make_params(**kw):
    return dict(kw)


def worker(rank, world, port, spp, results):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        accum = torch.zeros(H, W, 4)
        out = torch.full((H, W, 4), -1.0)
        stats = multi.render_sharded(fake_render_accumulate, fake_resolve, make_params, spp, accum, out, dist)
        results[rank] = (out.numpy().copy(), stats)
    finally:
        dist.destroy_process_group()


def free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


@pytest.mark.parametrize("world,spp", [(2, 8), (2, 7), (3, 4), (2, 1)])
def test_sharded_frame_equals_unsharded(world, spp):
    manager = mp.Manager()
    results = manager.dict()
    mp.spawn(worker, args=(world, free_port(), spp, results), nprocs=world, join=True)
    # unsharded frame
    accum = torch.zeros(H, W, 4)
    fake_render_accumulate(dict(spp=spp, sample_offset=0, sample_stride=1), accum)
    want = torch.empty(H, W, 4)
    fake_resolve(accum, want)
    out0, _ = results[0]
    assert np.allclose(out0, want.numpy(), rtol=1e-6, atol=1e-6)
    for rank in range(1, world):
        out, stats = results[rank]
        assert (out == -1.0).all()  # only rank 0 resolves
        count, _, _ = multi.shard_samples(spp, rank, world)
        assert stats == (count if count else None)
