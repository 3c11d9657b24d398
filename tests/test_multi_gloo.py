"""Host-side logic of the sample-sharded multi-GPU path on CPU: gloo backend, world size 2 (and 3,
ragged).  The partition rules are the library's own pure functions (ptb200_multi_shard / ptb200_multi_rows,
csrc/ptb_multi.cu — the ones ptb200_render_multi uses); the device calls are replaced by a deterministic
stand-in that adds f(pixel, global sample index) for every sample of the shard, and the NCCL reduce-scatter by
an all-reduce + slice (gloo has none).  What matters here: the shards partition the sample set, the row blocks
partition the frame, summing the shards and resolving row blocks per rank reproduces the unsharded frame, and
the 128-byte communicator id travels from rank 0 to every rank."""
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


def test_row_blocks_partition_the_frame():
    for height in (1, 2, 9, 54, 1080, 2160):
        for world in (1, 2, 3, 4, 8):
            rows, sizes = [], set()
            for rank in range(world):
                r0, r1 = multi.rows_of(height, rank, world)
                assert 0 <= r0 <= r1 <= height
                rows += list(range(r0, r1))
                if r1 - r0:
                    sizes.add(r1 - r0)
            assert rows == list(range(height)), (height, world)
            block = -(-height // world)
            assert max(sizes) == min(block, height)  # equal blocks (reduce-scatter), the last ones ragged or empty
    with pytest.raises(ValueError):
        multi.rows_of(0, 0, 2)
    with pytest.raises(ValueError):
        multi.rows_of(8, 2, 2)


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
        ident = multi.exchange_id(dist, lambda: bytes(range(128)))
        accum = torch.zeros(H, W, 4)
        out = torch.full((H, W, 4), -1.0)
        stats = multi.render_rows_emulated(fake_render_accumulate, fake_resolve, make_params, spp, accum, out, dist)
        results[rank] = (out.numpy().copy(), stats, ident)
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
    frame = np.full((H, W, 4), -1.0, np.float32)
    for rank in range(world):
        out, stats, ident = results[rank]
        assert ident == bytes(range(128))  # every rank got rank 0's communicator id
        r0, r1 = multi.rows_of(H, rank, world)
        assert (out[:r0] == -1.0).all() and (out[r1:] == -1.0).all()  # a rank delivers its own rows only
        frame[r0:r1] = out[r0:r1]
        count, _, _ = multi.shard_samples(spp, rank, world)
        assert stats == (count if count else None)
    assert np.allclose(frame, want.numpy(), rtol=1e-6, atol=1e-6)
