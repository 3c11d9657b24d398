"""ptb200_intersect_device_static (csrc/ptb_noalloc.cu): the ray-batch CSG kernels with their scratch — the span-pool
overflow area — in a static device array instead of a cudaMalloc'ed block, for callers that cannot allocate;
no counterpart in the reference.

Through the C ABI: its hits and span lists are bit-identical (the tolerance the ray-batch tests use for the same
kernels) to ptb200_intersect_device's, in both arithmetic policies, on two calls in a row (the array keeps the first
call's contents); the device's free memory does not move across the second call; a scene that needs more than the
stated 16 MiB is refused with a message naming both sizes.  Collected last."""
import ctypes as C

import numpy as np
import pytest

from ptb200 import capi, device
from ptb200.builder import product_builder

pytestmark = pytest.mark.gpu
MAX_SPANS = 24


def sphere_row(b, n, spacing=1.5, radius=0.5):
    """n disjoint spheres along z: a ray down the axis crosses all of them -> a root list of n spans, far beyond
    the shared-memory pool (12 slots fast / 16 exact), so the overflow area is really used."""
    mats = [b.material(b.color(0.5), b.color(1)) for _ in range(4)]
    return b.union_array([b.sphere((0.0, 0.0, -4.0 - spacing * i), radius, mats[i % 4]) for i in range(n)])


def rays(n, seed=5):
    rs = np.random.RandomState(seed)
    o = np.zeros((n, 4), np.float32)
    d = np.zeros((n, 4), np.float32)
    o[:, :2] = rs.uniform(-0.3, 0.3, (n, 2))
    d[:, :3] = rs.normal(size=(n, 3)) * 0.05
    d[:, 2] = -rs.choice([1.0, 2.0, 2400.0], n)  # mostly down the row of spheres
    third = n // 3
    o[:third, :3] = rs.uniform(-2, 2, (third, 3)) + np.array([0, 0, -20.0])  # and from inside the row, any direction
    d[:third, :3] = rs.normal(size=(third, 3))
    return o, d


def intersect(entry, scene, o4, d4, n, flags, torch):
    hits = torch.zeros((n, 8), dtype=torch.int32, device="cuda")
    spans = torch.zeros((n, MAX_SPANS, 12), dtype=torch.int32, device="cuda")
    capi.check(entry(scene.handle, o4.data_ptr(), d4.data_ptr(), n, hits.data_ptr(), spans.data_ptr(), MAX_SPANS, flags, None))
    torch.cuda.synchronize()
    return hits, spans


@pytest.mark.parametrize("flags", [capi.PTB_FLAG_INTERPRET, capi.PTB_FLAG_EXACT])
def test_static_scratch_entry_point_equals_the_allocating_one(flags):
    torch = pytest.importorskip("torch")
    lib = capi.product()
    assert lib.ptb200_static_scratch_bytes() == 16 << 20
    scene = device.DeviceScene(device.Flat(sphere_row(product_builder(), 24)))
    n = 1 << 16
    o, d = rays(n)
    o4, d4 = torch.from_numpy(o).cuda(), torch.from_numpy(d).cuda()
    want_hits, want_spans = intersect(lib.ptb200_intersect_device, scene, o4, d4, n, flags, torch)
    assert int(want_hits[:, 7].max()) > 16  # span lists longer than the shared pool: the overflow area was used
    for call in range(2):  # two calls in a row: the second starts on the first one's leftovers
        hits, spans = intersect(lib.ptb200_intersect_device_static, scene, o4, d4, n, flags, torch)
        assert torch.equal(hits, want_hits), "call %d: hits differ" % call
        assert torch.equal(spans, want_spans), "call %d: span lists differ" % call
    hits2 = torch.zeros((n, 8), dtype=torch.int32, device="cuda")
    before = torch.cuda.mem_get_info()[0]
    capi.check(lib.ptb200_intersect_device_static(scene.handle, o4.data_ptr(), d4.data_ptr(), n, hits2.data_ptr(), None, 0, flags, None))
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] == before  # nothing was allocated on the device
    assert torch.equal(hits2, want_hits)


def test_static_scratch_entry_point_refuses_what_does_not_fit():
    torch = pytest.importorskip("torch")
    lib = capi.product()
    scene = device.DeviceScene(device.Flat(sphere_row(product_builder(), 200, spacing=0.2, radius=0.05)))
    o4 = torch.zeros((32, 4), device="cuda")
    d4 = torch.ones((32, 4), device="cuda")
    hits = torch.zeros((32, 8), dtype=torch.int32, device="cuda")
    rc = lib.ptb200_intersect_device_static(scene.handle, o4.data_ptr(), d4.data_ptr(), 32, hits.data_ptr(), None, 0, 0, None)
    msg = lib.ptb200_last_error().decode()
    assert rc == -1 and "too large" in msg and str(16 << 20) in msg and "ptb200_intersect_device" in msg
    # the allocating entry point takes the same scene
    capi.check(lib.ptb200_intersect_device(scene.handle, o4.data_ptr(), d4.data_ptr(), 32, hits.data_ptr(), None, 0, 0, None))
    torch.cuda.synchronize()
    # a run-time compile allocates: refused here
    assert lib.ptb200_intersect_device_static(scene.handle, o4.data_ptr(), d4.data_ptr(), 32, hits.data_ptr(), None, 0, capi.PTB_FLAG_SPECIALIZE, None) == -1
