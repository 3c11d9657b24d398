"""Multi-GPU rendering inside the library (csrc/ptb_multi.cu) on a box with at least
two GPUs: the C++-only caller tests/cpp/multi_gpu_main.cpp (ptb200_comm_init_all: one process, n devices, NCCL
reduce-scatter inside libptb200.so) must reproduce the one-device frame; the same through the ctypes binding
with ragged sample counts and a frame height the ranks do not divide."""
import os
import subprocess

import numpy as np
import pytest

from ptb200 import capi, device, multi, scenes
from ptb200.builder import product_builder

pytestmark = pytest.mark.gpu

BIN = os.path.join(capi.PKG_ROOT, "build", "multi_gpu_main")


def need_gpus(n):
    have = capi.product().ptb200_device_count()
    if have < n:
        pytest.skip("%d GPU(s) on this box, the multi-GPU path needs %d" % (have, n))


def test_cpp_only_caller_matches_one_device():
    need_gpus(2)
    if not os.path.exists(BIN):
        subprocess.check_call(["make", "-C", capi.PKG_ROOT, BIN])
    r = subprocess.run([BIN, "2", "192", "108", "64"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr


@pytest.mark.parametrize("n,spp,height", [(2, 7, 55), (2, 64, 54)])
def test_init_all_ragged_shards_and_rows(n, spp, height):
    need_gpus(n)
    flat = device.Flat(scenes.config2(product_builder()))
    W = 96
    params = device.default_params(W, height, spp, seed=77)
    params.screen_width, params.screen_height, params.screen_distance = 1920.0, 1080.0, 2160.0
    one, st1 = device.DeviceScene(flat).render(params)
    comm = multi.LocalCommunicator(list(range(n)))
    try:
        scs = comm.create_scenes(flat)
        many = np.full((height, W, 4), -1.0, np.float32)
        stats = comm.render(scs, params, many, want_stats=True)
        assert sum(s.paths for s in stats) == st1.paths and sum(s.rays for s in stats) == st1.rays
        assert np.allclose(many, one, rtol=1e-4, atol=1e-5)
        for s in scs:
            s.close()
    finally:
        comm.close()
