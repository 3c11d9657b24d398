"""BASELINE.json's full sizes, through properties that need no oracle run of that size (the oracle
would take hours): config 3 at 2^26 rays, configs 4/5 at 3840x2160.

* a ray's result does not depend on the batch it travels in (whole batch == the two halves);
* the fast policy agrees with the exact one (which is bit-identical to the reference, see
  test_gpu_intersect.py) on hit/miss and on t within 1e-5, except for a small grazing fraction;
* a strided subsample of the big batch is still bit-identical to the oracle;
* a frame is linear in its sample shards, a rendered rectangle equals the same window of the whole
  frame (the RNG is keyed by pixel and sample, not by launch geometry), every path terminates, and the
  frame's mean does not depend on the pixel grid the image plane is sampled on.
"""
import numpy as np
import pytest

import common
import golden_cases as gc
from grazing import flag_grazing
from ptb200 import capi, device, scenes
from ptb200.builder import product_builder

pytestmark = pytest.mark.gpu


def config3_rays_fast(n, seed):
    """Same distribution as scenes.config3_rays, generated in float32 blocks."""
    rs = np.random.default_rng(seed)

    def ball(radius):
        v = rs.standard_normal((n, 3), dtype=np.float32)
        v /= np.linalg.norm(v, axis=1, keepdims=True)
        v *= radius * rs.random((n, 1), dtype=np.float32) ** np.float32(1 / 3)
        return v

    o = ball(4.0)
    d = ball(1.0)
    norm = np.linalg.norm(d, axis=1, keepdims=True)
    d[norm[:, 0] < 1e-3] = (0.0, 0.0, 1.0)
    norm = np.linalg.norm(d, axis=1, keepdims=True)
    cls = np.arange(n) % 4
    d[cls == 2] /= norm[cls == 2]
    d[cls == 3] *= np.float32(2400.0) / norm[cls == 3]
    return o, d


def test_config3_full_batch_of_2_pow_26_rays():
    n = 1 << 26
    flat = device.Flat(scenes.config3(product_builder()))
    sc = device.DeviceScene(flat)
    o, d = config3_rays_fast(n, 0xB200 + 31)
    exact, _ = sc.intersect(o, d, 0, capi.PTB_FLAG_EXACT)
    assert 0.05 < (exact["kind"] > 0).mean() < 0.5
    # batch independence, both policies
    half = n // 2
    for flags, whole in ((capi.PTB_FLAG_EXACT, exact), (0, None)):
        if whole is None:
            whole, _ = sc.intersect(o, d, 0, flags)
        lo, _ = sc.intersect(o[:half], d[:half], 0, flags)
        assert lo.tobytes() == whole[:half].tobytes()
        hi, _ = sc.intersect(o[half:], d[half:], 0, flags)
        assert hi.tobytes() == whole[half:].tobytes()
        del lo, hi
    fast = whole
    # the exact policy is still the reference's arithmetic: every 1024th ray against the oracle
    sub = slice(0, n, 1024)
    oh, _ = common.orc_intersect(flat, o[sub], d[sub], 8)
    assert exact[sub].tobytes() == oh.tobytes()
    # fast against exact over the whole batch
    same = (fast["kind"] == exact["kind"]) & (fast["prim"] == exact["prim"]) & (fast["n_spans"] == exact["n_spans"])
    assert same.mean() > 0.995, same.mean()  # the rest are grazing rays (tests/grazing.py flags <= 2 % here)
    both = same & (exact["kind"] > 0)
    rel = np.abs(fast["t"][both] - exact["t"][both]) / np.abs(exact["t"][both])
    assert (rel <= 1e-5).mean() > 0.998, (rel <= 1e-5).mean()  # the rest graze a surface or start on one (ill-conditioned t)
    assert np.quantile(rel, 0.99) <= 5e-6, np.quantile(rel, 0.99)
    # ... and strictly, with the float64 grazing analysis, on every 64th ray (2^20 of them)
    sub = slice(0, n, 64)
    topo, illcond = flag_grazing(flat.desc, o[sub], d[sub])
    eh, fh = exact[sub], fast[sub]
    ok = ~topo
    assert (~ok).mean() <= 0.02
    for f in ("kind", "prim", "n_spans"):
        assert np.array_equal(fh[f][ok], eh[f][ok]), f
    hit = ok & (eh["kind"] > 0)
    hit &= ~illcond[np.arange(len(eh)), np.maximum(eh["prim"], 0)]
    rel = np.abs(fh["t"][hit] - eh["t"][hit]) / np.abs(eh["t"][hit])
    assert rel.max() <= 1e-5, rel.max()


@pytest.mark.parametrize("name,spp", [("c5", 4), ("c4", 2)])
def test_full_4k_frame_properties(name, spp):
    torch = pytest.importorskip("torch")
    W, H = 3840, 2160
    sc = device.DeviceScene(device.Flat(gc.render_scene(product_builder(), name)))
    full = torch.zeros((H, W, 4), device="cuda")
    st = sc.render_accumulate(device.default_params(W, H, spp, seed=11), full)
    torch.cuda.synchronize()
    assert st.paths == W * H * spp and st.rays >= st.paths
    assert torch.all(full[..., 3] == spp)  # every path ended and was counted exactly once
    assert torch.isfinite(full).all() and (full[..., :3] >= 0).all()
    # linear in the sample shards
    parts = torch.zeros_like(full)
    for g in range(2):
        sc.render_accumulate(device.default_params(W, H, spp // 2, seed=11, sample_offset=g, sample_stride=2), parts)
    torch.cuda.synchronize()
    assert torch.allclose(full, parts, rtol=1e-4, atol=1e-4)
    del parts
    # a rectangle of the frame is that window of the whole frame
    x0, y0, x1, y1 = 1900, 1000, 2412, 1300
    tile = torch.zeros_like(full)
    sc.render_accumulate(device.default_params(W, H, spp, seed=11, rect=(x0, y0, x1, y1)), tile)
    torch.cuda.synchronize()
    assert torch.allclose(tile[y0:y1, x0:x1], full[y0:y1, x0:x1], rtol=1e-4, atol=1e-4)
    tile[y0:y1, x0:x1] = 0
    assert not tile.any()  # nothing outside the rectangle is touched
    del tile
    # the same image plane at 1/10 of the resolution and 100x the samples has the same mean.  (The
    # screen extents stay those of the 4K camera: the reference's rays are not normalised and its glossy
    # lobe depends on |d| (quirk Q11), so src/test.cpp:450's camera is NOT resolution independent.)
    lowres = device.default_params(W // 10, H // 10, spp * 100, seed=12)
    lowres.screen_width, lowres.screen_height, lowres.screen_distance = float(W), float(H), float(2 * H)
    small = torch.zeros((H // 10, W // 10, 4), device="cuda")
    sc.render_accumulate(lowres, small)
    torch.cuda.synchronize()
    mean_full = (full[..., :3].sum() / (W * H * spp)).item()
    mean_small = (small[..., :3].sum() / (W * H * spp)).item()
    assert abs(mean_full - mean_small) <= 0.03 * mean_small, (mean_full, mean_small)
