"""CUDA CSG kernel vs the oracle, through the C ABI (ptb200_intersect).

Exact policy: bit-identical hit records and span lists (t, normals, primitives, materials).
Fast policy : identical hit/miss + topology and |dt|/t <= 1e-5 on rays not flagged grazing."""
import os

import numpy as np
import pytest

import common
import golden_cases as gc
from grazing import flag_grazing
from ptb200 import capi, device, scenes
from ptb200.builder import product_builder

pytestmark = pytest.mark.gpu
SPANS = np.load(os.path.join(common.GOLDEN, "spans.npz"))
FIELDS = ["start", "end", "start_normal", "end_normal", "start_prim", "end_prim", "start_material", "end_material"]


def assert_exact(sc, flat, o, d, max_spans):
    oh, osp = common.orc_intersect(flat, o, d, max_spans)
    gh, gsp = sc.intersect(o, d, max_spans, capi.PTB_FLAG_EXACT)
    assert gh.tobytes() == oh.tobytes()
    valid = np.arange(max_spans)[None, :] < np.minimum(oh["n_spans"], max_spans)[:, None]
    for f in FIELDS:
        assert gsp[f][valid].tobytes() == osp[f][valid].tobytes(), f
    return oh, osp


def assert_fast(sc, flat, o, d, oh, max_flagged=0.02, flags=0):
    gh, _ = sc.intersect(o, d, 0, flags)
    topo, illcond = flag_grazing(flat.desc, o, d)
    ok = ~topo
    assert (~ok).mean() <= max_flagged, "too many rays flagged grazing: %.4f" % (~ok).mean()
    assert np.array_equal(gh["kind"][ok], oh["kind"][ok])
    assert np.array_equal(gh["prim"][ok], oh["prim"][ok])
    assert np.array_equal(gh["n_spans"][ok], oh["n_spans"][ok])
    hit = ok & (oh["kind"] > 0)
    hit &= ~illcond[np.arange(len(o)), np.maximum(oh["prim"], 0)]  # float32 cannot hold this t to 1e-5 (origin on the surface)
    rel = np.abs(gh["t"][hit] - oh["t"][hit]) / np.abs(oh["t"][hit])
    assert rel.max() <= 1e-5, rel.max()
    cos = (gh["normal"][hit] * oh["normal"][hit]).sum(1)
    assert cos.min() > 1 - 1e-4
    return (~ok).sum()


@pytest.mark.parametrize("seed,leaves", gc.FUZZ_CASES)
def test_fuzz_trees_exact_and_golden(seed, leaves):
    b = product_builder()
    b.materials_reset()
    flat = device.Flat(gc.fuzz_scene(b, seed, leaves))
    sc = device.DeviceScene(flat)
    o, d = gc.fuzz_rays(seed)
    oh, osp = assert_exact(sc, flat, o, d, gc.MAX_SPANS)
    # and directly against the reference's own output
    gold, counts = SPANS["fuzz%d_spans" % seed], SPANS["fuzz%d_counts" % seed]
    gh, gsp = sc.intersect(o, d, gc.MAX_SPANS, capi.PTB_FLAG_EXACT)
    assert np.array_equal(gh["n_spans"], counts)
    valid = np.arange(gc.MAX_SPANS)[None, :] < counts[:, None]
    for f in ["start", "end", "start_normal", "end_normal", "start_material", "end_material"]:
        assert gsp[f][valid].tobytes() == gold[f][valid].tobytes(), f


def test_config3_batch_exact_and_fast():
    b = product_builder()
    flat = device.Flat(scenes.config3(b))
    sc = device.DeviceScene(flat)
    o, d = scenes.config3_rays(1 << 18)
    oh, _ = assert_exact(sc, flat, o, d, 8)
    assert_fast(sc, flat, o, d, oh)
    assert 0.05 < (oh["kind"] > 0).mean() < 0.5


@pytest.mark.parametrize("name", ["c1", "c2", "c4", "c5"])
def test_camera_rays_on_image_configs(name):
    b = product_builder()
    flat = device.Flat(gc.render_scene(b, name, small=True))  # geometry only: the small stand-in images
    sc = device.DeviceScene(flat)
    rs = np.random.RandomState(8)
    n = 60000
    o = np.zeros((n, 3), np.float32)
    d = np.stack([rs.uniform(-1920, 1920, n), rs.uniform(-1080, 1080, n), np.full(n, -2160.0)], 1).astype(np.float32)
    o[n // 2:] = rs.uniform(-2, 2, (n - n // 2, 3)) + np.array([0, 0, -5.0])  # secondary-like rays
    d[n // 2:] = rs.normal(size=(n - n // 2, 3))
    oh, _ = assert_exact(sc, flat, o, d, 10)
    assert_fast(sc, flat, o, d, oh, max_flagged=0.05)
    # the camera-ray variant of the evaluator (origin folded into per-primitive constants)
    h = n // 2
    assert_fast(sc, flat, o[:h], d[:h], oh[:h], max_flagged=0.05, flags=capi.PTB_FLAG_ORIGIN0)


def test_pool_overflow_path_matches(monkeypatch):
    """Force a tiny shared-memory pool so lists spill to the global overflow area."""
    b = product_builder()
    flat = device.Flat(gc.render_scene(b, "c4", small=True))
    sc = device.DeviceScene(flat)
    rs = np.random.RandomState(9)
    n = 20000
    o = rs.uniform(-3, 3, (n, 3)).astype(np.float32) + np.array([0, 0, -9.0], np.float32)
    d = rs.normal(size=(n, 3)).astype(np.float32)
    monkeypatch.setenv("PTB200_POOL_CAP", "2")
    assert_exact(sc, flat, o, d, 12)


def test_edge_cases_empty_and_degenerate():
    b = product_builder()
    flat = device.Flat(scenes.config2(b))
    sc = device.DeviceScene(flat)
    hits, _ = sc.intersect(np.zeros((0, 3), np.float32), np.zeros((0, 3), np.float32))
    assert len(hits) == 0
    # rays parallel to / inside planes, huge and tiny directions, NaN-free outputs
    o = np.array([[0, -1, 0], [0, 5, 0], [0, 0, 0], [0, 0, 0], [0, 300, 0]], np.float32)
    d = np.array([[1, 0, 0], [0, 1e-9, 1], [0, 0, -1e-6], [0, 0, -3e7], [0, 1, 0]], np.float32)
    assert_exact(sc, flat, o, d, 8)


def test_device_resident_entry_point():
    torch = pytest.importorskip("torch")
    b = product_builder()
    flat = device.Flat(scenes.config3(b))
    sc = device.DeviceScene(flat)
    o, d = scenes.config3_rays(50000)
    o4 = torch.zeros((len(o), 4), device="cuda")
    d4 = torch.zeros((len(o), 4), device="cuda")
    o4[:, :3] = torch.from_numpy(o).cuda()
    d4[:, :3] = torch.from_numpy(d).cuda()
    hits = torch.zeros((len(o), 8), dtype=torch.int32, device="cuda")
    sc.intersect_device(o4, d4, hits, capi.PTB_FLAG_EXACT, torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    got = hits.cpu().numpy().view(capi.HIT_DTYPE).reshape(-1)
    want, _ = common.orc_intersect(flat, o, d, 1)
    assert got.tobytes() == want.tobytes()
