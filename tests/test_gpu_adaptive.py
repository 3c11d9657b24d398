"""ptb200_render_adaptive: the pixel-skipping quadtree of the reference's tile
renderer, RenderBlock::renderSquare (reference src/test.cpp:423-499), on the GPU.

Parity is checked against tests/adaptive_ref.py, a depth-first per-tile restatement of RenderBlock,
fed with the lattice values the GPU traced: same decisions, same interpolation arithmetic (bitwise),
same set of traced lattice points, same image."""
import ctypes as C

import numpy as np
import pytest

import adaptive_ref
from ptb200 import capi, device, scenes
from ptb200.builder import product_builder


def test_reference_walk_restatement_on_analytic_images():
    """CPU: the restatement itself on pixel functions with known outcomes."""
    asked = set()

    def flat(x, y):
        asked.add((x, y))
        return np.array([0.25, 0.5, 1.0], np.float32)

    img, src, traced = adaptive_ref.render_frame(flat, 70, 40, 32, 8, 0.003)
    # a constant image: every tile subdivides down to 8x8 squares (maximumSampleSize), which interpolate
    assert (img == np.array([0.25, 0.5, 1.0], np.float32)).all() and (src == 2).all()
    assert all(x % 8 == 0 and y % 8 == 0 for x, y in asked)
    # tiles do not share their edge points (each RenderBlock traces its own copies)
    assert traced > len(asked) and (0, 0) in asked and (64, 32) in asked and (96, 64) in asked

    def checker(x, y):
        return np.array([float((x + y) & 1), 0.0, 0.0], np.float32)

    img, src, traced = adaptive_ref.render_frame(checker, 33, 17, 16, 16, 0.003)
    # alternating pixels: red corners differ from black ones relatively by 1 > delta wherever a is red, and
    # black-vs-black squares interpolate; every traced pixel keeps its own value
    t = src == 1
    ys, xs = np.nonzero(t)
    assert (img[t][:, 0] == ((xs + ys) & 1)).all()


@pytest.mark.gpu
@pytest.mark.parametrize("name,width,height,block,mss,delta,spp", [
    ("c2", 200, 120, 32, 9, 0.003, 64),     # the reference's constants (src/test.cpp:38,50); ragged right/bottom tiles
    ("c2", 192, 128, 32, 16, 0.05, 16),     # exact multiples of the tile; looser threshold -> more interpolation
    ("c5", 150, 90, 16, 16, 0.2, 32),
    ("c4", 64, 64, 64, 64, 0.1, 16),        # one tile, interpolation allowed at every level
    ("c2", 40, 30, 2, 2, 0.003, 8),         # smallest tile
])
def test_adaptive_matches_reference_walk(name, width, height, block, mss, delta, spp):
    b = product_builder()
    world = {"c2": scenes.config2, "c4": scenes.config4, "c5": scenes.config5}[name](b)
    scene = device.DeviceScene(device.Flat(world))
    params = device.default_params(width, height, spp)
    img, source, lattice, stats, astats = scene.render_adaptive(params, block, mss, delta)
    asked = set()

    def pixel(x, y):
        assert lattice[y, x, 3] == 1.0, "the reference walk asks for lattice point (%d, %d) which the GPU did not trace" % (x, y)
        asked.add((x, y))
        return lattice[y, x, :3].copy()

    want, want_source, _ = adaptive_ref.render_frame(pixel, width, height, block, mss, delta)
    assert (source == want_source).all()
    assert img[..., :3].tobytes() == want.tobytes()
    assert (img[..., 3] == 1.0).all()
    ys, xs = np.nonzero(lattice[..., 3] == 1.0)
    assert set(zip(xs.tolist(), ys.tolist())) == asked, "GPU traced lattice points the reference walk never asks for"
    assert astats.pixels_traced == len(asked) and astats.pixels_interpolated == int((source == 2).sum())
    assert stats.paths == astats.pixels_traced * spp
    assert int((source == 1).sum()) + int((source == 2).sum()) == width * height


@pytest.mark.gpu
def test_adaptive_image_against_the_full_render():
    """Traced pixels are the same estimator as the full render (means agree); interpolated pixels lie
    within the interpolation threshold of it on smooth regions; with delta = 0 only squares with four
    bit-identical corners interpolate (the constant sky of config 2), and those pixels are exact."""
    b = product_builder()
    scene = device.DeviceScene(device.Flat(scenes.config2(b)))
    width, height, spp = 256, 144, 64
    params = device.default_params(width, height, spp)
    full, _ = scene.render(params)
    img, source, _, stats, astats = scene.render_adaptive(params, 32, 8, 0.0)
    interp = source == 2
    assert interp.any() and (source == 1).any()
    assert (img[interp][:, :3] == full[interp][:, :3]).all()
    traced = source == 1
    rel = abs(img[traced][:, :3].mean() - full[traced][:, :3].mean()) / full[traced][:, :3].mean()
    assert rel < 0.01, rel
    assert stats.paths < width * height * spp  # pixels were skipped
    img2, source2, _, _, astats2 = scene.render_adaptive(params, 32, 8, 0.1)
    assert astats2.pixels_interpolated > astats.pixels_interpolated
    err = np.abs(img2[..., :3] - full[..., :3]).mean() / full[..., :3].mean()
    assert err < 0.05, err


@pytest.mark.gpu
def test_adaptive_argument_errors():
    b = product_builder()
    scene = device.DeviceScene(device.Flat(scenes.config2(b)))
    params = device.default_params(64, 64, 4)
    for block in (0, 1, 3, 48, 2048):
        with pytest.raises(capi.PtbError) as e:
            scene.render_adaptive(params, block, 8, 0.003)
        assert e.value.status == -1
    with pytest.raises(capi.PtbError):
        scene.render_adaptive(device.default_params(64, 64, 4, rect=(0, 0, 32, 32)), 32, 8, 0.003)
    with pytest.raises(capi.PtbError):
        scene.render_adaptive(params, 32, -1, 0.003)
