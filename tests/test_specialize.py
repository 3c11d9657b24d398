"""Scene-specialised trace kernels (path-trace_b200/csrc/ptb_specialize.cu): the CSG tree of a scene printed
as CUDA — one span-list variable per node, single spans in registers (csrc/ptb_csg_reg.cuh), every leaf's
constants literals — and compiled with NVRTC at run time.  They must compute exactly what the interpreting
kernels compute: same leaf arithmetic, same merge state machines.  (tests/test_spec_emulation.py checks the same
on the CPU.)

CPU: the generated source compiles for sm_100a (NVRTC needs no GPU).  GPU: bitwise identity with the
interpreter on ray batches (hits and whole span lists) and on rendered frames (ray counts; pixels up to
the order of the float atomics)."""
import time

import numpy as np
import pytest

import golden_cases as gc
from ptb200 import capi, device, scenes
from ptb200.builder import product_builder

SPEC, INTERP = capi.PTB_FLAG_SPECIALIZE, capi.PTB_FLAG_INTERPRET


def test_generated_source_compiles_for_sm100a():
    b = product_builder()
    for world, which in ((scenes.config5(b), 0), (scenes.config3(b), 1), (scenes.random_tree(b, 7, 9), 0)):
        flat = device.Flat(world)
        source, cubin = device.specialize_probe(flat, which)
        assert ("specTrace" if which == 0 else "specList") in source and "rlUnionIntoLean(ev, " in source and "ptb_csg_reg.cuh" in source
        assert ("ptbSpecTracePrimary" in source) == (which == 0) and ("ptbSpecIntersectO0" in source) == (which == 1)
        assert cubin > 10000
    # every leaf of the tree has its code: the eight isomorphic groups of config 4 (six transformed spheres cut by
    # two planes) share ONE out-of-line function that reads its constants from the module's table, called per group
    import re
    flat = device.Flat(scenes.config4(b))
    source, _ = device.specialize_probe(flat, 0, compile=False)
    calls = len(re.findall(r"= ptbShape0<O0>\(ev\.poolBase", source))
    assert calls == 8 and source.count("xfDisc<O0>") == 6 and "ldsTab128" in source
    assert calls * 6 >= 48 and (source.count("planeFold<O0>") - 2) + calls * 2 >= 22


def test_sphere_runs_are_marked_in_the_compiled_stream(monkeypatch):
    """compileFastProgram flags the first leaf of every complete Union subtree of transformed spheres
    (FH_RUN, bit 30 of a header without tail, run length in bits 12-15): the eight groups of six
    spheres of config 4; config 5 (untransformed spheres in a mixed tree) has none.  (Read off the generated source
    with every leaf written out in line: no shape functions.)"""
    import re
    monkeypatch.setenv("PTB_SPEC_SHAPES", "0")
    b = product_builder()

    def run_headers(world):
        source, _ = device.specialize_probe(device.Flat(world), 0, compile=False)
        heads = [int(h, 16) for h in re.findall(r"const uint4 k0 = make_uint4\((0x[0-9a-f]{8})u[^\n]*\n\s*xfDisc<O0>\(k0, ", source)]
        return [(h >> 12) & 15 for h in heads if (h & 0x40000080) == 0x40000000]

    assert run_headers(scenes.config4(b)) == [6] * 8
    assert run_headers(scenes.config5(b)) == []


@pytest.mark.gpu
@pytest.mark.parametrize("case", ["c3", "c4_camera", "c5", "fuzz11", "fuzz12", "fuzz13"])
def test_specialised_ray_batches_equal_the_interpreter_bitwise(case):
    b = product_builder()
    flags = 0
    if case == "c3":
        world = scenes.config3(b)
        o, d = scenes.config3_rays(1 << 18)
    elif case == "c4_camera":
        world = scenes.config4(b)
        rs = np.random.RandomState(4)
        n = 1 << 17
        d = np.stack([(rs.rand(n) * 2 - 1) * 3840, (rs.rand(n) * 2 - 1) * 2160, np.full(n, -4320.0)], 1).astype(np.float32)
        o = np.zeros_like(d)
        flags = capi.PTB_FLAG_ORIGIN0
    elif case == "c5":
        world = scenes.config5(b)
        o, d = scenes.fuzz_rays(1 << 17, 55)
    else:
        seed = int(case[4:])
        world = scenes.random_tree(b, seed, 6 + seed % 7)
        o, d = scenes.fuzz_rays(1 << 16, seed)
    scene = device.DeviceScene(device.Flat(world))
    h0, s0 = scene.intersect(o, d, 8, flags | INTERP)
    h1, s1 = scene.intersect(o, d, 8, flags | SPEC)
    assert h0.tobytes() == h1.tobytes()
    assert s0.tobytes() == s1.tobytes()
    assert (h0["kind"] != 0).any()


@pytest.mark.gpu
def test_specialised_kernels_with_span_pool_overflow(monkeypatch):
    """A two-slot shared-memory pool: span lists of the specialised kernels spill to the global overflow
    area (SpecShared.ovf) and still equal the interpreter's with its default pool."""
    b = product_builder()
    scene = device.DeviceScene(device.Flat(gc.render_scene(b, "c4", small=True)))
    rs = np.random.RandomState(9)
    n = 20000
    o = rs.uniform(-3, 3, (n, 3)).astype(np.float32) + np.array([0, 0, -9.0], np.float32)
    d = rs.normal(size=(n, 3)).astype(np.float32)
    h0, s0 = scene.intersect(o, d, 12, INTERP)
    monkeypatch.setenv("PTB200_POOL_CAP", "2")
    h1, s1 = scene.intersect(o, d, 12, SPEC)
    assert h0.tobytes() == h1.tobytes() and s0.tobytes() == s1.tobytes()
    assert (h0["n_spans"] > 2).any()


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["c2", "c4", "c5"])
def test_specialised_render_equals_the_interpreter(name):
    b = product_builder()
    world = {"c2": scenes.config2, "c4": scenes.config4, "c5": scenes.config5}[name](b)
    scene = device.DeviceScene(device.Flat(world))
    p0 = device.default_params(192, 108, 32, flags=INTERP)
    p1 = device.default_params(192, 108, 32, flags=SPEC)
    img0, st0 = scene.render(p0)
    assert scene.specialized()[2] == 0
    img1, st1 = scene.render(p1)
    ready, ms, spec_passes, passes = scene.specialized()
    assert ready and spec_passes == passes > 0
    assert (st0.rays, st0.paths, st0.rays_bounce) == (st1.rays, st1.paths, st1.rays_bounce)
    # same paths, same contributions; only the order of the float atomicAdds per pixel differs
    np.testing.assert_allclose(img1, img0, rtol=2e-5, atol=1e-6)


@pytest.mark.gpu
def test_background_compile_switches_over_without_changing_the_image():
    """Default policy: a large render starts the compile on a background thread and keeps interpreting;
    later passes / frames run the specialised kernels.  The images agree."""
    b = product_builder()
    scene = device.DeviceScene(device.Flat(scenes.random_tree(b, 99, 8)))
    params = device.default_params(1024, 1024, 64, samples_per_pass=4)  # 2^26 paths, 16 passes
    ref, _ = scene.render(device.default_params(1024, 1024, 64, samples_per_pass=4, flags=INTERP))
    img, _ = scene.render(params)
    deadline = time.time() + 120
    while not scene.specialized()[0] and time.time() < deadline:
        time.sleep(0.2)
    assert scene.specialized()[0], "the background compile never finished"
    img2, _ = scene.render(params)
    assert scene.specialized()[2] == scene.specialized()[3] > 0
    np.testing.assert_allclose(img, ref, rtol=2e-5, atol=1e-6)
    np.testing.assert_allclose(img2, ref, rtol=2e-5, atol=1e-6)


def test_run_bounding_spheres_enclose_their_leaves():
    """Behind FK_END the compiled stream carries one bounding sphere per leaf of every sphere run (bits 16-29 of the
    run's first header: word index), as {C, sqrt(|C|^2 - R^2)}: what the camera-ray test needs.  Every point of a
    leaf {x : |M x + t'|^2 <= r^2} must lie inside its sphere — the interpreter skips a leaf whose sphere no ray of
    the warp meets — and the spheres must not be slack by more than a few per cent of the leaf's longest half-axis."""
    b = product_builder()
    text, _ = device.specialize_probe(device.Flat(scenes.config4(b)), 2, compile=False)
    words = np.array([int(w, 16) for w in text.split()], np.uint32).reshape(-1, 4)
    as_f = lambda u: np.array(u, np.uint32).view(np.float32).astype(np.float64)
    rs = np.random.RandomState(3)
    u = rs.normal(size=(4000, 3))
    u /= np.linalg.norm(u, axis=1, keepdims=True)
    runs = 0
    w = 0
    while (words[w, 0] & 0x3F) != 8:  # FK_END
        x = int(words[w, 0])
        if x & 1:  # FK_SPHERE_XF: header + 3 rows
            if (x & 0x40000080) == 0x40000000 and (x >> 16) & 0x3FFF:
                runs += 1
                n, table = (x >> 12) & 15, (x >> 16) & 0x3FFF
                for k in range(n):
                    rec = words[w + 4 * k:w + 4 * k + 4]
                    rows = as_f(rec[1:4].reshape(-1)).reshape(3, 4)
                    m, t, r = rows[:, :3], rows[:, 3], np.sqrt(-as_f(rec[0, 2:3])[0])
                    c = as_f(words[table + k, :3])
                    big_r = np.sqrt(c @ c - (as_f(words[table + k, 3:4])[0] / (1 - 1e-5)) ** 2)
                    surface = np.linalg.solve(m, (r * u - t).T).T
                    dist = np.linalg.norm(surface - c, axis=1)
                    assert dist.max() <= big_r, (w, k)
                    assert big_r <= 1.02 * r / np.linalg.svd(m, compute_uv=False).min() + 1e-4, (w, k)
            w += 4
        elif x & 2:
            w += 2 + 2 * int(words[w, 2])
        elif x & 4:
            w += 2
        else:
            w += 1
    assert runs == 8


@pytest.mark.gpu
@pytest.mark.parametrize("order", ["pixel", "random"])
def test_bounding_spheres_change_no_span(order, monkeypatch):
    """The interpreter's bounding-sphere votes (camera rays: PTB_FLAG_ORIGIN0) only skip leaves that would have
    missed: hits and whole span lists of config 4 are bit-identical with (default) and without (PTB_FAST_BOUNDS=0,
    read when the scene is created) the spheres — for rays in pixel order, where a warp covers a few pixels and whole
    runs are skipped, and in random order, where single leaves are."""
    b = product_builder()
    w, h = 1024, 576
    ys, xs = np.mgrid[0:h, 0:w]
    rs = np.random.RandomState(12)
    px = (xs + rs.rand(h, w)).reshape(-1) * (3840.0 / w)
    py = (ys + rs.rand(h, w)).reshape(-1) * (2160.0 / h)
    d = np.stack([(px * 2 / 3840.0 - 1) * 3840.0, (1 - py * 2 / 2160.0) * 2160.0, np.full(px.shape, -4320.0)], 1).astype(np.float32)
    if order == "random":
        d = d[rs.permutation(len(d))]
    o = np.zeros_like(d)
    flat = device.Flat(scenes.config4(b))
    with_bounds = device.DeviceScene(flat)
    monkeypatch.setenv("PTB_FAST_BOUNDS", "0")
    without = device.DeviceScene(flat)
    monkeypatch.delenv("PTB_FAST_BOUNDS")
    h1, s1 = with_bounds.intersect(o, d, 8, capi.PTB_FLAG_ORIGIN0 | INTERP)
    h0, s0 = without.intersect(o, d, 8, capi.PTB_FLAG_ORIGIN0 | INTERP)
    assert h0.tobytes() == h1.tobytes() and s0.tobytes() == s1.tobytes()
    on_sphere = (h0["kind"] != 0) & (h0["prim"] < 64) & (h0["prim"] % 8 < 6)
    assert 0.05 < on_sphere.mean() < 0.9  # the spheres are what a good part of the rays see first, and not all
    # and the scene without them really has none: its program is 48 words shorter
    text1, _ = device.specialize_probe(flat, 2, compile=False)
    monkeypatch.setenv("PTB_FAST_BOUNDS", "0")
    text0, _ = device.specialize_probe(flat, 2, compile=False)
    assert len(text1.split()) - len(text0.split()) == 48 * 4
