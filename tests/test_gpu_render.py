"""Shading, texture and whole-image parity of the CUDA wavefront integrator (through the C ABI)
against the reference's outputs committed under tests/golden/ and against the oracle.

The integrator replaces the reference's 10^4-way branching estimator by a one-sample estimator of
the same expectation and its LCG by a counter-based RNG, so image parity is
statistical: stated tolerances below."""
import ctypes as C
import os
import time

import numpy as np
import pytest

import common
import golden_cases as gc
from ptb200 import capi, device, scenes
from ptb200.builder import product_builder

pytestmark = pytest.mark.gpu
KAT = np.load(os.path.join(common.GOLDEN, "kat.npz"))
RENDERS = np.load(os.path.join(common.GOLDEN, "renders.npz"))

# Tolerances (relative to the golden image's mean radiance):
MEAN_TOL = 0.005      # |mean(gpu) - mean(ref)| / mean(ref)
RMSE_TOL = {"c1": 0.03, "c2": 0.01, "c4": 0.012, "c5": 0.05}  # measured on B200: 0.016, 0.0053, 0.0054, 0.025  # per-pixel RMS error / mean(ref); sample counts per case: golden_cases.RENDER_SETUP


@pytest.mark.parametrize("fold", ["1", "0"])
def test_texture_known_answers(fold, monkeypatch):
    """Point-sampled texture programs: same texel as the reference except where float rounding of
    atan2/asin/log moves a lookup across a texel edge (<= 1% of positions), values within 1e-5.
    Through the records chains are folded into at scene create (DevTexBaked: every chain here but `const`,
    `moved_const` and `double_affine`) and, with PTB_TEX_FOLD=0, through the interpreter."""
    monkeypatch.setenv("PTB_TEX_FOLD", fold)
    b = product_builder()
    pos = gc.kat_positions()
    for name, tex in gc.kat_textures(b).items():
        flat = device.Flat(texture=tex)
        sc = device.DeviceScene(flat)
        rgb, f = sc.texture_eval(flat.desc.materials[0].tex[0], pos)
        want_rgb, want_f = KAT["tex_%s_rgb" % name], KAT["tex_%s_f" % name]
        tol = 1e-5 * np.maximum(1.0, np.abs(want_rgb).max(1)) + 1e-7
        bad = (np.abs(rgb - want_rgb).max(1) > tol) | (np.abs(f - want_f) > tol)
        limit = 0.01 if name in ("spherical", "env", "mirrorball") else 0.0
        assert bad.mean() <= limit, (name, bad.mean())


def test_texture_api_through_cpp_classes():
    """Texture::getColor / getFloat of the host C++ classes evaluate on the GPU."""
    b = product_builder()
    tex = gc.kat_textures(b)["multiply"]
    rgb, f = b.tex_eval(tex, gc.kat_positions()[:8])
    assert np.allclose(rgb, KAT["tex_multiply_rgb"][:8], rtol=1e-5, atol=1e-7) and np.allclose(f, KAT["tex_multiply_f"][:8], rtol=1e-5, atol=1e-7)


def reference_lobe(normal, mirror, sc, n, seed):
    """The reference's rejection loops (include/path-trace.h:142-158) over its own ball sampler."""
    lib = common.oracle()
    k = 1.0 / sc - 1.0
    out = []
    need = n
    while need > 0:
        ball = np.zeros((4 * need, 3), np.float32)
        lib.orc_rng_ball(seed, 4 * need, 1.0, 0.0, ball.ctypes.data)
        seed += 1
        dvec = ball + k * np.asarray(mirror, np.float32)
        keep = dvec @ np.asarray(normal, np.float32) > 1e-3
        dvec = dvec[keep][:need]
        out.append(dvec / np.linalg.norm(dvec, axis=1, keepdims=True))
        need -= len(dvec)
    return np.concatenate(out)


@pytest.mark.parametrize("sc,mirror", [(1.0, (0.3, 0.2, 0.93)), (0.3, (0.6, 0.0, 0.8)), (0.3, (0.995, 0.0, 0.1)), (0.05, (0.0, 0.7, 0.7)), (0.5, (1200.0, 0.0, 2000.0))])
def test_scatter_lobe_matches_reference_distribution(sc, mirror):
    normal = (0.0, 0.0, 1.0)
    n = 400000
    want = reference_lobe(normal, mirror, sc, n, 11)
    for flags in (0, capi.PTB_FLAG_REJECTION):
        got = device.sample_scatter(normal, mirror, sc, n, seed=5, flags=flags)
        assert np.abs(np.linalg.norm(got, axis=1) - 1).max() < 1e-4
        se = 4.0 / np.sqrt(n)
        assert np.abs(got.mean(0) - want.mean(0)).max() < se + 2e-3, (flags, got.mean(0), want.mean(0))
        m2g, m2w = (got[:, :, None] * got[:, None, :]).mean(0), (want[:, :, None] * want[:, None, :]).mean(0)
        assert np.abs(m2g - m2w).max() < se + 2e-3
        # the quantity the integrator weights by: factor = 1 - (1 - n.d) * sc
        assert abs((1 - (1 - got[:, 2]) * sc).mean() - (1 - (1 - want[:, 2]) * sc).mean()) < se + 1e-3


def test_scatter_lobe_rotated_frame():
    rs = np.random.RandomState(2)
    normal = rs.normal(size=3)
    normal /= np.linalg.norm(normal)
    mirror = rs.normal(size=3) + 1.5 * normal
    want = reference_lobe(normal, mirror, 0.4, 300000, 3)
    got = device.sample_scatter(normal, mirror, 0.4, 300000, seed=9)
    assert np.abs(got.mean(0) - want.mean(0)).max() < 6e-3
    assert (got @ normal).min() > 0


def render_on_lattice(sc, spp, seed=1, flags=0, lattice=None):
    """The golden lattice is every 20th (40th at 4K) pixel of the image plane: ptb200_render_pixels traces
    exactly those pixels of the full-size frame."""
    W, H, x0, y0, step, nx, ny = lattice or gc.LATTICE
    ys, xs = np.meshgrid(y0 + step * np.arange(ny), x0 + step * np.arange(nx), indexing="ij")
    out, stats = sc.render_pixels(device.default_params(W, H, spp, seed=seed, flags=flags), np.stack([xs.ravel(), ys.ravel()], 1))
    assert (out[:, 3] == 1).all()
    return out[:, :3].reshape(ny, nx, 3), stats


def test_render_pixels_equals_the_frame_render():
    """A listed pixel has the value the frame render gives it (same samples, same RNG streams)."""
    sc = device.DeviceScene(device.Flat(gc.render_scene(product_builder(), "c2")))
    W, H = 192, 108
    p = device.default_params(W, H, 64, seed=5)
    frame, st = sc.render(p)
    rs = np.random.RandomState(1)
    xy = np.stack([rs.randint(0, W, 500), rs.randint(0, H, 500)], 1)
    got, st2 = sc.render_pixels(p, xy)
    assert np.allclose(got, frame[xy[:, 1], xy[:, 0]], rtol=1e-5, atol=1e-6)
    assert st2.paths == 500 * 64
    with pytest.raises(capi.PtbError):
        sc.render_pixels(p, [[W, 0]])


@pytest.mark.parametrize("name", gc.RENDER_CASES)
def test_image_matches_reference_golden(name):
    b = product_builder()
    flat = device.Flat(gc.render_scene(b, name))  # c4 / c5: the full-size assets and the 4K camera bench.py times
    sc = device.DeviceScene(flat)
    gold = RENDERS[name]
    lattice, _, gpu_spp = gc.RENDER_SETUP[name]
    img, _ = render_on_lattice(sc, gpu_spp, lattice=lattice)
    mean = gold.mean()
    rel_mean = abs(img.mean() - mean) / mean
    rmse = np.sqrt(((img - gold) ** 2).mean()) / mean
    print("%s: mean gpu %.5f ref %.5f (rel %.4f)  relRMSE %.4f" % (name, img.mean(), mean, rel_mean, rmse))
    assert rel_mean < MEAN_TOL
    assert rmse < RMSE_TOL[name]


def test_rejection_and_cap_samplers_render_the_same_image():
    b = product_builder()
    sc = device.DeviceScene(device.Flat(gc.render_scene(b, "c2")))
    a, _ = render_on_lattice(sc, 2048, seed=3)
    r, _ = render_on_lattice(sc, 2048, seed=4, flags=capi.PTB_FLAG_REJECTION)
    assert abs(a.mean() - r.mean()) / a.mean() < 0.004
    assert np.sqrt(((a - r) ** 2).mean()) / a.mean() < 0.03


def test_sample_sharding_reproduces_the_unsharded_image():
    """Rank g of G renders global samples g, g+G, ...: the union is the 1-GPU sample set, so the
    summed accumulation buffers equal the unsharded ones up to float summation order."""
    torch = pytest.importorskip("torch")
    b = product_builder()
    sc = device.DeviceScene(device.Flat(gc.render_scene(b, "c2")))
    W, H, spp, G = 160, 90, 32, 4
    full = torch.zeros((H, W, 4), device="cuda")
    sc.render_accumulate(device.default_params(W, H, spp, seed=77), full)
    parts = torch.zeros((H, W, 4), device="cuda")
    for g in range(G):
        sc.render_accumulate(device.default_params(W, H, spp // G, seed=77, sample_offset=g, sample_stride=G), parts)
    torch.cuda.synchronize()
    assert torch.all(full[..., 3] == spp) and torch.all(parts[..., 3] == spp)
    assert torch.allclose(full, parts, rtol=1e-4, atol=1e-4)
    out = torch.empty_like(full)
    device.resolve(full, out)
    torch.cuda.synchronize()
    assert torch.allclose(out[..., :3], full[..., :3] / spp, rtol=1e-6)


def test_passes_do_not_change_the_estimate():
    b = product_builder()
    sc = device.DeviceScene(device.Flat(gc.render_scene(b, "c2")))
    a, sa = sc.render(device.default_params(128, 72, 64, seed=5, samples_per_pass=64))
    c, sb = sc.render(device.default_params(128, 72, 64, seed=5, samples_per_pass=8))
    assert sa.paths == sb.paths == 128 * 72 * 64 and sa.rays == sb.rays
    assert np.allclose(a, c, rtol=1e-4, atol=1e-5)


def test_display8_matches_the_reference_conversion():
    """max(0, min(255, (int)floor(256 * c / count))), reference src/test.cpp:992-995 (x86-64: (int) of a
    NaN or out-of-range float is INT_MIN, hence 0 after the clamp)."""
    lib = capi.product()
    rs = np.random.RandomState(3)
    n = 100003
    rgba = rs.uniform(-0.2, 1.4, (n, 4)).astype(np.float32)
    rgba[:, 3] = rs.randint(0, 5, n)
    rgba[:16, 0] = [0.0, 1.0, 255.0 / 256.0, np.nextafter(np.float32(1), np.float32(0)), 1e-8, -1e-8, np.inf, -np.inf, np.nan, 1e7, 8.4e6, 3e38, -3e38, 0.5, 0.25, 0.999]
    rgba[:16, 3] = 1

    def want(divide):
        count = np.where(divide, rgba[:, 3:4], np.float32(1)).astype(np.float32)
        with np.errstate(all="ignore"):
            v = np.floor((np.float32(256) * rgba[:, :3]) / count)
            ok = np.isfinite(v) & (v < 2147483648.0) & (v >= -2147483648.0)
            out = np.where(ok, np.clip(np.where(ok, v, 0), 0, 255), 0).astype(np.uint8)
        out[(count <= 0)[:, 0]] = 0
        return out

    for divide in (0, 1):
        got = np.zeros((n, 4), np.uint8)
        assert lib.ptb200_display8(rgba.ctypes.data, n, divide, got.ctypes.data) == 0
        assert np.array_equal(got[:, :3], want(bool(divide)))
        assert (got[:, 3] == 255).all()
    torch = pytest.importorskip("torch")
    d_in = torch.from_numpy(rgba).cuda()
    d_out = torch.zeros((n, 4), dtype=torch.uint8, device="cuda")
    assert lib.ptb200_display8_device(d_in.data_ptr(), d_out.data_ptr(), n, 1, None) == 0
    torch.cuda.synchronize()
    assert np.array_equal(d_out.cpu().numpy()[:, :3], want(True))


def test_scratch_cache_survives_scene_lifetimes_and_trim():
    """Scratch blocks are cached across scenes (include/ptb200.h: ptb200_trim): a scene created after
    another one was destroyed renders the same image from recycled (dirty) blocks, and again after a trim."""
    lib = capi.product()
    flat = device.Flat(gc.render_scene(product_builder(), "c2"))
    params = device.default_params(256, 144, 16, seed=9)
    images = []
    for trim in (False, False, True):
        sc = device.DeviceScene(flat)
        img, st = sc.render(params)
        assert st.paths == 256 * 144 * 16
        images.append(img)
        sc.close()
        if trim:
            assert lib.ptb200_trim() == 0
    sc = device.DeviceScene(flat)
    images.append(sc.render(params)[0])
    for img in images[1:]:
        assert np.allclose(img, images[0], rtol=1e-4, atol=1e-5)  # float atomics: summation order only


def test_async_job_boundary():
    """BlockRenderer-style boundary: start, poll done(), progressive reads, final result."""
    lib = capi.product()
    b = product_builder()
    sc = device.DeviceScene(device.Flat(gc.render_scene(b, "c2")))
    W, H = 256, 144
    p = device.default_params(W, H, 256, seed=9, samples_per_pass=4, rect=(32, 16, 96, 80))
    job = C.c_void_p()
    capi.check(lib.ptb200_render_async(sc.handle, C.byref(p), C.byref(job)))
    other = C.c_void_p()
    assert lib.ptb200_render_async(sc.handle, C.byref(p), C.byref(other)) == -5  # PTB_EBUSY
    buf = np.full((H, W, 4), -1.0, np.float32)
    t0 = time.time()
    while lib.ptb200_job_done(job) == 0 and time.time() - t0 < 120:
        capi.check(lib.ptb200_job_read(job, buf.ctypes.data, W, H))
        time.sleep(0.001)
    stats = capi.PtbStats()
    capi.check(lib.ptb200_job_wait(job, C.byref(stats)))
    assert lib.ptb200_job_done(job) == 1
    capi.check(lib.ptb200_job_read(job, buf.ctypes.data, W, H))
    lib.ptb200_job_destroy(job)
    inside = buf[16:80, 32:96]
    assert np.all(inside[..., 3] == 1.0) and inside[..., :3].mean() > 0.1
    outside = buf.copy()
    outside[16:80, 32:96] = -1.0
    assert np.all(outside == -1.0)  # only valid pixels are written
    assert stats.paths == 64 * 64 * 256
    sync, _ = sc.render(device.default_params(W, H, 256, seed=9, samples_per_pass=4, rect=(32, 16, 96, 80)))
    assert np.allclose(sync[16:80, 32:96, :3], inside[..., :3], rtol=1e-4, atol=1e-5)


def test_cpp_host_render_and_span_iterator():
    """The C++ host layer end to end: ptb200::renderFrame and Object::makeSpanIterator()."""
    lib = capi.product()
    b = product_builder()
    b.materials_reset()
    world = gc.fuzz_scene(b, 3, 6)
    o, d = gc.fuzz_rays(3)
    spans, counts = b.trace_spans(world, o[:64], d[:64], gc.MAX_SPANS)
    gold = np.load(os.path.join(common.GOLDEN, "spans.npz"))
    assert np.array_equal(counts, gold["fuzz3_counts"][:64])
    valid = np.arange(gc.MAX_SPANS)[None, :] < counts[:, None]
    for f in ["start", "end", "start_normal", "end_normal", "start_material", "end_material"]:
        assert spans[f][valid].tobytes() == gold["fuzz3_spans"][f][:64][valid].tobytes()
    w2 = gc.render_scene(b, "c2")
    p = device.default_params(96, 54, 256, seed=2)
    out = np.zeros((54, 96, 4), np.float32)
    assert lib.ptbh_render(w2, C.byref(p), out.ctypes.data, None) == 0
    direct, _ = device.DeviceScene(device.Flat(w2)).render(device.default_params(96, 54, 256, seed=2))
    assert np.allclose(out, direct, rtol=1e-4, atol=1e-5)


def test_page_locked_image_buffers_give_the_same_scene():
    """ptb200_scene_create DMAs image pixels straight from page-locked memory (device.Flat.pin_images) and stages
    pageable ones through its own pinned buffer: the same scene either way."""
    pytest.importorskip("torch")
    b = product_builder()
    flat = device.Flat(gc.render_scene(b, "c5", small=True))
    p = device.default_params(96, 54, 16, seed=11)
    a, _ = device.DeviceScene(flat).render(p)
    flat.pin_images()
    c, _ = device.DeviceScene(flat).render(p)
    assert np.allclose(a, c, rtol=1e-5, atol=1e-6)  # (float atomics: the order of a pixel's sums is not fixed)


def test_large_programs_specialise_the_camera_ray_kernel_only():
    """Config 4 (70 leaves, 340 program words) is beyond the size that is specialised whole; its camera rays run a
    generated kernel whose shape-function calls are guarded by bounding-sphere votes, its queued rays the interpreter
    (ptb200_scene_specialized == 2), and the image is the interpreter's up to the order of the float atomics."""
    if os.environ.get("PTB_SPECIALIZE", "") == "0" or os.environ.get("PTB_SPEC_LARGE", "") == "0":
        pytest.skip("specialisation switched off by the environment")
    b = product_builder()
    flat = device.Flat(gc.render_scene(b, "c4", small=True))
    sc = device.DeviceScene(flat)
    sc.render(device.default_params(1920, 1080, 33, seed=1))  # >= 2^26 rays: starts the compile on a background thread
    deadline = time.time() + 180
    while not sc.specialized()[0] and time.time() < deadline:
        assert sc.spec_pending, "the library decided not to specialise config 4"
        time.sleep(0.1)
    assert sc.specialized()[0], "NVRTC compile of the camera-ray kernel did not finish in 180 s"
    a, _ = sc.render(device.default_params(160, 90, 32, seed=21))
    ready, ms, passes, total = sc.specialized()
    assert ready and sc.spec_camera_only and passes == total > 0 and ms > 0
    i, _ = device.DeviceScene(flat).render(device.default_params(160, 90, 32, seed=21, flags=capi.PTB_FLAG_INTERPRET))
    assert np.allclose(a, i, rtol=1e-4, atol=1e-5)
