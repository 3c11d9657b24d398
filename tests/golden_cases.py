"""The seeded inputs behind tests/golden/*.npz: one definition used by the generator
(tests/golden/make_golden.py, reference side) and by the tests (oracle / CUDA side)."""
import math

import numpy as np

from ptb200 import scenes

MAX_SPANS = 12
FUZZ_CASES = [(seed, 3 + seed % 7) for seed in range(16)]
FUZZ_RAYS = 600
TRACE_DEPTH = 3
TRACE_SEED = 4242
C3_RAYS = 4000
# every 20th pixel of the 1920x1080 image plane: the author's own 96x54-ish preview scale
LATTICE = (1920, 1080, 10, 10, 20, 96, 54)
# configs 4 and 5 are 3840x2160 (BASELINE.json): every 40th pixel of THAT image plane, camera of src/test.cpp:450 at 4K
LATTICE_4K = (3840, 2160, 20, 20, 40, 96, 54)
RENDER_SPP = 8
RENDER_SEED = 20260401
RENDER_CASES = ["c1", "c2", "c4", "c5"]
# per case: (lattice, reference camera samples per pixel, GPU samples per pixel).  A reference camera sample carries
# ~10^4 scatter rays, a GPU sample one: c5's diffuse surfaces see a 1.5-degree sun (3.4e-4 of the hemisphere), so it
# takes 64 reference samples (~220 sun hits per pixel) and 2^16 GPU samples for both sides' noise to fall to ~3 %.
RENDER_SETUP = {
    "c1": (LATTICE, 8, 4096),
    "c2": (LATTICE, 8, 4096),
    "c4": (LATTICE_4K, 16, 16384),
    "c5": (LATTICE_4K, 64, 65536),
}
KAT_N = 512
KAT_SEED = 977


def fuzz_scene(b, seed, leaves):
    return scenes.random_tree(b, seed, leaves)


def fuzz_rays(seed):
    return scenes.fuzz_rays(FUZZ_RAYS, 1000 + seed)


def small_images():
    """Reduced-size stand-ins so golden generation and tests stay fast; same code path."""
    return [scenes.pattern_image(160, 120, 11), scenes.pattern_image(320, 377, 12)]


def small_env():
    return scenes.env_latlong(800, 400)


def render_scene(b, name, small=False):
    """The scene of an image config.  c4 / c5 with the FULL-SIZE assets bench.py times (640x480 and 1280x1509
    textures, 3200x1600 environment) unless small (quick property tests that do not compare with a golden)."""
    if name == "c1":
        return scenes.config1(b, env=small_env())
    if name == "c2":
        return scenes.config2(b)
    if name == "c4":
        return scenes.config4(b, images=small_images()) if small else scenes.config4(b)
    if name == "c5":
        return scenes.config5(b, env=small_env(), images=small_images()) if small else scenes.config5(b)
    raise KeyError(name)


def kat_vectors():
    rs = np.random.RandomState(31)
    v = rs.normal(size=(256, 3)).astype(np.float32) * rs.choice([0.01, 1.0, 2400.0], (256, 1)).astype(np.float32)
    n = rs.normal(size=(256, 3)).astype(np.float32)
    ior = rs.choice([1.0 / 1.3, 1.3, 1.0, 2.419, 1 / 2.419, 0.0005, 2000.0], 256).astype(np.float32)
    v[0] = 0
    n[1] = 0
    return v, n, ior


def kat_positions():
    rs = np.random.RandomState(32)
    p = rs.uniform(-3, 3, (400, 3)).astype(np.float32)
    p[0] = 0
    p[1] = (0, 0, -1)   # mirror-ball pole
    p[2] = (1, 1, 0.5)  # cube-map tie |x| == |y|
    p[3] = (2, 0.5, 2)  # cube-map tie |x| == |z|
    p[4] = (0.5, 1.0, 1.0)
    p[5] = (1.0, 2.0, 0.25)  # integer coordinates: frac == 0 -> 1 - frac(y) == 1 -> row h -> black (Q8)
    return p


def kat_textures(b):
    img = b.image(scenes.pattern_image(37, 23, 5))
    faces = [b.image(scenes.pattern_image(16 + i, 16 + i, 20 + i)) for i in range(6)]
    rot = b.concat(b.rotate((1, 2, 3), 0.7), b.translate(0.1, -0.2, 0.3))
    return {
        "const": b.color(0.25, 0.5, 0.75),
        "image": b.tex_image(img),
        "image_alpha": b.tex_image_alpha(img),
        "skybox": b.tex_skybox(faces),
        "skybox_alpha": b.tex_skybox(faces, alpha=True),
        "affine_image": b.tex_transformed(rot, b.tex_image(img)),
        "multiply": b.tex_multiply((0.9, 0.6, 0.3), b.tex_image(img)),
        "log": b.tex_log(b.tex_image(img)),
        "mirrorball": b.tex_mirrorball(b.tex_image(img)),
        "spherical": b.tex_spherical(b.tex_image(img)),
        "env": b.tex_transform(b.rotate_x(2 * math.pi / 4), b.tex_multiply((0.01, 0.01, 0.01), b.tex_spherical(b.tex_image(img)))),
        "log_alpha": b.tex_log(b.tex_transformed(rot, b.tex_image_alpha(img))),
        "moved_const": b.tex_transform(rot, b.color(0.1, 0.2, 0.3)),
        "double_affine": b.tex_transform(rot, b.tex_transformed(b.scale(0.5, 2.0, 1.0), b.tex_multiply((2, 2, 2), b.tex_image(img)))),
    }


def kat_matrices(b):
    rot = b.rotate((1, 2, 3), 0.7)
    m = b.concat(b.concat(b.scale(0.5, 1.5, 2.0), rot), b.translate(1, -2, 3))
    return {
        "rotate": rot, "rotate_x": b.rotate_x(2 * math.pi / 4), "concat": m, "inverse": b.inverse(m), "det": [b.determinant(m)] * 12,
        "apply": b.apply(m, (0.3, -0.7, 1.1)) + b.apply(m, (0.3, -0.7, 1.1), False) + [0] * 6,
    }
