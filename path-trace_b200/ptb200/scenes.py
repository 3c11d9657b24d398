"""Synthetic scenes of BASELINE.json's configs, written once against the
Builder API so the GPU path, the oracle and the unmodified reference all get the same scene.

All random parameters come from numpy RandomState(0xB200 + config) so every side consumes
identical numbers.  The reference's own HDR assets are either missing
(Serpentine_Valley_3k.hdr, stpeters_probe.hdr: .MISSING_LARGE_BLOBS) or do not travel to the
GPU box (test.hdr, test2.hdr live under /root/reference), so images are synthetic, with the
sizes of the files they stand in for; every report says "synthetic".
"""
import math

import numpy as np


# ---- synthetic images -----------------------------------------------------------------------------
def env_latlong(width=3200, height=1600):
    """Analytic lat-long sky: horizon-to-zenith gradient, warm ground, a small very bright sun."""
    v, u = np.meshgrid((np.arange(height) + 0.5) / height, (np.arange(width) + 0.5) / width, indexing="ij")
    theta = (u - 0.5) * 2 * np.pi
    phi = (0.5 - v) * np.pi  # row 0 = up
    d = np.stack([np.cos(phi) * np.cos(theta), np.cos(phi) * np.sin(theta), np.sin(phi)], -1)
    up = np.clip(d[..., 2], 0, 1)[..., None]
    sky = np.array([60.0, 90.0, 160.0]) * (0.35 + 0.65 * up) + np.array([120.0, 110.0, 90.0]) * (1 - up) ** 4
    ground = np.array([40.0, 32.0, 24.0]) * np.ones_like(up)
    rgb = np.where(d[..., 2:3] >= 0, sky, ground)
    sun_dir = np.array([0.45, -0.35, 0.82])
    sun_dir /= np.linalg.norm(sun_dir)
    cosang = d @ sun_dir
    rgb = rgb + np.array([60000.0, 52000.0, 40000.0]) * (cosang > math.cos(math.radians(1.5)))[..., None]
    rgb = rgb + np.array([900.0, 700.0, 400.0]) * np.clip((cosang - 0.9) / 0.1, 0, 1)[..., None] ** 6
    out = np.ones((height, width, 4), np.float32)
    out[..., :3] = rgb
    return out


def pattern_image(width, height, seed):
    """Smooth colourful pattern + checker, values in [0.05, 2.5] (stands in for test*.hdr)."""
    rs = np.random.RandomState(seed)
    y, x = np.meshgrid(np.arange(height) / height, np.arange(width) / width, indexing="ij")
    out = np.ones((height, width, 4), np.float32)
    for c in range(3):
        f = rs.uniform(1.0, 6.0, 4)
        ph = rs.uniform(0, 2 * np.pi, 4)
        out[..., c] = 0.9 + 0.4 * np.sin(2 * np.pi * f[0] * x + ph[0]) * np.cos(2 * np.pi * f[1] * y + ph[1]) + 0.3 * np.sin(2 * np.pi * (f[2] * x + f[3] * y) + ph[2])
    checker = ((np.floor(x * 16) + np.floor(y * 12)) % 2).astype(np.float32)
    out[..., :3] *= (0.55 + 0.9 * checker)[..., None]
    out[..., :3] = np.clip(out[..., :3], 0.05, 2.5)
    out[..., 3] = 0.25 + 0.75 * checker
    return out.astype(np.float32)


# ---- shared pieces -----------------------------------------------------------------------------------
def sky_planes(b, sky, extent=200.0):
    """Six emissive far planes, the pattern of reference src/test.cpp:134-140."""
    return [b.plane(n, extent, sky) for n in ((0, 0, -1), (0, 0, 1), (0, -1, 0), (0, 1, 0), (1, 0, 0), (-1, 0, 0))]


def constant_sky(b, value=1.0):
    return b.material(b.color(0), b.color(0), b.color(value))


def env_sky(b, env_rgba, scale=0.01):
    """transform(rotateX(pi/2), makeSkySphericalCoordinates(file, Color(scale))), src/test.cpp:102-105,123."""
    img = b.image(env_rgba)
    inner = b.material(b.color(0), b.color(0), b.tex_multiply((scale, scale, scale), b.tex_spherical(b.tex_image(img))))
    return b.material_transform(b.rotate_x(2 * math.pi / 4), inner)


# ---- config 2: diffuse sphere on a plane under a constant sky ----------------------------------------
def config2(b):
    white = b.material(b.color(0.8), b.color(1))
    sky = constant_sky(b, 1.0)
    objs = [b.sphere((0, 0, -4), 1.0, white), b.plane_at((0, 1, 0), (0, -1, 0), white)] + sky_planes(b, sky)
    return b.union_array(objs)


# ---- config 3: 16-primitive Union/Difference tree for the ray-batch microbench ------------------------
def config3(b, seed=0xB200 + 3):
    rs = np.random.RandomState(seed)
    mats = [b.material(b.color(0.5 + 0.03 * i), b.color(1)) for i in range(16)]
    centres = rs.uniform(-2, 2, (12, 3))
    radii = rs.uniform(0.4, 1.0, 12)
    spheres = [b.sphere(tuple(float(v) for v in centres[i]), float(radii[i]), mats[i]) for i in range(12)]
    planes = []
    for i in range(4):
        u = rs.normal(size=3)
        u /= np.linalg.norm(u)
        planes.append(b.plane(tuple(float(-v) for v in u), 1.5, mats[12 + i]))
    return b.difference(b.difference(b.union_array(spheres[:8]), b.union_array(spheres[8:])), b.union_array(planes))


def config3_rays(n, seed=0xB200 + 30):
    """origin in a ball of radius 4; direction classes by i mod 4."""
    rs = np.random.RandomState(seed)

    def ball(k, radius):
        v = rs.normal(size=(k, 3))
        v /= np.linalg.norm(v, axis=1, keepdims=True)
        return v * (radius * rs.uniform(0, 1, (k, 1)) ** (1 / 3))

    o = ball(n, 4.0).astype(np.float32)
    d = ball(n, 1.0)
    d[np.linalg.norm(d, axis=1) < 1e-3] = (0.0, 0.0, 1.0)
    unit = d / np.linalg.norm(d, axis=1, keepdims=True)
    cls = np.arange(n) % 4
    d = np.where((cls == 2)[:, None], unit, d)
    d = np.where((cls == 3)[:, None], unit * 2400.0, d)
    return o, d.astype(np.float32)


# ---- config 4: deep CSG stress, 64 transformed primitives, textured materials --------------------------
def config4_materials(b, images=None):
    img_a = pattern_image(640, 480, 11) if images is None else images[0]
    img_b = pattern_image(1280, 1509, 12) if images is None else images[1]
    ia, ib = b.image(img_a), b.image(img_b)
    return [
        b.material(b.tex_transformed(b.scale(0.25), b.tex_image(ia)), b.color(1)),
        b.material(b.tex_multiply((0.9, 0.6, 0.3), b.tex_image(ib)), b.color(0.3)),
        b.material(b.color(0.99), b.color(0)),                                                # mirror, src/test.cpp:118
        b.material(b.color(0.7), b.color(0), b.color(0), b.color(0.9), 1.3, b.color(1)),      # glass, src/test.cpp:116
    ]


def config4(b, seed=0xB200 + 4, images=None, sky=None):
    rs = np.random.RandomState(seed)
    mats = config4_materials(b, images)
    sky = constant_sky(b, 1.0) if sky is None else sky

    def placed(obj):
        axis = rs.normal(size=3)
        axis /= np.linalg.norm(axis)
        rot = b.rotate(tuple(float(v) for v in axis), float(rs.uniform(0, 2 * np.pi)))
        sc = b.scale(*[float(v) for v in rs.uniform(0.5, 1.5, 3)])
        tr = b.translate(float(rs.uniform(-3, 3)), float(rs.uniform(-3, 3)), float(rs.uniform(-12, -6)))
        place = b.concat(b.concat(sc, rot), tr)  # object -> world
        return b.transformed(b.inverse(place), obj)  # TransformedObject wants world -> object (Q7)

    groups = []
    k = 0
    for g in range(8):
        spheres = []
        for _ in range(6):
            spheres.append(placed(b.sphere((0, 0, 0), float(rs.uniform(0.5, 1.0)), mats[k % 4])))
            k += 1
        planes = []
        for _ in range(2):
            n = rs.normal(size=3)
            n /= np.linalg.norm(n)
            planes.append(placed(b.plane(tuple(float(v) for v in n), float(rs.uniform(0.1, 0.5)), mats[k % 4])))
            k += 1
        groups.append(b.difference(b.union_array(spheres), b.union_array(planes)))
    return b.union_array(groups + sky_planes(b, sky))


# ---- config 5: HDR-environment-lit glossy scene with image textures ------------------------------------
def config5(b, env=None, images=None):
    env = env_latlong() if env is None else env
    sky = env_sky(b, env, 0.01)
    img_a = pattern_image(640, 480, 11) if images is None else images[0]
    ia = b.image(img_a)
    floor = b.material(b.tex_transformed(b.scale(0.2), b.tex_image(ia)), b.color(1))
    glossy = [b.material(b.tex_multiply((0.9, 0.85, 0.8), b.tex_transformed(b.scale(0.5), b.tex_image(ia))), b.color(s)) for s in (0.0, 0.1, 0.3)]
    glass = b.material(b.color(0.7), b.color(0), b.color(0), b.color(0.9), 1.3, b.color(1))
    diffuse = b.material(b.color(0.8), b.color(1))
    objs = [
        b.sphere((-2.2, 0, -6), 1.0, glossy[0]),
        b.sphere((0, 0, -6.5), 1.0, glossy[1]),
        b.sphere((2.2, 0, -6), 1.0, glossy[2]),
        b.sphere((-1.0, -0.55, -4.2), 0.45, glass),
        b.sphere((1.1, -0.6, -4.4), 0.4, diffuse),
        b.plane_at((0, 1, 0), (0, -1, 0), floor),
    ] + sky_planes(b, sky)
    return b.union_array(objs)


# ---- config 1: the reference's default scene, makeWorld() (reference src/test.cpp:107-145) -------------
def make_lens(b, position, orientation, radius, sphere_radius, mat):
    """makeLens, reference src/test.cpp:66-72 (float arithmetic as there)."""
    f = np.float32
    dist = f(np.sqrt(f(f(sphere_radius) * f(sphere_radius)) - f(f(radius) * f(radius))))
    o = np.array(orientation, np.float32)
    mag = f(np.sqrt(f(f(f(o[0] * o[0]) + f(o[1] * o[1])) + f(o[2] * o[2]))))
    o = (o / mag).astype(np.float32)
    p = np.array(position, np.float32)
    c1 = (p + (o * dist).astype(np.float32)).astype(np.float32)
    c2 = (p - (o * dist).astype(np.float32)).astype(np.float32)
    return b.intersection(b.sphere(tuple(float(v) for v in c1), sphere_radius, mat), b.sphere(tuple(float(v) for v in c2), sphere_radius, mat))


def config1(b, env=None):
    env = env_latlong() if env is None else env
    emit_w = b.material(b.color(0), b.color(0), b.color(2))
    diffuse_white = b.material(b.color(0.8), b.color(1))
    glass = b.material(b.color(0.7), b.color(0), b.color(0), b.color(0.9), 1.3, b.color(1))
    sky = env_sky(b, env, 0.01)
    shift = b.translate(-1, 0, 4)
    objs = [
        b.sphere((1, 0, -4), 0.2, b.material_transform(shift, diffuse_white)),
        b.intersection(b.sphere((1, 0, -4), 0.2 * 5, glass),
                       b.union(b.plane_at((-1, 0, -0.7), (1, 0, -4), glass), b.sphere((1, 0, -4), 0.2, b.material_transform(shift, emit_w)))),
        b.sphere((-1, 0, -4), 0.2, diffuse_white),
    ] + sky_planes(b, sky) + [make_lens(b, (-2.5 / 4, 0, -2.5), (-1, 0, -4), 0.5, 1.0, glass)]
    return b.union_array(objs)


# ---- fuzz scenes: random CSG trees over spheres / planes, optional nested transforms -------------------
def random_tree(b, seed, n_leaves=6, transforms=True, emissive=True):
    """Random Union/Difference/Intersection tree.  Every primitive gets its own material (so a
    material index identifies the primitive on the reference side, which has no primitive ids).
    Primitives crowd the z axis so axis-parallel rays produce every merge branch, including
    shared endpoints (quirk Q5) and the inverted Difference span (quirk Q4)."""
    rs = np.random.RandomState(seed)
    grid = [-2.0, -1.5, -1.0, -0.5, 0.0, 0.5, 1.0, 1.5, 2.0]

    def material(i):
        e = b.color(0.1 * (i + 1), 0.05 * (i + 1), 1.0 / (i + 1)) if emissive else b.color(0)
        kind = rs.randint(4)
        if kind == 0:
            return b.material(b.color(0.8), b.color(1), e)
        if kind == 1:
            return b.material(b.color(0.9), b.color(float(rs.choice([0.0, 0.05, 0.3]))), e)
        if kind == 2:
            return b.material(b.color(0.7), b.color(0), e, b.color(0.9), 1.3, b.color(1))
        return b.material(b.color(0.6, 0.7, 0.8), b.color(0.5), e, b.color(0.5), 1.5, b.color(0.4))

    def leaf(i):
        m = material(i)
        if rs.rand() < 0.75:
            c = (float(rs.choice([0.0, 0.0, 0.25])), float(rs.choice([0.0, 0.0, -0.25])), float(rs.choice(grid)) - 6.0)
            obj = b.sphere(c, float(rs.choice([0.5, 0.75, 1.0, 1.25])), m)
        else:
            n = [(0, 0, 1), (0, 0, -1), (0, 1, 0), (1, 0, 0.5)][rs.randint(4)]
            obj = b.plane_at(n, (0.0, 0.0, float(rs.choice(grid)) - 6.0), m)
        if transforms and rs.rand() < 0.4:
            obj = b.transformed(random_matrix(b, rs), obj)
        return obj

    def build(lo, hi):
        if hi - lo == 1:
            return leaf(lo)
        mid = lo + 1 + rs.randint(hi - lo - 1)
        a, c = build(lo, mid), build(mid, hi)
        node = [b.union, b.difference, b.intersection][rs.randint(3)](a, c)
        if transforms and rs.rand() < 0.2:
            node = b.transformed(random_matrix(b, rs), node)
        return node

    return build(0, n_leaves)


def random_matrix(b, rs):
    """A mild world->object matrix: small rotation about a random axis, near-unit scale, small shift."""
    axis = rs.normal(size=3)
    axis /= np.linalg.norm(axis)
    m = b.rotate(tuple(float(v) for v in axis), float(rs.uniform(-0.3, 0.3)))
    m = b.concat(m, b.scale(*[float(v) for v in rs.choice([1.0, 1.0, 0.8, 1.25], 3)]))
    return b.concat(m, b.translate(*[float(v) for v in rs.choice([0.0, 0.0, 0.25, -0.5], 3)]))


def fuzz_rays(n, seed):
    """Half axis-parallel rays down -z through the crowded axis (exact shared endpoints), half
    random rays from in and around the cluster; mixed direction magnitudes (quirk Q1)."""
    rs = np.random.RandomState(seed)
    o = np.zeros((n, 3), np.float32)
    d = np.zeros((n, 3), np.float32)
    half = n // 2
    o[:half] = np.stack([rs.choice([0.0, 0.25, 0.5], half), rs.choice([0.0, -0.25, 0.5], half), rs.choice([0.0, -4.0, -5.5, -6.0, -7.25], half)], 1)
    d[:half] = np.stack([np.zeros(half), np.zeros(half), rs.choice([-1.0, 1.0, -2.0, -0.5, -2400.0], half)], 1)
    o[half:] = rs.uniform(-3, 3, (n - half, 3)) + np.array([0, 0, -6.0])
    v = rs.normal(size=(n - half, 3))
    d[half:] = v * rs.choice([0.01, 1.0, 1.0, 100.0], (n - half, 1))
    return o, d
