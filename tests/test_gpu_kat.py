"""Known answers and statistics of the small device functions the shade and camera stages are made of
(SURVEY.md §8 rows A3, A5), through the C ABI:

* ptb200_vector_eval — vreflect / vrefract / refractStrength of csrc/ptb_shade.cuh against the reference's
  Vector3D::reflect / refract / refractStrength (include/vector3d.h:186-214) on the committed KAT vectors
  (tests/golden/kat.npz, generated from the unmodified reference), including the edge cases v = 0, n = 0,
  ior 5e-4 / 2000 (outside [eps, 1/eps]) and total internal reflection;
* ptb200_rng_words — the counter-based generators that replace the reference's shared LCG
  (DefaultRandomEngine, include/path-trace.h:21-54): the camera-jitter hash and Philox4x32-10.  Parity with the
  LCG is statistical: uniformity (chi-square), independence of the two jitter coordinates, of neighbouring
  pixels and of consecutive samples (correlations), per-pixel stratification-free variance, no collisions, and
  the same first two moments as the reference's own uniform stream (KAT rng_uniform)."""
import ctypes as C
import os

import numpy as np
import pytest

import common
import golden_cases as gc
from ptb200 import capi

pytestmark = pytest.mark.gpu
KAT = np.load(os.path.join(common.GOLDEN, "kat.npz"))


def vector_eval(op, v, n, ior):
    v = np.ascontiguousarray(v, np.float32)
    n = np.ascontiguousarray(n, np.float32)
    ior = np.ascontiguousarray(ior, np.float32)
    out = np.zeros(len(v) if op == 2 else (len(v), 3), np.float32)
    capi.check(capi.product().ptb200_vector_eval(op, v.ctypes.data, n.ctypes.data, ior.ctypes.data, len(v), out.ctypes.data))
    return out


def test_reflect_refract_known_answers_on_the_device():
    v, nrm, ior = gc.kat_vectors()
    refl, refr, strength = vector_eval(0, v, nrm, ior), vector_eval(1, v, nrm, ior), vector_eval(2, v, nrm, ior)
    want_refl, want_refr, want_s = KAT["reflect"], KAT["refract"], KAT["refract_strength"]
    # reflect keeps the input magnitude (2400 for camera-like rays): relative to |v|
    scale = np.maximum(np.linalg.norm(v, axis=1, keepdims=True), 1e-30)
    assert (np.abs(refl - want_refl) <= 4e-6 * scale + 1e-30).all(), np.abs((refl - want_refl) / scale).max()
    # refract returns unit vectors, or exactly (0,0,0) for v = 0, n = 0, ior outside [eps, 1/eps] and TIR
    zero = (want_refr == 0).all(1)
    assert zero.sum() >= 20 and (refr[zero] == 0).all()
    # Both go through k = 1 - ior^2 (1 - c^2): sqrt(k) in the refracted direction, k^(1/4) in the weight, so a
    # rounding difference dk ~ 4e-7 ior^2 in k (rsqrt normalisation, FMA) is amplified by 1 / (2 sqrt k) and
    # k^(-3/4) / 4 near total internal reflection (smallest k in the vectors: 7.3e-5; none sits on the k = 0 branch)
    v64, n64 = v.astype(np.float64), nrm.astype(np.float64)
    c = (v64 * n64).sum(1) / np.maximum(np.linalg.norm(v64, axis=1) * np.linalg.norm(n64, axis=1), 1e-300)
    k = np.maximum(1 - ior.astype(np.float64) ** 2 * (1 - c * c), 1e-12)
    dk = 4e-7 * np.maximum(1.0, ior.astype(np.float64) ** 2)
    tol_dir = 4e-6 + dk / (2 * np.sqrt(k))
    tol_s = 4e-6 + dk * 0.25 * k**-0.75
    err = np.abs(refr - want_refr).max(1)
    assert (err[~zero] <= tol_dir[~zero]).all(), (err[~zero] / tol_dir[~zero]).max()
    assert np.median(err[~zero]) <= 2e-6
    assert (np.abs(np.linalg.norm(refr[~zero], axis=1) - 1) < 1e-5).all()
    zs = want_s == 0
    assert (strength[zs] == 0).all() and zs.sum() >= 20
    err = np.abs(strength - want_s)
    assert (err[~zs] <= tol_s[~zs]).all(), (err[~zs] / tol_s[~zs]).max()
    assert np.median(err[~zs]) <= 2e-6
    # the documented edge cases, by index (tests/golden_cases.py: v[0] = 0, n[1] = 0)
    assert (refr[0] == 0).all() and strength[0] == 0 and (refr[1] == 0).all() and strength[1] == 0
    assert (refl[0] == 0).all()  # reflecting the zero vector
    assert np.allclose(refl[1], v[1])  # n = 0: normalize(0) = 0, reflect returns the input (vector3d.h:115-120,186-190)
    bad_ior = (ior < 1e-3) | (ior > 1e3)
    assert bad_ior.sum() >= 20 and (refr[bad_ior] == 0).all() and (strength[bad_ior] == 0).all()


def rng_words(which, seed, pixel, sample, block=None):
    pixel = np.ascontiguousarray(pixel, np.uint32)
    sample = np.ascontiguousarray(sample, np.uint32)
    out = np.zeros((len(pixel), 2 if which == 0 else 4), np.uint32)
    blk = np.ascontiguousarray(block, np.uint32) if block is not None else None
    capi.check(capi.product().ptb200_rng_words(which, seed, pixel.ctypes.data, sample.ctypes.data, blk.ctypes.data if blk is not None else None, len(pixel), out.ctypes.data))
    return out


def chi_square_uniform(u, bins=256):
    counts = np.bincount((u * bins).astype(np.int64).clip(0, bins - 1), minlength=bins)
    expect = len(u) / bins
    return ((counts - expect) ** 2 / expect).sum(), bins - 1


def corr(a, b):
    return abs(np.corrcoef(a, b)[0, 1])


@pytest.mark.parametrize("which", [0, 1])
def test_counter_based_generators_are_uniform_and_uncorrelated(which):
    W, H, S = 512, 256, 16  # a 512x256 window of a 4K frame, 16 consecutive global samples
    py, px, s = np.meshgrid(np.arange(H), np.arange(W), np.arange(S), indexing="ij")
    pixel = ((py + 700) * 3840 + px + 1500).ravel()
    sample = s.ravel()
    for seed in (0xB200, 1, 0xDEADBEEFCAFEF00D):
        w = rng_words(which, seed, pixel, sample)
        u = w.astype(np.float64) / 2.0**32
        n = len(u)
        # two (pixel, sample) keys drawing the same words: never for Philox (64 of its 128 bits compared: birthday
        # bound ~1e-7 for 2^21 draws); the jitter hash carries 32 bits of state, so its (x, y) pairs collide like
        # 32-bit random numbers — n^2 / 2^33 = 512 expected here — and no more often than that
        keys = w[:, 0].astype(np.uint64) << np.uint64(32) | w[:, 1].astype(np.uint64)
        dup = n - len(np.unique(keys))
        assert dup == 0 if which == 1 else dup < 2 * n * n / 2.0**33, dup
        for k in range(u.shape[1]):
            chi, dof = chi_square_uniform(u[:, k])
            assert chi < dof + 6 * np.sqrt(2 * dof), (which, seed, k, chi)  # 6 sigma
            assert abs(u[:, k].mean() - 0.5) < 5 / np.sqrt(12 * n) and abs(u[:, k].var() - 1 / 12) < 5 * np.sqrt(1 / 180 / n)
        bound = 5 / np.sqrt(n)
        # coordinates of one draw, neighbouring pixels (x and y), consecutive samples of a pixel
        assert corr(u[:, 0], u[:, 1]) < bound
        grid = u[:, 0].reshape(H, W, S)
        assert corr(grid[:, :-1].ravel(), grid[:, 1:].ravel()) < bound
        assert corr(grid[:-1].ravel(), grid[1:].ravel()) < bound
        assert corr(grid[..., :-1].ravel(), grid[..., 1:].ravel()) < bound
        assert corr(grid[:, :-1].ravel(), u[:, 1].reshape(H, W, S)[:, 1:].ravel()) < bound
        # the 16 samples of one pixel are plain independent uniforms (the reference jitters without strata):
        # their per-pixel sample variance averages to 1/12
        assert abs(grid.var(axis=2, ddof=1).mean() - 1 / 12) < 5 * np.sqrt(1 / 180 / (H * W))
        # 2-D uniformity of the jitter: 16x16 cells
        cells = np.bincount((u[:, 0] * 16).astype(int).clip(0, 15) * 16 + (u[:, 1] * 16).astype(int).clip(0, 15), minlength=256)
        chi = ((cells - n / 256) ** 2 / (n / 256)).sum()
        assert chi < 255 + 6 * np.sqrt(510), (which, seed, chi)
    # a different seed gives a different, uncorrelated stream
    a = rng_words(which, 5, pixel[:65536], sample[:65536])[:, 0] / 2.0**32
    b = rng_words(which, 6, pixel[:65536], sample[:65536])[:, 0] / 2.0**32
    assert corr(a, b) < 5 / np.sqrt(65536)


def test_philox_blocks_are_distinct_streams_and_match_the_reference_moments():
    n = 1 << 16
    pixel = np.full(n, 123456, np.uint32)
    sample = np.full(n, 7, np.uint32)
    w = rng_words(1, 0xB200, pixel, sample, np.arange(n))  # consecutive counter blocks of ONE path
    u = (w.astype(np.float64) / 2.0**32).ravel()
    chi, dof = chi_square_uniform(u)
    assert chi < dof + 6 * np.sqrt(2 * dof)
    assert corr(u[:-1], u[1:]) < 5 / np.sqrt(len(u))
    # same first two moments as the reference LCG's uniform stream on [-1, 1] (KAT rng_uniform, 512 draws)
    ref = KAT["rng_uniform"].astype(np.float64)
    ours = u[: len(ref)] * 2 - 1
    assert abs(ours.mean() - ref.mean()) < 5 * np.sqrt(2 / 3 / len(ref)) and abs(ours.var() - ref.var()) < 0.08
