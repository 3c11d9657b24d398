"""The oracle (oracle/pt_oracle.cpp, our CPU restatement) is pinned bit-for-bit against the
reference: always against the committed golden vectors generated from the reference, and
live against oracle/_ref/libptref.so where that was built."""
import ctypes as C
import os

import numpy as np
import pytest

import common
import golden_cases as gc
from ptb200 import device, scenes
from ptb200.builder import product_builder

SPANS = np.load(os.path.join(common.GOLDEN, "spans.npz"))
RENDERS = np.load(os.path.join(common.GOLDEN, "renders.npz"))
KAT = np.load(os.path.join(common.GOLDEN, "kat.npz"))
SPAN_FIELDS = ["start", "end", "start_normal", "end_normal", "start_material", "end_material"]


def same_bits(a, b):
    return np.ascontiguousarray(a).tobytes() == np.ascontiguousarray(b).tobytes()


def assert_spans_equal(spans, counts, gold_spans, gold_counts):
    assert np.array_equal(counts, gold_counts)
    valid = np.arange(spans.shape[1])[None, :] < np.minimum(counts, spans.shape[1])[:, None]
    for f in SPAN_FIELDS:
        assert same_bits(spans[f][valid], gold_spans[f][valid]), f


@pytest.mark.parametrize("seed,leaves", gc.FUZZ_CASES)
def test_fuzz_spans_and_trace_match_reference_golden(seed, leaves):
    b = product_builder()
    b.materials_reset()
    flat = device.Flat(gc.fuzz_scene(b, seed, leaves))
    o, d = gc.fuzz_rays(seed)
    hits, spans = common.orc_intersect(flat, o, d, gc.MAX_SPANS)
    assert_spans_equal(spans, hits["n_spans"], SPANS["fuzz%d_spans" % seed], SPANS["fuzz%d_counts" % seed])
    rgb = common.orc_trace_rays(flat, o, d, gc.TRACE_DEPTH, gc.TRACE_SEED)
    assert same_bits(rgb, SPANS["fuzz%d_rgb" % seed])


def test_config3_spans_match_reference_golden():
    b = product_builder()
    b.materials_reset()
    flat = device.Flat(scenes.config3(b))
    o, d = scenes.config3_rays(gc.C3_RAYS)
    hits, spans = common.orc_intersect(flat, o, d, gc.MAX_SPANS)
    assert_spans_equal(spans, hits["n_spans"], SPANS["c3_spans"], SPANS["c3_counts"])


def test_difference_quirk_q4_inverted_span():
    """A starts inside B and outlives it -> the reference emits [A.start, B.start] (inverted)."""
    b = product_builder()
    world = b.difference(b.sphere((0, 0, -5), 1, b.material()), b.sphere((0, 0, -4.5), 1, b.material()))
    hits, spans = common.orc_intersect(device.Flat(world), [[0, 0, 0]], [[0, 0, -1]], 4)
    assert hits["n_spans"][0] == 1 and spans["start"][0, 0] == 4.0 and spans["end"][0, 0] == 3.5
    assert hits["t"][0] == 4.0 and hits["prim"][0] == 0


@pytest.mark.parametrize("name", ["c1", "c4"])
def test_render_rows_match_reference_golden(name):
    """Same DefaultRandomEngine stream -> identical pixels (first rows of the golden lattice)."""
    b = product_builder()
    b.materials_reset()
    flat = device.Flat(gc.render_scene(b, name))
    (W, H, x0, y0, step, nx, ny), ref_spp, _ = gc.RENDER_SETUP[name]
    rows = 2 if name == "c1" else 1
    rgb, info = common.orc_render_mt(flat, W, H, x0, y0, step, nx, rows, ref_spp, 16, gc.RENDER_SEED, 2)
    assert same_bits(rgb, RENDERS[name][:rows])


def test_rng_and_vector_known_answers():
    lib = common.oracle()
    n = gc.KAT_N
    u32 = np.zeros(n, np.uint32)
    lib.orc_rng_u32(gc.KAT_SEED, n, u32.ctypes.data)
    uni = np.zeros(n, np.float32)
    lib.orc_rng_uniform(gc.KAT_SEED, n, -1.0, 1.0, uni.ctypes.data)
    ball = np.zeros((n, 3), np.float32)
    lib.orc_rng_ball(gc.KAT_SEED, n, 1.0, 0.0, ball.ctypes.data)
    assert same_bits(u32, KAT["rng_u32"]) and same_bits(uni, KAT["rng_uniform"]) and same_bits(ball, KAT["rng_ball"])
    v, nrm, ior = gc.kat_vectors()
    out = np.zeros(3, np.float32)
    for i in range(len(v)):
        lib.orc_vec_reflect(v[i].ctypes.data, nrm[i].ctypes.data, out.ctypes.data)
        assert same_bits(out, KAT["reflect"][i])
        lib.orc_vec_refract(v[i].ctypes.data, float(ior[i]), nrm[i].ctypes.data, out.ctypes.data)
        assert same_bits(out, KAT["refract"][i])
        s = np.float32(lib.orc_vec_refract_strength(v[i].ctypes.data, float(ior[i]), nrm[i].ctypes.data))
        assert same_bits(s, KAT["refract_strength"][i])


def test_texture_known_answers():
    b = product_builder()
    pos = gc.kat_positions()
    for name, tex in gc.kat_textures(b).items():
        flat = device.Flat(texture=tex)
        rgb, f = common.orc_tex_eval(flat, flat.desc.materials[0].tex[0], pos)
        assert same_bits(rgb, KAT["tex_%s_rgb" % name]), name
        assert same_bits(f, KAT["tex_%s_f" % name]), name


@pytest.mark.skipif(not common.have_ref(), reason="oracle/_ref not built (no /root/reference here)")
def test_live_reference_fuzz():
    b, rb = product_builder(), common.ref_builder()
    for seed in range(100, 112):
        b.materials_reset()
        rb.materials_reset()
        flat = device.Flat(scenes.random_tree(b, seed, 3 + seed % 8))
        rw = scenes.random_tree(rb, seed, 3 + seed % 8)
        o, d = scenes.fuzz_rays(1500, seed)
        hits, spans = common.orc_intersect(flat, o, d, gc.MAX_SPANS)
        rspans, rcounts = rb.trace_spans(rw, o, d, gc.MAX_SPANS)
        assert_spans_equal(spans, hits["n_spans"], rspans, rcounts)
        assert same_bits(common.orc_trace_rays(flat, o[:300], d[:300], 4, seed), common.ref_trace_rays(common.ref(), rw, o[:300], d[:300], 4, seed))
