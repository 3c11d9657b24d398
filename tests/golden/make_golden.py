"""Generates tests/golden/*.npz from the UNMODIFIED reference (oracle/_ref/libptref.so, built by
oracle/Makefile from /root/reference).  Run in the build container, where /root/reference exists:

    make -C oracle && python tests/golden/make_golden.py

The reference has no tests or golden vectors of its own; these are outputs of the
reference itself on seeded inputs, committed so the oracle and the CUDA path can be pinned on
machines where /root/reference is absent (the GPU box).
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "path-trace_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))

import common  # noqa: E402
import golden_cases as gc  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))


def main():
    lib = common.ref()
    rb = common.ref_builder()
    out = {}

    # 1. span dumps of fuzz trees + config 3
    for seed, leaves in gc.FUZZ_CASES:
        rb.materials_reset()
        world = gc.fuzz_scene(rb, seed, leaves)
        o, d = gc.fuzz_rays(seed)
        spans, counts = rb.trace_spans(world, o, d, gc.MAX_SPANS)
        out["fuzz%d_spans" % seed] = spans
        out["fuzz%d_counts" % seed] = counts
        out["fuzz%d_rgb" % seed] = common.ref_trace_rays(lib, world, o, d, gc.TRACE_DEPTH, gc.TRACE_SEED)
    rb.materials_reset()
    world = gc.scenes.config3(rb)
    o, d = gc.scenes.config3_rays(gc.C3_RAYS)
    spans, counts = rb.trace_spans(world, o, d, gc.MAX_SPANS)
    out["c3_spans"], out["c3_counts"] = spans, counts
    np.savez_compressed(os.path.join(OUT, "spans.npz"), **out)

    # 2. seeded renders of the image configs on a lattice of their image plane (gc.RENDER_SETUP: c1/c2 1920x1080,
    #    c4/c5 3840x2160 with the full-size assets bench.py times)
    out = {}
    for name in gc.RENDER_CASES:
        rb.materials_reset()
        world = gc.render_scene(rb, name)
        lattice, ref_spp, _ = gc.RENDER_SETUP[name]
        rgb, info = common.ref_render_mt(lib, world, *lattice, ref_spp, 16, gc.RENDER_SEED, os.cpu_count())
        out[name] = rgb
        out[name + "_rays"] = np.array([info["rays"]], np.uint64)
        print(name, "mean", rgb.mean(), "rays", info["rays"], "%.1fs" % info["seconds"])
    np.savez_compressed(os.path.join(OUT, "renders.npz"), **out)

    # 3. known answers of the helper functions
    out = {}
    n = gc.KAT_N
    u32 = np.zeros(n, np.uint32)
    lib.ref_rng_u32(gc.KAT_SEED, n, u32.ctypes.data)
    uni = np.zeros(n, np.float32)
    lib.ref_rng_uniform(gc.KAT_SEED, n, -1.0, 1.0, uni.ctypes.data)
    ball = np.zeros((n, 3), np.float32)
    lib.ref_rng_ball(gc.KAT_SEED, n, 1.0, 0.0, ball.ctypes.data)
    out.update(rng_u32=u32, rng_uniform=uni, rng_ball=ball)
    v, nrm, ior = gc.kat_vectors()
    out["reflect"] = np.array([rb.reflect(v[i], nrm[i]) for i in range(len(v))], np.float32)
    out["refract"] = np.array([rb.refract(v[i], float(ior[i]), nrm[i]) for i in range(len(v))], np.float32)
    out["refract_strength"] = np.array([rb.refract_strength(v[i], float(ior[i]), nrm[i]) for i in range(len(v))], np.float32)
    for name, tex in gc.kat_textures(rb).items():
        rgb, f = rb.tex_eval(tex, gc.kat_positions())
        out["tex_%s_rgb" % name], out["tex_%s_f" % name] = rgb, f
    mats = gc.kat_matrices(rb)
    for k, m in mats.items():
        out["mat_" + k] = np.array(m, np.float32)
    np.savez_compressed(os.path.join(OUT, "kat.npz"), **out)
    # 4. the reference's tile renderer and tile wire text (oracle/_ref/libptref_block.so) on the seeded lattice of
    #    tests/block_cases.py: digests of whole rows of tiles, and the socket text + screen pixels of two tiles
    import block_cases as bc
    blib = bc.block_lib()
    lat = bc.lattice()
    rgb, missing = bc.ref_frame(blib, lat)
    assert missing == 0
    out = {}
    for tr in bc.TILE_ROWS:
        y0 = tr * bc.BLOCK
        out["frame_rows_sha256_%d" % tr] = np.array(bc.digest(rgb[y0:min(bc.H, y0 + bc.BLOCK)]))
    for k, (x, y) in enumerate(bc.TEXT_TILES):
        first, _ = bc.ref_socket_text(blib, lat, x, y)
        out["tile_text_%d" % k] = np.frombuffer(first.encode(), np.uint8)
        tile = rgb[y:y + bc.BLOCK, x:x + bc.BLOCK]
        out["tile_rgba_%d" % k] = np.concatenate([tile, np.ones(tile.shape[:2] + (1,), np.float32)], 2)
    np.savez_compressed(os.path.join(OUT, "blocks.npz"), **out)
    print("golden vectors written to", OUT)


if __name__ == "__main__":
    main()
