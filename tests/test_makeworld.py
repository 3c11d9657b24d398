"""INTEGRATION.md Option A, tested: the reference's own scene-building code — src/test.cpp:52-145, unionArray /
makeLens / makeSkySphericalCoordinates / makeWorld, extracted at test time into a temporary directory and
compiled UNCHANGED against path-trace_b200/host/ptb200_scene.hpp — builds, runs, and flattens to the same
program, primitives, materials and texture programs as ptb200/scenes.py::config1 (BASELINE config 1), which the
parity and bench harness use.  Needs /root/reference (skipped on the GPU box)."""
import os
import subprocess

import numpy as np
import pytest

from ptb200 import capi, device, scenes
from ptb200.builder import product_builder

REF = "/root/reference"


def dump_words(name, ptr, n_bytes):
    import ctypes as C
    raw = C.string_at(ptr, n_bytes) if n_bytes else b""
    return name + "".join(" %08x" % v for v in np.frombuffer(raw, "<u4"))


@pytest.mark.skipif(not os.path.isdir(REF + "/src"), reason="/root/reference absent")
def test_reference_makeworld_compiles_unchanged_and_flattens_like_config1(tmp_path):
    import ctypes as C
    inc = tmp_path / "makeworld_extract.inc"
    lines = open(REF + "/src/test.cpp").read().splitlines(True)
    assert lines[51].startswith("Object *unionArray") and lines[106].startswith("Object *makeWorld") and lines[144].startswith("}")
    inc.write_text("".join(lines[51:145]))  # src/test.cpp:52-145
    exe = tmp_path / "makeworld_main"
    cmd = ["/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++", "-std=c++17", "-O1", "-w", "-I" + os.path.join(capi.PKG_ROOT, "host"), "-I" + os.path.join(capi.REPO_ROOT, "include"),
           '-DMAKEWORLD_INC="%s"' % inc, os.path.join(capi.REPO_ROOT, "tests", "cpp", "makeworld_main.cpp"), "-o", str(exe), "-L" + capi.PKG_ROOT, "-lptb200", "-Wl,-rpath," + capi.PKG_ROOT]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-3000:]
    # the two files makeWorld() opens in its working directory (src/test.cpp:121,123): the reference's own test2.hdr,
    # and a synthetic lat-long map written by OUR writeHDR in place of the missing Serpentine_Valley_3k.hdr
    b = product_builder()
    os.symlink(REF + "/test2.hdr", tmp_path / "test2.hdr")
    b.write_hdr(scenes.env_latlong(64, 32), str(tmp_path / "Serpentine_Valley_3k.hdr"))
    r = subprocess.run([str(exe)], capture_output=True, text=True, cwd=tmp_path)
    assert r.returncode == 0, r.stdout + r.stderr
    got = r.stdout.strip().splitlines()
    # the same scene through the Python builder: config1 with the pixels the C++ side read from that file
    env = b.image_pixels(b.image_load(str(tmp_path / "Serpentine_Valley_3k.hdr")))
    flat = device.Flat(scenes.config1(b, env=env))
    d = flat.desc
    want = ["counts %d %d %d %d %d %d %d %d %d" % (d.n_prims, d.n_xforms, d.n_chains, d.n_chain_xf, d.n_ops, d.n_materials, d.n_texops, d.n_texdata, d.n_images),
            dump_words("prims", d.prims, 32 * d.n_prims), dump_words("xforms", d.xforms, 96 * d.n_xforms), dump_words("chains", d.chains, 8 * d.n_chains),
            dump_words("chain_xf", d.chain_xf, 4 * d.n_chain_xf), dump_words("ops", d.ops, 4 * d.n_ops), dump_words("materials", d.materials, 32 * d.n_materials),
            dump_words("texops", d.texops, 8 * d.n_texops), dump_words("texdata", d.texdata, 4 * d.n_texdata)]
    for i in range(d.n_images):
        im = d.images[i]
        px = np.ctypeslib.as_array(im.rgba, shape=(im.height * im.width * 4,))
        want.append("image %d %d %.17g" % (im.width, im.height, float(px.astype(np.float64).sum())))
    assert d.n_prims == 13 and d.n_ops == 25  # 3 spheres + 6 sky planes + Intersection(S, Union(P, S)) + lens(S, S); 12 operators
    assert got == want
