"""The GENERATED register-resident evaluator (second-generation scene-specialised kernels: csrc/ptb_specialize.cu
prints it, csrc/ptb_csg_reg.cuh holds its list operations) against the interpreter, on the CPU: both are compiled
as host code through csrc/ptb_host_emu.h (tests/cpp/spec_emu_main.cpp) and run ray by ray.  They must agree bit
for bit — length of the root span list, every t, every tag, and the first boundary with t >= eps — on the
configs' scenes and on random Union / Difference / Intersection trees with nested transforms, for camera rays
(origin folded into per-leaf constants) and for general rays, including rays that make shared endpoints (quirk
Q5) and the inverted Difference span (quirk Q4).  The same comparison runs on the GPU in tests/test_specialize.py;
this one needs no GPU, so the generator is checked where it is developed."""
import os
import subprocess

import numpy as np
import pytest

from ptb200 import capi, device, scenes
from ptb200.builder import product_builder


def functions_only(src):
    """The generated source without its kernels: the evaluator function in namespace ptb."""
    a = src.index("namespace ptb\n{\n")
    b = src.index("} // namespace ptb\n#define PTB_SPEC_GEN 2")
    return src[a:b] + "} // namespace ptb\n"


def probe_text(flat, which):
    src, _ = device.specialize_probe(flat, which, compile=False)
    return src


def emulate(tmp_path, name, flat, rays_o, rays_d, origin0, cap=12):
    d = tmp_path / name
    d.mkdir(exist_ok=True)
    (d / "trace.inc").write_text(functions_only(probe_text(flat, 0)))
    # (both modules carry their own copy of the constant table and the shape functions: one program here, so rename)
    (d / "list.inc").write_text(functions_only(probe_text(flat, 1)).replace("ptbTab", "ptbTabList").replace("ptbShape", "ptbShapeList").replace("ptbCold", "ptbColdList").replace("specStageTable", "specStageTableList"))
    (d / "prog.hex").write_text(probe_text(flat, 2))
    rays = np.concatenate([np.asarray(rays_o, np.float32).reshape(-1, 3), np.asarray(rays_d, np.float32).reshape(-1, 3)], 1)
    rays.tofile(d / "rays.bin")
    exe = d / "emu"
    csrc = os.path.join(capi.PKG_ROOT, "csrc")
    cmd = ["/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++", "-std=c++17", "-O1", "-ffp-contract=off", "-w", "-I" + csrc, "-I" + os.path.join(capi.REPO_ROOT, "include"),
           '-DSPEC_LIST_INC="%s"' % (d / "list.inc"), '-DSPEC_TRACE_INC="%s"' % (d / "trace.inc"), os.path.join(capi.REPO_ROOT, "tests", "cpp", "spec_emu_main.cpp"), "-o", str(exe)]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-4000:]
    r = subprocess.run([str(exe), str(d / "prog.hex"), str(d / "rays.bin"), str(cap), "1" if origin0 else "0", "32"], capture_output=True, text=True)
    words = r.stdout.strip().splitlines()[-1].split()
    stats = dict(zip(words[0::2], [int(v) for v in words[1::2]]))
    assert r.returncode == 0 and stats["mismatches"] == 0, r.stdout[-2000:]
    return stats


def camera_rays(n, seed, w=1920.0, h=1080.0):
    rs = np.random.RandomState(seed)
    d = np.stack([rs.uniform(-w, w, n), rs.uniform(-h, h, n), np.full(n, -2.0 * min(w, h))], 1).astype(np.float32)
    return np.zeros((n, 3), np.float32), d


def bounce_rays(n, seed, centre=(0, 0, -5.0), spread=2.5):
    rs = np.random.RandomState(seed)
    o = (rs.uniform(-spread, spread, (n, 3)) + np.array(centre)).astype(np.float32)
    d = rs.normal(size=(n, 3)).astype(np.float32)
    return o, d


@pytest.mark.parametrize("name", ["c2", "c3", "c5", "c1", "c4"])
def test_generated_evaluator_equals_the_interpreter_on_the_configs(tmp_path, name):
    b = product_builder()
    env = scenes.env_latlong(64, 32)
    img = [scenes.pattern_image(16, 12, 11), scenes.pattern_image(20, 24, 12)]
    world = {"c2": lambda: scenes.config2(b), "c3": lambda: scenes.config3(b), "c5": lambda: scenes.config5(b, env=env, images=img), "c1": lambda: scenes.config1(b, env=env), "c4": lambda: scenes.config4(b, images=img)}[name]()
    flat = device.Flat(world)
    if name == "c4":  # 70 leaves: its eight groups share ONE shape function reading a constant table
        assert "ptbShape0<O0>" in probe_text(flat, 0) and "__constant__ uint4 ptbTab[" in probe_text(flat, 0)
    o, d = camera_rays(20000, 1)
    st = emulate(tmp_path, name + "_cam", flat, o, d, True, cap=16)  # 16 slots: what the kernels are launched with
    # (config 1 splits its sky planes over three clusters, so its lists are long: a few rays outgrow the shared pool
    # and are rerun by the kernel through the interpreter)
    assert st["rays"] == 20000 and st["overflowed"] <= (40 if name == "c1" else 0)
    if name == "c3":
        o, d = scenes.config3_rays(30000)
    elif name == "c4":
        o, d = bounce_rays(30000, 2, centre=(0, 0, -9.0), spread=4.0)
    else:
        o, d = bounce_rays(30000, 2, centre=(0, 0, -5.0))
    st = emulate(tmp_path, name + "_gen", flat, o, d, False, cap=16)
    assert st["hits"] > 1000 and st["overflowed"] < 0.01 * st["rays"]
    if name in ("c3", "c1", "c4"):
        assert st["multi_span"] > 100  # the general merges ran


@pytest.mark.parametrize("seed", range(12))
def test_generated_evaluator_equals_the_interpreter_on_random_trees(tmp_path, seed):
    b = product_builder()
    b.materials_reset()
    flat = device.Flat(scenes.random_tree(b, 200 + seed, 3 + seed % 8))
    o, d = scenes.fuzz_rays(6000, 300 + seed)  # half of them axis-parallel through shared endpoints
    st = emulate(tmp_path, "fuzz%d" % seed, flat, o, d, False)
    assert st["overflowed"] < 0.02 * st["rays"]
    o0, d0 = camera_rays(3000, seed, 400.0, 300.0)
    d0[:, 2] = -800.0
    emulate(tmp_path, "fuzz%d_cam" % seed, flat, o0, d0, True)
    # a pool of four slots: most multi-span rays give up (the kernel reruns them through the interpreter), none is wrong
    st = emulate(tmp_path, "fuzz%d_small" % seed, flat, o, d, False, cap=4)
    assert st["mismatches"] == 0
