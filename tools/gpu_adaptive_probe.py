"""GPU box: the adaptive quadtree (ptb200_render_adaptive) at full size against the full render."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "path-trace_b200"))
import numpy as np
from ptb200 import device, scenes
from ptb200.builder import product_builder

b = product_builder()
for name, W, H, spp in (("config2", 1920, 1080, 256), ("config5", 3840, 2160, 64), ("config4", 3840, 2160, 64)):
    scene = device.DeviceScene(device.Flat(getattr(scenes, name)(b)))
    params = device.default_params(W, H, spp)
    scene.render(device.default_params(W, H, 4))  # warm-up (allocations, kernels)
    t0 = time.perf_counter(); full, st = scene.render(params); t1 = time.perf_counter()
    for delta in (0.003, 0.03):
        t2 = time.perf_counter()
        img, source, _, ast_stats, ast = scene.render_adaptive(params, 32, H // 120, delta, want_lattice=False)
        t3 = time.perf_counter()
        err = np.abs(img[..., :3] - full[..., :3]).mean() / full[..., :3].mean()
        print("%s %dx%d %d spp: full %.0f ms (%d paths); adaptive delta=%.3f %.0f ms, traced %d lattice points (%.1f %% of the pixels), interpolated %d px, %d levels, mean abs diff / mean = %.4f"
              % (name, W, H, spp, (t1 - t0) * 1e3, st.paths, delta, (t3 - t2) * 1e3, ast.pixels_traced, 100.0 * ast.pixels_traced / (W * H), ast.pixels_interpolated, ast.levels, err))
