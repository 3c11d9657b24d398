"""GPU box: where does the end-to-end step (scene create + render to a host buffer + destroy) spend its time?"""
import sys, time, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "path-trace_b200"))
import numpy as np, torch
from ptb200 import device, scenes
from ptb200.builder import product_builder

b = product_builder()
flat = device.Flat(scenes.config5(b))
params = device.default_params(3840, 2160, int(sys.argv[1]) if len(sys.argv) > 1 else 64, seed=0xB200)
host = torch.empty((2160, 3840, 4), dtype=torch.float32).pin_memory().numpy()
for it in range(3):
    t0 = time.perf_counter()
    sc = device.DeviceScene(flat)
    torch.cuda.synchronize(); t1 = time.perf_counter()
    _, st = sc.render(params, host)
    t2 = time.perf_counter()
    sc.close()
    torch.cuda.synchronize(); t3 = time.perf_counter()
    print("iter %d: create %.1f ms, render %.1f ms (device %.1f ms), destroy %.1f ms" % (it, (t1 - t0) * 1e3, (t2 - t1) * 1e3, st.ms_total, (t3 - t2) * 1e3))
