"""Times ptb200_scene_create on C5 with pageable and with page-locked image buffers (tools/gpu_r02D.sh)."""
import sys, time, os
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "path-trace_b200"))
import torch
from ptb200 import device, scenes
from ptb200.builder import product_builder

flat = device.Flat(scenes.config5(product_builder()))
torch.cuda.init()
def create(n=4):
    out = []
    for _ in range(n):
        torch.cuda.synchronize()
        t = time.perf_counter()
        s = device.DeviceScene(flat)
        torch.cuda.synchronize()
        out.append(round((time.perf_counter() - t) * 1e3, 2))
        s.close()
    return out
print("pageable", create())
flat.pin_images()
print("pinned  ", create())
t = flat._pinned[0]
for _ in range(3):
    torch.cuda.synchronize(); a = time.perf_counter(); d = t.cuda(non_blocking=True); torch.cuda.synchronize()
    print("torch H2D of", t.numel() * 4 >> 20, "MiB:", round((time.perf_counter() - a) * 1e3, 2), "ms")
