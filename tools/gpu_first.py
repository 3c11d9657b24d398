"""First GPU shake-out: intersect parity vs oracle + a small render vs oracle render."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "path-trace_b200")); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
from ptb200 import scenes, device, capi
from ptb200.builder import product_builder
import common

b = product_builder()
imgs = [scenes.pattern_image(64, 48, 11), scenes.pattern_image(128, 151, 12)]
env = scenes.env_latlong(400, 200)
cases = {"c3": lambda: scenes.config3(b), "c2": lambda: scenes.config2(b), "c4": lambda: scenes.config4(b, images=imgs),
         "c5": lambda: scenes.config5(b, env=env, images=imgs), "c1": lambda: scenes.config1(b, env=env)}
for name, fn in cases.items():
    w = fn(); flat = device.Flat(w); sc = device.DeviceScene(flat)
    if name == "c3":
        o, d = scenes.config3_rays(200000)
    else:
        rs = np.random.RandomState(5)
        o = np.zeros((100000, 3), np.float32)
        d = np.stack([rs.uniform(-1000, 1000, 100000), rs.uniform(-600, 600, 100000), np.full(100000, -2160.0)], 1).astype(np.float32)
    oh, osp = common.orc_intersect(flat, o, d, 8)
    for flags, label in ((capi.PTB_FLAG_EXACT, "exact"), (0, "fast")):
        gh, gsp = sc.intersect(o, d, 8, flags)
        same_kind = (gh["kind"] == oh["kind"]).mean(); same_n = (gh["n_spans"] == oh["n_spans"]).mean()
        hit = (oh["kind"] > 0) & (gh["kind"] == oh["kind"])
        rel = np.abs(gh["t"][hit] - oh["t"][hit]) / np.abs(oh["t"][hit])
        bit = np.array_equal(gh.view(np.uint8), oh.view(np.uint8)) and np.array_equal(gsp.view(np.uint8), osp.view(np.uint8))
        print(name, label, "kind=%.6f nspans=%.6f prim=%.6f maxrel_t=%.3g bit_identical=%s hits=%.3f" % (same_kind, same_n, (gh["prim"] == oh["prim"]).mean(), rel.max() if len(rel) else 0, bit, (oh["kind"] > 0).mean()))
    # render vs oracle: small lattice
    W, H = 96, 54
    t = time.time(); ref_img, rays = common.orc_render_mt(flat, W, H, 0, 0, 1, W, H, 2, 16, 7, 8)[0:2]; tc = time.time() - t
    p = device.default_params(W, H, 4096)
    t = time.time(); img, st = sc.render(p); tg = time.time() - t
    g = img[..., :3]
    rmse = np.sqrt(((g - ref_img) ** 2).mean()) / max(ref_img.mean(), 1e-9)
    print(name, "render mean gpu %.5f oracle %.5f relRMSE %.4f  rays/path %.3f  gpu %.2fs (%.1f Mrays/s) cpu %.1fs" % (g.mean(), ref_img.mean(), rmse, st.rays / st.paths, tg, st.rays / st.ms_total / 1e3, tc))
    p2 = device.default_params(W, H, 4096, flags=capi.PTB_FLAG_REJECTION)
    img2, st2 = sc.render(p2)
    print(name, "  rejection-sampler mean %.5f relRMSE vs cap %.4f" % (img2[..., :3].mean(), np.sqrt(((img2[..., :3] - g) ** 2).mean()) / g.mean()))
print("fp32 peak", device.fp32_peak())
