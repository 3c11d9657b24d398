import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "path-trace_b200")); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
from ptb200 import scenes, device, capi
from ptb200.builder import product_builder
import common, golden_cases as gc
from grazing import flag_grazing
b = product_builder()
flat = device.Flat(gc.render_scene(b, "c4")); sc = device.DeviceScene(flat); desc = flat.desc
rs = np.random.RandomState(8); n = 60000
o = np.zeros((n, 3), np.float32)
d = np.stack([rs.uniform(-1920, 1920, n), rs.uniform(-1080, 1080, n), np.full(n, -2160.0)], 1).astype(np.float32)
o[n // 2:] = rs.uniform(-2, 2, (n - n // 2, 3)) + np.array([0, 0, -5.0]); d[n // 2:] = rs.normal(size=(n - n // 2, 3))
oh, _ = common.orc_intersect(flat, o, d, 1)
gh, _ = sc.intersect(o, d, 0, 0)
topo, unc = flag_grazing(desc, o, d)
hit = ~topo & (oh["kind"] > 0) & (gh["kind"] == oh["kind"]) & ~unc[np.arange(n), np.maximum(oh["prim"], 0)]
rel = np.zeros(n); rel[hit] = np.abs(gh["t"][hit] - oh["t"][hit]) / np.abs(oh["t"][hit])
for i in np.argsort(-rel)[:8]:
    pi = int(oh["prim"][i]); p = desc.prims[pi]
    lo, ld = o[i].astype(np.float64), d[i].astype(np.float64)
    if p.chain >= 0:
        c = desc.chains[p.chain]
        for k in range(c.count):
            m = np.array(list(desc.xforms[desc.chain_xf[c.first + k]].m), np.float64).reshape(3, 4)
            lo = m[:, :3] @ lo + m[:, 3]; ld = m[:, :3] @ ld
    g = np.array(list(p.p), np.float64)
    if p.type == 0:
        oc = lo - g[:3]; a = ld @ ld; bb = oc @ ld; cc = oc @ oc - g[3]; disc = bb * bb - a * cc
        print(i, "sphere prim", pi, "rel %.3g t_ref %.7g t_gpu %.7g kind %d" % (rel[i], oh["t"][i], gh["t"][i], oh["kind"][i]), "k=disc/b2 %.3g c/|oc|2 %.3g |ld| %.3g t_true %.8g %.8g" % (disc / (bb * bb), cc / (oc @ oc), np.sqrt(a), (-bb - np.sqrt(disc)) / a, (-bb + np.sqrt(disc)) / a))
    else:
        div = ld @ g[:3]; num = -g[3] - lo @ g[:3]
        print(i, "plane prim", pi, "rel %.3g t_ref %.7g t_gpu %.7g" % (rel[i], oh["t"][i], gh["t"][i]), "div %.4g terms %.4g num %.4g lo.n %.4g d %.4g t_true %.8g" % (div, np.abs(ld * g[:3]).sum(), num, lo @ g[:3], g[3], num / div))
print("count rel>1e-5:", (rel > 1e-5).sum(), "of", hit.sum())
