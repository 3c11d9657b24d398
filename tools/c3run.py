"""GPU box helper: the C3 leg of bench.py alone (PTB200_C3_LOG2 rays), for A/B runs and ncu."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "path-trace_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import bench  # noqa: E402

out = bench.run_c3(torch, 5, 72.4, 6557.8, False)
print(json.dumps({k: out[k] for k in ("value", "ms_per_step", "hit_fraction", "vs_exact_policy", "specialized")}))
