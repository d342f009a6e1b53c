"""Tuning aid: per-work-unit (half tile) start/end times of the backward blend at config A, from a library built
with -DB200_TRACE_UNITS (make -C 3d-gaussian-splatting_b200/csrc clean all EXTRA=-DB200_TRACE_UNITS).  Prints the
distribution of unit durations, the per-SM idle time at the end of the kernel, and what a heaviest-first order would
give under a greedy model.  Output -> gpurun_out/unit_trace.json"""
import ctypes as C
import importlib
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("3d-gaussian-splatting_b200")
from oracle import torch_oracle as orc  # scene generator only

N, W, H = 1_000_000, 1920, 1080
dev = "cuda"
sc = orc.synthetic_scene(N, W, H, seed=0)
act = {k: v.to(dev).requires_grad_(True) for k, v in orc.activate(sc).items()}
be = pkg.B200Backend(sc["K"].to(dev))
vm = orc.circle_cameras(64)[0].to(dev)
v = torch.randn(H, W, 3, device=dev)
for _ in range(3):
    res = be.render_with_depth(act, {"world_view_transform": vm, "image_width": W, "image_height": H},
                               bg_color=torch.zeros(3, device=dev), render_mode="RGB", device="cuda", sh_degree=3)
    res["render"].backward(v)
torch.cuda.synchronize()
lib = pkg._lib.lib
n_units = 2 * ((W + 15) // 16) * ((H + 15) // 16)
buf = (C.c_ulonglong * (3 * n_units))()
lib.b200_debug_unit_trace.argtypes = [C.c_void_p, C.c_int]
assert lib.b200_debug_unit_trace(buf, n_units) == 0
a = np.frombuffer(buf, dtype=np.uint64).reshape(n_units, 3).astype(np.int64)
t0, t1, sm = a[:, 0], a[:, 1], a[:, 2]
dur = (t1 - t0) / 1e3  # us
k0, k1 = t0.min(), t1.max()
total = (k1 - k0) / 1e3
sm_end = np.array([t1[sm == s].max() for s in np.unique(sm)])
sm_start = np.array([t0[sm == s].min() for s in np.unique(sm)])
out = {
    "units": int(n_units), "kernel_us": float(total),
    "unit_us": {q: float(np.percentile(dur, p)) for q, p in (("p1", 1), ("p10", 10), ("p50", 50), ("p90", 90), ("p99", 99), ("max", 100))},
    "unit_us_mean": float(dur.mean()),
    "sm_idle_at_end_us": {"mean": float(((k1 - sm_end) / 1e3).mean()), "max": float(((k1 - sm_end) / 1e3).max())},
    "sm_idle_at_start_us": {"mean": float(((sm_start - k0) / 1e3).mean()), "max": float(((sm_start - k0) / 1e3).max())},
    "last_unit_started_at_us": float((t0.max() - k0) / 1e3),
}
# concurrency over time: how many units are in flight in each 10-us bin
edges = np.arange(0, total + 10, 10)
conc = [(int(((t0 - k0) / 1e3 < e + 10).sum() - ((t1 - k0) / 1e3 < e).sum())) for e in edges[:-1]]
out["units_in_flight_per_10us"] = conc
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "unit_trace.json"), "w"), indent=1)
np.save(os.path.join(ROOT, "gpurun_out", "unit_trace.npy"), a)
print(json.dumps(out))
