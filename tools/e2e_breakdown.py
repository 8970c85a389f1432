"""Wall-clock breakdown of one end-to-end call sequence on K4 (tools only)."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bundleadjustment_b200 as bab
from bundleadjustment_b200 import synthetic as syn

import dataclasses
import numpy as np


def pinned(a):
    t = torch.from_numpy(np.ascontiguousarray(a)).pin_memory()
    return t.numpy()


pb = syn.flatten(syn.config("K4"))
pb = dataclasses.replace(pb, **{f.name: pinned(getattr(pb, f.name)) for f in dataclasses.fields(pb)})   # as bench.py's e2e leg
cam_out, pt_out = pinned(np.zeros_like(pb.cam)), pinned(np.zeros_like(pb.pt))
s = bab.Solver(bab.default_options(max_num_iterations=10))
for it in range(6):
    torch.cuda.synchronize()
    t0 = time.perf_counter(); s.set_problem(pb); torch.cuda.synchronize()
    t1 = time.perf_counter(); s.reduced_system_stats(); torch.cuda.synchronize()
    t2 = time.perf_counter(); sm = s.solve(); torch.cuda.synchronize()
    t3 = time.perf_counter(); s.cameras(cam_out); s.points(pt_out); torch.cuda.synchronize()
    t4 = time.perf_counter()
    print(f"set_problem {1e3*(t1-t0):7.2f} ms  ensure_dense {1e3*(t2-t1):7.2f} ms  solve {1e3*(t3-t2):7.2f} ms (device {1e3*sm.seconds_total:.2f})  get {1e3*(t4-t3):6.2f} ms  total {1e3*(t4-t0):7.2f}")
s.close()
