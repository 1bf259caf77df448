"""A few C4 runs (profiling driver): python scratch/c4_once.py [n]"""
import os, sys
import numpy as np
sys.path.insert(0, os.getcwd())
import prometheus_b200 as pb
from prometheus_b200 import workloads, scene, host
rows = np.load("tests/golden/linelist.npz")["rows"]
tr = pb.build_transit(workloads.get("C4"), nist_rows=rows)
ctx = host.get_context(0)
sc, keep, info = scene.flatten(tr)
ctx.set_scene(sc, keep)
import torch
out = torch.empty(info["n_orb"] * info["n_wav"], dtype=torch.float64, device="cuda")
for i in range(int(sys.argv[1]) if len(sys.argv) > 1 else 3):
    t = ctx.run_device(0, info["n_orb"], 0, info["n_wav"], out.data_ptr(), want_timing=True)
print(t)
