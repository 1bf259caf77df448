"""Scratch: time the transit kernel variants on C4 and compare R against variant 0 (run once per variant)."""
import os, sys, json, numpy as np
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import torch
from prometheus_b200 import builder, workloads, host
name = sys.argv[1] if len(sys.argv) > 1 else "C4"
import bench
tr = builder.build_transit(workloads.get(name), nist_rows=bench.nist_rows())
R = tr.sumOverChords()
ctx = host.get_context(0)
ts = []
for i in range(8):
    R2, t = ctx.run(0, R.shape[0], 0, R.shape[1])
    ts.append(t)
v = os.environ.get("PB200_VARIANT", "0")
np.save(f"gpurun_out/R_{name}_v{v}.npy", R)
print("variant", v, name, "timing", ts[-1])
if v != "0" and os.path.exists(f"gpurun_out/R_{name}_v0.npy"):
    R0 = np.load(f"gpurun_out/R_{name}_v0.npy")
    print("  max|dR| vs v0", float(np.max(np.abs(R - R0))), "min R", float(R.min()))
