"""One C2 run (kernel (m)), profiling driver."""
import os, sys
import numpy as np
sys.path.insert(0, os.getcwd())
import prometheus_b200 as pb
from prometheus_b200 import workloads
rows = np.load("tests/golden/linelist.npz")["rows"]
tr = pb.build_transit(workloads.get("C2"), nist_rows=rows)
R = tr.sumOverChords()
print(R.shape, R.min(), tr.last_timing)
