"""RET x 64 retrieval batch, a few calls (profiling driver)."""
import os, sys
import numpy as np
sys.path.insert(0, os.getcwd())
import prometheus_b200 as pb
from prometheus_b200 import retrieval, workloads
rows = np.load("tests/golden/linelist.npz")["rows"]
spec = workloads.get("RET")
template = pb.build_transit(spec, nist_rows=rows)
orb = template.spatialGrid.constructOrbphaseAxis()
band = retrieval.LightcurveBand.for_planet(template.planet, orb, spec["band"]["line_centres"], spec["band"]["bandwidth"])
batch = retrieval.RetrievalBatch(template)
fam, K = retrieval.make_family(template, workloads.retrieval_samples(64))
for i in range(int(sys.argv[1]) if len(sys.argv) > 1 else 3):
    lc, _, t = batch.run_family(fam, K, band)
print(lc.shape, float(lc.min()), t)
