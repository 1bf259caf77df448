"""C5 runs with and without sigma-sorted windows (timing + difference)."""
import os, sys
import numpy as np
sys.path.insert(0, os.getcwd())
import prometheus_b200 as pb
from prometheus_b200 import workloads, scene, host
rows = np.load("tests/golden/linelist.npz")["rows"]
name = sys.argv[1] if len(sys.argv) > 1 else "C5"
tr = pb.build_transit(workloads.get(name), nist_rows=rows)
ctx = host.get_context(0)
sc, keep, info = scene.flatten(tr)
ctx.set_scene(sc, keep)
res = {}
for flag in ("0", "1", "0", "1"):
    os.environ["PB200_SORT"] = flag
    R, t = ctx.run(0, info["n_orb"], 0, info["n_wav"])
    res[flag] = R.copy()
    print(name, "sort", flag, {k: round(t[k], 3) for k in ("column_ms", "transit_ms")}, "executed", t["executed_samples"] / t["samples"], "exp", t["exp_samples"] / t["samples"])
print("max |dR|", float(np.max(np.abs(res["0"] - res["1"]))))
