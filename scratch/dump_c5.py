import numpy as np, sys, os
sys.path.insert(0, os.getcwd())
import prometheus_b200 as pb
from prometheus_b200 import workloads, host
rows = np.load("tests/golden/linelist.npz")["rows"]
for name in ("C5", "C3"):
    tr = pb.build_transit(workloads.get(name), nist_rows=rows)
    ctx = host.get_context()
    ncol = tr.getColumnDensityGrid()          # (n_atoms, n_chords) chord order c = m*n_orb + p
    n_orb = tr.spatialGrid.constructOrbphaseAxis().size
    blocked, w = tr.getChordMasks()
    p = n_orb // 2
    sig = []
    for d in tr.atmosphere.densityDistributionList:
        for con in d.constituents:
            sig.append(ctx.sigma_lookup(con.lookupFunction.x, con.lookupFunction.y, con.lookupOffset, tr.wavelength))
    np.savez_compressed(f"gpurun_out/dump_{name}.npz", ncol=ncol[:, p::n_orb], blocked=blocked[p::n_orb], w=w[p::n_orb],
                        sigma=np.array(sig), wavelength=tr.wavelength)
    print(name, ncol.shape, len(sig))
