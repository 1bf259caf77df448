"""e2e of Transit.sumOverChords() at C4 with and without pb200_prefetch_tables, same process, interleaved; host-side
split of the call (prefetch / flatten / set_scene / run incl. the final synchronisation)."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import prometheus_b200 as pb
from prometheus_b200 import host, workloads, scene
rows = np.load(os.path.join(ROOT, "tests", "golden", "linelist.npz"))["rows"]
tr = pb.build_transit(workloads.get(sys.argv[1] if len(sys.argv) > 1 else "C4"), nist_rows=rows)
ctx = host.get_context(0)
t = {True: [], False: []}
for i in range(80):
    on = bool(i & 1)
    t0 = time.perf_counter()
    if on:
        ctx.prefetch_tables()
    t1 = time.perf_counter()
    sc, keep, info = scene.flatten(tr)
    t2 = time.perf_counter()
    ctx.set_scene(sc, keep)
    t3 = time.perf_counter()
    R, timing = ctx.run(0, info["n_orb"], 0, info["n_wav"])
    t4 = time.perf_counter()
    if i >= 20:
        t[on].append([(t1 - t0) * 1e3, (t2 - t1) * 1e3, (t3 - t2) * 1e3, (t4 - t3) * 1e3, (t4 - t0) * 1e3, timing["upload_ms"]])
for on in (True, False):
    m = np.median(np.array(t[on]), axis=0)
    print("prefetch %-5s: prefetch %.4f flatten %.4f set_scene %.4f run %.4f total %.4f ms (upload_ms %.4f)" % ((on,) + tuple(m)))
