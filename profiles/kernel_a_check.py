"""Kernel (a) on its own: speed (8 B per chord*x sample) and agreement with the libm kernel of pb200_number_density."""
import sys, numpy as np, torch
sys.path.insert(0, "/root/repo")
import prometheus_b200 as pb
from prometheus_b200 import host, scene, workloads
rows = np.load("/root/repo/tests/golden/linelist.npz")["rows"]
ctx = host.get_context(0)
ctx.set_stream(torch.cuda.current_stream().cuda_stream)     # the events below see this stream only
flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda:0")
for name in ("C4", "C5s", "C1", "HYDs", "THMs"):
    tr = pb.build_transit(workloads.get(name), nist_rows=rows)
    sc, keep, info = scene.flatten(tr); ctx.set_scene(sc, keep)
    n = info["n_chords"] * info["n_x"]
    dens = torch.empty(n, dtype=torch.float64, device="cuda:0")
    ms = []
    for i in range(6):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        flush.zero_()
        a.record(); ctx.number_density_device(0, dens.data_ptr()); b.record(); ctx.synchronize(); torch.cuda.synchronize()
        ms.append(a.elapsed_time(b))
    fast = dens.cpu().numpy().reshape(info["n_chords"], info["n_x"])
    if n <= 5e7 and name != "C4":
        ref = tr.getNumberDensityGrid(0)
        big = ref > 1e-250 * ref.max()
        rel = np.max(np.abs(fast[big] - ref[big]) / ref[big]) if big.any() else 0.0
        tiny_ok = np.all(np.abs(fast[~big] - ref[~big]) <= 1e-250 * ref.max() + 1e-300)
    else:
        rel, tiny_ok = float("nan"), True
    print(name, "ms %.4f  %.0f GB/s  rel %.2e tails %s" % (min(ms[2:]), 8 * n / min(ms[2:]) / 1e6, rel, tiny_ok))
