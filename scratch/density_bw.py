"""Kernel (a) standalone: achieved GB/s of the materialised number density (8 B per chord*x sample)."""
import copy, os, sys
import numpy as np, torch
sys.path.insert(0, os.getcwd())
import prometheus_b200 as pb
from prometheus_b200 import workloads, scene, host, constants as const
rows = np.load("tests/golden/linelist.npz")["rows"]
ctx = host.get_context(0)
ctx.set_stream(torch.cuda.current_stream().cuda_stream)
flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
def variants():
    c4 = workloads.get("C4"); yield "C4 torus", c4
    b = copy.deepcopy(c4); b["scenarios"] = [{"type": "barometric", "T": 3000.0, "P_0": 1e3, "mu": 2.3 * const.amu, "species": [{"name": "NaI", "chi": 1e-6}]}]; yield "C4 grid, barometric", b
    h = copy.deepcopy(c4); h["scenarios"] = [{"type": "hydrostatic", "T": 3000.0, "P_0": 1e3, "mu": 2.3 * const.amu, "species": [{"name": "NaI", "chi": 1e-6}]}]; yield "C4 grid, hydrostatic", h
    p = copy.deepcopy(c4); p["scenarios"] = [{"type": "powerLaw", "N": 1e33, "q_esc": 4.0, "species": [{"name": "NaI", "sigma_v": 1e6}]}]; yield "C4 grid, powerLaw", p
for name, spec in variants():
    tr = pb.build_transit(spec, nist_rows=rows)
    sc, keep, info = scene.flatten(tr)
    ctx.set_scene(sc, keep)
    n = info["n_chords"] * info["n_x"]
    out = torch.empty(n, dtype=torch.float64, device="cuda")
    ms = []
    for i in range(8):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); ctx.number_density_device(0, out.data_ptr()); b.record(); torch.cuda.synchronize()
        if i >= 3: ms.append(a.elapsed_time(b))
    t = float(np.mean(ms))
    ref = np.empty(n)
    ctx.check(ctx.lib.pb200_number_density(ctx.handle, 0, ref.ctypes.data_as(__import__("ctypes").POINTER(__import__("ctypes").c_double))))
    got = out.cpu().numpy()
    big = ref > ref.max() * 1e-250
    rel = float(np.max(np.abs(got[big] - ref[big]) / ref[big])) if big.any() else 0.0
    zeros_ok = bool(np.array_equal(got == 0.0, ref == 0.0))
    print(f"   vs exact kernel: max rel diff {rel:.2e}, zero pattern equal {zeros_ok}")
    print(f"{name}: {n:.3e} samples, {t:.4f} ms, {8 * n / t / 1e6:.0f} GB/s, sum {float(out.sum()):.6e}")
# write-only roofline of this box: torch fill of the same 367 MB buffer, and a device copy
ms = []
src = torch.empty_like(out)
for i in range(8):
    flush.zero_()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); out.fill_(1.0); b.record(); torch.cuda.synchronize()
    if i >= 3: ms.append(a.elapsed_time(b))
print(f"fill_ of {out.numel() * 8 / 1e6:.0f} MB: {np.mean(ms):.4f} ms, {out.numel() * 8 / np.mean(ms) / 1e6:.0f} GB/s written")
ms = []
for i in range(8):
    flush.zero_()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); out.copy_(src); b.record(); torch.cuda.synchronize()
    if i >= 3: ms.append(a.elapsed_time(b))
print(f"copy_ of {out.numel() * 8 / 1e6:.0f} MB: {np.mean(ms):.4f} ms, {2 * out.numel() * 8 / np.mean(ms) / 1e6:.0f} GB/s read+written")
