"""One workload, a few runs, for ncu (profiles/ holds the summaries):

    ncu --set full --clock-control none --import-source on -k regex:<kernel> -c <n> -o gpurun_out/<name> \
        python profiles/run_once.py <workload> [runs] [kernel_a] [opt=value ...]

`kernel_a`: also launch the materialised number density (pb200_number_density_device) of scenario 0."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import prometheus_b200 as pb  # noqa: E402
from prometheus_b200 import host, scene, workloads  # noqa: E402


def main(argv):
    name = argv[0]
    runs = int(argv[1]) if len(argv) > 1 and argv[1].isdigit() else 3
    opts = dict(a.split("=") for a in argv[1:] if "=" in a)
    rows = np.load(os.path.join(ROOT, "tests", "golden", "linelist.npz"))["rows"]
    ctx = host.get_context(0)
    for k, v in opts.items():
        if k not in ("batch", "ranks", "strong"):
            ctx.set_option(k, int(v))
    spec = workloads.get(name)
    strong = int(opts.pop("strong", 1))   # rank 0's share of the UNCHANGED grid split over `strong` ranks
    ranks = int(opts.pop("ranks", 1))
    if ranks > 1:             # one rank's block-cyclic tile of the `ranks`-fold refined grid (bench.py's weak-scaling rule)
        spec["wavelength"][3] /= ranks
        spec["wavelength"][4] /= ranks
    tr = pb.build_transit(spec, nist_rows=rows)
    if "batch" in opts:       # retrieval batch: `batch` parameter vectors stacked along the phase axis (pb200_run_batch)
        from prometheus_b200 import retrieval
        spec = workloads.get(name)
        n = int(opts.pop("batch"))
        band = retrieval.LightcurveBand.for_planet(tr.planet, tr.spatialGrid.constructOrbphaseAxis(), spec["band"]["line_centres"],
                                                   spec["band"]["bandwidth"])
        batch = retrieval.RetrievalBatch(tr)
        for _ in range(runs):
            family, K = retrieval.make_family(tr, workloads.retrieval_samples(n))
            lc, _, t = batch.run_family(family, K, band)
        print(name, "batch", n, {k: round(v, 4) if isinstance(v, float) else v for k, v in t.items()})
        return
    sc, keep, info = scene.flatten(tr)
    ctx.set_scene(sc, keep)
    out = torch.empty(info["n_orb"] * info["n_wav"], dtype=torch.float64, device="cuda:0")
    flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda:0")
    for _ in range(runs):
        flush.zero_()
        torch.cuda.synchronize()
        if strong > 1 and info["n_mols"] == 0:
            t = ctx.run_device_cyclic(0, info["n_orb"], 256, strong, 0, out.data_ptr(), want_timing=True)
        elif strong > 1:
            t = ctx.run_device(0, info["n_orb"], 0, -(-info["n_wav"] // strong), out.data_ptr(), want_timing=True)
        elif ranks > 1:
            t = ctx.run_device_cyclic(0, info["n_orb"], 256, ranks, 0, out.data_ptr(), want_timing=True)
        else:
            t = ctx.run_device(0, info["n_orb"], 0, info["n_wav"], out.data_ptr(), want_timing=True)
    print(name, {k: round(v, 4) if isinstance(v, float) else v for k, v in t.items()})
    if "kernel_a" in argv:
        dens = torch.empty(info["n_chords"] * info["n_x"], dtype=torch.float64, device="cuda:0")
        ctx.set_stream(torch.cuda.current_stream().cuda_stream)
        ms = []
        for _ in range(runs + 3):
            flush.zero_()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); ctx.number_density_device(0, dens.data_ptr()); b.record()
            torch.cuda.synchronize()
            ms.append(a.elapsed_time(b))
        t = float(np.mean(ms[3:]))
        print("kernel_a", name, "ms", round(t, 4), "GB/s", round(8.0 * dens.numel() / t / 1e6, 1))


if __name__ == "__main__":
    main(sys.argv[1:])
