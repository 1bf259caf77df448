#!/usr/bin/env python
"""bench.py -- chord*lambda samples/s of the transit hot path on N B200s, next to the CPU path.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload C4] [--impl reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P bench.py --gpus N --steps K --warmup W

One "step" = one complete Transit.sumOverChords of the workload (kernels (a') + (c') [+ split
reduction]) -> R(phase, lambda).  Workload at N GPUs (weak scaling): the named config with its
wavelength resolution refined N-fold, so n_wav ~ N x the single-GPU n_wav; rank r owns the r-th
contiguous wavelength chunk of every orbital phase; the only collective is the final
all-gather of the R tiles (NCCL).  Prints ONE JSON line on rank 0.

  value      whole-job chord*lambda samples/s, scene resident in HBM, CUDA-event timed, max over ranks
  e2e        same metric through the host API with HOST buffers: scene flattening + H2D upload from
             pinned memory + kernels + gather + D2H of R inside the timed region
  roofline   kernel (c') against the measured HBM peak on the north star's staged-equivalent
             convention (16 B per chord*lambda sample); plus its FP64-pipe fraction
  cpu_baseline / --impl reference
             the CPU restatement of the reference algorithm (oracle/, numpy+scipy like the
             reference itself) on all host cores via multiprocessing over wavelength chunks, on a
             bounded contiguous wavelength sample of the same workload
"""
from __future__ import annotations

import argparse
import json
import math
import os
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

BYTES_PER_SAMPLE_STAGED = 16.0      # one f64 tau store + one load (SURVEY 8(d), row (c'))


def flops_per_sample(n_atoms: int) -> float:
    return 2.0 * n_atoms + 29.0     # SURVEY 8(d) convention: S FMA + negate + exp(26) + 1 FMA


def executed_flops_per_sample(n_atoms: int) -> float:
    """FP64 flops kernel (c') really issues per sample (pb200_math.cuh "Big-table form"): one absorber
    2 FMA + 1 ADD + 3 FMA = 11; S absorbers 1 MUL + (S-1) FMA + 3 ADD + 3 FMA = 2S + 8."""
    return 11.0 if n_atoms == 1 else 2.0 * n_atoms + 8.0


# ------------------------------------------------------------------------------------------------
# workload
# ------------------------------------------------------------------------------------------------
def scaled_spec(name: str, n_gpus: int) -> dict:
    from prometheus_b200 import workloads
    spec = workloads.get(name)
    if n_gpus > 1:      # weak scaling along the wavelength axis: N-fold finer grid
        spec["wavelength"][3] /= n_gpus
        spec["wavelength"][4] /= n_gpus
        spec["name"] = f"{name}x{n_gpus}"
    return spec


def nist_rows():
    return np.load(os.path.join(ROOT, "tests", "golden", "linelist.npz"))["rows"]


# ------------------------------------------------------------------------------------------------
# clocks: sampled DURING the timed region (NVML in a thread)
# ------------------------------------------------------------------------------------------------
class ClockSampler:
    REASONS = {0x1: "gpu_idle", 0x2: "applications_clocks_setting", 0x4: "sw_power_cap", 0x8: "hw_slowdown",
               0x10: "sync_boost", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
               0x80: "hw_power_brake_slowdown", 0x100: "display_clock_setting"}

    def __init__(self, index: int, period: float = 0.005):
        self.samples, self.mask, self.max_mhz, self.power = [], 0, None, []
        self._stop = threading.Event()
        self._thr = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = int(vis.split(",")[index]) if vis and vis.replace(",", "").isdigit() else index
            self.h = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
        except Exception:
            self.nv = None
        self.period = period

    def _loop(self):
        nv = self.nv
        while not self._stop.is_set():
            try:
                self.samples.append(float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
                self.mask |= int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
                self.power.append(nv.nvmlDeviceGetPowerUsage(self.h) / 1000.0)
            except Exception:
                pass
            self._stop.wait(self.period)

    def start(self):
        if self.nv is not None:
            self._thr = threading.Thread(target=self._loop, daemon=True)
            self._thr.start()

    def stop(self) -> dict:
        if self._thr is not None:
            self._stop.set()
            self._thr.join()
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": ["unavailable"]}
        reasons = [n for b, n in self.REASONS.items() if self.mask & b and n != "gpu_idle"]
        return {"sm_mhz": float(np.median(self.samples)), "sm_max_mhz": self.max_mhz, "reasons": reasons,
                "samples": len(self.samples), "power_w_max": max(self.power) if self.power else None}


# ------------------------------------------------------------------------------------------------
# CPU arm.  kind "reference": the UNMODIFIED reference from oracle/_ref (oracle/make_ref.py) -- its own classes build
# the scene, its own Transit.sumOverChords(max_memory_gb=2.0) runs -- under multiprocessing over wavelength chunks
# (SURVEY 8(d) "Secondary": transit.wavelength = lambda[chunk] per worker, NUMBA/OMP threads = 1; exact, the columns
# of R are independent).  kind "port": the oracle (numpy/scipy restatement) the same way, used only where the
# reference copy is absent.
# ------------------------------------------------------------------------------------------------
_W = {}


def reference_available() -> bool:
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import make_ref
    return make_ref.available()


def _load_tables(tables_path):
    out = {}
    if tables_path:
        z = np.load(tables_path)
        for key in z.files:
            if key.startswith("x|"):
                _, i, sp = key.split("|")
                out[(int(i), sp)] = (z[key], z["y|" + i + "|" + sp])
    return out


def _cpu_init(name, n_gpus, tables_path, kind):
    for v in ("OMP_NUM_THREADS", "OPENBLAS_NUM_THREADS", "MKL_NUM_THREADS", "NUMBA_NUM_THREADS"):
        os.environ[v] = "1"
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    spec = scaled_spec(name, n_gpus)
    tables = _load_tables(tables_path)
    _W["kind"] = kind
    if kind == "reference":
        import make_ref
        from prometheus_b200 import builder
        ns = make_ref.ReferenceNamespace()
        tr = builder.build_transit(spec, ns=ns, nist_rows=ns.nist_rows, build_tables=not tables)
        for i, d in enumerate(tr.atmosphere.densityDistributionList):
            for con in d.constituents:
                if not con.isMolecule and tables:
                    con.lookupFunction = ns.table_1d(*tables[(i, con.species.name)])
        _W["transit"] = tr
        _W["n_orb"] = len(tr.spatialGrid.constructOrbphaseAxis())
        _W["wl_full"] = np.array(tr.wavelength)
        return
    from helpers import oracle, oracle_inputs
    inputs = oracle_inputs(spec, nist_rows() if spec.get("linelist", "nist") == "nist" else None)
    if tables:
        inputs["tables"] = tables
    _W["oracle"] = oracle
    _W["model"] = oracle.build_model(spec, inputs)


def _cpu_chunk(wl):
    if _W["kind"] == "reference":
        tr = _W["transit"]
        if wl.size == 0:                       # start-up call: absorbs the numba JIT of n_interp_log (~7 s cold)
            wl = _W["wl_full"][:2]
            tr.wavelength = wl
            tr.sumOverChords(max_memory_gb=2.0)
            return np.empty((_W["n_orb"], 0))
        for d in tr.atmosphere.densityDistributionList:       # the one thing the reference mutates (gasProperties.py:948-952)
            for con in d.constituents:
                if hasattr(con, "_batch_sigma_cache"):
                    del con._batch_sigma_cache
        tr.wavelength = wl                     # the attribute addWavelength sets (gasProperties.py:1105)
        return tr.sumOverChords(max_memory_gb=2.0)
    oracle, m = _W["oracle"], _W["model"]
    if wl.size == 0:
        return np.empty((len(m.orb), 0))
    has_mol = any(c.kind == "molecule" for s in m.scenarios for c in s.constituents)
    n_c = len(m.rho) * len(m.phi) * len(m.orb)
    bs = oracle.reference_batch_size(n_c, wl.size, len(m.x), has_mol, 2.0)     # memoryHandler.py:55-66
    return oracle.sum_over_chords(m, batch_size=None if bs >= n_c else bs, wavelength=wl)


class CpuArm:
    """multiprocessing.Pool(P) over wavelength chunks, each worker running Transit.sumOverChords with the
    reference's own RAM-bounded chord batching on its chunk."""

    def __init__(self, name, n_gpus, tables=None, procs=None, kind=None):
        import multiprocessing as mp
        self.kind = kind or ("reference" if reference_available() else "port")
        self.procs = procs or (os.cpu_count() or 1)
        self.tmp = None
        if tables:
            fd, self.tmp = tempfile.mkstemp(suffix=".npz")
            os.close(fd)
            flat = {}
            for (i, sp), (tx, ty) in tables.items():
                flat[f"x|{i}|{sp}"] = np.asarray(tx)
                flat[f"y|{i}|{sp}"] = np.asarray(ty)
            np.savez(self.tmp, **flat)
        self.pool = mp.get_context("spawn").Pool(self.procs, initializer=_cpu_init, initargs=(name, n_gpus, self.tmp, self.kind))
        self.pool.map(_cpu_chunk, [np.empty(0)] * self.procs, chunksize=1)      # every worker up and warm

    def run(self, wl):
        t0 = time.perf_counter()
        parts = self.pool.map(_cpu_chunk, np.array_split(wl, self.procs), chunksize=1)
        return np.hstack(parts), time.perf_counter() - t0

    def describe(self):
        if self.kind == "reference":
            return ("unmodified reference (oracle/_ref/Prometheus): Transit.sumOverChords(max_memory_gb=2.0) per worker on "
                    "transit.wavelength = lambda[chunk], NUMBA_NUM_THREADS=OMP_NUM_THREADS=1")
        return "oracle port (oracle/prometheus_oracle.py; the reference copy oracle/_ref is absent)"

    def close(self):
        self.pool.close()
        self.pool.join()
        if self.tmp:
            os.unlink(self.tmp)


def cost_model(arm, wl, centre):
    """Two probes -> t = a + b*n per step of n wavelengths over the arm's P workers; a is the wavelength-independent
    part (density and column density of every chord, repeated by every worker; set-up of the batches)."""
    n1, n2 = 8 * arm.procs, 48 * arm.procs
    t1 = min(arm.run(wl[slice(*pick_sample(wl, centre, n1))])[1] for _ in range(2))
    t2 = arm.run(wl[slice(*pick_sample(wl, centre, n2))])[1]
    b = max((t2 - t1) / (n2 - n1), 1e-9)
    a = max(t1 - b * n1, 0.0)
    return a, b


def calibrate_sample(arm, wl, centre, n_c, seconds, floor=0):
    """-> (lo, hi, a, b): a wavelength window around `centre` whose step takes ~`seconds`, at least `floor` wide."""
    a, b = cost_model(arm, wl, centre)
    n = int(max(48 * arm.procs, floor, (seconds - a) / b))
    lo, hi = pick_sample(wl, centre, n)
    return lo, hi, a, b


def pick_sample(wavelength, centre_index, n):
    n = int(min(max(n, 1), wavelength.size))
    lo = int(min(max(centre_index - n // 2, 0), wavelength.size - n))
    return lo, lo + n


def full_job_seconds(a, t_step, n_step, n_wav):
    """SURVEY 8(d) sub-sampling rule: the per-wavelength cost measured on the step scales linearly in n_wav, the
    wavelength-independent cost a is added once, unscaled."""
    return a + max(t_step - a, 0.0) * (n_wav / float(n_step))


# ------------------------------------------------------------------------------------------------
# reference arm: --impl reference
# ------------------------------------------------------------------------------------------------
def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from prometheus_b200 import builder
    n_gpus = args.gpus
    spec = scaled_spec(args.workload, n_gpus)
    tr = builder.build_transit(spec, nist_rows=nist_rows(), build_tables=False)     # host-only: axes + grid
    wl = tr.wavelength
    n_c = int(spec["grid"][4]) * int(spec["grid"][5]) * int(spec["grid"][7])
    arm = CpuArm(args.workload, n_gpus)          # the reference builds its own sigma tables (scipy) in every worker
    centre = wl.size // 2
    # a step = one bounded wavelength sample (>= 2048 wavelengths: line cores and continuum), sized so
    # that warm-up + steps end within --cpu-seconds
    budget = args.cpu_seconds / max(args.steps + args.warmup + 2, 1)
    lo, hi, a, b = calibrate_sample(arm, wl, centre, n_c, budget, floor=min(2048, wl.size))
    for _ in range(args.warmup):
        arm.run(wl[lo:hi])
    times = [arm.run(wl[lo:hi])[1] for _ in range(args.steps)]
    arm.close()
    t = float(np.mean(times))
    n_s = hi - lo
    t_full = full_job_seconds(a, t, n_s, wl.size)
    value = n_c * float(wl.size) / t_full
    sample = (f"{n_s} contiguous wavelengths [{lo},{hi}) of {wl.size} around the grid's centre, all {n_c} chords, per step; "
              f"{arm.describe()}")
    print(json.dumps({
        "impl": "reference", "metric": "chord*lambda samples/s", "value": value, "unit": "samples/s",
        "n_gpus": n_gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": t * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(args.workload, spec, n_c, wl.size, n_gpus),
        "cpu_baseline": {"value": value, "unit": "samples/s", "cores": arm.procs, "kind": arm.kind, "sample": sample},
        "e2e": {"value": value, "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "value_note": "whole-workload throughput by SURVEY 8(d)'s sub-sampling rule: per-wavelength cost of the timed step scaled "
                      "to n_wav, the wavelength-independent cost (density of every chord, repeated per worker) added once; "
                      "`sample_value` is the raw n_chords*n_sample/t_step, which over-weights that cost and under-reports the CPU",
        "sample_value": n_c * float(n_s) / t,
        "cost_model_s": {"lambda_independent": a, "per_wavelength": b, "step_mean": t, "step_min": float(np.min(times))},
        "full_run_seconds_extrapolated": t_full,
    }))


def workload_config(name, spec, n_c, n_wav, n_gpus):
    g = spec["grid"]
    return {"workload": f"{name}: {spec['title']}", "n_chords": n_c, "n_wav": int(n_wav), "n_x": int(g[2]),
            "n_orb": int(g[7]), "n_phi": int(g[5]), "n_rho": int(g[4]),
            "sharding": f"wavelengths over {n_gpus} rank(s) (blocks of 256 dealt round-robin), replicated scene; `value`: the R tiles stay on the GPU that computed them (independent units, no data-path collective); `e2e`: every rank's kernels store their tiles into one page-locked shared-memory result (pb200_run_scatter), no gather",
            "weak_scaling_rule": "wavelength resolution refined N-fold at N GPUs",
            "cache": "L2 flushed between timed steps (512 MiB write)"}


# ------------------------------------------------------------------------------------------------
# parity of the benchmarked result itself: the oracle (CPU, rank 0, as the checker) on a few columns over ALL chords
# ------------------------------------------------------------------------------------------------
def sigma_tables_of(tr):
    out = {}
    for i, d in enumerate(tr.atmosphere.densityDistributionList):
        for con in d.constituents:
            if not con.isMolecule:
                out[(i, con.species.name)] = (np.asarray(con.lookupFunction.x), np.asarray(con.lookupFunction.y))
    return out


def parity_check(spec, tr, R, n_cols=8):
    """R (n_orb, n_wav) of the timed workload against the oracle on the deepest line core, its neighbours and a
    strided set of columns (>= n_cols), every chord, the GPU's own sigma tables.  North-star criterion:
    |E - E_ref| <= 1e-10 + 1e-6 |E_ref| with E = R0 - R."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from helpers import oracle, oracle_inputs
    t0 = time.perf_counter()
    inputs = oracle_inputs(spec, nist_rows() if spec.get("linelist", "nist") == "nist" else None)
    inputs["tables"] = sigma_tables_of(tr)
    m = oracle.build_model(spec, inputs)
    n_orb, n_wav = R.shape
    assert np.array_equal(m.wavelength, np.asarray(tr.wavelength))
    core = int(np.argmin(R.min(axis=0)))
    sel = np.unique(np.clip(np.concatenate([[core - 1, core, core + 1], np.linspace(0, n_wav - 1, n_cols).astype(int)]), 0, n_wav - 1))
    has_mol = any(c.kind == "molecule" for sc_ in m.scenarios for c in sc_.constituents)
    n_c = len(m.rho) * len(m.phi) * len(m.orb)
    bs = oracle.reference_batch_size(n_c, len(sel), len(m.x), has_mol, 2.0)
    kw = dict(batch_size=None if bs >= n_c else bs, wavelength=m.wavelength[sel])
    R_ref = oracle.sum_over_chords(m, **kw)
    R0 = oracle.sum_over_chords(m, opaque_only=True, **kw)
    E, E_ref = R0 - R[:, sel], R0 - R_ref
    worst = float(np.max(np.abs(E - E_ref) / (1e-10 + 1e-6 * np.abs(E_ref))))
    return {"n_cols": int(sel.size), "n_chords": int(n_c), "columns": [int(i) for i in sel],
            "max_abs_diff": float(np.max(np.abs(R[:, sel] - R_ref))), "worst_err_over_bound": worst,
            "max_excess_depth": float(E_ref.max()), "checker": "oracle/prometheus_oracle.py on rank 0's host cores",
            "seconds": time.perf_counter() - t0}


# ------------------------------------------------------------------------------------------------
# FP64 work per unit, counted from the instruction streams (profiles/sass_transit_inner_loop.txt, profiles/sass_moment_item_loop.txt), add/mul = 1, FMA = 2
# ------------------------------------------------------------------------------------------------
FLOPS_MOMENT_ITEM = 33.0      # kernel (m), per (item, wavelength): d = V1 - V0 (1), fma(tc, d, V0) (2), table exp (12),
                              # Horner of 8 moments (16), e * p (1), accumulate (1)
FLOPS_CENTRE_PAIR = 49.0      # kernel (c'), sigma cells, per (chord, cell): table exp (11), E = T q (1), u (1), 13 sums (1 + 2 + 11 x 3)
FLOPS_MOMENT_BLOCK = 15.0     # kernel (m), per (chord pair, wavelength) besides its items and atoms: table exp (12) + clamp,
                              # subtract the offset sum (1), weight * exp accumulate (2)


def fp64_roofline(kernel, flops, kernel_ms, peak_tflops, clocks_mhz=None, **extra):
    ach = flops / (kernel_ms * 1e-3) / 1e12
    out = {"kernel": kernel, "bound": "fp64", "achieved": ach, "peak": peak_tflops, "unit": "TFLOP/s", "frac": ach / peak_tflops,
           "peak_source": "DFMA-chain probe on this GPU in this run (pb200_probe_fp64; MEASURED_PEAKS.json holds no FP64 figure)",
           "kernel_ms": kernel_ms, "flops_per_launch": flops}
    out.update(extra)
    return out


# ------------------------------------------------------------------------------------------------
# B200 arm
# ------------------------------------------------------------------------------------------------
def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def recorded_traffic(workload, kernel="transit_kernel"):
    """dram bytes per launch of the dominant kernel from the committed ncu capture, or None."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(p):
        return json.load(open(p)).get(workload, {}).get(kernel)
    return None


def quick_workload(name, ctx, torch, stream, flush, peak_tflops, steps=3, warmup=3):
    """Device-resident samples/s of another workload on this rank's GPU (3 warm-up + 3 timed steps)."""
    import prometheus_b200 as pb
    from prometheus_b200 import scene
    spec = scaled_spec(name, 1)
    t0 = time.perf_counter()
    tr = pb.build_transit(spec, nist_rows=nist_rows())
    t_setup = time.perf_counter() - t0
    sc, keep, info = scene.flatten(tr)
    ctx.set_scene(sc, keep)
    out = torch.empty(info["n_orb"] * info["n_wav"], dtype=torch.float64, device=flush.device)
    ms, kern, evs = [], [], []
    for i in range(warmup + steps):      # timed steps queue without a host synchronisation (see run_b200)
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        ctx.run_device(0, info["n_orb"], 0, info["n_wav"], out.data_ptr())
        b.record()
        if i >= warmup:
            evs.append((a, b))
    torch.cuda.synchronize()
    ms = [a.elapsed_time(b) for a, b in evs]
    for i in range(3):                   # instrumented steps: the library's per-kernel record
        flush.zero_()
        kern.append(ctx.run_device(0, info["n_orb"], 0, info["n_wav"], out.data_ptr(), want_timing=True))
    R = tr.sumOverChords()                       # warm-up: page-locked result block, scratch growth
    R = None                                     # give the block back to the pool: the timed call reuses it
    t0 = time.perf_counter()
    R = tr.sumOverChords()
    e2e_ms = (time.perf_counter() - t0) * 1e3
    samples = float(info["n_chords"]) * info["n_wav"]
    res = {"title": spec["title"], "n_chords": info["n_chords"], "n_wav": info["n_wav"], "n_x": info["n_x"],
           "n_atoms": info["n_atoms"], "n_molecules": info["n_mols"], "ms_per_step": float(np.mean(ms)),
           "samples_per_s": samples / (np.mean(ms) * 1e-3), "e2e_ms_one_call": e2e_ms,
           "kernels_ms": {k: float(np.mean([t[k] for t in kern])) for k in ("column_ms", "transit_ms", "finalize_ms")},
           "setup_seconds": t_setup, "R_min": float(R.min()), "R_max": float(R.max()),
           "executed_frac": float(np.mean([t["executed_samples"] for t in kern])) / samples}
    res["roofline"] = workload_roofline(name, tr, info, kern, peak_tflops)
    if info["n_mols"]:
        res["molecular_lookups_per_s"] = samples * info["n_x"] * info["n_mols"] / (np.mean(ms) * 1e-3)
    return res


def setup_comparison(name="C2"):
    """Set-up time of a workload: the mirror's build_transit -- line-list cache, refined wavelength
    grids (host numpy), Voigt tables (kernel (b2)) -- profiled by function, next to the reference's own classes building
    the same scene (AtmosphericConstituent.addLookupFunctionToConstituent: single-threaded scipy.special.voigt_profile,
    gasProperties.py:672-715), when the reference copy is on the box."""
    import cProfile
    import pstats
    import prometheus_b200 as pb
    from prometheus_b200 import workloads
    spec = workloads.get(name)
    pb.build_transit(spec, nist_rows=nist_rows())                   # warm: context, pinned pool
    prof = cProfile.Profile()
    t0 = time.perf_counter()
    prof.enable()
    tr = pb.build_transit(spec, nist_rows=nist_rows())
    prof.disable()
    total = time.perf_counter() - t0
    stats = pstats.Stats(prof).stats
    want = {"arangeWavelengthGrid": "refined / merged wavelength grids (host numpy)", "voigt_sigma": "Voigt tables incl. H2D/D2H (kernel b2)",
            "getLineParameters": "line selection", "species_lines": "line-list parsing (cached per list)",
            "resolve_line_list": "line list rows", "empty": "page-locked table buffers"}
    parts = {}
    for (fn, _, func), (_, _, _, cum, _) in stats.items():
        if func in want:
            parts[func] = parts.get(func, 0.0) + cum
    out = {"workload": name, "mirror_seconds": total, "mirror_by_function_seconds": {f"{k} ({want[k]})": v for k, v in parts.items()},
           "n_tables": sum(1 for d in tr.atmosphere.densityDistributionList for c in d.constituents if not c.isMolecule)}
    if reference_available():
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import make_ref
        from prometheus_b200 import builder
        ns = make_ref.ReferenceNamespace()
        t0 = time.perf_counter()
        ref = builder.build_transit(spec, ns=ns, nist_rows=ns.nist_rows)
        out["reference_seconds"] = time.perf_counter() - t0
        out["speedup"] = out["reference_seconds"] / total
        worst = 0.0
        for d_m, d_r in zip(tr.atmosphere.densityDistributionList, ref.atmosphere.densityDistributionList):
            for c_m, c_r in zip(d_m.constituents, d_r.constituents):
                if not c_m.isMolecule:
                    assert np.array_equal(np.asarray(c_m.lookupFunction.x), c_r.lookupFunction.x)
                    s_m, s_r = 10 ** np.asarray(c_m.lookupFunction.y), 10 ** c_r.lookupFunction.y
                    big = s_r > 1e-40
                    if big.any():
                        worst = max(worst, float(np.max(np.abs(s_m[big] - s_r[big]) / s_r[big])))
        out["max_rel_diff_sigma_tables"] = worst
    return out


def retrieval_workload(n_models=64, steps=5, warmup=3, cpu_models=2):
    """Retrieval batch: forward models/s of the mainRetrieval.py setup (RET: 60 000 chords x 25
    phases), n_models parameter vectors per call, end to end through RetrievalBatch.run -- host flattening of every
    model, H2D of the stacked per-phase arrays, one set of launches, light-curve kernel, D2H of the light curves --
    next to the oracle's time for the same forward model + light curve on one host core."""
    import prometheus_b200 as pb
    from prometheus_b200 import retrieval, workloads
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from helpers import oracle, oracle_inputs
    spec = workloads.get("RET")
    template = pb.build_transit(spec, nist_rows=nist_rows())
    orb = template.spatialGrid.constructOrbphaseAxis()
    band = retrieval.LightcurveBand.for_planet(template.planet, orb, spec["band"]["line_centres"], spec["band"]["bandwidth"])
    params = workloads.retrieval_samples(n_models)
    batch = retrieval.RetrievalBatch(template)
    info = batch.info
    wall, timings = [], []
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        family, K = retrieval.make_family(template, params)
        lc, _, t = batch.run_family(family, K, band)
        if i >= warmup:
            wall.append(time.perf_counter() - t0)
            timings.append(t)
    batch.reset()
    # CPU oracle: the same forward model and light curve for the first cpu_models parameter vectors
    t_cpu, err = [], 0.0
    for k in range(cpu_models):
        v = workloads.retrieval_variant(spec, *params[k])
        inp = oracle_inputs(v, nist_rows())
        con = template.atmosphere.densityDistributionList[0].constituents[0]
        inp["tables"] = {(0, con.species.name): (con.lookupFunction.x, con.lookupFunction.y)}
        m = oracle.build_model(v, inp)
        t0 = time.perf_counter()
        R = oracle.sum_over_chords(m, batch_size=oracle.reference_batch_size(info["n_chords"], info["n_wav"], info["n_x"], False))
        lc_ref = oracle.lightcurve(R, m.wavelength, oracle.band_shift(m), spec["band"]["line_centres"], spec["band"]["bandwidth"])
        t_cpu.append(time.perf_counter() - t0)
        err = max(err, float(np.max(np.abs(lc_ref - lc[k]))))
    sec = float(np.mean(wall))
    return {"title": spec["title"], "n_models": n_models, "n_chords": info["n_chords"], "n_wav": info["n_wav"],
            "models_per_s_e2e": n_models / sec, "ms_per_batch_e2e": sec * 1e3,
            "samples_per_s_e2e": n_models * float(info["n_chords"]) * info["n_wav"] / sec,
            "kernels_ms": {k: float(np.mean([t[k] for t in timings])) for k in ("upload_ms", "column_ms", "transit_ms", "download_ms")},
            "launches_per_batch": int(timings[-1]["launches"]) + 3,
            "h2d_bytes_per_model": int(info["n_orb"] * 8 * (2 + 2 + 4 + 8)), "d2h_bytes_per_model": int(info["n_orb"] * 8),
            "cpu_oracle_models_per_s_one_core": 1.0 / float(np.mean(t_cpu)),
            "max_abs_diff_lightcurve_gpu_vs_cpu": err}


def production_launch(kernel_ms, exp_samples, centre_pairs, fps, peak_tflops, workload):
    """The default launch of the headline workload: what it executed (library counters) against the FP64 peak."""
    flops = fps * (exp_samples - centre_pairs) + FLOPS_CENTRE_PAIR * centre_pairs
    ach = flops / (kernel_ms * 1e-3) / 1e12
    cells = centre_pairs > 0.0
    return {"kernel": "cell_moment_kernel + cell_eval_kernel (sigma cells)" if cells else "transit_kernel (c')",
            "kernel_ms": kernel_ms, "exp_samples": exp_samples - centre_pairs, "centre_pairs": centre_pairs, "flops": flops,
            "achieved": ach, "frac": ach / peak_tflops,
            "traffic": recorded_traffic(workload, "cell_moment_kernel" if cells else "transit_kernel_production"),
            "note": ("one table exp + 13 Taylor sums per (chord, sigma cell), 49 flops; the per-wavelength Horner evaluation "
                     "(12 FMAs) is not counted" if cells else
                     "flops of the samples that take the table exp only; the thin-chord expansions (4 FP64 instructions per chord "
                     "and thread) are not counted") + "; this launch is short and latency-bound, see kernels_ms and DESIGN.md"}


def workload_roofline(name, tr, info, timings, peak_tflops):
    """FP64-pipe roofline of the PRODUCTION launch of a workload's dominant kernel, from the library's own counters:
    atomic scenes -- kernel (c'): samples that took the table exp x flops per sample; scenes with molecules -- kernel (m)
    in moment form: items x wavelengths x 33 + blocks x wavelengths x (15 + 2 per atomic constituent)."""
    ms = float(np.mean([t["transit_ms"] for t in timings]))
    exp_s = float(np.mean([t["exp_samples"] for t in timings]))
    exe_s = float(np.mean([t["executed_samples"] for t in timings]))
    n_abs = sum(1 for d in tr.atmosphere.densityDistributionList if any(not c.isMolecule for c in d.constituents))
    if info["n_mols"]:      # the atoms of a scenario are ONE absorber in kernel (m) too (MolLayout::grouped): one FMA per group
        flops = FLOPS_MOMENT_ITEM * exp_s + (FLOPS_MOMENT_BLOCK + 2.0 * n_abs) * exe_s
        return fp64_roofline("transit_moment_kernel (m: molecular tau in moment form + atoms + exp + disk sum)", flops, ms, peak_tflops,
                             item_wavelengths=exp_s, block_wavelengths=exe_s, flops_per_item=FLOPS_MOMENT_ITEM,
                             flops_per_block=FLOPS_MOMENT_BLOCK + 2.0 * n_abs,
                             traffic=recorded_traffic(name, "transit_moment_kernel"))
    fps = executed_flops_per_sample(n_abs)
    pairs = float(np.mean([t.get("centre_pairs", 0) for t in timings]))
    if pairs > 0.0:     # sigma cells (or narrow sorted windows): one table exp + the Taylor sums per (chord, cell)
        flops = fps * (exp_s - pairs) + FLOPS_CENTRE_PAIR * pairs
        return fp64_roofline("cell_moment_kernel + cell_eval_kernel (c': R tabulated per phase on sigma cells), production launch",
                             flops, ms, peak_tflops, centre_pairs=pairs, flops_per_pair=FLOPS_CENTRE_PAIR,
                             exp_samples=exp_s - pairs, flops_per_sample=fps, executed_samples=exe_s,
                             note="the work no longer scales with the wavelength count: `executed_samples` (records x wavelengths) is what "
                                  "the per-sample kernel would have walked; the launch is short and partly latency-bound, see DESIGN.md 6b",
                             traffic=recorded_traffic(name, "cell_moment_kernel"))
    return fp64_roofline("transit_kernel (c'), production launch", fps * exp_s, ms, peak_tflops, exp_samples=exp_s,
                         flops_per_sample=fps, executed_samples=exe_s,
                         note="flops of the samples that take the table exp; thin-chord expansions and skipped (opaque) chords not counted",
                         traffic=recorded_traffic(name, "transit_kernel_production"))


def strong_scaling(name, torch, dist, ctx, flush, dev, rank, world, local, peak_tflops, steps=5, warmup=3):
    """Fixed-size (strong) scaling of one BASELINE config: the UNCHANGED workload split over the `world` ranks --
    block-cyclic wavelength tiles for atomic scenes (C5), contiguous wavelength chunks with trimmed sigma tables for
    scenes that run kernel (m) (C2).  `ms_per_step`: device-resident, CUDA events, max over ranks, INCLUDING the NCCL
    all-gather of the R tiles over NVLink and their placement into one (n_orb, n_wav) device array on every rank;
    `ms_no_gather` without; `ms_one_gpu`: the whole job on one GPU in the same run (max over ranks);
    efficiency = ms_one_gpu / (N * ms_per_step).  Parity: rank 0 checks the gathered result against the oracle."""
    import prometheus_b200 as pb
    from prometheus_b200 import distributed as pbd, scene
    spec = scaled_spec(name, 1)
    t0 = time.perf_counter()
    tr = pb.build_transit(spec, nist_rows=nist_rows())
    t_setup = time.perf_counter() - t0
    sc, keep, info = scene.flatten(tr)
    n_orb, n_wav, n_c = info["n_orb"], info["n_wav"], info["n_chords"]
    atomic = info["n_mols"] == 0 and not (sc.n_fstar > 0 and sc.vsini != 0.0)
    axis = "cyclic" if atomic else "wavelength"

    def timed(fn, n):
        ms = []
        for i in range(warmup + n):
            flush.zero_()
            if world > 1:
                dist.barrier()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            fn()
            b.record()
            torch.cuda.synchronize(dev)
            if i >= warmup:
                ms.append(a.elapsed_time(b))
        t = torch.tensor([float(np.mean(ms))], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # the whole job on one GPU (every rank does it; max over ranks)
    ctx.set_scene(sc, keep)
    whole = torch.empty(n_orb * n_wav, dtype=torch.float64, device=dev)
    kern1 = []
    ms_one = timed(lambda: kern1.append(ctx.run_device(0, n_orb, 0, n_wav, whole.data_ptr(), want_timing=True)), steps)
    res = {"title": spec["title"], "n_chords": n_c, "n_wav": n_wav, "axis": axis, "ms_one_gpu": ms_one,
           "kernels_ms_one_gpu": {k: float(np.mean([t[k] for t in kern1[warmup:]])) for k in ("column_ms", "transit_ms", "finalize_ms")},
           "setup_seconds": t_setup, "roofline_one_gpu": workload_roofline(name, tr, info, kern1[warmup:], peak_tflops)}
    if world == 1:
        res.update(ms_per_step=ms_one, ms_no_gather=ms_one, efficiency=1.0)
        R = whole.view(n_orb, n_wav).cpu().numpy()
    else:
        def measure(axis):
            if axis == "wavelength":     # this rank's chunk only needs the part of each sigma table its wavelengths reach
                sc_r, keep_r, _ = scene.flatten(tr, wav_range=pbd.partition(n_wav, world)[rank])
                ctx.set_scene(sc_r, keep_r)
            else:
                ctx.set_scene(sc, keep)
            st = pbd.ShardedTransit(n_orb, n_wav, None, axis, dev)
            kern = []
            st.compute_tile = pbd.tile_function(ctx, axis, world, rank, want_timing=True, sink=kern)
            ms_ng = timed(lambda: st.run_device(gather=False), steps)
            mine = {k: (float(np.mean([t[k] for t in kern[warmup:]])) if len(kern) > warmup else None)
                    for k in ("column_ms", "transit_ms", "finalize_ms")}
            st.compute_tile = pbd.tile_function(ctx, axis, world, rank)

            def step():
                st.run_device(gather=True)
                st.unpack_device()
            ms_n = timed(step, steps)
            per_rank = [None] * world
            dist.all_gather_object(per_rank, mine)
            return st, {"axis": axis, "ms_per_step": ms_n, "ms_no_gather": ms_ng, "efficiency": ms_one / (world * ms_n),
                        "efficiency_no_gather": ms_one / (world * ms_ng), "kernels_ms_per_rank": per_rank}
        # atomic scenes: wavelength tiles dealt round-robin (every rank still walks every chord of every phase) or whole
        # phases per rank (partitions ALL the work, but n_orb / N phases per rank may be uneven): measured both, the
        # faster one is reported on top, the other under `other_axis`
        axes = ["cyclic", "phase"] if (atomic and n_orb >= world) else [axis]
        tried = [measure(ax) for ax in axes]
        tried.sort(key=lambda t: t[1]["ms_per_step"])
        st, best = tried[0]
        axis = best["axis"]
        if len(tried) > 1:
            best["other_axis"] = {k: v for k, v in tried[1][1].items() if k != "kernels_ms_per_rank"}
            st, _ = measure(axis)        # leaves the context and the buffers in the state of the reported axis
        state = {}
        wall = []
        for i in range(2 + 3):
            dist.barrier()
            torch.cuda.synchronize(dev)
            t0 = time.perf_counter()
            R = pbd.sum_over_chords(tr, None, axis, local, 0, state)     # host buffers in, numpy out on rank 0
            wall.append((time.perf_counter() - t0) * 1e3)
        e2e = torch.tensor([float(np.mean(wall[2:]))], dtype=torch.float64, device=dev)
        dist.all_reduce(e2e, op=dist.ReduceOp.MAX)
        res.update(best)
        res.update(e2e_ms_per_call=float(e2e.item()), gather_bytes=int(n_orb * n_wav * 8))
        st.run_device(gather=True)
        R = st.to_host(0)
    if rank == 0:
        res["samples_per_s"] = float(n_c) * n_wav / (res["ms_per_step"] * 1e-3)
        res["parity_check"] = parity_check(spec, tr, R, 8)
        assert res["parity_check"]["worst_err_over_bound"] <= 1.0, res["parity_check"]
    return res


def run_b200(args):
    import torch
    import torch.distributed as dist
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py: no CUDA device; the transit hot path has no CPU implementation to fall back on")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    import prometheus_b200 as pb
    from prometheus_b200 import distributed as pbd, host, scene

    host.set_default_device(local)
    ctx = host.get_context(local)
    stream = torch.cuda.Stream(dev)
    spec = scaled_spec(args.workload, world)
    with torch.cuda.stream(stream):
        ctx.set_stream(stream.cuda_stream)
        probe = ctx.probe_fp64()
        peak_dfma = probe["dfma_tflops"]
        t0 = time.perf_counter()
        tr = pb.build_transit(spec, nist_rows=nist_rows())            # sigma tables: kernel (b2), set-up time
        t_setup = time.perf_counter() - t0
        voigt = ctx.last_timing()
        sc, keep, info = scene.flatten(tr)
        ctx.set_scene(sc, keep)
        n_orb, n_wav, n_c, n_atoms, n_mols = info["n_orb"], info["n_wav"], info["n_chords"], info["n_atoms"], info["n_mols"]
        # N > 1: block-cyclic wavelength split for atomic scenes (line cores and continuum spread evenly over the
        # ranks); contiguous chunks for scenes that run kernel (m)
        axis = "cyclic" if (world > 1 and n_mols == 0 and not (sc.n_fstar > 0 and sc.vsini != 0.0)) else "wavelength"
        st = pbd.ShardedTransit(n_orb, n_wav, None, axis, dev)
        tile_t = []
        st.compute_tile = pbd.tile_function(ctx, axis, world, rank, want_timing=True, sink=tile_t)
        flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
        samples_total = float(n_c) * float(n_wav)

        def barrier():
            if world > 1:
                dist.barrier()
            torch.cuda.synchronize(dev)

        # ---- device-resident timing --------------------------------------------------------------
        # The timed steps are enqueued WITHOUT asking for the library's timing record (that costs a host synchronisation
        # per step and would put every step's launch latency into the measurement); they are bracketed by CUDA events on
        # the launching stream, the L2 flush between them.  The per-kernel breakdown comes from instrumented steps after.
        tile_inst = st.compute_tile
        st.compute_tile = pbd.tile_function(ctx, axis, world, rank)
        for _ in range(args.warmup):
            st.run_device(gather=False)
        barrier()
        launches0 = ctx.launch_count()
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
        clocks = ClockSampler(local)
        clocks.start()
        for a, b in ev:
            flush.zero_()
            a.record()
            st.run_device(gather=False)       # the R tiles stay on the GPUs that computed them (independent units)
            b.record()
        barrier()
        launches = ctx.launch_count() - launches0
        st.compute_tile = tile_inst
        tile_t.clear()
        for _ in range(5):
            flush.zero_()
            st.run_device(gather=False)
        barrier()
        step_ms = [a.elapsed_time(b) for a, b in ev]
        total_ms = torch.tensor([sum(step_ms)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(total_ms, op=dist.ReduceOp.MAX)
        ms_per_step = float(total_ms.item()) / args.steps
        kernel_ms = float(np.mean([t["transit_ms"] for t in tile_t]))
        column_ms = float(np.mean([t["column_ms"] for t in tile_t]))
        finalize_ms = float(np.mean([t["finalize_ms"] for t in tile_t]))
        st.run_device()                       # gathered copy for the comparison with the e2e result below
        barrier()
        executed = float(np.mean([t["executed_samples"] for t in tile_t]))
        exp_samples = float(np.mean([t["exp_samples"] for t in tile_t]))
        centre_pairs = float(np.mean([t.get("centre_pairs", 0) for t in tile_t]))
        R_dev = st.to_host(0)

        # ---- the same steps with the transparent-chord culling switched off (every unblocked chord
        #      goes through the exp, as in the reference's loop): reported next to the headline ----------
        for opt in ("cull", "mirror", "thin"):
            ctx.set_option(opt, 0)
        ctx.set_scene(sc, keep)
        tile_t.clear()
        for _ in range(2):
            st.run_device()
        barrier()
        tile_t.clear()
        ev_nc = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(max(3, args.steps // 4))]
        for a, b in ev_nc:
            flush.zero_()
            a.record()
            st.run_device(gather=False)
            b.record()
        barrier()
        clk = clocks.stop()                   # sampled over both timed loops (production steps and the brute-force launches)
        nc_total = torch.tensor([sum(a.elapsed_time(b) for a, b in ev_nc)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(nc_total, op=dist.ReduceOp.MAX)
        nc_ms_per_step = float(nc_total.item()) / len(ev_nc)
        nc_kernel_ms = float(np.mean([t["transit_ms"] for t in tile_t]))
        nc_executed = float(np.mean([t["executed_samples"] for t in tile_t]))
        st.run_device()
        barrier()
        R_nc = st.to_host(0)
        for opt in ("cull", "mirror", "thin"):
            ctx.set_option(opt, 1)
        ctx.set_scene(sc, keep)
        tile_t.clear()
        st.run_device()
        barrier()

        # ---- end to end through the host API (host buffers, H2D + D2H inside) ---------------------
        state = {"timeline": True}
        e2e_ms = []
        R_e2e = None
        for i in range(args.warmup + args.steps):
            flush.zero_()
            barrier()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            if world == 1:
                R_e2e = tr.sumOverChords()                 # the user-facing call: numpy in, numpy out
            else:
                if axis == "cyclic" and not args.e2e_gather:
                    R_e2e = pbd.sum_over_chords_shared(tr, None, local, state)    # tiles stored straight into shared pinned memory
                else:
                    R_e2e = pbd.sum_over_chords(tr, None, axis, local, 0, state)   # all-gather + D2H on rank 0
            b.record()
            barrier()
            if i >= args.warmup:
                e2e_ms.append(a.elapsed_time(b))
        e2e_total = torch.tensor([sum(e2e_ms)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(e2e_total, op=dist.ReduceOp.MAX)
        e2e_ms_per_step = float(e2e_total.item()) / args.steps
        # ---- per-rank host timeline of the last e2e call, fixed-size scaling block (all ranks take part) --------
        breakdown = None
        if world > 1:
            mine = dict(state.get("host_ms") or {})
            mine["rank"] = rank
            breakdown = [None] * world
            dist.all_gather_object(breakdown, mine)
        strong = {}
        for name in [w for w in args.strong.split(",") if w]:
            strong[name] = strong_scaling(name, torch, dist, ctx, flush, dev, rank, world, local, peak_dfma)
        ctx.set_scene(sc, keep)
    # ---- parity of the benchmarked result (rank 0; the others wait in the collective below) -------------------
    pcheck = parity_check(spec, tr, R_e2e, 8) if rank == 0 else None
    h2d_local = state["info"]["h2d_bytes"] if (world > 1 and "info" in state) else info["h2d_bytes"]
    launch_total = torch.tensor([launches, h2d_local], dtype=torch.int64, device=dev)
    if world > 1:
        dist.all_reduce(launch_total)
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    assert R_dev.shape == (n_orb, n_wav) and np.all(np.isfinite(R_dev)) and np.array_equal(R_dev, R_e2e)
    cull_diff = float(np.max(np.abs(R_dev - R_nc)))
    assert cull_diff < 1e-12, cull_diff
    assert pcheck["worst_err_over_bound"] <= 1.0, pcheck

    hbm_peak, peak_src = measured_peaks()
    tile_samples = float(n_c) * float(st.samples_tile_shape[1])          # nominal samples of this rank's tile
    if n_mols:
        executed = nc_executed = exp_samples = tile_samples                 # kernel (m) has no culling
    n_abs = sum(1 for d in tr.atmosphere.densityDistributionList if any(not c.isMolecule for c in d.constituents))
    fps = executed_flops_per_sample(n_abs)
    # The bound of kernel (c') is the FP64 pipe (31 flops against 0.1 B per sample).  The headline figure is taken on the
    # launch that does the exp on EVERY unblocked sample (pb200_set_option CULL = MIRROR = THIN = 0: the kernel whose
    # per-sample work SURVEY 8(d) counts); the production launch of this workload -- most samples culled or summed as
    # thin-chord expansions -- is described beside it, and for the configs where a kernel dominates the production
    # launch itself carries the roofline (other_workloads.C5 / .C2).
    staged = BYTES_PER_SAMPLE_STAGED * nc_executed / (nc_kernel_ms * 1e-3) / 1e9
    roof = fp64_roofline("transit_kernel (c': tau + exp(-tau) + disk sum), every unblocked sample through the table exp",
                         fps * nc_executed, nc_kernel_ms, probe["dfma_tflops"],
                         flops_per_sample=fps, samples_per_launch=nc_executed,
                         flops_per_sample_survey_convention=flops_per_sample(n_abs),
                         launch="pb200_set_option CULL = MIRROR = THIN = 0 (measurement launch; same kernel, same tiles)",
                         traffic=recorded_traffic(args.workload),
                         traffic_note="dram bytes per launch from the committed ncu capture (profiles/traffic.json); below the algorithmic bytes: R stays in L2",
                         pipe_frac=((6.0 if n_abs == 1 else 6.0 + n_abs) * nc_executed / 32.0 * 2.0) /
                                   (148 * 4 * (clk["sm_mhz"] or 1965.0) * 1e6 * nc_kernel_ms * 1e-3) if n_mols == 0 else None,
                         pipe_frac_note="issued FP64 warp instructions (6 per sample for one absorber, else 6+S) x 2 issue cycles / SMSP cycles",
                         exp_probe_gops=probe["exp_gops"], exp_frac=nc_executed / (nc_kernel_ms * 1e-3) / 1e9 / probe["exp_gops"],
                         staged_hbm_equivalent={"achieved": staged, "peak": hbm_peak, "unit": "GB/s", "frac": staged / hbm_peak,
                                                "peak_source": peak_src,
                                                "convention": "the north star's staged-equivalent 16 B per chord*lambda sample (one f64 tau store + "
                                                              "one load, SURVEY 8(d)); tau is never staged, so this exceeds 1 by construction and "
                                                              "is not a physical roofline"},
                         production_launch=production_launch(kernel_ms, exp_samples, centre_pairs, fps, probe["dfma_tflops"], args.workload))
    out = {
        "metric": "chord*lambda samples/s", "value": samples_total / (ms_per_step * 1e-3), "unit": "samples/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(args.workload, spec, n_c, n_wav, world),
        "runs_per_s": 1e3 / ms_per_step,
        "e2e": {"value": samples_total / (e2e_ms_per_step * 1e-3), "unit": "samples/s", "ms_per_step": e2e_ms_per_step,
                "h2d_bytes_per_step": int(launch_total[1].item()), "d2h_bytes_per_step": int(n_orb * n_wav * 8),
                "runs_per_s": 1e3 / e2e_ms_per_step},
        "parity_check": pcheck,
        "e2e_breakdown_ms": (breakdown if world > 1 else
                             ({k: tr.last_timing[k] for k in ("upload_ms", "column_ms", "transit_ms", "finalize_ms", "download_ms")}
                              if tr.last_timing else None)),
        "strong": strong or None,
        "gpu_launches": int(launch_total[0].item()),
        "clocks": clk,
        "roofline": roof,
        "culling": {"nominal_samples_per_step": tile_samples * world, "executed_samples_per_rank_step": executed,
                    "executed_frac": executed / tile_samples,
                    "exp_samples_per_rank_step": exp_samples, "exp_frac": exp_samples / tile_samples,
                    "off": {"value": samples_total / (nc_ms_per_step * 1e-3), "ms_per_step": nc_ms_per_step,
                            "transit_kernel_ms": nc_kernel_ms, "executed_frac": nc_executed / tile_samples,
                            "max_abs_diff_R": cull_diff},
                    "note": "`value` counts the job's chord*lambda samples (BASELINE metric); chords whose optical depth is "
                            "below 2^-54 at every wavelength of the run contribute exactly w_in and skip the exp, and the chords "
                            "(phi, rho), (pi - phi, rho) -- equal column densities, the scene is symmetric in z -- share one record. "
                            "Of the remaining (`executed`) samples, those whose chord is optically thin over the warp's 256 wavelengths "
                            "(tau < 2.1e-5) are summed as a second-order expansion per chord; only `exp_samples` take the table exp. "
                            "With one absorber and >= 8192 wavelengths per run (C4, C5) the executed samples are not walked per wavelength "
                            "at all: R(p, .) is tabulated on sigma cells -- `exp_samples` then counts (chord, cell) pairs, one table exp "
                            "each (DESIGN.md 6b). `off` = same steps with pb200_set_option CULL = MIRROR = THIN = 0: every unblocked sample goes through the exp"},
        "kernels_ms": {"column": column_ms, "transit": kernel_ms, "finalize": finalize_ms},
        "setup": {"seconds": t_setup, "voigt_last_table_ms": voigt["transit_ms"], "voigt_last_table_evals": voigt["samples"]},
    }

    # ---- kernel (a) on its own: the materialised number density n(chord, x) of the headline scene (N = 1) ------
    if world == 1:
        with torch.cuda.stream(stream):
            ctx.set_scene(sc, keep)
            n_dens = n_c * info["n_x"]
            dens = torch.empty(n_dens, dtype=torch.float64, device=dev)
            d_ms, f_ms = [], []
            for i in range(8):
                flush.zero_()
                a, b, c2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
                a.record(); ctx.number_density_device(0, dens.data_ptr()); b.record()
                dens.fill_(0.0); c2.record()                       # write-only roofline of the same buffer
                torch.cuda.synchronize()
                if i >= 3:
                    d_ms.append(a.elapsed_time(b)); f_ms.append(b.elapsed_time(c2))
            d_t, f_t = float(np.mean(d_ms)), float(np.mean(f_ms))
            ka = (json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get("C4", {}).get("kernel_a", {})
                  if args.workload == "C4" else {})
            mhz = float((out.get("clocks") or {}).get("sm_mhz") or 1965.0)
            smsp = 4 * torch.cuda.get_device_properties(dev).multi_processor_count
            ach = 8.0 * n_dens / d_t / 1e6
            out["kernel_a"] = {"kernel": "number_density_rows_kernel (a: n(chord, x) materialised, scenario 0)",
                               "samples": n_dens, "ms": d_t, "bound": "issue", "achieved": ach, "unit": "GB/s",
                               "convention": "8 B per chord*x sample (one f64 store, SURVEY 8(d) row (a))",
                               "hbm": {"peak": hbm_peak, "frac": ach / hbm_peak, "write_only_fill_GBs": 8.0 * n_dens / f_t / 1e6,
                                       "traffic": ka.get("dram_bytes")}}
            if ka.get("warp_inst_per_32_samples"):     # ceilings from the committed ncu capture of this kernel (profiles/traffic.json)
                per_s = smsp * mhz * 1e6               # warp instructions per second the SMSPs can issue (1 per cycle)
                issue = 256.0 * per_s / ka["warp_inst_per_32_samples"] / 1e9
                fp64 = 256.0 * per_s / (2.0 * ka["fp64_inst_per_32_samples"]) / 1e9
                out["kernel_a"].update(peak=issue, frac=ach / issue, fp64_pipe_ceiling_GBs=fp64,
                                       warp_inst_per_32_samples=ka["warp_inst_per_32_samples"],
                                       fp64_inst_per_32_samples=ka["fp64_inst_per_32_samples"],
                                       peak_source="issue slots: SMSPs x clock / executed warp instructions per 32 samples (ncu, "
                                                   + ka.get("capture", "profiles/") + "); the FP64 pipe allows fp64_pipe_ceiling_GBs, "
                                                   "HBM hbm.peak: the kernel is instruction-bound, not bandwidth-bound (DESIGN.md 4(a'))")
            else:
                out["kernel_a"].update(peak=hbm_peak, frac=ach / hbm_peak, bound="hbm")
            del dens

    # ---- the other BASELINE configs, device-resident, few steps (N = 1 only; parity is in tests/) -----------
    if world == 1 and args.others:
        out["other_workloads"] = {}
        for name in [w for w in args.others.split(",") if w and w != args.workload]:
            with torch.cuda.stream(stream):
                out["other_workloads"][name] = quick_workload(name, ctx, torch, stream, flush, peak_dfma)
    if world == 1 and not args.no_retrieval:
        out["retrieval"] = retrieval_workload()

    if world == 1 and args.cpu_seconds > 0:
        out["setup_comparison"] = setup_comparison("C2")
    # ---- CPU baseline on the host cores of this box (rank 0, N = 1 only) -----------------------------
    if world == 1 and args.cpu_seconds > 0:
        tables = {}
        for i, d in enumerate(tr.atmosphere.densityDistributionList):
            for con in d.constituents:
                if not con.isMolecule:
                    tables[(i, con.species.name)] = (con.lookupFunction.x, con.lookupFunction.y)
        wl = tr.wavelength
        centre = wl.size // 2
        arm = CpuArm(args.workload, 1, tables)          # the timed scene's own sigma tables: times sumOverChords, not the set-up
        lo, hi, a, b = calibrate_sample(arm, wl, centre, n_c, args.cpu_seconds, floor=min(2048, wl.size))
        R_cpu, t_cpu = arm.run(wl[lo:hi])
        arm.close()
        err = float(np.max(np.abs(R_cpu - R_dev[:, lo:hi])))
        t_full = full_job_seconds(a, t_cpu, hi - lo, wl.size)
        out["cpu_baseline"] = {"value": n_c * float(wl.size) / t_full, "unit": "samples/s", "cores": arm.procs, "kind": arm.kind,
                               "sample": f"{hi - lo} contiguous wavelengths [{lo},{hi}) of {wl.size}, all {n_c} chords, "
                                         f"{t_cpu:.1f} s; {arm.describe()}; value = whole workload by SURVEY 8(d)'s sub-sampling "
                                         f"rule (wavelength-independent cost {a:.2f} s added once)",
                               "sample_value": n_c * float(hi - lo) / t_cpu, "full_run_seconds_extrapolated": t_full,
                               "max_abs_diff_R_gpu_vs_cpu_on_sample": err}
        if arm.kind == "reference":     # the oracle port on the same sample: how fast the restatement is next to the real thing
            port = CpuArm(args.workload, 1, tables, kind="port")
            _, t_port = port.run(wl[lo:hi])
            R_port, t_port = port.run(wl[lo:hi])
            port.close()
            out["cpu_baseline"]["port_over_reference_speed"] = t_cpu / t_port
            out["cpu_baseline"]["port_max_abs_diff_R_vs_reference"] = float(np.max(np.abs(R_port - R_cpu)))
    print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


def main():
    # stdout carries exactly ONE JSON line: anything libraries print there (NCCL's version banner) goes to stderr
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    sys.stdout = os.fdopen(real_stdout, "w")
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="C4")
    ap.add_argument("--others", default="C2,C3,C5", help="other BASELINE configs reported under other_workloads (N=1)")
    ap.add_argument("--no-retrieval", action="store_true", help="skip the retrieval-batch leg (N=1)")
    ap.add_argument("--strong", default="C5,C2", help="configs of the fixed-size (strong) scaling block, split over the N ranks")
    ap.add_argument("--e2e-gather", action="store_true", help="N>1 e2e through all-gather + rank-0 D2H instead of the shared result")
    ap.add_argument("--cpu-seconds", type=float, default=None,
                    help="CPU budget: cpu_baseline sample (default 15 s) / whole --impl reference run (default 120 s)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    if args.cpu_seconds is None:
        args.cpu_seconds = 15.0 if args.impl == "b200" else 120.0
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "b200" and world != args.gpus:
        if args.gpus == 1 or world > 1:
            raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torch.distributed.run --nproc-per-node {args.gpus}")
        # convenience: python bench.py --gpus N re-launches itself under torchrun
        import subprocess
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", "29517", os.path.abspath(__file__)] + sys.argv[1:]
        sys.stdout.flush()
        raise SystemExit(subprocess.call(cmd, stdout=real_stdout))   # the ranks' stdout is the real one, not our stderr alias
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
