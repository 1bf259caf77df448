"""Generates the golden fixtures in tests/golden/ by running the UNMODIFIED reference.

Run in the build container only (it needs /root/reference; the GPU box does not have it):

    python tests/golden/make_golden.py [case ...]

For every case of prometheus_b200.workloads.GOLDEN_CASES it builds the reference's own objects
through the reference's public classes (builder.build_transit with a namespace that points at the
reference modules), calls the reference's Transit.sumOverChords and friends, and stores inputs and
outputs as <case>.npz.  h5py / astropy are absent from this image and are never reached on this
path, so two empty stub modules stand in for them.
"""
from __future__ import annotations

import os
import sys
import time
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)


def import_reference():
    for m in ["h5py", "astropy", "astropy.io", "astropy.io.fits"]:
        sys.modules.setdefault(m, types.ModuleType(m))
    sys.modules["astropy"].io = sys.modules["astropy.io"]
    sys.modules["astropy.io"].fits = sys.modules["astropy.io.fits"]
    if "/root" not in sys.path:
        sys.path.insert(0, "/root")
    import reference.pythonScripts.celestialBodies as bodies
    import reference.pythonScripts.constants as consts
    import reference.pythonScripts.gasProperties as gasprop
    import reference.pythonScripts.geometryHandler as geom
    return gasprop, bodies, geom, consts


class ReferenceNamespace:
    def __init__(self):
        from scipy.interpolate import RegularGridInterpolator, interp1d
        self.gas, self.bodies, self.geom, self.consts = import_reference()
        self.Species = self.consts.Species
        self.findSpecies = lambda name: self.consts.AvailableSpecies().findSpecies(name)
        self._interp1d, self._rgi = interp1d, RegularGridInterpolator
        self.nist_rows = np.array(self.gas.LINE_LIST, dtype=str)

        # SerpensExosphere.calculateNumberDensity is scalar-only in the reference and raises for the batch
        # call: evaluate the reference's own scalar method chord by chord.
        scalar = self.gas.SerpensExosphere.calculateNumberDensity

        def batch(obj, x, phi, rho, orbphase):
            if np.ndim(phi) == 0:
                return scalar(obj, x, phi, rho, orbphase)
            return np.stack([scalar(obj, x, phi[i], rho[i], orbphase[i]) for i in range(len(phi))])

        self.gas.SerpensExosphere.calculateNumberDensity = batch

    def set_line_list(self, rows):
        self.gas.LINE_LIST = np.asarray(rows, dtype=str)

    def table_1d(self, x, y):
        return self._interp1d(x, y, kind="linear")

    def table_nd(self, grid, values, fill_value=None, bounds_error=False):
        if bounds_error:
            return self._rgi(grid, values)
        return self._rgi(grid, values, bounds_error=False, fill_value=fill_value)


def run_case(ns, name, out_dir=HERE, write=True):
    from prometheus_b200 import builder, workloads
    spec = workloads.get(name)
    t0 = time.time()
    tr = builder.build_transit(spec, ns=ns, nist_rows=ns.nist_rows)
    t_build = time.time() - t0
    t0 = time.time()
    R = tr.sumOverChords(2.0)
    t_run = time.time() - t0

    # opaque-disk value R0: the reference's own sumOverChords with tau == 0
    atm = tr.atmosphere
    real = atm.getLOSopticalDepth_Batch
    atm.getLOSopticalDepth_Batch = lambda x, phi, rho, orb, wl, dx: np.zeros((len(phi), len(wl)))
    R0 = tr.sumOverChords(2.0)
    atm.getLOSopticalDepth_Batch = real

    grid = tr.spatialGrid
    chords = grid.getChordGrid()
    phi, rho, orb = chords[:, 0], chords[:, 1], chords[:, 2]
    x = grid.constructXaxis()
    n_c = len(chords)
    # a deterministic subset of chords for the heavier intermediate quantities
    sub = np.unique(np.linspace(0, n_c - 1, min(n_c, 512, max(32, 100000 // len(tr.wavelength)))).astype(np.int64))
    out = {
        "R": R, "R0": R0, "wavelength": tr.wavelength,
        "x": x, "rho": grid.constructRhoAxis(), "phi": grid.constructPhiAxis(), "orb": grid.constructOrbphaseAxis(),
        "delta_x": np.float64(grid.getDeltaX()), "chord_subset": sub,
        "chord_grid_head": chords[:64], "chord_grid_tail": chords[-64:],
    }
    # blocking mask exactly as sumOverChords computes it (gasProperties.py:1224-1230)
    y, z = rho * np.sin(phi), rho * np.cos(phi)
    blocked = np.sqrt((y - tr.planet.a * np.sin(orb)) ** 2 + z ** 2) < tr.planet.R
    for dist in atm.densityDistributionList:
        if dist.hasMoon:
            blocked |= (y - dist.moon.getPosition(orb)[1]) ** 2 + z ** 2 < dist.moon.R ** 2
    out["blocked"] = np.packbits(blocked)
    out["n_blocked"] = np.int64(blocked.sum())

    k = 0
    for i, dist in enumerate(atm.densityDistributionList):
        n_sub = dist.calculateNumberDensity(x, phi[sub], rho[sub], orb[sub])
        out[f"density_{i}"] = n_sub
        n_all = dist.calculateNumberDensity(x, phi, rho, orb)
        for con in dist.constituents:
            if con.isMolecule:
                continue
            out[f"ncol_{k}"] = np.sum(n_all * con.chi, axis=1) * grid.getDeltaX()
            tx, ty = con.lookupFunction.x, con.lookupFunction.y
            out[f"table_n_{k}"] = np.int64(len(tx))
            stride = max(1, len(tx) // 20000)
            out[f"table_stride_{k}"] = np.int64(stride)
            out[f"table_x_{k}"] = tx[::stride]
            out[f"table_y_{k}"] = ty[::stride]
            out[f"table_xsum_{k}"] = np.float64(np.sum(tx))
            k += 1
    out["tau_sub"] = atm.getLOSopticalDepth_Batch(x, phi[sub], rho[sub], orb[sub], tr.wavelength, grid.getDeltaX())
    # (no wall-clock numbers in the fixture: every array of it is reproducible, tests/test_oracle_golden.py regenerates C1)
    if not write:
        return out
    path = os.path.join(out_dir, f"{name}.npz")
    np.savez_compressed(path, **out)
    print(f"{name}: n_c={n_c} n_wav={len(tr.wavelength)} build {t_build:.1f}s run {t_run:.1f}s "
          f"R in [{R.min():.6f}, {R.max():.6f}] -> {os.path.getsize(path) / 1e6:.2f} MB", flush=True)


def reference_lightcurve_block():
    """The 'Extract lightcurve' statements of the reference's mainRetrieval.py (from `filterBandwidth = ...` to
    `lightcurve = np.array(lightcurve)`), as source text: the golden light curves are produced by executing the
    reference's own lines, not a restatement."""
    lines = open("/root/reference/mainRetrieval.py").read().split("\n")
    i0 = next(i for i, l in enumerate(lines) if l.startswith("filterBandwidth ="))
    i1 = next(i for i, l in enumerate(lines) if l.startswith("lightcurve = np.array(lightcurve)"))
    return "\n".join(lines[i0:i1 + 1])


def run_retrieval_case(ns, name="RETs", n_models=4, out_dir=HERE):
    """Retrieval forward models: the setup of mainRetrieval.py for n_models parameter vectors
    (mainRetrieval.py:18-21), R from the reference's sumOverChords, light curves from the reference script's own
    statements (their band centres are the script's Na D literals)."""
    from prometheus_b200 import builder, workloads
    spec = workloads.get(name)
    params = workloads.retrieval_samples(n_models)
    block = reference_lightcurve_block()
    Rs, lcs = [], []
    t0 = time.time()
    for pv in params:
        tr = builder.build_transit(workloads.retrieval_variant(spec, *pv), ns=ns, nist_rows=ns.nist_rows)
        R = tr.sumOverChords(2.0)
        env = {"np": np, "const": ns.consts, "W49b": tr.planet, "R": R, "wavelength": tr.wavelength,
               "orbphase": tr.spatialGrid.constructOrbphaseAxis()}
        exec(block, env)
        Rs.append(R)
        lcs.append(env["lightcurve"])
        shift = env["shift"]
    out = {"params": params, "R": np.stack(Rs), "lightcurve": np.stack(lcs), "wavelength": tr.wavelength,
           "orb": tr.spatialGrid.constructOrbphaseAxis(), "band_shift": np.asarray(shift),
           "line_centres": np.array([env["NaD2_center"], env["NaD1_center"]]),
           "bandwidth": np.float64(env["filterBandwidth"])}
    path = os.path.join(out_dir, f"{name}.npz")
    np.savez_compressed(path, **out)
    print(f"{name}: {n_models} models, R {out['R'].shape}, lightcurve in [{out['lightcurve'].min():.6f}, "
          f"{out['lightcurve'].max():.6f}], {time.time() - t0:.1f}s -> {os.path.getsize(path) / 1e6:.2f} MB", flush=True)


def run_serpens_ingest_case(ns, name="SRPs", out_dir=HERE):
    """SerpensExosphere.addInterpolatedDensity of the reference (gasProperties.py:548-583) on a seeded synthetic
    particle file, without and with Gaussian smoothing."""
    import tempfile
    from prometheus_b200 import builder, workloads
    spec = workloads.get("SERs")
    tr = builder.build_transit(spec, ns=ns, nist_rows=ns.nist_rows, build_tables=False)
    grid, planet = tr.spatialGrid, tr.planet
    particles = workloads.synthetic_serpens_particles(spec["grid"])
    fd, path = tempfile.mkstemp(suffix=".txt")
    os.close(fd)
    out = {"N": np.float64(1e33), "sigmas": np.array([0.0, 1.3, 0.6])}
    try:
        np.savetxt(path, particles)
        for k, sigma in enumerate(out["sigmas"]):
            ex = ns.gas.SerpensExosphere(path, float(out["N"]), planet, float(sigma))
            ex.addInterpolatedDensity(grid)
            out[f"values_{k}"] = np.asarray(ex.InterpolatedDensity.values)
            gx, gy, gz = ex.InterpolatedDensity.grid
            out["gx"], out["gy"], out["gz"] = np.asarray(gx), np.asarray(gy), np.asarray(gz)
    finally:
        os.unlink(path)
    p = os.path.join(out_dir, f"{name}.npz")
    np.savez_compressed(p, **out)
    print(f"{name}: table {out['values_0'].shape}, {int(round(out['values_0'].sum() / out['values_0'][out['values_0'] > 0].min()))} particles binned "
          f"of {len(particles)} -> {os.path.getsize(p) / 1e6:.2f} MB", flush=True)


def main(argv):
    from prometheus_b200 import workloads
    ns = ReferenceNamespace()
    ll = os.path.join(HERE, "linelist.npz")
    np.savez_compressed(ll, rows=ns.nist_rows)
    print("line list rows:", ns.nist_rows.shape)
    for name in (argv or workloads.GOLDEN_CASES + ["RETs", "SRPs"]):
        if name == "RETs":
            run_retrieval_case(ns, name)
        elif name == "SRPs":
            run_serpens_ingest_case(ns, name)
        else:
            run_case(ns, name)


if __name__ == "__main__":
    main(sys.argv[1:])
