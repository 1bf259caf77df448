"""The tabulated-density producer, SerpensExosphere.addInterpolatedDensity (gasProperties.py:548-583; SURVEY 8(f)
rank 4): particle histogram + Gaussian smoothing.  Golden = the reference's own method run on a seeded synthetic
particle file (tests/golden/make_golden.py SRPs).  Binning is index work: bit-exact.  The smoothing follows scipy's
summation order; tolerance 1e-13 relative to the table's maximum."""
import os
import tempfile

import numpy as np
import pytest

from helpers import oracle
from prometheus_b200 import builder, gas, workloads


def _particles():
    spec = workloads.get("SERs")
    return spec, workloads.synthetic_serpens_particles(spec["grid"])


def test_oracle_matches_reference(golden):
    g = golden("SRPs")
    spec, particles = _particles()
    for k, sigma in enumerate(g["sigmas"]):
        (gx, gy, gz), vals = oracle.serpens_density(particles[:, 0:3] * 1e2, spec["grid"], float(g["N"]), float(sigma))
        assert np.array_equal(gx, g["gx"]) and np.array_equal(gy, g["gy"]) and np.array_equal(gz, g["gz"])
        assert np.array_equal(vals, g[f"values_{k}"])


def test_gaussian_weights_are_scipys():
    from scipy.ndimage import gaussian_filter1d
    for sigma in (0.6, 1.3, 2.0):
        w = gas.gaussian_weights(sigma)
        delta = np.zeros(w.size)
        delta[w.size // 2] = 1.0
        assert np.array_equal(gaussian_filter1d(delta, sigma, mode="constant"), w)


@pytest.mark.gpu
def test_gpu_ingest_matches_reference(golden, nist_rows, capsys):
    g = golden("SRPs")
    spec, particles = _particles()
    tr = builder.build_transit(spec, nist_rows=nist_rows, build_tables=False)
    fd, path = tempfile.mkstemp(suffix=".txt")
    os.close(fd)
    try:
        np.savetxt(path, particles)
        for k, sigma in enumerate(g["sigmas"]):
            ex = gas.SerpensExosphere(path, float(g["N"]), tr.planet, float(sigma))
            ex.addInterpolatedDensity(tr.spatialGrid)
            printed = capsys.readouterr().out
            assert "particles represented in all cells" in printed and "between the planetary and the stellar limb" in printed
            gx, gy, gz = ex.InterpolatedDensity.grid
            assert np.array_equal(gx, g["gx"]) and np.array_equal(gy, g["gy"]) and np.array_equal(gz, g["gz"])
            ref = g[f"values_{k}"]
            vals = np.asarray(ex.InterpolatedDensity.values)
            if sigma == 0.0:
                assert np.array_equal(vals, ref)                    # binning + scaling: bit-exact
            else:
                assert np.max(np.abs(vals - ref)) <= 1e-13 * ref.max()
            total = float(printed.split("if any):")[1].split()[0])     # the first diagnostic total, in particles
            assert abs(total - ref.sum() * (gx[1] - gx[0]) * (gy[1] - gy[0]) * (gz[1] - gz[0])) <= 1e-12 * total
    finally:
        os.unlink(path)


@pytest.mark.gpu
def test_gpu_ingested_table_runs_through_the_transit(nist_rows):
    """The table produced on the GPU feeds the serpens scenario like an injected one."""
    spec, particles = _particles()
    tr = builder.build_transit(spec, nist_rows=nist_rows)
    fd, path = tempfile.mkstemp(suffix=".txt")
    os.close(fd)
    try:
        np.savetxt(path, particles)
        dist = tr.atmosphere.densityDistributionList[0]
        dist.filename, dist.sigmaSmoothing = path, 1.3
        dist.addInterpolatedDensity(tr.spatialGrid)
    finally:
        os.unlink(path)
    R = tr.sumOverChords()
    (gx, gy, gz), vals = oracle.serpens_density(particles[:, 0:3] * 1e2, spec["grid"], dist.N, 1.3)
    assert np.all(np.isfinite(R)) and R.min() < 1.0
    assert np.max(np.abs(np.asarray(dist.InterpolatedDensity.values) - vals)) <= 1e-13 * vals.max()


@pytest.mark.gpu
def test_gpu_ingest_edge_cases(nist_rows):
    """No particles at all, particles only outside the grid, particles exactly on edges, a filter wider than an axis."""
    from prometheus_b200 import host
    ctx = host.get_context()
    xb = np.linspace(0.0, 4.0, 5)
    yb = np.linspace(-1.0, 1.0, 4)
    zb = np.linspace(-1.0, 1.0, 3)
    yp, zp = 0.5 * (yb[1:] + yb[:-1]), 0.5 * (zb[1:] + zb[:-1])

    def run(pos, sigma):
        w = gas.gaussian_weights(sigma) if sigma > 0 else None
        vals, sums = ctx.serpens_density(pos, xb, yb, zb, 7.0, 3.0, w, yp, zp, 0.1, 10.0)
        from scipy.ndimage import gaussian_filter
        ref = np.histogramdd(np.asarray(pos, dtype=float).reshape(-1, 3), bins=[xb, yb, zb])[0] * 7.0 / 3.0
        if sigma > 0:
            ref = gaussian_filter(ref, sigma=sigma)
        return vals, ref, sums

    vals, ref, sums = run(np.empty((0, 3)), 0.0)
    assert np.array_equal(vals, ref) and vals.sum() == 0.0 and sums[0] == 0.0
    vals, ref, _ = run(np.array([[9.0, 0.0, 0.0], [-1.0, 0.0, 0.0], [1.0, 5.0, 0.0], [np.nan, 0.0, 0.0]]), 0.0)
    assert np.array_equal(vals, ref) and vals.sum() == 0.0
    edges = np.array([[0.0, -1.0, -1.0], [4.0, 1.0, 1.0], [1.0, yb[1], 0.0], [2.0, yb[2], zb[1]], [4.0, 0.0, 0.0]])
    vals, ref, _ = run(edges, 0.0)
    assert np.array_equal(vals, ref) and np.count_nonzero(vals) == np.count_nonzero(ref) and vals.sum() == ref.sum()
    vals, ref, _ = run(edges, 2.5)          # radius 10 > every axis length: repeated reflection
    assert np.max(np.abs(vals - ref)) <= 1e-14 * ref.max()
