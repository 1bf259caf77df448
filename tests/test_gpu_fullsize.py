"""Parity at BASELINE.json's full sizes (C3, C4, C5) through the C ABI.

The oracle cannot do 1e10 chord*lambda samples in seconds, but every wavelength column of R is
independent (gasProperties.py:1243-1258), so a strided handful of columns computed by the oracle on
ALL chords is a genuine full-size check of kernels (a') and (c').  Plus size-independent properties:
the opaque-disk limit, R <= R0, tile independence, determinism."""
import numpy as np
import pytest

from helpers import excess_depth_ok, oracle, oracle_inputs
from prometheus_b200 import builder, host, workloads

pytestmark = pytest.mark.gpu


def gpu_tables(tr):
    out = {}
    for i, d in enumerate(tr.atmosphere.densityDistributionList):
        for con in d.constituents:
            if not con.isMolecule:
                out[(i, con.species.name)] = (con.lookupFunction.x, con.lookupFunction.y)
    return out


@pytest.mark.parametrize("name,n_cols", [("C3", 12), ("C4", 10), ("C5", 4)])
def test_full_size_columns_against_oracle(name, n_cols, nist_rows):
    spec = workloads.get(name)
    tr = builder.build_transit(spec, nist_rows=nist_rows)
    R = tr.sumOverChords()
    n_orb, n_wav = R.shape
    # (>= 0: in the cores of C5's strongest lines every chord has tau > 745 and exp(-tau) is exactly 0 in float64, in
    # the reference too; the sigma-sorted kernel also drops chords with tau > 50 on a whole window, < 2e-22 of the sum)
    assert np.all(np.isfinite(R)) and R.max() <= 1.0 + 1e-15 and R.min() >= 0.0
    # columns: the deepest one (a line core), its neighbours, and a strided set
    core = int(np.argmin(R.min(axis=0)))
    sel = np.unique(np.clip(np.concatenate([[core - 1, core, core + 1], np.linspace(0, n_wav - 1, n_cols).astype(int)]), 0, n_wav - 1))
    inputs = oracle_inputs(spec, nist_rows)
    inputs["tables"] = gpu_tables(tr)          # same sigma tables on both sides: isolates (a') + (c') at full size
    m = oracle.build_model(spec, inputs)
    assert np.array_equal(m.wavelength, tr.wavelength)
    n_c = len(m.rho) * len(m.phi) * len(m.orb)
    bs = oracle.reference_batch_size(n_c, len(sel), len(m.x), False, 2.0)
    R_ref = oracle.sum_over_chords(m, batch_size=None if bs >= n_c else bs, wavelength=m.wavelength[sel])
    R0 = oracle.sum_over_chords(m, batch_size=None if bs >= n_c else bs, wavelength=m.wavelength[sel], opaque_only=True)
    ok, worst = excess_depth_ok(R[:, sel], R_ref, R0)
    assert ok, f"{name}: worst err/bound {worst:.3g}"
    assert np.max(np.abs(R[:, sel] - R_ref)) < 5e-13
    assert np.all(R[:, sel] <= R0 + 1e-14)
    assert float((R0 - R[:, sel]).max()) > 1e-4, "the selected columns must see real absorption"


def test_c4_properties(nist_rows):
    ctx = host.get_context(0)
    tr = builder.build_transit(workloads.get("C4"), nist_rows=nist_rows)
    R = tr.sumOverChords()
    again = tr.sumOverChords()
    assert np.array_equal(R, again)
    n_orb, n_wav = R.shape
    # wavelength- and phase-tiles (what the multi-GPU path computes per rank) agree with the full run
    for (p0, p1, w0, w1) in [(0, n_orb, n_wav // 2, n_wav // 2 + 4000), (7, 19, 0, n_wav), (n_orb - 1, n_orb, n_wav - 3, n_wav)]:
        tile, _ = ctx.run(p0, p1, w0, w1)
        np.testing.assert_allclose(tile, R[p0:p1, w0:w1], rtol=0, atol=5e-15)
    # opaque-disk limit at full size: N -> 0
    spec0 = workloads.get("C4")
    spec0["scenarios"][0]["N"] = 0.0
    tr0 = builder.build_transit(spec0, nist_rows=nist_rows)
    R0 = tr0.sumOverChords()
    assert np.all(R0 >= R - 1e-15)
    assert np.ptp(R0, axis=1).max() < 1e-15, "without absorber R does not depend on wavelength"
    # doubling the absorber can only deepen the transit
    spec2 = workloads.get("C4")
    spec2["scenarios"][0]["N"] *= 2.0
    R2 = builder.build_transit(spec2, nist_rows=nist_rows).sumOverChords()
    assert np.all(R2 <= R + 1e-15)


# ---- C2 at full size: kernel (m), both forms (gasProperties.py:925-936, :789-818) -----------------
def _c2_columns(R, wavelength, n_strided=6):
    """Oracle columns for C2: the maximum of the synthetic molecular band, the deepest atomic line core with its
    neighbours, and a strided set."""
    lam_A = wavelength * 1e8
    band = (1.0 + np.sin(lam_A / 3.0 + 1) ** 2) + (1.0 + np.sin(lam_A / 3.0 + 2) ** 2)      # workloads.synthetic_molecule_table
    core = int(np.argmin(R.min(axis=0)))
    n_wav = wavelength.size
    cols = np.concatenate([[int(np.argmax(band)), int(np.argmin(band)), core - 1, core, core + 1],
                           np.linspace(0, n_wav - 1, n_strided).astype(int)])
    return np.unique(np.clip(cols, 0, n_wav - 1))


def test_c2_full_size_columns_against_oracle(nist_rows):
    """configs[1]: 33 000 chords x 50 x-samples x 95 897 wavelengths, 8 atoms + 2 molecules.  All chords, >= 8 columns,
    the moment form (default) and the per-record form of kernel (m) against the oracle's scipy RGI path."""
    ctx = host.get_context(0)
    spec = workloads.get("C2")
    tr = builder.build_transit(spec, nist_rows=nist_rows)
    R = tr.sumOverChords()
    with ctx.options(mol_records=1):
        R_rec = tr.sumOverChords()
    assert np.all(np.isfinite(R)) and R.max() <= 1.0 and R.min() > 0.0
    sel = _c2_columns(R, tr.wavelength)
    assert sel.size >= 8
    inputs = oracle_inputs(spec, nist_rows)
    inputs["tables"] = gpu_tables(tr)
    m = oracle.build_model(spec, inputs)
    assert np.array_equal(m.wavelength, tr.wavelength)
    n_c = len(m.rho) * len(m.phi) * len(m.orb)
    bs = oracle.reference_batch_size(n_c, len(sel), len(m.x), True, 2.0)                    # memoryHandler.py:42-47, 64 B per point
    kw = dict(batch_size=None if bs >= n_c else bs, wavelength=m.wavelength[sel])
    R_ref = oracle.sum_over_chords(m, **kw)
    R0 = oracle.sum_over_chords(m, opaque_only=True, **kw)
    for label, got in (("moments", R[:, sel]), ("records", R_rec[:, sel])):
        ok, worst = excess_depth_ok(got, R_ref, R0)
        assert ok, f"C2 {label}: worst err/bound {worst:.3g}"
        assert np.max(np.abs(got - R_ref)) < 5e-13, (label, float(np.max(np.abs(got - R_ref))))
    assert np.max(np.abs(R - R_rec)) < 5e-13
    assert float((R0 - R_ref).max()) > 1e-3


def test_steep_molecular_table_takes_the_per_record_form(nist_rows):
    """A table whose log sigma jumps by 12 dex between two pressure nodes needs more than 64 sub-intervals for the
    9e-13 truncation bound of the moment expansion: kernel (m) must fall back to the per-record form by itself
    (mol_bins == 0) and still match scipy's RGI."""
    spec = workloads.get("C2s")
    P = np.array([1e-4, 1e8])
    T = np.linspace(100.0, 3400.0, 12)
    lam = np.linspace(5700e-8, 6100e-8, 2001)
    logsig = np.empty((2, 12, lam.size))
    logsig[0] = -34.0 + 0.3 * np.sin(lam * 1e8 / 5.0)[None, :]
    logsig[1] = logsig[0] + 12.0
    tr = builder.build_transit(spec, nist_rows=nist_rows)
    tr_steep = builder.build_transit(spec, nist_rows=nist_rows)
    for con in tr_steep.atmosphere.densityDistributionList[0].constituents:
        if con.isMolecule:
            con.lookupFunction = host.GridFunction((P, T, lam), logsig, np.log10(1e-50))
    tr.sumOverChords()
    tr.sumOverChords()
    launches_moment = tr.last_timing["launches"]
    tr_steep.sumOverChords()
    R = tr_steep.sumOverChords()
    assert tr_steep.last_timing["launches"] < launches_moment, "the steep table must not run the moment kernels"
    inputs = oracle_inputs(spec, nist_rows)
    inputs["tables"] = gpu_tables(tr_steep)
    for key in inputs["molecules"]:
        inputs["molecules"][key] = (P, T, lam, logsig)
    m = oracle.build_model(spec, inputs)
    sel = np.unique(np.concatenate([[int(np.argmin(R.min(axis=0)))], np.linspace(0, R.shape[1] - 1, 24).astype(int)]))
    R_ref = oracle.sum_over_chords(m, wavelength=m.wavelength[sel])
    R0 = oracle.sum_over_chords(m, wavelength=m.wavelength[sel], opaque_only=True)
    ok, worst = excess_depth_ok(R[:, sel], R_ref, R0)
    assert ok, worst
    assert np.max(np.abs(R[:, sel] - R_ref)) < 5e-13 and float((R0 - R_ref).max()) > 1e-4


# ---- C5: columns chosen by the kernels' own decisions ---------------------------------------------------------------
def test_c5_regime_boundary_columns_against_oracle(nist_rows):
    """One absorber, many wavelengths: by default R(p, .) is tabulated on sigma cells (cell_*_kernel: Taylor sums about the
    centre of every cell of a geometric sigma grid, the order-2 totals below "thin for every chord"); with
    PB200_OPT_CENTRE = 2 the sigma-sorted windows of kernel (c') run, which decide per (chord, window of 256 sorted
    wavelengths): order-2 expansion below tau = 2.1e-5, order-8 below 0.1, skipped above 50, table exp between
    (ThinSums, ThinHi, transit_tile_sorted), and Taylor sums about the centre of narrow windows / 32-wavelength groups
    (transit_tile_centre).  All the boundaries sit at intermediate sigma, where a "deepest + strided" pick does not look:
    for two phases take the first, a middle and the last window of every regime (from the device's own sort order) and
    the two ends and the middle of the first, a middle and the last sigma cell, and check BOTH paths against the oracle."""
    ctx = host.get_context(0)
    spec = workloads.get("C5")
    tr = builder.build_transit(spec, nist_rows=nist_rows)
    R = tr.sumOverChords()
    n_orb, n_wav = R.shape
    assert ctx.last_sort_order(n_orb) is None, "the cell path sorts nothing"
    with ctx.options(centre=2):
        R_sorted = tr.sumOverChords()
        order = ctx.last_sort_order(n_orb)
    assert order is not None, "PB200_OPT_CENTRE = 2: C5 runs the sigma-sorted windows"
    assert np.max(np.abs(R - R_sorted)) < 5e-14
    perm, nsig = order
    ncol = tr.getColumnDensityGrid()                               # [constituent][chord], chord = m*n_orb + p
    chi = [con.chi for con in tr.atmosphere.densityDistributionList[0].constituents]
    scale = 4096.0 / np.log(2.0)
    thin_v, hi_v, opaque_v = 0.125, 0.1 * scale, 50.0 * scale
    sel = []
    for p in (0, n_orb // 2):
        N = ncol[0].reshape(-1, n_orb)[:, p] / chi[0]              # the absorber's column sum_x n dx; mixing ratios live in sigma
        N = N[N > 0.0]
        N_lo, N_hi = N.min(), N.max()
        a = np.abs(nsig[p])
        n_win = -(-n_wav // 256)
        w = np.concatenate([a, np.full(n_win * 256 - n_wav, np.nan)]).reshape(n_win, 256)
        smw, smin = np.nanmax(w, axis=1), np.nanmin(w, axis=1)
        narrow = (smw - smin) <= (smw + smin) / 40.0      # kCentreEps: Taylor sums about the window's centre (transit_tile_centre)
        import warnings
        with np.errstate(invalid="ignore"), warnings.catch_warnings():
            warnings.simplefilter("ignore", RuntimeWarning)
            g = w.reshape(n_win, 8, 32)                    # warp g of the CTA owns the window's g-th group of 32 sorted wavelengths
            gmx, gmn = np.nanmax(g, axis=2), np.nanmin(g, axis=2)
            by_group = np.all((gmx - gmn) <= (gmx + gmn) / 40.0, axis=1) & ~np.isnan(g).any(axis=(1, 2))
        needs_exp = (N_hi * smw >= hi_v) & (N_lo * smin <= opaque_v)
        regimes = {
            "thin2": N_lo * smw < thin_v,
            "thin8": (N_lo * smw < hi_v) & (N_hi * smw >= thin_v),
            "opaque": N_hi * smin > opaque_v,
            "exp": needs_exp & ~narrow & ~by_group,        # line cores: per-sample table exp
            "centre": needs_exp & narrow,                  # narrow windows: one exp per (chord, window)
            "groups": needs_exp & ~narrow & by_group,      # narrow 32-wavelength groups: one exp per (chord, group)
        }
        for name, mask in regimes.items():
            idx = np.flatnonzero(mask)
            if name == "exp" and not idx.size:
                continue                                   # every line core of the phase resolved by its groups
            assert idx.size, f"phase {p}: no window in regime {name}"
            for win in (idx[0], idx[len(idx) // 2], idx[-1]):
                for pos in (win * 256, win * 256 + 128, win * 256 + 255):     # both ends (|x| = 1) and the middle of the window
                    sel.append(int(perm[p, min(pos, n_wav - 1)]))
        # sigma cells: a in [g^k, g^(k+1)), g = (1 + 3/40)/(1 - 3/40); used from "thin for every chord" (N_max a < 0.125) upwards
        per_octave = 1.0 / np.log2((1.0 + 3.0 / 40.0) / (1.0 - 3.0 / 40.0))
        live = a[(a * N_hi >= thin_v) & np.isfinite(a)]
        kk = np.floor(np.log2(live) * per_octave).astype(np.int64)
        cells = np.unique(kk)
        assert cells.size > 100
        pos_of = {v: i for i, v in enumerate(a)}
        for k in (cells[0], cells[len(cells) // 3], cells[len(cells) // 2], cells[-2], cells[-1]):
            members = live[kk == k]
            lo, hi = 2.0 ** (k / per_octave), 2.0 ** ((k + 1) / per_octave)
            x = (members - 0.5 * (lo + hi)) / (0.5 * (hi - lo))
            for target in (-1.0, 0.0, 1.0):
                sel.append(int(perm[p, pos_of[members[np.argmin(np.abs(x - target))]]]))
        below = a[(a * N_hi < thin_v)]
        if below.size:
            sel.append(int(perm[p, pos_of[below.max()]]))                      # the last wavelength that takes the order-2 totals
    sel = np.unique(np.array(sel))
    assert sel.size >= 16
    inputs = oracle_inputs(spec, nist_rows)
    inputs["tables"] = gpu_tables(tr)
    m = oracle.build_model(spec, inputs)
    n_c = len(m.rho) * len(m.phi) * len(m.orb)
    bs = oracle.reference_batch_size(n_c, len(sel), len(m.x), False, 2.0)
    kw = dict(batch_size=None if bs >= n_c else bs, wavelength=m.wavelength[sel])
    R_ref = oracle.sum_over_chords(m, **kw)
    R0 = oracle.sum_over_chords(m, opaque_only=True, **kw)
    for label, Rx in (("cells", R), ("sorted windows", R_sorted)):
        ok, worst = excess_depth_ok(Rx[:, sel], R_ref, R0)
        assert ok, f"C5 regime columns ({label}): worst err/bound {worst:.3g}"
        assert np.max(np.abs(Rx[:, sel] - R_ref)) < 5e-13, (label, float(np.max(np.abs(Rx[:, sel] - R_ref))))
