"""GPU parity: the CUDA path (through the C ABI) against the committed reference outputs
(tests/golden/*.npz) and the CPU oracle on the same seeded inputs.

Tolerances (north star / SURVEY 8(c)):
  * chord/grid indexing and the blocked mask: bit-exact;
  * excess transit depth E = R0 - R:  |E - E_ref| <= 1e-10 + 1e-6 |E_ref|;
  * intermediates: Ncol rel 1e-12, density rel 1e-12, sigma lookup rel 1e-12 on an identical
    table, Voigt table rel 1e-9 where sigma > 1e-40.
"""
import numpy as np
import pytest

from helpers import excess_depth_ok, oracle, oracle_model
import prometheus_b200 as pb
from prometheus_b200 import builder, capi, host, workloads

pytestmark = pytest.mark.gpu

CASES = workloads.GOLDEN_CASES


@pytest.fixture(scope="module")
def ctx():
    return host.get_context(0)


@pytest.fixture(scope="module")
def transits(nist_rows, ctx):
    """Host-mirror Transit per case; sigma tables built by kernel (b2) on the GPU."""
    cache = {}

    def get(name):
        if name not in cache:
            cache[name] = builder.build_transit(workloads.get(name), nist_rows=nist_rows)
        return cache[name]

    return get


@pytest.fixture(scope="module")
def models(nist_rows):
    cache = {}

    def get(name):
        if name not in cache:
            cache[name] = oracle_model(name, nist_rows)
        return cache[name]

    return get


def with_oracle_tables(name, nist_rows, model):
    """Mirror Transit whose atomic sigma tables are the oracle's (scipy-built): isolates kernels
    (a') and (c') from the Voigt synthesis."""
    tr = builder.build_transit(workloads.get(name), nist_rows=nist_rows, build_tables=False)
    for dist, scn in zip(tr.atmosphere.densityDistributionList, model.scenarios):
        for con, ocon in zip(dist.constituents, scn.constituents):
            if not con.isMolecule:
                con.lookupFunction = host.TableFunction1D(ocon.tab_x, ocon.tab_y)
    return tr


# ---- the headline criterion --------------------------------------------------------------------
@pytest.mark.parametrize("name", CASES)
def test_R_matches_reference_end_to_end(name, transits, golden):
    """Transit.sumOverChords with GPU-built Voigt tables against the reference's R."""
    tr, g = transits(name), golden(name)
    R = tr.sumOverChords()
    assert R.shape == g["R"].shape and R.dtype == np.float64 and R.flags.c_contiguous
    ok, worst = excess_depth_ok(R, g["R"], g["R0"])
    assert ok, f"{name}: worst excess-depth error / bound = {worst:.3g}"
    assert tr.last_timing["launches"] >= 2


@pytest.mark.parametrize("name", CASES)
def test_R_with_identical_tables_is_tight(name, nist_rows, models, golden):
    tr, g = with_oracle_tables(name, nist_rows, models(name)), golden(name)
    R = tr.sumOverChords()
    assert np.max(np.abs(R - g["R"])) < 2e-13, float(np.max(np.abs(R - g["R"])))


# ---- a1/a4: indexing contract -----------------------------------------------------------------
@pytest.mark.parametrize("name", CASES)
def test_blocked_mask_and_weights_bit_exact(name, transits, models, golden):
    tr, g = transits(name), golden(name)
    blocked, weight = tr.getChordMasks()
    assert int(blocked.sum()) == int(g["n_blocked"])
    assert np.array_equal(np.packbits(blocked), g["blocked"])
    _, w_ref = oracle.chord_masks(models(name))
    assert np.array_equal(weight, w_ref)       # rho*clv in the reference's chord order


# ---- a6/a7/a8: density and column density --------------------------------------------------------
@pytest.mark.parametrize("name", CASES)
def test_number_density_and_column(name, transits, golden):
    tr, g = transits(name), golden(name)
    grid = tr.spatialGrid
    chords = grid.getChordGrid()
    sub = g["chord_subset"]
    x = grid.constructXaxis()
    blocked, _ = tr.getChordMasks()
    for i, dist in enumerate(tr.atmosphere.densityDistributionList):
        n = dist.calculateNumberDensity(x, chords[sub, 0], chords[sub, 1], chords[sub, 2])
        np.testing.assert_allclose(n, g[f"density_{i}"], rtol=1e-12, atol=1e-300)   # denormal tails
        one = dist.calculateNumberDensity(x, chords[sub[3], 0], chords[sub[3], 1], chords[sub[3], 2])
        assert one.shape == (x.size,)
        np.testing.assert_allclose(one, g[f"density_{i}"][3], rtol=1e-12, atol=1e-300)
        n_grid = tr.getNumberDensityGrid(i)
        np.testing.assert_allclose(n_grid[sub], g[f"density_{i}"], rtol=1e-12, atol=1e-300)
    ncol = tr.getColumnDensityGrid()
    act = ~blocked        # the reference evaluates only unblocked chords (gasProperties.py:1235-1240)
    for k in range(ncol.shape[0]):
        np.testing.assert_allclose(ncol[k][act], g[f"ncol_{k}"][act], rtol=1e-12, atol=1e-290)
        assert np.all(ncol[k][~act] == 0.0)


# ---- a8-a11: tau ----------------------------------------------------------------------------------
@pytest.mark.parametrize("name", CASES)
def test_optical_depth_rows(name, nist_rows, models, golden):
    tr, g = with_oracle_tables(name, nist_rows, models(name)), golden(name)
    blocked, _ = tr.getChordMasks()
    sub = g["chord_subset"]
    tau = tr.getLOSopticalDepthGrid()
    keep = ~blocked[sub]
    np.testing.assert_allclose(tau[sub][keep], g["tau_sub"][keep], rtol=2e-12, atol=1e-300)


@pytest.mark.parametrize("name", CASES)
def test_optical_depth_batch_of_caller_chords(name, nist_rows, models, golden):
    """Atmosphere.getLOSopticalDepth_Batch(x_grid, phi_batch, rho_batch, orbphase_batch, wavelength, delta_x) with the
    reference's signature (gasProperties.py:885) on the chords of the golden subset, in a scrambled order and with a
    wavelength subset: the golden rows are what the reference's own method returned for those chords."""
    tr, g = with_oracle_tables(name, nist_rows, models(name)), golden(name)
    grid = tr.spatialGrid
    chords = grid.getChordGrid()
    blocked, _ = tr.getChordMasks()
    sub = g["chord_subset"]
    keep = np.flatnonzero(~blocked[sub])
    rng = np.random.default_rng(2)
    pick = rng.permutation(keep)[: min(keep.size, 96)]
    cols = np.unique(np.linspace(0, tr.wavelength.size - 1, 57).astype(int))
    tau = tr.atmosphere.getLOSopticalDepth_Batch(grid.constructXaxis(), chords[sub[pick], 0], chords[sub[pick], 1],
                                                 chords[sub[pick], 2], tr.wavelength[cols], grid.getDeltaX())
    assert tau.shape == (pick.size, cols.size)
    np.testing.assert_allclose(tau, g["tau_sub"][pick][:, cols], rtol=2e-12, atol=1e-300)


# ---- a9: sigma lookup (n_interp_log) ---------------------------------------------------------------
def test_sigma_lookup_matches_np_interp(ctx, models):
    m = models("C5s")
    con = m.scenarios[0].constituents[0]
    rng = np.random.default_rng(5)
    lo, hi = con.tab_x[0], con.tab_x[-1]
    wl = np.concatenate([rng.uniform(lo * 0.999, hi * 1.001, 20000), con.tab_x[::37], [lo, hi, lo * 0.5, hi * 2, np.nan]])
    got = ctx.sigma_lookup(con.tab_x, con.tab_y, 1e-50, wl)
    ref = oracle.interp_log(wl, con.tab_x, con.tab_y, 1e-50)
    assert np.isnan(got[-1]) and np.isnan(ref[-1])
    np.testing.assert_allclose(got[:-1], ref[:-1], rtol=1e-12, atol=1e-60)
    # empty input and (n_orb, n_wav) shaped input
    assert ctx.sigma_lookup(con.tab_x, con.tab_y, 1e-50, np.empty(0)).size == 0
    wl2 = np.outer(np.linspace(0.9999, 1.0001, 7), m.wavelength)
    np.testing.assert_allclose(ctx.sigma_lookup(con.tab_x, con.tab_y, 1e-50, wl2),
                               oracle.interp_log(wl2, con.tab_x, con.tab_y, 1e-50), rtol=1e-12, atol=1e-60)


# ---- a10: Voigt synthesis ---------------------------------------------------------------------------
def test_voigt_profile_against_scipy(ctx):
    from scipy.special import voigt_profile
    rng = np.random.default_rng(11)
    n = 200000
    sigma = 10 ** rng.uniform(8, 10, n)
    y = 10 ** rng.uniform(-8, 1, n)                    # gamma/(sqrt2 sigma)
    gamma = y * sigma * np.sqrt(2.0)
    xs = np.concatenate([rng.uniform(-12, 12, n // 2), np.sign(rng.uniform(-1, 1, n - n // 2)) * 10 ** rng.uniform(-3, 5, n - n // 2)])
    x = xs * sigma * np.sqrt(2.0)
    got = ctx.voigt_profile(x, sigma, gamma)
    ref = voigt_profile(x, sigma, gamma)
    rel = np.abs(got - ref) / ref
    assert np.max(rel) < 1e-12, float(np.max(rel))
    # degenerate branches of scipy.special.voigt_profile
    xx = np.array([0.0, 1.0, -2.0, 0.5])
    np.testing.assert_allclose(ctx.voigt_profile(xx, 0.0, 1.5), voigt_profile(xx, 0.0, 1.5), rtol=1e-15)
    np.testing.assert_allclose(ctx.voigt_profile(xx, 2.0, 0.0), voigt_profile(xx, 2.0, 0.0), rtol=1e-14)


@pytest.mark.parametrize("name", ["C1", "C5s", "C2s"])
def test_voigt_tables_against_reference(name, transits, golden):
    tr, g = transits(name), golden(name)
    k = 0
    for dist in tr.atmosphere.densityDistributionList:
        for con in dist.constituents:
            if con.isMolecule:
                continue
            lf = con.lookupFunction
            assert lf.x.size == int(g[f"table_n_{k}"])
            st = int(g[f"table_stride_{k}"])
            assert np.array_equal(lf.x[::st], g[f"table_x_{k}"])
            assert np.sum(lf.x) == float(g[f"table_xsum_{k}"])
            s_got, s_ref = 10 ** lf.y[::st], 10 ** g[f"table_y_{k}"]
            big = s_ref > 1e-40
            if big.any():      # a species without lines in the range has sigma == 0 everywhere
                assert np.max(np.abs(s_got[big] - s_ref[big]) / s_ref[big]) < 1e-9
            np.testing.assert_allclose(lf.y[::st], g[f"table_y_{k}"], rtol=0, atol=1e-9)
            k += 1


def test_calculateVoigtProfile_signature(transits, models):
    tr, m = transits("C1"), models("C1")
    con = tr.atmosphere.densityDistributionList[0].constituents[0]
    wl = tr.wavelength
    got = con.calculateVoigtProfile(wl)
    el, ion, _ = oracle.SPECIES["NaI"]
    lam0, gamma, f = oracle.line_parameters(host.LINE_LIST, el, ion, wl)
    ref = oracle.voigt_sigma(lam0, gamma, f, con.sigma_v, wl)
    np.testing.assert_allclose(got, ref, rtol=1e-11)
    np.testing.assert_allclose(con.getSigmaAbs(wl), oracle.interp_log(wl, con.lookupFunction.x, con.lookupFunction.y, 1e-50),
                               rtol=1e-12)


# ---- a11: molecular lookup ----------------------------------------------------------------------------
def test_molecular_sigma_against_scipy_rgi(ctx, transits, models):
    tr, m = transits("C2s"), models("C2s")
    dist = tr.atmosphere.densityDistributionList[0]
    mol = [c for c in dist.constituents if c.isMolecule][0]
    omol = [c for c in m.scenarios[0].constituents if c.kind == "molecule"][0]
    rng = np.random.default_rng(3)
    n_c, n_x, n_wav = 6, 9, 301
    P = 10 ** rng.uniform(-6, 9, (n_c, n_x))          # below the clip, inside, above the grid
    wl = np.sort(rng.uniform(2850e-8, 8150e-8, (n_c, n_wav)), axis=1)
    got = mol.getSigmaAbs(P, 3000.0, wl)
    ref = oracle.molecular_sigma(omol, P, 3000.0, wl)
    assert got.shape == (n_c, n_x, n_wav)
    np.testing.assert_allclose(got, ref, rtol=1e-11, atol=1e-62)


# ---- tiles, chord splits, repeatability ------------------------------------------------------------------
def test_tiles_equal_full_run(ctx, transits):
    tr = transits("C3s")
    R = tr.sumOverChords()
    n_orb, n_wav = R.shape
    for (p0, p1, w0, w1) in [(0, 1, 0, n_wav), (2, 7, 5, 133), (n_orb - 1, n_orb, n_wav - 1, n_wav), (0, n_orb, 64, 65)]:
        tile, _ = ctx.run(p0, p1, w0, w1)
        np.testing.assert_allclose(tile, R[p0:p1, w0:w1], rtol=0, atol=3e-15)
    again = tr.sumOverChords()
    assert np.array_equal(R, again), "the chord reduction must be deterministic"


def test_opaque_disk_limit(nist_rows, golden):
    """tau == 0 (no absorber) => R is the opaque-disk value R0 of the reference."""
    spec = workloads.get("C3s")
    for s in spec["scenarios"]:
        s["N"] = 0.0
    tr = builder.build_transit(spec, nist_rows=nist_rows)
    R = tr.sumOverChords()
    np.testing.assert_allclose(R, golden("C3s")["R0"], rtol=0, atol=2e-15)


# ---- error conventions --------------------------------------------------------------------------------
def test_errors_are_loud(nist_rows):
    c = capi.Context(0)
    with pytest.raises(capi.Pb200Error) as ei:
        c.run(0, 1, 0, 1)
    assert ei.value.code == -5
    tr = builder.build_transit(workloads.get("C1"), nist_rows=nist_rows)
    tr._upload(c)
    for bad in [(0, 6, 0, 10), (-1, 2, 0, 10), (0, 1, 10, 10), (0, 1, 0, 10 ** 6)]:
        with pytest.raises(capi.Pb200Error) as ei:
            c.run(*bad)
        assert ei.value.code == -1
    c.close()


def test_serpens_out_of_grid_raises(nist_rows):
    """The reference's RegularGridInterpolator raises ValueError outside the histogram (:600)."""
    spec = workloads.get("SERs")
    spec["grid"][1] *= 1.5            # x axis now sticks out of the tabulated box
    tr = builder.build_transit(spec, nist_rows=nist_rows, build_tables=False)
    good = builder.build_transit(workloads.get("SERs"), nist_rows=nist_rows)
    dist, gdist = tr.atmosphere.densityDistributionList[0], good.atmosphere.densityDistributionList[0]
    dist.InterpolatedDensity = gdist.InterpolatedDensity
    dist.constituents[0].lookupFunction = gdist.constituents[0].lookupFunction
    with pytest.raises(capi.Pb200Error) as ei:
        tr.sumOverChords()
    assert ei.value.code == -6


def test_trimmed_sigma_tables_are_bit_identical(ctx, transits):
    """A wavelength-sharded rank uploads only the part of each sigma table its wavelengths can reach
    (scene.flatten(wav_range=...)); the tile it computes must not change by a single bit."""
    from prometheus_b200 import scene
    for name in ("C3s", "C5s", "C2s"):
        tr = transits(name)
        n_orb, n_wav = len(tr.spatialGrid.constructOrbphaseAxis()), tr.wavelength.size
        for lo, hi in [(0, n_wav // 3), (n_wav // 3, n_wav - 5), (n_wav - 5, n_wav), (n_wav // 2, n_wav // 2 + 1)]:
            sc, keep, info = scene.flatten(tr)
            ctx.set_scene(sc, keep)
            full, _ = ctx.run(0, n_orb, lo, hi)
            sc2, keep2, info2 = scene.flatten(tr, wav_range=(lo, hi))
            assert info2["h2d_bytes"] <= info["h2d_bytes"]
            ctx.set_scene(sc2, keep2)
            part, _ = ctx.run(0, n_orb, lo, hi)
            assert np.array_equal(full, part), (name, lo, hi)


# ---- transparent-chord culling and mirror pairing (compact_kernel) --------------------------------
ATOM_CASES = ["C1", "C3s", "C4s", "C5s", "PLNs", "THMs", "SERs", "HYDs"]


def run_with_env(tr, monkeypatch=None, **opts):
    """One Transit.sumOverChords with the given pb200_set_option values (capi.OPT names), defaults restored after."""
    with host.get_context(0).options(**opts):
        R = tr.sumOverChords()
        t = dict(tr.last_timing)
    return R, t


@pytest.mark.parametrize("name", ATOM_CASES)
def test_culling_and_mirror_leave_R_unchanged(name, transits, monkeypatch):
    """Transparent chords add exactly w_in (exp(-tau) == 1.0 for tau < 2^-54) and mirrored chords have equal
    column densities, so switching either off changes R only by summation order / 1e-12-level Ncol noise."""
    tr = transits(name)
    if any(con.isMolecule for d in tr.atmosphere.densityDistributionList for con in d.constituents):
        pytest.skip("kernel (m) has no culling")
    R, t = run_with_env(tr, monkeypatch)
    R_full, t_full = run_with_env(tr, monkeypatch, cull=0, mirror=0)
    R_nocull, t_nocull = run_with_env(tr, monkeypatch, cull=0)
    R_nomirror, t_nomirror = run_with_env(tr, monkeypatch, mirror=0)
    R_nothin, t_nothin = run_with_env(tr, monkeypatch, thin=0)
    R_brute, t_brute = run_with_env(tr, monkeypatch, cull=0, mirror=0, thin=0)
    for other in (R_full, R_nocull, R_nomirror, R_nothin, R_brute):
        assert np.max(np.abs(R - other)) < 5e-14
    assert t["exp_samples"] <= t_nothin["exp_samples"]
    # thin path off: every record of every window takes the table exp (windows are padded to whole warps)
    win = 32 * (8 if t["samples"] and _n_atoms(tr) <= 2 else 4 if _n_atoms(tr) <= 4 else 2)
    assert t_nothin["exp_samples"] == t_nothin["executed_samples"] // R.shape[1] * (-(-R.shape[1] // win) * win)
    n_orb, n_wav = R.shape
    assert t["samples"] == t_full["samples"]
    # with both switched off every unblocked chord has a record (tiles are padded to 256 chords per phase)
    M = t["samples"] // (n_wav * n_orb)
    assert t_full["executed_samples"] <= (M + 255) // 256 * 256 * n_orb * n_wav
    assert t_full["executed_samples"] >= 0.8 * t["samples"]
    assert t["executed_samples"] <= t_nocull["executed_samples"] <= t_full["executed_samples"]
    assert t["executed_samples"] <= t_nomirror["executed_samples"] <= t_full["executed_samples"]


def test_mirror_pairing_halves_the_records_when_the_scene_is_symmetric(transits, monkeypatch):
    tr = transits("C5s")           # power law around the planet, even n_phi
    _, t_pair = run_with_env(tr, monkeypatch, cull=0)
    _, t_full = run_with_env(tr, monkeypatch, cull=0, mirror=0)
    assert t_pair["executed_samples"] <= 0.6 * t_full["executed_samples"]


def test_mirror_pairing_is_off_for_tabulated_densities_and_odd_phi(transits, nist_rows, monkeypatch):
    tr = transits("SERs")          # SERPENS histogram: no z symmetry
    _, t_a = run_with_env(tr, monkeypatch, cull=0)
    _, t_b = run_with_env(tr, monkeypatch, cull=0, mirror=0)
    assert t_a["executed_samples"] == t_b["executed_samples"]
    spec = workloads.get("C4s")
    spec["grid"] = _with_phi(spec["grid"], 31)
    odd = builder.build_transit(spec, nist_rows=nist_rows)
    R_a, t_a = run_with_env(odd, monkeypatch, cull=0)
    R_b, t_b = run_with_env(odd, monkeypatch, cull=0, mirror=0)
    assert t_a["executed_samples"] == t_b["executed_samples"]
    assert np.array_equal(R_a, R_b)


def _n_atoms(tr):
    return sum(1 for d in tr.atmosphere.densityDistributionList for con in d.constituents if not con.isMolecule)


def _with_phi(grid, n_phi):
    g = list(grid)
    g[5] = n_phi          # Grid(x_midpoint, x_border, x_steps, rho_border, rho_steps, phi_steps, ...)
    return g


def test_exp_sample_count_for_an_optically_thick_scene(transits, monkeypatch):
    """C5s: the power-law envelope is optically thick in every wavelength window (every window holds a line),
    so every record of every window goes through the table exp: exp_samples counts (warp, chord) pairs x the
    window width, i.e. the records x the padded wavelength count."""
    tr = transits("C5s")
    R, t = run_with_env(tr, monkeypatch)
    n_orb, n_wav = R.shape
    records = t["executed_samples"] // n_wav
    win = 256
    padded = -(-n_wav // win) * win
    assert t["exp_samples"] <= records * padded
    assert t["exp_samples"] >= 0.5 * t["executed_samples"], (t["exp_samples"], t["executed_samples"], records, padded)


@pytest.mark.parametrize("name,n_ranks", [("C4s", 2), ("C3s", 3), ("C1", 8)])
def test_block_cyclic_tiles_reassemble_R(name, n_ranks, transits, ctx):
    """pb200_run_device_cyclic: the ranks' tiles (computed here one after the other on one GPU) put back in place
    give the full run, to summation-order noise (the window a wavelength falls in differs)."""
    import torch
    from prometheus_b200 import distributed as pbd
    tr = transits(name)
    R = tr.sumOverChords()
    n_orb, n_wav = R.shape
    out = np.full_like(R, np.nan)
    for r in range(n_ranks):
        idx = pbd.cyclic_indices(n_wav, n_ranks, r)
        assert capi.load().pb200_cyclic_count(n_wav, pbd.CYCLIC_BLOCK, n_ranks, r) == idx.size
        if idx.size == 0:
            continue
        buf = torch.empty(n_orb * idx.size, dtype=torch.float64, device="cuda:0")
        ctx.set_stream(torch.cuda.current_stream().cuda_stream)
        ctx.run_device_cyclic(0, n_orb, pbd.CYCLIC_BLOCK, n_ranks, r, buf.data_ptr())
        torch.cuda.synchronize()
        out[:, idx] = buf.cpu().numpy().reshape(n_orb, idx.size)
    assert np.max(np.abs(out - R)) < 5e-14


def test_absorber_grouping_leaves_R_unchanged(nist_rows, golden, monkeypatch):
    """One absorber per scenario (effective cross section sum_k chi_k sigma_k) against one per constituent
    (PB200_OPT_GROUP = 0): the same tau up to the order of two roundings.  C5s: Fe I + Fe II in one power-law
    atmosphere; HYDs: two scenarios with one constituent each (nothing to merge, identical bits)."""
    from prometheus_b200 import scene
    ctx = host.get_context(0)
    for name, merged in (("C5s", True), ("HYDs", False)):
        tr = builder.build_transit(workloads.get(name), nist_rows=nist_rows)
        out = {}
        for flag in ("0", "1"):
            with ctx.options(group=0 if flag == "1" else 1):
                sc, keep, info = scene.flatten(tr)
                ctx.set_scene(sc, keep)
                out[flag], _ = ctx.run(0, info["n_orb"], 0, info["n_wav"])
        if merged:
            assert np.max(np.abs(out["0"] - out["1"])) < 5e-14
        else:
            assert np.array_equal(out["0"], out["1"])
        g = golden(name)
        ok, worst = excess_depth_ok(out["0"], g["R"], g["R0"])
        assert ok, f"{name}: err/bound {worst}"


def test_moment_form_of_the_molecular_kernel(nist_rows, golden, monkeypatch):
    """Kernel (m) in moment form (items per pressure interval, computed once per chord) against its per-record form
    (PB200_OPT_MOL_RECORDS = 1): same tau up to the 9e-13 truncation of the expansion, and both within the parity
    criterion of the reference output."""
    from prometheus_b200 import scene
    ctx = host.get_context(0)
    tr = builder.build_transit(workloads.get("C2s"), nist_rows=nist_rows)
    sc, keep, info = scene.flatten(tr)
    out = {}
    for flag in ("1", "0"):
        with ctx.options(mol_records=int(flag)):
            ctx.set_scene(sc, keep)
            out[flag], _ = ctx.run(0, info["n_orb"], 0, info["n_wav"])
            # tiles: a sub-range of phases and wavelengths sees the same items
            tile, _ = ctx.run(1, 3, 100, 700)
            assert np.max(np.abs(tile - out[flag][1:3, 100:700])) < 1e-14
    assert np.max(np.abs(out["0"] - out["1"])) < 2e-13
    g = golden("C2s")
    ok, worst = excess_depth_ok(out["0"], g["R"], g["R0"])
    assert ok, f"err/bound {worst}"


@pytest.mark.parametrize("name", ["C1", "C4s", "PLNs", "THMs", "FSTs"])
def test_sigma_sorted_windows_leave_R_unchanged(name, nist_rows, golden, monkeypatch):
    """One absorber: wavelengths of a phase processed in the order of their cross section (PB200_OPT_SORT = 1 forces it on
    these small grids), with the order-8 thin expansion (tau < 0.1 on a whole window) and the opaque skip (tau > 50).
    FSTs runs kernel (m) (per-chord stellar spectrum): the switch must be ignored there."""
    from prometheus_b200 import scene
    ctx = host.get_context(0)
    tr = builder.build_transit(workloads.get(name), nist_rows=nist_rows)
    sc, keep, info = scene.flatten(tr)
    ctx.set_scene(sc, keep)
    out = {}
    for flag in ("0", "1"):
        with ctx.options(sort=int(flag)):
            out[flag], t = ctx.run(0, info["n_orb"], 0, info["n_wav"])
            tile, _ = ctx.run(1, 3, 5, info["n_wav"] - 7)                      # ragged tile: same values
            assert np.max(np.abs(tile - out[flag][1:3, 5:info["n_wav"] - 7])) < 5e-14
    with ctx.options(sort=1, thin=0):                                          # every active sample through the exp, sorted
        R_exp, _ = ctx.run(0, info["n_orb"], 0, info["n_wav"])
    assert np.max(np.abs(out["0"] - out["1"])) < 5e-14 and np.max(np.abs(R_exp - out["0"])) < 5e-14
    for centre in (0, 1, 2):  # the sorted windows (the default, 3, tabulates R on sigma cells instead and sorts nothing)
        with ctx.options(sort=1, centre=centre):
            R_c, _ = ctx.run(0, info["n_orb"], 0, info["n_wav"])
        assert np.max(np.abs(R_c - out["0"])) < 5e-14, centre
    g = golden(name)
    ok, worst = excess_depth_ok(out["1"], g["R"], g["R0"])
    assert ok, f"{name}: err/bound {worst}"


def test_sigma_sorted_windows_with_block_cyclic_tiles(nist_rows, monkeypatch):
    """The multi-GPU tiles (blocks of 256 wavelengths dealt round-robin, written in place) through the sorted kernel:
    the permutation maps a sorted position to the rank's LOCAL wavelength index, the scatter to the global one."""
    from prometheus_b200 import scene
    ctx = host.get_context(0)
    spec = workloads.get("C4s")
    spec["wavelength"][4] = 2e-11                      # ~2000 wavelengths: several blocks per rank
    tr = builder.build_transit(spec, nist_rows=nist_rows)
    sc, keep, info = scene.flatten(tr)
    ctx.set_scene(sc, keep)
    n_orb, n_wav = info["n_orb"], info["n_wav"]
    with ctx.options(sort=0):
        R_ref, _ = ctx.run(0, n_orb, 0, n_wav)
    full = capi.pinned_pool.empty((n_orb, n_wav))
    full[...] = -1.0
    with ctx.options(sort=1):
        for r in range(3):
            ctx.run_scatter(0, n_orb, 256, 3, r, full.ctypes.data, n_wav)
        ctx.synchronize()
    assert n_wav > 3 * 256 and np.max(np.abs(full - R_ref)) < 5e-14


def test_narrow_sorted_windows_as_taylor_sums(nist_rows):
    """On a fine wavelength grid most sorted windows (256 wavelengths) or their 32-wavelength groups span less than 1/40 in
    sigma: kernel (c') then needs one exp per (chord, window / group) and nine Taylor sums about the centre
    (transit_tile_centre).  Same R as the per-sample exp and as the unsorted run; far fewer exp evaluations."""
    from prometheus_b200 import scene
    ctx = host.get_context(0)
    spec = workloads.get("C4s")
    spec["wavelength"] = list(workloads.get("C4")["wavelength"])      # the high-resolution grid of C4: 31 969 wavelengths
    tr = builder.build_transit(spec, nist_rows=nist_rows)
    sc, keep, info = scene.flatten(tr)
    ctx.set_scene(sc, keep)
    n_orb, n_wav = info["n_orb"], info["n_wav"]
    assert n_wav > 8192
    with ctx.options(sort=0):
        R_ref, _ = ctx.run(0, n_orb, 0, n_wav)
    R, exps = {}, {}
    for centre in (0, 1, 2, 3):                        # 3 (default): sigma cells, no sort
        with ctx.options(sort=1, centre=centre):
            R[centre], t = ctx.run(0, n_orb, 0, n_wav)
            exps[centre] = t["exp_samples"]
            assert (ctx.last_sort_order(n_orb) is None) == (centre == 3)
        assert np.max(np.abs(R[centre] - R_ref)) < 5e-14, centre
    assert exps[3] < 0.5 * exps[0], exps             # cells: (chord, cell) pairs instead of samples through the exp
    for centre in (2, 3):                              # a ragged tile and a phase subset through the same paths
        with ctx.options(sort=1, centre=centre):
            tile, _ = ctx.run(1, 2, 3, n_wav - 11)
        assert np.max(np.abs(tile - R[centre][1:2, 3:n_wav - 11])) < 5e-14, centre
    # the same scene made opaque, transparent, and so opaque that tau spans more than the 512 cells of the arrays (2.6e33):
    # the cell grid follows the column densities; the wavelengths beyond the arrays take the exact sum of cell_eval_kernel
    for scale in (1e6, 1e-9, 1e31):
        spec2 = workloads.get("C4s")
        spec2["wavelength"] = list(workloads.get("C4")["wavelength"])
        for scn in spec2["scenarios"]:
            for key in ("Nparticles", "N", "P_0"):
                if key in scn:
                    scn[key] *= scale
        tr2 = builder.build_transit(spec2, nist_rows=nist_rows)
        sc2, keep2, info2 = scene.flatten(tr2)
        ctx.set_scene(sc2, keep2)
        with ctx.options(sort=0):
            R0_, _ = ctx.run(0, info2["n_orb"], 0, info2["n_wav"])
        with ctx.options(sort=1):
            R1_, _ = ctx.run(0, info2["n_orb"], 0, info2["n_wav"])
        assert np.max(np.abs(R0_ - R1_)) < 5e-14, scale


def test_power_law_with_non_positive_exponent_is_not_culled(nist_rows):
    """PowerLawAtmosphere accepts any q (gasProperties.py:228-244).  For q <= 0 the density grows along the chord, so
    the closest approach is no upper bound of the x-sum: the transparent-chord shortcut must not fire."""
    from helpers import oracle_inputs
    spec = workloads.get("HYDs")
    spec["scenarios"] = [{"type": "powerLaw", "T": 3000.0, "P_0": 3e-7, "q_esc": -0.5,
                          "species": [{"name": "NaI", "chi": 1e-5}]}]
    tr = builder.build_transit(spec, nist_rows=nist_rows)
    R = tr.sumOverChords()
    inputs = oracle_inputs(spec, nist_rows)
    con = tr.atmosphere.densityDistributionList[0].constituents[0]
    inputs["tables"] = {(0, "NaI"): (con.lookupFunction.x, con.lookupFunction.y)}
    m = oracle.build_model(spec, inputs)
    R_ref = oracle.sum_over_chords(m)
    R0 = oracle.sum_over_chords(m, opaque_only=True)
    ok, worst = excess_depth_ok(R, R_ref, R0)
    assert ok, worst
    assert np.max(np.abs(R - R_ref)) < 5e-13 and float((R0 - R).max()) > 1e-6
    R_nocull, _ = run_with_env(tr, cull=0, mirror=0, thin=0)
    assert np.max(np.abs(R - R_nocull)) < 5e-14


def test_options_are_per_context_and_validated(ctx):
    assert {k: ctx.get_option(k) for k in capi.OPT} == capi.OPT_DEFAULTS
    with ctx.options(cull=0, sort=1):
        assert ctx.get_option("cull") == 0 and ctx.get_option("sort") == 1
        other = capi.Context(0)
        assert other.get_option("cull") == 1
        other.close()
    assert ctx.get_option("cull") == 1 and ctx.get_option("sort") == -1
    for name, bad in (("cull", 2), ("sort", 5), ("min_tiles", 0)):
        with pytest.raises(capi.Pb200Error) as ei:
            ctx.set_option(name, bad)
        assert ei.value.code == -1
    with pytest.raises(capi.Pb200Error):
        ctx.check(ctx.lib.pb200_set_option(ctx.handle, 99, 0))


def test_prefetched_tables_follow_the_host_arrays(nist_rows, golden):
    """sum_over_chords starts the H2D copies of the previous call's page-locked sigma tables before it flattens the scene
    (pb200_prefetch_tables); pb200_set_scene then skips exactly those copies.  The result must follow the host arrays:
    same tables -> same bits; a table changed in place between two calls -> the new values; another scene in between ->
    its own result, and the first scene's again afterwards."""
    spec = workloads.get("C4s")
    spec["wavelength"] = list(workloads.get("C4")["wavelength"])      # C4's grid: a sigma table of 2.7 MB
    tr = builder.build_transit(spec, nist_rows=nist_rows)
    R1 = tr.sumOverChords().copy()
    R2 = tr.sumOverChords().copy()
    assert np.array_equal(R1, R2)
    con = tr.atmosphere.densityDistributionList[0].constituents[0]
    y = con.lookupFunction.y
    assert y.nbytes >= (64 << 10), "the table is large enough to be sent from where it lies"
    y += np.log10(2.0)                                   # sigma doubled, in place, same array
    R3 = tr.sumOverChords().copy()
    assert float((R1 - R3).max()) > 1e-6 and np.all(R3 <= R1 + 1e-15)
    tr_b = builder.build_transit(workloads.get("C1"), nist_rows=nist_rows)
    Rb = tr_b.sumOverChords().copy()
    g = golden("C1")
    ok, worst = excess_depth_ok(Rb, g["R"], g["R0"])
    assert ok, worst
    R4 = tr.sumOverChords().copy()
    assert np.array_equal(R3, R4)
    y -= np.log10(2.0)
    R5 = tr.sumOverChords().copy()
    assert np.max(np.abs(R5 - R1)) < 1e-13


def test_more_constituents_than_compiled_absorbers(nist_rows):
    """14 atomic constituents in one scenario: with the per-scenario grouping (default) they are ONE absorber and run; with
    PB200_OPT_GROUP = 0 they would be 14 absorbers -- more than kernel (c') is compiled for (PB200_MAX_ABSORBERS = 12) -- and
    the run must be refused with PB200_ERR_UNSUPPORTED before any kernel sizes its shared arrays by the absorber count.
    The context stays usable."""
    from prometheus_b200 import scene
    ctx = host.get_context(0)
    spec = workloads.get("C2s")
    names = ["NaI", "KI", "MgI", "MgII", "CaI", "CaII", "FeI", "TiI", "TiII", "SiI", "AlI", "CrI", "MnI", "NiI"]
    spec["scenarios"][0]["molecules"] = []
    spec["scenarios"][0]["species"] = [{"name": n, "chi": 1e-7} for n in names]
    tr = builder.build_transit(spec, nist_rows=nist_rows)
    sc, keep, info = scene.flatten(tr)
    ctx.set_scene(sc, keep)
    R, _ = ctx.run(0, info["n_orb"], 0, info["n_wav"])
    assert np.all(np.isfinite(R)) and R.min() < 1.0
    with ctx.options(group=0):
        with pytest.raises(capi.Pb200Error) as err:
            sc2, keep2, _ = scene.flatten(tr)
            ctx.set_scene(sc2, keep2)
            ctx.run(0, info["n_orb"], 0, info["n_wav"])
        assert err.value.code == -4 and "absorbers" in str(err.value)
    sc3, keep3, _ = scene.flatten(tr)
    ctx.set_scene(sc3, keep3)
    R3, _ = ctx.run(0, info["n_orb"], 0, info["n_wav"])
    assert np.array_equal(R, R3)
