"""GPU parity of the retrieval batch: pb200_set_models / pb200_run_batch / pb200_lightcurve
through the C ABI against the reference's outputs (tests/golden/RETs.npz: R from the reference's sumOverChords,
light curves from the reference script's own statements, mainRetrieval.py:69-93) and the CPU oracle.

Tolerances: the band selection is index work -> the light-curve kernel on a given R agrees with the oracle on the
same R to 1e-15 (sum order only); end to end the excess-depth criterion on R and 2e-10 absolute on the light curve
(the kernel-(b2) sigma tables differ from scipy's at 1e-9 relative)."""
import numpy as np
import pytest

from helpers import excess_depth_ok, oracle, oracle_inputs
from prometheus_b200 import builder, capi, host, retrieval, workloads

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def setup(nist_rows):
    spec = workloads.get("RETs")
    template = builder.build_transit(spec, nist_rows=nist_rows)
    orb = template.spatialGrid.constructOrbphaseAxis()
    band = retrieval.LightcurveBand.for_planet(template.planet, orb, spec["band"]["line_centres"], spec["band"]["bandwidth"])
    return spec, template, band


def test_batch_matches_reference(setup, golden):
    spec, template, band = setup
    g = golden("RETs")
    assert np.array_equal(band.shift, g["band_shift"])
    variants = retrieval.make_variants(template, g["params"])
    batch = retrieval.RetrievalBatch(template)
    lc, R, timing = batch.run(variants, band, want_R=True)
    n_models, n_orb, n_wav = g["R"].shape
    assert lc.shape == (n_models, n_orb) and R.shape == (n_models, n_orb, n_wav)
    for k in range(n_models):
        ok, worst = excess_depth_ok(R[k], g["R"][k], np.ones_like(R[k]))
        assert ok, f"model {k}: err/bound {worst}"
        assert np.max(np.abs(lc[k] - g["lightcurve"][k])) < 2e-10
        # the epilogue kernel on its own input
        lc_o = oracle.lightcurve(R[k], g["wavelength"], band.shift, band.line_centres, band.bandwidth)
        assert np.max(np.abs(lc[k] - lc_o)) < 1e-15
    # one set of launches for the whole batch: sky cells, F_out, sigma, column, compaction, transit, light curve
    assert timing["launches"] <= 8
    batch.reset()


def test_batch_rows_equal_single_runs(setup, nist_rows):
    """Stacking models along the phase axis must not change a bit of any of them."""
    spec, template, band = setup
    params = workloads.retrieval_samples(6)
    variants = retrieval.make_variants(template, params)
    batch = retrieval.RetrievalBatch(template)
    lc, R, _ = batch.run(variants, band, want_R=True)
    batch.reset()
    for k in (0, 3, 5):
        single = variants[k].sumOverChords()
        assert np.array_equal(single, R[k])
        lc1 = retrieval.extract_lightcurve(host.get_context(), band)
        assert np.array_equal(lc1, lc[k])


def test_family_path_agrees_with_object_path(setup):
    """run_family (one vectorised flattening) against run (one flattening per Transit object): the per-phase arrays
    agree to the last bit or two (numpy's sin/cos), the light curves to 1e-13."""
    spec, template, band = setup
    params = workloads.retrieval_samples(16, seed=2)
    batch = retrieval.RetrievalBatch(template)
    lc_obj, R_obj, _ = batch.run(retrieval.make_variants(template, params), band, want_R=True)
    fam, K = retrieval.make_family(template, params)
    lc_fam, R_fam, t = batch.run_family(fam, K, band, want_R=True)
    batch.reset()
    assert K == 16 and lc_fam.shape == lc_obj.shape
    assert np.max(np.abs(R_fam - R_obj)) < 1e-13 and np.max(np.abs(lc_fam - lc_obj)) < 1e-13


def test_batch_against_oracle_variants(setup, nist_rows):
    spec, template, band = setup
    params = workloads.retrieval_samples(3, seed=5)
    variants = retrieval.make_variants(template, params)
    batch = retrieval.RetrievalBatch(template)
    lc, R, _ = batch.run(variants, band, want_R=True)
    batch.reset()
    for k, pv in enumerate(params):
        v = workloads.retrieval_variant(spec, *pv)
        m = oracle.build_model(v, oracle_inputs(v, nist_rows))
        R_ref = oracle.sum_over_chords(m)
        R0 = oracle.sum_over_chords(m, opaque_only=True)
        ok, worst = excess_depth_ok(R[k], R_ref, R0)
        assert ok, f"model {k}: {worst}"
        lc_ref = oracle.lightcurve(R_ref, m.wavelength, oracle.band_shift(m), spec["band"]["line_centres"], spec["band"]["bandwidth"])
        assert np.max(np.abs(lc[k] - lc_ref)) < 2e-10


def test_inner_seams_address_stacked_rows(setup):
    """After pb200_set_models the context has n_models*n_orb phases; pb200_column_density shows them."""
    spec, template, band = setup
    params = workloads.retrieval_samples(2, seed=3)
    variants = retrieval.make_variants(template, params)
    batch = retrieval.RetrievalBatch(template)
    ctx = batch.ctx
    n_orb, M = batch.info["n_orb"], batch.info["n_rho"] * batch.info["n_phi"]
    batch.set_models(variants)
    stacked = np.empty((1, M * 2 * n_orb))
    ctx.check(ctx.lib.pb200_column_density(ctx.handle, capi.dptr(stacked)))
    stacked = stacked.reshape(M, 2, n_orb)
    batch.reset()
    for k in range(2):
        single = variants[k].getColumnDensityGrid().reshape(M, n_orb)
        assert np.array_equal(single, stacked[:, k, :])


def test_error_conventions(setup):
    spec, template, band = setup
    batch = retrieval.RetrievalBatch(template)
    ctx = batch.ctx
    variants = retrieval.make_variants(template, workloads.retrieval_samples(1))
    # no resident R after a fresh scene
    with pytest.raises(capi.Pb200Error) as ei:
        ctx.lightcurve(band.struct(), 0, batch.info["n_orb"])
    assert ei.value.code == -1
    # a model whose scenario type differs from the resident scene
    arr, keep = batch._models(variants)
    arr[0].scenarios[0].type = capi.SCN_TORUS
    with pytest.raises(capi.Pb200Error) as ei:
        ctx.set_models(arr, 1, keep)
    assert ei.value.code == -1
    # band without lines
    bad = band.struct()
    bad.n_lines = 0
    arr, keep = batch._models(variants)
    with pytest.raises(capi.Pb200Error) as ei:
        ctx.run_batch(arr, 1, bad, batch.info["n_orb"], batch.info["n_wav"], False, keep)
    assert ei.value.code == -1
    # a band that selects nothing: nan, as np.mean of an empty selection
    empty = retrieval.LightcurveBand([1.0], 1e-9, band.shift)
    lc, _, _ = batch.run(variants, empty)
    assert np.all(np.isnan(lc))
    batch.reset()


# ---- results written in place: zero-copy pb200_run, pb200_run_scatter ------------------------------------------
def test_scatter_tiles_and_zero_copy_equal_the_plain_run(nist_rows, monkeypatch):
    import ctypes as C
    from prometheus_b200 import scene
    tr = builder.build_transit(workloads.get("C3s"), nist_rows=nist_rows)
    ctx = host.get_context()
    sc, keep, info = scene.flatten(tr)
    ctx.set_scene(sc, keep)
    n_orb, n_wav = info["n_orb"], info["n_wav"]
    with ctx.options(zerocopy=0):
        R_copy, _ = ctx.run(0, n_orb, 0, n_wav)                  # device scratch + D2H
    R_zero, _ = ctx.run(0, n_orb, 0, n_wav)                      # kernels store into the page-locked result
    assert np.array_equal(R_copy, R_zero)
    pageable = np.empty((n_orb, n_wav))
    R_page, _ = ctx.run(0, n_orb, 0, n_wav, out=pageable)        # not page-locked: falls back to the copy
    assert np.array_equal(R_copy, R_page)
    # three "ranks" scatter their block-cyclic tiles into one mapped array with a padded row stride
    stride = n_wav + 5
    full = capi.pinned_pool.empty((n_orb, stride))
    full[...] = -1.0
    for r in range(3):
        ctx.run_scatter(0, n_orb, 256, 3, r, full.ctypes.data, stride)
    ctx.synchronize()
    assert np.array_equal(full[:, :n_wav], R_copy) and np.all(full[:, n_wav:] == -1.0)
    # a sub-range of phases, whole wavelength axis, one rank
    full[...] = -1.0
    ctx.run_scatter(1, 3, 256, 1, 0, full[1:].ctypes.data, stride)
    ctx.synchronize()
    assert np.array_equal(full[1:3, :n_wav], R_copy[1:3]) and np.all(full[0] == -1.0) and np.all(full[3:] == -1.0)
    # memory the caller owns, registered
    own = np.zeros(n_orb * n_wav + 512)
    base = (own.ctypes.data + 4095) // 4096 * 4096
    lib = capi.load()
    dev = C.c_void_p()
    assert lib.pb200_host_register(C.c_void_p(base), n_orb * n_wav * 8, C.byref(dev)) == 0
    ctx.run_scatter(0, n_orb, 256, 1, 0, base, n_wav)
    ctx.synchronize()
    off = (base - own.ctypes.data) // 8
    assert np.array_equal(own[off:off + n_orb * n_wav].reshape(n_orb, n_wav), R_copy)
    assert lib.pb200_host_unregister(C.c_void_p(base)) == 0
    with pytest.raises(capi.Pb200Error):
        ctx.run_scatter(0, n_orb, 256, 1, 0, full.ctypes.data, n_wav - 1)


def test_sharded_table_upload_fills_the_whole_table(nist_rows):
    """pb200_set_scene_sharded: rank r uploads slice r of every large sigma table.  On one GPU the three 'ranks'
    write into the same (grow-only) arena one after the other, which is what the all-gather does across GPUs."""
    from prometheus_b200 import scene
    spec = workloads.get("C4s")
    spec["wavelength"][4] = 2e-12           # finer grid: the refined sigma table exceeds the 1 MiB sharding threshold
    tr = builder.build_transit(spec, nist_rows=nist_rows)
    assert tr.atmosphere.densityDistributionList[0].constituents[0].lookupFunction.x.nbytes > (1 << 20)
    ctx = host.get_context()
    sc, keep, info = scene.flatten(tr)
    ctx.set_scene(sc, keep)
    R_ref, _ = ctx.run(0, info["n_orb"], 0, info["n_wav"])
    regions = None
    for r in (2, 0, 1):
        regions = ctx.set_scene_sharded(sc, keep, 3, r)
    assert len(regions) == 2 and all(chunk % 32 == 0 for _, chunk in regions)
    R, _ = ctx.run(0, info["n_orb"], 0, info["n_wav"])
    assert np.array_equal(R, R_ref)
    # a single rank's slice alone is not enough (the test would be vacuous otherwise)
    big = builder.build_transit(spec, nist_rows=nist_rows)
    lf = big.atmosphere.densityDistributionList[0].constituents[0].lookupFunction
    lf.y = lf.y + 1.0                       # a different table: fresh content in every slice
    sc2, keep2, _ = scene.flatten(big)
    ctx.set_scene_sharded(sc2, keep2, 3, 0)
    R_partial, _ = ctx.run(0, info["n_orb"], 0, info["n_wav"])
    ctx.set_scene(sc2, keep2)
    R_full, _ = ctx.run(0, info["n_orb"], 0, info["n_wav"])
    assert not np.array_equal(R_partial, R_full)


def test_lightcurve_and_batch_edge_cases(setup):
    """One model; a band that covers everything; a light curve of a sub-range of rows; more scatter ranks than
    wavelength blocks."""
    spec, template, band = setup
    batch = retrieval.RetrievalBatch(template)
    n_orb, n_wav = batch.info["n_orb"], batch.info["n_wav"]
    variants = retrieval.make_variants(template, workloads.retrieval_samples(1, seed=9))
    wide = retrieval.LightcurveBand([5.9e-5], 1.0, band.shift)            # selects every wavelength
    lc, R, _ = batch.run(variants, wide, want_R=True)
    assert lc.shape == (1, n_orb)
    ref = oracle.lightcurve(R[0], np.asarray(template.wavelength), wide.shift, wide.line_centres, wide.bandwidth)
    assert np.max(np.abs(lc[0] - ref)) < 1e-15
    # rows [2, 5) of the resident R
    part = batch.ctx.lightcurve(band.struct(), 2, 5)
    full = oracle.lightcurve(R[0], np.asarray(template.wavelength), band.shift, band.line_centres, band.bandwidth)
    assert np.max(np.abs(part - full[2:5])) < 1e-15
    batch.reset()
    # scatter with more ranks than 256-wavelength blocks: the extra ranks have nothing to do
    from prometheus_b200 import scene
    sc, keep, info = scene.flatten(template)
    batch.ctx.set_scene(sc, keep)
    R_ref, _ = batch.ctx.run(0, n_orb, 0, n_wav)
    full = capi.pinned_pool.empty((n_orb, n_wav))
    full[...] = -1.0
    n_blocks = -(-n_wav // 256)
    for r in range(n_blocks + 2):
        batch.ctx.run_scatter(0, n_orb, 256, n_blocks + 2, r, full.ctypes.data, n_wav)
    batch.ctx.synchronize()
    assert np.array_equal(full, R_ref)


def test_models_may_differ_in_temperature_and_planet_radius(nist_rows):
    """pb200_model carries the planet radius of the model, and the T of its scenarios is honoured (ABI v4): a batch whose
    models differ in them equals the single runs of the same objects, and the models really differ from one another.
    (The sigma tables are the scene's: sigma_v is fixed by the template, as the header says.)"""
    import copy
    template = builder.build_transit(workloads.get("HYDs"), nist_rows=nist_rows)
    variants = []
    for f_T, f_R in ((1.0, 1.0), (1.3, 1.0), (1.0, 1.08), (0.8, 0.95)):
        v = copy.deepcopy(template)
        v.planet.R *= f_R
        for dist in v.atmosphere.densityDistributionList:
            dist.T *= f_T
        variants.append(v)
    batch = retrieval.RetrievalBatch(template)
    _, R, _ = batch.run(variants, None, want_R=True)
    batch.reset()
    singles = [v.sumOverChords().copy() for v in variants]
    for k in range(len(variants)):
        assert np.max(np.abs(R[k] - singles[k])) < 5e-14, k
    for k in range(1, len(variants)):
        assert np.max(np.abs(singles[k] - singles[0])) > 1e-6, k        # T and R moved the transit depth
