"""Retrieval batch, CPU side: the oracle's restatement of mainRetrieval.py:69-93 against the
golden light curves (produced by executing the reference script's own statements, tests/golden/make_golden.py),
the host-side model flattening, and the data-parallel split of a batch over ranks (gloo, world size 2 and 3)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from helpers import oracle, oracle_inputs
from prometheus_b200 import builder, retrieval, scene, workloads


def test_oracle_lightcurve_matches_reference(golden, nist_rows):
    g = golden("RETs")
    spec = workloads.get("RETs")
    assert np.array_equal(g["line_centres"], spec["band"]["line_centres"]) and g["bandwidth"] == spec["band"]["bandwidth"]
    for k, pv in enumerate(g["params"]):
        v = workloads.retrieval_variant(spec, *pv)
        m = oracle.build_model(v, oracle_inputs(v, nist_rows))
        assert np.array_equal(m.wavelength, g["wavelength"])
        shift = oracle.band_shift(m)
        assert np.array_equal(shift, g["band_shift"])
        R = oracle.sum_over_chords(m)
        assert np.max(np.abs(R - g["R"][k])) < 1e-13
        # the light-curve statement itself, on the reference's own R: selection is index work -> exact
        lc = oracle.lightcurve(g["R"][k], g["wavelength"], shift, g["line_centres"], float(g["bandwidth"]))
        assert np.array_equal(lc, g["lightcurve"][k])
        lc2 = oracle.lightcurve(R, m.wavelength, shift, spec["band"]["line_centres"], spec["band"]["bandwidth"])
        assert np.max(np.abs(lc2 - g["lightcurve"][k])) < 1e-13


def test_oracle_lightcurve_empty_selection_is_nan():
    R = np.ones((2, 5))
    wl = np.linspace(1.0, 2.0, 5)
    lc = oracle.lightcurve(R, wl, np.ones(2), [3.0], 0.1)
    assert np.all(np.isnan(lc))


def test_flatten_model_agrees_with_flatten(nist_rows):
    spec = workloads.get("RETs")
    tr = builder.build_transit(spec, nist_rows=nist_rows, build_tables=False)
    pv = workloads.retrieval_samples(3)[2]
    (var,) = retrieval.make_variants(tr, [pv])
    ref = builder.build_transit(workloads.retrieval_variant(spec, *pv), nist_rows=nist_rows, build_tables=False)
    mdl, keep = scene.flatten_model(var)
    mdl_ref, keep_ref = scene.flatten_model(ref)
    n_orb = spec["grid"][7]
    for name in ("planet_y", "moon_y"):
        assert np.array_equal(np.ctypeslib.as_array(getattr(mdl, name), (n_orb,)), np.ctypeslib.as_array(getattr(mdl_ref, name), (n_orb,)))
    assert mdl.moon_R[0] == mdl_ref.moon_R[0] != spec["bodies"]["moon"]["R"]
    a, b = mdl.scenarios[0], mdl_ref.scenarios[0]
    assert list(a.p) == list(b.p) and a.type == b.type
    for name in ("centre_x", "centre_y", "shift", "amplitude"):
        assert np.array_equal(np.ctypeslib.as_array(getattr(a, name), (n_orb,)), np.ctypeslib.as_array(getattr(b, name), (n_orb,)))
    # the template itself is untouched
    assert tr.atmosphere.densityDistributionList[0].moon.R == spec["bodies"]["moon"]["R"]


def test_broadcast_family_matches_per_model(nist_rows):
    import ctypes as C
    from prometheus_b200 import capi
    spec = workloads.get("RETs")
    tr = builder.build_transit(spec, nist_rows=nist_rows, build_tables=False)
    params = workloads.retrieval_samples(7)
    phase = np.asarray(tr.spatialGrid.constructOrbphaseAxis())
    n_orb = phase.size
    fam, K = retrieval.make_family(tr, params)
    models, keep = scene.flatten_models_broadcast(fam, phase, K)
    assert models.dtype.itemsize == C.sizeof(capi.Model) and K == 7
    arr = models.ctypes.data_as(C.POINTER(capi.Model))
    view = lambda p: np.ctypeslib.as_array(p, (n_orb,))   # noqa: E731
    for k, var in enumerate(retrieval.make_variants(tr, params)):
        m, _ = scene.flatten_model(var, phase)
        b = arr[k]
        assert m.moon_R[0] == b.moon_R[0]
        np.testing.assert_allclose(view(b.planet_y), view(m.planet_y), rtol=0, atol=0)
        np.testing.assert_allclose(view(b.moon_y), view(m.moon_y), rtol=3e-16)
        sa, sb = m.scenarios[0], b.scenarios[0]
        assert (sa.type, sa.n_constituents, sa.first_constituent) == (sb.type, sb.n_constituents, sb.first_constituent)
        assert list(sa.p) == list(sb.p)
        for name in ("centre_x", "centre_y", "shift", "amplitude"):
            np.testing.assert_allclose(view(getattr(sb, name)), view(getattr(sa, name)), rtol=3e-16)


def test_model_partition():
    for n in (0, 1, 5, 64, 1000):
        for world in (1, 2, 3, 8):
            parts = retrieval.model_partition(n, world)
            assert len(parts) == world and parts[0][0] == 0 and parts[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(parts, parts[1:]))
            sizes = [h - l for l, h in parts]
            assert max(sizes) - min(sizes) <= 1


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, n_models, n_orb, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        models = list(range(n_models))
        seen = []

        def compute(ms):
            seen.extend(ms)
            return np.array([[m * 100.0 + p for p in range(n_orb)] for m in ms]).reshape(len(ms), n_orb)

        out = retrieval.run_sharded(models, None, n_orb, compute=compute)
        truth = np.array([[m * 100.0 + p for p in range(n_orb)] for m in models]).reshape(n_models, n_orb)
        q.put((rank, bool(np.array_equal(out, truth)), seen))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,n_models,n_orb", [(2, 5, 9), (3, 2, 4), (2, 8, 25)])
def test_sharded_batch_gathers_every_model_once(world, n_models, n_orb):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_models, n_orb, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    seen = sorted(m for _, _, s in results for m in s)
    assert seen == list(range(n_models))
    assert all(ok for _, ok, _ in results)
