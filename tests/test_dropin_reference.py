"""Drop-in check in the build container (needs /root/reference; skipped on the GPU box):
scene.flatten() applied to the UNMODIFIED reference's own Transit object yields the same POD the
host mirror yields, field by field and bit by bit -- i.e. the C ABI sees identical inputs whether it
sits behind the reference or behind prometheus_b200.host/.gas.  install()/uninstall() rebinding is
checked too (no compute call: there is no GPU here)."""
import ctypes as C
import importlib.util
import os
import sys

import numpy as np
import pytest

from prometheus_b200 import builder, capi, host, scene, workloads

HERE = os.path.dirname(os.path.abspath(__file__))
pytestmark = pytest.mark.skipif(not os.path.isdir("/root/reference/pythonScripts"), reason="reference tree not mounted")


@pytest.fixture(scope="module")
def ref_ns():
    spec = importlib.util.spec_from_file_location("make_golden", os.path.join(HERE, "golden", "make_golden.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod.ReferenceNamespace()


def arr(ptr, n):
    return np.ctypeslib.as_array(ptr, shape=(n,)).copy() if n else np.empty(0)


@pytest.mark.parametrize("name", ["C1", "C3s", "THMs", "HYDs", "FSTs", "PLNs"])
def test_flatten_reference_equals_flatten_mirror(name, ref_ns, nist_rows):
    spec = workloads.get(name)
    ref_tr = builder.build_transit(spec, ns=ref_ns, nist_rows=ref_ns.nist_rows)
    mir_tr = builder.build_transit(spec, nist_rows=nist_rows, build_tables=False)
    for rd, md in zip(ref_tr.atmosphere.densityDistributionList, mir_tr.atmosphere.densityDistributionList):
        for rc, mc in zip(rd.constituents, md.constituents):
            if not mc.isMolecule:       # sigma tables are inputs here: hand the mirror the reference's
                mc.lookupFunction = host.TableFunction1D(rc.lookupFunction.x, rc.lookupFunction.y)
    a, keep_a, info_a = scene.flatten(ref_tr)
    b, keep_b, info_b = scene.flatten(mir_tr)
    assert info_a == info_b
    for f in ("n_x", "n_rho", "n_phi", "n_orb", "n_scenarios", "n_constituents", "n_moons", "n_wav", "delta_x",
              "star_R", "clv_u1", "clv_u2", "vsini", "planet_R", "n_fstar"):
        assert getattr(a, f) == getattr(b, f), f
    n_orb = a.n_orb
    for f, n in (("x", a.n_x), ("rho", a.n_rho), ("sin_phi", a.n_phi), ("cos_phi", a.n_phi), ("cos_phi_rot", a.n_phi),
                 ("wavelength", a.n_wav), ("planet_y", n_orb), ("moon_R", a.n_moons), ("moon_y", a.n_moons * n_orb),
                 ("fstar_x", a.n_fstar), ("fstar_y", a.n_fstar)):
        assert np.array_equal(arr(getattr(a, f), n), arr(getattr(b, f), n)), f
    for i in range(a.n_scenarios):
        sa, sb = a.scenarios[i], b.scenarios[i]
        assert (sa.type, sa.n_constituents, sa.first_constituent, sa.T) == (sb.type, sb.n_constituents, sb.first_constituent, sb.T)
        assert list(sa.p) == list(sb.p)
        for f in ("centre_x", "centre_y", "shift", "amplitude"):
            pa, pb_ = getattr(sa, f), getattr(sb, f)
            assert bool(pa) == bool(pb_), f
            if pa:
                assert np.array_equal(arr(pa, n_orb), arr(pb_, n_orb)), f
    for k in range(a.n_constituents):
        ca, cb = a.constituents[k], b.constituents[k]
        assert (ca.kind, ca.chi, ca.offset, ca.n_tab) == (cb.kind, cb.chi, cb.offset, cb.n_tab)


def test_install_rebinds_and_restores(ref_ns):
    import prometheus_b200 as pb
    gas = ref_ns.gas
    orig_sum, orig_voigt = gas.Transit.sumOverChords, gas.AtmosphericConstituent.calculateVoigtProfile
    pb.install(gas)
    try:
        assert gas.Transit.sumOverChords is not orig_sum
        assert gas.AtmosphericConstituent.calculateVoigtProfile is not orig_voigt
        assert "max_memory_gb" in gas.Transit.sumOverChords.__code__.co_varnames
        if capi.load().pb200_device_count() == 0:       # no GPU here: the patched method must fail loudly, not fall back
            tr = builder.build_transit(workloads.get("C1"), ns=ref_ns, nist_rows=ref_ns.nist_rows, build_tables=False)
            with pytest.raises(capi.Pb200Error):
                tr.sumOverChords()
    finally:
        pb.uninstall()
    assert gas.Transit.sumOverChords is orig_sum
    assert gas.AtmosphericConstituent.calculateVoigtProfile is orig_voigt
