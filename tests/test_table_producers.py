"""The two table producers of SURVEY 8(f) rank 4 with file-format-free inputs, against the reference's OWN methods run on
the same arrays: Star.addFstarFunction (celestialBodies.py:242-265, with getSpectrum replaced by the arrays) and
MolecularConstituent.constructLookupFunction (gasProperties.py:765-781, with a stand-in for h5py.File that serves the
arrays).  Index work (range selection, axis reversal, grids) bit-exact; log10 within 4 ulp of numpy's."""
import os
import sys

import numpy as np
import pytest

from helpers import ROOT
from prometheus_b200 import gas, host

sys.path.insert(0, os.path.join(ROOT, "oracle"))
import make_ref  # noqa: E402

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not make_ref.available(), reason="oracle/_ref is absent")]


def _ulps(a, b):
    return np.max(np.abs(a - b) / np.spacing(np.abs(b)))


def test_fstar_table_equals_the_reference_method():
    gasprop, bodies, _, _ = make_ref.load()
    rng = np.random.default_rng(4)
    w = np.sort(rng.uniform(3000e-8, 9000e-8, 20000))
    F = 10 ** rng.uniform(4.0, 7.0, w.size)
    wavelength = np.linspace(5880e-8, 5905e-8, 300)
    for vsini in (0.0, 3e5):
        ref = bodies.Star(7e10, 0, 2e33, 0, 5000, 0, 4.5, 0, 0, 0, 0)
        ref.addRMparameters(vsini, 0.3)
        ref.getSpectrum = lambda: (w, F)                       # the one stubbed piece: network + FITS
        ref.addFstarFunction(wavelength)
        star = host.Star(7e10, 0, 2e33, 0, 5000, 0, 4.5, 0, 0, 0, 0)
        star.addRMparameters(vsini, 0.3)
        star.setSpectrum(w, F)
        star.addFstarFunction(wavelength)
        assert np.array_equal(star.Fstar_function.x, ref.Fstar_function.x)
        assert _ulps(np.asarray(star.Fstar_function.y), ref.Fstar_function.y) <= 4
    with pytest.raises(NotImplementedError):
        host.Star(7e10).addFstarFunction(wavelength)


def test_molecular_lookup_table_equals_the_reference_method(monkeypatch):
    gasprop = make_ref.load()[0]
    rng = np.random.default_rng(8)
    p_bar = np.logspace(-6, 2, 9)
    t = np.linspace(300.0, 3000.0, 7)
    bin_edges = np.sort(rng.uniform(1e4, 3e4, 501))            # wavenumbers, ascending
    xsec = 10 ** rng.uniform(-30, -18, (p_bar.size, t.size, bin_edges.size))
    xsec[0, 0, :5] = 0.0                                        # sigma = 0 -> log10(offset)
    data = {"p": p_bar, "t": t, "bin_edges": bin_edges, "xsecarr": xsec}

    class _File:                                                # what the reference needs of h5py.File
        def __init__(self, *a, **k): pass
        def __enter__(self): return data
        def __exit__(self, *exc): return False

    monkeypatch.setattr(gasprop, "h5py", type("h5", (), {"File": _File}), raising=False)
    ref = gasprop.MolecularConstituent("SYN", 1e-4).constructLookupFunction()
    mol = gas.MolecularConstituent("SYN", 1e-4)
    mol.setOpacityArrays(p_bar, t, bin_edges, xsec)
    got = mol.constructLookupFunction()
    for a, b in zip(got.grid, ref.grid):
        assert np.array_equal(a, b)
    assert got.fill_value == ref.fill_value
    assert _ulps(np.asarray(got.values), ref.values) <= 4
    with pytest.raises(NotImplementedError):
        gas.MolecularConstituent("NONE", 1.0).constructLookupFunction()
