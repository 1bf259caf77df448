"""prometheus_b200 -- B200 (sm_100a) implementation of the PROMETHEUS transit forward-model hot path.

    import prometheus_b200 as pb
    transit = pb.build_transit(pb.workloads.get("C1"), nist_rows=rows)   # host mirror objects
    R = transit.sumOverChords()                                          # CUDA, (n_orb, n_wav) float64

Drop-in behind the unchanged reference host code:

    import pythonScripts.gasProperties as gasprop
    prometheus_b200.install(gasprop)       # Transit.sumOverChords now runs on the GPU

The compute path is libprometheus_b200.so (csrc/, C ABI in include/prometheus_b200.h).  Importing this
package does not need a GPU; calling into it does, and fails loudly without one.
"""
from __future__ import annotations

import numpy as np

from . import capi, constants, workloads                      # noqa: F401
from . import host, gas, scene, builder                        # noqa: F401
from .builder import build_transit                             # noqa: F401
from .capi import Context, Pb200Error                          # noqa: F401

__all__ = ["install", "uninstall", "sum_over_chords", "build_transit", "Context", "Pb200Error", "workloads"]

_saved = {}


def sum_over_chords(transit, max_memory_gb: float = 2.0, device: int | None = None) -> np.ndarray:
    """GPU replacement of `Transit.sumOverChords(self, max_memory_gb)` for any Transit-like object
    (reference: pythonScripts/gasProperties.py:1160-1258)."""
    ctx = host.get_context(device)
    ctx.prefetch_tables()      # the sigma tables of the previous call start moving while the host flattens this one
    sc, keep, info = scene.flatten(transit)
    ctx.set_scene(sc, keep)
    R, timing = ctx.run(0, info["n_orb"], 0, info["n_wav"])
    try:
        transit.last_timing = timing
    except Exception:
        pass
    return R


def install(gasprop_module, voigt: bool = True) -> None:
    """Rebind the reference's hot-path methods to the CUDA library ("what calls it").

    `gasprop_module` is the reference's `pythonScripts.gasProperties` module.  After this call the
    unchanged `prometheus.py` produces the same output file through the GPU.
    """
    capi.load()   # fail now, not in the middle of a run, if the library is missing
    T = gasprop_module.Transit
    if "sumOverChords" not in _saved:
        _saved["sumOverChords"] = T.sumOverChords
        _saved["module"] = gasprop_module

    def sumOverChords(self, max_memory_gb: float = 2.0):
        return sum_over_chords(self, max_memory_gb)

    sumOverChords.__doc__ = "prometheus_b200: CUDA implementation (max_memory_gb is ignored)"
    T.sumOverChords = sumOverChords

    if voigt:
        A = gasprop_module.AtmosphericConstituent
        _saved.setdefault("calculateVoigtProfile", A.calculateVoigtProfile)

        def calculateVoigtProfile(self, wavelength):
            lam0, gamma, f = self.getLineParameters(wavelength)
            return host.get_context().voigt_sigma(lam0, gamma, f, self.sigma_v, np.asarray(wavelength, dtype=np.float64),
                                                  self.lookupOffset)

        A.calculateVoigtProfile = calculateVoigtProfile


def uninstall() -> None:
    mod = _saved.pop("module", None)
    if mod is None:
        return
    mod.Transit.sumOverChords = _saved.pop("sumOverChords")
    if "calculateVoigtProfile" in _saved:
        mod.AtmosphericConstituent.calculateVoigtProfile = _saved.pop("calculateVoigtProfile")
