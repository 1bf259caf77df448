"""Result file of `python prometheus.py <setup_name>` (reference: prometheus.py:142-156).

`save_output` writes output/<setup_name>.txt byte for byte as the reference's np.savetxt call does,
through the multi-threaded C writer of libprometheus_b200 (pb200_write_output): at C5 the table is
165 MB of '%.18e' text, tens of seconds in np.savetxt, which would dominate the run once R takes 0.2 s.
"""
from __future__ import annotations

import numpy as np

from . import capi

HEADER = ("Prometheus output file.\nFirst row: Orbital phases [1]\n"
          "All other rows: Wavelength [cm] (first column), Transit depth R(orbital phase, wavelength) [1] (other columns)")


def save_output(path: str, wavelength, orbphase, R, header: str = HEADER, threads: int = 0) -> None:
    """path <- [nan, orbphase/2pi] / [lambda_i, R[:, i]] rows (prometheus.py:149-156)."""
    wavelength, orbphase = capi.as_f64(wavelength), capi.as_f64(orbphase)
    R = capi.as_f64(R)
    if R.shape != (orbphase.size, wavelength.size):
        raise ValueError(f"R has shape {R.shape}, expected {(orbphase.size, wavelength.size)}")
    rc = capi.load().pb200_write_output(str(path).encode(), header.encode(), capi.dptr(wavelength), capi.dptr(orbphase),
                                        capi.dptr(R), orbphase.size, wavelength.size, int(threads))
    if rc != 0:
        raise capi.Pb200Error(rc, f"cannot write {path}")


def save_output_reference(path: str, wavelength, orbphase, R, header: str = HEADER) -> None:
    """The reference's own lines (prometheus.py:149-156), for parity tests and timing."""
    orbphaseOutput = np.insert(np.asarray(orbphase) / (2. * np.pi), 0, np.nan)
    mainOutput = np.vstack((wavelength, R))
    output = np.vstack((orbphaseOutput, mainOutput.T))
    np.savetxt(path, output, header=header)
