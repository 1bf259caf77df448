"""oracle/make_ref.py -- TEST INFRASTRUCTURE: puts the UNMODIFIED reference where the GPU box can run it.

    python oracle/make_ref.py            (build container only: needs /root/reference)

The reference is pure Python, so "building" it is a copy: /root/reference -> oracle/_ref/Prometheus (git-ignored, NOT
gpurun-ignored: it travels with the snapshot like the built .so; no reference source enters the history).  Next to it:
  oracle/_ref/stubs/          empty stand-ins for h5py and astropy.io.fits -- absent from this image, imported at module
                              level by the reference (gasProperties.py:14, celestialBodies.py), never reached on the
                              transit path
  oracle/_ref/setupFiles/     where prometheus.py looks for <setup_name>.txt   (prometheus.py:21, :56)
  oracle/_ref/output/         where prometheus.py writes <setup_name>.txt       (prometheus.py:156)
  oracle/_ref/MANIFEST.json   sha256 of every copied file

Users: bench.py --impl reference / cpu_baseline (kind "reference"), tests/test_dropin_cli.py.  `load()` imports the
reference's modules from the copy; nothing under prometheus_b200/ may call it.
"""
from __future__ import annotations

import hashlib
import json
import os
import shutil
import sys
import types

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SRC = "/root/reference"
REF_ROOT = os.path.join(HERE, "_ref")
REF_DIR = os.path.join(REF_ROOT, "Prometheus")
STUBS = os.path.join(REF_ROOT, "stubs")
KEEP = ("prometheus.py", "mainRetrieval.py", "__init__.py", "LICENSE", "README.md", "pythonScripts", "Resources")


def available() -> bool:
    return os.path.exists(os.path.join(REF_DIR, "pythonScripts", "gasProperties.py"))


def make(force: bool = False) -> str | None:
    """Copy the reference (when /root/reference exists); returns the directory or None."""
    if not os.path.isdir(REF_SRC):
        return REF_DIR if available() else None
    if available() and not force:
        return REF_DIR
    if os.path.isdir(REF_DIR):
        shutil.rmtree(REF_DIR)
    os.makedirs(REF_DIR)
    manifest = {}
    for name in KEEP:
        src = os.path.join(REF_SRC, name)
        dst = os.path.join(REF_DIR, name)
        if os.path.isdir(src):
            shutil.copytree(src, dst, ignore=shutil.ignore_patterns("__pycache__", "*.pyc"))
        elif os.path.exists(src):
            shutil.copy2(src, dst)
    for dirpath, _, files in os.walk(REF_DIR):
        for f in sorted(files):
            p = os.path.join(dirpath, f)
            manifest[os.path.relpath(p, REF_DIR)] = hashlib.sha256(open(p, "rb").read()).hexdigest()
    for d in ("setupFiles", "output", os.path.join("stubs", "astropy", "io")):
        os.makedirs(os.path.join(REF_ROOT, d), exist_ok=True)
    for rel in ("h5py.py", os.path.join("astropy", "__init__.py"), os.path.join("astropy", "io", "__init__.py"),
                os.path.join("astropy", "io", "fits.py")):
        with open(os.path.join(STUBS, rel), "w") as fh:
            fh.write("# empty stand-in: the package is absent from this image and never reached on the transit path\n")
    with open(os.path.join(REF_ROOT, "MANIFEST.json"), "w") as fh:
        json.dump({"source": REF_SRC, "files": manifest}, fh, indent=1, sort_keys=True)
    return REF_DIR


def pythonpath() -> list[str]:
    """sys.path entries a process needs to run the copy: the reference directory (prometheus.py imports
    `pythonScripts.*` top-level, prometheus.py:14-18) and the stubs -- the latter only for packages really missing."""
    import importlib.util
    out = [REF_DIR]
    for name in ("h5py", "astropy"):
        try:
            spec = importlib.util.find_spec(name)
        except (ImportError, ValueError):
            spec = None
        if spec is None or (spec.origin or "").startswith(STUBS):
            out.append(STUBS)
            break
    return out


def load():
    """-> (gasProperties, celestialBodies, geometryHandler, constants) of the UNMODIFIED reference copy."""
    if not available():
        raise FileNotFoundError("oracle/_ref/Prometheus is missing: run `python oracle/make_ref.py` in the build container")
    for p in reversed(pythonpath()):
        if p not in sys.path:
            sys.path.insert(0, p)
    import pythonScripts.celestialBodies as bodies
    import pythonScripts.constants as consts
    import pythonScripts.gasProperties as gasprop
    import pythonScripts.geometryHandler as geom
    return gasprop, bodies, geom, consts


class ReferenceNamespace:
    """Where prometheus_b200.builder.build_transit finds the classes when it is to build the REFERENCE's own object
    graph (the same role tests/golden/make_golden.py's namespace plays against /root/reference)."""

    def __init__(self):
        import numpy as np
        from scipy.interpolate import RegularGridInterpolator, interp1d
        self.gas, self.bodies, self.geom, self.consts = load()
        self.Species = self.consts.Species
        self.findSpecies = lambda name: self.consts.AvailableSpecies().findSpecies(name)
        self._interp1d, self._rgi, self._np = interp1d, RegularGridInterpolator, np
        if not hasattr(self.gas, "_pb200_nist_rows"):
            self.gas._pb200_nist_rows = np.array(self.gas.LINE_LIST, dtype=str)     # as loaded from Resources/LineList.txt
        self.nist_rows = self.gas._pb200_nist_rows

    def set_line_list(self, rows):
        self.gas.LINE_LIST = self._np.asarray(rows, dtype=str)

    def table_1d(self, x, y):
        return self._interp1d(x, y, kind="linear")

    def table_nd(self, grid, values, fill_value=None, bounds_error=False):
        if bounds_error:
            return self._rgi(grid, values)
        return self._rgi(grid, values, bounds_error=False, fill_value=fill_value)


if __name__ == "__main__":
    d = make(force="--force" in sys.argv)
    print(d if d else "reference not available here (no /root/reference and no oracle/_ref copy)")
