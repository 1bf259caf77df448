"""Shared test plumbing: oracle inputs for a workload spec, parity criterion."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)

import prometheus_oracle as oracle  # noqa: E402
from prometheus_b200 import builder, workloads  # noqa: E402


def oracle_inputs(spec, nist_rows):
    """Synthetic input tables of a workload in the form oracle.build_model expects."""
    inputs = {"linelist": builder.resolve_line_list(spec, nist_rows), "molecules": {}, "fstar": None,
              "serpens": None, "mdot": None}
    for s in spec["scenarios"]:
        for mol in s.get("molecules", []):
            inputs["molecules"][mol["table"]] = workloads.synthetic_molecule_table(mol["table"])
        if s["type"] == "serpens":
            inputs["serpens"] = workloads.synthetic_serpens_table(spec["grid"], None, s["table"]["synthetic"])
        if s["type"] == "tidallyHeated":
            inputs["mdot"] = workloads.synthetic_mdot_table(s["mdot"]["synthetic"])
    fs = spec["bodies"]["star"].get("fstar")
    if fs:
        inputs["fstar"] = workloads.synthetic_fstar_table(spec["wavelength"][0], spec["wavelength"][1], fs["synthetic"])
    return inputs


def oracle_model(name, nist_rows):
    spec = workloads.get(name)
    return oracle.build_model(spec, oracle_inputs(spec, nist_rows))


def excess_depth_ok(R, R_ref, R0_ref, rtol=1e-6, atol=1e-10):
    """north-star parity criterion: E = R0 - R, |E - E_ref| <= atol + rtol |E_ref|."""
    E, E_ref = R0_ref - R, R0_ref - R_ref
    err = np.abs(E - E_ref)
    bound = atol + rtol * np.abs(E_ref)
    return bool(np.all(err <= bound)), float(np.max(err / bound))
