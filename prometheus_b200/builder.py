"""spec dict (prometheus_b200.workloads) -> Transit object graph.

`build_transit(spec)` builds the host mirror (prometheus_b200.host / .gas).  The same routine, given a
namespace that points at the reference's modules, builds the reference's own objects through its
public classes -- that is how tests/golden/make_golden.py produces the fixtures, and it is the
constructive equivalent of what the reference's prometheus.py does with a setup file
(reference: prometheus.py:66-134).
"""
from __future__ import annotations

import os
import tempfile

import numpy as np

from . import constants as const
from . import workloads


class MirrorNamespace:
    """Where the classes live.  For the reference: gas = gasProperties, bodies = celestialBodies,
    geom = geometryHandler, consts = constants (see tests/golden/make_golden.py)."""

    def __init__(self):
        from . import gas, host
        self.gas, self.bodies, self.geom, self.host = gas, host, host, host
        self.Species = host.Species
        self.findSpecies = lambda name: host.AvailableSpecies().findSpecies(name)

    def set_line_list(self, rows):
        self.host.set_line_list(rows)

    def table_1d(self, x, y):
        return self.host.TableFunction1D(x, y)

    def table_nd(self, grid, values, fill_value=None, bounds_error=False):
        return self.host.GridFunction(grid, values, fill_value)


def resolve_line_list(spec, nist_rows=None):
    ll = spec.get("linelist", "nist")
    if ll == "nist":
        if nist_rows is None:
            raise ValueError("this workload needs the NIST line list rows (tests/golden/linelist.npz)")
        return np.asarray(nist_rows, dtype=str)
    return workloads.synthetic_fe_linelist(ll["synthetic_fe"], ll.get("n", 4096), ll.get("lo_A", 2000.0),
                                           ll.get("hi_A", 8000.0))


def build_transit(spec: dict, ns=None, nist_rows=None, build_tables: bool = True):
    ns = ns or MirrorNamespace()
    ns.set_line_list(resolve_line_list(spec, nist_rows))
    b = spec["bodies"]
    st = b["star"]
    star = ns.bodies.Star(st["R"], 0.0, st["M"], 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0)
    if st.get("u1") or st.get("u2"):
        star.addCLVparameters(st["u1"], st["u2"])
    if st.get("vsini"):
        star.addRMparameters(st["vsini"], st.get("phi_rot", 0.0))
    pl = b["planet"]
    planet = ns.bodies.Planet(b.get("planetName", "planet"), pl["R"], pl["M"], pl["a"], star, 0.0, 0.0, 0.0)
    moon = None
    if "moon" in b:
        mo = b["moon"]
        moon = ns.bodies.Moon(mo["phi0"], mo["R"], mo["a"], planet)

    wgrid = ns.gas.WavelengthGrid(*spec["wavelength"])
    g = spec["grid"]
    sgrid = ns.geom.Grid(g[0], g[1], int(g[2]), g[3], int(g[4]), int(g[5]), g[6], int(g[7]))

    scenarios = []
    for s in spec["scenarios"]:
        kind = s["type"]
        collisional = kind in ("barometric", "hydrostatic") or (kind == "powerLaw" and "P_0" in s)
        if kind == "barometric":
            dist = ns.gas.BarometricAtmosphere(s["T"], s["P_0"], s["mu"], planet)
        elif kind == "hydrostatic":
            dist = ns.gas.HydrostaticAtmosphere(s["T"], s["P_0"], s["mu"], planet)
        elif kind == "powerLaw" and collisional:
            dist = ns.gas.PowerLawAtmosphere(s["T"], s["P_0"], s["q_esc"], planet)
        elif kind == "powerLaw":
            dist = ns.gas.PowerLawExosphere(s["N"], s["q_esc"], planet)
        elif kind == "exomoon":
            dist = ns.gas.MoonExosphere(s["N"], s["q"], moon)
        elif kind == "tidallyHeated":
            dist = ns.gas.TidallyHeatedMoon(s["q"], moon)
        elif kind == "torus":
            dist = ns.gas.TorusExosphere(s["N"], s["a_torus"], s["v_ej"], planet)
        elif kind == "serpens":
            dist = ns.gas.SerpensExosphere("", s["N"], planet, 0.0)
            gx, gy, gz, vals = workloads.synthetic_serpens_table(spec["grid"], planet, s["table"]["synthetic"])
            dist.InterpolatedDensity = ns.table_nd((gx, gy, gz), vals, None, bounds_error=True)
        else:
            raise ValueError(f"unknown scenario type {kind}")

        for sp in s.get("species", []):
            name = sp["name"]
            if name in const.SPECIES_TABLE:
                if collisional:
                    dist.addConstituent(name, sp["chi"])
                else:
                    dist.addConstituent(name, sp["sigma_v"])
            else:   # not in the reference registry (Fe II): build the constituent directly, SURVEY 8(c).3
                el, ion, mass = const.EXTRA_SPECIES[name]
                species = ns.Species(name, el, ion, mass * const.amu)
                if collisional:
                    con = ns.gas.AtmosphericConstituent(species, sp["chi"], dist.getVelDispersion(species.mass))
                    dist.constituents.append(con)
                else:
                    dist.constituents = [ns.gas.AtmosphericConstituent(species, 1.0, sp["sigma_v"])]
        for mol in s.get("molecules", []):
            dist.addMolecularConstituent(mol["name"], mol["chi"])
            P, T, lam, logsig = workloads.synthetic_molecule_table(mol["table"])
            dist.constituents[-1].lookupFunction = ns.table_nd((P, T, lam), logsig, np.log10(const.LOOKUP_OFFSET))
        if kind == "tidallyHeated":
            mdot = workloads.synthetic_mdot_table(s["mdot"]["synthetic"])
            fd, path = tempfile.mkstemp(suffix=".txt")
            os.close(fd)
            try:
                np.savetxt(path, mdot)
                mass = ns.findSpecies(s["species"][0]["name"]).mass
                dist.addSourceRateFunction(path, s["tau_photoionization"], mass)
            finally:
                os.unlink(path)
        scenarios.append(dist)

    if build_tables:
        for dist in scenarios:
            for con in dist.constituents:
                if not con.isMolecule:
                    con.addLookupFunctionToConstituent(wgrid)

    if st.get("fstar"):
        w, logF = workloads.synthetic_fstar_table(spec["wavelength"][0], spec["wavelength"][1], st["fstar"]["synthetic"])
        star.Fstar_function = ns.table_1d(w, logF)

    atmos = ns.gas.Atmosphere(scenarios, bool(spec["doppler"]))
    transit = ns.gas.Transit(atmos, wgrid, sgrid)
    transit.addWavelength()
    return transit
