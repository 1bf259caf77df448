"""Synthetic transit setups ("workloads") of the shapes BASELINE.json names.

A workload is a plain, JSON-serialisable ``spec`` dict.  Its scenario / species /
grid keys follow the reference's setup-file schema (reference:
pythonScripts/setup.py:269-304, :436-479, :523-544); the body scalars are written out
(instead of a planets.csv lookup) so a spec is self-contained.  ``build_transit`` in
``prometheus_b200.host`` turns a spec into host objects; the same function, handed the
reference's modules, builds the reference's objects (tests/golden/make_golden.py).

SURVEY.md section 8(d) defines C1..C5; the ``*s`` twins are the CI-sized parity cases.
"""
from __future__ import annotations

import copy
import math

import numpy as np

from . import constants as const

# body scalars as the reference derives them from Resources/planets.csv / stars.csv
# (celestialBodies.py:597-638): R = R_J * col, M = M_J * col, a = AU * col, ...
BODIES = {
    "WASP-49b": {"planet": {"R": 8564741600.0, "M": 7.57302e+29, "a": 579400800000.0},
                 "star": {"R": 72244800000.0, "M": 1.993964e+33}},
    "HD189733b": {"planet": {"R": 8135789599.999999, "M": 2.159924e+30, "a": 466752000000.0},
                  "star": {"R": 54043286400.0, "M": 1.645768712962724e+33}},
    "KELT-9b": {"planet": {"R": 12747023600.0, "M": 4.63112e+30, "a": 517915200000.0},
                "star": {"R": 159226008000.0, "M": 4.832828e+33}},
}


def _bodies(name, **star_extra):
    b = copy.deepcopy(BODIES[name])
    b["planetName"] = name
    b["star"].update({"u1": 0.0, "u2": 0.0, "vsini": 0.0, "phi_rot": 0.0, "fstar": None})
    b["star"].update(star_extra)
    return b


def synthetic_molecule_table(k: int):
    """Deterministic ExoMol-shaped sigma(P, T, lambda) table.

    Returns (P [dyn/cm^2], T [K], lambda [cm], log10(sigma + 1e-50)) with the axis order
    the reference's interpolator uses (gasProperties.py:779).
    """
    P = np.logspace(-4, 8, 13)
    T = np.linspace(100.0, 3400.0, 12)
    lam = np.linspace(2900e-8, 8100e-8, 20001)
    lam_A = lam * 1e8
    sig = (1e-22 * (1.0 + np.sin(lam_A / 3.0 + k) ** 2)[None, None, :]
           * (P ** 0.1)[:, None, None] * (T / 1000.0)[None, :, None])
    return P, T, lam, np.log10(sig + const.LOOKUP_OFFSET)


def synthetic_fe_linelist(seed: int = 20260411, n: int = 4096, lo_A=2000.0, hi_A=8000.0):
    """Seeded synthetic Fe I / Fe II line list in the reference's (n, 5) str layout
    (columns element, sp_num, lambda_vac[A], A_ki, f_ik; gasProperties.py:29-30)."""
    rng = np.random.default_rng(seed)
    lam0 = np.sort(rng.uniform(lo_A, hi_A, n))
    A = 10 ** rng.uniform(6.0, 8.5, n)
    f = 10 ** rng.uniform(-3.0, 0.0, n)
    ion = np.where(rng.uniform(size=n) < 0.5, "1", "2")
    rows = [["Fe", ion[i], "%.4f" % lam0[i], "%.3e" % A[i], "%.3e" % f[i]] for i in range(n)]
    return np.array(rows, dtype=str)


def synthetic_fstar_table(lower_w, upper_w, n=4001):
    """Smooth synthetic log10 F*(lambda) table with two broad stellar lines (a3 test input)."""
    w = np.linspace(lower_w * 0.999, upper_w * 1.001, n)
    mid = 0.5 * (lower_w + upper_w)
    width = 0.05 * (upper_w - lower_w)
    F = 1e6 * (1.0 - 0.4 * np.exp(-((w - mid) / width) ** 2)
               - 0.2 * np.exp(-((w - (mid + 3 * width)) / (0.5 * width)) ** 2))
    return w, np.log10(F)


def synthetic_serpens_table(grid, planet, seed=7):
    """Smooth synthetic 3-D density histogram on the (x, y, z) midpoints the reference's
    SerpensExosphere builds (gasProperties.py:570-574); stands in for a SERPENS file."""
    x_mid, x_b, n_x, rho_b, n_rho = grid[0], grid[1], int(grid[2]), grid[3], int(grid[4])
    xP = np.linspace(x_mid - x_b, x_mid + x_b, n_x, endpoint=False) + x_b / float(n_x)
    yP = np.linspace(-rho_b, rho_b, 2 * n_rho, endpoint=False) + 2.0 * rho_b / (4.0 * n_rho)
    zP = yP.copy()
    X, Y, Z = np.meshgrid(xP, yP, zP, indexing="ij")
    rng = np.random.default_rng(seed)
    r2 = ((X - x_mid) / (0.5 * x_b)) ** 2 + (Y / (0.3 * rho_b)) ** 2 + (Z / (0.2 * rho_b)) ** 2
    vals = 1e4 * np.exp(-r2) * (1.0 + 0.1 * rng.standard_normal(r2.shape))
    return xP, yP, zP, np.abs(vals)


def synthetic_serpens_particles(grid, n=20000, seed=3):
    """(n, 6) rows in METRES like a SERPENS output file (positions, velocities; the reference reads columns 0-2 and
    multiplies by 1e2, gasProperties.py:554-555): a Gaussian cloud around the grid's x midpoint, some particles
    outside the grid, a few exactly on bin edges (first, inner and last edge)."""
    rng = np.random.default_rng(seed)
    x_mid, x_b, n_x, rho_b, n_rho = grid[0], grid[1], int(grid[2]), grid[3], int(grid[4])
    pos = np.empty((n, 3))
    pos[:, 0] = rng.normal(x_mid, 0.6 * x_b, n)
    pos[:, 1] = rng.normal(0.0, 0.4 * rho_b, n)
    pos[:, 2] = rng.normal(0.0, 0.4 * rho_b, n)
    xe = np.linspace(x_mid - x_b, x_mid + x_b, n_x + 1)
    ye = np.linspace(-rho_b, rho_b, 2 * n_rho + 1)
    for k, (i, j) in enumerate([(0, 0), (n_x, 2 * n_rho), (n_x // 2, n_rho), (1, 2 * n_rho), (n_x, 3)]):
        pos[k] = (xe[i], ye[j], ye[min(j + 1, 2 * n_rho)])
    out = np.zeros((n, 6))
    out[:, 0:3] = pos / 1e2
    return out


def synthetic_mdot_table(n=64):
    """Half-orbit mass-loss curve for TidallyHeatedMoon (gasProperties.py:418-423)."""
    ph = np.linspace(0.0, 1.0, n)
    return 1e7 * (1.0 + 0.8 * np.sin(math.pi * ph) ** 2)


def _w49(**kw):
    return _bodies("WASP-49b", **kw)


def _make():
    W = {}
    w49 = BODIES["WASP-49b"]
    Rp, ap, Rs = w49["planet"]["R"], w49["planet"]["a"], w49["star"]["R"]

    W["C1"] = {
        "name": "C1", "title": "barometric hot-Jupiter, Na I D, low-res, 5 phases",
        "bodies": _w49(), "doppler": True, "linelist": "nist",
        "scenarios": [{"type": "barometric", "T": 3000.0, "P_0": 1e3, "mu": 2.3 * const.amu,
                       "species": [{"name": "NaI", "chi": 1e-6}]}],
        "wavelength": [5888e-8, 5900e-8, 2e-8, 5e-9, 2e-10],
        "grid": [ap, 5 * Rp, 30, Rs, 40, 60, 0.25, 5],
    }

    hd = BODIES["HD189733b"]
    Rp2, ap2, Rs2 = hd["planet"]["R"], hd["planet"]["a"], hd["star"]["R"]
    atoms = [("NaI", 1e-6), ("KI", 1e-7), ("MgI", 1e-6), ("CaI", 1e-7), ("CaII", 1e-8),
             ("FeI", 1e-6), ("TiI", 1e-8), ("TiII", 1e-8)]
    W["C2"] = {
        "name": "C2", "title": "hydrostatic, 8 atoms/ions + 2 synthetic molecules, mid-res",
        "bodies": _bodies("HD189733b"), "doppler": True, "linelist": "nist",
        "scenarios": [{"type": "hydrostatic", "T": 3000.0, "P_0": 1e4, "mu": 2.3 * const.amu,
                       "species": [{"name": n, "chi": c} for n, c in atoms],
                       "molecules": [{"name": "SYN1", "chi": 1e-4, "table": 1},
                                     {"name": "SYN2", "chi": 1e-4, "table": 2}]}],
        "wavelength": [3000e-8, 8000e-8, 1e-8, 2e-9, 1e-10],
        "grid": [ap2, 5 * Rp2, 50, Rs2, 50, 60, 0.1, 11],
    }
    c2s = copy.deepcopy(W["C2"])
    c2s.update(name="C2s", title="C2 reduced twin (CI parity case)")
    c2s["wavelength"] = [5800e-8, 6000e-8, 1e-8, 2e-9, 1e-10]
    c2s["grid"] = [ap2, 5 * Rp2, 20, Rs2, 20, 24, 0.1, 5]
    W["C2s"] = c2s
    c2a = copy.deepcopy(W["C2"])          # atoms only: the part of C2 the atomic kernel runs
    c2a.update(name="C2a", title="C2 without the molecules (8 atoms/ions)")
    c2a["scenarios"][0]["molecules"] = []
    W["C2a"] = c2a

    moon = {"phi0": 0.65 * 2 * math.pi, "R": 0.1384 * const.R_Io, "a": 1.44 * Rp}
    c3b = _w49()
    c3b["moon"] = moon
    W["C3"] = {
        "name": "C3", "title": "exomoon Na I + torus K I, orbital Doppler, 401 phases",
        "bodies": c3b, "doppler": True, "linelist": "nist",
        "scenarios": [{"type": "exomoon", "N": 1e33, "q": 3.34,
                       "species": [{"name": "NaI", "sigma_v": 1e6}]},
                      {"type": "torus", "N": 1e33, "a_torus": 2 * Rp, "v_ej": 2e5,
                       "species": [{"name": "KI", "sigma_v": 1e6}]}],
        "wavelength": [5880e-8, 7710e-8, 2e-8, 5e-9, 2e-10],
        "grid": [ap, 5 * Rp, 30, Rs, 40, 60, 0.25, 401],
    }
    c3s = copy.deepcopy(W["C3"])
    c3s.update(name="C3s", title="C3 reduced twin with CLV (CI parity case)")
    c3s["bodies"]["star"].update(u1=0.34, u2=0.28)
    c3s["wavelength"] = [5885e-8, 5900e-8, 2e-8, 5e-9, 2e-10]
    c3s["scenarios"][1]["species"] = [{"name": "NaI", "sigma_v": 1e6}]
    c3s["grid"] = [ap, 5 * Rp, 30, Rs, 20, 24, 0.25, 9]
    W["C3s"] = c3s

    W["C4"] = {
        "name": "C4", "title": "torus Na I, high-res wavelength grid",
        "bodies": _w49(), "doppler": True, "linelist": "nist",
        "scenarios": [{"type": "torus", "N": 1e33, "a_torus": 2 * Rp, "v_ej": 2e5,
                       "species": [{"name": "NaI", "sigma_v": 1e6}]}],
        "wavelength": [5880e-8, 5905e-8, 10e-8, 5e-9, 5e-12],
        "grid": [ap, 5 * Rp, 100, Rs, 100, 90, 0.25, 51],
    }
    c4s = copy.deepcopy(W["C4"])
    c4s.update(name="C4s", title="C4 reduced twin (CI parity case)")
    c4s["wavelength"] = [5885e-8, 5900e-8, 2e-8, 5e-9, 1e-10]
    c4s["grid"] = [ap, 5 * Rp, 40, Rs, 20, 18, 0.25, 7]
    W["C4s"] = c4s

    k9 = BODIES["KELT-9b"]
    Rp5, ap5, Rs5 = k9["planet"]["R"], k9["planet"]["a"], k9["star"]["R"]
    W["C5"] = {
        "name": "C5", "title": "powerLaw atmosphere, 4096 synthetic Fe I/Fe II lines, fine chord grid",
        "bodies": _bodies("KELT-9b"), "doppler": True,
        "linelist": {"synthetic_fe": 20260411, "n": 4096, "lo_A": 2000.0, "hi_A": 8000.0},
        "scenarios": [{"type": "powerLaw", "T": 4000.0, "P_0": 1e3, "q_esc": 6.0,
                       "species": [{"name": "FeI", "chi": 1e-5}, {"name": "FeII", "chi": 1e-5}]}],
        "wavelength": [2000e-8, 8000e-8, 1e-8, 1e-8, 1e-10],
        "grid": [ap5, 8 * Rp5, 200, Rs5, 200, 180, 0.2, 21],
    }
    c5s = copy.deepcopy(W["C5"])
    c5s.update(name="C5s", title="C5 reduced twin: 96 synthetic Fe lines (CI parity case)")
    c5s["linelist"] = {"synthetic_fe": 20260411, "n": 96, "lo_A": 5000.0, "hi_A": 5200.0}
    c5s["wavelength"] = [5000e-8, 5200e-8, 1e-8, 1e-8, 2e-10]
    c5s["grid"] = [ap5, 8 * Rp5, 40, Rs5, 24, 20, 0.2, 5]
    W["C5s"] = c5s

    # ---- additional CI cases for the remaining rows of SURVEY 8(a) -----------------------
    W["PLNs"] = {
        "name": "PLNs", "title": "number-normalised powerLaw exosphere, Na I (a7 powerLaw(N))",
        "bodies": _w49(), "doppler": False, "linelist": "nist",
        "scenarios": [{"type": "powerLaw", "N": 1e34, "q_esc": 4.5,
                       "species": [{"name": "NaI", "sigma_v": 5e5}]}],
        "wavelength": [5885e-8, 5900e-8, 2e-8, 5e-9, 2e-10],
        "grid": [ap, 6 * Rp, 36, Rs, 22, 20, 0.2, 3],
    }
    thm = _w49()
    thm["moon"] = dict(moon)
    W["THMs"] = {
        "name": "THMs", "title": "tidally heated moon, phase-dependent N (a7 tidally heated)",
        "bodies": thm, "doppler": True, "linelist": "nist",
        "scenarios": [{"type": "tidallyHeated", "q": 3.34, "mdot": {"synthetic": 64},
                       "tau_photoionization": 240.0,
                       "species": [{"name": "NaI", "sigma_v": 1e6}]}],
        "wavelength": [5885e-8, 5900e-8, 2e-8, 5e-9, 2e-10],
        "grid": [ap, 5 * Rp, 30, Rs, 20, 24, 0.25, 9],
    }
    W["SERs"] = {
        "name": "SERs", "title": "serpens-style tabulated 3-D density, Na I (a7 serpens)",
        "bodies": _w49(), "doppler": True, "linelist": "nist",
        "scenarios": [{"type": "serpens", "N": 1e33, "table": {"synthetic": 7},
                       "species": [{"name": "NaI", "sigma_v": 1e6}]}],
        "wavelength": [5885e-8, 5900e-8, 2e-8, 5e-9, 2e-10],
        "grid": [ap, 5 * Rp, 24, Rs, 16, 20, 0.25, 5],
    }
    fst = _w49(u1=0.34, u2=0.28, vsini=3e5, phi_rot=0.3, fstar={"synthetic": 4001})
    W["FSTs"] = {
        "name": "FSTs", "title": "barometric Na I with F* table, CLV and stellar rotation (a2, a3)",
        "bodies": fst, "doppler": True, "linelist": "nist",
        "scenarios": [{"type": "barometric", "T": 3000.0, "P_0": 1e3, "mu": 2.3 * const.amu,
                       "species": [{"name": "NaI", "chi": 1e-6}]}],
        "wavelength": [5885e-8, 5900e-8, 2e-8, 5e-9, 2e-10],
        "grid": [ap, 5 * Rp, 30, Rs, 20, 24, 0.25, 5],
    }
    W["HYDs"] = {
        "name": "HYDs", "title": "hydrostatic Na I + powerLaw(P) K I, two collisional scenarios",
        "bodies": _w49(), "doppler": True, "linelist": "nist",
        "scenarios": [{"type": "hydrostatic", "T": 2500.0, "P_0": 1e2, "mu": 2.3 * const.amu,
                       "species": [{"name": "NaI", "chi": 1e-6}]},
                      {"type": "powerLaw", "T": 3000.0, "P_0": 1e1, "q_esc": 5.0,
                       "species": [{"name": "NaI", "chi": 1e-5}]}],
        "wavelength": [5885e-8, 5900e-8, 2e-8, 5e-9, 2e-10],
        "grid": [ap, 5 * Rp, 30, Rs, 20, 24, 0.25, 5],
    }
    # ---- retrieval forward model: the setup of mainRetrieval.py:18-67 ---------------
    ret = _w49(u1=0.34, u2=0.28)                                   # mainRetrieval.py:37 CLV of Wyttenbach+2017
    ret["moon"] = dict(moon)                                        # mainRetrieval.py:18-21, :42-43
    W["RET"] = {
        "name": "RET", "title": "retrieval forward model: tidally heated moon Na I, 25 phases (mainRetrieval.py)",
        "bodies": ret, "doppler": True, "linelist": "nist",
        "scenarios": [{"type": "tidallyHeated", "q": 3.34, "mdot": {"synthetic": 64},
                       "tau_photoionization": 241.0,
                       "species": [{"name": "NaI", "sigma_v": 1e6}]}],
        "wavelength": [5888e-8, 5900e-8, 2e-8, 5e-9, 2e-10],       # mainRetrieval.py:28
        "grid": [ap, 5 * Rp, 30, Rs, 40, 60, 0.25, 25],            # mainRetrieval.py:32-33
        "band": {"line_centres": [5891.583253e-8, 5897.558147e-8], "bandwidth": 0.75e-8},   # mainRetrieval.py:75-81
    }
    rets = copy.deepcopy(W["RET"])
    rets.update(name="RETs", title="RET reduced twin (CI parity case)")
    rets["grid"] = [ap, 5 * Rp, 30, Rs, 20, 24, 0.25, 9]
    W["RETs"] = rets
    return W


WORKLOADS = _make()

# free parameters of the retrieval sketch (mainRetrieval.py:18-21): moon phase at mid-transit [rad], moon orbit
# [planetary radii], moon radius [Io radii]
RETRIEVAL_PARAMS = ("midTransitOrbphase", "a_moon", "R_moon")


def retrieval_variant(spec: dict, midTransitOrbphase: float, a_moon: float, R_moon: float) -> dict:
    """The spec with the three free parameters of mainRetrieval.py:18-21 set."""
    v = copy.deepcopy(spec)
    v["bodies"]["moon"] = {"phi0": float(midTransitOrbphase), "R": float(R_moon) * const.R_Io,
                           "a": float(a_moon) * v["bodies"]["planet"]["R"]}
    return v


def retrieval_samples(n: int, seed: int = 11) -> np.ndarray:
    """(n, 3) parameter vectors around the values of mainRetrieval.py:18-21 (seeded; row 0 is the script's own)."""
    rng = np.random.default_rng(seed)
    out = np.empty((n, 3))
    out[:, 0] = rng.uniform(0.0, 2.0 * math.pi, n)
    out[:, 1] = rng.uniform(1.2, 2.5, n)
    out[:, 2] = rng.uniform(0.05, 0.5, n)
    out[0] = (0.65 * 2.0 * math.pi, 1.44, 0.1384)
    return out

# the CI-sized cases with committed golden fixtures (tests/golden/*.npz)
GOLDEN_CASES = ["C1", "C2s", "C3s", "C4s", "C5s", "PLNs", "THMs", "SERs", "FSTs", "HYDs"]


def get(name: str) -> dict:
    return copy.deepcopy(WORKLOADS[name])


def shape(spec: dict) -> dict:
    g = spec["grid"]
    return {"n_x": int(g[2]), "n_rho": int(g[4]), "n_phi": int(g[5]), "n_orb": int(g[7]),
            "n_chords": int(g[4]) * int(g[5]) * int(g[7])}


# ---------------------------------------------------------------------------------------------------
# the reference's setup-file schema (written by pythonScripts/setup.py:207-211, :258-304, :368-391, :424-479,
# :521-544 as JSON into <parent>/setupFiles/<name>.txt; read by prometheus.py:56-134)
# ---------------------------------------------------------------------------------------------------
_CLI_SCENARIOS = ("barometric", "hydrostatic", "powerLaw", "exomoon", "torus", "serpens")


def to_setup_file(spec: dict) -> dict:
    """The setup-file twin of a workload spec, for `python prometheus.py <name>`.  Raises ValueError for what the
    reference's CLI cannot express: CLV / stellar rotation / an F* table (prometheus.py never applies them), two
    scenarios of one type (the file is keyed by type), species outside the reference's registry, synthetic molecular
    tables, tidally heated moons, and the number-normalised power law (prometheus.py:85 indexes dict.keys())."""
    st = spec["bodies"]["star"]
    if st.get("u1") or st.get("u2") or st.get("vsini") or st.get("fstar"):
        raise ValueError("the reference CLI applies neither CLV nor stellar rotation nor an F* table")
    if spec.get("linelist", "nist") != "nist":
        raise ValueError("the reference CLI reads Resources/LineList.txt only")
    fundamentals = {"ExomoonSource": False, "DopplerPlanetRotation": False, "CLV_variations": False, "RM_effect": False,
                    "DopplerOrbitalMotion": bool(spec["doppler"])}
    scenarios, species = {}, {}
    arch = {"planetName": spec["bodies"]["planetName"]}
    for s in spec["scenarios"]:
        kind = s["type"]
        if kind not in _CLI_SCENARIOS or kind in scenarios:
            raise ValueError(f"scenario {kind!r}: not expressible in a setup file (unknown to the CLI, or given twice)")
        if s.get("molecules"):
            raise ValueError("synthetic molecular tables cannot be named in a setup file")
        collisional = kind in ("barometric", "hydrostatic") or (kind == "powerLaw" and "P_0" in s)
        if kind in ("barometric", "hydrostatic"):
            scenarios[kind] = {"T": s["T"], "P_0": s["P_0"], "mu": s["mu"]}
        elif kind == "powerLaw":
            if not collisional:
                raise ValueError("number-normalised power law: prometheus.py:85 fails on dict.keys()[0]")
            scenarios[kind] = {"q_esc": s["q_esc"], "P_0": s["P_0"], "T": s["T"]}
        elif kind == "exomoon":
            mo = spec["bodies"]["moon"]
            fundamentals["ExomoonSource"] = True
            scenarios[kind] = {"q_moon": s["q"]}
            arch.update(R_moon=mo["R"], a_moon=mo["a"], starting_orbphase_moon=mo["phi0"])
        elif kind == "torus":
            scenarios[kind] = {"a_torus": s["a_torus"], "v_ej": s["v_ej"]}
        else:
            scenarios[kind] = {"serpensPath": s["serpensPath"]}
        species[kind] = {}
        for sp in s.get("species", []):
            if sp["name"] not in const.SPECIES_TABLE:
                raise ValueError(f"species {sp['name']} is not in the reference's registry")
            species[kind][sp["name"]] = {"chi": sp["chi"]} if collisional else {"sigma_v": sp["sigma_v"], "Nparticles": s["N"]}
    w, g = spec["wavelength"], spec["grid"]
    grids = {"lower_w": w[0], "upper_w": w[1], "widthHighRes": w[2], "resolutionLow": w[3], "resolutionHigh": w[4],
             "x_midpoint": g[0], "x_border": g[1], "x_steps": int(g[2]), "upper_rho": g[3], "rho_steps": int(g[4]),
             "phi_steps": int(g[5]), "orbphase_border": g[6], "orbphase_steps": int(g[7])}
    return {"Fundamentals": fundamentals, "Scenarios": scenarios, "Architecture": arch, "Species": species, "Grids": grids}


def from_setup_file(param: dict, name: str = "setup", bodies: dict | None = None) -> dict:
    """A setup file (parsed JSON) -> workload spec.  The planet's scalars come from `bodies` (default: BODIES, the
    values the reference derives from its catalogue for the planets used here)."""
    arch, grids = param["Architecture"], param["Grids"]
    b = copy.deepcopy((bodies or BODIES)[arch["planetName"]])
    b["planetName"] = arch["planetName"]
    b["star"].update({"u1": 0.0, "u2": 0.0, "vsini": 0.0, "phi_rot": 0.0, "fstar": None})
    if "starting_orbphase_moon" in arch:
        b["moon"] = {"phi0": arch["starting_orbphase_moon"], "R": arch["R_moon"], "a": arch["a_moon"]}
    scenarios = []
    for kind, prm in param["Scenarios"].items():
        sp = param["Species"].get(kind, {})
        s = {"type": kind}
        if "T" in prm:                       # collisional (prometheus.py:112)
            s.update({k: prm[k] for k in ("T", "P_0", "mu", "q_esc") if k in prm})
            s["species"] = [{"name": n, "chi": d["chi"]} for n, d in sp.items()]
        else:
            first = next(iter(sp.values()))
            s["N"] = first["Nparticles"]
            if kind == "exomoon":
                s["q"] = prm["q_moon"]
            elif kind == "torus":
                s.update(a_torus=prm["a_torus"], v_ej=prm["v_ej"])
            elif kind == "powerLaw":
                s["q_esc"] = prm["q_esc"]
            elif kind == "serpens":
                s["serpensPath"] = prm["serpensPath"]
            s["species"] = [{"name": n, "sigma_v": d["sigma_v"]} for n, d in sp.items()]
        scenarios.append(s)
    return {"name": name, "title": f"setup file {name}", "bodies": b, "doppler": bool(param["Fundamentals"]["DopplerOrbitalMotion"]),
            "linelist": "nist", "scenarios": scenarios,
            "wavelength": [grids[k] for k in ("lower_w", "upper_w", "widthHighRes", "resolutionLow", "resolutionHigh")],
            "grid": [grids["x_midpoint"], grids["x_border"], int(grids["x_steps"]), grids["upper_rho"], int(grids["rho_steps"]),
                     int(grids["phi_steps"]), grids["orbphase_border"], int(grids["orbphase_steps"])]}


def cli_twin(name: str) -> dict:
    """Workload `name` in the form the reference's CLI can run: C3 / C3s without the CLV of the direct-API variant."""
    spec = get(name)
    spec["bodies"]["star"].update(u1=0.0, u2=0.0)
    return spec
