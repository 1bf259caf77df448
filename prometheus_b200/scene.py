"""Flattens a Transit object graph into the pb200_scene POD of the C ABI.

Works on anything that quacks like the reference's objects ("state read through the
seam"): the reference's own `Transit` (drop-in use) or the host mirror in
`prometheus_b200.host`.  Only O(axis) host work happens here -- axes, per-phase body positions and
Doppler factors are evaluated by the objects' own methods, so chord/grid indexing is bit-exact with
whatever host code produced them.  Everything O(chords) and above runs on the GPU.
"""
from __future__ import annotations

import ctypes as C
import math

import numpy as np

from . import capi
from . import constants as const

_PLANET_CENTRED = {"BarometricAtmosphere", "HydrostaticAtmosphere", "PowerLawAtmosphere",
                   "PowerLawExosphere", "TorusExosphere"}
_MOON_CENTRED = {"MoonExosphere", "TidallyHeatedMoon"}


def _kind(obj) -> str:
    """Scenario class name; subclasses of the reference classes resolve through the MRO."""
    for klass in type(obj).__mro__:
        if klass.__name__ in _PLANET_CENTRED or klass.__name__ in _MOON_CENTRED or klass.__name__ == "SerpensExosphere":
            return klass.__name__
    raise capi.Pb200Error(-4, f"unsupported density scenario {type(obj).__name__}")


def scenario_parameters(dist) -> tuple[int, list[float]]:
    """(pb200_scenario_type, p[8]) -- the scalar part of SURVEY 8(a) row a7."""
    name = _kind(dist)
    kB, G = const.k_B, const.G
    if name == "BarometricAtmosphere":          # gasProperties.py:156-158
        pl = dist.planet
        n0 = dist.P_0 / (kB * dist.T)
        H = kB * dist.T * pl.R ** 2 / (G * dist.mu * pl.M)
        return capi.SCN_BAROMETRIC, [n0, pl.R, H]
    if name == "HydrostaticAtmosphere":         # gasProperties.py:198-202
        pl = dist.planet
        n0 = dist.P_0 / (kB * dist.T)
        return capi.SCN_HYDROSTATIC, [n0, pl.R, G * dist.mu * pl.M, kB * dist.T]
    if name == "PowerLawAtmosphere":            # gasProperties.py:241-242
        pl = dist.planet
        return capi.SCN_POWERLAW, [dist.P_0 / (kB * dist.T), pl.R, dist.q]
    if name == "PowerLawExosphere":             # gasProperties.py:327
        pl = dist.planet
        return capi.SCN_POWERLAW, [(dist.q - 3.0) / (4.0 * np.pi * pl.R ** 3) * dist.N, pl.R, dist.q]
    if name == "MoonExosphere":                 # gasProperties.py:372
        mo = dist.moon
        return capi.SCN_POWERLAW, [(dist.q - 3.0) / (4.0 * np.pi * mo.R ** 3) * dist.N, mo.R, dist.q]
    if name == "TidallyHeatedMoon":             # gasProperties.py:459 (N enters per phase)
        mo = dist.moon
        return capi.SCN_POWERLAW, [(dist.q - 3.0) / (4.0 * np.pi * mo.R ** 3), mo.R, dist.q]
    if name == "TorusExosphere":                # gasProperties.py:507-514
        pl = dist.planet
        v_orbit = np.sqrt(G * pl.M / dist.a_torus)
        H = dist.a_torus * dist.v_ej / v_orbit
        term1 = 8.0 * H ** 2 * np.exp(-dist.a_torus ** 2 / (16.0 * H ** 2))
        term2 = 2.0 * np.sqrt(np.pi) * dist.a_torus * H * (math.erf(dist.a_torus / (4.0 * H)) + 1.0)
        n0 = 1.0 / (2.0 * np.pi ** 1.5 * H * (term1 + term2)) * dist.N
        return capi.SCN_TORUS, [n0, dist.a_torus, H]
    if name == "SerpensExosphere":
        return capi.SCN_SERPENS, []
    raise capi.Pb200Error(-4, f"unsupported density scenario {name}")


class SceneBuilder:
    """Collects numpy arrays (kept alive for the duration of pb200_set_scene) and fills the structs."""

    def __init__(self):
        self.keep = []

    def arr(self, a) -> np.ndarray:
        a = capi.as_f64(a)
        self.keep.append(a)
        return a

    def ptr(self, a):
        return capi.dptr(self.arr(a))

    def per_phase(self, v, n_orb):
        v = np.asarray(v, dtype=np.float64)
        return self.ptr(np.full(n_orb, v) if v.ndim == 0 else np.broadcast_to(v, (n_orb,)))


def _trim_table(tx, ty, q_lo, q_hi):
    """Sub-table that answers every np.interp query in [q_lo, q_hi] exactly like the full table:
    the queries stay strictly inside the kept range unless they were outside the full one."""
    n = tx.size
    i0 = max(int(np.searchsorted(tx, q_lo, side="right")) - 2, 0)
    i1 = min(int(np.searchsorted(tx, q_hi, side="left")) + 2, n)
    if i1 - i0 < 2:
        i0, i1 = max(min(i0, n - 2), 0), min(max(i1, i0 + 2), n)
    return tx[i0:i1], ty[i0:i1]


_AXES_CACHE: dict = {}
_GRID_FIELDS = ("x_midpoint", "x_border", "x_steps", "rho_border", "rho_steps", "phi_steps", "orbphase_border", "orbphase_steps")


def _axes(grid, phi_rot):
    """(x, rho, phase, sin phi, cos phi, cos(phi - phi_rot), delta_x) of a Grid, evaluated by the grid's own methods.
    The axes depend only on the grid's eight constructor arguments (geometryHandler.py:64-71) and phiStarrot, so they
    are kept per (class, arguments): a retrieval or a multi-GPU loop calls sumOverChords thousands of times on one
    grid, and the ~10 small numpy calls were a quarter of the host time of a call.  Objects without those attributes
    are evaluated every time."""
    try:
        key = (type(grid),) + tuple(float(getattr(grid, f)) for f in _GRID_FIELDS) + (float(phi_rot),)
    except (AttributeError, TypeError, ValueError):
        key = None
    hit = _AXES_CACHE.get(key) if key is not None else None
    if hit is None:
        x = np.asarray(grid.constructXaxis(), dtype=np.float64)
        rho = np.asarray(grid.constructRhoAxis(), dtype=np.float64)
        phi = np.asarray(grid.constructPhiAxis(), dtype=np.float64)
        phase = np.asarray(grid.constructOrbphaseAxis(), dtype=np.float64)
        hit = (x, rho, phase, np.sin(phi), np.cos(phi), np.cos(phi - phi_rot), float(grid.getDeltaX()))
        for a in hit[:6]:
            a.setflags(write=False)
        if key is not None:
            if len(_AXES_CACHE) > 64:
                _AXES_CACHE.clear()
            _AXES_CACHE[key] = hit
    return hit


def flatten(transit, wavelength=None, wav_range=None) -> tuple[capi.Scene, list, dict]:
    """Transit-like object -> (pb200_scene, keep-alive list, info).

    `wav_range = (lo, hi)`: the caller will only run wavelengths [lo, hi) of this scene (one rank of
    a wavelength-sharded run); the atomic sigma tables are then cut down to the part those
    wavelengths can reach through any phase's Doppler factor, which divides the upload by the
    number of ranks without changing a single result bit."""
    b = SceneBuilder()
    grid = transit.spatialGrid
    atmos = transit.atmosphere
    dists = list(atmos.densityDistributionList)
    planet = getattr(transit, "planet", None) or dists[0].planet
    star = planet.hostStar
    if wavelength is None:
        wavelength = transit.wavelength
    wavelength = np.asarray(wavelength, dtype=np.float64)

    x, rho, phase, sin_phi, cos_phi, cos_phi_rot, delta_x = _axes(grid, star.phiStarrot)
    n_orb = phase.size

    sc = capi.Scene()
    sc.abi_version = capi.ABI_VERSION
    sc.n_x, sc.n_rho, sc.n_phi, sc.n_orb, sc.n_wav = x.size, rho.size, sin_phi.size, n_orb, wavelength.size
    sc.x, sc.rho = b.ptr(x), b.ptr(rho)
    sc.sin_phi, sc.cos_phi = b.ptr(sin_phi), b.ptr(cos_phi)                      # geometryHandler.py:84-85
    sc.cos_phi_rot = b.ptr(cos_phi_rot)                                          # gasProperties.py:1182
    sc.wavelength = b.ptr(wavelength)
    sc.delta_x = delta_x
    sc.star_R, sc.clv_u1, sc.clv_u2 = float(star.R), float(star.CLV_u1), float(star.CLV_u2)
    sc.vsini = float(star.vsiniStarrot)
    fstar = getattr(star, "Fstar_function", None)
    if fstar is not None:
        sc.fstar_x, sc.fstar_y, sc.n_fstar = b.ptr(fstar.x), b.ptr(fstar.y), int(np.size(fstar.x))
    sc.planet_R = float(planet.R)
    sc.planet_y = b.ptr(planet.a * np.sin(phase))                                # gasProperties.py:1224

    moons_R, moons_y = [], []
    for dist in dists:
        if getattr(dist, "hasMoon", False):                                      # gasProperties.py:1226-1230
            moons_R.append(float(dist.moon.R))
            moons_y.append(np.broadcast_to(np.asarray(dist.moon.getPosition(phase)[1], dtype=np.float64), (n_orb,)))
    sc.n_moons = len(moons_R)
    if moons_R:
        sc.moon_R = b.ptr(np.array(moons_R))
        sc.moon_y = b.ptr(np.concatenate(moons_y))

    scn_arr = (capi.Scenario * len(dists))()
    cons_list = []
    shifts_of = {}
    for i, dist in enumerate(dists):
        q = scn_arr[i]
        q.type, params = scenario_parameters(dist)
        for j, v in enumerate(params):
            q.p[j] = float(v)
        q.T = float(getattr(dist, "T", 0.0) or 0.0)
        has_moon = bool(getattr(dist, "hasMoon", False))
        body = dist.moon if has_moon else dist.planet
        if q.type != capi.SCN_SERPENS:
            cx, cy = body.getPosition(phase)                                     # celestialBodies.py:370-383, :518-531
            q.centre_x, q.centre_y = b.per_phase(cx, n_orb), b.per_phase(cy, n_orb)
        else:
            tab = dist.InterpolatedDensity
            gx, gy, gz = tab.grid
            q.gx, q.gy, q.gz = b.ptr(gx), b.ptr(gy), b.ptr(gz)
            vals = np.asarray(tab.values, dtype=np.float64)
            q.values = b.ptr(vals)
            q.dims[0], q.dims[1], q.dims[2] = vals.shape
        if atmos.hasOrbitalDopplerShift:                                         # gasProperties.py:908-916
            v_los = body.getLOSvelocity(phase)
        else:
            v_los = np.zeros(n_orb)
        shift = np.broadcast_to(np.asarray(const.calculateDopplerShift(-np.asarray(v_los, dtype=np.float64)), dtype=np.float64), (n_orb,))
        shifts_of[id(dist)] = shift
        q.shift = b.per_phase(shift, n_orb)
        if _kind(dist) == "TidallyHeatedMoon":                                   # gasProperties.py:425-436
            q.amplitude = b.per_phase(dist.calculateAbsorberNumber(phase), n_orb)
        q.first_constituent = len(cons_list)
        q.n_constituents = len(dist.constituents)
        for con in dist.constituents:
            cons_list.append((con, dist))

    cons_arr = (capi.Constituent * max(len(cons_list), 1))()
    n_atoms = n_mols = 0
    for k, (con, dist) in enumerate(cons_list):
        q = cons_arr[k]
        q.chi, q.offset = float(con.chi), float(con.lookupOffset)
        lf = con.lookupFunction
        if con.isMolecule:
            q.kind = 1
            gP, gT, gW = lf.grid                                                 # gasProperties.py:779
            vals = np.asarray(lf.values, dtype=np.float64)
            q.grid_P, q.grid_T, q.grid_w, q.values = b.ptr(gP), b.ptr(gT), b.ptr(gW), b.ptr(vals)
            q.dims[0], q.dims[1], q.dims[2] = vals.shape
            fill = getattr(lf, "fill_value", None)
            q.fill = float(fill) if fill is not None else float("nan")
            n_mols += 1
        else:
            q.kind = 0
            tx, ty = np.asarray(lf.x), np.asarray(lf.y)                                  # gasProperties.py:730-735
            if wav_range is not None and tx.size > 4 and wav_range[1] > wav_range[0]:
                sh = shifts_of[id(dist)]
                lam_lo, lam_hi = wavelength[wav_range[0]], wavelength[wav_range[1] - 1]
                cand = np.outer(sh, [lam_lo, lam_hi])
                tx, ty = _trim_table(tx, ty, float(cand.min()), float(cand.max()))
            q.tab_x, q.tab_y, q.n_tab = b.ptr(tx), b.ptr(ty), int(tx.size)
            n_atoms += 1
    sc.n_scenarios, sc.n_constituents = len(dists), len(cons_list)
    sc.scenarios = C.cast(scn_arr, C.POINTER(capi.Scenario))
    sc.constituents = C.cast(cons_arr, C.POINTER(capi.Constituent))
    b.keep += [scn_arr, cons_arr]

    h2d = sum(a.nbytes for a in b.keep if isinstance(a, np.ndarray))
    info = {"n_x": x.size, "n_rho": rho.size, "n_phi": sin_phi.size, "n_orb": n_orb, "n_wav": wavelength.size,
            "n_chords": rho.size * sin_phi.size * n_orb, "n_atoms": n_atoms, "n_mols": n_mols,
            "h2d_bytes": int(h2d)}
    return sc, b.keep, info


def flatten_model(transit, phase=None) -> tuple[capi.Model, list]:
    """The part of a Transit-like object that a free parameter of a retrieval can move (pb200_model):
    scenario scalars and per-phase arrays, evaluated with the objects' own methods exactly as `flatten`
    does.  Tables, axes and the wavelength grid are NOT read -- they stay resident from the template's
    pb200_set_scene (the caller guarantees they are the same; the library checks types and counts)."""
    b = SceneBuilder()
    atmos = transit.atmosphere
    dists = list(atmos.densityDistributionList)
    planet = getattr(transit, "planet", None) or dists[0].planet
    if phase is None:
        phase = np.asarray(transit.spatialGrid.constructOrbphaseAxis(), dtype=np.float64)
    n_orb = phase.size
    mdl = capi.Model()
    mdl.planet_y = b.ptr(planet.a * np.sin(phase))                               # gasProperties.py:1224
    mdl.planet_R = float(planet.R)
    moons_R, moons_y = [], []
    for dist in dists:
        if getattr(dist, "hasMoon", False):                                      # gasProperties.py:1226-1230
            moons_R.append(float(dist.moon.R))
            moons_y.append(np.broadcast_to(np.asarray(dist.moon.getPosition(phase)[1], dtype=np.float64), (n_orb,)))
    if moons_R:
        mdl.moon_R = b.ptr(np.array(moons_R))
        mdl.moon_y = b.ptr(np.concatenate(moons_y))
    scn_arr = (capi.Scenario * len(dists))()
    first = 0
    for i, dist in enumerate(dists):
        q = scn_arr[i]
        q.type, params = scenario_parameters(dist)
        for j, v in enumerate(params):
            q.p[j] = float(v)
        q.T = float(getattr(dist, "T", 0.0) or 0.0)
        has_moon = bool(getattr(dist, "hasMoon", False))
        body = dist.moon if has_moon else dist.planet
        if q.type != capi.SCN_SERPENS:
            cx, cy = body.getPosition(phase)                                     # celestialBodies.py:370-383, :518-531
            q.centre_x, q.centre_y = b.per_phase(cx, n_orb), b.per_phase(cy, n_orb)
        v_los = body.getLOSvelocity(phase) if atmos.hasOrbitalDopplerShift else np.zeros(n_orb)   # gasProperties.py:908-916
        q.shift = b.per_phase(const.calculateDopplerShift(-np.asarray(v_los, dtype=np.float64)), n_orb)
        if _kind(dist) == "TidallyHeatedMoon":                                   # gasProperties.py:425-436
            q.amplitude = b.per_phase(dist.calculateAbsorberNumber(phase), n_orb)
        q.first_constituent = first
        q.n_constituents = len(dist.constituents)
        first += q.n_constituents
    mdl.scenarios = C.cast(scn_arr, C.POINTER(capi.Scenario))
    b.keep.append(scn_arr)
    return mdl, b.keep


def flatten_models_broadcast(transit, phase, n_models: int):
    """pb200_model structs of a whole family at once.  `transit` is ONE Transit-like object whose free
    attributes (moon.midTransitOrbphase, moon.a, moon.R, dist.N, dist.q ...) are arrays of shape
    (n_models, 1): the objects' own numpy expressions then broadcast against phase[None, :] to
    (n_models, n_orb) -- the same expressions as one flatten_model call per model at a tenth of the Python
    cost.  numpy's SIMD sin/cos are not position independent, so single values can differ from the per-model
    call in the last bit (as they do inside the reference between batch sizes); integers, types and scalars
    are identical (tests/test_retrieval_cpu.py::test_broadcast_family_matches_per_model).
    -> (models as a numpy struct array with the layout of pb200_model, keep-alive list)."""
    K = int(n_models)
    atmos = transit.atmosphere
    dists = list(atmos.densityDistributionList)
    planet = getattr(transit, "planet", None) or dists[0].planet
    phase = np.asarray(phase, dtype=np.float64)
    n_orb = phase.size
    ph = phase[None, :]
    keep = []

    def rows(v):
        a = np.ascontiguousarray(np.broadcast_to(np.asarray(v, dtype=np.float64), (K, n_orb)))
        keep.append(a)
        return a.ctypes.data + np.arange(K, dtype=np.uint64) * np.uint64(n_orb * 8)

    models = np.zeros(K, dtype=capi.struct_dtype(capi.Model))
    scn = np.zeros((K, len(dists)), dtype=capi.struct_dtype(capi.Scenario))
    keep.append(scn)
    models["scenarios"] = scn.ctypes.data + np.arange(K, dtype=np.uint64) * np.uint64(scn.strides[0])
    models["planet_y"] = rows(planet.a * np.sin(ph))                             # gasProperties.py:1224
    models["planet_R"] = np.broadcast_to(np.asarray(planet.R, dtype=np.float64).reshape(-1), (K,))
    moons_R, moons_y = [], []
    for dist in dists:
        if getattr(dist, "hasMoon", False):                                      # gasProperties.py:1226-1230
            moons_R.append(np.broadcast_to(np.asarray(dist.moon.R, dtype=np.float64), (K, 1)))
            moons_y.append(np.broadcast_to(np.asarray(dist.moon.getPosition(ph)[1], dtype=np.float64), (K, n_orb)))
    if moons_R:
        mR = np.ascontiguousarray(np.concatenate(moons_R, axis=1))               # [K][n_moons]
        mY = np.ascontiguousarray(np.stack(moons_y, axis=1))                     # [K][n_moons][n_orb]
        keep += [mR, mY]
        models["moon_R"] = mR.ctypes.data + np.arange(K, dtype=np.uint64) * np.uint64(mR.strides[0])
        models["moon_y"] = mY.ctypes.data + np.arange(K, dtype=np.uint64) * np.uint64(mY.strides[0])
    first = 0
    for i, dist in enumerate(dists):
        q = scn[:, i]
        typ, params = scenario_parameters(dist)
        q["type"] = typ
        for j, v in enumerate(params):
            q["p"][:, j] = np.broadcast_to(np.asarray(v, dtype=np.float64).reshape(-1), (K,))
        q["T"] = float(getattr(dist, "T", 0.0) or 0.0)
        body = dist.moon if getattr(dist, "hasMoon", False) else dist.planet
        if typ != capi.SCN_SERPENS:
            cx, cy = body.getPosition(ph)                                        # celestialBodies.py:370-383, :518-531
            q["centre_x"], q["centre_y"] = rows(cx), rows(cy)
        v_los = body.getLOSvelocity(ph) if atmos.hasOrbitalDopplerShift else np.zeros((1, n_orb))   # gasProperties.py:908-916
        q["shift"] = rows(const.calculateDopplerShift(-np.asarray(v_los, dtype=np.float64)))
        if _kind(dist) == "TidallyHeatedMoon":                                   # gasProperties.py:425-436
            q["amplitude"] = rows(dist.calculateAbsorberNumber(ph))
        q["first_constituent"] = first
        q["n_constituents"] = len(dist.constituents)
        first += len(dist.constituents)
    return models, keep
