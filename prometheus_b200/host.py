"""Host-side mirror of the reference's object model for the transit hot path.

Same class names, constructor arguments, attribute names and method names as the reference
(pythonScripts/gasProperties.py, celestialBodies.py, geometryHandler.py), so code and tests written
against the reference read the same here -- but every method that is on the hot path
 forwards to the CUDA library through `prometheus_b200.capi`:

    Transit.sumOverChords                      -> pb200_set_scene + pb200_run
    <Scenario>.calculateNumberDensity          -> pb200_number_density_chords
    AtmosphericConstituent.calculateVoigtProfile -> pb200_voigt_sigma
    AtmosphericConstituent.getSigmaAbs         -> pb200_sigma_lookup
    MolecularConstituent.getSigmaAbs           -> pb200_molecular_sigma
    Transit.getLOSopticalDepthGrid             -> pb200_optical_depth

Only O(axis) set-up work (axes, per-phase orbit positions, the wavelength grid, line selection)
is done in numpy here.  Nothing in this module computes tau, sigma, densities or R on the CPU.
"""
from __future__ import annotations

import copy
import math

import numpy as np

from . import capi
from . import constants as const
from . import scene as _scene

# ------------------------------------------------------------------------------------------------
# line list (reference: module global LINE_LIST, gasProperties.py:29-30) -- (n, 5) str array with
# columns element, ionisation stage, lambda_vac [A], A_ki, f_ik.  The reference ships it as
# Resources/LineList.txt; here it is injected (set_line_list) -- from the reference module in
# drop-in use, from tests/golden/linelist.npz or a synthetic generator otherwise.
# ------------------------------------------------------------------------------------------------
LINE_LIST = np.empty((0, 5), dtype=str)


_line_cache: dict = {}


def set_line_list(rows) -> None:
    global LINE_LIST
    LINE_LIST = np.asarray(rows, dtype=str).reshape(-1, 5)
    _line_cache.clear()


def species_lines(element: str, ionizationState: str):
    """(lambda0 [cm], gamma = A/(4 pi), f) of every line of one species, parsed ONCE per line list: the
    reference converts the whole str array on every getLineParameters call (gasProperties.py:640-670; SURVEY 8(f)
    rank 3) -- three calls per constituent and set-up."""
    key = (element, ionizationState)
    hit = _line_cache.get(key)
    if hit is None:
        rows = LINE_LIST
        if rows.shape[0] == 0:
            raise RuntimeError("line list is empty: call prometheus_b200.host.set_line_list(...) first")
        mine = (rows[:, 0] == element) & (rows[:, 1] == ionizationState)
        mine &= (rows[:, 2] != "") & (rows[:, 3] != "") & (rows[:, 4] != "")
        hit = (rows[mine, 2].astype(float) * 1e-8, rows[mine, 3].astype(float) / (4.0 * np.pi), rows[mine, 4].astype(float))
        for a in hit:
            a.setflags(write=False)
        _line_cache[key] = hit
    return hit


_contexts: dict[int, capi.Context] = {}


def get_context(device: int | None = None) -> capi.Context:
    """Process-wide pb200 context of a GPU (created on first use)."""
    if device is None:
        device = _default_device[0]
    ctx = _contexts.get(device)
    if ctx is None or ctx.handle is None:
        ctx = _contexts[device] = capi.Context(device)
    return ctx


_default_device = [0]


def set_default_device(device: int) -> None:
    _default_device[0] = int(device)


# ------------------------------------------------------------------------------------------------
# small containers standing in for scipy's interp1d / RegularGridInterpolator objects, which the
# reference only uses as (x, y) / (grid, values) holders on this path (gasProperties.py:730-735)
# ------------------------------------------------------------------------------------------------
def _table_array(a) -> np.ndarray:
    """float64, contiguous; a large table is moved to page-locked memory once, so that every pb200_set_scene sends it
    at DMA speed from where it lies instead of through a staging copy (C2: 50 MB of molecular tables, 30 -> 2 ms).
    Without a CUDA device (CPU-only tests of the host logic) it stays an ordinary array."""
    a = np.ascontiguousarray(a, dtype=np.float64)
    base = a
    while isinstance(base, np.ndarray) and base.base is not None:
        base = base.base
    if isinstance(base, capi._PinnedBlock):     # already page-locked (e.g. built by pb200_voigt_sigma)
        return a
    if a.nbytes >= (1 << 20):
        try:
            pinned = capi.pinned_pool.empty(a.shape)
        except Exception:
            return a
        pinned[...] = a
        return pinned
    return a


class TableFunction1D:
    def __init__(self, x, y, fill_value=None):
        self.x = _table_array(x)
        self.y = _table_array(y)
        self.fill_value = fill_value


class GridFunction:
    def __init__(self, grid, values, fill_value=None):
        self.grid = tuple(np.ascontiguousarray(g, dtype=np.float64) for g in grid)
        self.values = _table_array(values)
        self.fill_value = fill_value


# ------------------------------------------------------------------------------------------------
# species (constants.py:52-162)
# ------------------------------------------------------------------------------------------------
class Species:
    def __init__(self, name, element, ionizationState, mass):
        self.name, self.element, self.ionizationState, self.mass = name, element, ionizationState, mass


class AvailableSpecies:
    def __init__(self):
        self.speciesList = [Species(n, el, ion, m * const.amu) for n, (el, ion, m) in const.SPECIES_TABLE.items()]

    def findSpecies(self, nameSpecies):
        for sp in self.speciesList:
            if sp.name == nameSpecies:
                return sp
        print("Species", nameSpecies, "was not found.")
        return None

    def listSpeciesNames(self):
        return [sp.name for sp in self.speciesList]


# ------------------------------------------------------------------------------------------------
# bodies (celestialBodies.py): circular, edge-on orbits; the observer looks along +x
# ------------------------------------------------------------------------------------------------
class Star:
    def __init__(self, R, dR=0.0, M=0.0, dM=0.0, T_eff=0.0, dT_eff=0.0, log_g=0.0, dlog_g=0.0, Z=0.0, dZ=0.0,
                 alpha=0.0):
        self.R, self.dR, self.M, self.dM = R, dR, M, dM
        self.T_eff, self.dT_eff, self.log_g, self.dlog_g, self.Z, self.dZ, self.alpha = T_eff, dT_eff, log_g, dlog_g, Z, dZ, alpha
        self.CLV_u1 = 0.0
        self.CLV_u2 = 0.0
        self.vsiniStarrot = 0.0
        self.phiStarrot = 0.0
        self.Fstar_function = None

    def addCLVparameters(self, CLV_u1, CLV_u2):
        self.CLV_u1, self.CLV_u2 = CLV_u1, CLV_u2

    def addRMparameters(self, vsiniStarrot, phiStarrot):
        self.vsiniStarrot, self.phiStarrot = vsiniStarrot, phiStarrot

    def setSpectrum(self, wavelength_cm, flux):
        """Attach a stellar spectrum as arrays: what getSpectrum returns (wavelength [cm], flux / pi) -- the
        file-format-free door; the reference fetches a PHOENIX FITS pair over the network (celestialBodies.py:128-210)."""
        self.spectrum = (np.asarray(wavelength_cm, dtype=np.float64), np.asarray(flux, dtype=np.float64))

    def getSpectrum(self):
        if getattr(self, "spectrum", None) is None:
            raise NotImplementedError("no stellar spectrum attached: the PHOENIX download (celestialBodies.py:128-210) needs "
                                      "the network and astropy; use star.setSpectrum(wavelength_cm, flux)")
        return self.spectrum

    def addFstarFunction(self, wavelength):
        """log10 F*(lambda) table over the simulation's wavelength range widened by the rotational Doppler factors, one
        table point of margin below and two above (celestialBodies.py:242-265); the log10 runs on the GPU."""
        w_star, flux = self.getSpectrum()
        w_hi = np.max(wavelength) * const.calculateDopplerShift(-self.vsiniStarrot)
        w_lo = np.min(wavelength) * const.calculateDopplerShift(self.vsiniStarrot)
        inside = np.flatnonzero((w_star >= w_lo) & (w_star <= w_hi))
        first, last = max(int(inside[0]) - 1, 0), int(inside[-1]) + 2
        self.Fstar_function = TableFunction1D(w_star[first:last], get_context().log_table(flux[first:last]))


class Planet:
    def __init__(self, name, R, M, a, hostStar, transitDuration=0.0, orbitalPeriod=0.0, b=0.0):
        self.name, self.R, self.M, self.a, self.hostStar = name, R, M, a, hostStar
        self.transitDuration, self.orbitalPeriod, self.b = transitDuration, orbitalPeriod, b

    def getPosition(self, orbphase):            # celestialBodies.py:370-383
        return self.a * np.cos(orbphase), self.a * np.sin(orbphase)

    def getLOSvelocity(self, orbphase):         # celestialBodies.py:385-397
        return -np.sin(orbphase) * np.sqrt(const.G * self.hostStar.M / self.a)


class Moon:
    def __init__(self, midTransitOrbphase, R, a, hostPlanet):
        self.midTransitOrbphase, self.R, self.a, self.hostPlanet = midTransitOrbphase, R, a, hostPlanet

    def getOrbphase(self, orbphase):            # celestialBodies.py:499-516
        pl = self.hostPlanet
        a_p, M_p = np.float64(pl.a), np.float64(pl.M)
        a_m, M_s = np.float64(self.a), np.float64(pl.hostStar.M)
        ratio = np.sqrt((a_p ** 3 * M_p) / (a_m ** 3 * M_s))
        return self.midTransitOrbphase + np.float64(orbphase) * ratio

    def getPosition(self, orbphase):            # celestialBodies.py:518-531
        ph = self.getOrbphase(orbphase)
        x_p, y_p = self.hostPlanet.getPosition(orbphase)
        return x_p + self.a * np.cos(ph), y_p + self.a * np.sin(ph)

    def getLOSvelocity(self, orbphase):         # celestialBodies.py:533-550
        v_p = self.hostPlanet.getLOSvelocity(orbphase)
        ph = self.getOrbphase(orbphase)
        return v_p - np.sin(ph) * np.sqrt(const.G * self.hostPlanet.M / self.a)


# ------------------------------------------------------------------------------------------------
# spatial grid (geometryHandler.py:15-208): midpoint axes + the flattened chord grid
# ------------------------------------------------------------------------------------------------
def _cell_midpoints(lo, hi, n, half_step):
    return np.linspace(lo, hi, int(n), endpoint=False) + half_step


class Grid:
    def __init__(self, x_midpoint, x_border, x_steps, rho_border, rho_steps, phi_steps, orbphase_border, orbphase_steps):
        self.x_midpoint, self.x_border, self.x_steps = x_midpoint, x_border, x_steps
        self.rho_border, self.rho_steps, self.phi_steps = rho_border, rho_steps, phi_steps
        self.orbphase_border, self.orbphase_steps = orbphase_border, orbphase_steps

    @staticmethod
    def getCartesianFromCylinder(phi, rho):
        return rho * np.sin(phi), rho * np.cos(phi)

    def getDeltaX(self):
        return 2.0 * self.x_border / float(self.x_steps)

    def getDeltaRho(self):
        return self.rho_border / float(self.rho_steps)

    def getDeltaPhi(self):
        return 2.0 * np.pi / float(self.phi_steps)

    def constructXaxis(self, midpoints=True):
        lo, hi = self.x_midpoint - self.x_border, self.x_midpoint + self.x_border
        if midpoints:
            return _cell_midpoints(lo, hi, self.x_steps, self.x_border / float(self.x_steps))
        return np.linspace(lo, hi, int(self.x_steps) + 1)

    def constructRhoAxis(self, midpoints=True):
        if midpoints:
            return _cell_midpoints(0.0, self.rho_border, self.rho_steps, 0.5 * self.rho_border / float(self.rho_steps))
        return np.linspace(0.0, self.rho_border, int(self.rho_steps) + 1)

    def constructPhiAxis(self, midpoints=True):
        if midpoints:
            return _cell_midpoints(0, 2 * np.pi, self.phi_steps, np.pi / float(self.phi_steps))
        return np.linspace(0, 2 * np.pi, int(self.phi_steps) + 1)

    def constructOrbphaseAxis(self):
        return np.linspace(-self.orbphase_border, self.orbphase_border, int(self.orbphase_steps))

    def getChordGrid(self):
        """(n_phi*n_rho*n_orb, 3) rows (phi, rho, orbphase); flat index (i_phi*n_rho + i_rho)*n_orb + i_orb."""
        phi, rho, orb = self.constructPhiAxis(), self.constructRhoAxis(), self.constructOrbphaseAxis()
        out = np.empty((phi.size, rho.size, orb.size, 3))
        out[..., 0] = phi[:, None, None]
        out[..., 1] = rho[None, :, None]
        out[..., 2] = orb[None, None, :]
        return out.reshape(-1, 3)
