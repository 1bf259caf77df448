"""Gas properties side of the host mirror: density scenarios, constituents, wavelength grid,
Atmosphere and Transit (reference: pythonScripts/gasProperties.py).  See host.py for the rules:
same names and signatures as the reference, hot-path methods forward to the CUDA library.
"""
from __future__ import annotations

import copy

import numpy as np

from . import capi
from . import constants as const
from . import host
from . import scene as _scene


# ------------------------------------------------------------------------------------------------
# density scenarios (gasProperties.py:53-601)
# ------------------------------------------------------------------------------------------------
class _Scenario:
    hasMoon = False

    def _centre(self, orbphase):
        body = self.moon if self.hasMoon else self.planet
        return body.getPosition(orbphase)

    def _amplitude(self, orbphase):
        return None

    def calculateNumberDensity(self, x, phi, rho, orbphase):
        """n(x; phi, rho, orbphase) on the GPU (kernel (a), materialised).

        Scalar (phi, rho, orbphase) -> (n_x,); arrays of shape (n_chords,) -> (n_chords, n_x),
        as the reference's batch-broadcast methods (gasProperties.py:143-601).
        """
        x = np.atleast_1d(np.asarray(x, dtype=np.float64))
        scalar = np.ndim(phi) == 0 and np.ndim(rho) == 0 and np.ndim(orbphase) == 0
        phi_, rho_, orb_ = np.broadcast_arrays(np.atleast_1d(np.asarray(phi, dtype=np.float64)),
                                               np.atleast_1d(np.asarray(rho, dtype=np.float64)),
                                               np.atleast_1d(np.asarray(orbphase, dtype=np.float64)))
        n = phi_.size
        y, z = host.Grid.getCartesianFromCylinder(phi_, rho_)
        ctx = host.get_context()
        sc, keep = _mini_scene(self, x)
        ctx.set_scene(sc, keep)
        out = np.empty((n, x.size), dtype=np.float64)
        if _scene._kind(self) == "SerpensExosphere":
            cx = cy = None
        else:
            cx, cy = self._centre(orb_)
            cx = capi.as_f64(np.broadcast_to(cx, (n,)))
            cy = capi.as_f64(np.broadcast_to(cy, (n,)))
        amp = self._amplitude(orb_)
        amp = None if amp is None else capi.as_f64(np.broadcast_to(amp, (n,)))
        y, z = capi.as_f64(y), capi.as_f64(z)
        ctx.check(ctx.lib.pb200_number_density_chords(
            ctx.handle, 0, n, capi.dptr(y), capi.dptr(z),
            capi.dptr(cx) if cx is not None else None, capi.dptr(cy) if cy is not None else None,
            capi.dptr(amp) if amp is not None else None, capi.dptr(out)))
        return out[0] if scalar else out


class CollisionalAtmosphere(_Scenario):
    def __init__(self, T, P_0):
        self.T, self.P_0 = T, P_0
        self.constituents = []
        self.hasMoon = False

    def getReferenceNumberDensity(self):
        return self.P_0 / (const.k_B * self.T)

    def getVelDispersion(self, m):
        return np.sqrt(self.T * const.k_B / m)

    def addConstituent(self, speciesName, chi):
        species = host.AvailableSpecies().findSpecies(speciesName)
        self.constituents.append(AtmosphericConstituent(species, chi, self.getVelDispersion(species.mass)))

    def addMolecularConstituent(self, speciesName, chi):
        self.constituents.append(MolecularConstituent(speciesName, chi))


class BarometricAtmosphere(CollisionalAtmosphere):
    def __init__(self, T, P_0, mu, planet):
        super().__init__(T, P_0)
        self.mu, self.planet = mu, planet


class HydrostaticAtmosphere(CollisionalAtmosphere):
    def __init__(self, T, P_0, mu, planet):
        super().__init__(T, P_0)
        self.mu, self.planet = mu, planet


class PowerLawAtmosphere(CollisionalAtmosphere):
    def __init__(self, T, P_0, q, planet):
        super().__init__(T, P_0)
        self.q, self.planet = q, planet


class EvaporativeExosphere(_Scenario):
    def __init__(self, N):
        self.N = N
        self.hasMoon = False

    def addConstituent(self, speciesName, sigma_v):
        species = host.AvailableSpecies().findSpecies(speciesName)
        self.constituents = [AtmosphericConstituent(species, 1.0, sigma_v)]

    def addMolecularConstituent(self, speciesName, T):
        self.constituents = [MolecularConstituent(speciesName, 1.0)]
        self.T = T


class PowerLawExosphere(EvaporativeExosphere):
    def __init__(self, N, q, planet):
        super().__init__(N)
        self.q, self.planet = q, planet


class MoonExosphere(EvaporativeExosphere):
    def __init__(self, N, q, moon):
        super().__init__(N)
        self.q, self.moon = q, moon
        self.hasMoon = True
        self.planet = moon.hostPlanet


class TidallyHeatedMoon(EvaporativeExosphere):
    def __init__(self, q, moon):
        self.q, self.moon = q, moon
        self.hasMoon = True
        self.planet = moon.hostPlanet

    def addSourceRateFunction(self, filename, tau_photoionization, mass_absorber):
        """Half-orbit mass-loss curve mirrored to a full orbit; log10 N(phase) table (gasProperties.py:404-423)."""
        half = np.loadtxt(filename)
        full = np.concatenate((half, half[::-1]))
        phase = np.linspace(0.0, 2.0 * np.pi, len(full))
        self.N_function = host.TableFunction1D(phase, np.log10(full * tau_photoionization / mass_absorber))

    def calculateAbsorberNumber(self, orbphase):
        """N(orbphase) = 10**interp(moon phase mod 2 pi) -- O(n_orb) host set-up (gasProperties.py:425-436)."""
        ph = self.moon.getOrbphase(orbphase) % (2.0 * np.pi)
        return 10 ** np.interp(ph, self.N_function.x, self.N_function.y)

    def _amplitude(self, orbphase):
        return self.calculateAbsorberNumber(orbphase)


class TorusExosphere(EvaporativeExosphere):
    def __init__(self, N, a_torus, v_ej, planet):
        super().__init__(N)
        self.a_torus, self.v_ej, self.planet = a_torus, v_ej, planet


class SerpensExosphere(EvaporativeExosphere):
    def __init__(self, filename, N, planet, sigmaSmoothing):
        super().__init__(N)
        self.filename, self.planet, self.sigmaSmoothing = filename, planet, sigmaSmoothing

    def addInterpolatedDensity(self, spatialGrid):
        """Particle file -> tabulated density on the (x, y, z) cell midpoints (gasProperties.py:548-583).

        Host side: reading the file (positions in metres, first three columns) and the O(axis) bin edges -- x as the
        grid's own cell borders, y and z as 2 rho_steps equal cells across [-rho_border, +rho_border].  Device side
        (pb200_serpens_density): the histogram over the particles, its scaling to a number density, the optional
        Gaussian smoothing and the two diagnostic totals, which are reported in particle numbers as the reference does."""
        positions_cm = np.ascontiguousarray(np.loadtxt(self.filename)[:, 0:3] * 1e2)
        half_width, n_cells = spatialGrid.rho_border, 2 * int(spatialGrid.rho_steps)
        edges_x = spatialGrid.constructXaxis(midpoints=False)
        edges_y = np.linspace(-half_width, half_width, n_cells + 1)
        edges_z = np.linspace(-half_width, half_width, n_cells + 1)
        cell_volume = (edges_x[1] - edges_x[0]) * (edges_y[1] - edges_y[0]) * (edges_z[1] - edges_z[0])
        mid_x = spatialGrid.constructXaxis()
        # midpoints written as the reference writes them (left edges + a quarter of border/steps times two): the table axes
        # must be bit-identical to the reference's for the trilinear lookup to see the same cell fractions
        shift = 2. * half_width / (4. * spatialGrid.rho_steps)
        mid_y = np.linspace(-half_width, half_width, n_cells, endpoint=False) + shift
        mid_z = np.linspace(-half_width, half_width, n_cells, endpoint=False) + shift
        smoothing = gaussian_weights(self.sigmaSmoothing) if self.sigmaSmoothing > 0. else None
        table, totals = host.get_context().serpens_density(
            positions_cm, edges_x, edges_y, edges_z, self.N, positions_cm.shape[0] * cell_volume, smoothing,
            mid_y, mid_z, self.planet.R, self.planet.hostStar.R)
        print("SERPENS table: particles represented in all cells (after smoothing, if any):", totals[0] * cell_volume)
        print("SERPENS table: of those, between the planetary and the stellar limb:", totals[1] * cell_volume)
        self.InterpolatedDensity = host.GridFunction((mid_x, mid_y, mid_z), table)


def gaussian_weights(sigma, truncate=4.0):
    """The 1-D kernel of scipy.ndimage.gaussian_filter(order=0) -- scipy's _gaussian_kernel1d expression, O(radius)."""
    sd = float(sigma)
    radius = int(truncate * sd + 0.5)
    x = np.arange(-radius, radius + 1)
    phi_x = np.exp(-0.5 / (sd * sd) * x ** 2)
    return phi_x / phi_x.sum()


def _carrier_scene(atmosphere, x, wavelength, delta_x):
    """Scene with the atmosphere's scenarios, constituents and tables, the caller's x axis and wavelengths and dummy
    one-cell sky axes: the carrier of pb200_optical_depth_chords."""
    class _G:
        def constructXaxis(self): return x
        def constructRhoAxis(self): return np.array([1.0])
        def constructPhiAxis(self): return np.array([0.0])
        def constructOrbphaseAxis(self): return np.array([0.0])
        def getDeltaX(self): return delta_x

    class _T:
        pass

    t = _T()
    t.spatialGrid = _G()
    t.atmosphere = atmosphere
    t.planet = atmosphere.densityDistributionList[0].planet
    t.wavelength = wavelength
    sc, keep, _ = _scene.flatten(t)
    return sc, keep


def _mini_scene(dist, x):
    """Scene with one scenario and dummy 1-cell axes: carrier for pb200_number_density_chords."""
    class _G:
        def constructXaxis(self): return x
        def constructRhoAxis(self): return np.array([1.0])
        def constructPhiAxis(self): return np.array([0.0])
        def constructOrbphaseAxis(self): return np.array([0.0])
        def getDeltaX(self): return 1.0

    class _T:
        pass

    shadow = copy.copy(dist)
    shadow.constituents = []
    t = _T()
    t.spatialGrid = _G()
    t.atmosphere = Atmosphere([shadow], False)
    t.planet = dist.planet
    t.wavelength = np.array([1.0])
    sc, keep, _ = _scene.flatten(t)
    return sc, keep


# ------------------------------------------------------------------------------------------------
# constituents (gasProperties.py:609-818)
# ------------------------------------------------------------------------------------------------
class AtmosphericConstituent:
    def __init__(self, species, chi, sigma_v):
        self.isMolecule = False
        self.species, self.chi, self.sigma_v = species, chi, sigma_v
        self.wavelengthGridRefinement = 10.0
        self.wavelengthGridExtension = 0.01
        self.lookupOffset = const.LOOKUP_OFFSET

    def getLineParameters(self, wavelength):
        """(lambda0 [cm], gamma = A/(4 pi), f) of this species' lines strictly inside the range
        (gasProperties.py:640-670)."""
        lam0, gamma, f = host.species_lines(self.species.element, self.species.ionizationState)
        wavelength = np.asarray(wavelength)
        inside = (lam0 > wavelength.min()) & (lam0 < wavelength.max())   # the reference uses builtin min/max (0.13 s each at C5)
        return lam0[inside], gamma[inside], f[inside]

    def calculateVoigtProfile(self, wavelength):
        """sigma(lambda) = sum over lines of Voigt profiles: kernel (b2) (gasProperties.py:672-692)."""
        wavelength = np.asarray(wavelength, dtype=np.float64)
        lam0, gamma, f = self.getLineParameters(wavelength)
        return host.get_context().voigt_sigma(lam0, gamma, f, self.sigma_v, wavelength, self.lookupOffset)

    def constructLookupFunction(self, wavelengthGrid):
        """log10(sigma + offset) on the 10x refined, 1 % extended grid of this species
        (gasProperties.py:694-715); sigma and its log10 come from one launch of kernel (b2)."""
        fine = copy.deepcopy(wavelengthGrid)
        fine.resolutionHigh /= self.wavelengthGridRefinement
        fine.lower_w *= (1.0 - self.wavelengthGridExtension)
        fine.upper_w *= (1.0 + self.wavelengthGridExtension)
        grid = fine.constructWavelengthGridSingle(self)
        lam = capi.pinned_pool.empty(grid.shape)     # page-locked like log_sigma: uploaded without a staging copy
        lam[...] = grid
        lam0, gamma, f = self.getLineParameters(lam)
        _, log_sigma = host.get_context().voigt_sigma(lam0, gamma, f, self.sigma_v, lam, self.lookupOffset, want_log=True)
        return host.TableFunction1D(lam, log_sigma, fill_value=np.log10(self.lookupOffset))

    def addLookupFunctionToConstituent(self, wavelengthGrid):
        self.lookupFunction = self.constructLookupFunction(wavelengthGrid)

    def getSigmaAbs(self, wavelength):
        """10**interp(lambda) - offset on the GPU: kernel (b1) (gasProperties.py:727-735, :34-51)."""
        lf = self.lookupFunction
        return host.get_context().sigma_lookup(lf.x, lf.y, self.lookupOffset, wavelength)


class MolecularConstituent:
    def __init__(self, moleculeName, chi):
        self.isMolecule = True
        self.lookupOffset = const.LOOKUP_OFFSET
        self.moleculeName, self.chi = moleculeName, chi

    def setOpacityArrays(self, p_bar, t, bin_edges, xsecarr):
        """The four datasets of an ExoMol-style opacity file as arrays ('p' [bar], 't' [K], 'bin_edges' [1/cm, ascending],
        'xsecarr' [n_p][n_t][n_bins], cm^2): the file-format-free door of constructLookupFunction."""
        self.opacityArrays = tuple(np.asarray(a, dtype=np.float64) for a in (p_bar, t, bin_edges, xsecarr))

    def constructLookupFunction(self):
        """(P [dyn/cm^2], T, lambda [cm], ascending) grid and log10(sigma + offset) with fill value log10(offset)
        (gasProperties.py:765-781).  The arrays come from setOpacityArrays or, where h5py exists, from
        <molecularResources>/<moleculeName>.h5; the reversal of the wavenumber axis and the log10 run on the GPU."""
        arrays = getattr(self, "opacityArrays", None)
        if arrays is None:
            try:
                import h5py
                if not hasattr(h5py, "File"):      # an empty stand-in module (tests put one on the path for the reference)
                    raise ImportError("h5py stand-in")
            except ImportError:
                raise NotImplementedError("h5py is not installed: attach the datasets with setOpacityArrays(p, t, bin_edges, "
                                          "xsecarr) (gasProperties.py:774-778 reads exactly these four)") from None
            with h5py.File(getattr(self, "opacityPath", self.moleculeName + ".h5"), "r") as f:
                arrays = (f["p"][:], f["t"][:], f["bin_edges"][:], f["xsecarr"][:])
        p_bar, t, bin_edges, xsec = arrays
        log_sigma = host.get_context().log_table(xsec, self.lookupOffset, reverse_last_axis=True)
        return host.GridFunction((p_bar * 10., t, 1. / bin_edges[::-1]), log_sigma, np.log10(self.lookupOffset))

    def addLookupFunctionToConstituent(self):
        self.lookupFunction = self.constructLookupFunction()

    def getSigmaAbs(self, P, T, wavelength):
        """sigma(P(c,x), T, lambda(c,w)) -> (n_chords, n_x, n_wav) on the GPU (gasProperties.py:789-818)."""
        lf = self.lookupFunction
        return host.get_context().molecular_sigma(lf.grid, lf.values, lf.fill_value, self.lookupOffset, P, T, wavelength)


# ------------------------------------------------------------------------------------------------
# wavelength grid (gasProperties.py:959-1071): host-side, O(n_wav), runs once
# ------------------------------------------------------------------------------------------------
class WavelengthGrid:
    def __init__(self, lower_w, upper_w, widthHighRes, resolutionLow, resolutionHigh):
        self.lower_w, self.upper_w, self.widthHighRes = lower_w, upper_w, widthHighRes
        self.resolutionLow, self.resolutionHigh = resolutionLow, resolutionHigh

    def arangeWavelengthGrid(self, linesList):
        """Fine steps inside windows of +-widthHighRes/2 around (clusters of) line centres, coarse
        steps elsewhere; lines closer than widthHighRes share one window (gasProperties.py:990-1035)."""
        centres = np.sort(np.unique(linesList))
        if centres.size == 0:
            print("WARNING: No absorption lines from atoms/ions in the specified wavelength range!")
            return np.arange(self.lower_w, self.upper_w, self.resolutionLow)
        half = self.widthHighRes / 2.0
        gap = np.concatenate(([np.inf], np.diff(centres), [np.inf]))
        opens = [c - half for c, g in zip(centres, gap[:-1]) if g > self.widthHighRes]
        closes = [c + half for c, g in zip(centres, gap[1:]) if g > self.widthHighRes]
        pieces = []
        if self.lower_w < opens[0]:
            pieces.append(np.arange(self.lower_w, opens[0], self.resolutionLow))
        for k, (lo, hi) in enumerate(zip(opens, closes)):
            if k > 0:
                pieces.append(np.arange(closes[k - 1], lo, self.resolutionLow))
            pieces.append(np.arange(lo, hi, self.resolutionHigh))
        if self.upper_w > closes[-1]:
            pieces.append(np.arange(closes[-1], self.upper_w, self.resolutionLow))
        return np.sort(np.concatenate(pieces))

    def constructWavelengthGridSingle(self, constituent):
        return self.arangeWavelengthGrid(constituent.getLineParameters(np.array([self.lower_w, self.upper_w]))[0])

    def constructWavelengthGrid(self, densityDistributionList):
        centres = []
        for dist in densityDistributionList:
            for con in dist.constituents:
                if not con.isMolecule:
                    centres.extend(con.getLineParameters(np.array([self.lower_w, self.upper_w]))[0])
        if len(centres) == 0:
            return np.arange(self.lower_w, self.upper_w, self.resolutionLow)
        return self.arangeWavelengthGrid(centres)


# ------------------------------------------------------------------------------------------------
# Atmosphere and Transit (gasProperties.py:821-956, :1074-1258)
# ------------------------------------------------------------------------------------------------
class Atmosphere:
    def __init__(self, densityDistributionList, hasOrbitalDopplerShift):
        self.densityDistributionList = densityDistributionList
        self.hasOrbitalDopplerShift = hasOrbitalDopplerShift

    def getLOSopticalDepth_Batch(self, x_grid, phi_batch, rho_batch, orbphase_batch, wavelength, delta_x):
        """tau (n_chords, n_wav) of caller-supplied chords on the GPU -- same signature as the reference
        (gasProperties.py:885-956).  Host work is O(n_chords): the bodies' positions, line-of-sight velocities and
        amplitudes at each chord's orbital phase, through the objects' own methods (:908-916)."""
        x_grid = np.asarray(x_grid, dtype=np.float64)
        wavelength = np.asarray(wavelength, dtype=np.float64)
        phi, rho, orb = np.broadcast_arrays(np.atleast_1d(np.asarray(phi_batch, dtype=np.float64)),
                                            np.atleast_1d(np.asarray(rho_batch, dtype=np.float64)),
                                            np.atleast_1d(np.asarray(orbphase_batch, dtype=np.float64)))
        n = phi.size
        dists = list(self.densityDistributionList)
        ctx = host.get_context()
        sc, keep = _carrier_scene(self, x_grid, wavelength, float(delta_x))
        ctx.set_scene(sc, keep)
        y, z = host.Grid.getCartesianFromCylinder(phi, rho)
        cx, cy = np.zeros((len(dists), n)), np.zeros((len(dists), n))
        shift, amp = np.ones((len(dists), n)), None
        for i, dist in enumerate(dists):
            body = dist.moon if getattr(dist, "hasMoon", False) else dist.planet
            if _scene._kind(dist) != "SerpensExosphere":
                bx, by = body.getPosition(orb)
                cx[i], cy[i] = bx, by
            if self.hasOrbitalDopplerShift:
                shift[i] = const.calculateDopplerShift(-np.asarray(body.getLOSvelocity(orb), dtype=np.float64))
            a = dist._amplitude(orb) if hasattr(dist, "_amplitude") else None
            if a is not None:
                amp = np.ones((len(dists), n)) if amp is None else amp
                amp[i] = a
        out = np.empty((n, wavelength.size), dtype=np.float64)
        y, z = capi.as_f64(y), capi.as_f64(z)
        ctx.check(ctx.lib.pb200_optical_depth_chords(ctx.handle, n, capi.dptr(y), capi.dptr(z), capi.dptr(cx), capi.dptr(cy),
                                                     capi.dptr(shift), capi.dptr(amp) if amp is not None else None,
                                                     0, wavelength.size, capi.dptr(out)))
        return out


class Transit:
    def __init__(self, atmosphere, wavelengthGrid, spatialGrid):
        self.atmosphere, self.wavelengthGrid, self.spatialGrid = atmosphere, wavelengthGrid, spatialGrid
        self.planet = atmosphere.densityDistributionList[0].planet
        self.last_timing = None

    def addWavelength(self):
        self.wavelength = self.wavelengthGrid.constructWavelengthGrid(self.atmosphere.densityDistributionList)

    def _upload(self, ctx):
        ctx.prefetch_tables()      # the sigma tables of the previous call start moving while the host flattens this one
        sc, keep, info = _scene.flatten(self)
        ctx.set_scene(sc, keep)
        return info

    def sumOverChords(self, max_memory_gb: float = 2.0):
        """R(phase, lambda), float64 (n_orb, n_wav), computed by the CUDA library.

        `max_memory_gb` is accepted for signature compatibility and ignored: chord x lambda is never
        materialised, so there is nothing to batch (replaces memoryHandler.py:55-66).
        """
        ctx = host.get_context()
        info = self._upload(ctx)
        R, timing = ctx.run(0, info["n_orb"], 0, info["n_wav"])
        timing.update(h2d_bytes=info["h2d_bytes"], d2h_bytes=R.nbytes)
        self.last_timing = timing
        return R

    # ---- materialised inner seams on the chord grid (tests, bandwidth probes) ---------------------
    def getLOSopticalDepthGrid(self, wav_lo=0, wav_hi=None):
        """tau(chord, lambda) of every chord of the grid in reference chord order (the batch method
        Atmosphere.getLOSopticalDepth_Batch, gasProperties.py:885-956, applied to the whole grid)."""
        ctx = host.get_context()
        info = self._upload(ctx)
        wav_hi = info["n_wav"] if wav_hi is None else wav_hi
        out = np.empty((info["n_chords"], wav_hi - wav_lo), dtype=np.float64)
        ctx.check(ctx.lib.pb200_optical_depth(ctx.handle, wav_lo, wav_hi, capi.dptr(out)))
        return out

    def getColumnDensityGrid(self):
        ctx = host.get_context()
        info = self._upload(ctx)
        out = np.empty((max(info["n_atoms"], 1), info["n_chords"]), dtype=np.float64)
        ctx.check(ctx.lib.pb200_column_density(ctx.handle, capi.dptr(out)))
        return out[: info["n_atoms"]]

    def getNumberDensityGrid(self, scenario: int):
        ctx = host.get_context()
        info = self._upload(ctx)
        out = np.empty((info["n_chords"], info["n_x"]), dtype=np.float64)
        ctx.check(ctx.lib.pb200_number_density(ctx.handle, scenario, capi.dptr(out)))
        return out

    def getChordMasks(self):
        import ctypes as C
        ctx = host.get_context()
        info = self._upload(ctx)
        blocked = np.empty(info["n_chords"], dtype=np.uint8)
        weight = np.empty(info["n_chords"], dtype=np.float64)
        ctx.check(ctx.lib.pb200_chord_masks(ctx.handle, blocked.ctypes.data_as(C.POINTER(C.c_uint8)), capi.dptr(weight)))
        return blocked.astype(bool), weight
