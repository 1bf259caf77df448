"""Batched forward models for retrieval -- SURVEY 8(f) rank 1, the consumer sketched by the reference's
mainRetrieval.py:18-93 (free parameters -> forward model -> band-averaged light curve).

    batch = RetrievalBatch(template_transit)                # tables, axes, wavelength grid -> HBM, once
    band = LightcurveBand.for_planet(planet, orbphase, [5891.583253e-8, 5897.558147e-8], 0.75e-8)
    lightcurves, R, timing = batch.run(transits, band)      # (n_models, n_orb) float64

Each element of `transits` is a Transit-like object of the SAME setup (grid, wavelength grid, constituents and
their sigma tables) whose bodies / scenario scalars differ.  Only O(n_orb) numbers per model cross PCIe in
either direction; the models are stacked along the phase axis and run as one set of launches
(pb200_run_batch).  All compute is in libprometheus_b200.so; nothing here has a CPU fallback.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import capi, host
from . import constants as const
from . import scene as _scene


class LightcurveBand:
    """The filter of mainRetrieval.py:72-93: wavelengths within +-bandwidth/2 of any Doppler-shifted line centre."""

    def __init__(self, line_centres, bandwidth, shift):
        self.line_centres = capi.as_f64(np.atleast_1d(line_centres))
        self.shift = capi.as_f64(np.atleast_1d(shift))
        self.bandwidth = float(bandwidth)
        if not 1 <= self.line_centres.size <= 16:
            raise ValueError("1..16 line centres")

    @classmethod
    def for_planet(cls, planet, orbphase, line_centres, bandwidth):
        """shift = calculateDopplerShift(planet.getLOSvelocity(orbphase)), mainRetrieval.py:77-78."""
        v_los = np.asarray(planet.getLOSvelocity(np.asarray(orbphase, dtype=np.float64)), dtype=np.float64)
        return cls(line_centres, bandwidth, const.calculateDopplerShift(v_los))

    def struct(self) -> capi.Band:
        b = capi.Band()
        b.n_lines = self.line_centres.size
        b.line_centre, b.shift, b.bandwidth = capi.dptr(self.line_centres), capi.dptr(self.shift), self.bandwidth
        return b


class RetrievalBatch:
    def __init__(self, template, device: int | None = None):
        self.ctx = host.get_context(device)
        sc, keep, info = _scene.flatten(template)
        self.ctx.set_scene(sc, keep)
        self.info = info
        self._phase = np.asarray(template.spatialGrid.constructOrbphaseAxis(), dtype=np.float64)
        self.last_timing = None

    def _models(self, transits):
        arr = (capi.Model * len(transits))()
        keep = []
        for i, tr in enumerate(transits):
            mdl, k = _scene.flatten_model(tr, self._phase)
            arr[i] = mdl
            keep.append(k)
        return arr, keep

    def set_models(self, transits):
        """Stack the models on the resident scene; the context then has len(transits)*n_orb phases and every
        entry point (pb200_run, pb200_column_density, ...) addresses row m*n_orb + p."""
        arr, keep = self._models(transits)
        self.ctx.set_models(arr, len(transits), keep)

    def reset(self):
        self.ctx.set_models(None, 0)

    def run_family(self, family, n_models: int, band: LightcurveBand | None = None, want_R: bool = False):
        """As run(), for a broadcast family (make_family): one vectorised flattening for the whole batch."""
        models, keep = _scene.flatten_models_broadcast(family, self._phase, n_models)
        keep.append(models)
        lc, R, timing = self.ctx.run_batch(models.ctypes.data_as(C.POINTER(capi.Model)), n_models,
                                           band.struct() if band is not None else None,
                                           self.info["n_orb"], self.info["n_wav"], want_R, keep)
        self.last_timing = timing
        return lc, R, timing

    def run(self, transits, band: LightcurveBand | None = None, want_R: bool = False):
        """-> (lightcurves (n_models, n_orb) or None, R (n_models, n_orb, n_wav) or None, timing dict)."""
        if band is None and not want_R:
            raise ValueError("nothing to return: give a band and/or want_R=True")
        arr, keep = self._models(transits)
        if band is not None and band.shift.size != self.info["n_orb"]:
            raise ValueError("band.shift must have one entry per orbital phase")
        lc, R, timing = self.ctx.run_batch(arr, len(transits), band.struct() if band is not None else None,
                                           self.info["n_orb"], self.info["n_wav"], want_R, keep)
        self.last_timing = timing
        return lc, R, timing


def extract_lightcurve(transit_or_ctx, band: LightcurveBand, phase_lo: int = 0, phase_hi: int | None = None):
    """Light curve of the R left on the device by the last sumOverChords / run of this context
    (pb200_lightcurve with R_dev == NULL)."""
    ctx = transit_or_ctx if isinstance(transit_or_ctx, capi.Context) else host.get_context()
    if phase_hi is None:
        phase_hi = band.shift.size
    return ctx.lightcurve(band.struct(), phase_lo, phase_hi)


def make_variants(template, params):
    """Transit-like objects for parameter vectors (midTransitOrbphase [rad], a_moon [planet radii], R_moon [Io radii])
    -- the free parameters of mainRetrieval.py:18-21 -- sharing the template's grid, wavelength grid and
    constituents (hence its sigma tables).  Works on the reference's objects and on the host mirror alike: the
    moon is rebuilt with the template's own Moon class."""
    import copy
    out = []
    dists = list(template.atmosphere.densityDistributionList)
    planet = getattr(template, "planet", None) or dists[0].planet
    for phi0, a_moon, R_moon in np.asarray(params, dtype=np.float64).reshape(-1, 3):
        new_dists = []
        for d in dists:
            d2 = copy.copy(d)
            if getattr(d, "hasMoon", False):
                d2.moon = type(d.moon)(float(phi0), float(R_moon) * const.R_Io, float(a_moon) * planet.R, planet)
            new_dists.append(d2)
        tr = copy.copy(template)
        tr.atmosphere = copy.copy(template.atmosphere)
        tr.atmosphere.densityDistributionList = new_dists
        out.append(tr)
    return out


def make_family(template, params):
    """ONE Transit-like object standing for all parameter vectors: its moon carries (n_models, 1) arrays, so the
    objects' own expressions broadcast over models and phases (scene.flatten_models_broadcast)."""
    import copy
    P = np.asarray(params, dtype=np.float64).reshape(-1, 3)
    dists = list(template.atmosphere.densityDistributionList)
    planet = getattr(template, "planet", None) or dists[0].planet
    new_dists = []
    for d in dists:
        d2 = copy.copy(d)
        if getattr(d, "hasMoon", False):
            d2.moon = type(d.moon)(P[:, 0:1], P[:, 2:3] * const.R_Io, P[:, 1:2] * planet.R, planet)
        new_dists.append(d2)
    tr = copy.copy(template)
    tr.atmosphere = copy.copy(template.atmosphere)
    tr.atmosphere.densityDistributionList = new_dists
    return tr, P.shape[0]


def model_partition(n_models: int, world: int) -> list[tuple[int, int]]:
    """Contiguous [lo, hi) ranges of the models of a batch over `world` ranks (sizes differ by at most one)."""
    base, extra = divmod(int(n_models), int(world))
    out, lo = [], 0
    for r in range(world):
        hi = lo + base + (1 if r < extra else 0)
        out.append((lo, hi))
        lo = hi
    return out


def run_sharded(transits, band, n_orb: int, compute=None, group=None, template=None, device=None):
    """Data-parallel retrieval batch: forward models are independent, so rank r runs models
    model_partition(n, world)[r] on its GPU and ONE all-gather of the light curves (n_models*n_orb doubles)
    puts the result on every rank.  `compute(transits_of_rank) -> (k, n_orb) array` defaults to
    RetrievalBatch(template).run on this rank's GPU; CPU tests inject a host function."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    parts = model_partition(len(transits), world)
    lo, hi = parts[rank]
    if compute is None:
        batch = RetrievalBatch(template, device)
        compute = lambda trs: batch.run(trs, band)[0]   # noqa: E731
    width = max(h - l for l, h in parts)
    local = torch.zeros((max(width, 1), n_orb), dtype=torch.float64)
    if hi > lo:
        local[: hi - lo] = torch.from_numpy(np.ascontiguousarray(compute(transits[lo:hi])))
    if world == 1:
        return local[: hi - lo].numpy()
    use_cuda = dist.get_backend(group) == "nccl"
    buf = local.cuda() if use_cuda else local
    gathered = torch.empty(world * buf.numel(), dtype=torch.float64, device=buf.device)
    dist.all_gather_into_tensor(gathered, buf.reshape(-1), group=group)
    g = gathered.cpu().numpy().reshape((world,) + tuple(buf.shape))
    return np.concatenate([g[r, : h - l] for r, (l, h) in enumerate(parts)], axis=0)
