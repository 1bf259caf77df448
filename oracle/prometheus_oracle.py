"""CPU oracle for the PROMETHEUS transit hot path -- TEST INFRASTRUCTURE, NOT PRODUCT.

A numpy/scipy restatement of the reference algorithm (Transit.sumOverChords and everything it
calls), written from the reference's behaviour, each function citing the reference lines it follows
(paths relative to /root/reference).  Only tests/, __graft_entry__.smoke() and the cpu_baseline /
--impl reference legs of bench.py may import this module; prometheus_b200 never does.

Parity pin: the reference ships no tests or golden vectors, so this oracle is pinned
against outputs of the reference itself, generated in the build container by
tests/golden/make_golden.py and committed as tests/golden/*.npz; tests/test_oracle_golden.py checks
the oracle against every one of them (R, wavelength grids, sigma tables, column densities, masks).
Third-party arithmetic the reference delegates to (scipy.special.voigt_profile / erf,
scipy RegularGridInterpolator, np.interp) is called here exactly as the reference calls it; no
versions are pinned by the reference, the de-facto pin is this image (scipy 1.18 / numpy 2.3).
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field

import numpy as np
from scipy.interpolate import RegularGridInterpolator
from scipy.special import erf, voigt_profile

# constants.py:13-18 (values as in the reference)
E_CHARGE, M_E, C_LIGHT = 4.803e-10, 9.109e-28, 2.998e10
G_GRAV, K_B, AMU = 6.674 * 10 ** (-8), 1.381 * 10 ** (-16), 1.661 * 10 ** (-24)
OFFSET = 1e-50

SPECIES = {  # constants.py:139-162 (+ Fe II)
    "NaI": ("Na", "1", 22.99), "KI": ("K", "1", 39.0983), "SiI": ("Si", "1", 28.0855), "SiII": ("Si", "2", 28.0855),
    "SiIII": ("Si", "3", 28.0855), "SiIV": ("Si", "4", 28.0855), "MgI": ("Mg", "1", 24.305), "MgII": ("Mg", "2", 24.305),
    "AlI": ("Al", "1", 26.9815), "CaI": ("Ca", "1", 40.078), "CaII": ("Ca", "2", 40.078), "TiI": ("Ti", "1", 47.867),
    "TiII": ("Ti", "2", 47.867), "CrI": ("Cr", "1", 51.9961), "MnI": ("Mn", "1", 54.938), "FeI": ("Fe", "1", 55.845),
    "CoI": ("Co", "1", 58.933), "NiI": ("Ni", "1", 58.6934), "OI": ("O", "1", 15.999), "CII": ("C", "2", 12.011),
    "SIII": ("S", "3", 32.06), "SIV": ("S", "4", 32.06), "FeII": ("Fe", "2", 55.845),
}


def doppler_shift(v):
    """constants.py:31-45"""
    beta = v / C_LIGHT
    return np.sqrt((1.0 - beta) / (1.0 + beta))


# ---------------------------------------------------------------------------------------------------
# geometry -- geometryHandler.py:88-94, :112-208
# ---------------------------------------------------------------------------------------------------
def grid_axes(grid):
    x_mid, x_b, n_x, rho_b, n_rho, n_phi, orb_b, n_orb = grid
    n_x, n_rho, n_phi, n_orb = int(n_x), int(n_rho), int(n_phi), int(n_orb)
    x = np.linspace(x_mid - x_b, x_mid + x_b, n_x, endpoint=False) + x_b / float(n_x)
    rho = np.linspace(0.0, rho_b, n_rho, endpoint=False) + 0.5 * rho_b / float(n_rho)
    phi = np.linspace(0, 2 * np.pi, n_phi, endpoint=False) + np.pi / float(n_phi)
    orb = np.linspace(-orb_b, orb_b, n_orb)
    return x, rho, phi, orb, 2.0 * x_b / float(n_x)


def chord_grid(rho, phi, orb):
    """geometryHandler.py:188-208: meshgrid(phi, rho, orb, 'ij') flattened -> (phi, rho, orb) columns."""
    P, R, O = np.meshgrid(phi, rho, orb, indexing="ij")
    return P.ravel(), R.ravel(), O.ravel()


# ---------------------------------------------------------------------------------------------------
# bodies -- celestialBodies.py:370-397 (planet), :499-550 (moon)
# ---------------------------------------------------------------------------------------------------
def planet_position(pl, orb):
    return pl["a"] * np.cos(orb), pl["a"] * np.sin(orb)


def planet_vlos(pl, star, orb):
    return -np.sin(orb) * np.sqrt(G_GRAV * star["M"] / pl["a"])


def moon_phase(mo, pl, star, orb):
    a_p, M_p, a_m, M_s = np.float64(pl["a"]), np.float64(pl["M"]), np.float64(mo["a"]), np.float64(star["M"])
    return mo["phi0"] + np.float64(orb) * np.sqrt((a_p ** 3 * M_p) / (a_m ** 3 * M_s))


def moon_position(mo, pl, star, orb):
    ph = moon_phase(mo, pl, star, orb)
    x_p, y_p = planet_position(pl, orb)
    return x_p + mo["a"] * np.cos(ph), y_p + mo["a"] * np.sin(ph)


def moon_vlos(mo, pl, star, orb):
    return planet_vlos(pl, star, orb) - np.sin(moon_phase(mo, pl, star, orb)) * np.sqrt(G_GRAV * pl["M"] / mo["a"])


# ---------------------------------------------------------------------------------------------------
# lines, wavelength grid, Voigt table -- gasProperties.py:640-715, :990-1071
# ---------------------------------------------------------------------------------------------------
def line_parameters(rows, element, ion, wl_range):
    """gasProperties.py:640-670"""
    sel = (rows[:, 0] == element) & (rows[:, 1] == ion) & (rows[:, 2] != "") & (rows[:, 3] != "") & (rows[:, 4] != "")
    lam0 = rows[:, 2][sel].astype(float) * 1e-8
    gamma = rows[:, 3][sel].astype(float) / (4.0 * np.pi)
    f = rows[:, 4][sel].astype(float)
    keep = (lam0 > min(wl_range)) & (lam0 < max(wl_range))
    return lam0[keep], gamma[keep], f[keep]


def arange_wavelength_grid(lower_w, upper_w, width, res_low, res_high, lines):
    """gasProperties.py:990-1035"""
    peaks = np.sort(np.unique(lines))
    if len(peaks) == 0:
        return np.arange(lower_w, upper_w, res_low)
    diff = np.concatenate(([np.inf], np.diff(peaks), [np.inf]))
    left = [p - width / 2.0 for i, p in enumerate(peaks) if diff[i] > width]
    right = [p + width / 2.0 for i, p in enumerate(peaks) if diff[i + 1] > width]
    parts = []
    for i in range(len(left)):
        parts.append(np.arange(left[i], right[i], res_high))
        if i == 0:
            if lower_w < left[0]:
                parts.append(np.arange(lower_w, left[0], res_low))
            if len(left) == 1 and upper_w > right[-1]:
                parts.append(np.arange(right[0], upper_w, res_low))
        else:
            parts.append(np.arange(right[i - 1], left[i], res_low))
            if i == len(left) - 1 and upper_w > right[-1]:
                parts.append(np.arange(right[-1], upper_w, res_low))
    return np.sort(np.concatenate(parts))


def voigt_sigma(lam0, gamma, f, sigma_v, wl):
    """gasProperties.py:672-692"""
    out = np.zeros_like(wl)
    for i in range(len(lam0)):
        prof = voigt_profile(C_LIGHT / wl - C_LIGHT / lam0[i], sigma_v / lam0[i], gamma[i])
        out += np.pi * (E_CHARGE) ** 2 / (M_E * C_LIGHT) * f[i] * prof
    return out


def lookup_table(rows, element, ion, sigma_v, wgrid):
    """gasProperties.py:694-715: (refined wavelengths, log10(sigma + 1e-50))"""
    lower_w, upper_w, width, res_low, res_high = wgrid
    lo, hi, rh = lower_w * (1.0 - 0.01), upper_w * (1.0 + 0.01), res_high / 10.0
    own = line_parameters(rows, element, ion, np.array([lo, hi]))[0]
    fine = arange_wavelength_grid(lo, hi, width, res_low, rh, own)
    lam0, gamma, f = line_parameters(rows, element, ion, fine)
    return fine, np.log10(voigt_sigma(lam0, gamma, f, sigma_v, fine) + OFFSET)


def interp_log(x, xp, fp, offset):
    """gasProperties.py:34-51 (n_interp_log): 10**np.interp(x, xp, fp) - offset"""
    return 10 ** np.interp(x, xp, fp) - offset


# ---------------------------------------------------------------------------------------------------
# model assembly
# ---------------------------------------------------------------------------------------------------
@dataclass
class Constituent:
    kind: str                 # "atom" | "molecule"
    chi: float
    name: str = ""
    tab_x: np.ndarray | None = None
    tab_y: np.ndarray | None = None
    rgi: RegularGridInterpolator | None = None


@dataclass
class Scenario:
    type: str
    params: dict
    has_moon: bool = False
    T: float = 0.0
    constituents: list = field(default_factory=list)
    rgi: RegularGridInterpolator | None = None        # serpens
    n_table: tuple | None = None                      # tidally heated: (phase, log10 N)


@dataclass
class Model:
    spec: dict
    star: dict
    planet: dict
    moon: dict | None
    doppler: bool
    scenarios: list
    wavelength: np.ndarray
    x: np.ndarray
    rho: np.ndarray
    phi: np.ndarray
    orb: np.ndarray
    delta_x: float
    fstar: tuple | None = None


def build_model(spec, inputs) -> Model:
    """inputs: {"linelist": (n,5) str rows, "molecules": {table id: (P, T, lam, log10 sigma)},
    "fstar": (w, log10 F) | None, "serpens": (gx, gy, gz, values) | None, "mdot": array | None,
    "tables": optional {(scenario index, species name): (tab_x, tab_y)} -- ready-made sigma tables (bench.py hands over
    the tables of the timed scene so that the CPU baseline times sumOverChords, not the set-up)}"""
    rows = inputs["linelist"]
    b = spec["bodies"]
    star, planet, moon = b["star"], b["planet"], b.get("moon")
    wgrid = spec["wavelength"]
    scenarios, all_lines = [], []
    for i_scn, s in enumerate(spec["scenarios"]):
        kind = s["type"]
        collisional = kind in ("barometric", "hydrostatic") or (kind == "powerLaw" and "P_0" in s)
        scn = Scenario(kind, dict(s), has_moon=kind in ("exomoon", "tidallyHeated"), T=s.get("T", 0.0))
        if kind == "powerLaw":
            scn.type = "powerLawP" if collisional else "powerLawN"
        for sp in s.get("species", []):
            el, ion, mass = SPECIES[sp["name"]]
            if collisional:
                chi, sigma_v = sp["chi"], np.sqrt(s["T"] * K_B / (mass * AMU))       # gasProperties.py:95
            else:
                chi, sigma_v = 1.0, sp["sigma_v"]                                     # gasProperties.py:274
            if inputs.get("tables") and (i_scn, sp["name"]) in inputs["tables"]:
                tx, ty = inputs["tables"][(i_scn, sp["name"])]
            else:
                tx, ty = lookup_table(rows, el, ion, sigma_v, wgrid)
            scn.constituents.append(Constituent("atom", chi, sp["name"], tx, ty))
            all_lines.extend(line_parameters(rows, el, ion, np.array([wgrid[0], wgrid[1]]))[0])
        for mol in s.get("molecules", []):
            P, T, lam, logsig = inputs["molecules"][mol["table"]]
            rgi = RegularGridInterpolator((P, T, lam), logsig, bounds_error=False, fill_value=np.log10(OFFSET))
            scn.constituents.append(Constituent("molecule", mol["chi"] if collisional else 1.0, mol["name"], rgi=rgi))
        if kind == "serpens":
            gx, gy, gz, vals = inputs["serpens"]
            scn.rgi = RegularGridInterpolator((gx, gy, gz), vals)                     # gasProperties.py:580
        if kind == "tidallyHeated":                                                   # gasProperties.py:418-423
            mdot = np.concatenate((inputs["mdot"], inputs["mdot"][::-1]))
            mass = SPECIES[s["species"][0]["name"]][2] * AMU
            scn.n_table = (np.linspace(0.0, 2.0 * np.pi, len(mdot)), np.log10(mdot * s["tau_photoionization"] / mass))
        scenarios.append(scn)
    if len(all_lines) == 0:                                                           # gasProperties.py:1069-1071
        wavelength = np.arange(wgrid[0], wgrid[1], wgrid[3])
    else:
        wavelength = arange_wavelength_grid(*wgrid, all_lines)
    x, rho, phi, orb, dx = grid_axes(spec["grid"])
    return Model(spec, star, planet, moon, bool(spec["doppler"]), scenarios, wavelength, x, rho, phi, orb, dx,
                 fstar=inputs.get("fstar"))


# ---------------------------------------------------------------------------------------------------
# number density -- gasProperties.py:143-601, distances celestialBodies.py:399-470, :552-583
# ---------------------------------------------------------------------------------------------------
def number_density(m: Model, scn: Scenario, x, phi, rho, orb):
    """(n_chords, n_x) for chord arrays phi, rho, orb"""
    y, z = rho * np.sin(phi), rho * np.cos(phi)
    X = x[None, :]
    Y, Z = y[:, None], z[:, None]
    s = scn.params
    if scn.type == "serpens":   # gasProperties.py:585-601, evaluated per chord (the batch call crashes in the reference)
        out = np.empty((len(phi), len(x)))
        for i in range(len(phi)):
            pts = np.array([x, np.repeat(y[i], np.size(x)), np.repeat(z[i], np.size(x))]).T
            out[i] = scn.rgi(pts)
        return out
    if scn.has_moon:
        bx, by = moon_position(m.moon, m.planet, m.star, orb)
        Rb = m.moon["R"]
    else:
        bx, by = planet_position(m.planet, orb)
        Rb = m.planet["R"]
    bx, by = np.asarray(bx)[:, None], np.asarray(by)[:, None]
    if scn.type == "torus":     # gasProperties.py:506-515
        a = np.sqrt((X - bx) ** 2 + (Y - by) ** 2)
        a_t = s["a_torus"]
        H = a_t * s["v_ej"] / np.sqrt(G_GRAV * m.planet["M"] / a_t)
        n_a = np.exp(-((a - a_t) / (4.0 * H)) ** 2)
        n_z = np.exp(-(Z / H) ** 2)
        term1 = 8.0 * H ** 2 * np.exp(-a_t ** 2 / (16.0 * H ** 2))
        term2 = 2.0 * np.sqrt(np.pi) * a_t * H * (erf(a_t / (4.0 * H)) + 1.0)
        n0 = 1.0 / (2.0 * np.pi ** 1.5 * H * (term1 + term2)) * s["N"]
        return n0 * np.multiply(n_a, n_z)
    r = np.sqrt((X - bx) ** 2 + (Y - by) ** 2 + Z ** 2)
    step = np.heaviside(r - Rb, 1.0)
    if scn.type == "barometric":    # gasProperties.py:155-160
        n0 = s["P_0"] / (K_B * s["T"])
        H = K_B * s["T"] * Rb ** 2 / (G_GRAV * s["mu"] * m.planet["M"])
        return n0 * np.exp((Rb - r) / H) * step
    if scn.type == "hydrostatic":   # gasProperties.py:197-203
        n0 = s["P_0"] / (K_B * s["T"])
        J0 = G_GRAV * s["mu"] * m.planet["M"] / (K_B * s["T"] * Rb)
        J = G_GRAV * s["mu"] * m.planet["M"] / (K_B * s["T"] * r) * step
        return n0 * np.exp(J - J0)
    if scn.type == "powerLawP":     # gasProperties.py:240-243
        return s["P_0"] / (K_B * s["T"]) * (Rb / r) ** s["q_esc"] * step
    if scn.type == "powerLawN":     # gasProperties.py:326-329
        return (s["q_esc"] - 3.0) / (4.0 * np.pi * Rb ** 3) * s["N"] * (Rb / r) ** s["q_esc"] * step
    if scn.type == "exomoon":       # gasProperties.py:371-373
        return (s["q"] - 3.0) / (4.0 * np.pi * Rb ** 3) * s["N"] * (Rb / r) ** s["q"] * step
    if scn.type == "tidallyHeated":  # gasProperties.py:425-461
        ph = moon_phase(m.moon, m.planet, m.star, orb) % (2.0 * np.pi)
        N = 10 ** np.interp(ph, scn.n_table[0], scn.n_table[1])      # interp1d linear == np.interp inside the table
        n0 = (s["q"] - 3.0) / (4.0 * np.pi * Rb ** 3) * np.asarray(N)[:, None]
        return n0 * (Rb / r) ** s["q"] * step
    raise ValueError(scn.type)


def scenario_shift(m: Model, scn: Scenario, orb):
    """gasProperties.py:908-916"""
    if not m.doppler:
        v = np.zeros(len(orb))
    elif scn.has_moon:
        v = moon_vlos(m.moon, m.planet, m.star, orb)
    else:
        v = planet_vlos(m.planet, m.star, orb)
    return doppler_shift(-v)


def molecular_sigma(con: Constituent, P, T, wl):
    """gasProperties.py:789-818: P (n_c, n_x), wl (n_c, n_wav) -> (n_c, n_x, n_wav)"""
    n_c, n_wav = wl.shape
    n_x = P.shape[1]
    Pf = np.clip(np.broadcast_to(P[:, :, None], (n_c, n_x, n_wav)).reshape(-1), a_min=1e-4, a_max=None)
    wf = np.broadcast_to(wl[:, None, :], (n_c, n_x, n_wav)).reshape(-1)
    pts = np.column_stack([Pf, np.full(Pf.size, T), wf])
    return (10 ** con.rgi(pts) - OFFSET).reshape(n_c, n_x, n_wav)


def column_density(m: Model, phi, rho, orb):
    """sum_x(n*chi)*dx per atomic constituent (gasProperties.py:940) -> list of (n_chords,) arrays"""
    cols = []
    for scn in m.scenarios:
        n_tot = number_density(m, scn, m.x, phi, rho, orb)
        for con in scn.constituents:
            if con.kind == "atom":
                cols.append(np.sum(n_tot * con.chi, axis=1) * m.delta_x)
    return cols


def los_optical_depth(m: Model, phi, rho, orb, wavelength):
    """gasProperties.py:885-956"""
    tau = np.zeros((len(phi), len(wavelength)))
    for scn in m.scenarios:
        shifts = scenario_shift(m, scn, orb)
        shifted = shifts[:, None] * wavelength[None, :]
        n_tot = number_density(m, scn, m.x, phi, rho, orb)
        for con in scn.constituents:
            if con.kind == "molecule":
                P = n_tot * K_B * scn.T
                sig = molecular_sigma(con, P, scn.T, shifted)
                tau += np.einsum("cx,cxw->cw", n_tot * con.chi, sig) * m.delta_x
            else:
                col = np.sum(n_tot * con.chi, axis=1) * m.delta_x
                uniq, inv = np.unique(shifts, return_inverse=True)
                sig_u = interp_log(uniq[:, None] * wavelength[None, :], con.tab_x, con.tab_y, OFFSET)
                tau += col[:, None] * sig_u[inv]
    return tau


def chord_masks(m: Model):
    """blocked mask and rho*clv weights of every chord (gasProperties.py:1178-1180, :1224-1232)"""
    phi, rho, orb = chord_grid(m.rho, m.phi, m.orb)
    y, z = rho * np.sin(phi), rho * np.cos(phi)
    mu = np.sqrt(np.clip(1.0 - rho ** 2 / m.star["R"] ** 2, 0, 1))
    clv = 1.0 - m.star["u1"] * (1.0 - mu) - m.star["u2"] * (1.0 - mu) ** 2
    blocked = np.sqrt((y - m.planet["a"] * np.sin(orb)) ** 2 + z ** 2) < m.planet["R"]
    for scn in m.scenarios:
        if scn.has_moon:
            y_m = moon_position(m.moon, m.planet, m.star, orb)[1]
            blocked |= (y - y_m) ** 2 + z ** 2 < m.moon["R"] ** 2
    return blocked, rho * clv


def sum_over_chords(m: Model, batch_size: int | None = None, wavelength=None, opaque_only=False):
    """gasProperties.py:1160-1258.  batch_size None = one batch (reshape-sum path, :1243-1253);
    otherwise the chunked np.add.at path (:1254-1256).  opaque_only: tau == 0 (R0 of SURVEY 8(c))."""
    wavelength = m.wavelength if wavelength is None else wavelength
    phi, rho, orb = chord_grid(m.rho, m.phi, m.orb)
    n_c, n_wav, n_orb = len(phi), len(wavelength), len(m.orb)
    y, z = rho * np.sin(phi), rho * np.cos(phi)
    orb_idx = np.tile(np.arange(n_orb), n_c // n_orb)       # == argmin |orb - axis| (:1176) without the n_c x n_orb temp
    mu = np.sqrt(np.clip(1.0 - rho ** 2 / m.star["R"] ** 2, 0, 1))
    clv = 1.0 - m.star["u1"] * (1.0 - mu) - m.star["u2"] * (1.0 - mu) ** 2
    star_shift = doppler_shift(m.star["vsini"] * rho / m.star["R"] * np.cos(phi - m.star["phi_rot"]))
    F_in_sum, F_out_sum = np.zeros((n_orb, n_wav)), np.zeros((n_orb, n_wav))
    bs = n_c if batch_size is None else int(batch_size)
    for i in range(0, n_c, bs):
        sl = slice(i, i + bs)
        if m.fstar is None:
            F_star = np.ones((len(phi[sl]), n_wav))
        else:
            F_star = interp_log(wavelength[None, :] / star_shift[sl, None], m.fstar[0], m.fstar[1], 0.0)
        F_star *= clv[sl, None]
        blocked = np.sqrt((y[sl] - m.planet["a"] * np.sin(orb[sl])) ** 2 + z[sl] ** 2) < m.planet["R"]
        for scn in m.scenarios:
            if scn.has_moon:
                y_m = moon_position(m.moon, m.planet, m.star, orb[sl])[1]
                blocked |= (y[sl] - y_m) ** 2 + z[sl] ** 2 < m.moon["R"] ** 2
        F_out = rho[sl, None] * F_star
        F_in = np.zeros_like(F_out)
        act = ~blocked
        if np.any(act):
            if opaque_only:
                F_in[act] = F_out[act]
            else:
                tau = los_optical_depth(m, phi[sl][act], rho[sl][act], orb[sl][act], wavelength)
                F_in[act] = F_out[act] * np.exp(-tau)
        if bs == n_c:
            F_in_sum = F_in.reshape(n_c // n_orb, n_orb, n_wav).sum(axis=0)
            F_out_sum = F_out.reshape(n_c // n_orb, n_orb, n_wav).sum(axis=0)
        else:
            np.add.at(F_in_sum, orb_idx[sl], F_in)
            np.add.at(F_out_sum, orb_idx[sl], F_out)
    return F_in_sum / F_out_sum


# ---------------------------------------------------------------------------------------------------
# light curve of a retrieval forward model -- mainRetrieval.py:69-93
# ---------------------------------------------------------------------------------------------------
def band_shift(m: Model):
    """mainRetrieval.py:77-78: v_los = planet.getLOSvelocity(orbphase); shift = calculateDopplerShift(v_los)"""
    return doppler_shift(planet_vlos(m.planet, m.star, m.orb))


def lightcurve(R, wavelength, shift, line_centres, bandwidth):
    """mainRetrieval.py:83-93: per phase, mean of R over the wavelengths within +-bandwidth/2 of any shifted line
    centre, divided by the maximum of R over all wavelengths."""
    out = []
    for idx in range(R.shape[0]):
        sel = np.zeros(len(wavelength), dtype=bool)
        for centre in line_centres:
            sel = sel + ((wavelength >= centre * shift[idx] - bandwidth / 2.) * (wavelength <= centre * shift[idx] + bandwidth / 2.))
        with np.errstate(invalid="ignore", divide="ignore"):
            out.append(np.mean(R[idx, sel]) / np.max(R[idx, :]) if sel.any() else np.float64("nan"))
    return np.array(out)


# ---------------------------------------------------------------------------------------------------
# tabulated-density producer -- SerpensExosphere.addInterpolatedDensity, gasProperties.py:548-583
# ---------------------------------------------------------------------------------------------------
def serpens_density(particle_pos_cm, grid, N, sigma_smoothing):
    """-> ((xPoints, yPoints, zPoints), n_histogram), the same numpy / scipy calls as the reference."""
    from scipy.ndimage import gaussian_filter
    x_mid, x_b, n_x, rho_b, n_rho = grid[0], grid[1], int(grid[2]), grid[3], int(grid[4])
    xBins = np.linspace(x_mid - x_b, x_mid + x_b, n_x + 1)                                  # geometryHandler.py:112-128
    yBins = np.linspace(-rho_b, rho_b, 2 * n_rho + 1)
    zBins = np.linspace(-rho_b, rho_b, 2 * n_rho + 1)
    cellVolume = (xBins[1] - xBins[0]) * (yBins[1] - yBins[0]) * (zBins[1] - zBins[0])
    n_histogram = np.histogramdd(particle_pos_cm, bins=[xBins, yBins, zBins])[0] * N / (np.size(particle_pos_cm, axis=0) * cellVolume)
    if sigma_smoothing > 0.:
        n_histogram = gaussian_filter(n_histogram, sigma=sigma_smoothing)
    xPoints = np.linspace(x_mid - x_b, x_mid + x_b, n_x, endpoint=False) + x_b / float(n_x)
    yPoints = np.linspace(-rho_b, rho_b, 2 * n_rho, endpoint=False) + 2. * rho_b / (4. * n_rho)
    zPoints = np.linspace(-rho_b, rho_b, 2 * n_rho, endpoint=False) + 2. * rho_b / (4. * n_rho)
    return (xPoints, yPoints, zPoints), n_histogram


def reference_batch_size(n_chords, n_wav, n_x, has_molecules, max_memory_gb=2.0):
    """memoryHandler.py:34-66 without the psutil clamp"""
    per = (n_x * n_wav * 64 if has_molecules else n_wav * 16) * 2.0
    return min(max(1, int(max_memory_gb * 1e9 / per)), n_chords)
