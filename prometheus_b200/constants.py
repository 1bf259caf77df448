"""cgs constants of the transit forward model.

The numerical values are the (deliberately low-precision) literals the reference
uses (reference: pythonScripts/constants.py:13-26); results only match the
reference if exactly these values are used, so they are restated here and
mirrored in csrc/pb200_internal.cuh.
"""
import math

ELEMENTARY_CHARGE = 4.803e-10      # esu                      (constants.py:13)
ELECTRON_MASS = 9.109e-28          # g                        (constants.py:14)
LIGHT_SPEED = 2.998e10             # cm/s                     (constants.py:15)
GRAV = 6.674 * 10 ** (-8)          # cm^3 g^-1 s^-2           (constants.py:16)
BOLTZMANN = 1.381 * 10 ** (-16)    # erg/K                    (constants.py:17)
AMU = 1.661 * 10 ** (-24)          # g                        (constants.py:18)
R_JUPITER = 7.1492e9               # cm                       (constants.py:19)
M_JUPITER = 1.898e30               # g                        (constants.py:20)
M_EARTH = 5.974e27
R_SUN = 6.96e10
M_SUN = 1.988e33
R_IO = 1.822e8
AU = 1.496e13

# short aliases with the reference's own names, so host code reads like the reference
e = ELEMENTARY_CHARGE
m_e = ELECTRON_MASS
c = LIGHT_SPEED
G = GRAV
k_B = BOLTZMANN
amu = AMU
R_J = R_JUPITER
M_J = M_JUPITER
R_Io = R_IO

LOOKUP_OFFSET = 1e-50              # sigma-table offset       (gasProperties.py:638, :761)


def calculateDopplerShift(v):
    """Relativistic Doppler factor sqrt((1-b)/(1+b)), b = v/c (constants.py:31-45)."""
    import numpy as np
    beta = v / LIGHT_SPEED
    return np.sqrt((1.0 - beta) / (1.0 + beta))


# (name) -> (element, ionisation stage, mass [amu]); registry of constants.py:139-162.
SPECIES_TABLE = {
    "NaI": ("Na", "1", 22.99), "KI": ("K", "1", 39.0983),
    "SiI": ("Si", "1", 28.0855), "SiII": ("Si", "2", 28.0855),
    "SiIII": ("Si", "3", 28.0855), "SiIV": ("Si", "4", 28.0855),
    "MgI": ("Mg", "1", 24.305), "MgII": ("Mg", "2", 24.305),
    "AlI": ("Al", "1", 26.9815), "CaI": ("Ca", "1", 40.078), "CaII": ("Ca", "2", 40.078),
    "TiI": ("Ti", "1", 47.867), "TiII": ("Ti", "2", 47.867), "CrI": ("Cr", "1", 51.9961),
    "MnI": ("Mn", "1", 54.938), "FeI": ("Fe", "1", 55.845), "CoI": ("Co", "1", 58.933),
    "NiI": ("Ni", "1", 58.6934), "OI": ("O", "1", 15.999), "CII": ("C", "2", 12.011),
    "SIII": ("S", "3", 32.06), "SIV": ("S", "4", 32.06),
}
# species that are NOT in the reference registry but used by workload C5
EXTRA_SPECIES = {"FeII": ("Fe", "2", 55.845)}

assert math.isclose(GRAV, 6.674e-8)
