"""ORACLE (test infrastructure, not product code).

CPU restatement of the ISA atmosphere and speed conversions of
``bluesky/tools/aero.py`` (bluesky-simulator==1.0.3, un-vendored; SURVEY.md
Appendix A.2).  parity unpinned vs upstream: the upstream source is absent
from this container, the formulas below are the published ISA relations that
module implements.
"""
import numpy as np

kts = 0.514444          # m/s per knot
ft = 0.3048             # m per foot
fpm = ft / 60.0         # m/s per ft/min
nm = 1852.0             # m per nautical mile
g0 = 9.80665            # m/s2
R = 287.05287           # J/kg/K
p0 = 101325.0           # Pa
rho0 = 1.225            # kg/m3
T0 = 288.15             # K
gamma = 1.40
Rearth = 6371000.0      # m


def vtemp(h):
    return np.maximum(288.15 - 0.0065 * h, 216.65)


def vatmos(h):
    """ISA pressure, density, temperature for altitude h [m] (troposphere + lower stratosphere)."""
    T = vtemp(h)
    rhotrop = 1.225 * (T / 288.15) ** 4.256848030018761
    dhstrat = np.maximum(0.0, h - 11000.0)
    rho = rhotrop * np.exp(-dhstrat / 6341.552161)
    p = rho * R * T
    return p, rho, T


def vvsound(h):
    return np.sqrt(gamma * R * vtemp(h))


def vtas2mach(tas, h):
    return tas / vvsound(h)


def vmach2tas(M, h):
    return M * vvsound(h)


def vtas2cas(tas, h):
    p, rho, _ = vatmos(h)
    qdyn = p * ((1.0 + rho * tas * tas / (7.0 * p)) ** 3.5 - 1.0)
    cas = np.sqrt(7.0 * p0 / rho0 * ((qdyn / p0 + 1.0) ** (2.0 / 7.0) - 1.0))
    return np.where(tas < 0, -1 * cas, cas)


def vcas2tas(cas, h):
    p, rho, _ = vatmos(h)
    qdyn = p0 * ((1.0 + rho0 * cas * cas / (7.0 * p0)) ** 3.5 - 1.0)
    tas = np.sqrt(7.0 * p / rho * ((1.0 + qdyn / p) ** (2.0 / 7.0) - 1.0))
    return np.where(cas < 0, -1 * tas, tas)


def vcasormach2tas(spd, h):
    return np.where(np.abs(spd) < 1, vmach2tas(spd, h), vcas2tas(spd, h))


def vcasormach(spd, h):
    ismach = np.logical_and(0.1 < spd, spd < 1)
    tas = np.where(ismach, vmach2tas(spd, h), vcas2tas(spd, h))
    cas = np.where(ismach, vtas2cas(tas, h), spd)
    M = np.where(ismach, spd, vtas2mach(tas, h))
    return tas, cas, M


def tas2cas(tas, h):
    return float(vtas2cas(np.float64(tas), np.float64(h)))
