"""ORACLE (test infrastructure, not product code).

Documented, parameterised restatement of the OpenAP-style performance envelope
BlueSky applies each step (``bluesky/traffic/performance/openap``).
The exact OpenAP A320 coefficients are unknown here (package un-vendored):
the constants below are builder-chosen and calibrated to the anchors the
reference's shipped logs give (Horizontal at alt=0 cruises at ~80 m/s although
acspd=150; Vertical and Sector hold 150 m/s CAS).  parity unpinned vs upstream.

Phase codes and the phase -> (vmin, vmax) table are the contract shared with
the CUDA core (bluesky_gym_hallucination_b200/csrc/perf.cuh restates the same table).
"""
import numpy as np
from ..tools.aero import kts, ft, fpm, vtas2cas, vcas2tas, vtas2mach, vmach2tas

NA, GD, IC, CL, CR, DE, AP = 0, 1, 2, 3, 4, 5, 6

# A320-like envelope [m/s CAS], [m/s], [m]
ENVELOPE = dict(
    vminic=65.0, vmaxic=82.0,
    vminer=80.0, vmaxer=180.0,
    vminap=60.0, vmaxap=100.0,
    vsmin=-20.0, vsmax=20.0,
    hmax=12131.0, mmo=0.82, axmax=2.0,
)


def phase_fixwing(tas, vs, alt):
    """Flight phase from speed/roc/alt (SI in; thresholds in ft and ft/min)."""
    roc = vs / fpm
    h = alt / ft
    ph = np.zeros(len(tas), dtype=np.int64)
    ph[h <= 75.0] = GD
    ph[(h > 75.0) & (h <= 1000.0) & (roc >= 150.0)] = IC
    ph[(h > 75.0) & (h <= 1000.0) & (roc <= -150.0)] = AP
    ph[(h > 1000.0) & (roc >= 150.0)] = CL
    ph[(h > 1000.0) & (roc <= -150.0)] = DE
    ph[(h > 10000.0) & (roc > -150.0) & (roc < 150.0)] = CR
    return ph


class Perf:
    def __init__(self, traf):
        self.traf = traf
        self.reset()

    def reset(self):
        z = np.zeros(0)
        self.phase = np.zeros(0, dtype=np.int64)
        self.vmin = z.copy(); self.vmax = z.copy()
        self.vsmin = z.copy(); self.vsmax = z.copy()
        self.hmax = z.copy(); self.mmo = z.copy(); self.axmax = z.copy()

    def create(self, n):
        E = ENVELOPE
        self.phase = np.append(self.phase, np.zeros(n, dtype=np.int64))
        self.vmin = np.append(self.vmin, np.zeros(n))
        self.vmax = np.append(self.vmax, np.full(n, E['vmaxer']))
        self.vsmin = np.append(self.vsmin, np.full(n, E['vsmin']))
        self.vsmax = np.append(self.vsmax, np.full(n, E['vsmax']))
        self.hmax = np.append(self.hmax, np.full(n, E['hmax']))
        self.mmo = np.append(self.mmo, np.full(n, E['mmo']))
        self.axmax = np.append(self.axmax, np.full(n, E['axmax']))

    def delete(self, idx):
        for k in ('phase', 'vmin', 'vmax', 'vsmin', 'vsmax', 'hmax', 'mmo', 'axmax'):
            setattr(self, k, np.delete(getattr(self, k), idx))

    def update(self):
        E = ENVELOPE
        t = self.traf
        self.phase = ph = phase_fixwing(t.tas, t.vs, t.alt)
        enroute = (ph == CL) | (ph == CR) | (ph == DE)
        vmin = np.zeros(len(ph))
        vmin = np.where(ph == IC, E['vminic'], vmin)
        vmin = np.where(enroute, E['vminer'], vmin)
        vmin = np.where(ph == AP, E['vminap'], vmin)
        vmax = np.full(len(ph), E['vmaxer'])
        vmax = np.where((ph == IC) | (ph == GD), E['vmaxic'], vmax)
        vmax = np.where(ph == AP, E['vmaxap'], vmax)
        self.vmin, self.vmax = vmin, vmax

    def limits(self, intent_v_tas, intent_vs, intent_h, ax):
        allow_h = np.where(intent_h > self.hmax, self.hmax, intent_h)
        intent_v_cas = vtas2cas(intent_v_tas, allow_h)
        allow_v_cas = np.where(intent_v_cas < self.vmin, self.vmin, intent_v_cas)
        allow_v_cas = np.where(intent_v_cas > self.vmax, self.vmax, allow_v_cas)
        allow_v_tas = vcas2tas(allow_v_cas, allow_h)
        allow_v_tas = np.where(vtas2mach(allow_v_tas, allow_h) > self.mmo,
                               vmach2tas(self.mmo, allow_h), allow_v_tas)
        vs_max_with_acc = (1 - ax / self.axmax) * self.vsmax
        allow_vs = np.where((intent_vs > 0) & (intent_vs > self.vsmax), vs_max_with_acc, intent_vs)
        allow_vs = np.where((intent_vs < 0) & (intent_vs < self.vsmin), vs_max_with_acc, allow_vs)
        return allow_v_tas, allow_vs, allow_h
