"""ORACLE (test infrastructure, not product code; never imported by the product).

Minimal CPU stand-in for ``bluesky-simulator==1.0.3`` (reference pyproject.toml:14,
un-vendored and not installable offline) exposing exactly the surface the
reference env layer calls, so the UNMODIFIED reference
classes under /root/reference/bluesky_gym run on top of it.  The arithmetic
follows SURVEY.md Appendix A.  parity unpinned vs upstream BlueSky.
"""
from types import SimpleNamespace
from . import tools, simulation
from .tools import aero as _aero

settings = SimpleNamespace(simdt=0.05, asas_dt=1.0, asas_dtlookahead=300.0, asas_pzr=5.0,
                           asas_pzh=1000.0, asas_marh=1.05, asas_marv=1.05)
traf = None
sim = None
stack = None
scr = None
net = None
INIT, HOLD, OP, END = range(4)


def init(mode='sim', detached=True, **kwargs):
    """(Re)build the process-global singletons, like ``bs.init`` does."""
    global traf, sim, stack, scr
    from .traffic import Traffic
    from .stack import Stack
    sim = simulation.Simulation(settings)
    traf = Traffic(settings, sim)
    stack = Stack(sim, traf)
    sim.traf = traf
    sim.stack = stack
    scr = simulation.ScreenIO()
    tools.areafilter.reset()
