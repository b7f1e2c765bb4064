"""ORACLE shim: bluesky.tools (geo, aero, areafilter)."""
from . import geo, aero, areafilter
