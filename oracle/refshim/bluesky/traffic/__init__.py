"""ORACLE shim: bluesky.traffic."""
from .traffic import Traffic
