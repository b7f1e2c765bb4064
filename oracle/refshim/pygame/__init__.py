"""ORACLE shim: pygame is imported by the reference env modules but only used by
render_mode="human", which is unavailable on this box (no graphics stack)."""
