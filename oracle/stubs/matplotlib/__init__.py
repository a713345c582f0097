"""Inert shim: the reference imports matplotlib.pyplot in modules on the path but never plots there."""
