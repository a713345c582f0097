"""Inert shim for gym 0.26 (reference: Environment/Exoskeleton_env.py:5-7,34,74,111)."""
from . import spaces, utils  # noqa: F401


class Env(object):
    metadata = {}

    def __init__(self, *a, **k):
        pass
