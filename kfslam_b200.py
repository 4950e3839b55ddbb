"""Importable alias of the package directory `keyframe-based-slam_b200/` (hyphens are not identifiers)."""
import importlib as _il
import sys as _sys

_pkg = _il.import_module("keyframe-based-slam_b200")
_sys.modules[__name__] = _pkg
