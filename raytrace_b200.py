"""Alias so that `import raytrace_b200` works: the package directory `rust-raytrace_b200` is not an identifier."""
import importlib
import sys

_pkg = importlib.import_module("rust-raytrace_b200")
sys.modules[__name__] = _pkg
