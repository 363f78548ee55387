"""Importable alias for the package directory `rust-tensorflow-serving2_b200/` (a hyphen is not a
valid Python identifier, so `import tfs_b200` stands in for it)."""
import importlib
import sys

sys.modules[__name__] = importlib.import_module("rust-tensorflow-serving2_b200")
