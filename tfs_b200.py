"""Importable alias for the package directory `rust-tensorflow-serving2_b200/` (a hyphen is not a valid
Python identifier, so `import tfs_b200` stands in for it).  Submodules are aliased too, so that
`tfs_b200.codec` and `rust-tensorflow-serving2_b200.codec` are one module object, not two copies."""
import importlib
import sys

_REAL = "rust-tensorflow-serving2_b200"
_pkg = importlib.import_module(_REAL)
for _name, _mod in list(sys.modules.items()):
    if _name.startswith(_REAL + "."):
        sys.modules[__name__ + _name[len(_REAL):]] = _mod
sys.modules[__name__] = _pkg
