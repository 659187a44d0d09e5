"""Importable alias of the hyphen-named package directory (``import psrc_b200``).

Every submodule is imported once under its real name and registered under the alias too, so
``psrc_b200.step`` and ``<real name>.step`` are the same module object (no duplicate imports).
"""
import importlib
import os
import pkgutil
import sys

_REAL = "few-shot-prompt-learning-for-vision-language-models-in-imbalanced-datasets_b200"
_root = os.path.dirname(os.path.abspath(__file__))
if _root not in sys.path:
    sys.path.insert(0, _root)
_pkg = importlib.import_module(_REAL)
for _m in pkgutil.iter_modules(_pkg.__path__):
    if _m.name.startswith("_"):
        continue
    sys.modules[f"{__name__}.{_m.name}"] = importlib.import_module(f"{_REAL}.{_m.name}")
sys.modules[__name__] = _pkg
