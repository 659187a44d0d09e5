"""Importable alias of the hyphen-named package directory (``import psrc_b200``)."""
import importlib
import os
import sys

_root = os.path.dirname(os.path.abspath(__file__))
if _root not in sys.path:
    sys.path.insert(0, _root)
_pkg = importlib.import_module("few-shot-prompt-learning-for-vision-language-models-in-imbalanced-datasets_b200")
sys.modules[__name__] = _pkg
