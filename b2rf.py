"""Importable alias of the product package (its directory name carries hyphens): ``import b2rf``.
Every submodule is registered under both names so that ``b2rf.trainer`` and the package-internal relative imports
share one module object (one set of classes, one loaded library)."""
import importlib
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
_REAL = "scalable-machine-learning-with-apache-spark_b200"
_pkg = importlib.import_module(_REAL)
for _sub in ("_build", "_lib", "params", "forest", "synth", "trainer", "spark", "persistence"):
    try:
        _m = importlib.import_module(f"{_REAL}.{_sub}")
    except ModuleNotFoundError as _e:          # optional submodules (spark adapter needs pyspark or the stub)
        if _sub in str(_e):
            continue
        raise
    sys.modules[f"{__name__}.{_sub}"] = _m
sys.modules[__name__] = _pkg
