"""Importable alias of the product package (its directory name carries hyphens): ``import b2rf``."""
import importlib
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
_pkg = importlib.import_module("scalable-machine-learning-with-apache-spark_b200")
sys.modules[__name__] = _pkg
