"""Import alias: `import m3ae_b200` -> the package in ./mm-vqa-healthcare_b200/ (hyphenated directory)."""
import importlib
import os
import sys

_root = os.path.dirname(os.path.abspath(__file__))
if _root not in sys.path:
    sys.path.insert(0, _root)
_pkg = importlib.import_module("mm-vqa-healthcare_b200")
sys.modules[__name__] = _pkg
