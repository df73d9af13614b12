"""Import alias: `import m3ae_b200` -> the package in ./mm-vqa-healthcare_b200/ (hyphenated directory name).

Every submodule is imported once under its real name and registered under the alias too, so
`m3ae_b200.functional` and `mm-vqa-healthcare_b200.functional` are the SAME module object (a lazily resolved alias
would load second copies of the submodules)."""
import importlib
import os
import sys

_root = os.path.dirname(os.path.abspath(__file__))
if _root not in sys.path:
    sys.path.insert(0, _root)
_REAL = "mm-vqa-healthcare_b200"
_pkg = importlib.import_module(_REAL)
for _sub in ("_lib", "ops", "functional", "bert_cross_layer", "encoders", "m3ae_module", "decoder", "decoder_model", "ddp"):
    _m = importlib.import_module(f"{_REAL}.{_sub}")
    sys.modules[f"{__name__}.{_sub}"] = _m
sys.modules[__name__] = _pkg
