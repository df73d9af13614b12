"""B200-native implementation of M3AE's multimodal co-attention fusion hot path.

Host code (this package) mirrors the reference's module interface (`BertCrossLayer`,
`M3AETransformerSS.infer`, `Decoder`); all arithmetic on the path runs in hand-written sm_100a
kernels behind the C-ABI in include/m3ae_b200.h (libm3ae_b200.so).

The directory name carries a hyphen (project naming); import it as `m3ae_b200` (root-level alias)
or via importlib.import_module("mm-vqa-healthcare_b200").
"""
from . import _lib, ops  # noqa: F401

__all__ = ["_lib", "ops"]
