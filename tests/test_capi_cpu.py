"""CPU checks of the drop-in boundary: the C-ABI library loads without a GPU, exports every symbol include/m3ae_b200.h
declares, the ctypes signatures cover them all, and compute calls fail loudly (no CPU fallback)."""
import ctypes
import os
import re
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
HEADER = os.path.join(ROOT, "include", "m3ae_b200.h")


@pytest.fixture(scope="module")
def lib():
    import m3ae_b200
    if not os.path.exists(m3ae_b200._lib.LIB_PATH):
        subprocess.run(["make", "-C", os.path.join(ROOT, "mm-vqa-healthcare_b200", "csrc"), "-j8"], check=True)
    return m3ae_b200._lib.load()


def declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(m3ae_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_are_exported_and_bound(lib):
    import m3ae_b200
    names = declared_symbols()
    assert len(names) >= 15
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/m3ae_b200.h but not exported by libm3ae_b200.so"
        assert n in m3ae_b200._lib.SIGNATURES, f"{n} has no ctypes signature in _lib.py"
    assert lib.m3ae_version() == 100


def test_struct_layouts_match_header_sizes():
    # sizeof as the C compiler sees it (compile a tiny probe against the header)
    import tempfile
    import m3ae_b200
    with tempfile.TemporaryDirectory() as d:
        src = os.path.join(d, "p.c")
        open(src, "w").write('#include <stdio.h>\n#include "m3ae_b200.h"\nint main(){printf("%zu %zu %zu %zu %zu\\n", sizeof(m3ae_gemm_epilogue_t), sizeof(m3ae_gemm_problem_t), sizeof(m3ae_attn_params_t), sizeof(m3ae_skinny_params_t), sizeof(m3ae_dropbits_item_t));return 0;}\n')
        exe = os.path.join(d, "p")
        subprocess.run(["gcc", "-I", os.path.join(ROOT, "include"), src, "-o", exe], check=True)
        sizes = [int(x) for x in subprocess.run([exe], capture_output=True, text=True, check=True).stdout.split()]
    L = m3ae_b200._lib
    assert sizes == [ctypes.sizeof(L.GemmEpilogue), ctypes.sizeof(L.GemmProblem), ctypes.sizeof(L.AttnParams),
                     ctypes.sizeof(L.SkinnyParams), ctypes.sizeof(L.DropBitsItem)]


def test_no_cpu_fallback(lib):
    from m3ae_b200 import ops
    from m3ae_b200.bert_cross_layer import BertCrossLayer
    with pytest.raises(RuntimeError):
        ops.cast(torch.zeros(8), torch.zeros(8, dtype=torch.bfloat16))
    if not torch.cuda.is_available():
        rc = lib.m3ae_cast(None, 0, None, 1, 8, None)
        assert rc != 0 and lib.m3ae_last_error()


def test_product_path_does_not_import_oracle():
    pkg = os.path.join(ROOT, "mm-vqa-healthcare_b200")
    for fn in os.listdir(pkg):
        if fn.endswith(".py"):
            src = open(os.path.join(pkg, fn)).read()
            assert not re.search(r"^\s*(from|import)\s+\.*oracle|import_module\(.oracle", src, flags=re.M), fn


def test_state_dict_keys_match_reference_layout():
    from transformers import RobertaConfig
    from m3ae_b200.bert_cross_layer import BertCrossLayer
    from m3ae_b200.functional import PARAM_NAMES
    from oracle import m3ae_oracle as O
    cfg = RobertaConfig(hidden_size=768, num_attention_heads=12, intermediate_size=3072)
    layer = BertCrossLayer(cfg)
    keys = list(layer.state_dict().keys())
    assert sorted(keys) == sorted(O.make_cross_layer_state(0).keys()) and len(keys) == 26
    assert [n for n, _ in layer.named_parameters()] == keys
    assert sorted(PARAM_NAMES) == sorted(keys)
    assert sum(p.numel() for p in layer.parameters()) == 9451776
    for got, name in zip(layer.ordered_params(), PARAM_NAMES):
        assert got is dict(layer.named_parameters())[name]
