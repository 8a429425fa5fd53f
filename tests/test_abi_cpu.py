"""CPU checks of the C-ABI library: it loads, and exports every symbol include/splitq.h declares.
No compute call is made (no GPU here)."""
import ctypes
import os
import re

import pytest

import splitq
from splitq import _native as N

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "splitq.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(splitq_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    lib = N.load()
    names = declared_symbols()
    assert len(names) >= 11
    for name in names:
        assert hasattr(lib, name), f"{name} declared in splitq.h but not exported"
        assert name in N.SIGNATURES, f"{name} has no ctypes signature in _native.py"
    assert set(N.SIGNATURES) == set(names)


def test_version_and_size_queries_run_without_gpu():
    lib = N.load()
    assert lib.splitq_version() == 100
    import splitq_oracle as O
    for S, D, ladder, mode in [(512, 512, [(256, 4)], 0), (4096, 896, [(1024, 2), (1024, 4), (1024, 8)], 1),
                               (37, 5, [(3, 2), (10, 8)], 1), (1, 4, [], 0)]:
        assert splitq.payload_bytes(S, D, ladder, mode) == O.payload_bytes(S, D, ladder, mode)
    assert lib.splitq_quant_pack_workspace_bytes(3, 128, 64, 0) == 48
    assert lib.splitq_quant_pack_workspace_bytes(3, 128, 64, 1) == 0
    assert lib.splitq_importance_colmean_workspace_bytes(1, 8, 512, 512) > 0


def test_bad_ladder_is_rejected_with_message():
    with pytest.raises(N.SplitqError) as e:
        splitq.payload_bytes(16, 8, [(4, 3)], 0)
    assert "not in {2,4,8}" in str(e.value)
    with pytest.raises(N.SplitqError):
        splitq.payload_bytes(16, 8, [(10, 4), (10, 8)], 0)
    with pytest.raises(N.SplitqError):
        splitq.payload_bytes(16, 8, [(2, 4), (2, 4)], 0)


def test_cpu_tensors_are_refused():
    import torch
    with pytest.raises(N.SplitqError):
        splitq.importance_from_attention(torch.zeros(1, 1, 4, 4))


def test_unknown_method_raises_reference_message():
    with pytest.raises(ValueError) as e:
        splitq.get_importance_order("nope", [], 0)
    assert str(e.value) == "Unknown method: nope"
