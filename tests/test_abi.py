"""CPU-side checks of the drop-in boundary: the C-ABI library builds, loads and exports every symbol that
include/wkv6_b200.h declares; argument validation works without a GPU; the product never falls back to CPU."""
import ctypes
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def built():
    from paper_accurate_fast_cheap_b200 import build
    build.build_all()
    from paper_accurate_fast_cheap_b200 import _lib
    return _lib


def test_header_symbols_are_exported(built):
    hdr = open(os.path.join(ROOT, "include", "wkv6_b200.h")).read()
    declared = set(re.findall(r"\b(wkv6_b200_[a-z0-9_]+)\s*\(", hdr))
    assert declared == set(built.SYMBOLS)
    L = built.lib()
    for s in declared:
        assert hasattr(L, s), s
    assert L.wkv6_b200_abi_version() == 2


def test_argument_validation_without_gpu(built):
    L = built.lib()
    # C != H*64
    rc = L.wkv6_b200_forward(0, 1, 4, 96, 1, None, None, None, None, None, None, None, None, None, 0, None)
    assert rc == -2 and b"64" in L.wkv6_b200_last_error()
    rc = L.wkv6_b200_forward(7, 1, 4, 64, 1, None, None, None, None, None, None, None, None, None, 0, None)
    assert rc == -3
    # empty batch is a no-op; NULL tensors with work to do are rejected before any launch
    assert L.wkv6_b200_forward(0, 0, 4, 64, 1, None, None, None, None, None, None, None, None, None, 0, None) == 0
    assert L.wkv6_b200_forward(0, 1, 4, 64, 1, None, None, None, None, None, None, None, None, None, 0, None) == -1
    assert L.wkv6_b200_backward_workspace_bytes(1, 2, 10, 128, 2, 2) == 2 * 2 * 2 * 64 * 4
    assert L.wkv6_b200_backward_workspace_bytes(0, 2, 10, 128, 2, 2) == max(2 * 2 * 2 * 64 * 4 + 2 * 2 * 10 * 128 * 2,
                                                                            2 * L.wkv6_b200_checkpoint_bytes(2, 10, 128, 2))
    assert L.wkv6_b200_checkpoint_bytes(2, 130, 128, 2) == 2 * 2 * 3 * 64 * 64 * 2


def test_torch_shim_registers_reference_schemas(built):
    built.load_torch_ops()
    for name in ("forward", "backward", "forward_fp32", "backward_fp32", "forward_fp64", "backward_fp64"):
        op = getattr(torch.ops.wkv6, name)
        schema = str(op.default._schema)
        assert schema.startswith(f"wkv6::{name}(int _0, int _1, int _2, int _3, Tensor")
        assert schema.endswith("-> ()")
    # TorchScript can resolve the op by name, as the reference ScriptModule does (model.py:33-35, 132)
    @torch.jit.script
    def f(r: torch.Tensor, u: torch.Tensor, y: torch.Tensor):
        torch.ops.wkv6.forward_fp32(1, 1, 64, 1, r, r, r, r, u, y)
    assert "wkv6::forward_fp32" in str(f.graph)


def test_no_cpu_fallback(built):
    import paper_accurate_fast_cheap_b200 as pk
    c = torch.zeros(1, 4, 64)
    with pytest.raises(AssertionError):
        pk.wkv6_forward_raw(c, c, c, c, torch.zeros(1, 64))
    with pytest.raises(Exception):
        y = torch.empty_like(c)
        built.load_torch_ops()
        torch.ops.wkv6.forward_fp32(1, 4, 64, 1, c, c, c, c, torch.zeros(1, 64), y)


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, "paper_accurate_fast_cheap_b200")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                src = open(os.path.join(dp, f)).read()
                assert "import oracle" not in src and "from oracle" not in src and "oracle/" not in src, f
