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
    assert L.wkv6_b200_abi_version() == 4


def test_argument_validation_without_gpu(built):
    L = built.lib()
    # C != H*64
    rc = L.wkv6_b200_forward(0, 1, 4, 96, 1, None, None, None, None, None, None, None, None, None, 0, None, 0, None)
    assert rc == -2 and b"64" in L.wkv6_b200_last_error()
    rc = L.wkv6_b200_forward(7, 1, 4, 64, 1, None, None, None, None, None, None, None, None, None, 0, None, 0, None)
    assert rc == -3
    # empty batch is a no-op; NULL tensors with work to do are rejected before any launch
    assert L.wkv6_b200_forward(0, 0, 4, 64, 1, None, None, None, None, None, None, None, None, None, 0, None, 0, None) == 0
    assert L.wkv6_b200_forward(0, 1, 4, 64, 1, None, None, None, None, None, None, None, None, None, 0, None, 0, None) == -1
    assert L.wkv6_b200_backward_workspace_bytes(1, 2, 10, 128, 2, 2) == 2 * 2 * 2 * 64 * 4
    assert L.wkv6_b200_backward_workspace_bytes(0, 2, 10, 128, 2, 2) == max(2 * 2 * 2 * 64 * 4 + 2 * 2 * 10 * 128 * 2,
                                                                            2 * L.wkv6_b200_checkpoint_bytes(2, 10, 128, 2))
    assert L.wkv6_b200_checkpoint_bytes(2, 130, 128, 2) == 2 * 2 * 3 * 64 * 64 * 2
    # time-segmented backward: few, long sequences only; 148 one-CTA SMs, <= 40 segments, each >= 8 chunks
    assert L.wkv6_b200_backward_scratch_bytes(0, 16, 750, 512, 8, 2) == 0
    assert L.wkv6_b200_backward_scratch_bytes(1, 1, 30000, 512, 8, 2) == 0           # fp32: exact scan, not segmented
    assert L.wkv6_b200_backward_scratch_bytes(0, 1, 12499, 512, 8, 2) == 16 * 9 * (64 * 64 + 128) * 4
    assert L.wkv6_b200_backward_workspace_bytes(0, 1, 12499, 512, 8, 2) == (
        2 * L.wkv6_b200_checkpoint_bytes(1, 12499, 512, 8) + L.wkv6_b200_backward_scratch_bytes(0, 1, 12499, 512, 8, 2))
    # time segmentation: only few, long sequences; the size is what the planner will use (<= 40 segments per sequence)
    assert L.wkv6_b200_forward_scratch_bytes(0, 16, 750, 512, 8, 2) == 0            # short
    assert L.wkv6_b200_forward_scratch_bytes(0, 16, 30000, 512, 8, 2) == 0          # 256 sequences: SMs are full
    # per sequence: nseg x (state + decay product) + chunks x decay prefix (the one-pass scheme's correction kernel)
    assert L.wkv6_b200_forward_scratch_bytes(0, 1, 30000, 512, 8, 2) == 16 * (18 * (64 * 64 + 64) + 469 * 64) * 4
    assert L.wkv6_b200_forward_scratch_bytes(0, 1, 40000, 256, 4, 1) == 4 * (40 * (64 * 64 + 64) + 625 * 64) * 4
    assert L.wkv6_b200_forward_scratch_bytes(1, 1, 90000, 1024, 16, 2) > 0


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


def _header_params(name):
    hdr = open(os.path.join(ROOT, "include", "wkv6_b200.h")).read()
    m = re.search(r"\b" + name + r"\s*\(([^;]*?)\)\s*;", hdr, re.S)
    assert m, name
    return [a.strip() for a in m.group(1).replace("\n", " ").split(",")]


def test_integration_snippets_match_header(built, tmp_path):
    """INTEGRATION.md level 0: the C++ stub compiles against include/wkv6_b200.h (and the torch headers), and the ctypes
    stub declares exactly the header's parameter list (the round-1 snippets had drifted by one pointer)."""
    import subprocess
    from torch.utils import cpp_extension as ce
    md = open(os.path.join(ROOT, "INTEGRATION.md")).read()
    cpp = re.search(r"```cpp\n(.*?)```", md, re.S).group(1)
    src = tmp_path / "snippet.cpp"
    src.write_text(cpp)
    inc = []
    for p_ in ce.include_paths():
        inc += ["-isystem", p_]
    abi = int(getattr(torch._C, "_GLIBCXX_USE_CXX11_ABI", True))
    subprocess.run(["g++", "-std=c++17", "-fsyntax-only", f"-D_GLIBCXX_USE_CXX11_ABI={abi}", "-I/usr/local/cuda/include",
                    "-I" + os.path.join(ROOT, "include"), *inc, str(src)], check=True)
    py = re.search(r"```python\n(import ctypes, torch.*?)```", md, re.S).group(1)
    py = py.replace('ctypes.CDLL("libwkv6_b200.so")', f'ctypes.CDLL({built.LIB_PATH!r})')
    ns = {}
    exec(compile(py, "INTEGRATION.md", "exec"), ns)
    params = _header_params("wkv6_b200_forward")
    argtypes = ns["_L"].wkv6_b200_forward.argtypes
    assert len(argtypes) == len(params) == 18
    for a, t in zip(params, argtypes):
        want = ctypes.c_int if a.startswith("int ") else ctypes.c_size_t if a.startswith("size_t ") else ctypes.c_void_p
        assert t is want, (a, t)
    # the package's own binding agrees with the header as well, for every compute entry point
    L = built.lib()
    for name in ("wkv6_b200_forward", "wkv6_b200_backward", "wkv6_b200_bi_forward", "wkv6_b200_bi_backward",
                 "wkv6_b200_combine", "wkv6_b200_tmix_mix"):
        assert len(getattr(L, name).argtypes) == len(_header_params(name)), name


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
