"""wkv6_b200_forward_static_scratch (include/wkv6_b200.h): the forward for callers that cannot allocate device memory --
the time-segmentation scratch lives in a fixed `static __device__` array inside the library.  No counterpart in the
reference, so the checker is the existing entry point, wkv6_b200_forward
with a caller-owned scratch buffer, on the same inputs, at the tolerance of the other bf16 GPU tests.

This file is named to be collected LAST (pytest -x at the end of a round must not hide other tests behind it); it passes
alone in a fresh process and after the rest of the suite."""
import ctypes

import pytest
import torch

import oracle
from test_gpu_parity import TOL, _cuda, _inputs


def _p(t):
    return None if t is None else ctypes.c_void_p(t.data_ptr())


def test_static_scratch_rejects_shapes_beyond_the_array():
    """Validation happens before any CUDA call, so this runs without a GPU: 64 sequences x 1563 chunks would segment on
    29.9 MB of scratch, the array holds 16 MiB."""
    from paper_accurate_fast_cheap_b200 import _lib
    L = _lib.lib()
    need = L.wkv6_b200_forward_scratch_bytes(0, 8, 100000, 512, 8, 1)
    assert need > 16 << 20
    rc = L.wkv6_b200_forward_static_scratch(0, 8, 100000, 512, 8, *([None] * 9), 0, None)
    assert rc == -1
    msg = L.wkv6_b200_last_error().decode()
    assert str(need) in msg and str(16 << 20) in msg and "WKV6_STATIC_SCRATCH_BYTES" in msg
    # the other argument checks are those of wkv6_b200_forward
    assert L.wkv6_b200_forward_static_scratch(0, 1, 4096, 96, 1, *([None] * 9), 0, None) == -2


@pytest.mark.gpu
@pytest.mark.parametrize("rev", [False, True])
def test_static_scratch_forward_matches_forward_with_caller_scratch(rev):
    """Two calls in a row on different inputs (the second finds the first call's contents in the array), each against
    wkv6_b200_forward on a torch-allocated scratch buffer of wkv6_b200_forward_scratch_bytes: y and S_T within the bf16
    tolerance (they are the same kernels on the same plan, so in fact bit-equal), and against the fp64 oracle.  Both
    calls launch the three kernels of the time-segmented forward; free device memory does not move across the second."""
    from paper_accurate_fast_cheap_b200 import _lib
    L = _lib.lib()
    B, T, H = 1, 12499, 4
    C = 64 * H
    tol = TOL[torch.bfloat16]
    dev = torch.device("cuda:0")
    need = L.wkv6_b200_forward_scratch_bytes(0, B, T, C, H, 1)
    assert 0 < need <= 16 << 20
    stream = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    with torch.cuda.device(dev):
        scratch = torch.empty(need, device=dev, dtype=torch.uint8)
        for call, seed in enumerate((71, 72)):
            inp = _inputs(B, T, H, torch.bfloat16, with_state=(call == 1), seed=seed)
            g = _cuda(inp)
            s0 = g.get("s0")
            y0, y1 = torch.empty_like(g["r"]), torch.empty_like(g["r"])
            sT0 = torch.empty(B, H, 64, 64, device=dev)
            sT1 = torch.empty_like(sT0)
            args = (0, B, T, C, H, _p(g["r"]), _p(g["k"]), _p(g["v"]), _p(g["w"]), _p(g["u"]), _p(s0))
            _lib.check(L.wkv6_b200_forward(*args, _p(y0), _p(sT0), None, int(rev), _p(scratch), need, stream))
            torch.cuda.synchronize()
            free0 = torch.cuda.mem_get_info(dev)[0]
            n0 = L.wkv6_b200_launch_count()
            _lib.check(L.wkv6_b200_forward_static_scratch(*args, _p(y1), _p(sT1), None, int(rev), stream))
            torch.cuda.synchronize()
            assert L.wkv6_b200_launch_count() - n0 == 3, "the static-scratch call must run time-segmented"
            if call == 1:
                assert torch.cuda.mem_get_info(dev)[0] == free0, "the call changed the amount of free device memory"
            assert oracle.rel_err(y1.cpu(), y0.cpu().double()) <= tol
            assert oracle.rel_err(sT1.cpu(), sT0.cpu().double()) <= tol
            assert torch.equal(y1, y0) and torch.equal(sT1, sT0)
            y_ref, s_ref = oracle.fwd(inp["r"], inp["k"], inp["v"], inp["w"], inp["u"], inp.get("s0"), rev=rev,
                                      want_state=True, prec="f64")
            assert oracle.rel_err(y1, y_ref) <= tol and oracle.rel_err(sT1, s_ref) <= tol


@pytest.mark.gpu
def test_static_scratch_short_shape_runs_unsegmented():
    """A shape that would not be segmented goes through the same entry point without touching the array: one launch,
    same result as wkv6_b200_forward."""
    from paper_accurate_fast_cheap_b200 import _lib
    L = _lib.lib()
    B, T, H = 2, 200, 2
    C = 64 * H
    g = _cuda(_inputs(B, T, H, torch.bfloat16, seed=73))
    dev = g["r"].device
    stream = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    y0, y1 = torch.empty_like(g["r"]), torch.empty_like(g["r"])
    args = (0, B, T, C, H, _p(g["r"]), _p(g["k"]), _p(g["v"]), _p(g["w"]), _p(g["u"]), None)
    with torch.cuda.device(dev):
        _lib.check(L.wkv6_b200_forward(*args, _p(y0), None, None, 0, None, 0, stream))
        n0 = L.wkv6_b200_launch_count()
        _lib.check(L.wkv6_b200_forward_static_scratch(*args, _p(y1), None, None, 0, stream))
        assert L.wkv6_b200_launch_count() - n0 == 1
    torch.cuda.synchronize()
    assert torch.equal(y0, y1)
