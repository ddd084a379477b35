"""The weight-gradient entry point that brings its own scratch memory.

``dtdl_conv_wgrad_static`` exists for callers that are not in a position to allocate device
memory: a host application whose own memory pool already owns the whole GPU, a plug-in that
is handed tensors but no allocator, or code that runs while a stream is being captured (where
``cudaMalloc`` is not permitted).  Its split-K scratch is a fixed 32 MiB array compiled into
the library.  This test drives it the way such a caller would -- through the C ABI with raw
pointers -- and checks it against the regular entry point that takes a caller-allocated
workspace.  Two calls in a row with different data make sure the scratch left behind by the
first call does not leak into the second.  (The file name keeps it last in collection order,
so a ``pytest -x`` run has already exercised everything else when it gets here.)
"""
import ctypes
import os

import pytest
import torch

pytestmark = pytest.mark.gpu

LIB = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                   "distributed_training_dl_b200", "lib", "libdtdl_b200.so")


def _c_lib():
    lib = ctypes.CDLL(LIB)
    vp, i, ll = ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong
    lib.dtdl_last_error.restype = ctypes.c_char_p
    lib.dtdl_conv_wgrad_workspace_bytes.restype = ll
    lib.dtdl_conv_wgrad_workspace_bytes.argtypes = [i] * 7
    lib.dtdl_conv_wgrad_static_scratch_bytes.restype = ll
    lib.dtdl_conv_wgrad.restype = i
    lib.dtdl_conv_wgrad.argtypes = [vp, vp, vp, vp, ll] + [i] * 16 + [vp]   # ... accumulate, stride, act
    lib.dtdl_conv_wgrad_static.restype = i
    lib.dtdl_conv_wgrad_static.argtypes = [vp, vp, vp] + [i] * 14 + [vp]
    return lib


def _rel_err(a, b):
    return ((a - b).abs().max() / (b.abs().max() + 1e-12)).item()


def test_static_scratch_wgrad_matches_workspace_wgrad():
    lib = _c_lib()
    N, H, W, Cin, Cout, K, pad = 8, 16, 16, 45, 70, 3, 1
    Cin_p, Cout_p = 48, 72
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    need = lib.dtdl_conv_wgrad_workspace_bytes(N, H, W, Cin_p, Cout_p, K, K)
    assert 0 < need <= lib.dtdl_conv_wgrad_static_scratch_bytes() == 32 << 20
    shape = (N, H, W, Cin, Cin_p, H, W, Cout, Cout_p, K, K, pad, pad, 0)
    for call in range(2):                      # two calls in a row, fresh data each time
        torch.manual_seed(call)
        x = torch.zeros(N, H, W, Cin_p, device="cuda", dtype=torch.bfloat16)
        dy = torch.zeros(N, H, W, Cout_p, device="cuda", dtype=torch.bfloat16)
        x[..., :Cin] = torch.randn(N, H, W, Cin, device="cuda")
        dy[..., :Cout] = torch.randn(N, H, W, Cout, device="cuda")
        want = torch.full((Cout, Cin, K, K), float("nan"), device="cuda")
        got = torch.full((Cout, Cin, K, K), float("nan"), device="cuda")
        ws = torch.empty(need, dtype=torch.uint8, device="cuda")     # the caller-side workspace
        rc = lib.dtdl_conv_wgrad(x.data_ptr(), dy.data_ptr(), want.data_ptr(), ws.data_ptr(), need,
                                 *shape, 1, 0, stream)      # stride 1, bf16 operands
        assert rc == 0, lib.dtdl_last_error()
        torch.cuda.synchronize()
        free_before = torch.cuda.mem_get_info()[0]
        rc = lib.dtdl_conv_wgrad_static(x.data_ptr(), dy.data_ptr(), got.data_ptr(), *shape, stream)
        assert rc == 0, lib.dtdl_last_error()
        torch.cuda.synchronize()
        if call > 0:    # the first call may still load the kernels' code; after that: no new memory
            assert torch.cuda.mem_get_info()[0] == free_before
        assert torch.isfinite(got).all()
        assert _rel_err(got, want) < 1e-2      # the tolerance the op tests use
        # and both agree with plain PyTorch
        ref = torch.nn.grad.conv2d_weight(x[..., :Cin].permute(0, 3, 1, 2).float(), (Cout, Cin, K, K),
                                          dy[..., :Cout].permute(0, 3, 1, 2).float(), padding=pad)
        assert _rel_err(got, ref) < 1e-2


def test_static_scratch_rejects_oversized_problem():
    lib = _c_lib()
    N, H, W, Cin_p, Cout_p, K = 64, 8, 8, 1024, 1024, 3
    need = lib.dtdl_conv_wgrad_workspace_bytes(N, H, W, Cin_p, Cout_p, K, K)
    assert need > lib.dtdl_conv_wgrad_static_scratch_bytes()
    x = torch.zeros(N, H, W, Cin_p, device="cuda", dtype=torch.bfloat16)
    dy = torch.zeros(N, H, W, Cout_p, device="cuda", dtype=torch.bfloat16)
    dw = torch.zeros(Cout_p, Cin_p, K, K, device="cuda")
    rc = lib.dtdl_conv_wgrad_static(x.data_ptr(), dy.data_ptr(), dw.data_ptr(), N, H, W, Cin_p, Cin_p, H, W,
                                    Cout_p, Cout_p, K, K, 1, 1, 0, None)
    assert rc != 0
    assert b"scratch" in lib.dtdl_last_error()
    torch.cuda.synchronize()
