"""ftl_rmsnorm_bwd_static: the RMSNorm backward for callers that cannot allocate scratch.

Who needs it: the recovery path of a training job after a CUDA out-of-memory error (every cudaMalloc
fails from then on, yet a backward fragment or a sanity pass may still have to run before state is
snapshotted), and C programs that embed libftl_b200.so beside an allocator that already owns the whole
device.  The entry point keeps its work area in a fixed array inside the library, so it takes inputs and
outputs only.

The test talks to the library through the C ABI alone (ctypes, raw pointers, raw stream handle) and checks
the new entry point against ftl_rmsnorm_bwd with caller-provided scratch: two calls in a row with different
shapes (the shared array keeps the first call's contents when the second starts), plus the refusal of a
request larger than the built-in area.  This file sorts last on purpose: under `pytest -x` every other test
has already run when it starts.
"""
import ctypes
import os
from ctypes import c_char_p, c_int, c_int64, c_void_p

import pytest
import torch

pytestmark = pytest.mark.gpu

LIB = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                   "fault_tolerant_llm_training_b200", "_native", "libftl_b200.so")


def _load():
    L = ctypes.CDLL(LIB)
    P, I = c_void_p, c_int
    L.ftl_rmsnorm_partials.argtypes, L.ftl_rmsnorm_partials.restype = [I], I
    L.ftl_rmsnorm_bwd.argtypes, L.ftl_rmsnorm_bwd.restype = [P, P, P, P, P, P, P, P, I, I, P], I
    L.ftl_rmsnorm_bwd_static.argtypes, L.ftl_rmsnorm_bwd_static.restype = [P, P, P, P, P, P, P, I, I, P], I
    L.ftl_rmsnorm_bwd_static_capacity.argtypes, L.ftl_rmsnorm_bwd_static_capacity.restype = [], c_int64
    L.ftl_last_error.argtypes, L.ftl_last_error.restype = [], c_char_p
    return L


def _case(rows, dim, seed, with_res):
    g = torch.Generator(device="cuda").manual_seed(seed)
    mk = lambda *s: torch.randn(*s, device="cuda", generator=g).to(torch.bfloat16)
    x, dy, w = mk(rows, dim), mk(rows, dim), mk(dim) * 0.1 + 1
    dres = mk(rows, dim) if with_res else None
    rstd = torch.rsqrt(x.float().pow(2).mean(-1) + 1e-5).contiguous()
    return x, dy, w, dres, rstd


def _ptr(t):
    return None if t is None else c_void_p(t.data_ptr())


def test_static_scratch_entry_matches_scratch_entry_on_two_calls_in_a_row():
    L = _load()
    assert L.ftl_rmsnorm_bwd_static_capacity() == 296 * 4096 * 4
    stream = c_void_p(torch.cuda.current_stream().cuda_stream)
    # second shape is smaller than the first: stale partial rows from call 1 sit in the array during call 2
    for call, (rows, dim, with_res) in enumerate([(2048, 4096, True), (77, 1024, False)]):
        x, dy, w, dres, rstd = _case(rows, dim, seed=call, with_res=with_res)
        scratch = torch.empty(L.ftl_rmsnorm_partials(rows) * dim, device="cuda", dtype=torch.float32)
        dx_a, dw_a = torch.empty_like(x), torch.empty_like(w)
        dx_b, dw_b = torch.full_like(x, float("nan")), torch.full_like(w, float("nan"))
        rc = L.ftl_rmsnorm_bwd(_ptr(dy), _ptr(x), _ptr(w), _ptr(rstd), _ptr(dres), _ptr(dx_a), _ptr(dw_a),
                               _ptr(scratch), rows, dim, stream)
        assert rc == 0
        rc = L.ftl_rmsnorm_bwd_static(_ptr(dy), _ptr(x), _ptr(w), _ptr(rstd), _ptr(dres), _ptr(dx_b), _ptr(dw_b),
                                      rows, dim, stream)
        assert rc == 0, L.ftl_last_error()
        torch.cuda.synchronize()
        # same kernels, same launch shape, same reduction order: the tolerance of the other norm tests holds
        # with room to spare (the results are in fact bit-identical)
        torch.testing.assert_close(dx_b.float(), dx_a.float(), rtol=2e-2, atol=2e-2)
        torch.testing.assert_close(dw_b.float(), dw_a.float(), rtol=2e-2, atol=2e-2)
        assert torch.equal(dx_b, dx_a) and torch.equal(dw_b, dw_a)


def test_static_scratch_entry_refuses_oversize_requests():
    L = _load()
    stream = c_void_p(torch.cuda.current_stream().cuda_stream)
    x, dy, w, _, rstd = _case(16, 8192, seed=3, with_res=False)
    dx, dw = torch.empty_like(x), torch.empty_like(w)
    rc = L.ftl_rmsnorm_bwd_static(_ptr(dy), _ptr(x), _ptr(w), _ptr(rstd), None, _ptr(dx), _ptr(dw), 16, 8192, stream)
    assert rc != 0
    msg = L.ftl_last_error().decode()
    assert "8192" in msg and "4096" in msg and "ftl_rmsnorm_bwd" in msg
