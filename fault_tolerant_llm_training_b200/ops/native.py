"""ctypes binding of libftl_b200.so (C ABI in include/ftl_b200.h).

The library has no dependency on libtorch: tensors cross the boundary as raw device pointers plus the
current CUDA stream handle, so every launch lands on the stream PyTorch is using (and is captured by
CUDA graphs like any other launch).  Loading fails loudly: on a GPU box a missing or stale library
is an error, never a silent fallback to PyTorch ops.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import c_float, c_int, c_int64, c_void_p
from typing import Optional

import torch

_LIB: Optional[ctypes.CDLL] = None
LAUNCHES = 0  # native calls made through this module
KERNELS = 0   # kernels of libftl_b200.so launched through this module (bench.py reports the delta)

PKG_DIR = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB_PATH = os.path.join(PKG_DIR, "_native", "libftl_b200.so")


class NativeError(RuntimeError):
    pass


def _sig(lib, name, argtypes, restype=c_int):
    fn = getattr(lib, name)
    fn.argtypes, fn.restype = argtypes, restype
    return fn


def lib() -> ctypes.CDLL:
    global _LIB
    if _LIB is not None:
        return _LIB
    if not os.path.exists(LIB_PATH):
        if os.environ.get("FTL_NO_AUTOBUILD"):
            raise NativeError(f"{LIB_PATH} is missing; run `python -m fault_tolerant_llm_training_b200.build`")
        from ..build import build

        build()
    L = ctypes.CDLL(LIB_PATH)
    P, I, I64, F = c_void_p, c_int, c_int64, c_float
    _sig(L, "ftl_rmsnorm_fwd", [P, P, P, P, P, P, I, I, F, P])
    _sig(L, "ftl_rmsnorm_partials", [I])
    _sig(L, "ftl_rmsnorm_bwd", [P, P, P, P, P, P, P, P, I, I, P])
    _sig(L, "ftl_rope_inplace", [P, P, P, I64, I, I, I, I, I, P])
    _sig(L, "ftl_swiglu_fwd", [P, P, P, I64, P])
    _sig(L, "ftl_swiglu_bwd", [P, P, P, P, P, I64, P])
    _sig(L, "ftl_embedding_fwd", [P, P, P, I64, I, I, P])
    _sig(L, "ftl_embedding_bwd", [P, P, P, P, I64, I, I, P])
    _sig(L, "ftl_cross_entropy", [P, P, P, P, P, P, I64, I, I64, I, P])
    _sig(L, "ftl_grad_norm_partials", [])
    _sig(L, "ftl_grad_norm", [P, I64, P, P, P, I, P])
    _sig(L, "ftl_adamw_step", [P, P, P, P, I64, P, P, P])
    _sig(L, "ftl_gemm_bf16", [P, I64, I, P, I64, I, P, I64, I, I, I, P, I64, P, I64, I, I, P])
    _sig(L, "ftl_gemm_swiglu", [P, I64, P, P, I64, P, P, P, I64, I, I, I, I, P])
    _sig(L, "ftl_gemm_qkv_rope", [P, I64, P, I64, P, I64, P, P, I, I, I, I, I, I, I, P])
    PP = ctypes.POINTER(c_void_p)
    _sig(L, "ftl_symm_alloc", [PP, I64])
    _sig(L, "ftl_symm_free", [P])
    _sig(L, "ftl_ipc_export", [P, ctypes.c_char_p])
    _sig(L, "ftl_ipc_import", [ctypes.c_char_p, PP])
    _sig(L, "ftl_ipc_close", [P])
    _sig(L, "ftl_comm_max_ranks", [])
    _sig(L, "ftl_comm_flag_words", [])
    _sig(L, "ftl_barrier", [P, I, I, ctypes.c_uint32, P, I64, P])
    _sig(L, "ftl_raise_abort", [P, I, P])
    _sig(L, "ftl_dp_partials", [I])
    _sig(L, "ftl_dp_reduce_scatter", [P, I, I, I64, I64, P, P, I, P])
    _sig(L, "ftl_dp_reduce_scatter_to", [P, I, I, I64, I64, P, P, P, I, P])
    _sig(L, "ftl_dp_publish", [P, I, P, P, I, I, P])
    _sig(L, "ftl_dp_finalize", [P, I, P, P, P, I, P, P, P])
    _sig(L, "ftl_dp_adamw_allgather", [P, P, P, P, I, I, I64, I64, P, P, P, P, I, P])
    _sig(L, "ftl_ckpt_init", [I64])
    _sig(L, "ftl_ckpt_write", [ctypes.c_char_p, I, P, P, P, I, I, P])
    _sig(L, "ftl_ckpt_read", [ctypes.c_char_p, I, P, P, P, I, P])
    _sig(L, "ftl_ckpt_last_error", [], ctypes.c_char_p)
    _sig(L, "ftl_gemm_qkv_rope_gather", [P, I64, P, I64, P, I64, P, P, I, I, I, I, I, I, P, I64, I, I, I, P, P])
    _sig(L, "ftl_gemm_swiglu_gather", [P, I64, P, P, I64, P, P, P, I64, I, I, I, P, I64, I64, I, I, I, P, P])
    _sig(L, "ftl_fsdp_gather", [P, I64, P, I64, I, I, P, I, P])
    _sig(L, "ftl_adamw_shard", [P, P, P, P, I64, P, P, P, P, I, P])
    _sig(L, "ftl_attn_fwd", [P, I64, P, I64, P, I, I, I, I, I, P])
    _sig(L, "ftl_attn_bwd", [P, I64, P, P, I64, P, P, P, P, P, P, I, I, I, I, I, P])
    _sig(L, "ftl_attn_bwd_scratch_bytes", [I, I, I, I, I], c_int64)
    _sig(L, "ftl_gemm_sumsq_slots", [])
    _sig(L, "ftl_gemm_bf16_ex", [P, I64, I, P, I64, I, P, I64, I, I, I, P, I64, P, I64, I, P, I, P])
    _sig(L, "ftl_gemm_swiglu_bwd", [P, I64, P, I64, P, P, P, P, I64, I, I, I, I, P])
    _sig(L, "ftl_rmsnorm_sumsq_slots", [I])
    _sig(L, "ftl_rmsnorm_bwd_ex", [P, P, P, P, P, P, P, P, P, I, I, P])
    _sig(L, "ftl_embedding_bwd_direct", [P, P, P, P, I64, I, I, P, P])
    U64P, IP = ctypes.POINTER(ctypes.c_uint64), ctypes.POINTER(c_int)
    _sig(L, "ftl_vmm_supported", [IP, IP])
    _sig(L, "ftl_vmm_round", [I64, I], c_int64)
    _sig(L, "ftl_vmm_alloc", [I64, PP, U64P, IP])
    _sig(L, "ftl_vmm_import", [I, I64, PP, U64P])
    _sig(L, "ftl_vmm_unmap", [P, I64, ctypes.c_uint64])
    _sig(L, "ftl_fd_close", [I])
    _sig(L, "ftl_mc_create", [I, I64, U64P, IP])
    _sig(L, "ftl_mc_import", [I, U64P])
    _sig(L, "ftl_mc_add_device", [ctypes.c_uint64])
    _sig(L, "ftl_mc_bind_map", [ctypes.c_uint64, ctypes.c_uint64, I64, PP])
    _sig(L, "ftl_nvls_reduce_scatter", [P, P, I64, I64, P, P, I, P])
    _sig(L, "ftl_nvls_adamw_allgather", [P, P, P, P, P, I64, I64, P, P, P, P, I, P])
    _sig(L, "ftl_emb_tokens_push", [P, P, I64, I, I, I, I, P])
    _sig(L, "ftl_emb_sparse_reduce", [P, I, I, P, P, I, I, I64, I, I, I64, I64, P, P, P])
    _sig(L, "ftl_ce_chunk_stats", [P, I64, P, I, I, I, I, P, P, P, I64, P])
    _sig(L, "ftl_ce_combine", [P, P, P, P, I, I64, I, P, P, P, P, P])
    _sig(L, "ftl_ce_chunk_grad", [P, I64, P, I, I, I, P, P, I64, P])
    _sig(L, "ftl_gemm_bf16_acc32", [P, I64, P, I64, P, P, I64, I, I, I, I, I, P])
    for name, sig in _OPTIONAL.items():
        if hasattr(L, name):
            _sig(L, name, sig[0], sig[1] if len(sig) > 1 else c_int)
    _LIB = L
    return L


# entry points added by later build stages register here: name -> (argtypes[, restype])
_OPTIONAL: dict = {}


def has(name: str) -> bool:
    return hasattr(lib(), name)


def _ptr(t: Optional[torch.Tensor]):
    return None if t is None else c_void_p(t.data_ptr())


def _stream():
    return c_void_p(torch.cuda.current_stream().cuda_stream)


def _check(code: int, what: str, kernels: int = 1) -> None:
    global LAUNCHES, KERNELS
    LAUNCHES += 1
    KERNELS += kernels
    if code != 0:
        raise NativeError(f"{what} failed with CUDA error {code}")


def _bf16(*ts):
    for t in ts:
        if t is not None:
            assert t.is_cuda and t.dtype == torch.bfloat16 and t.is_contiguous(), (t.dtype, t.shape, t.stride())


# ------------------------------------------------------------------------------------------------
def rmsnorm_fwd(x, w, eps, y, rstd=None, delta=None, h_out=None):
    """y = rmsnorm(x [+ delta]); when delta is given, h_out receives x + delta."""
    _bf16(x, w, y, delta, h_out)
    rows, dim = x.shape
    _check(lib().ftl_rmsnorm_fwd(_ptr(x), _ptr(delta), _ptr(h_out), _ptr(w), _ptr(y), _ptr(rstd), rows, dim,
                                 float(eps), _stream()), "ftl_rmsnorm_fwd")
    return y


def rmsnorm_partials(rows: int) -> int:
    return lib().ftl_rmsnorm_partials(rows)


def rmsnorm_sumsq_slots(dim: int) -> int:
    return lib().ftl_rmsnorm_sumsq_slots(dim)


def rmsnorm_bwd(dy, x, w, rstd, dx, dw, scratch, dres=None, sumsq=None):
    """sumsq: optional fp32 [rmsnorm_sumsq_slots(dim)] receiving the pieces of sum(dw^2) (gradient-norm partials)."""
    _bf16(dy, x, w, dx, dw, dres)
    rows, dim = x.shape
    assert scratch.dtype == torch.float32 and scratch.numel() >= rmsnorm_partials(rows) * dim
    assert sumsq is None or (sumsq.dtype == torch.float32 and sumsq.numel() >= rmsnorm_sumsq_slots(dim))
    _check(lib().ftl_rmsnorm_bwd_ex(_ptr(dy), _ptr(x), _ptr(w), _ptr(rstd), _ptr(dres), _ptr(dx), _ptr(dw),
                                    _ptr(scratch), _ptr(sumsq), rows, dim, _stream()), "ftl_rmsnorm_bwd", 2)
    return dx


def rope_inplace(qkv, cos, sin, seq_len, rot_cols, head_dim, inverse=False):
    """qkv: [tokens, row] bf16; rotates columns [0, rot_cols) (the q heads then the k heads)."""
    assert qkv.dtype == torch.bfloat16 and qkv.stride(1) == 1 and cos.dtype == torch.float32
    tokens, _ = qkv.shape
    _check(lib().ftl_rope_inplace(_ptr(qkv), _ptr(cos), _ptr(sin), tokens, seq_len, qkv.stride(0), rot_cols,
                                  head_dim, int(inverse), _stream()), "ftl_rope_inplace")
    return qkv


def swiglu_fwd(gate, up, act):
    _bf16(gate, up, act)
    _check(lib().ftl_swiglu_fwd(_ptr(gate), _ptr(up), _ptr(act), gate.numel(), _stream()), "ftl_swiglu_fwd")
    return act


def swiglu_bwd(dact, gate, up, dgate, dup):
    _bf16(dact, gate, up, dgate, dup)
    _check(lib().ftl_swiglu_bwd(_ptr(dact), _ptr(gate), _ptr(up), _ptr(dgate), _ptr(dup), gate.numel(), _stream()),
           "ftl_swiglu_bwd")


def embedding_fwd(tokens, table, out):
    _bf16(table, out)
    assert tokens.dtype == torch.int64 and tokens.is_contiguous()
    _check(lib().ftl_embedding_fwd(_ptr(tokens), _ptr(table), _ptr(out), tokens.numel(), table.shape[1],
                                   table.shape[0], _stream()), "ftl_embedding_fwd")
    return out


def embedding_bwd(sorted_tokens, perm, dout, dtable):
    _bf16(dout, dtable)
    _check(lib().ftl_embedding_bwd(_ptr(sorted_tokens), _ptr(perm), _ptr(dout), _ptr(dtable), sorted_tokens.numel(),
                                   dtable.shape[1], dtable.shape[0], _stream()), "ftl_embedding_bwd")


def embedding_bwd_direct(tokens, prev_tokens, dout, dtable, sumsq=None):
    """Sort-free deterministic scatter-add of dout rows into the table gradient.  prev_tokens (or None): ids whose rows
    the previous step wrote; those not written by this step are cleared.  Rows outside both sets must already be zero."""
    _bf16(dout, dtable)
    assert tokens.dtype == torch.int64 and tokens.is_contiguous()
    assert prev_tokens is None or (prev_tokens.dtype == torch.int64 and prev_tokens.numel() == tokens.numel())
    assert sumsq is None or (sumsq.dtype == torch.float32 and sumsq.numel() >= tokens.numel())
    _check(lib().ftl_embedding_bwd_direct(_ptr(tokens), _ptr(prev_tokens), _ptr(dout), _ptr(dtable), tokens.numel(),
                                          dtable.shape[1], dtable.shape[0], _ptr(sumsq), _stream()),
           "ftl_embedding_bwd_direct")


def cross_entropy(logits, labels, inv_n, row_loss, loss_out, row_lse=None, write_grad=True):
    """In place: logits [rows, V] bf16 become d(loss)/d(logits).  loss_out[0] = sum(CE) * inv_n."""
    assert logits.dtype == torch.bfloat16 and logits.stride(1) == 1 and labels.dtype == torch.int64
    rows, vocab = logits.shape
    _check(lib().ftl_cross_entropy(_ptr(logits), _ptr(labels), _ptr(inv_n), _ptr(row_loss), _ptr(row_lse),
                                   _ptr(loss_out), rows, vocab, logits.stride(0), int(write_grad), _stream()),
           "ftl_cross_entropy", 2 if loss_out is not None else 1)


def ce_chunk_stats(chunk, labels, col0, n_chunks, c, cmax, csum, picked):
    """Pass 1 of the chunked head for chunk c = logits of columns [col0, col0 + chunk.shape[1])."""
    rows, width = chunk.shape
    _check(lib().ftl_ce_chunk_stats(_ptr(chunk), chunk.stride(0), _ptr(labels), col0, width, n_chunks, c, _ptr(cmax),
                                    _ptr(csum), _ptr(picked), rows, _stream()), "ftl_ce_chunk_stats")


def ce_combine(cmax, csum, picked, labels, n_chunks, vocab, inv_n, lse, row_loss, loss_out):
    _check(lib().ftl_ce_combine(_ptr(cmax), _ptr(csum), _ptr(picked), _ptr(labels), n_chunks, labels.numel(), vocab,
                                _ptr(inv_n), _ptr(lse), _ptr(row_loss), _ptr(loss_out), _stream()), "ftl_ce_combine",
           2 if loss_out is not None else 1)


def ce_chunk_grad(chunk, labels, col0, vocab, lse, inv_n):
    """Pass 2: the recomputed chunk of logits becomes (softmax - onehot) * inv_n in place."""
    rows, width = chunk.shape
    _check(lib().ftl_ce_chunk_grad(_ptr(chunk), chunk.stride(0), _ptr(labels), col0, width, vocab, _ptr(lse), _ptr(inv_n),
                                   rows, _stream()), "ftl_ce_chunk_grad")


def gemm_acc32(a, b, c32, mode, out=None, max_ctas=0):
    """c32 (+)= a [M, K] @ b [K, N] in fp32 across calls; mode bit 0 = add to c32 (else overwrite), bit 1 = also write
    the bf16 result to `out`."""
    _mat(a), _mat(b)
    M, K = a.shape
    Nn = b.shape[1]
    assert b.shape[0] == K and c32.dtype == torch.float32 and c32.shape == (M, Nn) and c32.is_contiguous()
    assert out is None or (_mat(out) is out and out.shape == (M, Nn) and out.stride(0) == Nn)
    _check(lib().ftl_gemm_bf16_acc32(_ptr(a), a.stride(0), _ptr(b), b.stride(0), _ptr(out), _ptr(c32), Nn, M, Nn, K, mode,
                                     max_ctas, _stream()), "ftl_gemm_bf16_acc32")


def grad_norm_partials() -> int:
    return lib().ftl_grad_norm_partials()


def grad_norm(g_flat, hyper, scratch, stats, clip=True):
    _bf16(g_flat)
    _check(lib().ftl_grad_norm(_ptr(g_flat), g_flat.numel(), _ptr(hyper), _ptr(scratch), _ptr(stats), int(clip),
                               _stream()), "ftl_grad_norm", 2)


def adamw_step(p, g, m, v, hyper, stats=None):
    _bf16(p, g, m, v)
    _check(lib().ftl_adamw_step(_ptr(p), _ptr(g), _ptr(m), _ptr(v), p.numel(), _ptr(hyper), _ptr(stats), _stream()),
           "ftl_adamw_step")


# ---- tcgen05 GEMMs (csrc/gemm_tc.cu) ---------------------------------------------------------------
def _mat(t):
    assert t.is_cuda and t.dtype == torch.bfloat16 and t.dim() == 2 and t.stride(1) == 1, (t.dtype, t.shape, t.stride())
    return t


def gemm_sumsq_slots() -> int:
    return lib().ftl_gemm_sumsq_slots()


def gemm(a, b, out, a_mn=False, b_mn=False, a2=None, b2=None, max_ctas=0, sumsq=None):
    """out[M, N] = A . B^T-style product on the tcgen05 path.

    a_mn=False: a is [M, K];  a_mn=True: a is [K, M] (A = a^T).
    b_mn=False: b is [N, K];  b_mn=True: b is [K, N].
    Optional second reduction segment (a2, b2) is added: out = A.B + A2.B2.
    sumsq: optional fp32 [gemm_sumsq_slots()]; receives partial sums of squares of the stored bf16 values (their
    total is sum(out^2); entries the launch does not use are left as they are — keep them zero)."""
    _mat(a), _mat(b), _mat(out)
    M, N = out.shape
    K = a.shape[0] if a_mn else a.shape[1]
    assert (a.shape[1] if a_mn else a.shape[0]) == M and (b.shape[1] if b_mn else b.shape[0]) == N
    assert (b.shape[0] if b_mn else b.shape[1]) == K
    K2 = 0
    if a2 is not None:
        _mat(a2), _mat(b2)
        K2 = a2.shape[0] if a_mn else a2.shape[1]
    assert sumsq is None or (sumsq.dtype == torch.float32 and sumsq.numel() >= gemm_sumsq_slots())
    _check(lib().ftl_gemm_bf16_ex(_ptr(a), a.stride(0), int(a_mn), _ptr(b), b.stride(0), int(b_mn), _ptr(out),
                                  out.stride(0), M, N, K, _ptr(a2), a2.stride(0) if a2 is not None else 0, _ptr(b2),
                                  b2.stride(0) if b2 is not None else 0, K2, _ptr(sumsq), max_ctas, _stream()),
           "ftl_gemm_bf16")
    return out


def gemm_swiglu_bwd(dx, w2, gate, up, dgate, dup, max_ctas=0):
    """dgate, dup = SwiGLU backward of dact = dx @ w2, computed in the GEMM epilogue (dact is never stored).
    dx [M, K]; w2 [K, H] (the Linear's [out, in] weight); gate / up / dgate / dup [M, H], H % 16 == 0."""
    _mat(dx), _mat(w2), _mat(gate), _mat(up), _mat(dgate), _mat(dup)
    M, K = dx.shape
    H = w2.shape[1]
    assert w2.shape[0] == K and gate.shape == (M, H) and gate.stride(0) == up.stride(0) == dgate.stride(0) == dup.stride(0)
    _check(lib().ftl_gemm_swiglu_bwd(_ptr(dx), dx.stride(0), _ptr(w2), w2.stride(0), _ptr(gate), _ptr(up), _ptr(dgate),
                                     _ptr(dup), gate.stride(0), M, H, K, max_ctas, _stream()), "ftl_gemm_swiglu_bwd")


def gemm_swiglu(x, w1, w3, act, gate=None, up=None, max_ctas=0):
    """act = silu(x @ w1^T) * (x @ w3^T) in one kernel; optionally also stores the two pre-activations."""
    _mat(x), _mat(w1), _mat(w3), _mat(act)
    M, K = x.shape
    H = w1.shape[0]
    assert w1.stride(0) == w3.stride(0) and act.shape == (M, H)
    for t in (gate, up):
        assert t is None or (t.shape == act.shape and t.stride(0) == act.stride(0))
    _check(lib().ftl_gemm_swiglu(_ptr(x), x.stride(0), _ptr(w1), _ptr(w3), w1.stride(0), _ptr(act), _ptr(gate), _ptr(up),
                                 act.stride(0), M, H, K, max_ctas, _stream()), "ftl_gemm_swiglu")
    return act


def gemm_qkv_rope(x, wqkv, qkv, cos, sin, seq_len, head_dim, rot_cols, max_ctas=0):
    """qkv = x @ wqkv^T with RoPE applied to the first rot_cols columns (q heads then k heads) in the epilogue."""
    _mat(x), _mat(wqkv), _mat(qkv)
    M, K = x.shape
    Nq = wqkv.shape[0]
    _check(lib().ftl_gemm_qkv_rope(_ptr(x), x.stride(0), _ptr(wqkv), wqkv.stride(0), _ptr(qkv), qkv.stride(0), _ptr(cos),
                                   _ptr(sin), M, Nq, K, seq_len, head_dim, rot_cols, max_ctas, _stream()),
           "ftl_gemm_qkv_rope")
    return qkv


# ---- multi-GPU data plane (csrc/comm.cu) -------------------------------------------------------------
def symm_alloc(nbytes: int) -> int:
    out = c_void_p()
    code = lib().ftl_symm_alloc(ctypes.byref(out), nbytes)
    if code != 0:
        raise NativeError(f"ftl_symm_alloc({nbytes}) failed with CUDA error {code}")
    return out.value


def ipc_export(ptr: int) -> bytes:
    buf = ctypes.create_string_buffer(64)
    code = lib().ftl_ipc_export(c_void_p(ptr), buf)
    if code != 0:
        raise NativeError(f"cudaIpcGetMemHandle failed with CUDA error {code}")
    return buf.raw


def ipc_import(handle: bytes) -> int:
    out = c_void_p()
    code = lib().ftl_ipc_import(handle, ctypes.byref(out))
    if code != 0:
        raise NativeError(f"cudaIpcOpenMemHandle failed with CUDA error {code}")
    return out.value


def barrier(peer_flags, rank, world, epoch, status, timeout_ms=120000):
    _check(lib().ftl_barrier(_ptr(peer_flags), rank, world, epoch & 0xffffffff, _ptr(status), timeout_ms, _stream()),
           "ftl_barrier")


def raise_abort(peer_flags, world):
    _check(lib().ftl_raise_abort(_ptr(peer_flags), world, _stream()), "ftl_raise_abort")


def dp_partials(max_ctas=0) -> int:
    return lib().ftl_dp_partials(max_ctas)


def dp_reduce_scatter(peer_g, rank, world, lo, n, partial, status, max_ctas=0):
    _check(lib().ftl_dp_reduce_scatter(_ptr(peer_g), rank, world, lo, n, _ptr(partial), _ptr(status), max_ctas, _stream()),
           "ftl_dp_reduce_scatter")


def emb_tokens_push(tokens, peer_toks, slot_off, rank, world, cap):
    """Store this rank's token ids of the step into every peer's id buffer (slot_off in elements)."""
    assert tokens.dtype == torch.int64 and tokens.is_contiguous()
    _check(lib().ftl_emb_tokens_push(_ptr(tokens), _ptr(peer_toks), slot_off, rank, world, tokens.numel(), cap, _stream()),
           "ftl_emb_tokens_push")


def emb_sparse_reduce(peer_g, rank, world, toks, toks_prev, T, cap, emb_off, dim, vocab, lo, hi, partial, status):
    """Reduce-scatter of the embedding-table gradient restricted to the rows that occur in any rank's token list."""
    assert partial.dtype == torch.float32 and partial.numel() >= world * T
    _check(lib().ftl_emb_sparse_reduce(_ptr(peer_g), rank, world, _ptr(toks), _ptr(toks_prev), T, cap, emb_off, dim, vocab,
                                       lo, hi, _ptr(partial), _ptr(status), _stream()), "ftl_emb_sparse_reduce")


def dp_reduce_scatter_to(peer_g, rank, world, lo, n, dst, partial, status, max_ctas=0):
    """Like dp_reduce_scatter, but the reduced range lands in the compact buffer `dst` (n elements)."""
    _bf16(dst)
    assert dst.numel() >= n
    _check(lib().ftl_dp_reduce_scatter_to(_ptr(peer_g), rank, world, lo, n, _ptr(dst), _ptr(partial), _ptr(status), max_ctas,
                                          _stream()), "ftl_dp_reduce_scatter_to")


def dp_publish(partial, nparts, loss_local, peer_slots, rank, world):
    _check(lib().ftl_dp_publish(_ptr(partial), nparts, _ptr(loss_local), _ptr(peer_slots), rank, world, _stream()),
           "ftl_dp_publish")


def dp_finalize(slots, world, hyper, stats, loss_out, clip=True, status=None, result=None):
    """status (optional): the sticky comm status word; a non-finite norm sets it to 3 so that no later update runs.
    result (optional): fp32[>=5] receiving loss, norm, clip coefficient, finite flag, status — the step's report."""
    _check(lib().ftl_dp_finalize(_ptr(slots), world, _ptr(hyper), _ptr(stats), _ptr(loss_out), int(clip), _ptr(status),
                                 _ptr(result), _stream()), "ftl_dp_finalize")


def dp_adamw_allgather(peer_p, g, m, v, rank, world, lo, n, hyper, stats, status, commits=None, max_ctas=0):
    _check(lib().ftl_dp_adamw_allgather(_ptr(peer_p), _ptr(g), _ptr(m), _ptr(v), rank, world, lo, n, _ptr(hyper),
                                        _ptr(stats), _ptr(status), _ptr(commits), max_ctas, _stream()),
           "ftl_dp_adamw_allgather")


# ---- VMM symmetric heap + NVLS multicast (csrc/nvls.cu) ------------------------------------------------------
def _drv(code: int, what: str) -> None:
    if code != 0:
        kind = f"CUDA driver error {code - 1000}" if code >= 1000 else f"CUDA error {code}"
        raise NativeError(f"{what} failed with {kind}")


def vmm_supported():
    """(vmm, multicast): the virtual-memory API with fd export / multicast objects are usable on the current device."""
    a, b = c_int(0), c_int(0)
    _drv(lib().ftl_vmm_supported(ctypes.byref(a), ctypes.byref(b)), "ftl_vmm_supported")
    return bool(a.value), bool(b.value)


def vmm_round(nbytes: int, world: int) -> int:
    r = lib().ftl_vmm_round(nbytes, world)
    if r < 0:
        raise NativeError("ftl_vmm_round: the CUDA virtual-memory API is not available")
    return r


def vmm_alloc(nbytes: int):
    """-> (device pointer, allocation handle, exportable file descriptor)"""
    ptr, h, fd = c_void_p(), ctypes.c_uint64(0), c_int(-1)
    _drv(lib().ftl_vmm_alloc(nbytes, ctypes.byref(ptr), ctypes.byref(h), ctypes.byref(fd)), f"ftl_vmm_alloc({nbytes})")
    return ptr.value, h.value, fd.value


def vmm_import(fd: int, nbytes: int):
    ptr, h = c_void_p(), ctypes.c_uint64(0)
    _drv(lib().ftl_vmm_import(fd, nbytes, ctypes.byref(ptr), ctypes.byref(h)), "ftl_vmm_import")
    return ptr.value, h.value


def vmm_unmap(ptr: int, nbytes: int, handle: int) -> None:
    _drv(lib().ftl_vmm_unmap(c_void_p(ptr), nbytes, ctypes.c_uint64(handle)), "ftl_vmm_unmap")


def mc_create(world: int, nbytes: int):
    h, fd = ctypes.c_uint64(0), c_int(-1)
    _drv(lib().ftl_mc_create(world, nbytes, ctypes.byref(h), ctypes.byref(fd)), "cuMulticastCreate")
    return h.value, fd.value


def mc_import(fd: int) -> int:
    h = ctypes.c_uint64(0)
    _drv(lib().ftl_mc_import(fd, ctypes.byref(h)), "multicast import")
    return h.value


def mc_add_device(mc: int) -> None:
    _drv(lib().ftl_mc_add_device(ctypes.c_uint64(mc)), "cuMulticastAddDevice")


def mc_bind_map(mc: int, mem: int, nbytes: int) -> int:
    ptr = c_void_p()
    _drv(lib().ftl_mc_bind_map(ctypes.c_uint64(mc), ctypes.c_uint64(mem), nbytes, ctypes.byref(ptr)), "cuMulticastBindMem / map")
    return ptr.value


def nvls_reduce_scatter(mc_g: int, g_local, lo, n, partial, status, max_ctas=0):
    """Reduce-scatter of shard [lo, lo+n) through the switch: multimem.ld_reduce from the multicast address mc_g."""
    _check(lib().ftl_nvls_reduce_scatter(c_void_p(mc_g), _ptr(g_local), lo, n, _ptr(partial), _ptr(status), max_ctas,
                                         _stream()), "ftl_nvls_reduce_scatter")


def nvls_adamw_allgather(mc_p: int, p_local, g, m, v, lo, n, hyper, stats, status, commits=None, max_ctas=0):
    """clip + AdamW on [lo, lo+n) with the all-gather as ONE multimem.st per vector through the multicast address mc_p."""
    _check(lib().ftl_nvls_adamw_allgather(c_void_p(mc_p), _ptr(p_local), _ptr(g), _ptr(m), _ptr(v), lo, n, _ptr(hyper),
                                          _ptr(stats), _ptr(status), _ptr(commits), max_ctas, _stream()),
           "ftl_nvls_adamw_allgather")


# ---- checkpoint snapshot engine (csrc/ckpt_io.cu) -----------------------------------------------------
def ckpt_init(stage_bytes: int = 256 << 20) -> None:
    code = lib().ftl_ckpt_init(stage_bytes)
    if code != 0:
        raise NativeError(f"ftl_ckpt_init failed: {lib().ftl_ckpt_last_error().decode()}")


def _seg_arrays(segments):
    n = len(segments)
    ptrs = (c_void_p * n)(*[c_void_p(p) for p, _, _ in segments])
    sizes = (c_int64 * n)(*[s for _, s, _ in segments])
    offs = (c_int64 * n)(*[o for _, _, o in segments])
    return n, ptrs, sizes, offs


def ckpt_write(path: str, segments, nthreads: int = 4, fsync: bool = True) -> dict:
    """segments: list of (device_ptr, nbytes, file_offset).  Blocks until the data is written (and fsynced)."""
    n, ptrs, sizes, offs = _seg_arrays(segments)
    t = (ctypes.c_double * 4)()
    code = lib().ftl_ckpt_write(path.encode(), n, ptrs, sizes, offs, nthreads, int(fsync), t)
    if code != 0:
        raise NativeError(f"ftl_ckpt_write({path}) failed: {lib().ftl_ckpt_last_error().decode()}")
    return {"d2h_s": t[0], "write_s": t[1], "fsync_s": t[2], "total_s": t[3]}


def ckpt_read(path: str, segments, nthreads: int = 4) -> dict:
    n, ptrs, sizes, offs = _seg_arrays(segments)
    t = (ctypes.c_double * 4)()
    code = lib().ftl_ckpt_read(path.encode(), n, ptrs, sizes, offs, nthreads, t)
    if code != 0:
        raise NativeError(f"ftl_ckpt_read({path}) failed: {lib().ftl_ckpt_last_error().decode()}")
    return {"total_s": t[3]}


# ---- fully-sharded path (csrc/gemm_tc.cu gather variants, csrc/comm.cu) -----------------------------------
def gemm_qkv_rope_gather(x, wqkv, qkv, cos, sin, seq_len, head_dim, rot_cols, peer_shards, shard_off, rows_per_rank,
                         rank, world, scratch):
    """QKV projection + RoPE whose weight operand is all-gathered from the ranks' shards inside the kernel.
    wqkv is the local [N, K] staging matrix the kernel fills while it computes."""
    _mat(x), _mat(wqkv), _mat(qkv)
    M, K = x.shape
    Nq = wqkv.shape[0]
    assert wqkv.is_contiguous() and scratch.dtype == torch.int32 and scratch.numel() >= 1 + Nq // 128
    _check(lib().ftl_gemm_qkv_rope_gather(_ptr(x), x.stride(0), _ptr(wqkv), wqkv.stride(0), _ptr(qkv), qkv.stride(0),
                                          _ptr(cos), _ptr(sin), M, Nq, K, seq_len, head_dim, rot_cols, _ptr(peer_shards),
                                          shard_off, rows_per_rank, rank, world, _ptr(scratch), _stream()),
           "ftl_gemm_qkv_rope_gather")
    return qkv


def gemm_swiglu_gather(x, w1, w3, act, gate, up, peer_shards, off_w1, off_w3, rows_per_rank, rank, world, scratch):
    _mat(x), _mat(w1), _mat(w3), _mat(act)
    M, K = x.shape
    H = w1.shape[0]
    assert w1.is_contiguous() and w3.is_contiguous() and scratch.numel() >= 1 + 2 * (H // 128)
    _check(lib().ftl_gemm_swiglu_gather(_ptr(x), x.stride(0), _ptr(w1), _ptr(w3), w1.stride(0), _ptr(act), _ptr(gate),
                                        _ptr(up), act.stride(0), M, H, K, _ptr(peer_shards), off_w1, off_w3, rows_per_rank,
                                        rank, world, _ptr(scratch), _stream()), "ftl_gemm_swiglu_gather")
    return act


def fsdp_gather(peer_shards, shard_off, dst, n_per_rank, rank, world, status, max_ctas=0):
    """dst[r * n : (r + 1) * n] = rank r's shard, for every r (plain parameter all-gather)."""
    _bf16(dst)
    _check(lib().ftl_fsdp_gather(_ptr(peer_shards), shard_off, _ptr(dst), n_per_rank, rank, world, _ptr(status), max_ctas,
                                 _stream()), "ftl_fsdp_gather")


def adamw_shard(p, g, m, v, hyper, stats, status, commits=None, max_ctas=0):
    _bf16(p, g, m, v)
    _check(lib().ftl_adamw_shard(_ptr(p), _ptr(g), _ptr(m), _ptr(v), p.numel(), _ptr(hyper), _ptr(stats), _ptr(status),
                                 _ptr(commits), max_ctas, _stream()), "ftl_adamw_shard")


# ---- tcgen05 flash attention (csrc/attn_tc.cu) -----------------------------------------------------------
def attn_fwd(qkv, out, lse, B, S, H, KV, hd):
    """Causal GQA attention read in place from the fused [B*S, (H+2KV)*hd] projection (RoPE already applied).
    out: [B*S, H*hd] bf16; lse: fp32 [B, H, S] (log2 domain), kept for the backward."""
    _mat(qkv), _mat(out)
    assert lse.dtype == torch.float32 and lse.numel() == B * H * S
    _check(lib().ftl_attn_fwd(_ptr(qkv), qkv.stride(0), _ptr(out), out.stride(0), _ptr(lse), B, S, H, KV, hd, _stream()),
           "ftl_attn_fwd")
    return out


_ATTN_SCRATCH = {}


def attn_bwd_scratch(B, S, H, KV, hd, device) -> torch.Tensor:
    """fp32 work area of the backward (chunk partials of dK/dV); cached per shape."""
    key = (B, S, H, KV, hd, str(device))
    if key not in _ATTN_SCRATCH:
        n = lib().ftl_attn_bwd_scratch_bytes(B, S, H, KV, hd) // 4
        _ATTN_SCRATCH[key] = torch.empty(max(n, 1), dtype=torch.float32, device=device)
    return _ATTN_SCRATCH[key]


def attn_bwd(qkv, out, d_out, lse, delta, dqkv, B, S, H, KV, hd, cos=None, sin=None, scratch=None):
    """Deterministic backward: writes dq | dk | dv into dqkv ([B*S, (H+2KV)*hd]); with cos/sin the inverse rotary
    embedding is applied to dq and dk, so dqkv is the gradient of the un-rotated projection."""
    _mat(qkv), _mat(out), _mat(d_out), _mat(dqkv)
    assert out.stride(0) == d_out.stride(0) and dqkv.stride(0) == qkv.stride(0) and delta.numel() == B * H * S
    if scratch is None:
        scratch = attn_bwd_scratch(B, S, H, KV, hd, qkv.device)
    _check(lib().ftl_attn_bwd(_ptr(qkv), qkv.stride(0), _ptr(out), _ptr(d_out), out.stride(0), _ptr(lse), _ptr(delta),
                              _ptr(dqkv), _ptr(cos), _ptr(sin), _ptr(scratch), B, S, H, KV, hd, _stream()), "ftl_attn_bwd", 4)
    return dqkv
