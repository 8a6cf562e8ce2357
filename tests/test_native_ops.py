"""T1: each sm_100a kernel against a plain fp32 PyTorch statement of the same op."""
import math

import pytest
import torch

from fault_tolerant_llm_training_b200.ops import torch_ops as T

pytestmark = pytest.mark.gpu

DEV = "cuda"


@pytest.fixture(scope="module")
def N():
    from fault_tolerant_llm_training_b200.ops import native
    native.lib()
    return native


def bf(*shape, scale=1.0, seed=0):
    g = torch.Generator(device=DEV).manual_seed(seed)
    return (torch.randn(*shape, device=DEV, generator=g) * scale).to(torch.bfloat16)


@pytest.mark.parametrize("rows,dim", [(64, 256), (300, 4096), (17, 1024), (5, 8192)])
def test_rmsnorm_fwd_matches_reference_rounding(N, rows, dim):
    x, w = bf(rows, dim, seed=1), bf(dim, seed=2) * 0.1 + 1
    y, rstd = torch.empty_like(x), torch.empty(rows, device=DEV)
    N.rmsnorm_fwd(x, w, 1e-5, y, rstd)
    ref = T.rmsnorm(x, w, 1e-5)
    torch.testing.assert_close(y.float(), ref.float(), rtol=1.6e-2, atol=1e-2)
    assert (y != ref).float().mean() < 0.02  # same rounding points: almost always bit-equal
    torch.testing.assert_close(rstd, torch.rsqrt(x.float().pow(2).mean(-1) + 1e-5), rtol=1e-5, atol=1e-6)
    # fused residual add
    d = bf(rows, dim, seed=3)
    h, y2 = torch.empty_like(x), torch.empty_like(x)
    N.rmsnorm_fwd(x, w, 1e-5, y2, rstd, delta=d, h_out=h)
    assert torch.equal(h, x + d)
    torch.testing.assert_close(y2.float(), T.rmsnorm(x + d, w, 1e-5).float(), rtol=1.6e-2, atol=1e-2)


@pytest.mark.parametrize("rows,dim", [(64, 256), (2048, 4096), (301, 1024)])
def test_rmsnorm_bwd(N, rows, dim):
    x, w, dy, dres = bf(rows, dim, seed=1), bf(dim, seed=2) * 0.1 + 1, bf(rows, dim, seed=3), bf(rows, dim, seed=4)
    xf = x.float().requires_grad_()
    wf = w.float().requires_grad_()
    inv = torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + 1e-5)
    (xf * inv * wf).backward(dy.float())
    rstd = inv.detach().squeeze(-1).contiguous()
    dx, dw = torch.empty_like(x), torch.empty_like(w)
    scratch = torch.empty(N.rmsnorm_partials(rows) * dim, device=DEV)
    N.rmsnorm_bwd(dy, x, w, rstd, dx, dw, scratch)
    torch.testing.assert_close(dx.float(), xf.grad, rtol=2e-2, atol=2e-2)
    torch.testing.assert_close(dw.float(), wf.grad, rtol=2e-2, atol=2e-2 * math.sqrt(rows))
    N.rmsnorm_bwd(dy, x, w, rstd, dx, dw, scratch, dres=dres)
    torch.testing.assert_close(dx.float(), xf.grad + dres.float(), rtol=2e-2, atol=3e-2)


@pytest.mark.parametrize("B,S,H,KV,hd", [(2, 64, 4, 2, 64), (1, 2048, 32, 8, 128)])
def test_rope_inplace_and_inverse(N, B, S, H, KV, hd):
    width = (H + 2 * KV) * hd
    qkv = bf(B * S, width, seed=5)
    orig = qkv.clone()
    cos, sin = T.rope_table(hd, S, 500000.0, device=DEV)
    N.rope_inplace(qkv, cos, sin, S, (H + KV) * hd, hd)
    q = orig[:, :H * hd].view(B, S, H, hd)
    k = orig[:, H * hd:(H + KV) * hd].view(B, S, KV, hd)
    # fp32 math rounded once to bf16 on both sides; nvcc contracts a*c - b*s into an FMA, so allow one bf16 ulp
    for got, want in ((qkv[:, :H * hd].view(B, S, H, hd), T.rope_apply(q, cos, sin)),
                      (qkv[:, H * hd:(H + KV) * hd].view(B, S, KV, hd), T.rope_apply(k, cos, sin))):
        torch.testing.assert_close(got.float(), want.float(), rtol=8e-3, atol=1e-3)
        assert (got != want).float().mean() < 0.01
    assert torch.equal(qkv[:, (H + KV) * hd:], orig[:, (H + KV) * hd:])  # v untouched
    N.rope_inplace(qkv, cos, sin, S, (H + KV) * hd, hd, inverse=True)
    torch.testing.assert_close(qkv.float(), orig.float(), rtol=2e-2, atol=2e-2)


def test_swiglu_fwd_bwd(N):
    g, u, d = bf(512, 1024, seed=1), bf(512, 1024, seed=2), bf(512, 1024, seed=3)
    act = torch.empty_like(g)
    N.swiglu_fwd(g, u, act)
    ref = T.swiglu(g, u)
    torch.testing.assert_close(act.float(), ref.float(), rtol=1.6e-2, atol=1e-2)
    gf, uf = g.float().requires_grad_(), u.float().requires_grad_()
    (torch.nn.functional.silu(gf) * uf).backward(d.float())
    dg, du = torch.empty_like(g), torch.empty_like(g)
    N.swiglu_bwd(d, g, u, dg, du)
    torch.testing.assert_close(dg.float(), gf.grad, rtol=2e-2, atol=2e-2)
    torch.testing.assert_close(du.float(), uf.grad, rtol=2e-2, atol=2e-2)


def test_embedding_fwd_bwd_deterministic(N):
    V, D, n = 1000, 256, 4096
    table = bf(V, D, seed=1)
    tok = torch.randint(0, V, (n,), device=DEV, generator=torch.Generator(device=DEV).manual_seed(3))
    out = torch.empty(n, D, device=DEV, dtype=torch.bfloat16)
    N.embedding_fwd(tok, table, out)
    assert torch.equal(out, table[tok])
    dout = bf(n, D, seed=2)
    ref = torch.zeros(V, D, device=DEV).index_add_(0, tok, dout.float())
    runs = []
    for _ in range(2):
        dt = torch.zeros(V, D, device=DEV, dtype=torch.bfloat16)
        s, p = torch.sort(tok, stable=True)
        N.embedding_bwd(s, p, dout, dt)
        runs.append(dt)
    torch.testing.assert_close(runs[0].float(), ref, rtol=2e-2, atol=3e-2)
    assert torch.equal(runs[0], runs[1])


@pytest.mark.parametrize("rows,V", [(64, 1024), (256, 131072)])
def test_cross_entropy_fwd_bwd(N, rows, V):
    logits = bf(rows, V, scale=2.0, seed=7)
    labels = torch.randint(0, V, (rows,), device=DEV)
    labels[::5] = -100
    n = int((labels != -100).sum())
    lf = logits.float().requires_grad_()
    ref_loss = torch.nn.functional.cross_entropy(lf, labels, reduction="sum") / n
    ref_loss.backward()
    inv_n = torch.tensor([1.0 / n], device=DEV)
    row_loss, loss = torch.empty(rows, device=DEV), torch.empty(1, device=DEV)
    work = logits.clone()
    N.cross_entropy(work, labels, inv_n, row_loss, loss)
    torch.testing.assert_close(loss[0], ref_loss.detach(), rtol=1e-4, atol=1e-4)
    torch.testing.assert_close(work.float(), lf.grad, rtol=2e-2, atol=1e-6)
    assert (work[labels == -100] == 0).all()


def test_grad_norm_and_adamw_match_torch(N):
    n = 1 << 20
    p, g = bf(n, seed=1), bf(n, scale=0.01, seed=2)
    m, v = torch.zeros_like(p), torch.zeros_like(p)
    ref_p = torch.nn.Parameter(p.float().clone())
    opt = torch.optim.AdamW([ref_p], lr=1e-2)
    scratch, stats = torch.empty(N.grad_norm_partials(), device=DEV), torch.zeros(4, device=DEV)
    for t in range(1, 4):
        hyper = torch.tensor([1e-2, 0.9, 0.999, 1e-8, 1e-2, 1 - 0.9 ** t, math.sqrt(1 - 0.999 ** t), 1.0], device=DEV)
        N.grad_norm(g, hyper, scratch, stats, clip=True)
        norm = g.float().norm()
        torch.testing.assert_close(stats[0], norm, rtol=1e-4, atol=1e-5)
        coef = min(1.0, 1.0 / (float(norm) + 1e-6))
        assert abs(float(stats[1]) - coef) < 1e-4 and float(stats[2]) == 1.0
        N.adamw_step(p, g, m, v, hyper, stats)
        ref_p.grad = g.float() * coef
        opt.step()
    # bf16 state vs fp32 reference: a few ulps of bf16
    torch.testing.assert_close(p.float(), ref_p.detach(), rtol=2e-2, atol=2e-2)
    st = opt.state[ref_p]
    torch.testing.assert_close(m.float(), st["exp_avg"], rtol=3e-2, atol=1e-4)
    torch.testing.assert_close(v.float(), st["exp_avg_sq"], rtol=3e-2, atol=1e-7)
    # a non-finite gradient makes the update a no-op
    g2 = g.clone(); g2[5] = float("inf")
    before = p.clone()
    N.grad_norm(g2, hyper, scratch, stats, clip=True)
    N.adamw_step(p, g2, m, v, hyper, stats)
    assert float(stats[2]) == 0.0 and torch.equal(p, before)


def test_attention_tc_model_shape_vs_sdpa(N):
    """Model shape (32 query / 8 kv heads, head_dim 128, S = 2048): more work items than SMs, so the dynamic work queues,
    the two-tile forward with lazy rescale and the chunked dK/dV units with their reduction all run.  Oracle: the library
    SDPA on the same bf16 inputs (the fp32 einsum oracle of the small cases would need 2 GB of scores)."""
    import torch.nn.functional as F
    B, S, H, KV, hd = 1, 2048, 32, 8, 128
    W = (H + 2 * KV) * hd
    qkv, dout = bf(B * S, W, scale=0.7, seed=11), bf(B * S, H * hd, scale=0.5, seed=12)
    out = torch.empty(B * S, H * hd, device=DEV, dtype=torch.bfloat16)
    lse, delta = torch.empty(B * H * S, device=DEV), torch.empty(B * H * S, device=DEV)
    dqkv = torch.zeros_like(qkv)
    N.attn_fwd(qkv, out, lse, B, S, H, KV, hd)
    N.attn_bwd(qkv, out, dout, lse, delta, dqkv, B, S, H, KV, hd)
    x = qkv.view(B, S, H + 2 * KV, hd)
    q, k, v = (t.transpose(1, 2).clone().requires_grad_() for t in (x[:, :, :H], x[:, :, H:H + KV], x[:, :, H + KV:]))
    ref = F.scaled_dot_product_attention(q, k.repeat_interleave(H // KV, 1), v.repeat_interleave(H // KV, 1), is_causal=True)
    ref.backward(dout.view(B, S, H, hd).transpose(1, 2))
    d4 = dqkv.view(B, S, -1, hd)
    assert _rel(out.view(B, S, H, hd), ref.detach().transpose(1, 2).float()) < 1e-2
    for got, want in ((d4[:, :, :H], q.grad), (d4[:, :, H:H + KV], k.grad), (d4[:, :, H + KV:], v.grad)):
        assert _rel(got, want.transpose(1, 2).float()) < 2e-2
    out2, dq2 = torch.empty_like(out), torch.zeros_like(dqkv)
    for _ in range(2):                                   # the queues reset themselves: a second and third launch agree bitwise
        N.attn_fwd(qkv, out2, lse, B, S, H, KV, hd)
        N.attn_bwd(qkv, out2, dout, lse, delta, dq2, B, S, H, KV, hd)
        assert torch.equal(out, out2) and torch.equal(dqkv, dq2)


def test_adamw_untouched_rows_exact(N):
    """Elements with g = m = v = 0 (embedding rows no batch has touched) only see the weight decay: the clamp that keeps
    sqrtf() off its zero/subnormal slow path (adam_vmin, csrc/common/util.cuh) must not change a bit of the result."""
    n = 1 << 16
    p = bf(n, seed=3)
    g, m, v = torch.zeros_like(p), torch.zeros_like(p), torch.zeros_like(p)
    v[: n // 2] = 1e-38                                   # tiny but non-zero second moments: below the clamp as well
    hyper = torch.tensor([1e-2, 0.9, 0.999, 1e-8, 1e-2, 1 - 0.9, math.sqrt(1 - 0.999), 1.0], device=DEV)
    want = (p.float() * (1.0 - 1e-2 * 1e-2)).to(torch.bfloat16)
    N.adamw_step(p, g, m, v, hyper, None)
    assert torch.equal(p, want) and not m.any()


# ---- tcgen05 GEMMs (csrc/gemm_tc.cu): every operand-major combination and fused epilogue vs fp32 matmul ----
def _rel(out, ref):
    return float((out.float() - ref).norm() / ref.norm())


@pytest.mark.parametrize("M,Nn,K", [(256, 512, 256), (200, 344, 136), (2048, 1024, 4096)])
def test_gemm_tc_nt(N, M, Nn, K):
    a, b = bf(M, K, scale=0.5, seed=1), bf(Nn, K, scale=0.5, seed=2)
    out = torch.empty(M, Nn, device=DEV, dtype=torch.bfloat16)
    N.gemm(a, b, out)
    assert _rel(out, a.float() @ b.float().t()) < 5e-3


def test_gemm_tc_dgrad_wgrad_and_split_reduction(N):
    M, Nn, K = 384, 512, 320
    dy, w, x = bf(M, Nn, scale=0.5, seed=1), bf(Nn, K, scale=0.5, seed=2), bf(M, K, scale=0.5, seed=3)
    dx = torch.empty(M, K, device=DEV, dtype=torch.bfloat16)
    N.gemm(dy, w, dx, b_mn=True)                       # dgrad: dY @ W
    assert _rel(dx, dy.float() @ w.float()) < 5e-3
    dw = torch.empty(Nn, K, device=DEV, dtype=torch.bfloat16)
    N.gemm(dy, x, dw, a_mn=True, b_mn=True)            # wgrad: dY^T @ X
    assert _rel(dw, dy.float().t() @ x.float()) < 5e-3
    dy2, w2 = bf(M, Nn, scale=0.5, seed=4), bf(Nn, K, scale=0.5, seed=5)
    N.gemm(dy, w, dx, b_mn=True, a2=dy2, b2=w2)        # dY @ W + dY2 @ W2 in one pass
    assert _rel(dx, dy.float() @ w.float() + dy2.float() @ w2.float()) < 5e-3


def test_gemm_tc_swiglu_epilogue(N):
    M, H, K = 256, 384, 256
    x, w1, w3 = bf(M, K, scale=0.5, seed=1), bf(H, K, scale=0.1, seed=2), bf(H, K, scale=0.1, seed=3)
    act, gate, up = (torch.empty(M, H, device=DEV, dtype=torch.bfloat16) for _ in range(3))
    N.gemm_swiglu(x, w1, w3, act, gate, up)
    g, u = x.float() @ w1.float().t(), x.float() @ w3.float().t()
    assert _rel(gate, g) < 5e-3 and _rel(up, u) < 5e-3
    assert _rel(act, torch.nn.functional.silu(g) * u) < 1e-2


def test_gemm_tc_qkv_rope_epilogue(N):
    B, S, H, KV, hd, K = 2, 128, 4, 2, 64, 256
    M, Nq, rot = B * S, (H + 2 * KV) * hd, (H + KV) * hd
    x, w = bf(M, K, scale=0.5, seed=1), bf(Nq, K, scale=0.1, seed=2)
    cos, sin = T.rope_table(hd, S, 500000.0, device=DEV)
    out = torch.empty(M, Nq, device=DEV, dtype=torch.bfloat16)
    N.gemm_qkv_rope(x, w, out, cos, sin, S, hd, rot)
    ref = (x.float() @ w.float().t()).to(torch.bfloat16)
    ref_rot = T.rope_apply(ref[:, :rot].reshape(B, S, H + KV, hd), cos, sin).reshape(M, rot)
    assert _rel(out[:, :rot], ref_rot.float()) < 5e-3
    assert _rel(out[:, rot:], ref[:, rot:].float()) < 5e-3


# ---- tcgen05 flash attention (csrc/attn_tc.cu) vs the explicit fp32 softmax(QK^T)V oracle and its autograd ----
@pytest.mark.parametrize("B,S,H,KV,hd", [(2, 256, 4, 2, 64), (1, 384, 4, 2, 128), (1, 256, 2, 2, 128)])
def test_attention_tc_fwd_bwd_deterministic(N, B, S, H, KV, hd):
    W = (H + 2 * KV) * hd
    qkv, dout = bf(B * S, W, scale=0.7, seed=1), bf(B * S, H * hd, scale=0.5, seed=2)
    out = torch.empty(B * S, H * hd, device=DEV, dtype=torch.bfloat16)
    lse, delta = torch.empty(B * H * S, device=DEV), torch.empty(B * H * S, device=DEV)
    dqkv = torch.zeros_like(qkv)
    N.attn_fwd(qkv, out, lse, B, S, H, KV, hd)
    N.attn_bwd(qkv, out, dout, lse, delta, dqkv, B, S, H, KV, hd)
    qf = qkv.float().view(B, S, H + 2 * KV, hd)
    q, k, v = (t.clone().requires_grad_() for t in (qf[:, :, :H], qf[:, :, H:H + KV], qf[:, :, H + KV:]))
    ref = T.attention_fp32_oracle(q, k, v)
    ref.backward(dout.float().view(B, S, H, hd))
    d4 = dqkv.view(B, S, -1, hd)
    assert _rel(out.view(B, S, H, hd), ref.detach()) < 1e-2
    assert _rel(d4[:, :, :H], q.grad) < 1e-2 and _rel(d4[:, :, H:H + KV], k.grad) < 1e-2 and _rel(d4[:, :, H + KV:], v.grad) < 1e-2
    # natural-log lse of the scaled scores
    sc = torch.einsum("bqhd,bkhd->bhqk", q.detach(), k.detach().repeat_interleave(H // KV, dim=2)) / hd ** 0.5
    sc = sc.masked_fill(~torch.ones(S, S, dtype=torch.bool, device=DEV).tril(), float("-inf"))
    torch.testing.assert_close(lse.view(B, H, S) * 0.6931471805599453, torch.logsumexp(sc, -1), rtol=1e-3, atol=1e-3)
    # bitwise repeatable, and the fused inverse RoPE equals rotating the gradients afterwards
    out2, dq2 = torch.empty_like(out), torch.zeros_like(dqkv)
    N.attn_fwd(qkv, out2, lse, B, S, H, KV, hd)
    N.attn_bwd(qkv, out2, dout, lse, delta, dq2, B, S, H, KV, hd)
    assert torch.equal(out, out2) and torch.equal(dqkv, dq2)
    cos, sin = T.rope_table(hd, S, 500000.0, device=DEV)
    dq3 = torch.zeros_like(dqkv)
    N.attn_bwd(qkv, out, dout, lse, delta, dq3, B, S, H, KV, hd, cos, sin)
    want = dqkv.clone()
    N.rope_inplace(want, cos, sin, S, (H + KV) * hd, hd, inverse=True)
    torch.testing.assert_close(dq3.float(), want.float(), rtol=2e-2, atol=2e-3)


@pytest.mark.parametrize("M,H,K", [(256, 512, 256), (200, 336, 192), (2048, 14336, 4096)])
def test_gemm_tc_swiglu_bwd_epilogue(N, M, H, K):
    """w2 dgrad GEMM with the SwiGLU backward in its epilogue == dgrad GEMM followed by the elementwise kernel."""
    dx, w2 = bf(M, K, scale=0.5, seed=1), bf(K, H, scale=0.1, seed=2)
    gate, up = bf(M, H, seed=3), bf(M, H, seed=4)
    dg, du = torch.empty_like(gate), torch.empty_like(gate)
    N.gemm_swiglu_bwd(dx, w2, gate, up, dg, du)
    dact = dx.float() @ w2.float()
    gf, uf = gate.float().requires_grad_(), up.float().requires_grad_()
    (torch.nn.functional.silu(gf) * uf).backward(dact)
    assert _rel(dg, gf.grad) < 1e-2 and _rel(du, uf.grad) < 1e-2
    # and bit-equal to the two-kernel path it replaces (same rounding points)
    dact_b, dg2, du2 = torch.empty_like(gate), torch.empty_like(gate), torch.empty_like(gate)
    N.gemm(dx, w2, dact_b, b_mn=True)
    N.swiglu_bwd(dact_b, gate, up, dg2, du2)
    assert torch.equal(dg, dg2) and torch.equal(du, du2)


@pytest.mark.parametrize("M,Nn,K", [(512, 768, 256), (200, 344, 136), (4096, 4096, 2048)])
def test_gemm_tc_sumsq_partials(N, M, Nn, K):
    """The weight-gradient GEMM also returns sum(out^2) in deterministic pieces."""
    dy, x = bf(K, M, scale=0.5, seed=1), bf(K, Nn, scale=0.5, seed=2)
    out = torch.empty(M, Nn, device=DEV, dtype=torch.bfloat16)
    parts = torch.zeros(N.gemm_sumsq_slots(), device=DEV)
    N.gemm(dy, x, out, a_mn=True, b_mn=True, sumsq=parts)
    ref = out.float().pow(2).sum()
    torch.testing.assert_close(parts.sum(), ref, rtol=1e-4, atol=1e-3)
    parts2 = torch.zeros_like(parts)
    N.gemm(dy, x, out, a_mn=True, b_mn=True, sumsq=parts2)
    assert torch.equal(parts, parts2)


def test_rmsnorm_bwd_sumsq(N):
    rows, dim = 300, 4096
    x, w, dy = bf(rows, dim, seed=1), bf(dim, seed=2) * 0.1 + 1, bf(rows, dim, seed=3)
    rstd = torch.rsqrt(x.float().pow(2).mean(-1) + 1e-5)
    dx, dw = torch.empty_like(x), torch.empty_like(w)
    scratch = torch.empty(N.rmsnorm_partials(rows) * dim, device=DEV)
    parts = torch.zeros(N.rmsnorm_sumsq_slots(dim), device=DEV)
    N.rmsnorm_bwd(dy, x, w, rstd, dx, dw, scratch, sumsq=parts)
    torch.testing.assert_close(parts.sum(), dw.float().pow(2).sum(), rtol=1e-5, atol=1e-5)
    dw2 = torch.empty_like(w)
    N.rmsnorm_bwd(dy, x, w, rstd, dx, dw2, scratch)
    assert torch.equal(dw, dw2)


@pytest.mark.parametrize("V,D,n", [(1000, 256, 4096), (50, 4096, 300), (131072, 4096, 2048)])
def test_embedding_bwd_direct(N, V, D, n):
    """Sort-free embedding backward: equals the fp32 index_add, is repeatable, equals the sorted kernel, clears the
    rows only the previous step touched, and reports sum(dtable^2)."""
    g = torch.Generator(device=DEV).manual_seed(3)
    tok = torch.randint(0, V, (n,), device=DEV, generator=g)
    tok[::7] = tok[0]                                   # a heavily repeated id
    dout = bf(n, D, seed=2)
    ref = torch.zeros(V, D, device=DEV).index_add_(0, tok, dout.float())
    dt = torch.zeros(V, D, device=DEV, dtype=torch.bfloat16)
    parts = torch.full((n,), 7.0, device=DEV)
    N.embedding_bwd_direct(tok, None, dout, dt, sumsq=parts)
    torch.testing.assert_close(dt.float(), ref, rtol=2e-2, atol=3e-2)
    torch.testing.assert_close(parts.sum(), dt.float().pow(2).sum(), rtol=1e-4, atol=1e-3)
    dt_sorted = torch.zeros_like(dt)
    s, p = torch.sort(tok, stable=True)
    N.embedding_bwd(s, p, dout, dt_sorted)
    assert torch.equal(dt, dt_sorted)
    # next step with other tokens: rows of the previous step that are not rewritten must end up zero
    tok2 = torch.randint(0, V, (n,), device=DEV, generator=g)
    dout2 = bf(n, D, seed=5)
    N.embedding_bwd_direct(tok2, tok, dout2, dt)
    want = torch.zeros_like(dt)
    N.embedding_bwd_direct(tok2, None, dout2, want)
    assert torch.equal(dt, want)


@pytest.mark.parametrize("rows,V,dim,chunk", [(256, 131072, 512, 8192), (192, 4096, 256, 1024)])
def test_chunked_head_matches_fp32_cross_entropy(N, rows, V, dim, chunk):
    """LM head + cross-entropy without the [T, V] logits (ops/fused_head.py) against fp32 F.cross_entropy on the full
    logits: loss rtol 1e-4-ish (bf16 logits), dX / dW relative error < 1e-2, gradient-norm partials, repeatable."""
    from fault_tolerant_llm_training_b200.ops.fused_head import ChunkedHead
    x, w = bf(rows, dim, scale=1.0, seed=1), bf(V, dim, scale=0.05, seed=2)
    g = torch.Generator(device=DEV).manual_seed(4)
    labels = torch.randint(0, V, (rows,), device=DEV, generator=g)
    labels[::9] = -100
    n = int(labels.ne(-100).sum())
    inv_n = torch.tensor([1.0 / n], device=DEV)
    head = ChunkedHead(rows, dim, V, torch.device(DEV), chunk=chunk)
    assert head.n_chunks == V // chunk and head.buf.numel() == rows * chunk
    row_loss, loss = torch.empty(rows, device=DEV), torch.empty(1, device=DEV)
    head.forward(x, w, labels, inv_n, row_loss, loss)
    dw, dx = torch.empty_like(w), torch.empty_like(x)
    slots = [torch.zeros(N.gemm_sumsq_slots(), device=DEV) for _ in range(head.n_chunks)]
    it = iter(slots)
    head.backward(x, w, labels, inv_n, dw, dx, sumsq=lambda: next(it))
    # fp32 oracle on the same bf16 inputs; logits rounded to bf16 like the Linear's output (model.py:379)
    xf, wf = x.float().requires_grad_(), w.float().requires_grad_()
    logits = (xf @ wf.t())
    ref = torch.nn.functional.cross_entropy(logits, labels, ignore_index=-100, reduction="sum") / n
    ref.backward()
    assert abs(float(loss) - float(ref)) < 2e-3 * abs(float(ref))
    assert _rel(dx, xf.grad) < 1e-2 and _rel(dw, wf.grad) < 1e-2
    total = sum(float(s.sum()) for s in slots)
    assert abs(total - float(dw.float().pow(2).sum())) < 1e-3 * total
    # the same thing again gives the same bits
    loss2, dw2, dx2 = torch.empty(1, device=DEV), torch.empty_like(w), torch.empty_like(x)
    head.forward(x, w, labels, inv_n, row_loss, loss2)
    head.backward(x, w, labels, inv_n, dw2, dx2)
    assert torch.equal(loss, loss2) and torch.equal(dw, dw2) and torch.equal(dx, dx2)
