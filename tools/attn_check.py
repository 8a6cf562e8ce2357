"""Correctness + timing of the tcgen05 flash attention (fwd + deterministic bwd) against an fp32 oracle and cuDNN SDPA.
Usage: python tools/attn_check.py all | <case>"""
import json, os, subprocess, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
CASES = {"s256_hd64": (2, 256, 4, 2, 64), "s512_hd128": (1, 512, 4, 2, 128), "s384_hd128_mha": (1, 384, 2, 2, 128),
         "model": (1, 2048, 32, 8, 128), "model8k": (1, 8192, 32, 8, 128)}


def timeit(fn, iters=8):
    import torch
    for _ in range(3):
        fn()
    ts = []
    for _ in range(iters):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return sorted(ts)[len(ts) // 2]


def run(case):
    import torch
    import torch.nn.functional as F
    from fault_tolerant_llm_training_b200.ops import native as N
    from fault_tolerant_llm_training_b200.ops import torch_ops as T
    B, S, H, KV, hd = CASES[case]
    torch.manual_seed(0)
    W = (H + 2 * KV) * hd
    qkv = (torch.randn(B * S, W, device="cuda") * 0.7).to(torch.bfloat16)
    dout = (torch.randn(B * S, H * hd, device="cuda") * 0.5).to(torch.bfloat16)
    out = torch.empty(B * S, H * hd, device="cuda", dtype=torch.bfloat16)
    lse = torch.empty(B * H * S, device="cuda"); delta = torch.empty_like(lse)
    dqkv = torch.zeros_like(qkv)
    N.attn_fwd(qkv, out, lse, B, S, H, KV, hd)
    N.attn_bwd(qkv, out, dout, lse, delta, dqkv, B, S, H, KV, hd)
    torch.cuda.synchronize()
    res = {"case": case}
    big = S > 1024
    qf = qkv.float().view(B, S, H + 2 * KV, hd)
    q, k, v = (t.clone().requires_grad_() for t in (qf[:, :, :H], qf[:, :, H:H + KV], qf[:, :, H + KV:]))
    if big:  # oracle via SDPA in fp32 (explicit S x S would not fit comfortably)
        o_ref = F.scaled_dot_product_attention(q.transpose(1, 2), k.transpose(1, 2), v.transpose(1, 2), is_causal=True,
                                               enable_gqa=True).transpose(1, 2)
    else:
        o_ref = T.attention_fp32_oracle(q, k, v)
    o_ref.backward(dout.float().view(B, S, H, hd))
    rel = lambda a, b: float((a.float() - b).norm() / b.norm())
    res["out_rel"] = rel(out.view(B, S, H, hd), o_ref.detach())
    res["dq_rel"] = rel(dqkv.view(B, S, -1, hd)[:, :, :H], q.grad)
    res["dk_rel"] = rel(dqkv.view(B, S, -1, hd)[:, :, H:H + KV], k.grad)
    res["dv_rel"] = rel(dqkv.view(B, S, -1, hd)[:, :, H + KV:], v.grad)
    # determinism: a second run must be bitwise identical
    out2, dqkv2 = torch.empty_like(out), torch.zeros_like(dqkv)
    N.attn_fwd(qkv, out2, lse, B, S, H, KV, hd)
    N.attn_bwd(qkv, out2, dout, lse, delta, dqkv2, B, S, H, KV, hd)
    res["bitwise_repeatable"] = bool(torch.equal(out, out2) and torch.equal(dqkv, dqkv2))
    res["ok"] = bool(max(res["out_rel"], res["dq_rel"], res["dk_rel"], res["dv_rel"]) < 2e-2 and res["bitwise_repeatable"])
    if case.startswith("model"):
        qb = qkv.view(B, S, -1, hd)
        qq, kk, vv = (qb[:, :, :H].transpose(1, 2).detach().requires_grad_(), qb[:, :, H:H + KV].transpose(1, 2).detach().requires_grad_(),
                      qb[:, :, H + KV:].transpose(1, 2).detach().requires_grad_())
        do4 = dout.view(B, S, H, hd).transpose(1, 2)

        def lib_fwd():
            return F.scaled_dot_product_attention(qq, kk, vv, is_causal=True, enable_gqa=True)

        def lib_fb():
            o = lib_fwd()
            torch.autograd.grad(o, (qq, kk, vv), do4)
        flops_f = 4 * B * H * S * S * hd / 2
        tf, tfb = timeit(lambda: N.attn_fwd(qkv, out, lse, B, S, H, KV, hd)), None
        tb = timeit(lambda: N.attn_bwd(qkv, out, dout, lse, delta, dqkv, B, S, H, KV, hd))
        lf = timeit(lib_fwd); lfb = timeit(lib_fb)
        res.update(ours_fwd_ms=round(tf, 3), ours_bwd_ms=round(tb, 3), cudnn_fwd_ms=round(lf, 3), cudnn_bwd_ms=round(lfb - lf, 3),
                   ours_fwd_tflops=round(flops_f / tf / 1e9, 1), cudnn_fwd_tflops=round(flops_f / lf / 1e9, 1))
    print(json.dumps(res), flush=True)


if __name__ == "__main__":
    which = sys.argv[1] if len(sys.argv) > 1 else "all"
    if which != "all":
        run(which)
    else:
        for c in (sys.argv[2:] or list(CASES)):
            try:
                r = subprocess.run([sys.executable, __file__, c], capture_output=True, text=True, timeout=150)
                line = [l for l in r.stdout.splitlines() if l.startswith("{")]
                print(line[-1] if line else json.dumps({"case": c, "rc": r.returncode, "err": r.stderr[-500:]}), flush=True)
            except subprocess.TimeoutExpired:
                print(json.dumps({"case": c, "timeout": True}), flush=True)
