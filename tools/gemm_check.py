"""Correctness + timing of the tcgen05 GEMM variants against cuBLAS (torch.mm) on the model's shapes.
Usage: python tools/gemm_check.py <case>   (each case runs in its own process: a device trap poisons the context)
       python tools/gemm_check.py all      (spawns every case with a timeout)"""
import json
import os
import subprocess
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

CASES = ["nt_small", "nt_tail", "nt_model", "nn_small", "nn_model", "tn_small", "tn_model", "split_small", "split_model",
         "swiglu_small", "swiglu_model", "rope_small", "rope_model", "lmhead"]


def timeit(fn, iters=10):
    import torch
    for _ in range(3):
        fn()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    ts = []
    for _ in range(iters):
        flush.zero_()  # flush L2 between timed iterations
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2]


def run(case):
    import torch
    from fault_tolerant_llm_training_b200.ops import native as N
    from fault_tolerant_llm_training_b200.ops import torch_ops as T
    torch.manual_seed(0)
    dev = "cuda"
    rnd = lambda *s: (torch.randn(*s, device=dev) * 0.5).to(torch.bfloat16)
    res = {"case": case}

    def report(out, ref, flops, ours_fn, lib_fn):
        err = (out.float() - ref).abs().max().item()
        rel = ((out.float() - ref).norm() / ref.norm()).item()
        res.update(max_abs_err=err, rel_err=rel, ok=bool(rel < 1e-2))
        if flops:
            t_o, t_l = timeit(ours_fn), timeit(lib_fn)
            res.update(ours_ms=round(t_o, 4), cublas_ms=round(t_l, 4), ours_tflops=round(flops / t_o / 1e9, 1),
                       cublas_tflops=round(flops / t_l / 1e9, 1), ratio=round(t_l / t_o, 3))

    kind, size = case.split("_") if "_" in case else (case, "model")
    if kind == "nt":
        M, Nn, K = {"small": (256, 512, 256), "tail": (200, 344, 136), "model": (2048, 14336, 4096)}[size]
        a, b = rnd(M, K), rnd(Nn, K)
        out = torch.empty(M, Nn, device=dev, dtype=torch.bfloat16)
        N.gemm(a, b, out)
        report(out, a.float() @ b.float().t(), 2 * M * Nn * K if size == "model" else 0,
               lambda: N.gemm(a, b, out), lambda: torch.mm(a, b.t(), out=out))
    elif kind == "nn":  # dgrad: dX[M,K] = dY[M,N] @ W[N,K]  ->  A = dY (K-major), B = W read MN-major
        M, Nn, K = {"small": (256, 512, 256), "model": (2048, 14336, 4096)}[size]
        dy, w = rnd(M, Nn), rnd(Nn, K)
        out = torch.empty(M, K, device=dev, dtype=torch.bfloat16)
        N.gemm(dy, w, out, b_mn=True)
        report(out, dy.float() @ w.float(), 2 * M * Nn * K if size == "model" else 0,
               lambda: N.gemm(dy, w, out, b_mn=True), lambda: torch.mm(dy, w, out=out))
    elif kind == "tn":  # wgrad: dW[N,K] = dY[M,N]^T @ X[M,K]
        M, Nn, K = {"small": (256, 512, 256), "model": (2048, 14336, 4096)}[size]
        dy, x = rnd(M, Nn), rnd(M, K)
        out = torch.empty(Nn, K, device=dev, dtype=torch.bfloat16)
        N.gemm(dy, x, out, a_mn=True, b_mn=True)
        report(out, dy.float().t() @ x.float(), 2 * M * Nn * K if size == "model" else 0,
               lambda: N.gemm(dy, x, out, a_mn=True, b_mn=True), lambda: torch.mm(dy.t(), x, out=out))
    elif kind == "split":  # dhn = dgate @ W1 + dup @ W3
        M, H, K = {"small": (256, 512, 256), "model": (2048, 14336, 4096)}[size]
        dg, du, w1, w3 = rnd(M, H), rnd(M, H), rnd(H, K), rnd(H, K)
        out = torch.empty(M, K, device=dev, dtype=torch.bfloat16)
        N.gemm(dg, w1, out, b_mn=True, a2=du, b2=w3)

        def lib():
            torch.mm(dg, w1, out=out)
            torch.addmm(out, du, w3, out=out)
        report(out, dg.float() @ w1.float() + du.float() @ w3.float(), 4 * M * H * K if size == "model" else 0,
               lambda: N.gemm(dg, w1, out, b_mn=True, a2=du, b2=w3), lib)
    elif kind == "swiglu":
        M, H, K = {"small": (256, 384, 256), "model": (2048, 14336, 4096)}[size]
        x, w1, w3 = rnd(M, K), rnd(H, K) * 0.1, rnd(H, K) * 0.1
        act, gate, up = (torch.empty(M, H, device=dev, dtype=torch.bfloat16) for _ in range(3))
        N.gemm_swiglu(x, w1, w3, act, gate, up)
        g_ref, u_ref = (x.float() @ w1.float().t()), (x.float() @ w3.float().t())
        ref = torch.nn.functional.silu(g_ref) * u_ref
        assert ((gate.float() - g_ref).norm() / g_ref.norm()) < 1e-2 and ((up.float() - u_ref).norm() / u_ref.norm()) < 1e-2

        def lib():
            torch.mm(x, w1.t(), out=gate); torch.mm(x, w3.t(), out=up); N.swiglu_fwd(gate, up, act)
        report(act, ref, 4 * M * H * K if size == "model" else 0, lambda: N.gemm_swiglu(x, w1, w3, act, gate, up), lib)
    elif kind == "rope":
        B, S, Hh, KV, hd, K = {"small": (2, 128, 4, 2, 64, 256), "model": (1, 2048, 32, 8, 128, 4096)}[size]
        M, Nq = B * S, (Hh + 2 * KV) * hd
        x, w = rnd(M, K), rnd(Nq, K) * 0.1
        cos, sin = T.rope_table(hd, S, 500000.0, device=dev)
        out = torch.empty(M, Nq, device=dev, dtype=torch.bfloat16)
        N.gemm_qkv_rope(x, w, out, cos, sin, S, hd, (Hh + KV) * hd)
        ref = (x.float() @ w.float().t()).to(torch.bfloat16)
        rot = (Hh + KV) * hd
        ref_rot = T.rope_apply(ref[:, :rot].reshape(B, S, Hh + KV, hd), cos, sin).reshape(M, rot)
        ref = torch.cat([ref_rot, ref[:, rot:]], dim=1).float()

        def lib():
            torch.mm(x, w.t(), out=out); N.rope_inplace(out, cos, sin, S, rot, hd)
        report(out, ref, 2 * M * Nq * K if size == "model" else 0,
               lambda: N.gemm_qkv_rope(x, w, out, cos, sin, S, hd, rot), lib)
    elif kind == "lmhead":
        M, Nn, K = 2048, 131072, 4096
        a, b = rnd(M, K), rnd(Nn, K)
        out = torch.empty(M, Nn, device=dev, dtype=torch.bfloat16)
        N.gemm(a, b, out)
        ref_cols = a.float() @ b[:4096].float().t()
        report(out[:, :4096], ref_cols, 2 * M * Nn * K, lambda: N.gemm(a, b, out), lambda: torch.mm(a, b.t(), out=out))
    torch.cuda.synchronize()
    print(json.dumps(res), flush=True)


if __name__ == "__main__":
    which = sys.argv[1] if len(sys.argv) > 1 else "all"
    if which != "all":
        run(which)
    else:
        for c in (sys.argv[2:] or CASES):
            try:
                r = subprocess.run([sys.executable, __file__, c], capture_output=True, text=True, timeout=120)
                line = [l for l in r.stdout.splitlines() if l.startswith("{")]
                print(line[-1] if line else json.dumps({"case": c, "rc": r.returncode, "err": r.stderr[-400:]}), flush=True)
            except subprocess.TimeoutExpired:
                print(json.dumps({"case": c, "timeout": True}), flush=True)
