"""One-GPU instances of the fused collective kernels, for `ncu` (which must never wrap a multi-rank command): the W = 1
templates of the reduce-scatter ⊕ norm² and clip ⊕ AdamW ⊕ all-gather kernels on a 256 MiB range, the plain gather, and
the gather-fused gate|up GEMM with every "owner" pointing at local memory (the fused mechanism without NVLink).

    python tools/comm_single_gpu_case.py            # prints CUDA-event timings (one JSON line)
    ncu --set full -k regex:'dp_|gather|gemm_tc2' ... python tools/comm_single_gpu_case.py
"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fault_tolerant_llm_training_b200.ops import native as N  # noqa: E402


ONCE = "--once" in sys.argv      # under ncu: one launch per kernel


def timed(fn, n=5):
    if ONCE:
        fn()
        torch.cuda.synchronize()
        return 0.0
    for _ in range(2):
        fn()
    ts = []
    for _ in range(n):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return sorted(ts)[len(ts) // 2]


def main():
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    n = 128 << 20                                   # elements (256 MiB of bf16)
    g = (torch.randn(n, device=dev) * 0.01).to(torch.bfloat16)
    p = (torch.randn(n, device=dev) * 0.02).to(torch.bfloat16)
    m, v = torch.zeros_like(p), torch.zeros_like(p)
    tab_g = torch.tensor([g.data_ptr()], dtype=torch.int64, device=dev)
    tab_p = torch.tensor([p.data_ptr()], dtype=torch.int64, device=dev)
    status = torch.zeros(1, dtype=torch.int32, device=dev)
    parts = torch.zeros(N.dp_partials(0), device=dev)
    hyper = torch.tensor([1e-4, 0.9, 0.999, 1e-8, 1e-2, 0.1, 0.0316, 1.0], device=dev)
    stats = torch.tensor([1.0, 1.0, 1.0, 1.0], device=dev)
    out = {}
    out["dp_reduce_scatter<1>_ms"] = timed(lambda: N.dp_reduce_scatter(tab_g, 0, 1, 0, n, parts, status))
    out["dp_adamw_allgather<1>_ms"] = timed(lambda: N.dp_adamw_allgather(tab_p, g, m, v, 0, 1, 0, n, hyper, stats, status))
    dst = torch.empty_like(p)
    out["fsdp_gather<1>_ms"] = timed(lambda: N.fsdp_gather(tab_p, 0, dst, n, 0, 1, status))
    T, d, F, W = 2048, 4096, 14336, 2               # two "owners", both this GPU
    own = (torch.randn(2 * (F // W) * d, device=dev) * 0.02).to(torch.bfloat16)
    tab_o = torch.tensor([own.data_ptr()] * W, dtype=torch.int64, device=dev)
    x = (torch.randn(T, d, device=dev) * 0.5).to(torch.bfloat16)
    w1, w3 = torch.empty(F, d, dtype=torch.bfloat16, device=dev), torch.empty(F, d, dtype=torch.bfloat16, device=dev)
    act, gate, up = (torch.empty(T, F, dtype=torch.bfloat16, device=dev) for _ in range(3))
    scratch = torch.zeros(1 + 2 * (F // 128) + 8, dtype=torch.int32, device=dev)
    out["gemm_swiglu_gather_local_ms"] = timed(lambda: N.gemm_swiglu_gather(x, w1, w3, act, gate, up, tab_o, 0, (F // W) * d,
                                                                            F // W, 0, W, scratch))
    out["gemm_swiglu_ms"] = timed(lambda: N.gemm_swiglu(x, w1, w3, act, gate, up))
    print(json.dumps({k: round(v, 4) for k, v in out.items()}))


if __name__ == "__main__":
    main()
