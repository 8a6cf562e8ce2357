"""Stand-alone run of the memory-bound kernels on flagship-sized buffers (for ncu captures and roofline numbers)."""
import math, os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fault_tolerant_llm_training_b200.ops import native as N
n = 1 << 30   # 1 Gi elements (2 GiB per bf16 buffer)
p, g, m, v = (torch.randn(n, device="cuda", dtype=torch.bfloat16) * 0.01 for _ in range(4))
hyper = torch.tensor([5e-5, 0.9, 0.999, 1e-8, 1e-2, 0.1, math.sqrt(1e-3), 1.0], device="cuda")
stats, scratch = torch.zeros(4, device="cuda"), torch.empty(N.grad_norm_partials(), device="cuda")
x = torch.randn(2048, 4096, device="cuda").to(torch.bfloat16); w = torch.ones(4096, device="cuda", dtype=torch.bfloat16)
y, rstd = torch.empty_like(x), torch.empty(2048, device="cuda")
logits = torch.randn(2048, 131072, device="cuda").to(torch.bfloat16); labels = torch.randint(0, 131072, (2048,), device="cuda")
inv_n, row_loss, loss = torch.tensor([1 / 2048.], device="cuda"), torch.empty(2048, device="cuda"), torch.empty(1, device="cuda")
def t(fn, it=5):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(it): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / it
ms = t(lambda: (N.grad_norm(g, hyper, scratch, stats), N.adamw_step(p, g, m, v, hyper, stats)))
print(f"grad_norm + adamw on {n} elements: {ms:.3f} ms -> {(n * 2 + n * 14) / ms / 1e6:.0f} GB/s")
ms = t(lambda: N.rmsnorm_fwd(x, w, 1e-5, y, rstd), 50)
print(f"rmsnorm_fwd 2048x4096: {ms * 1e3:.1f} us -> {x.numel() * 4 / ms / 1e6:.0f} GB/s")
ms = t(lambda: N.cross_entropy(logits, labels, inv_n, row_loss, loss), 10)
print(f"cross_entropy 2048x131072 (fwd+bwd in place): {ms * 1e3:.1f} us -> {logits.numel() * 4 / ms / 1e6:.0f} GB/s algorithmic")
