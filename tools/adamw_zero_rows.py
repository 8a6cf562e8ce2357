"""AdamW throughput on elements whose gradient / moments are exactly zero (embedding rows no batch has touched) against
dense state: sqrtf() and the IEEE division have slow fix-up paths for zero / tiny arguments (csrc/common/util.cuh)."""
import sys, math, os, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fault_tolerant_llm_training_b200.ops import native as N
n = 537 * 1000 * 1000 // 8 * 8
hyper = torch.tensor([1e-2, 0.9, 0.999, 1e-8, 1e-2, 1 - 0.9, math.sqrt(1 - 0.999), 1.0], device="cuda")
def rnd(scale): return (torch.randn(n, device="cuda") * scale).to(torch.bfloat16)
def zero(): return torch.zeros(n, device="cuda", dtype=torch.bfloat16)
for name, g, m, v in (("g=m=v=0", zero(), zero(), zero()), ("g=0 m=0 v dense", zero(), zero(), rnd(0.01).abs()),
                      ("g=0 m dense v=0", zero(), rnd(0.01), zero()), ("g=0 m,v dense", zero(), rnd(0.01), rnd(0.01).abs()),
                      ("all dense", rnd(0.01), rnd(0.01), rnd(0.01).abs())):
    p = rnd(1.0)
    for _ in range(2): N.adamw_step(p, g, m, v, hyper, None)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5): N.adamw_step(p, g, m, v, hyper, None)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    print(f"{name:18s} {ms:.3f} ms  {n * 14 / ms / 1e9:.2f} TB/s")
    del p, g, m, v
