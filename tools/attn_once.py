"""One forward + backward of the attention kernels at the model shape (for ncu captures)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from fault_tolerant_llm_training_b200.ops import native as N
B, S, H, KV, hd = 1, 2048, 32, 8, 128
torch.manual_seed(0)
qkv = (torch.randn(B * S, (H + 2 * KV) * hd, device="cuda") * 0.7).to(torch.bfloat16)
dout = (torch.randn(B * S, H * hd, device="cuda") * 0.5).to(torch.bfloat16)
out = torch.empty(B * S, H * hd, device="cuda", dtype=torch.bfloat16)
lse = torch.empty(B * H * S, device="cuda"); delta = torch.empty_like(lse)
dqkv = torch.zeros_like(qkv)
for _ in range(3):
    N.attn_fwd(qkv, out, lse, B, S, H, KV, hd); N.attn_bwd(qkv, out, dout, lse, delta, dqkv, B, S, H, KV, hd)
torch.cuda.synchronize()
print("ok")
