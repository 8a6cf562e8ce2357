"""Per-kernel device times of the attention forward + backward (torch profiler / CUPTI), model shape by default."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.profiler import profile, ProfilerActivity
from fault_tolerant_llm_training_b200.ops import native as N
from fault_tolerant_llm_training_b200.ops import torch_ops as T

B, S, H, KV, hd = 1, 2048, 32, 8, 128
if len(sys.argv) > 2:
    B, S = int(sys.argv[1]), int(sys.argv[2])
torch.manual_seed(0)
W = (H + 2 * KV) * hd
qkv = (torch.randn(B * S, W, device="cuda") * 0.7).to(torch.bfloat16)
dout = (torch.randn(B * S, H * hd, device="cuda") * 0.5).to(torch.bfloat16)
out = torch.empty(B * S, H * hd, device="cuda", dtype=torch.bfloat16)
lse = torch.empty(B * H * S, device="cuda"); delta = torch.empty_like(lse)
dqkv = torch.zeros_like(qkv)
cos, sin = T.rope_table(hd, S, 500000.0, device="cuda")   # the backward un-rotates dq / dk in its epilogues, as in training
for _ in range(3):
    N.attn_fwd(qkv, out, lse, B, S, H, KV, hd); N.attn_bwd(qkv, out, dout, lse, delta, dqkv, B, S, H, KV, hd, cos, sin)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(10):
        N.attn_fwd(qkv, out, lse, B, S, H, KV, hd); N.attn_bwd(qkv, out, dout, lse, delta, dqkv, B, S, H, KV, hd, cos, sin)
    torch.cuda.synchronize()
rows = [(e.key, e.device_time_total / e.count) for e in prof.key_averages() if e.device_time_total > 0]
for k, t in sorted(rows, key=lambda r: -r[1]):
    print(f"{t:9.1f} us  {k[:90]}")

# the library SDPA measured the same way (back to back, CUPTI kernel times) for a like-for-like comparison
import torch.nn.functional as F
x = qkv.view(B, S, H + 2 * KV, hd)
q, k, v = (t.transpose(1, 2).contiguous().requires_grad_() for t in (x[:, :, :H], x[:, :, H:H + KV], x[:, :, H + KV:]))
do = dout.view(B, S, H, hd).transpose(1, 2).contiguous()
def sdpa():
    o = F.scaled_dot_product_attention(q, k, v, is_causal=True, enable_gqa=True)
    o.backward(do)
for _ in range(3):
    sdpa()
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(10):
        sdpa()
    torch.cuda.synchronize()
rows = [(e.key, e.device_time_total / e.count, e.count // 10) for e in prof.key_averages() if e.device_time_total > 0]
print("library SDPA (forward + autograd backward), same method:")
for k_, t, n in sorted(rows, key=lambda r: -r[1])[:8]:
    print(f"{t:9.1f} us x{n}  {k_[:90]}")
