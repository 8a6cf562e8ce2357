import os, sys, torch
sys.path.insert(0, "/root/repo")
from fault_tolerant_llm_training_b200.ops import native as N
B,S,H,KV,hd = 1,2048,32,8,128
if len(sys.argv) > 1: B,S = int(sys.argv[1]), int(sys.argv[2])
qkv=(torch.randn(B*S,(H+2*KV)*hd,device="cuda")*0.7).to(torch.bfloat16)
out=torch.empty(B*S,H*hd,device="cuda",dtype=torch.bfloat16); lse=torch.empty(B*H*S,device="cuda")
for _ in range(3): N.attn_fwd(qkv,out,lse,B,S,H,KV,hd)
ts=[]
for _ in range(10):
    e0,e1=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
    e0.record(); N.attn_fwd(qkv,out,lse,B,S,H,KV,hd); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
print(os.environ.get("FTL_ATTN_DEBUG","0"), B, S, round(sorted(ts)[5]*1000,1), "us")
