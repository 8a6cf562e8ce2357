"""Where does run-to-run nondeterminism come from?  Two identical forward/backward passes, then two identical trainers."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fault_tolerant_llm_training_b200.config import ModelConfig
from fault_tolerant_llm_training_b200.parallel.env import DistEnv
from fault_tolerant_llm_training_b200.models.engine import NativeTrainer
from fault_tolerant_llm_training_b200.parallel.pipeline import PipelinedTrainer

S = int(os.environ.get("SEQ", "8192"))
CFG = ModelConfig(dim=256, n_layers=2, n_heads=4, n_kv_heads=2, multiple_of=64, seq_len=S, vocab_size=2048)
env = DistEnv(device=torch.device("cuda:0"))
kw = dict(dtype=torch.bfloat16, lr=1e-3, warmup_steps=2, grad_max_norm=1.0, seed=11)
g = torch.Generator().manual_seed(0)
tok = torch.randint(0, CFG.vocab_size, (1, S + 1), generator=g)
x, y = tok[:, :-1].contiguous(), tok[:, 1:].contiguous()

tr = NativeTrainer(CFG, env, **kw)
tr._ensure_engine(1, S)
e = tr.engine
inv_n = torch.tensor([1.0 / S], device="cuda")
snaps = []
for rep in range(2):
    e.forward(x.cuda(), y.cuda(), inv_n)
    acts = {"loss": e.loss.clone(), "qkv0": e.qkv[0].clone(), "attn0": e.attn[0].clone(), "act1": e.act[1].clone(),
            "dlogits": e.logits.clone()}
    e.backward(x.cuda())
    torch.cuda.synchronize()
    acts.update({"grad:" + n: p.grad.clone() for n, p in tr.model.named_parameters()})
    acts["dqkv_last"] = e.dqkv.clone()
    snaps.append(acts)
bad = [k for k in snaps[0] if not torch.equal(snaps[0][k], snaps[1][k])]
print("fwd/bwd repeat mismatches:", bad)

for cls in (NativeTrainer, PipelinedTrainer):
    outs = []
    for rep in range(2):
        t = cls(CFG, env, **kw)
        losses = [t.step(x, y).item() for _ in range(4)]
        t.drain(); torch.cuda.synchronize()
        outs.append((losses, torch.cat([p.detach().float().reshape(-1) for p in t.model.parameters()])))
        del t
    print(cls.__name__, "losses equal:", outs[0][0] == outs[1][0], outs[0][0], outs[1][0], "params equal:", torch.equal(outs[0][1], outs[1][1]))
