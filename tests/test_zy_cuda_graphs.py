"""Per-layer CUDA graphs (FTL_GRAPHS=1, models/engine.py): a block's forward / backward launches captured once and
replayed must reproduce the eager trajectory bit for bit — same kernels, same buffers, same order."""
import pytest
import torch

from fault_tolerant_llm_training_b200.config import ModelConfig
from fault_tolerant_llm_training_b200.parallel.env import DistEnv

pytestmark = pytest.mark.gpu

CFG = ModelConfig(dim=256, n_layers=3, n_heads=4, n_kv_heads=2, multiple_of=64, ffn_dim_multiplier=1.3,
                  seq_len=128, vocab_size=1024)


def _run(steps=6):
    from fault_tolerant_llm_training_b200.ops import native as N
    from fault_tolerant_llm_training_b200.parallel.pipeline import PipelinedTrainer
    env = DistEnv(device=torch.device("cuda:0"))
    tr = PipelinedTrainer(CFG, env, dtype=torch.bfloat16, lr=1e-3, warmup_steps=2, grad_max_norm=1.0, seed=5)
    g = torch.Generator().manual_seed(1)
    losses, k0 = [], N.KERNELS
    for _ in range(steps):
        tok = torch.randint(0, CFG.vocab_size, (2, CFG.seq_len + 1), generator=g)
        losses.append(tr.step(tok[:, :-1].contiguous(), tok[:, 1:].contiguous()).item())
    tr.drain()
    params = torch.cat([p.detach().float().reshape(-1) for p in tr.model.parameters()]).clone()
    return losses, params, tr, N.KERNELS - k0


def test_graph_replay_is_bitwise_equal_to_eager(monkeypatch):
    monkeypatch.delenv("FTL_GRAPHS", raising=False)
    eager_losses, eager_params, tr0, eager_kernels = _run()
    assert not tr0.engine.use_graphs
    monkeypatch.setenv("FTL_GRAPHS", "1")
    graph_losses, graph_params, tr1, graph_kernels = _run()
    assert tr1.engine.use_graphs and len(tr1.engine._graphs) == 2 * CFG.n_layers     # forward + backward of every block
    assert graph_losses == eager_losses
    assert torch.equal(graph_params, eager_params)
    assert graph_kernels == eager_kernels          # replayed kernels are counted like launched ones
