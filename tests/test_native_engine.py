"""The hand-scheduled B200 engine against autograd on the portable model (same weights, same batch)."""
import pytest
import torch

from fault_tolerant_llm_training_b200.config import ModelConfig
from fault_tolerant_llm_training_b200.models.llama import build_model
from fault_tolerant_llm_training_b200.ops import torch_ops as T
from fault_tolerant_llm_training_b200.parallel.env import DistEnv

pytestmark = pytest.mark.gpu

CFG = ModelConfig(dim=256, n_layers=2, n_heads=4, n_kv_heads=2, multiple_of=64, ffn_dim_multiplier=1.3,
                  seq_len=64, vocab_size=1024)


def _trainer(**kw):
    from fault_tolerant_llm_training_b200.models.engine import NativeTrainer
    env = DistEnv(device=torch.device("cuda:0"))
    args = dict(dtype=torch.bfloat16, lr=1e-3, warmup_steps=2, grad_max_norm=1.0, seed=11)
    args.update(kw)
    return NativeTrainer(CFG, env, **args)


def _batch(seed=0, B=2):
    g = torch.Generator().manual_seed(seed)
    tok = torch.randint(0, CFG.vocab_size, (B, CFG.seq_len + 1), generator=g)
    inputs, labels = tok[:, :-1].contiguous(), tok[:, 1:].clone()
    labels[:, :5] = -100
    return inputs, labels


@pytest.mark.parametrize("fused_head", [False, True])
def test_engine_loss_and_grads_match_autograd(fused_head, monkeypatch):
    if fused_head:   # vocabulary-chunked LM head + cross-entropy: 4 chunks of 256 columns, no [T, V] buffer
        monkeypatch.setenv("FTL_FUSED_HEAD", "1")
        monkeypatch.setenv("FTL_HEAD_CHUNK", "256")
    tr = _trainer()
    ref = build_model(CFG, torch.float32, "cuda", seed=0)
    ref.load_state_dict({k: v.float() for k, v in tr.model.state_dict().items()})
    inputs, labels = _batch()
    tr._ensure_engine(*inputs.shape)
    assert (tr.engine.head is not None and tr.engine.logits is None and tr.engine.head.n_chunks == 4) if fused_head \
        else tr.engine.head is None
    n = int(labels.ne(-100).sum())
    inv_n = torch.tensor([1.0 / n], device="cuda")
    loss = tr.engine.forward(inputs.cuda(), labels.cuda(), inv_n)
    tr.engine.backward(inputs.cuda())
    ref_loss = T.cross_entropy_sum(ref(inputs.cuda()), labels.cuda()) / n
    ref_loss.backward()
    assert abs(float(loss) - float(ref_loss)) < 3e-2 * max(1.0, abs(float(ref_loss)))
    for (name, p), (_, q) in zip(tr.model.named_parameters(), ref.named_parameters()):
        g, r = p.grad.float(), q.grad
        err = (g - r).norm() / (r.norm() + 1e-12)
        assert err < 5e-2, (name, float(err))


def test_trainer_steps_reduce_loss_and_checkpoint_roundtrip(tmp_path):
    from fault_tolerant_llm_training_b200.ckpt.layout import load_checkpoint, save_checkpoint
    tr = _trainer()
    inputs, labels = _batch(1)
    losses = [tr.step(inputs, labels).item() for _ in range(8)]
    assert losses[-1] < losses[0]
    assert tr.kernel_launches_per_step > 0
    save_checkpoint(str(tmp_path), "7", tr.model, tr.optimizer, tr.lr_scheduler, 8)
    ck = load_checkpoint(str(tmp_path), "7")
    assert len(ck["model"]) == 3 + 9 * CFG.n_layers
    assert all(int(s["step"]) == 8 for s in ck["optimizer"]["state"].values())
    tr2 = _trainer(seed=99)
    tr2.load_state(ck)
    a = tr.step(inputs, labels).item()
    b = tr2.step(inputs, labels).item()
    assert a == b  # bitwise-identical continuation after a save/load round trip
    for p, q in zip(tr.model.parameters(), tr2.model.parameters()):
        assert torch.equal(p, q)
    # the torch-AdamW layout loads into a stock torch optimizer too (what the reference does on resume)
    opt = torch.optim.AdamW(build_model(CFG, torch.bfloat16, "cpu").parameters(), lr=1e-3)
    opt.load_state_dict(ck["optimizer"])


def test_sharded_checkpoint_roundtrip_and_reference_layout(tmp_path):
    """Native snapshot engine: shards -> manifest (durable) -> consolidated reference-layout file; resume from
    either gives a bitwise-identical continuation."""
    import os
    from fault_tolerant_llm_training_b200.ckpt.layout import load_checkpoint
    from fault_tolerant_llm_training_b200.ckpt.sharded import MANIFEST, shard_dir
    tr = _trainer(keep_shards=True)
    inputs, labels = _batch(2)
    for _ in range(4):
        tr.step(inputs, labels)
    tr.drain()
    d = tr.save_fn()(str(tmp_path), "55", tr.model, tr.optimizer, tr.lr_scheduler, 4, data_state={"batches_consumed": 4})
    assert os.path.exists(os.path.join(d, MANIFEST)) and d == shard_dir(str(tmp_path), "55")
    assert tr.checkpointer.timings["bytes_this_rank"] == 3 * 2 * tr.flat.numel
    # (a) resume straight from the shard directory (as after a kill during consolidation)
    ck_sh = load_checkpoint(str(tmp_path), "55")
    assert "_sharded_dir" in ck_sh and ck_sh["training_step"] == 4
    tr_a = _trainer(seed=77)
    tr_a.load_state(ck_sh)
    # (b) consolidate, then resume from the reference-layout file
    out = tr.finalize_fn()()
    assert out and os.path.exists(out)
    ck = load_checkpoint(str(tmp_path), "55")
    assert list(ck.keys())[:4] == ["model", "optimizer", "lr_scheduler", "training_step"]
    assert ck["data_state"] == {"batches_consumed": 4} and len(ck["model"]) == 3 + 9 * CFG.n_layers
    tr_b = _trainer(seed=78)
    tr_b.load_state(ck)
    want = tr.step(inputs, labels).item()
    assert tr_a.step(inputs, labels).item() == want and tr_b.step(inputs, labels).item() == want
    for p, q, r in zip(tr.model.parameters(), tr_a.model.parameters(), tr_b.model.parameters()):
        assert torch.equal(p, q) and torch.equal(p, r)
    assert tr_a.optimizer.steps_done == tr_b.optimizer.steps_done == 5
    # stock torch AdamW accepts the consolidated optimizer state (reference resume path, train.py:70-72)
    opt = torch.optim.AdamW(build_model(CFG, torch.bfloat16, "cpu").parameters(), lr=1e-3)
    opt.load_state_dict(ck["optimizer"])


def test_pipelined_trainer_matches_plain_schedule():
    """The overlapped two-stream schedule (parallel/pipeline.py) computes the same training trajectory as the
    serial one; only the summation order of the gradient norm differs."""
    from fault_tolerant_llm_training_b200.parallel.pipeline import PipelinedTrainer
    env = DistEnv(device=torch.device("cuda:0"))
    kw = dict(dtype=torch.bfloat16, lr=1e-3, warmup_steps=2, grad_max_norm=1.0, seed=11)
    plain = _trainer()
    piped = PipelinedTrainer(CFG, env, **kw)
    a, b = [], []
    for s in range(6):
        inputs, labels = _batch(s)
        a.append(plain.step(inputs, labels).item())
        b.append(piped.step(inputs, labels).item())
    piped.drain()
    assert max(abs(x - y) for x, y in zip(a, b)) < 2e-2, (a, b)
    assert piped.committed_updates() == 6
    pa = torch.cat([p.float().reshape(-1) for p in plain.model.parameters()])
    pb = torch.cat([p.float().reshape(-1) for p in piped.model.parameters()])
    assert float((pa - pb).norm() / pa.norm()) < 1e-3
    # moments this "rank" maintains cover the whole flat space when W == 1
    assert sum(hi - lo for lo, hi in piped._moment_ranges()) == piped.flat.numel


def test_nonfinite_step_is_not_committed():
    from fault_tolerant_llm_training_b200.parallel.pipeline import PipelinedTrainer
    env = DistEnv(device=torch.device("cuda:0"))
    tr = PipelinedTrainer(CFG, env, dtype=torch.bfloat16, lr=1e-3, warmup_steps=2, grad_max_norm=1.0, seed=11)
    inputs, labels = _batch(0)
    tr.step(inputs, labels)
    tr.drain()
    before = [p.clone() for p in tr.model.parameters()]
    with torch.no_grad():
        tr.model.output.weight[0, 0] = float("inf")      # poisons the logits -> non-finite gradient norm
    tr.step(inputs, labels)
    with torch.no_grad():
        tr.model.output.weight[0, 0] = before[-1][0, 0]   # the host is already two steps ahead with healthy batches:
    tr.step(inputs, labels)                               # no update may be applied after the skipped one
    with pytest.raises(FloatingPointError):
        tr.drain()
    assert tr.committed_updates() == 1                    # the poisoned step and its successors left the state alone
    with torch.no_grad():
        tr.model.output.weight[0, 0] = before[-1][0, 0]
    for p, q in zip(tr.model.parameters(), before):
        assert torch.equal(p, q)


def test_loss_curve_tracks_the_reference_implementation(reference_modules):
    """T4: same init, same batches — the B200 engine and the reference's own model.py + torch AdamW (with the
    clipping the reference intended) follow the same loss curve within bf16 noise."""
    m = reference_modules["model"]
    tr = _trainer(lr=2e-3, warmup_steps=3)
    args = m.TransformerModelArgs(dim=CFG.dim, n_layers=CFG.n_layers, n_heads=CFG.n_heads, n_kv_heads=CFG.n_kv_heads,
                                  multiple_of=CFG.multiple_of, ffn_dim_multiplier=CFG.ffn_dim_multiplier,
                                  norm_eps=CFG.norm_eps, rope_theta=CFG.rope_theta, seq_len=CFG.seq_len,
                                  vocab_size=CFG.vocab_size)
    with reference_modules["utils"].set_default_dtype(torch.bfloat16):
        ref = m.Transformer(args)
    ref.load_state_dict(tr.model.state_dict(), strict=True)
    ref.cuda().train()
    opt = torch.optim.AdamW(ref.parameters(), lr=2e-3)
    sched = reference_modules["utils"].build_lr_scheduler(opt, 3)
    ours, theirs = [], []
    for s in range(12):
        inputs, labels = _batch(s % 3)
        ours.append(tr.step(inputs, labels).item())
        opt.zero_grad()
        logits = ref(inputs.cuda())
        loss = torch.nn.functional.cross_entropy(logits.flatten(0, 1).float(), labels.cuda().flatten(0, 1), reduction="sum")
        loss = loss / labels.ne(-100).sum()
        loss.backward()
        torch.nn.utils.clip_grad_norm_(list(ref.parameters()), 1.0)
        opt.step()
        sched.step()
        theirs.append(float(loss))
    assert theirs[-1] < theirs[0] - 0.5                      # it actually trains
    assert max(abs(a - b) for a, b in zip(ours, theirs)) < 0.08, (ours, theirs)
