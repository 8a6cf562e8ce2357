"""T1/T3 on CPU: the portable model equals the reference's model.py op for op, and checkpoints are
interchangeable with the reference's classes."""
import torch

from fault_tolerant_llm_training_b200.ckpt.layout import load_checkpoint, save_checkpoint
from fault_tolerant_llm_training_b200.config import ModelConfig
from fault_tolerant_llm_training_b200.models.llama import build_model, param_layout
from fault_tolerant_llm_training_b200.ops import torch_ops as T
from fault_tolerant_llm_training_b200.optim.lr import build_lr_scheduler

CFG = ModelConfig(dim=64, n_layers=2, n_heads=4, n_kv_heads=2, multiple_of=32, ffn_dim_multiplier=1.3,
                  seq_len=24, vocab_size=128, rope_theta=500000.0)


def _ref_model(reference_modules, cfg, dtype=torch.float32):
    m = reference_modules["model"]
    args = m.TransformerModelArgs(dim=cfg.dim, n_layers=cfg.n_layers, n_heads=cfg.n_heads, n_kv_heads=cfg.n_kv_heads,
                                  multiple_of=cfg.multiple_of, ffn_dim_multiplier=cfg.ffn_dim_multiplier,
                                  norm_eps=cfg.norm_eps, rope_theta=cfg.rope_theta, seq_len=cfg.seq_len,
                                  vocab_size=cfg.vocab_size)
    with reference_modules["utils"].set_default_dtype(dtype):
        return m.Transformer(args)


def test_state_dict_names_and_order_match_reference(reference_modules):
    ours = build_model(CFG, torch.float32)
    ref = _ref_model(reference_modules, CFG)
    assert list(ours.state_dict().keys()) == list(ref.state_dict().keys())
    assert [tuple(v.shape) for v in ours.state_dict().values()] == [tuple(v.shape) for v in ref.state_dict().values()]
    assert [n for n, _ in ours.named_parameters()] == [n for n, _ in ref.named_parameters()]
    assert len(param_layout(ours)) == 3 + 9 * CFG.n_layers
    full = ModelConfig()
    assert 3 + 9 * full.n_layers == 291  # the 291 tensors of SURVEY §2.4


def test_forward_backward_match_reference(reference_modules):
    torch.manual_seed(0)
    ours = build_model(CFG, torch.float32, seed=3)
    ref = _ref_model(reference_modules, CFG)
    ref.load_state_dict(ours.state_dict(), strict=True)
    tok = torch.randint(0, CFG.vocab_size, (2, CFG.seq_len))
    lo, lr = ours(tok), ref(tok)
    torch.testing.assert_close(lo, lr, rtol=1e-5, atol=1e-5)
    lo.square().mean().backward()
    lr.square().mean().backward()
    for (n, p), (_, q) in zip(ours.named_parameters(), ref.named_parameters()):
        torch.testing.assert_close(p.grad, q.grad, rtol=2e-4, atol=1e-6, msg=n)


def test_bf16_ops_match_reference_rounding(reference_modules):
    m = reference_modules["model"]
    torch.manual_seed(1)
    x = torch.randn(2, 8, 64, dtype=torch.bfloat16)
    w = torch.randn(64, dtype=torch.bfloat16)
    norm = m.RMSNorm(64, eps=1e-5)
    norm.weight.data = w.clone()
    assert torch.equal(T.rmsnorm(x, w, 1e-5), norm(x))
    q = torch.randn(2, 8, 4, 16, dtype=torch.bfloat16)
    k = torch.randn(2, 8, 2, 16, dtype=torch.bfloat16)
    cos, sin = T.rope_table(16, 8, 500000.0)
    fc = m.precompute_freqs_cis(16, 8, 500000.0)
    rq, rk = m.apply_rotary_emb(q, k, fc)
    assert torch.equal(T.rope_apply(q, cos, sin), rq) and torch.equal(T.rope_apply(k, cos, sin), rk)
    # inverse rotation is the transpose (the backward of RoPE)
    torch.testing.assert_close(T.rope_apply(T.rope_apply(q.float(), cos, sin), cos, sin, inverse=True), q.float(),
                               rtol=1e-5, atol=1e-5)


def test_attention_matches_oracle():
    torch.manual_seed(2)
    q, k, v = torch.randn(2, 12, 4, 16), torch.randn(2, 12, 2, 16), torch.randn(2, 12, 2, 16)
    torch.testing.assert_close(T.attention(q, k, v), T.attention_fp32_oracle(q, k, v), rtol=1e-5, atol=1e-5)


def test_checkpoint_loadable_by_reference_classes(reference_modules, tmp_path):
    """T3: keys, names, AdamW and LambdaLR layouts are what the reference's resume code expects
    (train.py:56-83)."""
    ours = build_model(CFG, torch.bfloat16, seed=5)
    opt = torch.optim.AdamW(ours.parameters(), lr=1e-3)
    sched = build_lr_scheduler(opt, 3)
    tok = torch.randint(0, CFG.vocab_size, (1, CFG.seq_len))
    for _ in range(2):
        opt.zero_grad()
        ours(tok).float().square().mean().backward()
        opt.step()
        sched.step()
    path = save_checkpoint(str(tmp_path), "123", ours, opt, sched, 2, data_state={"batches_consumed": 2})
    ck = torch.load(path, map_location="cpu", weights_only=False)
    assert list(ck.keys())[:4] == ["model", "optimizer", "lr_scheduler", "training_step"]
    assert ck["training_step"] == 2 and ck["data_state"]["batches_consumed"] == 2
    ref = _ref_model(reference_modules, CFG, torch.bfloat16)
    ref.load_state_dict(ck["model"], strict=True)
    ropt = torch.optim.AdamW(ref.parameters(), lr=1e-3)
    ropt.load_state_dict(ck["optimizer"])
    rs = reference_modules["utils"].build_lr_scheduler(ropt, 3)
    rs.load_state_dict(ck["lr_scheduler"])
    st = ck["optimizer"]["state"]
    assert sorted(st.keys()) == list(range(len(list(ours.parameters()))))
    assert set(st[0].keys()) == {"step", "exp_avg", "exp_avg_sq"}
    assert st[0]["exp_avg"].dtype == torch.bfloat16 and st[0]["step"].dtype == torch.float32
    # (the reference rebuilds its LambdaLR *after* loading the optimizer, train.py:70-77, which resets the
    # group's lr for one step; the schedule position itself round-trips)
    assert rs.get_last_lr() == sched.get_last_lr() and rs.last_epoch == sched.last_epoch == 2
    # and the other direction: a reference-written file resumes here
    torch.save({"model": ref.state_dict(), "optimizer": ropt.state_dict(), "lr_scheduler": rs.state_dict(),
                "training_step": 2}, tmp_path / "checkpoint_ref.ckpt")
    back = load_checkpoint(str(tmp_path), "ref")
    ours.load_state_dict(back["model"], strict=True)
    opt.load_state_dict(back["optimizer"])
    sched.load_state_dict(back["lr_scheduler"])
