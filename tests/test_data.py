"""Dataset tests: masking, packing, resumable position."""
import torch

from fault_tolerant_llm_training_b200.data.datasets import (BatchStream, CollatorForCLM, IterableParquetDataset,
                                                            ParquetDataset)
from fault_tolerant_llm_training_b200.data.synthetic import write_synthetic_parquet
from fault_tolerant_llm_training_b200.data.tokenizer import OfflineTokenizer


def _corpus(tmp_path, **kw):
    return write_synthetic_parquet(str(tmp_path / "d.parquet"), num_docs=16, min_len=5, max_len=60,
                                   vocab_size=512, **kw)


def test_offline_tokenizer_contract():
    tok = OfflineTokenizer(512)
    out = tok("t5 t6 t7", max_length=8, padding="max_length", truncation=True, padding_side="right")
    assert out["input_ids"] == [1, 9, 10, 11, 3, 3, 3, 3]
    assert tok.encode_plus("t5", max_length=2, truncation=True)["input_ids"] == [1, 9]
    assert tok("hello world")["input_ids"] == tok("hello world")["input_ids"]
    assert all(0 <= i < 512 for i in tok("some words here")["input_ids"])


def test_padded_dataset_and_collator(tmp_path):
    tok = OfflineTokenizer(512)
    ds = ParquetDataset(_corpus(tmp_path), tok, 32, training_samples=100)
    assert len(ds) == 100
    assert ds[3]["input_ids"] == ds[3 + 16]["input_ids"]  # wraps modulo file length (dataset.py:28)
    inputs, labels = CollatorForCLM(32, tok.pad_token_id)([ds[0], ds[1]])
    assert inputs.shape == labels.shape == (2, 32) and inputs.dtype == torch.long
    full = torch.tensor([ds[0]["input_ids"], ds[1]["input_ids"]])
    assert torch.equal(inputs, full[:, :-1])
    expect = full[:, 1:].clone()
    expect[expect == tok.pad_token_id] = -100
    assert torch.equal(labels, expect)


def test_packed_dataset_masks_bos_and_always_advances(tmp_path):
    tok = OfflineTokenizer(512)
    ds = IterableParquetDataset(_corpus(tmp_path), tok, 16, bos_token_id=tok.bos_token_id)
    last = -1
    for _ in range(20):  # documents longer than seq+1 do not stall it
        x, y = next(ds)
        assert x.shape == y.shape == (16,)
        assert (y[x == tok.bos_token_id] == -100).all()
        assert (y != tok.bos_token_id).all()
        assert ds.current_index > last or len(ds.token_buffer) > 0
        last = ds.current_index


def test_packed_state_roundtrip(tmp_path):
    tok = OfflineTokenizer(512)
    path = _corpus(tmp_path)
    a = IterableParquetDataset(path, tok, 16)
    for _ in range(5):
        next(a)
    st = a.state_dict()
    want = [next(a) for _ in range(4)]
    b = IterableParquetDataset(path, tok, 16)
    b.load_state_dict(st)
    got = [next(b) for _ in range(4)]
    for (x1, y1), (x2, y2) in zip(want, got):
        assert torch.equal(x1, x2) and torch.equal(y1, y2)


def test_batch_stream_resume_and_rank_sharding(tmp_path):
    tok = OfflineTokenizer(512)
    path = _corpus(tmp_path)
    mk = lambda r, w, b: BatchStream(ParquetDataset(path, tok, 16, 10_000), b, CollatorForCLM(16, tok.pad_token_id),
                                     rank=r, world_size=w)
    one = mk(0, 1, 4)
    ranks = [mk(r, 2, 2) for r in range(2)]
    for _ in range(3):  # 2 ranks x batch 2 see exactly what 1 rank x batch 4 sees
        X, Y = next(one)
        parts = [next(s) for s in ranks]
        assert torch.equal(X, torch.cat([p[0] for p in parts]))
        assert torch.equal(Y, torch.cat([p[1] for p in parts]))
    st = one.state_dict()
    want = next(one)
    fresh = mk(0, 1, 4)
    fresh.load_state_dict(st)
    got = next(fresh)
    assert torch.equal(want[0], got[0]) and st["batches_consumed"] == 3
    ff = mk(0, 1, 4)
    ff.fast_forward(3)
    assert torch.equal(next(ff)[0], want[0])
    for s in [one, fresh, ff] + ranks:
        s.close()


def test_batchstream_restore_validates_geometry(tmp_path):
    """Resume with another world size / batch size: padded mode converts the position exactly or refuses;
    packed mode refuses; the same geometry takes this rank's own state."""
    import pytest as _pytest

    from fault_tolerant_llm_training_b200.data.datasets import BatchStream, CollatorForCLM, ParquetDataset
    from fault_tolerant_llm_training_b200.data.synthetic import write_synthetic_parquet
    from fault_tolerant_llm_training_b200.data.tokenizer import OfflineTokenizer

    path = str(tmp_path / "d.parquet")
    write_synthetic_parquet(path, num_docs=64, min_len=40, max_len=80, vocab_size=512)
    tok = OfflineTokenizer(512)
    mk = lambda rank, world, batch: BatchStream(ParquetDataset(path, tok, 32, 4096), batch, CollatorForCLM(32, tok.pad_token_id),
                                                rank=rank, world_size=world, prefetch=0)
    # written by 2 ranks at batch 2 after 6 batches = 24 samples
    saved = []
    for r in range(2):
        s = mk(r, 2, 2)
        s.batches_consumed = 6
        saved.append(s.state_dict())
    same = mk(1, 2, 2)
    same.restore(saved)
    assert same.batches_consumed == 6
    four = mk(3, 4, 2)              # 24 samples = 3 global batches of 8
    four.restore(saved)
    assert four.batches_consumed == 3
    one = mk(0, 1, 4)               # 24 samples = 6 global batches of 4
    one.restore(saved)
    assert one.batches_consumed == 6
    # the converted stream continues exactly where the old geometry stopped: sample 24 comes next
    nxt, _ = next(one)
    ref = mk(0, 1, 1)
    ref.batches_consumed = 24
    want, _ = next(ref)
    assert torch.equal(nxt[0], want[0])
    bad = mk(0, 1, 5)               # 24 is not a multiple of 5
    with _pytest.raises(ValueError, match="whole number of global batches"):
        bad.restore(saved)
    packed_state = [dict(s, kind="packed") for s in saved]
    with _pytest.raises(ValueError, match="packed"):
        mk(0, 1, 4).restore(packed_state)
