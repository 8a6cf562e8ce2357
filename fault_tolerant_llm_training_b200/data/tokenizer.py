"""Tokenizer loading with an offline fallback.

The reference loads `unsloth/Mistral-Nemo-Base-2407-bnb-4bit` from the HF hub
(/root/reference/train.py:28, utils.py:132-137) and calls `encode_plus`
(dataset.py:29, 84), which transformers 5.x no longer has.  Here a
tokenizer is anything callable as `tok(text, max_length=, padding=, truncation=)`
that returns {"input_ids": [...]} and has `vocab_size`, `pad_token_id`,
`bos_token_id`.  `OfflineTokenizer` is a dependency-free word-hashing tokenizer
with the Mistral-Nemo vocabulary size (131072), used when the hub is unreachable
or when `--tokenizer-name-or-path offline` is given.
"""
from __future__ import annotations

import re
import zlib
from typing import Dict, List

_WORD = re.compile(r"\S+")


class OfflineTokenizer:
    """Deterministic word-level hashing tokenizer.

    ids 0..3 are reserved (0 = unk, 1 = bos, 2 = eos, 3 = pad) to mirror the
    bos_token_id=1 default of the reference's IterableParquetDataset
    (dataset.py:58).  A token of the form ``t<int>`` maps to that integer
    directly (mod vocab), which lets synthetic corpora hit exact token ids.
    """

    def __init__(self, vocab_size: int = 131072):
        if vocab_size < 8:
            raise ValueError("vocab_size too small")
        self.vocab_size = vocab_size
        self.unk_token_id, self.bos_token_id, self.eos_token_id, self.pad_token_id = 0, 1, 2, 3
        self.name_or_path = "offline"

    def _word_id(self, w: str) -> int:
        if len(w) > 1 and w[0] == "t" and w[1:].isdigit():
            return 4 + int(w[1:]) % (self.vocab_size - 4)
        return 4 + zlib.crc32(w.encode("utf-8")) % (self.vocab_size - 4)

    def encode(self, text: str, add_bos: bool = True) -> List[int]:
        ids = [self.bos_token_id] if add_bos else []
        ids.extend(self._word_id(w) for w in _WORD.findall(text))
        return ids

    def __call__(self, text: str, max_length: int | None = None, padding=False,
                 truncation: bool = False, padding_side: str = "right", **_) -> Dict[str, List[int]]:
        ids = self.encode(text)
        if truncation and max_length is not None:
            ids = ids[:max_length]
        mask = [1] * len(ids)
        if padding == "max_length" and max_length is not None and len(ids) < max_length:
            pad = [self.pad_token_id] * (max_length - len(ids))
            zeros = [0] * len(pad)
            if padding_side == "left":
                ids, mask = pad + ids, zeros + mask
            else:
                ids, mask = ids + pad, mask + zeros
        return {"input_ids": ids, "attention_mask": mask}

    # kept so that code written against transformers<5 (the reference) also works
    encode_plus = __call__

    def decode(self, ids) -> str:
        return " ".join(f"t{int(i) - 4}" if int(i) >= 4 else f"<{int(i)}>" for i in ids)


def load_tokenizer(name_or_path: str, vocab_size: int = 0, log=None):
    """HF tokenizer when it can be loaded locally, else the offline tokenizer."""
    if name_or_path not in ("offline", "", None):
        try:
            import os

            os.environ.setdefault("HF_HUB_OFFLINE", "1")  # never hang on DNS
            from transformers import AutoTokenizer

            tok = AutoTokenizer.from_pretrained(name_or_path)
            if tok.pad_token_id is None:
                tok.pad_token_id = tok.eos_token_id
            return tok
        except Exception as e:  # no hub, no cache
            if log is not None:
                log.info(f"Tokenizer '{name_or_path}' unavailable ({type(e).__name__}); "
                         f"using the offline tokenizer")
    return OfflineTokenizer(vocab_size or 131072)
