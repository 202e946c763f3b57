"""Tokenizers (ref: lit_llama/tokenizer.py:9-88).  CPU string work, outside the hot path; thin
wrappers so ``from lit_llama import Tokenizer, HFTokenizer`` keeps working.  The backing
libraries are imported lazily."""
from pathlib import Path
from typing import Optional

import torch


class Tokenizer:
    """SentencePiece tokenizer (ref: tokenizer.py:9-49)."""

    def __init__(self, model_path: Path) -> None:
        from sentencepiece import SentencePieceProcessor

        self.processor = SentencePieceProcessor(model_file=str(model_path))
        self.bos_id = self.processor.bos_id()
        self.eos_id = self.processor.eos_id()
        self.pad_id = self.processor.pad_id()

    @property
    def vocab_size(self) -> int:
        return self.processor.vocab_size()

    def encode(self, string: str, bos: bool = True, eos: bool = False, max_length: int = -1, pad: bool = False,
               device: Optional[torch.device] = None) -> torch.Tensor:
        ids = self.processor.encode(string)
        ids = ([self.bos_id] if bos else []) + ids + ([self.eos_id] if eos else [])
        if max_length > 0:
            ids = ids[:max_length]
        if pad and len(ids) < max_length:
            ids = ids + [self.pad_id] * (max_length - len(ids))
        return torch.tensor(ids, dtype=torch.int, device=device)

    def decode(self, tokens: torch.Tensor) -> str:
        return self.processor.decode(tokens.tolist())


class HFTokenizer:
    """HuggingFace ``tokenizers`` JSON tokenizer (ref: tokenizer.py:51-88; the fork's addition)."""

    def __init__(self, model_path: Path) -> None:
        from tokenizers import Tokenizer as _HF

        self.processor = _HF.from_file(str(model_path))
        # the fork's vocabulary starts ["<PAD>", "<BOS>", "<EOS>", "<UNK>", "<MASK>"]: fixed ids (ref tokenizer.py:56-61)
        self.bos_id = 1
        self.eos_id = 2
        self.pad_id = 0

    @property
    def vocab_size(self) -> int:
        return self.processor.get_vocab_size()

    def encode(self, string: str, bos: bool = True, eos: bool = False, max_length: int = -1, pad: bool = False,
               device: Optional[torch.device] = None) -> torch.Tensor:
        ids = self.processor.encode(string).ids
        if bos:
            ids = [self.bos_id] + ids
        if eos:
            ids = ids + [self.eos_id]
        if max_length > 0:
            ids = ids[:max_length]
        if pad and len(ids) < max_length:
            ids = ids + [self.pad_id] * (max_length - len(ids))
        return torch.tensor(ids, dtype=torch.int, device=device)

    def decode(self, tokens: torch.Tensor) -> str:
        return self.processor.decode(tokens.tolist())
