"""CLIP byte-pair tokenizer (77 tokens, BOS 49406 / EOS 49407, EOS padding).

With ``vocab.json`` + ``merges.txt`` from the upstream checkpoint directory this reproduces the
OpenAI CLIP BPE (lower-case, whitespace clean, ``</w>`` word ends). Offline — no checkpoint files can
be fetched here — ``HashTokenizer`` maps each word to a stable id with SHA-256 so prompts still
produce deterministic, prompt-dependent token streams for random-init benchmarking.
"""
from __future__ import annotations

import hashlib
import json
import re
from functools import lru_cache
from pathlib import Path
from typing import Dict, List, Optional, Tuple

BOS, EOS, MAX_LEN = 49406, 49407, 77
_WORD_RE = re.compile(r"<\|startoftext\|>|<\|endoftext\|>|'s|'t|'re|'ve|'m|'ll|'d|[a-zA-Z]+|[0-9]|[^\s a-zA-Z0-9]+")


@lru_cache()
def _bytes_to_unicode() -> Dict[int, str]:
    bs = list(range(ord("!"), ord("~") + 1)) + list(range(ord("¡"), ord("¬") + 1)) + list(range(ord("®"), ord("ÿ") + 1))
    cs = bs[:]
    n = 0
    for b in range(256):
        if b not in bs:
            bs.append(b)
            cs.append(256 + n)
            n += 1
    return dict(zip(bs, [chr(c) for c in cs]))


class ClipBPETokenizer:
    def __init__(self, vocab_path: str | Path, merges_path: str | Path, max_len: int = MAX_LEN):
        self.encoder: Dict[str, int] = json.loads(Path(vocab_path).read_text(encoding="utf-8"))
        lines = Path(merges_path).read_text(encoding="utf-8").split("\n")
        merges = [tuple(l.split()) for l in lines[1:] if l and not l.startswith("#")]
        self.ranks: Dict[Tuple[str, str], int] = {m: i for i, m in enumerate(merges) if len(m) == 2}
        self.byte_enc = _bytes_to_unicode()
        self.max_len = max_len
        self.bos = self.encoder.get("<|startoftext|>", BOS)
        self.eos = self.encoder.get("<|endoftext|>", EOS)
        self._cache: Dict[str, List[str]] = {}

    def _bpe(self, token: str) -> List[str]:
        if token in self._cache:
            return self._cache[token]
        word = list(token[:-1]) + [token[-1] + "</w>"]
        while len(word) > 1:
            pairs = [(word[i], word[i + 1]) for i in range(len(word) - 1)]
            best = min(pairs, key=lambda p: self.ranks.get(p, 1 << 30))
            if best not in self.ranks:
                break
            merged, i = [], 0
            while i < len(word):
                if i < len(word) - 1 and (word[i], word[i + 1]) == best:
                    merged.append(word[i] + word[i + 1])
                    i += 2
                else:
                    merged.append(word[i])
                    i += 1
            word = merged
        self._cache[token] = word
        return word

    def encode(self, text: str) -> List[int]:
        text = re.sub(r"\s+", " ", text.strip()).lower()
        ids: List[int] = []
        for tok in _WORD_RE.findall(text):
            tok = "".join(self.byte_enc[b] for b in tok.encode("utf-8"))
            ids.extend(self.encoder[t] for t in self._bpe(tok) if t in self.encoder)
        ids = ids[: self.max_len - 2]
        out = [self.bos] + ids + [self.eos]
        return out + [self.eos] * (self.max_len - len(out))


class HashTokenizer:
    """Checkpoint-free stand-in with the same framing (BOS, tokens, EOS, EOS padding)."""

    def __init__(self, vocab_size: int = 49408, max_len: int = MAX_LEN):
        self.vocab_size, self.max_len = vocab_size, max_len
        self.bos, self.eos = vocab_size - 2, vocab_size - 1

    def encode(self, text: str) -> List[int]:
        words = _WORD_RE.findall(re.sub(r"\s+", " ", text.strip()).lower())
        ids = [int.from_bytes(hashlib.sha256(w.encode()).digest()[:4], "big") % (self.vocab_size - 2) for w in words]
        ids = ids[: self.max_len - 2]
        out = [self.bos] + ids + [self.eos]
        return out + [self.eos] * (self.max_len - len(out))


class XlmrTokenizer:
    """XLM-RoBERTa tokenizer of Kandinsky 2.1's MultilingualCLIP text encoder: the upstream ``sentencepiece.bpe.model``
    run by the SentencePiece library, with fairseq's id layout on top — <s> 0, <pad> 1, </s> 2, <unk> 3, every piece id
    shifted by one — framed as <s> pieces </s> and padded with <pad> to 77 (reference: the container behind
    templates/kandinsky2.json tokenises with ``max_length=77, padding='max_length', truncation=True``)."""
    BOS, PAD, EOS, UNK = 0, 1, 2, 3

    def __init__(self, model_file: str | Path, max_len: int = MAX_LEN):
        import sentencepiece as spm
        self.sp = spm.SentencePieceProcessor(model_file=str(model_file))
        self.max_len = max_len
        self.vocab_size = self.sp.get_piece_size() + 2          # + the fairseq offset and <mask>
        self.bos, self.eos = self.BOS, self.EOS

    def encode(self, text: str) -> List[int]:
        ids = [self.UNK if i == self.sp.unk_id() else i + 1 for i in self.sp.encode(text)]
        ids = ids[: self.max_len - 2]
        out = [self.BOS] + ids + [self.EOS]
        return out + [self.PAD] * (self.max_len - len(out))


def load_xlmr_tokenizer(model_dir: Optional[str | Path], vocab_size: int, max_len: int = MAX_LEN):
    """``sentencepiece.bpe.model`` next to the checkpoints (or under ``text_encoder/``), else the hash stand-in."""
    if model_dir is not None:
        d = Path(model_dir)
        for cand in (d / "sentencepiece.bpe.model", d / "text_encoder" / "sentencepiece.bpe.model"):
            if cand.exists():
                return XlmrTokenizer(cand, max_len)
    return HashTokenizer(vocab_size, max_len)


def load_tokenizer(model_dir: Optional[str | Path] = None, vocab_size: int = 49408):
    if model_dir is not None:
        d = Path(model_dir)
        for sub in ("tokenizer", "."):
            v, m = d / sub / "vocab.json", d / sub / "merges.txt"
            if v.exists() and m.exists():
                return ClipBPETokenizer(v, m)
    return HashTokenizer(vocab_size)
