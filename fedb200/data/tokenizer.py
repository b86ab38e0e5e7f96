"""Offline tokenizer.

The reference calls ``AutoTokenizer.from_pretrained(CHECKPOINT)`` with ``truncation=True``
(src/Servercase/server_IID_IMDB.py:72-76).  No vocabulary file is reachable offline, so
words are mapped into the checkpoint's id space with a stable hash; the special ids match
BERT-cased ([PAD]=0, [UNK]=100, [CLS]=101, [SEP]=102) so batches have the same structure
(``[CLS] … [SEP]``, right-padded with 0).
"""
from __future__ import annotations

import re
import zlib
from typing import Iterable, List

_WORD = re.compile(r"[A-Za-z0-9]+|[^\sA-Za-z0-9]")


class HashTokenizer:
    pad_id, unk_id, cls_id, sep_id = 0, 100, 101, 102
    first_word_id = 1000

    def __init__(self, vocab_size: int = 28996, max_length: int = 512, lowercase: bool = False,
                 max_word_pieces: int = 3, piece_len: int = 6):
        assert vocab_size > self.first_word_id + 16
        self.vocab_size, self.max_length = vocab_size, max_length
        self.lowercase, self.max_word_pieces, self.piece_len = lowercase, max_word_pieces, piece_len

    def _piece_id(self, piece: str, cont: bool) -> int:
        h = zlib.crc32((("##" if cont else "") + piece).encode("utf-8"))
        return self.first_word_id + h % (self.vocab_size - self.first_word_id)

    def tokenize_words(self, text: str) -> List[int]:
        if self.lowercase:
            text = text.lower()
        out: List[int] = []
        for w in _WORD.findall(text):
            # long words are split into word-piece-like chunks so token counts resemble BERT's
            pieces = [w[i:i + self.piece_len] for i in range(0, len(w), self.piece_len)][: self.max_word_pieces]
            for j, p in enumerate(pieces):
                out.append(self._piece_id(p, j > 0))
        return out

    def encode(self, text: str, truncation: bool = True) -> List[int]:
        body = self.tokenize_words(text)
        if truncation:
            body = body[: self.max_length - 2]
        return [self.cls_id] + body + [self.sep_id]

    def __call__(self, texts: Iterable[str], truncation: bool = True) -> List[List[int]]:
        return [self.encode(t, truncation) for t in texts]


# ---------------------------------------------------------------------------------------------------------------
# WordPiece from a local ``vocab.txt`` (what ``AutoTokenizer.from_pretrained(CHECKPOINT)`` gives for BERT / BioBERT,
# src/Servercase/server_IID_IMDB.py:72-76): BERT's basic tokenisation (whitespace + punctuation split, optional
# lower-casing with accent stripping, CJK characters isolated) followed by greedy longest-match-first word pieces with the
# ``##`` continuation prefix.  Ids are the line numbers of ``vocab.txt``, so they index the checkpoint's embedding table.
# ---------------------------------------------------------------------------------------------------------------
import json
import os
import unicodedata
from typing import Dict, Optional


def _is_punct(ch: str) -> bool:
    cp = ord(ch)
    if 33 <= cp <= 47 or 58 <= cp <= 64 or 91 <= cp <= 96 or 123 <= cp <= 126:
        return True
    return unicodedata.category(ch).startswith("P")


def _is_cjk(cp: int) -> bool:
    return (0x4E00 <= cp <= 0x9FFF or 0x3400 <= cp <= 0x4DBF or 0x20000 <= cp <= 0x2A6DF or 0x2A700 <= cp <= 0x2B73F
            or 0x2B740 <= cp <= 0x2B81F or 0x2B820 <= cp <= 0x2CEAF or 0xF900 <= cp <= 0xFAFF or 0x2F800 <= cp <= 0x2FA1F)


class WordPieceTokenizer:
    """``tokenizer(text, truncation=True)`` of a BERT-family checkpoint, from its ``vocab.txt`` alone."""

    def __init__(self, vocab: Dict[str, int], lowercase: bool = False, max_length: int = 512,
                 max_chars_per_word: int = 100, strip_accents: Optional[bool] = None):
        self.vocab, self.lowercase, self.max_length = vocab, lowercase, max_length
        self.max_chars_per_word = max_chars_per_word
        self.strip_accents = lowercase if strip_accents is None else strip_accents
        self.vocab_size = max(vocab.values()) + 1
        for name, tok in (("pad_id", "[PAD]"), ("unk_id", "[UNK]"), ("cls_id", "[CLS]"), ("sep_id", "[SEP]")):
            if tok not in vocab:
                raise KeyError(f"vocab.txt has no {tok} entry")
            setattr(self, name, vocab[tok])
        self._never_split = {"[PAD]", "[UNK]", "[CLS]", "[SEP]", "[MASK]"}

    @classmethod
    def from_pretrained(cls, path: str, max_length: int = 512) -> "WordPieceTokenizer":
        """``path``: a checkpoint directory containing ``vocab.txt`` (and optionally ``tokenizer_config.json``), or the
        vocabulary file itself.  Cased / uncased comes from ``do_lower_case`` in the config; without a config the
        vocabulary decides (a cased vocabulary contains upper-case word starts)."""
        vocab_file = os.path.join(path, "vocab.txt") if os.path.isdir(path) else path
        vocab: Dict[str, int] = {}
        with open(vocab_file, encoding="utf-8") as fh:
            for i, line in enumerate(fh):
                vocab[line.rstrip("\n")] = i
        lower = None
        cfg = os.path.join(os.path.dirname(vocab_file), "tokenizer_config.json")
        if os.path.exists(cfg):
            with open(cfg) as fh:
                tc = json.load(fh)
            lower = tc.get("do_lower_case")
            max_length = min(max_length, int(tc.get("model_max_length", max_length) or max_length))
        if lower is None:
            lower = not any(t[:1].isupper() for t in vocab if t[:1].isalpha() and not t.startswith("["))
        return cls(vocab, lowercase=bool(lower), max_length=max_length)

    # ---- BERT BasicTokenizer ------------------------------------------------------------------------------
    def _basic(self, text: str) -> List[str]:
        out = []
        for ch in text:                                   # clean: drop control characters, normalise whitespace
            cp = ord(ch)
            if cp == 0 or cp == 0xFFFD or (unicodedata.category(ch) in ("Cc", "Cf") and ch not in "\t\n\r"):
                continue
            if _is_cjk(cp):
                out.append(f" {ch} ")
            else:
                out.append(" " if ch in "\t\n\r" or unicodedata.category(ch) == "Zs" else ch)
        words: List[str] = []
        for w in "".join(out).split():
            if w in self._never_split:
                words.append(w)
                continue
            if self.lowercase:
                w = w.lower()
            if self.strip_accents:
                w = "".join(c for c in unicodedata.normalize("NFD", w) if unicodedata.category(c) != "Mn")
            cur = []
            for ch in w:                                  # every punctuation character is its own token
                if _is_punct(ch):
                    if cur:
                        words.append("".join(cur))
                        cur = []
                    words.append(ch)
                else:
                    cur.append(ch)
            if cur:
                words.append("".join(cur))
        return words

    # ---- greedy longest-match-first word pieces ---------------------------------------------------------------
    def _wordpiece(self, word: str) -> List[int]:
        if len(word) > self.max_chars_per_word:
            return [self.unk_id]
        ids, start, n = [], 0, len(word)
        while start < n:
            end, cur = n, None
            while start < end:
                piece = word[start:end] if start == 0 else "##" + word[start:end]
                if piece in self.vocab:
                    cur = self.vocab[piece]
                    break
                end -= 1
            if cur is None:
                return [self.unk_id]                      # HF: the whole word becomes [UNK]
            ids.append(cur)
            start = end
        return ids

    def tokenize_words(self, text: str) -> List[int]:
        out: List[int] = []
        for w in self._basic(text):
            out.extend([self.vocab[w]] if w in self._never_split else self._wordpiece(w))
        return out

    def encode(self, text: str, truncation: bool = True) -> List[int]:
        body = self.tokenize_words(text)
        if truncation:
            body = body[: self.max_length - 2]
        return [self.cls_id] + body + [self.sep_id]

    def __call__(self, texts: Iterable[str], truncation: bool = True) -> List[List[int]]:
        return [self.encode(t, truncation) for t in texts]


def load_tokenizer(checkpoint: Optional[str], vocab_size: int, max_length: int = 512):
    """WordPiece when ``checkpoint`` (a directory or a vocab file) provides a ``vocab.txt``; the hash tokenizer otherwise
    (no vocabulary reachable offline — ids then do not correspond to any pretrained embedding table)."""
    if checkpoint:
        vf = os.path.join(checkpoint, "vocab.txt") if os.path.isdir(checkpoint) else checkpoint
        if os.path.exists(vf) and vf.endswith(".txt"):
            return WordPieceTokenizer.from_pretrained(checkpoint, max_length)
    return HashTokenizer(vocab_size=vocab_size, max_length=max_length)
