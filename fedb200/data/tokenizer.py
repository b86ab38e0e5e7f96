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
