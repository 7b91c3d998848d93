"""Deterministic synthetic text embedder.

The embedding model is out of scope (BASELINE.json north_star): vectors are synthetic and
deterministic.  `store_semantic(text, ...)` has no slot for a vector (reference signature,
pilott/memory/enhanced_memory.py:60-67), so when the caller passes no ``embedding=`` the
text is embedded here: every lower-cased character trigram is hashed (blake2b, NOT Python's
salted hash()) to a fixed pseudo-random +-1 direction and the directions are summed.  Texts
sharing substrings therefore have positive cosine similarity, which keeps the reference's
"query is contained in text" call sites meaningful, and identical texts embed identically.
"""
from __future__ import annotations

import hashlib
from functools import lru_cache

import numpy as np


class HashEmbedder:
    def __init__(self, dim: int = 768, ngram: int = 3):
        self.dim, self.ngram = int(dim), int(ngram)
        self._gram = lru_cache(maxsize=65536)(self._gram_uncached)

    def _gram_uncached(self, g: str) -> np.ndarray:
        need = (self.dim + 7) // 8
        buf = b""
        ctr = 0
        while len(buf) < need:
            buf += hashlib.blake2b(g.encode("utf-8"), digest_size=64, salt=ctr.to_bytes(8, "little")).digest()
            ctr += 1
        bits = np.unpackbits(np.frombuffer(buf[:need], dtype=np.uint8))[: self.dim]
        return bits.astype(np.float32) * 2.0 - 1.0

    def __call__(self, text: str) -> np.ndarray:
        t = " " + text.lower() + " "
        n = self.ngram
        v = np.zeros(self.dim, dtype=np.float32)
        for i in range(max(1, len(t) - n + 1)):
            v += self._gram(t[i:i + n])
        nrm = float(np.linalg.norm(v))
        return v / nrm if nrm > 0 else v

    def batch(self, texts) -> np.ndarray:
        return np.stack([self(t) for t in texts]) if len(texts) else np.zeros((0, self.dim), np.float32)
