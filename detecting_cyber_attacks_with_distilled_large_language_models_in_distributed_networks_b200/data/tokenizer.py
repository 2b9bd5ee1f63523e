"""Uncased BERT WordPiece tokenizer + synthetic vocabulary generator.

The reference tokenises with HF ``DistilBertTokenizer.from_pretrained('./distilbert-base-uncased')`` *per sample,
per epoch, inside* ``Dataset.__getitem__`` (client1.py:36-45, 364) with
``add_special_tokens=True, max_length=128, padding='max_length', truncation=True``.  Here the whole split is
tokenised ONCE into int32 id / length tensors (``data/pack.py``) that then live on the GPU.

The algorithm is BERT's: clean -> lower-case + strip accents -> split on whitespace and punctuation -> greedy
longest-match-first WordPiece with ``##`` continuation pieces -> ``[CLS] ... [SEP]`` + ``[PAD]``.  With a real
``vocab.txt`` it reproduces the HF tokenizer's ids (tested against HF when ``transformers`` is importable); with no
vocabulary on disk (this box has no network) ``synthetic_vocab()`` builds a 30,522-entry BERT-layout vocabulary that
covers the flow-sentence template words and all 1-3 digit number pieces.
"""
from __future__ import annotations

import os
import unicodedata
from typing import Dict, Iterable, List, Optional, Sequence, Tuple

import numpy as np

PAD, UNK, CLS, SEP, MASK = "[PAD]", "[UNK]", "[CLS]", "[SEP]", "[MASK]"

_TEMPLATE_WORDS = (
    "destination port is flow duration microseconds total forward packets are backward length of bytes maximum "
    "packet minimum per second e inf nan the a and to in for on with by at from ddos benign attack network traffic"
).split()


def synthetic_vocab(size: int = 30522) -> List[str]:
    """BERT-base-uncased layout: [PAD]=0, [unused*], [UNK]=100, [CLS]=101, [SEP]=102, [MASK]=103, then characters,
    template words, number pieces, filler."""
    v: List[str] = [PAD] + [f"[unused{i}]" for i in range(99)] + [UNK, CLS, SEP, MASK]
    v += [f"[unused{i}]" for i in range(99, 99 + 895)]  # ids 104..998, as in the real vocab
    seen = set(v)

    def add(tok: str) -> None:
        if tok not in seen and len(v) < size:
            seen.add(tok)
            v.append(tok)

    chars = "!\"#$%&'()*+,-./:;<=>?@[\\]^_`{|}~0123456789abcdefghijklmnopqrstuvwxyz"
    for c in chars:
        add(c)
    for c in "0123456789abcdefghijklmnopqrstuvwxyz":
        add("##" + c)
    for w in _TEMPLATE_WORDS:
        add(w)
    for width in (2, 3):  # every 2- and 3-digit piece, with and without the continuation prefix
        for i in range(10 ** width):
            s = str(i).zfill(width)
            add(s)
            add("##" + s)
    i = 0
    while len(v) < size:
        add(f"tok{i:05d}")
        i += 1
    return v[:size]


def write_vocab(path: str, vocab: Optional[Sequence[str]] = None) -> str:
    vocab = synthetic_vocab() if vocab is None else vocab
    os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
    with open(path, "w", encoding="utf-8") as f:
        f.write("\n".join(vocab) + "\n")
    return path


def _is_punct(ch: str) -> bool:
    cp = ord(ch)
    if 33 <= cp <= 47 or 58 <= cp <= 64 or 91 <= cp <= 96 or 123 <= cp <= 126:
        return True
    return unicodedata.category(ch).startswith("P")


def _is_whitespace(ch: str) -> bool:
    return ch in " \t\n\r" or unicodedata.category(ch) == "Zs"


def _is_control(ch: str) -> bool:
    if ch in "\t\n\r":
        return False
    return unicodedata.category(ch) in ("Cc", "Cf")


def _is_cjk(cp: int) -> bool:
    return (0x4E00 <= cp <= 0x9FFF or 0x3400 <= cp <= 0x4DBF or 0x20000 <= cp <= 0x2A6DF or 0x2A700 <= cp <= 0x2B73F
            or 0x2B740 <= cp <= 0x2B81F or 0x2B820 <= cp <= 0x2CEAF or 0xF900 <= cp <= 0xFAFF
            or 0x2F800 <= cp <= 0x2FA1F)


_NATIVE_LIB = None


def _load_native():
    """``_native/libdcad_tok.so`` (built by ``utils/build.py::build_host``); None when absent or ``DCAD_NATIVE_TOK=0``."""
    global _NATIVE_LIB
    if os.environ.get("DCAD_NATIVE_TOK", "1") == "0":
        return None
    if _NATIVE_LIB is None:
        import ctypes

        path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "_native", "libdcad_tok.so")
        if not os.path.exists(path):
            _NATIVE_LIB = False
            return None
        lib = ctypes.CDLL(path)
        lib.dcad_tok_create.restype = ctypes.c_void_p
        lib.dcad_tok_create.argtypes = [ctypes.c_char_p, ctypes.c_int64, ctypes.c_int, ctypes.c_int]
        lib.dcad_tok_destroy.argtypes = [ctypes.c_void_p]
        lib.dcad_tok_destroy.restype = None
        lib.dcad_tok_vocab_size.argtypes = [ctypes.c_void_p]
        lib.dcad_tok_vocab_size.restype = ctypes.c_int64
        lib.dcad_tok_encode_batch.restype = ctypes.c_int64
        lib.dcad_tok_encode_batch.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_void_p, ctypes.c_int64,
                                              ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
                                              ctypes.c_int]
        _NATIVE_LIB = lib
    return _NATIVE_LIB or None


class WordPieceTokenizer:
    """``encode_batch(texts, max_len)`` -> (ids [N, max_len] int32, lengths [N] int32)."""

    def __init__(self, vocab: Sequence[str], lower_case: bool = True, max_chars_per_word: int = 100):
        self.vocab = list(vocab)
        self.index: Dict[str, int] = {t: i for i, t in enumerate(self.vocab)}
        self.lower = lower_case
        self.max_chars = max_chars_per_word
        self.pad_id, self.unk_id = self.index[PAD], self.index[UNK]
        self.cls_id, self.sep_id = self.index[CLS], self.index[SEP]
        self._cache: Dict[str, List[int]] = {}

    # ---- constructors ----
    @classmethod
    def from_vocab_file(cls, path: str, **kw) -> "WordPieceTokenizer":
        with open(path, encoding="utf-8") as f:
            return cls([line.rstrip("\n") for line in f], **kw)

    @classmethod
    def from_dir_or_synthetic(cls, model_dir: Optional[str] = None, **kw) -> "WordPieceTokenizer":
        """Loads ``<model_dir>/vocab.txt`` when it exists (the reference hard-fails without the directory,
        client1.py:357-361); otherwise falls back to the synthetic vocabulary."""
        if model_dir and os.path.exists(os.path.join(model_dir, "vocab.txt")):
            return cls.from_vocab_file(os.path.join(model_dir, "vocab.txt"), **kw)
        return cls(synthetic_vocab(), **kw)

    @property
    def vocab_size(self) -> int:
        return len(self.vocab)

    # ---- BERT basic tokenizer ----
    def _basic(self, text: str) -> List[str]:
        out_chars = []
        for ch in text:
            cp = ord(ch)
            if cp == 0 or cp == 0xFFFD or _is_control(ch):
                continue
            if _is_cjk(cp):
                out_chars.append(f" {ch} ")
            elif _is_whitespace(ch):
                out_chars.append(" ")
            else:
                out_chars.append(ch)
        words: List[str] = []
        for tok in "".join(out_chars).split():
            if self.lower:
                tok = tok.lower()
                tok = "".join(c for c in unicodedata.normalize("NFD", tok) if unicodedata.category(c) != "Mn")
            cur = []
            for ch in tok:
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

    # ---- WordPiece ----
    def _wordpiece(self, word: str) -> List[int]:
        hit = self._cache.get(word)
        if hit is not None:
            return hit
        if len(word) > self.max_chars:
            ids = [self.unk_id]
        else:
            ids = []
            start, n = 0, len(word)
            while start < n:
                end = n
                cur = None
                while start < end:
                    piece = word[start:end]
                    if start > 0:
                        piece = "##" + piece
                    cur = self.index.get(piece)
                    if cur is not None:
                        break
                    end -= 1
                if cur is None:
                    ids = [self.unk_id]
                    break
                ids.append(cur)
                start = end
        if len(self._cache) < 1_000_000:
            self._cache[word] = ids
        return ids

    def tokenize_ids(self, text: str) -> List[int]:
        ids: List[int] = []
        for w in self._basic(text):
            ids.extend(self._wordpiece(w))
        return ids

    def encode(self, text: str, max_len: int = 128) -> Tuple[List[int], int]:
        body = self.tokenize_ids(text)[: max_len - 2]
        ids = [self.cls_id] + body + [self.sep_id]
        n = len(ids)
        return ids + [self.pad_id] * (max_len - n), n

    # ---- native batch path (csrc/host/tokenizer.cpp) ----
    def _native_handle(self):
        """ctypes handle of the C++ tokenizer for this vocabulary, or None (library not built / switched off)."""
        if getattr(self, "_native", None) is None:
            self._native = False
            lib = _load_native()
            if lib is not None:
                blob = "\n".join(self.vocab).encode("utf-8")
                h = lib.dcad_tok_create(blob, len(blob), 1 if self.lower else 0, self.max_chars)
                if h and lib.dcad_tok_vocab_size(h) == len(self.vocab):
                    self._native = (lib, h)
        return self._native or None

    def __del__(self):  # pragma: no cover - interpreter shutdown order
        nat = getattr(self, "_native", None)
        if nat:
            try:
                nat[0].dcad_tok_destroy(nat[1])
            except Exception:  # noqa: BLE001
                pass

    def encode_batch(self, texts: Iterable[str], max_len: int = 128, threads: int = 0) -> Tuple[np.ndarray, np.ndarray]:
        """Whole-batch encoding.  ASCII sentences go through the multi-threaded C++ tokenizer when it is built; rows
        with non-ASCII characters (and everything, when it is not) take the Python path — identical results."""
        texts = [str(t) for t in texts]
        ids = np.full((len(texts), max_len), self.pad_id, dtype=np.int32)
        lens = np.zeros(len(texts), dtype=np.int32)
        nat = self._native_handle() if texts else None
        if nat is not None:
            import ctypes

            lib, h = nat
            enc = [t.encode("utf-8") for t in texts]
            offsets = np.zeros(len(enc) + 1, dtype=np.int64)
            np.cumsum([len(e) for e in enc], out=offsets[1:])
            blob = b"".join(enc)
            fallback = np.zeros(len(texts), dtype=np.uint8)
            rc = lib.dcad_tok_encode_batch(h, blob, offsets.ctypes.data_as(ctypes.c_void_p), len(texts), max_len,
                                           ids.ctypes.data_as(ctypes.c_void_p), lens.ctypes.data_as(ctypes.c_void_p),
                                           fallback.ctypes.data_as(ctypes.c_void_p), threads)
            if rc >= 0:
                for i in np.nonzero(fallback)[0]:
                    row, n = self.encode(texts[i], max_len)
                    ids[i] = row
                    lens[i] = n
                return ids, lens
        for i, t in enumerate(texts):
            row, n = self.encode(str(t), max_len)
            ids[i] = row
            lens[i] = n
        return ids, lens
