"""CLIP byte-level BPE tokenizer (host side, integer-exact).

Replaces `CLIPTokenizer` as used by StableDiffusionPipeline.encode_prompt —
`tokenizer(prompt, padding="max_length", max_length=77, truncation=True)` — behind the reference's `pipe(prompt, ...)`
call (cluster-config/apps/sd15-api/configmap.yaml:105-106; transformers pinned at :12).  SURVEY.md §8(a) a6, A.4.

The real `vocab.json` / `merges.txt` are model-repo assets that are not available offline; when a directory holding
them is given they are used unchanged, otherwise a deterministic synthetic vocabulary with the real file's layout
(256 byte symbols, 256 byte symbols + "</w>", merges, BOS, EOS; padded to the model's vocab size) is built.
"""
from __future__ import annotations

import json
import os
import unicodedata
from collections import Counter
from functools import lru_cache

import regex

PAT = regex.compile(
    r"""<\|startoftext\|>|<\|endoftext\|>|'s|'t|'re|'ve|'m|'ll|'d|[\p{L}]+|[\p{N}]|[^\s\p{L}\p{N}]+""", regex.IGNORECASE)
BOS, EOS = "<|startoftext|>", "<|endoftext|>"

_WORDS = ("a an the of on in at with and or by for from to photo picture painting portrait drawing sketch render "
          "panda motorbike motorcycle riding cabin forest mountain lake river city street night day sunset sunrise "
          "astronaut horse cat dog robot dragon castle ocean beach desert snow rain cloud sky star moon sun tree flower "
          "red green blue yellow purple orange black white golden silver detailed realistic cinematic lighting sharp "
          "focus high quality resolution synthetic subject variant studio oil watercolor digital art concept fantasy "
          "science fiction ancient modern futuristic small large tiny huge beautiful dramatic colorful dark bright "
          "wooden stone metal glass water fire smoke light shadow portrait landscape macro wide angle lens").split()


@lru_cache()
def bytes_to_unicode() -> dict:
    """GPT-2 / CLIP reversible byte -> printable unicode map."""
    bs = list(range(ord("!"), ord("~") + 1)) + list(range(ord("¡"), ord("¬") + 1)) + list(range(ord("®"), ord("ÿ") + 1))
    cs = bs[:]
    n = 0
    for b in range(256):
        if b not in bs:
            bs.append(b)
            cs.append(256 + n)
            n += 1
    return dict(zip(bs, map(chr, cs)))


def _train_merges(num_merges: int) -> list:
    """Deterministic BPE training over a fixed word list (count = rank-weighted), ties broken lexicographically."""
    enc = bytes_to_unicode()
    words = Counter()
    for i, w in enumerate(_WORDS):
        sym = [enc[b] for b in w.encode("utf-8")]
        sym[-1] += "</w>"
        words[tuple(sym)] += 1 + (len(_WORDS) - i)
    merges = []
    while len(merges) < num_merges:
        pairs = Counter()
        for w, c in words.items():
            for a, b in zip(w, w[1:]):
                pairs[(a, b)] += c
        if not pairs:
            break
        best = min(pairs.items(), key=lambda kv: (-kv[1], kv[0]))[0]
        merges.append(best)
        new_words = Counter()
        for w, c in words.items():
            out, i = [], 0
            while i < len(w):
                if i + 1 < len(w) and (w[i], w[i + 1]) == best:
                    out.append(w[i] + w[i + 1])
                    i += 2
                else:
                    out.append(w[i])
                    i += 1
            new_words[tuple(out)] += c
        words = new_words
    return merges


@lru_cache()
def synthetic_vocab(vocab_size: int):
    """(vocab: token -> id, merges: list[(a, b)]) with the real CLIP file layout, BOS = V-2, EOS = V-1."""
    assert vocab_size >= 512 + 2
    base = list(bytes_to_unicode().values())
    tokens = base + [t + "</w>" for t in base]
    merges = _train_merges(min(400, vocab_size - 514))
    tokens += ["".join(m) for m in merges]
    i = 0
    while len(tokens) < vocab_size - 2:
        tokens.append(f"<|unused{i}|>")
        i += 1
    tokens += [BOS, EOS]
    assert len(tokens) == vocab_size and len(set(tokens)) == vocab_size
    return {t: i for i, t in enumerate(tokens)}, merges


def write_vocab_files(directory: str, vocab_size: int):
    """Write vocab.json / merges.txt in the upstream file format (used to build the transformers tokenizer in tests)."""
    vocab, merges = synthetic_vocab(vocab_size)
    os.makedirs(directory, exist_ok=True)
    with open(os.path.join(directory, "vocab.json"), "w", encoding="utf-8") as f:
        json.dump(vocab, f, ensure_ascii=False)
    with open(os.path.join(directory, "merges.txt"), "w", encoding="utf-8") as f:
        f.write("#version: 0.2\n")
        for a, b in merges:
            f.write(f"{a} {b}\n")
    return os.path.join(directory, "vocab.json"), os.path.join(directory, "merges.txt")


class CLIPTokenizer:
    model_max_length = 77

    def __init__(self, vocab: dict, merges: list):
        self.encoder = vocab
        self.decoder = {v: k for k, v in vocab.items()}
        self.bpe_ranks = {tuple(m): i for i, m in enumerate(merges)}
        self.byte_encoder = bytes_to_unicode()
        self.bos_token_id = vocab[BOS]
        self.eos_token_id = vocab[EOS]
        self.pad_token_id = self.eos_token_id  # CLIP pads with <|endoftext|>
        self.unk_token_id = self.eos_token_id
        self._cache = {BOS: BOS, EOS: EOS}

    @classmethod
    def synthetic(cls, vocab_size: int = 49408):
        return cls(*synthetic_vocab(vocab_size))

    @classmethod
    def from_files(cls, directory: str):
        with open(os.path.join(directory, "vocab.json"), encoding="utf-8") as f:
            vocab = json.load(f)
        with open(os.path.join(directory, "merges.txt"), encoding="utf-8") as f:
            lines = f.read().strip().split("\n")[1:]
        merges = [tuple(l.split()) for l in lines[: 49152 - 256 - 2 + 1]]
        return cls(vocab, merges)

    @classmethod
    def from_pretrained_or_synthetic(cls, vocab_size: int, directory: str | None = None):
        directory = directory or os.environ.get("B200SD_CLIP_VOCAB")
        if directory and os.path.exists(os.path.join(directory, "vocab.json")):
            return cls.from_files(directory)
        return cls.synthetic(vocab_size)

    def _bpe(self, token: str) -> str:
        if token in self._cache:
            return self._cache[token]
        word = tuple(token[:-1]) + (token[-1] + "</w>",)
        if len(word) == 1:
            return token + "</w>"
        while True:
            pairs = set(zip(word, word[1:]))
            if not pairs:
                break
            bigram = min(pairs, key=lambda p: self.bpe_ranks.get(p, float("inf")))
            if bigram not in self.bpe_ranks:
                break
            a, b = bigram
            out, i = [], 0
            while i < len(word):
                if i + 1 < len(word) and word[i] == a and word[i + 1] == b:
                    out.append(a + b)
                    i += 2
                else:
                    out.append(word[i])
                    i += 1
            word = tuple(out)
            if len(word) == 1:
                break
        res = " ".join(word)
        self._cache[token] = res
        return res

    def tokenize(self, text: str) -> list:
        text = unicodedata.normalize("NFC", text)
        text = regex.sub(r"\s+", " ", text).strip().lower()
        ids = []
        for tok in PAT.findall(text):
            sym = "".join(self.byte_encoder[b] for b in tok.encode("utf-8"))
            ids.extend(self.encoder.get(t, self.unk_token_id) for t in self._bpe(sym).split(" "))
        return ids

    def encode(self, text: str, max_length: int | None = None) -> list:
        """[BOS] + ids (truncated) + [EOS], padded with EOS to max_length (77)."""
        n = max_length or self.model_max_length
        ids = [self.bos_token_id] + self.tokenize(text)[: n - 2] + [self.eos_token_id]
        return ids + [self.pad_token_id] * (n - len(ids))

    def __call__(self, prompts, max_length: int | None = None):
        if isinstance(prompts, str):
            prompts = [prompts]
        return [self.encode(p, max_length) for p in prompts]
