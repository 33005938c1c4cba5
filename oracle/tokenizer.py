"""Oracle tokenizer — the upstream `transformers.CLIPTokenizer` itself (installed 5.5.0; the reference pins 4.44.2 at
configmap.yaml:12), constructed offline from a vocab dict + merges list (vocab.json / merges.txt contents).  Called
the way StableDiffusionPipeline.encode_prompt calls it behind the reference's `pipe(prompt, ...)`
(configmap.yaml:105-106): `tokenizer(prompt, padding="max_length", max_length=77, truncation=True)`.
TEST INFRASTRUCTURE — see oracle/__init__.py."""
from __future__ import annotations

import json


class OracleTokenizer:
    def __init__(self, vocab_file: str, merges_file: str):
        from transformers import CLIPTokenizer
        with open(vocab_file, encoding="utf-8") as f:
            vocab = json.load(f)
        with open(merges_file, encoding="utf-8") as f:
            merges = [tuple(l.split()) for l in f.read().strip().split("\n")[1:]]
        self.tok = CLIPTokenizer(vocab=vocab, merges=merges)

    def __call__(self, prompts, max_length: int = 77):
        if isinstance(prompts, str):
            prompts = [prompts]
        out = self.tok(prompts, padding="max_length", max_length=max_length, truncation=True)
        return [list(map(int, ids)) for ids in out["input_ids"]]
