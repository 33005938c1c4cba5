#!/usr/bin/env python3
"""Generate the committed golden fixtures from the UPSTREAM modules that are importable in the build container
(transformers 5.5.0; the reference pins 4.44.2 at cluster-config/apps/sd15-api/configmap.yaml:12):

  tokenizer.json  — input_ids of transformers.CLIPTokenizer (padding="max_length", max_length=77, truncation=True) on
                    a prompt corpus, for the deterministic synthetic vocabularies of size 49408 and 1000
  clip_tiny.pt    — transformers.CLIPTextModel (tiny config, weights from weights.random_state_dict seed 2)
                    last_hidden_state for two token rows

Run from the repo root:  python tests/golden/make_golden.py     (needs `transformers`; not needed to run the tests)
"""
import json
import os
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
HERE = os.path.dirname(os.path.abspath(__file__))

import torch  # noqa: E402

CORPUS = [
    "A Panda  on the\tMotorbike!", "", "panda " * 100,
    "a photo of an astronaut riding a horse on mars, 4k, highly-detailed",
    "it's the cat's 2nd birthday; don't!!", "naïve café Ünïcode – test", "  leading and trailing   ",
    "UPPER lower MiXeD 12345 6.7", "日本語 text", "a<|endoftext|>b", "é composed", "we'll they've I'm you're he'd",
    "a cabin in the forest at night, detailed", "a photo of a panda riding a motorbike",
    "x" * 300, "3.14159 2024-01-01 #hashtag @user $100 50%", "line one\nline two\r\nline three",
]


def main():
    import b200sd
    from oracle.tokenizer import OracleTokenizer
    T, W = b200sd.tokenizer, b200sd.weights
    gold = {"corpus": CORPUS, "ids": {}}
    for vs in (49408, 1000):
        d = tempfile.mkdtemp()
        vf, mf = T.write_vocab_files(d, vs)
        gold["ids"][str(vs)] = OracleTokenizer(vf, mf)(CORPUS)
    with open(os.path.join(HERE, "tokenizer.json"), "w", encoding="utf-8") as f:
        json.dump(gold, f, ensure_ascii=False)

    from transformers import CLIPTextConfig, CLIPTextModel
    cfg = W.load_config("tiny")["text_encoders"][0]
    sd = W.random_state_dict(W.clip_text_schema(cfg), 2)
    hf = CLIPTextModel(CLIPTextConfig(
        vocab_size=cfg["vocab_size"], hidden_size=cfg["hidden_size"], intermediate_size=cfg["intermediate_size"],
        num_hidden_layers=cfg["num_layers"], num_attention_heads=cfg["num_heads"], max_position_embeddings=77,
        hidden_act="quick_gelu", layer_norm_eps=cfg["eps"], bos_token_id=998, eos_token_id=999, pad_token_id=999)).eval()
    print(hf.load_state_dict(sd, strict=False))
    ids = torch.tensor([gold["ids"]["1000"][13], gold["ids"]["1000"][1]], dtype=torch.long)
    with torch.no_grad():
        out = hf(input_ids=ids).last_hidden_state
    torch.save({"ids": ids, "last_hidden_state": out, "weights_seed": 2}, os.path.join(HERE, "clip_tiny.pt"))
    print("wrote", os.listdir(HERE))


if __name__ == "__main__":
    main()
