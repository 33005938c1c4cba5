#!/usr/bin/env python3
"""Generate the committed golden fixtures from the UPSTREAM modules that are importable in the build container
(transformers 5.5.0; the reference pins 4.44.2 at cluster-config/apps/sd15-api/configmap.yaml:12):

  tokenizer.json  — input_ids of transformers.CLIPTokenizer (padding="max_length", max_length=77, truncation=True) on
                    a prompt corpus, for the deterministic synthetic vocabularies of size 49408 and 1000
  clip_tiny.pt    — transformers.CLIPTextModel (tiny config, weights from weights.random_state_dict seed 2)
                    last_hidden_state for two token rows
  clip_xl2_tiny.pt — transformers.CLIPTextModelWithProjection (tiny_xl second encoder, seed 3): hidden_states[-2] and
                    text_embeds for two token rows padded with id 0

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

    # SDXL's second encoder: CLIPTextModelWithProjection (gelu MLP), hidden_states[-2] and text_embeds, with the
    # second tokenizer's padding (id 0) so that the EOS position - not the padding - decides the pooled row
    from transformers import CLIPTextModelWithProjection
    cfg2 = W.load_config("tiny_xl")["text_encoders"][1]
    sd2 = W.random_state_dict(W.clip_text_schema(cfg2), 3)
    hf2 = CLIPTextModelWithProjection(CLIPTextConfig(
        vocab_size=cfg2["vocab_size"], hidden_size=cfg2["hidden_size"], intermediate_size=cfg2["intermediate_size"],
        num_hidden_layers=cfg2["num_layers"], num_attention_heads=cfg2["num_heads"], max_position_embeddings=77,
        hidden_act="gelu", layer_norm_eps=cfg2["eps"], projection_dim=cfg2["projection_dim"], bos_token_id=998,
        eos_token_id=999, pad_token_id=0)).eval()
    print(hf2.load_state_dict(sd2, strict=False))
    ids2 = ids.clone()
    for r in range(ids2.shape[0]):
        eos = int((ids2[r] == 999).int().argmax())
        ids2[r, eos + 1:] = 0
    with torch.no_grad():
        o2 = hf2(input_ids=ids2, output_hidden_states=True)
    torch.save({"ids": ids2, "penultimate": o2.hidden_states[-2], "text_embeds": o2.text_embeds, "weights_seed": 3},
               os.path.join(HERE, "clip_xl2_tiny.pt"))
    print("wrote", os.listdir(HERE))


if __name__ == "__main__":
    main()
