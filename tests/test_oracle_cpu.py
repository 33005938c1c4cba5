"""CPU tests (no GPU): the oracle against every golden vector / known answer available for this path, and the product's
host-side integer logic (tokenizer, timestep tables) against the oracle.

The reference holds no tests or fixtures of its own; what pins the oracle is
  - tests/golden/tokenizer.json, clip_tiny.pt: outputs of the upstream transformers modules (tests/golden/make_golden.py)
  - the closed-form anchors of SURVEY.md App. C
  - the public parameter counts and the diffusers state-dict key schema.
"""
import json
import math
import os

import pytest
import torch

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def gold():
    with open(os.path.join(HERE, "golden", "tokenizer.json"), encoding="utf-8") as f:
        return json.load(f)


# ------------------------------------------------------------------------------------------------ tokenizer
@pytest.mark.parametrize("vocab_size", [49408, 1000])
def test_product_tokenizer_matches_upstream_golden(b200sd, gold, vocab_size):
    tok = b200sd.tokenizer.CLIPTokenizer.synthetic(vocab_size)
    got = tok(gold["corpus"])
    assert got == gold["ids"][str(vocab_size)]          # bit-exact, every prompt
    for ids in got:
        assert len(ids) == 77 and ids[0] == vocab_size - 2 and ids[-1] == vocab_size - 1


def test_oracle_tokenizer_live_matches_golden(b200sd, gold, tmp_path):
    pytest.importorskip("transformers")
    from oracle.tokenizer import OracleTokenizer
    vf, mf = b200sd.tokenizer.write_vocab_files(str(tmp_path), 49408)
    assert OracleTokenizer(vf, mf)(gold["corpus"]) == gold["ids"]["49408"]


def test_tokenizer_edge_cases(b200sd):
    tok = b200sd.tokenizer.CLIPTokenizer.synthetic(49408)
    bos, eos = tok.bos_token_id, tok.eos_token_id
    assert tok.encode("") == [bos] + [eos] * 76                                     # empty prompt
    long = tok.encode("panda " * 100)
    assert long[0] == bos and long[76] == eos and eos not in long[1:76]             # truncation keeps EOS last
    assert tok.encode("A  PANDA") == tok.encode("a panda")                          # case + whitespace folding
    a = tok.tokenize("motorbike!")
    assert len(a) >= 2 and all(0 <= i < 49408 for i in a)


# ------------------------------------------------------------------------------------------------ CLIP text encoder
def test_oracle_clip_matches_upstream_golden(b200sd):
    from oracle.clip import CLIPTextModel
    g = torch.load(os.path.join(HERE, "golden", "clip_tiny.pt"))
    W = b200sd.weights
    cfg = W.load_config("tiny")["text_encoders"][0]
    m = CLIPTextModel(cfg).eval()
    m.load_state_dict(W.random_state_dict(W.clip_text_schema(cfg), g["weights_seed"]), strict=True)
    with torch.no_grad():
        out = m(g["ids"])
    assert (out - g["last_hidden_state"]).abs().max().item() < 1e-5


def test_oracle_clip_with_projection_matches_upstream_golden(b200sd):
    """SDXL's second text encoder: hidden_states[-2] and text_embeds of transformers.CLIPTextModelWithProjection (gelu
    MLP, pooled at the EOS token although the sequence is padded with id 0)."""
    from oracle.clip import CLIPTextModel
    g = torch.load(os.path.join(HERE, "golden", "clip_xl2_tiny.pt"))
    W = b200sd.weights
    cfg = W.load_config("tiny_xl")["text_encoders"][1]
    m = CLIPTextModel(cfg).eval()
    m.load_state_dict(W.random_state_dict(W.clip_text_schema(cfg), g["weights_seed"]), strict=True)
    with torch.no_grad():
        hidden, embeds = m.encode_xl(g["ids"], 999)
    assert (hidden - g["penultimate"]).abs().max().item() < 1e-5
    assert (embeds - g["text_embeds"]).abs().max().item() < 1e-5


def test_sdxl_host_pieces(b200sd):
    """Host-side SDXL logic of the product: size-id sinusoid == the oracle's timestep embedding; EOS rows; the second
    tokenizer pads with id 0; parameter count of the SDXL UNet schema."""
    from oracle.unet import timestep_embedding
    P, W, T = b200sd.pipeline, b200sd.weights, b200sd.tokenizer
    v = torch.tensor([1024.0, 1024.0, 0.0, 0.0, 1024.0, 768.0])
    assert torch.equal(P.sinusoid(v, 256), timestep_embedding(v, 256))
    tok = T.CLIPTokenizer.synthetic(1000)
    ids = torch.tensor(tok(["a cat", ""]), dtype=torch.int32)
    pos = [(row == tok.eos_token_id).nonzero()[0].item() for row in ids]
    assert pos[1] == 1 and pos[0] > 1


# ------------------------------------------------------------------------------------------------ schedulers (App. C)
def test_alphas_cumprod_anchors():
    from oracle.schedulers import alphas_cumprod
    ac = alphas_cumprod()
    for t, v in {0: 0.999149978, 1: 0.998296022, 250: 0.673792303, 251: 0.672151148, 500: 0.276332468,
                 501: 0.274998844, 750: 0.056169804, 751: 0.055718984, 921: 0.011337903, 961: 0.007281722,
                 999: 0.004660095}.items():
        assert abs(ac[t].item() - v) < 2e-7, (t, ac[t].item())


def test_timestep_tables_bit_exact(b200sd):
    from oracle.schedulers import DDIMScheduler, EulerDiscreteScheduler, PNDMScheduler
    S = b200sd.schedulers
    d = DDIMScheduler(steps_offset=0)
    d.set_timesteps(4)
    assert d.timesteps.tolist() == [750, 500, 250, 0]
    d = DDIMScheduler(steps_offset=1)
    d.set_timesteps(4)
    assert d.timesteps.tolist() == [751, 501, 251, 1]
    d.set_timesteps(25)
    assert d.timesteps.tolist() == list(range(961, 0, -40))
    p = PNDMScheduler(steps_offset=1)
    p.set_timesteps(25)
    assert p.timesteps.tolist() == [961, 921, 921] + list(range(881, 0, -40)) and len(p.timesteps) == 26
    p.set_timesteps(30)
    assert p.timesteps.tolist()[:5] == [958, 925, 925, 892, 859] and p.timesteps.tolist()[-3:] == [67, 34, 1]
    # product tables == oracle tables for every scheduler and several step counts
    for kind, cls in (("ddim", DDIMScheduler), ("pndm", PNDMScheduler), ("euler", EulerDiscreteScheduler)):
        for n in (1, 4, 25, 30, 50):
            o = cls(steps_offset=1)
            o.set_timesteps(n)
            assert S.make_schedule(kind, n, steps_offset=1).timesteps == [int(t) for t in o.timesteps], (kind, n)


def _emulate_device_step(row, x, e, hist, cur, L):
    """fp32 emulation of cfg_sched_kernel's scalar maths for one latent element (csrc/elementwise.cu)."""
    f = lambda v: torch.tensor(v, dtype=torch.float32)  # noqa: E731
    kind = int(row[L.ST_KIND])
    if kind == 0:
        x0 = (x - row[L.ST_C0] * e) / row[L.ST_C1]
        return row[L.ST_C2] * x0 + row[L.ST_C3] * e, hist, cur
    if kind == 1:
        s = row[L.ST_C1]
        x0 = x - s * e
        return x + ((x - x0) / s) * row[L.ST_C0], hist, cur
    if row[L.ST_PUSH] != 0:
        hist = [e] + hist[:3]
    mode = int(row[L.ST_PNDM_MODE])
    if mode == 1:
        cur, m = x, e
    elif mode == 2:
        m, x = (e + hist[0]) / 2, cur
    else:
        m = f(0.0)
        for j in range(4):
            if row[L.ST_W0 + j] != 0:
                m = m + row[L.ST_W0 + j] * hist[j]
    return row[L.ST_C0] * x - row[L.ST_C1] * m, hist, cur


@pytest.mark.parametrize("kind", ["ddim", "pndm", "euler"])
def test_step_table_reproduces_oracle_scheduler(b200sd, kind):
    """The per-step coefficient table uploaded to the GPU, applied with the kernel's formula, follows the oracle
    scheduler's latents (fp32, 1e-5)."""
    from oracle.schedulers import make_scheduler
    S, L = b200sd.schedulers, b200sd.lib
    cfg = {"type": kind, "steps_offset": 1, "beta_start": 0.00085, "beta_end": 0.012, "num_train_timesteps": 1000}
    o = make_scheduler(cfg)
    o.set_timesteps(25)
    sched = S.make_schedule(kind, 25, steps_offset=1)
    g = torch.Generator().manual_seed(0)
    x_o = torch.randn(64, generator=g) * o.init_noise_sigma
    assert abs(sched.init_noise_sigma - o.init_noise_sigma) < 1e-6
    x_d, hist, cur = x_o.clone(), [torch.zeros(64)] * 4, torch.zeros(64)
    for i, t in enumerate(o.timesteps):
        e = torch.randn(64, generator=g) * 0.5
        if kind == "euler":
            scaled = o.scale_model_input(x_o, t)
            assert torch.allclose(scaled, x_d * sched.table[i, L.ST_IN_SCALE], atol=1e-6)
        assert float(sched.table[i, L.ST_TIMESTEP]) == float(t)
        x_o = o.step(e, t, x_o)
        x_d, hist, cur = _emulate_device_step(sched.table[i], x_d, e, hist, cur, L)
        assert torch.allclose(x_o, x_d, rtol=2e-5, atol=2e-5), (kind, i)


def test_one_ddim_step_anchor():
    from oracle.schedulers import DDIMScheduler
    d = DDIMScheduler(steps_offset=1)
    d.set_timesteps(25)
    out = d.step(torch.tensor([0.5]), 961, torch.tensor([1.0]))
    assert abs(out.item() - 1.12333941) < 1e-6


def test_timestep_embedding_anchor():
    from oracle.unet import timestep_embedding
    e = timestep_embedding(torch.tensor([961.0]), 320)[0]
    assert torch.allclose(e[0:3], torch.tensor([0.94689709, -0.77896410, -0.39694813]), atol=2e-5)
    assert torch.allclose(e[160:163], torch.tensor([-0.32153678, 0.62706852, 0.91784102]), atol=2e-5)
    assert abs(e[319].item() - 0.10161857) < 2e-5


# ------------------------------------------------------------------------------------------------ structure
@pytest.mark.parametrize("name,unet_m,vae_m,clip_m", [("sd15", 859.5, 49.5, 123.1), ("sdxl", 2567.5, 49.5, 123.1)])
def test_parameter_counts_and_key_schema(b200sd, name, unet_m, vae_m, clip_m):
    from oracle.clip import CLIPTextModel
    from oracle.unet import UNet2DConditionModel
    from oracle.vae import AutoencoderKLDecoder
    W = b200sd.weights
    cfg = W.load_config(name)
    with torch.device("meta"):
        mods = [(UNet2DConditionModel(cfg["unet"]), W.unet_schema(cfg["unet"]), unet_m),
                (AutoencoderKLDecoder(cfg["vae"]), W.vae_decoder_schema(cfg["vae"]), vae_m),
                (CLIPTextModel(cfg["text_encoders"][0]), W.clip_text_schema(cfg["text_encoders"][0]), clip_m)]
    for mod, schema, millions in mods:
        sd = {k: tuple(v.shape) for k, v in mod.state_dict().items()}
        assert sd == {k: tuple(s[1]) for k, s in schema.items()}
        assert abs(W.num_params(schema) / 1e6 - millions) < 0.06


def test_oracle_pipeline_tiny_runs_and_is_deterministic(b200sd):
    from oracle.pipeline import OraclePipeline
    W, T = b200sd.weights, b200sd.tokenizer
    cfg = W.load_config("tiny")
    sds = [W.random_state_dict(W.unet_schema(cfg["unet"]), 0), W.random_state_dict(W.vae_decoder_schema(cfg["vae"]), 1),
           W.random_state_dict(W.clip_text_schema(cfg["text_encoders"][0]), 2)]
    pipe = OraclePipeline(cfg, *sds, T.CLIPTokenizer.synthetic(1000))
    kw = dict(num_inference_steps=2, guidance_scale=7.5, width=64, height=64, scheduler="ddim")
    a = pipe(["a panda", "a cabin"], generator=[torch.Generator().manual_seed(1234 + i) for i in range(2)], **kw)
    b = pipe("a cabin", generator=torch.Generator().manual_seed(1235), **kw)
    assert a["images"].shape == (2, 64, 64, 3) and a["images"].dtype.name == "uint8"
    assert a["timesteps"] == [500, 0]
    # batch invariance: image 1 of the batch == the same prompt/seed generated alone (sharding invariance on CPU)
    assert (a["latents"][1] - b["latents"][0]).abs().max().item() < 1e-4
    assert math.isfinite(float(a["latents"].abs().max()))


# ------------------------------------------------------------------------------------------------ attention anchors
def test_oracle_attention_matches_torch_sdpa():
    """oracle.unet.Attention (self- and cross-attention of BasicTransformerBlock) against an independent statement of
    the same layer built from torch's own F.scaled_dot_product_attention + F.linear (torch is the reference's runtime,
    configmap.yaml:10): same weights, fp32, 1e-5."""
    import torch.nn.functional as F
    from oracle.unet import Attention
    torch.manual_seed(0)
    for dim, heads, ctx_dim, n, m in ((320, 8, None, 96, 96), (640, 8, 768, 50, 77), (64, 2, 32, 33, 7)):
        att = Attention(dim, heads, ctx_dim).eval()
        x = torch.randn(2, n, dim)
        ctx = None if ctx_dim is None else torch.randn(2, m, ctx_dim)
        src = x if ctx is None else ctx
        with torch.no_grad():
            got = att(x, ctx)
            d = dim // heads
            q = F.linear(x, att.to_q.weight).view(2, n, heads, d).transpose(1, 2)
            k = F.linear(src, att.to_k.weight).view(2, src.shape[1], heads, d).transpose(1, 2)
            v = F.linear(src, att.to_v.weight).view(2, src.shape[1], heads, d).transpose(1, 2)
            o = F.scaled_dot_product_attention(q, k, v)  # default scale = d ** -0.5, no mask
            ref = F.linear(o.transpose(1, 2).reshape(2, n, dim), att.to_out[0].weight, att.to_out[0].bias)
        assert (got - ref).abs().max().item() < 1e-5 * max(1.0, ref.abs().max().item())


def test_oracle_vae_attention_matches_torch_sdpa():
    """oracle.vae.VaeAttention (single head, d = C, GroupNorm in front, residual behind) against
    F.group_norm + F.scaled_dot_product_attention + F.linear."""
    import torch.nn.functional as F
    from oracle.vae import VaeAttention
    torch.manual_seed(1)
    for c, groups, h, w in ((64, 8, 6, 5), (512, 32, 8, 8)):
        att = VaeAttention(c, groups).eval()
        x = torch.randn(2, c, h, w)
        with torch.no_grad():
            got = att(x)
            t = F.group_norm(x, groups, att.group_norm.weight, att.group_norm.bias, 1e-6).flatten(2).transpose(1, 2)
            q = F.linear(t, att.to_q.weight, att.to_q.bias)[:, None]
            k = F.linear(t, att.to_k.weight, att.to_k.bias)[:, None]
            v = F.linear(t, att.to_v.weight, att.to_v.bias)[:, None]
            o = F.scaled_dot_product_attention(q, k, v)[:, 0]
            o = F.linear(o, att.to_out[0].weight, att.to_out[0].bias)
            ref = o.transpose(1, 2).reshape(2, c, h, w) + x
        assert (got - ref).abs().max().item() < 1e-5 * max(1.0, ref.abs().max().item())
