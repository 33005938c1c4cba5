"""GPU parity for the SDXL-architecture path (BASELINE.json configs[3] scaled down as
configs/tiny_xl.json): two text encoders (penultimate hidden states + pooled/projected output), text_time
conditioning, linear proj_in/out, transformer depth > 1, head dims 16, Euler-discrete — the product pipeline (CUDA
kernels through the C ABI) against the fp32 CPU oracle.  Tolerances as for configs[0]: tokens / timesteps bit-exact,
per-step latents <= 1e-2 relative, PSNR >= 30 dB."""
import copy
import math

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

PROMPTS = ["a photo of a panda riding a motorbike", "a cabin in the forest at night, detailed"]


def _rel(a, b):
    return ((a.float() - b.float()).norm() / (b.float().norm() + 1e-12)).item()


def _psnr(a, b):
    mse = np.mean((a.astype(np.float64) - b.astype(np.float64)) ** 2)
    return 99.0 if mse == 0 else 10 * math.log10(255.0 ** 2 / mse)


@pytest.fixture(scope="module")
def tiny_xl(b200sd, cuda_lib):
    W, T, P = b200sd.weights, b200sd.tokenizer, b200sd.pipeline
    from oracle.pipeline import OraclePipeline
    cfg = W.load_config("tiny_xl")
    usd = W.random_state_dict(W.unet_schema(cfg["unet"]), 0)
    vsd = W.random_state_dict(W.vae_decoder_schema(cfg["vae"]), 1)
    tsd = [W.random_state_dict(W.clip_text_schema(c), 2 + i) for i, c in enumerate(cfg["text_encoders"])]
    tok = T.CLIPTokenizer.synthetic(cfg["text_encoders"][0]["vocab_size"])
    tok2 = copy.copy(tok)
    tok2.pad_token_id = 0
    oracle = OraclePipeline(cfg, usd, vsd, tsd, [tok, tok2], eos_token_id=tok.eos_token_id)
    pipe = P.StableDiffusionXLPipeline(cfg, usd, vsd, tsd, tokenizer=tok)
    return cfg, oracle, pipe


def _gens():
    return [torch.Generator().manual_seed(1234 + i) for i in range(len(PROMPTS))]


def test_text_encoders_and_added_conditioning(tiny_xl):
    cfg, oracle, pipe = tiny_xl
    n = len(PROMPTS)
    neg = ["blurry"] * n
    ids = pipe.tokenize(PROMPTS, neg, True)
    job = pipe.job(n, 32, 32, 5.0, True, zero_negative=False)
    job.ids_dev.copy_(ids)
    job.eos_dev.copy_(pipe.eos_rows(ids, job.n_text))
    job.run_text()
    torch.cuda.synchronize()
    ref_ids = [t(neg + PROMPTS) for t in oracle.tokenizers]
    assert ids.view(2, -1, 77).tolist() == ref_ids  # bit-exact tokenisation, second tokenizer pads with id 0
    assert ids.view(2, -1, 77)[1, 0, -1].item() == 0
    with torch.no_grad():
        ref_neg, ref_neg_pooled = oracle._encode_xl(neg)
        ref_pos, ref_pos_pooled = oracle._encode_xl(PROMPTS)
        ref = torch.cat([ref_neg, ref_pos])
        pooled = torch.cat([ref_neg_pooled, ref_pos_pooled])
        got = torch.cat([c.float().cpu() for c in job.ctx], dim=-1)
        r_ctx, r_pool = _rel(got, ref), _rel(job.pooled.cpu(), pooled)
        # aug_emb = add_embedding(concat(text_embeds, sinusoid(time ids)))
        U = oracle.unet
        from oracle.unet import timestep_embedding
        tid = torch.tensor([[64, 64, 0, 0, 64, 64]], dtype=torch.float32).repeat(2 * n, 1)
        te = timestep_embedding(tid.flatten(), cfg["unet"]["addition_time_embed_dim"]).reshape(2 * n, -1)
        ref_aug = U.add_embedding(torch.cat([pooled, te], dim=-1))
        r_aug = _rel(job.aug_emb.cpu(), ref_aug)
    print("xl text rel err: hidden %.2e pooled %.2e aug_emb %.2e" % (r_ctx, r_pool, r_aug))
    assert r_ctx < 1e-2 and r_pool < 1e-2 and r_aug < 1e-2


@pytest.mark.parametrize("negative", [None, "blurry, low quality"])
def test_sdxl_denoise_and_image_parity(tiny_xl, negative):
    """negative=None exercises force_zeros_for_empty_prompt (zero unconditional embeddings)."""
    cfg, oracle, pipe = tiny_xl
    kw = dict(num_inference_steps=4, guidance_scale=5.0, width=64, height=64, negative_prompt=negative)
    ref = oracle(PROMPTS, generator=_gens(), **kw)
    got_steps, got_t = [], []
    out = pipe(PROMPTS, generator=_gens(), output_type="u8",
               callback_on_step_end=lambda p, i, t, k: (got_steps.append(k["latents"]), got_t.append(int(t))), **kw)
    assert got_t == ref["timesteps"]  # bit-exact Euler timestep table (leading spacing, steps_offset 1)
    rels = [_rel(a, b) for a, b in zip(got_steps, ref["per_step"])]
    print("xl per-step latent rel err", ["%.2e" % r for r in rels])
    assert max(rels) <= 1e-2, rels
    psnr = [_psnr(out.images[i], ref["images"][i]) for i in range(len(PROMPTS))]
    print("xl PSNR", psnr)
    assert min(psnr) >= 30.0, psnr


def test_sdxl_bf16_vae_beside_fp16_unet(tiny_xl, b200sd):
    """Per-plan 16-bit type: the VAE decoder bound to the bfloat16 image of the library (what configs/sdxl.json and a
    force_upcast checkpoint select - fp32's exponent range), the UNet and text encoders on the fp16 image, in one
    pipeline.  Weights rounded to bf16 for both sides; image PSNR >= 30 dB against the fp32 oracle, latents as before."""
    W, T, P = b200sd.weights, b200sd.tokenizer, b200sd.pipeline
    from oracle.pipeline import OraclePipeline
    cfg, oracle0, _ = tiny_xl
    usd = W.random_state_dict(W.unet_schema(cfg["unet"]), 0)
    vsd = W.random_state_dict(W.vae_decoder_schema(cfg["vae"]), 1, round_dtype=torch.bfloat16)
    tsd = [W.random_state_dict(W.clip_text_schema(c), 2 + i) for i, c in enumerate(cfg["text_encoders"])]
    oracle = OraclePipeline(cfg, usd, vsd, tsd, oracle0.tokenizers, eos_token_id=oracle0.eos_token_id)
    pipe = P.StableDiffusionXLPipeline(cfg, usd, vsd, tsd, tokenizer=oracle0.tokenizers[0], vae_dtype="bf16")
    assert pipe.vae_engine.variant == "bf16"
    kw = dict(num_inference_steps=3, guidance_scale=5.0, width=64, height=64)
    ref = oracle(PROMPTS, generator=_gens(), **kw)
    out = pipe(PROMPTS, generator=_gens(), output_type="u8", **kw)
    psnr = [_psnr(out.images[i], ref["images"][i]) for i in range(len(PROMPTS))]
    print("xl bf16-VAE PSNR", psnr)
    assert min(psnr) >= 30.0, psnr
    # large-magnitude latents: fp16 would overflow inside the decoder (activations ~1e5), bf16 must stay finite
    job = pipe.job(len(PROMPTS), 32, 32, 5.0, True, zero_negative=True)
    job.latents.normal_(0, 3000.0)
    job.vae_plan.run()
    torch.cuda.synchronize()
    assert torch.isfinite(pipe.vae_engine.last_rgb).all()


def test_sdxl_full_size_one_step_vs_oracle(b200sd, cuda_lib):
    """BASELINE.json configs[3] at FULL size: SDXL-base architecture (2.57 B-parameter UNet, CLIP-L + OpenCLIP-bigG,
    VAE at 1024x1024 with its 16 384-token d=512 attention, bf16 decoder), one prompt, one Euler step + decode, against
    the fp32 CPU oracle (about 8x the SD1.5 cost per evaluation - hence one step)."""
    W, T, P = b200sd.weights, b200sd.tokenizer, b200sd.pipeline
    from oracle.pipeline import OraclePipeline
    cfg = W.load_config("sdxl")
    usd = W.random_state_dict(W.unet_schema(cfg["unet"]), 0)
    vsd = W.random_state_dict(W.vae_decoder_schema(cfg["vae"]), 1, round_dtype=torch.bfloat16)
    tsd = [W.random_state_dict(W.clip_text_schema(c), 2 + i) for i, c in enumerate(cfg["text_encoders"])]
    tok = T.CLIPTokenizer.synthetic(cfg["text_encoders"][0]["vocab_size"])
    tok2 = copy.copy(tok)
    tok2.pad_token_id = 0
    pipe = P.StableDiffusionXLPipeline(cfg, usd, vsd, tsd, tokenizer=tok)
    assert pipe.vae_engine.variant == "bf16"
    kw = dict(num_inference_steps=1, guidance_scale=5.0, width=1024, height=1024)
    steps = []
    out = pipe(PROMPTS[0], generator=torch.Generator().manual_seed(1234), output_type="u8",
               callback_on_step_end=lambda p, i, t, k: steps.append(k["latents"]), **kw)
    img = out.images[0].copy()
    del pipe
    torch.cuda.empty_cache()
    oracle = OraclePipeline(cfg, usd, vsd, tsd, [tok, tok2], eos_token_id=tok.eos_token_id)
    ref = oracle(PROMPTS[0], generator=torch.Generator().manual_seed(1234), **kw)
    rel = _rel(steps[0], ref["per_step"][0])
    psnr = _psnr(img, ref["images"][0])
    print("sdxl full size: latent rel err %.2e, PSNR %.1f dB" % (rel, psnr))
    assert rel <= 1e-2 and psnr >= 30.0
