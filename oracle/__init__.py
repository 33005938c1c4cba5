"""CPU oracle — TEST INFRASTRUCTURE ONLY.

A plain PyTorch fp32 restatement of the reference's CPU branch for the hot path
(`pipe(prompt, num_inference_steps, guidance_scale, width, height, generator)`,
reference cluster-config/apps/sd15-api/configmap.yaml:105-112 with TORCH_DTYPE=float32 and a CPU
torch.Generator, :31, :90-92, :103).

The arithmetic of that call lives in un-vendored third-party packages pinned at configmap.yaml:10-15
(diffusers==0.30.2, transformers==4.44.2, torch==2.3.1); none of their sources are under /root/reference and
diffusers is not installable offline.  Each module here restates the published algorithm of the upstream class it
names and cites the reference call site it serves.

PARITY PINNING
  - tokenizer  : pinned bit-exactly against transformers.CLIPTokenizer (installed 5.5.0) — tests/golden/tokenizer.json
  - CLIP text  : pinned against transformers.CLIPTextModel (installed 5.5.0) — tests/golden/clip_tiny.pt
  - schedulers / timestep embedding: pinned against closed-form anchors — tests/test_oracle.py
  - UNet / VAE / end-to-end pipeline: **parity unpinned** — the reference holds no golden vectors, tests or fixtures
    for this path and diffusers cannot be imported here; structure is checked through the
    public parameter counts (859.5 M / 49.5 M) and the diffusers state-dict key schema only.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this package.
The product (k8s-nvidia-gpus_b200/) never does.
"""
