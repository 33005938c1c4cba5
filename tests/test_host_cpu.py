"""CPU tests (no GPU) of the host side: the C-ABI library loads and exports every symbol include/b200sd.h declares (no
compute calls), ctypes mirrors match the C structs, weight packing, the CLI / HTTP boundaries (B1/B2) with a fake
pipeline, and the multi-GPU sharding logic on world_size-2 gloo."""
import ctypes
import io
import json
import os
import re
import socket
import subprocess
import sys
import threading
from http.server import BaseHTTPRequestHandler, HTTPServer
from pathlib import Path

import numpy as np
import pytest
import torch

ROOT = Path(__file__).resolve().parent.parent


# ------------------------------------------------------------------------------------------------ C ABI
def test_library_exports_every_declared_symbol(b200sd):
    lib = b200sd.lib.load()          # loads without a GPU; raises if the .so is missing (no fallback)
    header = (ROOT / "include" / "b200sd.h").read_text()
    declared = set(re.findall(r"\b(b200sd_[a-z0-9_]+)\s*\(", header)) - {"b200sd_stream"}
    assert len(declared) >= 20
    nm = subprocess.run(["nm", "-D", "--defined-only", str(b200sd.lib.lib_path())], capture_output=True, text=True).stdout
    exported = set(re.findall(r"\bT (b200sd_[a-z0-9_]+)", nm))
    assert declared <= exported, declared - exported
    assert declared == set(b200sd.lib.ENTRY_POINTS), declared ^ set(b200sd.lib.ENTRY_POINTS)
    for name in declared:
        assert getattr(lib, name) is not None
    assert lib.b200sd_version() >= 100
    assert lib.b200sd_act_dtype() in (0, 1)


def test_ctypes_mirrors_match_c_structs(b200sd):
    assert b200sd.lib.struct_sizes() == [ctypes.sizeof(s) for s in b200sd.lib.STRUCTS]


def test_no_gpu_means_loud_failure(b200sd):
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(b200sd.lib.B200SDError):
        b200sd.lib.device_info()                      # CUDA error surfaces, nothing falls back
    W = b200sd.weights
    cfg = W.load_config("tiny")
    with pytest.raises(RuntimeError, match="no CPU path"):
        b200sd.StableDiffusionPipeline(cfg, {}, {}, {})


def test_product_never_imports_the_oracle():
    for p in (ROOT / "k8s-nvidia-gpus_b200").glob("*.py"):
        src = p.read_text()
        assert "import oracle" not in src and "from oracle" not in src, p
    for p in (ROOT / "scripts").glob("*.py"):
        assert "oracle" not in p.read_text(), p


# ------------------------------------------------------------------------------------------------ weight packing
def test_weight_packing_layouts(b200sd):
    ops = b200sd.ops
    w = torch.arange(2 * 6 * 3 * 3, dtype=torch.float32).reshape(2, 6, 3, 3) / 64
    p = ops.pack_conv_weight(w, (4, 2)).float()
    assert p.shape == (2, 9, 128)
    assert torch.equal(p[:, :, :4], w.permute(0, 2, 3, 1).reshape(2, 9, 6)[:, :, :4])
    assert torch.equal(p[:, :, 64:66], w.permute(0, 2, 3, 1).reshape(2, 9, 6)[:, :, 4:])
    assert p[:, :, 4:64].abs().sum() == 0 and p[:, :, 66:].abs().sum() == 0
    perm = ops.geglu_perm(128)
    assert perm[:32].tolist() == list(range(32)) and perm[32:64].tolist() == list(range(64, 96))
    assert sorted(perm.tolist()) == list(range(128))


def test_bn_independent_flop_model(b200sd):
    """bench.py's algorithmic-FLOP model over a built plan equals SURVEY's 803.27 GF/sample (needs no GPU: counts only)."""
    sys.path.insert(0, str(ROOT))
    import bench
    p = b200sd.lib.GemmParams()
    p.nimg, p.h, p.w_, p.c0, p.c1, p.n, p.ksize = 16, 64, 64, 320, 0, 320, 3

    class C_:
        name, params = "b200sd_gemm", p
    fl, kind = bench.flops_of_call(C_)
    assert kind == "gemm" and fl == 2.0 * 16 * 4096 * 320 * 9 * 320


def test_fastdiv_constants_are_exact():
    """The multiply-high + shift division the GEMM's tile decode uses (csrc/gemm.cu make_fastdiv / fdiv), replayed in
    Python: exact for every dividend < 2^31 - checked on edge dividends for a spread of divisors."""
    def make(d):
        if d <= 1:
            return 0, -1
        s = 0
        while (1 << s) < d:
            s += 1
        return ((1 << (31 + s)) + d - 1) // d, s - 1

    def fdiv(n, mul, shift):
        return n if shift < 0 else ((n * mul) >> 32) >> shift

    rng = np.random.default_rng(0)
    divisors = list(range(1, 70)) + [100, 127, 128, 129, 148, 255, 256, 257, 511, 512, 1000, 4096, 65535, 65536, 99991,
                                     2 ** 20 + 7, 2 ** 30 - 1, 2 ** 30, 2 ** 31 - 1]
    for d in divisors:
        mul, shift = make(d)
        assert mul < 2 ** 32
        ns = [0, 1, d - 1, d, d + 1, 2 * d - 1, 2 * d, 2 ** 31 - 1, 2 ** 31 - 2, (2 ** 31 - 1) // d * d,
              max(0, (2 ** 31 - 1) // d * d - 1)] + [int(x) for x in rng.integers(0, 2 ** 31, size=200)]
        for n in ns:
            if 0 <= n < 2 ** 31:
                assert fdiv(n, mul, shift) == n // d, (n, d)


# ------------------------------------------------------------------------------------------------ fake pipeline
class FakePipe:
    """Stands in for the GPU pipeline in boundary tests: deterministic image from (prompt, seed)."""
    device = torch.device("cpu")

    def __init__(self):
        self.calls = []

    def __call__(self, prompt, num_inference_steps=50, guidance_scale=7.5, width=512, height=512, generator=None,
                 output_type="pil", **kw):
        prompts = [prompt] if isinstance(prompt, str) else list(prompt)
        gens = generator if isinstance(generator, (list, tuple)) else [generator] * len(prompts)
        self.calls.append(dict(prompts=prompts, steps=num_inference_steps, gs=guidance_scale, w=width, h=height))
        imgs = []
        for p_, g in zip(prompts, gens):
            seed = int(torch.randint(0, 256, (1,), generator=g).item()) if g is not None else 7
            imgs.append(np.full((height, width, 3), (len(p_) * 3 + seed) % 256, dtype=np.uint8))
        arr = np.stack(imgs)
        if output_type == "u8":
            images = arr
        else:
            from PIL import Image
            images = [Image.fromarray(a) for a in arr]
        return type("Out", (), {"images": images})()


# ------------------------------------------------------------------------------------------------ B2: HTTP server
def test_http_contract_matches_reference(b200sd):
    from fastapi.testclient import TestClient
    import importlib
    server = importlib.import_module("k8s-nvidia-gpus_b200.server")
    fake = FakePipe()
    c = TestClient(server.create_app(pipe=fake))
    assert c.get("/healthz").json() == {"ok": True}
    assert c.get("/last").status_code == 404
    assert "No image generated yet" in c.get("/").text
    r = c.post("/generate", json={"prompt": "   "})
    assert r.status_code == 400 and r.json()["detail"] == "prompt is required"
    assert c.post("/generate", json={}).status_code == 422                      # prompt is a required field
    r = c.post("/generate", json={"prompt": "a panda", "steps": 0, "guidance_scale": 0.0, "seed": 3, "width": 64,
                                  "height": 64})
    assert r.status_code == 200 and r.headers["content-type"] == "image/png"
    assert re.fullmatch(r"\d+\.\d\ds", r.headers["X-Gen-Time"])
    assert r.content[:8] == b"\x89PNG\r\n\x1a\n"
    assert fake.calls[-1]["steps"] == 30 and fake.calls[-1]["gs"] == 7.5      # falsy -> defaults (configmap.yaml:107-110)
    assert fake.calls[-1]["w"] == 64 and fake.calls[-1]["h"] == 64
    r2 = c.post("/generate", json={"prompt": "a panda", "steps": 25})
    assert fake.calls[-1]["w"] == 512 and fake.calls[-1]["steps"] == 25
    assert c.get("/last").content == r2.content
    assert "data:image/png;base64," in c.get("/").text


def test_dynamic_batcher_coalesces_routes_and_pads(b200sd):
    """concurrent same-shape requests become one batched pipe() call (padded to a bucket size),
    each waiter gets ITS image (same as a lone request with that seed), other shapes are not mixed in, and a pipeline
    failure reaches every waiter of the batch."""
    import importlib
    import threading
    B = importlib.import_module("k8s-nvidia-gpus_b200.batcher")
    fake = FakePipe()
    gate = threading.Event()

    def slow_pipe(*a, **kw):
        gate.wait(5)  # holds the worker so that the following requests pile up in the queue
        return fake(*a, **kw)

    bt = B.DynamicBatcher(lambda: slow_pipe, max_batch=4, max_wait_ms=1, buckets=(1, 2, 4, 8), output_type="u8")
    first = bt.submit("warm", 5, 7.5, 8, 8, seed=1)          # occupies the worker
    import time
    time.sleep(0.05)
    futs = [bt.submit(f"prompt {i}", 5, 7.5, 8, 8, seed=10 + i) for i in range(3)]
    other = bt.submit("other shape", 9, 7.5, 8, 8, seed=3)
    gate.set()
    imgs = [f.result(10) for f in futs]
    first.result(10), other.result(10)
    solo = FakePipe()
    for i, im in enumerate(imgs):  # same image as an unbatched call with that prompt and seed
        ref = solo(f"prompt {i}", num_inference_steps=5, width=8, height=8,
                   generator=torch.Generator().manual_seed(10 + i), output_type="u8").images[0]
        assert np.array_equal(im, ref)
    sizes = [len(c["prompts"]) for c in fake.calls]
    assert sizes == [1, 4, 1]                               # 3 requests padded to the bucket of 4; the odd shape alone
    assert fake.calls[1]["prompts"][:3] == [f"prompt {i}" for i in range(3)] and fake.calls[2]["steps"] == 9
    assert bt.stats["requests"] == 5 and bt.stats["batches"] == 3 and bt.stats["padded"] == 1

    def broken(*a, **kw):
        raise RuntimeError("device fell over")
    bt2 = B.DynamicBatcher(lambda: broken, max_batch=2, max_wait_ms=20, output_type="u8")
    fa, fb = bt2.submit("a", 5, 7.5, 8, 8, 1), bt2.submit("b", 5, 7.5, 8, 8, 2)
    for f in (fa, fb):
        with pytest.raises(RuntimeError, match="fell over"):
            f.result(10)
    bt.shutdown(), bt2.shutdown()


def test_http_server_batches_concurrent_requests(b200sd):
    from fastapi.testclient import TestClient
    from concurrent.futures import ThreadPoolExecutor
    import importlib
    server = importlib.import_module("k8s-nvidia-gpus_b200.server")
    fake = FakePipe()
    c = TestClient(server.create_app(pipe=fake, max_batch=8, max_wait_ms=200))
    with ThreadPoolExecutor(6) as ex:
        rs = list(ex.map(lambda i: c.post("/generate", json={"prompt": f"p{i}", "seed": i, "width": 16, "height": 16,
                                                             "steps": 4}), range(6)))
    assert all(r.status_code == 200 and r.content[:4] == b"\x89PNG" for r in rs)
    assert len(fake.calls) < 6 and sum(len(c_["prompts"]) for c_ in fake.calls) >= 6   # coalesced


# ------------------------------------------------------------------------------------------------ real checkpoints
@pytest.mark.parametrize("name,sched", [("tiny", "PNDMScheduler"), ("tiny_xl", "EulerDiscreteScheduler")])
def test_diffusers_checkpoint_directory_loads(b200sd, tmp_path, name, sched):
    """a checkpoint directory in diffusers layout (written with the upstream safetensors
    library, upstream config key names, legacy VAE attention names, fp16 storage, encoder-side and position_ids
    extras) comes back as this repo's config + exactly the stored tensors."""
    sys.path.insert(0, str(ROOT / "tests"))
    from ckpt_util import write_diffusers_dir
    W, T, C = b200sd.weights, b200sd.tokenizer, b200sd.checkpoint
    cfg = W.load_config(name)
    usd = W.random_state_dict(W.unet_schema(cfg["unet"]), 0)
    vsd = W.random_state_dict(W.vae_decoder_schema(cfg["vae"]), 1)
    tsd = [W.random_state_dict(W.clip_text_schema(c), 2 + i) for i, c in enumerate(cfg["text_encoders"])]
    root = write_diffusers_dir(str(tmp_path / "ckpt"), cfg, usd, vsd, tsd, T, scheduler_class=sched)
    assert C.is_checkpoint_dir(root)
    cfg2, usd2, vsd2, tsd2, tok_dir = C.load_diffusers_dir(root)
    for key in ("unet", "vae", "text_encoders"):
        a, b = cfg[key], cfg2[key]
        if key in ("unet", "vae"):
            a = {k: v for k, v in a.items()}
            b = {k: v for k, v in b.items() if k in a}
        assert a == b, key
    assert cfg2["vae"]["dtype"] == "fp16"      # no force_upcast in this checkpoint (SDXL's VAE config sets it -> bf16)
    assert cfg2["scheduler"]["type"] == {"PNDMScheduler": "pndm", "EulerDiscreteScheduler": "euler"}[sched]
    assert cfg2["scheduler"]["steps_offset"] == 1
    for ref, got in [(usd, usd2), (vsd, vsd2)] + list(zip(tsd, tsd2)):
        assert list(ref) == list(got)
        for k in ref:  # random_state_dict values are fp16-representable: the fp16 file holds them exactly
            assert got[k].dtype == torch.float16 and torch.equal(got[k].float(), ref[k]), k
    tok = T.CLIPTokenizer.from_files(tok_dir)
    assert tok(["a panda"]) == T.CLIPTokenizer.synthetic(cfg["text_encoders"][0]["vocab_size"])(["a panda"])
    # a file that does not match the architecture is refused with the offending keys named
    bad = dict(usd)
    bad.pop("conv_in.weight")
    from safetensors.torch import save_file
    save_file({k: v.half() for k, v in bad.items()}, os.path.join(root, "unet", "diffusion_pytorch_model.safetensors"))
    with pytest.raises(ValueError, match="conv_in.weight"):
        C.load_diffusers_dir(root)


def test_safetensors_reader_matches_upstream(b200sd, tmp_path):
    from safetensors.torch import save_file
    t = {"a": torch.randn(3, 5), "b": torch.randn(7).to(torch.bfloat16), "c": torch.randn(2, 2, 2).half(),
         "d": torch.arange(6).reshape(2, 3), "e": torch.tensor(3.5)}
    path = str(tmp_path / "x.safetensors")
    save_file(t, path, metadata={"format": "pt"})
    got = b200sd.checkpoint.read_safetensors(path)
    assert set(got) == set(t)
    for k in t:
        assert got[k].dtype == t[k].dtype and got[k].shape == t[k].shape and torch.equal(got[k], t[k])
    assert set(b200sd.checkpoint.read_safetensors(path, keep=lambda k: k in "ab")) == {"a", "b"}


# ------------------------------------------------------------------------------------------------ B1: CLI
class _Stub(BaseHTTPRequestHandler):
    bodies = []

    def do_POST(self):
        n = int(self.headers["Content-Length"])
        _Stub.bodies.append((self.headers["Content-Type"], json.loads(self.rfile.read(n))))
        self.send_response(200)
        self.send_header("Content-Type", "image/png")
        self.send_header("X-Gen-Time", "1.23s")
        self.end_headers()
        self.wfile.write(b"\x89PNG fake")

    def log_message(self, *a):
        pass


def _load_cli():
    import importlib.util
    spec = importlib.util.spec_from_file_location("batch_generate_cli", ROOT / "scripts" / "batch_generate.py")
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def test_cli_http_mode_is_wire_compatible(tmp_path, capsys):
    cli = _load_cli()
    _Stub.bodies.clear()
    srv = HTTPServer(("127.0.0.1", 0), _Stub)
    threading.Thread(target=srv.serve_forever, daemon=True).start()
    try:
        rc = cli.main(["a panda on a motorbike", "3", "panda", str(tmp_path / "out"), "--steps", "25", "--url",
                       f"http://127.0.0.1:{srv.server_port}/generate"])
    finally:
        srv.shutdown()
    assert rc == 0
    assert [b for _, b in _Stub.bodies] == [{"prompt": "a panda on a motorbike", "steps": 25}] * 3
    assert all(ct == "application/json" for ct, _ in _Stub.bodies)
    assert sorted(p.name for p in (tmp_path / "out").iterdir()) == ["panda_01.png", "panda_02.png", "panda_03.png"]
    out = capsys.readouterr().out
    assert "[*] Generating panda_01.png -> " in out and "✔ done in 1.23ss" in out and "All done. Images saved under" in out


def test_cli_defaults_and_failure_keeps_going(tmp_path, capsys):
    cli = _load_cli()
    a = cli.build_parser().parse_args(["p", "2", "x"])
    assert (a.steps, a.url, a.delay, a.out_dir) == (40, "http://192.168.0.176:30800/generate", 0, "outputs")
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()                                                     # nothing listens: every request fails
    rc = cli.main(["p", "2", "x", str(tmp_path), "--url", f"http://127.0.0.1:{port}/generate"])
    out = capsys.readouterr().out
    assert rc == 0 and out.count("✖ Request failed") == 2 and "Generation loop finished" in out


def test_cli_local_mode_batches_and_seeds(tmp_path, capsys):
    cli = _load_cli()
    fake = FakePipe()
    cli.generate_local(["a", "bb"], 4, tmp_path, "img", 5, batch_size=2, seed=100, guidance_scale=7.5, width=64,
                       height=64, scheduler="ddim", model="tiny", pipe=fake)
    assert [c["prompts"] for c in fake.calls] == [["a", "bb"], ["a", "bb"], ["a"]]
    assert sorted(p.name for p in tmp_path.iterdir()) == [f"img_{i:02d}.png" for i in range(1, 6)]


# ------------------------------------------------------------------------------------------------ multi-GPU logic
def test_shard_ranges(b200sd):
    import importlib
    D = importlib.import_module("k8s-nvidia-gpus_b200.dist")
    for n in (0, 1, 7, 8, 64, 65):
        for w in (1, 2, 4, 8):
            spans = [D.shard_range(n, w, r) for r in range(w)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = D.shard_sizes(n, w)
            assert max(sizes) - min(sizes) <= 1 and sizes == sorted(sizes, reverse=True)


def _gloo_worker(rank, world, port, n_prompts, out_dir):
    import importlib
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    sys.path.insert(0, str(ROOT))
    import torch.distributed as dist
    dist.init_process_group("gloo", rank=rank, world_size=world)
    D = importlib.import_module("k8s-nvidia-gpus_b200.dist")
    prompts = [f"prompt {'x' * i}" for i in range(n_prompts)]
    imgs = D.generate_sharded(FakePipe(), prompts, base_seed=1234, max_batch=2, num_inference_steps=2, width=16,
                              height=16)
    np.save(os.path.join(out_dir, f"r{rank}.npy"), imgs.numpy())
    dist.destroy_process_group()


def _gloo_cli_worker(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), LOCAL_RANK=str(rank),
                      WORLD_SIZE=str(world))
    sys.path.insert(0, str(ROOT))
    sys.path.insert(0, str(ROOT / "scripts"))
    import importlib
    cli = importlib.import_module("batch_generate")
    from pathlib import Path
    cli.generate_local(["a panda", "a cabin in the woods"], 3, Path(out_dir), "img", 5, batch_size=2, seed=40,
                       guidance_scale=7.5, width=16, height=16, scheduler=None, model="tiny", pipe=FakePipe())


def test_cli_local_mode_shards_under_torchrun_env(tmp_path):
    """`batch_generate.py --local` with WORLD_SIZE=2 (what torchrun sets): prompts sharded over the ranks through
    dist.generate_sharded, rank 0 writes the same files a single process writes."""
    import torch.multiprocessing as mp
    from PIL import Image
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    two, one = tmp_path / "w2", tmp_path / "w1"
    mp.spawn(_gloo_cli_worker, args=(2, port, str(two)), nprocs=2, join=True)
    sys.path.insert(0, str(ROOT / "scripts"))
    import importlib
    cli = importlib.import_module("batch_generate")
    for k in ("RANK", "LOCAL_RANK", "WORLD_SIZE"):
        os.environ.pop(k, None)
    cli.generate_local(["a panda", "a cabin in the woods"], 3, one, "img", 5, batch_size=2, seed=40,
                       guidance_scale=7.5, width=16, height=16, scheduler=None, model="tiny", pipe=FakePipe())
    names = sorted(p_.name for p_ in two.iterdir())
    assert names == [f"img_{i:02d}.png" for i in range(1, 6)] == sorted(p_.name for p_ in one.iterdir())
    for n in names:
        assert np.array_equal(np.asarray(Image.open(two / n)), np.asarray(Image.open(one / n)))


@pytest.mark.parametrize("n_prompts", [5, 1])
def test_sharded_generation_is_world_size_invariant(tmp_path, n_prompts):
    import torch.multiprocessing as mp
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    mp.spawn(_gloo_worker, args=(2, port, n_prompts, str(tmp_path)), nprocs=2, join=True)
    r0, r1 = np.load(tmp_path / "r0.npy"), np.load(tmp_path / "r1.npy")
    assert r0.shape == (n_prompts, 16, 16, 3) and np.array_equal(r0, r1)
    # single-process result (world 1) is identical: per-image seed = base + global index
    import importlib
    D = importlib.import_module("k8s-nvidia-gpus_b200.dist")
    prompts = [f"prompt {'x' * i}" for i in range(n_prompts)]
    one = D.generate_sharded(FakePipe(), prompts, base_seed=1234, max_batch=2, num_inference_steps=2, width=16, height=16)
    assert np.array_equal(one.numpy(), r0)


# ------------------------------------------------------------------------------------------------ robustness
def test_server_rejects_bad_sizes_and_steps_before_the_gpu(b200sd):
    from fastapi.testclient import TestClient
    import importlib
    server = importlib.import_module("k8s-nvidia-gpus_b200.server")
    fake = FakePipe()
    c = TestClient(server.create_app(pipe=fake))
    for body in ({"prompt": "x", "width": 60}, {"prompt": "x", "height": 100000}, {"prompt": "x", "steps": 5000},
                 {"prompt": "x", "guidance_scale": 1e9}):
        assert c.post("/generate", json=body).status_code == 400, body
    assert fake.calls == []                                   # nothing reached the pipeline
    assert c.post("/generate", json={"prompt": "x", "width": 16, "height": 16, "steps": 2}).status_code == 200


def test_scheduler_config_with_unsupported_options_is_refused(b200sd):
    C = b200sd.checkpoint
    base = {"_class_name": "PNDMScheduler", "beta_start": 0.00085, "beta_end": 0.012, "beta_schedule": "scaled_linear",
            "skip_prk_steps": True, "steps_offset": 1, "set_alpha_to_one": False}
    assert C.scheduler_config(base)["type"] == "pndm"
    for k, v in (("timestep_spacing", "trailing"), ("skip_prk_steps", False), ("set_alpha_to_one", True),
                 ("prediction_type", "v_prediction")):
        with pytest.raises(NotImplementedError):
            C.scheduler_config(dict(base, **{k: v}))
    ddim = dict(base, _class_name="DDIMScheduler")
    with pytest.raises(NotImplementedError):      # upstream DDIM default clip_sample=True must not load silently
        C.scheduler_config(ddim)
    assert C.scheduler_config(dict(ddim, clip_sample=False))["type"] == "ddim"
    with pytest.raises(NotImplementedError):
        C.scheduler_config(dict(base, _class_name="EulerDiscreteScheduler", use_karras_sigmas=True))


def test_from_pretrained_never_falls_back_to_random_weights_silently(b200sd, monkeypatch):
    """A hub id or a mistyped / unmounted path must raise, not serve noise (the pod's MODEL_ID default was a hub id).
    The check runs before any CUDA work, so it is testable here."""
    P = b200sd.pipeline
    monkeypatch.delenv("B200SD_RANDOM_WEIGHTS", raising=False)
    for bad in ("runwayml/stable-diffusion-v1-5", "/models/not-mounted"):
        with pytest.raises(FileNotFoundError, match="random_init"):
            P.StableDiffusionPipeline.from_pretrained(bad)


def test_integration_md_ctypes_stub_runs(b200sd):
    """The ctypes stub printed in INTEGRATION.md §4 is executed (up to the GPU call) against the built library: its
    struct mirror must match sizeof(b200sd_gemm_params)."""
    text = (ROOT / "INTEGRATION.md").read_text()
    sec = text[text.index("## 4."):]
    code = sec[sec.index("```python") + len("```python"):]
    code = code[:code.index("```")]
    head = code[:code.index("# --- GPU part")]
    ns = {}
    cwd = os.getcwd()
    os.chdir(ROOT)
    try:
        exec(compile(head, "INTEGRATION.md#4", "exec"), ns)
    finally:
        os.chdir(cwd)
    assert ns["ctypes"].sizeof(ns["GemmParams"]) == b200sd.lib.struct_sizes()[0]


def test_bench_exclusive_kernel_times():
    """bench.py's CUPTI pass: a kernel that became resident under its predecessor's tail (programmatic dependent launch)
    is counted from the predecessor's end; without overlap every kernel keeps its whole duration; times add up to the
    busy time of the stream."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("bench_mod", os.path.join(ROOT, "bench.py"))
    bench = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(bench)
    assert bench.exclusive_us([(0.0, 10.0), (12.0, 20.0), (20.0, 21.0)]) == [10.0, 8.0, 1.0]
    # second kernel resident from t = 6 but waiting until 10; third entirely inside the second (cannot happen, but
    # must not go negative)
    assert bench.exclusive_us([(0.0, 10.0), (6.0, 18.0), (7.0, 17.0)]) == [10.0, 8.0, 0.0]
    spans = [(0.0, 5.0), (4.0, 9.0), (8.5, 30.0), (31.0, 32.0)]
    assert sum(bench.exclusive_us(spans)) == 5.0 + 4.0 + 21.0 + 1.0
