"""Prompt-level data parallelism over the GPUs of one node.

The path shards by independent units (one prompt = one image): rank r of W takes a contiguous slice of the prompts
(remainder to the low ranks), holds the full weights and keeps each prompt's cond/uncond CFG pair local, so no
collective runs inside the denoise loop.  The only exchange is one gather of the uint8 images after VAE decode
(NCCL on GPUs; gloo in the CPU tests).  Image i always uses seed `base_seed + i` (global index; the per-output seed
convention of the reference's own video client, cluster-config/apps/llm/scripts/generate_wan_t2v.py:338-342), so the
result does not depend on W.  The reference itself scales out as W independent 1-GPU pods (README.md:301-349).
"""
from __future__ import annotations

import torch


def shard_range(n: int, world: int, rank: int) -> tuple:
    """[begin, end) of rank's contiguous slice of n items."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError(f"bad rank/world {rank}/{world}")
    base, rem = divmod(n, world)
    begin = rank * base + min(rank, rem)
    return begin, begin + base + (1 if rank < rem else 0)


def shard_sizes(n: int, world: int) -> list:
    return [shard_range(n, world, r)[1] - shard_range(n, world, r)[0] for r in range(world)]


def gather_images(local: torch.Tensor, counts: list, group=None) -> torch.Tensor | None:
    """local: uint8 [n_local, H, W, 3] on this rank -> uint8 [sum(counts), H, W, 3] on every rank (all_gather of
    equal-size padded shards; rank order = prompt order)."""
    import torch.distributed as dist
    world = dist.get_world_size(group)
    mx = max(counts)
    pad = local
    if local.shape[0] < mx:
        pad = torch.zeros((mx,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
        pad[: local.shape[0]] = local
    out = torch.empty((world * mx,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    dist.all_gather_into_tensor(out, pad.contiguous(), group=group)
    out = out.view((world, mx) + tuple(local.shape[1:]))
    return torch.cat([out[r, : counts[r]] for r in range(world)], dim=0)


def generate_sharded(pipe, prompts: list, *, base_seed: int = 1234, max_batch: int = 8, group=None, **call_kwargs):
    """Run `pipe` on this rank's slice of `prompts` (in chunks of max_batch) and gather all images on every rank.
    Returns uint8 [len(prompts), H, W, 3] (torch tensor on the pipe's device, or CPU for CPU fakes)."""
    import torch.distributed as dist
    world = dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1
    rank = dist.get_rank(group) if world > 1 else 0
    b, e = shard_range(len(prompts), world, rank)
    chunks = []
    for s in range(b, e, max_batch):
        idx = list(range(s, min(e, s + max_batch)))
        gens = [torch.Generator().manual_seed(base_seed + i) for i in idx]
        out = pipe([prompts[i] for i in idx], generator=gens, output_type="u8", **call_kwargs)
        chunks.append(torch.as_tensor(out.images))
    dev = getattr(pipe, "device", torch.device("cpu"))
    if chunks:
        local = torch.cat(chunks, dim=0).to(dev)
    else:
        local = None
    if world == 1:
        return local
    counts = shard_sizes(len(prompts), world)
    if local is None:  # a rank with no prompts still takes part in the collective
        local = torch.zeros((0, 1, 1, 3), dtype=torch.uint8, device=dev)
    # agree on the image shape (ranks with an empty shard do not know it)
    hw = torch.tensor(list(local.shape[1:3]), dtype=torch.int64, device=dev)
    dist.all_reduce(hw, op=dist.ReduceOp.MAX, group=group)
    if local.shape[0] == 0:
        local = torch.zeros((0, int(hw[0]), int(hw[1]), 3), dtype=torch.uint8, device=dev)
    return gather_images(local, counts, group)
